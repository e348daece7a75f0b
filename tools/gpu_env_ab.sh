#!/bin/bash
# A/B of environment settings on one GPU: gpu_env_ab.sh <tag> "<VAR=val ...>" ["<VAR=val ...>" ...]   ("-" = no setting)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
tag=$1; shift
for round in 1 2; do
i=0
for v in "$@"; do
  i=$((i+1))
  if [ "$v" = "-" ]; then pre=""; else pre="env $v"; fi
  timeout 600 $pre python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${tag}_${i}_${round}.json 2> gpurun_out/${tag}_${i}_${round}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_${i}_${round}.json").read().strip().splitlines()[-1])
print("[$v]", "round $round", "value %.1f  ms/frame %.4f  primary %.0f  incoherent %.1f  e2e %.1f" % (d["value"], d["ms_per_frame"], d["primary_mrays_s"], d["incoherent_mrays_s"], d["e2e"]["value"]))
PY
done
done
