#!/bin/bash
# A/B of librt variants on one GPU: gpu_ab.sh <tag> <variant> [<variant> ...]   ("default" = the in-tree librt.so)
# each: parity subset (-k "parity or configs") once for the last variant only if AB_TEST=1; bench 8 steps
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
tag=$1; shift
for round in 1 2; do
for v in "$@"; do
  if [ "$v" = default ]; then unset RT_LIBRT_PATH; else export RT_LIBRT_PATH=$PWD/raytracer_b200/csrc/variants/librt_$v.so; fi
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${tag}_${v}_${round}.json 2> gpurun_out/${tag}_${v}_${round}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_${v}_${round}.json").read().strip().splitlines()[-1])
print("$v", "round $round", "value %.1f  ms/frame %.4f  primary %.0f  incoherent %.1f  e2e %.1f" % (d["value"], d["ms_per_frame"], d["primary_mrays_s"], d["incoherent_mrays_s"], d["e2e"]["value"]))
PY
done
done
