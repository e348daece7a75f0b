#!/bin/bash
# round-2 GPU call O (1 GPU): the first-hit filter — full -m gpu suite, then A/B RT_FILTER=0 vs default on the bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2o}
timeout 1000 python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/${T}_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_gpu.log
tail -n 25 gpurun_out/${T}_gpu.log | cut -c1-300
for round in 1 2; do
for v in RT_FILTER=0 -; do
  if [ "$v" = "-" ]; then pre=""; tag=on; else pre="env $v"; tag=off; fi
  timeout 600 $pre python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_${tag}_${round}.json 2> gpurun_out/${T}_${tag}_${round}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_${tag}_${round}.json").read().strip().splitlines()[-1])
print("[$v]", "round $round", "value %.1f  ms/frame %.4f  primary %.0f  incoherent %.1f  e2e %.1f" % (d["value"], d["ms_per_frame"], d["primary_mrays_s"], d["incoherent_mrays_s"], d["e2e"]["value"]), d["roofline"]["kernel_ms"], d["gpu_launches"])
PY
done
done
bash tools/gpu_list.sh ${T}_list
