#!/bin/bash
# quick A/B without the test suite: gpu_ab2.sh <tag>  (RT_FILTER=0 vs default, twice) + launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-ab2}
timeout 600 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -p no:cacheprovider -k "filter or config3" 2>&1 | tail -n 3
for round in 1 2; do
for v in RT_FILTER=0 -; do
  if [ "$v" = "-" ]; then pre=""; tag=on; else pre="env $v"; tag=off; fi
  timeout 600 $pre python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_${tag}_${round}.json 2> gpurun_out/${T}_${tag}_${round}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_${tag}_${round}.json").read().strip().splitlines()[-1])
print("[$v]", "round $round", "value %.1f  ms/frame %.4f  primary %.0f  incoherent %.1f  e2e %.1f" % (d["value"], d["ms_per_frame"], d["primary_mrays_s"], d["incoherent_mrays_s"], d["e2e"]["value"]), d["roofline"]["kernel_ms"])
PY
done
done
bash tools/gpu_list.sh ${T}_list
