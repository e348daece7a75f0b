#!/bin/bash
# N-GPU A/B of the tail kernel under torchrun: gpu_tail8.sh <N>
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-8}
for t in 0 x; do
  if [ $t = 0 ]; then export RT_TAIL=0; else unset RT_TAIL; fi
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/tail${N}_$t.json 2> gpurun_out/tail${N}_$t.err
  python - "$t" "$N" <<'PY'
import json,sys
t,n=sys.argv[1],sys.argv[2]
j=json.loads(open(f"gpurun_out/tail{n}_{t}.json").read().strip().splitlines()[-1])
print("RT_TAIL=0" if t=="0" else "default", round(j["value"],1), round(j["ms_per_step"],3), "e2e", round(j["e2e"]["value"],1), j["ranks"]["slowest_over_mean"], (j.get("verify") or {}).get("equal"))
PY
done
