#!/bin/bash
# one point of the scaling table for the r2u build: torchrun bench at N GPUs (as the driver launches it)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2u_bench_n$N.json 2> gpurun_out/r2u_bench_n$N.err; echo "bench rc=$?"
python - "$N" <<'PY'
import json,sys
n=sys.argv[1]
j=json.loads(open(f"gpurun_out/r2u_bench_n{n}.json").read().strip().splitlines()[-1])
print({k:j[k] for k in ("value","ms_per_step","n_gpus")}, "e2e", round(j["e2e"]["value"],1), j.get("ranks"), j.get("verify"))
PY
