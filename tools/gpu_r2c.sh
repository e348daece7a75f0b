#!/bin/bash
# round-2 GPU call C (N GPUs): scaling of the torchrun and in-process paths, balanced vs round-robin strips
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-8}
TAG=${2:-r2c}
show() { python - "$1" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], {k:j[k] for k in ("value","ms_per_step","n_gpus")}, "e2e", round(j["e2e"]["value"],1), round(j["e2e"]["ms_per_step"],3), j.get("ranks"), (j.get("verify") or {}).get("equal"))
except Exception as e: print("no json", sys.argv[1], e)
PY
}
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_host.py -m gpu -q -p no:cacheprovider 2>&1 | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 16 --warmup 3 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "torchrun bench rc=$?"
tail -c 600 gpurun_out/${TAG}_bench_n$N.err
show gpurun_out/${TAG}_bench_n$N.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 16 --warmup 3 --balance round_robin --no-verify > gpurun_out/${TAG}_bench_rr_n$N.json 2> gpurun_out/${TAG}_bench_rr_n$N.err; echo "torchrun rr rc=$?"
show gpurun_out/${TAG}_bench_rr_n$N.json
timeout 600 python bench.py --gpus $N --inproc --steps 16 --warmup 3 > gpurun_out/${TAG}_inproc_n$N.json 2> gpurun_out/${TAG}_inproc_n$N.err; echo "inproc rc=$?"
tail -c 600 gpurun_out/${TAG}_inproc_n$N.err
show gpurun_out/${TAG}_inproc_n$N.json
