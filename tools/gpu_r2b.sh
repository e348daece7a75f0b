#!/bin/bash
# round-2 GPU call B (2 GPUs): multi-GPU tests, torchrun bench N=2, in-process bench N=2
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi -L > gpurun_out/r2b_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_configs.py -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/r2b_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_gpu.log
tail -30 gpurun_out/r2b_gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/r2b_bench_n$N.json 2> gpurun_out/r2b_bench_n$N.err; echo "torchrun bench rc=$?"
tail -c 1500 gpurun_out/r2b_bench_n$N.err
python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/r2b_bench_n$N.json").read().strip().splitlines()[-1])
    print({k:j[k] for k in ("value","ms_per_step","n_gpus","gpu_launches")}, j["e2e"], j.get("ranks"), j.get("verify"))
except Exception as e: print("no json", e)
PY
timeout 600 python bench.py --gpus $N --inproc --steps 8 --warmup 3 > gpurun_out/r2b_inproc_n$N.json 2> gpurun_out/r2b_inproc_n$N.err; echo "inproc bench rc=$?"
tail -c 800 gpurun_out/r2b_inproc_n$N.err
python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/r2b_inproc_n$N.json").read().strip().splitlines()[-1])
    print({k:j[k] for k in ("value","ms_per_step","n_gpus","gpu_launches")}, j["e2e"], j.get("ranks"), j.get("verify"))
except Exception as e: print("no json", e)
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 8 --warmup 3 --balance round_robin --no-verify > gpurun_out/r2b_bench_rr_n$N.json 2> gpurun_out/r2b_bench_rr_n$N.err; echo "torchrun rr bench rc=$?"
python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/r2b_bench_rr_n$N.json").read().strip().splitlines()[-1])
    print("round robin:", {k:j[k] for k in ("value","ms_per_step","n_gpus")}, j["e2e"], j.get("ranks"))
except Exception as e: print("no json", e)
PY
