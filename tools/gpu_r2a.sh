#!/bin/bash
# round-2 GPU call A: full -m gpu suite, gather peaks, bench, launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_gpus.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/r2a_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_gpu.log
tail -30 gpurun_out/r2a_gpu.log
timeout 120 tools/mb/gather gpurun_out/gather_peaks.json > gpurun_out/r2a_gather.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r2a_bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err; echo "ref rc=$?"
