#!/bin/bash
# round-2 final record run (1 GPU): full -m gpu suite, smoke, bench + reference arm, configs table
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/r2k_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2k_gpu.log
tail -n 6 gpurun_out/r2k_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2k_bench.json 2> gpurun_out/r2k_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/r2k_ref.json 2> gpurun_out/r2k_ref.err; echo "ref rc=$?"
timeout 900 python tools/run_configs.py --out gpurun_out/configs_r2k > gpurun_out/configs_r2k.log 2>&1; tail -n 9 gpurun_out/configs_r2k.log | cut -c1-160
