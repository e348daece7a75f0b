#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "raster or c_host or restir_image or host_decided" -p no:cacheprovider -s 2>&1 | tail -25
timeout 300 python tools/fit_balance.py --out gpurun_out/fit_balance.json > gpurun_out/fit_balance.log 2>&1; tail -20 gpurun_out/fit_balance.log
timeout 900 python tools/run_configs.py --out gpurun_out/configs_r2 > gpurun_out/configs_r2.log 2>&1; tail -12 gpurun_out/configs_r2.log
