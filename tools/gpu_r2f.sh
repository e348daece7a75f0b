#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_host.py tests/test_gpu_parity.py -m gpu -q -k "refit or incremental or update or restir" -p no:cacheprovider 2>&1 | tail -15
timeout 120 python tools/build_time.py --scene sponza-proxy 2>&1 | tail -9
timeout 120 python tools/build_time.py --scene soup1m --rebuilds 3 2>&1 | tail -7
timeout 600 ncu --set full --clock-control none --import-source on -k regex:restir_ -s 6 -c 3 -o gpurun_out/r2f_restir -f python tools/profile_frame.py --scene conference --restir --frames 4 --depth 3 > gpurun_out/r2f_ncu.log 2>&1; tail -3 gpurun_out/r2f_ncu.log
