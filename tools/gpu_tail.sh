#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -p no:cacheprovider -k "tail" 2>&1 | tail -n 6
bash tools/gpu_env_ab.sh tail RT_TAIL=0 - RT_TAIL=1
