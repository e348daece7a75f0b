#!/bin/bash
# ncu --set full of one launch of a kernel inside the bench (1 GPU): gpu_ncu_k.sh <tag> <kernel regex> [skip]
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=$1; K=$2; S=${3:-0}
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k "regex:$K" -s $S -c 1 -f -o gpurun_out/${T} python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/${T}.ncu-rep
