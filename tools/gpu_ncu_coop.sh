#!/bin/bash
# ncu --set full of the bounce-1 wf_extend_coop_kernel launch of the first timed batch (1 GPU).  usage: gpu_ncu_coop.sh <tag>
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-x}
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_extend_coop_kernel -s 1 -c 1 -f -o gpurun_out/${T}_coop python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_coop.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/${T}_coop.ncu-rep
