#!/bin/bash
# launch list of the bench (1 GPU): gpu_list.sh <tag> [ENV=val ...]
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=$1; shift
env "$@" timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_list.log 2>&1; echo "launch list rc=$?"
python tools/launch_summary.py gpurun_out/${T}_launches.csv
