#!/bin/bash
# round-2 GPU call U (1 GPU): record set of the build with the first-hit filter — full -m gpu suite, smoke, bench + reference arm,
# launch list, ncu --set full of the cooperative extend kernel (bounce-1 launch) and of the first shade launch
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2u}
timeout 1000 python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/${T}_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_gpu.log
tail -n 6 gpurun_out/${T}_gpu.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -n 1 gpurun_out/${T}_smoke.log | cut -c1-200
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err; echo "ref rc=$?"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_list.log 2>&1; echo "launch list rc=$?"
bash tools/gpu_ncu_k.sh ${T}_coop wf_extend_coop_kernel 1
bash tools/gpu_ncu_k.sh ${T}_shade wf_shade_kernel 0
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_frame","primary_mrays_s","incoherent_mrays_s","gpu_launches")}, d["e2e"]["value"], d["roofline"]["frac"], d["clocks"])
PY
