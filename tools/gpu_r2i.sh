#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/r2i_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2i_gpu.log
tail -n 8 gpurun_out/r2i_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
j=json.loads(open("gpurun_out/r2i_bench.json").read().strip().splitlines()[-1])
print({k:j[k] for k in ("value","ms_per_step","ms_per_frame","gpu_launches")}, j["e2e"]["value"], j["verify"])
PY
for r in 0 3 7; do timeout 200 python tools/strip_rank_time.py --world 8 --rank $r --steps 16 | tail -n 1; done
