#!/bin/bash
# quick A/B loop: parity subset + bench (1 GPU).  usage: gpu_quick.sh <tag> [pytest -k expression]
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
tag=${1:-q}; kexpr=${2:-}
if [ -n "$kexpr" ]; then timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "$kexpr" 2>&1 | tail -n 4
else timeout 1500 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -n 4; fi
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_frame","primary_mrays_s","incoherent_mrays_s")}, d["e2e"]["value"], d["roofline"])
print(d.get("kernel_ms") or d.get("breakdown") or "")
PY
