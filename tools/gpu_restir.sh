#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "restir" -p no:cacheprovider 2>&1 | tail -5
timeout 300 python tools/profile_frame.py --scene conference --restir --frames 32 --depth 3 --counters --repeat 2 2>&1 | tail -2
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:restir -c 30 --csv --log-file gpurun_out/restir_launches.csv python tools/profile_frame.py --scene conference --restir --frames 10 --depth 3 > gpurun_out/restir_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/restir_launches.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); mi=hdr.index("Metric Name"); vi=hdr.index("Metric Value")
from collections import defaultdict
d=defaultdict(list)
for r in rows[1:]:
    if r[mi]=="gpu__time_duration.sum": d[r[ki][:40]].append(float(r[vi].replace(",","")))
for k,v in d.items(): print(k, len(v), "avg us", sum(v)/len(v)/1e3)
PY
