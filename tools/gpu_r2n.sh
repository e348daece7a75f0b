#!/bin/bash
# round-2 GPU call N (1 GPU): A/B of the per-group primary slots in the packet kernel, the tail kernel on an eighth share of the
# frame (what one rank of N = 8 traces), then the record set of the winning build: full -m gpu suite, smoke, bench + reference
# arm, launch list, ncu --set full of the packet kernel.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2n}
V=$PWD/raytracer_b200/csrc/variants
for round in 1 2; do
for v in default noslot; do
  if [ "$v" = default ]; then unset RT_LIBRT_PATH; else export RT_LIBRT_PATH=$V/librt_$v.so; fi
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ab_${v}_${round}.json 2> gpurun_out/${T}_ab_${v}_${round}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_ab_${v}_${round}.json").read().strip().splitlines()[-1])
print("$v", "round $round", "value %.1f  ms/frame %.4f  primary %.0f  incoherent %.1f  e2e %.1f" % (d["value"], d["ms_per_frame"], d["primary_mrays_s"], d["incoherent_mrays_s"], d["e2e"]["value"]))
PY
done
done
unset RT_LIBRT_PATH
WIN=$(python - <<PY
import json
def val(v):
    return sum(json.loads(open(f"gpurun_out/${T}_ab_{v}_{r}.json").read().strip().splitlines()[-1])["value"] for r in (1, 2)) / 2
print("default" if val("default") >= val("noslot") else "noslot")
PY
)
echo "winner: $WIN" | tee gpurun_out/${T}_winner.txt
if [ "$WIN" != default ]; then export RT_LIBRT_PATH=$V/librt_$WIN.so; fi
# the tail kernel where it is aimed: rank 0's share of the N = 8 step, and the whole frame at N = 1
for round in 1 2; do
  for t in 0 x; do
    if [ $t = 0 ]; then export RT_TAIL=0; else unset RT_TAIL; fi
    echo "RT_TAIL=${t} round $round: $(timeout 300 python tools/strip_rank_time.py --world 8 --rank 0 --steps 16 2>&1 | tail -n 1)"
  done
done 2>&1 | tee gpurun_out/${T}_tail_share.txt
unset RT_TAIL
for v in tail2 tail4; do
  echo "$v: $(RT_LIBRT_PATH=$V/librt_$v.so timeout 300 python tools/strip_rank_time.py --world 8 --rank 0 --steps 16 2>&1 | tail -n 1)"
done 2>&1 | tee -a gpurun_out/${T}_tail_share.txt
# ---- record set of the winning build
timeout 1000 python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/${T}_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_gpu.log
tail -n 6 gpurun_out/${T}_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -n 1 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err; echo "ref rc=$?"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_list.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_extend_packet_kernel -c 1 -f -o gpurun_out/${T}_packet python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_packet.log 2>&1; echo "ncu packet rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_frame","primary_mrays_s","incoherent_mrays_s","gpu_launches")}, d["e2e"]["value"], d["roofline"]["frac"], d["clocks"])
PY
