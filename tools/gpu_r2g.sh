#!/bin/bash
# round-2 GPU call G (1 GPU): the record set — full -m gpu suite, bench + reference arm, launch list, ncu --set full of the three big kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2g}
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/${T}_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_gpu.log
tail -6 gpurun_out/${T}_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err; echo "ref rc=$?"
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_list.log 2>&1; echo "launch list rc=$?"
# rays of the bounce-1 launch of the first timed batch (frames 768..895): depth-2 minus depth-1 extend rays
python - > gpurun_out/${T}_rays.txt <<'PY'
import bench
from raytracer_b200 import librt
scene, cam = bench.make_scene("sponza-proxy", 1920, 1080)
out = []
for d in (1, 2):
    c = librt.Context(device=0, counters=True, max_depth=d, rr_start_depth=3, frames_in_flight=256)
    c.build_scene(scene); c.resize(1920, 1080)
    c.trace_frames(cam.ubo(), bench.CLEAR, 768, 64)  # one batch of the bench: 64 frames per stream share
    out.append(c.counters()["extend_rays"]); c.close()
print(out[1] - out[0])
PY
cat gpurun_out/${T}_rays.txt
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_extend_coop_kernel -s 1 -c 1 -f -o gpurun_out/${T}_coop python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_coop.log 2>&1; echo "ncu coop rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_extend_packet_kernel -c 1 -f -o gpurun_out/${T}_packet python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_packet.log 2>&1; echo "ncu packet rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_shade_kernel -c 2 -f -o gpurun_out/${T}_shade python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/${T}_ncu_shade.log 2>&1; echo "ncu shade rc=$?"
ls -la gpurun_out/*.ncu-rep
