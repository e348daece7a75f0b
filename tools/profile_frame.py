#!/usr/bin/env python
"""Render a few frames of one workload through the C ABI — the short command ncu wraps.

    python tools/profile_frame.py [--scene sponza-proxy] [--frames 2] [--batch 1] [--depth 8] [--pipeline wavefront]

Launch order per frame (wavefront, depth D): raygen, D x (reset, extend, shade, shadow), accumulate, so with
`--frames 2 --batch 1` the extend launches 0 and D are the depth-0 launches of frames 0 and 1.
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="sponza-proxy")
    ap.add_argument("--frames", type=int, default=2)
    ap.add_argument("--batch", type=int, default=1, help="frames per rt_trace_frames call")
    ap.add_argument("--depth", type=int, default=8)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--pipeline", default="wavefront")
    ap.add_argument("--first-frame", type=int, default=1)
    ap.add_argument("--counters", action="store_true")
    ap.add_argument("--repeat", type=int, default=1)
    ap.add_argument("--kprof", action="store_true", help="per-kernel CUDA-event times (serialises launches)")
    ap.add_argument("--lanes", type=int, default=0, help="frames_in_flight (0 = library default)")
    ap.add_argument("--per-ray", action="store_true", help="RT_FLAG_PER_RAY_PRIMARY: no primary-ray packets")
    ap.add_argument("--restir", action="store_true", help="rt_restir_frame (ReSTIR DI) instead of rt_trace_frames")
    a = ap.parse_args()
    import bench
    from raytracer_b200 import librt

    if a.scene in ("sponza-proxy", "soup1m"):
        scene, cam = bench.make_scene(a.scene, a.width, a.height)
    else:  # a fixture scene of the configs table (cornell_box, cornell_pbr, conference_mirror_chairs, multi_dispersion ...)
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        import run_configs

        scene, cam = run_configs.load(a.scene), run_configs.camera_for(a.scene, a.width / a.height)
    pipe = librt.RT_PIPELINE_WAVEFRONT if a.pipeline == "wavefront" else librt.RT_PIPELINE_MEGAKERNEL
    ctx = librt.Context(counters=a.counters, device=0, max_depth=a.depth, rr_start_depth=3, pipeline=pipe, frames_in_flight=a.lanes,
                        per_ray_primary=a.per_ray)
    info = ctx.build_scene(scene)
    ctx.resize(a.width, a.height)
    ubo = cam.ubo()
    for rep in range(a.repeat):
        if a.kprof and rep == a.repeat - 1:
            ctx.set_profiling(True)
        ctx.reset_counters()
        ctx.event_record(0)
        f = a.first_frame
        while f < a.first_frame + a.frames:
            n = 1 if a.restir else min(a.batch, a.first_frame + a.frames - f)
            if a.restir:
                ctx.restir_frame(ubo, bench.CLEAR, f)
            else:
                ctx.trace_frames(ubo, bench.CLEAR, f, n)
            f += n
        ctx.event_record(1)
        ms = ctx.event_elapsed_ms()
        c = ctx.counters()
        rays = c["extend_rays"] + c["shadow_rays"]
        print(f"rep {rep}: {a.frames} frames in {ms:.3f} ms = {ms / a.frames:.3f} ms/frame; tris {info.n_triangles} nodes {info.n_nodes}; "
              f"launches {c['launches']}" + (f"; rays {rays} = {rays / ms / 1e3:.1f} Mrays/s; nodes/ray "
              f"{c['nodes'] / max(rays, 1):.2f} tris/ray {c['tris'] / max(rays, 1):.2f}" if a.counters else ""), flush=True)
    if a.kprof:
        kp = ctx.kernel_profile()
        print("   per-kernel ms/frame: " + ", ".join(f"{k} {v['ms'] / a.frames:.3f} ({v['launches']})" for k, v in kp.items()), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
