#!/usr/bin/env python
"""Time ONE rank's share of the N-GPU headline step on a single GPU (no NCCL): strip_count = N, strip_index = R.

    python tools/strip_rank_time.py --world 8 [--rank 0] [--steps 16]      (RT_LIBRT_PATH selects a librt variant)

The per-frame path has no collective, so this is what rank R's GPU does in `bench.py --gpus N`; it lets a variant's effect on
the small per-rank images of N = 8 be measured without an 8-GPU box.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--spp", type=int, default=256)
    ap.add_argument("--strip-rows", type=int, default=8)
    a = ap.parse_args()
    import bench
    from raytracer_b200 import librt

    W, H = 1920, 1080
    scene, cam = bench.make_scene("sponza-proxy", W, H)
    ctx = librt.Context(device=0, counters=False, max_depth=8, rr_start_depth=3, strip_rows=a.strip_rows, strip_index=a.rank,
                        strip_count=a.world, frames_in_flight=a.spp)
    ctx.build_scene(scene)
    ctx.resize(W, H)
    ubo = cam.ubo()
    for k in range(a.warmup):
        ctx.trace_frames(ubo, bench.CLEAR, k * a.spp, a.spp)
    ctx.synchronize()
    ctx.event_record(0)
    for k in range(a.steps):
        ctx.trace_frames(ubo, bench.CLEAR, (a.warmup + k) * a.spp, a.spp)
    ctx.event_record(1)
    ctx.synchronize()
    ms = ctx.event_elapsed_ms() / a.steps
    print(f"world {a.world} rank {a.rank}: {ms:.3f} ms/step, {ms / a.spp:.5f} ms/frame; x world = {ms / a.spp * a.world:.4f} "
          f"(single-GPU frame time this would correspond to)")


if __name__ == "__main__":
    main()
