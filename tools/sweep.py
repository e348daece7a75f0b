#!/usr/bin/env python
"""synthetic triangle-soup sweep, N in {1e4, 1e5, 1e6, 1e7} — BVH-build Mtris/s and traversal
scaling on one B200 (1920x1080, 16 spp, depth 3, sky only).

    python tools/sweep.py [--sizes 10000 100000 1000000 10000000] [--out gpurun_out/sweep.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CLEAR = (0.53, 0.81, 0.92)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", type=int, nargs="+", default=[10_000, 100_000, 1_000_000, 10_000_000])
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    a = ap.parse_args()
    import raytracer_b200 as rb
    from raytracer_b200 import librt, proxy_scenes

    W, H, spp, depth = 1920, 1080, 16, 3
    cam = rb.Camera(aspect=W / H, **proxy_scenes.SOUP_CAMERA)
    rows = []
    for n in a.sizes:
        scene = proxy_scenes.soup(n)
        cctx = librt.Context(device=0, max_depth=depth, counters=True, per_ray_primary=True, frames_in_flight=16)
        cctx.build_scene(scene)
        cctx.resize(W, H)
        cctx.trace_frames(cam.ubo(), CLEAR, 0, spp)
        c = cctx.counters()
        cctx.close()
        ctx = librt.Context(device=0, max_depth=depth, counters=False, frames_in_flight=16)
        info = ctx.build_scene(scene)
        first_ms = info.build_ms
        m = np.asarray(scene.objects[0].transform.model_matrix, np.float32).copy()
        rebuilds = []
        ctx.resize(W, H)
        for k in range(3):  # rebuilds with the scratch pool warm: what an edit costs
            m2 = m.copy()
            m2.reshape(-1)[12] += 1e-3 * (k + 1)
            ctx.update_object_transform(0, m2)
            ctx.trace_frame(cam.ubo(), CLEAR, 0)
            ctx.synchronize()
            rebuilds.append(ctx.scene_info().build_ms)
        ctx.update_object_transform(0, m)
        ctx.trace_frames(cam.ubo(), CLEAR, 0, spp)
        ctx.event_record(0)
        ctx.trace_frames(cam.ubo(), CLEAR, 0, spp)
        ctx.event_record(1)
        ms = ctx.event_elapsed_ms()
        ctx.close()
        rays = c["extend_rays"] + c["shadow_rays"]
        row = {"triangles": n, "bvh8_nodes": info.n_nodes, "bvh_depth": info.bvh_depth, "first_build_ms": round(first_ms, 3),
               "rebuild_ms": round(min(rebuilds), 3), "rebuild_mtris_s": round(n / min(rebuilds) / 1e3, 1),
               "ms_per_frame": round(ms / spp, 4), "mrays_s": round(rays / ms / 1e3, 1),
               "nodes_per_ray": round(c["nodes"] / max(rays, 1), 2), "tris_per_ray": round(c["tris"] / max(rays, 1), 2)}
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    json.dump(rows, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
