#!/usr/bin/env python
"""BVH build time: first build and rebuilds after rt_update_object_transform (the editor's drag), per scene.

    python tools/build_time.py [--scene sponza-proxy|soup1m|cornell_pbr] [--rebuilds 5]
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="sponza-proxy")
    ap.add_argument("--rebuilds", type=int, default=5)
    a = ap.parse_args()
    import bench
    from raytracer_b200 import librt

    scene, cam = bench.make_scene(a.scene, 256, 256)
    ctx = librt.Context(device=0, counters=False)
    t0 = time.perf_counter()
    info = ctx.build_scene(scene)
    ctx.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    print(f"{a.scene}: {info.n_triangles} triangles, {info.n_nodes} nodes; first rt_build_scene {wall:.2f} ms wall, "
          f"BVH build {info.build_ms:.2f} ms on the stream = {info.n_triangles / info.build_ms / 1e3:.1f} Mtris/s")
    ctx.resize(256, 256)
    m = np.asarray(scene.objects[0].transform.model_matrix, np.float32).copy()
    for k in range(a.rebuilds):
        m2 = m.copy()
        m2.reshape(-1)[12] += 0.01 * (k + 1)  # translate x (column-major)
        t0 = time.perf_counter()
        ctx.update_object_transform(0, m2)
        ctx.trace_frame(cam.ubo(), bench.CLEAR, 0)  # flushes the scene: tables upload + BVH rebuild, then one frame
        ctx.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        bi = ctx.scene_info()
        print(f"  update {k}: rt_update_object_transform + flush (refit #{bi.refits}) + one 256x256 frame {wall:.2f} ms wall "
              f"(the last full build took {bi.build_ms:.2f} ms on the stream)")
    for k in range(2):   # and a full rebuild of the same scene for comparison
        t0 = time.perf_counter()
        bi = ctx.build_scene(scene)
        ctx.trace_frame(cam.ubo(), bench.CLEAR, 0)
        ctx.synchronize()
        print(f"  rt_build_scene {k}: upload + build + one frame {(time.perf_counter() - t0) * 1e3:.2f} ms wall; BVH build {bi.build_ms:.2f} ms "
              f"= {bi.n_triangles / bi.build_ms / 1e3:.1f} Mtris/s")
    ctx.close()


if __name__ == "__main__":
    main()
