#!/usr/bin/env python
"""CPU estimate for DESIGN.md 10's "doomed rays" lever: traversal work of an any-hit walk against a closest-hit walk for rays
like those of bounce 1 on the sponza-proxy (from the first hits of a 480x270 frame, mirror direction + a GGX-like perturbation),
counted by the oracle's traversal of the same BVH8.  python tools/anyhit_estimate.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_b200 as rb  # noqa: E402
from oracle.binding import OracleScene  # noqa: E402
from raytracer_b200 import proxy_scenes  # noqa: E402


def main():
    s = proxy_scenes.sponza_proxy()
    W, H = 480, 270
    cam = rb.Camera(aspect=W / H, **proxy_scenes.SPONZA_CAMERA)
    orc = OracleScene(s).build_bvh()
    g = orc.gbuffer(cam.ubo(), W, H)
    pos = g["position_metal"][..., :3].reshape(-1, 3).astype(np.float32)
    nrm = g["normal_rough"].view(np.float16).astype(np.float32).reshape(-1, 4)[:, :3]
    o = np.asarray(cam.ubo(), np.float32)[32:48].reshape(4, 4)[3, :3]
    d = pos - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    n = nrm / np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-9)
    n[(n * d).sum(1) > 0] *= -1
    r = d - 2 * (d * n).sum(1, keepdims=True) * n
    r = r + 0.35 * np.random.default_rng(1).normal(size=r.shape).astype(np.float32)
    r /= np.linalg.norm(r, axis=1, keepdims=True)
    ok = (r * n).sum(1) > 0.05
    rays = np.zeros((int(ok.sum()), 8), np.float32)
    rays[:, 0:3], rays[:, 3], rays[:, 4:7], rays[:, 7] = pos[ok], 0.001, r[ok], 1e4
    c1, c2 = np.zeros(6, np.uint64), np.zeros(6, np.uint64)
    _, ids = orc.trace_rays(rays, counters=c1)
    orc.trace_rays(rays, any_hit=True, ignore_inst=np.full(len(rays), 0xFFFFFFFF, np.uint32), counters=c2)
    k = len(rays)
    print(f"{k} rays, {100 * (ids != 0xFFFFFFFF).mean():.1f} % hit")
    for name, c in (("closest-hit", c1), ("any-hit", c2)):
        print(f"{name:12s} {c[2] / k:6.2f} nodes/ray {c[3] / k:5.2f} tris/ray {(4 * c[2] + 2 * c[3]) / k:6.1f} thread-loads/ray")


if __name__ == "__main__":
    main()
