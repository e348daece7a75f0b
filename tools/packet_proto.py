#!/usr/bin/env python
"""CPU estimate of what a shared-stack ("packet") traversal of a 4x8 primary-ray tile would visit, against the
per-ray traversal the extend kernel does today.  float64 numpy, statistics only (no parity claim).

    python tools/packet_proto.py [--tiles 400] [--tw 4 --th 8]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def camera_rays(ubo, W, H, xs, ys):
    inv_view = ubo[32:48].reshape(4, 4).T.astype(np.float64)
    inv_proj = ubo[48:64].reshape(4, 4).T.astype(np.float64)
    u = (xs + 0.5) / W * 2 - 1
    v = (ys + 0.5) / H * 2 - 1
    tgt = (inv_proj @ np.stack([u, v, np.ones_like(u), np.ones_like(u)])).T[:, :3]
    tgt /= np.linalg.norm(tgt, axis=1, keepdims=True)
    d = (inv_view[:3, :3] @ tgt.T).T
    o = inv_view[:3, 3]
    return o, d


def decode(nodes):
    n = nodes.shape[0]
    p = nodes[:, 0:12].copy().view(np.float32).reshape(n, 3).astype(np.float64)
    e = nodes[:, 12:15].astype(np.int32)
    imask = nodes[:, 15]
    child_base = nodes[:, 16:20].copy().view(np.uint32).reshape(n)
    tri_base = nodes[:, 20:24].copy().view(np.uint32).reshape(n)
    meta = nodes[:, 24:32]
    q = nodes[:, 32:128].copy().view(np.float16).reshape(n, 3, 2, 8).astype(np.float64)
    scale = np.ldexp(1.0, e - 127)  # u2f(e << 23)
    lo = p[:, :, None] + q[:, :, 0, :] * scale[:, :, None]
    hi = p[:, :, None] + q[:, :, 1, :] * scale[:, :, None]
    return dict(lo=lo, hi=hi, imask=imask, child_base=child_base, tri_base=tri_base, meta=meta)


def tri_hits(o, d, t):
    """Moeller-Trumbore for all rays against one triangle; returns t (inf when missed)."""
    p0, p1, p2 = t
    e1, e2 = p1 - p0, p2 - p0
    pv = np.cross(d, e2)
    det = pv @ e1
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / det
        tv = o - p0
        u = (pv @ tv) * inv
        qv = np.cross(tv, e1)
        v = (d @ qv) * inv
        tt = (qv @ e2) * inv
    ok = (np.abs(det) > 0) & (u >= 0) & (v >= 0) & (u + v <= 1) & (tt > 1e-3)
    return np.where(ok, tt, np.inf)


def packet_traverse(B, tris, o, d):
    idir = 1.0 / d
    sgn = d < 0
    if np.any(sgn != sgn[0]):
        return None
    s = sgn[0]
    idlo, idhi = idir.min(0), idir.max(0)
    octant = int(s[0]) | int(s[1]) << 1 | int(s[2]) << 2
    best = np.full(d.shape[0], 1e4)
    stack = [0]
    n_nodes = n_tris = 0
    while stack:
        node = stack.pop()
        n_nodes += 1
        lo, hi = B["lo"][node], B["hi"][node]  # [3, 8]
        near = np.where(s[:, None], hi, lo) - o[:, None]
        far = np.where(s[:, None], lo, hi) - o[:, None]
        tn = np.minimum(near * idlo[:, None], near * idhi[:, None]).max(0)
        tf = np.maximum(far * idlo[:, None], far * idhi[:, None]).min(0)
        tn = np.maximum(tn, 1e-3)
        tf = np.minimum(tf, best.max())
        meta = B["meta"][node]
        hit = (tn <= tf) & (meta != 0)
        inner = []
        rel = 0
        for sl in range(8):
            is_inner = (B["imask"][node] >> sl) & 1
            if is_inner:
                if hit[sl]:
                    inner.append((sl ^ octant, B["child_base"][node] + rel))
                rel += 1
            elif hit[sl]:
                cnt = {1: 1, 3: 2, 7: 3}[int(meta[sl]) >> 5]
                off = int(meta[sl]) & 31
                for k in range(cnt):
                    n_tris += 1
                    t = tris[B["tri_base"][node] + off + k]
                    best = np.minimum(best, tri_hits(o, d, t))
        for _, c in sorted(inner):  # highest priority (slot ^ octant) popped first
            stack.append(int(c))
    return n_nodes, n_tris


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tiles", type=int, default=400)
    ap.add_argument("--tw", type=int, default=4)
    ap.add_argument("--th", type=int, default=8)
    ap.add_argument("--scene", default="sponza-proxy")
    a = ap.parse_args()
    import bench
    from oracle.binding import OracleScene

    W, H = 1920, 1080
    scene, cam = bench.make_scene(a.scene, W, H)
    osc = OracleScene(scene).build_bvh()
    nodes, btris, _, _ = osc.bvh_arrays()
    B = decode(nodes)
    tv = btris.view(np.float32).reshape(-1, 12)
    tris = np.stack([tv[:, 0:3], tv[:, 4:7], tv[:, 8:11]], 1).astype(np.float64)
    rng = np.random.default_rng(1)
    ubo = cam.ubo()
    pn = pt = 0
    rn = rt = 0
    fallback = done = 0
    for _ in range(a.tiles):
        tx = rng.integers(0, W // a.tw) * a.tw
        ty = rng.integers(0, H // a.th) * a.th
        xs, ys = np.meshgrid(np.arange(tx, tx + a.tw), np.arange(ty, ty + a.th))
        o, d = camera_rays(ubo, W, H, xs.ravel().astype(np.float64), ys.ravel().astype(np.float64))
        r = packet_traverse(B, tris, o, d)
        if r is None:
            fallback += 1
            continue
        done += 1
        pn += r[0]
        pt += r[1]
        rays = np.zeros((d.shape[0], 8), np.float32)
        rays[:, 0:3] = o
        rays[:, 3] = 1e-3
        rays[:, 4:7] = d
        rays[:, 7] = 1e4
        cnt = np.zeros(6, np.uint64)
        osc.trace_rays(rays, mode="bvh", counters=cnt)
        rn += int(cnt[2])
        rt += int(cnt[3])
    nr = a.tw * a.th
    print(f"tiles {done} (+{fallback} mixed-sign): packet nodes/tile {pn / done:.1f} tris/tile {pt / done:.1f} | "
          f"per-ray nodes/ray {rn / done / nr:.1f} tris/ray {rt / done / nr:.1f}")
    print(f"  union factor: nodes {pn / (rn / nr):.2f}x  tris {pt / (rt / nr):.2f}x")


if __name__ == "__main__":
    main()
