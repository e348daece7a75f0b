"""Deterministic procedural stand-ins for scenes whose geometry blobs are absent from the reference
mount (`/root/reference/.MISSING_LARGE_BLOBS`): `models/sponza/sponza.bin`.

`sponza_proxy()` keeps everything the reference's loader would have produced that does NOT live in
the missing blob — 103 objects (one per glTF primitive, `scene_loader.odin:101-187`), their exact
triangle counts (262,267 in total) and material indices, the 25 materials as the loader reads them
(factors only: albedo 0.588, metallic 1 unless stated, roughness 1; no lights, so the scene is sky-lit)
and the node scale 0.008 — and replaces the vertex data by a procedural two-storey colonnaded atrium
of the same extent ([-15.4,-1,-9.5]..[14.4,11.4,8.8] after the node transform).  Every result
obtained on it is labelled "sponza-proxy".  The layout table is `data/sponza_layout.npz`
(tools/make_fixtures.py extracts it from `models/sponza/sponza.gltf`, which is JSON and present).

`soup()` is the synthetic sweep scene of SURVEY §8d (config S).
"""
from __future__ import annotations

import os

import numpy as np

from .scene import F32, Material, Object, Scene, Transform, make_mesh

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def _grid_dims(n_tris: int, aspect: float):
    """a x b quads with 2ab <= n_tris, a/b ~ aspect, a,b >= 3."""
    q = max(n_tris // 2, 9)
    a = max(3, int(np.sqrt(q * aspect)))
    b = max(3, q // a)
    while 2 * a * b > n_tris and b > 3:
        b -= 1
    while 2 * a * b > n_tris and a > 3:
        a -= 1
    return a, b


def parametric_mesh(name: str, fn, n_tris: int, aspect: float = 1.0, wrap_u: bool = False):
    """Triangulate fn(u, v) -> (N,3) on a grid so that the mesh has EXACTLY n_tris triangles.

    The remainder after 2*a*b is made up by splitting triangles at an edge midpoint (+1 each)."""
    assert n_tris >= 1, n_tris
    drop = 0
    if n_tris < 18:  # a one-quad-high band around the surface; an odd count drops the last triangle
        a, b = (n_tris + 1) // 2, 1
        drop = 2 * a - n_tris
    else:
        a, b = _grid_dims(n_tris, aspect)
    us = np.linspace(0.0, 1.0, a + 1)
    vs = np.linspace(0.0, 1.0, b + 1)
    U, V = np.meshgrid(us, vs, indexing="ij")
    P = fn(U.ravel(), V.ravel()).astype(np.float64)
    e = 1e-4
    du = fn(np.clip(U.ravel() + e, 0, 1), V.ravel()) - fn(np.clip(U.ravel() - e, 0, 1), V.ravel())
    dv = fn(U.ravel(), np.clip(V.ravel() + e, 0, 1)) - fn(U.ravel(), np.clip(V.ravel() - e, 0, 1))
    N = np.cross(du, dv)
    ln = np.linalg.norm(N, axis=1, keepdims=True)
    N = np.where(ln > 1e-12, N / np.maximum(ln, 1e-12), np.array([[0.0, 1.0, 0.0]]))
    i, j = np.meshgrid(np.arange(a), np.arange(b), indexing="ij")
    i, j = i.ravel(), j.ravel()
    v00 = i * (b + 1) + j
    v10 = (i + 1) * (b + 1) + j
    v01, v11 = v00 + 1, v10 + 1
    tris = np.concatenate([np.stack([v00, v10, v11], 1), np.stack([v00, v11, v01], 1)], 0)
    if drop:
        tris = tris[:-drop]
    r = n_tris - len(tris)
    assert 0 <= r <= len(tris), (n_tris, a, b, r)
    if r:
        A, B, C = tris[:r, 0], tris[:r, 1], tris[:r, 2]
        M = len(P) + np.arange(r)
        P = np.concatenate([P, 0.5 * (P[B] + P[C])], 0)
        nm = N[B] + N[C]
        nm /= np.maximum(np.linalg.norm(nm, axis=1, keepdims=True), 1e-12)
        N = np.concatenate([N, nm], 0)
        tris = np.concatenate([np.stack([A, B, M], 1), np.stack([A, M, C], 1), tris[r:]], 0)
    assert len(tris) == n_tris
    return make_mesh(name, P.astype(F32), N.astype(F32), tris.astype(np.uint32).ravel())


# ---- surface generators (all closed or sheet-like, world units = metres) ---------------------------
def _superbox(centre, half, e=0.25):
    cx, cy, cz = centre
    hx, hy, hz = half

    def f(u, v):
        th = (u - 0.5) * 2 * np.pi
        ph = (v - 0.5) * np.pi
        sp = lambda x: np.sign(x) * np.abs(x) ** e  # noqa: E731
        return np.stack([cx + hx * sp(np.cos(ph)) * sp(np.cos(th)), cy + hy * sp(np.sin(ph)),
                         cz + hz * sp(np.cos(ph)) * sp(np.sin(th))], 1)
    return f


def _relief_wall(p0, du, dv, normal, amp, fu, fv):
    p0, du, dv, normal = (np.asarray(x, np.float64) for x in (p0, du, dv, normal))

    def f(u, v):
        h = amp * (np.sin(2 * np.pi * fu * u) * np.sin(2 * np.pi * fv * v)) ** 2
        return p0 + np.outer(u, du) + np.outer(v, dv) + np.outer(h, normal)
    return f


def _column(base, radius, height, flutes=12):
    bx, by, bz = base

    def f(u, v):
        th = 2 * np.pi * u
        prof = 1.0 + 0.35 * np.exp(-((v - 0.02) / 0.04) ** 2) + 0.45 * np.exp(-((v - 0.97) / 0.035) ** 2)
        r = radius * prof * (1.0 + 0.04 * np.cos(flutes * th))
        return np.stack([bx + r * np.cos(th), by + height * v, bz + r * np.sin(th)], 1)
    return f


def _arch(c0, c1, y0, rise, thick):
    c0, c1 = np.asarray(c0, np.float64), np.asarray(c1, np.float64)
    mid, half = 0.5 * (c0 + c1), 0.5 * np.linalg.norm(c1 - c0)
    ax = (c1 - c0) / (2 * half)
    side = np.array([-ax[1], ax[0]])

    def f(u, v):
        a = np.pi * u
        t = 2 * np.pi * v
        R = half + thick * np.cos(t)
        x = mid[0] - ax[0] * R * np.cos(a) + side[0] * thick * 1.6 * np.sin(t)
        z = mid[1] - ax[1] * R * np.cos(a) + side[1] * thick * 1.6 * np.sin(t)
        y = y0 + (rise / half) * R * np.sin(a)
        return np.stack([x, y, z], 1)
    return f


def _drape(p0, width_vec, drop, waves, depth):
    p0, width_vec = np.asarray(p0, np.float64), np.asarray(width_vec, np.float64)
    n = np.array([-width_vec[2], 0.0, width_vec[0]])
    n /= np.linalg.norm(n)

    def f(u, v):
        sag = depth * np.sin(2 * np.pi * waves * u) * (0.3 + v) + 0.25 * depth * np.sin(7 * np.pi * v + 3 * u)
        return p0 + np.outer(u, width_vec) + np.outer(-drop * v, [0, 1, 0]) + np.outer(sag, n)
    return f


def _vase(base, radius, height):
    bx, by, bz = base

    def f(u, v):
        th = 2 * np.pi * u
        r = radius * (0.35 + 0.65 * np.sin(np.pi * np.clip(v, 0, 1)) ** 0.8 * (1 + 0.3 * np.sin(3 * np.pi * v)))
        r = r * (1 + 0.05 * np.sin(9 * th) * np.sin(np.pi * v))
        return np.stack([bx + r * np.cos(th), by + height * v, bz + r * np.sin(th)], 1)
    return f


def _sponza_elements():
    """103 (name, surface fn, aspect) entries, largest structures first (they get the largest triangle
    counts of the layout table)."""
    E = []
    # outer shell: x in [-15.4, 14.4], z in [-9.5, 8.8], y in [-1, 11.4]; courtyard x in [-11, 10], z in [-4, 3.3]
    x0, x1, z0, z1 = -15.4, 14.4, -9.5, 8.8
    E.append(("floor", _relief_wall((x0, 0.0, z0), (x1 - x0, 0, 0), (0, 0, z1 - z0), (0, -1, 0), 0.03, 30, 18), 1.6))
    E.append(("wall_n", _relief_wall((x0, -1.0, z1), (x1 - x0, 0, 0), (0, 12.4, 0), (0, 0, -1), 0.12, 24, 10), 2.4))
    E.append(("wall_s", _relief_wall((x0, -1.0, z0), (x1 - x0, 0, 0), (0, 12.4, 0), (0, 0, 1), 0.12, 24, 10), 2.4))
    E.append(("wall_e", _relief_wall((x1, -1.0, z0), (0, 0, z1 - z0), (0, 12.4, 0), (-1, 0, 0), 0.12, 14, 10), 1.5))
    E.append(("wall_w", _relief_wall((x0, -1.0, z0), (0, 0, z1 - z0), (0, 12.4, 0), (1, 0, 0), 0.12, 14, 10), 1.5))
    cx0, cx1, cz0, cz1 = -11.0, 10.0, -4.0, 3.3
    for y, tag in ((5.0, "gallery"), (10.6, "roof")):
        E.append((f"{tag}_n", _superbox(((x0 + x1) / 2, y, (cz1 + z1) / 2), ((x1 - x0) / 2, 0.25, (z1 - cz1) / 2)), 3.0))
        E.append((f"{tag}_s", _superbox(((x0 + x1) / 2, y, (cz0 + z0) / 2), ((x1 - x0) / 2, 0.25, (cz0 - z0) / 2)), 3.0))
        E.append((f"{tag}_e", _superbox(((cx1 + x1) / 2, y, (cz0 + cz1) / 2), ((x1 - cx1) / 2, 0.25, (cz1 - cz0) / 2)), 1.0))
        E.append((f"{tag}_w", _superbox(((cx0 + x0) / 2, y, (cz0 + cz1) / 2), ((cx0 - x0) / 2, 0.25, (cz1 - cz0) / 2)), 1.0))
    xs = np.linspace(cx0, cx1, 7)
    for k in range(12):  # drapes between the upper columns
        xa = xs[k % 6]
        z = cz1 + 0.3 if k < 6 else cz0 - 0.3
        E.append((f"drape{k}", _drape((xa + 0.4, 9.6, z), (xs[1] - xs[0] - 0.8, 0, 0), 3.6, 3 + (k % 3), 0.35), 1.0))
    cols = []
    for storey, (yb, h, r) in enumerate(((0.0, 4.75, 0.42), (5.25, 5.1, 0.34))):
        for row, z in enumerate((cz0, cz1)):
            for xa in np.linspace(cx0, cx1, 10):
                cols.append((f"col{storey}{row}_{xa:.1f}", _column((xa, yb, z), r, h), 0.35))
    E += cols  # 40 columns
    for row, z in enumerate((cz0, cz1)):  # 20 arches on the ground floor
        xa = np.linspace(cx0, cx1, 10)
        for k in range(9):
            E.append((f"arch{row}_{k}", _arch((xa[k], z), (xa[k + 1], z), 3.2, 1.1, 0.16), 3.0))
        E.append((f"arch{row}_end", _arch((cx0 if row else cx1, cz0), (cx0 if row else cx1, cz1), 3.2, 1.4, 0.16), 3.0))
    k = 0
    while len(E) < 103:  # vases along the aisles and in the courtyard
        x = -12.5 + 1.45 * k
        z = (-6.6, 6.0, -1.5, 1.0)[k % 4]
        E.append((f"vase{k}", _vase((x if k % 4 < 2 else -9.0 + 1.1 * k, 0.0, z), 0.38, 1.1 + 0.1 * (k % 3)), 1.0))
        k += 1
    assert len(E) == 103
    return E


def sponza_layout():
    z = np.load(os.path.join(_DATA, "sponza_layout.npz"), allow_pickle=False)
    return z["tri_counts"].astype(np.int64), z["material_index"].astype(np.int64), z["materials"].astype(F32)


def sponza_proxy() -> Scene:
    tri_counts, mat_idx, mats = sponza_layout()
    elements = _sponza_elements()
    order = np.argsort(-tri_counts, kind="stable")  # largest primitive -> first (largest) element
    s = Scene()
    for i, r in enumerate(mats):
        s.materials.append(Material(f"sponza_mat{i}", tuple(r[0:3]), tuple(r[3:6]), F32(r[6]), F32(r[7]), F32(r[8])))
    slot_of = np.empty(len(order), np.int64)
    slot_of[order] = np.arange(len(order))
    scale = F32(0.00800000037997961)  # the glTF node's scale; the mesh is stored in the node's units
    model = np.zeros(16, F32)
    model[0] = model[5] = model[10] = scale
    model[15] = 1
    for prim in range(len(tri_counts)):  # objects keep the primitive order of the glTF
        name, fn, aspect = elements[int(slot_of[prim])]
        n = int(tri_counts[prim])
        mesh = parametric_mesh(name, lambda u, v, fn=fn: fn(u, v) / float(scale), n, aspect)
        mi = s.add_mesh(mesh)
        tr = Transform(scale=np.full(3, scale, F32), model_matrix=model.copy())
        s.objects.append(Object(name, tr, mi, int(mat_idx[prim])))
    return s


SPONZA_CAMERA = dict(position=(-10.0, 2.0, 0.0), target=(0.0, 3.0, 0.0))


def soup(n_tris: int, seed: int = 0) -> Scene:
    """Config S: uniformly scattered triangles in the unit cube, one grey rough material, sky only."""
    rng = np.random.default_rng(seed)
    s = Scene()
    s.materials.append(Material("grey", albedo=(0.6, 0.6, 0.6), roughness=1.0))
    edge = 0.5 * float(n_tris) ** (-1.0 / 3.0)
    c = rng.random((n_tris, 1, 3), dtype=np.float32)
    v = (c + (rng.random((n_tris, 3, 3), dtype=np.float32) - 0.5) * np.float32(2 * edge)).reshape(-1, 3)
    n = np.cross(v[1::3] - v[0::3], v[2::3] - v[0::3])
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    mi = s.add_mesh(make_mesh("soup", v, np.repeat(n, 3, axis=0).astype(F32), np.arange(3 * n_tris, dtype=np.uint32)))
    s.objects.append(Object("soup", Transform(), mi, 0))
    return s


SOUP_CAMERA = dict(position=(0.5, 0.5, -2.0), target=(0.5, 0.5, 0.5))
