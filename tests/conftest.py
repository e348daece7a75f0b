import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
CLEAR = (0.53, 0.81, 0.92)  # raytracing_renderer.odin:77


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def fixture_scene(name):
    from raytracer_b200.scene_io import load_scene_npz

    return load_scene_npz(os.path.join(GOLDEN, "scenes", name + ".npz"))


def soup_scene(n_tris, seed=0, n_objects=3, lights=1):
    """Deterministic triangle soup split over a few objects/materials (small)."""
    import raytracer_b200 as rb

    rng = np.random.default_rng(seed)
    s = rb.Scene()
    s.add_material(rb.Material("grey", albedo=(0.6, 0.6, 0.6), roughness=1.0))
    s.add_material(rb.Material("gold", albedo=(0.9, 0.7, 0.3), roughness=0.3, metallic=1.0))
    s.add_material(rb.Material("lamp", albedo=(1, 1, 1), emission_color=(1.0, 0.8, 0.6), emission_power=4.0, roughness=0.5))
    per = max(1, n_tris // n_objects)
    edge = 0.5 * float(n_tris) ** (-1.0 / 3.0)
    for o in range(n_objects):
        c = rng.random((per, 1, 3), dtype=np.float32)
        v = (c + (rng.random((per, 3, 3), dtype=np.float32) - 0.5) * np.float32(2 * edge)).reshape(-1, 3)
        e1, e2 = v[1::3] - v[0::3], v[2::3] - v[0::3]
        n = np.cross(e1, e2)
        n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
        nrm = np.repeat(n, 3, axis=0).astype(np.float32)
        mi = s.add_mesh(rb.make_mesh(f"soup{o}", v, nrm, np.arange(3 * per, dtype=np.uint32)))
        mat = 2 if o < lights else (o % 2)
        s.add_object(f"obj{o}", mi, mat, position=(0.1 * o, 0, 0), rotation=(10 * o, 20 * o, 0), scale=(1, 1 + 0.1 * o, 1))
    s.changes.clear()
    return s


def random_rays(n, seed, lo=(-3, -3, -3), hi=(3, 3, 3), tmin=1e-3, tmax=1e4, target_lo=None, target_hi=None):
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    o = (rng.random((n, 3), dtype=np.float32) * (hi - lo) + lo).astype(np.float32)
    if target_lo is not None:
        tl, th = np.asarray(target_lo, np.float32), np.asarray(target_hi, np.float32)
        tgt = rng.random((n, 3), dtype=np.float32) * (th - tl) + tl
        d = tgt - o
    else:
        d = rng.standard_normal((n, 3)).astype(np.float32)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3], rays[:, 3], rays[:, 4:7], rays[:, 7] = o, tmin, d, tmax
    return rays


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import binding

    binding.build()
    return binding
