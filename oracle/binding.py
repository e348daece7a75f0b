"""ORACLE — test infrastructure only: ctypes binding of oracle/liboracle.so.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs ONLY.  The product package (raytracer_b200/) never imports this module.
Parity status: the shader code is pinned to the reference's SHIPPED SPIR-V binaries executed by tools/spirv_interp.py
(tests/golden/spirv_frames.npz, tests/test_oracle.py); what Vulkan leaves to the driver — built-in precision, BVH traversal,
the hit a terminate-on-first-hit ray query commits — is defined by the arithmetic contract and stays unpinned.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

u32p = C.POINTER(C.c_uint32)
f32p = C.POINTER(C.c_float)


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (g++, no GPU needed)."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    stale = (not os.path.exists(_LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, capture_output=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_build_bvh.argtypes = [C.c_void_p]
        for name in ("orc_num_nodes", "orc_num_tris", "orc_num_lights", "orc_bvh_depth"):
            getattr(L, name).argtypes = [C.c_void_p]
            getattr(L, name).restype = C.c_uint32
        L.orc_get_nodes.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_get_btris.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_get_morton.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_get_lights.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_set_light_index.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.orc_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                 C.c_uint32, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_gbuffer.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_direct_reference.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p]
        L.orc_direct_reference_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                                C.c_int, C.c_int, C.c_void_p]
        L.orc_raster_gbuffer.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32] + [C.c_void_p] * 5
        L.orc_unit_raster_vertex.argtypes = [C.c_void_p] * 5
        L.orc_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p]
        L.orc_tea.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_tea.restype = C.c_uint32
        L.orc_rnd_seq.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_sincos.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_unit_sample_brdf.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_unit_eval_brdf.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_unit_specular_probability.argtypes = [C.c_void_p]
        L.orc_unit_specular_probability.restype = C.c_float
        L.orc_set_num_threads.argtypes = [C.c_int]
        L.orc_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


COUNTER_NAMES = ("extend_rays", "shadow_rays", "nodes", "tris", "hits", "nee")


class OracleScene:
    """CPU restatement of gpu_scene_init + the shader path, driven from a `raytracer_b200.Scene`."""

    def __init__(self, scene):
        L = lib()
        nm = len(scene.meshes)
        self._keep = []
        vtx = (C.c_void_p * nm)()
        idx = (u32p * nm)()
        nv = np.zeros(nm, np.uint32)
        ni = np.zeros(nm, np.uint32)
        for i, m in enumerate(scene.meshes):
            v = np.ascontiguousarray(m.vertices)
            ix = np.ascontiguousarray(m.indices, dtype=np.uint32)
            self._keep += [v, ix]
            vtx[i] = v.ctypes.data
            idx[i] = ix.ctypes.data_as(u32p)
            nv[i], ni[i] = len(v), len(ix)
        om, omat, model = scene.packed_objects()
        mats = scene.packed_materials()
        err = C.c_int(0)
        self.h = L.orc_scene_create(C.c_uint32(nm), vtx, _ptr(nv), idx, _ptr(ni), C.c_uint32(len(om)), _ptr(om),
                                    _ptr(omat), _ptr(model), C.c_uint32(len(mats)), _ptr(mats), C.byref(err))
        if not self.h:
            raise ValueError(f"oracle scene rejected: {err.value}")
        self.h = C.c_void_p(self.h)
        self.L = L

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_scene_destroy(self.h)
            self.h = None

    # ---- BVH -------------------------------------------------------------------------------
    def build_bvh(self):
        self.L.orc_build_bvh(self.h)
        return self

    @property
    def num_tris(self):
        return self.L.orc_num_tris(self.h)

    @property
    def num_lights(self):
        return self.L.orc_num_lights(self.h)

    @property
    def bvh_depth(self):
        return self.L.orc_bvh_depth(self.h)

    def bvh_arrays(self):
        """(nodes uint8[n,128], tris uint8[m,48], morton u64[T], sorted_ids u32[T])."""
        nn = self.L.orc_num_nodes(self.h)
        nt = self.num_tris
        nodes = np.zeros((nn, 128), np.uint8)
        tris = np.zeros((nt, 48), np.uint8)
        keys = np.zeros(nt, np.uint64)
        ids = np.zeros(nt, np.uint32)
        self.L.orc_get_nodes(self.h, _ptr(nodes))
        if nt:
            self.L.orc_get_btris(self.h, _ptr(tris))
            self.L.orc_get_morton(self.h, _ptr(keys), _ptr(ids))
        return nodes, tris, keys, ids

    def set_light_index(self, obj, li):
        self.L.orc_set_light_index(self.h, obj, li)

    # ---- rendering --------------------------------------------------------------------------
    def render(self, cam_ubo, clear, frame, W, H, image=None, y0=0, y1=None, max_depth=3, clamp=5.0, use_bvh=True,
               counters=None):
        if image is None:
            image = np.zeros((H, W, 4), np.float32)
        assert image.dtype == np.float32 and image.shape == (H, W, 4) and image.flags.c_contiguous
        cam = np.ascontiguousarray(cam_ubo, np.float32)
        clr = np.ascontiguousarray(clear, np.float32)
        cnt = _ptr(counters) if counters is not None else None
        self.L.orc_render(self.h, _ptr(cam), _ptr(clr), frame, W, H, y0, H if y1 is None else y1, max_depth, clamp,
                          int(use_bvh), _ptr(image), cnt)
        return image

    def gbuffer(self, cam_ubo, W, H, use_bvh=True):
        cam = np.ascontiguousarray(cam_ubo, np.float32)
        out = dict(albedo=np.zeros((H, W, 4), np.uint8), normal_rough=np.zeros((H, W, 4), np.uint16),
                   position_metal=np.zeros((H, W, 4), np.float32), emission=np.zeros((H, W, 4), np.uint16),
                   output=np.zeros((H, W, 4), np.uint8))
        self.L.orc_gbuffer(self.h, _ptr(cam), W, H, 0, H, int(use_bvh), _ptr(out["albedo"]), _ptr(out["normal_rough"]),
                           _ptr(out["position_metal"]), _ptr(out["emission"]), _ptr(out["output"]))
        return out

    def raster_gbuffer(self, cam_ubo, W, H):
        """CPU twin of rt_raster_gbuffer (gbuffer.vert / gbuffer.frag + the stated fixed-function rules)."""
        cam = np.ascontiguousarray(cam_ubo, np.float32)
        names = ("albedo", "normal_metal", "position_rough", "emission", "props")
        out = {k: np.zeros((H, W, 4), np.float32) for k in names}
        self.L.orc_raster_gbuffer(self.h, _ptr(cam), W, H, *[_ptr(out[k]) for k in names])
        return out

    def direct_reference(self, cam_ubo, clear, W, H, samples_per_axis=4, use_bvh=True, y0=0, y1=None):
        """Brute-force direct lighting of the primary hits (ground truth of the ReSTIR DI tests); rows y0 .. y1-1."""
        cam = np.ascontiguousarray(cam_ubo, np.float32)
        clr = np.ascontiguousarray(clear, np.float32)
        img = np.zeros((H, W, 4), np.float32)
        self.L.orc_direct_reference_rows(self.h, _ptr(cam), _ptr(clr), W, H, y0, H if y1 is None else y1, int(use_bvh),
                                         int(samples_per_axis), _ptr(img))
        return img

    def trace_rays(self, rays, mode="bvh", any_hit=False, ignore_inst=None, counters=None):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        m = (1 if mode == "bvh" else 0) + (2 if any_hit else 0)
        tuv = np.zeros((n, 3), np.float32)
        ids = np.zeros(n, np.uint32)
        ig = None
        if ignore_inst is not None:
            ig = np.ascontiguousarray(ignore_inst, np.uint32)
        self.L.orc_trace_rays(self.h, _ptr(rays), n, m, _ptr(ig) if ig is not None else None, _ptr(tuv), _ptr(ids),
                              _ptr(counters) if counters is not None else None)
        return tuv, ids


def tea(a, b):
    return lib().orc_tea(a, b)


def rnd_seq(seed, n):
    out = np.zeros(n, np.float32)
    so = C.c_uint32(0)
    lib().orc_rnd_seq(seed, n, _ptr(out), C.byref(so))
    return out, so.value


def sincos(a):
    a = np.ascontiguousarray(a, np.float32)
    s = np.zeros_like(a)
    c = np.zeros_like(a)
    lib().orc_sincos(_ptr(a), a.size, _ptr(s), _ptr(c))
    return s, c
