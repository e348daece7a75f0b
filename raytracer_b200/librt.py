"""ctypes binding of `raytracer_b200/csrc/librt.so` — the C ABI declared in `include/rt_api.h`.

This is the only route from Python to pixels: there is no CPU fallback.  Loading fails loudly when
the library has not been built (`python -c "import __graft_entry__ as g; g.build()"` or
`make -C raytracer_b200/csrc`), and `rt_create` fails with RT_ERR_NO_DEVICE without a B200.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_LIBRT_PATH") or os.path.join(_HERE, "csrc", "librt.so")  # override: kernel A/B experiments

RT_OK = 0
STATUS_NAMES = {
    0: "RT_OK", -1: "RT_ERR_INVALID_ARGUMENT", -2: "RT_ERR_NO_DEVICE", -3: "RT_ERR_CUDA", -4: "RT_ERR_OUT_OF_MEMORY",
    -5: "RT_ERR_NO_SCENE", -6: "RT_ERR_NO_TARGET", -7: "RT_ERR_BVH_TOO_DEEP", -8: "RT_ERR_INDEX_RANGE",
}
RT_FLAG_DIRECT_LIGHTING, RT_FLAG_MIS, RT_FLAG_COUNTERS, RT_FLAG_PER_RAY_PRIMARY = 1, 2, 4, 8
RT_FLAG_DEFAULT = 7
RT_PIPELINE_WAVEFRONT, RT_PIPELINE_MEGAKERNEL = 0, 1
RT_RESTIR_TEMPORAL, RT_RESTIR_SPATIAL = 1, 2
RT_BALANCE_ROUND_ROBIN, RT_BALANCE_COST = 0, 1
NODE_BYTES, TRI_BYTES = 128, 48  # Node8 / BvhTri records of rt_get_bvh

# every entry point include/rt_api.h declares (checked against the header and `nm -D` by the tests)
EXPORTS = (
    "rt_create", "rt_destroy", "rt_default_config", "rt_last_error", "rt_version", "rt_build_scene",
    "rt_scene_info_get", "rt_update_material", "rt_update_object_transform", "rt_update_object_material",
    "rt_update_object_mesh", "rt_add_material", "rt_remove_material", "rt_add_object", "rt_remove_object", "rt_resize", "rt_trace_frame", "rt_trace_frames", "rt_trace_frames_sliced", "rt_synchronize", "rt_readback",
    "rt_readback_srgb8", "rt_device_image", "rt_upload_accum", "rt_get_strip_owners", "rt_rebalance", "rt_measure_row_costs", "rt_set_strip_owners", "rt_trace_gbuffer", "rt_readback_gbuffer", "rt_raster_gbuffer", "rt_readback_raster_gbuffer",
    "rt_default_restir_params", "rt_restir_frame", "rt_readback_reservoirs", "rt_get_counters", "rt_reset_counters",
    "rt_last_trace_ms", "rt_event_record", "rt_event_elapsed_ms", "rt_event_elapsed_per_device", "rt_set_profiling", "rt_get_kernel_profile",
    "rt_trace_rays_closest", "rt_trace_rays_any", "rt_get_bvh", "rt_static_max_rays", "rt_trace_rays_closest_static",
    "rt_filter_qualifies",
)


class RtConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_depth", C.c_int32), ("rr_start_depth", C.c_int32),
                ("radiance_clamp", C.c_float), ("flags", C.c_uint32), ("pipeline", C.c_uint32),
                ("strip_rows", C.c_uint32), ("strip_index", C.c_uint32), ("strip_count", C.c_uint32),
                ("frames_in_flight", C.c_uint32), ("n_devices", C.c_uint32), ("device_ids", C.c_int32 * 8),
                ("strip_balance", C.c_uint32), ("reserved", C.c_uint32 * 4)]


class RtMesh(C.Structure):
    _fields_ = [("vertices", C.c_void_p), ("n_vertices", C.c_uint32), ("indices", C.c_void_p),
                ("n_indices", C.c_uint32)]


class RtObject(C.Structure):
    _fields_ = [("mesh_index", C.c_uint32), ("material_index", C.c_uint32), ("model", C.c_float * 16)]


class RtCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("extend_rays", "shadow_rays", "nodes", "tris", "hits", "nee", "frames",
                                           "shadow_nodes", "shadow_tris", "launches")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


KERNEL_NAMES = ("raygen", "extend", "shade", "shadow", "accumulate", "extend_packet")


class RtKernelProfile(C.Structure):
    _fields_ = [("ms", C.c_float * 6), ("launches", C.c_uint32 * 6)]


class RtSceneInfo(C.Structure):
    _fields_ = [("n_triangles", C.c_uint32), ("n_nodes", C.c_uint32), ("bvh_depth", C.c_uint32),
                ("n_lights", C.c_uint32), ("n_objects", C.c_uint32), ("n_materials", C.c_uint32),
                ("build_ms", C.c_float), ("refits", C.c_uint32)]


class RtRestirParams(C.Structure):
    _fields_ = [("flags", C.c_uint32), ("initial_candidates", C.c_uint32), ("spatial_samples", C.c_uint32),
                ("spatial_radius", C.c_float), ("history_cap", C.c_uint32), ("reserved", C.c_uint32 * 3)]


RESERVOIR_DTYPE = np.dtype([("weight_sum", np.float32), ("target_pdf", np.float32), ("M", np.uint32),
                            ("light_index", np.uint32), ("uv", np.float32, 2), ("distance", np.float32),
                            ("W", np.float32)])
assert RESERVOIR_DTYPE.itemsize == 32


class RtError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status


_lib = None


def load() -> C.CDLL:
    """dlopen librt.so and declare the prototypes.  Raises if the library is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is not built; run __graft_entry__.build() (nvcc, sm_100a). "
                          "There is no CPU fallback for the ray-tracing path.")
    L = C.CDLL(LIB_PATH)
    vp, u32, f32p = C.c_void_p, C.c_uint32, C.c_void_p
    L.rt_create.argtypes = [C.POINTER(RtConfig), C.POINTER(vp)]
    L.rt_destroy.argtypes = [vp]
    L.rt_destroy.restype = None
    L.rt_default_config.argtypes = [C.POINTER(RtConfig)]
    L.rt_default_config.restype = None
    L.rt_last_error.argtypes = [vp]
    L.rt_last_error.restype = C.c_char_p
    L.rt_version.restype = C.c_char_p
    L.rt_build_scene.argtypes = [vp, C.POINTER(RtMesh), u32, C.POINTER(RtObject), u32, vp, u32]
    L.rt_scene_info_get.argtypes = [vp, C.POINTER(RtSceneInfo)]
    L.rt_update_material.argtypes = [vp, u32, vp]
    L.rt_update_object_transform.argtypes = [vp, u32, f32p]
    L.rt_update_object_material.argtypes = [vp, u32, u32]
    L.rt_update_object_mesh.argtypes = [vp, u32, u32]
    L.rt_add_material.argtypes = [vp, vp, C.POINTER(u32)]
    L.rt_remove_material.argtypes = [vp, u32]
    L.rt_add_object.argtypes = [vp, C.POINTER(RtObject), C.POINTER(u32)]
    L.rt_remove_object.argtypes = [vp, u32]
    L.rt_resize.argtypes = [vp, u32, u32]
    L.rt_trace_frame.argtypes = [vp, f32p, f32p, u32]
    L.rt_trace_frames.argtypes = [vp, f32p, f32p, u32, u32]
    L.rt_trace_frames_sliced.argtypes = [vp, f32p, f32p, u32, u32, u32, u32]
    L.rt_synchronize.argtypes = [vp]
    L.rt_readback.argtypes = [vp, vp, C.c_size_t]
    L.rt_readback_srgb8.argtypes = [vp, vp, C.c_size_t]
    L.rt_device_image.argtypes = [vp, C.POINTER(vp), C.POINTER(u32)]
    L.rt_upload_accum.argtypes = [vp, vp, C.c_size_t]
    L.rt_get_strip_owners.argtypes = [vp, vp, u32, C.POINTER(u32)]
    L.rt_rebalance.argtypes = [vp]
    L.rt_measure_row_costs.argtypes = [vp, f32p, u32, vp, u32]
    L.rt_set_strip_owners.argtypes = [vp, vp, u32]
    L.rt_trace_gbuffer.argtypes = [vp, f32p]
    L.rt_readback_gbuffer.argtypes = [vp, vp, vp, vp, vp, vp]
    L.rt_raster_gbuffer.argtypes = [vp, f32p]
    L.rt_readback_raster_gbuffer.argtypes = [vp, vp, vp, vp, vp, vp]
    L.rt_default_restir_params.argtypes = [C.POINTER(RtRestirParams)]
    L.rt_default_restir_params.restype = None
    L.rt_restir_frame.argtypes = [vp, f32p, f32p, u32, C.POINTER(RtRestirParams)]
    L.rt_readback_reservoirs.argtypes = [vp, vp, C.c_size_t]
    L.rt_get_counters.argtypes = [vp, C.POINTER(RtCounters)]
    L.rt_reset_counters.argtypes = [vp]
    L.rt_last_trace_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.rt_event_record.argtypes = [vp, C.c_int]
    L.rt_event_elapsed_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.rt_event_elapsed_per_device.argtypes = [vp, vp, u32, C.POINTER(u32)]
    L.rt_set_profiling.argtypes = [vp, C.c_int]
    L.rt_get_kernel_profile.argtypes = [vp, C.POINTER(RtKernelProfile)]
    L.rt_trace_rays_closest.argtypes = [vp, f32p, u32, vp, vp]
    L.rt_trace_rays_any.argtypes = [vp, f32p, u32, vp, vp]
    L.rt_static_max_rays.argtypes = []
    L.rt_static_max_rays.restype = u32
    L.rt_trace_rays_closest_static.argtypes = [vp, f32p, u32, vp, vp]
    L.rt_filter_qualifies.argtypes = [C.POINTER(RtMesh), vp, vp]
    L.rt_filter_qualifies.restype = C.c_int
    L.rt_get_bvh.argtypes = [vp, vp, C.c_size_t, vp, C.c_size_t, vp, vp]
    _lib = L
    return L


def filter_qualifies(mesh, model_matrix, material) -> bool:
    """rt_filter_qualifies (host-only, needs no GPU): would hits on an object made of this Mesh, 4x4 column-major model matrix and
    Material qualify for the first-hit filter?"""
    L = load()
    v = np.ascontiguousarray(mesh.vertices)
    ix = np.ascontiguousarray(mesh.indices, dtype=np.uint32)
    m = RtMesh()
    m.vertices, m.n_vertices, m.indices, m.n_indices = v.ctypes.data, len(v), ix.ctypes.data, len(ix)
    mm = np.ascontiguousarray(np.asarray(model_matrix, np.float32).reshape(16))
    rec = np.ascontiguousarray(material.packed().reshape(1))
    return bool(L.rt_filter_qualifies(C.byref(m), mm.ctypes.data, rec.ctypes.data))


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Context:
    """Thin object wrapper over one `rt_ctx`; every method raises RtError on a non-zero status."""

    def __init__(self, device: int = -1, max_depth: int = 3, rr_start_depth: int = 3, radiance_clamp: float = 5.0,
                 counters: bool = True, pipeline: int = RT_PIPELINE_WAVEFRONT, strip_rows: int = 8, strip_index: int = 0,
                 strip_count: int = 1, frames_in_flight: int = 0, per_ray_primary: bool = False, devices=None,
                 strip_balance: int = RT_BALANCE_ROUND_ROBIN):
        self.L = load()
        cfg = RtConfig()
        self.L.rt_default_config(C.byref(cfg))
        cfg.device, cfg.max_depth, cfg.rr_start_depth = device, max_depth, rr_start_depth
        cfg.radiance_clamp = radiance_clamp
        cfg.flags = (RT_FLAG_DIRECT_LIGHTING | RT_FLAG_MIS | (RT_FLAG_COUNTERS if counters else 0) |
                     (RT_FLAG_PER_RAY_PRIMARY if per_ray_primary else 0))
        cfg.pipeline = pipeline
        cfg.strip_rows, cfg.strip_index, cfg.strip_count = strip_rows, strip_index, strip_count
        if frames_in_flight:
            cfg.frames_in_flight = frames_in_flight
        cfg.strip_balance = strip_balance
        if devices is not None and len(devices) >= 2:  # multi-GPU fan-out inside the library: one context, all devices
            cfg.n_devices = len(devices)
            for k, d in enumerate(devices):
                cfg.device_ids[k] = int(d)
        self.cfg = cfg
        h = C.c_void_p()
        rc = self.L.rt_create(C.byref(cfg), C.byref(h))
        if rc != RT_OK:
            raise RtError(rc, (self.L.rt_last_error(None) or b"").decode())
        self.h = h
        self.width = self.height = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.rt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != RT_OK:
            raise RtError(rc, (self.L.rt_last_error(self.h) or b"").decode())

    # ---- build ------------------------------------------------------------------------------
    def build_scene(self, scene) -> RtSceneInfo:
        """`scene` is a raytracer_b200.Scene (host mirror of scene.odin)."""
        nm, no = len(scene.meshes), len(scene.objects)
        meshes = (RtMesh * max(nm, 1))()
        keep = []
        for i, m in enumerate(scene.meshes):
            v = np.ascontiguousarray(m.vertices)
            ix = np.ascontiguousarray(m.indices, dtype=np.uint32)
            keep += [v, ix]
            meshes[i].vertices, meshes[i].n_vertices = v.ctypes.data, len(v)
            meshes[i].indices, meshes[i].n_indices = ix.ctypes.data, len(ix)
        objects = (RtObject * max(no, 1))()
        for i, o in enumerate(scene.objects):
            objects[i].mesh_index, objects[i].material_index = o.mesh_index, o.material_index
            objects[i].model[:] = [float(x) for x in np.asarray(o.transform.model_matrix, np.float32)]
        mats = scene.packed_materials()
        self._check(self.L.rt_build_scene(self.h, meshes, nm, objects, no, _p(mats), len(mats)))
        return self.scene_info()

    def scene_info(self) -> RtSceneInfo:
        info = RtSceneInfo()
        self._check(self.L.rt_scene_info_get(self.h, C.byref(info)))
        return info

    def update_material(self, index, material):
        rec = np.ascontiguousarray(material.packed().reshape(1))
        self._check(self.L.rt_update_material(self.h, index, _p(rec)))

    def update_object_transform(self, index, model16):
        m = np.ascontiguousarray(model16, np.float32).reshape(16)
        self._check(self.L.rt_update_object_transform(self.h, index, _p(m)))

    def update_object_material(self, index, material_index):
        self._check(self.L.rt_update_object_material(self.h, index, material_index))

    def update_object_mesh(self, index, mesh_index):
        self._check(self.L.rt_update_object_mesh(self.h, index, mesh_index))

    def add_material(self, material) -> int:
        rec = np.ascontiguousarray(material.packed().reshape(1))
        idx = C.c_uint32()
        self._check(self.L.rt_add_material(self.h, _p(rec), C.byref(idx)))
        return idx.value

    def remove_material(self, index):
        self._check(self.L.rt_remove_material(self.h, index))

    def add_object(self, mesh_index, material_index, model16) -> int:
        o = RtObject()
        o.mesh_index, o.material_index = mesh_index, material_index
        o.model[:] = [float(x) for x in np.asarray(model16, np.float32).reshape(16)]
        idx = C.c_uint32()
        self._check(self.L.rt_add_object(self.h, C.byref(o), C.byref(idx)))
        return idx.value

    def remove_object(self, index):
        self._check(self.L.rt_remove_object(self.h, index))

    # ---- target / trace -----------------------------------------------------------------------
    def resize(self, width, height):
        self._check(self.L.rt_resize(self.h, width, height))
        self.width, self.height = width, height

    def trace_frame(self, camera_ubo, clear_color, frame_number):
        cam = np.ascontiguousarray(camera_ubo, np.float32).reshape(64)
        clr = np.ascontiguousarray(clear_color, np.float32).reshape(3)
        self._check(self.L.rt_trace_frame(self.h, _p(cam), _p(clr), frame_number))

    def trace_frames(self, camera_ubo, clear_color, first_frame, n_frames):
        cam = np.ascontiguousarray(camera_ubo, np.float32).reshape(64)
        clr = np.ascontiguousarray(clear_color, np.float32).reshape(3)
        self._check(self.L.rt_trace_frames(self.h, _p(cam), _p(clr), first_frame, n_frames))

    def trace_frames_sliced(self, camera_ubo, clear_color, first_frame, n_frames, frame_stride, first_sample):
        """Frames first_frame + k * frame_stride folded in as samples first_sample + k (spp slices)."""
        cam = np.ascontiguousarray(camera_ubo, np.float32).reshape(64)
        clr = np.ascontiguousarray(clear_color, np.float32).reshape(3)
        self._check(self.L.rt_trace_frames_sliced(self.h, _p(cam), _p(clr), first_frame, n_frames, frame_stride, first_sample))

    def synchronize(self):
        self._check(self.L.rt_synchronize(self.h))

    def last_trace_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self.L.rt_last_trace_ms(self.h, C.byref(ms)))
        return ms.value

    def readback(self, out=None) -> np.ndarray:
        if out is None:
            out = np.zeros((self.height, self.width, 4), np.float32)
        self._check(self.L.rt_readback(self.h, _p(out), out.nbytes))
        return out

    def readback_srgb8(self, out=None) -> np.ndarray:
        if out is None:
            out = np.zeros((self.height, self.width, 4), np.uint8)
        self._check(self.L.rt_readback_srgb8(self.h, _p(out), out.nbytes))
        return out

    def device_image(self):
        ptr, rows = C.c_void_p(), C.c_uint32()
        self._check(self.L.rt_device_image(self.h, C.byref(ptr), C.byref(rows)))
        return ptr.value, rows.value

    def strip_owners(self) -> np.ndarray:
        """owners[s] = rank (strip_index / device slot) that renders strip s under the assignment in force."""
        n = C.c_uint32()
        self._check(self.L.rt_get_strip_owners(self.h, None, 0, C.byref(n)))
        out = np.zeros(n.value, np.uint32)
        self._check(self.L.rt_get_strip_owners(self.h, _p(out), n.value, C.byref(n)))
        return out

    def measure_row_costs(self, camera_ubo, frame_number=0) -> np.ndarray:
        """(H, 4) counters of the strip-cost calibration pass: primary work, continuation work, hits, continuation rays."""
        cam = np.ascontiguousarray(camera_ubo, np.float32).reshape(64)
        out = np.zeros((self.height, 4), np.uint32)
        self._check(self.L.rt_measure_row_costs(self.h, _p(cam), frame_number, _p(out), self.height))
        return out

    def set_strip_owners(self, owners):
        """Install a host-decided strip -> rank table; restart the accumulation (frame 0) afterwards."""
        o = np.ascontiguousarray(owners, np.uint32)
        self._check(self.L.rt_set_strip_owners(self.h, _p(o), len(o)))

    def rebalance(self):
        self._check(self.L.rt_rebalance(self.h))

    def upload_accum(self, image):
        img = np.ascontiguousarray(image, np.float32)
        self._check(self.L.rt_upload_accum(self.h, _p(img), img.nbytes))

    # ---- G-buffer / ReSTIR ------------------------------------------------------------------
    def trace_gbuffer(self, camera_ubo):
        cam = np.ascontiguousarray(camera_ubo, np.float32).reshape(64)
        self._check(self.L.rt_trace_gbuffer(self.h, _p(cam)))

    def readback_gbuffer(self):
        H, W = self.height, self.width
        out = dict(albedo=np.zeros((H, W, 4), np.uint8), normal_rough=np.zeros((H, W, 4), np.uint16),
                   position_metal=np.zeros((H, W, 4), np.float32), emission=np.zeros((H, W, 4), np.uint16),
                   output=np.zeros((H, W, 4), np.uint8))
        self._check(self.L.rt_readback_gbuffer(self.h, _p(out["albedo"]), _p(out["normal_rough"]),
                                               _p(out["position_metal"]), _p(out["emission"]), _p(out["output"])))
        return out

    def raster_gbuffer(self, camera_ubo) -> dict:
        """The rasterised G-buffer variant (gbuffer.vert / gbuffer.frag): five (H, W, 4) float32 attachments."""
        cam = np.ascontiguousarray(camera_ubo, np.float32).reshape(64)
        self._check(self.L.rt_raster_gbuffer(self.h, _p(cam)))
        names = ("albedo", "normal_metal", "position_rough", "emission", "props")
        out = {k: np.zeros((self.height, self.width, 4), np.float32) for k in names}
        self._check(self.L.rt_readback_raster_gbuffer(self.h, *[_p(out[k]) for k in names]))
        return out

    def restir_frame(self, camera_ubo, clear_color, frame_number, params: RtRestirParams = None):
        cam = np.ascontiguousarray(camera_ubo, np.float32).reshape(64)
        clr = np.ascontiguousarray(clear_color, np.float32).reshape(3)
        if params is None:
            params = RtRestirParams()
            self.L.rt_default_restir_params(C.byref(params))
        self._check(self.L.rt_restir_frame(self.h, _p(cam), _p(clr), frame_number, C.byref(params)))

    def readback_reservoirs(self) -> np.ndarray:
        out = np.zeros((self.height, self.width), RESERVOIR_DTYPE)
        self._check(self.L.rt_readback_reservoirs(self.h, _p(out), out.nbytes))
        return out

    # ---- instrumentation / verification -------------------------------------------------------
    def counters(self) -> dict:
        c = RtCounters()
        self._check(self.L.rt_get_counters(self.h, C.byref(c)))
        return c.as_dict()

    def event_record(self, which: int):
        self._check(self.L.rt_event_record(self.h, which))

    def event_elapsed_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self.L.rt_event_elapsed_ms(self.h, C.byref(ms)))
        return ms.value

    def event_elapsed_per_device(self) -> list:
        n = C.c_uint32()
        self._check(self.L.rt_event_elapsed_per_device(self.h, None, 0, C.byref(n)))
        out = np.zeros(n.value, np.float32)
        self._check(self.L.rt_event_elapsed_per_device(self.h, _p(out), n.value, C.byref(n)))
        return [float(x) for x in out]

    def set_profiling(self, enable: bool):
        self._check(self.L.rt_set_profiling(self.h, int(enable)))

    def kernel_profile(self) -> dict:
        p = RtKernelProfile()
        self._check(self.L.rt_get_kernel_profile(self.h, C.byref(p)))
        return {n: {"ms": float(p.ms[i]), "launches": int(p.launches[i])} for i, n in enumerate(KERNEL_NAMES)}

    def reset_counters(self):
        self._check(self.L.rt_reset_counters(self.h))

    def trace_rays_closest(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        tuv = np.zeros((len(rays), 3), np.float32)
        ids = np.zeros(len(rays), np.uint32)
        self._check(self.L.rt_trace_rays_closest(self.h, _p(rays), len(rays), _p(tuv), _p(ids)))
        return tuv, ids

    def trace_rays_closest_static(self, rays):
        """rt_trace_rays_closest_static: same results, no allocation (static device scratch, <= rt_static_max_rays rays)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        tuv = np.zeros((len(rays), 3), np.float32)
        ids = np.zeros(len(rays), np.uint32)
        self._check(self.L.rt_trace_rays_closest_static(self.h, _p(rays), len(rays), _p(tuv), _p(ids)))
        return tuv, ids

    def trace_rays_any(self, rays, ignore_inst=None):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        occ = np.zeros(len(rays), np.uint32)
        ig = None if ignore_inst is None else np.ascontiguousarray(ignore_inst, np.uint32)
        self._check(self.L.rt_trace_rays_any(self.h, _p(rays), len(rays), _p(ig), _p(occ)))
        return occ

    def get_bvh(self):
        info = self.scene_info()
        nodes = np.zeros((info.n_nodes, NODE_BYTES), np.uint8)
        tris = np.zeros((info.n_triangles, 48), np.uint8)
        keys = np.zeros(info.n_triangles, np.uint64)
        ids = np.zeros(info.n_triangles, np.uint32)
        self._check(self.L.rt_get_bvh(self.h, _p(nodes), nodes.nbytes, _p(tris), tris.nbytes, _p(keys), _p(ids)))
        return nodes, tris, keys, ids
