#!/usr/bin/env python3
"""Golden frames from the reference's SHIPPED shader binaries, executed by tools/spirv_interp.py — test infrastructure.

    python tools/spirv_golden.py [--ref /root/reference] [--out tests/golden/spirv_frames.npz]

`ShaderVM` stands in for the Vulkan ray-tracing pipeline of `raytracing_renderer.odin:62-97,204-270`: it loads `rgen.spv`,
`rchit.spv` and `rmiss.spv` (the three modules the shader binding table dispatches to: raygen, hit group 0, miss 0),
fills the descriptor sets from a `raytracer_b200.Scene` the way `gpu_scene.odin:317-428` does, and runs the ray-generation
shader once per pixel, like `vkCmdTraceRaysKHR(W, H, 1)`.  `OpTraceRayKHR` finds the closest hit with the oracle's BVH
traversal (the driver's part: no file of the reference defines it) and then runs the closest-hit or the miss BINARY on the
shared payload.  With prefix "restir_" the same machine runs the G-buffer pass (restir_rgen / restir_rchit.spv).

The committed fixture holds, per scene, the RGBA32F image after every frame (and the five G-buffer images).  Pins:
  tests/test_oracle.py      the CPU oracle reproduces every frame bit for bit (CPU suite; no reference mount needed) and,
                            where the mount exists, a sample of pixels is re-executed from the binaries;
  tests/test_gpu_parity.py  the CUDA path reproduces them bit for bit through the C ABI (-m gpu).
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import spirv_interp as si  # noqa: E402

F32 = np.float32
CLEAR = (0.53, 0.81, 0.92)  # raytracing_renderer.odin:77

# what the fixture holds: scene fixture name -> (W, H, frames, camera kwargs)
CASES = {
    "cornell_box": (32, 32, 3, dict()),
    "cornell_pbr": (32, 32, 3, dict(position=(0.0, 1.0, 4.5), target=(0.0, 1.0, 0.0))),
    "conference": (32, 20, 2, dict(position=(-3.5, -2.0, 1.6), target=(1.0, 0.5, 1.0), up=(0.0, 0.0, 1.0))),
    "multi_dispersion": (32, 18, 2, dict(position=(0.0, 2.0, 8.0), target=(0.0, 0.5, 0.0))),
}
GBUFFER_CASES = ("cornell_pbr", "conference")


def _vec(a):
    return [F32(x) for x in a]


def _mat4(m16):
    m = np.asarray(m16, np.float32).reshape(4, 4)  # column-major: row c of the reshape is column c
    return [_vec(m[c]) for c in range(4)]


class ShaderVM:
    def __init__(self, ref_dir, scene, oracle_scene, sincos, prefix="", query_policy="all"):
        """prefix "" = the path tracer (rgen/rchit/rmiss.spv); "restir_" = the G-buffer pass (restir_rgen/rchit/rmiss.spv).
        query_policy: which hit the inline ray query of isVisibleRQ commits (simple.rchit:350-385; with
        TerminateOnFirstHit that is implementation defined): "all" = the defined semantics of this repo — the sample is
        visible only if NO other instance lies in range (si.CommittedInstances); "closest" = a driver that happens to
        commit the nearest hit (differs where a light's own geometry hides an occluder behind it)."""
        self.prefix, self.query_policy = prefix, query_policy
        sh = os.path.join(ref_dir, "shaders")
        self.mod = {k: si.Module(os.path.join(sh, prefix + k + ".spv")) for k in ("rgen", "rchit", "rmiss")}
        self.rgen = si.Machine(self.mod["rgen"], sincos)
        self.rchit = si.Machine(self.mod["rchit"], sincos)
        self.rmiss = si.Machine(self.mod["rmiss"], sincos)
        self.scene, self.orc = scene, oracle_scene
        self.rays = {"closest": 0, "query": 0, "hits": 0}
        # ---- descriptor set 0 of the hit shader (gpu_scene.odin:317-428) ----
        VBASE, IBASE = 0x10_0000_0000, 0x20_0000_0000
        for k, m in enumerate(scene.meshes):
            verts = [[_vec(v["pos"]), _vec(v["normal"])] for v in m.vertices]
            tris = [[int(i) for i in m.indices[3 * t:3 * t + 3]] for t in range(len(m.indices) // 3)]
            self.rchit.addresses[VBASE + (k << 24)] = [[verts]]   # buffer Vertices { Vertex v[]; }
            self.rchit.addresses[IBASE + (k << 24)] = [[tris]]    # buffer Indices { ivec3 indices[]; }
        objects, lights, self.tri_first = [], [], [0]
        for i, o in enumerate(scene.objects):
            # light_index is written to a copy in gpu_scene_compile_objects_data (:333-336): 0 on the GPU after a bake
            objects.append([VBASE + (o.mesh_index << 24), IBASE + (o.mesh_index << 24), int(o.material_index), int(o.mesh_index), 0])
            ntri = scene.meshes[o.mesh_index].num_triangles
            self.tri_first.append(self.tri_first[-1] + ntri)
            if float(scene.materials[o.material_index].emission_power) > 0.0:
                lights.append([_mat4(o.transform.model_matrix), i, ntri, 1 << (i & 7)])
        mats = []
        for m in scene.materials:
            mats.append([_vec(m.albedo), _vec(m.emission_color), F32(m.emission_power), F32(m.roughness), F32(m.metallic),
                         F32(m.transmission), F32(m.ior)])
        self.rchit.vars["objects_data"].store([objects])
        self.rchit.vars["materials_data"].store([mats])
        if "lights_data" in self.rchit.vars:  # the G-buffer hit shader binds no lights
            self.rchit.vars["lights_data"].store([lights])
        self.tri_first = np.asarray(self.tri_first, np.int64)
        self.rgen.trace_ray = self._trace_ray
        self.rchit.ray_query = self._ray_query

    # ---- the driver's part: traversal (oracle BVH), then the shader binding table's dispatch ----
    @staticmethod
    def _ray(origin, tmin, direction, tmax):
        return np.asarray([[origin[0], origin[1], origin[2], tmin, direction[0], direction[1], direction[2], tmax]], np.float32)

    def _closest(self, origin, tmin, direction, tmax):
        tuv, ids = self.orc.trace_rays(self._ray(origin, tmin, direction, tmax), mode="bvh")
        if ids[0] == 0xFFFFFFFF:
            return None
        obj = int(np.searchsorted(self.tri_first, int(ids[0]), side="right") - 1)
        return obj, int(ids[0]) - int(self.tri_first[obj]), tuv[0]

    def _trace_ray(self, machine, origin, tmin, direction, tmax, payload_ptr):
        self.rays["closest"] += 1
        hit = self._closest(origin, tmin, direction, tmax)
        if hit is None and self.prefix == "restir_":
            # restir_rmiss.spv declares the PATH TRACER's RayPayload (restir.rmiss:5-16), not RestirPayload: its `hit = false`
            # lands outside the struct the ray-gen shader reads.  restir.rgen:59 has already cleared its own flag, so the
            # miss stage changes nothing the ray-gen shader looks at; it is not run here.
            return
        stage = self.rmiss if hit is None else self.rchit
        stage.vars["payload"].root[0] = payload_ptr.root[0]  # one payload, shared by reference (location 0)
        if hit is not None:
            self.rays["hits"] += 1
            obj, prim, tuv = hit
            m = np.asarray(self.scene.objects[obj].transform.model_matrix, np.float32).reshape(4, 4)
            inputs = {"InstanceCustomIndex": obj,                  # instanceCustomIndex = object index (gpu_scene.odin:135)
                      "PrimitiveId": prim, "WorldRayDirection": _vec(direction),
                      "ObjectToWorld": [_vec(m[c][:3]) for c in range(4)]}  # mat4x3: four columns of three rows
            for k, val in inputs.items():
                if k in self.rchit.builtin:                          # a module declares only the built-ins it reads
                    self.rchit.builtin[k].store(val)
            self.rchit.vars["attribs"].store([F32(tuv[1]), F32(tuv[2])])
        stage.run()
        payload_ptr.root[0] = stage.vars["payload"].root[0]

    def _ray_query(self, origin, tmin, direction, tmax):
        self.rays["query"] += 1
        hit = self._closest(origin, tmin, direction, tmax)
        if hit is None:
            return None
        if self.query_policy == "closest":
            return hit[0]
        _, occ = self.orc.trace_rays(self._ray(origin, tmin, direction, tmax), mode="bvh", any_hit=True,
                                     ignore_inst=np.asarray([hit[0]], np.uint32))
        return si.CommittedInstances(hit[0], occ[0] != 0)

    # ---- vkCmdTraceRaysKHR(W, H, 1) ----
    def _setup(self, camera_ubo, clear, frame, W, H):
        cam = np.asarray(camera_ubo, np.float32).reshape(4, 16)
        g = self.rgen
        g.vars["cam"].store([_mat4(cam[k]) for k in range(4)])  # proj, view, inverse_view, inverse_proj
        g.vars["push"].store([_vec(clear), int(frame)])
        g.builtin["LaunchSize"].store([W, H, 1])
        return g

    def _pixel(self, g, x, y):
        g.builtin["LaunchId"].store([x, y, 0])
        g.vars["payload"].store(self.mod["rgen"].default(self._payload_type()))
        g.run()

    def render(self, camera_ubo, clear, frame, W, H, image=None, pixels=None):
        if image is None:
            image = np.zeros((H, W, 4), np.float32)
        g = self._setup(camera_ubo, clear, frame, W, H)
        g.images["image"] = image
        for (x, y) in (pixels if pixels is not None else [(x, y) for y in range(H) for x in range(W)]):
            self._pixel(g, x, y)
        return image

    GBUFFER_IMAGES = ("outputImage", "gBufferAlbedo", "gBufferNormal", "gBufferPosition", "gBufferEmission")

    def render_gbuffer(self, camera_ubo, W, H):
        """restir.rgen over W x H (prefix "restir_"): the five storage images as float texels, then converted the way the
        image formats store them (rgba8: unorm RNE; rgba16f: binary16 RNE; rgba32f as is) — the layout of rt_readback_gbuffer."""
        g = self._setup(camera_ubo, CLEAR, 0, W, H)
        tex = {k: np.zeros((H, W, 4), np.float32) for k in self.GBUFFER_IMAGES}
        g.images.update(tex)
        for y in range(H):
            for x in range(W):
                self._pixel(g, x, y)

        def unorm8(a):
            c = np.clip(np.nan_to_num(a, nan=0.0), 0.0, 1.0).astype(np.float32)
            return np.rint(c * np.float32(255.0)).astype(np.uint8)

        return dict(albedo=unorm8(tex["gBufferAlbedo"]), output=unorm8(tex["outputImage"]),
                    normal_rough=tex["gBufferNormal"].astype(np.float16).view(np.uint16),
                    emission=tex["gBufferEmission"].astype(np.float16).view(np.uint16), position_metal=tex["gBufferPosition"])

    def _payload_type(self):
        m = self.mod["rgen"]
        for vid, tid, _ in m.globals_decl:
            if m.names.get(vid) == "payload":
                return m.types[tid][2]
        raise KeyError("payload")


def unit_vectors(ref_dir, sincos, n=256, seed=7):
    """Function-level vectors from rchit.spv: sampleBRDF (simple.rchit:403-449; reaches getSpecularProbability, sampleGGX,
    generateCosineWeightedDirection, microfacetF, microfacetPDF, D_GGX, G_Smith, F_Schlick) and evaluateFullBRDF (:118-166)
    on random materials and directions, grazing angles and the roughness extremes included — inputs the small frames
    need not contain.  Returned as arrays: inputs, then what the BINARY computes."""
    mod = si.Module(os.path.join(ref_dir, "shaders", "rchit.spv"))
    vm = si.Machine(mod, sincos)
    f_sample, f_eval = mod.function_named("sampleBRDF"), mod.function_named("evaluateFullBRDF")
    rng = np.random.default_rng(seed)
    mats = np.zeros((n, 11), np.float32)      # the 44-byte Material record
    mats[:, 0:3] = rng.random((n, 3))
    mats[:, 7] = np.where(rng.random(n) < 0.2, rng.choice([0.0, 0.02, 1.0], n), rng.random(n))   # roughness
    mats[:, 8] = np.where(rng.random(n) < 0.5, rng.choice([0.0, 1.0], n), rng.random(n))         # metallic
    def dirs(k):
        d = rng.standard_normal((k, 3)).astype(np.float32)
        d[:, 2] = np.abs(d[:, 2]) * np.where(rng.random(k) < 0.25, 0.02, 1.0)   # a quarter of them grazing
        return (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    wo, wi = dirs(n), dirs(n)
    xi = rng.random((n, 2)).astype(np.float32)
    seeds = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    ident = [_vec(c) for c in np.eye(3, dtype=np.float32)]
    sample_out, eval_out, seed_out = np.zeros((n, 8), np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.uint32)
    for k in range(n):
        m = mats[k]
        mat = [_vec(m[0:3]), _vec(m[3:6]), F32(m[6]), F32(m[7]), F32(m[8]), F32(m[9]), F32(m[10])]
        vm.vars["payload"].at([4]).store(int(seeds[k]))   # RayPayload.seed: the lobe choice draws from it (simple.rchit:408)
        r = vm.run(f_sample, [si.Ptr([_vec(wo[k])]), si.Ptr([mat]), si.Ptr([_vec(xi[k])]), si.Ptr([ident])])
        sample_out[k] = list(r[0]) + list(r[1]) + [r[2], F32(1.0 if r[3] else 0.0)]
        seed_out[k] = vm.vars["payload"].at([4]).load()
        eval_out[k] = vm.run(f_eval, [si.Ptr([_vec(wo[k])]), si.Ptr([_vec(wi[k])]), si.Ptr([mat])])
    return {"units/materials": mats, "units/wo": wo, "units/wi": wi, "units/xi": xi, "units/seed_in": seeds,
            "units/sampleBRDF": sample_out, "units/seed_out": seed_out, "units/evaluateFullBRDF": eval_out}


def raster_stage_vectors(ref_dir, sincos, n=128, seed=11):
    """gbuffer_vert.spv and gbuffer_frag.spv (shaders/restir/gbuffer.vert:22-37, gbuffer.frag:26-39) run on random inputs: the two
    programmable stages of the rasterised G-buffer variant.  Inputs and what the binaries compute, as arrays."""
    sh = os.path.join(ref_dir, "shaders", "restir")
    vs = si.Machine(si.Module(os.path.join(sh, "gbuffer_vert.spv")), sincos)
    fs = si.Machine(si.Module(os.path.join(sh, "gbuffer_frag.spv")), sincos)
    import raytracer_b200 as rb

    rng = np.random.default_rng(seed)
    cam = rb.Camera(position=(0.3, 1.2, 4.0), target=(0.0, 0.8, 0.0), aspect=16 / 9).ubo()
    camt = np.asarray(cam, np.float32).reshape(4, 16)
    vs.vars["camera"].store([_mat4(camt[k]) for k in range(4)])
    models = np.zeros((n, 16), np.float32)
    pos = (rng.standard_normal((n, 3)) * 2).astype(np.float32)
    nrm = rng.standard_normal((n, 3)).astype(np.float32)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    vout = np.zeros((n, 10), np.float32)
    for k in range(n):
        m = rb.model_matrix_from_trs(rng.standard_normal(3), rng.uniform(-180, 180, 3), rng.uniform(0.3, 2.0, 3))
        models[k] = m
        cols = _mat4(m)
        nm = si.inverse3([c[:3] for c in cols[:3]])                    # inverse(mat3(model)) ...
        nm = [[nm[c][r] for c in range(3)] for r in range(3)]           # ... transposed: columns of the normal matrix
        nm4 = [nm[0] + [F32(0)], nm[1] + [F32(0)], nm[2] + [F32(0)], [F32(0), F32(0), F32(0), F32(1)]]
        vs.vars["push"].store([cols, nm4, 0])
        vs.vars["aPos"].store(_vec(pos[k]))
        vs.vars["aNormal"].store(_vec(nrm[k]))
        vs.run()
        vout[k] = list(vs.vars["outPos"].load()) + list(vs.vars["outNormal"].load()) + list(vs.vars[""].load()[0])
    mats = np.zeros((n, 11), np.float32)
    mats[:, 0:6] = rng.random((n, 6))
    mats[:, 6] = np.where(rng.random(n) < 0.5, 0.0, rng.random(n) * 10)
    mats[:, 7:9] = rng.random((n, 2))
    in_pos = (rng.standard_normal((n, 3)) * 3).astype(np.float32)
    in_nrm = (rng.standard_normal((n, 3))).astype(np.float32)
    fout = np.zeros((n, 5, 4), np.float32)
    fs.vars[""].store([[[_vec(m[0:3]), _vec(m[3:6]), F32(m[6]), F32(m[7]), F32(m[8]), F32(m[9]), F32(m[10])] for m in mats]])
    for k in range(n):
        fs.vars["push"].store([_mat4(np.eye(4, dtype=np.float32).reshape(16)), _mat4(np.eye(4, dtype=np.float32).reshape(16)), k])
        fs.vars["inPos"].store(_vec(in_pos[k]))
        fs.vars["inNormal"].store(_vec(in_nrm[k]))
        fs.run()
        for j, name in enumerate(("outAlbedo", "outNormal", "outPos", "outEmission", "outMaterialProps")):
            o = list(fs.vars[name].load())
            fout[k, j, :len(o)] = o
    return {"raster/camera": np.asarray(cam, np.float32), "raster/models": models, "raster/pos": pos, "raster/normal": nrm, "raster/vertex_out": vout,
            "raster/materials": mats, "raster/in_pos": in_pos, "raster/in_normal": in_nrm, "raster/fragment_out": fout}


def make_sincos():
    from oracle import binding

    def sincos(a):
        s, c = binding.sincos(np.asarray([a], np.float32))
        return F32(s[0]), F32(c[0])
    return sincos


def load_case(name):
    import raytracer_b200 as rb
    from raytracer_b200.scene_io import load_scene_npz

    W, H, n, camkw = CASES[name]
    scene = load_scene_npz(os.path.join(ROOT, "tests", "golden", "scenes", name + ".npz"))
    return scene, rb.Camera(aspect=W / H, **camkw), W, H, n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "spirv_frames.npz"))
    ap.add_argument("--cases", default=",".join(CASES))
    ap.add_argument("--only-units", action="store_true", help="regenerate the function vectors only")
    a = ap.parse_args()
    from oracle.binding import OracleScene

    sincos = make_sincos()
    out = {}
    if a.only_units:
        a.cases = ""
    for name in [c for c in a.cases.split(",") if c]:
        scene, cam, W, H, n = load_case(name)
        orc = OracleScene(scene).build_bvh()
        vm = ShaderVM(a.ref, scene, orc, sincos)
        frames = np.zeros((n, H, W, 4), np.float32)
        img = None
        t0 = time.time()
        for f in range(n):
            img = vm.render(cam.ubo(), CLEAR, f, W, H, image=img)
            frames[f] = img
        ins = sum(m.stats["instructions"] for m in vm.mod.values())
        print(f"{name}: {W}x{H} x {n} frames, {vm.rays}, {ins} SPIR-V instructions executed, {time.time() - t0:.1f} s", flush=True)
        out[name + "/frames"] = frames
        out[name + "/camera"] = cam.ubo()
        out[name + "/rays"] = np.asarray([vm.rays["closest"], vm.rays["query"], vm.rays["hits"]], np.uint64)
        if name == "conference":  # box lights: the one fixture where the committed hit of the ray query matters
            vm2 = ShaderVM(a.ref, scene, orc, sincos, query_policy="closest")
            img = None
            alt = np.zeros_like(frames)
            for f in range(n):
                img = vm2.render(cam.ubo(), CLEAR, f, W, H, image=img)
                alt[f] = img
            out[name + "/frames_closest_commit"] = alt
            print(f"   a driver committing the CLOSEST hit instead: {int((alt.view(np.uint32) != frames.view(np.uint32)).any(axis=3).sum())} "
                  f"of {n * W * H} pixels differ", flush=True)
    for name in GBUFFER_CASES:   # the G-buffer pass: restir_rgen / restir_rchit / restir_rmiss.spv
        if name not in a.cases.split(","):
            continue
        scene, cam, W, H, n = load_case(name)
        vm = ShaderVM(a.ref, scene, OracleScene(scene).build_bvh(), sincos, prefix="restir_")
        t0 = time.time()
        gb = vm.render_gbuffer(cam.ubo(), W, H)
        print(f"{name}: G-buffer {W}x{H}, {vm.rays}, {time.time() - t0:.1f} s", flush=True)
        for k, v in gb.items():
            out[f"{name}/gbuffer/{k}"] = v
    t0 = time.time()
    out.update(unit_vectors(a.ref, sincos))
    out.update(raster_stage_vectors(a.ref, sincos))
    print(f"function vectors: {len(out['units/wo'])} x (sampleBRDF, evaluateFullBRDF), {time.time() - t0:.1f} s", flush=True)
    out["clear"] = np.asarray(CLEAR, np.float32)
    if a.only_units:   # keep the frames of the existing fixture
        old = dict(np.load(a.out))
        old.update({k: v for k, v in out.items() if k.startswith(("units/", "raster/"))})
        out = old
    np.savez_compressed(a.out, **out)
    print("wrote", a.out)


if __name__ == "__main__":
    main()
