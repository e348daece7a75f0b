"""CPU suite, part 1: the oracle itself (runs with -m "not gpu").

The reference holds no tests, golden vectors or CPU path, so the pins available are:
integer-exact RNG against an independent NumPy restatement of shaders/random.glsl:23-53, analytic facts
(create_sphere counts, furnace bound, brute force vs BVH), and the committed golden frames."""
import os

import numpy as np
import pytest

from conftest import CLEAR, GOLDEN, fixture_scene, random_rays, soup_scene

import raytracer_b200 as rb


def _tea_numpy(a, b):  # shaders/random.glsl:23-37, written independently of oracle_math.h
    v0, v1, s = np.uint32(a), np.uint32(b), np.uint32(0)
    with np.errstate(over="ignore"):
        for _ in range(16):
            s = np.uint32(s + np.uint32(0x9E3779B9))
            v0 = np.uint32(v0 + (((v1 << np.uint32(4)) + np.uint32(0xA341316C)) ^ (v1 + s) ^ ((v1 >> np.uint32(5)) + np.uint32(0xC8013EA4))))
            v1 = np.uint32(v1 + (((v0 << np.uint32(4)) + np.uint32(0xAD90777D)) ^ (v0 + s) ^ ((v0 >> np.uint32(5)) + np.uint32(0x7E95761E))))
    return int(v0)


def test_tea_lcg_against_numpy(oracle_lib):
    rng = np.random.default_rng(1)
    for a, b in [(0, 0), (1, 0), (0, 1), (262143, 7), (0xFFFFFFFF, 0xFFFFFFFF)] + [tuple(int(x) for x in rng.integers(0, 2**32, 2)) for _ in range(200)]:
        assert oracle_lib.tea(a, b) == _tea_numpy(a, b)
    seed = _tea_numpy(12345, 3)
    vals, state = oracle_lib.rnd_seq(seed, 64)
    s = seed
    for v in vals:  # random.glsl:41-53
        s = (1664525 * s + 1013904223) & 0xFFFFFFFF
        assert v == np.float32(s & 0x00FFFFFF) / np.float32(16777216.0)
    assert state == s
    assert vals.min() >= 0.0 and vals.max() < 1.0


def test_sincos_accuracy(oracle_lib):
    a = np.linspace(0.0, 2.0 * np.pi, 20001).astype(np.float32)
    s, c = oracle_lib.sincos(a)
    assert np.abs(s - np.sin(a.astype(np.float64))).max() < 3e-7
    assert np.abs(c - np.cos(a.astype(np.float64))).max() < 3e-7


def test_create_sphere_and_plane_counts():
    s = rb.create_sphere()  # scene.odin:250-320: 32 x 32 UV sphere
    assert len(s.vertices) == 994 and len(s.indices) // 3 == 1984
    p = rb.create_plane()
    assert len(p.indices) // 3 == 2


@pytest.mark.parametrize("name", ["cornell_box", "cornell_pbr"])
def test_golden_frames_oracle(oracle_lib, name):
    """The oracle reproduces the committed golden frames and counters (tools/make_fixtures.py)."""
    z = np.load(os.path.join(GOLDEN, f"oracle_{name}.npz"))
    orc = oracle_lib.OracleScene(fixture_scene(name)).build_bvh()
    H, W = z["frames"].shape[1:3]
    img, cnt = None, np.zeros(6, np.uint64)
    for f in range(z["frames"].shape[0]):
        img = orc.render(z["camera"], CLEAR, f, W, H, image=img, counters=cnt)
        assert np.array_equal(img.view(np.uint32), z["frames"][f].view(np.uint32))
    assert [int(x) for x in cnt] == [int(x) for x in z["counters"]]
    nodes, tris, keys, ids = orc.bvh_arrays()
    assert len(nodes) == int(z["n_nodes"]) and orc.bvh_depth == int(z["bvh_depth"])
    assert np.array_equal(ids[:256], z["sorted_ids_head"]) and np.array_equal(keys[:256], z["morton_head"])


@pytest.mark.parametrize("name", ["cornell_box", "cornell_pbr", "conference_mirror_chairs"])
def test_bvh_traversal_equals_brute_force(oracle_lib, name):
    orc = oracle_lib.OracleScene(fixture_scene(name)).build_bvh()
    rays = random_rays(4000, seed=3, lo=(-3, -1, -4), hi=(3, 3, 4))
    t0, i0 = orc.trace_rays(rays, mode="brute")
    t1, i1 = orc.trace_rays(rays, mode="bvh")
    assert np.array_equal(i0, i1) and np.array_equal(t0.view(np.uint32), t1.view(np.uint32))
    assert (i0 != 0xFFFFFFFF).sum() > 100
    rays[:, 7] = 2.5
    ig = np.full(len(rays), 5, np.uint32)
    _, o0 = orc.trace_rays(rays, mode="brute", any_hit=True, ignore_inst=ig)
    _, o1 = orc.trace_rays(rays, mode="bvh", any_hit=True, ignore_inst=ig)
    assert np.array_equal(o0, o1)


def test_bvh_render_equals_brute_force_render(oracle_lib):
    scene = fixture_scene("cornell_pbr")
    orc = oracle_lib.OracleScene(scene).build_bvh()
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=1.0, target=(0.0, 1.0, 0.0))
    a = orc.render(cam.ubo(), CLEAR, 1, 32, 32, use_bvh=True)
    b = orc.render(cam.ubo(), CLEAR, 1, 32, 32, use_bvh=False)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_furnace_bound(oracle_lib):
    """White diffuse closed box, no lights, camera inside: every path ends on the sky term or dies, so the
    radiance is bounded by the clear colour (simple.rgen:106-109 with throughput <= 1 for albedo <= 1)."""
    s = rb.Scene()
    s.add_material(rb.Material("white", albedo=(1.0, 1.0, 1.0), roughness=1.0, metallic=0.0))
    plane = s.add_mesh(rb.create_plane())
    for pos, rot in (((0, -2, 0), (-90, 0, 0)), ((0, 2, 0), (90, 0, 0)), ((0, 0, 2), (0, 180, 0)), ((0, 0, -5), (0, 0, 0)),
                     ((-2, 0, 0), (0, 90, 0)), ((2, 0, 0), (0, -90, 0))):
        s.add_object("wall", plane, 0, pos, rot, (12, 12, 1))
    s.changes.clear()
    orc = oracle_lib.OracleScene(s).build_bvh()
    img = orc.render(rb.Camera(aspect=1.0).ubo(), CLEAR, 2, 48, 48)
    assert np.isfinite(img).all()
    assert (img[..., :3] <= np.asarray(CLEAR, np.float32) * 1.0001 + 1e-6).all()


def test_light_tables_follow_reference_rules(oracle_lib):
    """gpu_scene.odin:366-379: a light is any object whose material has emission_power > 0."""
    scene = fixture_scene("conference")
    orc = oracle_lib.OracleScene(scene)
    assert orc.num_lights == len(scene.lights()) == 4
    assert oracle_lib.OracleScene(fixture_scene("conference_room_large_pbr")).num_lights == 0


def test_soup_oracle_counters_are_sane(oracle_lib):
    s = soup_scene(3000)
    orc = oracle_lib.OracleScene(s).build_bvh()
    cnt = np.zeros(6, np.uint64)
    orc.render(rb.Camera(position=(0.5, 0.5, -2.0), target=(0.5, 0.5, 0.5)).ubo(), CLEAR, 1, 32, 32, counters=cnt)
    ext, sh, nodes, tris, hits, nee = (int(x) for x in cnt)
    assert 32 * 32 <= ext <= 3 * 32 * 32 and hits <= ext and nee <= hits and sh <= nee and nodes >= ext


# ---- pins against the reference's SHIPPED SPIR-V (tests/golden/spirv_constants.json, made by tools/spirv_constants.py) ----
def _source_literals(paths):
    """f32 bit patterns and u32 values of every numeric literal in the given C++ sources (comments stripped)."""
    import re
    import struct
    src = "".join(open(p).read() for p in paths)
    src = re.sub(r"//.*", "", src)
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    f32, u32 = set(), set()
    for m in re.finditer(r"(?<![\w.])(\d+\.\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)f?\b", src):
        f32.add(struct.unpack("<I", struct.pack("<f", float(m.group(1))))[0])
    for m in re.finditer(r"(?<![\w.])(0[xX][0-9a-fA-F]+|\d+)[uU]?\b", src):
        u32.add(int(m.group(1), 0) & 0xFFFFFFFF)
        f32.add(struct.unpack("<I", struct.pack("<f", float(int(m.group(1), 0) & 0xFFFFFFFF)))[0])
    return f32, u32


def _spirv():
    import json
    return json.load(open(os.path.join(GOLDEN, "spirv_constants.json")))


@pytest.mark.parametrize("where", ["oracle", "cuda"])
def test_constants_are_the_shipped_ones(where):
    """SURVEY §8(c) pin (2): every 32-bit float / unsigned literal of the shipped rgen.spv, rchit.spv, restir_rgen.spv
    and restir_rchit.spv (the binaries raytracing_renderer.odin:62-71 loads) appears, bit for bit, as a literal of the
    CPU oracle — and of the CUDA headers, which restate the same math a second time."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if where == "oracle":
        files = [os.path.join(root, "oracle", f) for f in ("oracle_render.cpp", "oracle_math.h", "oracle_scene.h", "oracle_api.cpp")]
    else:
        files = [os.path.join(root, "raytracer_b200", "csrc", f) for f in
                 ("rt_math.cuh", "rt_shade.cuh", "rt_frame.cu", "rt_types.cuh", "rt_internal.h", "rt_trav.cuh", "rt_api.cu", "rt_restir.cu")]
    f32, u32 = _source_literals(files)
    # the firefly clamp min(color, 5) of simple.rgen:121 is a run-time field on both sides (rt_config.radiance_clamp,
    # orc_render's `clamp`), default 5.0 in both bindings: checked below instead of as a literal
    configurable = {0x40A00000}
    spv = _spirv()
    for mod in ("rgen.spv", "rchit.spv", "restir_rgen.spv", "restir_rchit.spv"):
        r = spv[mod]
        missing = [v for v, b in zip(r["f32"], r["f32_bits"])
                   if b not in f32 and (b ^ 0x80000000) not in f32 and b not in configurable]
        assert not missing, f"{mod}: float constants of the shipped shader with no literal in the {where} sources: {missing}"
        missing = [v for v in r["u32"] if v not in u32]
        assert not missing, f"{mod}: unsigned constants of the shipped shader with no literal in the {where} sources: {missing}"
    import inspect
    from oracle.binding import OracleScene
    from raytracer_b200.librt import Context
    assert inspect.signature(OracleScene.render).parameters["clamp"].default == 5.0
    assert inspect.signature(Context.__init__).parameters["radiance_clamp"].default == 5.0
    assert inspect.signature(Context.__init__).parameters["max_depth"].default == 3      # MAX_DEPTH, rgen.spv i32 3


def test_rng_constants_are_the_shipped_ones(oracle_lib):
    """The TEA key schedule / LCG multipliers the NumPy restatement above uses are the ones inside rgen.spv, and the
    24-bit mantissa scale of rnd() (random.glsl:49-53) is there as 16777215 & / 16777216.0."""
    r = _spirv()["rgen.spv"]
    for c in (0x9E3779B9, 0xA341316C, 0xC8013EA4, 0xAD90777D, 0x7E95761E, 1664525, 1013904223, 0x00FFFFFF, 16):
        assert c in r["u32"], hex(c)
    assert 16777216.0 in r["f32"]
    vals, state = oracle_lib.rnd_seq(0x12345678, 1)
    assert state == (1664525 * 0x12345678 + 1013904223) & 0xFFFFFFFF


def test_shipped_call_graph_is_the_restated_one():
    """Which GLSL functions the shipped closest-hit / ray-gen `main` reaches (OpFunctionCall walk of the binaries).
    SURVEY §8(c): brdfHitsSelectedLightRQ, misWeight, lightPdfFromTriangleHitParams and calculatePDF are source-only,
    and NEE goes through estimateDirectLightingMIS -> ..._PowerImportance -> evaluateLightMIS -> sampleLight / isVisibleRQ."""
    spv = _spirv()
    chit = spv["rchit.spv"]
    assert chit["entry_points"] == [["ClosestHitKHR", "main"]]
    for dead in ("brdfHitsSelectedLightRQ", "misWeight", "lightPdfFromTriangleHitParams", "calculatePDF"):
        assert dead not in chit["calls"], dead
    assert chit["calls"]["main"] == ["computeLightSelectionPdf", "cosTheta", "createBasis", "estimateDirectLightingMIS",
                                     "localToWorld", "misWeightPower", "rnd", "sampleBRDF", "worldToLocal"]
    assert chit["calls"]["estimateDirectLightingMIS"] == ["estimateDirectLightingMIS_PowerImportance"]
    assert chit["calls"]["estimateDirectLightingMIS_PowerImportance"] == ["computeLightWeight", "evaluateLightMIS", "rnd"]
    assert {"isVisibleRQ", "sampleLight", "misWeightPower", "evaluateFullBRDF"} <= set(chit["calls"]["evaluateLightMIS"])
    assert chit["calls"]["sampleBRDF"] == ["F_Schlick", "cosTheta", "generateCosineWeightedDirection", "getSpecularProbability",
                                           "microfacetF", "microfacetPDF", "rnd", "sampleGGX"]
    assert sorted(chit["reachable"]) == sorted(chit["calls"])          # nothing else was compiled in
    gen = spv["rgen.spv"]
    assert gen["calls"]["main"] == ["calculateCameraRay", "getSampleOffset", "russianRoulette", "tea"]
    assert gen["calls"]["russianRoulette"] == ["rnd"] and gen["calls"]["getSampleOffset"] == ["rnd"]
    # every function the binaries reach is cited by name in the oracle's sources
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = "".join(open(os.path.join(root, "oracle", f)).read() for f in ("oracle_render.cpp", "oracle_math.h"))
    uncited = [f for f in set(chit["reachable"]) | set(gen["reachable"]) if f != "main" and f not in text]
    assert not uncited, f"functions of the shipped shaders the oracle does not cite: {uncited}"


# ---- the reference's SHIPPED shader binaries, executed (tools/spirv_interp.py) --------------------------------------------
# tests/golden/spirv_frames.npz holds what rgen.spv / rchit.spv / rmiss.spv (and restir_rgen / restir_rchit.spv) compute when
# an interpreter runs them instruction by instruction on the reference's own scene files (tools/spirv_golden.py).  That pins
# evaluation order, association, branch structure and the order of the random draws of the binary — not just its constants.
def _spirv_fixture():
    return np.load(os.path.join(GOLDEN, "spirv_frames.npz"))


SPIRV_CASES = {"cornell_box": (32, 32, 3), "cornell_pbr": (32, 32, 3), "conference": (32, 20, 2), "multi_dispersion": (32, 18, 2)}


@pytest.mark.parametrize("name", sorted(SPIRV_CASES))
def test_oracle_reproduces_the_frames_of_the_shipped_binaries(oracle_lib, name):
    z = _spirv_fixture()
    W, H, n = SPIRV_CASES[name]
    orc = oracle_lib.OracleScene(fixture_scene(name)).build_bvh()
    cnt = np.zeros(6, np.uint64)
    img = None
    for f in range(n):
        img = orc.render(z[name + "/camera"], z["clear"], f, W, H, image=img, counters=cnt)
        want = z[name + "/frames"][f]
        bad = (img.view(np.uint32) != want.view(np.uint32)).any(axis=2)
        assert not bad.any(), f"{name} frame {f}: {int(bad.sum())} pixels differ from what the shipped SPIR-V computes"
    # the same rays: OpTraceRayKHR executions, ray queries, closest-hit shader invocations
    assert [int(cnt[0]), int(cnt[1]), int(cnt[4])] == [int(x) for x in z[name + "/rays"]]
    assert img[..., :3].max() > 0.05


@pytest.mark.parametrize("name", ["cornell_pbr", "conference"])
def test_oracle_reproduces_the_gbuffer_of_the_shipped_binaries(oracle_lib, name):
    z = _spirv_fixture()
    W, H, _ = SPIRV_CASES[name]
    g = oracle_lib.OracleScene(fixture_scene(name)).build_bvh().gbuffer(z[name + "/camera"], W, H)
    for k in ("albedo", "output", "normal_rough", "position_metal", "emission"):
        assert np.array_equal(g[k].view(np.uint8), z[f"{name}/gbuffer/{k}"].view(np.uint8)), k


def test_oracle_reproduces_the_function_vectors_of_rchit_spv(oracle_lib):
    """sampleBRDF / evaluateFullBRDF of rchit.spv (and everything they reach: D_GGX, G_Smith, F_Schlick, microfacetF,
    microfacetPDF, sampleGGX, getSpecularProbability, generateCosineWeightedDirection) on 256 random inputs incl. grazing
    directions and the roughness / metallic extremes: same bits, same seed afterwards."""
    import ctypes as C

    z = _spirv_fixture()
    L = oracle_lib.lib()
    n = len(z["units/wo"])
    assert n >= 256 and 0.1 < z["units/sampleBRDF"][:, 7].mean() < 0.9  # both lobes occur
    for k in range(n):
        m, wo, wi = (np.ascontiguousarray(z[f"units/{q}"][k]) for q in ("materials", "wo", "wi"))
        seed = C.c_uint32(int(z["units/seed_in"][k]))
        out = np.zeros(8, np.float32)
        L.orc_unit_sample_brdf(m.ctypes.data, wo.ctypes.data, float(z["units/xi"][k, 0]), float(z["units/xi"][k, 1]), C.byref(seed), out.ctypes.data)
        assert np.array_equal(out.view(np.uint32), z["units/sampleBRDF"][k].view(np.uint32)), k
        assert seed.value == int(z["units/seed_out"][k])
        o4 = np.zeros(4, np.float32)
        L.orc_unit_eval_brdf(m.ctypes.data, wo.ctypes.data, wi.ctypes.data, o4.ctypes.data)
        assert np.array_equal(o4[:3].view(np.uint32), z["units/evaluateFullBRDF"][k].view(np.uint32)), k


def test_dead_diffuse_rule_of_the_first_hit_filter(oracle_lib):
    """The rule the CUDA path's first-hit filter rests on (rt_frame.cu, wf_shade_kernel DENSE): on a material with
    metallic == 1, a lobe draw that picks the Lambert lobe (`rnd >= specularWeight`, simple.rchit:408) with ry > 0 returns a
    direction above the surface, a value of exactly zero and a finite positive pdf — so the hit shader leaves payload.hit true
    with throughput 0 and the path ends without a colour term.  Checked (a) on the vectors the SHIPPED rchit.spv computed
    (tests/golden/spirv_frames.npz) and (b) on 20 000 random inputs through the oracle, with the filter's own predicate (the
    oracle's specular probability, the LCG draw) deciding which samples the rule is claimed for; incl. grazing and
    below-horizon view directions, roughness 0 and 1, albedo 0."""
    import ctypes as C

    z = _spirv_fixture()
    mats, out = z["units/materials"], z["units/sampleBRDF"]
    metal = mats[:, 8] == 1.0
    lambert = metal & (out[:, 7] == 0.0) & (out[:, 2] > 0.0)
    assert lambert.sum() >= 8  # the fixture holds such draws
    assert (out[lambert, 3:6] == 0.0).all() and np.isfinite(out[lambert, 6]).all() and (out[lambert, 6] > 0.0).all()

    L = oracle_lib.lib()
    rng = np.random.default_rng(7)
    n_claimed = 0
    for k in range(20000):
        m = np.zeros(11, np.float32)
        # every third sample at the bounds rt_filter_qualifies admits (|albedo|, |roughness| <= 1e6), the rest in glTF's [0, 1]
        m[0:3] = rng.choice([0.0, 1.0, 1e6, -1e6, rng.random()], 3) if k % 3 == 0 else rng.random(3)
        m[7] = rng.choice([0.0, 1.0, 1e6, -1e6, rng.random()]) if k % 3 == 0 else rng.choice([0.0, 1.0, rng.random()])   # roughness
        m[8] = 1.0                                       # metallic
        th = rng.choice([0.0, 1e-7, np.pi / 2 - 1e-6, np.pi / 2, np.pi / 2 + 0.3, rng.random() * np.pi / 2])
        ph = rng.random() * 2 * np.pi
        wo = np.array([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)], np.float32)
        rx, ry = np.float32(rng.random()), np.float32(rng.choice([0.0, 2.0 ** -24, rng.random()]))
        seed_in = int(rng.integers(0, 2 ** 32))
        lobe, _ = oracle_lib.rnd_seq(seed_in, 1)
        sw = np.float32(L.orc_unit_specular_probability(m.ctypes.data))
        claimed = (not (lobe[0] < sw)) and ry > 0.0        # the filter's predicate
        seed = C.c_uint32(seed_in)
        o8 = np.zeros(8, np.float32)
        L.orc_unit_sample_brdf(m.ctypes.data, wo.ctypes.data, float(rx), float(ry), C.byref(seed), o8.ctypes.data)
        if claimed:
            n_claimed += 1
            assert o8[7] == 0.0 and o8[2] > 0.0, (k, o8)                      # Lambert lobe, direction above the surface
            assert (o8[3:6] == 0.0).all(), (k, o8)                             # value exactly zero
            assert np.isfinite(o8[6]) and o8[6] > 0.0, (k, o8)                # finite positive pdf -> payload.hit stays true
            thr = np.float32(1.0) * (o8[3:6] * o8[2] / o8[6])                  # simple.rchit:694
            assert (thr == 0.0).all()
    assert n_claimed > 2500


@pytest.mark.skipif(not os.path.isdir("/root/reference/shaders"), reason="reference mount absent (the fixture itself is committed)")
def test_the_fixture_is_what_the_binaries_compute(oracle_lib):
    """Re-executes a sample of pixels from /root/reference/shaders/*.spv: the committed frames really come from the binaries."""
    import sys

    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), os.pardir, "tools"))
    import spirv_golden as G

    z = _spirv_fixture()
    sincos = G.make_sincos()
    for name, px in (("cornell_pbr", [(16, 16), (20, 12), (5, 5), (10, 28), (27, 19)]), ("conference", [(13, 0), (24, 5), (6, 14), (18, 9)])):
        scene, cam, W, H, n = G.load_case(name)
        vm = G.ShaderVM("/root/reference", scene, oracle_lib.OracleScene(scene).build_bvh(), sincos)
        img = z[name + "/frames"][0].copy()   # frame 1 reads the previous mean of the pixels it touches
        vm.render(z[name + "/camera"], z["clear"], 1, W, H, image=img, pixels=px)
        for (x, y) in px:
            assert np.array_equal(img[y, x].view(np.uint32), z[name + "/frames"][1][y, x].view(np.uint32)), (name, x, y)
    # the ray-query policy matters where the fixture says it does: a driver committing the closest hit lights more pixels
    a, b = z["conference/frames"], z["conference/frames_closest_commit"]
    d = (a.view(np.uint32) != b.view(np.uint32)).any(axis=3)
    assert 0 < d.sum() < 0.1 * d.size and (b[d][:, :3].sum(axis=1) >= a[d][:, :3].sum(axis=1) - 1e-6).mean() > 0.9


def test_raster_gbuffer_stages_match_gbuffer_vert_and_frag_spv(oracle_lib):
    """The two programmable stages of the rasterised G-buffer variant against the shipped gbuffer_vert.spv / gbuffer_frag.spv
    (vectors in tests/golden/spirv_frames.npz): the oracle's vertex stage bit for bit — incl. (projection * view) as a matrix
    product FIRST (OpMatrixTimesMatrix in the binary) and the y mirror of gbuffer.vert:24-25 — and the fragment outputs."""
    z = _spirv_fixture()
    L = oracle_lib.lib()
    cam = np.ascontiguousarray(z["raster/camera"])
    for k in range(len(z["raster/pos"])):
        m, p, nrm = (np.ascontiguousarray(z[q][k]) for q in ("raster/models", "raster/pos", "raster/normal"))
        out = np.zeros(10, np.float32)
        L.orc_unit_raster_vertex(m.ctypes.data, p.ctypes.data, nrm.ctypes.data, cam.ctypes.data, out.ctypes.data)
        assert np.array_equal(out.view(np.uint32), z["raster/vertex_out"][k].view(np.uint32)), k
    mats, fo = z["raster/materials"], z["raster/fragment_out"]
    v = z["raster/in_normal"]
    inv = np.float32(1.0) / np.sqrt((v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2])   # normalize of the contract
    emissive = (mats[:, 6] > 0).astype(np.float32)
    want = np.zeros_like(fo)
    want[:, 0, :3], want[:, 0, 3] = mats[:, 0:3], 1.0
    want[:, 1, :3], want[:, 1, 3] = v * inv[:, None], mats[:, 8]
    want[:, 2, :3], want[:, 2, 3] = z["raster/in_pos"], mats[:, 7]
    want[:, 3, :3], want[:, 3, 3] = mats[:, 6:7] * mats[:, 3:6], emissive
    want[:, 4, 0], want[:, 4, 1], want[:, 4, 2] = mats[:, 7], mats[:, 8], emissive
    assert np.array_equal(want.view(np.uint32), fo.view(np.uint32))
    assert 0.2 < emissive.mean() < 0.8


def test_raster_gbuffer_twin_agrees_with_the_ray_traced_gbuffer_on_a_mirrored_scene(oracle_lib):
    """Sanity of the fixed-function rules this repo states for the raster variant: gbuffer.vert mirrors the geometry in y, so
    rasterising a scene must show what the ray-traced G-buffer shows of the y-mirrored scene.  restir.rgen shoots through pixel
    CORNERS; at twice the resolution the corners of the odd pixels are the centres the rasteriser samples, so the two passes
    look at the same points: same coverage, same surface positions to float accuracy."""
    scene = fixture_scene("cornell_pbr")
    W = H = 96
    cam = rb.Camera(position=(0.0, -1.0, 4.5), aspect=1.0, target=(0.0, -1.0, 0.0))
    ras = oracle_lib.OracleScene(scene).raster_gbuffer(cam.ubo(), W, H)
    mirrored = fixture_scene("cornell_pbr")
    for m in mirrored.meshes:
        m.vertices["pos"][:, 1] *= -1
        m.vertices["normal"][:, 1] *= -1
    # mirroring object space == mirroring the world only for objects whose model matrix commutes with the y flip; cornell_pbr's do
    rt_img = {k: v[1::2, 1::2] for k, v in oracle_lib.OracleScene(mirrored).build_bvh().gbuffer(cam.ubo(), 2 * W, 2 * H).items()}
    cov_r = ras["albedo"][..., 3] > 0
    cov_t = rt_img["position_metal"][..., :3].any(axis=2)
    assert (cov_r == cov_t).mean() > 0.995 and cov_r.mean() > 0.3
    both = cov_r & cov_t
    d = np.linalg.norm(ras["position_rough"][..., :3] - rt_img["position_metal"][..., :3], axis=2)[both]
    assert np.percentile(d, 99) < 1e-3 and np.median(d) < 1e-5, (np.median(d), np.percentile(d, 99))
    assert np.abs(ras["albedo"][..., :3][both] - rt_img["albedo"][..., :3][both] / 255.0).max(axis=1).mean() < 0.01
