"""CPU suite, part 2: the host-side mirror of the reference's scene / camera / loader interfaces."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, fixture_scene

import raytracer_b200 as rb
from raytracer_b200 import multi_gpu, proxy_scenes
from raytracer_b200.scene_io import load_scene_npz, save_scene_npz

REF = "/root/reference"  # present in the build container only; tests that read it skip elsewhere


def test_record_layouts_match_the_reference():
    assert rb.MATERIAL_DTYPE.itemsize == 44  # gpu_scene.odin:17-21 / ray_common.glsl:3-11
    assert rb.VERTEX_DTYPE.itemsize == 24    # scene.odin:10-13 / ray_common.glsl:48-51
    s = fixture_scene("cornell_box")
    mesh_idx, mat_idx, model = s.packed_objects()
    assert model.dtype == np.float32 and model.shape == (len(s.objects), 16)
    assert s.packed_materials().nbytes == 44 * len(s.materials)


def test_default_camera_is_the_applications():
    c = rb.Camera(aspect=1280 / 1020)  # application.odin:52, camera.odin:45-85
    ubo = c.ubo().reshape(4, 4, 4)     # proj, view, inverse_view, inverse_proj; column-major
    assert np.allclose(ubo[2][3], [0, 0, -3, 1])          # camera position = translation column of inverse_view
    assert ubo[0][1][1] < 0                                # proj[1][1] *= -1 (camera.odin:80)
    assert np.allclose(ubo[0].T @ ubo[3].T, np.eye(4), atol=1e-5)
    assert np.allclose(ubo[1].T @ ubo[2].T, np.eye(4), atol=1e-5)
    t = np.tan(np.radians(45.0) / 2)
    assert np.isclose(ubo[0][1][1], -1 / t, rtol=1e-6) and np.isclose(ubo[0][0][0], 1 / (t * 1280 / 1020), rtol=1e-6)


def test_model_matrix_order_is_t_rx_ry_rz_s():
    m = rb.model_matrix_from_trs((1, 2, 3), (90, 0, 0), (2, 2, 2)).reshape(4, 4).T  # scene.odin:213-224
    p = m @ np.array([0, 1, 0, 1], np.float32)
    assert np.allclose(p, [1, 2, 5, 1], atol=1e-5)  # scale 2, rotate +90 deg about x: y -> z, then translate


def test_change_log_mirrors_scene_odin():
    s = rb.Scene()
    a = s.add_material(rb.Material("a"))
    m = s.add_mesh(rb.create_plane())
    s.add_object("o", m, a)
    s.changes.clear()
    s.update_material(a, rb.Material("b", roughness=0.2))
    s.update_object_position(0, (1, 0, 0))
    s.update_object_material(0, a)
    s.update_object_mesh(0, m)
    kinds = [c.type for c in s.changes]
    T = rb.SceneChangeType
    assert kinds == [T.Material_Changed, T.Object_Transform_Changed, T.Object_Material_Changed, T.Object_Mesh_Changed]


@pytest.mark.parametrize("name,objects,tris,lights", [("cornell_box", 9, 5964, 1), ("cornell_pbr", 9, 8096, 1),
                                                       ("conference_room_large_pbr", 39, 468, 0), ("conference", None, None, 4),
                                                       ("multi_dispersion", 51, 9002, 0)])
def test_fixture_scenes_have_the_surveyed_shape(name, objects, tris, lights):
    s = fixture_scene(name)
    if objects is not None:
        assert len(s.objects) == objects and s.num_instanced_triangles() == tris
    assert len(s.lights()) == lights


def test_scene_npz_round_trip(tmp_path):
    s = fixture_scene("cornell_pbr")
    p = str(tmp_path / "x.npz")
    save_scene_npz(p, s)
    b = load_scene_npz(p)
    assert np.array_equal(b.packed_materials(), s.packed_materials())
    for x, y in zip(b.packed_objects(), s.packed_objects()):
        assert np.array_equal(x, y)
    assert all(np.array_equal(m0.vertices, m1.vertices) and np.array_equal(m0.indices, m1.indices) for m0, m1 in zip(b.meshes, s.meshes))


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference mount absent")
@pytest.mark.parametrize("name,rel", [("cornell_box", "scenes/cornell_box.json"), ("cornell_pbr", "scenes/cornell_pbr.glb")])
def test_loaders_reproduce_the_committed_fixtures(name, rel):
    """scene_loader.odin:37-190 (glTF/GLB) and :193-253 (JSON) restated; fixtures were made with them."""
    s = rb.load_scene(os.path.join(REF, rel))
    f = fixture_scene(name)
    assert np.array_equal(s.packed_materials(), f.packed_materials())
    for x, y in zip(s.packed_objects(), f.packed_objects()):
        assert np.array_equal(x, y)


def test_loader_errors():
    with pytest.raises(rb.SceneLoadError):
        rb.load_scene_from_gltf("/nonexistent/file.glb")
    with pytest.raises(rb.SceneLoadError):
        rb.load_scene_from_gltf(os.path.join(GOLDEN, "oracle_cornell_box.npz"))  # wrong extension -> Invalid_File


def test_sponza_proxy_keeps_the_gltf_layout():
    s = proxy_scenes.sponza_proxy()
    assert len(s.objects) == 103 and s.num_instanced_triangles() == 262267 and len(s.materials) == 25 and not s.lights()


@pytest.mark.parametrize("world,height,strip", [(1, 17, 8), (2, 1080, 8), (4, 1080, 8), (8, 1080, 8), (8, 61, 4), (3, 5, 8)])
def test_strip_partition_is_a_partition(world, height, strip):
    rows = [multi_gpu.strip_rows_of(r, world, height, strip) for r in range(world)]
    allr = np.sort(np.concatenate(rows))
    assert np.array_equal(allr, np.arange(height))
    parts = [np.full((len(r), 3, 1), k, np.int32) for k, r in enumerate(rows)]
    img = multi_gpu.assemble(parts, height, strip)
    for k, r in enumerate(rows):
        assert (img[r] == k).all()


def test_png_round_trip_and_ssim_restatement(tmp_path):
    """image_io.ssim restates skimage's structural_similarity(channel_axis=2) (ssim_compare.py:20); checked here
    against a direct windowed evaluation of the same formula."""
    from raytracer_b200.image_io import read_png, ssim, write_pfm, write_png

    rng = np.random.default_rng(0)
    a = (rng.random((24, 31, 3)) * 255).astype(np.uint8)
    b = a.copy()
    b[5:15, 8:20] = (b[5:15, 8:20].astype(np.int32) * 2 // 3).astype(np.uint8)
    p = str(tmp_path / "a.png")
    write_png(p, a)
    assert np.array_equal(read_png(p), a)
    write_pfm(str(tmp_path / "a.pfm"), a.astype(np.float32) / 255)
    assert ssim(a, a) == 1.0
    c1, c2, acc = (0.01 * 255) ** 2, (0.03 * 255) ** 2, []
    for ch in range(3):
        x, y = a[..., ch].astype(np.float64), b[..., ch].astype(np.float64)
        for i in range(3, 24 - 3):
            for j in range(3, 31 - 3):
                wx, wy = x[i - 3:i + 4, j - 3:j + 4].ravel(), y[i - 3:i + 4, j - 3:j + 4].ravel()
                ux, uy = wx.mean(), wy.mean()
                vx, vy = wx.var(ddof=1), wy.var(ddof=1)
                vxy = ((wx - ux) * (wy - uy)).sum() / 48.0
                acc.append(((2 * ux * uy + c1) * (2 * vxy + c2)) / ((ux * ux + uy * uy + c1) * (vx + vy + c2)))
    assert abs(ssim(a, b) - float(np.mean(acc))) < 1e-9
    assert 0.5 < ssim(a, b) < 1.0


class _RecorderCtx:
    """Stands in for librt.Context in the CPU test of the frame logic: records the frame numbers the controller asks for."""

    def __init__(self):
        self.calls = []

    def resize(self, w, h):
        self.calls.append(("resize", w, h))

    def build_scene(self, scene):
        self.calls.append(("build", len(scene.objects)))

    def trace_frame(self, ubo, clear, frame):
        self.calls.append(("trace", frame, 1))

    def trace_frames(self, ubo, clear, first, n):
        self.calls.append(("trace", first, n))

    def upload_accum(self, image):
        self.calls.append(("upload", image.shape))

    def __getattr__(self, name):  # update_material, update_object_* ...
        return lambda *a: self.calls.append((name,) + tuple(x for x in a if isinstance(x, int)))


def test_accumulation_controller_mirrors_the_reference():
    """raytracing_renderer.odin:128-205,346-350 and the limit dialog ui.odin:519-531, on a recorder (no GPU)."""
    from raytracer_b200.renderer import RaytracingRenderer

    rec = _RecorderCtx()
    r = RaytracingRenderer(32, 16, ctx=rec)
    scene = fixture_scene("cornell_box")
    cam = rb.Camera(aspect=2.0)
    r.set_scene(scene)
    r.set_camera(cam)

    def loop(n):
        for _ in range(n):
            r.begin_frame()
            r.render_scene()
            r.end_frame()

    def frames():
        out = [c[1] for c in rec.calls if c[0] == "trace"]
        rec.calls.clear()
        return out

    loop(3)
    assert frames() == [0, 1, 2]            # the camera starts dirty: frame 0 first
    cam.move("Forward", 0.1)
    loop(2)
    assert frames() == [0, 1]
    r.set_accumulation_limit(3)
    loop(4)
    assert frames() == [2] and r.accumulation_frame == 3   # stops AT the limit, counter frozen
    r.set_accumulation_limit(2)                               # limit <= current frame: restart
    loop(3)
    assert frames() == [0, 1]
    r.set_accumulation_limit(-1)
    assert r.accumulation_limit is None
    loop(1)
    assert frames() == [2]
    scene.update_object_position(0, (0.1, 0, 0))              # a scene edit restarts too (:141-187)
    loop(1)
    assert ("update_object_transform", 0) in rec.calls and frames() == [0]
    r.on_resize(40, 20)
    loop(1)
    assert rec.calls[0] == ("resize", 40, 20) and frames() == [0]
    r.resume_from(np.zeros((20, 40, 4), np.float32), 100)     # checkpoint/resume: continue with frame 100
    r.render_frames(5)
    assert frames() == [100] and r.accumulation_frame == 105
    r.set_accumulation_limit(107)
    r.render_frames(10)                                        # clipped at the limit
    assert [c for c in rec.calls if c[0] == "trace"] == [("trace", 105, 2)]


def test_strip_balancer_keeps_the_partition_and_converges():
    """multi_gpu.StripBalancer on a synthetic cost field with 8 % per-strip model error: every table is a partition with
    the round-robin strip counts, the partial last strip never moves, and the measured slowest/mean ratio drops."""
    H, world, sr = 1084, 8, 8   # 135 full strips + a partial one
    rng = np.random.default_rng(3)
    n_strips, n_full = (H + sr - 1) // sr, H // sr
    true = 1.0 + 0.5 * np.sin(np.arange(n_strips) / 7.0) + 0.2 * rng.random(n_strips)
    model = true * (1.0 + 0.08 * rng.standard_normal(n_strips))
    b = multi_gpu.StripBalancer(H, world, sr, costs=model)
    rr = np.arange(n_strips) % world
    assert np.array_equal(b.owners, rr)
    ratios = []
    for _ in range(8):
        t = np.array([true[b.owners == r].sum() for r in range(world)])
        ratios.append(t.max() / t.mean())
        o = b.update(t)
        assert [int((o[:n_full] == r).sum()) for r in range(world)] == [int((rr[:n_full] == r).sum()) for r in range(world)]
        assert o[n_full] == n_full % world
    assert b.best[0] < 1.012 < ratios[0]
    assert [int((b.best_owners() == r).sum()) for r in range(world)] == [int((rr == r).sum()) for r in range(world)]


def test_which_objects_the_first_hit_filter_may_end_early():
    """rt_filter_qualifies (host-only): the filter ends a path at a Lambert-lobe draw only on objects for which
    every intermediate of the hit shader is provably finite — a NaN normal would turn `payload.hit` false and the pixel would get
    the sky colour instead of nothing (simple.rchit:690-700, simple.rgen:106-118).  The checks are conservative: what fails them
    simply goes through the hit shader."""
    import raytracer_b200 as rb
    from raytracer_b200 import librt, proxy_scenes

    metal = rb.Material("metal", albedo=(0.9, 0.7, 0.4), roughness=0.6, metallic=1.0)
    eye = np.eye(4, dtype=np.float32).T.reshape(16)
    sphere, plane = rb.create_sphere(), rb.create_plane()
    assert librt.filter_qualifies(sphere, eye, metal) and librt.filter_qualifies(plane, eye, metal)
    # materials: anything but metallic == 1 keeps a non-zero Lambert lobe; non-finite parameters are left to the shader
    for kw in (dict(metallic=0.999), dict(metallic=0.0), dict(metallic=1.0, roughness=float("nan")), dict(metallic=1.0, albedo=(0.5, float("inf"), 0.5)),
               dict(metallic=1.0, roughness=1e20), dict(metallic=1.0, albedo=(0.5, 0.5, -3e7))):  # alpha^2 would overflow D_GGX
        assert not librt.filter_qualifies(sphere, eye, rb.Material("m", **{**dict(albedo=(0.5, 0.5, 0.5), roughness=0.5), **kw}))
    # meshes: no normals, non-finite normals, normals that can cancel inside a triangle, normals of the wrong length
    for edit in (lambda n: n.__setitem__(slice(None), 0.0), lambda n: n.__setitem__((3, 1), np.nan), lambda n: n.__setitem__(slice(None, None, 2), -n[::2]),
                 lambda n: n.__setitem__(slice(None), n * 3.0)):
        m = rb.create_sphere()
        edit(m.vertices["normal"])
        assert not librt.filter_qualifies(m, eye, metal)
    # matrices: uniform scales of ordinary size pass, a squash by 1e-5, a singular, a huge and a non-finite matrix do not
    def scaled(sx, sy, sz):
        return np.diag([sx, sy, sz, 1.0]).astype(np.float32).reshape(16)
    assert librt.filter_qualifies(sphere, scaled(0.008, 0.008, 0.008), metal) and librt.filter_qualifies(sphere, scaled(50, 20, 80), metal)
    for mm in (scaled(0.5, 1e-5, 0.5), scaled(1, 0, 1), scaled(1e7, 1e7, 1e7), scaled(np.nan, 1, 1)):
        assert not librt.filter_qualifies(sphere, mm, metal)
    # the benchmarked scene: the reference's loader leaves every glTF material metallic, most of the proxy qualifies
    s = proxy_scenes.sponza_proxy()
    q = [librt.filter_qualifies(s.meshes[o.mesh_index], o.transform.model_matrix, s.materials[o.material_index]) for o in s.objects]
    assert 0.75 * len(q) < sum(q) < len(q)
