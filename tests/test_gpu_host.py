"""GPU tests (-m gpu) of the rows SURVEY §8(f) ranks next to the hot path, through the C ABI and the host mirror:
the camera / accumulation controller (raytracing_renderer.odin:108-126,194-205,346-350, ui.odin:519-531), checkpoint /
resume of a long accumulation (rt_upload_accum) and the incremental scene updates (gpu_scene.odin:430-601).
Every image is compared bit for bit with the CPU oracle replaying the same sequence of frame numbers.
"""
import numpy as np
import pytest

import raytracer_b200 as rb
from conftest import CLEAR, fixture_scene
from oracle.binding import OracleScene
from raytracer_b200 import librt
from raytracer_b200.renderer import RaytracingRenderer

pytestmark = pytest.mark.gpu


def _cam(aspect=1.0):
    return rb.Camera(position=(0.0, 1.0, 4.5), aspect=aspect, target=(0.0, 1.0, 0.0))


def _oracle_replay(orc, cam, W, H, frames):
    img = None
    for f in frames:
        img = orc.render(cam.ubo(), CLEAR, f, W, H, image=img)
    return img


def _loop(r, n):
    for _ in range(n):  # application.odin:109-115
        r.begin_frame()
        r.render_scene()
        r.end_frame()


def test_renderer_replay_dirty_camera_limit_and_resize():
    scene = fixture_scene("cornell_pbr")
    orc = OracleScene(scene).build_bvh()
    W, H = 64, 48
    cam = _cam(W / H)
    r = RaytracingRenderer(W, H, device=0)
    try:
        r.set_scene(scene)
        r.set_camera(cam)
        assert r.accumulation_frame == 1  # raytracing_renderer.odin:45; the first render resets it (camera.dirty)
        _loop(r, 3)
        assert r.accumulation_frame == 3
        assert np.array_equal(r.output_image().view(np.uint32), _oracle_replay(orc, cam, W, H, range(3)).view(np.uint32))
        # camera move -> dirty -> the next frame is frame 0 again (raytracing_renderer.odin:196-199)
        cam.move("Right", 0.05)
        assert cam.dirty
        _loop(r, 2)
        assert r.accumulation_frame == 2 and not cam.dirty
        assert np.array_equal(r.output_image().view(np.uint32), _oracle_replay(orc, cam, W, H, range(2)).view(np.uint32))
        # accumulation limit: frames stop at the limit, the image stays, the counter does not advance (:204-205, :346-350)
        r.set_accumulation_limit(4)
        _loop(r, 5)
        assert r.accumulation_frame == 4
        at_limit = r.output_image().copy()
        assert np.array_equal(at_limit.view(np.uint32), _oracle_replay(orc, cam, W, H, range(4)).view(np.uint32))
        launches = r.ctx.counters()["launches"]
        _loop(r, 3)
        assert r.ctx.counters()["launches"] == launches and r.accumulation_frame == 4
        assert np.array_equal(r.output_image().view(np.uint32), at_limit.view(np.uint32))
        # a limit at or below the current frame restarts the accumulation (ui.odin:523-526); zero removes it
        r.set_accumulation_limit(2)
        assert r.accumulation_frame == 0
        _loop(r, 3)
        assert r.accumulation_frame == 2
        assert np.array_equal(r.output_image().view(np.uint32), _oracle_replay(orc, cam, W, H, range(2)).view(np.uint32))
        r.set_accumulation_limit(0)
        assert r.accumulation_limit is None
        # resize: new target, accumulation restarts, the camera's aspect follows (raytracing_renderer.odin:108-126)
        W2, H2 = 80, 40
        r.on_resize(W2, H2)
        cam.on_resize(W2 / H2)
        assert r.accumulation_frame == 0
        _loop(r, 2)
        got = r.output_image()
        assert got.shape == (H2, W2, 4)
        assert np.array_equal(got.view(np.uint32), _oracle_replay(orc, cam, W2, H2, range(2)).view(np.uint32))
        # the window's resize flag takes the same route at the head of the next frame (:128-136)
        r.width, r.height, r.framebuffer_resized = W, H, True
        cam.on_resize(W / H)
        _loop(r, 1)
        assert r.output_image().shape == (H, W, 4) and r.accumulation_frame == 1
        assert np.array_equal(r.output_image().view(np.uint32), _oracle_replay(orc, cam, W, H, range(1)).view(np.uint32))
    finally:
        r.destroy()


def test_scene_edits_through_the_change_log_restart_the_accumulation():
    """begin_frame drains Scene.changes into rt_update_* / a rebuild (raytracing_renderer.odin:141-187)."""
    scene = fixture_scene("cornell_pbr")
    W = H = 48
    cam = _cam()
    r = RaytracingRenderer(W, H, device=0)
    try:
        r.set_scene(scene)
        r.set_camera(cam)
        _loop(r, 2)
        scene.update_object_position(3, (0.2, 0.0, 0.1))
        m = scene.materials[2]
        scene.update_material(2, rb.Material(m.name, (0.1, 0.8, 0.2), m.emission_color, m.emission_power, 0.4, 0.0))
        scene.add_material(rb.Material("extra", albedo=(0.9, 0.9, 0.1), roughness=0.7))  # structural: rebuild path
        scene.add_object("extra plane", scene.add_mesh(rb.create_plane()), len(scene.materials) - 1, (0.0, 0.6, 0.0), (90, 0, 0), (0.5, 0.5, 1))
        _loop(r, 2)
        assert r.accumulation_frame == 2 and not scene.changes
        ref = _oracle_replay(OracleScene(scene).build_bvh(), cam, W, H, range(2))
        assert np.array_equal(r.output_image().view(np.uint32), ref.view(np.uint32))
    finally:
        r.destroy()


def test_upload_accum_resumes_a_long_accumulation():
    """trace 8 frames == trace 4, read back, upload into a NEW context, trace frames 4..7 — bit for bit."""
    scene = fixture_scene("cornell_pbr")
    W, H = 96, 64
    cam = _cam(W / H)
    a = librt.Context(device=0)
    b = librt.Context(device=0)
    try:
        for c in (a, b):
            c.build_scene(scene)
            c.resize(W, H)
        a.trace_frames(cam.ubo(), CLEAR, 0, 8)
        want = a.readback().copy()
        a.trace_frames(cam.ubo(), CLEAR, 0, 4)
        checkpoint = a.readback().copy()
        b.upload_accum(checkpoint)
        assert np.array_equal(b.readback().view(np.uint32), checkpoint.view(np.uint32))
        b.trace_frames(cam.ubo(), CLEAR, 4, 4)
        assert np.array_equal(b.readback().view(np.uint32), want.view(np.uint32))
    finally:
        a.close()
        b.close()
    # the same through the renderer mirror, on strips (each context keeps only its own rows of the checkpoint)
    got = np.zeros_like(want)
    for rank in range(2):
        r = RaytracingRenderer(W, H, device=0, strip_rows=8, strip_index=rank, strip_count=2)
        try:
            r.set_scene(scene)
            r.set_camera(cam)
            r.resume_from(checkpoint, 4)
            r.render_frames(4)
            assert r.accumulation_frame == 8
            r.ctx.readback(got)
        finally:
            r.destroy()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_update_object_mesh_matches_rebuild_and_oracle():
    """gpu_scene_update_object_mesh (gpu_scene.odin:485-515): the object points at another mesh of the scene."""
    scene = fixture_scene("cornell_pbr")
    extra = scene.add_mesh(rb.create_sphere(12, 16))
    scene.changes.clear()
    W = H = 64
    cam = _cam()
    c = librt.Context(device=0)
    try:
        c.build_scene(scene)
        c.resize(W, H)
        c.trace_frames(cam.ubo(), CLEAR, 0, 2)
        before = c.readback().copy()
        victim = max(range(len(scene.objects)), key=lambda i: scene.meshes[scene.objects[i].mesh_index].num_triangles)
        scene.update_object_mesh(victim, extra)
        c.update_object_mesh(victim, extra)
        c.trace_frames(cam.ubo(), CLEAR, 0, 2)
        got = c.readback().copy()
        info = c.scene_info()
        assert info.n_triangles == scene.num_instanced_triangles()
        ref = _oracle_replay(OracleScene(scene).build_bvh(), cam, W, H, range(2))
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
        assert not np.array_equal(got, before)
        c.build_scene(scene)  # the same scene from scratch
        c.trace_frames(cam.ubo(), CLEAR, 0, 2)
        assert np.array_equal(c.readback().view(np.uint32), got.view(np.uint32))
        with pytest.raises(librt.RtError):
            c.update_object_mesh(victim, 10 ** 6)
    finally:
        c.close()


def test_transform_update_is_a_refit_and_matches_rebuild_and_oracle():
    """rt_update_object_transform refits the boxes over the existing topology (TLAS update of the reference,
    acceleration_structure.odin:91, gpu_scene.odin:457-483): refit + trace == rebuild + trace == oracle, for a drag of
    several steps on the 262 k-triangle benchmark scene and on an instanced scene; a move that blows the tree up rebuilds."""
    from raytracer_b200 import proxy_scenes

    W, H = 160, 90
    scene = proxy_scenes.sponza_proxy()
    cam = rb.Camera(aspect=W / H, **proxy_scenes.SPONZA_CAMERA)
    c = librt.Context(device=0, max_depth=4)
    try:
        info = c.build_scene(scene)
        c.resize(W, H)
        n_nodes = info.n_nodes
        for step in range(3):   # drag two objects
            for obj in (5, 40):
                mm = np.asarray(scene.objects[obj].transform.model_matrix, np.float32).copy()
                mm[12] += 0.15
                mm[13] += 0.05 * (step + 1)
                scene.objects[obj].transform.model_matrix = mm
                c.update_object_transform(obj, mm)
            c.trace_frames(cam.ubo(), CLEAR, 0, 2)
            info = c.scene_info()
            assert info.refits == step + 1 and info.n_nodes == n_nodes     # no rebuild happened
        got = c.readback().copy()
        ref = _oracle_replay_depth(OracleScene(scene).build_bvh(), cam, W, H, range(2), 4)
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
        # traversal hooks see the refitted tree: closest hits == brute force over the moved triangles
        from conftest import random_rays
        rays = random_rays(4000, 3, lo=(-14, 0, -8), hi=(14, 10, 8), target_lo=(-14, 0, -8), target_hi=(14, 10, 8))
        tuv, ids = c.trace_rays_closest(rays)
        o_tuv, o_ids = OracleScene(scene).build_bvh().trace_rays(rays, mode="bvh")
        assert np.array_equal(ids, o_ids) and np.array_equal(tuv.view(np.uint32), o_tuv.view(np.uint32))
        c.build_scene(scene)    # from scratch
        c.trace_frames(cam.ubo(), CLEAR, 0, 2)
        assert np.array_equal(c.readback().view(np.uint32), got.view(np.uint32))
        assert c.scene_info().refits == 0
        # far away: the root box grows a hundredfold -> the refit reports a degraded tree and the flush rebuilds
        mm = np.asarray(scene.objects[5].transform.model_matrix, np.float32).copy()
        mm[12] += 500.0
        scene.objects[5].transform.model_matrix = mm
        c.update_object_transform(5, mm)
        c.trace_frames(cam.ubo(), CLEAR, 0, 2)
        assert c.scene_info().refits == 0
        ref = _oracle_replay_depth(OracleScene(scene).build_bvh(), cam, W, H, range(2), 4)
        assert np.array_equal(c.readback().view(np.uint32), ref.view(np.uint32))
    finally:
        c.close()
    # instanced meshes (scenes/cornell_box.json shares one plane and one sphere mesh): only the moved instance's triangles change
    scene = fixture_scene("cornell_box")
    cam = rb.Camera(aspect=1.0)
    c = librt.Context(device=0)
    try:
        c.build_scene(scene)
        c.resize(64, 64)
        scene.update_object_position(7, (0.3, -0.2, 0.1))
        c.update_object_transform(7, scene.objects[7].transform.model_matrix)
        c.trace_frames(cam.ubo(), CLEAR, 0, 3)
        assert c.scene_info().refits == 1
        ref = _oracle_replay(OracleScene(scene).build_bvh(), cam, 64, 64, range(3))
        assert np.array_equal(c.readback().view(np.uint32), ref.view(np.uint32))
    finally:
        c.close()


def _oracle_replay_depth(orc, cam, W, H, frames, depth):
    img = None
    for f in frames:
        img = orc.render(cam.ubo(), CLEAR, f, W, H, image=img, max_depth=depth)
    return img


def test_add_and_remove_objects_and_materials_match_a_fresh_build_and_the_oracle():
    """rt_add_object / rt_remove_object / rt_add_material / rt_remove_material (gpu_scene.odin:430-455, dispatched at
    raytracing_renderer.odin:163-183) through the renderer mirror's change log: after every edit the frames equal the oracle
    on the edited scene — and so a context that built that scene from scratch."""
    scene = fixture_scene("cornell_pbr")
    W = H = 56
    cam = _cam()
    r = RaytracingRenderer(W, H, device=0)

    def check(what):
        r.render_frames(2)
        ref = _oracle_replay(OracleScene(scene).build_bvh(), cam, W, H, range(2))
        assert r.accumulation_frame == 2 and not scene.changes
        assert np.array_equal(r.output_image().view(np.uint32), ref.view(np.uint32)), what
        info = r.ctx.scene_info()
        assert (info.n_objects, info.n_materials, info.n_triangles, info.n_lights) == (
            len(scene.objects), len(scene.materials), scene.num_instanced_triangles(), len(scene.lights())), what

    try:
        r.set_scene(scene)
        r.set_camera(cam)
        check("as loaded")
        glow = scene.add_material(rb.Material("glow", albedo=(1, 1, 1), emission_color=(0.3, 0.6, 1.0), emission_power=4.0, roughness=0.8))
        scene.add_object("second lamp", scene.objects[2].mesh_index, glow, (0.3, 0.9, 0.2), (0, 30, 0), (0.3, 0.3, 0.3))
        check("material + emissive object added")           # a second light: NEE now picks between two
        scene.delete_object(1)
        check("object removed")                               # later objects (and lights) move down one index
        victim = scene.objects[3].material_index
        scene.delete_material(victim)                         # last material takes the slot; its users fall back to material 0
        check("material removed")
        scene.delete_object(len(scene.objects) - 1)           # the added lamp goes again
        scene.update_object_position(2, (0.1, 0.0, 0.0))
        check("lamp removed, object moved (refit after rebuild)")
        with pytest.raises(librt.RtError):
            r.ctx.add_object(10 ** 6, 0, np.eye(4, dtype=np.float32).reshape(16))
        with pytest.raises(librt.RtError):
            r.ctx.remove_object(10 ** 6)
        with pytest.raises(librt.RtError):
            r.ctx.remove_material(10 ** 6)
        fresh = librt.Context(device=0)
        fresh.build_scene(scene)
        fresh.resize(W, H)
        fresh.trace_frames(cam.ubo(), CLEAR, 0, 2)
        assert np.array_equal(fresh.readback().view(np.uint32), r.output_image().view(np.uint32))
        fresh.close()
    finally:
        r.destroy()
