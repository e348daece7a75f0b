"""GPU tests (-m gpu) of the multi-GPU path BEHIND the C ABI: one rt_ctx created with
rt_config.device_ids[] fans every call out to one sub-context per device and rt_readback gathers the strips, so the host's
three calls (build / trace / consume: gpu_scene.odin:45-47, raytracing_renderer.odin:204-270, :271-339) stay three calls.

A device may be listed twice, which puts the whole fan-out (two sub-contexts, two strip sets, the gather) under test on a
one-GPU box; the two-device test proper skips unless the box has two GPUs.  Every image is compared bit for bit with a
single-device context and, for the benchmarked scene, with the CPU oracle.
"""
import numpy as np
import pytest

import raytracer_b200 as rb
from conftest import CLEAR, fixture_scene
from oracle.binding import OracleScene
from raytracer_b200 import librt, proxy_scenes

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch

    return torch.cuda.device_count()


def _pinned(shape, dtype=np.float32):
    import torch

    t = torch.zeros(shape, dtype=getattr(torch, np.dtype(dtype).name), pin_memory=True)
    return t, t.numpy()


def _single(scene, cam, W, H, n, depth=3):
    c = librt.Context(device=0, max_depth=depth, rr_start_depth=3, frames_in_flight=n)
    try:
        c.build_scene(scene)
        c.resize(W, H)
        c.trace_frames(cam.ubo(), CLEAR, 0, n)
        return c.readback().copy(), c.counters(), c.readback_srgb8().copy()
    finally:
        c.close()


@pytest.mark.parametrize("devices", [[0, 0], [0, 0, 0]], ids=["2 slots", "3 slots"])
def test_fanout_context_equals_single_context(devices):
    scene = fixture_scene("cornell_pbr")
    W, H, n = 96, 52, 3  # not a multiple of the strip height: one rank gets the partial strip
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
    want, cnt, want8 = _single(scene, cam, W, H, n)
    c = librt.Context(devices=devices, frames_in_flight=n)
    try:
        info = c.build_scene(scene)
        assert info.n_triangles == scene.num_instanced_triangles()
        c.resize(W, H)
        c.event_record(0)
        c.trace_frames(cam.ubo(), CLEAR, 0, n)
        c.event_record(1)
        assert c.event_elapsed_ms() > 0.0 and c.last_trace_ms() > 0.0
        got = c.readback()                      # pageable destination: staged copies per device
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        keep, pinned = _pinned((H, W, 4))       # pinned destination: every device scatters its rows itself
        c.readback(pinned)
        assert np.array_equal(pinned.view(np.uint32), want.view(np.uint32))
        dev = c.counters()
        for k in ("extend_rays", "shadow_rays", "hits", "nee"):
            assert dev[k] == cnt[k], k           # the same rays, summed over the devices
        assert np.array_equal(c.readback_srgb8(), want8)
        owners = c.strip_owners()
        assert list(owners) == [s % len(devices) for s in range((H + 7) // 8)]
        # running mean continues across calls; single-frame dispatch == batch
        c.trace_frame(cam.ubo(), CLEAR, n)
        one = librt.Context(device=0)
        one.build_scene(scene)
        one.resize(W, H)
        one.trace_frames(cam.ubo(), CLEAR, 0, n + 1)
        assert np.array_equal(c.readback().view(np.uint32), one.readback().view(np.uint32))
        # G-buffer and the verification hooks through the same context
        c.trace_gbuffer(cam.ubo())
        one.trace_gbuffer(cam.ubo())
        g, g1 = c.readback_gbuffer(), one.readback_gbuffer()
        for k in g:
            assert np.array_equal(g[k].view(np.uint8), g1[k].view(np.uint8)), k
        nodes, _, _, _ = c.get_bvh()
        nodes1, _, _, _ = one.get_bvh()
        assert np.array_equal(nodes, nodes1)
        one.close()
        with pytest.raises(librt.RtError):
            c.device_image()
    finally:
        c.close()


def test_fanout_incremental_updates_and_resume():
    scene = fixture_scene("cornell_pbr")
    W = H = 64
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=1.0, target=(0.0, 1.0, 0.0))
    c = librt.Context(devices=[0, 0])
    one = librt.Context(device=0)
    try:
        for x in (c, one):
            x.build_scene(scene)
            x.resize(W, H)
        mm = np.asarray(scene.objects[3].transform.model_matrix, np.float32).copy()
        mm[12] += 0.25
        m = scene.materials[1]
        m2 = rb.Material(m.name, m.albedo, (0.2, 0.5, 1.0), 3.0, m.roughness, m.metallic)
        for x in (c, one):
            x.update_object_transform(3, mm)
            x.update_material(1, m2)
            x.update_object_material(4, 2)
            x.trace_frames(cam.ubo(), CLEAR, 0, 4)
        half = one.readback().copy()
        assert np.array_equal(c.readback().view(np.uint32), half.view(np.uint32))
        c.upload_accum(half)   # every device takes its own rows of the checkpoint
        for x in (c, one):
            x.trace_frames(cam.ubo(), CLEAR, 4, 2)
        assert np.array_equal(c.readback().view(np.uint32), one.readback().view(np.uint32))
    finally:
        c.close()
        one.close()


def test_fanout_benchmark_scene_vs_oracle():
    """sponza-proxy depth 8 / roulette 3 through a 2-slot multi-GPU context against the CPU oracle."""
    W, H, n = 320, 180, 4
    scene = proxy_scenes.sponza_proxy()
    cam = rb.Camera(aspect=W / H, **proxy_scenes.SPONZA_CAMERA)
    orc = OracleScene(scene).build_bvh()
    ref = None
    for f in range(n):
        ref = orc.render(cam.ubo(), CLEAR, f, W, H, image=ref, max_depth=8)
    for balance in (librt.RT_BALANCE_ROUND_ROBIN, librt.RT_BALANCE_COST):
        c = librt.Context(devices=[0, 0], max_depth=8, rr_start_depth=3, frames_in_flight=n, strip_balance=balance)
        try:
            c.build_scene(scene)
            c.resize(W, H)
            c.trace_frames(cam.ubo(), CLEAR, 0, n)
            got = c.readback()
            assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), balance
        finally:
            c.close()


def test_cost_balanced_strips_are_a_partition_and_bit_identical():
    """RT_BALANCE_COST on strip contexts (the torchrun path: one context per rank): every rank derives the same owner table,
    keeps its round-robin strip count, and the union of the ranks' rows is the single-context image."""
    W, H, n, world = 480, 270, 3, 4
    scene = proxy_scenes.sponza_proxy()
    cam = rb.Camera(aspect=W / H, **proxy_scenes.SPONZA_CAMERA)
    want, _, _ = _single(scene, cam, W, H, n, depth=8)
    got = np.zeros_like(want)
    tables = []
    for rank in range(world):
        c = librt.Context(device=0, max_depth=8, rr_start_depth=3, strip_index=rank, strip_count=world, frames_in_flight=n,
                          strip_balance=librt.RT_BALANCE_COST)
        try:
            c.build_scene(scene)
            c.resize(W, H)
            assert list(c.strip_owners()) == [s % world for s in range((H + 7) // 8)]  # before the first frame: round robin
            c.trace_frames(cam.ubo(), CLEAR, 0, n)
            c.readback(got)
            tables.append(c.strip_owners())
            rows = c.device_image()[1]
            assert rows == sum(min(8, H - 8 * s) for s in range((H + 7) // 8) if s % world == rank)
            # the table stays put while the accumulation runs, and a resume keeps it too
            c.trace_frames(cam.ubo(), CLEAR, n, 1)
            assert np.array_equal(c.strip_owners(), tables[-1])
        finally:
            c.close()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    for t in tables[1:]:
        assert np.array_equal(t, tables[0])
    owners = tables[0]
    n_strips = (H + 7) // 8
    assert len(owners) == n_strips and owners.max() == world - 1
    assert [int((owners == r).sum()) for r in range(world)] == [len(range(r, n_strips, world)) for r in range(world)]
    assert owners[-1] == (n_strips - 1) % world          # the partial last strip stays with its round-robin owner
    assert not np.array_equal(owners, np.arange(n_strips) % world)  # the costs of this scene are not uniform


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs on the box")
def test_two_devices_through_the_c_abi_only():
    """Two physical devices driven through the C ABI from this one host thread: build on both, trace, gather."""
    n_dev = min(_n_gpus(), 8)
    W, H, n = 640, 360, 4
    scene = proxy_scenes.sponza_proxy()
    cam = rb.Camera(aspect=W / H, **proxy_scenes.SPONZA_CAMERA)
    want, cnt, _ = _single(scene, cam, W, H, n, depth=8)
    for balance in (librt.RT_BALANCE_ROUND_ROBIN, librt.RT_BALANCE_COST):
        c = librt.Context(devices=list(range(n_dev)), max_depth=8, rr_start_depth=3, frames_in_flight=n, strip_balance=balance)
        try:
            c.build_scene(scene)
            c.resize(W, H)
            c.trace_frames(cam.ubo(), CLEAR, 0, n)
            keep, pinned = _pinned((H, W, 4))
            c.readback(pinned)
            assert np.array_equal(pinned.view(np.uint32), want.view(np.uint32))
            assert np.array_equal(c.readback().view(np.uint32), want.view(np.uint32))
            dev = c.counters()
            assert dev["extend_rays"] == cnt["extend_rays"] and dev["hits"] == cnt["hits"]
        finally:
            c.close()


def test_host_decided_strip_table_is_bit_identical_and_validated():
    """rt_set_strip_owners: any table that keeps every rank's round-robin strip count gives the single-context image (the
    multi-GPU context and strip contexts alike); a table that changes a rank's row count is refused."""
    scene = fixture_scene("cornell_pbr")
    W, H, n, world = 96, 84, 3, 3   # 10 full strips + a partial one
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
    want, _, _ = _single(scene, cam, W, H, n)
    table = np.array([2, 0, 1, 0, 1, 2, 0, 2, 1, 0, 1], np.uint32)  # counts 4 / 3 / 3 as round robin; strip 10 (partial) stays with rank 1
    c = librt.Context(devices=[0, 0, 0], frames_in_flight=n)
    try:
        c.build_scene(scene)
        c.resize(W, H)
        c.trace_frames(cam.ubo(), CLEAR, 0, n)
        c.set_strip_owners(table)
        assert np.array_equal(c.strip_owners(), table)
        c.trace_frames(cam.ubo(), CLEAR, 0, n)   # the accumulation restarts: rows changed hands
        assert np.array_equal(c.readback().view(np.uint32), want.view(np.uint32))
        ms = c.event_elapsed_per_device() if False else None
        bad = table.copy()
        bad[0] = 0                                # rank 0 would own 5 strips
        with pytest.raises(librt.RtError):
            c.set_strip_owners(bad)
        with pytest.raises(librt.RtError):
            c.set_strip_owners(table[:-1])
        cost = c.measure_row_costs(cam.ubo(), 0)
        assert cost.shape == (H, 4) and cost[:, 0].min() > 0 and cost[:, 2].max() <= 3 * W
    finally:
        c.close()
    got = np.zeros_like(want)
    for rank in range(world):
        s = librt.Context(device=0, strip_index=rank, strip_count=world, frames_in_flight=n)
        try:
            s.build_scene(scene)
            s.resize(W, H)
            s.set_strip_owners(table)
            s.event_record(0)
            s.trace_frames(cam.ubo(), CLEAR, 0, n)
            s.event_record(1)
            assert len(s.event_elapsed_per_device()) == 1
            s.readback(got)
        finally:
            s.close()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def _restir_frames(c, cam, n, prm=None):
    for f in range(n):
        c.restir_frame(cam.ubo(), CLEAR, f, prm)
    return c.readback().copy(), c.readback_reservoirs().copy()


@pytest.mark.parametrize("mode", ["strip contexts", "multi-GPU context", "multi-GPU context, strips by cost"])
def test_restir_image_does_not_depend_on_the_partition(mode):
    """ReSTIR DI with temporal + spatial reuse on a strip partition == the one-context frame, bit for bit, reservoirs
    included: every rank recomputes surfaces, candidates and temporal reuse for the halo of ceil(spatial_radius) rows
    around its strips, so spatial reuse crosses strip borders exactly as it does on one GPU."""
    scene = fixture_scene("conference")
    W, H, n, world = 96, 84, 5, 3
    cam = rb.Camera(position=(-3.5, -2.0, 1.6), aspect=W / H, target=(1.0, 0.5, 1.0), up=(0.0, 0.0, 1.0))
    one = librt.Context(device=0)
    one.build_scene(scene)
    one.resize(W, H)
    want, want_res = _restir_frames(one, cam, n)
    one.close()
    assert want_res["M"].max() > 8 and want[..., :3].max() > 0.05  # history and reuse were at work
    if mode == "strip contexts":
        got, got_res = np.zeros_like(want), np.zeros_like(want_res)
        for rank in range(world):
            c = librt.Context(device=0, strip_index=rank, strip_count=world)
            try:
                c.build_scene(scene)
                c.resize(W, H)
                for f in range(n):
                    c.restir_frame(cam.ubo(), CLEAR, f)
                c.readback(got)
                c.L.rt_readback_reservoirs(c.h, got_res.ctypes.data, got_res.nbytes)
            finally:
                c.close()
    else:
        c = librt.Context(devices=[0] * world, strip_balance=librt.RT_BALANCE_COST if "cost" in mode else librt.RT_BALANCE_ROUND_ROBIN)
        try:
            c.build_scene(scene)
            c.resize(W, H)
            if "cost" in mode:   # a path-traced frame 0 deals the strips by cost first; ReSTIR then runs on that table
                c.trace_frame(cam.ubo(), CLEAR, 0)
                assert not np.array_equal(c.strip_owners(), np.arange((H + 7) // 8) % world)
            got, got_res = _restir_frames(c, cam, n)
        finally:
            c.close()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(got_res.view(np.uint8), want_res.view(np.uint8))
    # a smaller radius after a larger one keeps working (the halo list shrinks), and a larger one restarts the history
    c = librt.Context(device=0, strip_index=1, strip_count=world)
    try:
        c.build_scene(scene)
        c.resize(W, H)
        prm = librt.RtRestirParams()
        c.L.rt_default_restir_params(prm)
        for radius in (4.0, 24.0, 8.0):
            prm.spatial_radius = radius
            c.restir_frame(cam.ubo(), CLEAR, 0, prm)
        assert np.isfinite(c.readback()).all()
    finally:
        c.close()


def test_a_plain_c_host_drives_the_gpus_through_the_abi(tmp_path):
    """examples/host_three_calls.c (gcc, no Python in the process): build / trace / consume with the GPUs listed in
    rt_config.device_ids.  Its image file must equal the ctypes single-context image bit for bit."""
    import subprocess

    from test_abi import _build_c_host

    scene = fixture_scene("cornell_pbr")
    W, H, n, depth = 96, 52, 3, 3
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
    want, _, _ = _single(scene, cam, W, H, n, depth)
    blob = tmp_path / "scene.bin"
    with open(blob, "wb") as f:
        f.write(np.asarray([len(scene.meshes), len(scene.objects), len(scene.materials), W, H, n, depth], np.uint32).tobytes())
        for m in scene.meshes:
            f.write(np.asarray([len(m.vertices), len(m.indices)], np.uint32).tobytes())
            f.write(np.ascontiguousarray(m.vertices).tobytes())
            f.write(np.ascontiguousarray(m.indices, np.uint32).tobytes())
        mesh, mat, model = scene.packed_objects()
        for i in range(len(scene.objects)):
            f.write(np.asarray([mesh[i], mat[i]], np.uint32).tobytes() + np.ascontiguousarray(model[i], np.float32).tobytes())
        f.write(scene.packed_materials().tobytes())
        f.write(np.ascontiguousarray(cam.ubo(), np.float32).tobytes() + np.asarray(CLEAR, np.float32).tobytes())
    exe = _build_c_host(tmp_path)
    devices = ["0", "1"] if _n_gpus() >= 2 else ["0", "0"]
    for devs in (["0"], devices, ["0", "0", "0"]):
        out = tmp_path / "out.rgba32f"
        r = subprocess.run([exe, str(blob), str(out)] + devs, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        got = np.fromfile(out, np.float32).reshape(H, W, 4)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), devs
        assert f"{max(len(devs), 1)} devices" in r.stdout
