"""CPU suite, part 4: the N > 1 host path on world_size-2 gloo.  Each rank renders its own screen strips
(with the CPU oracle standing in for the GPU, which this container lacks), the strips are gathered on rank 0
with the same StripGather the GPU path uses, and the assembled image must equal a one-rank frame bit for bit
(seeds and camera rays use the full launch size)."""
import os
import socket

import numpy as np
import pytest

from conftest import CLEAR, fixture_scene


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, W, H, strip, out_path):
    import torch
    import torch.distributed as dist

    from oracle.binding import OracleScene
    from raytracer_b200 import multi_gpu
    import raytracer_b200 as rb

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc = OracleScene(fixture_scene("cornell_pbr")).build_bvh()
        cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
        rows = multi_gpu.strip_rows_of(rank, world, H, strip)
        full = np.zeros((H, W, 4), np.float32)
        for f in range(2):  # two accumulated frames, each rank only touches its own rows
            for y0 in rows[::strip] if len(rows) else []:
                y1 = min(H, int(y0) + strip)
                orc.render(cam.ubo(), CLEAR, f, W, H, image=full, y0=int(y0), y1=y1)
        local = torch.from_numpy(full[rows].copy())
        g = multi_gpu.StripGather(dist, H, W, 4, strip)
        parts = g.gather(local)
        if rank == 0:
            a, b = g.assemble_host(parts), g.assemble_device(parts).numpy()
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
            np.save(out_path, a)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,H,strip", [(2, 40, 8), (2, 21, 4)])
def test_two_rank_strips_assemble_to_the_single_rank_frame(tmp_path, oracle_lib, world, H, strip):
    import torch.multiprocessing as mp

    import raytracer_b200 as rb

    W = 32
    out = str(tmp_path / "img.npy")
    mp.spawn(_worker, args=(world, _free_port(), W, H, strip, out), nprocs=world, join=True)
    got = np.load(out)
    orc = oracle_lib.OracleScene(fixture_scene("cornell_pbr")).build_bvh()
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
    ref = None
    for f in range(2):
        ref = orc.render(cam.ubo(), CLEAR, f, W, H, image=ref)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def _slice_worker(rank, world, port, W, H, n_frames, out_path):
    import torch
    import torch.distributed as dist

    from oracle.binding import OracleScene
    from raytracer_b200 import multi_gpu
    import raytracer_b200 as rb

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc = OracleScene(fixture_scene("cornell_pbr")).build_bvh()
        cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
        # this rank's slice: frames rank, rank + world, ...; each frame rendered on its own (the oracle's frame
        # number is seed and blend index at once), then folded into the rank's running mean
        mean = np.zeros((H, W, 4), np.float32)
        for k, f in enumerate(range(rank, n_frames, world)):
            img = orc.render(cam.ubo(), CLEAR, f, W, H, image=None) if f == 0 else _single_frame(orc, cam, f, W, H)
            mean = img if k == 0 else (mean + (img - mean) * np.float32(1.0 / (k + 1))).astype(np.float32)
        red = multi_gpu.SppSliceReduce(dist, H, W, 4)
        out = red.reduce(torch.from_numpy(mean), n_frames)
        if rank == 0:
            np.save(out_path, out.numpy())
    finally:
        dist.destroy_process_group()


def _single_frame(orc, cam, f, W, H):
    """The clamped single-sample image of frame f: x with mix(0, x, 1/(f+1)) == what the oracle stores, undone."""
    img = orc.render(cam.ubo(), CLEAR, f, W, H, image=np.zeros((H, W, 4), np.float32))
    out = img * np.float32(f + 1)
    out[..., 3] = 1.0
    return out


def test_two_rank_spp_slices_reduce_to_the_sequential_mean(tmp_path, oracle_lib):
    import torch.multiprocessing as mp

    import raytracer_b200 as rb

    W, H, n = 24, 16, 5
    out = str(tmp_path / "img.npy")
    mp.spawn(_slice_worker, args=(2, _free_port(), W, H, n, out), nprocs=2, join=True)
    got = np.load(out)
    orc = oracle_lib.OracleScene(fixture_scene("cornell_pbr")).build_bvh()
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
    ref = None
    for f in range(n):
        ref = orc.render(cam.ubo(), CLEAR, f, W, H, image=ref)
    assert np.allclose(got[..., :3], ref[..., :3], rtol=1e-5, atol=1e-6)
    assert np.all(got[..., 3] == 1.0)


def _shared_worker(rank, world, port, W, H, strip, owners, out_path):
    import torch.distributed as dist

    from oracle.binding import OracleScene
    from raytracer_b200 import multi_gpu
    import raytracer_b200 as rb

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # the consume step of bench.py at N > 1: one host frame shared by the ranks, each writes its own strips into it
        # (the GPU path pins the mapping and rt_readback writes the rows; here the oracle writes them), then a barrier
        shared = multi_gpu.SharedHostImage(dist, H, W, 4, register=False)
        orc = OracleScene(fixture_scene("cornell_pbr")).build_bvh()
        cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
        for f in range(2):
            for s, owner in enumerate(owners):   # an arbitrary (cost-balanced) strip -> rank table, not round robin
                if owner == rank:
                    orc.render(cam.ubo(), CLEAR, f, W, H, image=shared.array, y0=s * strip, y1=min(H, (s + 1) * strip))
        dist.barrier()
        if rank == 0:
            np.save(out_path, np.array(shared.array, copy=True))
        dist.barrier()
        shared.close()
    finally:
        dist.destroy_process_group()


def test_two_ranks_write_their_strips_into_one_shared_host_frame(tmp_path, oracle_lib):
    import torch.multiprocessing as mp

    import raytracer_b200 as rb

    W, H, strip = 32, 44, 8
    owners = [1, 1, 0, 1, 0, 0]  # 6 strips (the last one partial), dealt unevenly on purpose
    out = str(tmp_path / "img.npy")
    mp.spawn(_shared_worker, args=(2, _free_port(), W, H, strip, owners, out), nprocs=2, join=True)
    got = np.load(out)
    orc = oracle_lib.OracleScene(fixture_scene("cornell_pbr")).build_bvh()
    cam = rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))
    ref = None
    for f in range(2):
        ref = orc.render(cam.ubo(), CLEAR, f, W, H, image=ref)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
