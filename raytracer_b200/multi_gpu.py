"""Screen-tile sharding across the GPUs of one box — host-side plumbing only.

The reference renders on one GPU.  Here every rank holds the whole scene and BVH and traces the image rows
y with (y // strip_rows) % world == rank; seeds and camera rays use the full launch size
(shaders/simple.rgen:71, :41-53), so the union of the ranks' rows is bit-identical to a one-GPU frame.
The accumulation stays on each GPU across frames; rows only move at readback: one `gather` of the
compact strip images to rank 0 (NCCL over NVLink on the GPU box, gloo in the CPU tests).
`torch.distributed` is plumbing here: there is no collective on the per-frame data path.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

DEFAULT_STRIP_ROWS = 8


def strip_rows_of(rank: int, world: int, height: int, strip_rows: int = DEFAULT_STRIP_ROWS) -> np.ndarray:
    """Global row indices owned by `rank`, in the order of its compact image (rt_config.strip_*)."""
    y = np.arange(height, dtype=np.int64)
    return y[(y // strip_rows) % world == rank]


def local_row_count(rank: int, world: int, height: int, strip_rows: int = DEFAULT_STRIP_ROWS) -> int:
    return int(len(strip_rows_of(rank, world, height, strip_rows)))


def assemble(parts, height: int, strip_rows: int = DEFAULT_STRIP_ROWS) -> np.ndarray:
    """Full (H, W, C) image from per-rank compact strip images (rank order)."""
    world = len(parts)
    first = np.asarray(parts[0])
    out = np.zeros((height,) + first.shape[1:], first.dtype)
    for r, p in enumerate(parts):
        rows = strip_rows_of(r, world, height, strip_rows)
        out[rows] = np.asarray(p)[: len(rows)]
    return out


class StripGather:
    """Gathers the ranks' compact strip images on rank 0 with one collective.

    `local` tensors may live on the GPU (nccl) or the CPU (gloo).  Ranks own at most one strip more than
    others, so every rank pads to the largest count and a single fixed-size gather moves everything."""

    def __init__(self, dist, height: int, width: int, channels: int = 4, strip_rows: int = DEFAULT_STRIP_ROWS, device="cpu",
                 dtype=None):
        import torch

        self.dist, self.torch = dist, torch
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.height, self.strip_rows = height, strip_rows
        self.rows = [local_row_count(r, self.world, height, strip_rows) for r in range(self.world)]
        self.max_rows = max(self.rows)
        dtype = dtype or torch.float32
        self.send = torch.zeros((self.max_rows, width, channels), dtype=dtype, device=device)
        self.recv = [torch.empty_like(self.send) for _ in range(self.world)] if self.rank == 0 else None

    def gather(self, local) -> Optional[list]:
        """local: (rows[rank], W, C) tensor.  Returns the list of per-rank tensors on rank 0, else None."""
        n = self.rows[self.rank]
        if n:
            self.send[:n].copy_(local[:n])
        self.dist.gather(self.send, self.recv, dst=0)
        return self.recv

    def assemble_device(self, parts):
        """Full (H, W, C) image on rank 0's device from the gathered parts (one scatter per rank, no host pass)."""
        torch = self.torch
        if getattr(self, "_full", None) is None:
            self._full = torch.empty((self.height,) + tuple(self.send.shape[1:]), dtype=self.send.dtype, device=self.send.device)
            self._row_idx = [torch.as_tensor(strip_rows_of(r, self.world, self.height, self.strip_rows), device=self.send.device)
                             for r in range(self.world)]
        for r, p in enumerate(parts):
            if self.rows[r]:
                self._full.index_copy_(0, self._row_idx[r], p[: self.rows[r]])
        return self._full

    def assemble_host(self, parts) -> np.ndarray:
        return assemble([p[: self.rows[r]].cpu().numpy() for r, p in enumerate(parts)], self.height, self.strip_rows)


def slice_frames(rank: int, world: int, n_frames: int) -> int:
    """How many of the frames 0 .. n_frames-1 rank `rank` traces when frame f goes to rank f % world."""
    return (n_frames - rank + world - 1) // world if n_frames > rank else 0


class SppSliceReduce:
    """spp slices (the alternative to screen strips for low-resolution / high-spp runs): every rank
    traces whole frames f with f % world == rank (`rt_trace_frames_sliced(first_frame=rank, frame_stride=world)`)
    into its own running mean; the estimate over all n frames is sum_r (n_r / n) * mean_r — ONE reduce(SUM) of the
    pre-scaled images to rank 0 (NCCL over NVLink on the GPU box, gloo in the CPU tests).  Same estimator as the
    sequential mix of simple.rgen:127-136, float sums in another order (not bit-identical; ~1e-6 relative)."""

    def __init__(self, dist, height: int, width: int, channels: int = 4, device="cpu"):
        import torch

        self.dist, self.torch = dist, torch
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.buf = torch.zeros((height, width, channels), dtype=torch.float32, device=device)

    def reduce(self, local_mean, n_frames: int):
        """local_mean: (H, W, C) mean over this rank's slice_frames(rank, world, n_frames) frames.  Returns the
        combined image on rank 0 (alpha forced to 1), None elsewhere."""
        torch = self.torch
        n_r = slice_frames(self.rank, self.world, n_frames)
        if n_r:
            torch.mul(local_mean, float(n_r) / float(n_frames), out=self.buf)
        else:
            self.buf.zero_()
        self.dist.reduce(self.buf, dst=0, op=self.dist.ReduceOp.SUM)
        if self.rank != 0:
            return None
        self.buf[..., 3] = 1.0
        return self.buf


class SharedHostImage:
    """One full-size host image shared by the ranks of a box: the consume step of N one-GPU processes.

    The accumulation image only ever has to reach HOST memory (the reference blits it; `rt_readback` hands it to the
    caller), so the ranks do not gather on a GPU first: rank 0 creates a /dev/shm file, every rank maps it and pins the
    mapping (`cudaHostRegister`), and `rt_readback(ctx, image)` on each rank writes that rank's strips straight into it
    — N PCIe links side by side, no collective, no staging.  One `barrier()` later rank 0 holds the whole frame.
    `dist` is only used for the file name and the barrier (NCCL on the GPU box, gloo in the CPU tests, where
    `register=False` leaves the mapping unpinned and the ranks write with numpy)."""

    def __init__(self, dist, height: int, width: int, channels: int = 4, dtype=np.float32, register: bool = True):
        import mmap
        import os
        import tempfile

        self.dist = dist
        self.rank = dist.get_rank() if dist is not None else 0
        nbytes = int(height) * int(width) * int(channels) * np.dtype(dtype).itemsize
        name = [None]
        if self.rank == 0:
            fd, path = tempfile.mkstemp(prefix="rt_frame_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
            os.ftruncate(fd, nbytes)
            os.close(fd)
            name[0] = path
        if dist is not None and dist.get_world_size() > 1:
            dist.broadcast_object_list(name, src=0)
        self.path = name[0]
        self._f = open(self.path, "r+b")
        self._mm = mmap.mmap(self._f.fileno(), nbytes)
        self.array = np.frombuffer(self._mm, dtype=dtype).reshape(height, width, channels)
        self._registered = False
        if register:
            import torch

            rc = torch.cuda.cudart().cudaHostRegister(self.array.ctypes.data, nbytes, 3)  # portable | mapped
            if int(rc) != 0:
                raise RuntimeError(f"cudaHostRegister of the shared frame failed: {rc}")
            self._registered = True
        if dist is not None and dist.get_world_size() > 1:
            dist.barrier()
        if self.rank == 0:
            os.unlink(self.path)  # every rank holds its mapping; the name can go

    def close(self):
        if self._registered:
            import torch

            torch.cuda.cudart().cudaHostUnregister(self.array.ctypes.data)
            self._registered = False
        self.array = None
        try:
            self._mm.close()
        except BufferError:
            pass  # a numpy view is still alive somewhere; the mapping goes with it
        self._f.close()
