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


# the cost model of RT_BALANCE_COST (rt_api.cu: RT_BALANCE_W_*): tenths of a picosecond per pixel, per unit of primary /
# continuation traversal work, per closest-hit invocation (tools/fit_balance.py)
MODEL_WEIGHTS = (267.0, 9.0, 12.0, 1572.0)


class StripBalancer:
    """Strip -> rank assignment from MEASURED step times (host logic; the library only installs the table:
    `rt_set_strip_owners`).  The cost model of RT_BALANCE_COST predicts a rank's time to ~1 %, but single strips are off by
    more, and at N = 8 the slowest rank sets the step.  Feedback closes the gap: after a step, the cost estimates of every
    rank's strips are scaled so that they add up to the time the rank really took, and the strips are dealt again —
    longest processing time first, every rank keeping the NUMBER of strips the round-robin deal gives it (so no buffer
    changes size), the partial last strip staying with its round-robin owner.  Deterministic: ranks that feed it the same
    times (one all_gather) derive the same table without further talk.  The image is bit-identical under any table."""

    def __init__(self, height: int, world: int, strip_rows: int = DEFAULT_STRIP_ROWS, owners=None, costs=None):
        self.world, self.strip_rows = int(world), int(strip_rows)
        self.n_strips = (height + strip_rows - 1) // strip_rows
        self.n_full = height // strip_rows
        rr = np.arange(self.n_strips) % self.world
        self.owners = np.asarray(owners if owners is not None else rr, np.int64).copy()
        self.capacity = [int((rr[: self.n_full] == r).sum()) for r in range(self.world)]
        self.cost = np.asarray(costs, np.float64).copy() if costs is not None else np.ones(self.n_strips, np.float64)
        self.best = (float("inf"), self.owners.copy())

    def deal(self) -> np.ndarray:
        """Capacity-constrained LPT over the current cost estimates (ties: lower strip, lower rank)."""
        order = sorted(range(self.n_full), key=lambda s: (-self.cost[s], s))
        load, cap = [0.0] * self.world, list(self.capacity)
        owners = np.zeros(self.n_strips, np.int64)
        for s in order:
            r = min((r for r in range(self.world) if cap[r]), key=lambda r: (load[r], r))
            owners[s] = r
            load[r] += self.cost[s]
            cap[r] -= 1
        if self.n_strips > self.n_full:
            owners[self.n_full] = self.n_full % self.world
        return owners

    def update(self, rank_ms) -> np.ndarray:
        """rank_ms[r] = the step time rank r measured under the current table.  Returns the next table to try: the estimates
        of each rank's strips are scaled to add up to its measured time, then the slowest rank trades ONE strip with the
        fastest, the second slowest with the second fastest, ... — the pair of strips whose estimated cost difference
        comes closest to half the time difference.  A trade moves a fraction of one strip's cost, so an error of the
        per-strip estimates costs a fraction of a per cent of a rank's time, not a re-deal of everything."""
        t = np.asarray(rank_ms, np.float64)
        ratio = float(t.max() / t.mean())
        if ratio < self.best[0]:
            self.best = (ratio, self.owners.copy())
        for r in range(self.world):
            mine = np.nonzero(self.owners[: self.n_full] == r)[0]
            est = self.cost[mine].sum()
            if len(mine) and est > 0 and t[r] > 0:
                self.cost[mine] *= t[r] / est
        order = np.argsort(-t, kind="stable")
        owners = self.owners.copy()
        for k in range(self.world // 2):
            slow, fast = int(order[k]), int(order[self.world - 1 - k])
            want = 0.5 * (t[slow] - t[fast])
            if want <= 0.001 * t.mean():
                break
            a = np.nonzero(owners[: self.n_full] == slow)[0]
            b = np.nonzero(owners[: self.n_full] == fast)[0]
            if not len(a) or not len(b):
                continue
            diff = self.cost[a][:, None] - self.cost[b][None, :]     # what the slow rank sheds by trading a[i] for b[j]
            i, j = np.unravel_index(np.argmin(np.abs(diff - want)), diff.shape)
            if 0.0 < diff[i, j] < 2.0 * want:
                owners[a[i]], owners[b[j]] = fast, slow
        self.owners = owners
        return self.owners

    def best_owners(self) -> np.ndarray:
        return self.best[1]
