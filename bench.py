#!/usr/bin/env python
"""bench.py — Mrays/s and ms/frame of the path-tracing hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (config.workload): BASELINE.json configs[2] — Sponza 1920x1080, depth 8 (roulette from depth 3),
one step = `--spp-per-step` accumulated frames (default 256 = the config's whole 256-spp image, traced as
batches of as many frames in flight as the path-state arrays hold in 56 GiB: two batches of 128 at N = 1, all 256 at N >= 2).  The Sponza geometry blob is absent from the reference mount, so the scene is the
deterministic "sponza-proxy" (raytracer_b200/proxy_scenes.py: same object/triangle/material counts).
N > 1: screen tiles (8-row strips, rank r owns strip s when s % N == r), scene replicated, no
collective on the data path; the strips are gathered over NCCL only for the end-to-end number.

  value        Mrays/s (extend + visibility rays actually traced, from device counters of an identical untimed
               pass) over K steps timed with CUDA events on the library's stream, inputs resident in HBM
  e2e          the same through the C ABI with HOST buffers: camera UBO from host memory each step and the
               RGBA32F image read back to (pinned) host memory each step
  roofline     dominant kernel (wf_extend_coop_kernel: per-ray traversal of bounces >= 1): algorithmic bytes / its
               CUDA-event time; roofline_packet: the same for the first-bounce packet kernel
  cpu_baseline the CPU oracle (a port: the reference has no CPU path) on a bounded sample, rank 0, N = 1
  --impl reference   times that CPU port with all host threads (the reference's Vulkan path cannot run here)
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CLEAR = (0.53, 0.81, 0.92)
METRIC = "Mrays/s @1080p Sponza(proxy), depth 8"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--spp-per-step", type=int, default=256)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--max-depth", type=int, default=8)
    ap.add_argument("--scene", default="sponza-proxy", choices=["sponza-proxy", "cornell_pbr", "soup1m"])
    ap.add_argument("--pipeline", default="wavefront", choices=["wavefront", "megakernel"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-rows", type=int, default=0, help="rows of one frame the CPU sample renders (0 = auto)")
    return ap.parse_args()


def make_scene(name, W, H):
    import raytracer_b200 as rb
    from raytracer_b200 import proxy_scenes

    if name == "sponza-proxy":
        return proxy_scenes.sponza_proxy(), rb.Camera(aspect=W / H, **proxy_scenes.SPONZA_CAMERA)
    if name == "soup1m":
        return proxy_scenes.soup(1_000_000), rb.Camera(aspect=W / H, **proxy_scenes.SOUP_CAMERA)
    from raytracer_b200.scene_io import load_scene_npz

    scene = load_scene_npz(os.path.join(ROOT, "tests", "golden", "scenes", "cornell_pbr.npz"))
    return scene, rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H, target=(0.0, 1.0, 0.0))


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def cpu_sample(scene, cam, W, H, max_depth, rows, frames=(1,)):
    """The CPU oracle on `rows` image rows (centred band) of the same workload. Returns (Mrays/s, s, rays, threads)."""
    from oracle.binding import OracleScene, lib

    orc = OracleScene(scene).build_bvh()
    y0 = max(0, (H - rows) // 2)
    y1 = min(H, y0 + rows)
    img = np.zeros((H, W, 4), np.float32)
    cnt = np.zeros(6, np.uint64)
    t0 = time.perf_counter()
    for f in frames:
        orc.render(cam.ubo(), CLEAR, f, W, H, image=img, y0=y0, y1=y1, max_depth=max_depth, counters=cnt)
    dt = time.perf_counter() - t0
    rays = int(cnt[0] + cnt[1])
    return rays / dt / 1e6, dt, rays, lib().orc_max_threads(), (y0, y1)


def run_reference(args):
    """--impl reference: the reference's Vulkan/GLSL path cannot run on this box (no graphics stack) and the
    reference has no CPU implementation; what is timed is the CPU port (oracle/), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    W, H = args.width, args.height
    scene, cam = make_scene(args.scene, W, H)
    rows = args.cpu_rows or H  # one whole frame per step (~0.5 s on 16 cores): K = 16 steps stay well within minutes
    for _ in range(min(args.warmup, 1)):
        cpu_sample(scene, cam, W, H, args.max_depth, min(rows, 32))
    tot_rays, tot_s = 0, 0.0
    threads = 1
    for k in range(args.steps):
        _, dt, rays, threads, band = cpu_sample(scene, cam, W, H, args.max_depth, rows, frames=(1 + k,))
        tot_rays += rays
        tot_s += dt
    v = tot_rays / tot_s / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / max(args.steps, 1), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.scene} {W}x{H} depth {args.max_depth} rr 3, {args.spp_per_step} spp per step",
                       "sample": f"rows {band[0]}..{band[1]} of one frame per step"},
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port",
                             "sample": f"{args.steps} steps x {rows} rows x {W} px, frames 1..{args.steps}"},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference GLSL/Vulkan path unavailable on this box; CPU port of the same shader math (oracle/)"}
    print(json.dumps(line), flush=True)


class _DevArray:
    """__cuda_array_interface__ view of a device pointer owned by librt (the strip image)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 3,
                                         "strides": None}


def run_ours(args):
    import torch

    import raytracer_b200 as rb  # noqa: F401
    from raytracer_b200 import librt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist

        # NCCL announces its version on stdout when the first communicator comes up; keep stdout for the JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_scalar(x, op):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    W, H, B, K, Wm = args.width, args.height, args.spp_per_step, args.steps, args.warmup
    scene, cam = make_scene(args.scene, W, H)
    pipe = librt.RT_PIPELINE_WAVEFRONT if args.pipeline == "wavefront" else librt.RT_PIPELINE_MEGAKERNEL
    kw = dict(device=local_rank, max_depth=args.max_depth, rr_start_depth=3, pipeline=pipe, strip_rows=8, strip_index=rank,
              strip_count=world, frames_in_flight=max(B, 1))
    ctx = librt.Context(counters=False, **kw)
    info = ctx.build_scene(scene)
    ctx.resize(W, H)
    ubo = cam.ubo()
    f0 = Wm * B

    # ---- ray counts of the timed frames: identical untimed passes on counting contexts -------------------
    # (a) per-ray primary traversal, full depth: the ray counts of `value` and the per-ray work definition of
    #     SURVEY §8d; (b) the same at depth 1: (a) - (b) = nodes / triangles of the bounces >= 1, which is what
    #     wf_extend_kernel traces; (c) depth 1 with packets: what the packet kernel really fetches.
    def count(per_ray, depth):
        k2 = dict(kw, max_depth=depth)
        c = librt.Context(counters=True, per_ray_primary=per_ray, **k2)
        c.build_scene(scene)
        c.resize(W, H)
        c.trace_frames(ubo, CLEAR, f0, K * B)
        out = c.counters()
        c.close()
        return out

    cnt = count(True, args.max_depth)
    cnt1 = count(True, 1)
    cntp = count(False, 1)
    rays_local = cnt["extend_rays"] + cnt["shadow_rays"]
    rays_total = reduce_scalar(float(rays_local), "SUM")

    # ---- device-resident throughput ---------------------------------------------------------------------
    for s in range(Wm):
        ctx.trace_frames(ubo, CLEAR, s * B, B)
    ctx.synchronize()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ctx.reset_counters()
    torch.cuda.profiler.start()  # `ncu --profile-from-start off` then lists exactly the launches of the timed region
    ctx.event_record(0)
    for s in range(K):
        ctx.trace_frames(ubo, CLEAR, f0 + s * B, B)
    ctx.event_record(1)
    ms = ctx.event_elapsed_ms()
    ctx.synchronize()
    torch.cuda.profiler.stop()
    barrier()
    sampler.stop_flag = True
    launches = ctx.counters()["launches"]
    ms_max = reduce_scalar(ms, "MAX")
    value = rays_total / (ms_max * 1e-3) / 1e6

    # ---- per-kernel times (CUDA events around every launch, separate pass over the same frames) -----------
    roof = None
    roof_pkt = None
    kprof = None
    if pipe == librt.RT_PIPELINE_WAVEFRONT:
        ctx.set_profiling(True)
        for s in range(K):
            ctx.trace_frames(ubo, CLEAR, f0 + s * B, B)
        kprof = ctx.kernel_profile()
        ctx.set_profiling(False)
        ext, pkt = kprof["extend"], kprof["extend_packet"]
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        total_k = sum(v["ms"] for v in kprof.values())

        def alg(c):  # BVH8 nodes 128 B + triangles 48 B fetched by closest-hit rays, ray in 32 B, hit out 16 B
            return (c["nodes"] - c["shadow_nodes"]) * 128 + (c["tris"] - c["shadow_tris"]) * 48 + c["extend_rays"] * 48

        sec_rays = cnt["extend_rays"] - cnt1["extend_rays"]
        alg_sec = alg(cnt) - alg(cnt1)
        achieved = alg_sec / (ext["ms"] * 1e-3) / 1e9 if ext["ms"] > 0 else 0.0
        traffic, l1_stage = None, None
        try:  # dram__bytes_read+write per ray of the last committed `ncu --set full` capture of this kernel
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["wf_extend_kernel"]
            traffic = tj["dram_bytes_per_ray"] * sec_rays / max(ext["launches"], 1)
            l1_stage = tj.get("l1_data_stage")
        except Exception:
            pass
        roof = {"bound": "hbm", "kernel": "wf_extend_coop_kernel (one ray per lane, pooled triangle tests; bounces 1..%d)" % (args.max_depth - 1),
                "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": "measured" if "hbm_gbs" in peaks else "fallback",
                "launches": ext["launches"], "avg_launch_ms": ext["ms"] / max(ext["launches"], 1),
                "alg_bytes_per_launch": alg_sec / max(ext["launches"], 1),
                "bytes_per_ray": alg_sec / max(sec_rays, 1), "rays_per_launch": sec_rays / max(ext["launches"], 1),
                "share_of_step": ext["ms"] / total_k if total_k > 0 else None,
                "note": "the BVH (< 20 MB) is L2/L1-resident, so algorithmic bytes / time is a bookkeeping ratio against the "
                        "HBM copy peak, not a DRAM utilisation; `traffic` is the DRAM figure",
                "l1_data_stage": l1_stage,  # what bounds the kernel in practice (ncu, committed capture): L1 wavefronts, not DRAM
                "kernel_ms": {k: round(v["ms"], 3) for k, v in kprof.items()}}
        if pkt["launches"]:
            a_ray, a_pkt = alg(cnt1), alg(cntp)
            roof_pkt = {"kernel": "wf_extend_packet_kernel (first bounce, 128 rays per warp)", "launches": pkt["launches"],
                        "avg_launch_ms": pkt["ms"] / pkt["launches"], "share_of_step": pkt["ms"] / total_k if total_k > 0 else None,
                        "per_ray_alg_bytes_per_launch": a_ray / pkt["launches"],
                        "achieved_per_ray_definition": a_ray / (pkt["ms"] * 1e-3) / 1e9,
                        "fetched_bytes_per_launch": a_pkt / pkt["launches"],
                        "achieved_fetched": a_pkt / (pkt["ms"] * 1e-3) / 1e9, "unit": "GB/s", "peak": peak,
                        "rays_per_launch": cnt1["extend_rays"] / pkt["launches"]}
        else:
            roof_pkt = None

    # ---- end to end through the C ABI with host buffers ------------------------------------------------
    host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True).numpy()
    gather = None
    if dist is not None:
        from raytracer_b200 import multi_gpu

        ptr, rows = ctx.device_image()
        view = torch.as_tensor(_DevArray(ptr, (rows, W, 4)), device="cuda") if rows else torch.empty((0, W, 4), device="cuda")
        sg = multi_gpu.StripGather(dist, H, W, 4, 8, device="cuda")
        host_t = torch.from_numpy(host)

        def gather():
            ctx.synchronize()
            parts = sg.gather(view)
            if rank == 0:
                host_t.copy_(sg.assemble_device(parts), non_blocking=True)  # one D2H of the assembled frame
                torch.cuda.synchronize()

    def e2e_step(s):
        ctx.trace_frames(np.array(ubo, copy=True), CLEAR, f0 + s * B, B)  # camera UBO + clear colour from host memory
        if gather is None:
            ctx.readback(host)
        else:
            gather()

    for s in range(min(Wm, 3)):
        e2e_step(s)
    barrier()
    t0 = time.perf_counter()
    for s in range(K):
        e2e_step(s)
    ctx.synchronize()
    barrier()
    e2e_s = reduce_scalar(time.perf_counter() - t0, "MAX")
    e2e_value = rays_total / e2e_s / 1e6
    sampler.join(timeout=1.0)

    # ---- CPU baseline (rank 0, N = 1 only) --------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # bounded sample: whole frames 1..n of the same workload, n sized for ~15 s of CPU work
        _, dt, _, _, _ = cpu_sample(scene, cam, W, H, args.max_depth, args.cpu_rows or H)
        n = int(max(1, min(64, round(15.0 / max(dt, 1e-3)))))
        v, dt, rays, threads, band = cpu_sample(scene, cam, W, H, args.max_depth, args.cpu_rows or H, frames=tuple(range(1, n + 1)))
        cpu = {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port",
               "sample": f"frames 1..{n}, rows {band[0]}..{band[1]} x {W} px, {rays} rays in {dt:.2f} s "
                         f"(oracle/ CPU restatement, OpenMP; the reference has no CPU path)"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": Wm,
                "ms_per_step": ms_max / K, "ms_per_frame": ms_max / (K * B), "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{args.scene} {W}x{H} depth {args.max_depth} rr 3, {B} spp per step",
                           "scene": {"triangles": info.n_triangles, "objects": info.n_objects, "lights": info.n_lights,
                                     "bvh8_nodes": info.n_nodes, "bvh_depth": info.bvh_depth, "build_ms": round(info.build_ms, 3)},
                           "pipeline": args.pipeline, "partition": f"8-row screen strips over {world} GPU(s), scene replicated",
                           "l2": "per-step working set (path state + queues + image, > 300 MB at 1080p) exceeds the 126 MB L2; "
                                 "the BVH (< 20 MB) stays L2-resident frame to frame, as in the reference's use",
                           "rays_per_step": rays_total / K,
                           "e2e_includes": "camera UBO from host + trace + RGBA32F image to pinned host memory every step "
                                           "(NCCL gather of strips to rank 0 first when N > 1); scene build is outside, "
                                           "as gpu_scene_init is in the reference"},
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 64 * 4 + 12,
                        "d2h_bytes_per_step": W * H * 16, "ms_per_step": 1e3 * e2e_s / K},
                "gpu_launches": int(launches), "clocks": sampler.summary()}
        if roof:
            line["roofline"] = roof
            if roof_pkt:
                line["roofline_packet"] = roof_pkt
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
