#!/usr/bin/env python
"""BASELINE.json configs 1-5 on one B200: Mrays/s, ms/frame, algorithmic bytes per ray, CPU-port throughput, parity.

    python tools/run_configs.py [--out gpurun_out/configs] [--quick]

Per config: the full-size run is timed with CUDA events on the library's stream (rt_trace_frames, up to 64 frames in
flight); parity is judged at a reduced size the CPU oracle finishes in seconds — identical RGBA32F bits, hence
"% of pixels within 1e-3" = 100 and SSIM of the sRGB-8 images = 1 whenever it holds.  sRGB PNGs of both renderers
land in --out.  Sponza / white-room geometry blobs are absent from the reference mount: config 3 runs on the
labelled sponza-proxy, config 5 on multi-dispersion.gltf only.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CLEAR = (0.53, 0.81, 0.92)


def camera_for(name, aspect):
    import raytracer_b200 as rb
    from raytracer_b200 import proxy_scenes

    if name == "cornell_pbr":
        return rb.Camera(position=(0.0, 1.0, 4.5), aspect=aspect, target=(0.0, 1.0, 0.0))
    if name.startswith("conference"):
        return rb.Camera(position=(-3.5, -2.0, 1.6), aspect=aspect, target=(1.0, 0.5, 1.0), up=(0.0, 0.0, 1.0))
    if name == "multi_dispersion":
        return rb.Camera(position=(0.0, 2.0, 8.0), aspect=aspect, target=(0.0, 0.5, 0.0))
    if name == "sponza-proxy":
        return rb.Camera(aspect=aspect, **proxy_scenes.SPONZA_CAMERA)
    return rb.Camera(aspect=aspect)  # application.odin:52 default


def load(name):
    from conftest import fixture_scene
    from raytracer_b200 import proxy_scenes

    return proxy_scenes.sponza_proxy() if name == "sponza-proxy" else fixture_scene(name)


def run_config(tag, name, W, H, spp, depth, out_dir, restir=False, small=(96, 54), cpu_frames=2):
    from oracle.binding import OracleScene
    from raytracer_b200 import librt
    from raytracer_b200.image_io import ssim, write_png

    scene = load(name)
    cam = camera_for(name, W / H)
    row = {"config": tag, "scene": name, "res": f"{W}x{H}", "spp": spp, "depth": depth}
    # ---- full size: throughput (timed context without counters; ray / node / triangle counts from an identical
    # pass on a counting context with per-ray primary traversal = the per-ray work definition) -----------------
    ubo = cam.ubo()
    cctx = librt.Context(device=0, max_depth=depth, rr_start_depth=3, counters=True, frames_in_flight=64, per_ray_primary=True)
    info = cctx.build_scene(scene)
    cctx.resize(W, H)
    row.update(triangles=info.n_triangles, lights=info.n_lights, build_ms=round(info.build_ms, 2))
    if restir:
        for f in range(spp):
            cctx.restir_frame(ubo, CLEAR, f)
    else:
        cctx.trace_frames(ubo, CLEAR, 0, spp)
    c = cctx.counters()
    cctx.close()
    # frames in flight: lit scenes, where every path continues and takes a light sample at every bounce, run best at 32
    # (config 4b: 2.11 ms/frame at 32, 2.20 at 64); light-less ones keep gaining up to 128 (config 3: 0.540 at 41, 0.525 at 128)
    lanes = 32 if len(scene.lights()) else 128
    ctx = librt.Context(device=0, max_depth=depth, rr_start_depth=3, counters=False, frames_in_flight=lanes)
    ctx.build_scene(scene)
    ctx.resize(W, H)
    if restir:
        for f in range(4):
            ctx.restir_frame(ubo, CLEAR, f)
        ctx.event_record(0)
        for f in range(spp):
            ctx.restir_frame(ubo, CLEAR, f)
        ctx.event_record(1)
    else:
        ctx.trace_frames(ubo, CLEAR, 0, min(spp, 64))
        ctx.event_record(0)
        ctx.trace_frames(ubo, CLEAR, 0, spp)
        ctx.event_record(1)
    ms = ctx.event_elapsed_ms()
    rays = c["extend_rays"] + c["shadow_rays"]
    alg = c["nodes"] * 128 + c["tris"] * 48 + c["hits"] * 160 + W * H * 32 * spp
    # bytes_per_ray: the §8(d) byte model (bookkeeping: the BVH is cache-resident, so no fraction of the HBM peak is derived
    # from it any more); loads_per_ray: (thread, load instruction) pairs of the BVH gather, the quantity that binds traversal
    row.update(mrays_s=round(rays / ms / 1e3, 1), ms_per_frame=round(ms / spp, 4), rays_per_frame=rays // spp,
               bytes_per_ray=round(alg / max(rays, 1), 1), loads_per_ray=round((4 * c["nodes"] + 2 * c["tris"]) / max(rays, 1), 1))
    # ---- the CPU port at THIS resolution: a centred band of rows of frame 1, sized for about a second, scaled to the frame
    if not restir:
        from oracle.binding import OracleScene as _O, lib as _lib
        _lib().orc_set_num_threads(max(1, len(os.sched_getaffinity(0))))
        orc_f = _O(scene).build_bvh()
        band = max(8, min(H, int(150000 // W)))
        y0 = (H - band) // 2
        cnt_f = np.zeros(6, np.uint64)
        img_f = np.zeros((H, W, 4), np.float32)
        t0 = time.perf_counter()
        orc_f.render(ubo, CLEAR, 1, W, H, image=img_f, y0=y0, y1=y0 + band, max_depth=depth, counters=cnt_f)
        dt = time.perf_counter() - t0
        cpu_ms = dt * 1e3 * H / band
        row.update(cpu_full=f"{(int(cnt_f[0]) + int(cnt_f[1])) / dt / 1e6:.2f} Mrays/s, ~{cpu_ms:.0f} ms/frame ({_lib().orc_max_threads()} threads, rows {y0}..{y0 + band} of {W}x{H})",
                   speedup=round(cpu_ms / (ms / spp), 0))
    write_png(os.path.join(out_dir, f"{tag}_gpu_full.png"), ctx.readback_srgb8())
    ctx.close()
    # ---- reduced size: parity against the CPU oracle + its throughput --------------------------------
    w, h = small
    cam_s = camera_for(name, w / h)
    ctx = librt.Context(device=0, max_depth=depth, rr_start_depth=3, counters=True)
    ctx.build_scene(scene)
    ctx.resize(w, h)
    orc = OracleScene(scene).build_bvh()
    n_par = 4
    if restir:
        g = ctx.gbuffer(cam_s.ubo()) if hasattr(ctx, "gbuffer") else None
        truth = orc.direct_reference(cam_s.ubo(), CLEAR, w, h, 6)
        for f in range(256):
            ctx.restir_frame(cam_s.ubo(), CLEAR, f)
        got = ctx.readback()
        lit = truth[..., :3].sum(axis=2) > 1e-3
        ratio = got[..., :3].sum(axis=2)[lit] / truth[..., :3].sum(axis=2)[lit]
        row.update(parity=f"no reference algorithm: median ratio to brute-force direct lighting {np.median(ratio):.4f} (256 frames); "
                          "image independent of the strip partition bit for bit (tests/test_gpu_multi.py)")
        ref8 = (np.clip(truth[..., :3], 0, 1) ** (1 / 2.2) * 255).astype(np.uint8)
        got8 = (np.clip(got[..., :3], 0, 1) ** (1 / 2.2) * 255).astype(np.uint8)
        row.update(ssim=round(ssim(ref8, got8), 4))
        t0 = time.perf_counter()
        orc.direct_reference(cam_s.ubo(), CLEAR, w, h, 2)
        row.update(cpu="n/a (brute-force ground truth only)")
    else:
        ref = None
        cnt = np.zeros(6, np.uint64)
        t0 = time.perf_counter()
        for f in range(n_par):
            ref = orc.render(cam_s.ubo(), CLEAR, f, w, h, image=ref, max_depth=depth, counters=cnt)
        dt = time.perf_counter() - t0
        ctx.trace_frames(cam_s.ubo(), CLEAR, 0, n_par)
        got = ctx.readback()
        same = np.array_equal(got.view(np.uint32), ref.view(np.uint32))
        rel = np.abs(got[..., :3] - ref[..., :3]) / np.maximum(np.abs(ref[..., :3]), 1e-6)
        within = float((rel.max(axis=2) <= 1e-3).mean() * 100)
        # sRGB-8 of the oracle image through the same OETF table the library uses: upload + read back
        ctx.upload_accum(ref) if hasattr(ctx, "upload_accum") else None
        ref8 = ctx.readback_srgb8() if hasattr(ctx, "upload_accum") else None
        ctx.upload_accum(got)
        got8 = ctx.readback_srgb8()
        s = ssim(ref8[..., :3], got8[..., :3]) if ref8 is not None else float("nan")
        write_png(os.path.join(out_dir, f"{tag}_gpu_small.png"), got8)
        if ref8 is not None:
            write_png(os.path.join(out_dir, f"{tag}_oracle_small.png"), ref8)
        from oracle.binding import lib
        row.update(parity=f"{'bit-identical' if same else 'DIFFERS'}; {within:.2f} % px within 1e-3 ({w}x{h}, {n_par} frames)",
                   ssim=round(s, 6), cpu=f"{(int(cnt[0]) + int(cnt[1])) / dt / 1e6:.2f} Mrays/s ({lib().orc_max_threads()} threads, {w}x{h})")
    ctx.close()
    return row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs"))
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    q = a.quick
    specs = [
        ("1_cornell_box", "cornell_box", 512, 512, 1, 3, dict(small=(128, 128))),
        ("1b_cornell_box_d4", "cornell_box", 512, 512, 16, 4, dict(small=(128, 128))),
        ("2_cornell_pbr", "cornell_pbr", 1024, 1024, 64, 3, dict(small=(128, 128))),
        ("3_sponza_proxy", "sponza-proxy", 1920, 1080, 32 if q else 256, 8, dict(small=(160, 90))),
        ("4_conference_restir", "conference", 1920, 1080, 16, 1, dict(restir=True, small=(96, 54))),
        ("4b_conference_mirror_chairs", "conference_mirror_chairs", 1920, 1080, 64, 3, dict(small=(160, 90))),
        ("5_multi_dispersion_4k", "multi_dispersion", 3840, 2160, 64 if q else 1024, 3, dict(small=(160, 90))),
    ]
    rows = []
    for tag, name, W, H, spp, depth, kw in specs:
        rows.append(run_config(tag, name, W, H, spp, depth, a.out, **kw))
        print(json.dumps(rows[-1]), flush=True)  # a row is on record as soon as it exists
        with open(os.path.join(a.out, "configs_partial.json"), "w") as f:
            json.dump(rows, f, indent=1)
    with open(os.path.join(a.out, "configs.json"), "w") as f:
        json.dump(rows, f, indent=1)
    cols = ["config", "res", "spp", "depth", "triangles", "lights", "mrays_s", "ms_per_frame", "loads_per_ray", "bytes_per_ray", "cpu_full", "speedup", "parity", "ssim"]
    lines = ["| " + " | ".join(cols) + " |", "|" + "---|" * len(cols)]
    for r in rows:
        lines.append("| " + " | ".join(str(r.get(c, "")) for c in cols) + " |")
    open(os.path.join(a.out, "configs.md"), "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
