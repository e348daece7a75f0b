#!/usr/bin/env python
"""Fit the weights of the strip-cost model (RT_BALANCE_W_* in rt_api.cu, multi_gpu.MODEL_WEIGHTS) on ONE GPU.

    python tools/fit_balance.py [--world 8] [--steps 8] [--out gpurun_out/fit_balance.json]

For every rank of an N-way strip partition — under the round-robin deal AND under the deal the current weights produce —
this times the rank's share of the headline step on this GPU (the per-frame path has no collective, so that is what the
rank's GPU does in `bench.py --gpus N`) and sums the calibration counters (rt_measure_row_costs) over the rank's rows.
Non-negative least squares over those 2 N equations gives device picoseconds per pixel / per unit of primary traversal
work / per unit of continuation traversal work / per hit; measured and modelled times are printed side by side.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--spp", type=int, default=256)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import bench
    from raytracer_b200 import librt
    from scipy.optimize import nnls

    W, H, sr = 1920, 1080, 8
    scene, cam = bench.make_scene("sponza-proxy", W, H)
    ubo = cam.ubo()
    rows_feat = None
    eqs, times, labels, tables = [], [], [], {}
    for balance, tag in ((librt.RT_BALANCE_ROUND_ROBIN, "round_robin"), (librt.RT_BALANCE_COST, "cost")):
        per_rank = []
        for r in range(a.world):
            ctx = librt.Context(device=0, counters=False, max_depth=8, rr_start_depth=3, strip_rows=sr, strip_index=r,
                                strip_count=a.world, frames_in_flight=a.spp, strip_balance=balance)
            ctx.build_scene(scene)
            ctx.resize(W, H)
            if rows_feat is None:
                rows_feat = ctx.measure_row_costs(ubo, 0).astype(np.float64)
            for k in range(a.warmup):
                ctx.trace_frames(ubo, bench.CLEAR, k * a.spp, a.spp)   # frame 0 first: the cost deal is made here
            ctx.synchronize()
            owners = ctx.strip_owners()
            ctx.event_record(0)
            for k in range(a.steps):
                ctx.trace_frames(ubo, bench.CLEAR, (a.warmup + k) * a.spp, a.spp)
            ctx.event_record(1)
            ctx.synchronize()
            ms = ctx.event_elapsed_ms() / a.steps
            ctx.close()
            mine = np.repeat(owners == r, sr)[:H]
            f = rows_feat[mine].sum(axis=0)
            eqs.append([float(mine.sum() * W), float(f[0]), float(f[1]), float(f[2])])
            times.append(float(ms / a.spp * 1e9))  # ps per frame
            labels.append((tag, r))
            per_rank.append(ms)
            tables[tag] = [int(x) for x in owners]
            print(f"{tag:11s} rank {r}: {ms:.3f} ms/step, rows {int(mine.sum())}, features {f.astype(np.int64).tolist()}", flush=True)
        per_rank = np.asarray(per_rank)
        print(f"{tag}: slowest / mean = {per_rank.max() / per_rank.mean():.4f}  (step = {per_rank.max():.3f} ms)", flush=True)
    A, b = np.asarray(eqs, np.float64), np.asarray(times, np.float64)
    w, _ = nnls(A, b)
    pred = A @ w
    print("fitted ps per {pixel, primary work, continuation work, hit}:", np.round(w, 2).tolist(),
          f"rms relative error {np.sqrt(np.mean(((pred - b) / b) ** 2)):.4f}")
    for (tag, r), p, t in zip(labels, pred, b):
        print(f"   {tag:11s} rank {r}: measured {t / 1e6:.4f} us/frame, model {p / 1e6:.4f}")
    if a.out:
        json.dump({"world": a.world, "weights_ps": dict(zip(("pixel", "primary_work", "continuation_work", "hit"), (float(x) for x in w))),
                   "equations": [dict(assign=t, rank=int(r), features=e, ps_per_frame=y) for (t, r), e, y in zip(labels, eqs, times)],
                   "owners": tables}, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
