#!/usr/bin/env python
"""profiles/ncu_traffic.json from an `ncu --set full` capture of the dominant kernel — by script, not by hand.

    ncu -i X.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_traffic.py raw.csv --kernel wf_extend_coop_kernel --rays N --source "..." [--out profiles/ncu_traffic.json]

Takes the LAST launch of the kernel in the capture (the capture command skips the tiny hand-back launch), reads
dram__bytes_read.sum + dram__bytes_write.sum, the L1 data-stage wavefronts and the active cycles, and divides the DRAM bytes
by the launch's ray count (`--rays`, counted by librt's device counters on the same frames).  bench.py multiplies the
per-ray figure by the rays per launch of the run it reports.
"""
import argparse
import csv
import json
import os


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("raw")
    ap.add_argument("--kernel", default="wf_extend_coop_kernel")
    ap.add_argument("--rays", type=float, required=True)
    ap.add_argument("--source", default="")
    ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json"))
    a = ap.parse_args()
    rows = list(csv.reader(open(a.raw)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[2:] if len(r) == len(hdr) and a.kernel in r[ix["Kernel Name"]]]
    if not data:
        raise SystemExit(f"no launch of {a.kernel} in {a.raw}")
    r = data[-1]

    def get(name, scale=True):
        v = num(r[ix[name]])
        u = units[ix[name]].lower()
        if scale and v is not None:
            v *= {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "byte": 1.0}.get(u, 1.0)
        return v

    rd, wr = get("dram__bytes_read.sum"), get("dram__bytes_write.sum")
    wname = next((h for h in hdr if h.endswith("l1tex__data_pipe_lsu_wavefronts.avg")), None)  # "SM_A.TriageCompute." prefix in ncu 2025
    wave = get(wname, False) if wname else None
    act = get("smsp__cycles_active.avg", False)
    out = {}
    if os.path.exists(a.out):
        try:
            out = json.load(open(a.out))
        except Exception:
            out = {}
    out[a.kernel] = {
        "dram_bytes_per_ray": (rd + wr) / a.rays,
        "dram_bytes_read": rd, "dram_bytes_write": wr, "rays_in_launch": a.rays,
        "gpu_time_us": get("gpu__time_duration.sum", False),
        "grid": r[ix["Grid Size"]], "registers": num(r[ix["launch__registers_per_thread"]]),
        "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active", False),
        "l1tex_throughput_pct": get("l1tex__throughput.avg.pct_of_peak_sustained_active", False),
        "l1_hit_rate_pct": get("l1tex__t_sector_hit_rate.pct", False), "l2_hit_rate_pct": get("lts__t_sector_hit_rate.pct", False),
        "l1_data_stage": ({"frac": wave / act, "wavefronts_per_sm": wave, "active_cycles_per_sm": act} if wave and act else None),
        "source": a.source,
    }
    json.dump(out, open(a.out, "w"), indent=1)
    print(json.dumps(out[a.kernel], indent=1))


if __name__ == "__main__":
    main()
