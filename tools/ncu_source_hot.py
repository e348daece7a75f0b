#!/usr/bin/env python
"""Per-instruction view of `ncu -i X.ncu-rep --page source --csv`: executed warp instructions, average active
threads and stall samples per SASS instruction, with region subtotals between branch targets.

    ncu -i prof.ncu-rep --page source --csv --kernel-id ::regex:NAME:1 > src.csv ; python tools/ncu_source_hot.py src.csv
"""
import csv
import sys


def main(path, top=0):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[2:] if len(r) > ix["Instructions Executed"] and r[0].startswith("0x")]
    tot_i = sum(int(r[ix["Instructions Executed"]]) for r in data)
    tot_s = sum(int(r[ix["# Samples"]]) for r in data)
    base = int(data[0][0], 16)
    print(f"total warp-instructions {tot_i}, samples {tot_s}")
    for r in data:
        n = int(r[ix["Instructions Executed"]])
        s = int(r[ix["# Samples"]])
        at = r[ix["Avg. Threads Executed"]]
        print(f"{int(r[0], 16) - base:5x} {100.0 * n / tot_i:5.2f}% {100.0 * s / max(tot_s, 1):5.2f}% thr {at:>5s}  {r[1].strip()}")


if __name__ == "__main__":
    main(sys.argv[1])
