#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (time, count, share)."""
import collections
import csv
import sys


def main(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        try:
            v = float(row["Metric Value"].replace(",", ""))
        except (KeyError, ValueError):
            continue
        k = row["Kernel Name"].split("(")[0].replace("void ", "")
        a = agg.setdefault(k, [0, 0.0, row["Grid Size"], row["Block Size"]])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{'kernel':40s} {'launches':>8s} {'total_us':>12s} {'avg_us':>10s} {'share':>7s}  grid block")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:40s} {v[0]:8d} {v[1] / 1e3:12.1f} {v[1] / 1e3 / v[0]:10.2f} {v[1] / tot:7.3f}  {v[2]} {v[3]}")
    print(f"{'TOTAL':40s} {sum(v[0] for v in agg.values()):8d} {tot / 1e3:12.1f}")


if __name__ == "__main__":
    main(sys.argv[1])
