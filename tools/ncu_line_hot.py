#!/usr/bin/env python
"""Executed warp instructions and stall samples of one kernel per SOURCE LINE: joins the SASS page of an ncu report with the
line table of the object file the report was taken from (same build).

    ncu -i prof.ncu-rep --page source --csv > sass.csv
    python tools/ncu_line_hot.py sass.csv raytracer_b200/csrc/rt_frame.o wf_extend_packet_kernelILb0 [top]
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile


def line_table(obj, func):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, check=True, capture_output=True)
    cubin = os.path.join(d, sorted(os.listdir(d))[0])
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
    table, cur, inside = {}, ("?", 0), False
    for l in txt:
        if l.startswith(".text.") and l.endswith(":"):
            inside = func in l
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "(.*)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            table[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return table


def main(sass_csv, obj, func, top=40):
    rows = list(csv.reader(open(sass_csv)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[2:] if len(r) > ix["Instructions Executed"] and r[0].startswith("0x")]
    base = int(data[0][0], 16)
    table = line_table(obj, func)
    agg = collections.defaultdict(lambda: [0, 0, 0])
    tot_i = tot_s = 0
    for r in data:
        off = int(r[0], 16) - base
        n, s = int(r[ix["Instructions Executed"]]), int(r[ix["# Samples"]])
        key = table.get(off, (("?", 0), ""))[0]
        a = agg[key]
        a[0] += n
        a[1] += s
        a[2] += 1
        tot_i += n
        tot_s += s
    print(f"total warp-instructions {tot_i}, samples {tot_s}, {len(data)} SASS instructions, {len(table)} in the line table")
    src = {}
    for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[: int(top)]:
        text = ""
        for cand in (os.path.join(os.path.dirname(obj), f), ):
            if os.path.exists(cand):
                src.setdefault(cand, open(cand).read().splitlines())
                if 0 < ln <= len(src[cand]):
                    text = src[cand][ln - 1].strip()[:90]
        print(f"{100.0 * a[0] / tot_i:5.2f}% inst {100.0 * a[1] / max(tot_s, 1):5.2f}% samples {a[2]:4d} sass  {f}:{ln}  {text}")


if __name__ == "__main__":
    main(*sys.argv[1:])
