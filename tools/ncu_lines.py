#!/usr/bin/env python
"""Aggregate the warp-stall samples of an .ncu-rep by CUDA source line (needs -lineinfo and --import-source on).
usage: python tools/ncu_lines.py prof.ncu-rep [top_n]"""
import csv, subprocess, sys
from collections import defaultdict
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr = None, None
agg = {}
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) or r[2] != "-":
        continue
    try:
        ln, s, e = int(r[0]), int(r[hdr.index("# Samples")]), int(r[hdr.index("Instructions Executed")])
    except ValueError:
        continue
    st = {}
    for i, h in enumerate(hdr):
        if h.startswith("stall_") and "Not Issued" not in h and r[i] not in ("", "0"):
            st[h[6:]] = int(r[i])
    agg[(cur, ln)] = (s, e, r[1], st)
tot = sum(v[0] for v in agg.values())
print("total samples", tot)
for key, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    st = sorted(v[3].items(), key=lambda kv: -kv[1])[:3]
    print(f"{v[0]:6d} {100 * v[0] / max(tot, 1):5.1f}% inst {v[1]:9d} {key[0]}:{key[1]:<4d} {v[2].strip()[:64]:64s} {st}")
