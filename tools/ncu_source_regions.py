#!/usr/bin/env python
"""Per-instruction view of one kernel from `ncu --page source --csv --print-source sass`: prints every SASS
instruction of the first kernel with executions, average active lanes and samples, so that regions (node step,
triangle test, refill ...) can be read off.  Usage: ncu_source_regions.py src.csv [kernel_index]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = int(sys.argv[2]) if len(sys.argv) > 2 else 0
k = -1
out = []
hdr = None
for r in rows:
    if r and r[0] == "Kernel Name":
        k += 1
        continue
    if r and r[0] == "Address":
        hdr = r
        continue
    if k == want and hdr and len(r) >= len(hdr) - 2:
        out.append(r)
ia, isrc, iex, ith, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Avg. Threads Executed"), hdr.index("# Samples")
base = int(out[0][ia], 16)
tot = sum(int(r[iex]) for r in out)
tots = sum(int(r[ismp]) for r in out)
print(f"total warp instructions {tot}, samples {tots}")
for r in out:
    print(f"{int(r[ia],16)-base:5x} {int(r[iex]):>11} {float(r[ith]):5.1f} {int(r[ismp]):>6}  {r[isrc].strip()}")
