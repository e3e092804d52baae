#!/usr/bin/env python
"""Top stalled SASS instructions of an ncu report (source page): python tools/ncu_hot.py rep.ncu-rep [n]"""
import csv, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = rows[1]
isrc, isamp = hdr.index("Source"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
out = []
for k, r in enumerate(rows[2:]):
    if len(r) < len(hdr) or r[0].startswith("Kernel"):
        continue
    try:
        s = int(r[isamp])
    except ValueError:
        continue
    top = sorted(((int(r[i] or 0), hdr[i]) for i in stall_cols), reverse=True)[:2]
    out.append((s, k, r[isrc], top))
tot = sum(o[0] for o in out)
print("total samples", tot)
for s, k, t, top in sorted(out, reverse=True)[:n]:
    print("%6d %5.1f%%  #%5d  %-60s %s" % (s, 100.0 * s / tot, k, t[:60], top))
