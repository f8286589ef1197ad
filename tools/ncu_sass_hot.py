#!/usr/bin/env python
"""Top SASS instructions of a captured kernel by executed count (ncu --import-source on capture).
  python tools/ncu_sass_hot.py x.ncu-rep [min_share_pct]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = rows[1]
ia, isrc, iex, ism, iwf, iwfi = (hdr.index(x) for x in ("Address", "Source", "Instructions Executed", "# Samples", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal"))
data = [r for r in rows[2:] if len(r) > iex and r[iex].isdigit()]
tot = sum(int(r[iex]) for r in data)
tots = sum(int(r[ism]) for r in data)
print("total warp-instr", tot, "samples", tots)
# print contiguous listing with counts, collapsed: only lines with >= thr share
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
for n, r in enumerate(data):
    sh = 100.0 * int(r[iex]) / tot
    if sh >= thr:
        print(f"{n:5d} {sh:5.2f}% ex={int(r[iex]):>10d} smp={100.0*int(r[ism])/tots:5.2f}% wf={r[iwf]:>9s}/{r[iwfi]:>9s} {r[isrc].strip()}")
