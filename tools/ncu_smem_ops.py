#!/usr/bin/env python
"""Shared-memory wavefronts per opcode and stall samples per opcode class for each captured launch.
  python tools/ncu_smem_ops.py x.ncu-rep [launch_index]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
# the csv holds one block per launch: a "Kernel Name" line, a header line, rows
blocks, cur = [], None
for row in csv.reader(io.StringIO(txt)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": []}; blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(row)
sel = int(sys.argv[2]) if len(sys.argv) > 2 else None
for bi, b in enumerate(blocks):
    if sel is not None and bi != sel: continue
    hdr = b["rows"][0]
    isrc, iex, iwf, iwfi, ism = (hdr.index(x) for x in ("Source", "Instructions Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal", "# Samples"))
    data = [r for r in b["rows"][1:] if len(r) > iex and r[iex].isdigit()]
    tot = sum(int(r[iex]) for r in data); tots = sum(int(r[ism]) for r in data)
    print(f"== launch {bi}: {b['name'][:60]}  warp-instr {tot}  samples {tots}")
    c = collections.defaultdict(lambda: [0, 0, 0, 0])
    for r in data:
        t = r[isrc].split()
        op = t[1] if t[0].startswith("@") else t[0]
        e = c[op]; e[0] += int(r[iex]); e[1] += int(r[iwf] or 0); e[2] += int(r[iwfi] or 0); e[3] += int(r[ism])
    for op, e in sorted(c.items(), key=lambda kv: -kv[1][3])[:22]:
        print(f"  {op:28s} instr {100*e[0]/tot:5.1f}%  samples {100*e[3]/tots:5.1f}%  wavefronts {e[1]:>10d} (ideal {e[2]:>10d}) per-instr {e[1]/max(e[0],1):.2f}")
