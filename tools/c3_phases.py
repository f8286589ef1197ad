#!/usr/bin/env python
"""Phase times of the C3 radix join (|R|=2^27 PK, |S|=2^30 FK, device resident), median of reps.

  python tools/c3_phases.py [--log-r 27 --log-s 30 --bits 15 --passes 2 --reps 5]
Kernel variants are picked by environment variables read in the library (HJ_J32_VARIANT, HJ_NO_JOIN32, ...),
so one process = one variant; the shell loop around this script is the sweep.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import htm_hashjoin_b200 as hj  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--log-r", type=int, default=27)
ap.add_argument("--log-s", type=int, default=30)
ap.add_argument("--bits", type=int, default=0)
ap.add_argument("--passes", type=int, default=0)
ap.add_argument("--pass1-bits", type=int, default=0)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--tag", default="")
a = ap.parse_args()
n, m = 1 << a.log_r, 1 << a.log_s
kw = {}
if a.bits:
    kw["num_partitions"] = 1 << a.bits
if a.passes:
    kw["passes"] = a.passes
if a.pass1_bits:
    kw["pass1_bits"] = a.pass1_bits
with hj.HashJoin("radix", **kw) as h:
    dR, dS = h.device_alloc(n * 8), h.device_alloc(m * 8)
    h.generate(1, 0, n, n, 12345, dR)
    h.generate(1, 0, m, n, 54321, dS)
    rows = []
    for i in range(a.reps + 2):
        r = h.join((dR, n), (dS, m))
        assert r.matches == m, (r.matches, m)
        if i >= 2:
            rows.append(r.as_dict())
    out = {k: round(float(np.median([x[k] for x in rows])), 1) for k in rows[0] if k.startswith("us_") and rows[0][k]}
    out["Mtuples/s"] = round((n + m) / out["us_total"], 1)
    out["checksum"] = rows[-1]["checksum"]
    out["tag"] = a.tag or " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("HJ_"))
    out["bits"], out["passes"] = rows[-1]["radix_bits"], rows[-1]["passes"]
    print(json.dumps(out), flush=True)
