#!/usr/bin/env python
"""Materialising join at the C3 shape (|R|=2^27 PK, |S|=2^30 FK, device resident): time and bandwidth.

  python tools/mat_bench.py [--log-r 27 --log-s 30 --reps 3]
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
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
n, m = 1 << a.log_r, 1 << a.log_s
with hj.HashJoin("radix") as h:
    dR, dS = h.device_alloc(n * 8), h.device_alloc(m * 8)
    h.generate(1, 0, n, n, 12345, dR)
    h.generate(1, 0, m, n, 54321, dS)
    d_out = h.device_alloc(m * 16)
    rows, cnt = [], []
    for i in range(a.reps + 1):
        r, need = h.join_materialize((dR, n), (dS, m), d_out, m)
        assert need == m == r.matches
        c, _ = h.join_materialize((dR, n), (dS, m))
        p = h.join((dR, n), (dS, m))
        assert p.checksum == r.checksum
        if i:
            rows.append(r.as_dict()); cnt.append(c.as_dict())
    med = lambda rs, k: round(float(np.median([x[k] for x in rs])), 1)
    out = {"R": n, "S": m, "pairs": m, "us_total_materialize": med(rows, "us_total"), "us_join_count+scan+write": med(rows, "us_join"),
           "us_join_count_only": med(cnt, "us_join"), "us_scatter": med(rows, "us_scatter1") + med(rows, "us_scatter2"),
           "plain_join_us_total": round(float(p.us_total), 1), "plain_join_us_join": round(float(p.us_join), 1)}
    out["Mtuples/s_materialize"] = round((n + m) / out["us_total_materialize"], 1)
    wr = out["us_join_count+scan+write"] - out["us_join_count_only"]
    out["write_pass_us"] = round(wr, 1)
    out["write_pass_GB/s"] = round((8.0 * (n + m) + 16.0 * m) / wr / 1e3, 1)
    print(json.dumps(out))
