#!/usr/bin/env python
"""Radix join under a zipf-skewed probe side (mc --skew): R = PK 1..n shuffled, S = zipf(z) FK.

  python tools/skew_bench.py [--log-r 24 --log-s 27 --z 0,0.5,0.75,1.0,1.25]
The zipf keys are drawn on the host (numpy inverse-cdf, like genzipf.c but not bit-identical).
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import htm_hashjoin_b200 as hj  # noqa: E402


def zipf_fk(m, n, z, seed):
    rng = np.random.default_rng(seed)
    if z == 0:
        return rng.integers(1, n + 1, m, dtype=np.uint64)
    alphabet = rng.permutation(n).astype(np.uint64) + np.uint64(1)
    cdf = np.cumsum(1.0 / np.arange(1, n + 1, dtype=np.float64) ** z)
    cdf /= cdf[-1]
    out = np.empty(m, dtype=np.uint64)
    step = 1 << 24
    for i in range(0, m, step):
        k = min(step, m - i)
        out[i:i + k] = alphabet[np.minimum(np.searchsorted(cdf, rng.random(k)), n - 1)]
    return out


ap = argparse.ArgumentParser()
ap.add_argument("--log-r", type=int, default=24)
ap.add_argument("--log-s", type=int, default=27)
ap.add_argument("--z", default="0,0.5,0.75,1.0,1.25")
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
n, m = 1 << a.log_r, 1 << a.log_s
with hj.HashJoin("radix") as h:
    dR, dS = h.device_alloc(n * 8), h.device_alloc(m * 8)
    h.generate(1, 0, n, n, 12345, dR)
    for z in [float(x) for x in a.z.split(",")]:
        S = zipf_fk(m, n, z, 3)
        top = int(np.bincount((S & np.uint64(0xFFFF)).astype(np.int64)).max())
        h.copy(dS, S.ctypes.data, m * 8)
        rows = []
        for i in range(a.reps + 2):
            r = h.join((dR, n), (dS, m))
            assert r.matches == m
            if i >= 2:
                rows.append(r.as_dict())
        o = {k: round(float(np.median([x[k] for x in rows])), 1) for k in rows[0] if k.startswith("us_") and rows[0][k]}
        o.update({"z": z, "Mtuples/s": round((n + m) / o["us_total"], 1), "bits": rows[-1]["radix_bits"],
                  "largest_16bit_bucket_share": round(top / m, 4)})
        print(json.dumps(o), flush=True)
