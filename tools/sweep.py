#!/usr/bin/env python
"""Sweeps for DESIGN.md / profiles: device-resident timings of both paths.

  python tools/sweep.py global [--log-n 24]   dist x W x algo on the global table (BASELINE configs 1, 2, 4)
  python tools/sweep.py radix  [--log-r 24 --log-s 27]   fan-out x passes (BASELINE config 3)
  python tools/sweep.py tsize  [--log-n 24]   --transactionSize sweep (config 4)
One JSON line per point (median of reps, CUDA-event phase times from the library).
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import htm_hashjoin_b200 as hj  # noqa: E402

PEAK = 6552.6
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def med(h, dR, dS, reps=5, warm=2):
    rows = []
    for i in range(reps + warm):
        r = h.join(dR, dS)
        if i >= warm:
            rows.append(r.as_dict())
    out = {k: float(np.median([x[k] for x in rows])) for k in rows[0] if k.startswith("us_")}
    last = rows[-1]
    out.update({k: last[k] for k in ("matches", "conflicts", "path_taken", "radix_bits", "passes", "hbm_bytes",
                                     "contention", "cas_retries", "sampled", "max_partition", "kernel_launches")})
    return out


def emit(tag, n_tuples, o):
    us = o["us_total"]
    o2 = {"tag": tag, "Mtuples/s": round(n_tuples / us, 1), "roofline_frac": round(o["hbm_bytes"] / (us * 1e-6) / (PEAK * 1e9), 4)}
    if o.get("path_taken") == 1 and o.get("us_build", 0) + o.get("us_probe", 0) > 0:
        bp = o["us_build"] + o["us_probe"]  # the reference's timed region: build + probe, table zeroed before
        o2["us_build+probe"] = round(bp, 1)
        o2["roofline_frac_build+probe"] = round(o["hbm_bytes"] / (bp * 1e-6) / (PEAK * 1e9), 4)
    o2.update({k: (round(v, 1) if isinstance(v, float) else v) for k, v in o.items() if not (k.startswith("us_") and v == 0)})
    print(json.dumps(o2), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["global", "radix", "tsize"])
    ap.add_argument("--log-n", type=int, default=24)
    ap.add_argument("--log-r", type=int, default=24)
    ap.add_argument("--log-s", type=int, default=27)
    ap.add_argument("--algos", default="atomic,nocc,radix,adaptive")
    a = ap.parse_args()
    if a.what == "global":
        n = 1 << a.log_n
        S = hj.generate_data("sorted", n)
        points = [("sorted", 16), ("uniform", 16)] + [("local_shuffle", 1 << w) for w in (4, 8, 12, 16, 20, 24)] + [("shuffle", 16)]
        for dist, rng in points:
            R = hj.generate_data(dist, n, n, rng)
            for algo in a.algos.split(","):
                with hj.HashJoin(algo) as h:
                    dR, dS = h.to_device(R), h.to_device(S)
                    emit(f"{algo}/{dist}/W={rng}/n=2^{a.log_n}", 2 * n, med(h, dR, dS))
                    h.device_free(dR[0]); h.device_free(dS[0])
    elif a.what == "tsize":
        n = 1 << a.log_n
        S = hj.generate_data("sorted", n)
        for dist in ("sorted", "shuffle"):
            R = hj.generate_data(dist, n)
            for ts in (1, 2, 4, 8, 16):
                with hj.HashJoin("atomic", transaction_size=ts) as h:
                    dR, dS = h.to_device(R), h.to_device(S)
                    emit(f"atomic/{dist}/tsize={ts}/n=2^{a.log_n}", 2 * n, med(h, dR, dS))
                    h.device_free(dR[0]); h.device_free(dS[0])
    else:
        n, m = 1 << a.log_r, 1 << a.log_s
        with hj.HashJoin("radix") as g:
            dR, dS = g.device_alloc(n * 8), g.device_alloc(m * 8)
            g.generate(1, 0, n, n, 12345, dR)
            g.generate(1, 0, m, n, 54321, dS)
            for bits, passes in [(7, 1), (8, 1), (9, 1), (10, 1), (11, 1), (12, 1), (10, 2), (11, 2), (12, 2), (13, 2),
                                 (14, 2), (15, 2)]:
                with hj.HashJoin("radix", num_partitions=1 << bits, passes=passes) as h:
                    o = med(h, (dR, n), (dS, m), reps=3, warm=1)
                    assert o["matches"] == m
                    emit(f"radix/bits={bits}/passes={passes}/R=2^{a.log_r}/S=2^{a.log_s}", n + m, o)


if __name__ == "__main__":
    main()
