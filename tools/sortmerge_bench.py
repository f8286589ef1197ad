#!/usr/bin/env python
"""Sort-merge comparator vs the radix hash join on the same device-resident FK join: one JSON
line per size.  The reference's figure compares its hash builds with SortMerge the same way (figs/perf.png).

  python tools/sortmerge_bench.py [--sizes 24:27,27:30 --reps 3]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import htm_hashjoin_b200 as hj  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", default="24:27,27:30", help="log2|R|:log2|S| pairs")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--cpu-log-n", type=int, default=0,
                help="also time the reference's own SortMerge.cpp (oracle/_ref) on the host cores against the GPU "
                     "entry point on the same DataGen arrays of 2^n tuples (checker code: only this tool runs it)")
a = ap.parse_args()
if a.cpu_log_n:
    import time
    from tests import oracle_lib
    ref = oracle_lib.Reference(probe=True)
    n = 1 << a.cpu_log_n
    S = hj.generate_data("sorted", n)
    for dist in ("sorted", "shuffle"):
        R = hj.generate_data(dist, n)
        row = {"comparator": "SortMerge", "rSize": n, "dataDistr": dist, "host_cores": os.cpu_count()}
        if ref.available and ref.sort_merge(S[:64], S[:64]) is not None:
            j = ref.sort_merge(R, S)
            row["reference_cpu"] = {k: j[k] for k in ("hashBuildTimeInMicroseconds", "sortTimeInMicroseconds",
                                                      "mergeTimeInMicroseconds", "totalMatches")}
        hj.SortMerge(R, n, S, n, 64)  # warm-up: context, buffers
        t0 = time.perf_counter()
        g = hj.SortMerge(R, n, S, n, 64)
        wall = (time.perf_counter() - t0) * 1e6
        row["gpu"] = {"hashBuildTimeInMicroseconds": g["hashBuildTimeInMicroseconds"],
                      "deviceTimeInMicroseconds": g["deviceTimeInMicroseconds"],
                      "sortTimeInMicroseconds": g["sortTimeInMicroseconds"],
                      "mergeTimeInMicroseconds": g["mergeTimeInMicroseconds"], "h2dMicroseconds": g["h2dMicroseconds"],
                      "totalMatches": g["totalMatches"],
                      "call_wall_us_incl_create_reserve": round(wall, 1)}
        print(json.dumps(row), flush=True)
for pair in a.sizes.split(","):
    lr, ls = (int(x) for x in pair.split(":"))
    n, m = 1 << lr, 1 << ls
    out = {"R": n, "S": m}
    want = None
    for algo in ("radix", "sortmerge"):
        with hj.HashJoin(algo, key_mask=0xFFFFFFFF) as h:
            dR, dS = h.device_alloc(n * 8), h.device_alloc(m * 8)
            h.generate(1, 0, n, n, 12345, dR)
            h.generate(1, 0, m, n, 54321, dS)
            rows = []
            for i in range(a.reps + 1):
                r = h.join((dR, n), (dS, m))
                assert r.matches == m, (algo, r.matches)
                want = want if want is not None else r.checksum
                assert r.checksum == want, algo
                if i:
                    rows.append(r.as_dict())
            med = lambda k: round(float(np.median([x[k] for x in rows])), 1)
            out[algo] = {"us_total": med("us_total"), "us_partition": round(med("us_total") - med("us_join"), 1),
                         "us_join": med("us_join"), "radix_bits": int(r.radix_bits), "passes": int(r.passes),
                         "Mtuples/s": round((n + m) / med("us_total"), 1)}
            h.device_free(dR)
            h.device_free(dS)
    print(json.dumps(out), flush=True)
