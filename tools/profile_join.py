#!/usr/bin/env python
"""Small fixed workload for ncu captures: two identical joins (the first is warm-up).

  python tools/profile_join.py radix   # |R|=2^25 PK, |S|=2^28 FK, 2^11 partitions in 2 passes (C3's partition size)
  python tools/profile_join.py global [log_n]  # atomic table, |R|=|S|=2^log_n (24), uniform, sorted, local_shuffle 256
  python tools/profile_join.py mat     # hj_join_materialize on the radix workload (count / scan / write kernels)
"""
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import htm_hashjoin_b200 as hj  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "radix"
if what == "radix":
    n, m = 1 << 25, 1 << 28
    # 2^11 partitions: R partitions of 16384 tuples like the default plan at C3 (2^27 / 2^13)
    with hj.HashJoin("radix", num_partitions=1 << 11, passes=2) as h:
        dR, dS = h.device_alloc(n * 8), h.device_alloc(m * 8)
        h.generate(1, 0, n, n, 12345, dR)
        h.generate(1, 0, m, n, 54321, dS)
        for _ in range(2):
            r = h.join((dR, n), (dS, m))
            assert r.matches == m
            print({k: round(v, 1) for k, v in r.as_dict().items() if k.startswith("us_") and v})
elif what == "mat":
    n, m = 1 << 25, 1 << 28
    with hj.HashJoin("radix", num_partitions=1 << 11, passes=2) as h:
        dR, dS = h.device_alloc(n * 8), h.device_alloc(m * 8)
        h.generate(1, 0, n, n, 12345, dR)
        h.generate(1, 0, m, n, 54321, dS)
        d_out = h.device_alloc(m * 16)
        for _ in range(2):
            r, pairs = h.join_materialize((dR, n), (dS, m), d_out, m)
            assert r.matches == m == pairs
            print({k: round(v, 1) for k, v in r.as_dict().items() if k.startswith("us_") and v})
else:
    n = 1 << (int(sys.argv[2]) if len(sys.argv) > 2 else 24)
    S = hj.generate_data("sorted", n)
    for dist in ("uniform", "sorted", "local_shuffle"):
        R = hj.generate_data(dist, n, n, 256 if dist == "local_shuffle" else 16)
        with hj.HashJoin("atomic") as h:
            dR, dS = h.to_device(R), h.to_device(S)
            for _ in range(2):
                r = h.join(dR, dS)
                assert r.matches == n
                print(dist, {k: round(v, 1) for k, v in r.as_dict().items() if k.startswith("us_") and v})
