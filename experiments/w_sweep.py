#!/usr/bin/env python
"""The reference's locality sweep against the GPU library.

Reproduces the family of runs behind the reference's figs/perf.png -- experiments/probe.sh and
experiments/motivation.sh: |R| = |S| = 2^27, R = local_shuffle with --shuffleRange 2^0 .. 2^27,
S = sorted 1..|R| (main.cpp:93), probeLength 4, N repetitions (runner.sh: N=5) -- with the algorithms of
this library: nocc and atomic (global table), radix (the PRJ counterpart) and adaptive (the HTM-Adaptive
counterpart: sampled contention decides between the two).  One process, inputs generated once per W by
the bit-identical DataGen port and shared by all algorithms; times are CUDA-event times of device
resident joins (the reference times the in-memory join after generation too).

  python experiments/w_sweep.py [--log-n 27] [--w-step 1] [--reps 5] [--out experiments/results/w_sweep.jsonl]
  python experiments/w_sweep.py --sweep tsize      # SizeToAbortsAndTime{Sorted,Shuffled}.sh counterpart

One JSON line per (algo, W): the reference's keys first (algo, rSize, probeLength,
hashBuildTimeInMicroseconds, conflicts, totalMatches, inputSum, outputSum), then ours.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import htm_hashjoin_b200 as hj  # noqa: E402

PATHS = {0: "none", 1: "global", 2: "radix", 3: "sortmerge"}


def timed(h, dR, dS, reps, warm=2):
    rows = []
    for i in range(reps + warm):
        r = h.join(dR, dS)
        if i >= warm:
            rows.append(r)
    us = sorted(float(r.us_total) for r in rows)
    return rows[-1], us[len(us) // 2], us[0]


def line(algo, n, m, dist, w, ts, res, med_us, best_us, check=None):
    o = {"algo": algo, "rSize": n, "probeLength": 4, "hashBuildTimeInMicroseconds": round(med_us, 1),
         "conflicts": int(res.conflicts), "totalMatches": int(res.matches), "inputSum": int(res.input_sum),
         "outputSum": int(res.table_sum + res.overflow_sum) & (2**64 - 1),
         "sSize": m, "dataDistr": dist, "shuffleRange": w, "transactionSize": ts, "bestMicroseconds": round(best_us, 1),
         "checksum": int(res.checksum), "tuplesPerSec": round((n + m) / med_us * 1e6), "pathTaken": PATHS.get(res.path_taken, str(res.path_taken)),
         "firstRoundFailureFraction": round(float(res.contention), 5), "radixBits": int(res.radix_bits), "passes": int(res.passes),
         "buildMicroseconds": round(float(res.us_build), 1), "probeMicroseconds": round(float(res.us_probe), 1),
         "sampleMicroseconds": round(float(res.us_sample), 1)}
    if check is not None:
        o["verified"] = bool(check)
    return o


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", choices=["w", "tsize"], default="w")
    ap.add_argument("--log-n", type=int, default=27)
    ap.add_argument("--w-step", type=int, default=1, help="W = 2^0, 2^step, 2^(2 step) ... (the reference uses 1)")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--algos", default="nocc,atomic,radix,adaptive")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    n = 1 << a.log_n
    out = open(a.out, "w") if a.out else None

    def emit(o):
        s = json.dumps(o)
        print(s, flush=True)
        if out:
            out.write(s + "\n")
            out.flush()

    S = hj.generate_data("sorted", n)
    # unique dense keys joined with 1..n: every S tuple finds its R tuple
    want_matches, want_sum = n, (n * (n + 1)) & (2**64 - 1)  # checksum = sum over pairs of (r + s)
    with hj.HashJoin("atomic") as mem:
        dS = mem.to_device(S)
        dR = (mem.device_alloc(n * 8), n)
        if a.sweep == "w":
            ws = sorted(set(list(range(0, a.log_n + 1, a.w_step)) + [a.log_n]))
            handles = {al: hj.HashJoin(al, transaction_size=16) for al in a.algos.split(",")}
            for lw in ws:
                w = 1 << lw
                R = hj.generate_data("local_shuffle", n, n, w)
                mem.copy(dR[0], R.ctypes.data, n * 8)
                for al, h in handles.items():
                    res, med_us, best_us = timed(h, dR, dS, a.reps)
                    ok = res.matches == want_matches and res.checksum == want_sum
                    emit(line(al, n, n, "local_shuffle", w, 16, res, med_us, best_us, ok))
            for h in handles.values():
                h.close()
        else:
            for dist in ("sorted", "shuffle"):
                R = hj.generate_data(dist, n)
                mem.copy(dR[0], R.ctypes.data, n * 8)
                for lt in range(0, 13 if dist == "sorted" else 10):
                    for al in ("atomic", "adaptive"):
                        with hj.HashJoin(al, transaction_size=1 << lt) as h:
                            res, med_us, best_us = timed(h, dR, dS, a.reps)
                            emit(line(al, n, n, dist, 0, 1 << lt, res, med_us, best_us, res.matches == want_matches and res.checksum == want_sum))
        mem.device_free(dR[0])
        mem.device_free(dS[0])


if __name__ == "__main__":
    main()
