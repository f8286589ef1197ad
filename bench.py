#!/usr/bin/env python
"""bench.py -- join throughput of the build/probe hot path on N B200s.

  python bench.py --gpus N --steps K --warmup W            (N>1: under torchrun)
  python bench.py --impl reference ...                     (the reference's CPU code)

Workload (BASELINE.json configs[2] per GPU, configs[4] at N=8): the radix join of
|R| = N * 2^27 with |S| = N * 2^30 8-byte tuples, each rank holding a 2^27 / 2^30
slice.  A step = one whole join: [N>1: range-partition both slices by destination
GPU, all-to-all over NCCL/NVLink,] histogram -> scan -> scatter (x passes) ->
per-partition shared-memory build+probe -> {matches, checksum} [N>1: all-reduce].
`value` times steps with inputs resident in HBM; `e2e` times the same call with
HOST (pinned) inputs, H2D inside the timed region.  One JSON line on stdout.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "join M tuples/s (|R|+|S|)"
UNIT = "Mtuples/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log-r", type=int, default=27, help="log2 |R| per GPU")
    ap.add_argument("--log-s", type=int, default=30, help="log2 |S| per GPU")
    ap.add_argument("--radix-bits", type=int, default=0, help="log2 fan-out (0 = automatic)")
    ap.add_argument("--passes", type=int, default=0)
    ap.add_argument("--algo", default="radix", choices=["radix", "atomic", "adaptive", "nocc"])
    ap.add_argument("--r-dist", default="shuffle", help="DataGen distribution of R at N=1 (host generated)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--extras", action="store_true", help="also time configs[0] (atomic 16M uniform) etc.")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# =============================================================================
# reference arm / cpu_baseline: the reference's own CPU code from oracle/_ref
# (the only place besides tests/ and smoke() that may execute oracle/)
# =============================================================================
def _mc_run(exe, algo, nthreads, r, s, timeout=600):
    out = subprocess.run([exe, f"--algo={algo}", f"--nthreads={nthreads}", f"--r-size={r}", f"--s-size={s}"],
                         capture_output=True, text=True, timeout=timeout)
    lines = (out.stdout + "\n" + out.stderr).splitlines()
    usecs = result = None
    for i, l in enumerate(lines):
        if l.startswith("TOTAL-TIME-USECS") and i + 1 < len(lines):
            usecs = float(lines[i + 1].split()[0])
        if "Results =" in l:
            result = int(l.split("=")[1].split(".")[0])
    if usecs is None:
        raise RuntimeError("mc output not understood: " + "\n".join(lines[-8:]))
    return usecs, result


def cpu_reference_join(log_r, log_s, steps, warmup):
    """mc NPO (no_partitioning_join.c:537), the reference's complete build+probe with an FK
    S, all host threads, on a bounded sample with the workload's |S|/|R| ratio.  Returns
    (Mtuples/s, descriptor)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "mc_b14_p2")
    cores = os.cpu_count() or 1
    if not os.path.exists(exe):
        return None, {"kind": "reference", "cores": cores, "sample": "oracle/_ref not built"}
    r, s = 1 << log_r, 1 << log_s
    times = []
    for i in range(warmup + steps):
        us, result = _mc_run(exe, "NPO", cores, r, s)
        assert result == s, f"reference NPO returned {result}, expected {s}"
        if i >= warmup:
            times.append(us)
    med = float(np.median(times))
    desc = {"kind": "reference", "cores": cores,
            "sample": f"mc NPO build+probe (latched bucket chaining, {cores} pthreads), |R|=2^{log_r} PK, "
                      f"|S|=2^{log_s} FK, median of {steps} runs after {warmup} warm-ups; the reference's "
                      f"PRO never probes S (parallel_radix_join.c:262-276) so NPO is its only complete FK join"}
    return (r + s) / med, desc, med


def cpu_reference_extras():
    """Other reference CPU paths on bounded samples (informational)."""
    extra = {}
    cores = os.cpu_count() or 1
    try:
        exe = os.path.join(ROOT, "oracle", "_ref", "mc_b14_p2")
        us, _ = _mc_run(exe, "PRO", cores, 1 << 24, 2)
        extra["PRO_partition+build_R_only_2^24_Mtuples/s"] = round((1 << 24) / us, 1)
    except Exception as e:  # noqa: BLE001
        extra["PRO_error"] = str(e)[:80]
    try:
        from tests.oracle_lib import Reference
        ref = Reference(probe=True)
        if ref.available:
            n = 1 << 24
            R = ref.generate_data("uniform", n, n, 16)
            S = ref.generate_data("sorted", n, n, 16)
            for algo in ("atomic", "nocc"):
                j = ref.hash_build(algo, R, S, num_partitions=max(64, cores))
                extra[f"{algo}_build+probe_uniform_2^24_Mtuples/s"] = round(2 * n / j["hashBuildTimeInMicroseconds"], 1)
            extra["htm"] = "not run: host CPU has no rtm flag" if "rtm" not in open("/proc/cpuinfo").read() else "rtm present"
    except Exception as e:  # noqa: BLE001
        extra["atomic_error"] = str(e)[:80]
    return extra


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    log_r, log_s = min(args.log_r, 24), min(args.log_s, 27)
    t0 = time.time()
    val, desc, med_us = cpu_reference_join(log_r, log_s, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 2), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(med_us / 1000.0, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "radix join |R|=2^27 PK x |S|=2^30 FK per GPU (BASELINE configs[2]; configs[4] at 8 GPUs)",
                   "reference_sample": desc["sample"]},
        "cpu_baseline": dict(desc, value=round(val, 2), unit=UNIT),
        "e2e": {"value": round(val, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": round(time.time() - t0, 1),
    }
    print(json.dumps(line))
    return 0
