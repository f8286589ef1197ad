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
# one hardware queue per stream (compute + one copy stream per peer): no false dependencies between them
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")  # no kernel is loaded (a device-wide sync) in mid-pipeline

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
    ap.add_argument("--exchange", default="xchg", choices=["xchg", "regions", "fused", "nccl"],
                    help="N>1: xchg = local scatter + copy-engine transfers pipelined under pass 2 + join (default); "
                         "regions / fused = SM stores over NVLink fused with radix pass 1 (histogram-free / exact "
                         "counts first); nccl = partition + NCCL all-to-all")
    ap.add_argument("--chunks", type=int, default=0,
                    help="split S into this many chunks so the exchange of chunk c+1 overlaps pass 2 + join of chunk c "
                         "(a join is additive over disjoint parts of S); 0 = 8 for xchg, 1 otherwise")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every GPU holds 2^log-r x 2^log-s; strong: those are the GLOBAL sizes, split over the GPUs "
                         "(BASELINE configs[4]: --scaling strong --log-r 30 --log-s 33)")
    ap.add_argument("--peer-ctas", type=int, default=48, help="CTAs of an overlapped peer scatter (NVLink-bound)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-global-table", action="store_true", help="skip the global_table key (BASELINE configs[0]/[1])")
    ap.add_argument("--cpu-also-log-n", type=int, default=27, help="log2 |R| of the cpu_baseline.also runs (atomic / nocc / PRO)")
    ap.add_argument("--cpu-curves", default="", help="write the reference's CPU W-sweep on this host to this file and exit")
    ap.add_argument("--cpu-log-n", type=int, default=24)
    ap.add_argument("--cpu-w-step", type=int, default=4)
    ap.add_argument("--cpu-reps", type=int, default=3)
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
def _mc_bench(algo, nthreads, r, s, reps, warmup, timeout=1700):
    """oracle/_ref/mc_bench_b14_p2 (oracle/ref_bench.c around the reference's unmodified mc/src): relations
    generated ONCE with the reference's generators and seeds, the entry point called warmup + reps times.
    Returns (list of the callee's own TOTAL-TIME-USECS of the timed calls, list of results, generation seconds)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "mc_bench_b14_p2")
    out = subprocess.run([exe, f"--algo={algo}", f"--nthreads={nthreads}", f"--r-size={r}", f"--s-size={s}",
                          f"--reps={reps}", f"--warmup={warmup}"], capture_output=True, text=True, timeout=timeout)
    if out.returncode != 0:
        raise RuntimeError("mc_bench failed: " + (out.stdout + out.stderr)[-400:])
    # the callee prints half of its timing block to stderr (no_partitioning_join.c:322-330): parse stdout only
    lines = out.stdout.splitlines()
    usecs, results, gen_s, last = [], [], None, None
    for i, l in enumerate(lines):
        if l.startswith("GENERATED"):
            gen_s = float(l.split("GEN_US")[1]) / 1e6
        if l.startswith("TOTAL-TIME-USECS") and i + 1 < len(lines):
            last = float(lines[i + 1].split()[0])
        if l.startswith("RUN "):
            f = l.split()
            if f[-1] == "TIMED":
                usecs.append(last)
                results.append(int(f[5]))
    if not usecs:
        raise RuntimeError("mc_bench output not understood: " + "\n".join(lines[-8:]))
    return usecs, results, gen_s


def host_desc():
    model = "unknown CPU"
    try:
        for l in open("/proc/cpuinfo"):
            if l.startswith("model name"):
                model = l.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return model


def mem_available_gib():
    try:
        for l in open("/proc/meminfo"):
            if l.startswith("MemAvailable"):
                return int(l.split()[1]) / (1 << 20)
    except OSError:
        pass
    return 0.0


def cpu_reference_join(log_r, log_s, steps, warmup):
    """mc NPO (no_partitioning_join.c:537), the reference's complete build+probe with an FK S, on all host
    threads.  Returns (Mtuples/s, descriptor, median us); (None, descriptor, None) when oracle/_ref is absent."""
    exe = os.path.join(ROOT, "oracle", "_ref", "mc_bench_b14_p2")
    cores = os.cpu_count() or 1
    if not os.path.exists(exe):
        return None, {"kind": "reference", "cores": cores, "sample": "oracle/_ref not built"}, None
    r, s = 1 << log_r, 1 << log_s
    usecs, results, gen_s = _mc_bench("NPO", cores, r, s, steps, warmup)
    assert all(x == s for x in results), f"reference NPO returned {results}, expected {s}"
    med = float(np.median(usecs))
    desc = {"kind": "reference", "cores": cores, "cpu": host_desc(),
            "sample": f"mc NPO build+probe (latched bucket chaining, {cores} pthreads), |R|=2^{log_r} PK, "
                      f"|S|=2^{log_s} FK from the reference's generators (seeds 12345/54321, generated once in "
                      f"{gen_s:.0f} s), median of {steps} calls after {warmup} warm-up calls in one process "
                      f"(oracle/ref_bench.c); the reference's PRO never probes S (parallel_radix_join.c:262-276) "
                      f"so NPO is its only complete FK join",
            "runs_us": [round(x) for x in usecs]}
    return (r + s) / med, desc, med


def cpu_reference_extras(log_n=27, reps=5):
    """The reference's other CPU paths (north_star: atomic, nocc, PRJ) on the host cores: AtomicHashBuild /
    NoCCHashBuild compiled unmodified (build+probe, S = sorted 1..n as main.cpp:93), sorted and fully shuffled
    R, and PRO 14 bits / 2 passes (partition + build of R only, as the reference measured it).  BASELINE.md
    section 2.4: `reps` repetitions, median."""
    extra = {"n": f"2^{log_n}", "reps": reps, "cores": os.cpu_count() or 1, "cpu": host_desc()}
    cores = os.cpu_count() or 1
    n = 1 << log_n
    try:
        usecs, _, _ = _mc_bench("PRO", cores, n, 2, reps, 1)
        extra["PRO_b14_p2_partition+build_R_only_Mtuples/s"] = round(n / float(np.median(usecs)), 1)
    except Exception as e:  # noqa: BLE001
        extra["PRO_error"] = str(e)[:120]
    try:
        from tests.oracle_lib import Reference
        import htm_hashjoin_b200 as hj
        ref = Reference(probe=True)
        if ref.available:
            S = hj.generate_data("sorted", n, n, 16)
            for dist in ("sorted", "shuffle"):
                R = hj.generate_data(dist, n, n, 16)
                for algo in ("atomic", "nocc"):
                    ts = []
                    for _ in range(reps):
                        j = ref.hash_build(algo, R, S, num_partitions=max(64, cores))
                        assert j["totalMatches"] == n
                        ts.append(j["hashBuildTimeInMicroseconds"])
                    extra[f"{algo}_build+probe_{dist}_Mtuples/s"] = round(2 * n / float(np.median(ts)), 1)
                del R
            extra["scheduler"] = "tbb/tbb.h stand-in (oracle/tbb_shim: one std::thread per hardware thread), TBB is not installed"
        extra["htm"] = ("not run: host CPU has no rtm flag" if " rtm" not in open("/proc/cpuinfo").read()
                        else "rtm present (HTMHashBuild.hpp's probe does not compile: not timed)")
    except Exception as e:  # noqa: BLE001
        extra["atomic_error"] = str(e)[:120]
    return extra


def cpu_w_sweep(args):
    """--cpu-curves FILE: the reference's CPU paths over the shuffle window W on THIS host (the CPU side of
    figs/perf.svg): AtomicHashBuild / NoCCHashBuild build+probe compiled unmodified (R = DataGen
    local_shuffle W, S = sorted 1..n, main.cpp:93) and mc's NPO / PRO with --local-shuffle-range W, all host
    threads, median of --cpu-reps runs.  One JSON line per (algo, W), experiments/plot_perf.py --cpu-measured reads it."""
    import htm_hashjoin_b200 as hj
    from tests.oracle_lib import Reference, mc_binary
    n, cores = 1 << args.cpu_log_n, os.cpu_count() or 1
    ref = Reference(probe=True)
    exe = mc_binary(14, 2)
    out = open(args.cpu_curves, "w")

    def emit(o):
        o.update({"rSize": n, "cores": cores, "cpu": host_desc(), "reps": args.cpu_reps})
        out.write(json.dumps(o) + "\n")
        out.flush()
        print(json.dumps(o), flush=True)

    S = hj.generate_data("sorted", n)
    for lw in sorted(set(list(range(0, args.cpu_log_n + 1, args.cpu_w_step)) + [args.cpu_log_n])):
        w = 1 << lw
        if ref.available:
            R = hj.generate_data("local_shuffle", n, n, w)
            for algo in ("nocc", "atomic"):
                ts = []
                for _ in range(args.cpu_reps):
                    j = ref.hash_build(algo, R, S, num_partitions=max(64, cores))
                    assert j["totalMatches"] == n and j["inputSum"] == n * (n + 1) // 2
                    ts.append(j["hashBuildTimeInMicroseconds"])
                emit({"algo": f"{algo} (CPU, this host)", "dataDistr": "local_shuffle", "shuffleRange": w,
                      "hashBuildTimeInMicroseconds": float(np.median(ts))})
        if exe:
            for algo, s_size in (("NPO", n), ("PRO", 2)):   # experiments/motivation.sh:11, npo_probe logs
                ts = []
                for _ in range(max(1, args.cpu_reps // 2)):
                    o = subprocess.run([exe, f"--algo={algo}", f"--nthreads={cores}", f"--r-size={n}", f"--s-size={s_size}",
                                        f"--local-shuffle-range={w}"], capture_output=True, text=True, timeout=900)
                    lines = o.stdout.splitlines()
                    for i, l in enumerate(lines):
                        if l.startswith("TOTAL-TIME-USECS") and i + 1 < len(lines):
                            ts.append(float(lines[i + 1].split()[0]))
                name = "NPJ / NPO build+probe" if algo == "NPO" else "PRJ / PRO partition+build of R"
                emit({"algo": f"{name} (CPU, this host)", "dataDistr": "local_shuffle", "shuffleRange": w,
                      "hashBuildTimeInMicroseconds": float(np.median(ts))})
    out.close()
    return 0


def run_reference(args):
    """--impl reference: the reference's own CPU join on the bench's config (2^27 PK x 2^30 FK per GPU unless
    --log-r/--log-s say otherwise), all host threads.  A step = one NPO call on relations generated once."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    log_r, log_s = args.log_r, args.log_s
    same = True
    need_gib = (8.0 * ((1 << log_r) + (1 << log_s)) + 32.0 * (1 << log_r)) / (1 << 30) + 2
    avail = mem_available_gib()
    while avail and need_gib > 0.8 * avail and log_s > 20:  # only if host RAM allows
        log_r, log_s, same = log_r - 1, log_s - 1, False
        need_gib = (8.0 * ((1 << log_r) + (1 << log_s)) + 32.0 * (1 << log_r)) / (1 << 30) + 2
    if log_r > 30 or log_s > 30:  # the reference's sizes are int32_t (mc/src/generator.h)
        log_r, log_s, same = min(log_r, 27), min(log_s, 30), False
    t0 = time.time()
    val, desc, med_us = cpu_reference_join(log_r, log_s, args.steps, args.warmup)
    if val is None:
        print(json.dumps({"impl": "reference", "unavailable": desc["sample"]}))
        return 0
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 2), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(med_us / 1000.0, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": f"join |R|=2^{log_r} PK x |S|=2^{log_s} FK "
                               + ("(BASELINE configs[2], the per-GPU workload of the B200 arm)" if same else
                                  f"(REDUCED from 2^{args.log_r} x 2^{args.log_s}: host RAM {avail:.0f} GiB / int32 sizes)"),
                   "same_config_as_b200_arm": same,
                   "note": "one host runs ONE GPU's share of the weak-scaling workload; at N GPUs the B200 arm joins N times as much",
                   "reference_sample": desc["sample"]},
        "cpu_baseline": dict(desc, value=round(val, 2), unit=UNIT),
        "e2e": {"value": round(val, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": round(time.time() - t0, 1),
    }
    print(json.dumps(line))
    return 0


# =============================================================================
# B200 arm
# =============================================================================
def expected_result(D, m):
    """FK S (m tuples) over PK R = a permutation of 1..D: every S tuple matches exactly one R tuple with
    r == s, so matches = |S| and checksum = 2 * sum(S); S = |S|/D permutations of 1..D."""
    blocks, rest = divmod(m, D)
    assert rest == 0
    return m, (2 * blocks * (D * (D + 1) // 2)) % (1 << 64)


def bind_to_gpu_numa_node(local_rank):
    """Pin this process (and so the first touch of its pinned host buffers) to the CPUs of the NUMA node its GPU
    hangs off: with 4-8 ranks on one host the H2D copies otherwise cross sockets (round 1: 52 -> 22 GB/s per GPU
    at N=8).  Best effort; returns a description for the JSON line."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id  # torch >= 2.x
        bus = "%04x:%02x:%02x.0" % (torch.cuda.get_device_properties(local_rank).pci_domain_id, bus,
                                   torch.cuda.get_device_properties(local_rank).pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return f"gpu {bus}: no NUMA node reported"
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & os.sched_getaffinity(0))
        if not allowed:
            return f"gpu {bus}: node {node} has no allowed CPU"
        os.sched_setaffinity(0, allowed)
        return f"gpu {bus} -> NUMA node {node}, {len(allowed)} CPUs"
    except Exception as e:  # noqa: BLE001
        return f"not bound ({type(e).__name__}: {str(e)[:60]})"


class Rank:
    """One process = one GPU: inputs, exchange buffers and the hj handle."""

    def __init__(self, args):
        import torch
        import htm_hashjoin_b200 as hj
        self.torch, self.hj, self.args = torch, hj, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dist = None
        self.numa = bind_to_gpu_numa_node(self.local) if self.world > 1 else "single rank: not bound"
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
            from htm_hashjoin_b200 import exchange
            self.exchange = exchange
        G = self.world
        self.n, self.m = 1 << args.log_r, 1 << args.log_s
        if args.scaling == "strong":  # global sizes, split over the ranks
            assert self.n % G == 0 and self.m % G == 0
            self.n, self.m = self.n // G, self.m // G
        self.domain = G * self.n
        if args.chunks == 0:
            # measured with the page-aligned window (profiles/r02_n2_pg_c*.json, r02_n4_pg_c*.json): N=2 4 chunks
            # 12.0 ms vs 12.7 with 8; N=4 13.5 vs 13.9 (more, smaller copies: more per-copy gaps and token waits)
            args.chunks = 4 if (G > 1 and args.exchange == "xchg") else 1
        self.xchg = G > 1 and args.exchange == "xchg"
        self.h = hj.HashJoin(args.algo, device=self.local, num_partitions=(1 << args.radix_bits) if args.radix_bits else 0,
                             passes=args.passes)
        dev = torch.device("cuda", self.local)
        self.dR = torch.empty(self.n, dtype=torch.int64, device=dev)
        self.dS = torch.empty(self.m, dtype=torch.int64, device=dev)
        t0 = time.time()
        if G == 1 and args.r_dist != "feistel":
            # R: the reference's own DataGen distribution (bit-identical, host generated)
            self.hR = hj.pinned_empty(self.n)
            self.hR[:] = hj.generate_data(args.r_dist, self.n, self.n, 16)
            self.h.copy(self.dR.data_ptr(), self.hR.ctypes.data, self.n * 8)
            self.r_desc = f"DataGen.hpp '{args.r_dist}' srand(0), host generated"
        else:
            self.h.generate(1, self.rank * self.n, self.n, self.domain, 12345, self.dR.data_ptr())
            self.hR = None
            self.r_desc = "counter-based PK permutation (hj_generate kind 1, seed 12345), slice per rank"
        # S: FK relation = |S|/|R| independent permutations of 1..|R| (structure of
        # mc/src/generator.c:409-445), generated on the device (rand() is too slow at 2^30)
        self.h.generate(1, self.rank * self.m, self.m, self.domain, 54321, self.dS.data_ptr())
        self.s_desc = "counter-based FK (hj_generate kind 1, seed 54321: |S|/|R| permutations of 1..|R|)"
        self.hS = None
        self.gen_s = time.time() - t0
        self.fused = G > 1 and args.exchange in ("fused", "regions")
        self.regions = G > 1 and args.exchange == "regions"
        if G > 1:
            cap_r = self.n + self.n // 32 + 65536
            cap_s = self.m + self.m // 32 + 65536
        if self.xchg:
            # pipelined exchange over the copy engines (include/hj_b200.h, hj_xchg_*): local scatter by
            # (owner GPU, low b1 bits) into G << b1 <= 256 regions, pass 2 takes the remaining bits
            bits, _, _, passes = self.h.radix_plan(self.n)
            assert passes == 2, "the exchange needs the 2-pass plan"
            self.b1 = min(6, (256 // G).bit_length() - 1, bits - 1)
            self.h.close()
            self.h = hj.HashJoin(args.algo, device=self.local, num_partitions=1 << bits, passes=2, pass1_bits=self.b1)
            self.f1 = 1 << self.b1
            self.chunks = max(1, args.chunks)
            dom_bits = self.exchange.domain_bits_for(self.domain)
            win, self.window_bytes = self.h.xchg_create(G, self.rank, self.n, self.m, self.chunks, dom_bits)
            handles = self.exchange.exchange_ipc_handles(self.h.ipc_export(win))
            peers = [win if g == self.rank else self.h.ipc_open(handles[g]) for g in range(G)]
            self.h.xchg_connect(peers)
            self.epoch = 0
            if os.environ.get("HJ_BENCH_COPYPROBE"):
                # diagnostic: what the copy engines give THIS process (torch + NCCL initialised, the exchange's own
                # window mapping) for plain ring-ordered block copies in lock step, no compute beside them
                import torch.distributed as dist
                blk = 512 << 20
                fresh_src = self.h.device_alloc(G * blk)
                fresh_win = self.h.device_alloc(G * blk)
                hs2 = self.exchange.exchange_ipc_handles(self.h.ipc_export(fresh_win))
                peers2 = [fresh_win if g == self.rank else self.h.ipc_open(hs2[g]) for g in range(G)]
                variants = [("window<-dS", peers, self.dS.data_ptr()), ("window<-fresh", peers, fresh_src),
                            ("fresh<-dS", peers2, self.dS.data_ptr()), ("fresh<-fresh", peers2, fresh_src)]
                print(f"[copyprobe r{self.rank}] mod 2MiB: window {win % (2 << 20)} dS {self.dS.data_ptr() % (2 << 20)} "
                      f"peer windows {[q % (2 << 20) for q in peers]} window_bytes {self.window_bytes}", file=sys.stderr, flush=True)
                for name, pw, src in variants:
                    torch.cuda.synchronize()
                    dist.barrier()
                    t0 = time.perf_counter()
                    for _ in range(4):
                        for k in range(1, G):
                            p = (self.rank + k) % G
                            self.h.copy(pw[p] + self.rank * blk, src + p * blk, blk)
                    dt = time.perf_counter() - t0
                    print(f"[copyprobe r{self.rank}] {name}: {4 * (G - 1) * blk / dt / 1e9:.1f} GB/s "
                          f"({dt * 1e3 / (4 * (G - 1)):.3f} ms per 512 MiB block)", file=sys.stderr, flush=True)
                    dist.barrier()
        elif self.fused:
            # exchange == radix pass 1: every rank stores straight into the owners' pass-1 buffers
            # (CUDA IPC over NVLink).  The fan-out is fixed so all ranks agree on b1.
            bits, self.b1, _, passes = self.h.radix_plan(self.n)
            assert passes == 2, "the fused exchange needs the 2-pass plan"
            # the fused pass scatters into G << b1 bins: keep that near 128 (NVLink wants long runs,
            # measured 434 GB/s at 512 bins vs 565 at 128), pass 2 takes the remaining bits
            while self.b1 > 3 and (G << self.b1) > 128:
                self.b1 -= 1
            self.h.close()
            self.h = hj.HashJoin(args.algo, device=self.local, num_partitions=1 << bits, passes=2,
                                 pass1_bits=self.b1)
            self.f1 = 1 << self.b1
            self.recvR, self.recvS = self.h.device_alloc(cap_r * 8), self.h.device_alloc(cap_s * 8)
            self.cap = (cap_r, cap_s)
            handles = self.exchange.exchange_ipc_handles((self.h.ipc_export(self.recvR), self.h.ipc_export(self.recvS)))
            self.peerR = [self.recvR if g == self.rank else self.h.ipc_open(handles[g][0]) for g in range(G)]
            self.peerS = [self.recvS if g == self.rank else self.h.ipc_open(handles[g][1]) for g in range(G)]
            if self.regions:
                ex = self.exchange
                self.reg_cap = (ex.region_cap(self.n, G, self.f1), ex.region_cap(self.m, G, self.f1))
                self.reg_trash = (self.f1 * G * self.reg_cap[0], self.f1 * G * self.reg_cap[1])
                assert self.reg_trash[0] + 16386 <= cap_r and self.reg_trash[1] + 16386 <= cap_s
                self.reg_init = (ex.region_cursor_init(self.rank, G, self.f1, self.reg_cap[0], dev),
                                 ex.region_cursor_init(self.rank, G, self.f1, self.reg_cap[1], dev))
                self.reg_vec = torch.zeros(2 * G * self.f1 + 1, dtype=torch.int64, device=dev)
            self.chunks = max(1, args.chunks)
            self.peer_ctas = args.peer_ctas
            self.cnt = torch.zeros((1 + self.chunks) * G * self.f1, dtype=torch.int64, device=dev)
            if self.chunks > 1:
                assert self.m % (2 * self.chunks) == 0
                # second handle = second stream: the exchange of chunk c+1 runs beside the join of chunk c
                self.hx = hj.HashJoin(args.algo, device=self.local, num_partitions=1 << bits, passes=2,
                                      pass1_bits=self.b1)
                self.cap_c = (cap_s // self.chunks) & ~1  # per-chunk receive buffer (16-byte aligned)
        elif G > 1:
            self.sendR = torch.empty(self.n, dtype=torch.int64, device=dev)
            self.sendS = torch.empty(self.m, dtype=torch.int64, device=dev)
            self.recvR = torch.empty(cap_r, dtype=torch.int64, device=dev)
            self.recvS = torch.empty(cap_s, dtype=torch.int64, device=dev)
            self.offR = torch.empty(G + 1, dtype=torch.int64, device=dev)
            self.offS = torch.empty(G + 1, dtype=torch.int64, device=dev)
        if G == 1:
            self.h.reserve(self.n, self.m)
            self.h.reserve_staging(self.n, self.m)
        elif self.xchg:
            pass  # hj_xchg_create sized everything
        else:
            self.h.reserve(self.n + self.n // 16 + 65536, self.m + self.m // 16 + 65536)
        self.last = None
        self.launches = 0
        self.nvlink_bytes = 0

    def host_inputs(self):
        """Pinned host copies of this rank's slices (for the e2e leg)."""
        if self.hR is None:
            self.hR = self.hj.pinned_empty(self.n)
            self.h.copy(self.hR.ctypes.data, self.dR.data_ptr(), self.n * 8)
        if self.hS is None:
            self.hS = self.hj.pinned_empty(self.m)
            self.h.copy(self.hS.ctypes.data, self.dS.data_ptr(), self.m * 8)

    def step(self, host=False):
        """One whole join.  host=True: inputs start in pinned host memory."""
        torch, G = self.torch, self.world
        if G == 1:
            R = self.hR if host else self.dR
            S = self.hS if host else self.dS
            res = self.h.join(R, S)
            self.launches += res.kernel_launches
            self.last = res
            return res.matches, res.checksum
        if host:  # H2D of this rank's slices, then the same device pipeline
            self.h.copy(self.dR.data_ptr(), self.hR.ctypes.data, self.n * 8)
            self.h.copy(self.dS.data_ptr(), self.hS.ctypes.data, self.m * 8)
        ex = self.exchange
        dom_bits = ex.domain_bits_for(self.domain)
        if self.xchg:
            return self.step_xchg()
        if self.regions and self.chunks == 1:
            out = self.step_regions(dom_bits)
            if out is not None:
                return out
            self.region_fallbacks = getattr(self, "region_fallbacks", 0) + 1  # skewed keys: exact protocol
        if self.fused:
            return self.step_fused_chunked(dom_bits) if self.chunks > 1 else self.step_fused(dom_bits)
        # 1. split both slices by destination GPU (high key bits of the key domain)
        self.h.exchange_partition(self.dR.data_ptr(), self.n, G, dom_bits, self.sendR.data_ptr(), self.offR.data_ptr())
        self.h.exchange_partition(self.dS.data_ptr(), self.m, G, dom_bits, self.sendS.data_ptr(), self.offS.data_ptr())
        self.launches += 6  # hist + scan + scatter per relation
        # 2. counts, then the tuples, all-to-all over NCCL (NVLink 5 / NVSwitch)
        sc_r = self.offR[1:] - self.offR[:-1]
        sc_s = self.offS[1:] - self.offS[:-1]
        rc_r, rc_s = ex.exchange_counts(sc_r), ex.exchange_counts(sc_s)
        counts = torch.stack([sc_r, sc_s, rc_r, rc_s]).cpu().tolist()
        nr = ex.exchange_tuples(self.sendR, counts[0], self.recvR, counts[2])
        ns = ex.exchange_tuples(self.sendS, counts[1], self.recvS, counts[3])
        torch.cuda.synchronize()
        self.nvlink_bytes += 8 * (self.n - counts[0][self.rank] + self.m - counts[1][self.rank])
        # 3. local join of the received partitions
        res = self.h.join((self.recvR.data_ptr(), nr), (self.recvS.data_ptr(), ns))
        self.launches += res.kernel_launches
        self.last = res
        # 4. {matches, checksum} summed over ranks
        return ex.reduce_result(res.matches, res.checksum, "cuda")

    def step_xchg(self):
        """Pipelined exchange over the copy engines: one enqueue, one wait, one all-reduce (which is also the
        barrier the next epoch needs)."""
        t = [time.perf_counter()]
        self.epoch += 1
        self.h.xchg_start(self.dR.data_ptr(), self.dS.data_ptr(), self.epoch)
        t.append(time.perf_counter())
        ovf = 0
        try:
            res = self.h.xchg_finish()
            matches, checksum = res.matches, res.checksum
            self.launches += res.kernel_launches
            self.last = res
        except self.hj.HJError as e:
            if e.code != self.hj.HJ_EOVERFLOW:
                raise
            ovf, matches, checksum = 1, 0, 0
        t.append(time.perf_counter())
        out = self.exchange.reduce_result3(matches, checksum, ovf, "cuda")
        t.append(time.perf_counter())
        self.fused_ms = [round((b - a) * 1e3, 3) for a, b in zip(t, t[1:])]  # enqueue, pipeline, allreduce
        if out[2]:
            raise RuntimeError("hj_xchg: a region overflowed on some rank (skewed keys): run with --exchange fused")
        self.nvlink_bytes += 8 * (self.n + self.m) * (self.world - 1) // self.world  # uniform keys: (G-1)/G leaves
        return out[0], out[1]

    def step_fused(self, dom_bits):
        """Exchange fused with radix pass 1 (include/hj_b200.h, hj_peer_*)."""
        torch, G, ex, gf = self.torch, self.world, self.exchange, self.world * self.f1
        t = [time.perf_counter()]
        # 1. per (owner GPU, low b1 bits) counts of the local slices; all ranks learn all counts
        self.h.peer_hist(self.dR.data_ptr(), self.n, G, dom_bits, self.b1, self.cnt.data_ptr())
        self.h.peer_hist(self.dS.data_ptr(), self.m, G, dom_bits, self.b1, self.cnt.data_ptr() + 8 * gf)
        t.append(time.perf_counter())
        allc = ex.gather_counts(self.cnt)
        cur_r, recv_r, off_r = ex.peer_cursors(allc[:, :gf].contiguous(), self.rank, G, self.f1)
        cur_s, recv_s, off_s = ex.peer_cursors(allc[:, gf:].contiguous(), self.rank, G, self.f1)
        own = allc[self.rank, self.rank * self.f1:(self.rank + 1) * self.f1].sum() + \
            allc[self.rank, gf + self.rank * self.f1:gf + (self.rank + 1) * self.f1].sum()
        nr, ns, own = torch.stack([recv_r[self.rank], recv_s[self.rank], own]).tolist()  # one sync: cursors ready too
        if nr > self.cap[0] or ns > self.cap[1]:
            raise RuntimeError("exchange receive buffer too small")
        t.append(time.perf_counter())
        # 2. scatter into the owners' pass-1 buffers: plain stores over NVLink
        self.h.peer_scatter(self.dR.data_ptr(), self.n, G, dom_bits, self.b1, self.peerR, cur_r.data_ptr())
        self.h.peer_scatter(self.dS.data_ptr(), self.m, G, dom_bits, self.b1, self.peerS, cur_s.data_ptr())
        t.append(time.perf_counter())
        self.launches += 4
        self.nvlink_bytes += 8 * (self.n + self.m - own)
        # 3. everybody's stores have landed -> histogram + pass 2 + join of what arrived
        self.dist.barrier()
        torch.cuda.synchronize()
        t.append(time.perf_counter())
        res = self.h.join_pass1(self.recvR, nr, self.recvS, ns, off_r.data_ptr(), off_s.data_ptr())
        t.append(time.perf_counter())
        self.launches += res.kernel_launches
        self.last = res
        out = ex.reduce_result(res.matches, res.checksum, "cuda")
        t.append(time.perf_counter())
        self.fused_ms = [round((b - a) * 1e3, 3) for a, b in zip(t, t[1:])]  # hist, counts, scatter, barrier, join, reduce
        return out

    def step_regions(self, dom_bits):
        """Histogram-free exchange (hj_peer_scatter_regions / hj_join_regions): no counting pass, the
        all-gather of the final cursors is the only collective before the join and doubles as the barrier.
        Returns None when a region overflowed anywhere (the caller redoes the step with exact counts)."""
        torch, G, ex, gf = self.torch, self.world, self.exchange, self.world * self.f1
        t = [time.perf_counter()]
        v = self.reg_vec
        v[:gf].copy_(self.reg_init[0])
        v[gf:2 * gf].copy_(self.reg_init[1])
        torch.cuda.synchronize()
        ovf = self.h.peer_scatter_regions(self.dR.data_ptr(), self.n, G, dom_bits, self.b1, self.peerR, v.data_ptr(),
                                          self.rank, self.reg_cap[0], self.reg_trash[0])
        ovf |= self.h.peer_scatter_regions(self.dS.data_ptr(), self.m, G, dom_bits, self.b1, self.peerS,
                                           v.data_ptr() + 8 * gf, self.rank, self.reg_cap[1], self.reg_trash[1])
        v[2 * gf] = int(ovf)
        t.append(time.perf_counter())
        allf = ex.gather_counts(v)  # everybody's stores have landed once this returns
        cur_r, nr = ex.regions_received(allf[:, :gf].contiguous(), self.rank, G, self.f1, self.reg_cap[0])
        cur_s, ns = ex.regions_received(allf[:, gf:2 * gf].contiguous(), self.rank, G, self.f1, self.reg_cap[1])
        own = (v[self.rank * self.f1:(self.rank + 1) * self.f1] - self.reg_init[0][:self.f1]).sum() + \
            (v[gf + self.rank * self.f1:gf + (self.rank + 1) * self.f1] - self.reg_init[1][:self.f1]).sum()
        nr, ns, own, any_ovf = torch.stack([nr, ns, own, allf[:, 2 * gf].sum()]).tolist()
        t.append(time.perf_counter())
        self.launches += 2
        if any_ovf:
            return None
        self.nvlink_bytes += 8 * (self.n + self.m - own)
        res = self.h.join_regions(self.recvR, nr, cur_r.data_ptr(), self.reg_cap[0], self.recvS, ns, cur_s.data_ptr(),
                                  self.reg_cap[1], G)
        t.append(time.perf_counter())
        self.launches += res.kernel_launches
        self.last = res
        out = ex.reduce_result(res.matches, res.checksum, "cuda")
        t.append(time.perf_counter())
        self.fused_ms = [round((b - a) * 1e3, 3) for a, b in zip(t, t[1:])]  # scatter, gather, join, reduce
        return out

    def step_fused_chunked(self, dom_bits):
        """step_fused with S cut into chunks by position: join(R, S) = sum over chunks of join(R, S_c), so
        chunk c can be joined (handle h) while chunk c+1 still travels (handle hx, a few CTAs: NVLink-bound)."""
        torch, G, ex, gf, C = self.torch, self.world, self.exchange, self.world * self.f1, self.chunks
        mc = self.m // C
        t = [time.perf_counter()]
        self.hx.peer_hist(self.dR.data_ptr(), self.n, G, dom_bits, self.b1, self.cnt.data_ptr())
        for c in range(C):
            self.hx.peer_hist(self.dS.data_ptr() + 8 * c * mc, mc, G, dom_bits, self.b1,
                              self.cnt.data_ptr() + 8 * gf * (1 + c))
        t.append(time.perf_counter())
        allc = ex.gather_counts(self.cnt)
        cur, recv, off, own = [], [], [], 0
        for c in range(1 + C):
            a = allc[:, c * gf:(c + 1) * gf].contiguous()
            cu, rc, of = ex.peer_cursors(a, self.rank, G, self.f1)
            cur.append(cu); recv.append(rc[self.rank]); off.append(of)
            own = own + a[self.rank, self.rank * self.f1:(self.rank + 1) * self.f1].sum()
        sizes = torch.stack(recv + [own]).tolist()  # one sync: cursors ready too
        nr, ns, own = sizes[0], sizes[1:1 + C], sizes[-1]
        if nr > self.cap[0] or max(ns) > self.cap_c:
            raise RuntimeError("exchange receive buffer too small")
        t.append(time.perf_counter())
        peer_c = lambda c: [p + 8 * c * self.cap_c for p in self.peerS]
        self.hx.peer_scatter_async(self.dR.data_ptr(), self.n, G, dom_bits, self.b1, self.peerR, cur[0].data_ptr())
        self.hx.peer_scatter_async(self.dS.data_ptr(), mc, G, dom_bits, self.b1, peer_c(0), cur[1].data_ptr())
        self.hx.sync()
        self.dist.barrier()
        torch.cuda.synchronize()
        t.append(time.perf_counter())
        matches = checksum = 0
        for c in range(C):
            if c + 1 < C:  # next chunk leaves on a few CTAs while this one is joined on the rest
                self.hx.peer_scatter_async(self.dS.data_ptr() + 8 * (c + 1) * mc, mc, G, dom_bits, self.b1,
                                           peer_c(c + 1), cur[2 + c].data_ptr(), self.peer_ctas)
            res = self.h.join_pass1(self.recvR, nr, self.recvS + 8 * c * self.cap_c, ns[c], off[0].data_ptr(),
                                    off[1 + c].data_ptr())
            matches += res.matches
            checksum = (checksum + res.checksum) & (2**64 - 1)
            self.launches += res.kernel_launches
            if c + 1 < C:
                self.hx.sync()
                self.dist.barrier()
                torch.cuda.synchronize()
        t.append(time.perf_counter())
        self.launches += 2 + 2 * C
        self.nvlink_bytes += 8 * (self.n + self.m - own)
        self.last = res
        out = ex.reduce_result(matches, checksum, "cuda")
        t.append(time.perf_counter())
        self.fused_ms = [round((b - a) * 1e3, 3) for a, b in zip(t, t[1:])]  # hist, counts, R + chunk 0, pipeline, reduce
        return out

    def barrier(self):
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t[0])

    def timed(self, steps, host, want):
        """K steps bracketed by barrier + synchronize, CUDA events, max over ranks (ms)."""
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        phases = {}
        self.launches = 0
        self.nvlink_bytes = 0
        self.barrier()
        e0.record()
        for _ in range(steps):
            got = self.step(host)
            assert got == want, f"join result {got} != expected {want}"
            r = self.last
            for k in ("us_total", "us_hist", "us_scan", "us_scatter1", "us_scatter2", "us_join", "us_build",
                      "us_probe", "us_clear", "us_h2d"):
                phases[k] = phases.get(k, 0.0) + getattr(r, k)
            phases["us_wait"] = phases.get("us_wait", 0.0) + r.us_reserved[0]
        e1.record()
        torch.cuda.synchronize()
        ms = self.max_over_ranks(e0.elapsed_time(e1))
        self.barrier()
        return ms, {k: v / steps for k, v in phases.items()}


# ncu --set full, profiles/r02_ncu_global_*: DRAM bytes per tuple of the two kernels (read + write) -- filled from the
# committed captures; None = not captured for that input
GT_TRAFFIC = {
    # profiles/r02_ncu_global_n27_raw.csv, sorted 2^27: gt_build_kernel 1.091 GB read + 1.035 GB written (0.670 ms under
    # ncu), gt_probe_home_kernel 2.148 GB read + 0.004 GB written (0.306 ms = 7.0 TB/s): no re-reads on either side
    "c2_atomic_sorted_2^27": 4277000000,
}


def global_table_lines(rk, reps=5, warm=2):
    """BASELINE configs[0] / [1] on the global table (north_star's first subsystem), device resident:
    C1 = --algo atomic --rSize 16M --dataDistr uniform; sorted and local_shuffle W=2^8 at |R|=|S|=2^27.
    Time = build + probe (the reference's timed region, AtomicHashBuild.hpp:34-88: the table is zeroed
    before the timer there too); algorithmic bytes 16 B per R tuple + 16 B per S tuple."""
    hj = rk.hj
    peak, _ = load_peaks()
    out = {}
    for name, dist, rng, log_n in (("c1_atomic_uniform_16M", "uniform", 16, 24),
                                   ("c2_atomic_sorted_2^27", "sorted", 16, 27),
                                   ("c2_atomic_local_shuffle_W256_2^27", "local_shuffle", 256, 27)):
        n = 1 << log_n
        R = hj.generate_data(dist, n, n, rng)
        with hj.HashJoin("atomic", device=rk.local) as h:
            dR = h.to_device(R)
            del R
            dS = h.device_alloc(n * 8)
            h.generate(0, 0, n, 1, 0, dS)  # S = sorted 1..n (main.cpp:93)
            h.reserve(n, n)
            rows = []
            for i in range(reps + warm):
                r = h.join(dR, (dS, n))
                assert r.matches == n, (name, r.matches)
                if i >= warm:
                    rows.append((r.us_build, r.us_probe, r.us_clear))
            ub, up, uc = (float(np.median([x[k] for x in rows])) for k in range(3))
            us = ub + up
            b = 32.0 * n
            out[name] = {"Mtuples/s": round(2 * n / us, 1), "us_build": round(ub, 1), "us_probe": round(up, 1),
                         "us_table_clear_untimed": round(uc, 1), "algorithmic_bytes": int(b),
                         "achieved_GBs": round(b / (us * 1e-6) / 1e9, 1), "frac": round(b / (us * 1e-6) / 1e9 / peak, 4),
                         "traffic": GT_TRAFFIC.get(name), "conflicts": int(r.conflicts), "matches": int(r.matches),
                         "launches_per_join": int(r.kernel_launches)}
            h.device_free(dR[0]); h.device_free(dS)
    return out


def run_b200(args):
    rk = Rank(args)
    G = rk.world
    want = expected_result(G * rk.n, G * rk.m) if args.algo in ("radix", "adaptive", "atomic") else None
    peak, peak_src = load_peaks()
    tuples = G * (rk.n + rk.m)
    # ---- device-resident: `value` ----
    for _ in range(args.warmup):
        got = rk.step(False)
        assert got == want, f"warm-up join result {got} != expected {want}"
    sampler = ClockSampler(rk.local) if rk.rank == 0 else None
    if sampler:
        sampler.start()
    ms, ph = rk.timed(args.steps, False, want)
    launches, nvl = rk.launches, rk.nvlink_bytes // max(args.steps, 1)
    last = rk.last
    clocks = sampler.stop() if sampler else None
    value = tuples * args.steps / (ms * 1e-3) / 1e6
    # ---- end to end from pinned host memory: `e2e` ----
    e2e = None
    if not args.no_e2e:
        rk.host_inputs()
        for _ in range(1):
            rk.step(True)
        ems, eph = rk.timed(max(2, min(args.steps, 3)), True, want)
        esteps = max(2, min(args.steps, 3))
        e2e = {"value": round(tuples * esteps / (ems * 1e-3) / 1e6, 2), "unit": UNIT,
               "h2d_bytes_per_step": 8 * tuples, "d2h_bytes_per_step": 16 * 8,
               "ms_per_step": round(ems / esteps, 3), "steps": esteps,
               "note": "inputs in pinned host memory, H2D inside hj_join (copy stream), result counters read back"}
    if rk.rank != 0:
        return 0
    # ---- roofline of the dominant kernel ----
    if last.path_taken == rk.hj.HJ_PATH_RADIX:
        P = last.passes
        n_loc, m_loc = rk.n, rk.m  # per-GPU tuples (exact at N=1, +-0.1% after the exchange)
        sc_us = ph["us_scatter1"] + ph["us_scatter2"]
        sc_bytes = 16.0 * P * (n_loc + m_loc)  # every scatter launch reads 8 B and writes 8 B per tuple
        n_launch = 2 * P
        local_bytes = float(last.hbm_bytes)  # what ran on this GPU inside us_total
        if getattr(rk, "xchg", False):
            # both passes are local kernels (the copy engines move the blocks in between); they share HBM with
            # the DMA traffic of the exchange (8 B read + 8 B written per tuple that travels)
            n_launch = 2 * (1 + rk.chunks)
        if getattr(rk, "fused", False):
            # pass 1 ran as the peer scatter (NVLink-bound, timed on the host as fused_ms[2]); the
            # HBM roofline line is pass 2 alone
            P = 1
            n_launch = 2
            sc_us, sc_bytes = ph["us_scatter2"], 16.0 * (n_loc + m_loc)
            local_bytes = 24.0 * (n_loc + m_loc)  # us_total of hj_join_regions/_pass1 holds pass 2 + join only
        roof = {"bound": "hbm", "kernel": f"rx_scatter_kernel<16> ({n_launch} launches/step: R and S (chunks) x {P} passes)",
                "achieved": round(sc_bytes / (sc_us * 1e-6) / 1e9, 1), "peak": peak, "unit": "GB/s",
                "frac": round(sc_bytes / (sc_us * 1e-6) / 1e9 / peak, 4),
                # ncu --set full, profiles/r01_ncu_radix_v4_raw.csv: the S launches over 2^28 tuples moved
                # dram__bytes_read 2.148 GB + dram__bytes_write 2.100 GB = 15.83 B/tuple (algorithmic: 16)
                "traffic": int(15.83 * sc_bytes / 16.0 / n_launch),
                "traffic_source": "15.83 B/tuple measured by ncu at |S|=2^28 (profiles/r01_ncu_radix_v4_raw.csv), "
                                  "scaled to this launch size",
                "peak_source": peak_src, "avg_launch_us": round(sc_us / n_launch, 1),
                "algorithmic_bytes_per_launch": int(sc_bytes / n_launch),
                "whole_join": {"algorithmic_bytes": int(local_bytes),
                               "achieved": round(local_bytes / (ph["us_total"] * 1e-6) / 1e9, 1),
                               "frac": round(local_bytes / (ph["us_total"] * 1e-6) / 1e9 / peak, 4),
                               "note": "algorithmic bytes of the kernels inside us_total on one GPU / that time"}}
    else:
        us = ph["us_build"] + ph["us_probe"]
        b = 16.0 * (rk.n + rk.m)
        roof = {"bound": "hbm", "kernel": "gt_build_kernel + gt_probe_kernel", "achieved": round(b / (us * 1e-6) / 1e9, 1),
                "peak": peak, "unit": "GB/s", "frac": round(b / (us * 1e-6) / 1e9 / peak, 4), "traffic": None,
                "peak_source": peak_src}
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": G, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "u64", "data": f"synthetic: R = {rk.r_desc}; S = {rk.s_desc}",
        "config": {"workload": (f"radix join |R|=2^{args.log_r} PK x |S|=2^{args.log_s} FK per GPU "
                                f"(BASELINE configs[2] at N=1; x8 = configs[4])" if args.scaling == "weak" else
                                f"radix join |R|=2^{args.log_r} PK x |S|=2^{args.log_s} FK in total, split over {G} GPUs "
                                f"(BASELINE configs[4] is 2^30 x 2^33)"),
                   "algo": args.algo, "radix_bits": int(last.radix_bits), "passes": int(last.passes),
                   "tuples_per_step": tuples, "l2": f"inputs ({8 * (rk.n + rk.m) / 2**30:.0f} GiB/GPU) larger than L2, no flush needed",
                   "exchange": "none" if G == 1 else (f"pipelined over the copy engines: local scatter by (owner GPU, low {rk.b1} bits) into "
                                                      f"fixed-capacity regions, one device-to-device copy per (peer, stage) over NVLink "
                                                      f"(CUDA IPC) under pass 2 + join of the previous stage, ring order kept in lock step by tokens; S in {rk.chunks} chunks"
                                                      if args.exchange == "xchg" else
                                                      "fused with radix pass 1, histogram-free: TMA stores into fixed-capacity regions of the owners' pass-1 buffers over NVLink (CUDA IPC)"
                                                      if args.exchange == "regions" else
                                                      "fused with radix pass 1: stores into the owners' pass-1 buffers over NVLink (CUDA IPC)"
                                                      if args.exchange == "fused" else
                                                      "range partition + torch.distributed all_to_all_single (NCCL)")},
        "phases_us": {k: round(v, 1) for k, v in ph.items()},
        "roofline": roof, "gpu_launches": int(launches), "clocks": clocks, "e2e": e2e,
        "result": {"matches": want[0], "checksum": want[1], "verified_every_step": True},
        "host": {"cpus": os.cpu_count(), "numa_binding_rank0": rk.numa},
    }
    if G > 1:
        line["nvlink_bytes_per_gpu_per_step"] = int(nvl)
        if getattr(rk, "xchg", False):
            line["xchg_ms_last_step"] = dict(zip(["enqueue", "pipeline (device)", "allreduce"], rk.fused_ms))
            line["xchg_wait_us_per_step"] = round(float(ph.get("us_wait", 0.0)), 1)
            line["nvlink_gbs_over_step"] = round(nvl / (ms / args.steps * 1e-3) / 1e9, 1)
            line["window_bytes"] = int(rk.window_bytes)
        if getattr(rk, "fused", False):
            if getattr(rk, "regions", False) and rk.chunks == 1 and not getattr(rk, "region_fallbacks", 0):
                names, sc = ["peer_scatter", "cursor all-gather (= barrier)", "pass2+join", "allreduce"], 0
            elif rk.chunks > 1:
                names, sc = ["peer_hist", "counts+cursors", "scatter R + chunk 0", "pipeline", "allreduce"], 2
            else:
                names, sc = ["peer_hist", "counts+cursors", "peer_scatter", "barrier", "pass2+join", "allreduce"], 2
            line["fused_exchange_ms_last_step"] = dict(zip(names, rk.fused_ms))
            if rk.chunks == 1:
                line["nvlink_gbs_in_peer_scatter"] = round(nvl / (rk.fused_ms[sc] * 1e-3) / 1e9, 1)
            line["region_fallbacks"] = getattr(rk, "region_fallbacks", 0)
    if G == 1 and not args.no_global_table:
        try:
            rk.h.close()  # its 28 GiB of radix buffers are not needed any more
            line["global_table"] = global_table_lines(rk)
        except Exception as e:  # noqa: BLE001  (the headline line is printed no matter what)
            line["global_table"] = {"error": str(e)[:200]}
    if G == 1 and not args.no_cpu_baseline:
        try:
            # bounded sample (1/8 of the workload, same |S|/|R|); `--impl reference` runs the full config
            v, desc, _ = cpu_reference_join(24, 27, 3, 1)
            line["cpu_baseline"] = dict(desc, value=None if v is None else round(v, 2), unit=UNIT)
            if v is not None:
                line["cpu_baseline"]["also"] = cpu_reference_extras(args.cpu_also_log_n)
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference",
                                    "sample": "failed: " + str(e)[:200]}
    print(json.dumps(line))
    return 0


def main():
    args = parse_args()
    if args.cpu_curves:
        return cpu_w_sweep(args)
    if args.impl == "reference":
        return run_reference(args)
    rc = run_b200(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass
    return rc


if __name__ == "__main__":
    sys.exit(main())
