"""htm_hashjoin_b200 -- B200-native build/probe hash-join hot path.

Python is only the test/bench harness over the C ABI (include/hj_b200.h,
csrc/libhj_b200.so).  The functions below mirror the reference's interface for
this path with the same names and argument meaning:

  generate_data          include/DataGen.hpp:26
  AtomicHashBuild        AtomicHashBuild.hpp:14-20
  NoCCHashBuild          NoCCHashBuild.hpp:13-19
  HTMHashBuild           HTMHashBuild.hpp:54-60  (adaptive path on the GPU)
  SortMerge              SortMerge.hpp:5-6  (the sort-merge comparator)
  PRO                    mc/src/parallel_radix_join.h:33-34

Each *HashBuild returns the dict of the JSON line the reference prints
(AtomicHashBuild.hpp:133-152) plus the extra keys.  Importing this package
fails loudly when the CUDA library has not been built; nothing here computes
a join on the CPU.
"""
import ctypes as C
import json

import numpy as np

from . import _capi
from ._capi import (HJ_ALGO_ADAPTIVE, HJ_ALGO_ATOMIC, HJ_ALGO_NOCC, HJ_ALGO_RADIX, HJ_ALGO_SORTMERGE, HJ_EINVAL,
                    HJ_EOVERFLOW, HJ_OK, HJ_PATH_GLOBAL, HJ_PATH_RADIX, HJ_PATH_SORTMERGE, hj_config, hj_result)

lib = _capi.load()  # raises ImportError if libhj_b200.so is missing

ALGOS = {"atomic": HJ_ALGO_ATOMIC, "nocc": HJ_ALGO_NOCC, "radix": HJ_ALGO_RADIX,
         "adaptive": HJ_ALGO_ADAPTIVE, "htm": HJ_ALGO_ADAPTIVE, "sortmerge": HJ_ALGO_SORTMERGE}


class HJError(RuntimeError):
    def __init__(self, code, where, detail=""):
        self.code = code
        msg = lib.hj_strerror(code).decode()
        super().__init__(f"{where}: {msg} [{code}] {detail}")


def _ptr(x):
    """(address, element count) of a host numpy array, a torch tensor, or a raw int pointer."""
    if x is None:
        return None, 0
    if isinstance(x, tuple):  # (raw device pointer, count)
        return int(x[0]), int(x[1])
    if isinstance(x, np.ndarray):
        if x.dtype != np.uint64 or not x.flags["C_CONTIGUOUS"]:
            raise TypeError("host relations must be C-contiguous uint64 arrays")
        return x.ctypes.data, x.size
    if hasattr(x, "data_ptr"):  # torch tensor (int64/uint64, cpu or cuda)
        if x.element_size() != 8 or not x.is_contiguous():
            raise TypeError("tensors must be contiguous with 8-byte elements")
        return x.data_ptr(), x.numel()
    raise TypeError(f"unsupported relation type {type(x)}")


# ---- inputs (csrc/datagen.cpp) ------------------------------------------------
def generate_data(dist, size, distinct_keys=None, local_shuffle_range=16):
    """include/DataGen.hpp:26 -- same libc calls, bit-identical arrays."""
    out = np.empty(int(size), dtype=np.uint64)
    rc = lib.hj_generate_data(dist.encode(), int(size), int(distinct_keys if distinct_keys is not None else size),
                              int(local_shuffle_range), out.ctypes.data)
    if rc != 0:
        raise ValueError("Unknown distribution")  # DataGen.hpp:117-120
    return out


def mc_relation_pk(n, seed=12345):
    out = np.empty(int(n), dtype=np.uint64)
    lib.hj_mc_relation_pk(out.ctypes.data, int(n), int(seed))
    return out


def mc_relation_fk(m, maxid, seed=54321):
    out = np.empty(int(m), dtype=np.uint64)
    lib.hj_mc_relation_fk(out.ctypes.data, int(m), int(maxid), int(seed))
    return out


def mc_relation_nonunique(n, maxid, seed=12345, bias=0):
    """mc --non-unique / --full-range R (generator.c:493): keys in [0, maxid) + bias (bias=1: no key 0)."""
    out = np.empty(int(n), dtype=np.uint64)
    lib.hj_mc_relation_nonunique(out.ctypes.data, int(n), int(maxid), int(seed), int(bias))
    return out


def mc_relation_fk_from_pk(m, pk, seed=54321):
    """mc --full-range S (generator.c:448): copies of pk, Knuth-shuffled."""
    pk = np.ascontiguousarray(pk, dtype=np.uint64)
    out = np.empty(int(m), dtype=np.uint64)
    lib.hj_mc_relation_fk_from_pk(out.ctypes.data, int(m), pk.ctypes.data, pk.size, int(seed))
    return out


def mc_relation_pk_lshuffle(n, local_shuffle_range, seed=12345):
    out = np.empty(int(n), dtype=np.uint64)
    lib.hj_mc_relation_pk_lshuffle(out.ctypes.data, int(n), int(local_shuffle_range), int(seed))
    return out


def mc_relation_zipf(m, maxid, z, seed=54321):
    """mc --skew z (genzipf.c): zipf-distributed foreign key over a permuted alphabet 1..maxid."""
    out = np.empty(int(m), dtype=np.uint64)
    lib.hj_mc_relation_zipf(out.ctypes.data, int(m), int(maxid), float(z), int(seed))
    return out


# ---- handle --------------------------------------------------------------------
class HashJoin:
    """One hj_handle: owns the device buffers of one GPU."""

    def __init__(self, algo="atomic", device=0, probe_length=4, scale_output=2, num_partitions=0, passes=0,
                 transaction_size=16, enable_probe=True, sample_rounds=5, key_mask=0, adaptive_threshold=0.0,
                 pass1_bits=0, tuple_bytes=8):
        cfg = hj_config()
        lib.hj_config_default(C.byref(cfg))
        cfg.algo = ALGOS[algo] if isinstance(algo, str) else int(algo)
        cfg.device = device
        cfg.probe_length = probe_length
        cfg.scale_output = scale_output
        cfg.num_partitions = num_partitions
        cfg.passes = passes
        cfg.transaction_size = transaction_size
        cfg.enable_probe = 1 if enable_probe else 0
        cfg.sample_rounds = sample_rounds
        cfg.key_mask = key_mask if key_mask else 0xFFFFFFFFFFFFFFFF
        cfg.adaptive_threshold = adaptive_threshold
        cfg.pass1_bits = pass1_bits
        cfg.tuple_bytes = tuple_bytes
        self.words = 2 if tuple_bytes == 16 else 1   # 8-byte words per tuple
        self.cfg = cfg
        self._h = C.c_void_p()
        rc = lib.hj_create(C.byref(cfg), C.byref(self._h))
        if rc != HJ_OK:
            self._h = None
            raise HJError(rc, "hj_create", "(no usable CUDA device; there is no CPU fallback)")

    def _check(self, rc, where):
        if rc != HJ_OK:
            raise HJError(rc, where, lib.hj_last_error(self._h).decode())

    def close(self):
        if self._h:
            lib.hj_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def reserve(self, max_r, max_s=0):
        self._check(lib.hj_reserve(self._h, int(max_r), int(max_s)), "hj_reserve")

    def reserve_staging(self, max_r, max_s=0):
        self._check(lib.hj_reserve_staging(self._h, int(max_r), int(max_s)), "hj_reserve_staging")

    def alloc_count(self):
        return int(lib.hj_alloc_count(self._h))

    def build(self, R):
        p, n = _ptr(R)
        res = hj_result()
        self._check(lib.hj_build(self._h, p, n // self.words, C.byref(res)), "hj_build")
        return res

    def probe(self, S):
        p, m = _ptr(S)
        res = hj_result()
        self._check(lib.hj_probe(self._h, p, m // self.words, C.byref(res)), "hj_probe")
        return res

    def join(self, R, S=None):
        """R, S: uint64 arrays (tuple_bytes=16: 2 words per tuple, e.g. shape (n, 2)), tensors or (pointer, words)."""
        pr, n = _ptr(R)
        ps, m = _ptr(S)
        res = hj_result()
        self._check(lib.hj_join(self._h, pr, n // self.words, ps, m // self.words, C.byref(res)), "hj_join")
        return res

    def join_materialize(self, R, S, d_out=None, capacity=0):
        """hj_join_materialize: every matching pair {r tuple, s tuple} into the device buffer d_out
        (raw pointer, `capacity` pairs of 16 bytes).  d_out=None: count only.  Returns (result, n_pairs);
        raises HJError(HJ_EOVERFLOW) when the buffer is too small (e.n_pairs tells the size needed)."""
        pr, n = _ptr(R)
        ps, m = _ptr(S)
        res = hj_result()
        need = C.c_uint64(0)
        rc = lib.hj_join_materialize(self._h, pr, n, ps, m, int(d_out) if d_out else None, int(capacity),
                                     C.byref(need), C.byref(res))
        if rc != HJ_OK:
            e = HJError(rc, "hj_join_materialize", lib.hj_last_error(self._h).decode())
            e.n_pairs = need.value
            raise e
        return res, need.value

    def sample(self, R):
        p, n = _ptr(R)
        res = hj_result()
        self._check(lib.hj_sample(self._h, p, n, C.byref(res)), "hj_sample")
        return res

    def checksums(self, R=None):
        p, n = _ptr(R)
        res = hj_result()
        self._check(lib.hj_checksums(self._h, p, n // self.words, C.byref(res)), "hj_checksums")
        return res

    def radix_partition(self, d_in, n, shift, bits, d_out, d_offsets):
        self._check(lib.hj_radix_partition(self._h, int(d_in), int(n), shift, bits, int(d_out), int(d_offsets)),
                    "hj_radix_partition")

    def exchange_partition(self, d_in, n, n_dest, domain_bits, d_out, d_offsets):
        self._check(lib.hj_exchange_partition(self._h, int(d_in), int(n), n_dest, domain_bits, int(d_out),
                                              int(d_offsets)), "hj_exchange_partition")

    # ---- exchange fused with radix pass 1 (multi-GPU; see include/hj_b200.h) ----
    def radix_plan(self, n_build):
        """(bits, b1, b2, passes) the radix join would use for a build side of n_build tuples."""
        v = [C.c_uint32() for _ in range(4)]
        self._check(lib.hj_radix_plan(self._h, int(n_build), *[C.byref(x) for x in v]), "hj_radix_plan")
        return tuple(x.value for x in v)

    def ipc_export(self, dptr):
        buf = C.create_string_buffer(64)
        self._check(lib.hj_ipc_export(self._h, int(dptr), buf), "hj_ipc_export")
        return buf.raw

    def ipc_open(self, handle):
        p = C.c_void_p()
        self._check(lib.hj_ipc_open(self._h, C.create_string_buffer(handle, 64), C.byref(p)), "hj_ipc_open")
        return p.value

    def ipc_close(self, dptr):
        self._check(lib.hj_ipc_close(self._h, dptr), "hj_ipc_close")

    def peer_hist(self, d_in, n, n_dest, domain_bits, b1, d_counts):
        self._check(lib.hj_peer_hist(self._h, int(d_in), int(n), n_dest, domain_bits, b1, int(d_counts)),
                    "hj_peer_hist")

    def peer_scatter(self, d_in, n, n_dest, domain_bits, b1, peer_bufs, d_cursors):
        arr = (C.c_uint64 * n_dest)(*[int(p) for p in peer_bufs])
        self._check(lib.hj_peer_scatter(self._h, int(d_in), int(n), n_dest, domain_bits, b1, arr, int(d_cursors)),
                    "hj_peer_scatter")

    def peer_scatter_async(self, d_in, n, n_dest, domain_bits, b1, peer_bufs, d_cursors, max_ctas=0):
        """hj_peer_scatter_async: enqueued only (sync() waits), on at most max_ctas CTAs."""
        arr = (C.c_uint64 * n_dest)(*[int(p) for p in peer_bufs])
        self._check(lib.hj_peer_scatter_async(self._h, int(d_in), int(n), n_dest, domain_bits, b1, arr, int(d_cursors),
                                              int(max_ctas)), "hj_peer_scatter_async")

    def peer_scatter_regions(self, d_in, n, n_dest, domain_bits, b1, peer_bufs, d_cursors, rank, region_cap, trash_off):
        """hj_peer_scatter_regions (histogram-free exchange); returns True when a region overflowed."""
        arr = (C.c_uint64 * n_dest)(*[int(p) for p in peer_bufs])
        ovf = C.c_uint32(0)
        self._check(lib.hj_peer_scatter_regions(self._h, int(d_in), int(n), n_dest, domain_bits, b1, arr, int(d_cursors),
                                                int(rank), int(region_cap), int(trash_off), C.byref(ovf)),
                    "hj_peer_scatter_regions")
        return bool(ovf.value)

    def join_regions(self, d_R1, n, d_curR, cap_r, d_S1, m, d_curS, cap_s, n_src):
        res = hj_result()
        self._check(lib.hj_join_regions(self._h, int(d_R1), int(n), int(d_curR), int(cap_r), int(d_S1), int(m),
                                        int(d_curS), int(cap_s), int(n_src), C.byref(res)), "hj_join_regions")
        return res

    def join_pass1(self, d_R1, n, d_S1, m, d_offR1=None, d_offS1=None):
        res = hj_result()
        self._check(lib.hj_join_pass1(self._h, int(d_R1), int(n), int(d_offR1) if d_offR1 else None, int(d_S1), int(m),
                                      int(d_offS1) if d_offS1 else None, C.byref(res)), "hj_join_pass1")
        return res

    # ---- pipelined exchange over the copy engines (include/hj_b200.h, hj_xchg_*) ----
    def xchg_create(self, n_ranks, rank, n_local, m_local, s_chunks, domain_bits):
        """Returns (window device pointer, window bytes): export the window with ipc_export()."""
        w, nb = C.c_void_p(), C.c_uint64()
        self._check(lib.hj_xchg_create(self._h, int(n_ranks), int(rank), int(n_local), int(m_local), int(s_chunks),
                                       int(domain_bits), C.byref(w), C.byref(nb)), "hj_xchg_create")
        return w.value, nb.value

    def xchg_connect(self, peer_windows):
        arr = (C.c_uint64 * len(peer_windows))(*[int(p) if p else 0 for p in peer_windows])
        self._check(lib.hj_xchg_connect(self._h, arr), "hj_xchg_connect")

    def xchg_start(self, d_R, d_S, epoch):
        self._check(lib.hj_xchg_start(self._h, int(d_R), int(d_S), int(epoch)), "hj_xchg_start")

    def xchg_finish(self):
        res = hj_result()
        self._check(lib.hj_xchg_finish(self._h, C.byref(res)), "hj_xchg_finish")
        return res

    def generate(self, kind, first, count, domain, seed, d_out):
        self._check(lib.hj_generate(self._h, kind, int(first), int(count), int(domain), int(seed), int(d_out)),
                    "hj_generate")

    def device_alloc(self, nbytes):
        p = C.c_void_p()
        self._check(lib.hj_device_alloc(self._h, int(nbytes), C.byref(p)), "hj_device_alloc")
        return p.value

    def device_free(self, p):
        self._check(lib.hj_device_free(self._h, p), "hj_device_free")

    def copy(self, dst, src, nbytes):
        self._check(lib.hj_copy(self._h, int(dst), int(src), int(nbytes)), "hj_copy")

    def to_device(self, arr):
        """Copy a host uint64 array to a fresh device buffer; returns (ptr, count)."""
        p = self.device_alloc(max(arr.nbytes, 8))
        if arr.size:
            self.copy(p, arr.ctypes.data, arr.nbytes)
        return (p, arr.size)

    def to_host(self, dptr, count):
        out = np.empty(int(count), dtype=np.uint64)
        if count:
            self.copy(out.ctypes.data, dptr, out.nbytes)
        return out

    def flush_l2(self):
        self._check(lib.hj_flush_l2(self._h), "hj_flush_l2")

    def sync(self):
        self._check(lib.hj_sync(self._h), "hj_sync")


NOALLOC_MAX_R = 1 << 16  # HJ_NOALLOC_MAX_R (include/hj_b200.h)


def join_noalloc(d_R, n, d_S, m, d_out, probe_length=4, device=0, stream=0):
    """hj_join_noalloc: the global-table join on the library's static device scratch -- no handle, no allocation,
    no host synchronisation.  d_R, d_S, d_out: raw device pointers; d_out receives {matches, checksum} once
    `stream` (a cudaStream_t value, 0 = default stream) reaches that point.  Raises HJError (HJ_EINVAL) when
    n > NOALLOC_MAX_R."""
    err = C.create_string_buffer(256)
    rc = lib.hj_join_noalloc(int(device), int(stream) or None, int(d_R), int(n), int(d_S) if d_S else None, int(m),
                             int(probe_length), int(d_out), err, 256)
    if rc != HJ_OK:
        raise HJError(rc, "hj_join_noalloc", err.value.decode())


def pinned_empty(n):
    """uint64 numpy array over pinned host memory (hj_host_alloc)."""
    p = C.c_void_p()
    rc = lib.hj_host_alloc(int(n) * 8, C.byref(p))
    if rc != HJ_OK:
        raise HJError(rc, "hj_host_alloc")
    buf = (C.c_uint64 * int(n)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=np.uint64)
    _PINNED[arr.ctypes.data] = p.value
    return arr


_PINNED = {}


def pinned_free(arr):
    p = _PINNED.pop(arr.ctypes.data, None)
    if p is not None:
        lib.hj_host_free(p)


# ---- the reference's entry points ----------------------------------------------
def _ref_run(algo, relR, rSize, relS, sSize, transactionSize, scaleOutput, numPartitions, probeLength, passes=0):
    pr, n = _ptr(relR)
    ps, m = _ptr(relS)
    if rSize is not None:
        n = int(rSize)
    if relS is not None and sSize is not None:
        m = int(sSize)
    buf = C.create_string_buffer(4096)
    rc = lib.hj_ref_run(algo.encode(), pr, n, ps, m, int(transactionSize), int(scaleOutput), int(numPartitions),
                        int(probeLength), int(passes), buf, 4096)
    line = buf.value.decode()
    out = json.loads(line)
    if rc != HJ_OK:
        raise HJError(rc, "hj_ref_run", out.get("error", ""))
    out["_json"] = line
    return out


def AtomicHashBuild(relR, rSize=None, relS=None, sSize=None, scaleOutput=2, numPartitions=64, probeLength=4):
    return _ref_run("atomic", relR, rSize, relS, sSize, 0, scaleOutput, numPartitions, probeLength)


def NoCCHashBuild(relR, rSize=None, relS=None, sSize=None, scaleOutput=2, numPartitions=64, probeLength=4):
    return _ref_run("nocc", relR, rSize, relS, sSize, 0, scaleOutput, numPartitions, probeLength)


def HTMHashBuild(relR, rSize=None, relS=None, sSize=None, transactionSize=16, scaleOutput=2, numPartitions=64,
                 probeLength=4):
    return _ref_run("htm", relR, rSize, relS, sSize, transactionSize, scaleOutput, numPartitions, probeLength)


def SortMerge(build, buildSize=None, probe=None, probeSize=None, numPartitions=64):
    """SortMerge.hpp:5-6 -- the sort-merge comparator (partitioned sort-merge join on the GPU)."""
    return _ref_run("sortmerge", build, buildSize, probe, probeSize, 0, 2, numPartitions, 4)


def PRO(relR, relS=None, nthreads=1, radix_bits=14, passes=2):
    """mc PRO (mc/src/parallel_radix_join.h:33-34) through the reference-named C++ entry point
    (csrc/ref_entry.cpp: PRO): the join cardinality, or -- relS None or a dummy of at most 2 tuples, as
    experiments/motivation.sh:11 passes -- the reference's build-only "Results" checksum
    (parallel_radix_join.c:231-283).  NUM_RADIX_BITS / NUM_PASSES are hj_ref_set_radix()."""
    pr, n = _ptr(relR)
    ps, m = _ptr(relS)
    lib.hj_ref_set_radix(int(radix_bits), int(passes))
    try:
        out = lib.hj_ref_pro(pr, n, ps, m, int(nthreads))
    finally:
        lib.hj_ref_set_radix(0, 0)
    if out < 0 and lib.hj_ref_last_status() != HJ_OK:
        raise HJError(lib.hj_ref_last_status(), "PRO")
    return int(out)
