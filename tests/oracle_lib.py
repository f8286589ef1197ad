"""ctypes loader for oracle/liboracle.so and oracle/_ref (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product package never does.
"""
import ctypes as C
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


class orc_gt_result(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("conflicts", "total_matches", "input_sum", "table_sum", "overflow_sum", "ref_output_sum")]


def _p(a):
    return a.ctypes.data if a is not None else None


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(ORACLE_SO)
        L = self.lib
        vp, u64 = C.c_void_p, C.c_uint64
        L.orc_generate_data.restype = C.c_int
        L.orc_generate_data.argtypes = [C.c_char_p, u64, u64, C.c_int, vp]
        L.orc_mc_pk.argtypes = [vp, u64, C.c_uint]
        L.orc_mc_fk.argtypes = [vp, u64, u64, C.c_uint]
        L.orc_gt_build_probe_serial.restype = C.c_int
        L.orc_gt_build_probe_serial.argtypes = [vp, u64, vp, u64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                                C.POINTER(orc_gt_result)]
        L.orc_true_join.argtypes = [vp, u64, vp, u64, u64, C.POINTER(u64), C.POINTER(u64)]
        if hasattr(L, "orc_true_join16"):
            L.orc_true_join16.argtypes = [vp, u64, vp, u64, u64, C.POINTER(u64), C.POINTER(u64)]
        L.orc_radix_partition.argtypes = [vp, u64, C.c_uint32, C.c_uint32, vp, vp]
        L.orc_pro_checksum.restype = u64
        L.orc_pro_checksum.argtypes = [vp, u64, C.c_uint32]
        L.orc_radix_join.restype = u64
        L.orc_radix_join.argtypes = [vp, u64, vp, u64, C.c_uint32, C.c_uint32, C.POINTER(u64)]
        L.orc_htm_build_serial.restype = u64
        L.orc_htm_build_serial.argtypes = [vp, u64, C.POINTER(u64), C.POINTER(u64)]
        L.orc_sum.restype = u64
        L.orc_sum.argtypes = [vp, u64]
        L.orc_possum.restype = u64
        L.orc_possum.argtypes = [vp, u64]
        if hasattr(L, "orc_feistel_gen"):
            L.orc_feistel_gen.argtypes = [C.c_int, u64, u64, u64, u64, vp]

    def generate_data(self, dist, n, distinct=None, rng=16):
        out = np.empty(n, dtype=np.uint64)
        rc = self.lib.orc_generate_data(dist.encode(), n, distinct if distinct is not None else n, rng, _p(out))
        if rc != 0:
            raise ValueError("Unknown distribution")
        return out

    def mc_pk(self, n, seed=12345):
        out = np.empty(n, dtype=np.uint64)
        self.lib.orc_mc_pk(_p(out), n, seed)
        return out

    def mc_fk(self, m, maxid, seed=54321):
        out = np.empty(m, dtype=np.uint64)
        self.lib.orc_mc_fk(_p(out), m, maxid, seed)
        return out

    def gt_serial(self, R, S=None, probe_length=4, nocc=False):
        res = orc_gt_result()
        rc = self.lib.orc_gt_build_probe_serial(_p(R), R.size, _p(S), 0 if S is None else S.size, 2, 64,
                                                probe_length, 1 if nocc else 0, C.byref(res))
        assert rc == 0
        return {n: getattr(res, n) for n, _ in res._fields_}

    def true_join(self, R, S, key_mask=0xFFFFFFFFFFFFFFFF):
        m, c = C.c_uint64(), C.c_uint64()
        self.lib.orc_true_join(_p(R), R.size, _p(S), S.size, key_mask, C.byref(m), C.byref(c))
        return m.value, c.value

    def true_join_pairs(self, R, S, key_bits=32):
        """Every matching pair (r tuple, s tuple) of the true equi-join on the low key_bits, as an
        (n_pairs, 2) uint64 array sorted lexicographically -- the checker of hj_join_materialize (the
        reference never materialises: no_partitioning_join.c:301).  numpy; count and checksum are
        cross-checked against orc_true_join by the caller."""
        mask = np.uint64((1 << key_bits) - 1)
        R = R[R != 0]
        S = S[S != 0]
        Rs = R[np.argsort(R & mask, kind="stable")]
        rk = Rs & mask
        lo = np.searchsorted(rk, S & mask, side="left")
        hi = np.searchsorted(rk, S & mask, side="right")
        cnt = (hi - lo).astype(np.int64)
        s_rep = np.repeat(S, cnt)
        start = np.repeat(lo, cnt)
        within = np.arange(int(cnt.sum()), dtype=np.int64) - np.repeat(np.cumsum(cnt) - cnt, cnt)
        pairs = np.stack([Rs[start + within], s_rep], axis=1) if cnt.sum() else np.empty((0, 2), np.uint64)
        return pairs[np.lexsort((pairs[:, 1], pairs[:, 0]))]

    def radix_partition(self, a, shift, bits):
        out = np.empty_like(a)
        off = np.empty((1 << bits) + 1, dtype=np.uint64)
        self.lib.orc_radix_partition(_p(a), a.size, shift, bits, _p(out), _p(off))
        return out, off

    def pro_checksum(self, R, bits):
        return self.lib.orc_pro_checksum(_p(R), R.size, bits)

    def radix_join(self, R, S, bits, passes):
        c = C.c_uint64()
        m = self.lib.orc_radix_join(_p(R), R.size, _p(S), S.size, bits, passes, C.byref(c))
        return m, c.value

    def htm_serial(self, R):
        ts, cs = C.c_uint64(), C.c_uint64()
        conflicts = self.lib.orc_htm_build_serial(_p(R), R.size, C.byref(ts), C.byref(cs))
        return conflicts, ts.value, cs.value

    def sum(self, a):
        return self.lib.orc_sum(_p(a), a.size)

    def possum(self, a):
        return self.lib.orc_possum(_p(a), a.size)

    def feistel_gen(self, kind, first, count, domain, seed):
        out = np.empty(count, dtype=np.uint64)
        self.lib.orc_feistel_gen(kind, first, count, domain, seed, _p(out))
        return out


def _true_join16(self, R, S, key_mask=0xFFFFFFFFFFFFFFFF):
    """KEY_8B tuples: R, S uint64 arrays of shape (n, 2) = {key, payload}."""
    R = np.ascontiguousarray(R, dtype=np.uint64).reshape(-1, 2)
    S = np.ascontiguousarray(S, dtype=np.uint64).reshape(-1, 2)
    mt, ck = C.c_uint64(0), C.c_uint64(0)
    self.lib.orc_true_join16(_p(R), R.shape[0], _p(S), S.shape[0], key_mask, C.byref(mt), C.byref(ck))
    return mt.value, ck.value


Oracle.true_join16 = _true_join16


class Reference:
    """The reference's own code compiled unmodified into oracle/_ref (may be absent)."""

    def __init__(self, probe=True):
        path = os.path.join(REF_DIR, "libref_probe.so" if probe else "libref_build.so")
        self.available = os.path.exists(path)
        if not self.available:
            return
        self.lib = C.CDLL(path)
        self.lib.ref_generate_data.restype = C.c_int
        self.lib.ref_generate_data.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]
        self.lib.ref_hash_build.restype = C.c_int
        self.lib.ref_hash_build.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32,
                                            C.c_uint32, C.c_uint32, C.c_char_p, C.c_uint32]

    def sort_merge(self, build, probe, num_partitions=64):
        """The reference's SortMerge (SortMerge.cpp:8-39) on copies of the inputs; None if this build lacks it."""
        if not hasattr(self.lib, "ref_sort_merge"):
            return None
        self.lib.ref_sort_merge.restype = C.c_int
        self.lib.ref_sort_merge.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_char_p,
                                            C.c_uint32]
        b, p = build.copy(), probe.copy()
        buf = C.create_string_buffer(4096)
        rc = self.lib.ref_sort_merge(_p(b), b.size, _p(p), p.size, num_partitions, buf, 4096)
        assert rc >= 0
        return json.loads(buf.value.decode())

    def generate_data(self, dist, n, distinct=None, rng=16):
        out = np.empty(n, dtype=np.uint64)
        rc = self.lib.ref_generate_data(dist.encode(), n, distinct if distinct is not None else n, rng, _p(out))
        if rc != 0:
            raise ValueError("Unknown distribution")
        return out

    def hash_build(self, algo, R, S=None, num_partitions=64, probe_length=4):
        buf = C.create_string_buffer(4096)
        rc = self.lib.ref_hash_build(algo.encode(), _p(R), R.size, _p(S), 0 if S is None else S.size, 2,
                                     num_partitions, probe_length, buf, 4096)
        assert rc >= 0
        return json.loads(buf.value.decode())


def mc_binary(bits, passes):
    p = os.path.join(REF_DIR, f"mc_b{bits}_p{passes}")
    return p if os.path.exists(p) else None
