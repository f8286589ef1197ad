"""hj_join_noalloc (include/hj_b200.h): the global-table join on the library's static device scratch.

Collected last on purpose.  The entry point has no counterpart in the reference; it
is checked against the existing entry point for the same work -- hj_join on an HJ_ALGO_ATOMIC handle -- bit for bit
(integer work: no tolerance), on two calls in a row that share the scratch array."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(hj, which):
    rng = np.random.default_rng(100 + which)
    if which == 0:
        # the largest build the scratch holds, heavy duplication: windows fill up, the overflow table is in use
        n = hj.NOALLOC_MAX_R
        R = hj.generate_data("uniform", n, n // 4, 16)
        S = np.concatenate([hj.generate_data("sorted", n // 2), R[: n // 8]])
    elif which == 1:
        # a smaller table than the call before (stale slots beyond it and a stale overflow table stay behind),
        # unique shuffled keys: every tuple in its home slot, nothing overflows
        n = 5000
        R = rng.permutation(np.arange(1, n + 1, dtype=np.uint64))
        S = rng.integers(1, 2 * n, 40000, dtype=np.uint64)
    else:
        # keys that collide in the table (same low bits), payload-free 64-bit keys
        n = 3000
        R = (rng.integers(0, 64, n, dtype=np.uint64) << np.uint64(20)) + rng.integers(1, 50, n, dtype=np.uint64)
        S = np.concatenate([R[::3], rng.integers(1, 1 << 26, 10000, dtype=np.uint64)])
    return np.ascontiguousarray(R), np.ascontiguousarray(S)


def test_join_noalloc_matches_hj_join_twice_in_a_row(hj):
    import torch
    with hj.HashJoin("atomic") as h:                     # the existing entry point; also lends its copy helpers
        d_out = h.device_alloc(16)
        cases = []
        for which in (0, 1, 2):
            R, S = _inputs(hj, which)
            cases.append((R, S, h.to_device(R)[0], h.to_device(S)[0]))
        try:
            for rep in range(2):                         # the whole sequence twice: 6 calls on the shared scratch
                for R, S, dR, dS in cases:
                    want = h.join(R, S)
                    h.copy(d_out, np.full(2, 0xDEADBEEF, dtype=np.uint64).ctypes.data, 16)
                    torch.cuda.synchronize()
                    free_before = torch.cuda.mem_get_info(0)[0]
                    hj.join_noalloc(dR, R.size, dS, S.size, d_out)
                    torch.cuda.synchronize()             # the call only enqueues (default stream); hj_copy runs on
                    got = h.to_host(d_out, 2)            # the handle's own non-blocking stream
                    assert (int(got[0]), int(got[1])) == (want.matches, want.checksum), (rep, R.size)
                    assert want.matches > 0
                    if rep:                              # (the very first call may load the kernels' module)
                        assert torch.cuda.mem_get_info(0)[0] == free_before
        finally:
            for _, _, dR, dS in cases:
                h.device_free(dR)
                h.device_free(dS)
            h.device_free(d_out)


def test_join_noalloc_probe_lengths_and_empty_s(hj, oracle):
    import torch
    R, S = _inputs(hj, 0)
    with hj.HashJoin("atomic") as h:
        dR, dS, d_out = h.to_device(R)[0], h.to_device(S)[0], h.device_alloc(16)
        try:
            for L in (1, 4, 9):
                hj.join_noalloc(dR, R.size, dS, S.size, d_out, probe_length=L)
                torch.cuda.synchronize()
                got = h.to_host(d_out, 2)
                assert (int(got[0]), int(got[1])) == oracle.true_join(R, S)
            hj.join_noalloc(dR, 100, 0, 0, d_out)        # |S| = 0: a build only
            torch.cuda.synchronize()
            assert h.to_host(d_out, 2).tolist() == [0, 0]
        finally:
            for p in (dR, dS, d_out):
                h.device_free(p)


def test_join_noalloc_rejects_sizes_beyond_the_scratch(hj):
    with hj.HashJoin("atomic") as h:
        d = h.device_alloc(16)
        try:
            with pytest.raises(hj.HJError) as e:
                hj.join_noalloc(d, hj.NOALLOC_MAX_R + 1, d, 1, d)
            assert "HJ_NOALLOC_MAX_R" in str(e.value)
            with pytest.raises(hj.HJError):
                hj.join_noalloc(d, 0, d, 1, d)           # |R| = 0
            host = np.zeros(2, dtype=np.uint64)
            with pytest.raises(hj.HJError):
                hj.join_noalloc(host.ctypes.data, 2, d, 1, d)   # a host pointer
        finally:
            h.device_free(d)
