"""GPU parity tests of the sort-merge comparator (reference SortMerge.cpp:8-39): the partitioned
sort-merge join through the C ABI (HJ_ALGO_SORTMERGE) vs the oracle's true join, bit-exact (match count +
order-independent 64-bit checksum), on the reference's input distributions, duplicates on both sides, payloads,
ragged sizes, skew, wide keys, R partitions larger than one shared-memory sort, and against the radix hash join."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DISTS = [("sorted", 16), ("shuffle", 16), ("local_shuffle", 4096), ("uniform", 16), ("random", 16)]


@pytest.mark.parametrize("dist,rng", DISTS)
def test_sortmerge_matches_oracle(hj, oracle, dist, rng):
    n = 1 << 18
    R = hj.generate_data(dist, n, n, rng)
    S = R.copy() if dist == "random" else hj.generate_data("sorted", n, n, 16)  # main.cpp:92-97
    want = oracle.true_join(R, S)
    with hj.HashJoin("sortmerge") as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        assert res.path_taken == hj.HJ_PATH_SORTMERGE and res.kernel_launches > 0


@pytest.mark.parametrize("bits,passes", [(3, 1), (6, 2), (9, 1), (0, 0)])
def test_sortmerge_payloads_duplicates_and_chunked_partitions(hj, oracle, bits, passes):
    """mc tuples (key low 32 bits, payload high 32), duplicates on both sides, odd sizes.  2^3 partitions of
    12500 R tuples exceed one 8192-element sort: R_p is merged chunk by chunk."""
    rng = np.random.default_rng(5)
    n, m = 100003, 400007
    R = rng.integers(1, 30000, n, dtype=np.uint64) | (np.arange(n, dtype=np.uint64) << np.uint64(32))
    S = rng.integers(1, 40000, m, dtype=np.uint64) | (np.arange(m, dtype=np.uint64) << np.uint64(32))
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    with hj.HashJoin("sortmerge", num_partitions=(1 << bits) if bits else 0, passes=passes, key_mask=0xFFFFFFFF) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        if bits:
            assert res.radix_bits == bits
        if bits == 3:
            assert res.max_partition > 8192


@pytest.mark.parametrize("bits,passes", [(4, 1), (10, 2), (15, 2)])
def test_sortmerge_fk_ragged(hj, oracle, bits, passes):
    n, m = 1 << 17, (5 << 17) + 77
    R, S = hj.mc_relation_pk(n), hj.mc_relation_fk(m, n)
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    assert want[0] == m
    with hj.HashJoin("sortmerge", num_partitions=1 << bits, passes=passes, key_mask=0xFFFFFFFF) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


def test_sortmerge_tiny_and_empty_partitions(hj, oracle):
    """About one tuple per partition: many partitions are empty on one side or both."""
    rng = np.random.default_rng(3)
    R = rng.integers(1, 400, 301, dtype=np.uint64)
    S = rng.integers(1, 400, 517, dtype=np.uint64)
    want = oracle.true_join(R, S)
    assert want[0] > 0
    with hj.HashJoin("sortmerge", num_partitions=1 << 8, passes=1) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


def test_sortmerge_wide_keys(hj, oracle):
    """64-bit keys (key_mask = all bits): the order is on the whole word."""
    rng = np.random.default_rng(9)
    n, m = 50021, 150001
    hi = np.uint64(1) << np.uint64(40)
    R = rng.integers(1, 20000, n, dtype=np.uint64) | (rng.integers(0, 2, n, dtype=np.uint64) * hi)
    S = rng.integers(1, 20000, m, dtype=np.uint64) | (rng.integers(0, 2, m, dtype=np.uint64) * hi)
    want = oracle.true_join(R, S)
    with hj.HashJoin("sortmerge") as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


def test_sortmerge_zipf_skew(hj, oracle):
    """mc --skew: heavy hitters make some S partitions many tiles long and their equal runs long."""
    n, m, z = 1 << 16, 1 << 19, 1.0
    rng = np.random.default_rng(11)
    alphabet = rng.permutation(n).astype(np.uint64) + np.uint64(1)
    cdf = np.cumsum(1.0 / np.arange(1, n + 1, dtype=np.float64) ** z)
    cdf /= cdf[-1]
    S = alphabet[np.minimum(np.searchsorted(cdf, rng.random(m)), n - 1)] | (np.arange(m, dtype=np.uint64) << np.uint64(32))
    R = hj.mc_relation_pk(n)
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    assert want[0] == m
    with hj.HashJoin("sortmerge", key_mask=0xFFFFFFFF) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


def test_sortmerge_equals_radix_join_at_size(hj):
    """2^22 x 2^24 FK join: closed-form cardinality, and the same checksum as the radix hash join."""
    n, m = 1 << 22, 1 << 24
    R, S = hj.mc_relation_pk(n), hj.mc_relation_fk(m, n)
    with hj.HashJoin("radix", key_mask=0xFFFFFFFF) as h:
        ref = h.join(R, S)
    with hj.HashJoin("sortmerge", key_mask=0xFFFFFFFF) as h:
        res = h.join(R, S)
    assert res.matches == ref.matches == m
    assert res.checksum == ref.checksum


@pytest.mark.parametrize("key_mask", [0, 0xFFFFFFFF])
def test_sortmerge_generic_order_equals_encoded(hj, oracle, monkeypatch, key_mask):
    """The register sort orders encoded words for the two usual key masks; HJ_SM_MODE=0 forces the generic
    (key, tuple) comparison that any other mask takes.  Same result either way."""
    rng = np.random.default_rng(21)
    n, m = 70001, 300017
    R = rng.integers(1, 50000, n, dtype=np.uint64) | (np.arange(n, dtype=np.uint64) << np.uint64(32))
    S = rng.integers(1, 50000, m, dtype=np.uint64) | (np.arange(m, dtype=np.uint64) << np.uint64(32))
    want = oracle.true_join(R, S, key_mask) if key_mask else oracle.true_join(R, S)
    for mode in ("1", "0"):
        monkeypatch.setenv("HJ_SM_MODE", mode)
        with hj.HashJoin("sortmerge", key_mask=key_mask) as h:
            res = h.join(R, S)
            assert (res.matches, res.checksum) == want, mode


def test_sortmerge_needs_a_probe_side(hj):
    R = hj.generate_data("sorted", 1 << 12)
    with hj.HashJoin("sortmerge") as h:
        with pytest.raises(hj.HJError) as e:
            h.join(R, None)
        assert e.value.code == hj.HJ_EINVAL


def test_sortmerge_entry_point_and_cli(hj):
    n = 1 << 20
    R, S = hj.generate_data("shuffle", n), hj.generate_data("sorted", n)
    j = hj.SortMerge(R, n, S, n, 64)
    assert list(j.keys())[:7] == ["algo", "rSize", "hashBuildTimeInMicroseconds", "sortTimeInMicroseconds",
                                  "mergeTimeInMicroseconds", "totalMatches", "inputSum"]  # SortMerge.cpp:49-68
    assert (j["totalMatches"], j["inputSum"]) == (n, n * (n + 1) // 2)
    exe = os.path.join(ROOT, "htm_hashjoin_b200", "csrc", "main")
    out = subprocess.run([exe, "--algo", "sortmerge", "--rSize", "1M", "--dataDistr", "shuffle"], capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["totalMatches"] == n and line["path"] == "sortmerge"


def test_sortmerge_equals_the_reference_binary(hj):
    """The reference's SortMerge.cpp itself (oracle/_ref, compiled unmodified) on the same arrays: same totalMatches
    and inputSum as the GPU entry point, for the unique-key inputs on which the reference's merge is a join."""
    from tests import oracle_lib
    ref = oracle_lib.Reference(probe=True)
    n = 1 << 18
    S = hj.generate_data("sorted", n)
    if not ref.available or ref.sort_merge(S[:64], S[:64]) is None:
        pytest.skip("oracle/_ref without SortMerge")
    for dist, rng in (("shuffle", 16), ("local_shuffle", 256), ("sorted", 16)):
        R = hj.generate_data(dist, n, n, rng)
        want = ref.sort_merge(R, S)
        got = hj.SortMerge(R, n, S, n, 64)
        assert (got["totalMatches"], got["inputSum"], got["rSize"]) == (want["totalMatches"], want["inputSum"],
                                                                        want["rSize"]), dist
