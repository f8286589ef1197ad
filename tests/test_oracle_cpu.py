"""CPU tests: the oracle against the reference's golden vectors / KATs,
against the reference itself compiled into oracle/_ref, and the host-side mirror
(csrc/datagen.cpp) against both.  No GPU needed."""
import json
import os
import subprocess

import numpy as np
import pytest

from tests import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "kats.json")


@pytest.fixture(scope="module")
def kats():
    with open(GOLDEN) as f:
        return json.load(f)


# ---- generator KATs (probed from the reference's DataGen.hpp with glibc rand) ----
@pytest.mark.parametrize("name", ["sorted_2^20", "shuffle_2^20", "local_shuffle_2^20_w16", "local_shuffle_2^20_w4096",
                                  "uniform_2^20"])
def test_generator_kats(oracle, hj, kats, name):
    k = kats["generators"][name]
    for gen in (oracle.generate_data, hj.generate_data):
        a = gen(k["dist"], k["n"], k["n"], k["range"])
        assert oracle.possum(a) == k["possum"]
        if "sum" in k:
            assert oracle.sum(a) == k["sum"]
        if "first8" in k:
            assert a[:8].tolist() == k["first8"]


def test_generator_matches_reference_binary(oracle, hj):
    ref = oracle_lib.Reference(probe=True)
    if not ref.available:
        pytest.skip("oracle/_ref not built (no /root/reference)")
    for dist, rng in (("uniform", 16), ("shuffle", 16), ("local_shuffle", 64), ("sorted", 16), ("random", 16)):
        n = 1 << 16
        want = ref.generate_data(dist, n, n, rng)
        assert np.array_equal(oracle.generate_data(dist, n, n, rng), want), dist
        assert np.array_equal(hj.generate_data(dist, n, n, rng), want), dist


def test_unknown_distribution(oracle, hj):
    with pytest.raises(ValueError):
        oracle.generate_data("zipf", 16)
    with pytest.raises(ValueError):
        hj.generate_data("zipf", 16)


def test_mc_generators_host_vs_oracle(oracle, hj):
    assert np.array_equal(hj.mc_relation_pk(5000, 12345), oracle.mc_pk(5000, 12345))
    fk_h, fk_o = hj.mc_relation_fk(23000, 5000, 54321), oracle.mc_fk(23000, 5000, 54321)
    assert np.array_equal(fk_h, fk_o)
    # 4 full permutations of 1..5000 and one of 1..3000
    for i in range(4):
        assert np.array_equal(np.sort(fk_o[i * 5000:(i + 1) * 5000]), np.arange(1, 5001, dtype=np.uint64))
    assert np.array_equal(np.sort(fk_o[20000:]), np.arange(1, 3001, dtype=np.uint64))


# ---- closed forms pinned by the reference's logs (scaled down) --------------------
@pytest.mark.parametrize("n", [1 << 10, 1 << 16])
def test_global_table_unique_keys_closed_form(oracle, n):
    R = oracle.generate_data("shuffle", n)
    S = oracle.generate_data("sorted", n)
    r = oracle.gt_serial(R, S)
    assert r["input_sum"] == n * (n + 1) // 2
    assert r["conflicts"] == 0 and r["total_matches"] == n
    assert r["table_sum"] + r["overflow_sum"] == r["input_sum"]
    # NoCC prints inputSum - n (NoCCHashBuild.hpp:94 sums half the table), as every nocc log line shows
    assert oracle.gt_serial(R, S, nocc=True)["ref_output_sum"] == r["input_sum"] - n
    assert oracle.true_join(R, S) == (n, 2 * r["input_sum"])


def test_logged_golden_values_are_the_closed_forms(kats):
    g = kats["reference_logs_2^27"]
    n = 1 << 27
    assert g["inputSum"] == n * (n + 1) // 2
    assert g["totalMatches"] == n
    assert g["nocc_outputSum"] == g["inputSum"] - n
    assert g["PRO_Results_b14"] == (n // 2) * (n // (1 << 14) - 1)


def test_serial_reference_numbers_uniform(oracle, kats):
    k = kats["ref_serial_uniform_2^20"]
    R = oracle.generate_data("uniform", 1 << 20)
    S = oracle.generate_data("sorted", 1 << 20)
    r = oracle.gt_serial(R, S)
    assert r["conflicts"] == k["conflicts"]
    assert r["total_matches"] == k["totalMatches"]
    assert r["table_sum"] + r["overflow_sum"] == r["input_sum"]
    assert oracle.true_join(R, S)[0] == 1 << 20


def test_oracle_vs_reference_library(oracle):
    """The reference's own AtomicHashBuild/NoCCHashBuild, run with one partition so its
    schedule is the serial one the oracle restates."""
    ref = oracle_lib.Reference(probe=True)
    if not ref.available:
        pytest.skip("oracle/_ref not built")
    n = 1 << 16
    S = oracle.generate_data("sorted", n)
    for dist in ("uniform", "shuffle", "local_shuffle"):
        R = oracle.generate_data(dist, n)
        o = oracle.gt_serial(R, S)
        j = ref.hash_build("atomic", R.copy(), S.copy(), num_partitions=1)
        assert j["conflicts"] == o["conflicts"], dist
        assert j["totalMatches"] == o["total_matches"], dist
        assert j["inputSum"] == o["input_sum"], dist
        assert j["outputSum"] == o["ref_output_sum"], dist


@pytest.mark.parametrize("bits", [7, 10, 14])
def test_pro_checksum_closed_form(oracle, bits):
    n = 1 << 20
    R = oracle.mc_pk(n)
    want = (n // 2) * (n // (1 << bits) - 1)
    assert oracle.pro_checksum(R, bits) == want
    if bits == 14:
        assert want == 33030144


def test_pro_checksum_vs_reference_binary(oracle):
    exe = oracle_lib.mc_binary(10, 1)
    if not exe:
        pytest.skip("oracle/_ref mc binaries not built")
    n = 1 << 20
    out = subprocess.run([exe, "--algo=PRO", "--nthreads=2", f"--r-size={n}", "--s-size=2"], capture_output=True,
                         text=True, timeout=120).stdout
    got = [l for l in out.splitlines() if "Results" in l][0]
    val = int(got.split("=")[1].split(".")[0])
    assert val == oracle.pro_checksum(oracle.mc_pk(n), 10)


def test_radix_join_equals_true_join(oracle):
    n, m = 1 << 14, 5 << 14
    R = oracle.mc_pk(n)
    S = oracle.mc_fk(m, n)
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    assert want[0] == m
    for bits, passes in ((4, 1), (8, 2), (9, 2)):
        assert oracle.radix_join(R, S, bits, passes) == want
    Ru = oracle.generate_data("uniform", n)
    Ss = oracle.generate_data("sorted", n)
    assert oracle.radix_join(Ru, Ss, 6, 2) == oracle.true_join(Ru, Ss)


def test_radix_partition_properties(oracle):
    a = oracle.generate_data("shuffle", 1 << 12)
    out, off = oracle.radix_partition(a, 3, 5)
    assert off[0] == 0 and off[-1] == a.size
    for p in range(32):
        seg = out[off[p]:off[p + 1]]
        assert np.all(((seg >> np.uint64(3)) & np.uint64(31)) == p)
    assert np.array_equal(np.sort(out), np.sort(a))


def test_feistel_generator_is_a_permutation(oracle):
    for k in (1, 2, 7, 10, 13):
        d = 1 << k
        a = oracle.feistel_gen(1, 0, d, d, 42)
        assert np.array_equal(np.sort(a), np.arange(1, d + 1, dtype=np.uint64)), k
    # consecutive blocks are different permutations; slices agree with the whole
    a = oracle.feistel_gen(1, 0, 3 << 10, 1 << 10, 7)
    assert not np.array_equal(a[:1024], a[1024:2048])
    assert np.array_equal(oracle.feistel_gen(1, 1500, 700, 1 << 10, 7), a[1500:2200])
    assert np.array_equal(oracle.feistel_gen(0, 10, 5, 0, 0), np.arange(11, 16, dtype=np.uint64))


def test_htm_serial_no_conflicts_on_dense_keys(oracle):
    R = oracle.generate_data("shuffle", 1 << 14)
    conflicts, ts, cs = oracle.htm_serial(R)
    assert conflicts == 0 and cs == 0 and ts == oracle.sum(R)


def _ref_sortmerge_restated(build, probe, slices=64):
    """SortMerge.cpp:13-35 restated: sort the build side, then a two-pointer merge of build[a:b) with probe[a:b)
    for each of the 64 equal index slices (blocked_range grain = n/64, simple partitioner), both pointers advancing
    on equality -- the true join only when both sides hold the same unique keys."""
    b = np.sort(build)
    n, total = b.size, 0
    for k in range(slices):
        lo, hi = k * n // slices, (k + 1) * n // slices
        li = ri = lo
        bl, pl = b[lo:hi].tolist(), probe[lo:hi].tolist()
        li = ri = 0
        m = hi - lo
        while li < m and ri < m:
            left, right = bl[li], pl[ri]
            total += left == right
            li += left <= right
            ri += left >= right
    return total


def test_reference_sortmerge_pins_the_comparator(oracle):
    """The reference's own SortMerge.cpp compiled into oracle/_ref.  Unique dense keys against S = sorted 1..n
    (main.cpp:93, the only probe side the reference runs it on): totalMatches = n = the true join, which is what
    the GPU comparator returns (tests/test_sortmerge_gpu.py).  With duplicate keys ("uniform") its slice-wise
    lock-step merge is not a join: the restatement above reproduces its count, the true join is larger."""
    ref = oracle_lib.Reference(probe=True)
    if not ref.available or ref.sort_merge(np.arange(1, 65, dtype=np.uint64), np.arange(1, 65, dtype=np.uint64)) is None:
        pytest.skip("oracle/_ref without SortMerge")
    n = 1 << 16
    S = oracle.generate_data("sorted", n)
    for dist in ("sorted", "shuffle", "local_shuffle"):
        R = oracle.generate_data(dist, n)
        j = ref.sort_merge(R, S)
        assert list(j.keys()) == ["algo", "rSize", "hashBuildTimeInMicroseconds", "sortTimeInMicroseconds",
                                  "mergeTimeInMicroseconds", "totalMatches", "inputSum"]
        assert j["totalMatches"] == oracle.true_join(R, S)[0] == n, dist
        assert j["inputSum"] == oracle.sum(R) == n * (n + 1) // 2, dist
        assert j["totalMatches"] == _ref_sortmerge_restated(R, S), dist
    R = oracle.generate_data("uniform", n)
    j = ref.sort_merge(R, S)
    assert j["totalMatches"] == _ref_sortmerge_restated(R, S)
    assert j["totalMatches"] < oracle.true_join(R, S)[0] == n


# ---- mc's skew / non-unique / full-range / local-shuffle generators: the ports are the reference's arrays ----
def _mc_gen(*args):
    """Keys the reference's own generator code produces (oracle/_ref/mc_gen = oracle/ref_gen.c around
    mc/src/generator.c + genzipf.c compiled unmodified)."""
    import subprocess
    exe = os.path.join(ROOT, "oracle", "_ref", "mc_gen")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/mc_gen not built (needs /root/reference once)")
    out = subprocess.run([exe] + [str(a) for a in args], capture_output=True, timeout=120)
    assert out.returncode == 0
    return np.frombuffer(out.stdout, dtype=np.int32).astype(np.int64)


@pytest.mark.parametrize("case", [("pk", 10007, 12345), ("lshuffle", 10007, 16, 12345), ("lshuffle", 4096, 4096, 7),
                                  ("fk", 23000, 5000, 54321), ("nonunique", 20011, 20011, 12345),
                                  ("nonunique", 5003, 2147483647, 99), ("zipf", 30011, 4001, 0.5, 54321),
                                  ("zipf", 30011, 4001, 1.25, 54321), ("zipf", 1000, 1, 1.0, 5),
                                  ("fullrange_s", 9001, 3000, 12345, 54321)])
def test_mc_generator_ports_equal_the_reference(hj, case):
    want = _mc_gen(*case)
    kind = case[0]
    if kind == "pk":
        got = hj.mc_relation_pk(case[1], case[2])
    elif kind == "lshuffle":
        got = hj.mc_relation_pk_lshuffle(case[1], case[2], case[3])
    elif kind == "fk":
        got = hj.mc_relation_fk(case[1], case[2], case[3])
    elif kind == "nonunique":
        got = hj.mc_relation_nonunique(case[1], case[2], case[3])
        assert np.array_equal(hj.mc_relation_nonunique(case[1], case[2], case[3], bias=1), got + np.uint64(1))
    elif kind == "zipf":
        got = hj.mc_relation_zipf(case[1], case[2], case[3], case[4])
        assert got.min() >= 1 and got.max() <= case[2]          # genzipf.c:39: 0 is not in the alphabet
    else:
        pk = hj.mc_relation_nonunique(case[2], 2147483647, case[3])
        got = hj.mc_relation_fk_from_pk(case[1], pk, case[4])
    assert got.dtype == np.uint64 and np.array_equal(got.astype(np.int64), want)


def test_mc_nonunique_emits_key_zero(hj):
    """RAND_RANGE(maxid) can emit 0, the empty-slot sentinel -- that is why the join paths
    take these relations with bias=1."""
    a = hj.mc_relation_nonunique(200000, 1000, 12345)
    assert (a == 0).sum() > 0 and a.max() <= 999


def test_oracle_join16_is_pinned_on_the_reference_key8b_build(oracle, hj):
    """16-byte tuples: mc compiled with -DKEY_8B (mc/src/types.h:22-37) runs NPO on its own PK / FK relations
    (same generator code, so the keys are csrc/datagen.cpp's); its match count is orc_true_join16's."""
    import subprocess
    exe = os.path.join(ROOT, "oracle", "_ref", "mc_key8b")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/mc_key8b not built")
    n, m = 100000, 350000
    out = subprocess.run([exe, "--algo=NPO", "--nthreads=4", f"--r-size={n}", f"--s-size={m}"], capture_output=True,
                         text=True, timeout=300)
    ref_matches = int([l for l in out.stdout.splitlines() if "Results =" in l][0].split("=")[1].split(".")[0])
    rng = np.random.default_rng(1)
    R = np.stack([hj.mc_relation_pk(n), rng.integers(0, 1 << 62, n, dtype=np.uint64)], axis=1)
    S = np.stack([hj.mc_relation_fk(m, n), rng.integers(0, 1 << 62, m, dtype=np.uint64)], axis=1)
    matches, checksum = oracle.true_join16(R, S)
    assert matches == ref_matches == m
    # every S tuple meets the R tuple with its key: checksum = sum of all S words + the R words of the partners
    pay_r = np.zeros(n + 1, dtype=np.uint64)
    pay_r[R[:, 0]] = R[:, 1]
    want = (int(S.sum(dtype=np.uint64)) + int(S[:, 0].sum(dtype=np.uint64)) + int(pay_r[S[:, 0]].sum(dtype=np.uint64))) % (1 << 64)
    assert checksum == want
    # duplicates on both sides, zero keys ignored, key mask honoured
    Rd = np.array([[5, 1], [5, 2], [7, 3], [0, 9], [(1 << 40) | 5, 4]], dtype=np.uint64)
    Sd = np.array([[5, 10], [5, 20], [7, 30], [0, 1], [8, 1]], dtype=np.uint64)
    assert oracle.true_join16(Rd, Sd) == (5, (2 * (5 + 10 + 5 + 20) + 2 * (5 + 1 + 5 + 2) + (7 + 3 + 7 + 30)))
    assert oracle.true_join16(Rd, Sd, 0xFFFFFFFF)[0] == 7
