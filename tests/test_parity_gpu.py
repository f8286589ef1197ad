"""GPU parity tests: the CUDA path through the C ABI vs the oracle (bit-exact: match
count + order-independent 64-bit checksum), vs the golden KATs, and size-independent
properties at larger sizes."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DISTS = [("sorted", 16), ("shuffle", 16), ("local_shuffle", 16), ("local_shuffle", 4096), ("uniform", 16),
         ("random", 16)]


def _inputs(hj, dist, rng, n):
    R = hj.generate_data(dist, n, n, rng)
    S = R.copy() if dist == "random" else hj.generate_data("sorted", n, n, 16)  # main.cpp:92-97
    return R, S


@pytest.mark.parametrize("dist,rng", DISTS)
@pytest.mark.parametrize("algo", ["atomic", "radix", "adaptive"])
def test_join_matches_oracle(hj, oracle, algo, dist, rng):
    n = 1 << 18
    R, S = _inputs(hj, dist, rng, n)
    want = oracle.true_join(R, S)
    with hj.HashJoin(algo) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        assert res.kernel_launches > 0
        if res.path_taken == hj.HJ_PATH_GLOBAL:
            sums = h.checksums(R)
            assert sums.input_sum == oracle.sum(R)
            assert sums.table_sum + sums.overflow_sum == sums.input_sum  # no tuple lost
            if dist in ("sorted", "shuffle", "local_shuffle"):
                assert res.conflicts == 0 and res.window_matches == n   # pinned by the logs


@pytest.mark.parametrize("tsize", [1, 2, 4, 16])
def test_transaction_size_variants(hj, oracle, tsize):
    n = (1 << 17) + 13  # odd size: tails of the vector loads
    R = hj.generate_data("uniform", n, 1 << 17, 16)
    S = hj.generate_data("sorted", 1 << 17)
    want = oracle.true_join(R, S)
    with hj.HashJoin("atomic", transaction_size=tsize) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


def test_nocc_unique_keys(hj, oracle):
    n = 1 << 18
    R, S = _inputs(hj, "shuffle", 16, n)
    with hj.HashJoin("nocc") as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == oracle.true_join(R, S)
        sums = h.checksums(R)
        assert sums.table_sum == sums.input_sum == n * (n + 1) // 2


def test_serial_reference_numbers_bracket_gpu_conflicts(hj, oracle):
    """conflicts/totalMatches on duplicate keys are schedule dependent: not
    bit-pinned, but window matches + what the overflow table adds must be the true join,
    and the GPU's conflict count must be of the order of the serial one."""
    n = 1 << 20
    R, S = _inputs(hj, "uniform", 16, n)
    serial = oracle.gt_serial(R, S)
    with hj.HashJoin("atomic") as h:
        res = h.join(R, S)
        assert res.matches == n
        assert res.window_matches <= res.matches
        assert 0.5 * serial["conflicts"] < res.conflicts < 2.0 * serial["conflicts"]
        assert res.window_matches >= n - res.conflicts


def test_device_resident_inputs_and_split_calls(hj, oracle):
    n = 1 << 18
    R, S = _inputs(hj, "uniform", 16, n)
    want = oracle.true_join(R, S)
    with hj.HashJoin("atomic") as h:
        dR, dS = h.to_device(R), h.to_device(S)
        b = h.build(dR)
        p = h.probe(dS)
        assert (p.matches, p.checksum) == want
        assert b.us_h2d == 0.0
        # probing twice gives the same answer (idempotent), and an empty probe is fine
        p2 = h.probe(dS)
        assert (p2.matches, p2.checksum) == want
        h.device_free(dR[0]); h.device_free(dS[0])
    with hj.HashJoin("atomic") as h:
        with pytest.raises(hj.HJError) as e:
            h.probe(S)
        assert e.value.code == hj._capi.HJ_ESTATE


@pytest.mark.parametrize("bits,passes", [(4, 1), (7, 1), (10, 1), (12, 1), (8, 2), (11, 2), (14, 2), (15, 2)])
def test_radix_fanout_sweep_fk(hj, oracle, bits, passes):
    n, m = 1 << 17, (5 << 17) + 77
    R = hj.mc_relation_pk(n)
    S = hj.mc_relation_fk(m, n)
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    assert want[0] == m
    with hj.HashJoin("radix", num_partitions=1 << bits, passes=passes, key_mask=0xFFFFFFFF) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        assert (res.radix_bits, res.passes) == (bits, passes)


def test_radix_payloads_and_duplicates(hj, oracle):
    """mc tuples: key in the low 32 bits, payload (row id) in the high 32."""
    rng = np.random.default_rng(5)
    n, m = 100003, 400007
    keys_r = rng.integers(1, 30000, n, dtype=np.uint64)
    keys_s = rng.integers(1, 40000, m, dtype=np.uint64)
    R = keys_r | (np.arange(n, dtype=np.uint64) << np.uint64(32))
    S = keys_s | (np.arange(m, dtype=np.uint64) << np.uint64(32))
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    for algo, kw in (("radix", dict(num_partitions=1 << 6, passes=2)), ("radix", dict(num_partitions=1 << 9, passes=1)),
                     ("atomic", {})):
        with hj.HashJoin(algo, key_mask=0xFFFFFFFF, **kw) as h:
            res = h.join(R, S)
            assert (res.matches, res.checksum) == want, algo


def test_radix_oversized_partitions_l2_table(hj, oracle):
    """Fan-out too small for shared memory: partition-local table in global memory (rx_joinl2.cuh)."""
    n = 1 << 18
    R, S = _inputs(hj, "shuffle", 16, n)
    with hj.HashJoin("radix", num_partitions=2, passes=1) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == oracle.true_join(R, S)
        assert res.max_partition == n // 2


@pytest.mark.parametrize("dist", ["shuffle", "uniform"])
def test_radix_partitions_of_two_chunks_stay_in_shared_memory(hj, oracle, dist):
    """R partitions of up to three table chunks (here 32768 tuples = 2 chunks) take the shared-memory join by
    default, S_p probed once per chunk; the L2 table starts beyond that."""
    n = 1 << 18
    R, S = _inputs(hj, dist, 16, n)
    with hj.HashJoin("radix", num_partitions=8, passes=2) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == oracle.true_join(R, S)
        if dist == "shuffle":
            assert res.max_partition == n // 8
        res2, pairs = h.join_materialize(R, S)  # count pass over chunked partitions
        assert pairs == res.matches


def test_radix_oversized_partitions_chunked(hj, oracle, monkeypatch):
    """Same with the L2 path switched off: R partitions go through shared memory in chunks."""
    monkeypatch.setenv("HJ_NO_L2JOIN", "1")
    n = 1 << 18
    R, S = _inputs(hj, "shuffle", 16, n)
    with hj.HashJoin("radix", num_partitions=2, passes=1) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == oracle.true_join(R, S)
        assert res.max_partition == n // 2


@pytest.mark.parametrize("dist,rng", [("uniform", 16), ("random", 16), ("shuffle", 16)])
@pytest.mark.parametrize("bits,passes", [(3, 1), (8, 2)])
def test_radix_l2_join_duplicates_and_payloads(hj, oracle, monkeypatch, dist, rng, bits, passes):
    """The L2 join forced at fan-outs that would fit shared memory: duplicate keys on both sides, payloads."""
    monkeypatch.setenv("HJ_FORCE_L2JOIN", "1")
    n = (1 << 17) + 5
    R, S = _inputs(hj, dist, rng, n)
    R = R | (np.arange(n, dtype=np.uint64) << np.uint64(32))
    S = S | ((np.arange(n, dtype=np.uint64) * np.uint64(7)) << np.uint64(32))
    want = oracle.true_join(R, S, key_mask=0xFFFFFFFF)
    with hj.HashJoin("radix", key_mask=0xFFFFFFFF, num_partitions=1 << bits, passes=passes) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


@pytest.mark.parametrize("force_l2", [False, True])
def test_radix_wide_keys_take_generic_kernels(hj, oracle, monkeypatch, force_l2):
    """Keys above 2^32: rx_hist flags them, rx_join32 steps aside, rx_join_kernel / the L2 join do the work."""
    if force_l2:
        monkeypatch.setenv("HJ_FORCE_L2JOIN", "1")
    n = 1 << 16
    R = hj.generate_data("shuffle", n) + (np.uint64(1) << np.uint64(40))
    S = np.concatenate([R[: n // 2], hj.generate_data("sorted", n // 2)])
    want = oracle.true_join(R, S)
    assert want[0] == n // 2
    with hj.HashJoin("radix", num_partitions=1 << 6, passes=1) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


def test_radix_histogram_free_overflow_falls_back(hj, oracle):
    """2-pass joins partition without a histogram into fixed-capacity regions (ScOpt, rx_kernels.cuh).
    Skewed keys overflow a region: the library must notice and redo the join on the exact path."""
    n = 1 << 17
    k = np.arange(1, n + 1, dtype=np.uint64)
    R = k << np.uint64(10)                 # every key in radix partition 0 of 2^10
    S = np.concatenate([R, R[: n // 2], k])  # plus keys spread over all partitions
    want = oracle.true_join(R, S)
    assert want[0] == n + n // 2 + np.count_nonzero((k & np.uint64(1023)) == 0)
    with hj.HashJoin("radix", num_partitions=1 << 10, passes=2) as h:
        for _ in range(2):                 # second call: the handle stays on the exact path
            res = h.join(R, S)
            assert (res.matches, res.checksum) == want
        # a well-behaved input on the same handle still joins correctly
        R2, S2 = _inputs(hj, "shuffle", 16, n)
        res = h.join(R2, S2)
        assert (res.matches, res.checksum) == oracle.true_join(R2, S2)


def test_radix_histogram_free_equals_exact(hj, oracle, monkeypatch):
    n = (1 << 18) + 77
    R = hj.generate_data("uniform", n, n, 16)
    S = hj.generate_data("shuffle", 3 * n)
    want = oracle.true_join(R, S)
    for env in (None, "1"):
        if env:
            monkeypatch.setenv("HJ_NO_OPTIMISTIC", env)
        with hj.HashJoin("radix", num_partitions=1 << 9, passes=2) as h:
            res = h.join(R, S)
            assert (res.matches, res.checksum) == want


def test_pro_build_only_checksum(hj, oracle):
    """PRO as shipped by the reference: partition + build, Results = sum idx
    (parallel_radix_join.c:231-283); closed form (n/2)(n/2^B - 1), SURVEY 8c."""
    n = 1 << 20
    R = hj.mc_relation_pk(n)
    for bits, passes in ((14, 2), (10, 1), (7, 1)):
        got = hj.PRO(R, None, radix_bits=bits, passes=passes)
        assert got == (n // 2) * (n // (1 << bits) - 1) == oracle.pro_checksum(R, bits)
    assert hj.PRO(R, None, radix_bits=14, passes=2) == 33030144


def test_radix_partition_building_block(hj, oracle):
    n = 300001
    a = hj.generate_data("shuffle", n)
    with hj.HashJoin("radix") as h:
        d_in = h.to_device(a)
        d_out = h.device_alloc(n * 8)
        for shift, bits in ((0, 7), (3, 5), (5, 11)):
            d_off = h.device_alloc(((1 << bits) + 1) * 8)
            h.radix_partition(d_in[0], n, shift, bits, d_out, d_off)
            out, off = h.to_host(d_out, n), h.to_host(d_off, (1 << bits) + 1)
            want_out, want_off = oracle.radix_partition(a, shift, bits)
            assert np.array_equal(off, want_off)
            for p in range(1 << bits):  # order inside a partition is unspecified
                assert np.array_equal(np.sort(out[off[p]:off[p + 1]]), np.sort(want_out[off[p]:off[p + 1]]))
            h.device_free(d_off)


def test_exchange_partition_ranges(hj):
    n, G, dom = 1 << 18, 8, 18
    a = hj.generate_data("shuffle", n)
    with hj.HashJoin("radix") as h:
        d_in = h.to_device(a)
        d_out, d_off = h.device_alloc(n * 8), h.device_alloc((G + 1) * 8)
        h.exchange_partition(d_in[0], n, G, dom, d_out, d_off)
        out, off = h.to_host(d_out, n), h.to_host(d_off, G + 1)
        assert off.tolist() == [g * (n // G) for g in range(G + 1)]
        for g in range(G):
            seg = out[off[g]:off[g + 1]]
            assert seg.min() == g * (n // G) + 1 and seg.max() == (g + 1) * (n // G)


def test_fused_exchange_logical_ranks(hj, oracle):
    """hj_peer_hist / hj_peer_scatter / hj_join_pass1 with two logical ranks on one GPU: both
    'peer' buffers are local, the protocol and kernels are the ones the NVLink path runs."""
    import torch
    from htm_hashjoin_b200 import exchange
    G, n, m = 2, 1 << 18, 3 << 18
    R = hj.generate_data("shuffle", n)
    S = np.concatenate([hj.generate_data("shuffle", n), hj.generate_data("sorted", n), hj.generate_data("uniform", n)])
    want = oracle.true_join(R, S)
    dom_bits = exchange.domain_bits_for(n)
    with hj.HashJoin("radix", num_partitions=1 << 8, passes=2) as h:
        bits, b1, b2, passes = h.radix_plan(n // G)
        assert (bits, passes) == (8, 2) and b1 + b2 == 8
        f1 = 1 << b1
        total = [0, 0]
        recv_bufs, recv_n, recv_off = {}, {}, {}
        for name, rel in (("R", R), ("S", S)):
            slices = [rel[g * rel.size // G:(g + 1) * rel.size // G] for g in range(G)]
            d_slices = [h.to_device(sl) for sl in slices]
            counts = []
            for g in range(G):
                c = torch.zeros(G * f1, dtype=torch.int64, device="cuda")
                h.peer_hist(d_slices[g][0], slices[g].size, G, dom_bits, b1, c.data_ptr())
                counts.append(c)
            allc = torch.stack(counts)
            assert int(allc.sum()) == rel.size
            bufs = None
            recv_off[name] = []
            for g in range(G):
                cur, recv, own_off = exchange.peer_cursors(allc, g, G, f1)
                recv_off[name].append(own_off.contiguous())
                if bufs is None:
                    bufs = [h.device_alloc((int(recv[d]) + 2) * 8) for d in range(G)]
                cur = cur.contiguous()
                torch.cuda.synchronize()
                if g == 0:
                    h.peer_scatter(d_slices[g][0], slices[g].size, G, dom_bits, b1, bufs, cur.data_ptr())
                else:  # the enqueue-only variant on a limited number of CTAs
                    h.peer_scatter_async(d_slices[g][0], slices[g].size, G, dom_bits, b1, bufs, cur.data_ptr(), 16)
                    h.sync()
            recv_bufs[name], recv_n[name] = bufs, [int(x) for x in recv]
            # what arrived at logical rank 0 is pass-1 partitioned and in its key range
            got = h.to_host(bufs[0], recv_n[name][0])
            assert (np.diff((got & np.uint64(f1 - 1)).astype(np.int64)) >= 0).all()
            assert got.min() >= 1 and got.max() <= n // G
        for with_offsets in (False, True):   # histogram of what arrived / histogram-free pass 2
            total = [0, 0]
            for g in range(G):
                offs = (recv_off["R"][g].data_ptr(), recv_off["S"][g].data_ptr()) if with_offsets else (None, None)
                torch.cuda.synchronize()
                res = h.join_pass1(recv_bufs["R"][g], recv_n["R"][g], recv_bufs["S"][g], recv_n["S"][g], *offs)
                total[0] += res.matches
                total[1] = (total[1] + res.checksum) % (1 << 64)
            assert tuple(total) == want


def test_device_generator_matches_oracle(hj, oracle):
    with hj.HashJoin("radix") as h:
        d = h.device_alloc(5000 * 8)
        for kind, first, count, domain, seed in ((0, 10, 5000, 0, 0), (1, 0, 4096, 4096, 9), (1, 3000, 5000, 2048, 3)):
            h.generate(kind, first, count, max(domain, 1), seed, d)
            assert np.array_equal(h.to_host(d, count), oracle.feistel_gen(kind, first, count, max(domain, 1), seed))


def test_adaptive_switch_direction(hj):
    """Sorted input stays on the global table; a fully shuffled one goes to radix."""
    n = 1 << 24
    S = hj.generate_data("sorted", n)
    with hj.HashJoin("adaptive") as h:
        r_sorted = h.join(S, S)
        assert r_sorted.path_taken == hj.HJ_PATH_GLOBAL and r_sorted.matches == n
        R = hj.generate_data("shuffle", n)
        r_shuf = h.join(R, S)
        assert r_shuf.matches == n and r_shuf.checksum == n * (n + 1)
        assert r_shuf.contention > r_sorted.contention
        assert r_shuf.sampled > 0


def test_golden_2_24_and_full_size_properties(hj, oracle):
    """BASELINE config 1 size: generator KAT, closed-form join result, and the
    no-tuple-lost invariant; idempotence of the whole join."""
    kats = json.load(open(os.path.join(ROOT, "tests", "golden", "kats.json")))
    n = 1 << 24
    R = hj.generate_data("uniform", n, n, 16)
    assert oracle.sum(R) == kats["generators"]["uniform_2^24"]["sum"]
    assert oracle.possum(R) == kats["generators"]["uniform_2^24"]["possum"]
    S = hj.generate_data("sorted", n)
    in_sum = oracle.sum(R)
    # every R key is in [1, n] and S = 1..n: matches = n, checksum = sum(r) + sum(key of r) = 2*inputSum
    for algo in ("atomic", "radix"):
        with hj.HashJoin(algo) as h:
            a = h.join(R, S)
            b = h.join(R, S)
            assert (a.matches, a.checksum) == (n, (2 * in_sum) % (1 << 64)) == (b.matches, b.checksum)
            if algo == "atomic":
                sums = h.checksums(R)
                assert sums.input_sum == in_sum == sums.table_sum + sums.overflow_sum


def test_reference_named_entry_points_and_cli(hj, oracle):
    n = 1 << 20
    R = hj.generate_data("shuffle", n)
    S = hj.generate_data("sorted", n)
    j = hj.AtomicHashBuild(R, n, S, n, 2, 64, 4)
    keys = list(j.keys())
    assert keys[:8] == ["algo", "rSize", "probeLength", "hashBuildTimeInMicroseconds", "conflicts", "totalMatches",
                        "inputSum", "outputSum"]  # AtomicHashBuild.hpp:133-152 order
    assert (j["conflicts"], j["totalMatches"], j["inputSum"], j["outputSum"]) == (0, n, n * (n + 1) // 2,
                                                                                   n * (n + 1) // 2)
    jb = hj.NoCCHashBuild(R, n)  # build-only form (ENABLE_PROBE=0)
    assert "totalMatches" not in jb and jb["outputSum"] == jb["inputSum"]
    jh = hj.HTMHashBuild(R, n, S, n, 16)
    assert jh["algo"] == "htm" and jh["transactionSize"] == 16 and jh["matches"] == n
    exe = os.path.join(ROOT, "htm_hashjoin_b200", "csrc", "main")
    out = subprocess.run([exe, "--algo", "atomic", "--rSize", "1M", "--dataDistr", "uniform"], capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    Ru = hj.generate_data("uniform", n)
    assert line["inputSum"] == oracle.sum(Ru) and line["matches"] == n
    assert line["outputSum"] == line["inputSum"]


# ---- result materialisation and skewed inputs (8f-3) ---------------------------------
def _zipf_fk(m, n, z, seed):
    """S keys drawn from a zipf(z) distribution over a random permutation of R's keys 1..n, payload =
    position (mc's --skew workload, genzipf.c: alphabet permutation + inverse-cdf lookup; numpy instead
    of serial rand(), so not bit-identical to the reference's generator)."""
    rng = np.random.default_rng(seed)
    alphabet = rng.permutation(n).astype(np.uint64) + np.uint64(1)
    cdf = np.cumsum(1.0 / np.arange(1, n + 1, dtype=np.float64) ** z)
    cdf /= cdf[-1]
    keys = alphabet[np.minimum(np.searchsorted(cdf, rng.random(m)), n - 1)]
    return keys | (np.arange(m, dtype=np.uint64) << np.uint64(32))


def _sorted_pairs(h, d_out, n_pairs):
    flat = h.to_host(d_out, 2 * n_pairs).reshape(-1, 2)
    return flat[np.lexsort((flat[:, 1], flat[:, 0]))]


@pytest.mark.parametrize("case", ["pk_fk", "duplicates", "zipf", "ragged"])
def test_materialize_pairs_equal_oracle(hj, oracle, case):
    """hj_join_materialize: the emitted pairs, sorted, are exactly the oracle's pairs; count and checksum
    agree with the counting join; pairs are dense (no gaps)."""
    if case == "pk_fk":
        R, S = hj.mc_relation_pk(1 << 16), hj.mc_relation_fk(1 << 19, 1 << 16)
    elif case == "duplicates":  # both sides repeat keys: several pairs per probe tuple
        R = hj.generate_data("uniform", 1 << 16, 1 << 13, 16) | (np.arange(1 << 16, dtype=np.uint64) << np.uint64(32))
        S = hj.generate_data("uniform", 1 << 17, 1 << 13, 16) | (np.arange(1 << 17, dtype=np.uint64) << np.uint64(40))
    elif case == "zipf":
        R, S = hj.mc_relation_pk(1 << 16), _zipf_fk(1 << 19, 1 << 16, 1.0, 7)
    else:  # odd sizes, keys without partner
        R = hj.generate_data("shuffle", (1 << 15) + 77)
        S = hj.generate_data("shuffle", (1 << 16) + 5)
    want = oracle.true_join_pairs(R, S)
    cnt, chk = oracle.true_join(R, S, 0xFFFFFFFF)
    assert want.shape[0] == cnt
    with hj.HashJoin("radix", key_mask=0xFFFFFFFF) as h:
        res, need = h.join_materialize(R, S)  # count only
        assert need == cnt and res.matches == cnt
        d_out = h.device_alloc(max(need, 1) * 16)
        res, n_pairs = h.join_materialize(R, S, d_out, need)
        assert n_pairs == cnt and (res.matches, res.checksum) == (cnt, chk)
        got = _sorted_pairs(h, d_out, n_pairs)
        assert np.array_equal(got, want)
        plain = h.join(R, S)
        assert (plain.matches, plain.checksum) == (cnt, chk)
        h.device_free(d_out)


def test_materialize_buffer_too_small_reports_size(hj):
    R, S = hj.mc_relation_pk(1 << 14), hj.mc_relation_fk(1 << 16, 1 << 14)
    with hj.HashJoin("radix", key_mask=0xFFFFFFFF) as h:
        d_out = h.device_alloc(1000 * 16)
        with pytest.raises(hj.HJError) as e:
            h.join_materialize(R, S, d_out, 1000)
        assert e.value.code == hj.HJ_EOVERFLOW and e.value.n_pairs == 1 << 16
        h.device_free(d_out)


def test_materialize_large_properties(hj):
    """2^24 x 2^26 FK join: every S tuple appears exactly once with its R partner (size-independent
    properties: pair count, key equality of the halves, sum of the s halves == sum of S)."""
    n, m = 1 << 24, 1 << 26
    with hj.HashJoin("radix") as h:
        dR, dS = h.device_alloc(n * 8), h.device_alloc(m * 8)
        h.generate(1, 0, n, n, 12345, dR)
        h.generate(1, 0, m, n, 54321, dS)
        d_out = h.device_alloc(m * 16)
        res, n_pairs = h.join_materialize((dR, n), (dS, m), d_out, m)
        assert n_pairs == m == res.matches
        plain = h.join((dR, n), (dS, m))
        assert plain.checksum == res.checksum
        pairs = h.to_host(d_out, 2 * m).reshape(-1, 2)
        assert np.array_equal(pairs[:, 0] & np.uint64(0xFFFFFFFF), pairs[:, 1] & np.uint64(0xFFFFFFFF))
        S = h.to_host(dS, m)
        assert int(pairs[:, 1].sum(dtype=np.uint64)) == int(S.sum(dtype=np.uint64))
        assert int((pairs[:, 0].sum(dtype=np.uint64) + pairs[:, 1].sum(dtype=np.uint64)) & np.uint64(2**64 - 1)) == res.checksum
        for p in (dR, dS, d_out):
            h.device_free(p)


@pytest.mark.parametrize("z", [0.5, 1.0, 1.5])
@pytest.mark.parametrize("algo", ["atomic", "radix", "adaptive"])
def test_zipf_skewed_probe_side(hj, oracle, algo, z):
    """mc --skew: a zipf-distributed foreign key.  Heavy hitters make S partitions very uneven (the
    histogram-free regions overflow and the exact layout takes over); the join result must not change."""
    n, m = 1 << 18, 1 << 21
    R, S = hj.mc_relation_pk(n), _zipf_fk(m, n, z, 11)
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    assert want[0] == m
    with hj.HashJoin(algo, key_mask=0xFFFFFFFF) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want


@pytest.mark.parametrize("skewed", [False, True])
def test_region_exchange_logical_ranks(hj, oracle, skewed):
    """hj_peer_scatter_regions / hj_join_regions (histogram-free exchange) with two logical ranks on one
    GPU.  Uniform keys: the regions hold everything and the join equals the oracle.  Skewed keys (one
    key repeated): the overflow is reported, nothing is written outside the region or the scratch zone."""
    import torch
    from htm_hashjoin_b200 import exchange
    G, n, m = 2, 1 << 18, 3 << 18
    R = hj.generate_data("shuffle", n)
    S = np.concatenate([hj.generate_data("shuffle", n), hj.mc_relation_fk(n, n) & np.uint64(0xFFFFFFFF),
                        hj.generate_data("shuffle", n)])  # random order: every (source, dest) pair gets its share
    if skewed:
        S = S.copy()
        S[: m // 4] = 12345
    want = oracle.true_join(R, S)
    dom_bits = exchange.domain_bits_for(n)
    with hj.HashJoin("radix", num_partitions=1 << 8, passes=2) as h:
        bits, b1, b2, passes = h.radix_plan(n // G)
        f1 = 1 << b1
        state = {}
        any_ovf = False
        for name, rel in (("R", R), ("S", S)):
            slices = [rel[g * rel.size // G:(g + 1) * rel.size // G] for g in range(G)]
            cap = exchange.region_cap(slices[0].size, G, f1)
            trash = f1 * G * cap
            total_elems = trash + 16384 + 2 + 2
            bufs = [h.device_alloc(total_elems * 8) for _ in range(G)]
            guard = np.uint64(0xDEADBEEFDEADBEEF)
            fill = np.full(total_elems, guard, dtype=np.uint64)
            for b in bufs:
                h.copy(b, fill.ctypes.data, fill.nbytes)
            finals = []
            for g in range(G):
                d_sl = h.to_device(slices[g])
                cur = exchange.region_cursor_init(g, G, f1, cap, "cuda")
                torch.cuda.synchronize()
                any_ovf |= h.peer_scatter_regions(d_sl[0], slices[g].size, G, dom_bits, b1, bufs, cur.data_ptr(), g, cap, trash)
                finals.append(cur)
                h.device_free(d_sl[0])
            final_all = torch.stack(finals)
            state[name] = (bufs, final_all, cap)
            # nothing is written outside [region start, fill level) and the scratch zone: every other word of
            # every destination buffer still holds the guard value
            for d in range(G):
                got = h.to_host(bufs[d], total_elems)
                touched = np.zeros(total_elems, dtype=bool)
                touched[trash:trash + 16384 + 2] = True                      # scratch zone (diverted claims)
                for src in range(G):
                    cur = final_all[src].cpu().numpy().reshape(G, f1)[d]     # rank src's cursors for dest d
                    for low in range(f1):
                        start = (low * G + src) * cap
                        end = min(int(cur[low]), start + cap)
                        assert end >= start
                        touched[start:end] = True
                assert (got[~touched] == guard).all(), f"{name}: a word outside regions/scratch of buffer {d} changed"
                if not skewed:  # every claimed slot was filled with a tuple of the right owner and low bits
                    vals = got[touched & (np.arange(total_elems) < trash)]
                    assert (vals != guard).all()
        assert any_ovf == skewed
        if skewed:
            return
        total = [0, 0]
        for g in range(G):
            curR, nr = exchange.regions_received(state["R"][1], g, G, f1, state["R"][2])
            curS, ns = exchange.regions_received(state["S"][1], g, G, f1, state["S"][2])
            torch.cuda.synchronize()
            res = h.join_regions(state["R"][0][g], int(nr), curR.data_ptr(), state["R"][2],
                                 state["S"][0][g], int(ns), curS.data_ptr(), state["S"][2], G)
            total[0] += res.matches
            total[1] = (total[1] + res.checksum) % (1 << 64)
        assert sum(int(exchange.regions_received(state["S"][1], g, G, f1, state["S"][2])[1]) for g in range(G)) == m
        assert tuple(total) == want
