"""GPU tests added in round 2: the bitmap-claimed global table (edge cases of the claim runs,
the home-only and the sector probe), misaligned device inputs, a reserved handle that allocates
nothing, stale-table protection, 1-pass plans beyond 2^12, wide pass 1."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


# ---- global table: bitmap claims ---------------------------------------------------------------
@pytest.mark.parametrize("tsize", [1, 2, 4, 16])
@pytest.mark.parametrize("case", ["dups_adjacent", "dups_far", "two_homes_collide", "wrap", "all_same", "tiny"])
def test_global_table_claim_edge_cases(hj, oracle, case, tsize):
    """Inputs built to hit every branch of claim_slot / the loser ranking: duplicate keys in neighbouring
    lanes (a run that asks for one bit twice), duplicates far apart (different warps), keys whose windows
    overlap, homes at the very end of the table (wrap-around), one key repeated (everything overflows),
    and a relation smaller than one warp."""
    rng = np.random.default_rng(3)
    if case == "dups_adjacent":
        R = np.repeat(np.arange(1, 20001, dtype=np.uint64), 3)          # k k k, sorted
    elif case == "dups_far":
        R = np.tile(np.arange(1, 20001, dtype=np.uint64), 3)            # k ... k ... k
    elif case == "two_homes_collide":
        base = np.arange(1, 30001, dtype=np.uint64)
        R = np.concatenate([base, base + np.uint64(1), base[::2]])      # windows overlap everywhere
    elif case == "wrap":
        n = 1 << 12                                                     # table = 2^13 slots
        R = np.concatenate([np.full(7, (1 << 13) - 1, dtype=np.uint64), np.arange(1, n - 6, dtype=np.uint64)])
    elif case == "all_same":
        R = np.full(5000, 77, dtype=np.uint64)
    else:
        R = np.array([5, 5, 6, 900, 5], dtype=np.uint64)
    R = R | (rng.integers(0, 1 << 20, R.size, dtype=np.uint64) << np.uint64(32))   # payloads: tuples differ
    S = np.concatenate([np.arange(1, 40001, dtype=np.uint64), np.array([(1 << 13) - 1, 77, 5], dtype=np.uint64)])
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    with hj.HashJoin("atomic", key_mask=0xFFFFFFFF, transaction_size=tsize) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        sums = h.checksums(R)
        assert (sums.table_sum + sums.overflow_sum) % (1 << 64) == sums.input_sum == oracle.sum(R)   # no tuple lost
        assert res.window_matches <= res.matches
        if case == "all_same":
            assert res.conflicts == R.size - 4            # a window of probeLength slots, the rest overflows
        # the probe is idempotent and independent of S's order
        p = h.probe(S[::-1].copy())
        assert (p.matches, p.checksum) == want


@pytest.mark.parametrize("probe_length", [1, 2, 3, 4, 7, 9, 33])
def test_global_table_probe_lengths(hj, oracle, probe_length):
    """--probeLength other than 4: windows of 1..33 slots (several sectors, several bitmap words)."""
    n = 1 << 16
    R = hj.generate_data("uniform", n, n // 4, 16)       # heavy duplication: long displacement chains
    S = hj.generate_data("sorted", n // 4)
    want = oracle.true_join(R, S)
    with hj.HashJoin("atomic", probe_length=probe_length) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        sums = h.checksums(R)
        assert sums.table_sum + sums.overflow_sum == sums.input_sum


def test_home_only_probe_is_taken_and_exact(hj, oracle):
    """Unique keys: the build displaces nothing (probe_steps == 0), the probe reads one slot per tuple.
    Keys that differ only above the table mask must not match (same home, different key)."""
    n = 1 << 18
    R = hj.generate_data("shuffle", n)
    with hj.HashJoin("atomic") as h:
        b = h.build(R)
        assert b.probe_steps == 0 and b.conflicts == 0
        S = np.concatenate([hj.generate_data("sorted", n), R[:1000] + np.uint64(1 << 19)])  # same home, other key
        p = h.probe(S)
        assert (p.matches, p.checksum) == oracle.true_join(R, S) and p.matches == n
        assert p.window_matches == n


def test_nocc_and_atomic_agree_on_unique_keys(hj, oracle):
    n = (1 << 18) + 3
    R = hj.generate_data("local_shuffle", n, n, 256)
    S = hj.generate_data("sorted", n)
    want = oracle.true_join(R, S)
    for algo in ("atomic", "nocc"):
        with hj.HashJoin(algo) as h:
            res = h.join(R, S)
            assert (res.matches, res.checksum) == want
            sums = h.checksums(R)
            assert sums.table_sum == sums.input_sum and sums.overflow_sum == 0


# ---- ADVICE round 1 -----------------------------------------------------------------------------
@pytest.mark.parametrize("algo", ["radix", "adaptive", "atomic", "sortmerge"])
def test_misaligned_device_inputs(hj, oracle, algo):
    """Device pointers that are only 8-byte aligned (ptr + 8): hj_join re-stages them for the radix
    kernels' 16-byte loads; the building blocks refuse them with HJ_EINVAL instead of faulting."""
    n, m = (1 << 17) + 1, (1 << 18) + 1
    R, S = hj.generate_data("shuffle", n), hj.generate_data("shuffle", m)
    want = oracle.true_join(R[1:], S[1:])
    with hj.HashJoin(algo) as h:
        dR, dS = h.to_device(R), h.to_device(S)
        res = h.join((dR[0] + 8, n - 1), (dS[0] + 8, m - 1))
        assert (res.matches, res.checksum) == want
        if algo == "radix":
            d_out, d_off = h.device_alloc(n * 8), h.device_alloc(129 * 8)
            with pytest.raises(hj.HJError) as e:
                h.radix_partition(dR[0] + 8, n - 1, 0, 7, d_out, d_off)
            assert e.value.code == hj.HJ_EINVAL and "16-byte" in str(e.value)
            res = h.join((dR[0], n), (dS[0], m))          # the context is still healthy
            assert (res.matches, res.checksum) == oracle.true_join(R, S)
            h.device_free(d_out); h.device_free(d_off)
        h.device_free(dR[0]); h.device_free(dS[0])


@pytest.mark.parametrize("algo,kw", [("radix", {}), ("radix", dict(num_partitions=1 << 7, passes=1)), ("atomic", {}),
                                     ("adaptive", {}), ("sortmerge", {})])
def test_reserved_handle_allocates_nothing(hj, algo, kw):
    """hj_reserve + hj_reserve_staging size what hj_join uses: no cudaMalloc inside the (timed) call."""
    n, m = 1 << 18, 1 << 20
    R, S = hj.mc_relation_pk(n), hj.mc_relation_fk(m, n)
    with hj.HashJoin(algo, key_mask=0xFFFFFFFF, **kw) as h:
        h.reserve(n, m)
        h.reserve_staging(n, m)
        before = h.alloc_count()
        assert before > 0
        res = h.join(R, S)
        assert res.matches == m
        assert h.alloc_count() == before, "hj_join allocated after hj_reserve"
        dR, dS = h.to_device(R), h.to_device(S)
        res = h.join(dR, dS)
        assert res.matches == m and h.alloc_count() == before
        h.device_free(dR[0]); h.device_free(dS[0])


def test_stale_table_is_not_probed(hj, oracle):
    """hj_build, then a radix call on the same handle (counters zeroed, maybe the table reused), then
    hj_probe: must be refused, not silently wrong.  Same after hj_reserve grew the table."""
    n = 1 << 16
    R = hj.generate_data("uniform", n, n // 8, 16)
    S = hj.generate_data("sorted", n // 8)
    with hj.HashJoin("atomic") as h:
        h.build(R)
        p = h.probe(S)
        assert (p.matches, p.checksum) == oracle.true_join(R, S)
        _, pairs = h.join_materialize(R, S)
        assert pairs == p.matches
        with pytest.raises(hj.HJError) as e:
            h.probe(S)
        assert e.value.code == hj._capi.HJ_ESTATE
        h.build(R)
        h.reserve(8 * n)                       # a larger table: the old contents are gone
        with pytest.raises(hj.HJError) as e:
            h.probe(S)
        assert e.value.code == hj._capi.HJ_ESTATE
        h.build(R)
        p = h.probe(S)
        assert (p.matches, p.checksum) == oracle.true_join(R, S)


# ---- BASELINE configs[2]: 1 pass up to 2^14, any split of two passes -----------------------------
@pytest.mark.parametrize("bits,passes,b1", [(13, 1, 0), (14, 1, 0), (15, 1, 0), (13, 2, 9), (14, 2, 3), (15, 2, 12)])
def test_radix_one_pass_beyond_4096_and_wide_pass1(hj, oracle, bits, passes, b1):
    n, m = 1 << 17, (3 << 17) + 11
    R = hj.mc_relation_pk(n)
    S = hj.mc_relation_fk(m, n)
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    with hj.HashJoin("radix", num_partitions=1 << bits, passes=passes, key_mask=0xFFFFFFFF, pass1_bits=b1) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        assert (res.radix_bits, res.passes) == (bits, passes)
    if passes == 1:   # PRO build-only checksum (closed form) through the direct scatter
        assert hj.PRO(R, None, radix_bits=bits, passes=1) == (n // 2) * (n // (1 << bits) - 1)


def test_pass1_bits_out_of_range_is_einval(hj):
    with hj.HashJoin("radix", num_partitions=1 << 14, passes=2, pass1_bits=13) as h:
        with pytest.raises(hj.HJError) as e:
            h.join(hj.mc_relation_pk(1 << 12), hj.mc_relation_pk(1 << 12))
        assert e.value.code == hj.HJ_EINVAL


# ---- pipelined exchange over the copy engines (hj_xchg_*) with logical ranks on one GPU ----------
@pytest.mark.parametrize("G,chunks,skewed", [(2, 1, False), (2, 3, False), (4, 2, False), (2, 2, True), (4, 4, "mc")])
def test_xchg_pipeline_logical_ranks(hj, oracle, G, chunks, skewed):
    """G logical ranks = G handles on one device; the 'peer windows' are local pointers, the copies are
    device-to-device on the same GPU -- the protocol (stages, flags, fill levels, pass 2 over (source, low)
    regions, join per chunk) is the one the NVLink path runs.  Two epochs on the same windows; skewed keys
    must report the region overflow instead of a wrong result."""
    from htm_hashjoin_b200 import exchange
    n, m = 1 << 18, 3 << 18          # global sizes
    R = hj.generate_data("shuffle", n)
    S = np.concatenate([hj.generate_data("shuffle", n), hj.mc_relation_fk(n, n) & np.uint64(0xFFFFFFFF),
                        hj.generate_data("shuffle", n)])
    if skewed == "mc":   # mc --non-unique: duplicate keys on both sides, spread evenly over the key range
        R, S, skewed = hj.mc_relation_nonunique(n, n, 12345, bias=1), hj.mc_relation_nonunique(m, n, 54321, bias=1), False
    elif skewed:
        S = S.copy()
        S[: m // 4] = 12345
    want = oracle.true_join(R, S)
    dom_bits = exchange.domain_bits_for(n)
    nl, ml = n // G, m // G
    hs = [hj.HashJoin("radix", num_partitions=1 << 8, passes=2, pass1_bits=3) for _ in range(G)]
    try:
        wins = [h.xchg_create(G, g, nl, ml, chunks, dom_bits)[0] for g, h in enumerate(hs)]
        for h in hs:
            h.xchg_connect(wins)
        dR = [hs[g].to_device(R[g * nl:(g + 1) * nl]) for g in range(G)]
        dS = [hs[g].to_device(S[g * ml:(g + 1) * ml]) for g in range(G)]
        for epoch in (1, 2):
            for g in range(G):
                hs[g].xchg_start(dR[g][0], dS[g][0], epoch)
            total, ovf = [0, 0], False
            for g in range(G):
                try:
                    res = hs[g].xchg_finish()
                    total[0] += res.matches
                    total[1] = (total[1] + res.checksum) % (1 << 64)
                    assert res.us_total > 0 and res.kernel_launches > 0
                except hj.HJError as e:
                    assert e.code == hj.HJ_EOVERFLOW
                    ovf = True
            assert ovf == skewed
            if not skewed:
                assert tuple(total) == want
        with pytest.raises(hj.HJError):          # epochs must increase
            hs[0].xchg_start(dR[0][0], dS[0][0], 2)
    finally:
        for h in hs:
            h.close()


def test_pro_entry_point_with_the_reference_scripts_dummy_s(hj):
    """experiments/motivation.sh:11 runs `--algo=PRO --s-size=2` and logs the build checksum: the reference's PRO
    never looks at S.  The drop-in PRO treats an S of <= 2 tuples as that dummy (Results = sum idx, closed form
    (n/2)(n/2^B - 1)) and joins a real S."""
    n = 1 << 20
    R = hj.mc_relation_pk(n)
    dummy = hj.mc_relation_fk(2, n)
    assert hj.PRO(R, dummy, nthreads=8) == hj.PRO(R, None) == 33030144 == (n // 2) * (n // (1 << 14) - 1)
    S = hj.mc_relation_fk(4 * n, n)
    assert hj.PRO(R, S, nthreads=8) == 4 * n


# ---- mc's own skew / non-unique / full-range workloads, generated by the bit-identical ports ----
@pytest.mark.parametrize("algo", ["atomic", "radix", "adaptive", "sortmerge"])
@pytest.mark.parametrize("workload", ["skew0.5", "skew1.0", "skew1.5", "non_unique", "full_range", "lshuffle"])
def test_mc_workloads_join_equals_oracle(hj, oracle, algo, workload):
    """mc --skew z / --non-unique / --full-range / --local-shuffle-range (mc/src/main.c:362-411), inputs from
    csrc/datagen.cpp's ports of generator.c / genzipf.c (pinned on the reference's code in test_oracle_cpu.py)."""
    n, m = 1 << 17, 1 << 19
    if workload.startswith("skew"):
        R, S = hj.mc_relation_pk(n), hj.mc_relation_zipf(m, n, float(workload[4:]))
    elif workload == "non_unique":      # keys in [0, n) + 1: no tuple carries the empty-slot sentinel
        R, S = hj.mc_relation_nonunique(n, n, 12345, bias=1), hj.mc_relation_nonunique(m, n, 54321, bias=1)
    elif workload == "full_range":      # 31-bit keys, S = copies of R shuffled
        R = hj.mc_relation_nonunique(n, 2147483647, 12345, bias=1)
        S = hj.mc_relation_fk_from_pk(m, R, 54321)
    else:
        R, S = hj.mc_relation_pk_lshuffle(n, 64), hj.mc_relation_fk(m, n)
    want = oracle.true_join(R, S, 0xFFFFFFFF)
    assert want[0] >= m if workload != "non_unique" else want[0] > 0
    with hj.HashJoin(algo, key_mask=0xFFFFFFFF) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        assert res.zero_keys == 0


@pytest.mark.parametrize("algo,kw", [("atomic", {}), ("nocc", {}), ("radix", {}), ("radix", dict(num_partitions=64, passes=1)),
                                     ("sortmerge", {})])
def test_zero_keys_are_reported_not_silently_dropped(hj, oracle, algo, kw):
    """mc --non-unique as the reference generates it: RAND_RANGE(maxid) emits key 0 (generator.c:20,231-238), the
    empty-slot sentinel.  Such tuples take no part in the join -- and the result says how many there were."""
    n, m = 100003, 300007
    R, S = hj.mc_relation_nonunique(n, 5000, 12345), hj.mc_relation_nonunique(m, 5000, 54321)
    zr, zs = int((R == 0).sum()), int((S == 0).sum())
    assert zr > 0 and zs > 0
    want = oracle.true_join(R[R != 0], S[S != 0], 0xFFFFFFFF)
    with hj.HashJoin(algo, key_mask=0xFFFFFFFF, **kw) as h:
        res = h.join(R, S)
        if algo != "nocc":   # nocc races by design on duplicate keys
            assert (res.matches, res.checksum) == want
        assert res.zero_keys == zr + zs


def test_experiment_harness_w_sweep(hj, tmp_path):
    """experiments/w_sweep.py (the reference's probe.sh / motivation.sh sweep against this library)
    at 2^20 for four shuffle windows: every point verified against the closed-form result, the adaptive path on
    the global table for small windows, and the figure writer takes its output."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "w.jsonl"
    subprocess.run([sys.executable, os.path.join(root, "experiments", "w_sweep.py"), "--log-n", "20", "--w-step", "7",
                    "--reps", "2", "--out", str(out)], check=True, capture_output=True, timeout=600)
    rows = [json.loads(l) for l in open(out)]
    assert len(rows) == 4 * 4   # W = 2^0, 2^7, 2^14, 2^20 x nocc, atomic, radix, adaptive
    n = 1 << 20
    for r in rows:
        assert r["verified"] is True and r["totalMatches"] == n and r["hashBuildTimeInMicroseconds"] > 0
    by = {(r["algo"], r["shuffleRange"]): r for r in rows}
    assert by[("adaptive", 1)]["pathTaken"] == "global" and by[("atomic", 1)]["pathTaken"] == "global"
    assert by[("radix", 1 << 20)]["pathTaken"] == "radix"
    # locality is measured: the contention signal grows with the window
    assert by[("adaptive", 1 << 20)]["firstRoundFailureFraction"] > by[("adaptive", 1)]["firstRoundFailureFraction"]
    svg = tmp_path / "perf.svg"
    subprocess.run([sys.executable, os.path.join(root, "experiments", "plot_perf.py"), "--in", str(out), "--out", str(svg)],
                   check=True, capture_output=True)
    assert svg.read_text().startswith("<svg")


# ---- 16-byte tuples (mc's KEY_8B, mc/src/types.h:22-37) ---------------------------------------------
def _wide(keys, seed):
    rng = np.random.default_rng(seed)
    return np.ascontiguousarray(np.stack([keys.astype(np.uint64), rng.integers(0, 1 << 63, keys.size, dtype=np.uint64)], axis=1))


@pytest.mark.parametrize("algo,kw", [("radix", {}), ("radix", dict(num_partitions=1 << 9, passes=1)),
                                     ("radix", dict(num_partitions=1 << 12, passes=2)), ("atomic", {})])
@pytest.mark.parametrize("case", ["pk_fk", "duplicates", "wide_keys", "zeros"])
def test_wide_tuples_join_equals_oracle(hj, oracle, algo, kw, case):
    n, m = (1 << 16) + 3, (3 << 16) + 5
    if case == "pk_fk":
        R, S = _wide(hj.mc_relation_pk(n), 1), _wide(hj.mc_relation_fk(m, n), 2)
    elif case == "duplicates":
        R, S = _wide(hj.generate_data("uniform", n, 1 << 13, 16), 3), _wide(hj.generate_data("uniform", m, 1 << 13, 16), 4)
    elif case == "wide_keys":   # keys above 2^32 (and above the table mask): the whole 64-bit key is compared
        k = hj.mc_relation_pk(n) * np.uint64(0x100000001)
        R, S = _wide(k, 5), _wide(np.concatenate([k, k[: m - n] + np.uint64(1 << 40)]), 6)
    else:
        R, S = _wide(hj.mc_relation_nonunique(n, 4096, 12345), 7), _wide(hj.mc_relation_nonunique(m, 4096, 54321), 8)
    want = oracle.true_join16(R, S)
    assert want[0] > 0
    with hj.HashJoin(algo, tuple_bytes=16, **kw) as h:
        res = h.join(R, S)
        assert (res.matches, res.checksum) == want
        assert res.zero_keys == int((R[:, 0] == 0).sum() + (S[:, 0] == 0).sum())
        dR, dS = h.to_device(R.reshape(-1)), h.to_device(S.reshape(-1))   # device-resident inputs
        res2 = h.join(dR, dS)
        assert (res2.matches, res2.checksum) == want and res2.us_h2d == 0.0
        if algo == "atomic":
            sums = h.checksums(R)
            assert sums.input_sum == int(R.sum(dtype=np.uint64))
            zero_sum = int(R[R[:, 0] == 0].sum(dtype=np.uint64))     # tuples with key 0 are in neither table
            assert (sums.table_sum + sums.overflow_sum + zero_sum) % (1 << 64) == sums.input_sum
            h.build(R)
            p = h.probe(S)
            assert (p.matches, p.checksum) == want
        h.device_free(dR[0]); h.device_free(dS[0])


def test_wide_tuples_refuse_what_is_not_built(hj):
    with pytest.raises(hj.HJError) as e:
        hj.HashJoin("adaptive", tuple_bytes=16)
    assert e.value.code == hj.HJ_EINVAL


# ---- the reference's own inputs at sizes beyond the small parity cases ("what's weak") ----
def test_mc_pk_fk_at_2_22_by_2_25(hj):
    """create_relation_pk / create_relation_fk (mc/src/generator.c:241-282,409-445; seeds of mc/src/main.c) at
    2^22 x 2^25: every S tuple finds its one R tuple.  Expected pair sum computed with numpy (R indexed by key):
    the C oracle needs 15 s at this size and is pinned against the same definition on the small cases."""
    n, m = 1 << 22, 1 << 25
    R, S = hj.mc_relation_pk(n), hj.mc_relation_fk(m, n)
    key = np.uint64(0xFFFFFFFF)
    by_key = np.zeros(n + 1, dtype=np.uint64)
    by_key[(R & key).astype(np.int64)] = R                       # PK: keys 1..n, each once
    assert (by_key[1:] != 0).all()
    with np.errstate(over="ignore"):
        want = (m, int((by_key[(S & key).astype(np.int64)] + S).sum(dtype=np.uint64)))
    for algo in ("radix", "adaptive", "atomic"):
        with hj.HashJoin(algo, key_mask=0xFFFFFFFF) as h:
            res = h.join(R, S)
            assert (res.matches, res.checksum) == want, algo
            assert res.zero_keys == 0


@pytest.mark.parametrize("log_w", [8, 12, 16, 20, 22])
def test_local_shuffle_windows_at_2_22(hj, log_w):
    """DataGen local_shuffle with --shuffleRange 2^8 .. 2^22 (experiments/motivation.sh:21-24) at |R| = |S| = 2^22,
    S = sorted 1..n (main.cpp:93): unique keys, so matches = n and the checksum has the closed form n(n+1); every
    algorithm, and the adaptive switch whichever way it goes."""
    n = 1 << 22
    R = hj.generate_data("local_shuffle", n, n, 1 << log_w)
    S = hj.generate_data("sorted", n)
    for algo in ("atomic", "nocc", "radix", "adaptive"):
        with hj.HashJoin(algo) as h:
            res = h.join(R, S)
            assert (res.matches, res.checksum) == (n, n * (n + 1)), (algo, log_w)
            if algo in ("atomic", "nocc"):
                sums = h.checksums(R)
                assert sums.input_sum == n * (n + 1) // 2
                assert (sums.table_sum + sums.overflow_sum) % (1 << 64) == sums.input_sum
