"""World-size-2 gloo test of the multi-GPU host logic (htm_hashjoin_b200/exchange.py).
The CUDA pieces (range split, local join) are stood in for by numpy / the oracle here;
on the GPU box bench.py runs the same exchange code over NCCL."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _split_by_dest(a, G, dom_bits):
    """numpy statement of hj_exchange_partition's digit (rx_kernels.cuh digit_of mode 2)."""
    dest = np.minimum(((a - np.uint64(1)) * np.uint64(G)) >> np.uint64(dom_bits), np.uint64(G - 1)).astype(np.int64)
    order = np.argsort(dest, kind="stable")
    return a[order], np.bincount(dest, minlength=G)


def _worker(rank, world, port, n, m, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from htm_hashjoin_b200 import exchange
    from tests.oracle_lib import Oracle
    orc = Oracle()
    # global relations: every rank can rebuild them, and keeps only its slice
    R = orc.generate_data("shuffle", n)
    S = orc.mc_fk(m, n)
    r_slice = R[rank * n // world:(rank + 1) * n // world]
    s_slice = S[rank * m // world:(rank + 1) * m // world]
    dom_bits = exchange.domain_bits_for(n)
    recv = []
    for sl in (r_slice, s_slice):
        send, counts = _split_by_dest(sl, world, dom_bits)
        sc = torch.from_numpy(counts.astype(np.int64))
        rc = exchange.exchange_counts(sc)
        buf = torch.empty(int(rc.sum()) + 5, dtype=torch.int64)
        got = exchange.exchange_tuples(torch.from_numpy(send.view(np.int64).copy()), sc.tolist(), buf, rc.tolist())
        recv.append(buf[:got].numpy().view(np.uint64).copy())
    # every key received belongs to this rank's range
    lo, hi = rank * n // world, (rank + 1) * n // world
    for a in recv:
        assert a.size == 0 or (a.min() >= lo + 1 and a.max() <= hi)
    mt, ck = orc.true_join(recv[0], recv[1])
    total = exchange.reduce_result(mt, ck, "cpu")
    if rank == 0:
        q.put((total, orc.true_join(R, S)))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_exchange_matches_global_join(built):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    n, m = 1 << 14, 5 << 14
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, m, q)) for r in range(2)]
    for p in procs:
        p.start()
    total, want = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert total == want
    assert want[0] == m


def test_reduce_wraps_mod_2_64():
    """int64 two's-complement addition is u64 addition mod 2^64."""
    a = np.array([2 ** 63 + 5, 2 ** 64 - 1], dtype=np.uint64).view(np.int64)
    b = np.array([2 ** 63 + 7, 3], dtype=np.uint64).view(np.int64)
    s = (torch.from_numpy(a.copy()) + torch.from_numpy(b.copy())).numpy().view(np.uint64)
    assert s.tolist() == [12, 2]


def _peer_digit(a, G, dom_bits, b1):
    """numpy statement of digit_of mode 3 (exchange fused with radix pass 1)."""
    dest = np.minimum(((a - np.uint64(1)) * np.uint64(G)) >> np.uint64(dom_bits), np.uint64(G - 1)).astype(np.int64)
    return (dest << b1) | (a & np.uint64((1 << b1) - 1)).astype(np.int64)


def _fused_worker(rank, world, port, n, m, q):
    """The fused protocol with numpy standing in for hj_peer_hist / hj_peer_scatter: every rank
    'stores' its runs into the owners' buffers (gathered over gloo here instead of NVLink stores)
    at the offsets exchange.peer_cursors hands out; the owners' buffers must come out exactly
    tiled, pass-1 partitioned, and join to the global result."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from htm_hashjoin_b200 import exchange
    from tests.oracle_lib import Oracle
    orc = Oracle()
    R = orc.generate_data("shuffle", n)
    S = orc.mc_fk(m, n)
    b1, f1 = 3, 8
    dom_bits = exchange.domain_bits_for(n)
    bufs = []
    for rel, tot in ((R, n), (S, m)):
        sl = rel[rank * tot // world:(rank + 1) * tot // world]
        dig = _peer_digit(sl, world, dom_bits, b1)
        counts = torch.from_numpy(np.bincount(dig, minlength=world * f1).astype(np.int64))
        allc = exchange.gather_counts(counts)
        cur, recv, own_off = exchange.peer_cursors(allc, rank, world, f1)
        # what this rank would store where: (dest, absolute offset, tuples)
        writes = []
        for d in range(world * f1):
            t = sl[dig == d]
            if t.size:
                writes.append((d // f1, int(cur[d]), t))
        gathered = [None] * world
        dist.all_gather_object(gathered, writes)
        buf = np.zeros(int(recv[rank]), dtype=np.uint64)
        filled = np.zeros(buf.size, dtype=np.int64)
        for w in gathered:
            for dest, off, t in w:
                if dest == rank:
                    buf[off:off + t.size] = t
                    filled[off:off + t.size] += 1
        assert (filled == 1).all()                                   # exact tiling, no overlap, no gap
        low = (buf & np.uint64(f1 - 1)).astype(np.int64)
        assert (np.diff(low) >= 0).all()                             # partition-major by the low b1 bits
        assert np.array_equal(np.searchsorted(low, np.arange(f1 + 1)), own_off.numpy())  # offsets for hj_join_pass1
        lo, hi = rank * n // world, (rank + 1) * n // world
        assert buf.min() >= lo + 1 and buf.max() <= hi               # only this rank's key range
        bufs.append(buf)
    mt, ck = orc.true_join(bufs[0], bufs[1])
    total = exchange.reduce_result(mt, ck, "cpu")
    if rank == 0:
        q.put((total, orc.true_join(R, S)))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_fused_exchange_protocol(built):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    n, m = 1 << 13, 3 << 13
    procs = [ctx.Process(target=_fused_worker, args=(r, 2, port, n, m, q)) for r in range(2)]
    for p in procs:
        p.start()
    total, want = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert total == want and want[0] == m


def _regions_worker(rank, world, port, n, m, q):
    """Histogram-free exchange on CPU: every rank appends to ITS regions (exchange.region_cap /
    region_cursor_init), the final cursors are all-gathered, exchange.regions_received gives every receiver
    the fill levels of its regions; what sits inside them joins to the global result."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from htm_hashjoin_b200 import exchange
    from tests.oracle_lib import Oracle
    orc = Oracle()
    R = orc.generate_data("shuffle", n)
    S = orc.mc_fk(m, n)
    b1, f1 = 3, 8
    dom_bits = exchange.domain_bits_for(n)
    bufs = []
    for rel, tot in ((R, n), (S, m)):
        sl = rel[rank * tot // world:(rank + 1) * tot // world]
        cap = exchange.region_cap(sl.size, world, f1)
        assert cap % 2 == 0
        dig = _peer_digit(sl, world, dom_bits, b1)
        cur = exchange.region_cursor_init(rank, world, f1, cap, "cpu")
        assert cur.numel() == world * f1
        writes = []
        for d in range(world * f1):
            t = sl[dig == d]
            low = d % f1
            assert int(cur[d]) == (low * world + rank) * cap
            assert t.size <= cap                                     # uniform keys never overflow 8 sigma
            writes.append((d // f1, int(cur[d]), t))
            cur[d] += t.size
        final_all = exchange.gather_counts(cur)                      # doubles as the barrier
        mine, n_recv = exchange.regions_received(final_all, rank, world, f1, cap)
        gathered = [None] * world
        dist.all_gather_object(gathered, writes)
        buf = np.zeros(f1 * world * cap, dtype=np.uint64)
        for w in gathered:
            for dest, off, t in w:
                if dest == rank:
                    assert (buf[off:off + t.size] == 0).all()        # regions never overlap
                    buf[off:off + t.size] = t
        got = []
        for p in range(f1 * world):
            seg = buf[p * cap:int(mine[p])]
            assert (seg != 0).all() and ((seg & np.uint64(f1 - 1)) == np.uint64(p // world)).all()
            got.append(seg)
        got = np.concatenate(got)
        assert got.size == int(n_recv) == np.count_nonzero(buf)
        lo, hi = rank * n // world, (rank + 1) * n // world
        assert got.min() >= lo + 1 and got.max() <= hi
        bufs.append(got)
    mt, ck = orc.true_join(bufs[0], bufs[1])
    total = exchange.reduce_result(mt, ck, "cpu")
    if rank == 0:
        q.put((total, orc.true_join(R, S)))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_region_exchange_protocol(built):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    n, m = 1 << 13, 3 << 13
    procs = [ctx.Process(target=_regions_worker, args=(r, 2, port, n, m, q)) for r in range(2)]
    for p in procs:
        p.start()
    total, want = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert total == want
