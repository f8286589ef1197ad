"""Multi-GPU exchange plumbing: one process per GPU over torch.distributed.

The join shards on the key: every rank splits its slices of R and S by
destination rank (hj_exchange_partition: dest = min(G-1, ((key-1)*G) >> domain_bits)),
the pieces are swapped with an all-to-all, each rank joins what it received, and
{matches, checksum} are summed.  This module is only the collective part -- counts,
tuples, result -- and is backend agnostic (NCCL on the GPU box, gloo in the CPU tests);
it never computes a join.
"""
import numpy as np
import torch
import torch.distributed as dist


def exchange_counts(send_counts):
    """send_counts: int64 tensor [G] (tuples this rank sends to each rank), on the
    collective's device.  Returns recv_counts [G]."""
    recv = torch.empty_like(send_counts)
    dist.all_to_all_single(recv, send_counts)
    return recv


def exchange_tuples(send_buf, send_counts, recv_buf, recv_counts):
    """All-to-all of 8-byte tuples.  send_buf holds the rank-ordered pieces back to back
    (layout produced by hj_exchange_partition); recv_buf receives rank-ordered pieces.
    Counts are host lists.  Returns the number of tuples received."""
    n_recv = int(sum(recv_counts))
    n_send = int(sum(send_counts))
    if n_recv > recv_buf.numel():
        raise RuntimeError(f"exchange receive buffer too small: {n_recv} > {recv_buf.numel()}")
    dist.all_to_all_single(recv_buf[:n_recv], send_buf[:n_send], list(map(int, recv_counts)),
                           list(map(int, send_counts)))
    return n_recv


def reduce_result(matches, checksum, device):
    """Sum of (matches, checksum) over ranks, both mod 2^64 (two's complement int64 add)."""
    t = torch.from_numpy(np.array([matches, checksum], dtype=np.uint64).view(np.int64).copy()).to(device)
    dist.all_reduce(t)
    out = t.cpu().numpy().view(np.uint64)
    return int(out[0]), int(out[1])


def reduce_result3(matches, checksum, flag, device):
    """reduce_result plus a flag (e.g. 'a region overflowed here') summed over ranks."""
    t = torch.from_numpy(np.array([matches, checksum, flag], dtype=np.uint64).view(np.int64).copy()).to(device)
    dist.all_reduce(t)
    out = t.cpu().numpy().view(np.uint64)
    return int(out[0]), int(out[1]), int(out[2])


def domain_bits_for(domain):
    """Bits of the key domain [1, domain] used by the range split."""
    return max(1, int(domain - 1).bit_length())


# ---- exchange fused with radix pass 1 (hj_peer_hist / hj_peer_scatter / hj_join_pass1) ----

def peer_cursors(counts_all, rank, n_dest, f1):
    """counts_all: int64 tensor [G, G*f1], row s = hj_peer_hist of rank s
    (entry dest*f1 + d1 = tuples rank s holds for pass-1 partition d1 of GPU dest).

    Layout of GPU dest's pass-1 buffer: partitions d1 = 0..f1-1 back to back, inside a partition
    the contributions of source ranks 0..G-1 back to back.  Returns (cursors, recv_counts, part_off):
    cursors [G*f1] = first element of `rank`'s exclusive range for every (dest, d1);
    recv_counts [G] = tuples every GPU receives in total;
    part_off [f1+1] = offsets of the pass-1 partitions inside `rank`'s own buffer (for hj_join_pass1)."""
    G = n_dest
    c = counts_all.view(G, G, f1)                      # [src, dest, d1]
    per_part = c.sum(0)                                # [dest, d1]
    part_off = per_part.cumsum(1) - per_part           # exclusive over d1
    src_prefix = c.cumsum(0) - c                       # exclusive over src
    start = part_off.unsqueeze(0) + src_prefix         # [src, dest, d1]
    recv = per_part.sum(1)
    own_off = torch.cat([part_off[rank], recv[rank:rank + 1]]).contiguous()
    return start[rank].reshape(-1).contiguous(), recv, own_off


# ---- histogram-free exchange (hj_peer_scatter_regions / hj_join_regions) -------------------------
def region_cap(n_local, n_dest, f1):
    """Capacity of one (source, dest, low bits) region: expected tuples + 8 sigma of a random split + 64,
    even (16-byte alignment).  The same number on every rank: all ranks hold n_local tuples."""
    mean = n_local // (n_dest * f1) + 1
    r = 1
    while r * r < mean:
        r += 1
    return (mean + 8 * r + 64 + 1) & ~1


def region_cursor_init(rank, n_dest, f1, cap, device):
    """cursors[(dest * f1) + low] = first element of region low * n_dest + rank (the same in every
    destination buffer)."""
    low = torch.arange(f1, dtype=torch.int64, device=device)
    return ((low * n_dest + rank) * cap).repeat(n_dest).contiguous()


def regions_received(final_all, rank, n_dest, f1, cap):
    """final_all: int64 [G, G*f1], row s = rank s's cursors after hj_peer_scatter_regions.  Returns
    (cur [f1*G], n): the fill levels of `rank`'s own regions, index low * G + src (what hj_join_regions
    takes), and the number of tuples that arrived."""
    G = n_dest
    mine = final_all.view(G, G, f1)[:, rank, :]          # [src, low]
    cur = mine.t().contiguous().view(-1)                 # index low * G + src
    start = torch.arange(f1 * G, dtype=torch.int64, device=final_all.device) * cap
    return cur, (cur - start).sum()


def gather_counts(local_counts):
    """All-gather of the per-rank hj_peer_hist vectors -> [G, len]."""
    G = dist.get_world_size()
    out = torch.empty(G * local_counts.numel(), dtype=local_counts.dtype, device=local_counts.device)
    dist.all_gather_into_tensor(out, local_counts.contiguous().view(-1))
    return out.view(G, local_counts.numel())


def exchange_ipc_handles(handle_bytes):
    """All-gather of one 64-byte CUDA IPC handle per rank -> list of G handles."""
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, handle_bytes)
    return out
