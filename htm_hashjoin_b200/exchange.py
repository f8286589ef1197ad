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


def domain_bits_for(domain):
    """Bits of the key domain [1, domain] used by the range split."""
    return max(1, int(domain - 1).bit_length())
