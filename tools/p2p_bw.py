#!/usr/bin/env python
"""NVLink ceilings on this box, both directions at once (run with torchrun, 2 ranks):
copy-engine peer copy vs an SM kernel storing into the peer's buffer (8 B and 16 B per lane).
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/p2p_bw.py"""
import os, sys, time
import torch, torch.distributed as dist
from torch.utils.cpp_extension import load_inline
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import htm_hashjoin_b200 as hj
from htm_hashjoin_b200 import exchange

rank = int(os.environ["RANK"]); torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
N = 1 << 29  # u64 elements = 4 GiB
h = hj.HashJoin("radix", device=rank)
buf = h.device_alloc(N * 8)
handles = exchange.exchange_ipc_handles(h.ipc_export(buf))
peer = h.ipc_open(handles[1 - rank])
src = torch.arange(N, dtype=torch.int64, device="cuda")

def timed(fn, reps=3):
    fn(); torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); t = (time.perf_counter() - t0) / reps
    dist.barrier()
    return N * 8 / t / 1e9

# 1. copy engine (cudaMemcpy D2D to the peer mapping through the library's hj_copy)
bw = timed(lambda: h.copy(peer, src.data_ptr(), N * 8))
if rank == 0: print(f"hj_copy (cudaMemcpy) to peer, both directions at once: {bw:.0f} GB/s per direction", flush=True)
# 2. NCCL all_to_all_single of the same volume
recv = torch.empty_like(src); 
def a2a():
    dist.all_to_all_single(recv.view(2, -1), torch.stack([src[:N//2], src[:N//2]]) if False else src.view(2, -1))
bw = timed(a2a)
if rank == 0: print(f"NCCL all_to_all_single 4 GiB/rank (half stays local): {bw/2:.0f} GB/s per direction over NVLink", flush=True)
