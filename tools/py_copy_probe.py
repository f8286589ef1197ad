"""Copy-engine rate of ring-ordered 512 MiB block copies between one-process-per-GPU ranks, from inside a Python
process with torch.distributed initialised -- the environment bench.py runs hj_xchg_* in.  Compares backends /
NCCL settings (tools/ce_probe_ipc.cu is the same measurement without Python, torch or NCCL).
  torchrun --nproc-per-node 4 tools/py_copy_probe.py [gloo|nccl]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import htm_hashjoin_b200 as hj

backend = sys.argv[1] if len(sys.argv) > 1 else "nccl"
use_torch_src = len(sys.argv) > 2 and sys.argv[2] == "torchsrc"
rank, G = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group(backend, device_id=torch.device("cuda", rank) if backend == "nccl" else None)
if backend == "nccl":
    t = torch.ones(1, device="cuda")
    dist.all_reduce(t)          # creates the communicator
    torch.cuda.synchronize()
blk = 512 << 20
with hj.HashJoin("radix", device=rank) as h:
    win_extra = int(os.environ.get("PROBE_WINDOW_EXTRA_GIB", "0")) << 30   # a window as large as the exchange's
    win_off = int(os.environ.get("PROBE_WINDOW_OFFSET_GIB", "0")) << 30
    win_bytes = int(os.environ.get("PROBE_WINDOW_BYTES", "0")) or G * blk + win_extra   # an exact size, e.g. an odd one
    win = h.device_alloc(win_bytes)
    if use_torch_src:
        keep = torch.empty(G * blk // 8, dtype=torch.int64, device="cuda")
        src = keep.data_ptr()
    else:
        src = h.device_alloc(G * blk)
    fill = os.environ.get("PROBE_SRC_FILL", "none")   # does the payload's content matter?
    if fill != "none":
        words = G * blk // 8
        if fill == "random":
            data = torch.randint(-(1 << 62), 1 << 62, (words,), dtype=torch.int64, device="cuda")
        elif fill == "tuples":   # what the join ships: 32-bit key, zero payload half
            data = torch.randint(1, 1 << 29, (words,), dtype=torch.int64, device="cuda")
        else:
            data = torch.full((words,), 0x0101010101010101, dtype=torch.int64, device="cuda")
        h.copy(src, data.data_ptr(), G * blk)
        del data
    handles = [None] * G
    dist.all_gather_object(handles, h.ipc_export(win))
    peers = [win if g == rank else h.ipc_open(handles[g]) for g in range(G)]
    for rep in range(2):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(4):
            for k in range(1, G):
                p = (rank + k) % G
                h.copy(peers[p] + win_off + rank * blk, src + p * blk, blk)
        dt = time.perf_counter() - t0
        dist.barrier()
    env = {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}
    if rank == 0:
        print(f'{{"window_bytes": {win_bytes}, "src_fill": "{fill}", "window_extra_gib": {win_extra >> 30}, "window_offset_gib": {win_off >> 30}, "backend": "{backend}", "torch_src": {str(use_torch_src).lower()}, "nccl_env": "{env}", "gpus": {G}, '
              f'"rank0_gbs": {4 * (G - 1) * blk / dt / 1e9:.1f}, "ms_per_512MiB_block": {dt * 1e3 / (4 * (G - 1)):.3f}}}', flush=True)
    for g in range(G):
        if g != rank:
            h.ipc_close(peers[g])
dist.destroy_process_group()
