#!/usr/bin/env python3
"""Plain pinned cudaMemcpyAsync bandwidth, D2H and H2D, one process per GPU, all processes at once: the ceiling of the
host-buffer (e2e) numbers of bench.py on this box.  usage: [torchrun --nproc-per-node N] python profiles/pcie_bw.py [GB]"""
import json
import os
import sys

import torch
import torch.distributed as dist

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(gb * 2 ** 30)
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
res = {}
for name, (dst, src) in (("d2h", (h, d)), ("h2d", (d, h))):
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        dst.copy_(src, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 5], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[name + "_GBps_per_gpu"] = n / (float(t.item()) * 1e-3) / 1e9
if rank == 0:
    res.update(n_gpus=world, bytes=n, total_d2h_GBps=res["d2h_GBps_per_gpu"] * world, total_h2d_GBps=res["h2d_GBps_per_gpu"] * world,
               what="pinned cudaMemcpyAsync, 5 copies, slowest rank, all ranks concurrently")
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
