#!/usr/bin/env python
"""Concurrent pinned host->device bandwidth of the box: every rank copies the bench step's packed block
([V,P,43] floats + cameras = 34.4 MB at the C3 shape) from its own pinned buffer, all ranks at once.
Explains the e2e curve of bench.py at N = 8.  Launch with torch.distributed.run;
rank 0 prints one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    bind = "--bind" in sys.argv
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    cpus = None
    if bind:
        from guava_renderer_b200 import sharding
        cpus = sharding.bind_host_to_gpu(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n = 2 * 100_000 * 43 + 64
    host = [torch.randn(n).pin_memory() for _ in range(3)]
    devb = [torch.empty(n, device=dev) for _ in range(3)]
    for i in range(6):
        devb[i % 3].copy_(host[i % 3], non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    reps = 60
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        devb[i % 3].copy_(host[i % 3], non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    gbs = n * 4 * reps / (a.elapsed_time(b) * 1e-3) / 1e9
    t = torch.tensor([gbs], device=dev, dtype=torch.float64)
    if world > 1:
        allv = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
        vals = [float(x) for x in allv]
    else:
        vals = [gbs]
    if rank == 0:
        print(json.dumps({"ranks": world, "bytes_per_copy": n * 4, "numa_bind": bool(cpus), "per_rank_GBps": [round(v, 2) for v in vals],
                          "aggregate_GBps": round(sum(vals), 1)}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
