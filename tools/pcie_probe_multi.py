#!/usr/bin/env python
"""Aggregate host<->device copy rate with one process per GPU copying at the same time (run under
torchrun): the platform bound of the end-to-end number at N GPUs.  Rank 0 prints one JSON line."""
import json, os, time
import torch
import torch.distributed as dist
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n = 1 << 30
x = torch.empty(n, dtype=torch.uint8).pin_memory(); y = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device=dev); e = torch.empty(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps=4):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1): d.copy_(x, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): y.copy_(e, non_blocking=True)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return reps * n / float(dt.item()) / 1e9     # GB/s per GPU per direction (slowest rank)


run(True, True, 1)
out = {"n_gpus": world, "h2d_gbs_per_gpu": run(True, False), "d2h_gbs_per_gpu": run(False, True),
       "bidir_gbs_per_gpu_per_direction": run(True, True), "cpus": os.cpu_count()}
nodes = [p for p in os.listdir("/sys/devices/system/node") if p.startswith("node")] if os.path.isdir("/sys/devices/system/node") else []
out["numa_nodes"] = len(nodes)
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
