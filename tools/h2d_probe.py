"""Aggregate pinned host <-> device copy bandwidth with one process per GPU (torchrun): what the box's host side can
deliver when N ranks move their step's bytes at the same time.  Prints one JSON line from rank 0."""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = 149317120
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(43043200, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(43043200, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
res = {}
for mode in ("h2d", "d2h", "both"):
    for it in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(20):
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    nbytes = 20 * ((n if mode != "d2h" else 0) + (43043200 if mode != "h2d" else 0))
    t = torch.tensor([dt], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[mode] = {"per_gpu_GBps": nbytes / float(t.item()) / 1e9, "aggregate_GBps": world * nbytes / float(t.item()) / 1e9}
if rank == 0:
    print(json.dumps({"n_gpus": world, **res}))
