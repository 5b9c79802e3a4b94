"""Multi-GPU partitioning of the path.

* Frame batches: frame i -> rank i % world; no collective (``frame_shard``).
* Large brute-force matcher query sets (BASELINE.json config 5): query rows are split in contiguous
  blocks of ceil(Nq / world) per rank, the train set is replicated, every rank runs the kNN-2 kernel on
  its block and ONE all-gather of (idx0, idx1, dist0, dist1) int32 rows returns the full result
  (NCCL over NVLink on GPUs; gloo in the CPU tests).  The exchange is <= 16 B per query: latency bound.
"""
from __future__ import annotations

import numpy as np


def frame_shard(n_frames: int, world: int, rank: int):
    """Indices of the frames rank `rank` extracts (round robin, no exchange needed)."""
    return list(range(rank, n_frames, world))


def row_shard(nq: int, world: int, rank: int):
    per = -(-nq // world) if nq else 0
    a = min(rank * per, nq)
    return a, min(a + per, nq), per


def knn_match_sharded(query: np.ndarray, train: np.ndarray, group=None, lib_path=None, device=0):
    """kNN-2 of replicated host arrays with query rows sharded over the process group.
    Returns (idx[nq,2], dist[nq,2]) on every rank; equal to matcher.knn_match(query, train)."""
    import torch
    import torch.distributed as dist
    from . import matcher

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    nq = query.shape[0]
    a, b, per = row_shard(nq, world, rank)
    local = np.full((max(per, 1), 4), -1, np.int32)
    if b > a:
        idx, dst = matcher.knn_match(query[a:b], train, device=device, lib_path=lib_path)
        local[: b - a, 0:2] = idx
        local[: b - a, 2:4] = dst
    if world == 1:
        full = local
    else:
        on_gpu = dist.get_backend(group) == "nccl"
        t = torch.from_numpy(local)
        if on_gpu:
            t = t.cuda(device)
        out = torch.empty((world * local.shape[0], 4), dtype=torch.int32, device=t.device)
        dist.all_gather_into_tensor(out, t, group=group)
        full = out.cpu().numpy()
    full = full[:nq] if per * world >= nq else full
    return np.ascontiguousarray(full[:, 0:2]), np.ascontiguousarray(full[:, 2:4])


def knn_match_sharded_device(q_dev, t_dev, group=None):
    """Device-resident variant used by the sweep benchmark: q_dev is THIS rank's block of query rows
    (torch uint8 [per, 32] on the GPU, rows beyond the valid count may be padding), t_dev the replicated
    train set.  Launches the kNN-2 kernel on the current torch stream and all-gathers the packed rows with
    NCCL.  Returns the gathered int32 [world * per, 4] tensor (idx0, idx1, dist0, dist1)."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from . import _lib
    from .extractor import _check

    lib = _lib.load()
    per = q_dev.shape[0]
    idx = torch.empty((per, 2), dtype=torch.int32, device=q_dev.device)
    dst = torch.empty((per, 2), dtype=torch.int32, device=q_dev.device)
    stream = torch.cuda.current_stream().cuda_stream
    _check(lib, lib.viorb_knn2_device(q_dev.data_ptr(), per, t_dev.data_ptr(), t_dev.shape[0], idx.data_ptr(),
                                      dst.data_ptr(), C.c_void_p(stream)))
    packed = torch.cat([idx, dst], dim=1).contiguous()
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return packed
    out = torch.empty((dist.get_world_size(group) * per, 4), dtype=torch.int32, device=q_dev.device)
    dist.all_gather_into_tensor(out, packed, group=group)
    return out
