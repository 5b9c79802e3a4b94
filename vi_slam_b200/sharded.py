"""Multi-GPU partitioning of the path, host side over the C ABI.

* Frame batches: frame i -> rank i % world; no collective (``frame_shard``).
* Large brute-force matcher query sets (BASELINE.json config 5): ``ShardedMatcher`` is one rank's end of
  ``viorb_knn2_sharded`` (include/viorb.h): query rows in contiguous blocks of ceil(Nq / world) per rank, train set
  replicated, the kNN-2 kernel writes 16-byte (idx0, idx1, dist0, dist1) rows straight into its block of the full result
  and ONE in-place all-gather completes it on every rank.  Under a NCCL process group the exchange is libviorb's own
  NCCL communicator (NVLink / NVSwitch); torch.distributed only carries the 128-byte unique id.  Under gloo (the CPU
  tests with the emulation build) the exchange is a callback into torch.distributed.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .extractor import _check


def frame_shard(n_frames: int, world: int, rank: int):
    """Indices of the frames rank `rank` extracts (round robin, no exchange needed)."""
    return list(range(rank, n_frames, world))


def row_shard(nq: int, world: int, rank: int, lib_path=None):
    """(row0, row_end, rows_per_rank) of `rank` — viorb_shard_rows, the partition the C side uses."""
    a, n, per = C.c_int32(), C.c_int32(), C.c_int32()
    _lib.load(lib_path).viorb_shard_rows(int(nq), int(world), int(rank), C.byref(a), C.byref(n), C.byref(per))
    return a.value, a.value + n.value, per.value


class ShardedMatcher:
    """One rank of the row-sharded kNN-2.  Collective: construct it and call it on every rank of `group`."""

    def __init__(self, group=None, device=None, lib_path=None):
        import torch
        import torch.distributed as dist
        self._lib = _lib.load(lib_path)
        self._group = group
        self._h = C.c_void_p()
        self._cb = None
        inited = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if inited else 1
        self.rank = dist.get_rank(group) if inited else 0
        backend = dist.get_backend(group) if inited else None
        self.on_gpu = backend == "nccl" or (backend is None and lib_path is None and torch.cuda.is_available())
        self.device = int(device if device is not None else (torch.cuda.current_device() if self.on_gpu else 0))
        if self.world > 1 and backend == "nccl":
            ident = torch.zeros(128, dtype=torch.uint8)
            if self.rank == 0:
                _check(self._lib, self._lib.viorb_comm_unique_id(C.c_void_p(ident.data_ptr())))
            ident = ident.cuda(self.device)
            src = dist.get_global_rank(group, 0) if group is not None else 0
            dist.broadcast(ident, src=src, group=group)
            ident = ident.cpu()
            _check(self._lib, self._lib.viorb_comm_create_nccl(C.c_void_p(ident.data_ptr()), self.rank, self.world, self.device,
                                                               C.byref(self._h)))
        else:
            # world 1, or a CPU process group: the in-place all-gather is done by torch.distributed on the buffer
            def gather(_user, buf, nbytes, rank, world, _stream):
                try:
                    arr = np.ctypeslib.as_array(C.cast(buf, C.POINTER(C.c_uint8)), shape=(world * nbytes,))
                    full = torch.from_numpy(arr)
                    mine = full[rank * nbytes:(rank + 1) * nbytes].clone()
                    dist.all_gather_into_tensor(full, mine, group=self._group)
                    return 0
                except Exception:          # never let an exception cross the C boundary
                    return 1
            self._cb = _lib.ALLGATHER_FN(gather)
            _check(self._lib, self._lib.viorb_comm_create_callback(C.cast(self._cb, C.c_void_p), None, self.rank, self.world,
                                                                   self.device, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.viorb_comm_destroy(self._h)
            self._h = None

    __del__ = close

    def rows(self, nq):
        a, n, per = C.c_int32(), C.c_int32(), C.c_int32()
        self._lib.viorb_shard_rows(int(nq), self.world, self.rank, C.byref(a), C.byref(n), C.byref(per))
        return a.value, a.value + n.value, per.value

    def match_device(self, q_local, nq_total, t_dev, out=None):
        """q_local: THIS rank's block of query rows (torch uint8 [rows, 32] on the GPU), t_dev the replicated train set.
        Asynchronous on the current torch stream.  Returns int32 [world * per, 4] (idx0, idx1, dist0, dist1); rows past
        nq_total are -1."""
        import torch
        _, _, per = self.rows(nq_total)
        if out is None:
            out = torch.empty((self.world * per, 4), dtype=torch.int32, device=t_dev.device)
        stream = torch.cuda.current_stream().cuda_stream
        _check(self._lib, self._lib.viorb_knn2_sharded(self._h, C.c_void_p(q_local.data_ptr() if q_local.numel() else 0), int(nq_total),
                                                       C.c_void_p(t_dev.data_ptr()), int(t_dev.shape[0]), C.c_void_p(out.data_ptr()),
                                                       C.c_void_p(stream)))
        return out

    def match(self, query: np.ndarray, train: np.ndarray):
        """Replicated host arrays in, (idx[nq,2], dist[nq,2]) out on every rank; equal to matcher.knn_match(query, train)."""
        query = np.ascontiguousarray(query, np.uint8).reshape(-1, 32)
        train = np.ascontiguousarray(train, np.uint8).reshape(-1, 32)
        nq = query.shape[0]
        a, b, per = self.rows(nq)
        if nq == 0:
            return np.zeros((0, 2), np.int32), np.zeros((0, 2), np.int32)
        if self.on_gpu:
            import torch
            dev = torch.device("cuda", self.device)
            ql = torch.from_numpy(query[a:b]).to(dev)
            tt = torch.from_numpy(train).to(dev) if len(train) else torch.zeros((1, 32), dtype=torch.uint8, device=dev)[:0]
            full = self.match_device(ql, nq, tt).cpu().numpy()
        else:           # emulation build: "device" pointers are host pointers
            full = np.empty((self.world * per, 4), np.int32)
            ql = np.ascontiguousarray(query[a:b])
            tr = train if len(train) else np.zeros((1, 32), np.uint8)
            _check(self._lib, self._lib.viorb_knn2_sharded(self._h, C.c_void_p(ql.ctypes.data if b > a else 0), nq, C.c_void_p(tr.ctypes.data),
                                                           len(train), C.c_void_p(full.ctypes.data), None))
        full = full[:nq]
        return np.ascontiguousarray(full[:, 0:2]), np.ascontiguousarray(full[:, 2:4])


def knn_match_sharded(query: np.ndarray, train: np.ndarray, group=None, lib_path=None, device=None):
    """Convenience wrapper: a ShardedMatcher for one call (collective)."""
    m = ShardedMatcher(group, device, lib_path)
    try:
        return m.match(query, train)
    finally:
        m.close()
