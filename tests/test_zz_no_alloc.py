"""viorb_knn2_device_static: the kNN-2 entry point that allocates nothing (its split scratch is a file-scope device array;
include/viorb.h).  No counterpart in the reference: added for callers that cannot call cudaMalloc.  Checked through the C
ABI against the existing entry point viorb_knn2_device (integer results: bit-exact, the tolerance of every kNN test here)
on two calls in a row that share the static scratch, and against the oracle.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(lib, q, t):
    import torch
    from vi_slam_b200.extractor import _check
    nq, nt = len(q), len(t)
    dq = torch.from_numpy(q).cuda()
    dt = torch.from_numpy(t).cuda() if nt else torch.zeros((1, 32), dtype=torch.uint8, device="cuda")
    out = []
    for fn in (lib.viorb_knn2_device, lib.viorb_knn2_device_static):
        idx = torch.full((nq, 2), -7, dtype=torch.int32, device="cuda")
        dist = torch.full((nq, 2), -7, dtype=torch.int32, device="cuda")
        _check(lib, fn(C.c_void_p(dq.data_ptr()), nq, C.c_void_p(dt.data_ptr()), nt, C.c_void_p(idx.data_ptr()),
                       C.c_void_p(dist.data_ptr()), None))
        torch.cuda.synchronize()
        out.append((idx.cpu().numpy(), dist.cpu().numpy()))
    return out


def test_static_scratch_knn2_equals_allocating_entry_point(oracle_mod):
    import torch
    from vi_slam_b200 import _lib, synthetic
    lib = _lib.load()
    assert b"sm_100a" in lib.viorb_version()
    # two calls in a row on the shared scratch: a long train set (train-range split: the scratch is used), then a shape
    # with other tile counts, then shapes that do not split, a single row, and the first shape again
    shapes = [(300, 20000), (2000, 2000), (5000, 300), (1, 4000), (16384, 700), (300, 20000)]
    for i, (nq, nt) in enumerate(shapes):
        q, t = synthetic.make_descriptors(nq, nt, seed=900 + i % 5)
        if nt > 600:
            t[nt - 1] = t[3]                     # a tie across two splits: the lower index must win
        (i0, d0), (i1, d1) = _both(lib, q, t)
        assert np.array_equal(i0, i1) and np.array_equal(d0, d1), (nq, nt)
        if nq <= 5000:
            oi, od = oracle_mod.knn2(q, t, nthreads=4)
            assert np.array_equal(i1, oi) and np.array_equal(d1, od), (nq, nt)
    # forced split counts reach the scratch on small inputs too (same hook as tests/test_gpu_parity.py)
    try:
        for split in (2, 5, 32):
            lib.viorb_debug_knn_split(split)
            q, t = synthetic.make_descriptors(129, 2570, seed=split)
            (i0, d0), (i1, d1) = _both(lib, q, t)
            assert np.array_equal(i0, i1) and np.array_equal(d0, d1), split
    finally:
        lib.viorb_debug_knn_split(0)
    # empty train set, and on a caller stream
    (i0, d0), (i1, d1) = _both(lib, synthetic.make_descriptors(5, 4, seed=1)[0], np.zeros((0, 32), np.uint8))
    assert (i1 == -1).all() and (d1 == -1).all() and np.array_equal(i0, i1)
    s = torch.cuda.Stream()
    q, t = synthetic.make_descriptors(700, 9000, seed=77)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    idx = torch.empty((700, 2), dtype=torch.int32, device="cuda")
    dist = torch.empty((700, 2), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for _ in range(2):
        assert lib.viorb_knn2_device_static(C.c_void_p(dq.data_ptr()), 700, C.c_void_p(dt.data_ptr()), 9000, C.c_void_p(idx.data_ptr()),
                                            C.c_void_p(dist.data_ptr()), C.c_void_p(s.cuda_stream)) == 0
    s.synchronize()
    oi, od = oracle_mod.knn2(q, t, nthreads=4)
    assert np.array_equal(idx.cpu().numpy(), oi) and np.array_equal(dist.cpu().numpy(), od)


def test_static_scratch_knn2_rejects_sizes_beyond_its_limit():
    import torch
    from vi_slam_b200 import _lib
    lib = _lib.load()
    n = 16384 + 1
    dq = torch.zeros((n, 32), dtype=torch.uint8, device="cuda")
    out = torch.zeros((n, 2), dtype=torch.int32, device="cuda")
    rc = lib.viorb_knn2_device_static(C.c_void_p(dq.data_ptr()), n, C.c_void_p(dq.data_ptr()), 100, C.c_void_p(out.data_ptr()),
                                      C.c_void_p(out.data_ptr()), None)
    assert rc == -4                                  # VIORB_E_CAPACITY
    msg = lib.viorb_last_error().decode()
    assert "16384" in msg and "static scratch" in msg
