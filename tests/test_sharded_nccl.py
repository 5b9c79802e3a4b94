"""The NCCL transport of the row-sharded matcher (BASELINE.json configs[4]) on real GPUs: two ranks, the C entry point
viorb_knn2_sharded of the CUDA build, libviorb's own NCCL communicator.  EVERY row is compared with the oracle.
Needs >= 2 GPUs (gpurun --gpus 2); skipped on a 1-GPU box."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu

NQ, NT = 30001, 20000            # ragged last shard


def _data():
    from vi_slam_b200 import synthetic
    q, t = synthetic.make_descriptors(NQ, NT, seed=99)
    t[NT - 1] = t[5]              # planted ties: the lower train index must win on every rank
    t[NT // 2] = t[6]
    q[17] = t[5]
    q[NQ - 1] = t[6]              # ... also in the last (short) shard
    return q, t


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from vi_slam_b200 import sharded, synthetic
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    q, t = _data()
    m = sharded.ShardedMatcher()
    idx, dst = m.match(q, t)                                       # host arrays in, gathered rows out
    # device-resident form, twice into the same buffer (stream order, buffer reuse)
    a, b, per = m.rows(NQ)
    ql, tt = torch.from_numpy(q[a:b]).cuda(), torch.from_numpy(t).cuda()
    out = m.match_device(ql, NQ, tt)
    out = m.match_device(ql, NQ, tt, out=out)
    torch.cuda.synchronize()
    full = out.cpu().numpy()
    # fewer query rows than ranks, and an empty train set
    idx1, dst1 = m.match(q[:1], t)
    idx0, dst0 = m.match(q[:5], t[:0])
    # a second size through the same communicator
    q2, t2 = synthetic.make_descriptors(1000, 70000, seed=3)
    idx2, dst2 = m.match(q2, t2)
    m.close()
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), idx=idx, dist=dst, full=full, per=per, idx1=idx1, dist1=dst1, idx0=idx0,
             dist0=dst0, idx2=idx2, dist2=dst2)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_knn_world2_nccl(oracle_mod, tmp_path):
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from vi_slam_b200 import synthetic
    port = 29600 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    q, t = _data()
    oidx, odist = oracle_mod.knn2(q, t, nthreads=8)
    q2, t2 = synthetic.make_descriptors(1000, 70000, seed=3)
    oidx2, odist2 = oracle_mod.knn2(q2, t2, nthreads=8)
    assert oidx[17, 0] == 5 and odist[17, 0] == 0 and oidx[17, 1] == NT - 1 and odist[17, 1] == 0
    for r in range(2):
        z = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(z["idx"], oidx) and np.array_equal(z["dist"], odist), r           # every row
        full, per = z["full"], int(z["per"])
        assert full.shape == (2 * per, 4)
        assert np.array_equal(full[:NQ, 0:2], oidx) and np.array_equal(full[:NQ, 2:4], odist)
        assert (full[NQ:] == -1).all()                                                              # padding rows
        assert np.array_equal(z["idx1"], oidx[:1]) and np.array_equal(z["dist1"], odist[:1])
        assert (z["idx0"] == -1).all() and (z["dist0"] == -1).all()
        assert np.array_equal(z["idx2"], oidx2) and np.array_equal(z["dist2"], odist2)


def test_packed_rows_equal_pairs(oracle_mod):
    """viorb_knn2_packed_device (the row format the shards exchange) == viorb_knn2_device, incl. the split / merge path."""
    import ctypes as C
    import torch
    from vi_slam_b200 import _lib, synthetic
    from vi_slam_b200.extractor import _check
    lib = _lib.load()
    for nq, nt in ((300, 5000), (129, 257), (2000, 2000), (5, 1), (1000, 70000)):
        q, t = synthetic.make_descriptors(nq, nt, seed=nq + nt)
        dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
        rows = torch.empty((nq, 4), dtype=torch.int32, device="cuda")
        _check(lib, lib.viorb_knn2_packed_device(dq.data_ptr(), nq, dt.data_ptr(), nt, rows.data_ptr(), None))
        torch.cuda.synchronize()
        oidx, odist = oracle_mod.knn2(q, t, nthreads=4)
        r = rows.cpu().numpy()
        assert np.array_equal(r[:, 0:2], oidx) and np.array_equal(r[:, 2:4], odist), (nq, nt)
