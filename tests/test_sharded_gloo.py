"""world_size-2 gloo test of the row-sharded matcher (BASELINE.json config 5 partitioning), on CPU: the C entry point
viorb_knn2_sharded of the emulation build partitions, runs the local kernel and pads; the in-place all-gather is the
callback transport (torch.distributed/gloo).  The NCCL transport is tested on GPUs (tests/test_sharded_nccl.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, emu, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from vi_slam_b200 import sharded, synthetic
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    q, t = synthetic.make_descriptors(301, 200, seed=21)          # 301 rows: ragged last shard
    m = sharded.ShardedMatcher(lib_path=emu)
    idx, dst = m.match(q, t)
    idx1, dst1 = m.match(q[:1], t)                                 # fewer rows than ranks: rank 1 owns nothing
    m.close()
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), idx=idx, dist=dst, idx1=idx1, dist1=dst1)
    dist.barrier()
    dist.destroy_process_group()


def test_row_shard_bounds():
    import __graft_entry__ as ge
    from vi_slam_b200 import sharded
    emu = ge.build_emu()
    rs = lambda n, w, r: sharded.row_shard(n, w, r, lib_path=emu)
    assert rs(10, 4, 0) == (0, 3, 3) and rs(10, 4, 3) == (9, 10, 3)
    assert rs(2, 4, 3) == (2, 2, 1) and rs(0, 4, 1) == (0, 0, 0)
    assert sharded.frame_shard(10, 4, 1) == [1, 5, 9]
    cover = sorted(i for r in range(8) for i in range(*rs(200_000, 8, r)[:2]))
    assert cover == list(range(200_000))


def test_sharded_knn_world2_gloo(oracle_mod, tmp_path):
    import torch.multiprocessing as mp
    import __graft_entry__ as ge
    from vi_slam_b200 import synthetic
    emu = ge.build_emu()
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, emu, str(tmp_path)), nprocs=2, join=True)
    q, t = synthetic.make_descriptors(301, 200, seed=21)
    oidx, odist = oracle_mod.knn2(q, t)
    for r in range(2):
        z = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(z["idx"], oidx) and np.array_equal(z["dist"], odist)
        assert np.array_equal(z["idx1"], oidx[:1]) and np.array_equal(z["dist1"], odist[:1])


def test_sharded_world1_callback_transport(oracle_mod):
    import __graft_entry__ as ge
    from vi_slam_b200 import sharded, synthetic
    emu = ge.build_emu()
    q, t = synthetic.make_descriptors(130, 77, seed=5)
    idx, dst = sharded.knn_match_sharded(q, t, lib_path=emu)
    oidx, odist = oracle_mod.knn2(q, t)
    assert np.array_equal(idx, oidx) and np.array_equal(dst, odist)
