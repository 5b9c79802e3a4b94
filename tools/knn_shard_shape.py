#!/usr/bin/env python3
"""One rank's share of the row-sharded 200k x 200k sweep on one GPU: nq x nt kNN-2 through viorb_knn2_device, device time of
the C call alone (CUDA events, 10 repeats).  VIORB_KNN_SPLIT=1 disables the train-range split for an A/B."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vi_slam_b200 import _lib, synthetic  # noqa: E402

lib = _lib.load()
for nq, nt in ((25000, 200000), (2000, 2000), (2000, 20000), (8000, 8000), (32000, 32000)):
    q, t = synthetic.make_descriptors(nq, nt, seed=3)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    idx = torch.empty((nq, 2), dtype=torch.int32, device="cuda")
    dst = torch.empty((nq, 2), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    call = lambda: lib.viorb_knn2_device(dq.data_ptr(), nq, dt.data_ptr(), nt, idx.data_ptr(), dst.data_ptr(), C.c_void_p(st))
    for _ in range(3):
        assert call() == 0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(json.dumps({"nq": nq, "nt": nt, "ms": ms, "pairs_per_s": nq * nt / ms * 1e3, "split": os.environ.get("VIORB_KNN_SPLIT", "auto")}), flush=True)
