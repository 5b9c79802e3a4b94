"""Resident frame-to-frame matching (viorb_f2f_batch, the monocular path of BASELINE.json configs[0]) against the oracle:
SearchForInitialization (fmatcher.cpp:983-1098) with vbPrevMatched = F1's keypoints (tracking.cpp:2286-2288) and the
brute-force kNN-2 + ratio (matchFeatures method 2) between consecutive frames of one batch."""
import numpy as np
import pytest

import oracle
from vi_slam_b200 import FExtractor, matcher, synthetic


def _sequence(seed, n, w, h):
    f = [synthetic.make_frame(seed, w, h)]
    for i in range(1, n):
        f.append(synthetic.make_next_frame(f[-1], seed + i))
    return f


def _check(lib_path, seed, n, w, h, nf, window=100, ratio=0.8):
    frames = _sequence(seed, n, w, h)
    ex = FExtractor(nf, 1.2, 8, 20, 7, max_size=(w, h), max_batch=n, lib_path=lib_path)
    mono, kps, desc = ex.compute_batch(frames, (0, 1000))           # frame.cpp:289: mono frames pass {0, 1000}
    matcher.f2f_batch(ex, n, None, window, 0.9, True, ratio, 1)
    out = matcher.f2f_fetch_all(ex, n - 1)
    bounds = (0.0, float(w), 0.0, float(h))
    total = 0
    for p in range(n - 1):
        k1, d1, k2, d2 = kps[p], desc[p], kps[p + 1], desc[p + 1]
        pm = np.stack([k1["x"], k1["y"]], 1)
        on, om12, _ = oracle.search_for_initialization(k1, d1, k2, d2, bounds, pm, window, 0.9, True)
        assert out["nmatches"][p] == on, p
        assert np.array_equal(out["matches12"][p, :len(k1)], om12), p
        oidx, odist = oracle.knn2(d1, d2)
        om = oracle.ratio_filter(oidx, odist, ratio, 1)
        want = np.full(len(k1), -1, np.int32)
        want[om[:, 0]] = om[:, 1]
        assert np.array_equal(out["knn_match"][p, :len(k1)], want) and out["n_ratio"][p] == len(om), p
        total += on
    return total


def test_f2f_emulation(oracle_mod):
    import __graft_entry__ as ge
    emu = ge.build_emu()
    assert _check(emu, 40, 4, 320, 240, 500) > 50
    assert _check(emu, 41, 2, 400, 200, 700, window=30) > 20


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n,w,h,nf", [(50, 6, 1241, 376, 2000), (51, 3, 1241, 376, 10000), (52, 2, 1280, 720, 4000)])
def test_f2f_cuda(oracle_mod, seed, n, w, h, nf):
    assert _check(None, seed, n, w, h, nf) > 100
