"""Shared parity checks: the library under test (CUDA build on the GPU box, or the g++ emulation
build of the same sources on CPU) against the oracle.  Integer outputs must be bit-exact; the angle
must agree within 1e-4 rad (north star) - it is in fact compared bit-exact and the tolerance only
decides pass/fail; descriptor rows that differ because cosf/sinf rounding differs are counted."""
from __future__ import annotations

import numpy as np

import oracle
from vi_slam_b200 import FExtractor, matcher

ANGLE_TOL_DEG = 1e-4 * 180.0 / np.pi      # 1e-4 rad


def check_extract(img, nfeatures, lapping, lib_path=None, stages=True, max_desc_diff=0, ex=None, ox=None):
    h, w = img.shape
    ox = ox or oracle.OracleExtractor(nfeatures, 1.2, 8, 20, 7)
    om, ok, od = ox.compute(img, lapping)
    gx = ex or FExtractor(nfeatures, 1.2, 8, 20, 7, max_size=(w, h), max_batch=1, lib_path=lib_path)
    gm, gk, gd = gx.compute(img, None, lapping)
    if stages:
        for l in range(8):
            assert np.array_equal(ox.level_image(l), gx.pyramid_level(l)), f"pyramid level {l}"
            assert np.array_equal(ox.level_candidates(l), gx.debug_candidates(l)), f"FAST candidates level {l}"
            okk, oang = ox.level_keypoints(l)
            gkk, gang = gx.debug_kept(l)
            assert np.array_equal(okk, gkk), f"quadtree level {l}"
            assert np.array_equal(oang, gang), f"angle level {l}"
            bl = ox.level_blurred(l)
            if bl is not None:
                assert np.array_equal(bl, gx.debug_blurred(l)), f"blur level {l}"
    assert gm == om and len(gk) == len(ok)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(gk[f], ok[f]), f
    assert np.abs(gk["angle"] - ok["angle"]).max(initial=0) <= ANGLE_TOL_DEG
    ndiff = int((gd != od).any(axis=1).sum())
    assert ndiff <= max_desc_diff, f"{ndiff} descriptor rows differ"
    return gx, ox, (gm, gk, gd), ndiff


def check_knn(q, t, lib_path=None):
    idx, dist = matcher.knn_match(q, t, lib_path=lib_path)
    oidx, odist = oracle.knn2(q, t, nthreads=4)
    assert np.array_equal(idx, oidx)
    assert np.array_equal(dist, odist)
    return idx, dist


def check_ratio(q, t, lib_path=None):
    oidx, odist = oracle.knn2(q, t, nthreads=4)
    m0 = matcher.fisheye_ratio_match(q, t, 0.7, lib_path=lib_path)
    o0 = oracle.ratio_filter(oidx, odist, 0.7, 0)
    assert np.array_equal(np.stack([m0["query_idx"], m0["train_idx"], m0["distance"]], 1).reshape(-1, 3), o0)
    for ratio in (0.8, 1.0):        # 1.0 = what the shipped yaml yields through Config::get<int>
        m1 = matcher.match_features(q, t, 2, ratio, lib_path=lib_path)
        o1 = oracle.dedup_by_train(oracle.ratio_filter(oidx, odist, ratio, 1))
        assert np.array_equal(np.stack([m1["query_idx"], m1["train_idx"], m1["distance"]], 1).reshape(-1, 3), o1)
    return len(o0)


def check_stereo(left, right, nfeatures=2000, lib_path=None, bf=386.1448, fx=718.856):
    h, w = left.shape
    b = float(np.float32(bf) / np.float32(fx))
    gl, ol, (_, kl, dl), nd1 = check_extract(left, nfeatures, (0, 0), lib_path, stages=False, max_desc_diff=2)
    gr, orr, (_, kr, dr), nd2 = check_extract(right, nfeatures, (0, 0), lib_path, stages=False, max_desc_diff=2)
    # feed the oracle the library's own keypoints / descriptors so a (counted) cosf-rounding descriptor
    # difference cannot masquerade as a matcher difference
    our, odp, obi, obd = oracle.stereo_match(ol, orr, kl, dl, kr, dr, bf, b)
    ur, dp, bi, bd = matcher.compute_stereo_matches(gl, gr, bf, b)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd)
    assert np.array_equal(ur.view(np.uint32), our.view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), odp.view(np.uint32))
    # batched resident path gives the same rows + the fisheye-style ratio matches
    matcher.stereo_batch(gl, gr, 1, bf, b, 0.7)
    r = matcher.stereo_fetch(gl, 0)
    assert np.array_equal(r["u_right"].view(np.uint32), our.view(np.uint32))
    assert np.array_equal(r["depth"].view(np.uint32), odp.view(np.uint32))
    assert r["n_stereo"] == int((odp > 0).sum())
    oidx, odist = oracle.knn2(dl, dr)
    om = oracle.ratio_filter(oidx, odist, 0.7, 0)
    want = np.full(len(kl), -1, np.int32)
    want[om[:, 0]] = om[:, 1]
    assert np.array_equal(r["lr_match"], want) and r["n_ratio"] == len(om)
    mm = matcher.match_resident(gl, 0, gr, 0, 0.7, 0)
    assert np.array_equal(np.stack([mm["query_idx"], mm["train_idx"], mm["distance"]], 1).reshape(-1, 3), om)
    return int((odp > 0).sum()), len(om), nd1 + nd2
