"""Pins the CPU oracle (oracle/orb_oracle.cpp) against golden vectors produced by
real OpenCV code (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def prim(golden_dir):
    return np.load(os.path.join(golden_dir, "primitives.npz"))


def test_resize_matches_cv2(oracle_mod, prim):
    for k in ("resize", "resize2"):
        src, want = prim[k + "_src"], prim[k + "_dst"]
        got = oracle_mod.resize_linear(src, want.shape[1], want.shape[0])
        assert np.array_equal(got, want), k


def test_resize_exact_half_is_area(oracle_mod):
    rng = np.random.default_rng(0)
    src = rng.integers(0, 256, (40, 60), dtype=np.uint8)
    got = oracle_mod.resize_linear(src, 30, 20)
    s = src.astype(np.int32)
    want = ((s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2).astype(np.uint8)
    assert np.array_equal(got, want)


def test_gaussian_matches_cv2(oracle_mod, prim):
    assert np.array_equal(oracle_mod.gaussian7(prim["blur_src"]), prim["blur_dst"])
    assert np.array_equal(oracle_mod.gaussian7(prim["resize_src"]), prim["blur2_dst"])


@pytest.mark.parametrize("th", [7, 20])
def test_fast_matches_cv2(oracle_mod, prim, th):
    got = oracle_mod.fast(prim["fast_src"], th)
    want = prim["fast_th%d" % th]
    assert want.shape[0] > 50
    assert np.array_equal(got, want)          # coordinates, scores and raster order


def test_fast_score_is_threshold_independent(oracle_mod, prim):
    lo = oracle_mod.fast(prim["fast_src"], 7)
    hi = oracle_mod.fast(prim["fast_src"], 20)
    assert np.array_equal(lo[lo[:, 2] >= 20], hi)


def test_fast_atan2_matches_cv2(oracle_mod, prim):
    got = oracle_mod.fast_atan2(prim["atan_y"], prim["atan_x"])
    assert np.array_equal(got.view(np.uint32), prim["atan_deg"].view(np.uint32))


def test_knn2_matches_bfmatcher(oracle_mod, prim):
    idx, dist = oracle_mod.knn2(prim["knn_q"], prim["knn_t"])
    assert np.array_equal(idx, prim["knn_idx"])
    assert np.array_equal(dist, prim["knn_dist"])
    idx2, dist2 = oracle_mod.knn2(prim["knn_q"], prim["knn_t"], nthreads=3)
    assert np.array_equal(idx, idx2) and np.array_equal(dist, dist2)


def test_descriptor_distance_is_popcount(oracle_mod, prim):
    q, t = prim["knn_q"], prim["knn_t"]
    for i in range(20):
        want = int(np.unpackbits(q[i] ^ t[i]).sum())
        assert oracle_mod.descriptor_distance(q[i], t[i]) == want


def test_knn2_small_train_sets(oracle_mod):
    q = np.arange(64, dtype=np.uint8).reshape(2, 32)
    idx, dist = oracle_mod.knn2(q, q[:1])
    assert idx.tolist() == [[0, -1], [0, -1]]
    assert dist[0].tolist() == [0, -1]
    idx, dist = oracle_mod.knn2(q, np.zeros((0, 32), np.uint8))
    assert idx.tolist() == [[-1, -1], [-1, -1]]


@pytest.mark.parametrize("name", ["qvga", "vga_lap", "kitti"])
def test_extract_matches_reference_flow(oracle_mod, golden_dir, name):
    g = np.load(os.path.join(golden_dir, "extract_%s.npz" % name))
    nf, nl, ini, mn = [int(v) for v in g["params"]]
    ex = oracle_mod.OracleExtractor(nf, float(g["scale_factor"]), nl, ini, mn)
    mono, kps, desc = ex.compute(g["image"], tuple(int(v) for v in g["lapping"]))
    assert mono == int(g["mono"])
    assert kps.shape == g["kps"].shape
    for f in kps.dtype.names:
        assert np.array_equal(kps[f], g["kps"][f]), f
    assert np.array_equal(desc, g["desc"])
    for l in range(nl):
        im = ex.level_image(l)
        assert list(im.shape) == g["level_shapes"][l].tolist()
        assert int(im.astype(np.int64).sum()) == int(g["level_sums"][l])
        assert len(ex.level_candidates(l)) == int(g["n_candidates"][l])


def test_extract_tables(oracle_mod):
    ex = oracle_mod.OracleExtractor(2000, 1.2, 8, 20, 7)
    assert ex.features_per_level.tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    assert ex.umax.tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert ex.scale[1] == np.float32(1.2) and ex.scale[2] == np.float32(1.2) * np.float32(1.2)


def test_extract_edge_cases(oracle_mod):
    ex = oracle_mod.OracleExtractor(500, 1.2, 8, 20, 7)
    mono, kps, desc = ex.compute(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(kps) == 0
    mono, kps, desc = ex.compute(np.zeros((120, 160), np.uint8))      # all black: no keypoints
    assert mono == 0 and len(kps) == 0 and desc.shape == (0, 32)
    rng = np.random.default_rng(1)
    noise = rng.integers(0, 256, (120, 160), dtype=np.uint8)          # pure noise
    mono, kps, desc = ex.compute(noise)
    assert len(kps) > 50 and mono == len(kps)
    assert (kps["octave"][1:] >= kps["octave"][:-1]).all()            # levels in order when nothing laps
    # non-contiguous input (row step > width) gives the same answer
    big = np.zeros((120, 200), np.uint8)
    big[:, :160] = noise
    mono2, kps2, desc2 = ex.compute(big[:, :160])
    assert mono2 == mono and np.array_equal(desc2, desc)


def test_stereo_matches_reference_flow(oracle_mod, golden_dir):
    """ComputeStereoMatches + kNN-2/0.7-ratio against the cv2-based transliteration."""
    g = np.load(os.path.join(golden_dir, "stereo_kitti.npz"))
    exl, exr = oracle_mod.OracleExtractor(), oracle_mod.OracleExtractor()
    _, kl, dl = exl.compute(g["left"], (0, 0))
    _, kr, dr = exr.compute(g["right"], (0, 0))
    for f in kl.dtype.names:
        assert np.array_equal(kl[f], g["kps_l"][f]) and np.array_equal(kr[f], g["kps_r"][f])
    assert np.array_equal(dl, g["desc_l"]) and np.array_equal(dr, g["desc_r"])
    ur, dp, bidx, bdist = oracle_mod.stereo_match(exl, exr, kl, dl, kr, dr, float(g["bf"]), float(g["b"]))
    assert int((dp > 0).sum()) > 300
    assert np.array_equal(ur.view(np.uint32), g["u_right"].view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), g["depth"].view(np.uint32))
    idx, dist = oracle_mod.knn2(dl, dr)
    m = oracle_mod.ratio_filter(idx, dist, 0.7, 0)
    assert np.array_equal(m, g["ratio_matches"])


def test_dedup_by_train(oracle_mod):
    m = np.array([[0, 5, 10], [1, 3, 7], [2, 5, 4], [3, 3, 9], [4, 1, 1]], np.int32)
    out = oracle_mod.dedup_by_train(m)
    assert out.tolist() == [[4, 1, 1], [1, 3, 7], [0, 5, 10]]


def _window_cases(golden_dir):
    g = np.load(os.path.join(golden_dir, "window_match.npz"))
    return g


def test_search_for_initialization_matches_transliteration(oracle_mod, golden_dir):
    """oracle search_for_initialization == tests/golden/ref_match.py (statement-by-statement transliteration of
    fmatcher.cpp:983-1098 + frame.cpp:386-414,678-757)."""
    g = _window_cases(golden_dir)
    pm = np.stack([g["k1"]["x"], g["k1"]["y"]], 1).astype(np.float32)
    for tag in "abc":
        win, ratio, chk = g[f"ini_{tag}_args"]
        n, m12, pm2 = oracle_mod.search_for_initialization(g["k1"], g["d1"], g["k2"], g["d2"], g["bounds"], pm, int(win), ratio, bool(chk))
        assert n == int(g[f"ini_{tag}_n"])
        assert np.array_equal(m12, g[f"ini_{tag}_m12"])
        assert np.array_equal(pm2.view(np.uint32), g[f"ini_{tag}_pm"].view(np.uint32))


def test_search_by_projection_matches_transliteration(oracle_mod, golden_dir):
    g = _window_cases(golden_dir)
    for tag in "abcd":
        th, direction, chk, stereo = g[f"proj_{tag}_args"]
        n, asg = oracle_mod.search_by_projection(g["uv"], g["invz"], g["k1"]["octave"], g["k1"]["angle"], g["d1"], g["valid"],
                                                 g["observed"], g["k2"], g["d2"], g["uright2"] if stereo else None, g["blocked2"],
                                                 g["bounds"], g["sf"], th, 386.1448, int(direction), bool(chk))
        assert n == int(g[f"proj_{tag}_n"])
        assert np.array_equal(asg, g[f"proj_{tag}_asg"])
