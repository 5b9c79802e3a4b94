"""KannalaBrandt8::TriangulateMatches and the match loop of Frame::ComputeStereoFishEyeMatches (src/geometry/cameramodels/
kannalabrandt8.cpp:324-393, src/datastructures/frame.cpp:1148-1189), SURVEY.md section 8 row f-4.  The library (CUDA
build under -m gpu, the g++ emulation build otherwise) against tests/golden/kb8.npz (tests/golden/ref_kb8.py: the
reference's arithmetic with the real cv::SVD of OpenCV 4.13) and against the oracle.

Floating point, so a tolerance - written here: REL = 1e-3 relative on depth and on the 3D point (observed: 6e-5), and the
accept / reject decision must be identical unless the oracle reports that a tested quantity lies within MARGIN = 1e-3 of
its threshold (at most 1 % of the pairs may be such borderline cases)."""
import os

import numpy as np
import pytest

import __graft_entry__ as ge
import oracle
from tests.golden.make_kb8_golden import PRM1, PRM2, make_matches, rig_pose
from vi_slam_b200 import matcher

REL, MARGIN = 1e-3, 1e-3
SIGMA2 = ((np.float32(1.2) ** np.arange(8, dtype=np.float32)) ** 2).astype(np.float32)


def _compare(depth, p3d, ref_depth, ref_p3d, margin):
    acc, ref_acc = depth > 0.0001, ref_depth > 0.0001
    mism = acc != ref_acc
    if margin is not None:
        assert (margin[mism] < MARGIN).all(), "decision differs away from every threshold"
    assert mism.sum() <= max(1, len(depth) // 100)
    both = acc & ref_acc
    assert both.sum() > 0.3 * len(depth)
    assert (np.abs(depth[both] - ref_depth[both]) <= REL * ref_depth[both]).all()
    scale = np.linalg.norm(ref_p3d[both], axis=1)
    assert (np.abs(p3d[both] - ref_p3d[both]).max(1) <= REL * scale).all()
    assert (depth[~acc] == -1).all() and (p3d[~acc] == 0).all()


def test_oracle_matches_cv_svd_golden(oracle_mod, golden_dir):
    g = np.load(os.path.join(golden_dir, "kb8.npz"))
    rig = oracle.kb8_rig(g["prm1"], g["prm2"], g["R12"], g["t12"])
    d, p, m = oracle.kb8_triangulate(rig, g["pt1"], g["pt2"], g["sigma1"], g["sigma2"])
    _compare(d, p, g["depth"], g["p3d"], m)
    rej = g["depth"] < 0
    assert rej.sum() > 500 and (~rej).sum() > 500                      # both outcomes are well represented


def _check_golden(golden_dir, lib):
    g = np.load(os.path.join(golden_dir, "kb8.npz"))
    rig = matcher.kb8_rig(g["prm1"], g["prm2"], g["R12"], g["t12"])
    d, p = matcher.triangulate_matches(rig, g["pt1"], g["pt2"], g["sigma1"], g["sigma2"], lib_path=lib)
    _, _, m = oracle.kb8_triangulate(rig, g["pt1"], g["pt2"], g["sigma1"], g["sigma2"])
    _compare(d, p, g["depth"], g["p3d"], m)


def _fisheye_frames(n, seed, mono_left, mono_right):
    """Two keypoint / descriptor sets of a synthetic fisheye stereo frame: the first mono_* keypoints of each side lie outside
    the lapping area (never matched), the others are projections of common 3D points with noisy descriptors + distractors."""
    rng = np.random.default_rng(seed)
    p1, p2, _, _ = make_matches(n, seed)
    nl, nr = mono_left + n, mono_right + n + n // 4
    kl, kr = np.zeros(nl, oracle.KP_DTYPE), np.zeros(nr, oracle.KP_DTYPE)
    for k in (kl, kr):
        k["x"], k["y"] = rng.uniform(0, 512, len(k)), rng.uniform(0, 512, len(k))
        k["octave"] = rng.integers(0, 8, len(k))
        k["size"], k["angle"], k["class_id"] = 31, rng.uniform(0, 360, len(k)), -1
    perm = rng.permutation(n + n // 4)[:n]
    kl["x"][mono_left:], kl["y"][mono_left:] = p1[:, 0], p1[:, 1]
    kr["x"][mono_right + perm], kr["y"][mono_right + perm] = p2[:, 0], p2[:, 1]
    dl = rng.integers(0, 256, (nl, 32), dtype=np.uint8)
    dr = rng.integers(0, 256, (nr, 32), dtype=np.uint8)
    noisy = dl[mono_left:].copy()
    for i in range(n):
        for bit in rng.choice(256, int(rng.integers(0, 90)), replace=False):
            noisy[i, bit >> 3] ^= 1 << (bit & 7)
    dr[mono_right + perm] = noisy
    dup = rng.choice(n, n // 10, replace=False)                         # two left keypoints with one right partner:
    dl[mono_left + dup] = dl[mono_left + (dup + 1) % n]                 # mvRightToLeftMatch keeps the later one
    kl["x"][mono_left + dup] = kl["x"][mono_left + (dup + 1) % n] + 0.25
    kl["y"][mono_left + dup] = kl["y"][mono_left + (dup + 1) % n]
    return kl, dl, kr, dr


def _check_fused(lib, n, seed, mono_left, mono_right):
    R12, t12 = rig_pose()
    rig = matcher.kb8_rig(PRM1, PRM2, R12, t12)
    kl, dl, kr, dr = _fisheye_frames(n, seed, mono_left, mono_right)
    exp = oracle.stereo_fisheye_matches(rig, kl, dl, mono_left, kr, dr, mono_right, SIGMA2)
    got = matcher.compute_stereo_fisheye_matches(rig, kl, dl, mono_left, kr, dr, mono_right, SIGMA2, lib_path=lib)
    assert got["desc_matches"] == exp["desc_matches"] > n // 4          # the integer part (kNN + ratio) is exact
    border = exp["margin"] < MARGIN
    acc, eacc = got["depth"] > 0.0001, exp["depth"] > 0.0001
    assert (acc == eacc)[~border].all() and border.sum() <= max(1, n // 100)
    same = acc & eacc
    assert same.sum() > n // 10
    assert np.array_equal(got["left_to_right"][~border], exp["left_to_right"][~border])
    if not border.any():
        assert np.array_equal(got["right_to_left"], exp["right_to_left"]) and got["matches"] == exp["matches"]
    assert (np.abs(got["depth"][same] - exp["depth"][same]) <= REL * exp["depth"][same]).all()
    assert (np.abs(got["p3d"][same] - exp["p3d"][same]).max(1) <= REL * np.linalg.norm(exp["p3d"][same], axis=1)).all()
    assert (got["left_to_right"][:mono_left] == -1).all() and (got["depth"][~acc] == -1).all()
    # two left keypoints sharing a right partner: the later left index stays in mvRightToLeftMatch
    r2l = got["right_to_left"]
    for ir in np.nonzero(r2l >= 0)[0][:200]:
        assert r2l[ir] == np.nonzero(got["left_to_right"] == ir)[0].max()


def _check_edges(lib):
    R12, t12 = rig_pose()
    rig = matcher.kb8_rig(PRM1, PRM2, R12, t12)
    kl, dl, kr, dr = _fisheye_frames(40, 3, 5, 7)
    r = matcher.compute_stereo_fisheye_matches(rig, kl[:5], dl[:5], 5, kr, dr, 7, SIGMA2, lib_path=lib)        # empty lapping area
    assert r["matches"] == 0 and (r["left_to_right"] == -1).all() and (r["right_to_left"] == -1).all()
    r = matcher.compute_stereo_fisheye_matches(rig, kl, dl, 5, kr[:8], dr[:8], 7, SIGMA2, lib_path=lib)        # one right candidate:
    assert r["matches"] == 0 and r["desc_matches"] == 0                                                          # knnMatch gives < 2
    d, p = matcher.triangulate_matches(rig, np.zeros((0, 2)), np.zeros((0, 2)), np.zeros(0), np.zeros(0), lib_path=lib)
    assert len(d) == 0
    # the principal point unprojects to the optical axis (theta_d <= 1e-8 branch): identical rays, parallax gate rejects
    c1, c2 = PRM1[2:4][None], PRM2[2:4][None]
    d, _ = matcher.triangulate_matches(matcher.kb8_rig(PRM1, PRM2, np.eye(3), t12), c1, c2, [1.0], [1.0], lib_path=lib)
    assert d[0] == -1


@pytest.fixture(scope="module")
def emu(oracle_mod):
    return ge.build_emu()


def test_kb8_golden_emulation(emu, golden_dir):
    _check_golden(golden_dir, emu)


def test_kb8_fused_emulation(emu):
    _check_fused(emu, 300, 5, 20, 35)


def test_kb8_edges_emulation(emu):
    _check_edges(emu)


@pytest.mark.gpu
def test_kb8_golden_cuda(oracle_mod, golden_dir):
    _check_golden(golden_dir, None)


@pytest.mark.gpu
@pytest.mark.parametrize("n,seed,ml,mr", [(1500, 11, 500, 400), (4000, 12, 0, 0)])
def test_kb8_fused_cuda(oracle_mod, n, seed, ml, mr):
    _check_fused(None, n, seed, ml, mr)


@pytest.mark.gpu
def test_kb8_edges_cuda(oracle_mod):
    _check_edges(None)


@pytest.mark.gpu
def test_kb8_many_pairs_cuda(oracle_mod):
    R12, t12 = rig_pose()
    rig = matcher.kb8_rig(PRM1, PRM2, R12, t12)
    p1, p2, s1, s2 = make_matches(20000, 99)
    d, p = matcher.triangulate_matches(rig, p1, p2, s1, s2)
    od, op, m = oracle.kb8_triangulate(rig, p1, p2, s1, s2)
    _compare(d, p, od, op, m)
