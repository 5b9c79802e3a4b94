"""Frame::UndistortKeyPoints / ComputeImageBounds (src/datastructures/frame.cpp:758-821), SURVEY.md section 8 row f-2.
The library (CUDA build under -m gpu, the g++ emulation build of the same sources otherwise) against
cv2.undistortPoints golden vectors (tests/golden/undistort.npz, OpenCV 4.13) and against the oracle.

Tolerance: none.  The kernel evaluates OpenCV's double-precision iteration operation by operation with
round-to-nearest intrinsics (no FMA contraction), so the float results are compared bit for bit."""
import os

import numpy as np
import pytest

import __graft_entry__ as ge
import oracle
from vi_slam_b200 import FExtractor, frame, synthetic

CAMERAS = ("zed0", "zed1", "barrel4", "barrel5", "negative")


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _golden(golden_dir):
    return np.load(os.path.join(golden_dir, "undistort.npz"))


def test_oracle_matches_cv2_golden(oracle_mod, golden_dir):
    g = _golden(golden_dir)
    for name in CAMERAS:
        cam, dist = g[name + "_cam"], g[name + "_dist"]
        assert np.array_equal(_bits(oracle.undistort_points(g[name + "_xy"], cam, dist)), _bits(g[name + "_und"])), name
        w, h = g[name + "_size"]
        assert np.array_equal(_bits(oracle.image_bounds(w, h, cam, dist)), _bits(g[name + "_bounds"])), name
    # the regression-14583 branch (icdist < 0) is exercised: the corners of that camera come back unchanged
    assert np.array_equal(g["negative_und"][:4], g["negative_xy"][:4])


def _check_golden(golden_dir, lib):
    g = _golden(golden_dir)
    for name in CAMERAS:
        cam, dist = g[name + "_cam"], g[name + "_dist"]
        xy = g[name + "_xy"]
        assert np.array_equal(_bits(frame.undistort_points(xy, cam, dist, lib_path=lib)), _bits(g[name + "_und"])), name
        w, h = g[name + "_size"]
        assert np.array_equal(_bits(frame.compute_image_bounds(w, h, cam, dist, lib_path=lib)), _bits(g[name + "_bounds"])), name
        kps = np.zeros(len(xy), oracle.KP_DTYPE)
        kps["x"], kps["y"] = xy[:, 0], xy[:, 1]
        kps["size"], kps["angle"], kps["response"] = 31.0, np.arange(len(xy)) % 360, 20 + np.arange(len(xy)) % 50
        kps["octave"], kps["class_id"] = np.arange(len(xy)) % 8, -1
        un = frame.undistort_keypoints(kps, cam, dist, lib_path=lib)
        assert np.array_equal(_bits(np.stack([un["x"], un["y"]], 1)), _bits(g[name + "_und"]))
        for fld in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(un[fld], kps[fld])
        assert un.tobytes() == oracle.undistort_keypoints(kps, cam, dist).tobytes()


def _check_rules(lib):
    cam = np.array([500, 510, 320, 240], np.float32)
    kps = np.zeros(7, oracle.KP_DTYPE)
    kps["x"], kps["y"] = np.arange(7) * 90.5, np.arange(7) * 60.25
    # k1 == 0: keypoints are copied even when the other coefficients are not zero (frame.cpp:762-766) ...
    d0 = np.array([0, 0.1, 0.01, 0.02], np.float32)
    assert frame.undistort_keypoints(kps, cam, d0, lib_path=lib).tobytes() == kps.tobytes()
    assert np.array_equal(frame.compute_image_bounds(640, 480, cam, d0, lib_path=lib), np.array([0, 640, 0, 480], np.float32))
    # ... while the bare cv::undistortPoints call still applies them
    xy = np.stack([kps["x"], kps["y"]], 1)
    assert np.array_equal(_bits(frame.undistort_points(xy, cam, d0, lib_path=lib)), _bits(oracle.undistort_points(xy, cam, d0)))
    assert not np.array_equal(frame.undistort_points(xy, cam, d0, lib_path=lib), xy)
    # empty input, 3x3 K argument
    assert len(frame.undistort_keypoints(kps[:0], cam, d0, lib_path=lib)) == 0
    K = np.array([[500, 0, 320], [0, 510, 240], [0, 0, 1]], np.float32)
    d = np.array([-0.2, 0.05, 0.001, -0.002, 0.01], np.float32)
    assert np.array_equal(_bits(frame.undistort_points(xy, K, d, lib_path=lib)), _bits(oracle.undistort_points(xy, cam, d)))
    with pytest.raises(Exception):
        frame.undistort_points(xy, cam, d[:3], lib_path=lib)          # 3 coefficients: not a radtan set the reference can hold


def _check_fused(lib, w, h, nf, seeds):
    """The extractor epilogue's ukeypoints_ == UndistortKeyPoints(oracle keypoints), frame by frame."""
    cam = np.array([669.2387507702717, 669.6062139634853, 647.4136147885813 * w / 1280, 348.40757297218505 * h / 720], np.float32)
    dist = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32)
    imgs = np.stack([synthetic.make_frame(s, w, h) for s in seeds])
    ex = FExtractor(nf, 1.2, 8, 20, 7, max_size=(w, h), max_batch=len(seeds), lib_path=lib)
    ox = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    ex.set_camera(cam, dist)
    ex.compute_batch(imgs, (0, 0))
    for i in range(len(seeds)):
        _, kps, _ = ex.fetch(i)
        _, okps, _ = ox.compute(imgs[i], (0, 0))
        un = ex.fetch_undistorted(i)
        assert len(kps) == len(okps) > 50
        assert un.tobytes() == oracle.undistort_keypoints(okps, cam, dist).tobytes()
        assert np.abs(un["x"] - kps["x"]).max() > 0.5                  # it did move the points
    ex.set_camera(cam, np.array([0, 0, 0, 0], np.float32))            # k1 == 0: copy
    ex.compute_batch(imgs[:1], (0, 0))
    assert ex.fetch_undistorted(0).tobytes() == ex.fetch(0)[1].tobytes()
    ex.set_camera(None, None)                                           # off again: fetching is an error, not stale data
    ex.compute_batch(imgs[:1], (0, 0))
    with pytest.raises(Exception):
        ex.fetch_undistorted(0)
    ex.close()


@pytest.fixture(scope="module")
def emu(oracle_mod):
    return ge.build_emu()


def test_undistort_golden_emulation(emu, golden_dir):
    _check_golden(golden_dir, emu)


def test_undistort_rules_emulation(emu):
    _check_rules(emu)


def test_undistort_fused_emulation(emu):
    _check_fused(emu, 320, 240, 500, (3, 4))


@pytest.mark.gpu
def test_undistort_golden_cuda(oracle_mod, golden_dir):
    _check_golden(golden_dir, None)


@pytest.mark.gpu
def test_undistort_rules_cuda(oracle_mod):
    _check_rules(None)


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,nf", [(1280, 720, 4000), (1241, 376, 2000)])
def test_undistort_fused_cuda(oracle_mod, w, h, nf):
    _check_fused(None, w, h, nf, (31, 32, 33))


@pytest.mark.gpu
def test_undistort_large_random_cuda(oracle_mod):
    """One million points over a wide range (inside and far outside the image) against the oracle."""
    rng = np.random.default_rng(8)
    xy = (rng.random((1_000_000, 2)) * [3000, 2000] - [800, 600]).astype(np.float32)
    cam = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
    for dist in ([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], [0.1, -0.3, 0.002, 0.001, 0.05]):
        d = np.array(dist, np.float32)
        assert np.array_equal(_bits(frame.undistort_points(xy, cam, d)), _bits(oracle.undistort_points(xy, cam, d)))
