"""Host-side mirror of the per-frame geometry helpers of the reference's Frame over the C ABI.

* ``undistort_keypoints``   Frame::UndistortKeyPoints   (src/datastructures/frame.cpp:758-791)
* ``undistort_points``      the cv::undistortPoints(pts, pts, mK, mDistCoef, Mat(), mK) call inside it (frame.cpp:779)
* ``compute_image_bounds``  Frame::ComputeImageBounds   (frame.cpp:793-821)
All compute happens in libviorb.so (CUDA); no CPU path.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .extractor import KP_DTYPE, _check


def _cam(K, dist):
    K = np.asarray(K, np.float32)
    cam = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], np.float32) if K.ndim == 2 else np.ascontiguousarray(K, np.float32)
    d = np.ascontiguousarray(dist, np.float32).ravel()
    return cam, d


def undistort_keypoints(kps, K, dist, device=0, lib_path=None):
    lib = _lib.load(lib_path)
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    cam, d = _cam(K, dist)
    out = np.empty_like(kps)
    _check(lib, lib.viorb_undistort_keypoints(kps.ctypes.data, len(kps), cam.ctypes.data_as(_lib.f32p), d.ctypes.data_as(_lib.f32p),
                                              len(d), out.ctypes.data, device))
    return out


def undistort_points(xy, K, dist, device=0, lib_path=None):
    lib = _lib.load(lib_path)
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    cam, d = _cam(K, dist)
    out = np.empty_like(xy)
    _check(lib, lib.viorb_undistort_points(xy.ctypes.data_as(_lib.f32p), len(xy), cam.ctypes.data_as(_lib.f32p),
                                           d.ctypes.data_as(_lib.f32p), len(d), out.ctypes.data_as(_lib.f32p), device))
    return out


def compute_image_bounds(cols, rows, K, dist, device=0, lib_path=None):
    """-> np.float32[4] = (mnMinX, mnMaxX, mnMinY, mnMaxY), the ``bounds`` argument of the windowed matchers."""
    lib = _lib.load(lib_path)
    cam, d = _cam(K, dist)
    out = np.zeros(4, np.float32)
    _check(lib, lib.viorb_image_bounds(int(cols), int(rows), cam.ctypes.data_as(_lib.f32p), d.ctypes.data_as(_lib.f32p), len(d),
                                       out.ctypes.data, device))
    return out
