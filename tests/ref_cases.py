"""Inputs on which the oracle and the CUDA path are pinned to the reference's own FExtractor object code
(oracle/_ref/liborb_ref.so = /root/reference/src/geometry/fextractor.cpp compiled unmodified, oracle/ref.mk)."""
from __future__ import annotations

import numpy as np

from vi_slam_b200 import synthetic


def checkerboard(w, h, sq=24):
    yy, xx = np.mgrid[0:h, 0:w]
    return (((xx // sq + yy // sq) & 1) * 200 + 30).astype(np.uint8)


def noise(seed, w, h):
    return np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)


# name -> (image factory, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, lapping)
CASES = {
    "kitti_mono": (lambda: synthetic.make_frame(1000), 2000, 1.2, 8, 20, 7, (0, 1000)),        # config 1 (frame.cpp:289)
    "kitti_stereo": (lambda: synthetic.make_frame(1001), 2000, 1.2, 8, 20, 7, (0, 0)),         # config 2 (frame.cpp:107)
    "kitti_ini_10000": (lambda: synthetic.make_frame(1002), 10000, 1.2, 8, 20, 7, (0, 1000)),  # tracking.cpp:1092
    "zed_4000": (lambda: synthetic.make_frame(1003, 1280, 720), 4000, 1.2, 8, 20, 7, (0, 0)),  # config 3
    "lap_100_250": (lambda: synthetic.make_frame(2, 400, 200), 700, 1.2, 8, 20, 7, (100, 250)),
    "qvga": (lambda: synthetic.make_frame(1, 320, 240), 500, 1.2, 8, 20, 7, (0, 0)),
    # (levels narrower than 62 px make the reference divide by zero in its cell grid, fextractor.cpp:775-778: not a case)
    "odd_259x233": (lambda: synthetic.make_frame(3, 259, 233), 300, 1.2, 8, 20, 7, (0, 0)),
    "black": (lambda: np.zeros((240, 320), np.uint8), 500, 1.2, 8, 20, 7, (0, 0)),             # 0 kp -> descriptors.release()
    "noise": (lambda: noise(5, 640, 360), 1500, 1.2, 8, 20, 7, (0, 0)),
    "noise_kitti": (lambda: noise(6, 1241, 376), 2000, 1.2, 8, 20, 7, (0, 0)),
    "checkerboard": (lambda: checkerboard(798, 798), 1000, 1.2, 8, 20, 7, (0, 0)),             # test/images/chessboard_798_798.png shape
    "low_texture": (lambda: (synthetic.make_frame(9, 640, 360).astype(np.float32) * 0.08 + 100).astype(np.uint8), 1000, 1.2, 8, 20, 7, (0, 0)),
    "scale_1p5_4lv": (lambda: synthetic.make_frame(11, 640, 480), 1200, 1.5, 4, 20, 7, (0, 0)),
    "th_40_12": (lambda: synthetic.make_frame(12, 640, 360), 1000, 1.2, 8, 40, 12, (0, 0)),
}

# the ones whose reference outputs are also committed as fixtures (tests/golden/ref_*.npz)
GOLDEN = ["kitti_mono", "kitti_ini_10000", "zed_4000", "lap_100_250", "noise", "checkerboard", "scale_1p5_4lv"]
