#!/usr/bin/env python3
"""Generate the golden vectors that pin the oracle (run in the build container,
where the cv2 4.13 wheel imports; the outputs are committed).

    python tests/golden/make_golden.py

Writes tests/golden/primitives.npz and tests/golden/extract_<name>.npz.
Every output comes from real OpenCV code: cv2 primitives directly, or the
reference control flow transliterated on top of them (ref_cv2.RefExtractor).
The reference pins nothing for this path, and its pinned
OpenCV 4.2 binary is not available; these vectors are from cv2 4.13.0.
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from tests.golden.ref_cv2 import RefExtractor, compute_stereo_matches, load_pattern  # noqa: E402
from vi_slam_b200 import synthetic  # noqa: E402

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])

EXTRACT_CASES = {
    # name: (w, h, nfeatures, lapping, seed)
    "qvga": (320, 240, 500, (0, 0), 1),
    "vga_lap": (640, 480, 1000, (100, 300), 2),
    "kitti": (1241, 376, 2000, (0, 1000), 1000),
}


def primitives():
    rng = np.random.default_rng(42)
    out = {}
    src = synthetic.make_frame(5, 413, 251)
    out["resize_src"] = src
    out["resize_dst"] = cv2.resize(src, (344, 209), interpolation=cv2.INTER_LINEAR)
    src2 = rng.integers(0, 256, (97, 131), dtype=np.uint8)
    out["resize2_src"] = src2
    out["resize2_dst"] = cv2.resize(src2, (109, 81), interpolation=cv2.INTER_LINEAR)
    out["blur_src"] = src2
    out["blur_dst"] = cv2.GaussianBlur(src2.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    out["blur2_dst"] = cv2.GaussianBlur(src.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    fimg = synthetic.make_frame(9, 160, 120)
    out["fast_src"] = fimg
    for th in (7, 20):
        k = cv2.FastFeatureDetector_create(th, True).detect(fimg)
        out["fast_th%d" % th] = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in k], np.int32).reshape(-1, 3)
    y = np.concatenate([rng.integers(-3_000_000, 3_000_000, 4000), [0, 0, 5, -5, 0, 7, -7]]).astype(np.float32)
    x = np.concatenate([rng.integers(-3_000_000, 3_000_000, 4000), [0, 5, 0, 0, -5, 7, -7]]).astype(np.float32)
    out["atan_y"], out["atan_x"] = y, x
    out["atan_deg"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    q, t = synthetic.make_descriptors(300, 257, seed=3)
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)
    out["knn_q"], out["knn_t"] = q, t
    out["knn_idx"] = np.array([[a.trainIdx, b.trainIdx] for a, b in m], np.int32)
    out["knn_dist"] = np.array([[a.distance, b.distance] for a, b in m], np.float32).astype(np.int32)
    np.savez_compressed(os.path.join(HERE, "primitives.npz"), **out)
    print("primitives.npz written")


def extract_cases():
    pat = load_pattern(os.path.join(ROOT, "vi_slam_b200", "csrc", "orb_pattern.inc"))
    for name, (w, h, nf, lap, seed) in EXTRACT_CASES.items():
        img = synthetic.make_frame(seed, w, h)
        rx = RefExtractor(nf, 1.2, 8, 20, 7, pat)
        mono, kps, desc = rx.compute(img, lap)
        k = np.array([tuple(r) for r in kps], dtype=KP_DTYPE)
        np.savez_compressed(
            os.path.join(HERE, "extract_%s.npz" % name),
            image=img, params=np.array([nf, 8, 20, 7], np.int32), scale_factor=np.float32(1.2),
            lapping=np.array(lap, np.int32), mono=np.int32(mono), kps=k, desc=desc,
            n_candidates=np.array([len(c) for c in rx.cands], np.int32),
            level_shapes=np.array([p.shape for p in rx.pyr], np.int32),
            level_sums=np.array([int(p.astype(np.int64).sum()) for p in rx.pyr], np.int64),
        )
        print(name, "n =", len(k), "mono =", mono)


def stereo_case():
    """KITTI00-Stereo geometry: bf = 386.1448, fx = 718.856 -> b = bf / fx
    (reference config/KITTI00-Stereo.yaml, frame.cpp: mb = mbf / fx)."""
    pat = load_pattern(os.path.join(ROOT, "vi_slam_b200", "csrc", "orb_pattern.inc"))
    left, right = synthetic.make_stereo_pair(1000)
    out = {"left": left, "right": right}
    ex = {}
    for side, img in (("l", left), ("r", right)):
        rx = RefExtractor(2000, 1.2, 8, 20, 7, pat)
        mono, kps, desc = rx.compute(img, (0, 0))
        ex[side] = (rx, np.array([tuple(r) for r in kps], dtype=KP_DTYPE), desc)
        out["kps_" + side], out["desc_" + side] = ex[side][1], desc
    bf = np.float32(386.1448)
    b = np.float32(bf / np.float32(718.856))
    rl, kl, dl = ex["l"]
    rr, kr, dr = ex["r"]
    ur, dp = compute_stereo_matches(rl.pyr, rr.pyr, rl.scale, rl.inv_scale, kl, dl, kr, dr, bf, b)
    out["bf"], out["b"], out["u_right"], out["depth"] = bf, b, ur, dp
    # fisheye-style kNN-2 + 0.7 ratio (frame.cpp:1167-1170) through cv2.BFMatcher
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(dl, dr, k=2)
    good = [(p[0].queryIdx, p[0].trainIdx, int(p[0].distance)) for p in m
            if len(p) >= 2 and p[0].distance < p[1].distance * 0.7]
    out["ratio_matches"] = np.array(good, np.int32).reshape(-1, 3)
    np.savez_compressed(os.path.join(HERE, "stereo_kitti.npz"), **out)
    print("stereo: matched", int((dp > 0).sum()), "of", len(kl), "; ratio matches", len(good))


if __name__ == "__main__":
    cv2.setNumThreads(1)
    primitives()
    extract_cases()
    stereo_case()
