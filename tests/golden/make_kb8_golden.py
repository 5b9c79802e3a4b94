"""Generates tests/golden/kb8.npz with tests/golden/ref_kb8.py (real cv::SVD through cv2.SVDecomp).
Rig: a TUM-VI-like pair of 512x512 Kannala-Brandt cameras, 10 cm baseline with a small relative rotation.  Matches:
projections of random 3D points (0.3 .. 30 m) with pixel noise, plus outliers, far (low-parallax) points and points
behind a camera, so that every rejection branch of TriangulateMatches fires.
usage: python tests/golden/make_kb8_golden.py"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_kb8                      # noqa: E402

PRM1 = np.array([190.97847715128717, 190.9733070521226, 254.93170605935475, 256.8974428996504,
                 0.0034823894022493434, 0.0007150348452162257, -0.0020532361418706202, 0.00020293673591811182], np.float32)
PRM2 = np.array([190.44236969414825, 190.4344384721956, 252.59949716835982, 254.91723064636983,
                 0.0034003170790442797, 0.001766278153469831, -0.00266312569781606, 0.0003299517423931039], np.float32)


def rig_pose():
    rvec = np.array([0.004, -0.012, 0.007])
    R12, _ = cv2.Rodrigues(rvec)
    return R12.astype(np.float32), np.array([0.1009, 0.0019, -0.0011], np.float32)


def make_matches(n, seed):
    rng = np.random.default_rng(seed)
    R12, t12 = rig_pose()
    R21 = R12.T.astype(np.float64)
    t21 = -R21 @ t12.astype(np.float64)
    d = np.exp(rng.uniform(np.log(0.3), np.log(30.0), n))
    d[rng.random(n) < 0.1] *= 30                                      # far: parallax gate
    ang = rng.uniform(0, 2 * np.pi, n)
    th = rng.uniform(0, 1.2, n)
    dirs = np.stack([np.sin(th) * np.cos(ang), np.sin(th) * np.sin(ang), np.cos(th)], 1)
    X1 = dirs * d[:, None]
    X1[rng.random(n) < 0.03, 2] *= -1                                 # behind camera 1
    X2 = X1 @ R21.T + t21
    p1 = np.array([ref_kb8.project(x, PRM1) for x in X1], np.float32)
    p2 = np.array([ref_kb8.project(x, PRM2) for x in X2], np.float32)
    noise = rng.normal(0, 0.4, (n, 4))
    big = rng.random(n) < 0.15
    noise[big] *= 8                                                   # outliers: reprojection gates
    p1 += noise[:, :2].astype(np.float32)
    p2 += noise[:, 2:].astype(np.float32)
    lv = rng.integers(0, 8, (n, 2))
    sig = (np.float32(1.2) ** lv.astype(np.float32)) ** 2
    return p1, p2, sig[:, 0].astype(np.float32), sig[:, 1].astype(np.float32)


def main():
    R12, t12 = rig_pose()
    p1, p2, s1, s2 = make_matches(3000, 7)
    depth = np.zeros(len(p1), np.float32)
    p3d = np.zeros((len(p1), 3), np.float32)
    for i in range(len(p1)):
        z, x = ref_kb8.triangulate_matches(PRM1, PRM2, p1[i], p2[i], R12, t12, s1[i], s2[i])
        depth[i] = z
        if x is not None:
            p3d[i] = x
    print("accepted", int((depth > 0.0001).sum()), "of", len(depth))
    np.savez_compressed(os.path.join(HERE, "kb8.npz"), prm1=PRM1, prm2=PRM2, R12=R12, t12=t12, pt1=p1, pt2=p2, sigma1=s1, sigma2=s2,
                        depth=depth, p3d=p3d, version=np.array(cv2.__version__))


if __name__ == "__main__":
    main()
