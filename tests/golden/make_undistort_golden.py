"""Generates tests/golden/undistort.npz: cv2.undistortPoints (OpenCV 4.13 in this container) called exactly as
Frame::UndistortKeyPoints / Frame::ComputeImageBounds call it (src/datastructures/frame.cpp:779, 806):
undistortPoints(mat, mat, mK, mDistCoef, Mat(), mK) with CV_32F K and distortion, default 5 iterations.
Cameras: the reference's config/zed_camera.yaml cam0 / cam1 (radtan, 1280x720), an EuRoC-like strong barrel
distortion (4 and 5 coefficients) and a pathological one that drives icdist negative (OpenCV regression 14583 branch).
usage: python tests/golden/make_undistort_golden.py"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

CAMERAS = {
    "zed0": ((669.2387507702717, 669.6062139634853, 647.4136147885813, 348.40757297218505),
             (0.0018645604002542789, -0.009206711115906055, -0.001490842343490958, 0.0047045781898403), (1280, 720)),
    "zed1": ((669.7077049723667, 669.7830132578491, 648.2500643003343, 348.45508924255745),
             (-0.0011120079644645446, -0.006192062533471337, -0.0011416874899672696, 0.004836945809987094), (1280, 720)),
    "barrel4": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05), (752, 480)),
    "barrel5": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0123), (752, 480)),
    "negative": ((300.0, 300.0, 320.0, 240.0), (-1.9, 0.3, 0.01, -0.02), (640, 480)),
}


def undistort(xy, cam, dist):
    K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
    D = np.array(dist, np.float32).reshape(-1, 1)
    mat = np.ascontiguousarray(xy, np.float32).reshape(-1, 1, 2)
    return cv2.undistortPoints(mat, K, D, None, K).reshape(-1, 2)


def main():
    rng = np.random.default_rng(2024)
    out = {"version": np.array(cv2.__version__)}
    for name, (cam, dist, (w, h)) in CAMERAS.items():
        # keypoint-like coordinates: level-0 integers and level-l integers times 1.2^l (what k_emit produces)
        lv = rng.integers(0, 8, 4000)
        sc = np.float32(1.2) ** lv.astype(np.float32)
        xy = np.stack([np.floor(rng.random(4000) * w / sc), np.floor(rng.random(4000) * h / sc)], 1).astype(np.float32)
        xy = (xy * sc[:, None].astype(np.float32)).astype(np.float32)
        xy[:4] = [[0, 0], [w, 0], [0, h], [w, h]]
        und = undistort(xy, cam, dist)
        b = np.array([min(und[0, 0], und[2, 0]), max(und[1, 0], und[3, 0]), min(und[0, 1], und[1, 1]), max(und[2, 1], und[3, 1])],
                     np.float32)
        out[name + "_cam"] = np.array(cam, np.float32)
        out[name + "_dist"] = np.array(dist, np.float32)
        out[name + "_size"] = np.array([w, h], np.int32)
        out[name + "_xy"] = xy
        out[name + "_und"] = und
        out[name + "_bounds"] = b
    np.savez_compressed(os.path.join(HERE, "undistort.npz"), **out)
    print("wrote undistort.npz", {k: v.shape for k, v in out.items() if k.endswith("_und")})


if __name__ == "__main__":
    main()
