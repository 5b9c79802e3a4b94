"""Inputs of tests/test_ref_pin_matchers.py and tests/golden/make_ref_matcher_golden.py: stereo pairs and consecutive
frames on which the matcher rows are pinned to the reference's own function bodies (oracle/ref_compat/ref_slices.cpp)."""
from __future__ import annotations

import numpy as np

from vi_slam_b200 import synthetic

# name -> (seed, nfeatures, width, height, bf, fx)
STEREO = {
    "kitti": (41, 2000, 1241, 376, 386.1448, 718.856),
    "vga": (42, 1000, 640, 480, 120.0, 500.0),
}
# name -> (seed, nfeatures, width, height, window)
INI = {
    "kitti": (51, 2000, 1241, 376, 100),
    "qvga": (52, 500, 320, 240, 30),
}


def stereo_images(name):
    seed, _, w, h, _, _ = STEREO[name]
    return synthetic.make_stereo_pair(seed, w, h)


def stereo_inputs(name):
    """Reference and oracle extractors after compute() on the pair, and the reference's keypoints / descriptors."""
    import oracle
    from oracle import ref
    _, nf, _, _, bf, fx = STEREO[name]
    left, right = stereo_images(name)
    rl, rr = ref.RefExtractor(nf, monotone=True), ref.RefExtractor(nf, monotone=True)
    _, kl, dl = rl.compute(left, (0, 0))
    _, kr, dr = rr.compute(right, (0, 0))
    ol, orr = oracle.OracleExtractor(nf), oracle.OracleExtractor(nf)
    _, okl, odl = ol.compute(left, (0, 0))
    _, okr, odr = orr.compute(right, (0, 0))
    assert np.array_equal(dl, odl) and np.array_equal(dr, odr) and np.array_equal(kl, okl) and np.array_equal(kr, okr)
    return {"rl": rl, "rr": rr, "ol": ol, "orr": orr, "kl": kl, "dl": dl, "kr": kr, "dr": dr, "bf": bf,
            "b": float(np.float32(bf) / np.float32(fx))}


def ini_images(name):
    seed, _, w, h, _ = INI[name]
    f0 = synthetic.make_frame(seed, w, h)
    return f0, synthetic.make_next_frame(f0, seed + 1)


def ini_inputs(name, extractor=None):
    """(k1, d1, k2, d2, bounds, vbPrevMatched, windowSize) of two consecutive frames, extracted with lapping {0, 1000} as the
    monocular Frame constructor does (frame.cpp:289); vbPrevMatched starts as the first frame's keypoint positions
    (tracking.cpp:2290-2292)."""
    import oracle
    _, nf, w, h, win = INI[name]
    f0, f1 = ini_images(name)
    ex = extractor or oracle.OracleExtractor(nf)
    _, k1, d1 = ex.compute(f0, (0, 1000))
    k1, d1 = k1.copy(), d1.copy()
    _, k2, d2 = ex.compute(f1, (0, 1000))
    bounds = np.array([0, w, 0, h], np.float32)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    return k1, d1, k2.copy(), d2.copy(), bounds, prev, win


# ---- SearchByProjection(CurrentFrame, LastFrame): synthetic scenes --------------------------------------------------------
SF = np.float32(1.2) ** np.arange(8, dtype=np.float32)
# name -> (seed, mono, camera motion along z in metres, rig?, th, check_orientation, use mvuRight, use blocked)
PROJ = {
    "mono": (61, True, 0.05, False, 15.0, True, False, True),
    "stereo_forward": (62, False, 0.9, False, 7.0, True, True, True),
    "stereo_backward": (63, False, -0.9, False, 15.0, False, True, False),
    "rig": (64, True, 0.05, True, 15.0, True, False, True),
    "rig_forward": (65, False, 0.9, True, 10.0, True, False, False),
}


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    rxm = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    rym = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rzm = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return rzm @ rym @ rxm


def proj_scene(name):
    """A last frame with map points, a current frame a small motion later (one camera or a two-camera rig), keypoints of
    the current frame near the projections with perturbed copies of the map points' descriptors."""
    from oracle import KP_DTYPE
    seed, mono, dz, rig, th, chk, use_ur, use_bl = PROJ[name]
    rng = np.random.default_rng(seed)
    w, h = 640, 480
    cam = np.array([500, 500, 320, 240], np.float32)
    n1 = 1500
    # last camera at the origin; world points in its frustum
    z = rng.uniform(4, 30, n1)
    u, v = rng.uniform(0, w, n1), rng.uniform(0, h, n1)
    xw = np.stack([(u - cam[2]) / cam[0] * z, (v - cam[3]) / cam[1] * z, z], 1).astype(np.float32)
    xw[rng.random(n1) < 0.15, 0] = np.nan                        # no map point
    outlier = (rng.random(n1) < 0.05).astype(np.uint8)
    observed = (rng.random(n1) < 0.5).astype(np.uint8)
    k_last = np.zeros(n1, KP_DTYPE)
    k_last["x"], k_last["y"] = u, v
    k_last["octave"] = rng.choice(8, n1, p=[.25, .2, .15, .12, .1, .08, .06, .04])
    k_last["angle"] = rng.uniform(0, 360, n1)
    k_last["size"] = 31 * SF[k_last["octave"]]
    desc_mp = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    tcw_last = np.eye(4, dtype=np.float32)
    r = _rot(*(rng.uniform(-0.01, 0.01, 3)))
    cpos = np.array([rng.uniform(-0.05, 0.05), rng.uniform(-0.05, 0.05), dz])       # current camera centre in the world
    tcw_cur = np.eye(4, dtype=np.float32)
    tcw_cur[:3, :3] = r
    tcw_cur[:3, 3] = -r @ cpos
    trl = None
    if rig:
        rr = _rot(0.0, -0.15, 0.0)
        trl = np.concatenate([rr, (rr @ np.array([[-0.12], [0.0], [0.0]]))], 1).astype(np.float32)
    # current keypoints: near the projections (double precision here; the reference's float projection is exported by the pin)
    xc = (tcw_cur[:3, :3].astype(np.float64) @ np.nan_to_num(xw, nan=0.0).T.astype(np.float64)).T + tcw_cur[:3, 3]
    ks, ds = [], []
    nleft2 = -1

    def spawn(x3, frac):
        zc = x3[:, 2]
        uu, vv = cam[0] * x3[:, 0] / zc + cam[2], cam[1] * x3[:, 1] / zc + cam[3]
        ok = np.isfinite(xw[:, 0]) & (zc > 0.5) & (uu > 4) & (uu < w - 4) & (vv > 4) & (vv < h - 4) & (rng.random(n1) < frac)
        idx = np.flatnonzero(ok)
        k = np.zeros(len(idx), KP_DTYPE)
        k["x"] = uu[idx] + rng.normal(0, 2.0, len(idx))
        k["y"] = vv[idx] + rng.normal(0, 2.0, len(idx))
        k["octave"] = np.clip(k_last["octave"][idx] + rng.integers(-1, 2, len(idx)), 0, 7)
        k["angle"] = np.where(rng.random(len(idx)) < 0.85, (k_last["angle"][idx] + rng.normal(0, 4, len(idx))) % 360,
                              rng.uniform(0, 360, len(idx)))
        d = desc_mp[idx].copy()
        flips = rng.integers(0, 256, (len(idx), 14))
        for j in range(14):                                           # up to 14 flipped bits
            d[np.arange(len(idx)), flips[:, j] // 8] ^= (1 << (flips[:, j] % 8)).astype(np.uint8)
        # clutter: unrelated keypoints
        nc = len(idx) // 2
        kc = np.zeros(nc, KP_DTYPE)
        kc["x"], kc["y"] = rng.uniform(0, w, nc), rng.uniform(0, h, nc)
        kc["octave"] = rng.integers(0, 8, nc)
        kc["angle"] = rng.uniform(0, 360, nc)
        k = np.concatenate([k, kc])
        d = np.concatenate([d, rng.integers(0, 256, (nc, 32), dtype=np.uint8)])
        p = rng.permutation(len(k))
        return k[p], d[p]

    kl2, dl2 = spawn(xc, 0.8)
    if rig:
        xr = (trl[:, :3].astype(np.float64) @ xc.T).T + trl[:, 3]
        kr2, dr2 = spawn(xr, 0.7)
        nleft2 = len(kl2)
        k2, d2 = np.concatenate([kl2, kr2]), np.concatenate([dl2, dr2])
    else:
        k2, d2 = kl2, dl2
    k2["size"] = 31 * SF[k2["octave"]]
    bf, b = 40.0, 0.08
    u_right2 = None
    if use_ur:
        u_right2 = np.where(rng.random(len(k2)) < 0.6, k2["x"] - bf / rng.uniform(4, 30, len(k2)), -1).astype(np.float32)
    blocked2 = (rng.random(len(k2)) < 0.08).astype(np.uint8) if use_bl else None
    return {"k_last": k_last, "xw": xw, "desc_mp": desc_mp, "outlier": outlier, "observed": observed, "k2": k2, "d2": d2,
            "u_right2": u_right2, "blocked2": blocked2, "nleft2": nleft2, "bounds": np.array([0, w, 0, h], np.float32),
            "tcw_cur": tcw_cur, "tcw_last": tcw_last, "trl": trl, "cam": cam, "sf": SF, "th": th, "mono": mono, "bf": bf, "b": b,
            "check": chk}
