"""Generates tests/golden/window_match.npz with tests/golden/ref_match.py (see its header).
Inputs: keypoints / descriptors of two seeded synthetic frames (extracted with the oracle extractor; they are just
input DATA and are stored in the file), outputs: the transliteration's results.
usage: python tests/golden/make_window_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle                                         # noqa: E402
from tests.golden import ref_match                    # noqa: E402
from vi_slam_b200 import synthetic                    # noqa: E402


def main():
    w, h = 640, 360
    f1 = synthetic.make_frame(901, w, h)
    f2 = synthetic.make_next_frame(f1, 901)
    ex = oracle.OracleExtractor(1500, 1.2, 8, 20, 7)
    _, k1, d1 = ex.compute(f1, (0, 0))
    _, k2, d2 = ex.compute(f2, (0, 0))
    bounds = np.array([0, w, 0, h], np.float32)
    # make the data ambiguous so that every rule fires: random bit flips (distances around TH_LOW / the ratio),
    # duplicated descriptors (ties, one-to-one conflicts), a share of keypoints with a rotated angle
    rng0 = np.random.default_rng(77)
    d2 = d2.copy()
    k2 = k2.copy()
    for i in range(len(d2)):
        nflip = int(rng0.integers(0, 70))
        bits = rng0.choice(256, nflip, replace=False)
        for b in bits:
            d2[i, b >> 3] ^= 1 << (b & 7)
    dup = rng0.choice(len(d2), len(d2) // 6, replace=False)
    d2[dup] = d2[(dup + 1) % len(d2)]
    rot = rng0.random(len(k2))
    k2["angle"] = np.where(rot < 0.2, (k2["angle"] + 95) % 360, np.where(rot < 0.3, (k2["angle"] + 200) % 360, k2["angle"])).astype(np.float32)
    out = {"k1": k1, "d1": d1, "k2": k2, "d2": d2, "bounds": bounds}
    rng = np.random.default_rng(5)
    pm = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    for tag, win, ratio, chk in (("a", 100, 0.9, True), ("b", 40, 0.9, False), ("c", 100, 0.6, True)):
        n, m12, pm2 = ref_match.search_for_initialization(k1, d1, k2, d2, bounds, pm, win, ratio, chk)
        out.update({f"ini_{tag}_args": np.array([win, ratio, chk], np.float64), f"ini_{tag}_n": n, f"ini_{tag}_m12": m12,
                    f"ini_{tag}_pm": pm2})
    # projection search: "last frame" = frame 1 keypoints predicted into frame 2 with a small error
    n1 = len(k1)
    uv = np.stack([k1["x"] + 7 + rng.normal(0, 2, n1), k1["y"] + 2 + rng.normal(0, 2, n1)], 1).astype(np.float32)
    uv[::17] += 2000                                                     # some projections far outside the grid
    invz = rng.uniform(0.02, 0.5, n1).astype(np.float32)
    valid = (rng.random(n1) < 0.8).astype(np.uint8)
    observed = (rng.random(n1) < 0.7).astype(np.uint8)
    blocked2 = (rng.random(len(k2)) < 0.1).astype(np.uint8)
    uright2 = np.where(rng.random(len(k2)) < 0.5, k2["x"] - 386.1448 * 0.1, -1).astype(np.float32)
    sf = (1.2 ** np.arange(8)).astype(np.float32)
    out.update({"uv": uv, "invz": invz, "valid": valid, "observed": observed, "blocked2": blocked2, "uright2": uright2, "sf": sf})
    for tag, th, direction, chk, stereo in (("a", 15.0, 0, True, False), ("b", 7.0, 1, True, True), ("c", 15.0, 2, False, True),
                                            ("d", 30.0, 0, True, True)):
        n, asg = ref_match.search_by_projection(uv, invz, k1["octave"], k1["angle"], d1, valid, observed, k2, d2,
                                                uright2 if stereo else None, blocked2, bounds, sf, th, 386.1448, direction, chk)
        out.update({f"proj_{tag}_args": np.array([th, direction, chk, stereo], np.float64), f"proj_{tag}_n": n, f"proj_{tag}_asg": asg})
    np.savez_compressed(os.path.join(HERE, "window_match.npz"), **out)
    print({k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items() if k.endswith("_n")})


if __name__ == "__main__":
    main()
