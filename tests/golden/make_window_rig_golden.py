"""Generates tests/golden/window_rig.npz with the transliteration tests/golden/ref_match.py::search_by_projection_rig
(fmatcher.cpp:2471-2688 with CurrentFrame.Nleft != -1; the right-camera block is :2593-2660).
Inputs (stored in the file): keypoints / descriptors of a seeded synthetic two-camera frame pair extracted with the
oracle extractor, perturbed so every rule fires.  usage: python tests/golden/make_window_rig_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle                                         # noqa: E402
from tests.golden import ref_match                    # noqa: E402
from vi_slam_b200 import synthetic                    # noqa: E402


def rig_case(seed, w=640, h=360, nf=1200):
    """-> dict of inputs for a two-camera SearchByProjection call (LastFrame -> CurrentFrame)."""
    l1, r1 = synthetic.make_stereo_pair(seed, w, h)
    l2, r2 = synthetic.make_next_frame(l1, seed), synthetic.make_next_frame(r1, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kl1, dl1 = ex.compute(l1, (0, 0))
    _, kr1, dr1 = ex.compute(r1, (0, 0))
    _, kl2, dl2 = ex.compute(l2, (0, 0))
    _, kr2, dr2 = ex.compute(r2, (0, 0))
    rng = np.random.default_rng(seed + 5)
    # LastFrame: N = Nleft + Nright sources (keypoints_ then keypointsRight_), each with a map point descriptor
    ks = np.concatenate([kl1, kr1])
    ds = np.concatenate([dl1, dr1])
    n1 = len(ks)
    # CurrentFrame: keypoints_ then keypointsRight_
    k2 = np.concatenate([kl2, kr2])
    d2 = np.concatenate([dl2, dr2]).copy()
    for i in range(len(d2)):                               # spread distances around TH_HIGH
        nflip = int(rng.integers(0, 90))
        for b in rng.choice(256, nflip, replace=False):
            d2[i, b >> 3] ^= 1 << (b & 7)
    rot = rng.random(len(k2))
    k2 = k2.copy()
    k2["angle"] = np.where(rot < 0.25, (k2["angle"] + 120) % 360, k2["angle"]).astype(np.float32)
    uv = np.stack([ks["x"] + rng.normal(0, 3, n1), ks["y"] + rng.normal(0, 3, n1)], 1).astype(np.float32)
    uvr = np.stack([ks["x"] - 20 + rng.normal(0, 3, n1), ks["y"] + rng.normal(0, 3, n1)], 1).astype(np.float32)
    uv[::13] += 3000                                       # left window off the grid: the right pass is skipped too
    uvr[::11] -= 3000                                      # right window off the grid only
    return {"uv": uv, "uv_right": uvr, "invz": rng.uniform(0.02, 0.5, n1).astype(np.float32), "octave": ks["octave"].astype(np.int32),
            "angle": ks["angle"].astype(np.float32), "dmp": ds, "valid": (rng.random(n1) < 0.8).astype(np.uint8),
            "observed": (rng.random(n1) < 0.7).astype(np.uint8), "k2": k2, "d2": d2, "nleft2": np.int32(len(kl2)),
            "blocked2": (rng.random(len(k2)) < 0.1).astype(np.uint8), "bounds": np.array([0, w, 0, h], np.float32),
            "sf": (np.float32(1.2) ** np.arange(8)).astype(np.float32)}


def main():
    c = rig_case(903)
    out = dict(c)
    for tag, th, direction, chk, use_blocked in (("a", 15.0, 0, True, True), ("b", 7.0, 1, True, False), ("c", 15.0, 2, False, True),
                                                 ("d", 30.0, 0, True, True)):
        n, asg = ref_match.search_by_projection_rig(c["uv"], c["uv_right"], c["invz"], c["octave"], c["angle"], c["dmp"], c["valid"],
                                                    c["observed"], c["k2"], c["d2"], int(c["nleft2"]), c["blocked2"] if use_blocked else None,
                                                    c["bounds"], c["sf"], th, direction, chk)
        out.update({f"rig_{tag}_args": np.array([th, direction, chk, use_blocked], np.float64), f"rig_{tag}_n": n, f"rig_{tag}_asg": asg})
        print(tag, n, int((asg[:int(c["nleft2"])] >= 0).sum()), int((asg[int(c["nleft2"]):] >= 0).sum()), int((asg == -2).sum()))
    np.savez_compressed(os.path.join(HERE, "window_rig.npz"), **out)


if __name__ == "__main__":
    main()
