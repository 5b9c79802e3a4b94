"""Windowed (grid) matchers, SURVEY.md section 8 row a-11: Frame::AssignFeaturesToGrid / GetFeaturesInArea,
FMatcher::SearchForInitialization, FMatcher::SearchByProjection(CurrentFrame, LastFrame).  The library (CUDA build
under -m gpu, the g++ emulation build of the same sources otherwise) against the golden vectors of
tests/golden/window_match.npz and against the oracle on seeded inputs.  Everything is integer: bit-exact."""
import os

import numpy as np
import pytest

import __graft_entry__ as ge
import oracle
from vi_slam_b200 import matcher, synthetic

SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


def _golden(golden_dir):
    return np.load(os.path.join(golden_dir, "window_match.npz"))


def _check_golden(golden_dir, lib):
    g = _golden(golden_dir)
    k1, d1, k2, d2, b = g["k1"], g["d1"], g["k2"], g["d2"], g["bounds"]
    pm = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    for tag in "abc":
        win, ratio, chk = g[f"ini_{tag}_args"]
        n, m12, pm2 = matcher.search_for_initialization(k1, d1, k2, d2, b, pm, int(win), ratio, bool(chk), lib_path=lib)
        assert n == int(g[f"ini_{tag}_n"]) and np.array_equal(m12, g[f"ini_{tag}_m12"])
        assert np.array_equal(pm2.view(np.uint32), g[f"ini_{tag}_pm"].view(np.uint32))
    for tag in "abcd":
        th, direction, chk, stereo = g[f"proj_{tag}_args"]
        n, asg = matcher.search_by_projection(g["uv"], g["invz"], k1["octave"], k1["angle"], d1, g["valid"], g["observed"], k2, d2,
                                              g["uright2"] if stereo else None, g["blocked2"], b, g["sf"], th, 386.1448,
                                              int(direction), bool(chk), lib_path=lib)
        assert n == int(g[f"proj_{tag}_n"]) and np.array_equal(asg, g[f"proj_{tag}_asg"])


def _frames(seed, w, h, nf):
    f1 = synthetic.make_frame(seed, w, h)
    f2 = synthetic.make_next_frame(f1, seed)
    ox = oracle.OracleExtractor(nf)
    _, k1, d1 = ox.compute(f1)
    _, k2, d2 = ox.compute(f2)
    rng = np.random.default_rng(seed)
    d2 = d2.copy()
    flips = rng.integers(0, 60, len(d2))
    for i, nfl in enumerate(flips):                                   # spread the distances around TH_LOW / the ratio
        for bit in rng.choice(256, int(nfl), replace=False):
            d2[i, bit >> 3] ^= 1 << (bit & 7)
    return k1, d1, k2, d2, np.array([0, w, 0, h], np.float32), rng


def _check_vs_oracle(lib, seed, w, h, nf):
    k1, d1, k2, d2, b, rng = _frames(seed, w, h, nf)
    # GetFeaturesInArea: windows inside, across and outside the grid, all level conventions
    for x, y, r, lo, hi in ((w / 2, h / 2, 40.0, -1, -1), (3.0, 2.0, 25.0, 0, 0), (w - 1.0, h - 1.0, 60.0, 1, -1),
                            (-500.0, 10.0, 30.0, -1, -1), (w + 400.0, 10.0, 30.0, 0, 7), (w / 3, h / 3, 1000.0, 2, 3),
                            (w / 2, -300.0, 20.0, -1, -1), (w / 2, h / 2, 0.0, -1, -1)):
        assert np.array_equal(matcher.features_in_area(k2, b, x, y, r, lo, hi, lib_path=lib), oracle.features_in_area(k2, b, x, y, r, lo, hi))
    pm = np.stack([k1["x"], k1["y"]], 1).astype(np.float32) + rng.normal(0, 3, (len(k1), 2)).astype(np.float32)
    for win, ratio, chk in ((100, 0.9, True), (30, 0.7, False)):
        on, om, opm = oracle.search_for_initialization(k1, d1, k2, d2, b, pm, win, ratio, chk)
        n, m12, pm2 = matcher.search_for_initialization(k1, d1, k2, d2, b, pm, win, ratio, chk, lib_path=lib)
        assert n == on and np.array_equal(m12, om) and np.array_equal(pm2.view(np.uint32), opm.view(np.uint32))
        assert on > 10
    n1 = len(k1)
    uv = (np.stack([k1["x"], k1["y"]], 1) + rng.normal(0, 4, (n1, 2))).astype(np.float32)
    invz = rng.uniform(0.02, 0.5, n1).astype(np.float32)
    valid = (rng.random(n1) < 0.85).astype(np.uint8)
    observed = (rng.random(n1) < 0.6).astype(np.uint8)
    blocked = (rng.random(len(k2)) < 0.15).astype(np.uint8)
    ur = np.where(rng.random(len(k2)) < 0.5, k2["x"] - 38.0, -1).astype(np.float32)
    for th, direction, chk, stereo, bl in ((15.0, 0, True, False, None), (7.0, 1, True, True, blocked), (15.0, 2, False, True, blocked)):
        on, oa = oracle.search_by_projection(uv, invz, k1["octave"], k1["angle"], d1, valid, observed, k2, d2, ur if stereo else None,
                                             bl, b, SF, th, 386.1448, direction, chk)
        n, a = matcher.search_by_projection(uv, invz, k1["octave"], k1["angle"], d1, valid, observed, k2, d2, ur if stereo else None,
                                            bl, b, SF, th, 386.1448, direction, chk, lib_path=lib)
        assert n == on and np.array_equal(a, oa)
        assert on > 10
    return k1, d1, k2, d2, b, pm


def _check_edges(lib):
    k1, d1, k2, d2, b, rng = _frames(5, 200, 160, 300)
    pm = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    e_k, e_d = k1[:0], d1[:0]
    n, m12, _ = matcher.search_for_initialization(e_k, e_d, k2, d2, b, pm[:0], lib_path=lib)       # no source keypoints
    assert n == 0 and len(m12) == 0
    n, m12, _ = matcher.search_for_initialization(k1, d1, e_k, e_d, b, pm, lib_path=lib)            # empty target frame
    assert n == 0 and (m12 == -1).all()
    n, m12, _ = matcher.search_for_initialization(k1, d1, k2, d2, b, pm + 5000, lib_path=lib)       # every window off the grid
    assert n == 0 and (m12 == -1).all()
    same = np.repeat(d2[:1], len(d2), axis=0)                                                        # all candidates tie
    on, om, _ = oracle.search_for_initialization(k1, d1, k2, same, b, pm, 100, 0.9, True)
    n, m12, _ = matcher.search_for_initialization(k1, d1, k2, same, b, pm, 100, 0.9, True, lib_path=lib)
    assert n == on and np.array_equal(m12, om)
    # batch of ragged pairs == the pairs one by one
    ks = [(k1, d1, k2, d2, pm), (k1[:50], d1[:50], k2[:70], d2[:70], pm[:50]), (k2, d2, k1, d1, np.stack([k2["x"], k2["y"]], 1))]
    nm, ms, pms = matcher.search_for_initialization_batch([k[0] for k in ks], [k[1] for k in ks], [k[2] for k in ks], [k[3] for k in ks],
                                                          b, [k[4] for k in ks], 60, 0.9, True, lib_path=lib)
    for p, k in enumerate(ks):
        on, om, opm = oracle.search_for_initialization(k[0], k[1], k[2], k[3], b, k[4], 60, 0.9, True)
        assert nm[p] == on and np.array_equal(ms[p], om) and np.array_equal(pms[p], opm)


def _check_rig_golden(golden_dir, lib):
    """Two-camera SearchByProjection (CurrentFrame.Nleft != -1, fmatcher.cpp:2593-2660) against the transliteration golden."""
    g = np.load(os.path.join(golden_dir, "window_rig.npz"))
    for tag in "abcd":
        th, direction, chk, ub = g[f"rig_{tag}_args"]
        n, asg = matcher.search_by_projection(g["uv"], g["invz"], g["octave"], g["angle"], g["dmp"], g["valid"], g["observed"], g["k2"],
                                              g["d2"], None, g["blocked2"] if ub else None, g["bounds"], g["sf"], th, 0.0, int(direction),
                                              bool(chk), lib_path=lib, uv_right=g["uv_right"], nleft2=int(g["nleft2"]))
        assert n == int(g[f"rig_{tag}_n"]) and np.array_equal(asg, g[f"rig_{tag}_asg"]), tag
        nl = int(g["nleft2"])
        assert (asg[:nl] >= 0).sum() > 100 and (asg[nl:] >= 0).sum() > 100          # both cameras matched


def _kb8_rig_case(seed, n_pts=1500):
    """A synthetic fisheye rig (the KB8 pair of tests/test_kb8.py): 3D points seen by the LastFrame are projected into the
    CurrentFrame's left camera and, through mTrl, into its right camera (what the caller of the ABI does); the current
    frame's keypoints are those projections with pixel noise plus clutter."""
    from tests.golden import ref_kb8
    from tests.golden.make_kb8_golden import PRM1, PRM2, rig_pose
    rng = np.random.default_rng(seed)
    R12, t12 = rig_pose()
    R21 = R12.T.astype(np.float64)
    t21 = -R21 @ t12.astype(np.float64)
    d = np.exp(rng.uniform(np.log(0.5), np.log(20.0), n_pts))
    ang, th = rng.uniform(0, 2 * np.pi, n_pts), rng.uniform(0, 1.1, n_pts)
    X1 = np.stack([np.sin(th) * np.cos(ang), np.sin(th) * np.sin(ang), np.cos(th)], 1) * d[:, None]
    X2 = X1 @ R21.T + t21
    uv = np.array([ref_kb8.project(x, PRM1) for x in X1], np.float32)
    uvr = np.array([ref_kb8.project(x, PRM2) for x in X2], np.float32)
    octave = rng.integers(0, 8, n_pts).astype(np.int32)
    angle = rng.uniform(0, 360, n_pts).astype(np.float32)
    dmp = rng.integers(0, 256, (n_pts, 32), dtype=np.uint8)

    def camera(proj, n_clutter):
        seen = rng.random(n_pts) < 0.7
        k = np.zeros(int(seen.sum()) + n_clutter, oracle.KP_DTYPE)
        m = int(seen.sum())
        k["x"][:m], k["y"][:m] = proj[seen, 0] + rng.normal(0, 1.5, m), proj[seen, 1] + rng.normal(0, 1.5, m)
        k["octave"][:m] = np.clip(octave[seen] + rng.integers(-1, 2, m), 0, 7)
        k["angle"][:m] = (angle[seen] + np.where(rng.random(m) < 0.2, 140.0, rng.normal(0, 3, m))) % 360
        k["x"][m:], k["y"][m:] = rng.uniform(0, 512, n_clutter), rng.uniform(0, 512, n_clutter)
        k["octave"][m:], k["angle"][m:] = rng.integers(0, 8, n_clutter), rng.uniform(0, 360, n_clutter)
        k["size"], k["class_id"] = 31, -1
        dd = rng.integers(0, 256, (len(k), 32), dtype=np.uint8)
        noisy = dmp[seen].copy()
        for i in range(m):
            for bit in rng.choice(256, int(rng.integers(0, 80)), replace=False):
                noisy[i, bit >> 3] ^= 1 << (bit & 7)
        dd[:m] = noisy
        perm = rng.permutation(len(k))
        return k[perm], dd[perm]

    kl, dl = camera(uv, 400)
    kr, dr = camera(uvr, 300)
    k2, d2 = np.concatenate([kl, kr]), np.concatenate([dl, dr])
    invz = (1.0 / X1[:, 2]).astype(np.float32)
    valid = ((rng.random(n_pts) < 0.9) & (uv[:, 0] >= 0) & (uv[:, 0] <= 512) & (uv[:, 1] >= 0) & (uv[:, 1] <= 512)).astype(np.uint8)
    return {"uv": uv, "uv_right": uvr, "invz": invz, "octave": octave, "angle": angle, "dmp": dmp, "valid": valid,
            "observed": (rng.random(n_pts) < 0.6).astype(np.uint8), "k2": k2, "d2": d2, "nleft2": len(kl),
            "blocked2": (rng.random(len(k2)) < 0.1).astype(np.uint8), "bounds": np.array([0, 512, 0, 512], np.float32)}


def _check_rig_vs_oracle(lib, seed):
    c = _kb8_rig_case(seed)
    for th, direction, chk, bl in ((15.0, 0, True, c["blocked2"]), (7.0, 1, True, None), (25.0, 2, False, c["blocked2"])):
        on, oa = oracle.search_by_projection_rig(c["uv"], c["uv_right"], c["invz"], c["octave"], c["angle"], c["dmp"], c["valid"],
                                                 c["observed"], c["k2"], c["d2"], c["nleft2"], bl, c["bounds"], SF, th, direction, chk)
        n, a = matcher.search_by_projection(c["uv"], c["invz"], c["octave"], c["angle"], c["dmp"], c["valid"], c["observed"], c["k2"], c["d2"],
                                            None, bl, c["bounds"], SF, th, 0.0, direction, chk, lib_path=lib, uv_right=c["uv_right"],
                                            nleft2=c["nleft2"])
        assert n == on and np.array_equal(a, oa)
        assert (oa[:c["nleft2"]] >= 0).sum() > 50 and (oa[c["nleft2"]:] >= 0).sum() > 50
    # Nleft = 0 / Nleft = N: one camera of the rig without keypoints
    for nl in (0, len(c["k2"])):
        on, oa = oracle.search_by_projection_rig(c["uv"], c["uv_right"], c["invz"], c["octave"], c["angle"], c["dmp"], c["valid"],
                                                 c["observed"], c["k2"], c["d2"], nl, None, c["bounds"], SF, 15.0, 0, True)
        n, a = matcher.search_by_projection(c["uv"], c["invz"], c["octave"], c["angle"], c["dmp"], c["valid"], c["observed"], c["k2"], c["d2"],
                                            None, None, c["bounds"], SF, 15.0, 0.0, 0, True, lib_path=lib, uv_right=c["uv_right"], nleft2=nl)
        assert n == on and np.array_equal(a, oa)


@pytest.fixture(scope="module")
def emu(oracle_mod):
    return ge.build_emu()


def test_window_rig_golden_emulation(emu, golden_dir):
    _check_rig_golden(golden_dir, emu)


def test_window_rig_vs_oracle_emulation(emu):
    _check_rig_vs_oracle(emu, 31)


@pytest.mark.gpu
def test_window_rig_golden_cuda(oracle_mod, golden_dir):
    _check_rig_golden(golden_dir, None)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [32, 33])
def test_window_rig_vs_oracle_cuda(oracle_mod, seed):
    _check_rig_vs_oracle(None, seed)


def test_window_golden_emulation(emu, golden_dir):
    _check_golden(golden_dir, emu)


def test_window_vs_oracle_emulation(emu):
    _check_vs_oracle(emu, 11, 320, 240, 800)


def test_window_edges_emulation(emu):
    _check_edges(emu)


@pytest.mark.gpu
def test_window_golden_cuda(oracle_mod, golden_dir):
    _check_golden(golden_dir, None)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,w,h,nf", [(21, 1241, 376, 2000), (22, 1241, 376, 10000), (23, 1280, 720, 4000)])
def test_window_vs_oracle_cuda(oracle_mod, seed, w, h, nf):
    """nf = 10000 is the monocular initialiser's extractor (5 x nFeatures, tracking.cpp:1092-1093)."""
    _check_vs_oracle(None, seed, w, h, nf)


@pytest.mark.gpu
def test_window_edges_cuda(oracle_mod):
    _check_edges(None)
