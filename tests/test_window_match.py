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


@pytest.fixture(scope="module")
def emu(oracle_mod):
    return ge.build_emu()


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
