"""Parity pin of the matcher rows against the reference's OWN function bodies.

oracle/_ref/liborb_ref.so also holds Frame::ComputeStereoMatches (frame.cpp:823-997), Frame::AssignFeaturesToGrid /
GetFeaturesInArea / PosInGrid (frame.cpp:386-414, 678-756), FMatcher::SearchForInitialization (fmatcher.cpp:983-1098),
ComputeThreeMaxima (:2813-2854), DescriptorDistance (:2859-2875) and the free matchFeatures / removeDuplicatedMatches
(:132-266): oracle/ref.mk cuts those definitions out of the
reference files by line range at build time and compiles them unmodified against stand-in class declarations
(oracle/ref_compat/ref_slices.cpp); the pyramids the stereo matcher reads are the reference FExtractor's own.
Compared here: the oracle (CPU), the emulation build (CPU) and - marked gpu - the CUDA path; committed fixtures
(tests/golden/refm_*.npz, written by tests/golden/make_ref_matcher_golden.py) let the comparisons run where
/root/reference is absent."""
import os

import numpy as np
import pytest

import oracle
from oracle import ref
from tests import ref_matcher_cases as cases



@pytest.fixture(scope="module")
def emu_lib(oracle_mod):
    import __graft_entry__ as ge
    return ge.build_emu()


needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/liborb_ref.so not built and /root/reference absent")


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


# --------------------------------------------------------------------------------------------- row a-8
@needs_ref
def test_descriptor_distance_reference_text():
    rng = np.random.default_rng(3)
    d = rng.integers(0, 256, (400, 32), dtype=np.uint8)
    d[0] = 0
    d[1] = 255
    d[2] = d[3]
    for i in range(0, 400, 2):
        want = ref.descriptor_distance(d[i], d[i + 1])
        assert want == int(np.unpackbits(d[i] ^ d[i + 1]).sum())
        assert oracle.descriptor_distance(d[i], d[i + 1]) == want
    assert ref.descriptor_distance(d[0], d[1]) == 256 and ref.descriptor_distance(d[2], d[3]) == 0


# --------------------------------------------------------------------------------------------- row a-9 (matchFeatures method 2)
def _match_features_case(seed, nq, nt):
    """queries = noisy copies of train rows, several per train row, plus unrelated rows: exercises the ratio rule and
    removeDuplicatedMatches."""
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    src = rng.integers(0, max(nt // 2, 1), nq)
    q = t[src].copy()
    flips = rng.integers(0, 60, nq)
    for i in range(nq):
        pos = rng.choice(256, flips[i], replace=False)
        np.bitwise_xor.at(q[i], pos >> 3, (1 << (pos & 7)).astype(np.uint8))
    q[::7] = rng.integers(0, 256, (len(q[::7]), 32), dtype=np.uint8)
    return q, t


def _check_match_features(got, want, q, t):
    """want = the reference's matchFeatures output.  The surviving trainIdx set is exact; which query survives on a
    duplicated trainIdx is std::sort's choice in the reference (unstable) and the lowest query index in ours
    (include/viorb.h: stable) - both must come from the pre-dedup candidates of that trainIdx."""
    idx, dist = oracle.knn2(q, t)
    cand = oracle.ratio_filter(idx, dist, 1.0, 1)
    np.testing.assert_array_equal(got[:, 1], want[:, 1])
    assert np.all(np.diff(want[:, 1]) > 0)
    by_train = {}
    for qi, ti, d in cand:
        by_train.setdefault(int(ti), []).append((int(qi), int(d)))
    n_dup = 0
    for g, w in zip(got, want):
        c = by_train[int(w[1])]
        assert (int(w[0]), int(w[2])) in c and (int(g[0]), int(g[2])) in c
        if len(c) == 1:
            assert tuple(g) == tuple(w)
        else:
            n_dup += 1
            assert int(g[0]) == min(x[0] for x in c)
    return n_dup


@needs_ref
@pytest.mark.parametrize("seed,nq,nt", [(1, 900, 700), (2, 12, 40), (3, 300, 2)])
def test_oracle_match_features_equals_reference_text(seed, nq, nt):
    q, t = _match_features_case(seed, nq, nt)
    want = ref.match_features(q, t, 1)
    idx, dist = oracle.knn2(q, t)
    got = oracle.dedup_by_train(oracle.ratio_filter(idx, dist, 1.0, 1))
    n_dup = _check_match_features(got, want, q, t)
    assert len(want) > 0 and (nq < 100 or n_dup > 0)
    if nq <= 16:                      # std::sort is an insertion sort (stable) up to 16 elements: everything is exact
        np.testing.assert_array_equal(got, want)


def test_oracle_match_features_equals_committed_reference_fixture(golden_dir):
    z = np.load(os.path.join(golden_dir, "refm_matchfeatures.npz"))
    idx, dist = oracle.knn2(z["q"], z["t"])
    got = oracle.dedup_by_train(oracle.ratio_filter(idx, dist, 1.0, 1))
    assert _check_match_features(got, z["matches"], z["q"], z["t"]) > 0


@pytest.mark.gpu
def test_cuda_match_features_equals_committed_reference_fixture(golden_dir):
    from vi_slam_b200 import matcher
    z = np.load(os.path.join(golden_dir, "refm_matchfeatures.npz"))
    m = matcher.match_features(z["q"], z["t"], method_index=2, lowe_method_dist_ratio=1.0)
    got = np.stack([m["query_idx"], m["train_idx"], m["distance"]], 1).astype(np.int32)
    assert _check_match_features(got, z["matches"], z["q"], z["t"]) > 0


# --------------------------------------------------------------------------------------------- row a-10
@needs_ref
@pytest.mark.parametrize("name", list(cases.STEREO))
def test_oracle_stereo_equals_reference_text(name):
    """ComputeStereoMatches: the reference's text on the reference extractor's pyramids == oracle.stereo_match on the
    oracle's pyramids (which test_ref_pin.py shows to be the same bytes), mvuRight / mvDepth bit for bit."""
    c = cases.stereo_inputs(name)
    ur, dp = ref.stereo_match(c["rl"], c["rr"], c["kl"], c["dl"], c["kr"], c["dr"], c["bf"], c["b"])
    our, odp, _, _ = oracle.stereo_match(c["ol"], c["orr"], c["kl"], c["dl"], c["kr"], c["dr"], c["bf"], c["b"])
    assert np.array_equal(_bits(ur), _bits(our)) and np.array_equal(_bits(dp), _bits(odp))
    n = int((dp > 0).sum())
    print(f"{name}: {len(c['kl'])} left / {len(c['kr'])} right keypoints, {n} stereo matches")
    assert n > 50


def _check_stereo_fixture(golden_dir, name, run):
    g = np.load(os.path.join(golden_dir, f"refm_stereo_{name}.npz"))
    ur, dp = run(g)
    assert np.array_equal(_bits(ur), g["u_right"].view(np.uint32)) and np.array_equal(_bits(dp), g["depth"].view(np.uint32))


@pytest.mark.parametrize("name", list(cases.STEREO))
def test_oracle_stereo_equals_committed_reference_fixture(golden_dir, name):
    """Runs anywhere: the oracle on the same images must reproduce the reference's committed mvuRight / mvDepth (and, on
    the way, the reference's keypoints and descriptors)."""
    def run(g):
        left, right = cases.stereo_images(name)
        nf = cases.STEREO[name][1]
        ol, orr = oracle.OracleExtractor(nf), oracle.OracleExtractor(nf)
        _, kl, dl = ol.compute(left, (0, 0))
        _, kr, dr = orr.compute(right, (0, 0))
        assert np.array_equal(dl, g["dl"]) and np.array_equal(dr, g["dr"]) and np.array_equal(kl["x"], g["klx"])
        return oracle.stereo_match(ol, orr, kl, dl, kr, dr, float(g["bf"]), float(g["b"]))[:2]
    _check_stereo_fixture(golden_dir, name, run)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(cases.STEREO))
def test_cuda_stereo_equals_committed_reference_fixture(golden_dir, name):
    """The CUDA path (k_stereo_rows / coarse / refine / cut behind viorb_compute_stereo_matches and the batched resident
    form) against the reference's own ComputeStereoMatches output."""
    from vi_slam_b200 import FExtractor, matcher

    def run(g):
        left, right = cases.stereo_images(name)
        nf = cases.STEREO[name][1]
        h, w = left.shape
        gl, gr = FExtractor(nf, max_size=(w, h)), FExtractor(nf, max_size=(w, h))
        _, kl, dl = gl.compute(left, None, (0, 0))
        _, kr, dr = gr.compute(right, None, (0, 0))
        assert np.array_equal(dl, g["dl"]) and np.array_equal(dr, g["dr"]) and np.array_equal(kl["x"], g["klx"])
        ur, dp, _, _ = matcher.compute_stereo_matches(gl, gr, float(g["bf"]), float(g["b"]))
        matcher.stereo_batch(gl, gr, 1, float(g["bf"]), float(g["b"]), 0.7)
        r = matcher.stereo_fetch(gl, 0)
        assert np.array_equal(_bits(r["u_right"]), _bits(ur)) and np.array_equal(_bits(r["depth"]), _bits(dp))
        return ur, dp
    _check_stereo_fixture(golden_dir, name, run)


# --------------------------------------------------------------------------------------------- row a-11 / f-1 (grid, initialiser)
@needs_ref
def test_features_in_area_reference_text():
    """GetFeaturesInArea on the AssignFeaturesToGrid grid: same indices in the same order, incl. windows that leave the
    image, level filters and degenerate radii."""
    k1, _, k2, _, bounds = cases.ini_inputs("kitti")[:5]
    rng = np.random.default_rng(11)
    for i in range(300):
        x, y = float(rng.uniform(-80, bounds[1] + 80)), float(rng.uniform(-80, bounds[3] + 80))
        r = float(rng.choice([0.0, 3.0, 15.0, 100.0, 400.0]))
        lo, hi = [(-1, -1), (0, 0), (2, -1), (1, 3), (0, 7)][i % 5]
        want = ref.features_in_area(k2, bounds, x, y, r, lo, hi)
        assert np.array_equal(oracle.features_in_area(k2, bounds, x, y, r, lo, hi), want), (x, y, r, lo, hi)


@needs_ref
@pytest.mark.parametrize("name", list(cases.INI))
@pytest.mark.parametrize("nn_ratio,check", [(0.9, True), (0.6, False)])
def test_oracle_search_for_initialization_equals_reference_text(name, nn_ratio, check):
    k1, d1, k2, d2, bounds, prev, win = cases.ini_inputs(name)
    n, m12, pm = ref.search_for_initialization(k1, d1, k2, d2, bounds, prev, win, nn_ratio, check)
    on, om12, opm = oracle.search_for_initialization(k1, d1, k2, d2, bounds, prev, win, nn_ratio, check)
    assert n == on and np.array_equal(m12, om12) and np.array_equal(_bits(pm), _bits(opm))
    print(f"{name} nn {nn_ratio} orientation {check}: {n} matches of {len(k1)} keypoints")
    assert n > 20


def _check_ini_fixture(golden_dir, name, run):
    g = np.load(os.path.join(golden_dir, f"refm_ini_{name}.npz"))
    n, m12, pm = run(g["k1"], g["d1"], g["k2"], g["d2"], g["bounds"], g["prev"], int(g["window"]))
    assert n == int(g["nmatches"]) and np.array_equal(m12, g["matches12"]) and np.array_equal(_bits(pm), g["prev_out"].view(np.uint32))


@pytest.mark.parametrize("name", list(cases.INI))
def test_oracle_search_for_initialization_equals_committed_reference_fixture(golden_dir, name):
    _check_ini_fixture(golden_dir, name, lambda *a: oracle.search_for_initialization(*a, 0.9, True))


@pytest.mark.parametrize("name", ["qvga"])
def test_emulation_search_for_initialization_equals_committed_reference_fixture(golden_dir, emu_lib, name):
    from vi_slam_b200 import matcher
    _check_ini_fixture(golden_dir, name, lambda *a: matcher.search_for_initialization(*a, 0.9, True, lib_path=emu_lib))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(cases.INI))
def test_cuda_search_for_initialization_equals_committed_reference_fixture(golden_dir, name):
    from vi_slam_b200 import matcher
    _check_ini_fixture(golden_dir, name, lambda *a: matcher.search_for_initialization(*a, 0.9, True))


# --------------------------------------------------------------------------------------------- row f-1 (tracking matcher, both camera branches)
def _ours_projection(run_single, run_rig, c, host):
    kl = c["k_last"]
    if int(c["nleft2"]) < 0:
        n, a = run_single(host["uv"], host["invz"], kl["octave"], kl["angle"], c["desc_mp"], host["valid"], c["observed"], c["k2"], c["d2"],
                          c["u_right2"], c["blocked2"], c["bounds"], c["sf"], float(c["th"]), float(c["bf"]), int(host["direction"]),
                          bool(c["check"]))
    else:
        n, a = run_rig(host["uv"], host["uv_right"], host["invz"], kl["octave"], kl["angle"], c["desc_mp"], host["valid"], c["observed"],
                       c["k2"], c["d2"], int(c["nleft2"]), c["blocked2"], c["bounds"], c["sf"], float(c["th"]), int(host["direction"]),
                       bool(c["check"]))
    return int(n), np.where(a >= 0, a, -1)           # -2 (cleared by the rotation check) and -1 are both "no map point"


def _ref_projection(c):
    return ref.search_by_projection(c["k_last"], c["xw"], c["desc_mp"], c["outlier"], c["observed"], c["k2"], c["d2"], c["u_right2"],
                                    c["blocked2"], c["nleft2"], c["bounds"], c["tcw_cur"], c["tcw_last"], c["trl"], c["cam"], c["sf"],
                                    c["th"], c["mono"], c["bf"], c["b"], c["check"])


@needs_ref
@pytest.mark.parametrize("name", list(cases.PROJ))
def test_oracle_search_by_projection_equals_reference_text(name):
    """SearchByProjection(CurrentFrame, LastFrame) - pose algebra, Pinhole::project, window query, distance loop, the
    mvuRight predicate, the two-camera pass with its Nleft offsets, the joint rotation histogram - run as the reference's own
    text; the oracle gets the uv / invz / valid / direction that text computed and must give the same assignment."""
    c = cases.proj_scene(name)
    n, asg, host = _ref_projection(c)
    on, oa = _ours_projection(oracle.search_by_projection, oracle.search_by_projection_rig, c, host)
    assert n == on and np.array_equal(asg, oa)
    print(f"{name}: direction {host['direction']}, {int(host['valid'].sum())} projected points, {n} matches")
    assert n > 300
    if c["nleft2"] >= 0:
        assert (asg[:c["nleft2"]] >= 0).sum() > 100 and (asg[c["nleft2"]:] >= 0).sum() > 100


def _proj_fixture(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"refm_proj_{name}.npz"))
    c = {k: g[k] for k in g.files}
    for k in ("u_right2", "blocked2"):
        if c[k].size == 0:
            c[k] = None
    host = {k: g["host_" + k] for k in ("uv", "uv_right", "invz", "valid")}
    host["direction"] = int(g["host_direction"])
    return c, host, int(g["nmatches"]), g["assigned"]


@pytest.mark.parametrize("name", list(cases.PROJ))
def test_oracle_search_by_projection_equals_committed_reference_fixture(golden_dir, name):
    c, host, n, asg = _proj_fixture(golden_dir, name)
    on, oa = _ours_projection(oracle.search_by_projection, oracle.search_by_projection_rig, c, host)
    assert on == n and np.array_equal(oa, asg)


def _device_runs(lib_path):
    from vi_slam_b200 import matcher
    single = lambda *a: matcher.search_by_projection(*a, lib_path=lib_path)
    rig = lambda uv, uvr, invz, lo, la, dmp, va, ob, k2, d2, nl, bl, bd, sf, th, direction, chk: matcher.search_by_projection(
        uv, invz, lo, la, dmp, va, ob, k2, d2, None, bl, bd, sf, th, 0.0, direction, chk, lib_path=lib_path, uv_right=uvr, nleft2=nl)
    return single, rig


@pytest.mark.parametrize("name", ["mono", "rig"])
def test_emulation_search_by_projection_equals_committed_reference_fixture(golden_dir, emu_lib, name):
    c, host, n, asg = _proj_fixture(golden_dir, name)
    on, oa = _ours_projection(*_device_runs(emu_lib), c, host)
    assert on == n and np.array_equal(oa, asg)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(cases.PROJ))
def test_cuda_search_by_projection_equals_committed_reference_fixture(golden_dir, name):
    c, host, n, asg = _proj_fixture(golden_dir, name)
    on, oa = _ours_projection(*_device_runs(None), c, host)
    assert on == n and np.array_equal(oa, asg)


# --------------------------------------------------------------------------------------------- row f-3 (DBoW3 transform)
def _same_bow(a, b):
    for x, y in zip(a, b):
        x, y = np.ascontiguousarray(x), np.ascontiguousarray(y)
        assert x.shape == y.shape and np.array_equal(x.view(np.uint8), y.view(np.uint8))


@needs_ref
def test_dbow3_own_transform_equals_committed_golden(golden_dir):
    """DBoW3's own Vocabulary::transform / BowVector / FeatureVector / DescManip::distance (compiled from thirdparty/DBoW3)
    reproduce tests/golden/bow.npz - the fixture the oracle, the emulation build and the CUDA kernels are held to in
    tests/test_bow.py (it was written by the Python transliteration tests/golden/ref_bow.py): word ids, node ids, feature
    lists and the double values bit for bit, every weighting and every normalisation."""
    from tests.golden.make_bow_golden import CASES
    from vi_slam_b200 import synthetic
    g = np.load(os.path.join(golden_dir, "bow.npz"))
    for tag, (kw, levelsup, _) in CASES.items():
        rv = ref.RefVocabulary(synthetic.make_vocabulary(**kw))
        got = rv.transform(g[tag + "_desc"], levelsup)
        _same_bow(got, (g[tag + "_bow_word"], g[tag + "_bow_value"], g[tag + "_fv_node"], g[tag + "_fv_start"], g[tag + "_fv_feature"]))
        assert len(got[0]) > 30, tag


@needs_ref
@pytest.mark.parametrize("k,levels,n,seed,prune", [(10, 4, 1500, 5, 0.0), (6, 5, 800, 6, 0.2), (40, 2, 600, 7, 0.0), (3, 6, 400, 8, 0.3)])
def test_oracle_bow_equals_dbow3_own_transform(k, levels, n, seed, prune):
    from vi_slam_b200 import synthetic
    for weighting in range(4):
        for scoring in (0, 1, 5):                       # L1, L2, dot product (no normalisation)
            v = synthetic.make_vocabulary(seed, k=k, levels=levels, weighting=weighting, scoring=scoring, stop_fraction=0.03, prune=prune)
            rv, ov = ref.RefVocabulary(v), oracle.OracleVocabulary(v)
            rng = np.random.default_rng(seed)
            for d in (synthetic.descriptors_near_words(v, n, seed + 1, 20), rng.integers(0, 256, (n // 3, 32), dtype=np.uint8),
                      np.repeat(synthetic.descriptors_near_words(v, 3, seed + 3), 40, axis=0), np.zeros((0, 32), np.uint8)):
                for levelsup in (0, 2, levels, levels + 3):
                    a, b = rv.transform(d, levelsup), ov.transform(d, levelsup)
                    if prune == 0.0:
                        _same_bow(a, b)
                        continue
                    # An irregular tree has leaves above level L - levelsup: DBoW3 then never writes *nid (Vocabulary.cpp:868)
                    # and the FeatureVector gets whatever the stack held - the one place where the reference is undefined
                    # (the oracle and the kernels report node 0).  BowVector: identical; FeatureVector:
                    # identical on every feature whose node is defined.
                    _same_bow(a[:2], b[:2])
                    node_of = lambda r: {int(f): int(r[2][k]) for k in range(len(r[2])) for f in r[4][r[3][k]:r[3][k + 1]]}
                    na, nb = node_of(a), node_of(b)
                    assert na.keys() == nb.keys()
                    undefined = [f for f in nb if nb[f] == 0 and levelsup < levels]
                    assert all(na[f] == nb[f] for f in nb if f not in set(undefined))


# --------------------------------------------------------------------------------------------- the timed CPU baseline body
@needs_ref
def test_reference_bench_body_does_the_same_work_as_the_port():
    """bench.py's CPU arm (oracle/_ref, std::threads, glibc malloc) against the oracle port on the same frames: keypoint and
    match totals agree up to the reference's own pointer tie-breaks (a few keypoints per frame)."""
    from vi_slam_b200 import synthetic
    left = np.stack([synthetic.make_stereo_pair(70 + i, 640, 360)[0] for i in range(4)])
    right = np.stack([synthetic.make_stereo_pair(70 + i, 640, 360)[1] for i in range(4)])
    frames = np.empty((8, 360, 640), np.uint8)
    frames[0::2], frames[1::2] = left, right
    rt, rm = ref.bench_frames(frames, 1000, mode=1, bf=120.0, b=0.24, nthreads=3)
    ot, om = oracle.bench_frames(frames, 1000, stereo=True, bf=120.0, b=0.24, nthreads=3)
    assert abs(rt - ot) <= 0.01 * ot and abs(rm - om) <= 0.03 * om and rm > 500
    rt, _ = ref.bench_frames(frames, 1000, lapping=(0, 1000), mode=0, nthreads=2)
    ot, _ = oracle.bench_frames(frames, 1000, lapping=(0, 1000), stereo=False, nthreads=2)
    assert abs(rt - ot) <= 0.01 * ot
    rt, rm = ref.bench_frames(frames[0::2], 1000, lapping=(0, 1000), mode=2, nthreads=2)
    assert rt > 3000 and rm > 0


# --------------------------------------------------------------------------------------------- row f-4 (fisheye stereo epilogue)
@needs_ref
@pytest.mark.parametrize("seed", [3, 17])
def test_oracle_kb8_triangulation_equals_reference_text(seed):
    """KannalaBrandt8::TriangulateMatches with unproject (Newton), project, Triangulate - the reference's own text - against the
    oracle: every accept / reject decision, depth and 3D point bit for bit.  (cv::SVD::compute is the same Jacobi stand-in on
    both sides; the real cv::SVD is pinned with a tolerance through tests/golden/kb8.npz, tests/test_kb8.py.)"""
    from tests.golden.make_kb8_golden import PRM1, PRM2, make_matches, rig_pose
    R12, t12 = rig_pose()
    rig = oracle.kb8_rig(PRM1, PRM2, R12, t12)
    p1, p2, s1, s2 = make_matches(6000, seed)
    rd, rp = ref.kb8_triangulate(rig, p1, p2, s1, s2)
    od, op, _ = oracle.kb8_triangulate(rig, p1, p2, s1, s2)
    assert np.array_equal(_bits(rd), _bits(od)) and np.array_equal(_bits(rp), _bits(op))
    assert 0.2 * len(rd) < (rd > 0).sum() < 0.8 * len(rd)            # both outcomes are well represented


@needs_ref
@pytest.mark.parametrize("n,seed,mono_left,mono_right", [(1200, 5, 0, 0), (900, 6, 150, 90)])
def test_oracle_fisheye_stereo_equals_reference_text(n, seed, mono_left, mono_right):
    """Frame::ComputeStereoFishEyeMatches - lapping-area slices, knnMatch, the 0.7 rule, TriangulateMatches, the three output
    vectors (a right keypoint claimed twice keeps the later left index) - as the reference's own text."""
    from tests.golden.make_kb8_golden import PRM1, PRM2, rig_pose
    from tests.test_kb8 import SIGMA2, _fisheye_frames
    R12, t12 = rig_pose()
    rig = oracle.kb8_rig(PRM1, PRM2, R12, t12)
    kl, dl, kr, dr = _fisheye_frames(n, seed, mono_left, mono_right)
    got = ref.stereo_fisheye_matches(rig, kl, dl, mono_left, kr, dr, mono_right, SIGMA2)
    exp = oracle.stereo_fisheye_matches(rig, kl, dl, mono_left, kr, dr, mono_right, SIGMA2)
    assert np.array_equal(got["left_to_right"], exp["left_to_right"]) and np.array_equal(got["right_to_left"], exp["right_to_left"])
    assert np.array_equal(_bits(got["depth"]), _bits(exp["depth"])) and np.array_equal(_bits(got["p3d"]), _bits(exp["p3d"]))
    assert (got["left_to_right"] >= 0).sum() > n // 10


# --------------------------------------------------------------------------------------------- row f-2 (undistortion, image bounds)
@needs_ref
@pytest.mark.parametrize("cam,dist", [([500, 510, 320, 240], [-0.28, 0.07, 0.0002, 0.0001]),
                                      ([500, 510, 320, 240], [0.0, 0.07, 0.0002, 0.0001]),            # k1 == 0: the copy branch (frame.cpp:762)
                                      ([458.654, 457.296, 367.215, 248.375], [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.01])])
def test_oracle_undistortion_equals_reference_text(cam, dist):
    """Frame::UndistortKeyPoints / ComputeImageBounds as the reference's own text (the k1 == 0 branch, the point matrix, the
    corner order of the bounds); cv::undistortPoints underneath is the restatement that tests/test_undistort.py pins against cv2."""
    from vi_slam_b200 import synthetic
    _, k, _ = oracle.OracleExtractor(1500).compute(synthetic.make_frame(5, 640, 480), (0, 0))
    a, b = ref.undistort_keypoints(k, cam, dist), oracle.undistort_keypoints(k, cam, dist)
    for f in a.dtype.names:
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f
    assert np.array_equal(_bits(ref.image_bounds(640, 480, cam, dist)), _bits(oracle.image_bounds(640, 480, cam, dist)))
