"""Host-side mirror of the reference's descriptor matching calls over the C ABI.

* ``descriptor_distance``      FMatcher::DescriptorDistance   (src/geometry/fmatcher.cpp:2859-2875)
* ``knn_match``                cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, 2)  (frame.cpp:1167)
* ``match_features``           matchFeatures(..., method_index=2)             (fmatcher.cpp:204-229,247-266)
* ``fisheye_ratio_match``      the 0.7 ratio rule of ComputeStereoFishEyeMatches (frame.cpp:1170)
* ``compute_stereo_matches``   Frame::ComputeStereoMatches                    (frame.cpp:823-997)
* ``stereo_batch``             both stereo matchers on a resident batch
* ``features_in_area``         Frame::GetFeaturesInArea on Frame::AssignFeaturesToGrid   (frame.cpp:386-414,678-757)
* ``search_for_initialization`` FMatcher::SearchForInitialization                        (fmatcher.cpp:983-1098)
* ``search_by_projection``     FMatcher::SearchByProjection(CurrentFrame, LastFrame, ..) (fmatcher.cpp:2471-2688)
* ``kb8_rig`` / ``triangulate_matches``  KannalaBrandt8::TriangulateMatches (src/geometry/cameramodels/kannalabrandt8.cpp:324-393)
* ``compute_stereo_fisheye_matches``      Frame::ComputeStereoFishEyeMatches     (frame.cpp:1148-1189)
All compute happens in libviorb.so (CUDA); no CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .extractor import KP_DTYPE, MATCH_DTYPE, FExtractor, _check

TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30      # fmatcher.cpp:313-315


def _desc(a):
    a = np.ascontiguousarray(a, np.uint8)
    assert a.ndim == 2 and a.shape[1] == 32
    return a


def descriptor_distance(a, b, device=0, lib_path=None):
    lib = _lib.load(lib_path)
    a, b = _desc(np.atleast_2d(a)), _desc(np.atleast_2d(b))
    out = np.zeros(a.shape[0], np.int32)
    _check(lib, lib.viorb_descriptor_distance(a.ctypes.data, b.ctypes.data, a.shape[0], out.ctypes.data_as(_lib.i32p), device))
    return out


def knn_match(query, train, device=0, lib_path=None):
    """-> (idx int32[nq,2], dist int32[nq,2]); ascending distance, ties to the lower train index, -1 if absent."""
    lib = _lib.load(lib_path)
    q, t = _desc(query), _desc(train)
    idx = np.full((q.shape[0], 2), -1, np.int32)
    dist = np.full((q.shape[0], 2), -1, np.int32)
    _check(lib, lib.viorb_knn2(q.ctypes.data, q.shape[0], t.ctypes.data, t.shape[0], idx.ctypes.data_as(_lib.i32p),
                               dist.ctypes.data_as(_lib.i32p), device))
    return idx, dist


def _ratio(query, train, ratio, rule, dedup, device, lib_path):
    lib = _lib.load(lib_path)
    q, t = _desc(query), _desc(train)
    out = np.zeros(max(q.shape[0], 1), MATCH_DTYPE)
    n = C.c_int32()
    _check(lib, lib.viorb_match_ratio(q.ctypes.data, q.shape[0], t.ctypes.data, t.shape[0], float(ratio), rule, dedup,
                                      out.ctypes.data, out.shape[0], C.byref(n), device))
    return out[:n.value].copy()


def match_features(descriptors_1, descriptors_2, method_index=2, lowe_method_dist_ratio=0.8, device=0, lib_path=None):
    """matchFeatures method 2: kNN-2, Lowe ratio, removeDuplicatedMatches.  Other method indices raise like the
    reference's std::runtime_error (methods 1/3 are FLANN-LSH based and out of scope)."""
    if method_index != 2:
        raise RuntimeError("feature_match.cpp::matchFeatures: wrong method index.")
    return _ratio(descriptors_1, descriptors_2, lowe_method_dist_ratio, 1, 1, device, lib_path)


def fisheye_ratio_match(desc_left, desc_right, ratio=0.7, device=0, lib_path=None):
    return _ratio(desc_left, desc_right, ratio, 0, 0, device, lib_path)


def compute_stereo_matches(left: FExtractor, right: FExtractor, bf: float, b: float, frame=0):
    """-> (mvuRight, mvDepth, bestIdx, bestDist) for the frames last extracted by ``left`` / ``right``."""
    lib = left._lib
    cap = left.capacity
    ur = np.zeros(cap, np.float32)
    dp = np.zeros(cap, np.float32)
    bi = np.zeros(cap, np.int32)
    bd = np.zeros(cap, np.int32)
    n = C.c_int32()
    _check(lib, lib.viorb_stereo_match(left._h, right._h, frame, bf, b, ur.ctypes.data_as(_lib.f32p),
                                       dp.ctypes.data_as(_lib.f32p), bi.ctypes.data_as(_lib.i32p),
                                       bd.ctypes.data_as(_lib.i32p), cap, C.byref(n)))
    k = n.value
    return ur[:k].copy(), dp[:k].copy(), bi[:k].copy(), bd[:k].copy()


def stereo_batch(left: FExtractor, right: FExtractor, n: int, bf: float, b: float, ratio=0.7):
    _check(left._lib, left._lib.viorb_stereo_batch(left._h, right._h, n, bf, b, float(ratio)))


def stereo_fetch(left: FExtractor, frame: int):
    """-> dict(u_right, depth, lr_match, n_stereo, n_ratio) of one frame after ``stereo_batch``."""
    lib = left._lib
    cap = left.capacity
    ur = np.zeros(cap, np.float32)
    dp = np.zeros(cap, np.float32)
    lr = np.zeros(cap, np.int32)
    nl, ns, nr = C.c_int32(), C.c_int32(), C.c_int32()
    _check(lib, lib.viorb_stereo_fetch(left._h, frame, ur.ctypes.data_as(_lib.f32p), dp.ctypes.data_as(_lib.f32p),
                                       lr.ctypes.data_as(_lib.i32p), cap, C.byref(nl), C.byref(ns), C.byref(nr)))
    k = nl.value
    return {"u_right": ur[:k].copy(), "depth": dp[:k].copy(), "lr_match": lr[:k].copy(), "n_stereo": ns.value,
            "n_ratio": nr.value}


def match_resident(cur: FExtractor, frame_q: int, prev: FExtractor, frame_t: int, ratio=0.7, rule=0):
    lib = cur._lib
    out = np.zeros(cur.capacity, MATCH_DTYPE)
    n = C.c_int32()
    _check(lib, lib.viorb_match_resident(cur._h, frame_q, prev._h, frame_t, float(ratio), rule, out.ctypes.data,
                                         out.shape[0], C.byref(n)))
    return out[:n.value].copy()


STEREO_STAGES = ("stereo_coarse", "stereo_refine", "stereo_cut", "knn2", "ratio")


def stereo_stage_ms(left: FExtractor):
    ms = np.zeros(5, np.float32)
    _check(left._lib, left._lib.viorb_stereo_stage_ms(left._h, ms.ctypes.data_as(_lib.f32p)))
    return dict(zip(STEREO_STAGES, ms.tolist()))


def stereo_fetch_all(left: FExtractor, n: int, out=None):
    """Whole-batch read-back after ``stereo_batch``: dict of [n, cap] arrays and [n] counts."""
    lib = left._lib
    cap = left.capacity
    if out is None:
        out = {"u_right": np.zeros((n, cap), np.float32), "depth": np.zeros((n, cap), np.float32),
               "lr_match": np.zeros((n, cap), np.int32), "n_left": np.zeros(n, np.int32),
               "n_stereo": np.zeros(n, np.int32), "n_ratio": np.zeros(n, np.int32)}
    _check(lib, lib.viorb_stereo_fetch_all(left._h, n, out["u_right"].ctypes.data_as(_lib.f32p),
                                           out["depth"].ctypes.data_as(_lib.f32p), out["lr_match"].ctypes.data_as(_lib.i32p),
                                           cap, out["n_left"].ctypes.data_as(_lib.i32p),
                                           out["n_stereo"].ctypes.data_as(_lib.i32p), out["n_ratio"].ctypes.data_as(_lib.i32p)))
    return out


# ---- windowed (grid) matchers -------------------------------------------------------------------

def _bounds(b):
    return np.ascontiguousarray(b, np.float32)          # (mnMinX, mnMaxX, mnMinY, mnMaxY)


def features_in_area(kps, bounds, x, y, r, min_level=-1, max_level=-1, device=0, lib_path=None):
    lib = _lib.load(lib_path)
    kps = np.ascontiguousarray(kps)
    b = _bounds(bounds)
    out = np.zeros(max(len(kps), 1), np.int32)
    n = C.c_int32()
    _check(lib, lib.viorb_features_in_area(kps.ctypes.data, len(kps), b.ctypes.data, x, y, r, min_level, max_level,
                                           out.ctypes.data_as(_lib.i32p), len(out), C.byref(n), device))
    return out[:n.value].copy()


def search_for_initialization(k1, d1, k2, d2, bounds, prev_matched, window_size=100, nn_ratio=0.9, check_orientation=True,
                              device=0, lib_path=None):
    """FMatcher(nn_ratio, check_orientation).SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
    -> (nmatches, vnMatches12, updated vbPrevMatched).  One frame pair; see ``search_for_initialization_batch``."""
    n, m12, pm = search_for_initialization_batch([k1], [d1], [k2], [d2], bounds, [prev_matched], window_size, nn_ratio,
                                                 check_orientation, device, lib_path)
    return int(n[0]), m12[0], pm[0]


def search_for_initialization_batch(k1s, d1s, k2s, d2s, bounds, prev_matched, window_size=100, nn_ratio=0.9,
                                    check_orientation=True, device=0, lib_path=None):
    lib = _lib.load(lib_path)
    P = len(k1s)
    n1 = np.array([len(k) for k in k1s], np.int32)
    n2 = np.array([len(k) for k in k2s], np.int32)
    c1, c2 = int(n1.max(initial=0)), int(n2.max(initial=0))
    K1, K2 = np.zeros((P, max(c1, 1)), KP_DTYPE), np.zeros((P, max(c2, 1)), KP_DTYPE)
    D1, D2 = np.zeros((P, max(c1, 1), 32), np.uint8), np.zeros((P, max(c2, 1), 32), np.uint8)
    PM = np.zeros((P, max(c1, 1), 2), np.float32)
    for p in range(P):
        K1[p, :n1[p]], D1[p, :n1[p]], PM[p, :n1[p]] = k1s[p], d1s[p], prev_matched[p]
        K2[p, :n2[p]], D2[p, :n2[p]] = k2s[p], d2s[p]
    m12 = np.full((P, max(c1, 1)), -1, np.int32)
    nm = np.zeros(P, np.int32)
    b = _bounds(bounds)
    _check(lib, lib.viorb_search_for_initialization(P, c1, c2, K1.ctypes.data, D1.ctypes.data, n1.ctypes.data_as(_lib.i32p),
                                                    K2.ctypes.data, D2.ctypes.data, n2.ctypes.data_as(_lib.i32p), b.ctypes.data,
                                                    PM.ctypes.data_as(_lib.f32p), int(window_size), float(nn_ratio),
                                                    int(check_orientation), m12.ctypes.data_as(_lib.i32p),
                                                    nm.ctypes.data_as(_lib.i32p), device))
    return nm, [m12[p, :n1[p]].copy() for p in range(P)], [PM[p, :n1[p]].copy() for p in range(P)]


def search_by_projection(uv, invz, last_octave, last_angle, desc_mp, valid, observed, k2, d2, u_right2, blocked2, bounds,
                         scale_factors, th, bf, direction=0, check_orientation=True, device=0, lib_path=None, uv_right=None,
                         nleft2=None):
    """FMatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono); the caller projects.  ``uv_right`` / ``nleft2``
    select the two-camera branch (CurrentFrame.Nleft != -1): k2 / d2 / blocked2 = keypoints_ then keypointsRight_.
    -> (nmatches, assigned[n2]) with assigned: -1 untouched, -2 cleared by the rotation check, i >= 0 matched source."""
    lib = _lib.load(lib_path)
    uv = np.ascontiguousarray(uv, np.float32)
    invz = np.ascontiguousarray(invz, np.float32)
    lo = np.ascontiguousarray(last_octave, np.int32)
    la = np.ascontiguousarray(last_angle, np.float32)
    dm = np.ascontiguousarray(desc_mp, np.uint8)
    va = np.ascontiguousarray(valid, np.uint8)
    ob = np.ascontiguousarray(observed, np.uint8)
    k2 = np.ascontiguousarray(k2)
    d2 = np.ascontiguousarray(d2, np.uint8)
    ur = None if u_right2 is None else np.ascontiguousarray(u_right2, np.float32)
    bl = None if blocked2 is None else np.ascontiguousarray(blocked2, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    b = _bounds(bounds)
    n1 = np.array([len(lo)], np.int32)
    n2 = np.array([len(k2)], np.int32)
    uvr = None if uv_right is None else np.ascontiguousarray(uv_right, np.float32)
    nl2 = None if uv_right is None else np.array([int(nleft2)], np.int32)
    out = np.full(max(len(k2), 1), -1, np.int32)
    nm = np.zeros(1, np.int32)
    _check(lib, lib.viorb_search_by_projection(
        1, len(lo), len(k2), uv.ctypes.data_as(_lib.f32p), invz.ctypes.data_as(_lib.f32p), lo.ctypes.data_as(_lib.i32p),
        la.ctypes.data_as(_lib.f32p), dm.ctypes.data, va.ctypes.data, ob.ctypes.data, n1.ctypes.data_as(_lib.i32p),
        k2.ctypes.data, d2.ctypes.data, None if ur is None else ur.ctypes.data, None if bl is None else bl.ctypes.data,
        n2.ctypes.data_as(_lib.i32p), b.ctypes.data, sf.ctypes.data_as(_lib.f32p), len(sf), float(th), float(bf),
        int(direction), int(check_orientation), None if uvr is None else uvr.ctypes.data, None if nl2 is None else nl2.ctypes.data,
        out.ctypes.data_as(_lib.i32p), nm.ctypes.data_as(_lib.i32p), device))
    return int(nm[0]), out[:len(k2)].copy()


def f2f_batch(ex: FExtractor, n: int, bounds=None, window_size=100, nn_ratio=0.9, check_orientation=True, ratio=0.8, rule=1):
    """Consecutive resident frames p -> p + 1 of ``ex``'s last batch: SearchForInitialization (vbPrevMatched = F1's own
    keypoints, tracking.cpp:2286-2288) and brute-force kNN-2 + ratio.  Asynchronous; read back with ``f2f_fetch_all``."""
    b = None if bounds is None else _bounds(bounds)
    _check(ex._lib, ex._lib.viorb_f2f_batch(ex._h, int(n), None if b is None else b.ctypes.data, int(window_size), float(nn_ratio),
                                            int(check_orientation), float(ratio), int(rule)))


def f2f_fetch_all(ex: FExtractor, npairs: int, out=None):
    """-> dict(matches12 [npairs, cap], nmatches [npairs], knn_match [npairs, cap], n_ratio [npairs])."""
    cap = ex.capacity
    if out is None:
        out = {"matches12": np.zeros((npairs, cap), np.int32), "nmatches": np.zeros(npairs, np.int32),
               "knn_match": np.zeros((npairs, cap), np.int32), "n_ratio": np.zeros(npairs, np.int32)}
    _check(ex._lib, ex._lib.viorb_f2f_fetch_all(ex._h, int(npairs), out["matches12"].ctypes.data, out["nmatches"].ctypes.data,
                                                out["knn_match"].ctypes.data, out["n_ratio"].ctypes.data, cap))
    return out


def kb8_rig(cam1, cam2, R12, t12, precision=1e-6):
    """viorb_kb8_rig as 29 floats: mpCamera / mpCamera2 parameters (fx fy cx cy k0..k3), mRlr (row-major), mtlr, precision."""
    r = np.concatenate([np.asarray(cam1, np.float32).ravel(), np.asarray(cam2, np.float32).ravel(), np.asarray(R12, np.float32).ravel(),
                        np.asarray(t12, np.float32).ravel(), np.array([precision], np.float32)]).astype(np.float32)
    assert r.size == 29
    return r


def triangulate_matches(rig, pt1, pt2, sigma1, sigma2, device=0, lib_path=None):
    """-> (depth float32[n] (-1 = rejected), p3d float32[n, 3])"""
    lib = _lib.load(lib_path)
    rig = np.ascontiguousarray(rig, np.float32)
    pt1, pt2 = np.ascontiguousarray(pt1, np.float32).reshape(-1, 2), np.ascontiguousarray(pt2, np.float32).reshape(-1, 2)
    s1, s2 = np.ascontiguousarray(sigma1, np.float32), np.ascontiguousarray(sigma2, np.float32)
    n = len(pt1)
    depth, p3d = np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
    f = _lib.f32p
    _check(lib, lib.viorb_kb8_triangulate_matches(rig.ctypes.data_as(f), pt1.ctypes.data_as(f), pt2.ctypes.data_as(f), s1.ctypes.data_as(f),
                                                  s2.ctypes.data_as(f), n, depth.ctypes.data_as(f), p3d.ctypes.data_as(f), device))
    return depth, p3d


def compute_stereo_fisheye_matches(rig, kps_left, desc_left, mono_left, kps_right, desc_right, mono_right, level_sigma2, device=0,
                                   lib_path=None):
    """-> dict(left_to_right, right_to_left, depth, p3d, desc_matches, matches) = mvLeftToRightMatch, mvRightToLeftMatch, mvDepth,
    mvStereo3Dpoints, descMatches, nMatches."""
    lib = _lib.load(lib_path)
    rig = np.ascontiguousarray(rig, np.float32)
    kl, kr = np.ascontiguousarray(kps_left, KP_DTYPE), np.ascontiguousarray(kps_right, KP_DTYPE)
    dl, dr = _desc(desc_left) if len(kl) else np.zeros((0, 32), np.uint8), _desc(desc_right) if len(kr) else np.zeros((0, 32), np.uint8)
    sig = np.ascontiguousarray(level_sigma2, np.float32)
    l2r, r2l = np.zeros(len(kl), np.int32), np.zeros(len(kr), np.int32)
    depth, p3d = np.zeros(len(kl), np.float32), np.zeros((len(kl), 3), np.float32)
    nd, nm = C.c_int32(), C.c_int32()
    f, i = _lib.f32p, _lib.i32p
    _check(lib, lib.viorb_fisheye_stereo_matches(rig.ctypes.data_as(f), kl.ctypes.data, dl.ctypes.data, len(kl), int(mono_left),
                                                 kr.ctypes.data, dr.ctypes.data, len(kr), int(mono_right), sig.ctypes.data_as(f), len(sig),
                                                 l2r.ctypes.data_as(i), r2l.ctypes.data_as(i), depth.ctypes.data_as(f),
                                                 p3d.ctypes.data_as(f), C.byref(nd), C.byref(nm), device))
    return {"left_to_right": l2r, "right_to_left": r2l, "depth": depth, "p3d": p3d, "desc_matches": nd.value, "matches": nm.value}
