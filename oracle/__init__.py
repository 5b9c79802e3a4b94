"""ctypes binding of the CPU oracle (oracle/orb_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(vi_slam_b200) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "orb_oracle.cpp")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-s"] + (["-B"] if force else []), check=True)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        u8p, i32p, f32p = C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float)
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_extractor_destroy.argtypes = [C.c_void_p]
        L.orc_tables.argtypes = [C.c_void_p, f32p, f32p, f32p, f32p, i32p, i32p]
        L.orc_extract.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p, u8p, C.c_int, i32p]
        L.orc_level_size.argtypes = [C.c_void_p, C.c_int, i32p, i32p]
        L.orc_level_image.argtypes = [C.c_void_p, C.c_int, u8p]
        L.orc_level_blurred.argtypes = [C.c_void_p, C.c_int, u8p]
        L.orc_level_candidates.argtypes = [C.c_void_p, C.c_int, i32p, C.c_int]
        L.orc_level_keypoints.argtypes = [C.c_void_p, C.c_int, i32p, f32p, C.c_int]
        L.orc_tie_events.restype = C.c_long
        L.orc_tie_events.argtypes = [C.c_void_p]
        L.orc_resize_linear.argtypes = [u8p, C.c_int, C.c_int, C.c_int, u8p, C.c_int, C.c_int, C.c_int]
        L.orc_gaussian7.argtypes = [u8p, C.c_int, C.c_int, C.c_int, u8p, C.c_int]
        L.orc_fast.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, i32p, C.c_int]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_fast_atan2_array.argtypes = [f32p, f32p, f32p, C.c_int]
        L.orc_sincos_deg.argtypes = [f32p, f32p, f32p, C.c_int]
        L.orc_descriptor_distance.argtypes = [u8p, u8p]
        L.orc_knn2.argtypes = [u8p, C.c_int, u8p, C.c_int, i32p, i32p, C.c_int]
        L.orc_ratio_filter.argtypes = [i32p, i32p, C.c_int, C.c_double, C.c_int, i32p]
        L.orc_dedup_by_train.argtypes = [i32p, C.c_int]
        L.orc_stereo_match.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, u8p, C.c_int, C.c_void_p, u8p, C.c_int,
                                       C.c_float, C.c_float, f32p, f32p, i32p, i32p]
        L.orc_bench_frames.restype = C.c_long
        L.orc_bench_frames.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, u8p, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.POINTER(C.c_long)]
        L.orc_features_in_area.argtypes = [C.c_void_p, C.c_int, f32p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int,
                                           i32p, C.c_int]
        L.orc_search_for_initialization.argtypes = [C.c_void_p, u8p, C.c_int, C.c_void_p, u8p, C.c_int, f32p, f32p, C.c_int,
                                                    C.c_float, C.c_int, i32p]
        L.orc_search_by_projection.argtypes = [f32p, f32p, i32p, f32p, u8p, u8p, u8p, C.c_int, C.c_void_p, u8p, f32p, u8p,
                                               C.c_int, f32p, f32p, C.c_float, C.c_float, C.c_int, C.c_int, i32p]
        L.orc_search_by_projection_rig.argtypes = [f32p, f32p, f32p, i32p, f32p, u8p, u8p, u8p, C.c_int, C.c_void_p, u8p, u8p, C.c_int,
                                                   C.c_int, f32p, f32p, C.c_float, C.c_int, C.c_int, i32p]
        L.orc_undistort_points.argtypes = [f32p, C.c_int, f32p, f32p, C.c_int, f32p]
        L.orc_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, f32p, f32p, C.c_int, C.c_void_p]
        L.orc_image_bounds.argtypes = [C.c_int, C.c_int, f32p, f32p, C.c_int, f32p]
        f64p, u32p = C.POINTER(C.c_double), C.POINTER(C.c_uint32)
        L.orc_vocab_create.restype = C.c_void_p
        L.orc_vocab_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, i32p, i32p, u8p, f64p, i32p]
        L.orc_vocab_destroy.argtypes = [C.c_void_p]
        L.orc_bow_transform.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, u32p, f64p, i32p, u32p, i32p, u32p, i32p]
        L.orc_kb8_triangulate.argtypes = [f32p, f32p, f32p, f32p, f32p, C.c_int, f32p, f32p, f32p]
        _lib = L
    return _lib


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _i32(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def _f32(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class OracleExtractor:
    """CPU restatement of FExtractor (reference include/vi_slam/geometry/fextractor.h:26-64)."""

    def __init__(self, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = lib().orc_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        n = nlevels
        self.scale = np.zeros(n, np.float32)
        self.inv_scale = np.zeros(n, np.float32)
        self.sigma2 = np.zeros(n, np.float32)
        self.inv_sigma2 = np.zeros(n, np.float32)
        self.features_per_level = np.zeros(n, np.int32)
        self.umax = np.zeros(16, np.int32)
        lib().orc_tables(self._h, _f32(self.scale), _f32(self.inv_scale), _f32(self.sigma2), _f32(self.inv_sigma2),
                         _i32(self.features_per_level), _i32(self.umax))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h)
            self._h = None

    def compute(self, image: np.ndarray, lapping=(0, 0)):
        """-> (mono_index, keypoints[KP_DTYPE], descriptors uint8[n,32]); mono_index -1 on empty image."""
        if image is None or image.size == 0:
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2
        h, w = image.shape
        step = image.strides[0]
        assert image.strides[1] == 1
        cap = self.nfeatures + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mono = C.c_int32(0)
        n = lib().orc_extract(self._h, _u8(image), w, h, step, int(lapping[0]), int(lapping[1]),
                              kps.ctypes.data, _u8(desc), cap, C.byref(mono))
        if n < 0:
            return -1, kps[:0], desc[:0]
        assert n <= cap
        return mono.value, kps[:n].copy(), desc[:n].copy()

    def level_image(self, l):
        w, h = C.c_int32(), C.c_int32()
        lib().orc_level_size(self._h, l, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        lib().orc_level_image(self._h, l, _u8(out))
        return out

    def level_blurred(self, l):
        w, h = C.c_int32(), C.c_int32()
        lib().orc_level_size(self._h, l, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        if lib().orc_level_blurred(self._h, l, _u8(out)) != 0:
            return None
        return out

    def level_candidates(self, l):
        cap = 1 << 20
        buf = np.zeros((cap, 3), np.int32)
        n = lib().orc_level_candidates(self._h, l, _i32(buf), cap)
        return buf[:n].copy()

    def level_keypoints(self, l):
        cap = self.nfeatures + 64
        buf = np.zeros((cap, 3), np.int32)
        ang = np.zeros(cap, np.float32)
        n = lib().orc_level_keypoints(self._h, l, _i32(buf), _f32(ang), cap)
        return buf[:n].copy(), ang[:n].copy()

    @property
    def tie_events(self):
        return lib().orc_tie_events(self._h)


def resize_linear(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    src = np.ascontiguousarray(src)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orc_resize_linear(_u8(src), src.shape[1], src.shape[0], src.shape[1], _u8(dst), dw, dh, dw)
    return dst


def gaussian7(src: np.ndarray) -> np.ndarray:
    src = np.ascontiguousarray(src)
    dst = np.zeros_like(src)
    lib().orc_gaussian7(_u8(src), src.shape[1], src.shape[0], src.shape[1], _u8(dst), src.shape[1])
    return dst


def fast(img: np.ndarray, threshold: int) -> np.ndarray:
    img = np.ascontiguousarray(img)
    cap = img.size
    buf = np.zeros((cap, 3), np.int32)
    n = lib().orc_fast(_u8(img), img.shape[1], img.shape[0], img.shape[1], threshold, _i32(buf), cap)
    return buf[:n].copy()


def fast_atan2(y: np.ndarray, x: np.ndarray) -> np.ndarray:
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(y)
    lib().orc_fast_atan2_array(_f32(y), _f32(x), _f32(out), y.size)
    return out


def sincos_deg(deg: np.ndarray):
    deg = np.ascontiguousarray(deg, np.float32)
    c = np.zeros_like(deg)
    s = np.zeros_like(deg)
    lib().orc_sincos_deg(_f32(deg), _f32(c), _f32(s), deg.size)
    return c, s


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> int:
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().orc_descriptor_distance(_u8(a), _u8(b))


def knn2(q: np.ndarray, t: np.ndarray, nthreads: int = 1):
    q = np.ascontiguousarray(q, np.uint8)
    t = np.ascontiguousarray(t, np.uint8)
    idx = np.full((q.shape[0], 2), -1, np.int32)
    dist = np.full((q.shape[0], 2), -1, np.int32)
    lib().orc_knn2(_u8(q), q.shape[0], _u8(t), t.shape[0], _i32(idx), _i32(dist), nthreads)
    return idx, dist


def ratio_filter(idx, dist, ratio: float, rule: int) -> np.ndarray:
    idx = np.ascontiguousarray(idx, np.int32)
    dist = np.ascontiguousarray(dist, np.int32)
    out = np.zeros((idx.shape[0], 3), np.int32)
    n = lib().orc_ratio_filter(_i32(idx), _i32(dist), idx.shape[0], ratio, rule, _i32(out))
    return out[:n].copy()


def dedup_by_train(m: np.ndarray) -> np.ndarray:
    m = np.ascontiguousarray(m, np.int32).copy()
    n = lib().orc_dedup_by_train(_i32(m), m.shape[0])
    return m[:n].copy()


def stereo_match(ex_l: OracleExtractor, ex_r: OracleExtractor, kl, dl, kr, dr, bf: float, b: float):
    kl = np.ascontiguousarray(kl)
    kr = np.ascontiguousarray(kr)
    dl = np.ascontiguousarray(dl)
    dr = np.ascontiguousarray(dr)
    n = kl.shape[0]
    u_right = np.zeros(n, np.float32)
    depth = np.zeros(n, np.float32)
    bidx = np.zeros(n, np.int32)
    bdist = np.zeros(n, np.int32)
    lib().orc_stereo_match(ex_l._h, ex_r._h, kl.ctypes.data, _u8(dl), n, kr.ctypes.data, _u8(dr), kr.shape[0],
                           bf, b, _f32(u_right), _f32(depth), _i32(bidx), _i32(bdist))
    return u_right, depth, bidx, bdist


def bench_frames(frames: np.ndarray, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7,
                 lapping=(0, 0), stereo=False, bf=386.1448, b=0.5371657, nthreads=1):
    """Timed CPU baseline body: returns (total keypoints, matches)."""
    frames = np.ascontiguousarray(frames)
    n, h, w = frames.shape
    nm = C.c_long(0)
    tot = lib().orc_bench_frames(nfeatures, scale_factor, nlevels, ini_th, min_th, _u8(frames), n, w, h,
                                 int(lapping[0]), int(lapping[1]), int(stereo), bf, b, nthreads, C.byref(nm))
    return tot, nm.value


def features_in_area(kps, bounds, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea (frame.cpp:678-744) on a grid built by AssignFeaturesToGrid (frame.cpp:386-414)."""
    kps = np.ascontiguousarray(kps)
    b = np.ascontiguousarray(bounds, np.float32)
    out = np.zeros(max(len(kps), 1), np.int32)
    n = lib().orc_features_in_area(kps.ctypes.data, len(kps), _f32(b), x, y, r, min_level, max_level, _i32(out), len(out))
    return out[:n].copy()


def search_for_initialization(k1, d1, k2, d2, bounds, prev_matched, window_size=100, nn_ratio=0.9, check_orientation=True):
    """FMatcher::SearchForInitialization (fmatcher.cpp:983-1098) -> (nmatches, vnMatches12, updated vbPrevMatched)."""
    k1, k2 = np.ascontiguousarray(k1), np.ascontiguousarray(k2)
    d1, d2 = np.ascontiguousarray(d1, np.uint8), np.ascontiguousarray(d2, np.uint8)
    b = np.ascontiguousarray(bounds, np.float32)
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.full(len(k1), -1, np.int32)
    n = lib().orc_search_for_initialization(k1.ctypes.data, _u8(d1), len(k1), k2.ctypes.data, _u8(d2), len(k2), _f32(b), _f32(pm),
                                            int(window_size), float(nn_ratio), int(check_orientation), _i32(m12))
    return n, m12, pm


def search_by_projection(uv, invz, last_octave, last_angle, desc_mp, valid, observed, k2, d2, u_right2, blocked2, bounds,
                         scale_factors, th, bf, direction=0, check_orientation=True):
    """FMatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (fmatcher.cpp:2471-2688), single-camera branch;
    the projection is done by the caller.  -> (nmatches, assigned[n2])."""
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
    b = np.ascontiguousarray(bounds, np.float32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    out = np.full(len(k2), -1, np.int32)
    n = lib().orc_search_by_projection(_f32(uv), _f32(invz), _i32(lo), _f32(la), _u8(dm), _u8(va), _u8(ob), len(lo),
                                       k2.ctypes.data, _u8(d2), None if ur is None else _f32(ur),
                                       None if bl is None else _u8(bl), len(k2), _f32(b), _f32(sf), float(th), float(bf),
                                       int(direction), int(check_orientation), _i32(out))
    return n, out


def search_by_projection_rig(uv, uv_right, invz, last_octave, last_angle, desc_mp, valid, observed, k2, d2, nleft2, blocked2, bounds,
                             scale_factors, th, direction=0, check_orientation=True):
    """FMatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) for a two-camera rig (CurrentFrame.Nleft != -1,
    fmatcher.cpp:2471-2688 incl. :2593-2660): k2 / d2 / blocked2 = left keypoints followed by the right ones, nleft2 = Nleft.
    -> (nmatches, assigned[N])."""
    uv = np.ascontiguousarray(uv, np.float32)
    uvr = np.ascontiguousarray(uv_right, np.float32)
    invz = np.ascontiguousarray(invz, np.float32)
    lo = np.ascontiguousarray(last_octave, np.int32)
    la = np.ascontiguousarray(last_angle, np.float32)
    dm = np.ascontiguousarray(desc_mp, np.uint8)
    va = np.ascontiguousarray(valid, np.uint8)
    ob = np.ascontiguousarray(observed, np.uint8)
    k2 = np.ascontiguousarray(k2)
    d2 = np.ascontiguousarray(d2, np.uint8)
    bl = None if blocked2 is None else np.ascontiguousarray(blocked2, np.uint8)
    b = np.ascontiguousarray(bounds, np.float32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    out = np.full(len(k2), -1, np.int32)
    n = lib().orc_search_by_projection_rig(_f32(uv), _f32(uvr), _f32(invz), _i32(lo), _f32(la), _u8(dm), _u8(va), _u8(ob), len(lo),
                                           k2.ctypes.data, _u8(d2), None if bl is None else _u8(bl), len(k2), int(nleft2), _f32(b),
                                           _f32(sf), float(th), int(direction), int(check_orientation), _i32(out))
    return n, out


def undistort_points(xy, cam, dist):
    """cv::undistortPoints(pts, pts, mK, mDistCoef, Mat(), mK) as Frame::UndistortKeyPoints calls it (frame.cpp:779)."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    cam = np.ascontiguousarray(cam, np.float32)
    dist = np.ascontiguousarray(dist, np.float32)
    out = np.empty_like(xy)
    lib().orc_undistort_points(_f32(xy), len(xy), _f32(cam), _f32(dist), len(dist), _f32(out))
    return out


def undistort_keypoints(kps, cam, dist):
    """Frame::UndistortKeyPoints (frame.cpp:758-791): keypoints_ -> ukeypoints_."""
    kps = np.ascontiguousarray(kps)
    cam = np.ascontiguousarray(cam, np.float32)
    dist = np.ascontiguousarray(dist, np.float32)
    out = np.empty_like(kps)
    lib().orc_undistort_keypoints(kps.ctypes.data, len(kps), _f32(cam), _f32(dist), len(dist), out.ctypes.data)
    return out


def image_bounds(cols, rows, cam, dist):
    """Frame::ComputeImageBounds (frame.cpp:793-821) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    cam = np.ascontiguousarray(cam, np.float32)
    dist = np.ascontiguousarray(dist, np.float32)
    out = np.zeros(4, np.float32)
    lib().orc_image_bounds(int(cols), int(rows), _f32(cam), _f32(dist), len(dist), _f32(out))
    return out


class OracleVocabulary:
    """DBoW3::Vocabulary restricted to transform() (thirdparty/DBoW3/DBoW3/src/Vocabulary.cpp:754-876) on a tree given as flat
    arrays (see vi_slam_b200.synthetic.make_vocabulary)."""

    def __init__(self, voc):
        self._keep = [np.ascontiguousarray(voc["child_start"], np.int32), np.ascontiguousarray(voc["children"], np.int32),
                      np.ascontiguousarray(voc["descriptors"], np.uint8), np.ascontiguousarray(voc["weight"], np.float64),
                      np.ascontiguousarray(voc["word_id"], np.int32)]
        cs, ch, de, we, wi = self._keep
        self._h = lib().orc_vocab_create(len(we), int(voc["levels"]), int(voc["weighting"]), int(voc["scoring"]), _i32(cs), _i32(ch),
                                         _u8(de), we.ctypes.data_as(C.POINTER(C.c_double)), _i32(wi))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_vocab_destroy(self._h)
            self._h = None

    def transform(self, desc, levelsup=4):
        """-> (bow_word uint32[m], bow_value float64[m], fv_node uint32[k], fv_start int32[k+1], fv_feature uint32[...])"""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        bw, bv = np.zeros(n + 1, np.uint32), np.zeros(n + 1, np.float64)
        fn, fs, ff = np.zeros(n + 1, np.uint32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.uint32)
        nb, nf = C.c_int32(), C.c_int32()
        u32 = C.POINTER(C.c_uint32)
        lib().orc_bow_transform(self._h, _u8(desc), n, int(levelsup), bw.ctypes.data_as(u32), bv.ctypes.data_as(C.POINTER(C.c_double)),
                                C.byref(nb), fn.ctypes.data_as(u32), _i32(fs), ff.ctypes.data_as(u32), C.byref(nf))
        return bw[:nb.value].copy(), bv[:nb.value].copy(), fn[:nf.value].copy(), fs[:nf.value + 1].copy(), ff[:fs[nf.value]].copy()


def kb8_rig(prm1, prm2, R12, t12, precision=1e-6):
    """The 29 floats the KB8 functions take: cam1 (fx fy cx cy k0..k3), cam2, R12 row-major, t12, Newton precision."""
    return np.concatenate([np.asarray(prm1, np.float32).ravel(), np.asarray(prm2, np.float32).ravel(), np.asarray(R12, np.float32).ravel(),
                           np.asarray(t12, np.float32).ravel(), np.array([precision], np.float32)]).astype(np.float32)


def kb8_triangulate(rig, pt1, pt2, sigma1, sigma2):
    """KannalaBrandt8::TriangulateMatches (kannalabrandt8.cpp:324-393) per pair -> (depth, p3d, margin)."""
    rig = np.ascontiguousarray(rig, np.float32)
    pt1, pt2 = np.ascontiguousarray(pt1, np.float32).reshape(-1, 2), np.ascontiguousarray(pt2, np.float32).reshape(-1, 2)
    s1, s2 = np.ascontiguousarray(sigma1, np.float32), np.ascontiguousarray(sigma2, np.float32)
    n = len(pt1)
    depth, p3d, margin = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.float32)
    lib().orc_kb8_triangulate(_f32(rig), _f32(pt1), _f32(pt2), _f32(s1), _f32(s2), n, _f32(depth), _f32(p3d), _f32(margin))
    return depth, p3d, margin


def stereo_fisheye_matches(rig, kps_left, desc_left, mono_left, kps_right, desc_right, mono_right, level_sigma2):
    """Frame::ComputeStereoFishEyeMatches (frame.cpp:1148-1189) composed from the oracle's knnMatch and TriangulateMatches.
    -> dict like vi_slam_b200.matcher.compute_stereo_fisheye_matches, plus the per-query triangulation margin."""
    nl, nr = len(kps_left), len(kps_right)
    out = {"left_to_right": np.full(nl, -1, np.int32), "right_to_left": np.full(nr, -1, np.int32), "depth": np.full(nl, -1, np.float32),
           "p3d": np.zeros((nl, 3), np.float32), "desc_matches": 0, "matches": 0, "margin": np.full(nl, np.inf, np.float32)}
    nq, nt = nl - mono_left, nr - mono_right
    if nq == 0 or nt < 2:
        return out
    idx, dist = knn2(desc_left[mono_left:], desc_right[mono_right:])
    sig = np.asarray(level_sigma2, np.float32)
    for q in range(nq):
        if not (np.float32(dist[q, 0]) < float(np.float32(dist[q, 1])) * 0.7):
            continue
        out["desc_matches"] += 1
        il, ir = q + mono_left, int(idx[q, 0]) + mono_right
        a, b = kps_left[il], kps_right[ir]
        d, p, m = kb8_triangulate(rig, [[a["x"], a["y"]]], [[b["x"], b["y"]]], [sig[a["octave"]]], [sig[b["octave"]]])
        out["margin"][il] = m[0]
        if d[0] > np.float32(0.0001):
            out["left_to_right"][il] = ir
            out["right_to_left"][ir] = il
            out["p3d"][il] = p[0]
            out["depth"][il] = d[0]
            out["matches"] += 1
    return out
