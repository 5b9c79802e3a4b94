"""ctypes binding of oracle/_ref/liborb_ref.so: the reference's OWN, unmodified
src/geometry/fextractor.cpp compiled against the OpenCV stand-in of oracle/ref_compat/ (oracle/ref.mk).

TEST INFRASTRUCTURE ONLY, like the rest of oracle/.  `build()` needs /root/reference (this container);
on the GPU box the prebuilt .so travels with the snapshot and `available()` just finds it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import KP_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "liborb_ref.so")
REFERENCE = os.environ.get("VIORB_REFERENCE_DIR", "/root/reference")
REF_SOURCE = os.path.join(REFERENCE, "src", "geometry", "fextractor.cpp")


def build(force: bool = False) -> str | None:
    """Compile the reference translation unit where it lies. Returns the library path, or None when the
    reference checkout is absent (GPU box) and no prebuilt library travelled."""
    if os.path.exists(REF_SOURCE):
        deps = [REF_SOURCE, os.path.join(_HERE, "orb_oracle.cpp"), os.path.join(_HERE, "ref.mk")] + \
               [os.path.join(_HERE, "ref_compat", f) for f in ("cv_compat.hpp", "cv_compat.cpp", "ref_driver.cpp", "ref_slices.cpp", "ref_dbow3.cpp", "ref_kb8.cpp", "ref_matchfeatures.cpp")] + \
               [os.path.join(REFERENCE, "src", "datastructures", "frame.cpp"), os.path.join(REFERENCE, "src", "geometry", "fmatcher.cpp")]
        stale = not os.path.exists(LIB_PATH) or any(os.path.getmtime(d) > os.path.getmtime(LIB_PATH) for d in deps)
        if force or stale:
            subprocess.run(["make", "-C", _HERE, "-f", "ref.mk", "-s", "REF=" + REFERENCE] + (["-B"] if force else []), check=True)
    return LIB_PATH if os.path.exists(LIB_PATH) else None


def available() -> bool:
    return build() is not None


_lib = None


def lib():
    global _lib
    if _lib is None:
        if build() is None:
            raise RuntimeError("oracle/_ref/liborb_ref.so is absent and /root/reference is not here to build it")
        L = C.CDLL(LIB_PATH)
        u8p, i32p, f32p = C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float)
        L.orb_ref_create.restype = C.c_void_p
        L.orb_ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orb_ref_destroy.argtypes = [C.c_void_p]
        L.orb_ref_tables.argtypes = [C.c_void_p, f32p, f32p, f32p, f32p, i32p, i32p]
        L.orb_ref_compute.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, u8p,
                                      C.c_int, i32p, C.c_int]
        L.orb_ref_compute_plain.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, u8p, C.c_int, i32p]
        L.orb_ref_consistent.argtypes = [C.c_void_p]
        L.orb_ref_level_size.argtypes = [C.c_void_p, C.c_int, i32p, i32p]
        L.orb_ref_level_image.argtypes = [C.c_void_p, C.c_int, u8p]
        L.orb_ref_level_candidates.argtypes = [C.c_void_p, C.c_int, i32p, C.c_int]
        L.orb_ref_level_keypoints.argtypes = [C.c_void_p, C.c_int, i32p, f32p, C.c_int]
        L.orb_ref_arena_fallbacks.restype = C.c_long
        # ref_slices.cpp: function bodies of frame.cpp / fmatcher.cpp
        L.orb_refs_descriptor_distance.argtypes = [u8p, u8p]
        L.orb_refs_stereo.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, u8p, C.c_void_p, C.c_int, u8p, C.c_float, C.c_float,
                                      f32p, f32p]
        L.orb_refs_features_in_area.argtypes = [C.c_void_p, C.c_int, f32p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, i32p, C.c_int]
        L.orb_refs_search_for_initialization.argtypes = [C.c_void_p, C.c_int, u8p, C.c_void_p, C.c_int, u8p, f32p, f32p, C.c_int,
                                                         C.c_float, C.c_int, i32p]
        L.orb_refs_search_by_projection.argtypes = [C.c_int, C.c_void_p, f32p, u8p, u8p, u8p, C.c_void_p, C.c_int, u8p, f32p, u8p, C.c_int,
                                                    f32p, f32p, f32p, f32p, f32p, f32p, C.c_int, C.c_float, C.c_int, C.c_float, C.c_float,
                                                    C.c_float, C.c_int, i32p, f32p, f32p, f32p, u8p, i32p]
        L.orb_refs_bench_frames.restype = C.c_long
        L.orb_refs_bench_frames.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_int, C.c_float, C.c_float, C.c_int, C.POINTER(C.c_long)]
        L.orb_refs_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, f32p, f32p, C.c_int, C.c_void_p]
        L.orb_refs_image_bounds.argtypes = [C.c_int, C.c_int, f32p, f32p, C.c_int, f32p]
        # ref_kb8.cpp: KannalaBrandt8 / fisheye stereo
        L.orb_refk_triangulate.argtypes = [f32p, f32p, f32p, f32p, f32p, C.c_int, f32p, f32p]
        L.orb_refk_fisheye_matches.argtypes = [f32p, C.c_void_p, u8p, C.c_int, C.c_int, C.c_void_p, u8p, C.c_int, C.c_int, f32p, C.c_int,
                                               i32p, i32p, f32p, f32p]
        # ref_dbow3.cpp: DBoW3's own transform
        L.orb_refd_vocab_create.restype = C.c_void_p
        L.orb_refd_vocab_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, i32p, i32p, u8p, C.POINTER(C.c_double), i32p]
        L.orb_refd_vocab_destroy.argtypes = [C.c_void_p]
        u32p = C.POINTER(C.c_uint32)
        L.orb_refd_transform.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, u32p, C.POINTER(C.c_double), i32p, u32p, i32p, u32p, i32p]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class RefExtractor:
    """vi_slam::geometry::FExtractor itself (reference include/vi_slam/geometry/fextractor.h:26-64).

    monotone=True serves the reference's allocations from a never-reusing bump arena, so the (size, pointer)
    sort of DistributeOctTree (fextractor.cpp:675) breaks ties by allocation order; monotone=False is glibc malloc."""

    def __init__(self, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, monotone=True):
        self.nfeatures, self.nlevels, self.monotone = nfeatures, nlevels, monotone
        self._h = lib().orb_ref_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        n = nlevels
        self.scale = np.zeros(n, np.float32)
        self.inv_scale = np.zeros(n, np.float32)
        self.sigma2 = np.zeros(n, np.float32)
        self.inv_sigma2 = np.zeros(n, np.float32)
        self.features_per_level = np.zeros(n, np.int32)
        self.umax = np.zeros(16, np.int32)
        lib().orb_ref_tables(self._h, _p(self.scale, C.c_float), _p(self.inv_scale, C.c_float), _p(self.sigma2, C.c_float),
                             _p(self.inv_sigma2, C.c_float), _p(self.features_per_level, C.c_int32), _p(self.umax, C.c_int32))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_ref_destroy(self._h)
            self._h = None

    def compute(self, image: np.ndarray, lapping=(0, 0)):
        """-> (mono_index, keypoints[KP_DTYPE], descriptors uint8[n,32]); mono_index -1 on an empty image."""
        if image is None or image.size == 0:
            mono = C.c_int32(0)
            n = lib().orb_ref_compute(self._h, None, 0, 0, 0, int(lapping[0]), int(lapping[1]), None, None, 0,
                                      C.byref(mono), int(self.monotone))
            assert n == -1 and mono.value == -1
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2 and image.strides[1] == 1
        h, w = image.shape
        cap = self.nfeatures + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mono = C.c_int32(0)
        n = lib().orb_ref_compute(self._h, _p(image, C.c_uint8), w, h, image.strides[0], int(lapping[0]), int(lapping[1]),
                                  kps.ctypes.data, _p(desc, C.c_uint8), cap, C.byref(mono), int(self.monotone))
        assert 0 <= n <= cap
        return mono.value, kps[:n].copy(), desc[:n].copy()

    def compute_plain(self, image: np.ndarray, lapping=(0, 0)):
        """FExtractor::compute only (no probes, glibc malloc, thread-safe across instances): the timed CPU baseline."""
        h, w = image.shape
        cap = self.nfeatures + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mono = C.c_int32(0)
        n = lib().orb_ref_compute_plain(self._h, _p(image, C.c_uint8), w, h, image.strides[0], int(lapping[0]), int(lapping[1]),
                                        kps.ctypes.data, _p(desc, C.c_uint8), cap, C.byref(mono))
        return mono.value, kps[:n], desc[:n]

    @property
    def consistent(self):
        """False if the second (probe) pass kept a different number of keypoints than compute() did — only
        possible with monotone=False, where pointer ties may fall differently between the two passes."""
        return bool(lib().orb_ref_consistent(self._h))

    def level_image(self, l):
        w, h = C.c_int32(), C.c_int32()
        lib().orb_ref_level_size(self._h, l, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        lib().orb_ref_level_image(self._h, l, _p(out, C.c_uint8))
        return out

    def level_candidates(self, l):
        cap = 1 << 20
        buf = np.zeros((cap, 3), np.int32)
        n = lib().orb_ref_level_candidates(self._h, l, _p(buf, C.c_int32), cap)
        return buf[:n].copy()

    def level_keypoints(self, l):
        cap = self.nfeatures + 64
        buf = np.zeros((cap, 3), np.int32)
        ang = np.zeros(cap, np.float32)
        n = lib().orb_ref_level_keypoints(self._h, l, _p(buf, C.c_int32), _p(ang, C.c_float), cap)
        return buf[:n].copy(), ang[:n].copy()


def arena_fallbacks() -> int:
    return lib().orb_ref_arena_fallbacks()


# ---- the reference's own matcher functions (oracle/ref_compat/ref_slices.cpp) -------------------------------------------

def descriptor_distance(a: np.ndarray, b: np.ndarray) -> int:
    """FMatcher::DescriptorDistance (fmatcher.cpp:2859-2875), the reference's own text."""
    a = np.ascontiguousarray(a, np.uint8).reshape(32)
    b = np.ascontiguousarray(b, np.uint8).reshape(32)
    return int(lib().orb_refs_descriptor_distance(_p(a, C.c_uint8), _p(b, C.c_uint8)))


def stereo_match(ref_l: RefExtractor, ref_r: RefExtractor, kl, dl, kr, dr, bf: float, b: float):
    """Frame::ComputeStereoMatches (frame.cpp:823-997), the reference's own text, on the pyramids the two reference
    extractors hold after their compute() calls -> (mvuRight, mvDepth).  The caller must make sure at least one match
    survives (the reference reads an empty vector otherwise, frame.cpp:984)."""
    kl = np.ascontiguousarray(kl, KP_DTYPE); kr = np.ascontiguousarray(kr, KP_DTYPE)
    dl = np.ascontiguousarray(dl, np.uint8).reshape(-1, 32); dr = np.ascontiguousarray(dr, np.uint8).reshape(-1, 32)
    ur = np.zeros(len(kl), np.float32)
    dp = np.zeros(len(kl), np.float32)
    rc = lib().orb_refs_stereo(ref_l._h, ref_r._h, kl.ctypes.data, len(kl), _p(dl, C.c_uint8), kr.ctypes.data, len(kr), _p(dr, C.c_uint8),
                               float(bf), float(b), _p(ur, C.c_float), _p(dp, C.c_float))
    assert rc == 0
    return ur, dp


def features_in_area(kps, bounds, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea (frame.cpp:678-744) on the grid Frame::AssignFeaturesToGrid (:386-414) builds."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    bd = np.ascontiguousarray(bounds, np.float32)
    out = np.zeros(max(len(kps), 1), np.int32)
    n = lib().orb_refs_features_in_area(kps.ctypes.data, len(kps), _p(bd, C.c_float), float(x), float(y), float(r), int(min_level),
                                        int(max_level), _p(out, C.c_int32), len(out))
    return out[:n].copy()


def search_for_initialization(k1, d1, k2, d2, bounds, prev_matched, window_size=100, nn_ratio=0.9, check_orientation=True):
    """FMatcher(nn_ratio, check_orientation).SearchForInitialization (fmatcher.cpp:983-1098 with :2813-2854, :2859-2875 and
    the frame grid of frame.cpp:386-414, 678-756) -> (nmatches, vnMatches12, vbPrevMatched)."""
    k1 = np.ascontiguousarray(k1, KP_DTYPE); k2 = np.ascontiguousarray(k2, KP_DTYPE)
    d1 = np.ascontiguousarray(d1, np.uint8).reshape(-1, 32); d2 = np.ascontiguousarray(d2, np.uint8).reshape(-1, 32)
    bd = np.ascontiguousarray(bounds, np.float32)
    prev = np.ascontiguousarray(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.zeros(len(k1), np.int32)
    n = lib().orb_refs_search_for_initialization(k1.ctypes.data, len(k1), _p(d1, C.c_uint8), k2.ctypes.data, len(k2), _p(d2, C.c_uint8),
                                                 _p(bd, C.c_float), _p(prev, C.c_float), int(window_size), float(nn_ratio),
                                                 int(bool(check_orientation)), _p(m12, C.c_int32))
    return int(n), m12, prev


def search_by_projection(k_last, xw, desc_mp, outlier, observed, k2, d2, u_right2, blocked2, nleft2, bounds, tcw_cur, tcw_last, trl, cam,
                         scale_factors, th, mono, bf, b, check_orientation=True):
    """FMatcher(0.9, check_orientation).SearchByProjection(CurrentFrame, LastFrame, th, bMono) (fmatcher.cpp:2471-2687, the
    reference's own text incl. the two-camera branch :2593-2660) on a synthetic scene: last-frame keypoints with world points
    `xw` (NaN = no map point), world -> camera poses (4x4), a pinhole camera (Pinhole::project, pinhole.cpp:13-26), `nleft2` < 0
    for one camera or Nleft of a rig with `trl` (3x4).  Returns (nmatches, assigned[n2], host) where `host` holds what the
    caller of the device API computes on the host - uv, uv_right, invz, valid, direction - as the reference's own
    expressions evaluate them."""
    kl = np.ascontiguousarray(k_last, KP_DTYPE); k2 = np.ascontiguousarray(k2, KP_DTYPE)
    n1, n2 = len(kl), len(k2)
    f = lambda a: np.ascontiguousarray(a, np.float32)
    u = lambda a: np.ascontiguousarray(a, np.uint8)
    xw, dmp, d2 = f(xw).reshape(n1, 3), u(desc_mp).reshape(n1, 32), u(d2).reshape(n2, 32)
    out, obs = u(outlier), u(observed)
    ur = None if u_right2 is None else f(u_right2)
    bl = None if blocked2 is None else u(blocked2)
    bd, tc, tl, cm, sf = f(bounds), f(tcw_cur).reshape(16), f(tcw_last).reshape(16), f(cam), f(scale_factors)
    tr = None if trl is None else f(trl).reshape(12)
    asg = np.zeros(n2, np.int32)
    uv, uvr, invz = np.zeros((n1, 2), np.float32), np.zeros((n1, 2), np.float32), np.zeros(n1, np.float32)
    valid = np.zeros(n1, np.uint8)
    direction = C.c_int32(0)
    n = lib().orb_refs_search_by_projection(
        n1, kl.ctypes.data, _p(xw, C.c_float), _p(dmp, C.c_uint8), _p(out, C.c_uint8), _p(obs, C.c_uint8), k2.ctypes.data, n2,
        _p(d2, C.c_uint8), None if ur is None else _p(ur, C.c_float), None if bl is None else _p(bl, C.c_uint8), int(nleft2),
        _p(bd, C.c_float), _p(tc, C.c_float), _p(tl, C.c_float), None if tr is None else _p(tr, C.c_float), _p(cm, C.c_float),
        _p(sf, C.c_float), len(sf), float(th), int(bool(mono)), float(bf), float(b), 0.9, int(bool(check_orientation)),
        _p(asg, C.c_int32), _p(uv, C.c_float), _p(uvr, C.c_float), _p(invz, C.c_float), _p(valid, C.c_uint8), C.byref(direction))
    return int(n), asg, {"uv": uv, "uv_right": uvr, "invz": invz, "valid": valid, "direction": int(direction.value)}


class RefVocabulary:
    """DBoW3::Vocabulary::transform - DBoW3's own text (Vocabulary.cpp:754-876, DescManip.cpp:92-130, BowVector.cpp and
    FeatureVector.cpp whole; oracle/ref_compat/ref_dbow3.cpp) on a tree given as flat arrays (synthetic.make_vocabulary)."""

    def __init__(self, voc):
        self._keep = [np.ascontiguousarray(voc["child_start"], np.int32), np.ascontiguousarray(voc["children"], np.int32),
                      np.ascontiguousarray(voc["descriptors"], np.uint8), np.ascontiguousarray(voc["weight"], np.float64),
                      np.ascontiguousarray(voc["word_id"], np.int32)]
        cs, ch, de, we, wi = self._keep
        self._h = lib().orb_refd_vocab_create(len(we), int(voc["levels"]), int(voc["weighting"]), int(voc["scoring"]), _p(cs, C.c_int32),
                                              _p(ch, C.c_int32), _p(de, C.c_uint8), _p(we, C.c_double), _p(wi, C.c_int32))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_refd_vocab_destroy(self._h)
            self._h = None

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        bw, bv = np.zeros(n + 1, np.uint32), np.zeros(n + 1, np.float64)
        fn, fs, ff = np.zeros(n + 1, np.uint32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.uint32)
        nb, nf = C.c_int32(), C.c_int32()
        lib().orb_refd_transform(self._h, _p(desc, C.c_uint8), n, int(levelsup), _p(bw, C.c_uint32), _p(bv, C.c_double), C.byref(nb),
                                 _p(fn, C.c_uint32), _p(fs, C.c_int32), _p(ff, C.c_uint32), C.byref(nf))
        return bw[:nb.value].copy(), bv[:nb.value].copy(), fn[:nf.value].copy(), fs[:nf.value + 1].copy(), ff[:fs[nf.value]].copy()


def bench_frames(frames: np.ndarray, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, lapping=(0, 0), mode=0,
                 bf=386.1448, b=0.5371657, nthreads=1):
    """Timed CPU baseline body on the reference's own code, std::threads inside the library: mode 0 extraction per frame,
    1 stereo pairs (frames 2i, 2i+1: compute x 2 + ComputeStereoMatches + kNN-2 / 0.7), 2 consecutive frames (compute +
    SearchForInitialization + kNN-2 / 0.8 per pair).  -> (total keypoints, matches)."""
    frames = np.ascontiguousarray(frames)
    n, h, w = frames.shape
    nm = C.c_long(0)
    tot = lib().orb_refs_bench_frames(nfeatures, scale_factor, nlevels, ini_th, min_th, _p(frames, C.c_uint8), n, w, h, int(lapping[0]),
                                      int(lapping[1]), int(mode), bf, b, nthreads, C.byref(nm))
    return tot, nm.value


def match_features(desc1, desc2, lowe_ratio_as_int=1):
    """geometry::matchFeatures method 2 + removeDuplicatedMatches (fmatcher.cpp:132-245, 247-266), the reference's own
    text; knnMatch is the oracle's kNN-2.  The reference reads the ratio with Config::get<int> and caches it in a
    function-local static, so the FIRST call of a process fixes it.  desc2 needs >= 2 rows (the reference indexes
    knn_matches[i][1] unconditionally).  -> int32[n][3] (queryIdx, trainIdx, distance), sorted by trainIdx; which of
    several matches onto one trainIdx survives follows std::sort (unstable beyond 16 elements)."""
    a, b = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32), np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    out = np.zeros((max(len(a), 1), 3), np.int32)
    n = lib().orb_refm_match_features(_p(a, C.c_uint8), len(a), _p(b, C.c_uint8), len(b), int(lowe_ratio_as_int), _p(out, C.c_int32))
    return out[:n].copy()


def kb8_triangulate(rig, pt1, pt2, sigma1, sigma2):
    """KannalaBrandt8::TriangulateMatches (kannalabrandt8.cpp:324-393 with :11-27, 33-37, 70-74, 82-109, 485-498), the
    reference's own text; cv::SVD::compute is the oracle's Jacobi stand-in.  -> (depth, p3d)."""
    f = lambda a: np.ascontiguousarray(a, np.float32)
    rig, pt1, pt2, s1, s2 = f(rig), f(pt1).reshape(-1, 2), f(pt2).reshape(-1, 2), f(sigma1), f(sigma2)
    n = len(pt1)
    depth, p3d = np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
    lib().orb_refk_triangulate(_p(rig, C.c_float), _p(pt1, C.c_float), _p(pt2, C.c_float), _p(s1, C.c_float), _p(s2, C.c_float), n,
                               _p(depth, C.c_float), _p(p3d, C.c_float))
    return depth, p3d


def stereo_fisheye_matches(rig, kps_left, desc_left, mono_left, kps_right, desc_right, mono_right, level_sigma2):
    """Frame::ComputeStereoFishEyeMatches (frame.cpp:1149-1189), the reference's own text (BFMatcher.knnMatch = the oracle's kNN-2)."""
    kl = np.ascontiguousarray(kps_left, KP_DTYPE); kr = np.ascontiguousarray(kps_right, KP_DTYPE)
    dl = np.ascontiguousarray(desc_left, np.uint8).reshape(-1, 32); dr = np.ascontiguousarray(desc_right, np.uint8).reshape(-1, 32)
    rig = np.ascontiguousarray(rig, np.float32)
    sig = np.ascontiguousarray(level_sigma2, np.float32)
    nl, nr = len(kl), len(kr)
    l2r, r2l = np.zeros(nl, np.int32), np.zeros(max(nr, 1), np.int32)
    depth, p3d = np.zeros(nl, np.float32), np.zeros((nl, 3), np.float32)
    lib().orb_refk_fisheye_matches(_p(rig, C.c_float), kl.ctypes.data, _p(dl, C.c_uint8), nl, int(mono_left), kr.ctypes.data,
                                   _p(dr, C.c_uint8), nr, int(mono_right), _p(sig, C.c_float), len(sig), _p(l2r, C.c_int32),
                                   _p(r2l, C.c_int32), _p(depth, C.c_float), _p(p3d, C.c_float))
    return {"left_to_right": l2r, "right_to_left": r2l[:nr], "depth": depth, "p3d": p3d}


def undistort_keypoints(kps, cam, dist):
    """Frame::UndistortKeyPoints (frame.cpp:758-791), the reference's own text; cv::undistortPoints = the oracle's restatement."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    cam, dist = np.ascontiguousarray(cam, np.float32), np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(kps)
    lib().orb_refs_undistort_keypoints(kps.ctypes.data, len(kps), _p(cam, C.c_float), _p(dist, C.c_float), len(dist), out.ctypes.data)
    return out


def image_bounds(cols, rows, cam, dist):
    """Frame::ComputeImageBounds (frame.cpp:793-821), the reference's own text -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    cam, dist = np.ascontiguousarray(cam, np.float32), np.ascontiguousarray(dist, np.float32)
    out = np.zeros(4, np.float32)
    lib().orb_refs_image_bounds(int(cols), int(rows), _p(cam, C.c_float), _p(dist, C.c_float), len(dist), _p(out, C.c_float))
    return out
