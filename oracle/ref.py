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
               [os.path.join(_HERE, "ref_compat", f) for f in ("cv_compat.hpp", "cv_compat.cpp", "ref_driver.cpp")]
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
        L.orb_ref_consistent.argtypes = [C.c_void_p]
        L.orb_ref_level_size.argtypes = [C.c_void_p, C.c_int, i32p, i32p]
        L.orb_ref_level_image.argtypes = [C.c_void_p, C.c_int, u8p]
        L.orb_ref_level_candidates.argtypes = [C.c_void_p, C.c_int, i32p, C.c_int]
        L.orb_ref_level_keypoints.argtypes = [C.c_void_p, C.c_int, i32p, f32p, C.c_int]
        L.orb_ref_arena_fallbacks.restype = C.c_long
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
