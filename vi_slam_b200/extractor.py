"""Host-side mirror of the reference's extractor interface over the C ABI.

``FExtractor`` keeps the names and argument meaning of
reference include/vi_slam/geometry/fextractor.h:26-64 (the reference is C++;
its toolchain dependencies are absent here, so this mirror is what the parity
tests drive; the C++ shim with the identical signature is include/viorb_fextractor.hpp).
All compute happens in libviorb.so (CUDA, sm_100a); there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
MATCH_DTYPE = np.dtype([("query_idx", "<i4"), ("train_idx", "<i4"), ("distance", "<i4")])


class ViorbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"viorb error {code}: {msg}")
        self.code = code


def pinned_empty(shape, dtype, lib_path=None):
    """numpy array in page-locked host memory (viorb_host_alloc): images / result buffers placed here are
    DMA'd by the library without a staging copy.  The memory is released when the array is collected."""
    lib = _lib.load(lib_path)
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    ptr = lib.viorb_host_alloc(max(n, 1))
    if not ptr:
        raise ViorbError(-3, lib.viorb_last_error().decode())
    buf = (C.c_uint8 * max(n, 1)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[arr.__array_interface__["data"][0]] = (lib, ptr)
    import weakref
    weakref.finalize(buf, lib.viorb_host_free, ptr)
    return arr


_PINNED = {}


def _check(lib, rc):
    if rc != 0:
        raise ViorbError(rc, lib.viorb_last_error().decode())


class HostBatch:
    """Pointer table of n equally sized CV_8UC1 host images (fextractor.cpp:1041) as viorb_extract_batch takes it."""

    def __init__(self, images):
        n = len(images)
        h, w = images[0].shape
        step = images[0].strides[0]
        self.keep = []
        self.ptrs = (C.c_void_p * n)()
        for i, im in enumerate(images):
            assert im.dtype == np.uint8 and im.ndim == 2 and im.shape == (h, w)
            if im.strides[1] != 1 or im.strides[0] != step:
                im = np.ascontiguousarray(im)
                assert im.strides[0] == step or n == 1
                step = im.strides[0]
            self.keep.append(im)
            self.ptrs[i] = im.ctypes.data
        self.n, self.w, self.h, self.step = n, w, h, step


class FExtractor:
    """FExtractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) — fextractor.cpp:401-461.

    Extra keyword arguments size the device arenas: ``max_size=(w, h)``, ``max_batch``, ``device``.
    """

    def __init__(self, nfeatures=2000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, *,
                 max_size=(1280, 720), max_batch=1, device=0, lib_path=None):
        self._lib = _lib.load(lib_path)
        self._h = C.c_void_p()
        p = _lib.ViorbParams(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_size[0], max_size[1],
                             max_batch, device)
        _check(self._lib, self._lib.viorb_extractor_create(C.byref(p), C.byref(self._h)))
        self.nfeatures, self.nlevels, self.scaleFactor = nfeatures, nlevels, scaleFactor
        self.max_batch = max_batch
        self.capacity = self._lib.viorb_extractor_capacity(self._h)
        nl = C.c_int32()
        t = [np.zeros(nlevels, np.float32) for _ in range(4)]
        self.mnFeaturesPerLevel = np.zeros(nlevels, np.int32)
        _check(self._lib, self._lib.viorb_extractor_tables(
            self._h, C.byref(nl), *[a.ctypes.data_as(_lib.f32p) for a in t],
            self.mnFeaturesPerLevel.ctypes.data_as(_lib.i32p)))
        self.mvScaleFactor, self.mvInvScaleFactor, self.mvLevelSigma2, self.mvInvLevelSigma2 = t
        self._last_n = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.viorb_extractor_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- getters (fextractor.h:40-62)
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self.scaleFactor

    def GetScaleFactors(self):
        return self.mvScaleFactor.copy()

    def GetInverseScaleFactors(self):
        return self.mvInvScaleFactor.copy()

    def GetScaleSigmaSquares(self):
        return self.mvLevelSigma2.copy()

    def GetInverseScaleSigmaSquares(self):
        return self.mvInvLevelSigma2.copy()

    # --- compute (fextractor.cpp:1034-1133)
    def compute(self, image, mask=None, vLappingArea=(0, 0)):
        """-> (monoIndex, keypoints[KP_DTYPE], descriptors uint8[n,32]); monoIndex == -1 for an empty image.
        ``mask`` is ignored like in the reference."""
        if image is None or getattr(image, "size", 0) == 0:
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        mono, kps, desc = self.compute_batch([image], vLappingArea)
        return mono[0], kps[0], desc[0]

    def alloc_outputs(self, n, pinned=True):
        """Reusable result buffers for ``compute_batch(..., out=...)``."""
        cap = self.capacity
        mk = (lambda sh, dt: pinned_empty(sh, dt)) if pinned else (lambda sh, dt: np.zeros(sh, dt))
        return {"kps": mk((n, cap), KP_DTYPE), "desc": mk((n, cap, 32), np.uint8), "count": np.zeros(n, np.int32),
                "mono": np.zeros(n, np.int32)}

    def host_batch(self, images):
        """Pointer table of a list of equally sized uint8 images for ``compute_batch`` - build it once when the same host
        buffers (a capture ring, page-locked staging memory) are handed over repeatedly."""
        return HostBatch(images)

    def compute_batch(self, images, vLappingArea=(0, 0), out=None, views=True):
        """Batch form of ``compute``; ``images`` is a list of arrays or a ``HostBatch``.  With ``out`` (from
        ``alloc_outputs``) results are written in place and the return value holds views (rows beyond count[i] are
        undefined) instead of fresh copies; ``views=False`` returns only (mono, count) and leaves slicing to the caller."""
        hb = images if isinstance(images, HostBatch) else HostBatch(images)
        n, w, h, step = hb.n, hb.w, hb.h, hb.step
        cap = self.capacity
        if out is None:
            kps = np.zeros((n, cap), KP_DTYPE)
            desc = np.zeros((n, cap, 32), np.uint8)
            cnt = np.zeros(n, np.int32)
            mono = np.zeros(n, np.int32)
            addr = (kps.ctypes.data, desc.ctypes.data, cnt.ctypes.data_as(_lib.i32p), mono.ctypes.data_as(_lib.i32p))
        else:
            kps, desc, cnt, mono = out["kps"], out["desc"], out["count"], out["mono"]
            assert kps.shape[0] >= n and kps.shape[1] == cap
            addr = out.get("_addr")
            if addr is None:
                addr = out["_addr"] = (kps.ctypes.data, desc.ctypes.data, cnt.ctypes.data_as(_lib.i32p), mono.ctypes.data_as(_lib.i32p))
        _check(self._lib, self._lib.viorb_extract_batch(
            self._h, hb.ptrs, n, w, h, step, int(vLappingArea[0]), int(vLappingArea[1]), addr[0], addr[1], cap, addr[2], addr[3]))
        self._last_n = n
        if out is not None:
            if not views:
                return mono[:n], cnt[:n]
            return mono[:n], [kps[i, :cnt[i]] for i in range(n)], [desc[i, :cnt[i]] for i in range(n)]
        return ([int(m) for m in mono], [kps[i, :cnt[i]].copy() for i in range(n)],
                [desc[i, :cnt[i]].copy() for i in range(n)])

    # --- device-resident path
    def extract_device(self, ptr, n, width, height, step, frame_stride, vLappingArea=(0, 0)):
        _check(self._lib, self._lib.viorb_extract_device(self._h, ptr, n, width, height, step, frame_stride,
                                                         int(vLappingArea[0]), int(vLappingArea[1])))
        self._last_n = n

    def fetch(self, frame):
        cap = self.capacity
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        cnt, mono = C.c_int32(), C.c_int32()
        _check(self._lib, self._lib.viorb_fetch(self._h, frame, kps.ctypes.data, desc.ctypes.data, cap,
                                                C.byref(cnt), C.byref(mono)))
        return mono.value, kps[:cnt.value].copy(), desc[:cnt.value].copy()

    def set_camera(self, K, dist):
        """mK (fx, fy, cx, cy or a 3x3 matrix) and mDistCoef (k1 k2 p1 p2 [k3]) of the Frame this extractor feeds: from the
        next compute on, the epilogue also produces Frame::UndistortKeyPoints' ukeypoints_ (frame.cpp:758-791).
        ``dist=None`` switches it off."""
        if dist is None:
            _check(self._lib, self._lib.viorb_extractor_set_camera(self._h, None, None, 0))
            return
        K = np.asarray(K, np.float32)
        cam = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], np.float32) if K.ndim == 2 else np.ascontiguousarray(K, np.float32)
        d = np.ascontiguousarray(dist, np.float32).ravel()
        _check(self._lib, self._lib.viorb_extractor_set_camera(self._h, cam.ctypes.data_as(_lib.f32p), d.ctypes.data_as(_lib.f32p),
                                                               len(d)))

    def fetch_undistorted(self, frame=0):
        """ukeypoints_ of a resident frame (same order as ``fetch``'s keypoints)."""
        cap = self.capacity
        kps = np.zeros(cap, KP_DTYPE)
        cnt = C.c_int32()
        _check(self._lib, self._lib.viorb_fetch_undistorted(self._h, frame, kps.ctypes.data, cap, C.byref(cnt)))
        return kps[:cnt.value].copy()

    def sync(self):
        _check(self._lib, self._lib.viorb_extractor_sync(self._h))

    STAGES = ("resize", "fast", "octree", "orient", "emit", "blur", "describe")

    def set_timing(self, enable=True):
        _check(self._lib, self._lib.viorb_extractor_set_timing(self._h, int(enable)))

    def stage_ms(self):
        ms = np.zeros(7, np.float32)
        _check(self._lib, self._lib.viorb_extractor_stage_ms(self._h, ms.ctypes.data_as(_lib.f32p)))
        return dict(zip(self.STAGES, ms.tolist()))

    def set_stream(self, stream_ptr):
        _check(self._lib, self._lib.viorb_extractor_set_stream(self._h, stream_ptr))

    # --- mvImagePyramid (fextractor.h:64)
    def pyramid_level(self, level, frame=0):
        w, h = C.c_int32(), C.c_int32()
        _check(self._lib, self._lib.viorb_pyramid_level(self._h, frame, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        _check(self._lib, self._lib.viorb_pyramid_level(self._h, frame, level, out.ctypes.data, w.value,
                                                        C.byref(w), C.byref(h)))
        return out

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    # --- stage inspection (parity tests)
    def debug_candidates(self, level, frame=0):
        n = C.c_int32()
        _check(self._lib, self._lib.viorb_debug_candidates(self._h, frame, level, None, 0, C.byref(n)))
        buf = np.zeros((max(n.value, 1), 3), np.int32)
        _check(self._lib, self._lib.viorb_debug_candidates(self._h, frame, level, buf.ctypes.data_as(_lib.i32p),
                                                           n.value, C.byref(n)))
        return buf[:n.value]

    def debug_kept(self, level, frame=0):
        cap = self.capacity
        buf = np.zeros((cap, 3), np.int32)
        ang = np.zeros(cap, np.float32)
        n = C.c_int32()
        _check(self._lib, self._lib.viorb_debug_kept(self._h, frame, level, buf.ctypes.data_as(_lib.i32p),
                                                     ang.ctypes.data_as(_lib.f32p), cap, C.byref(n)))
        return buf[:n.value], ang[:n.value]

    def debug_blurred(self, level, frame=0):
        w, h = C.c_int32(), C.c_int32()
        _check(self._lib, self._lib.viorb_pyramid_level(self._h, frame, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        _check(self._lib, self._lib.viorb_debug_blurred(self._h, frame, level, out.ctypes.data, w.value))
        return out
