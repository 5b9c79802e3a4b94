"""Loader for libviorb.so (the C ABI of include/viorb.h).

The product has no CPU path: if the CUDA library is missing this raises.  No
environment variable redirects the loader; tests that check host logic without
a GPU pass ``lib_path=`` (the g++ emulation build under tests/cuda_emu)
explicitly to the classes they construct.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.path.join(_HERE, "libviorb.so")

class ViorbVocabDesc(C.Structure):
    _fields_ = [("nnodes", C.c_int32), ("levels", C.c_int32), ("weighting", C.c_int32), ("scoring", C.c_int32),
                ("child_start", C.c_void_p), ("children", C.c_void_p), ("descriptors", C.c_void_p), ("weight", C.c_void_p),
                ("word_id", C.c_void_p)]


ALLGATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p)
u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)


class ViorbParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("max_width", C.c_int32),
                ("max_height", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32)]


SYMBOLS = {
    # name: (restype, argtypes)
    "viorb_last_error": (C.c_char_p, []),
    "viorb_version": (C.c_char_p, []),
    "viorb_extractor_create": (C.c_int, [C.POINTER(ViorbParams), C.POINTER(C.c_void_p)]),
    "viorb_extractor_destroy": (None, [C.c_void_p]),
    "viorb_extractor_tables": (C.c_int, [C.c_void_p, i32p, f32p, f32p, f32p, f32p, i32p]),
    "viorb_extractor_capacity": (C.c_int, [C.c_void_p]),
    "viorb_host_alloc": (C.c_void_p, [C.c_size_t]),
    "viorb_host_free": (None, [C.c_void_p]),
    "viorb_extractor_stream": (C.c_void_p, [C.c_void_p]),
    "viorb_extractor_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "viorb_extractor_sync": (C.c_int, [C.c_void_p]),
    "viorb_extractor_set_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "viorb_extractor_stage_ms": (C.c_int, [C.c_void_p, f32p]),
    "viorb_stereo_stage_ms": (C.c_int, [C.c_void_p, f32p]),
    "viorb_stereo_fetch_all": (C.c_int, [C.c_void_p, C.c_int, f32p, f32p, i32p, C.c_int, i32p, i32p, i32p]),
    "viorb_extract": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_void_p, C.c_void_p, C.c_int, i32p, i32p]),
    "viorb_extract_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, i32p, i32p]),
    "viorb_extract_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                       C.c_int, C.c_int]),
    "viorb_fetch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, i32p, i32p]),
    "viorb_pyramid_level": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, i32p, i32p]),
    "viorb_debug_candidates": (C.c_int, [C.c_void_p, C.c_int, C.c_int, i32p, C.c_int, i32p]),
    "viorb_debug_kept": (C.c_int, [C.c_void_p, C.c_int, C.c_int, i32p, f32p, C.c_int, i32p]),
    "viorb_debug_blurred": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]),
    "viorb_debug_knn_split": (None, [C.c_int]),
    "viorb_descriptor_distance": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, i32p, C.c_int]),
    "viorb_knn2": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, i32p, i32p, C.c_int]),
    "viorb_knn2_device": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "viorb_knn2_device_static": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "viorb_knn2_packed_device": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "viorb_comm_unique_id": (C.c_int, [C.c_void_p]),
    "viorb_comm_create_nccl": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "viorb_comm_adopt_nccl": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "viorb_comm_create_callback": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "viorb_comm_destroy": (None, [C.c_void_p]),
    "viorb_comm_rank": (C.c_int, [C.c_void_p]),
    "viorb_comm_world": (C.c_int, [C.c_void_p]),
    "viorb_shard_rows": (None, [C.c_int, C.c_int, C.c_int, i32p, i32p, i32p]),
    "viorb_knn2_sharded": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "viorb_match_ratio": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_int,
                                    C.c_void_p, C.c_int, i32p, C.c_int]),
    "viorb_stereo_match": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, f32p, f32p, i32p, i32p,
                                     C.c_int, i32p]),
    "viorb_stereo_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_double]),
    "viorb_stereo_fetch": (C.c_int, [C.c_void_p, C.c_int, f32p, f32p, i32p, C.c_int, i32p, i32p, i32p]),
    "viorb_match_resident": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p,
                                       C.c_int, i32p]),
    "viorb_features_in_area": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int,
                                         i32p, C.c_int, i32p, C.c_int]),
    "viorb_search_for_initialization": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, i32p, C.c_void_p,
                                                  C.c_void_p, i32p, C.c_void_p, f32p, C.c_int, C.c_float, C.c_int, i32p,
                                                  i32p, C.c_int]),
    "viorb_search_by_projection": (C.c_int, [C.c_int, C.c_int, C.c_int, f32p, f32p, i32p, f32p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, i32p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, i32p,
                                             C.c_void_p, f32p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                             i32p, i32p, C.c_int]),
    "viorb_f2f_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_double, C.c_int]),
    "viorb_f2f_fetch_all": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "viorb_extractor_set_camera": (C.c_int, [C.c_void_p, f32p, f32p, C.c_int]),
    "viorb_fetch_undistorted": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, i32p]),
    "viorb_undistort_keypoints": (C.c_int, [C.c_void_p, C.c_int, f32p, f32p, C.c_int, C.c_void_p, C.c_int]),
    "viorb_undistort_points": (C.c_int, [f32p, C.c_int, f32p, f32p, C.c_int, f32p, C.c_int]),
    "viorb_image_bounds": (C.c_int, [C.c_int, C.c_int, f32p, f32p, C.c_int, C.c_void_p, C.c_int]),
    "viorb_vocabulary_create": (C.c_int, [C.POINTER(ViorbVocabDesc), C.c_int, C.POINTER(C.c_void_p)]),
    "viorb_vocabulary_destroy": (None, [C.c_void_p]),
    "viorb_bow_transform": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, i32p, C.c_int, C.c_void_p, C.c_void_p, i32p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, i32p]),
    "viorb_bow_resident": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "viorb_bow_fetch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, i32p, C.c_void_p, C.c_void_p, C.c_void_p, i32p]),
    "viorb_kb8_triangulate_matches": (C.c_int, [f32p, f32p, f32p, f32p, f32p, C.c_int, f32p, f32p, C.c_int]),
    "viorb_fisheye_stereo_matches": (C.c_int, [f32p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                               C.c_int, f32p, C.c_int, i32p, i32p, f32p, f32p, i32p, i32p, C.c_int]),
}

_cache = {}


def load(path: str | None = None):
    path = path or DEFAULT_LIB
    if path in _cache:
        return _cache[path]
    if not os.path.exists(path):
        raise ImportError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). vi_slam_b200 has no CPU fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the ABI is incomplete
        fn.restype = res
        fn.argtypes = args
    _cache[path] = lib
    return lib
