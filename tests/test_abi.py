"""CPU-only checks of the drop-in boundary: the CUDA library loads, exports every symbol that
include/viorb.h declares, and refuses to compute without a GPU (no silent fallback)."""
import ctypes as C
import os
import re

import pytest

import __graft_entry__ as ge
from vi_slam_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cuda_lib():
    ge.build_cuda()
    return C.CDLL(ge.LIB)


def _declared():
    text = open(os.path.join(ROOT, "include", "viorb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(viorb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(cuda_lib):
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(cuda_lib, n), n


def test_python_binding_covers_header():
    assert sorted(_lib.SYMBOLS) == _declared()


def test_keypoint_layout_is_cv_keypoint():
    from vi_slam_b200 import KP_DTYPE, MATCH_DTYPE
    assert KP_DTYPE.itemsize == 28 and KP_DTYPE.names == ("x", "y", "size", "angle", "response", "octave", "class_id")
    assert MATCH_DTYPE.itemsize == 12


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(ImportError):
        _lib.load(str(tmp_path / "nope.so"))


def test_no_gpu_means_error_not_fallback(cuda_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from vi_slam_b200 import FExtractor, ViorbError, matcher
    import numpy as np
    with pytest.raises(ViorbError):
        FExtractor(500, max_size=(320, 240))
    with pytest.raises(ViorbError):
        matcher.knn_match(np.zeros((4, 32), np.uint8), np.zeros((4, 32), np.uint8))


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "vi_slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "orb_oracle" not in src, f
                assert "cuda_emu::" not in src or f == "viorb_common.cuh", f
