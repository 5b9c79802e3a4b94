"""The C++ shim with FExtractor's interface (include/viorb_fextractor.hpp) over the C ABI: compiled with
g++ and run against the emulation build here, against the CUDA build on the GPU box."""
import os
import subprocess

import numpy as np
import pytest

import __graft_entry__ as ge
import oracle
from vi_slam_b200 import synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(lib, tmp_path, w, h, nf, lap):
    exe = str(tmp_path / "shim_driver")
    libdir, libname = os.path.dirname(lib), os.path.basename(lib)[3:-3]
    subprocess.run(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "cpp", "shim_driver.cpp"), "-L" + libdir, "-l" + libname,
                    "-Wl,-rpath," + libdir, "-pthread", "-o", exe], check=True)
    img = synthetic.make_frame(21, w, h)
    raw, out = str(tmp_path / "in.raw"), str(tmp_path / "out.bin")
    img.tofile(raw)
    subprocess.run([exe, raw, str(w), str(h), str(nf), str(lap[0]), str(lap[1]), out], check=True)
    blob = open(out, "rb").read()
    mono, n = np.frombuffer(blob, np.int32, 2)
    kps = np.frombuffer(blob, oracle.KP_DTYPE, n, 8)
    desc = np.frombuffer(blob, np.uint8, n * 32, 8 + 28 * n).reshape(n, 32)
    ox = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    om, ok, od = ox.compute(img, lap)
    assert mono == om and n == len(ok)
    assert np.array_equal(kps, ok) and np.array_equal(desc, od)
    off = 8 + 60 * n
    for l in range(8):
        lw, lh = np.frombuffer(blob, np.int32, 2, off)
        lv = np.frombuffer(blob, np.uint8, lw * lh, off + 8).reshape(lh, lw)
        assert np.array_equal(lv, ox.level_image(l)), f"mvImagePyramid[{l}]"
        off += 8 + lw * lh


def test_cpp_shim_on_emulation_build(oracle_mod, tmp_path):
    _run(ge.build_emu(), tmp_path, 320, 240, 500, (100, 200))


@pytest.mark.gpu
def test_cpp_shim_on_cuda_build(oracle_mod, tmp_path):
    _run(ge.build_cuda(), tmp_path, 1241, 376, 2000, (0, 1000))
