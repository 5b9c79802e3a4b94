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


def _dropin(lib, tmp_path, w, h, nf, lap):
    """The same caller source against the reference's own FExtractor (liborb_ref.so) and against the shim over `lib`."""
    from oracle import ref
    if not (ref.available() and os.path.exists(ref.REF_SOURCE)):
        pytest.skip("needs /root/reference (its headers) and oracle/_ref/liborb_ref.so")
    src = os.path.join(ROOT, "tests", "cpp", "dropin_driver.cpp")
    compat = os.path.join(ROOT, "oracle", "ref_compat")
    exe_ref, exe_new = str(tmp_path / "dropin_ref"), str(tmp_path / "dropin_viorb")
    rdir = os.path.dirname(ref.LIB_PATH)
    subprocess.run(["g++", "-std=c++17", "-O1", "-w", "-I" + os.path.join(ref.REFERENCE, "include"), "-I" + compat, src, "-L" + rdir,
                    "-lorb_ref", "-Wl,-rpath," + rdir, "-pthread", "-o", exe_ref], check=True)
    libdir, libname = os.path.dirname(lib), os.path.basename(lib)[3:-3]
    subprocess.run(["g++", "-std=c++17", "-O1", "-w", "-DUSE_VIORB", "-DVIORB_WITH_OPENCV", "-I" + os.path.join(ROOT, "include"),
                    "-I" + compat, src, os.path.join(compat, "cv_compat.cpp"), os.path.join(ROOT, "oracle", "orb_oracle.cpp"),
                    "-L" + libdir, "-l" + libname, "-Wl,-rpath," + libdir, "-pthread", "-o", exe_new], check=True)
    img = synthetic.make_frame(31, w, h)
    raw = str(tmp_path / "in.raw")
    img.tofile(raw)
    outs = []
    for exe in (exe_ref, exe_new):
        out = exe + ".bin"
        subprocess.run([exe, raw, str(w), str(h), str(nf), str(lap[0]), str(lap[1]), out], check=True)
        outs.append(open(out, "rb").read())
    a, b = outs
    assert len(b) == len(a) + 4                      # the shim build appends its matchFeatures count
    assert a == b[:len(a)], "drop-in FExtractor differs from the reference's own class"
    mono, n = np.frombuffer(a, np.int32, 2)
    assert n > 100 and 0 <= mono <= n


def test_dropin_same_caller_source_reference_vs_shim_emulation(oracle_mod, tmp_path):
    _dropin(ge.build_emu(), tmp_path, 320, 240, 500, (100, 200))


@pytest.mark.gpu
def test_matcher_shims_on_cuda_build(oracle_mod, tmp_path):
    """include/viorb_fmatcher.hpp + the VIORB_WITH_OPENCV branch of viorb_fextractor.hpp on the CUDA build.  On the GPU box
    the reference headers are absent, so only the shim build of the drop-in driver runs; its extractor output is checked
    against the oracle."""
    lib = ge.build_cuda()
    src = os.path.join(ROOT, "tests", "cpp", "dropin_driver.cpp")
    compat = os.path.join(ROOT, "oracle", "ref_compat")
    exe = str(tmp_path / "dropin_viorb")
    libdir, libname = os.path.dirname(lib), os.path.basename(lib)[3:-3]
    subprocess.run(["g++", "-std=c++17", "-O1", "-w", "-DUSE_VIORB", "-DVIORB_WITH_OPENCV", "-I" + os.path.join(ROOT, "include"),
                    "-I" + compat, src, os.path.join(compat, "cv_compat.cpp"), os.path.join(ROOT, "oracle", "orb_oracle.cpp"),
                    "-L" + libdir, "-l" + libname, "-Wl,-rpath," + libdir, "-pthread", "-o", exe], check=True)
    w, h, nf, lap = 1241, 376, 2000, (0, 1000)
    img = synthetic.make_frame(31, w, h)
    raw, out = str(tmp_path / "in.raw"), str(tmp_path / "out.bin")
    img.tofile(raw)
    subprocess.run([exe, raw, str(w), str(h), str(nf), str(lap[0]), str(lap[1]), out], check=True)
    blob = open(out, "rb").read()
    mono, n = np.frombuffer(blob, np.int32, 2)
    off = 16 + 4 * 4 * 8
    kps = np.frombuffer(blob, oracle.KP_DTYPE, n, off)
    desc = np.frombuffer(blob, np.uint8, n * 32, off + 28 * n).reshape(n, 32)
    om, ok, od = oracle.OracleExtractor(nf).compute(img, lap)
    assert mono == om and np.array_equal(kps, ok) and np.array_equal(desc, od)
