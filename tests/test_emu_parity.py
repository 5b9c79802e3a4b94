"""Logic parity on CPU: the CUDA sources compiled by g++ on the execution shim (tests/cuda_emu) must
reproduce the oracle.  This checks index math, barriers and host orchestration without a GPU; the
real parity tests (-m gpu, tests/test_gpu_parity.py) run the nvcc build on the B200."""
import numpy as np
import pytest

import __graft_entry__ as ge
from tests import parity
from vi_slam_b200 import synthetic


@pytest.fixture(scope="module")
def emu(oracle_mod):
    return ge.build_emu()


@pytest.mark.parametrize("w,h,nf,lap,seed", [(320, 240, 500, (0, 0), 1), (400, 200, 700, (100, 250), 2),
                                             (640, 360, 3000, (0, 1000), 3), (131, 97, 300, (0, 0), 4), (258, 130, 400, (0, 0), 5),
                                             (123, 91, 200, (0, 0), 6), (700, 96, 600, (0, 0), 7)])   # one 59-px cell column; strips of 8 cells, 2 cell rows
def test_extract_stages(emu, w, h, nf, lap, seed):
    parity.check_extract(synthetic.make_frame(seed, w, h), nf, lap, lib_path=emu)


@pytest.mark.parametrize("seed", range(8))
def test_extract_random_geometries(emu, seed):
    """Random image sizes, feature counts and frame classes: strip geometry of k_fast (cell width 30..59, clipped last cells),
    quadtree root counts 1..4+ and the per-axis leaf tables, patches next to the 19-pixel border."""
    rng = np.random.default_rng(100 + seed)
    w, h = int(rng.integers(96, 460)), int(rng.integers(80, 300))
    if seed % 3 == 2:
        w, h = h, w + 40                                            # portrait: one quadtree root
    nf = int(rng.integers(150, 1800))
    img = synthetic.make_frame(200 + seed, w, h)
    if seed % 4 == 1:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)          # noise: the dense quadtree launch, long FAST queues
    if seed % 4 == 3:
        img = (img.astype(np.float32) * 0.1 + 90).astype(np.uint8)  # low texture: minThFAST cells
    parity.check_extract(img, nf, (0, 0) if seed % 2 else (w // 4, w // 2), lib_path=emu)


def test_extract_edge_cases(emu):
    from vi_slam_b200 import FExtractor
    ex = FExtractor(500, 1.2, 8, 20, 7, max_size=(320, 240), lib_path=emu)
    mono, kps, desc = ex.compute(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(kps) == 0
    parity.check_extract(np.zeros((120, 160), np.uint8), 500, (0, 0), lib_path=emu, ex=ex)          # no keypoints
    rng = np.random.default_rng(1)
    parity.check_extract(rng.integers(0, 256, (120, 160), dtype=np.uint8), 500, (0, 0), lib_path=emu, ex=ex)
    parity.check_extract(synthetic.make_frame(4, 320, 240), 500, (0, 0), lib_path=emu, ex=ex)       # size change
    big = np.zeros((120, 200), np.uint8)
    big[:, :160] = rng.integers(0, 256, (120, 160), dtype=np.uint8)
    parity.check_extract(big[:, :160], 500, (0, 0), lib_path=emu, ex=ex)                            # strided input
    with pytest.raises(Exception):
        ex.compute(np.zeros((300, 400), np.uint8))                                                  # above max_size


def test_knn_and_ratio(emu):
    q, t = synthetic.make_descriptors(300, 257, seed=3)
    parity.check_knn(q, t, emu)
    assert parity.check_ratio(q, t, emu) > 10
    parity.check_knn(q[:5], t[:1], emu)                 # second neighbour missing
    parity.check_knn(q[:130], t[:129], emu)             # tile boundary


def test_stereo(emu):
    left, right = synthetic.make_stereo_pair(5, 480, 270)
    n_stereo, n_ratio, _ = parity.check_stereo(left, right, 1000, emu)
    assert n_stereo > 50 and n_ratio > 50


def test_batch_matches_single(emu):
    from vi_slam_b200 import FExtractor
    frames = [synthetic.make_frame(10 + i, 320, 240) for i in range(3)]
    ex = FExtractor(500, 1.2, 8, 20, 7, max_size=(320, 240), max_batch=3, lib_path=emu)
    mono, kps, desc = ex.compute_batch(frames, (0, 0))
    for i, f in enumerate(frames):
        _, _, (m1, k1, d1), _ = parity.check_extract(f, 500, (0, 0), lib_path=emu, stages=False)
        assert m1 == mono[i] and np.array_equal(k1, kps[i]) and np.array_equal(d1, desc[i])


def test_chunked_batch(emu):
    """More than 32 frames: viorb_extract_batch cuts the batch into chunks (frame0 offsets in every kernel)."""
    from vi_slam_b200 import FExtractor
    frames = [synthetic.make_frame(40 + i, 96, 80) for i in range(81)]
    ex = FExtractor(100, 1.2, 4, 20, 7, max_size=(96, 80), max_batch=81, lib_path=emu)
    one = FExtractor(100, 1.2, 4, 20, 7, max_size=(96, 80), max_batch=1, lib_path=emu)
    mono, kps, desc = ex.compute_batch(frames, (0, 0))
    for i, f in enumerate(frames):
        m1, k1, d1 = one.compute(f, None, (0, 0))
        assert m1 == mono[i] and np.array_equal(k1, kps[i]) and np.array_equal(d1, desc[i])
    assert sum(len(k) for k in kps) > 300
