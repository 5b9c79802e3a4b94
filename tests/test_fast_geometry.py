"""The multiply-shift reciprocals k_fast takes from the host (CellDesc::invNC, ginv, LevelGeom::cinv; vi_slam_b200/csrc/
viorb_api.cu build_geometry) are exact over the ranges the kernel uses them on (vi_slam_b200/csrc/kernels_extract.cuh)."""
import numpy as np

TP = 336                 # VIORB_FAST_TP
MAXRW = TP - 48          # VIORB_FAST_MAXRW
MAXRH = 127              # build_geometry rejects taller strips


def test_chunk_index_to_tile_row():
    # i / nC = (i * invNC) >> 20 for every chunk of a tile: nC <= TP / 16 chunks per row, <= MAXRH + 6 rows
    for nC in range(1, TP // 16 + 1):
        inv = (1 << 20) // nC + 1
        i = np.arange((MAXRH + 6) * nC, dtype=np.uint64)
        assert i.max() * inv < 2 ** 32
        assert np.array_equal((i * inv) >> 20, i // nC)


def test_word_index_to_region_row():
    # i / ng = (i * ginv) >> 22 for every aligned word of a region: ng <= MAXRW / 4 + 1 words per row, <= MAXRH rows
    for ng in range(1, MAXRW // 4 + 2):
        inv = (1 << 22) // ng + 1
        i = np.arange(MAXRH * ng, dtype=np.uint64)
        assert i.max() * inv < 2 ** 32
        assert np.array_equal((i * inv) >> 22, i // ng)


def test_strip_pixel_to_cell():
    # px / wCell = (px * cinv) >> 20 inside a strip; a cell is 30..59 pixels wide (width / floor(width / 30), fextractor.cpp:775-778)
    for wcell in range(1, 60):
        inv = (1 << 20) // wcell + 1
        px = np.arange(MAXRW, dtype=np.uint64)
        assert np.array_equal((px * inv) >> 20, px // wcell)


def test_queue_entries_fit_16_bits_and_strips_fit_the_tile():
    assert (MAXRH - 1) * TP + MAXRW - 1 < 2 ** 16                      # linear queue entry py * TP + px
    # tile row: alignment shift (<= 19) + 3-pixel halo + region + 2 + the word right of the halo, in whole 16-byte chunks
    for shift in range(4, 20):
        nB = (((shift + 3 + MAXRW + 2) >> 2) + 2) << 2
        assert ((nB + 15) >> 4) * 16 <= TP
    for wcell in range(30, 60):                                        # the host keeps >= 1 cell per strip
        assert MAXRW // wcell >= 4
