"""CPU check of the arithmetic k_knn2_f4 relies on (vi_slam_b200/csrc/kernels_knn_f4.cuh): with the e2m1 codes, block scales
and key-step digits the kernel writes, the fp32 accumulator of every (Hamming distance, column) is a number in [2^23, 2^24)
whose low 16 pattern bits are 8192 + 64 * Hamming + column, padding rows sort above every real key, and float32 addition in
any order is exact for these sums.  The constants are read from the kernel source so the test follows the code.
(That the tensor core itself forms the sum exactly was measured on the GPU: tools/probe/mxf4_probe.cu; the kernel's results are
compared with the oracle in tests/test_gpu_parity.py.)"""
import os
import re

import numpy as np

SRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vi_slam_b200", "csrc", "kernels_knn_f4.cuh")
E2M1 = np.array([0, .5, 1, 1.5, 2, 3, 4, 6, -0., -.5, -1, -1.5, -2, -3, -4, -6], np.float64)


def _const(name, text):
    m = re.search(r"constexpr\s+\w+\s+" + name + r"\s*=\s*([^;]+);", text)
    assert m, name
    return m.group(1)


def _col_code(col6):
    """col_code() of the kernel."""
    d0, d1, d2 = col6 & 3, (col6 >> 2) & 3, col6 >> 4
    small, xs, ys = 0x5420, 0x7660, 0x7600
    return ((small >> (4 * d0)) & 15) | (((small >> (4 * d1)) & 15) << 4) | (((xs >> (4 * d2)) & 15) << 8) | (((ys >> (4 * d2)) & 15) << 12)


def _nibbles(word, n):
    return [(word >> (4 * i)) & 15 for i in range(n)]


def test_key_arithmetic_matches_the_kernel_constants():
    text = open(SRC).read()
    sfa = eval(_const("SFA_WORD", text).replace("u", ""))
    assert eval(_const("SFB_WORD", text).replace("u", "")) == 0x7f7f7f7f
    key0, pad_key = int(_const("KEY0", text)), int(_const("PAD_KEY", text))
    assert "0x6662u" in text and "0x7777u | (1u << 16) | (1u << 20)" in text and "(valid ? 0u : (2u << 20))" in text
    s_main = [2.0 ** (((sfa >> (8 * i)) & 0xff) - 127) for i in (0, 1)]
    s_x, s_y = (2.0 ** (((sfa >> (8 * i)) & 0xff) - 127) for i in (2, 3))
    assert s_main == [32.0, 32.0] and s_x == 1.0 and s_y == 65536.0
    qa_x = E2M1[_nibbles(0x6662, 4)]                                   # query side of block X
    qa_y = E2M1[_nibbles(0x7777 | (1 << 16) | (1 << 20), 6)]           # query side of block Y
    for pad in (False, True):
        tb_y = E2M1[_nibbles(0x7777 | (1 << 16) | ((2 << 20) if pad else 0), 6)]
        const = float(np.dot(qa_y, tb_y)) * s_y
        for col in range(64):
            colsum = float(np.dot(qa_x, E2M1[_nibbles(_col_code(col), 4)])) * s_x
            assert colsum == col
            for h in ((0,) if pad else range(257)):
                main = 0.0 if pad else (2 * h - 256) * s_main[0]       # +1 per differing bit, -1 per equal bit, times 2^5
                # float32 accumulation in two different orders
                a = np.float32(0)
                for term in (main, colsum, const):
                    a = np.float32(a + np.float32(term))
                b = np.float32(0)
                for term in (const, colsum, main):
                    b = np.float32(b + np.float32(term))
                assert a == b and 2.0 ** 23 <= float(a) < 2.0 ** 24
                low16 = int(a.view(np.uint32)) & 0xFFFF
                if pad:
                    assert low16 == pad_key + col and low16 < 65536
                else:
                    assert low16 == key0 + 64 * h + col < pad_key
                    assert (low16 >> 6) - (key0 >> 6) == h and (low16 & 63) == col


def test_sign_nibbles_position_map():
    """sign_nibbles(): nibble j of output word s carries bit 4 j + 3 - s of the descriptor word, as 0x2 (+1.0) or 0xA (-1.0);
    with the query side built from ~w the dot product over a word is (differing bits) - (equal bits)."""
    rng = np.random.default_rng(5)

    def expand(w):
        out = []
        for s, mask in enumerate((0x88888888, 0x44444444, 0x22222222, 0x11111111)):
            out.append((((w & mask) << s) & 0xFFFFFFFF) + 0x22222222)
        return out

    for _ in range(200):
        a, b = (int(x) for x in rng.integers(0, 2 ** 32, 2, dtype=np.uint64))
        ea, eb = expand(~a & 0xFFFFFFFF), expand(b)
        dot = 0.0
        for wa, wb in zip(ea, eb):
            dot += float(np.dot(E2M1[_nibbles(wa, 8)], E2M1[_nibbles(wb, 8)]))
        ham = bin(a ^ b).count("1")
        assert dot == 2 * ham - 32
        for s, w in enumerate(eb):
            for j, nib in enumerate(_nibbles(w, 8)):
                assert nib == (0xA if (b >> (4 * j + 3 - s)) & 1 else 0x2)
