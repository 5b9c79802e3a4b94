// kernels_extract.cuh — sm_100a kernels of the ORB extraction path.
//
// Stage map (reference src/geometry/fextractor.cpp):
//   k_resize    ComputePyramid              :1135-1160  (cv::resize INTER_LINEAR, fixed point)
//   k_fast      ComputeKeyPointsOctTree     :756-820    (per-cell cv::FAST ini/min threshold + NMS)
//   k_octree    DistributeOctTree           :530-754    (quadtree culling, list order preserved)
//   k_orient    IC_Angle / fastAtan2        :68-95
//   k_emit      compute() output ordering   :1074-1129  (scale, mono / lapping split)
//   k_blur      GaussianBlur 7x7 s=2        :1086       (8-bit fixed point, reflect-101)
//   k_describe  computeOrbDescriptor        :99-138     (steered BRIEF-256)
#pragma once
#include "viorb_common.cuh"
#include "kernels_undistort.cuh"

static __device__ __align__(16) const int8_t d_orb_pattern[1024] = {
#include "orb_pattern.inc"
};

__device__ __forceinline__ int reflect101_dev(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
    return p;
}

// reflect-101 of a position at most 3 outside [0, n): branch-free for n >= 4
__device__ __forceinline__ int reflect101_near(int p, int n) {
    if (n < 4) return reflect101_dev(p, n);
    p = abs(p);
    return min(p, 2 * (n - 1) - p);
}

// ---------------------------------------------------------------------------------------------
// Pyramid level l from level l-1.  One thread = 4 adjacent output pixels x VIORB_RS_ROWS rows: the
// horizontal coefficients (11-bit, built on the host exactly like cv::resize does) are fetched once per
// thread with two 128-bit loads, and a source row's horizontal pass is reused by the next output row when
// both need it (the source advances 1.2 rows per output row).
// grid (ceil(w/256), ceil(h/(4*VIORB_RS_ROWS)), B), block (64,4).
// ---------------------------------------------------------------------------------------------
#ifndef VIORB_RS_ROWS
#define VIORB_RS_ROWS 8
#endif
#ifndef VIORB_RS_MINB
#define VIORB_RS_MINB 6          // <= 40 registers (88 B of spills) keeps 6 CTAs per SM: 0.491 ms against 0.573 ms at the 59 registers of the unrolled rows
#endif
__global__ void __launch_bounds__(256, VIORB_RS_MINB) k_resize(ExtractGeom g, FrameSrc fs, int l, const int* __restrict__ xofs,
                                                const short2* __restrict__ xalpha, const int* __restrict__ yofs,
                                                const short2* __restrict__ ybeta) {
    const LevelGeom& L = g.lv[l];
    const LevelGeom& S = g.lv[l - 1];
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int dy0 = (blockIdx.y * blockDim.y + threadIdx.y) * VIORB_RS_ROWS;
    const int f = blockIdx.z + g.frame0;
    if (x4 >= L.w || dy0 >= L.h) return;
    int sp;
    const uint8_t* src = level_ptr(g, fs, l - 1, f, sp);
    uint8_t* dst = fs.pyr + (size_t)f * g.arenaStride + L.off;
    const int nrow = min(VIORB_RS_ROWS, L.h - dy0);
    if (L.areaHalf) {
        for (int r = 0; r < nrow; ++r) {
            const int dy = dy0 + r;
            const uint8_t* a = src + (size_t)(2 * dy) * sp;
            const uint8_t* b = a + sp;
            for (int i = 0; i < 4 && x4 + i < L.w; ++i) {
                const int x = x4 + i;
                dst[(size_t)dy * L.pitch + x] = (uint8_t)((a[2 * x] + a[2 * x + 1] + b[2 * x] + b[2 * x + 1] + 2) >> 2);
            }
        }
        return;
    }
    // tables are padded to a multiple of 4 entries per level, so the vector loads stay inside them
    const int4 sx = __ldg(reinterpret_cast<const int4*>(xofs + L.xTab + x4));
    const int4 alr = __ldg(reinterpret_cast<const int4*>(xalpha + L.xTab + x4));
    const int o0[4] = {sx.x, sx.y, sx.z, sx.w};
    const int ar[4] = {alr.x, alr.y, alr.z, alr.w};
    // the host tables never point at the last source column (it is expressed as (w-2, weight 0|2048)),
    // so column o0 + 1 is always inside the row
    int a0[4], a1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        a0[i] = (short)(ar[i] & 0xFFFF);
        a1[i] = ar[i] >> 16;
    }
    if (L.rsWords) {
        // word path: a source row contributes three aligned 32-bit loads; the (left, right) byte pair of each output
        // pixel is picked with one PRMT and weighted with one IDP.2A (alpha pair as u16x2) - instead of two byte
        // loads and two multiply-adds per pixel and row
        const int wb = o0[0] >> 2;
        const int lastWord = (sp >> 2) - 1;
        const int w0i = min(wb, lastWord), w1i = min(wb + 1, lastWord), w2i = min(wb + 2, lastWord);
        unsigned sel[4];
        bool hi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int o = o0[i] - 4 * wb;                 // 0 .. 10
            hi[i] = o > 6;
            const int b = hi[i] ? o - 4 : o;              // first byte inside the chosen register pair
            sel[i] = (unsigned)b | ((unsigned)(b + 1) << 4);
        }
        unsigned g0[4], g1[4];
        int haveRow = -0x40000000;
        // h = (row sum) >> 4, the form the vertical pass consumes (each row sum is used by up to two output rows)
        auto hpass = [&](int sy, unsigned h[4]) {
            const uint32_t* r = reinterpret_cast<const uint32_t*>(src + (unsigned)(sy * sp));
            const unsigned a = __ldg(r + w0i), b = __ldg(r + w1i), c = __ldg(r + w2i);
#pragma unroll
            for (int i = 0; i < 4; ++i)
                h[i] = __dp2a_lo((unsigned)ar[i], __byte_perm(hi[i] ? b : a, hi[i] ? c : b, sel[i]), 0u) >> 4;
        };
        const int srcH1 = S.h - 1, opitch = L.pitch;
        uint8_t* d = dst + (size_t)dy0 * opitch + x4;
        // the rows' source offsets and weights: the y tables are padded to whole groups of VIORB_RS_ROWS entries per level
        static_assert(VIORB_RS_ROWS == 8, "two int4 loads per table");
        const int4 syA = __ldg(reinterpret_cast<const int4*>(yofs + L.yTab + dy0)), syB = __ldg(reinterpret_cast<const int4*>(yofs + L.yTab + dy0) + 1);
        const int4 btA = __ldg(reinterpret_cast<const int4*>(ybeta + L.yTab + dy0)), btB = __ldg(reinterpret_cast<const int4*>(ybeta + L.yTab + dy0) + 1);
        const int syv[8] = {syA.x, syA.y, syA.z, syA.w, syB.x, syB.y, syB.z, syB.w};
        const int btv[8] = {btA.x, btA.y, btA.z, btA.w, btB.x, btB.y, btB.z, btB.w};
#pragma unroll
        for (int r = 0; r < VIORB_RS_ROWS; ++r) {
            if (r >= nrow) break;
            const int sy = syv[r];
            const int btr = btv[r];
            const int sy0 = min(max(sy, 0), srcH1), sy1 = min(max(sy + 1, 0), srcH1);
            if (haveRow == sy0) {
#pragma unroll
                for (int i = 0; i < 4; ++i) g0[i] = g1[i];
            } else {
                hpass(sy0, g0);
            }
            hpass(sy1, g1);
            haveRow = sy1;
            // (beta * h) >> 16 as the high half of (beta << 16) * h: beta in [0, 2048], h < 2^20, one IMAD.HI each
            const unsigned B0 = (unsigned)btr << 16, B1 = (unsigned)btr & 0xFFFF0000u;
            uint32_t packed = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const unsigned v = (__umulhi(B0, g0[i]) + __umulhi(B1, g1[i]) + 2u) >> 2;      // <= 255
                packed = v * (1u << (8 * i)) + packed;
            }
            // rows are padded to 64 bytes: the last word of a row may spill into the padding
            *reinterpret_cast<uint32_t*>(d) = packed;
            d += opitch;
        }
        return;
    }
    int h0[4], h1[4], have = -0x40000000;       // source row whose horizontal pass sits in h1
    for (int r = 0; r < nrow; ++r) {
        const int dy = dy0 + r;
        const int sy = __ldg(yofs + L.yTab + dy);
        const short2 bt = __ldg(ybeta + L.yTab + dy);
        const int sy0 = min(max(sy, 0), S.h - 1), sy1 = min(max(sy + 1, 0), S.h - 1);
        if (have == sy0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) h0[i] = h1[i];
        } else {
            const uint8_t* r0 = src + (size_t)sy0 * sp;
#pragma unroll
            for (int i = 0; i < 4; ++i) { const uint8_t* q = r0 + o0[i]; h0[i] = q[0] * a0[i] + q[1] * a1[i]; }
        }
        {
            const uint8_t* r1 = src + (size_t)sy1 * sp;
#pragma unroll
            for (int i = 0; i < 4; ++i) { const uint8_t* q = r1 + o0[i]; h1[i] = q[0] * a0[i] + q[1] * a1[i]; }
            have = sy1;
        }
        uint32_t packed = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int v = ((((int)bt.x * (h0[i] >> 4)) >> 16) + (((int)bt.y * (h1[i] >> 4)) >> 16) + 2) >> 2;
            packed |= (uint32_t)(v & 0xFF) << (8 * i);
        }
        uint8_t* d = dst + (size_t)dy * L.pitch + x4;
        if (x4 + 3 < L.w) *reinterpret_cast<uint32_t*>(d) = packed;
        else for (int i = 0; i < 4 && x4 + i < L.w; ++i) d[i] = (uint8_t)(packed >> (8 * i));
    }
}

// ---------------------------------------------------------------------------------------------
// FAST-9 + cell-local 3x3 NMS + per-cell ini/min threshold.  One CTA = one strip (up to
// VIORB_FAST_STRIP_CELLS consecutive cells of one cell row) of one frame: per-CTA fixed costs and the
// five barriers are amortised over ~8k pixels, and rows are loaded as long coalesced runs.
//   phase 0  stage the cell region + 3-px halo in shared memory with 16-byte cp.async
//   phase A  opposite-pair pre-test per pixel -> queue of pixels that may be corners
//   phase B  dense corner score (cornerScore<16>) on the queue -> 8-bit score tile
//   phase C  strict 3x3 NMS inside the cell (outside = 0), survivors to shared
//   phase D  keep score >= iniTh if any survivor reaches it, else all (>= minTh); append to the
//            (frame, level) candidate list with one global atomic per warp
// Candidate word: x | y << 12 | score << 24, x/y relative to the 16-px border (as vToDistributeKeys).
// grid (totalStrips, B), block 256.
// ---------------------------------------------------------------------------------------------
#ifndef VIORB_FAST_THREADS
#define VIORB_FAST_THREADS 256
#endif

// cornerScore<16> of TWO pixels at once in s16x2 lanes (pixel A low, pixel B high): the 96 min/max of the
// arc network are shared.  Every lane carries its value + 256: centre - ring pixel + 256 lies in [1, 511], so ONE 32-bit
// multiply-add per pixel forms both lanes' differences without a borrow between them (the packed subtraction
// __vsub2 costs three ALU-pipe instructions on sm_100a, and that pipe bounds the kernel; the multiply-adds go to the
// FMA pipe, the multipliers come from constant memory so that ptxas keeps them there).  min / max commute with the
// bias.  Returns, per lane, score + 257 (a value below 257 + threshold = not a corner).
#define VIORB_FAST_SCORE_BIAS 257
#ifndef VIORB_EMU
__constant__ int c_fastMul[2] = {-1, -65536};
__device__ __forceinline__ unsigned fast_mad(unsigned a, int which, unsigned c) {
    unsigned r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"((unsigned)c_fastMul[which]), "r"(c));
    return r;
}
#else
__device__ __forceinline__ unsigned fast_mad(unsigned a, int which, unsigned c) { return a * (which ? 0xFFFF0000u : 0xFFFFFFFFu) + c; }
#endif
__device__ __forceinline__ unsigned fast_score16x2(const uint8_t* pa, const uint8_t* pb, const int* o) {
    const unsigned v = ((unsigned)pa[0] + ((unsigned)pb[0] << 16)) | 0x01000100u;         // centre + 256 per lane
    unsigned d[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = fast_mad((unsigned)pb[o[k]], 1, fast_mad((unsigned)pa[o[k]], 0, v));
    unsigned t[16], u[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        t[k] = __vimin3_s16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
        u[k] = __vimax3_s16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
    }
    // arc minima / maxima, then their max / min as 3-input trees (8 instead of 16 dependent 2-input steps each)
    unsigned m[16], n[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        m[k] = __vimin3_s16x2(t[k], t[(k + 3) & 15], t[(k + 6) & 15]);
        n[k] = __vimax3_s16x2(u[k], u[(k + 3) & 15], u[(k + 6) & 15]);
    }
    unsigned a = __vimax3_s16x2(__vimax3_s16x2(m[0], m[1], m[2]), __vimax3_s16x2(m[3], m[4], m[5]), __vimax3_s16x2(m[6], m[7], m[8]));
    a = __vimax3_s16x2(a, __vimax3_s16x2(m[9], m[10], m[11]), __vimax3_s16x2(m[12], m[13], m[14]));
    a = __vimax3_s16x2(a, m[15], 0u);                                               // -256 (+ bias) per lane
    unsigned b = __vimin3_s16x2(__vimin3_s16x2(n[0], n[1], n[2]), __vimin3_s16x2(n[3], n[4], n[5]), __vimin3_s16x2(n[6], n[7], n[8]));
    b = __vimin3_s16x2(b, __vimin3_s16x2(n[9], n[10], n[11]), __vimin3_s16x2(n[12], n[13], n[14]));
    b = __vimin3_s16x2(b, n[15], 0x02000200u);                                      // +256 (+ bias) per lane
    // max(a, -b) - 1 without the bias = max(a + 256, 512 - (b + 256)) - 257; 512 - (b + 256) >= 0 in both lanes: plain subtraction
    return __vmaxs2(a, 0x02000200u - b);
}

// bytes 0,1 (resp. 2,3) of w zero-extended into the two 16-bit lanes
__device__ __forceinline__ unsigned lo16x2(unsigned w) { return __byte_perm(w, 0, 0x4140); }
__device__ __forceinline__ unsigned hi16x2(unsigned w) { return __byte_perm(w, 0, 0x4342); }

// opposite-pair pre-test of 2 pixels packed as s16x2: a 9-arc brighter (darker) than centre +/- th
// contains one pixel of each opposite pair, so min over pairs of max(pair) > c + th (resp. max over
// pairs of min(pair) < c - th) is necessary.  Returns 0x4000 per lane that may be a corner.
// The comparisons run on biased lanes: with K = 0x4000 and kc = (K - th - 1) per lane, mb + kc - c >= K <=> bright
// possible and kc + c - md >= K <=> dark possible; every lane stays inside (0, 2K), so the two 3-input 32-bit adds
// never carry between lanes and bit 14 of each lane is the answer (3 instructions instead of 7 packed ones).
__device__ __forceinline__ unsigned fast_pretest2(unsigned c, unsigned up, unsigned dn, unsigned lf, unsigned rt,
                                                  unsigned d22, unsigned dm2m2, unsigned d2m2, unsigned dm22, unsigned kc) {
    const unsigned mb = __vmins2(__vimin3_s16x2(__vmaxs2(up, dn), __vmaxs2(lf, rt), __vmaxs2(d22, dm2m2)), __vmaxs2(d2m2, dm22));
    const unsigned md = __vmaxs2(__vimax3_s16x2(__vmins2(up, dn), __vmins2(lf, rt), __vmins2(d22, dm2m2)), __vmins2(d2m2, dm22));
    const unsigned xb = mb + kc - c;
    const unsigned yd = kc + c - md;
    return (xb | yd) & 0x40004000u;
}

// ---- k_fast: every pitch is a compile-time constant ---------------------------------------------------------------------
// The host cuts strips so that a region is at most VIORB_FAST_MAXRW pixels wide (wCell < 60, so a strip always holds >= 4
// cells); pixel tile and score tile share ONE pitch, VIORB_FAST_TP, so a queue entry is the linear offset py * TP + px and
// addresses both tiles without decoding, and the 16 ring offsets of the score network are immediates.
#ifndef VIORB_FAST_TP
#define VIORB_FAST_TP 336
#endif
#define VIORB_FAST_MAXRW (VIORB_FAST_TP - 48)               // 19 alignment shift + 6 halo + word right of the halo + chunk rounding
#ifndef VIORB_FAST_MAXAREA
#define VIORB_FAST_MAXAREA 7168                             // the host shortens strips of tall cell rows to this many pixels
#endif
__host__ __device__ inline size_t fast_smem_bytes(int tileH, int area) {
    return (size_t)tileH * VIORB_FAST_TP + (size_t)(tileH - 4) * VIORB_FAST_TP +   // pixel tile (rh + 6 rows), score tile (rh + 2 rows)
           (size_t)2 * area + 32;                                                   // queue (2 B / px); the survivors reuse the pixel tile
}

__global__ void __launch_bounds__(VIORB_FAST_THREADS) k_fast(ExtractGeom g, FrameSrc fs, const CellDesc* __restrict__ cells,
                                                             uint32_t* __restrict__ cand, int* __restrict__ candCount,
                                                             int* __restrict__ status) {
    constexpr int TP = VIORB_FAST_TP, TPW = TP >> 2;
    VIORB_DYN_SMEM(smem);
    __shared__ int qCount, sCount;
    __shared__ int hasIni[32];                               // per cell of the strip
    __shared__ uint2 wordMask[VIORB_FAST_MAXRW / 4 + 4];
    const int tid = threadIdx.x, lane = tid & 31;
    const int f = blockIdx.y + g.frame0;
    const CellDesc c = cells[blockIdx.x];
    const int l = c.level;
    const LevelGeom& L = g.lv[l];
    int pitch;
    const uint8_t* img = level_ptr(g, fs, l, f, pitch);
    const int rw = c.x1 - c.x0, rh = c.y1 - c.y0, wCell = L.wCell, ncell = c.ncell;
    uint8_t* T = smem;
    uint8_t* S = T + g.tileH * TP;
    uint16_t* Q = reinterpret_cast<uint16_t*>(S + (g.tileH - 4) * TP);          // entries: py * TP + px
    uint32_t* surv = reinterpret_cast<uint32_t*>(T);        // NMS survivors overwrite the pixel tile, dead after the score phase
    if (tid == 0) { qCount = 0; sCount = 0; }
    if (tid < 32) hasIni[tid] = 0;

    // phase 0: tile column 0 <-> image x = ax (16-byte aligned, at least one pad word left of the halo).  The chunks of
    // all rows form one flat index space (row = i / nC by the host's reciprocal): three 16-byte cp.async per thread.
    const int ox = c.x0 - 3, oy = c.y0 - 3;
    const int ax = ((ox & ~3) - 4) & ~15, shift = ox - ax;  // shift in [4,19]
    if (tid < c.ng) {
        // pixels of the first / last word of a region row that lie inside the region, as 0x4000-per-lane masks
        // (x: px 0,1; y: px 2,3); every word between them is all ones
        const int col0 = shift + 3, g0 = col0 >> 2, g1 = (col0 + rw - 1) >> 2;
        const int s0 = col0 - 4 * g0, e1 = col0 + rw - 1 - 4 * g1;
        uint2 m = make_uint2(0x40004000u, 0x40004000u);
        if (tid == 0) { m.x &= (s0 <= 0 ? 0x4000u : 0u) | (s0 <= 1 ? 0x40000000u : 0u); m.y &= (s0 <= 2 ? 0x4000u : 0u) | 0x40000000u; }
        if (tid == g1 - g0) { m.x &= 0x4000u | (e1 >= 1 ? 0x40000000u : 0u); m.y &= (e1 >= 2 ? 0x4000u : 0u) | (e1 >= 3 ? 0x40000000u : 0u); }
        wordMask[tid] = m;
    }
    {
        const int nC = c.nC, nChunk = (rh + 6) * nC;
        const uint8_t* src = img + (size_t)oy * pitch + ax;
        for (int i = tid; i < nChunk; i += VIORB_FAST_THREADS) {
            const int r = (int)(((unsigned)i * c.invNC) >> 20);
            const int k16 = (i - r * nC) << 4;
            const uint8_t* s = src + r * pitch + k16;
            uint8_t* d = T + r * TP + k16;
            if (ax + k16 + 16 <= pitch) {
                __pipeline_memcpy_async(d, s, 16);
            } else {                                         // never read past the end of an image row
                for (int j = 0; j < 4; ++j)
                    if (ax + k16 + 4 * j + 4 <= pitch) __pipeline_memcpy_async(d + 4 * j, s + 4 * j, 4);
            }
        }
        __pipeline_commit();
    }
    for (int i = tid; i < (((rh + 2) * TP) >> 4); i += VIORB_FAST_THREADS) reinterpret_cast<uint4*>(S)[i] = make_uint4(0, 0, 0, 0);
    __pipeline_wait_prior(0);
    __syncthreads();

    const int th = g.minTh;
    const int col0 = shift + 3;                             // tile column of region pixel 0
    const uint8_t* T0 = T + 3 * TP + col0;                  // region pixel (0,0)
    uint8_t* S0 = S + TP + 1;                               // its score

    // phase A: thread = aligned tile word (4 horizontally adjacent pixels) of the flattened region, packed 16x2
    // pre-test.  Pixels that may be corners go to the queue; its order is irrelevant, so a thread claims its
    // slots with one shared-memory atomic (MIO pipe) instead of a warp scan (the kernel is ALU-pipe bound).
    {
        const int g0 = col0 >> 2;
        const unsigned kc = (0x4000u - (unsigned)(th + 1)) * 0x10001u;                     // K - th - 1 in both lanes
        const int ng = c.ng, total = ng * rh;
        const unsigned ginv = c.ginv;                            // i / ng = (i * ginv) >> 22, exact while i * ng < 2^22 (ng <= 74, rh <= 127)
        const uint32_t* TW = reinterpret_cast<const uint32_t*>(T + 3 * TP) + g0;
        for (int i = tid; i < total; i += VIORB_FAST_THREADS) {
            const int py = (int)(((unsigned)i * ginv) >> 22);
            const int gi = i - py * ng;                          // word of the row, 0 = g0
            const uint32_t* R = TW + py * TPW + gi;
            const unsigned cw = R[0], wl = R[-1], wr = R[1];
            const unsigned up = R[-3 * TPW], dn = R[3 * TPW];
            const unsigned rt = __funnelshift_r(cw, wr, 24), lf = __funnelshift_r(wl, cw, 8);
            const unsigned d22 = __funnelshift_r(R[2 * TPW], R[2 * TPW + 1], 16), dm22 = __funnelshift_r(R[2 * TPW - 1], R[2 * TPW], 16);
            const unsigned d2m2 = __funnelshift_r(R[-2 * TPW], R[-2 * TPW + 1], 16), dm2m2 = __funnelshift_r(R[-2 * TPW - 1], R[-2 * TPW], 16);
            unsigned a = fast_pretest2(lo16x2(cw), lo16x2(up), lo16x2(dn), lo16x2(lf), lo16x2(rt), lo16x2(d22),
                                       lo16x2(dm2m2), lo16x2(d2m2), lo16x2(dm22), kc);
            unsigned b = fast_pretest2(hi16x2(cw), hi16x2(up), hi16x2(dn), hi16x2(lf), hi16x2(rt), hi16x2(d22),
                                       hi16x2(dm2m2), hi16x2(d2m2), hi16x2(dm22), kc);
            const uint2 wm = wordMask[gi];
            a &= wm.x; b &= wm.y;
            const unsigned ab = a | (b >> 1);
            if (ab) {
                const int q = atomicAdd(&qCount, __popc(ab));
                const int e0 = py * TP + ((g0 + gi) * 4 - col0);  // valid pixels have e0 + j >= py * TP
                const int f0 = (a >> 14) & 1, f1 = a >> 30, f2 = (b >> 14) & 1;
                if (f0) Q[q] = (uint16_t)e0;
                if (f1) Q[q + f0] = (uint16_t)(e0 + 1);
                if (f2) Q[q + f0 + f1] = (uint16_t)(e0 + 2);
                if (b >> 30) Q[q + f0 + f1 + f2] = (uint16_t)(e0 + 3);
            }
        }
    }
    __syncthreads();
    const int nq = qCount;

    // phase B: dense corner score on the queue, two queued pixels per thread (s16x2 lanes)
    {
        constexpr int o[16] = {3 * TP, 3 * TP + 1, 2 * TP + 2, TP + 3, 3, -TP + 3, -2 * TP + 2, -3 * TP + 1,
                               -3 * TP, -3 * TP - 1, -2 * TP - 2, -TP - 3, -3, TP - 3, 2 * TP - 2, 3 * TP - 1};
        for (int i = 2 * tid; i < nq; i += 2 * VIORB_FAST_THREADS) {
            const int ea = Q[i], eb = Q[min(i + 1, nq - 1)];
            const unsigned sc = fast_score16x2(T0 + ea, T0 + eb, o);
            const int sa = (int)(sc & 0xFFFF) - VIORB_FAST_SCORE_BIAS, sb = (int)(sc >> 16) - VIORB_FAST_SCORE_BIAS;
            if (sa >= th) S0[ea] = (uint8_t)sa;
            if (sb >= th) S0[eb] = (uint8_t)sb;
        }
    }
    __syncthreads();

    // phase C: strict 3x3 NMS run densely over the queue (only queued pixels can hold a score).  Neighbours in
    // another cell do not count (cv::FAST ran per cell): rows outside the strip are the zero ring of S, columns
    // across a cell boundary are masked here.  A thread keeps its first two survivors in registers and appends them
    // after the loop (10 % of the queue survives: the append, not the test, was the cost of this phase).
    const unsigned cinv = L.cinv;                                                 // px / wCell = (px * cinv) >> 20 for px < 16384
    {
        // pending survivors as e | score << 16 (never 0: score >= minTh >= 1), turned into candidate words at the end
        auto candidate = [&](uint32_t p) {
            const int e = (int)(p & 0xFFFF), py = (int)((unsigned)e / (unsigned)TP), px = e - py * TP;
            return (uint32_t)(c.x0 + px - VIORB_MINB) | ((uint32_t)(c.y0 + py - VIORB_MINB) << 12) | ((p >> 16) << 24);
        };
        uint32_t pend0 = 0, pend1 = 0;
        for (int i = tid; i < nq; i += VIORB_FAST_THREADS) {
            const int e = Q[i];
            const uint8_t* s = S0 + e;
            const int v = s[0];
            if (v == 0) continue;                                  // not a corner
            const int px = e - (int)((unsigned)e / (unsigned)TP) * TP;
            const int cell = min((int)(((unsigned)px * cinv) >> 20), ncell - 1);
            const int cl = px - cell * wCell;
            const bool hasL = cl > 0, hasR = cl < wCell - 1;      // the strip's own ends fall on the zero ring
            const int mc = max((int)s[-TP], (int)s[TP]);
            const int ml = hasL ? __vimax3_s32(s[-TP - 1], s[-1], s[TP - 1]) : 0;
            const int mr = hasR ? __vimax3_s32(s[-TP + 1], s[1], s[TP + 1]) : 0;
            if (v > __vimax3_s32(mc, ml, mr)) {
                if (v >= g.iniTh) hasIni[cell] = 1;
                const uint32_t p = (uint32_t)e | ((uint32_t)v << 16);
                if (pend1) surv[atomicAdd(&sCount, 1)] = candidate(p);   // third and later survivors of a thread: rare
                else if (pend0) pend1 = p;
                else pend0 = p;
            }
        }
        const int np = (pend0 != 0) + (pend1 != 0);
        if (np) {
            const int q = atomicAdd(&sCount, np);
            surv[q] = candidate(pend0);
            if (pend1) surv[q + 1] = candidate(pend1);
        }
    }
    __syncthreads();

    // phase D: per cell keep score >= iniTh if any survivor of that cell reaches it, else all of them
    const int ns = sCount;
    uint32_t* out = cand + (size_t)f * g.candPerFrame + L.candOff;
    int* cnt = candCount + f * g.nlevels + l;
    for (int base = 0; base < ns; base += VIORB_FAST_THREADS) {
        const int i = base + tid;
        uint32_t e = 0;
        bool keep = false;
        if (i < ns) {
            e = surv[i];
            const int px = (int)(e & 0xFFF) + VIORB_MINB - c.x0;
            const int cell = min((int)(((unsigned)px * cinv) >> 20), ncell - 1);
            keep = (int)(e >> 24) >= (hasIni[cell] ? g.iniTh : 0);
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (m) {
            const int leader = __ffs(m) - 1;
            int gb = 0;
            if (lane == leader) gb = atomicAdd(cnt, __popc(m));
            gb = __shfl_sync(0xffffffffu, gb, leader);
            if (keep) {
                const int pos = gb + __popc(m & ((1u << lane) - 1));
                if (pos < L.candCap) out[pos] = e;
                else atomicOr(status, 1);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Quadtree culling.  One CTA per (level, frame).  The std::list of the reference becomes an index
// order: position 0 = list front.  Every round splits a set of nodes "in push order"; children are
// pushed to the front (so they appear reversed), unsplit nodes keep their relative order behind.
//   full round    : every node with > 1 key, in list order                       (:614-662)
//   careful round : nodes of the last round, biggest first, until N is reached   (:664-728)
// The reference breaks (size) ties by heap pointer; here, as in the oracle, by creation order
// (later created first), which for nodes of one round equals "smaller list index first".
// Final pick per node: max response, first in cell-major raster order wins ties          (:735-751).
// grid (B, nlevels), two launches (k_octree / k_octree_dense, below), dynamic smem = octree_smem_bytes(maxNodes, cells).
// ---------------------------------------------------------------------------------------------
// Two launches share the work by candidate count n of the (level, frame):
//   k_octree        n <= VIORB_OCT_DENSE_KEYS   256-thread CTAs, several per SM: the many light jobs are barrier-latency
//                                               bound, what they need is CTAs in flight;
//   k_octree_dense  n >  VIORB_OCT_DENSE_KEYS   1024-thread CTAs (as many warps as n needs stay): a noise-like frame holds
//                                               50 times the candidates of a textured one and the launch ends with it.
// A CTA whose job belongs to the other launch exits at once.
#ifndef VIORB_OCT_THREADS
#define VIORB_OCT_THREADS 1024
#endif
#define VIORB_OCT_LIGHT_THREADS 256
#ifndef VIORB_OCT_LIGHT_MINB
#define VIORB_OCT_LIGHT_MINB 4
#endif
#ifndef VIORB_OCT_DENSE_KEYS
#define VIORB_OCT_DENSE_KEYS 4096
#endif
#ifndef VIORB_OCT_KEYS_PER_THREAD
#define VIORB_OCT_KEYS_PER_THREAD 8
#endif
#ifndef VIORB_OCT_AGG
#define VIORB_OCT_AGG 0
#endif
// count pyramid: per root a full 4-ary tree of key counts down to depth D0, D0 the largest depth with
// nIni * 4^D0 <= min(cells of the launch, n / VIORB_OCT_KEYS_PER_CELL); all levels together stay below 4/3 of the cells
#ifndef VIORB_OCT_KEYS_PER_CELL
#define VIORB_OCT_KEYS_PER_CELL 8
#endif
#define VIORB_OCT_PYR_CELLS 4096
#define VIORB_OCT_LIGHT_PYR_CELLS 512

struct OctBounds { short ulx, uly, brx, bry; };

// lutInts: the per-axis leaf tables of the counting pass (4 B per key column + 2 B per key row of the widest level, in ints)
__host__ __device__ inline int octree_lut_ints(int nx, int ny) { return nx + (ny + 1) / 2 + 4; }
__host__ __device__ inline size_t octree_smem_bytes(int nc, int pyrCells, int lutInts) {
    // bounds[2][nc] (8 B) + cnt[2][nc] + child[4 nc] + rank[nc] + aux0[nc] + aux1[nc] + newIdx[nc] + child2[4 nc] + code[2][nc]
    // + the count pyramid (VIORB_OCT_PYR_INTS)
    return (size_t)nc * (16 + 8 + 16 + 4 + 4 + 4 + 4 + 16 + 8) + 64 + (size_t)(pyrCells * 4 / 3 + 8 + lutInts) * 4;
}

// exclusive scan of a[0..m) in place by the whole CTA; returns the total (all threads).
// nthOverride: number of participating threads when the trailing warps of the CTA have already exited (k_octree)
__device__ __forceinline__ int block_excl_scan(int* a, int m, int* warpSums /* >= 33 ints */, int nthOverride = 0) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nth = nthOverride ? nthOverride : (int)blockDim.x;
    const int chunk = (m + nth - 1) / nth;
    const int st = min(tid * chunk, m), en = min(st + chunk, m);
    int sum = 0;
    for (int i = st; i < en; ++i) sum += a[i];
    int inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) warpSums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int v = lane < (nth >> 5) ? warpSums[lane] : 0;
        int vi = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, vi, d);
            if (lane >= d) vi += t;
        }
        if (lane < (nth >> 5)) warpSums[lane] = vi - v;
        if (lane == 31) warpSums[32] = vi;
    }
    __syncthreads();
    int run = warpSums[wid] + inc - sum;
    for (int i = st; i < en; ++i) { int t = a[i]; a[i] = run; run += t; }
    const int total = warpSums[32];
    __syncthreads();
    return total;
}

template <int THREADS, int PYR_CELLS>
__device__ __forceinline__ void octree_body(const ExtractGeom& g, const uint32_t* __restrict__ cand,
                                            const int* __restrict__ candCount, uint16_t* __restrict__ keyNode,
                                            KeptKp* __restrict__ kept, int* __restrict__ keptCount,
                                            int* __restrict__ status, int nLo, int nHi) {
    VIORB_DYN_SMEM(smem);
    __shared__ int warpSums[36];
    __shared__ int sh_live, sh_nExp, sh_cut;
    __shared__ OctBounds sh_rootB[64];
    const int tid = threadIdx.x;
    // grid (frames, levels): the CTAs of level 0 - the ones with the most keys - are dispatched first and the short ones
    // of the upper levels fill the tail (longest job first)
    const int l = blockIdx.y, f = blockIdx.x + g.frame0;
    const LevelGeom& L = g.lv[l];
    const int NC = g.maxNodes;
    const int n = min(candCount[f * g.nlevels + l], L.candCap);
    if (n < nLo || n > nHi) return;                        // the other launch's job
    // Of a dense CTA only as many warps as the job has work for stay; the others exit here (an exited warp counts as
    // arrived at every later barrier).
    const int nth = min(THREADS, max(VIORB_OCT_LIGHT_THREADS, ((n + VIORB_OCT_KEYS_PER_THREAD * 32 - 1) / (VIORB_OCT_KEYS_PER_THREAD * 32)) * 32));
    if (tid >= nth) return;
    const uint32_t* keys = cand + (size_t)f * g.candPerFrame + L.candOff;
    uint16_t* kn = keyNode + (size_t)f * g.candPerFrame + L.candOff;
    KeptKp* out = kept + (size_t)f * g.keptPerFrame + L.keptOff;
    if (n == 0) {
        if (tid == 0) keptCount[f * g.nlevels + l] = 0;
        return;
    }
    OctBounds* bnd[2];
    int* cnt[2];
    bnd[0] = reinterpret_cast<OctBounds*>(smem);
    bnd[1] = bnd[0] + NC;
    cnt[0] = reinterpret_cast<int*>(bnd[1] + NC);
    cnt[1] = cnt[0] + NC;
    int* child = cnt[1] + NC;          // [4 NC] child key counts, later child new index
    int* rankv = child + 4 * NC;       // push rank of a split node, -1 otherwise
    int* aux0 = rankv + NC;            // scans
    int* aux1 = aux0 + NC;
    int* newIdx = aux1 + NC;           // new index of an unsplit node
    int* child2 = newIdx + NC;         // [4 NC] child key counts of the NEXT round, filled while the keys move
    int* code[2];                      // per node: depth << 24 | root << 16 | path (2 bits per level), while depth <= D0
    code[0] = child2 + 4 * NC;
    code[1] = code[0] + NC;
    int* pyr = code[1] + NC;           // count pyramid; later the (depth, root, path) -> node table
    const int N = L.nWant;
    const int lane = tid & 31;
    // one shared-memory atomic per distinct counter of a warp (neighbouring candidates fall into the same node, and on
    // dense frames a node holds hundreds of keys: unaggregated, the 32 lanes serialise on one address)
    auto warp_add = [&](int* ctr, int c) {
#if VIORB_OCT_AGG
        const unsigned m = __match_any_sync(0xffffffffu, c);
        if (c >= 0 && lane == __ffs(m) - 1) atomicAdd(&ctr[c], __popc(m));
#else
        (void)lane;
        if (c >= 0) atomicAdd(&ctr[c], 1);
#endif
    };

    // ---- roots (:535-578)
    const int nIni = L.nIni;
    const float hX = L.hX;
    // The first D0 rounds never touch the keys.  Every expandable node of round r was created in round r - 1, so it has
    // depth r, and its children's key counts are entries of a count pyramid built ONCE: each key walks D0 levels down
    // from its root (the same ceil-half comparisons as DivideNode, :474-475, :506-517) and counts itself in its depth-D0
    // cell; the upper levels are sums.  Only if the tree must grow below depth D0 (sparse, clustered levels - few keys)
    // do the keys get their node and the per-round key passes start.  A dense level (tens of thousands of keys, a few
    // hundred nodes) finishes within the pyramid: 3 passes over the keys instead of 2 + one per round.
    // The pyramid's fixed cost (clearing, summing, the node table) is proportional to its cells: its depth follows the
    // key count, so that a light level (a few hundred keys) pays next to nothing and switches to key passes early.
    const int cellCap = min(PYR_CELLS, max(nIni, n / VIORB_OCT_KEYS_PER_CELL));
    int D0 = 0;
    while (D0 < 6 && (nIni << (2 * (D0 + 1))) <= cellCap) ++D0;
    auto lvl_off = [&](int d) { return nIni * (((1 << (2 * d)) - 1) / 3); };                 // nIni * (4^d - 1) / 3
    auto cell_idx = [&](int d, int r, int pth) { return lvl_off(d) + (r << (2 * d)) + pth; };
    const int pyrTotal = lvl_off(D0 + 1);
    for (int i = tid; i < nIni; i += nth) {
        OctBounds b;
        b.ulx = (short)(int)__fmul_rn(hX, (float)i);
        b.brx = (short)(int)__fmul_rn(hX, (float)(i + 1));
        b.uly = 0;
        b.bry = (short)(L.maxBY - VIORB_MINB);
        sh_rootB[i] = b;
    }
    for (int i = tid; i < pyrTotal; i += nth) pyr[i] = 0;
    // A key's leaf cell = (root << 2 D0) + path, two path bits per level: bit 0 from the x comparisons, bit 1 from the y
    // comparisons of DivideNode (:474-475, :506-517).  The x bits (and the root) depend on x alone, the y bits on y alone (all
    // roots span the full height), so the D0-level descent is done once per key COLUMN and once per key ROW into two tables,
    // and the counting pass is two table reads and an OR per key.
    uint32_t* lutX = reinterpret_cast<uint32_t*>(pyr + PYR_CELLS * 4 / 3 + 8);
    const int nX = L.maxBX - VIORB_MINB, nY = L.maxBY - VIORB_MINB;
    uint16_t* lutY = reinterpret_cast<uint16_t*>(lutX + nX);
    __syncthreads();                                   // sh_rootB
    for (int x = tid; x < nX; x += nth) {
        const int r = min((int)__fdiv_rn((float)x, hX), nIni - 1);
        int ul = sh_rootB[r].ulx, br = sh_rootB[r].brx, bits = 0;
        for (int d = 0; d < D0; ++d) {
            const int m = ul + ((br - ul + 1) >> 1);
            bits <<= 2;
            if (x < m) br = m; else { ul = m; bits |= 1; }
        }
        lutX[x] = (uint32_t)((r << (2 * D0)) | bits);
    }
    for (int y = tid; y < nY; y += nth) {
        int ul = 0, br = nY, bits = 0;
        for (int d = 0; d < D0; ++d) {
            const int m = ul + ((br - ul + 1) >> 1);
            bits <<= 2;
            if (y < m) br = m; else { ul = m; bits |= 2; }
        }
        lutY[y] = (uint16_t)bits;
    }
    __syncthreads();
    const int offD0 = lvl_off(D0);
    for (int base = 0; base < n; base += nth) {
        const int k = base + tid;
        int c = -1;
        if (k < n) {
            const uint32_t e = keys[k];
            const int leaf = (int)(lutX[e & 0xFFF] | lutY[(e >> 12) & 0xFFF]);
            c = offD0 + leaf;
            kn[k] = (uint16_t)leaf;                    // assign_keys finds the key's node from its leaf cell
        }
        warp_add(pyr, c);
    }
    __syncthreads();
    for (int d = D0 - 1; d >= 0; --d) {
        const int cells = nIni << (2 * d), o = lvl_off(d), oc = lvl_off(d + 1);
        for (int c = tid; c < cells; c += nth) pyr[o + c] = pyr[oc + 4 * c] + pyr[oc + 4 * c + 1] + pyr[oc + 4 * c + 2] + pyr[oc + 4 * c + 3];
        __syncthreads();
    }
    if (tid == 0) {
        int m = 0;
        for (int i = 0; i < nIni; ++i)
            if (pyr[i] > 0) { bnd[1][m] = sh_rootB[i]; cnt[1][m] = pyr[i]; code[1][m] = i << 16; ++m; }
        sh_live = m;
    }
    __syncthreads();
    int cur = 1;
    int live = sh_live;
    bool careful = false;
    bool usePyr = true;
    // (depth, root, path) -> node table over the pyramid's cells (-1 = split further), then every key finds its node among the
    // ancestors of its leaf cell (recorded by the counting pass: ancestor at depth d = leaf >> 2 (D0 - d), no second walk over
    // the midpoints and no second read of the key)
    auto assign_keys = [&](const OctBounds* /*B*/, const int* codes, int nLive) {
        for (int i = tid; i < pyrTotal; i += nth) pyr[i] = -1;
        __syncthreads();
        for (int i = tid; i < nLive; i += nth) {
            const int c = codes[i];
            pyr[cell_idx(c >> 24, (c >> 16) & 0xFF, c & 0xFFFF)] = i;
        }
        __syncthreads();
        for (int k = tid; k < n; k += nth) {
            const int leaf = kn[k];
            int d = 0, off = 0, t;
            while ((t = pyr[off + (leaf >> (2 * (D0 - d)))]) < 0 && d < D0) {
                off += nIni << (2 * d);
                ++d;
            }
            if (t < 0) { atomicOr(status, 8); t = 0; }          // cannot happen: the leaves partition the keys
            kn[k] = (uint16_t)t;
        }
        __syncthreads();
    };

    for (int round = 0; round < 64; ++round) {
        const OctBounds* B = bnd[cur];
        const int* Cn = cnt[cur];
        OctBounds* B2 = bnd[cur ^ 1];
        int* Cn2 = cnt[cur ^ 1];
        const int prevSize = live;
        // 1. child[4 i ..] = child key counts of every expandable node i: read from the count pyramid while the nodes
        //    are not deeper than D0, counted when the keys moved into the node afterwards
        if (usePyr && round == D0) {
            // the tree grows below the pyramid: give the keys their nodes and count the children of the expandable ones
            assign_keys(B, code[cur], live);
            usePyr = false;
            for (int i = tid; i < 4 * live; i += nth) child[i] = 0;
            __syncthreads();
            for (int k = tid; k < n; k += nth) {
                const int nd = kn[k];
                if (Cn[nd] > 1) {
                    const uint32_t e = keys[k];
                    const int x = e & 0xFFF, y = (e >> 12) & 0xFFF;
                    const OctBounds b = B[nd];
                    const int mx = b.ulx + ((b.brx - b.ulx + 1) >> 1), my = b.uly + ((b.bry - b.uly + 1) >> 1);
                    atomicAdd(&child[nd * 4 + (x < mx ? 0 : 1) + (y < my ? 0 : 2)], 1);
                }
            }
            __syncthreads();
        } else if (usePyr) {
            const int* cd = code[cur];
            for (int i = tid; i < live; i += nth)
                if (Cn[i] > 1) {
                    const int c = cd[i], d = c >> 24;
                    const int o = lvl_off(d + 1) + 4 * ((((c >> 16) & 0xFF) << (2 * d)) + (c & 0xFFFF));
                    child[4 * i] = pyr[o]; child[4 * i + 1] = pyr[o + 1]; child[4 * i + 2] = pyr[o + 2]; child[4 * i + 3] = pyr[o + 3];
                }
            __syncthreads();
        }
        // 2. push rank of the nodes to split
        int nSplit;
        if (!careful) {
            for (int i = tid; i < live; i += nth) aux0[i] = Cn[i] > 1 ? 1 : 0;
            __syncthreads();
            nSplit = block_excl_scan(aux0, live, warpSums, nth);
            for (int i = tid; i < live; i += nth) rankv[i] = Cn[i] > 1 ? aux0[i] : -1;
            __syncthreads();
        } else {
            // rank by (key count desc, list index asc); gains in rank order -> cut where N is reached
            for (int i = tid; i < live; i += nth) {
                int r = -1;
                const int ci = Cn[i];
                if (ci > 1) {
                    r = 0;
                    for (int j = 0; j < live; ++j) {
                        const int cj = Cn[j];
                        r += (cj > ci) || (cj == ci && j < i);
                    }
                }
                rankv[i] = r;
            }
            for (int i = tid; i < live; i += nth) aux0[i] = 0;
            if (tid == 0) sh_cut = 0x7fffffff;
            __syncthreads();
            for (int i = tid; i < live; i += nth)
                if (rankv[i] >= 0) {
                    const int* ch = child + 4 * i;
                    aux0[rankv[i]] = (ch[0] > 0) + (ch[1] > 0) + (ch[2] > 0) + (ch[3] > 0) - 1;
                }
            __syncthreads();
            int nE = 0;
            for (int i = tid; i < live; i += nth) nE += rankv[i] >= 0;
            if (tid == 0) sh_nExp = 0;
            __syncthreads();
            if (nE) atomicAdd(&sh_nExp, nE);
            __syncthreads();
            nE = sh_nExp;                           // number of expandable nodes
            __syncthreads();
            block_excl_scan(aux0, nE, warpSums, nth);    // aux0[r] = gain of ranks < r
            for (int r = tid; r < nE; r += nth) {
                // inclusive gain after splitting rank r = aux0[r+1] (or total); recompute from neighbours
                const int before = live + aux0[r];
                // gain of rank r itself: find via next exclusive value
                int after;
                if (r + 1 < nE) after = live + aux0[r + 1];
                else after = 0x7ffffffe;   // last rank: everything is split if no earlier cut
                if (before < N && after >= N) atomicMin(&sh_cut, r);
            }
            __syncthreads();
            const int cut = sh_cut;     // ranks <= cut are split (cut >= nE-1 means all)
            for (int i = tid; i < live; i += nth)
                if (rankv[i] > cut) rankv[i] = -1;
            __syncthreads();
            nSplit = min(nE, cut == 0x7fffffff ? nE : cut + 1);
        }
        if (nSplit == 0) break;            // nothing to split: size == prevSize (:655, :724)
        // 3. pushes: nonempty children in rank order
        for (int i = tid; i < nSplit; i += nth) aux0[i] = 0;
        __syncthreads();
        for (int i = tid; i < live; i += nth)
            if (rankv[i] >= 0) {
                const int* ch = child + 4 * i;
                aux0[rankv[i]] = (ch[0] > 0) + (ch[1] > 0) + (ch[2] > 0) + (ch[3] > 0);
            }
        __syncthreads();
        const int P = block_excl_scan(aux0, nSplit, warpSums, nth);      // aux0[rank] = first push position
        for (int i = tid; i < live; i += nth) aux1[i] = rankv[i] < 0 ? 1 : 0;
        __syncthreads();
        const int nKeep = block_excl_scan(aux1, live, warpSums, nth);    // aux1[i] = #unsplit before i
        const int newLive = P + nKeep;
        if (newLive > NC) {                 // cannot happen for maxNodes = max(N+3, 4 nIni); guard anyway
            if (tid == 0) atomicOr(status, 2);
            break;
        }
        if (tid == 0) sh_nExp = 0;
        __syncthreads();
        int myExp = 0;
        for (int i = tid; i < live; i += nth) {
            if (rankv[i] >= 0) {
                const OctBounds b = B[i];
                const int mx = b.ulx + ((b.brx - b.ulx + 1) >> 1), my = b.uly + ((b.bry - b.uly + 1) >> 1);
                int pp = aux0[rankv[i]];
                int* ch = child + 4 * i;
                const int cdi = code[cur][i];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = ch[q];
                    if (c > 0) {
                        const int ni = P - 1 - pp;
                        ++pp;
                        OctBounds nb;
                        nb.ulx = (q & 1) ? (short)mx : b.ulx;
                        nb.brx = (q & 1) ? b.brx : (short)mx;
                        nb.uly = (q & 2) ? (short)my : b.uly;
                        nb.bry = (q & 2) ? b.bry : (short)my;
                        B2[ni] = nb;
                        Cn2[ni] = c;
                        code[cur ^ 1][ni] = (((cdi >> 24) + 1) << 24) | (cdi & 0xFF0000) | (((cdi & 0xFFFF) << 2) + q);
                        myExp += c > 1;
                        ch[q] = ni;
                        child2[4 * ni] = 0; child2[4 * ni + 1] = 0; child2[4 * ni + 2] = 0; child2[4 * ni + 3] = 0;
                    }
                }
            } else {
                const int ni = P + aux1[i];
                B2[ni] = B[i];
                Cn2[ni] = Cn[i];
                code[cur ^ 1][ni] = code[cur][i];
                newIdx[i] = ni;
            }
        }
        if (myExp) atomicAdd(&sh_nExp, myExp);
        __syncthreads();
        // 4. (below the pyramid only) move the keys; a key that lands in a node with more than one key is counted for that
        //    node's next split
        for (int base = 0; !usePyr && base < n; base += nth) {
            const int k = base + tid;
            int c = -1;
            if (k < n) {
                const int nd = kn[k];
                if (rankv[nd] >= 0) {
                    const uint32_t e = keys[k];
                    const int x = e & 0xFFF, y = (e >> 12) & 0xFFF;
                    const OctBounds b = B[nd];
                    const int mx = b.ulx + ((b.brx - b.ulx + 1) >> 1), my = b.uly + ((b.bry - b.uly + 1) >> 1);
                    const int ni = child[nd * 4 + (x < mx ? 0 : 1) + (y < my ? 0 : 2)];
                    kn[k] = (uint16_t)ni;
                    if (Cn2[ni] > 1) {
                        const OctBounds nb = B2[ni];
                        const int mx2 = nb.ulx + ((nb.brx - nb.ulx + 1) >> 1), my2 = nb.uly + ((nb.bry - nb.uly + 1) >> 1);
                        c = ni * 4 + (x < mx2 ? 0 : 1) + (y < my2 ? 0 : 2);
                    }
                } else {
                    kn[k] = (uint16_t)newIdx[nd];
                }
            }
            warp_add(child2, c);
        }
        const int nToExpand = sh_nExp;
        __syncthreads();
        if (!usePyr) { int* t = child; child = child2; child2 = t; }
        cur ^= 1;
        live = newLive;
        // 5. termination (:655-662, :724-725)
        if (live >= N || live == prevSize) break;
        if (!careful && live + 3 * nToExpand > N) careful = true;
    }

    // ---- best key per node (:735-751): max score, ties -> first in (cell row, cell col, y, x) order
    if (usePyr) assign_keys(bnd[cur], code[cur], live);
    // two steps: the node's best score with a native 32-bit atomic, then only the keys that hold that score compete for
    // "first in the reference's order" with the 64-bit one
    unsigned long long* best = reinterpret_cast<unsigned long long*>(child);
    int* bestScore = aux0;
    for (int i = tid; i < live; i += nth) { best[i] = 0ull; bestScore[i] = 0; }
    __syncthreads();
    for (int k = tid; k < n; k += nth) {
        const int sc = (int)(keys[k] >> 24), nd = kn[k];
        if (sc > bestScore[nd]) atomicMax(&bestScore[nd], sc);
    }
    __syncthreads();
    const unsigned long long OMASK = (1ull << 40) - 1;
    for (int k = tid; k < n; k += nth) {
        const uint32_t e = keys[k];
        if ((int)(e >> 24) != bestScore[kn[k]]) continue;
        const int x = e & 0xFFF, y = (e >> 12) & 0xFFF;
        const unsigned long long sc = e >> 24;
        const int cj = min(max(x - 3, 0) / L.wCell, L.lastCol);
        const int ci = min(max(y - 3, 0) / L.hCell, L.lastRow);
        const unsigned long long okey = ((unsigned long long)ci << 32) | ((unsigned long long)cj << 24) |
                                        ((unsigned long long)y << 12) | (unsigned long long)x;
        atomicMax(&best[kn[k]], (sc << 40) | (OMASK - okey));
    }
    __syncthreads();
    for (int i = tid; i < live; i += nth) {
        const unsigned long long v = best[i];
        const unsigned long long okey = OMASK - (v & OMASK);
        KeptKp kp;
        kp.x = (short)((int)(okey & 0xFFF) + VIORB_MINB);
        kp.y = (short)((int)((okey >> 12) & 0xFFF) + VIORB_MINB);
        kp.score = (int)(v >> 40) | (l << 16);
        kp.angle = -1.f; kp.c = 1.f; kp.s = 0.f;
        if (i < L.keptCap) out[i] = kp;
    }
    if (tid == 0) {
        if (live > L.keptCap) atomicOr(status, 4);
        keptCount[f * g.nlevels + l] = min(live, L.keptCap);
    }
}

__global__ void __launch_bounds__(VIORB_OCT_LIGHT_THREADS, VIORB_OCT_LIGHT_MINB) k_octree(ExtractGeom g, const uint32_t* __restrict__ cand,
                                                                       const int* __restrict__ candCount, uint16_t* __restrict__ keyNode,
                                                                       KeptKp* __restrict__ kept, int* __restrict__ keptCount,
                                                                       int* __restrict__ status) {
    octree_body<VIORB_OCT_LIGHT_THREADS, VIORB_OCT_LIGHT_PYR_CELLS>(g, cand, candCount, keyNode, kept, keptCount, status, 0, VIORB_OCT_DENSE_KEYS);
}
__global__ void __launch_bounds__(VIORB_OCT_THREADS) k_octree_dense(ExtractGeom g, const uint32_t* __restrict__ cand,
                                                                    const int* __restrict__ candCount, uint16_t* __restrict__ keyNode,
                                                                    KeptKp* __restrict__ kept, int* __restrict__ keptCount,
                                                                    int* __restrict__ status) {
    octree_body<VIORB_OCT_THREADS, VIORB_OCT_PYR_CELLS>(g, cand, candCount, keyNode, kept, keptCount, status, VIORB_OCT_DENSE_KEYS + 1, 0x7fffffff);
}

// ---------------------------------------------------------------------------------------------
// Intensity-centroid orientation.  One warp per retained keypoint: lane = column u+15, loop over
// the 31 rows of the radius-15 disc (coalesced row segments), integer moments reduced exactly with
// __reduce_add_sync; cv::fastAtan2's float32 polynomial evaluated without FMA contraction.
// Also stores cos/sin of the angle for k_describe: double-precision sincos rounded to float.
// grid (ceil(keptPerFrame/8), B), block 256.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float sc = (float)(180.0 / 3.14159265358979323846);
    const float p1 = __fmul_rn(0.9997878412794807f, sc), p3 = __fmul_rn(-0.3258083974640975f, sc);
    const float p5 = __fmul_rn(0.1555786518463281f, sc), p7 = __fmul_rn(-0.04432655554792128f, sc);
    const float ax = fabsf(x), ay = fabsf(y);
    const float eps = 2.220446049250313e-16f;     // (float)DBL_EPSILON
    float a;
    if (ax >= ay) {
        const float c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        const float c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        const float c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        const float c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}


// dot product of 4 unsigned bytes with 4 signed bytes (IDP.4A.U8.S8)
__device__ __forceinline__ int dp4a_us(unsigned a, int b, int c) {
#ifdef VIORB_EMU
    for (int i = 0; i < 4; ++i) c += (int)((a >> (8 * i)) & 0xFF) * (int)(int8_t)((unsigned)b >> (8 * i));
    return c;
#else
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#endif
}

// One warp per retained keypoint.  lane = (row group rg = lane >> 3, word column wc = lane & 7): the lane handles
// bytes [4 wc, 4 wc + 4) of rows rg, rg + 4, ... of the 31x31 patch.  Every load instruction fetches 4 rows x 32
// contiguous bytes (coalesced sectors), all 16 loads of a lane are independent, and each row word costs one funnel
// shift and two IDP.4A against per-row weight tables built from umax: sum(I) and sum(u * I) over the row's disc
// segment.  m01 = sum_v v * sum(I), m10 = sum_v sum(u * I); integer, so the order of the adds is free.
// Weight tables of k_orient, built once on the host from umax: [0][|v|][k] = bytes 1 inside the disc segment of row
// v (word k of the 32-byte row window starting at column -15), [1][|v|][k] = the column offsets u as signed bytes.
struct OrientTab { int w[2][16][8]; };
__host__ inline void build_orient_tab(const int* umax, OrientTab& t) {
    for (int tb = 0; tb < 2; ++tb)
        for (int av = 0; av < 16; ++av)
            for (int k = 0; k < 8; ++k) {
                unsigned w = 0;
                for (int j = 0; j < 4; ++j) {
                    const int col = 4 * k + j, u = col - VIORB_HALF_PATCH;
                    if (col < 31 && (u < 0 ? -u : u) <= umax[av]) w |= (unsigned)((tb ? u : 1) & 0xFF) << (8 * j);
                }
                t.w[tb][av][k] = (int)w;
            }
}

// sin and cos of x in [0, 2 pi] in double precision (Cody-Waite reduction by pi/2 in two parts, fdlibm kernel
// polynomials, |error| < 1e-15): rounded to float32 they stand in for the reference's cosf / sinf.
__device__ __forceinline__ void sincos_small(double x, double& sn, double& cs) {
    const double q = rint(x * 6.36619772367581382433e-01);                     // 2 / pi
    double r = fma(-q, 1.57079632679489655800e+00, x);                         // pi/2 high
    r = fma(-q, 6.12323399573676603587e-17, r);                                // pi/2 low
    const double z = r * r;
    double ps = fma(z, 1.58969099521155010221e-10, -2.50507602534068634195e-08);
    ps = fma(z, ps, 2.75573137070700676789e-06);
    ps = fma(z, ps, -1.98412698298579493134e-04);
    ps = fma(z, ps, 8.33333333332248946124e-03);
    ps = fma(z, ps, -1.66666666666666324348e-01);
    const double s = fma(z * r, ps, r);
    double pc = fma(z, -1.13596475577881948265e-11, 2.08757232129817482790e-09);
    pc = fma(z, pc, -2.75573143513906633035e-07);
    pc = fma(z, pc, 2.48015872894767294178e-05);
    pc = fma(z, pc, -1.38888888888741095749e-03);
    pc = fma(z, pc, 4.16666666666666019037e-02);
    const double c = fma(z * z, pc, fma(z, -0.5, 1.0));
    const int n = (int)q & 3;
    sn = (n & 1) ? c : s;
    cs = (n & 1) ? s : c;
    if (n & 2) sn = -sn;
    if ((n + 1) & 2) cs = -cs;
}

#ifndef VIORB_ORIENT_KPW
#define VIORB_ORIENT_KPW 8                  // consecutive kept-arena slots handled by one warp
#endif
__global__ void __launch_bounds__(256) k_orient(ExtractGeom g, FrameSrc fs, const int* __restrict__ orientTab,
                                                const uint8_t* __restrict__ slotLevel, KeptKp* __restrict__ kept,
                                                const int* __restrict__ keptCount) {
    __shared__ int wOnes[16][8], wU[16][8];           // [|v|][word]: bytes = 1 resp. u inside the disc, 0 outside
    if (threadIdx.x < 128) (&wOnes[0][0])[threadIdx.x] = orientTab[threadIdx.x];
    else (&wU[0][0])[threadIdx.x - 128] = orientTab[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int slot0 = (blockIdx.x * 8 + (threadIdx.x >> 5)) * VIORB_ORIENT_KPW;
    const int f = blockIdx.y + g.frame0;
    if (slot0 >= g.keptPerFrame) return;
    // The warp sums the moments of VIORB_ORIENT_KPW keypoints one after the other; lane j keeps keypoint j's pair, and the
    // angle + sincos arithmetic (a third of the kernel when one lane did it per keypoint) runs once, on KPW lanes in parallel.
    int my10 = 0, my01 = 0;
    bool mine = false;
    const int wc = lane & 7, rg = lane >> 3;
#pragma unroll 1
    for (int j = 0; j < VIORB_ORIENT_KPW; ++j) {
        const int slot = slot0 + j;
        if (slot >= g.keptPerFrame) break;
        const int l = slotLevel[slot];                // the level whose slice of the kept arena holds this slot
        if (slot - g.lv[l].keptOff >= keptCount[f * g.nlevels + l]) continue;
        const KeptKp* kp = kept + (size_t)f * g.keptPerFrame + slot;
        int pitch;
        const uint8_t* img = level_ptr(g, fs, l, f, pitch);
        const int x = kp->x, y = kp->y;
        int m10 = 0, m01 = 0;
        // keypoints stay >= 19 px inside the level, so rows y - 15 .. y + 15 exist and none is the last one: the
        // aligned words around [x - 15, x + 15] never leave the image buffer.  pitch % 4 == 0: one shift for all rows.
        const uint8_t* p = img + (unsigned)((y - VIORB_HALF_PATCH) * pitch + (x - VIORB_HALF_PATCH));
        const unsigned sh = ((unsigned)(uintptr_t)p & 3u) * 8u;
        const uint8_t* a = p - ((uintptr_t)p & 3) + 4 * wc;
        unsigned lo[8], hi[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int r = rg + 4 * i;
            if (r < 31) {
                const uint32_t* q = reinterpret_cast<const uint32_t*>(a + (unsigned)(r * pitch));
                lo[i] = __ldg(q);
                hi[i] = __ldg(q + 1);
            } else { lo[i] = hi[i] = 0; }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int r = rg + 4 * i;
            if (r < 31) {
                const int v = r - VIORB_HALF_PATCH, av = v < 0 ? -v : v;
                const unsigned b = __funnelshift_r(lo[i], hi[i], sh);
                const int sumI = (int)__dp4a(b, (unsigned)wOnes[av][wc], 0u);
                m10 = dp4a_us(b, wU[av][wc], m10);
                m01 += v * sumI;
            }
        }
        m10 = __reduce_add_sync(0xffffffffu, m10);
        m01 = __reduce_add_sync(0xffffffffu, m01);
        if (lane == j) { my10 = m10; my01 = m01; mine = true; }
    }
    if (mine) {
        KeptKp* kp = kept + (size_t)f * g.keptPerFrame + slot0 + lane;
        const float ang = fast_atan2_deg((float)my01, (float)my10);
        const float factorPI = (float)(3.14159265358979323846 / 180.0);
        double sn, cs;
        sincos_small((double)__fmul_rn(ang, factorPI), sn, cs);
        kp->angle = ang;
        kp->c = (float)cs;
        kp->s = (float)sn;
    }
}

// ---------------------------------------------------------------------------------------------
// Output ordering (:1074-1129).  One CTA per frame walks the retained keypoints level by level in
// list order; a keypoint whose scaled x lies in [lap0, lap1] goes to the back (reverse order), the
// others to the front.  Writes cv::KeyPoint records and the slot -> kept-index map for k_describe.
// grid (B), block 256.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_emit(ExtractGeom g, const KeptKp* __restrict__ kept, const int* __restrict__ keptCount,
                                              int lap0, int lap1, viorb_keypoint* __restrict__ kps, int* __restrict__ slot2kept,
                                              int* __restrict__ nOut, int* __restrict__ monoOut, CamModel cam,
                                              viorb_keypoint* __restrict__ kpsUn) {
    __shared__ int warpCnt[8];
    __shared__ int sh_mono, sh_lap;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int f = blockIdx.x + g.frame0;
    int total = 0;
    for (int l = 0; l < g.nlevels; ++l) total += keptCount[f * g.nlevels + l];
    if (tid == 0) { sh_mono = 0; sh_lap = 0; }
    __syncthreads();
    viorb_keypoint* okp = kps + (size_t)f * g.outCap;
    int* s2k = slot2kept + (size_t)f * g.outCap;
    for (int l = 0; l < g.nlevels; ++l) {
        const LevelGeom& L = g.lv[l];
        const int cnt = keptCount[f * g.nlevels + l];
        const KeptKp* src = kept + (size_t)f * g.keptPerFrame + L.keptOff;
        for (int base = 0; base < cnt; base += 256) {
            const int i = base + tid;
            const bool valid = i < cnt;
            viorb_keypoint kp;
            bool lap = false;
            if (valid) {
                const KeptKp k = src[i];
                kp.x = (float)k.x; kp.y = (float)k.y;
                if (l != 0) { kp.x = __fmul_rn(kp.x, L.scale); kp.y = __fmul_rn(kp.y, L.scale); }
                kp.size = L.kpSize;
                kp.angle = k.angle;
                kp.response = (float)(k.score & 0xFFFF);
                kp.octave = l;
                kp.class_id = -1;
                lap = kp.x >= (float)lap0 && kp.x <= (float)lap1;
            }
            const unsigned mLap = __ballot_sync(0xffffffffu, valid && lap);
            const unsigned mMono = __ballot_sync(0xffffffffu, valid && !lap);
            if (lane == 0) warpCnt[wid] = (__popc(mLap) << 16) | __popc(mMono);
            __syncthreads();
            int lapBefore = sh_lap, monoBefore = sh_mono;
            for (int w = 0; w < wid; ++w) { lapBefore += warpCnt[w] >> 16; monoBefore += warpCnt[w] & 0xFFFF; }
            if (valid) {
                const unsigned lt = (1u << lane) - 1;
                const int slot = lap ? total - 1 - (lapBefore + __popc(mLap & lt)) : monoBefore + __popc(mMono & lt);
                okp[slot] = kp;
                s2k[slot] = L.keptOff + i;
                if (cam.mode) {   // Frame::UndistortKeyPoints fused into the epilogue (frame.cpp:758-791)
                    if (cam.mode == 2) {
                        const float2 p = undistort_point_dev(kp.x, kp.y, cam);
                        kp.x = p.x; kp.y = p.y;
                    }
                    kpsUn[(size_t)f * g.outCap + slot] = kp;
                }
            }
            __syncthreads();
            if (tid == 0) {
                int a = 0, b = 0;
                for (int w = 0; w < 8; ++w) { a += warpCnt[w] >> 16; b += warpCnt[w] & 0xFFFF; }
                sh_lap += a; sh_mono += b;
            }
            __syncthreads();
        }
    }
    if (tid == 0) { nOut[f] = total; monoOut[f] = sh_mono; }
}

// ---------------------------------------------------------------------------------------------
// 7x7 sigma-2 Gaussian, 8-bit fixed point: taps {18,34,48,56,48,34,18}/256, row pass Q8.8, column
// pass Q16.16, +0.5, >>16; BORDER_REFLECT_101 from the level image itself.  All levels in one launch:
// ---------------------------------------------------------------------------------------------
#define VIORB_BLUR_TW 256        // tile width: 64 threads x 4 pixels
#ifndef VIORB_BLUR_TH
#define VIORB_BLUR_TH 32         // output rows walked by one thread
#endif
struct BlurTiles { int base[VIORB_MAX_LEVELS + 1]; int tilesX[VIORB_MAX_LEVELS]; };

// One image row -> the four Q8.8 row sums of the thread's 4 columns.  w0,w1,w2 = the aligned words
// holding columns [4t-4, 4t+8); column j needs bytes 1+j .. 7+j of that 12-byte window.
__device__ __forceinline__ void blur_row4(unsigned w0, unsigned w1, unsigned w2, unsigned h[4]) {
    const unsigned K0 = 18u | (34u << 8) | (48u << 16) | (56u << 24), K1 = 48u | (34u << 8) | (18u << 16);
    h[0] = __dp4a(__funnelshift_r(w0, w1, 8), K0, __dp4a(__funnelshift_r(w1, w2, 8), K1, 0u));
    h[1] = __dp4a(__funnelshift_r(w0, w1, 16), K0, __dp4a(__funnelshift_r(w1, w2, 16), K1, 0u));
    h[2] = __dp4a(__funnelshift_r(w0, w1, 24), K0, __dp4a(__funnelshift_r(w1, w2, 24), K1, 0u));
    h[3] = __dp4a(w1, K0, __dp4a(w2, K1, 0u));
}

// One CTA = one 256 x 32 output tile.  The tile plus a 3-row / 16-byte apron is staged in shared memory with
// 16-byte cp.async copies (every load of the CTA is in flight at once; ncu showed the previous register-only
// version stalled on long-scoreboard global loads at 43 % occupancy and on instruction fetch of its three
// border paths).  Chunks that touch the image border are filled byte-wise with BORDER_REFLECT_101, so the
// compute phase is branch-free: each thread owns a 4-pixel column strip and walks down the rows keeping the
// last seven row sums in registers (unrolled by 7, the ring needs no moves): per row 3 LDS, 6 funnel shifts +
// 8 IDP.4A for the row pass, 28 multiply-adds for the column pass and one 32-bit store.
// grid (totalBlurTiles, B), block 64.
#define VIORB_BLUR_SP (VIORB_BLUR_TW + 32)       // shared row pitch: 16-byte apron left and right
__global__ void __launch_bounds__(64) k_blur(ExtractGeom g, FrameSrc fs, BlurTiles bt) {
    __shared__ __align__(16) uint8_t tile[(VIORB_BLUR_TH + 6) * VIORB_BLUR_SP];
    const int f = blockIdx.y + g.frame0;
    int l = 0;
    while (l + 1 < g.nlevels && (int)blockIdx.x >= bt.base[l + 1]) ++l;
    const LevelGeom& L = g.lv[l];
    const int t = blockIdx.x - bt.base[l];
    const int ty = t / bt.tilesX[l], tx = t - ty * bt.tilesX[l];
    const int X0 = tx * VIORB_BLUR_TW, y0 = ty * VIORB_BLUR_TH;
    int pitch;
    const uint8_t* img = level_ptr(g, fs, l, f, pitch);
    uint8_t* dst = fs.blur + (size_t)f * g.arenaStride + L.off;
    const int Lw = L.w, Lh = L.h;                    // (L indexes kernel parameters by a run-time level: keep the loops off the constant bank)
    const int yEnd = min(y0 + VIORB_BLUR_TH, Lh);
    const int nrows = yEnd - y0 + 6;
    const int NCH = VIORB_BLUR_SP / 16;              // 18 chunks per staged row
    for (int i = threadIdx.x; i < nrows * NCH; i += 64) {
        const int r = (i * 3641) >> 16, c = i - r * NCH;            // i / 18 for i < 2^13
        const int cx = X0 - 16 + 16 * c;
        if (cx >= 0 && cx + 16 <= Lw) {
            const int sy = reflect101_near(y0 - 3 + r, Lh);
            __pipeline_memcpy_async(tile + r * VIORB_BLUR_SP + 16 * c, img + (unsigned)(sy * pitch + cx), 16);
        }
    }
    // Columns that no whole in-image chunk covers: the 3 reflected columns left of the image (first tile column) and
    // the tail [w & ~15, w + 3) (last tile column), at most 3 + 18 bytes per row.  One lane per such column, two rows per
    // pass: a lane-serial byte loop inside the chunk loop above made the whole warp wait ~300 instructions per chunk.
    {
        const int nL = X0 == 0 ? 3 : 0;
        const int cR = max(Lw & ~15, X0 - 16), eR = min(Lw + 3, X0 + VIORB_BLUR_TW + 16);
        const int nE = nL + max(eR - cR, 0);
        const int e = threadIdx.x & 31;
        if (e < nE) {
            const int colx = e < nL ? e - 3 : cR + (e - nL);
            const uint8_t* scol = img + reflect101_near(colx, Lw);
            uint8_t* tcol = tile + (colx - X0 + 16);
            for (int r = threadIdx.x >> 5; r < nrows; r += 2)
                tcol[r * VIORB_BLUR_SP] = scol[(unsigned)(reflect101_near(y0 - 3 + r, Lh) * pitch)];
        }
    }
    __pipeline_commit();
    __pipeline_wait_prior(0);
    __syncthreads();
    const int x0 = X0 + threadIdx.x * 4;
    if (x0 >= Lw) return;
    const uint32_t* col = reinterpret_cast<const uint32_t*>(tile + 16 + threadIdx.x * 4);
    const int SPW = VIORB_BLUR_SP / 4;
    unsigned ring[7][4];
#pragma unroll
    for (int r = 0; r < 6; ++r) blur_row4(col[r * SPW - 1], col[r * SPW], col[r * SPW + 1], ring[r]);
    const uint32_t* nxt = col + 6 * SPW;             // staged row of image row y + 3
    const int opitch = L.pitch;
    uint8_t* orow = dst + (size_t)y0 * opitch + x0;
    for (int yb = y0; yb < yEnd; yb += 7) {
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const int y = yb + k;
            if (y < yEnd) {
                // rows y-3 .. y+2 sit in ring[(k+0..5) % 7]; the new row y+3 goes to ring[(k+6) % 7]
                blur_row4(nxt[-1], nxt[0], nxt[1], ring[(k + 6) % 7]);
                nxt += SPW;
                unsigned out = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const unsigned acc = 18u * (ring[k % 7][j] + ring[(k + 6) % 7][j]) + 34u * (ring[(k + 1) % 7][j] + ring[(k + 5) % 7][j]) +
                                         48u * (ring[(k + 2) % 7][j] + ring[(k + 4) % 7][j]) + 56u * ring[(k + 3) % 7][j];
                    out |= ((acc + 32768u) >> 16) << (8 * j);
                }
                // rows are padded to 64 bytes, so the last word of a row may spill into the padding
                *reinterpret_cast<uint32_t*>(orow) = out;
                orow += opitch;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Steered BRIEF-256.  One warp per output slot, lane = descriptor byte (8 tests, 16 samples).
// Sample offset = cvRound(x*b + y*a), cvRound(x*a - y*b): float32 products and sums rounded
// separately (no FMA), round-half-even.  grid (ceil(outCap/8), B), block 256.
// ---------------------------------------------------------------------------------------------
#define VIORB_DESC_R 18                     // largest rounded sample offset: |(13,13)| = 18.4
#define VIORB_DESC_PW 64                    // staged patch row: 37 bytes + alignment shift (<= 15) in four 16-byte chunks
// round-half-even of |v| < 2^22 on the FMA pipe (F2I would go to the quarter-rate XU pipe): adding 1.5 * 2^23
// leaves the integer in the low mantissa bits.  Same result as cvRound / lrintf.
__device__ __forceinline__ int round_rn_magic(float v) {
    return __float_as_int(__fadd_rn(v, 12582912.0f)) - 0x4B400000;
}

#ifndef VIORB_DESC_KPW
#define VIORB_DESC_KPW 8                    // consecutive output slots handled by one warp
#endif
#ifndef VIORB_DESC_MINB
#define VIORB_DESC_MINB 4                   // <= 64 registers: 4 CTAs per SM (measured: 0.37 ms against 0.39 at 79 registers)
#endif
__global__ void __launch_bounds__(256, VIORB_DESC_MINB) k_describe(ExtractGeom g, FrameSrc fs, const KeptKp* __restrict__ kept,
                                                  const int* __restrict__ slot2kept, const int* __restrict__ nOut,
                                                  uint8_t* __restrict__ desc) {
    // two patch buffers per warp: the next keypoint's patch is in flight (cp.async) while this one is sampled
    __shared__ __align__(16) uint8_t patch[8][2][(2 * VIORB_DESC_R + 1) * VIORB_DESC_PW];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int f = blockIdx.y + g.frame0;
    const int n = nOut[f];
    const int slot0 = (blockIdx.x * 8 + wid) * VIORB_DESC_KPW;
    if (slot0 >= n) return;
    // the lane's 8 tests (descriptor byte `lane`) stay in registers for all of the warp's keypoints: the shared-memory
    // pattern table of the previous version cost 8 LDS.128 (32 L1 wavefronts) per keypoint in a kernel that ncu showed
    // bound by L1 wavefronts
    float4 pt[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        const char4 q = *reinterpret_cast<const char4*>(d_orb_pattern + 4 * (lane * 8 + t));
        pt[t] = make_float4((float)q.x, (float)q.y, (float)q.z, (float)q.w);
    }
    const KeptKp* kf = kept + (size_t)f * g.keptPerFrame;
    const int* s2k = slot2kept + (size_t)f * g.outCap;
    const int NCH = (2 * VIORB_DESC_R + 1) * (VIORB_DESC_PW / 16);
    // stage the (2R+1)-row neighbourhood of keypoint k with aligned 16-byte asynchronous copies (the 512 samples would
    // otherwise be 512 scattered byte requests to L1): 5 copies per lane.  A keypoint lies at least 19 pixels inside the
    // level, so the rows exist; the last chunk of a row may run a few bytes into the next row of the level (never sampled).
    auto stage = [&](const KeptKp& k, uint8_t* P) {
        const LevelGeom& L = g.lv[k.score >> 16];
        const int lp = L.pitch;
        const uint8_t* src = fs.blur + (size_t)f * g.arenaStride + L.off + (size_t)((k.y - VIORB_DESC_R) * lp + ((k.x - VIORB_DESC_R) & ~15));
#pragma unroll
        for (int i0 = 0; i0 < NCH; i0 += 32) {
            const int i = i0 + lane;
            if (i < NCH) __pipeline_memcpy_async(P + 16 * i, src + (unsigned)((i >> 2) * lp + 16 * (i & 3)), 16);
        }
        __pipeline_commit();
    };
    KeptKp k = kf[s2k[slot0]];
    stage(k, patch[wid][0]);
#pragma unroll 1
    for (int j = 0; j < VIORB_DESC_KPW; ++j) {
        const int slot = slot0 + j;
        const bool more = j + 1 < VIORB_DESC_KPW && slot + 1 < n;
        KeptKp kn = k;
        if (more) {
            kn = kf[s2k[slot + 1]];
            stage(kn, patch[wid][(j + 1) & 1]);
            __pipeline_wait_prior(1);
        } else {
            __pipeline_wait_prior(0);
        }
        __syncwarp();
        const int shift = (k.x - VIORB_DESC_R) & 15;
        const uint8_t* ctr = patch[wid][j & 1] + VIORB_DESC_R * VIORB_DESC_PW + VIORB_DESC_R + shift;
        const float a = k.c, b = k.s;
        int val = 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const float4 p = pt[t];
            const int r0 = round_rn_magic(__fadd_rn(__fmul_rn(p.x, b), __fmul_rn(p.y, a)));
            const int c0 = round_rn_magic(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, b)));
            const int r1 = round_rn_magic(__fadd_rn(__fmul_rn(p.z, b), __fmul_rn(p.w, a)));
            const int c1 = round_rn_magic(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, b)));
            const int t0 = ctr[r0 * VIORB_DESC_PW + c0], t1 = ctr[r1 * VIORB_DESC_PW + c1];
            val |= (t0 < t1) << t;
        }
        desc[((size_t)f * g.outCap + slot) * 32 + lane] = (uint8_t)val;
        if (!more) break;
        __syncwarp();                       // everyone is done with buffer j & 1 before it is refilled
        k = kn;
    }
}
