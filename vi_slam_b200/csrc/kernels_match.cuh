// kernels_match.cuh — sm_100a kernels of the descriptor matching path.
//
//   k_knn2            cv::BFMatcher(NORM_HAMMING).knnMatch(k=2)   frame.cpp:1167, fmatcher.cpp:214
//   k_ratio           Lowe ratio rules                             frame.cpp:1170, fmatcher.cpp:220-223
//   k_dedup_*         removeDuplicatedMatches                      fmatcher.cpp:247-266
//   k_stereo_coarse   ComputeStereoMatches row-band Hamming        frame.cpp:833-906
//   k_stereo_refine   11x11 SAD slide + parabola sub-pixel         frame.cpp:909-979
//   k_stereo_cut      median-based outlier cut                     frame.cpp:983-996
//   k_desc_dist       FMatcher::DescriptorDistance                 fmatcher.cpp:2859-2875
#pragma once
#include "viorb_common.cuh"

#define VIORB_KNN_THREADS 128
#define VIORB_KNN_IDX_BITS 22          // train index bits in the packed (distance, index) key
#define VIORB_TH_HIGH 100              // FMatcher::TH_HIGH (fmatcher.cpp:313)
#define VIORB_TH_LOW 50                // FMatcher::TH_LOW  (fmatcher.cpp:314)

// A batch of descriptor sets: set p lives at base + p*stride, its row count is n[p] (device array) or
// nFixed when n == nullptr.
struct DescSet {
    const uint8_t* base;
    size_t stride;
    const int* n;
    int nFixed;
};
__device__ __forceinline__ int descset_count(const DescSet& s, int p) { return s.n ? s.n[p] : s.nFixed; }

// One result row of the kNN-2 kernels.  dist != NULL: idx[row][2] and dist[row][2] (cv::DMatch-like pairs).
// dist == NULL: ONE 16-byte row (idx0, idx1, dist0, dist1) at idx[row][4] - the row format the row-sharded matcher
// all-gathers (viorb_knn2_sharded), written by the kernel itself so that no packing pass sits before the exchange.
__device__ __forceinline__ void store_knn_row(int32_t* __restrict__ idx, int32_t* __restrict__ dist, size_t row, int i0, int i1,
                                              int d0, int d1) {
    if (dist) {
        *reinterpret_cast<int2*>(idx + row * 2) = make_int2(i0, i1);
        *reinterpret_cast<int2*>(dist + row * 2) = make_int2(d0, d1);
    } else {
        *reinterpret_cast<int4*>(idx + row * 4) = make_int4(i0, i1, d0, d1);
    }
}

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// 256-bit Hamming distance with a carry-save adder tree.  POPC issues on the XU pipe (one warp instruction per
// 8 clocks per SM sub-partition: ncu showed xu 91 % busy with 8 POPC per pair), LOP3 / IADD3 on the ALU pipe (one
// per 2 clocks).  Three 3:2 compressors (6 LOP3) over seven of the eight XOR words leave five words to count:
// XU 5 x 8 = 40 clocks and ALU (8 XOR + 6 CSA + 2 IADD3 + 3 min/max) x 2 = 38 clocks per warp-pair - the balance
// point between the two pipes (4 POPC + 8 LOP3 is ALU bound at 44, 6 POPC is XU bound at 48).
// Returns (ones, twos) with distance = ones + 2 * twos.
__device__ __forceinline__ int2 hamming256_csa(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    const unsigned s0 = x0 ^ x1 ^ x2, c0 = (x0 & x1) | (x2 & (x0 ^ x1));
    const unsigned s1 = x3 ^ x4 ^ x5, c1 = (x3 & x4) | (x5 & (x3 ^ x4));
    const unsigned s2 = s0 ^ s1 ^ x6, c2 = (s0 & s1) | (x6 & (s0 ^ s1));
    return make_int2(__popc(s2) + __popc(x7), __popc(c0) + __popc(c1) + __popc(c2));
}

// The two kernels below are build-time variants (VIORB_KNN_VARIANT 2 / 1; viorb_api.cu): the product library holds
// k_knn2_tc5 only.
#if VIORB_KNN_VARIANT == 2
// ---------------------------------------------------------------------------------------------
// Brute-force Hamming 2-NN.  One thread = one query row (256 bits in 8 registers); train rows are
// staged 128 at a time in shared memory and read as warp-wide broadcasts (2 x LDS.128 per row).
// Per pair: 8 XOR + 8 LOP3 (carry-save tree) + 4 POPC + adds, then a 3-instruction running top-2 on the packed key
// (distance << 22 | train index), which also yields cv::BFMatcher's tie rule (lower index first).
// grid (ceil(maxNq/128), nSets), block 128.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(VIORB_KNN_THREADS) k_knn2(DescSet Q, DescSet T, int32_t* __restrict__ idx,
                                                            int32_t* __restrict__ dist, size_t outStride) {
    __shared__ uint4 tile[VIORB_KNN_THREADS * 2];
    const int tid = threadIdx.x, p = blockIdx.y;
    const int nq = descset_count(Q, p), nt = descset_count(T, p);
    const int q0 = blockIdx.x * VIORB_KNN_THREADS;
    if (q0 >= nq) return;
    const int qi = q0 + tid;
    const bool valid = qi < nq;
    const uint4* qrow = reinterpret_cast<const uint4*>(Q.base + (size_t)p * Q.stride) + (size_t)(valid ? qi : q0) * 2;
    const uint4 a0 = __ldg(qrow), a1 = __ldg(qrow + 1);
    const uint4* tb = reinterpret_cast<const uint4*>(T.base + (size_t)p * T.stride);
    unsigned best0 = 0xFFFFFFFFu, best1 = 0xFFFFFFFFu;
    for (int t0 = 0; t0 < nt; t0 += VIORB_KNN_THREADS) {
        const int m = min(VIORB_KNN_THREADS, nt - t0);
        __syncthreads();
        if (tid < m) {
            tile[tid * 2] = __ldg(tb + (size_t)(t0 + tid) * 2);
            tile[tid * 2 + 1] = __ldg(tb + (size_t)(t0 + tid) * 2 + 1);
        }
        __syncthreads();
        const unsigned kb = (unsigned)t0;
#pragma unroll 4
        for (int j = 0; j < m; ++j) {
            const uint4 b0 = tile[2 * j], b1 = tile[2 * j + 1];
            // key = (ones + 2 twos) << 22 | index, built with two multiply-adds (FMA pipe, not the ALU pipe)
            const int2 h = hamming256_csa(a0, a1, b0, b1);
            const unsigned key = (unsigned)h.y * (2u << VIORB_KNN_IDX_BITS) + ((unsigned)h.x * (1u << VIORB_KNN_IDX_BITS) + (kb + j));
            const unsigned lo = min(key, best0);
            best1 = min(best1, max(key, best0));
            best0 = lo;
        }
    }
    if (valid) {
        const unsigned mask = (1u << VIORB_KNN_IDX_BITS) - 1;
        store_knn_row(idx, dist, (size_t)p * outStride + qi, best0 == 0xFFFFFFFFu ? -1 : (int)(best0 & mask),
                      best1 == 0xFFFFFFFFu ? -1 : (int)(best1 & mask), best0 == 0xFFFFFFFFu ? -1 : (int)(best0 >> VIORB_KNN_IDX_BITS),
                      best1 == 0xFFFFFFFFu ? -1 : (int)(best1 >> VIORB_KNN_IDX_BITS));
    }
}
#endif  // VIORB_KNN_VARIANT == 2

#if VIORB_KNN_VARIANT == 1
// ---------------------------------------------------------------------------------------------
// Brute-force Hamming 2-NN on the tensor cores.  Hamming(a, b) = popc(a) + popc(b) - 2 a.b, and the dot product
// of two 256-bit rows is an integer GEMM once the bits are expanded to {0,1} bytes.  sm_100a has no native
// 1-bit MMA (mma.sync ... b1 lowers to 8x IMMA.16832.U8 plus software expansion); the warp
// level IMMA.16832.U8.U8 itself runs at 0.5 MMA / clk / SM (measured, tools/probe/imma_probe.cu) = 8 descriptor
// pairs / clk / SM against 3.2 for the LOP3/POPC kernel at its pipe limit.
//   warp = 32 query rows (two 16-row A tiles, bits expanded once into 64 registers)
//   CTA  = 4 warps = 128 query rows; train rows are staged 128 at a time as BITS (4 KB) and a thread expands
//          the nibble it needs with one multiply and one AND ((n * 0x00204081) & 0x01010101) right before the
//          MMA, so shared-memory traffic stays 1/8 of an expanded tile
//   per 8 train rows: 16 IMMA (8 k-steps x 2 A tiles), then per accumulator one multiply-add builds the packed
//          key (distance << 22 | train index) and three min/max keep the running top-2, exactly as in k_knn2.
// Results are bit-identical to k_knn2 (same keys, same tie rule).  grid (ceil(maxNq/128), nSets), block 128.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned expand_nibble(unsigned n) { return (n * 0x00204081u) & 0x01010101u; }

#define VIORB_KNN_NONE 0x7F000000      // key of a padding train row: above every real key ((256 << 22) | index < 2^30 + 2^22)

__global__ void __launch_bounds__(VIORB_KNN_THREADS) k_knn2_mma(DescSet Q, DescSet T, int32_t* __restrict__ idx,
                                                                int32_t* __restrict__ dist, size_t outStride) {
    // train tile, 128 rows: as bits (staging), and expanded to {0,1} bytes in B-fragment order
    // [n-block 16][chunk 4][lane 32] x 16 bytes: lane (g, t) of n-block nb finds its b0 / b1 words of k-steps 2c, 2c+1
    // in chunk c, so a warp reads its fragments with four conflict-free LDS.128 per n-block
    __shared__ uint4 bits[VIORB_KNN_THREADS * 2];
    __shared__ uint4 frag[16 * 4 * 32];
    __shared__ int kc[VIORB_KNN_THREADS];                  // popc(train row) << 22 | train index
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, p = blockIdx.y;
    const int g = lane >> 2, t = lane & 3;
    const int nq = descset_count(Q, p), nt = descset_count(T, p);
    const int q0 = blockIdx.x * VIORB_KNN_THREADS;
    if (q0 >= nq) return;
    const uint4* qb = reinterpret_cast<const uint4*>(Q.base + (size_t)p * Q.stride);
    const uint4* tb = reinterpret_cast<const uint4*>(T.base + (size_t)p * T.stride);
    // A fragments: rows q0 + 32 warp + 16 a + {g, g + 8}
    unsigned A[2][8][4];
    int pqk[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int row = min(q0 + 32 * warp + 16 * a + g + 8 * h, nq - 1);
            const uint4 w0 = __ldg(qb + (size_t)row * 2), w1 = __ldg(qb + (size_t)row * 2 + 1);
            const unsigned w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
            int pc = 0;
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                pc += __popc(w[s]);
                A[a][s][h] = expand_nibble((w[s] >> (4 * t)) & 15u);
                A[a][s][2 + h] = expand_nibble((w[s] >> (4 * t + 16)) & 15u);
            }
            pqk[a][h] = pc << VIORB_KNN_IDX_BITS;
        }
    }
    // running top-2 per query row on keys WITHOUT the query term popc(q) << 22 (constant per row: it does not change
    // the order and is added at the end); the keys are signed: (popc(t) - 2 q.t) << 22 | index
    int best0[2][2], best1[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int h = 0; h < 2; ++h) best0[a][h] = best1[a][h] = 0x7FFFFFFF;
    for (int t0 = 0; t0 < nt; t0 += VIORB_KNN_THREADS) {
        const int m = min(VIORB_KNN_THREADS, nt - t0);
        __syncthreads();                                   // the previous tile's fragments are no longer read
        if (tid < m) {
            const uint4 r0 = __ldg(tb + (size_t)(t0 + tid) * 2), r1 = __ldg(tb + (size_t)(t0 + tid) * 2 + 1);
            bits[tid * 2] = r0;
            bits[tid * 2 + 1] = r1;
            const int pc = __popc(r0.x) + __popc(r0.y) + __popc(r0.z) + __popc(r0.w) + __popc(r1.x) + __popc(r1.y) + __popc(r1.z) + __popc(r1.w);
            kc[tid] = (pc << VIORB_KNN_IDX_BITS) | (t0 + tid);
        } else {
            bits[tid * 2] = make_uint4(0, 0, 0, 0);
            bits[tid * 2 + 1] = make_uint4(0, 0, 0, 0);
            kc[tid] = VIORB_KNN_NONE;
        }
        __syncthreads();
        // expand once per CTA: producer thread (warp pw, lane (g, t)) writes the chunks of n-blocks pw, pw + 4, ...
        const unsigned* bw = reinterpret_cast<const unsigned*>(bits);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int nb = warp + 4 * i;
            if (nb * 8 < m) {
                const unsigned* row = bw + (nb * 8 + g) * 8;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const unsigned wa = row[2 * c], wb = row[2 * c + 1];
                    frag[(nb * 4 + c) * 32 + lane] = make_uint4(expand_nibble((wa >> (4 * t)) & 15u), expand_nibble((wa >> (4 * t + 16)) & 15u),
                                                                expand_nibble((wb >> (4 * t)) & 15u), expand_nibble((wb >> (4 * t + 16)) & 15u));
                }
            }
        }
        __syncthreads();
        for (int nb = 0; nb * 8 < m; ++nb) {
            int acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint4 f = frag[(nb * 4 + c) * 32 + lane];
                mma_u8_16832(acc[0], A[0][2 * c], f.x, f.y);
                mma_u8_16832(acc[1], A[1][2 * c], f.x, f.y);
                mma_u8_16832(acc[0], A[0][2 * c + 1], f.z, f.w);
                mma_u8_16832(acc[1], A[1][2 * c + 1], f.z, f.w);
            }
            const int k0 = kc[nb * 8 + 2 * t], k1 = kc[nb * 8 + 2 * t + 1];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int key = (j ? k1 : k0) - acc[a][2 * h + j] * (2 << VIORB_KNN_IDX_BITS);
                        best1[a][h] = min(best1[a][h], max(key, best0[a][h]));
                        best0[a][h] = min(best0[a][h], key);
                    }
        }
    }
    // the candidates of one query row sit in the four lanes of a quad: merge the (best, second) pairs
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int b0 = best0[a][h], b1 = best1[a][h];
#pragma unroll
            for (int d = 1; d <= 2; d <<= 1) {
                const int o0 = __shfl_xor_sync(0xffffffffu, b0, d), o1 = __shfl_xor_sync(0xffffffffu, b1, d);
                b1 = min(max(b0, o0), min(b1, o1));
                b0 = min(b0, o0);
            }
            const int qi = q0 + 32 * warp + 16 * a + g + 8 * h;
            if (t == 0 && qi < nq) {
                const int mask = (1 << VIORB_KNN_IDX_BITS) - 1;
                const bool h0 = b0 < VIORB_KNN_NONE, h1 = b1 < VIORB_KNN_NONE;
                const int f0 = b0 + pqk[a][h], f1 = b1 + pqk[a][h];
                store_knn_row(idx, dist, (size_t)p * outStride + qi, h0 ? (f0 & mask) : -1, h1 ? (f1 & mask) : -1,
                              h0 ? (f0 >> VIORB_KNN_IDX_BITS) : -1, h1 ? (f1 >> VIORB_KNN_IDX_BITS) : -1);
            }
        }
}
#endif  // VIORB_KNN_VARIANT == 1

// ---------------------------------------------------------------------------------------------
// Ratio test + ordered compaction.  One CTA per set; accepted matches are written in query order.
// rule 0: d0 < d1 * ratio with d as float, product in double (frame.cpp:1170).
// rule 1: (double)d0 < ratio * d1 (fmatcher.cpp:220-223).
// lrMatch (optional): per query the accepted train index or -1.
// grid (nSets), block 256.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_ratio(DescSet Q, const int32_t* __restrict__ idx, const int32_t* __restrict__ dist,
                                               size_t knnStride, double ratio, int rule, viorb_match* __restrict__ out,
                                               size_t outStride, int* __restrict__ nOut, int32_t* __restrict__ lrMatch) {
    __shared__ int warpCnt[8];
    __shared__ int sh_base;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, p = blockIdx.x;
    const int nq = descset_count(Q, p);
    const int32_t* I = idx + (size_t)p * knnStride * 2;
    const int32_t* D = dist + (size_t)p * knnStride * 2;
    viorb_match* o = out + (size_t)p * outStride;
    if (tid == 0) sh_base = 0;
    __syncthreads();
    for (int base = 0; base < nq; base += 256) {
        const int i = base + tid;
        bool ok = false;
        int i0 = -1, d0i = 0;
        if (i < nq) {
            i0 = I[2 * i];
            const int i1 = I[2 * i + 1];
            d0i = D[2 * i];
            if (i0 >= 0 && i1 >= 0) {
                const double d0 = (double)(float)d0i, d1 = (double)(float)D[2 * i + 1];
                ok = rule == 0 ? (d0 < d1 * ratio) : (d0 < ratio * d1);
            }
            if (lrMatch) lrMatch[(size_t)p * knnStride + i] = ok ? i0 : -1;
        }
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) warpCnt[wid] = __popc(m);
        __syncthreads();
        int before = sh_base;
        for (int w = 0; w < wid; ++w) before += warpCnt[w];
        if (ok) {
            viorb_match mm;
            mm.query_idx = i; mm.train_idx = i0; mm.distance = d0i;
            const int pos = before + __popc(m & ((1u << lane) - 1));
            if ((size_t)pos < outStride) o[pos] = mm;
        }
        __syncthreads();
        if (tid == 0) { int s = 0; for (int w = 0; w < 8; ++w) s += warpCnt[w]; sh_base += s; }
        __syncthreads();
    }
    if (tid == 0) nOut[p] = sh_base;
}

// removeDuplicatedMatches, deterministic form: per train index keep the accepted match with the
// smallest query index; output ordered by train index.  Single set.
__global__ void k_dedup_mark(const viorb_match* __restrict__ m, int n, int* __restrict__ firstQ) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicMin(&firstQ[m[i].train_idx], m[i].query_idx);
}
__global__ void __launch_bounds__(256) k_dedup_emit(const int* __restrict__ firstQ, int nt, const int32_t* __restrict__ dist,
                                                    viorb_match* __restrict__ out, int* __restrict__ nOut) {
    __shared__ int warpCnt[8];
    __shared__ int sh_base;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) sh_base = 0;
    __syncthreads();
    for (int base = 0; base < nt; base += 256) {
        const int t = base + tid;
        const int q = t < nt ? firstQ[t] : 0x7fffffff;
        const bool ok = q != 0x7fffffff;
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) warpCnt[wid] = __popc(m);
        __syncthreads();
        int before = sh_base;
        for (int w = 0; w < wid; ++w) before += warpCnt[w];
        if (ok) {
            viorb_match mm;
            mm.query_idx = q; mm.train_idx = t; mm.distance = dist[2 * q];
            out[before + __popc(m & ((1u << lane) - 1))] = mm;
        }
        __syncthreads();
        if (tid == 0) { int s = 0; for (int w = 0; w < 8; ++w) s += warpCnt[w]; sh_base += s; }
        __syncthreads();
    }
    if (tid == 0) *nOut = sh_base;
}

__global__ void k_desc_dist(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int n, int32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* pa = reinterpret_cast<const uint4*>(a) + (size_t)i * 2;
    const uint4* pb = reinterpret_cast<const uint4*>(b) + (size_t)i * 2;
    out[i] = hamming256(pa[0], pa[1], pb[0], pb[1]);
}

// ---------------------------------------------------------------------------------------------
// Rectified stereo, coarse stage.  One warp per left keypoint; the right keypoints' row band
// [floor(y - 2 s), ceil(y + 2 s)], octave and x are staged 256 at a time in shared memory, lanes
// stride over them, and only candidates passing the row / octave / disparity predicate pay for the
// 256-bit Hamming distance.  Winner = min (distance, right index) with distance < TH_HIGH, which is
// the reference's "first strictly smaller wins" over its ascending-index row lists.
// grid (ceil(outCap/8), nFrames), block 256.
// ---------------------------------------------------------------------------------------------
struct StereoParams {
    float bf, maxD;                         // mbf, mbf / mb
    float scale[VIORB_MAX_LEVELS];
    float invScale[VIORB_MAX_LEVELS];
    int nlevels, outCap;
};

// Row table (frame.cpp:833-850): for every image row the right keypoints whose band
// [floor(y - 2 s), ceil(y + 2 s)] covers it, as a CSR structure built by one CTA per frame
// (count, block scan, fill).  Order inside a row is arbitrary; the matcher's winner rule
// (min over (distance, index)) does not depend on it.  A list entry carries what the coarse matcher filters on
// (index | octave << 20, x as float bits), so a candidate costs one 8-byte load before the descriptor is touched.
// grid (nFrames), block 256, dynamic smem 2 * (rows + 1) ints.
__global__ void __launch_bounds__(256) k_stereo_rows(StereoParams sp, int rows, int listCap, const viorb_keypoint* __restrict__ kr,
                                                     const int* __restrict__ nR, int* __restrict__ rowStart,
                                                     int2* __restrict__ rowList, int* __restrict__ status) {
    VIORB_DYN_SMEM(smem);
    __shared__ int warpSums[36];
    int* cnt = reinterpret_cast<int*>(smem);
    int* cur = cnt + rows + 1;
    const int tid = threadIdx.x, f = blockIdx.x;
    const int nr = nR[f];
    const viorb_keypoint* KR = kr + (size_t)f * sp.outCap;
    for (int i = tid; i <= rows; i += 256) { cnt[i] = 0; cur[i] = 0; }
    __syncthreads();
    for (int i = tid; i < nr; i += 256) {
        const viorb_keypoint k = KR[i];
        const float r = __fmul_rn(2.0f, sp.scale[k.octave]);
        const int maxr = min(__float2int_ru(__fadd_rn(k.y, r)), rows - 1);
        const int minr = max(__float2int_rd(__fsub_rn(k.y, r)), 0);
        for (int y = minr; y <= maxr; ++y) atomicAdd(&cnt[y], 1);
    }
    __syncthreads();
    const int total = block_excl_scan(cnt, rows + 1, warpSums);      // cnt[y] = start of row y, cnt[rows] = total
    int* RS = rowStart + (size_t)f * (rows + 1);
    for (int i = tid; i <= rows; i += 256) RS[i] = min(cnt[i], listCap);
    if (total > listCap && tid == 0) atomicOr(status, 8);
    int2* RL = rowList + (size_t)f * listCap;
    for (int i = tid; i < nr; i += 256) {
        const viorb_keypoint k = KR[i];
        const float r = __fmul_rn(2.0f, sp.scale[k.octave]);
        const int maxr = min(__float2int_ru(__fadd_rn(k.y, r)), rows - 1);
        const int minr = max(__float2int_rd(__fsub_rn(k.y, r)), 0);
        for (int y = minr; y <= maxr; ++y) {
            const int pos = cnt[y] + atomicAdd(&cur[y], 1);
            if (pos < listCap) RL[pos] = make_int2(i | (k.octave << 20), __float_as_int(k.x));
        }
    }
}

// One warp per left keypoint: lanes stride over the candidates of its row, test octave and disparity
// range, and only then pay for the 256-bit Hamming distance.
#ifndef VIORB_COARSE_MINB
#define VIORB_COARSE_MINB 8      // 32 registers, all 64 warps resident: 0.160 -> 0.139 ms (load-latency bound)
#endif
__global__ void __launch_bounds__(256, VIORB_COARSE_MINB) k_stereo_coarse(StereoParams sp, int rows, int listCap, const viorb_keypoint* __restrict__ kl,
                                                       const uint8_t* __restrict__ dl, const int* __restrict__ nL,
                                                       const uint8_t* __restrict__ dr, const int* __restrict__ rowStart,
                                                       const int2* __restrict__ rowList,
                                                       int32_t* __restrict__ bestIdx, int32_t* __restrict__ bestDist) {
    const int lane = threadIdx.x & 31, f = blockIdx.y;
    const int iL = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (iL >= nL[f]) return;
    const viorb_keypoint k = kl[(size_t)f * sp.outCap + iL];
    const uint4* DL = reinterpret_cast<const uint4*>(dl + (size_t)f * sp.outCap * 32);
    const uint4* DR = reinterpret_cast<const uint4*>(dr + (size_t)f * sp.outCap * 32);
    const float uL = k.x, minU = __fsub_rn(k.x, sp.maxD);
    const int row = (int)k.y, levelL = k.octave;
    int best = VIORB_TH_HIGH << 20;
    if (!(uL < 0.f) && row >= 0 && row < rows) {          // maxU = uL - minD, minD = 0 (frame.cpp:872-876)
        const uint4 a0 = DL[(size_t)iL * 2], a1 = DL[(size_t)iL * 2 + 1];
        const int* RS = rowStart + (size_t)f * (rows + 1);
        const int2* RL = rowList + (size_t)f * listCap;
        const int e = RS[row + 1];
        for (int j = RS[row] + lane; j < e; j += 32) {
            const int2 m = RL[j];
            const int iR = m.x & 0xFFFFF, oc = m.x >> 20;
            const float uR = __int_as_float(m.y);
            if (oc >= levelL - 1 && oc <= levelL + 1 && uR >= minU && uR <= uL) {
                const int d = hamming256(a0, a1, DR[(size_t)iR * 2], DR[(size_t)iR * 2 + 1]);
                best = min(best, (d << 20) | iR);
            }
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0) {
        const bool found = best < (VIORB_TH_HIGH << 20);
        bestIdx[(size_t)f * sp.outCap + iL] = found ? (best & 0xFFFFF) : -1;
        bestDist[(size_t)f * sp.outCap + iL] = found ? (best >> 20) : -1;
    }
}

// ---------------------------------------------------------------------------------------------
// Stereo refinement.  One warp per left keypoint whose coarse distance is < (TH_HIGH+TH_LOW)/2:
// 11x11 centre-subtracted L1 over 11 horizontal shifts on pyramid level octave(left) of both images
// (integer arithmetic: identical to the reference's float sums, which are exact below 2^24), then
// the parabola fit and the disparity / depth arithmetic in float32 without FMA contraction.
// grid (ceil(outCap/8), nFrames), block 256.
// ---------------------------------------------------------------------------------------------
#ifndef VIORB_REFINE_MINB
#define VIORB_REFINE_MINB 6      // <= 40 registers (36 B of spills): the kernel is a chain of dependent loads per keypoint, occupancy pays (0.197 -> 0.180 ms against 5 CTAs per SM; 8 CTAs at 32 registers: 0.199)
#endif
__global__ void __launch_bounds__(256, VIORB_REFINE_MINB) k_stereo_refine(StereoParams sp, ExtractGeom g, FrameSrc fsL, FrameSrc fsR,
                                                       const viorb_keypoint* __restrict__ kl, const int* __restrict__ nL,
                                                       const viorb_keypoint* __restrict__ kr, const int32_t* __restrict__ bestIdx,
                                                       const int32_t* __restrict__ bestDist, float* __restrict__ uRight,
                                                       float* __restrict__ depth, int32_t* __restrict__ sad) {
    const int lane = threadIdx.x & 31, f = blockIdx.y;
    const int iL = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (iL >= nL[f]) return;
    const size_t o = (size_t)f * sp.outCap + iL;
    float outU = -1.f, outD = -1.f;
    int outS = -1;
    const int bd = bestDist[o];
    if (bd >= 0 && bd < (VIORB_TH_HIGH + VIORB_TH_LOW) / 2) {
        const viorb_keypoint kpL = kl[o];
        const int lv = kpL.octave;
        const float uR0 = kr[(size_t)f * sp.outCap + bestIdx[o]].x;
        const float sf = sp.invScale[lv];
        const float su = roundf(__fmul_rn(kpL.x, sf)), sv = roundf(__fmul_rn(kpL.y, sf)), sr = roundf(__fmul_rn(uR0, sf));
        const int cu = (int)su, cv = (int)sv, cr = (int)sr;
        const LevelGeom& L = g.lv[lv];
        const float endu = __fadd_rn(__fadd_rn(__fadd_rn(sr, 5.f), 5.f), 1.f);
        if (!(sr < 0.f) && !(endu >= (float)L.w)) {
            int pl, pr;
            const uint8_t* imL = level_ptr(g, fsL, lv, f, pl);
            const uint8_t* imR = level_ptr(g, fsR, lv, f, pr);
            const int cL = imL[(size_t)cv * pl + cu];
            // |(a - cL) - (b - cR_k)| = |(a + cR_k - cL) - b|: one add and one VABSDIFF (absolute difference + accumulate) per
            // (pixel, shift) instead of two subtractions, IABS, select and add
            int tk[11];
            unsigned acc[11];
#pragma unroll
            for (int k = 0; k < 11; ++k) { tk[k] = (int)imR[(unsigned)(cv * pr + cr + k - 5)] - cL; acc[k] = 0u; }
            // the four passes over the 121 patch pixels are unrolled so that all 48 byte loads of a lane are in flight
            // together (the kernel is load-latency bound)
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int p = min(lane + 32 * it, 120);
                const bool on = lane + 32 * it < 121;
                const int dy = p / 11 - 5, dx = p % 11 - 5;
                const int a = (int)imL[(unsigned)((cv + dy) * pl + cu + dx)];
                const uint8_t* rr = imR + (unsigned)((cv + dy) * pr + cr + dx - 5);
                int b[11];
#pragma unroll
                for (int k = 0; k < 11; ++k) b[k] = (int)rr[k];
                if (on) {
#pragma unroll
                    for (int k = 0; k < 11; ++k) acc[k] = __sad(a + tk[k], b[k], acc[k]);
                }
            }
#pragma unroll
            for (int k = 0; k < 11; ++k) acc[k] = __reduce_add_sync(0xffffffffu, acc[k]);
            int bestS = 0x7fffffff, bestInc = 0;
#pragma unroll
            for (int k = 0; k < 11; ++k)
                if ((int)acc[k] < bestS) { bestS = (int)acc[k]; bestInc = k - 5; }
            if (bestInc != -5 && bestInc != 5) {
                float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
                for (int k = 1; k < 10; ++k)
                    if (k - 5 == bestInc) { d1 = (float)acc[k - 1]; d2 = (float)acc[k]; d3 = (float)acc[k + 1]; }
                const float deltaR = __fdiv_rn(__fsub_rn(d1, d3),
                                               __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                if (!(deltaR < -1.f) && !(deltaR > 1.f)) {
                    float bestuR = __fmul_rn(sp.scale[lv], __fadd_rn(__fadd_rn(sr, (float)bestInc), deltaR));
                    float disparity = __fsub_rn(kpL.x, bestuR);
                    if (disparity >= 0.f && disparity < sp.maxD) {
                        if (disparity <= 0.f) {
                            disparity = 0.01f;
                            bestuR = (float)((double)kpL.x - 0.01);
                        }
                        outD = __fdiv_rn(sp.bf, disparity);
                        outU = bestuR;
                        outS = bestS;
                    }
                }
            }
        }
    }
    if (lane == 0) { uRight[o] = outU; depth[o] = outD; sad[o] = outS; }
}

// ---------------------------------------------------------------------------------------------
// Outlier cut: median of the accepted SAD values (element size/2 of the ascending list), reject
// everything with sad >= 1.5 * 1.4 * median.  16-bit radix select (SAD <= 121 * 510 < 2^16) with
// two 256-bin shared histograms.  grid (nFrames), block 256.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_stereo_cut(int outCap, const int* __restrict__ nL, const int32_t* __restrict__ sad,
                                                    float* __restrict__ uRight, float* __restrict__ depth, int* __restrict__ nStereo) {
    __shared__ int hist[256];
    __shared__ int sh_cnt, sh_bin, sh_k, sh_kept;
    const int tid = threadIdx.x, f = blockIdx.x;
    const int n = nL[f];
    const int32_t* S = sad + (size_t)f * outCap;
    hist[tid] = 0;
    if (tid == 0) { sh_cnt = 0; sh_kept = 0; }
    __syncthreads();
    int loc = 0;
    for (int i = tid; i < n; i += 256) {
        const int v = S[i];
        if (v >= 0) { atomicAdd(&hist[(v >> 8) & 0xFF], 1); ++loc; }
    }
    if (loc) atomicAdd(&sh_cnt, loc);
    __syncthreads();
    const int cnt = sh_cnt;
    if (cnt == 0) { if (tid == 0) nStereo[f] = 0; return; }
    if (tid == 0) {
        int k = cnt / 2, b = 0;
        while (k >= hist[b]) { k -= hist[b]; ++b; }
        sh_bin = b; sh_k = k;
    }
    __syncthreads();
    const int hiBin = sh_bin;
    __syncthreads();
    hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
        const int v = S[i];
        if (v >= 0 && ((v >> 8) & 0xFF) == hiBin) atomicAdd(&hist[v & 0xFF], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int k = sh_k, b = 0;
        while (k >= hist[b]) { k -= hist[b]; ++b; }
        sh_bin = (hiBin << 8) | b;
    }
    __syncthreads();
    const float median = (float)sh_bin;
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), median);
    int kept = 0;
    for (int i = tid; i < n; i += 256) {
        const int v = S[i];
        if (v >= 0) {
            if ((float)v >= thDist) { uRight[(size_t)f * outCap + i] = -1.f; depth[(size_t)f * outCap + i] = -1.f; }
            else ++kept;
        }
    }
    if (kept) atomicAdd(&sh_kept, kept);
    __syncthreads();
    if (tid == 0) nStereo[f] = sh_kept;
}
