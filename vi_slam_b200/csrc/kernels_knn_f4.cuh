// kernels_knn_f4.cuh — brute-force Hamming kNN-2 on the 5th-generation tensor cores in 4-bit floating point:
// tcgen05.mma kind::mxf4.block_scale (packed e2m1 operands, ue8m0 scales per 32 elements, fp32 accumulators in tensor
// memory).  Same contract as every other kNN-2 kernel here: cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, 2)
// (frame.cpp:1167, fmatcher.cpp:214), ties -> lower train index.  Results are bit-identical to the kind::i8 kernel
// (kernels_knn_tc5.cuh, build variant 3) and to k_knn2 / k_knn2_mma.
//
// Arithmetic.  Bits become signs: query bit a -> (a ? +1 : -1), train bit b -> (b ? -1 : +1), so the product of a
// position is +1 where the bits differ and -1 where they agree and the sum over 256 positions is 2 Hamming - 256.
// e2m1 holds +-1 exactly, one descriptor word becomes one 16-byte chunk of nibbles with four logic ops, and the
// popcount terms of the i8 formulation are gone.  The scale of the bit blocks is 2^5, i.e. 64 Hamming - 8192.
// A fifth K step carries two more blocks of 32 elements with their own scales:
//     block X (scale 2^0)   column index within its 64-column group, as 1 d0 + 4 d1 + 4 (x + y), x + y = 4 d2
//     block Y (scale 2^16)  the constant 4 * 6 * 6 + 0.5 * 0.5 = 144.25 (and + 0.5 for rows past the end of the train set)
// so that the fp32 accumulator is 144 * 2^16 + 8192 + 64 Hamming + column, a number in [2^23, 2^24): its unit in
// the last place is 1, the low 16 bits of its bit pattern ARE the sort key (8192 + 64 Hamming + column < 2^16) and
// tcgen05.ld ... .pack::16b delivers two keys per register exactly as in the i8 kernel.  tools/probe/mxf4_probe.cu
// measured that the tensor core forms this sum exactly (no truncation across the 2^23 dynamic range), which scale
// bytes a_sf_id selects (0 -> bytes 0, 1; 2 -> bytes 2, 3 of the word) and that padding rows come out at 49152 + column.
// All scale factors are the same for every row, so the scale-factor regions of tensor memory are filled once with one
// word per side: A (132, 132, 127, 143) = 2^5, 2^5 for the bit steps, 2^0, 2^16 for the key step; B all 127.
//
// Shape.  As the i8 kernel: persistent CTAs of 416 threads, one per SM, work items = (descriptor set, 128-row query
// tile[, train split]); warps 4-7 expand (A 128 x 160 B double-buffered by item, B 240 x 160 B in VIORB_F4_STAGES stages),
// one thread of warp 8 issues 5 MMAs (M 128, N 240, K 64) per train tile into one of two accumulators, warps 0-3 and
// 9-12 keep the running top-2.  The train tile is 240 rows, not 256: two 240-column accumulators leave the 12 columns
// of tensor memory the scale factors need (columns 240-251; the accumulators start at 0 and 256).
// Against the i8 kernel per train row: 160 instead of 288 bytes of shared memory written and read, and 5 K-64 steps
// instead of 9 K-32 steps at the same MMA rate per step.
#pragma once
#include "kernels_match.cuh"

#ifndef VIORB_EMU
#ifndef TC5_DIAG
#define TC5_DIAG 0
#endif
#ifndef VIORB_F4_STAGES
#define VIORB_F4_STAGES 3
#endif
namespace tc5 {

constexpr int QT = 128;                 // query rows per work item (MMA M)
constexpr int TT = 240;                 // train rows per tile (MMA N)
constexpr int KCH = 10;                 // 16-byte K chunks per row: 8 chunks of sign nibbles + 2 of the key step
constexpr int LBO = 128;                // bytes between core matrices adjacent in K
constexpr int SBO = KCH * 128;          // bytes between 8-row groups
constexpr int A_BYTES = QT / 8 * SBO;   // 20 480
constexpr int B_BYTES = TT / 8 * SBO;   // 38 400
constexpr int NST = VIORB_F4_STAGES;    // train-tile stages
constexpr int NBAR = 8 + 2 * NST;
constexpr int XCH_OFF = 2 * A_BYTES + NST * B_BYTES + NBAR * 8 + 16;   // uint2[2][QT]: the upper column half's (best, second)
constexpr int SMEM_BYTES = XCH_OFF + 2 * QT * 8;
constexpr int THREADS = 416;
constexpr int ACC1 = 256;               // tensor-memory column of the second accumulator
constexpr int SFA_COL = 240, SFB_COL = 244;
constexpr uint32_t SFA_WORD = 132u | (132u << 8) | (127u << 16) | (143u << 24);
constexpr uint32_t SFB_WORD = 0x7f7f7f7fu;
constexpr unsigned KEY0 = 8192;         // key of Hamming 0, column 0
constexpr unsigned PAD_KEY = 49152;     // keys of padding rows start here
constexpr int DIST_BIAS = KEY0 >> 6;    // (key >> 6) - DIST_BIAS = Hamming
// e2m1 x e2m1 (format 1), ue8m0 scales, K-major A and B, N >> 3, M >> 4; bit 29.. = a_sf_id
constexpr uint32_t IDESC = (1u << 7) | (1u << 10) | ((uint32_t)(TT >> 3) << 17) | (1u << 23) | ((uint32_t)(QT >> 4) << 24);
constexpr uint32_t IDESC_KEY = IDESC | (2u << 29);

enum { BAR_A_FULL = 0, BAR_A_EMPTY = 2, BAR_D_FULL = 4, BAR_D_EMPTY = 6, BAR_B_FULL = 8, BAR_B_EMPTY = 8 + NST };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    unsigned long long t0 = 0;
    while (true) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        if ((++spins & 0xFFF) == 0) {       // a lost arrival must not hang the device: give up after 2 s
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > 2000000000ull) __trap();
        }
    }
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// shared-memory matrix descriptor: K-major, no swizzle, 8 x 16 B core matrices (version 1 = sm_100)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
    return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)(LBO >> 4) << 16) | ((uint64_t)(SBO >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void mma_f4(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t sfa, uint32_t sfb, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb) : "memory");
}
// 64 columns, the low halves of two adjacent columns packed into one register (column 2j in the low half)
__device__ __forceinline__ void tmem_ld64_pack16(uint32_t taddr, unsigned (&v)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
                   "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
                   "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
}
// 32 and 16 columns, packed the same way
__device__ __forceinline__ void tmem_ld32_pack16(uint32_t taddr, unsigned (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16_pack16(uint32_t taddr, unsigned (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// One descriptor word -> one 16-byte chunk of e2m1 nibbles: nibble j of output word s holds bit 4 j + 3 - s of w as the
// sign of 1.0 (0x2 / 0xA).  Query and train rows use the same position map, which is all the dot product needs.
// multipliers 1, 2, 4, 8 read from constant memory: ptxas cannot turn the multiply-add into a shift-add on the ALU pipe
__constant__ unsigned c_nibbleShift[4] = {1u, 2u, 4u, 8u};
__device__ __forceinline__ unsigned mad_u32(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// (the AND runs on the ALU pipe, shift + OR as one multiply-add on the FMA pipe: the ALU pipe is this kernel's busiest)
__device__ __forceinline__ uint4 sign_nibbles(unsigned w) {
    return make_uint4(mad_u32(w & 0x88888888u, c_nibbleShift[0], 0x22222222u), mad_u32(w & 0x44444444u, c_nibbleShift[1], 0x22222222u),
                      mad_u32(w & 0x22222222u, c_nibbleShift[2], 0x22222222u), mad_u32(w & 0x11111111u, c_nibbleShift[3], 0x22222222u));
}

// nibble codes of the column index of a train row inside its 64-column group (block X): 1 * d0 + 4 * d1 + 4 * x + 4 * y
__device__ __forceinline__ unsigned col_code(int col6) {
    const int d0 = col6 & 3, d1 = (col6 >> 2) & 3, d2 = col6 >> 4;
    const unsigned small = 0x5420u;                          // codes of 0, 1, 2, 3
    const unsigned xs = 0x7660u, ys = 0x7600u;               // x + y = 4 d2: (0,0) (4,0) (4,4) (6,6)
    return ((small >> (4 * d0)) & 15u) | (((small >> (4 * d1)) & 15u) << 4) | (((xs >> (4 * d2)) & 15u) << 8) | (((ys >> (4 * d2)) & 15u) << 12);
}

// Train row -> 10 chunks at `dst` (= tile base + row offset, chunk stride LBO): bit b -> (b ? -1 : +1)
__device__ __forceinline__ void expand_train_row(unsigned char* dst, const uint4& r0, const uint4& r1, bool valid, unsigned colCode) {
    const unsigned w[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    if (valid) {
#pragma unroll
        for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4*>(dst + i * LBO) = sign_nibbles(w[i]);
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4*>(dst + i * LBO) = make_uint4(0u, 0u, 0u, 0u);
    }
    *reinterpret_cast<uint4*>(dst + 8 * LBO) = make_uint4(colCode, 0u, 0u, 0u);
    // block Y: 6, 6, 6, 6, 0.5 and 1.0 for a padding row (against the query's 6, 6, 6, 6, 0.5, 0.5)
    *reinterpret_cast<uint4*>(dst + 9 * LBO) = make_uint4(0x7777u | (1u << 16) | (valid ? 0u : (2u << 20)), 0u, 0u, 0u);
}

// Query row -> 10 chunks: bit a -> (a ? +1 : -1), then the constant factors of the key step
__device__ __forceinline__ void expand_query_row(unsigned char* dst, const uint4& r0, const uint4& r1) {
    const unsigned w[8] = {~r0.x, ~r0.y, ~r0.z, ~r0.w, ~r1.x, ~r1.y, ~r1.z, ~r1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4*>(dst + i * LBO) = sign_nibbles(w[i]);
    *reinterpret_cast<uint4*>(dst + 8 * LBO) = make_uint4(0x6662u, 0u, 0u, 0u);                    // 1, 4, 4, 4
    *reinterpret_cast<uint4*>(dst + 9 * LBO) = make_uint4(0x7777u | (1u << 16) | (1u << 20), 0u, 0u, 0u);   // 6, 6, 6, 6, 0.5, 0.5
}

__device__ __forceinline__ void top2(unsigned& b0, unsigned& b1, unsigned k) {
    b1 = min(b1, max(b0, k));
    b0 = min(b0, k);
}

// (best, second) keys ((Hamming + popc(query)) << 22 | train index, 0xFFFFFFFF = none) -> idx[2], dist[2] of one query row
__device__ __forceinline__ void write_knn_row(int32_t* __restrict__ idx, int32_t* __restrict__ dist, size_t rowIdx, unsigned gb0,
                                              unsigned gb1, int pq) {
    const unsigned mask = (1u << VIORB_KNN_IDX_BITS) - 1;
    const bool h0 = gb0 != 0xFFFFFFFFu, h1 = gb1 != 0xFFFFFFFFu;
    store_knn_row(idx, dist, rowIdx, h0 ? (int)(gb0 & mask) : -1, h1 ? (int)(gb1 & mask) : -1,
                  h0 ? (int)(gb0 >> VIORB_KNN_IDX_BITS) - pq : -1, h1 ? (int)(gb1 >> VIORB_KNN_IDX_BITS) - pq : -1);
}

// Merge of the per-split partial pairs when a query tile's train range was cut over several CTAs (few query tiles,
// many train rows: a single stereo pair, the row-sharded sweep).  grid (ceil(qTiles * 128 / 256), nSets), block 256.
__global__ void __launch_bounds__(256) k_knn2_tc5_merge(DescSet Q, const uint2* __restrict__ partial, int32_t* __restrict__ idx,
                                                        int32_t* __restrict__ dist, size_t outStride, int qTiles, int nSplit) {
    const int p = blockIdx.y, qi = blockIdx.x * 256 + threadIdx.x;
    if (qi >= descset_count(Q, p)) return;
    unsigned gb0 = 0xFFFFFFFFu, gb1 = 0xFFFFFFFFu;
    for (int j = 0; j < nSplit; ++j) {
        const uint2 o = partial[((size_t)p * nSplit + j) * ((size_t)qTiles * QT) + qi];
        top2(gb0, gb1, o.x);
        top2(gb0, gb1, o.y);
    }
    write_knn_row(idx, dist, (size_t)p * outStride + qi, gb0, gb1, DIST_BIAS);
}

// The two smallest of the 2 * NR packed 16-bit keys in v[] folded into the running pair (gb0, gb1) of this query row.
// Pass 1 takes the minimum with 3-input min (NR / 2 instructions).  Train rows arrive in index order, so a key can enter
// the pair only with a distance strictly below the current second's: if the group's minimum fails that test - the
// common case once a few tiles have passed - the group is done.  Otherwise pass 2 takes the runner-up as the minimum of
// (key - (best + 1)) in wrapping 16-bit arithmetic, one fused add-min per register: keys are unique inside a group (they
// carry the column), so exactly the best key wraps to 0xFFFF and every other key keeps its order.
template <int NR>
__device__ __forceinline__ void reduce_keys(const unsigned (&v)[NR], unsigned base, unsigned& gb0, unsigned& gb1) {
#if TC5_DIAG == 1
    unsigned x = v[0];
#pragma unroll
    for (int j = 1; j < NR; j += 8) x &= v[j];
    if (x == 0x12345u) top2(gb0, gb1, base);
    return;
#endif
    static_assert(NR % 8 == 0, "four chains of an even number of registers");
    constexpr int C = NR / 4;
    unsigned mc[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        unsigned mm = v[c * C];
#pragma unroll
        for (int j = 1; j + 1 < C; j += 2) mm = __vimin3_u16x2(mm, v[c * C + j], v[c * C + j + 1]);
        mc[c] = __vminu2(mm, v[c * C + C - 1]);
    }
    const unsigned mm = __vminu2(__vimin3_u16x2(mc[0], mc[1], mc[2]), mc[3]);
    const unsigned x0 = min(mm & 0xFFFFu, mm >> 16);
    if ((x0 >> 6) >= min(gb1 >> VIORB_KNN_IDX_BITS, PAD_KEY >> 6)) return;      // cannot enter the pair / only padding rows
    const unsigned nk = __vsub2(0u, (x0 + 1u) * 0x10001u);
    unsigned s0 = 0xFFFFFFFFu, s1 = 0xFFFFFFFFu, s2 = 0xFFFFFFFFu, s3 = 0xFFFFFFFFu;
#pragma unroll
    for (int j = 0; j < NR; j += 4) {
        s0 = __viaddmin_u16x2(v[j], nk, s0);
        s1 = __viaddmin_u16x2(v[j + 1], nk, s1);
        s2 = __viaddmin_u16x2(v[j + 2], nk, s2);
        s3 = __viaddmin_u16x2(v[j + 3], nk, s3);
    }
    const unsigned ss = __vminu2(__vimin3_u16x2(s0, s1, s2), s3);
    const unsigned x1 = x0 + 1u + min(ss & 0xFFFFu, ss >> 16);
    top2(gb0, gb1, ((x0 >> 6) << VIORB_KNN_IDX_BITS) | (base + (x0 & 63u)));
    if (x1 < PAD_KEY) top2(gb0, gb1, ((x1 >> 6) << VIORB_KNN_IDX_BITS) | (base + (x1 & 63u)));
}

__global__ void __launch_bounds__(THREADS, 1) k_knn2_f4(DescSet Q, DescSet T, int32_t* __restrict__ idx, int32_t* __restrict__ dist,
                                                         size_t outStride, int qTiles, int nSets, int nItems, int nSplit, uint2* __restrict__ partial) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* sA = smem;
    unsigned char* sB = smem + 2 * A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * A_BYTES + NST * B_BYTES);
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(bars + NBAR);
    const uint32_t bar0 = smem_u32(bars);
    const int tid = threadIdx.x, warp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar0 + 8 * (BAR_A_FULL + s), 128);
            mbar_init(bar0 + 8 * (BAR_A_EMPTY + s), 1);
            mbar_init(bar0 + 8 * (BAR_D_FULL + s), 1);
            mbar_init(bar0 + 8 * (BAR_D_EMPTY + s), 256);
        }
        for (int s = 0; s < NST; ++s) {
            mbar_init(bar0 + 8 * (BAR_B_FULL + s), 128);
            mbar_init(bar0 + 8 * (BAR_B_EMPTY + s), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) {       // the whole tensor memory of this SM: two 128 x 240 fp32 accumulators + the scale factors
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmemSlot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmemSlot;
    if (warp < 4) {        // scale factors: one word per side in every lane of the 4 (A) and 8 (B) columns
        const uint32_t lane0 = tmem + ((uint32_t)(warp * 32) << 16);
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%1,%1,%1};" ::"r"(lane0 + SFA_COL), "r"(SFA_WORD) : "memory");
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(lane0 + SFB_COL), "r"(SFB_WORD) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (warp >= 4 && warp < 8) {
        // ------------------------------------------------------------------ producers
        const int pt = tid - 128;
        const int rowOff = (pt >> 3) * SBO + (pt & 7) * 16;
        const unsigned colCode = col_code(pt & 63);          // rows pt and pt + 128 share the column code
        uint32_t g = 0, it = 0;
        for (int item = blockIdx.x; item < nItems; item += gridDim.x) {
            const int js = item % nSplit, rest = item / nSplit, qt = rest / nSets, p = rest - qt * nSets, q0 = qt * QT;
            const int nq = descset_count(Q, p), nt = descset_count(T, p);
            if (q0 >= nq) continue;
            const uint4* qb = reinterpret_cast<const uint4*>(Q.base + (size_t)p * Q.stride);
            const uint4* tb = reinterpret_cast<const uint4*>(T.base + (size_t)p * T.stride);
            const int nTilesAll = (nt + TT - 1) / TT, tBeg = js * nTilesAll / nSplit, tEnd = (js + 1) * nTilesAll / nSplit;
            uint4 n0[2], n1[2];
            auto fetch = [&](int t) {
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    const int row = min(t * TT + pt + 128 * r, nt - 1);
                    n0[r] = __ldg(tb + (size_t)row * 2);
                    n1[r] = __ldg(tb + (size_t)row * 2 + 1);
                }
            };
            if (tBeg < tEnd) fetch(tBeg);
            {
                const int row = min(q0 + pt, nq - 1);
                const uint4 r0 = __ldg(qb + (size_t)row * 2), r1 = __ldg(qb + (size_t)row * 2 + 1);
                // the query tile is double-buffered by item parity: it is expanded while the previous item's MMAs still run
                mbar_wait(bar0 + 8 * (BAR_A_EMPTY + (it & 1)), ((it >> 1) & 1) ^ 1);
                expand_query_row(sA + (it & 1) * A_BYTES + rowOff, r0, r1);
                fence_async_smem();
                mbar_arrive(bar0 + 8 * (BAR_A_FULL + (it & 1)));
            }
            for (int t = tBeg; t < tEnd; ++t, ++g) {
                const uint32_t s = g % NST, ph = (g / NST) & 1;
                const uint4 c0[2] = {n0[0], n0[1]}, c1[2] = {n1[0], n1[1]};
                if (t + 1 < tEnd) fetch(t + 1);
                mbar_wait(bar0 + 8 * (BAR_B_EMPTY + s), ph ^ 1);
                unsigned char* dstB = sB + s * B_BYTES + rowOff;
#if TC5_DIAG != 3
                expand_train_row(dstB, c0[0], c1[0], t * TT + pt < nt, colCode);
                if (pt + 128 < TT) expand_train_row(dstB + 16 * SBO, c0[1], c1[1], t * TT + pt + 128 < nt, colCode);
#else
                (void)dstB; (void)c0; (void)c1;
#endif
                fence_async_smem();
                mbar_arrive(bar0 + 8 * (BAR_B_FULL + s));
            }
            ++it;
        }
    } else if (warp == 8) {
        // ------------------------------------------------------------------ MMA issue (one thread)
        if ((tid & 31) == 0) {
            const uint64_t descA0 = smem_desc(smem_u32(sA));
            const uint64_t descB0 = smem_desc(smem_u32(sB));
            const uint32_t sfa = tmem + SFA_COL, sfb = tmem + SFB_COL;
            uint32_t g = 0, it = 0;
            for (int item = blockIdx.x; item < nItems; item += gridDim.x) {
                const int js = item % nSplit, rest = item / nSplit, qt = rest / nSets, p = rest - qt * nSets, q0 = qt * QT;
                const int nq = descset_count(Q, p), nt = descset_count(T, p);
                if (q0 >= nq) continue;
                const int nTilesAll = (nt + TT - 1) / TT, tBeg = js * nTilesAll / nSplit, tEnd = (js + 1) * nTilesAll / nSplit;
                mbar_wait(bar0 + 8 * (BAR_A_FULL + (it & 1)), (it >> 1) & 1);
                const uint64_t descA = descA0 + (uint64_t)(((it & 1) * A_BYTES) >> 4);
                for (int t = tBeg; t < tEnd; ++t, ++g) {
                    const uint32_t s = g % NST, d = g & 1;
                    mbar_wait(bar0 + 8 * (BAR_B_FULL + s), (g / NST) & 1);
                    mbar_wait(bar0 + 8 * (BAR_D_EMPTY + d), ((g >> 1) & 1) ^ 1);
                    tc_fence_after();
                    const uint64_t descB = descB0 + (uint64_t)((s * B_BYTES) >> 4);
#pragma unroll
                    for (int k = 0; k < KCH / 2; ++k)
                        mma_f4(tmem + d * ACC1, descA + (uint64_t)((k * 2 * LBO) >> 4), descB + (uint64_t)((k * 2 * LBO) >> 4),
                               k == KCH / 2 - 1 ? IDESC_KEY : IDESC, sfa, sfb, k > 0);
                    tc_commit(bar0 + 8 * (BAR_B_EMPTY + s));
                    tc_commit(bar0 + 8 * (BAR_D_FULL + d));
                }
                tc_commit(bar0 + 8 * (BAR_A_EMPTY + (it & 1)));
                ++it;
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------------ epilogue: thread = query row = TMEM lane
        // warps 0-3 reduce columns 0-127 of every tile, warps 9-12 columns 128-239 (a warp reaches the 32 lanes
        // of its SM sub-partition, warp % 4); the two halves of a row meet in shared memory at the end of the item
        const int half = warp >= 9, quarter = warp & 3, row = quarter * 32 + (tid & 31);
        const uint32_t laneBase = tmem + ((uint32_t)(quarter * 32) << 16) + half * 128;
        uint32_t g = 0, it = 0;
        for (int item = blockIdx.x; item < nItems; item += gridDim.x) {
            const int js = item % nSplit, rest = item / nSplit, qt = rest / nSets, p = rest - qt * nSets, q0 = qt * QT;
            const int nq = descset_count(Q, p), nt = descset_count(T, p);
            if (q0 >= nq) continue;
            const int nTilesAll = (nt + TT - 1) / TT, tBeg = js * nTilesAll / nSplit, tEnd = (js + 1) * nTilesAll / nSplit;
            const int qi = q0 + row;
            unsigned gb0 = 0xFFFFFFFFu, gb1 = 0xFFFFFFFFu;
            for (int t = tBeg; t < tEnd; ++t, ++g) {
                const uint32_t d = g & 1;
                mbar_wait(bar0 + 8 * (BAR_D_FULL + d), (g >> 1) & 1);
                tc_fence_after();
                // packed 16-bit keys: two columns per min / max; the load of the next columns is in flight while the
                // first 64 are reduced
                const uint32_t ta = laneBase + d * ACC1;
                const unsigned base = (unsigned)(t * TT + half * 128);
#if TC5_DIAG >= 2
                tc_fence_before();
                mbar_arrive(bar0 + 8 * (BAR_D_EMPTY + d));
                (void)ta; (void)base;
#else
                unsigned va[32];
                tmem_ld64_pack16(ta, va);
                tmem_ld_wait();
                if (!half) {
                    unsigned vb[32];
                    tmem_ld64_pack16(ta + 64, vb);
                    reduce_keys<32>(va, base, gb0, gb1);
                    tmem_ld_wait();
                    tc_fence_before();
                    mbar_arrive(bar0 + 8 * (BAR_D_EMPTY + d));
                    reduce_keys<32>(vb, base + 64, gb0, gb1);
                } else {                                  // columns 192-239: 32 + 16
                    unsigned vb[16], vc[8];
                    tmem_ld32_pack16(ta + 64, vb);
                    tmem_ld16_pack16(ta + 96, vc);
                    reduce_keys<32>(va, base, gb0, gb1);
                    tmem_ld_wait();
                    tc_fence_before();
                    mbar_arrive(bar0 + 8 * (BAR_D_EMPTY + d));
                    // one group of 48 columns
                    unsigned vd[24];
#pragma unroll
                    for (int j = 0; j < 16; ++j) vd[j] = vb[j];
#pragma unroll
                    for (int j = 0; j < 8; ++j) vd[16 + j] = vc[j];
                    reduce_keys<24>(vd, base + 64, gb0, gb1);
                }
#endif
            }
            // the upper half hands its pair to the lower half (double-buffered by item parity; the named barrier
            // of the next item orders the reuse)
            uint2* xch = reinterpret_cast<uint2*>(smem + XCH_OFF) + (it & 1) * QT;
            if (half) xch[row] = make_uint2(gb0, gb1);
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (!half && qi < nq) {
                const uint2 o = xch[row];
                top2(gb0, gb1, o.x);
                top2(gb0, gb1, o.y);
                if (nSplit > 1) {
                    // the train tiles of this query tile are split over nSplit CTAs: raw keys go to the scratch rows
                    // [set][split][query], k_knn2_tc5_merge finishes
                    partial[((size_t)p * nSplit + js) * ((size_t)qTiles * QT) + qi] = make_uint2(gb0, gb1);
                } else {
                    write_knn_row(idx, dist, (size_t)p * outStride + qi, gb0, gb1, DIST_BIAS);
                }
            }
            ++it;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}

}  // namespace tc5
#endif  // !VIORB_EMU
