// kernels_knn_tc5.cuh — brute-force Hamming kNN-2 on the 5th-generation tensor cores (tcgen05.mma kind::i8,
// accumulators in tensor memory).  Same contract as k_knn2 / k_knn2_mma in kernels_match.cuh:
// cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, 2) (frame.cpp:1167, fmatcher.cpp:214), ties -> lower train index.
//
// Arithmetic.  For 256-bit rows a (query) and b (train): Hamming(a, b) = 2 popc(a & ~b) + popc(b) - popc(a).
// popc(a & ~b) is an integer GEMM over expanded bits.  Train bits are expanded with ONE logic op per four
// output bytes: word w of ~b and mask 0x01010101 << j give four bytes that are 0 or 2^j; the query side carries
// 0 or 2^(7-j) at the same K positions, so every matched bit contributes exactly 128.  A ninth K block carries,
// per train row, (popc(b) << 6) + (column & 63) (bytes [col, min(pt,255), pt-min(pt,255)] against the constant
// query bytes [1, 64, 64]), so the s32 accumulator IS the sort key of its 64-column group:
//     acc = 128 popc(a & ~b) + 64 popc(b) + col  =  ((Hamming + popc(a)) << 6) | col
// and the epilogue is nothing but the running top-2 (min / max / min per pair).  Rows past the end of the train
// set carry 255 x 255 in a fourth byte and lose against every real key.
//
// Shape.  Persistent CTAs of 416 threads, one per SM, work items = (descriptor set, 128-row query tile), sets
// fastest, so that the empty trailing tiles of ragged sets (rows past the set's count) all sit at the end of the list:
//   warps 4-7  producers: expand the query tile (A, 128 x 288 B, double-buffered by item) and the train tiles (B, 256 x 288 B, two stages)
//              into the K-major no-swizzle core-matrix layout, fence.proxy.async, arrive on the full barriers
//   warp 8     one thread issues 9 x tcgen05.mma.cta_group::1.kind::i8 (M 128, N 256, K 32) per train tile into one
//              of two 256-column accumulators and commits to the barriers that free the stage / publish the tile
//   warps 0-3, 9-12  epilogue: thread = TMEM lane = query row, each warp set takes one 128-column half of the
//              tile.  Every key is below 2^16, so tcgen05.ld ... .pack::16b delivers two columns per register
//              and the running top-2 of 64 columns is 16 x 3 VIMNMX.U16x2 on two independent chains; the two
//              survivors become (distance' << 22 | train index) keys; idx / dist are written at the end
// Results are bit-identical to k_knn2 / k_knn2_mma.  grid = min(items, SMs), block 416, dynamic smem 222 KB.
// Measured bound (nold / nocomp variants, tools/gpu_tc5_variants.sh): shared-memory bandwidth - per train tile
// the producers store 72 KB and the tensor core reads 108 KB (A is re-read for every tile) at 128 B / clk.
#pragma once
#include "kernels_match.cuh"

#ifndef VIORB_EMU
#ifndef TC5_DIAG
#define TC5_DIAG 0
#endif
namespace tc5 {

constexpr int QT = 128;                 // query rows per work item (MMA M)
constexpr int TT = 256;                 // train rows per tile (MMA N)
constexpr int KCH = 18;                 // 16-byte K chunks per row: 16 expanded-bit chunks + 2 of the key block
constexpr int LBO = 128;                // bytes between core matrices adjacent in K
constexpr int SBO = KCH * 128;          // bytes between 8-row groups
constexpr int A_BYTES = QT / 8 * SBO;   // 36 864
constexpr int B_BYTES = TT / 8 * SBO;   // 73 728
constexpr int NBAR = 12;
constexpr int XCH_OFF = 2 * A_BYTES + 2 * B_BYTES + NBAR * 8 + 16;   // uint2[2][QT]: the upper column half's (best, second)
constexpr int SMEM_BYTES = XCH_OFF + 2 * QT * 8;
constexpr int THREADS = 416;
constexpr int PAD_KEY = 255 * 255;      // key contribution of a padding train row
constexpr uint32_t IDESC = (2u << 4) | ((uint32_t)(TT >> 3) << 17) | ((uint32_t)(QT >> 4) << 24);   // s32 += u8 x u8, K-major A and B

enum { BAR_A_FULL = 0, BAR_A_EMPTY = 2, BAR_B_FULL = 4, BAR_B_EMPTY = 6, BAR_D_FULL = 8, BAR_D_EMPTY = 10 };   // two of each

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
__device__ __forceinline__ void mma_i8(uint32_t d, uint64_t a, uint64_t b, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(IDESC), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, int (&v)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
                   "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
                   "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
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
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ int popc256(const uint4& a, const uint4& b) {
    return __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w) + __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
}

// Train row -> 18 chunks at `dst` (= tile base + row offset, chunk stride LBO).  col6 = row & 63.
__device__ __forceinline__ void expand_train_row(unsigned char* dst, const uint4& r0, const uint4& r1, bool valid, int col6) {
    const unsigned w[8] = {~r0.x, ~r0.y, ~r0.z, ~r0.w, ~r1.x, ~r1.y, ~r1.z, ~r1.w};
    if (valid) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            *reinterpret_cast<uint4*>(dst + (2 * i) * LBO) =
                make_uint4(w[i] & 0x01010101u, w[i] & 0x02020202u, w[i] & 0x04040404u, w[i] & 0x08080808u);
            *reinterpret_cast<uint4*>(dst + (2 * i + 1) * LBO) =
                make_uint4(w[i] & 0x10101010u, w[i] & 0x20202020u, w[i] & 0x40404040u, w[i] & 0x80808080u);
        }
        const int pt = popc256(r0, r1), lo = min(pt, 255);
        *reinterpret_cast<uint4*>(dst + 16 * LBO) = make_uint4((unsigned)col6 | ((unsigned)lo << 8) | ((unsigned)(pt - lo) << 16), 0u, 0u, 0u);
    } else {
#pragma unroll
        for (int c = 0; c < 16; ++c) *reinterpret_cast<uint4*>(dst + c * LBO) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(dst + 16 * LBO) = make_uint4(255u << 24, 0u, 0u, 0u);
    }
    *reinterpret_cast<uint4*>(dst + 17 * LBO) = make_uint4(0u, 0u, 0u, 0u);
}

// Query row -> 18 chunks: bit j of byte b of word i becomes the byte 2^(7-j) at K position 32 i + 4 j + b.
__device__ __forceinline__ void expand_query_row(unsigned char* dst, const uint4& r0, const uint4& r1) {
    const unsigned w[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        *reinterpret_cast<uint4*>(dst + (2 * i) * LBO) =
            make_uint4((w[i] & 0x01010101u) << 7, (w[i] & 0x02020202u) << 5, (w[i] & 0x04040404u) << 3, (w[i] & 0x08080808u) << 1);
        *reinterpret_cast<uint4*>(dst + (2 * i + 1) * LBO) =
            make_uint4((w[i] & 0x10101010u) >> 1, (w[i] & 0x20202020u) >> 3, (w[i] & 0x40404040u) >> 5, (w[i] & 0x80808080u) >> 7);
    }
    *reinterpret_cast<uint4*>(dst + 16 * LBO) = make_uint4(1u | (64u << 8) | (64u << 16) | (255u << 24), 0u, 0u, 0u);
    *reinterpret_cast<uint4*>(dst + 17 * LBO) = make_uint4(0u, 0u, 0u, 0u);
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
    const uint4* qb = reinterpret_cast<const uint4*>(Q.base + (size_t)p * Q.stride);
    const int pq = popc256(__ldg(qb + (size_t)qi * 2), __ldg(qb + (size_t)qi * 2 + 1));
    write_knn_row(idx, dist, (size_t)p * outStride + qi, gb0, gb1, pq);
}

__global__ void __launch_bounds__(THREADS, 1) k_knn2_tc5(DescSet Q, DescSet T, int32_t* __restrict__ idx, int32_t* __restrict__ dist,
                                                          size_t outStride, int qTiles, int nSets, int nItems, int nSplit, uint2* __restrict__ partial) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* sA = smem;
    unsigned char* sB = smem + 2 * A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * A_BYTES + 2 * B_BYTES);
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(bars + NBAR);
    const uint32_t bar0 = smem_u32(bars);
    const int tid = threadIdx.x, warp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar0 + 8 * (BAR_A_FULL + s), 128);
            mbar_init(bar0 + 8 * (BAR_A_EMPTY + s), 1);
            mbar_init(bar0 + 8 * (BAR_B_FULL + s), 128);
            mbar_init(bar0 + 8 * (BAR_B_EMPTY + s), 1);
            mbar_init(bar0 + 8 * (BAR_D_FULL + s), 1);
            mbar_init(bar0 + 8 * (BAR_D_EMPTY + s), 256);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) {       // the whole 512-column tensor memory of this SM: two 128 x 256 s32 accumulators
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmemSlot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmemSlot;

    if (warp >= 4 && warp < 8) {
        // ------------------------------------------------------------------ producers
        const int pt = tid - 128;
        const int rowOff = (pt >> 3) * SBO + (pt & 7) * 16;
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
                const uint32_t s = g & 1;
                const uint4 c0[2] = {n0[0], n0[1]}, c1[2] = {n1[0], n1[1]};
                if (t + 1 < tEnd) fetch(t + 1);
                mbar_wait(bar0 + 8 * (BAR_B_EMPTY + s), ((g >> 1) & 1) ^ 1);
                unsigned char* dstB = sB + s * B_BYTES + rowOff;
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    const int n = pt + 128 * r;
                    expand_train_row(dstB + r * 16 * SBO, c0[r], c1[r], t * TT + n < nt, n & 63);
                }
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
            uint32_t g = 0, it = 0;
            for (int item = blockIdx.x; item < nItems; item += gridDim.x) {
                const int js = item % nSplit, rest = item / nSplit, qt = rest / nSets, p = rest - qt * nSets, q0 = qt * QT;
                const int nq = descset_count(Q, p), nt = descset_count(T, p);
                if (q0 >= nq) continue;
                const int nTilesAll = (nt + TT - 1) / TT, tBeg = js * nTilesAll / nSplit, tEnd = (js + 1) * nTilesAll / nSplit;
                mbar_wait(bar0 + 8 * (BAR_A_FULL + (it & 1)), (it >> 1) & 1);
                const uint64_t descA = descA0 + (uint64_t)(((it & 1) * A_BYTES) >> 4);
                for (int t = tBeg; t < tEnd; ++t, ++g) {
                    const uint32_t s = g & 1, ph = (g >> 1) & 1;
                    mbar_wait(bar0 + 8 * (BAR_B_FULL + s), ph);
                    mbar_wait(bar0 + 8 * (BAR_D_EMPTY + s), ph ^ 1);
                    tc_fence_after();
                    const uint64_t descB = descB0 + (uint64_t)((s * B_BYTES) >> 4);
#pragma unroll
                    for (int k = 0; k < KCH / 2; ++k)
                        mma_i8(tmem + s * TT, descA + (uint64_t)((k * 2 * LBO) >> 4), descB + (uint64_t)((k * 2 * LBO) >> 4), k > 0);
                    tc_commit(bar0 + 8 * (BAR_B_EMPTY + s));
                    tc_commit(bar0 + 8 * (BAR_D_FULL + s));
                }
                tc_commit(bar0 + 8 * (BAR_A_EMPTY + (it & 1)));
                ++it;
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------------ epilogue: thread = query row = TMEM lane
        // warps 0-3 reduce columns 0-127 of every tile, warps 9-12 columns 128-255 (a warp reaches the 32 lanes
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
                const uint32_t s = g & 1;
                mbar_wait(bar0 + 8 * (BAR_D_FULL + s), (g >> 1) & 1);
                tc_fence_after();
                // packed 16-bit keys (every key is below 2^16): two columns per min / max, four chains per 64 columns;
                // the load of the second 64 columns is in flight while the first 64 are reduced
                unsigned va[32], vb[32];
                auto reduce64 = [&](const unsigned (&v)[32], int grp) {
#if TC5_DIAG != 1
                    unsigned a0 = 0xFFFFFFFFu, a1 = 0xFFFFFFFFu, c0 = 0xFFFFFFFFu, c1 = 0xFFFFFFFFu;
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        a1 = __vminu2(a1, __vmaxu2(a0, v[j]));
                        a0 = __vminu2(a0, v[j]);
                        c1 = __vminu2(c1, __vmaxu2(c0, v[16 + j]));
                        c0 = __vminu2(c0, v[16 + j]);
                    }
                    // four chains -> two (packed), even / odd columns -> the two smallest keys of the group
                    const unsigned m1 = __vminu2(__vminu2(a1, c1), __vmaxu2(a0, c0)), m0 = __vminu2(a0, c0);
                    const unsigned l0 = m0 & 0xFFFFu, h0 = m0 >> 16, l1 = m1 & 0xFFFFu, h1 = m1 >> 16;
                    const unsigned x0 = min(l0, h0), x1 = min(max(l0, h0), min(l1, h1));
                    const unsigned base = (unsigned)(t * TT + half * 128 + grp * 64);
                    if (x0 < (unsigned)PAD_KEY) top2(gb0, gb1, ((x0 >> 6) << VIORB_KNN_IDX_BITS) | (base + (x0 & 63u)));
                    if (x1 < (unsigned)PAD_KEY) top2(gb0, gb1, ((x1 >> 6) << VIORB_KNN_IDX_BITS) | (base + (x1 & 63u)));
#endif
                };
                const uint32_t ta = laneBase + s * TT;
#if TC5_DIAG != 2
                tmem_ld64_pack16(ta, va);
                tmem_ld_wait();
                tmem_ld64_pack16(ta + 64, vb);
                reduce64(va, 0);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(bar0 + 8 * (BAR_D_EMPTY + s));
                reduce64(vb, 1);
#else
                tc_fence_before();
                mbar_arrive(bar0 + 8 * (BAR_D_EMPTY + s));
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
                    const uint4* qb = reinterpret_cast<const uint4*>(Q.base + (size_t)p * Q.stride);
                    const int pq = popc256(__ldg(qb + (size_t)qi * 2), __ldg(qb + (size_t)qi * 2 + 1));
                    write_knn_row(idx, dist, (size_t)p * outStride + qi, gb0, gb1, pq);
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
