// kernels_window.cuh — sm_100a kernels of the windowed (grid-accelerated) frame-to-frame matchers.
//
//   k_grid_build        Frame::AssignFeaturesToGrid / PosInGrid      frame.cpp:386-414, 746-757
//   k_window_src_*      the per-keypoint search window of            fmatcher.cpp:997-1003 / :2519-2530
//   k_window_list       Frame::GetFeaturesInArea + DescriptorDistance frame.cpp:678-744, fmatcher.cpp:2859
//   k_resolve_ini       FMatcher::SearchForInitialization            fmatcher.cpp:983-1098
//   k_resolve_proj      FMatcher::SearchByProjection(Cur, Last)      fmatcher.cpp:2471-2688 (Nleft == -1)
//   (both)              FMatcher::ComputeThreeMaxima                 fmatcher.cpp:2813-2854
//
// The reference loops are sequential in the source keypoint: whether a candidate counts depends on what
// earlier keypoints matched (vMatchedDistance / mvpMapPoints).  The expensive, order-independent part - the
// window query and the Hamming distances - is done by one warp per source keypoint into a CSR candidate list
// that keeps GetFeaturesInArea's order; the order-dependent bookkeeping then runs in ONE warp per frame pair
// that walks the source keypoints in order with the candidate scan spread over its lanes.
#pragma once
#include "kernels_match.cuh"

#define VIORB_GRID_COLS 64            // FRAME_GRID_COLS (frame.h:43)
#define VIORB_GRID_ROWS 48            // FRAME_GRID_ROWS (frame.h:42)
#define VIORB_GRID_CELLS (VIORB_GRID_COLS * VIORB_GRID_ROWS)
#define VIORB_HISTO_LENGTH 30         // fmatcher.cpp:315

struct GridBounds { float minX, maxX, minY, maxY, invW, invH; };

// One search window: centre, half size, octave range (GetFeaturesInArea's minLevel / maxLevel), and for the
// projection matcher the predicted right-image coordinate.  r < 0: this source keypoint is skipped.
struct WinSrc {
    float x, y, r, ur;
    int minLevel, maxLevel;
};

// A batch of frame pairs: pair p uses rows [p*cap1, p*cap1 + n1[p]) of the source arrays and
// [p*cap2, p*cap2 + n2[p]) of the target arrays.
struct WinBatch {
    int cap1, cap2;
    const int* n1;
    const int* n2;
    const int* n1fill;      // optional: source count of the fill pass (0 for a pair whose list did not fit the budget)
    const int* off2;        // optional: first target row of the pair inside its cap2 rows (the right camera of a two-camera
                            // frame: keypointsRight_ follow keypoints_, offset Nleft; indices stay relative to it, frame.cpp:409)
};
__device__ __forceinline__ size_t winbatch_base2(const WinBatch& wb, int p) { return (size_t)p * wb.cap2 + (wb.off2 ? wb.off2[p] : 0); }

// ---------------------------------------------------------------------------------------------
// Grid: CSR table over the 64x48 cells, cell id = ix * 48 + iy (so the cells of one column are
// contiguous, which is GetFeaturesInArea's inner loop), entries of a cell in ascending keypoint index
// (= push_back order).  grid (nPairs), block 256.
// ---------------------------------------------------------------------------------------------
// cellKp[j] = (x, y, octave, index) of the keypoint at position j of the cell-ordered list: a window query then reads
// the entries of a grid column as ONE contiguous run of 16-byte records instead of gathering 28-byte keypoints.
__global__ void __launch_bounds__(256) k_grid_build(WinBatch wb, GridBounds gb, const viorb_keypoint* __restrict__ kps2,
                                                    int* __restrict__ cellStart, int* __restrict__ cellList,
                                                    float4* __restrict__ cellKp) {
    __shared__ int cnt[VIORB_GRID_CELLS + 1];
    __shared__ int cur[VIORB_GRID_CELLS];
    __shared__ int warpSums[36];
    const int tid = threadIdx.x, p = blockIdx.x;
    const int n2 = wb.n2[p];
    const viorb_keypoint* K = kps2 + winbatch_base2(wb, p);
    int* CS = cellStart + (size_t)p * (VIORB_GRID_CELLS + 1);
    int* CL = cellList + (size_t)p * wb.cap2;
    for (int i = tid; i <= VIORB_GRID_CELLS; i += 256) cnt[i] = 0;
    for (int i = tid; i < VIORB_GRID_CELLS; i += 256) cur[i] = 0;
    __syncthreads();
    auto cell_of = [&](const viorb_keypoint& k) {
        const int px = (int)roundf(__fmul_rn(__fsub_rn(k.x, gb.minX), gb.invW));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(k.y, gb.minY), gb.invH));
        if (px < 0 || px >= VIORB_GRID_COLS || py < 0 || py >= VIORB_GRID_ROWS) return -1;
        return px * VIORB_GRID_ROWS + py;
    };
    for (int i = tid; i < n2; i += 256) {
        const int c = cell_of(K[i]);
        if (c >= 0) atomicAdd(&cnt[c], 1);
    }
    __syncthreads();
    block_excl_scan(cnt, VIORB_GRID_CELLS + 1, warpSums);
    for (int i = tid; i <= VIORB_GRID_CELLS; i += 256) CS[i] = cnt[i];
    for (int i = tid; i < n2; i += 256) {
        const int c = cell_of(K[i]);
        if (c >= 0) CL[cnt[c] + atomicAdd(&cur[c], 1)] = i;
    }
    __syncthreads();
    // cells hold a handful of entries: insertion sort per cell restores ascending index order
    for (int c = tid; c < VIORB_GRID_CELLS; c += 256) {
        const int s = cnt[c], e = cnt[c + 1];
        for (int a = s + 1; a < e; ++a) {
            const int v = CL[a];
            int b = a - 1;
            while (b >= s && CL[b] > v) { CL[b + 1] = CL[b]; --b; }
            CL[b + 1] = v;
        }
    }
    __syncthreads();
    float4* CK = cellKp + (size_t)p * wb.cap2;
    const int total = cnt[VIORB_GRID_CELLS];
    for (int j = tid; j < total; j += 256) {
        const int i2 = CL[j];
        const viorb_keypoint k = K[i2];
        CK[j] = make_float4(k.x, k.y, __int_as_float(k.octave), __int_as_float(i2));
    }
}

// SearchForInitialization: level-0 keypoints of F1 search a square of half size windowSize around
// vbPrevMatched[i1], level 0 only (fmatcher.cpp:997-1003).  grid (ceil(cap1/256), nPairs).
// initPrev != 0: vbPrevMatched starts as F1's own keypoint positions, as Tracking::MonocularInitialization sets it
// (src/core/tracking.cpp:2286-2288), and is written to prevMatched first.
__global__ void k_window_src_ini(WinBatch wb, const viorb_keypoint* __restrict__ kps1, float* __restrict__ prevMatched,
                                 float windowSize, WinSrc* __restrict__ src, int initPrev) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
    if (i >= wb.n1[p]) return;
    const size_t o = (size_t)p * wb.cap1 + i;
    WinSrc s;
    if (initPrev) { prevMatched[2 * o] = kps1[o].x; prevMatched[2 * o + 1] = kps1[o].y; }
    s.x = prevMatched[2 * o];
    s.y = prevMatched[2 * o + 1];
    s.r = kps1[o].octave > 0 ? -1.f : windowSize;
    s.ur = 0.f;
    s.minLevel = 0;
    s.maxLevel = 0;
    src[o] = s;
}

// SearchByProjection: radius = th * mvScaleFactors[nLastOctave]; octave window by direction (fmatcher.cpp:2519-2530).
struct ProjParams {
    float th, bf;
    int direction, nlevels;            // 0: nLastOctave-1..+1, 1: forward (>= nLastOctave), 2: backward (<= nLastOctave)
    float scale[VIORB_MAX_LEVELS];
};
__global__ void k_window_src_proj(WinBatch wb, ProjParams pp, const float* __restrict__ uv, const float* __restrict__ invz,
                                  const int32_t* __restrict__ lastOctave, const uint8_t* __restrict__ valid,
                                  WinSrc* __restrict__ src) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
    if (i >= wb.n1[p]) return;
    const size_t o = (size_t)p * wb.cap1 + i;
    WinSrc s;
    const int oc = min(max(lastOctave[o], 0), pp.nlevels - 1);
    s.x = uv[2 * o];
    s.y = uv[2 * o + 1];
    s.r = valid[o] ? __fmul_rn(pp.th, pp.scale[oc]) : -1.f;
    s.ur = __fsub_rn(s.x, __fmul_rn(pp.bf, invz[o]));
    if (pp.direction == 1) { s.minLevel = oc; s.maxLevel = -1; }
    else if (pp.direction == 2) { s.minLevel = 0; s.maxLevel = oc; }
    else { s.minLevel = oc - 1; s.maxLevel = oc + 1; }
    src[o] = s;
}

// ---------------------------------------------------------------------------------------------
// Window query.  One warp per source keypoint walks the covered grid columns; the cells of one column are
// one contiguous run of the CSR list, scanned by the lanes with a ballot prefix so that accepted candidates
// keep GetFeaturesInArea's order.  fill == 0: count only (listStart receives the counts);
// fill == 1: write (distance << 16 | target index) at listStart[i] + position.
// uRight2 != nullptr adds the stereo consistency predicate of fmatcher.cpp:2549-2555 (it does not depend
// on earlier matches, so it is applied here).  grid (ceil(cap1/8), nPairs), block 256.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_window_list(WinBatch wb, GridBounds gb, const WinSrc* __restrict__ src,
                                                     const uint8_t* __restrict__ desc1, const viorb_keypoint* __restrict__ kps2,
                                                     const uint8_t* __restrict__ desc2, const float* __restrict__ uRight2,
                                                     const int* __restrict__ cellStart, const float4* __restrict__ cellKp,
                                                     int* __restrict__ listStart, uint32_t* __restrict__ list, int fill) {
    const int lane = threadIdx.x & 31, p = blockIdx.y;
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= (fill && wb.n1fill ? wb.n1fill[p] : wb.n1[p])) return;
    const size_t o = (size_t)p * wb.cap1 + i;
    const WinSrc s = src[o];
    int total = 0;
    const int* LS = listStart + (size_t)p * (wb.cap1 + 1);
    uint32_t* out = fill ? list + LS[i] : nullptr;
    if (s.r >= 0.f) {
        const float x = s.x, y = s.y, r = s.r;
        const int c0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, gb.minX), r), gb.invW)));
        const int c1 = min(VIORB_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, gb.minX), r), gb.invW)));
        const int r0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, gb.minY), r), gb.invH)));
        const int r1 = min(VIORB_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, gb.minY), r), gb.invH)));
        if (c0 < VIORB_GRID_COLS && c1 >= 0 && r0 < VIORB_GRID_ROWS && r1 >= 0) {
            const bool check = s.minLevel > 0 || s.maxLevel >= 0;
            const int* CS = cellStart + (size_t)p * (VIORB_GRID_CELLS + 1);
            const float4* CK = cellKp + (size_t)p * wb.cap2;
            const uint4* D2 = reinterpret_cast<const uint4*>(desc2 + winbatch_base2(wb, p) * 32);
            uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
            if (fill && desc1) {
                const uint4* q = reinterpret_cast<const uint4*>(desc1 + o * 32);
                a0 = q[0]; a1 = q[1];
            }
            // one entry of the CSR list per lane: test it, and append the accepted ones in lane order
            auto visit = [&](int j, bool live) {
                bool ok = false;
                int i2 = 0;
                if (live) {
                    const float4 k = CK[j];
                    const int oct = __float_as_int(k.z);
                    i2 = __float_as_int(k.w);
                    ok = true;
                    if (check) {
                        if (oct < s.minLevel) ok = false;
                        if (s.maxLevel >= 0 && oct > s.maxLevel) ok = false;
                    }
                    if (!(fabsf(__fsub_rn(k.x, x)) < r && fabsf(__fsub_rn(k.y, y)) < r)) ok = false;
                    if (ok && uRight2) {
                        const float u2 = uRight2[(size_t)p * wb.cap2 + i2];
                        if (u2 > 0.f && fabsf(__fsub_rn(s.ur, u2)) > r) ok = false;
                    }
                }
                const unsigned m = __ballot_sync(0xffffffffu, ok);
                if (fill && ok) {
                    const int d = desc1 ? hamming256(a0, a1, D2[(size_t)i2 * 2], D2[(size_t)i2 * 2 + 1]) : 0;
                    out[total + __popc(m & ((1u << lane) - 1))] = ((uint32_t)d << 16) | (uint32_t)i2;
                }
                total += __popc(m);
            };
            const int ncols = c1 - c0 + 1;
            if (ncols <= 32) {
                // All covered columns at once: lane c reads the run [b, e) of column c0 + c, an inclusive scan of the
                // run lengths flattens the runs into one index space [0, T) in the same (column, entry) order, and the
                // lanes walk it 32 entries at a time — one dependent round trip for the bounds instead of one per column.
                int b = 0, len = 0;
                if (lane < ncols) {
                    b = CS[(c0 + lane) * VIORB_GRID_ROWS + r0];
                    len = CS[(c0 + lane) * VIORB_GRID_ROWS + r1 + 1] - b;
                }
                int incl = len;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += v;
                }
                const int T = __shfl_sync(0xffffffffu, incl, 31);
                for (int t0 = 0; t0 < T; t0 += 32) {
                    const int t = min(t0 + lane, T - 1);
                    int lo = 0;                      // first column whose inclusive prefix exceeds t
#pragma unroll
                    for (int step = 16; step; step >>= 1) {
                        const int v = __shfl_sync(0xffffffffu, incl, lo + step - 1);
                        if (v <= t) lo += step;
                    }
                    const int cb = __shfl_sync(0xffffffffu, b, lo), ci = __shfl_sync(0xffffffffu, incl, lo),
                              cl = __shfl_sync(0xffffffffu, len, lo);
                    visit(cb + (t - (ci - cl)), t0 + lane < T);
                }
            } else {
                for (int ix = c0; ix <= c1; ++ix) {
                    const int b = CS[ix * VIORB_GRID_ROWS + r0], e = CS[ix * VIORB_GRID_ROWS + r1 + 1];
                    for (int j0 = b; j0 < e; j0 += 32) visit(j0 + lane, j0 + lane < e);
                }
            }
        }
    }
    if (!fill && lane == 0) listStart[(size_t)p * (wb.cap1 + 1) + i] = total;
}

// exclusive scan of the per-source counts of one pair -> list offsets relative to the pair's base, which
// is pairBase[p] (filled by the host from the totals).  grid (nPairs), block 256.
__global__ void __launch_bounds__(256) k_window_scan(WinBatch wb, int* __restrict__ listStart, int* __restrict__ totals) {
    __shared__ int warpSums[36];
    const int p = blockIdx.x;
    int* LS = listStart + (size_t)p * (wb.cap1 + 1);
    const int n1 = wb.n1[p];
    const int total = block_excl_scan(LS, n1, warpSums);
    if (threadIdx.x == 0) { LS[n1] = total; totals[p] = total; }
}
// Device-side replacement of the host prefix sum over pairs (resident, asynchronous callers): list offsets of the
// pairs and the budget check.  A pair whose list does not fit gets n1eff = 0 (it produces no matches) and status bit 4.
// grid 1, block 1.
__global__ void k_window_pairbase(int npairs, const int* __restrict__ n1, const int* __restrict__ totals, int* __restrict__ pairBase,
                                  int* __restrict__ n1eff, size_t listCap, int* __restrict__ status) {
    size_t sum = 0;
    for (int p = 0; p < npairs; ++p) {
        pairBase[p] = (int)sum;
        if (sum + (size_t)totals[p] > listCap) { n1eff[p] = 0; atomicOr(status, 4); }
        else { n1eff[p] = n1[p]; sum += (size_t)totals[p]; }
    }
}
__global__ void k_window_rebase(WinBatch wb, int* __restrict__ listStart, const int* __restrict__ pairBase) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
    if (i <= wb.n1[p]) listStart[(size_t)p * (wb.cap1 + 1) + i] += pairBase[p];
}

// ---------------------------------------------------------------------------------------------
// Order-dependent part.  One warp per frame pair.
// ---------------------------------------------------------------------------------------------
#define VIORB_WIN_EMPTY 0xFFFFFFFFu

__device__ __forceinline__ void three_maxima_dev(const int* count, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < VIORB_HISTO_LENGTH; ++i) {
        const int s = count[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
}

__device__ __forceinline__ int rot_bin_dev(float a1, float a2) {
    const float factor = __fdiv_rn(1.0f, (float)VIORB_HISTO_LENGTH);
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, factor));
    if (bin == VIORB_HISTO_LENGTH) bin = 0;
    return bin;
}

// The candidate lists of 32 consecutive source keypoints are one contiguous range of `list`; the sequential kernels
// below copy the range of the NEXT chunk of 32 sources into shared memory (cp.async) while they work on the current one,
// and know the range of the chunk after that (LS values in registers), so that no global-memory round trip sits on the
// per-keypoint critical path.  Entries of a range beyond VIORB_WIN_STAGE are read from global memory.
#define VIORB_WIN_STAGE 2048
struct WinChunk { int sA, cN, r0, r1; };       // per lane: list start / count of source c0 + lane; range of the chunk
__device__ __forceinline__ WinChunk win_chunk_load(const int* __restrict__ LS, int c0, int n1, int lane) {
    WinChunk w{0, 0, 0, 0};
    if (c0 < n1) {
        if (c0 + lane < n1) { w.sA = LS[c0 + lane]; w.cN = LS[c0 + lane + 1] - w.sA; }
        w.r0 = LS[c0];
        w.r1 = LS[min(c0 + 32, n1)];
    }
    return w;
}
__device__ __forceinline__ void win_chunk_stage(uint32_t* buf, const uint32_t* __restrict__ list, const WinChunk& w, int lane) {
    const int m = min(w.r1 - w.r0, VIORB_WIN_STAGE);
    for (int j = lane; j < m; j += 32) __pipeline_memcpy_async(buf + j, list + w.r0 + j, 4);
    __pipeline_commit();
}

// SearchForInitialization (fmatcher.cpp:983-1098).  matchedDist / m21: per-pair scratch of cap2 entries, used only when
// they do not fit the CTA's dynamic shared memory (useSmem == 0); binOf / accepted: per-pair scratch of cap1 entries.
// matches12 [cap1] and prevMatched are outputs.
// The loop over the source keypoints is sequential (a candidate counts only if no earlier keypoint matched it better,
// :1024), so its latency per keypoint is what matters: vMatchedDistance / vnMatches21 live in shared memory, the warp
// skips the keypoints without candidates 32 at a time (only octave-0 keypoints search, :997), the first 32 candidates of
// the next keypoint are loaded while the current one is reduced, and everything that does not feed the next decision
// (vnMatches12, the rotation bins) is written and forgotten or computed after the loop.
// grid (nPairs), block 32, dynamic smem = useSmem ? cap2 * 8 : 0.
__global__ void __launch_bounds__(32) k_resolve_ini(WinBatch wb, const viorb_keypoint* __restrict__ kps1,
                                                    const viorb_keypoint* __restrict__ kps2, const int* __restrict__ listStart,
                                                    const uint32_t* __restrict__ list, float nnRatio, int checkOrientation,
                                                    int* __restrict__ matchedDist, int* __restrict__ m21, int8_t* __restrict__ binOf,
                                                    int* __restrict__ accepted, int32_t* __restrict__ matches12,
                                                    float* __restrict__ prevMatched, int* __restrict__ nMatches, int useSmem) {
    VIORB_DYN_SMEM(dsm);
    __shared__ int hist[VIORB_HISTO_LENGTH];
    __shared__ int keep[3];
    const int lane = threadIdx.x, p = blockIdx.x;
    const int n1 = wb.n1[p], n2 = wb.n2[p];
    const int* LS = listStart + (size_t)p * (wb.cap1 + 1);
    int* MD = useSmem ? reinterpret_cast<int*>(dsm) : matchedDist + (size_t)p * wb.cap2;
    int* M21 = useSmem ? reinterpret_cast<int*>(dsm) + wb.cap2 : m21 + (size_t)p * wb.cap2;
    int8_t* BO = binOf + (size_t)p * wb.cap1;
    int* ACC = accepted + (size_t)p * wb.cap1;           // the target a source keypoint was matched to when it was accepted
    int32_t* M12 = matches12 + (size_t)p * wb.cap1;
    const viorb_keypoint* K1 = kps1 + (size_t)p * wb.cap1;
    const viorb_keypoint* K2 = kps2 + (size_t)p * wb.cap2;
    for (int i = lane; i < n2; i += 32) { MD[i] = 0x7fffffff; M21[i] = -1; }
    for (int i = lane; i < n1; i += 32) { M12[i] = -1; ACC[i] = -1; }
    if (lane < VIORB_HISTO_LENGTH) hist[lane] = 0;
    __syncwarp();
    __shared__ uint32_t stage[2][VIORB_WIN_STAGE];
    WinChunk wc = win_chunk_load(LS, 0, n1, lane);
    win_chunk_stage(stage[0], list, wc, lane);
    WinChunk wn = win_chunk_load(LS, 32, n1, lane);
    for (int c0 = 0, cb = 0; c0 < n1; c0 += 32, cb ^= 1) {
        win_chunk_stage(stage[cb ^ 1], list, wn, lane);                     // next chunk's candidates
        const WinChunk wnn = win_chunk_load(LS, c0 + 64, n1, lane);          // the chunk after that: its LS values
        __pipeline_wait_prior(1);                                           // this chunk's candidates have landed
        __syncwarp();
        const uint32_t* SB = stage[cb];
        unsigned act = __ballot_sync(0xffffffffu, wc.cN > 0);
        while (act) {
            const int b = __ffs(act) - 1;
            act &= act - 1;
            const int s0 = __shfl_sync(0xffffffffu, wc.sA, b), cnt = __shfl_sync(0xffffffffu, wc.cN, b), i1 = c0 + b;
            const int so = s0 - wc.r0;
            auto entry = [&](int j) { return so + j < VIORB_WIN_STAGE ? SB[so + j] : list[s0 + j]; };
            unsigned b1 = VIORB_WIN_EMPTY, b2 = VIORB_WIN_EMPTY;          // keys: distance << 16 | list position
            for (int j = lane; j < cnt; j += 32) {
                const uint32_t e = entry(j);
                const int dist = (int)(e >> 16), i2 = (int)(e & 0xFFFF);
                if (MD[i2] <= dist) continue;                               // fmatcher.cpp:1024
                const unsigned key = ((unsigned)dist << 16) | (unsigned)j;
                b2 = min(b2, max(key, b1));
                b1 = min(b1, key);
            }
            // best and second best of the warp: keys are unique (they hold the list position)
            const unsigned B1 = __reduce_min_sync(0xffffffffu, b1);
            const unsigned B2 = __reduce_min_sync(0xffffffffu, b1 == B1 ? b2 : b1);
            if (B1 != VIORB_WIN_EMPTY) {
                const int bestDist = (int)(B1 >> 16);
                const float second = B2 == VIORB_WIN_EMPTY ? (float)0x7fffffff : (float)(int)(B2 >> 16);
                if (bestDist <= VIORB_TH_LOW && (float)bestDist < __fmul_rn(second, nnRatio)) {
                    const int bestIdx2b = (int)(entry((int)(B1 & 0xFFFF)) & 0xFFFF);
                    if (lane == 0) {
                        const int old = M21[bestIdx2b];
                        if (old >= 0) M12[old] = -1;
                        M12[i1] = bestIdx2b;
                        M21[bestIdx2b] = i1;
                        MD[bestIdx2b] = bestDist;
                        ACC[i1] = bestIdx2b;                                 // its rotation bin is computed after the loop
                    }
                    __syncwarp();
                }
            }
        }
        __syncwarp();                                                       // everyone is done with stage[cb] before it is refilled
        wc = wn;
        wn = wnn;
    }
    __pipeline_wait_prior(0);
    __syncwarp();
    if (checkOrientation) {
        // rotHist holds every keypoint that was accepted at its turn, displaced later or not (:1059-1072)
        for (int i = lane; i < n1; i += 32) {
            const int a2 = ACC[i];
            int bin = -1;
            if (a2 >= 0) {
                bin = rot_bin_dev(K1[i].angle, K2[a2].angle);
                atomicAdd(&hist[bin], 1);
            }
            BO[i] = (int8_t)bin;
        }
        __syncwarp();
        if (lane == 0) three_maxima_dev(hist, keep[0], keep[1], keep[2]);
        __syncwarp();
        const int a = keep[0], b = keep[1], c = keep[2];
        for (int i = lane; i < n1; i += 32) {
            const int bin = BO[i];
            if (bin >= 0 && bin != a && bin != b && bin != c && M12[i] >= 0) M12[i] = -1;
        }
        __syncwarp();
    }
    int nm = 0;
    for (int i = lane; i < n1; i += 32) {
        const int m = M12[i];
        if (m >= 0) {
            ++nm;
            prevMatched[2 * ((size_t)p * wb.cap1 + i)] = K2[m].x;
            prevMatched[2 * ((size_t)p * wb.cap1 + i) + 1] = K2[m].y;
        }
    }
    nm = __reduce_add_sync(0xffffffffu, nm);
    if (lane == 0) nMatches[p] = nm;
}

// SearchByProjection(CurrentFrame, LastFrame) (fmatcher.cpp:2471-2688).
// blocked [cap2]: in: current keypoint already holds an observed map point; updated as matches are made.
// assigned [cap2]: -1 untouched, -2 cleared by the rotation check, >= 0 the source index now held.
// pushIdx: per-pair scratch of 4 * cap1 ints ((target, source) per accepted match = the rotHist lists), pushBin 2 * cap1.
// Two-camera frames (CurrentFrame.Nleft != -1, :2593-2660): listStartR / listR are the candidate lists of the SAME sources
// in the right camera's grid (indices relative to keypointsRight_), nleft[p] = Nleft.  Per source, in the reference's
// order: the left window (an EMPTY left window skips the right one too: the `continue` at :2534 precedes the right
// block), then the right window with indices, blocked flags and the rotation histogram offset by Nleft.
// Sequential in the source like k_resolve_ini, and built the same way: the blocked flags live in shared memory, sources
// without candidates are skipped 32 at a time, the next source's first candidates and `observed` flags are loaded ahead,
// the rotation bins are computed after the loop.  grid (nPairs), block 32, dynamic smem = n2cap bytes.
__global__ void __launch_bounds__(32) k_resolve_proj(WinBatch wb, const float* __restrict__ lastAngle,
                                                     const uint8_t* __restrict__ observed, const viorb_keypoint* __restrict__ kps2,
                                                     const int* __restrict__ listStart, const uint32_t* __restrict__ list,
                                                     const int* __restrict__ listStartR, const uint32_t* __restrict__ listR,
                                                     const int* __restrict__ nleft, int n2cap,
                                                     int checkOrientation, const uint8_t* __restrict__ blocked,
                                                     int* __restrict__ pushIdx, int8_t* __restrict__ pushBin,
                                                     int32_t* __restrict__ assigned, int* __restrict__ nMatches,
                                                     const int* __restrict__ n2total) {
    VIORB_DYN_SMEM(dsm);
    __shared__ int hist[VIORB_HISTO_LENGTH];
    __shared__ int keep[3];
    const int lane = threadIdx.x, p = blockIdx.x;
    const int n1 = wb.n1[p], n2 = n2total ? n2total[p] : wb.n2[p];
    const bool rig = listStartR != nullptr;
    const int nL = rig ? nleft[p] : 0;
    const int* LS = listStart + (size_t)p * (wb.cap1 + 1);
    const int* LSR = rig ? listStartR + (size_t)p * (wb.cap1 + 1) : nullptr;
    uint8_t* BL = dsm;                                                      // mvpMapPoints[i2] && Observations() > 0, as it evolves
    int32_t* AS = assigned + (size_t)p * n2cap;
    int* PI = pushIdx + (size_t)p * wb.cap1 * 4;
    int8_t* PB = pushBin + (size_t)p * wb.cap1 * 2;
    const viorb_keypoint* K2 = kps2 + (size_t)p * n2cap;
    const uint8_t* OB = observed + (size_t)p * wb.cap1;
    for (int i = lane; i < n2; i += 32) { AS[i] = -1; BL[i] = blocked[(size_t)p * n2cap + i]; }
    if (lane < VIORB_HISTO_LENGTH) hist[lane] = 0;
    __syncwarp();
    int nPush = 0;
    // best unblocked candidate of one list; on acceptance lane 0 records it.  off: index offset of this camera; SB / so:
    // the staged copy of the list range and this source's offset in it (so >= VIORB_WIN_STAGE: not staged).
    auto pass = [&](const uint32_t* L, const uint32_t* SB, int so, int s0, int cnt, int off, int i, int obs) {
        auto entry = [&](int j) { return so + j < VIORB_WIN_STAGE ? SB[so + j] : L[s0 + j]; };
        unsigned b1 = VIORB_WIN_EMPTY;
        for (int j = lane; j < cnt; j += 32) {
            const uint32_t e = entry(j);
            if (BL[(e & 0xFFFF) + off]) continue;                           // fmatcher.cpp:2545-2547 / :2621-2623
            b1 = min(b1, (e & 0xFFFF0000u) | (unsigned)j);
        }
        b1 = __reduce_min_sync(0xffffffffu, b1);
        if (b1 != VIORB_WIN_EMPTY && (int)(b1 >> 16) <= VIORB_TH_HIGH) {
            const int bestIdx2 = (int)(entry((int)(b1 & 0xFFFF)) & 0xFFFF) + off;
            if (lane == 0) {
                AS[bestIdx2] = i;
                BL[bestIdx2] = (uint8_t)obs;
                PI[2 * nPush] = bestIdx2;                                   // its rotation bin is computed after the loop
                PI[2 * nPush + 1] = i;
            }
            ++nPush;
            __syncwarp();
        }
    };
    __shared__ uint32_t stage[2][VIORB_WIN_STAGE];
    WinChunk wc = win_chunk_load(LS, 0, n1, lane);
    win_chunk_stage(stage[0], list, wc, lane);
    WinChunk wn = win_chunk_load(LS, 32, n1, lane);
    int ob = lane < n1 ? OB[lane] : 0, obn = 32 + lane < n1 ? OB[32 + lane] : 0;
    for (int c0 = 0, cb = 0; c0 < n1; c0 += 32, cb ^= 1) {
        win_chunk_stage(stage[cb ^ 1], list, wn, lane);
        const WinChunk wnn = win_chunk_load(LS, c0 + 64, n1, lane);
        const int obnn = c0 + 64 + lane < n1 ? OB[c0 + 64 + lane] : 0;
        __pipeline_wait_prior(1);
        __syncwarp();
        const uint32_t* SB = stage[cb];
        unsigned act = __ballot_sync(0xffffffffu, wc.cN > 0);               // vIndices2.empty(): next map point
        while (act) {
            const int b = __ffs(act) - 1;
            act &= act - 1;
            const int s0 = __shfl_sync(0xffffffffu, wc.sA, b), cnt = __shfl_sync(0xffffffffu, wc.cN, b), i = c0 + b;
            const int obs = __shfl_sync(0xffffffffu, ob, b);
            pass(list, SB, s0 - wc.r0, s0, cnt, 0, i, obs);
            if (rig) {
                const int r0 = LSR[i], rc = LSR[i + 1] - r0;
                if (rc > 0) pass(listR, nullptr, VIORB_WIN_STAGE, r0, rc, nL, i, obs);
            }
        }
        __syncwarp();
        wc = wn; wn = wnn;
        ob = obn; obn = obnn;
    }
    __pipeline_wait_prior(0);
    __syncwarp();
    int removed = 0;
    if (checkOrientation) {
        const float* LA = lastAngle + (size_t)p * wb.cap1;
        for (int k = lane; k < nPush; k += 32) {
            const int bin = rot_bin_dev(LA[PI[2 * k + 1]], K2[PI[2 * k]].angle);
            PB[k] = (int8_t)bin;
            atomicAdd(&hist[bin], 1);
        }
        __syncwarp();
        if (lane == 0) three_maxima_dev(hist, keep[0], keep[1], keep[2]);
        __syncwarp();
        const int a = keep[0], b = keep[1], c = keep[2];
        // an entry of a bin that is not kept clears its target: CurrentFrame.mvpMapPoints[rotHist[i][j]] = NULL (:2676)
        for (int k = lane; k < nPush; k += 32) {
            const int bin = PB[k];
            if (bin != a && bin != b && bin != c) { AS[PI[2 * k]] = -2; ++removed; }
        }
    }
    removed = __reduce_add_sync(0xffffffffu, removed);
    if (lane == 0) nMatches[p] = nPush - removed;
}
