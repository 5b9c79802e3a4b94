// viorb_api.cu — host side of libviorb.so: handles, arenas, geometry tables, launches, and the C ABI
// declared in include/viorb.h.  No OpenCV, no torch; CUDA runtime only.
//
// There is deliberately NO CPU fallback: every entry point that computes launches the sm_100a
// kernels in kernels_extract.cuh / kernels_match.cuh and fails with VIORB_E_CUDA / VIORB_E_NO_DEVICE
// if that is impossible.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

// kNN-2 kernel of this build: 0 = k_knn2_f4 (the product: tcgen05.mma kind::mxf4), 1 = k_knn2_mma, 2 = k_knn2,
// 3 = k_knn2_tc5 (tcgen05.mma kind::i8, the round-1 product); 1-3 are experiments, the CPU emulation build has no
// tcgen05 and takes 1
#ifndef VIORB_KNN_VARIANT
#ifdef VIORB_EMU
#define VIORB_KNN_VARIANT 1
#else
#define VIORB_KNN_VARIANT 0
#endif
#endif

#include "kernels_extract.cuh"
#include "kernels_match.cuh"
#if VIORB_KNN_VARIANT == 0
#include "kernels_knn_f4.cuh"
#define VIORB_KNN_TC_KERNEL tc5::k_knn2_f4
#elif VIORB_KNN_VARIANT == 3
#include "kernels_knn_tc5.cuh"
#define VIORB_KNN_TC_KERNEL tc5::k_knn2_tc5
#endif
#define VIORB_KNN_TENSOR (VIORB_KNN_VARIANT == 0 || VIORB_KNN_VARIANT == 3)
#include "kernels_window.cuh"
#include "kernels_bow.cuh"
#include "kernels_kb8.cuh"

#define VIORB_MAX_CHUNKS 8

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(VIORB_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));         \
    } while (0)

inline int round_up(int v, int a) { return (v + a - 1) / a * a; }
inline int cv_round_f(float v) { return (int)lrintf(v); }
inline short sat_short(float v) { return (short)std::min(std::max(cv_round_f(v), -32768), 32767); }

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count) {
        release();
        n = count;
        return cudaMalloc((void**)&p, std::max<size_t>(count, 1) * sizeof(T));
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};
template <class T>
struct PinBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count) {
        release();
        n = count;
        return cudaMallocHost((void**)&p, std::max<size_t>(count, 1) * sizeof(T));
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; n = 0; }
};

// Results of viorb_bow_* for `frames` sets of `cap` slots (device side).
struct BowBufs {
    DevBuf<int> leaf, nid, nBow, nFv, fvStart;
    DevBuf<uint32_t> bowWord, fvNode, fvFeature;
    DevBuf<double> bowValue;
    size_t frames = 0, cap = 0;
    cudaError_t ensure(size_t f, size_t c) {
        if (f <= frames && c == cap && leaf.p) return cudaSuccess;
        release();
        cudaError_t e = cudaSuccess;
        auto ok = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
        ok(leaf.alloc(f * c)); ok(nid.alloc(f * c)); ok(nBow.alloc(f)); ok(nFv.alloc(f)); ok(fvStart.alloc(f * (c + 1)));
        ok(bowWord.alloc(f * c)); ok(fvNode.alloc(f * c)); ok(fvFeature.alloc(f * c)); ok(bowValue.alloc(f * c));
        frames = f; cap = c;
        return e;
    }
    void release() {
        leaf.release(); nid.release(); nBow.release(); nFv.release(); fvStart.release(); bowWord.release(); fvNode.release();
        fvFeature.release(); bowValue.release();
        frames = cap = 0;
    }
};

}  // namespace

struct viorb_vocabulary {
    int device = 0;
    VocabDev d{};
    DevBuf<int> childStart, children, wordId;
    DevBuf<uint4> desc;
    DevBuf<double> weight;
};

// buffers of viorb_f2f_batch (host_f2f.inc)
struct viorb_f2f_state {
    DevBuf<int> cellStart, cellList, listStart, totals, pairBase, n1eff, matchedDist, m21, nMatches, accepted;
    DevBuf<float4> cellKp;
    DevBuf<uint32_t> list;
    DevBuf<WinSrc> src;
    DevBuf<float> prevMatched;
    DevBuf<int8_t> binOf;
    DevBuf<int32_t> matches12;
    size_t listCap = 0;
    bool ready = false;
    int lastPairs = 0;
    void release() {
        cellStart.release(); cellList.release(); cellKp.release(); listStart.release(); totals.release(); pairBase.release(); n1eff.release();
        matchedDist.release(); m21.release(); nMatches.release(); accepted.release(); list.release(); src.release(); prevMatched.release();
        binOf.release(); matches12.release();
        ready = false;
    }
};

struct viorb_extractor {
    viorb_params prm;
    viorb_f2f_state* f2f = nullptr;   // buffers of viorb_f2f_batch (host_f2f.inc), created on first use
    // FExtractor ctor tables (fextractor.cpp:401-461)
    float scale[VIORB_MAX_LEVELS], invScale[VIORB_MAX_LEVELS], sigma2[VIORB_MAX_LEVELS], invSigma2[VIORB_MAX_LEVELS];
    int featPerLevel[VIORB_MAX_LEVELS];
    int umaxv[16];
    // geometry of the current image size
    ExtractGeom g;
    BlurTiles blurTiles;
    int totalBlurTiles = 0;
    std::vector<CellDesc> cells;
    size_t fastSmem = 0, octSmem = 0, octSmemDense = 0;
    // limits the arenas were sized for
    size_t capArena = 0, capCand = 0, capKept = 0, capCells = 0, capXTab = 0, capYTab = 0;
    int capOut = 0;
    // device state
    cudaStream_t stream = nullptr;
    bool ownStream = false;
    // host-batch pipelining (viorb_extract_batch): images go up on cpStream and results come down on outStream
    // chunk by chunk while the kernels of the neighbouring chunks run on `stream`
    cudaStream_t cpStream = nullptr, outStream = nullptr;
    cudaStream_t sideOct = nullptr;                       // k_octree_dense runs here, beside k_octree on `stream`
    cudaEvent_t evOctFork = nullptr, evOctJoin = nullptr;
    cudaEvent_t evIn[VIORB_MAX_CHUNKS] = {}, evDone[VIORB_MAX_CHUNKS] = {}, evCnt[VIORB_MAX_CHUNKS] = {}, evEntry = nullptr;
    cudaEvent_t evJoin = nullptr;     // orders this handle's stream against a partner handle's (stereo)
    cudaEvent_t evBlock = nullptr;    // blocking-sync event: batch calls sleep on it instead of spinning in the driver
    DevBuf<uint8_t> dIn, dPyr, dBlur, dDesc;
    DevBuf<uint32_t> dCand;
    DevBuf<uint16_t> dKeyNode;
    DevBuf<int> dCandCount, dKeptCount, dSlot2Kept, dNOut, dMono, dStatus, dXofs, dYofs;
    DevBuf<short2> dXalpha, dYbeta;
    DevBuf<KeptKp> dKept;
    DevBuf<viorb_keypoint> dKps;
    DevBuf<viorb_keypoint> dKpsUn;    // ukeypoints_, filled by k_emit once viorb_extractor_set_camera was called
    CamModel cam{};
    BowBufs bow;                      // viorb_bow_resident results of the last batch
    int bowN = 0;
    DevBuf<CellDesc> dCells;
    DevBuf<int> dOrientTab;
    DevBuf<uint8_t> dSlotLevel;
    // stereo / matching results of the last viorb_stereo_batch on this (left) handle
    DevBuf<int32_t> dBestIdx, dBestDist, dSad, dKnnIdx, dKnnDist, dLrMatch;
    DevBuf<float> dURight, dDepth;
    DevBuf<int> dNStereo, dNRatio, dRowStart;
    DevBuf<int2> dRowList;
    int rowListCap = 0;
    bool matchReady = false;
    DevBuf<viorb_match> dMatches;
    // pinned staging
    PinBuf<uint8_t> hIn, hDesc;
    PinBuf<viorb_keypoint> hKps;
    PinBuf<int> hCounts;
    // stage timing (CUDA events on the handle's stream; read back with viorb_*_stage_ms)
    cudaEvent_t evX[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t evS[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool timing = false, haveX = false, haveS = false;
    // last run
    FrameSrc src{};
    int lastN = 0;
    int lastMaxCount = 0;             // largest keypoint count of the last host-batch extraction (0 = not known on the host)
    int inPitch = 0;
};

namespace {

// ---- FExtractor::FExtractor tables ----------------------------------------------------------
void build_tables(viorb_extractor* ex) {
    const viorb_params& p = ex->prm;
    const int nl = p.nlevels;
    ex->scale[0] = 1.0f;
    ex->sigma2[0] = 1.0f;
    for (int i = 1; i < nl; ++i) {
        ex->scale[i] = ex->scale[i - 1] * p.scale_factor;
        ex->sigma2[i] = ex->scale[i] * ex->scale[i];
    }
    for (int i = 0; i < nl; ++i) {
        ex->invScale[i] = 1.0f / ex->scale[i];
        ex->invSigma2[i] = 1.0f / ex->sigma2[i];
    }
    const float factor = 1.0f / p.scale_factor;
    float want = p.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        ex->featPerLevel[l] = cv_round_f(want);
        sum += ex->featPerLevel[l];
        want *= factor;
    }
    ex->featPerLevel[nl - 1] = std::max(p.nfeatures - sum, 0);
    // umax: end of each row of the radius-15 disc, made symmetric (fextractor.cpp:444-460)
    int* um = ex->umaxv;
    const int hp = VIORB_HALF_PATCH;
    const int vmax = (int)std::floor(hp * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(hp * std::sqrt(2.f) / 2);
    for (int v = 0; v <= vmax; ++v) um[v] = (int)lrint(std::sqrt((double)hp * hp - v * v));
    for (int v = hp, v0 = 0; v >= vmin; --v) {
        while (um[v0] == um[v0 + 1]) ++v0;
        um[v] = v0;
        ++v0;
    }
}

// ---- geometry for one image size ------------------------------------------------------------
struct ResizeTabs {
    std::vector<int> xofs, yofs;
    std::vector<short2> xalpha, ybeta;
};

// cv::resize INTER_LINEAR coefficient tables (OpenCV resize.cpp, linear 8U path).
void build_resize_tables(int sw, int sh, int dw, int dh, ResizeTabs& t) {
    const double scale_x = 1.0 / ((double)dw / sw), scale_y = 1.0 / ((double)dh / sh);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)std::floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        short a0 = sat_short((1.f - fx) * 2048), a1 = sat_short(fx * 2048);
        // k_resize reads columns sx and sx+1 unconditionally: express "last column, weight 2048|0" as
        // "(last-1), weight 0|2048" (same value, both columns inside the row)
        if (sx == sw - 1 && sw >= 2) { sx = sw - 2; a1 = a0; a0 = 0; }
        t.xofs.push_back(sx);
        t.xalpha.push_back(make_short2(a0, a1));
    }
    while (t.xofs.size() % 4) { t.xofs.push_back(0); t.xalpha.push_back(make_short2(0, 0)); }   // int4 loads in k_resize
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)std::floor(fy);
        fy -= sy;
        t.yofs.push_back(sy);
        t.ybeta.push_back(make_short2(sat_short((1.f - fy) * 2048), sat_short(fy * 2048)));
    }
    // k_resize fetches the VIORB_RS_ROWS entries of a thread's rows with vector loads
    while (t.yofs.size() % VIORB_RS_ROWS) { t.yofs.push_back(0); t.ybeta.push_back(make_short2(2048, 0)); }
}

int build_geometry(viorb_extractor* ex, int W, int H, ResizeTabs& tabs) {
    ExtractGeom& g = ex->g;
    std::memset(&g, 0, sizeof(g));
    const int nl = ex->prm.nlevels;
    g.nlevels = nl; g.W = W; g.H = H;
    g.iniTh = std::min(std::max(ex->prm.ini_th_fast, 0), 255);
    g.minTh = std::min(std::max(ex->prm.min_th_fast, 0), 255);
    ex->cells.clear();
    tabs = ResizeTabs();
    size_t off = 0;
    int candOff = 0, keptOff = 0, maxRW = 1, maxRH = 1, maxArea = 1, maxNodes = 8, blurBase = 0;
    for (int l = 0; l < nl; ++l) {
        LevelGeom& L = g.lv[l];
        L.w = cv_round_f((float)W * ex->invScale[l]);           // fextractor.cpp:1139-1140
        L.h = cv_round_f((float)H * ex->invScale[l]);
        if (L.w < 1 || L.h < 1) return fail(VIORB_E_ARG, "image too small for the configured number of levels");
        if (L.w > 4096 + 32 || L.h > 4096 + 32) return fail(VIORB_E_ARG, "image larger than 4128 px is not supported");
        L.pitch = round_up(L.w, 64);
        L.off = (int)off;
        off += (size_t)round_up(L.pitch * L.h, 256);
        L.scale = ex->scale[l];
        L.kpSize = (float)(int)(31 * ex->scale[l]);             // fextractor.cpp:828,837
        L.maxBX = L.w - VIORB_EDGE + 3;
        L.maxBY = L.h - VIORB_EDGE + 3;
        L.xTab = (int)tabs.xofs.size();
        L.yTab = (int)tabs.yofs.size();
        if (l > 0) {
            const LevelGeom& S = g.lv[l - 1];
            L.areaHalf = (S.w == 2 * L.w && S.h == 2 * L.h) ? 1 : 0;
            build_resize_tables(S.w, S.h, L.w, L.h, tabs);
            // word path of k_resize: the two source columns of each of the 4 pixels of a group lie within 12 bytes
            // of the aligned word holding the group's first column (true for any scale factor <= 2)
            L.rsWords = 1;
            for (size_t k = (size_t)L.xTab; k + 3 < tabs.xofs.size(); k += 4) {
                const int wb = tabs.xofs[k] & ~3;
                for (int i = 0; i < 4 && (int)(k - L.xTab) + i < L.w; ++i) {      // the table's padding entries do not count
                    const int o = tabs.xofs[k + i] - wb;
                    if (o < 0 || o + 1 > 11) L.rsWords = 0;
                }
            }
        }
        // FAST cell grid (fextractor.cpp:769-806)
        const float width = (float)(L.maxBX - VIORB_MINB), height = (float)(L.maxBY - VIORB_MINB);
        L.nCols = (int)(width / 30.f);
        L.nRows = (int)(height / 30.f);
        L.cellBase = (int)ex->cells.size();
        L.lastCol = L.lastRow = 0;
        if (L.nCols > 0 && L.nRows > 0 && width > 0 && height > 0) {
            L.wCell = (int)std::ceil(width / L.nCols);
            L.hCell = (int)std::ceil(height / L.nRows);
            L.cinv = (1u << 20) / (unsigned)L.wCell + 1;
            for (int i = 0; i < L.nRows; ++i) {
                const float iniY = (float)(VIORB_MINB + i * L.hCell);
                float maxY = iniY + L.hCell + 6;
                if (iniY >= L.maxBY - 3) continue;
                if (maxY > L.maxBY) maxY = (float)L.maxBY;
                const int y0 = (int)iniY + 3, y1 = (int)maxY - 3;
                if (y1 <= y0) continue;                            // ROI lower than 7: cv::FAST finds nothing
                // the cells of this row: [x0_j, x1_j), all wCell wide except possibly the last
                std::vector<std::pair<int, int>> row;
                for (int j = 0; j < L.nCols; ++j) {
                    const float iniX = (float)(VIORB_MINB + j * L.wCell);
                    float maxX = iniX + L.wCell + 6;
                    if (iniX >= L.maxBX - 6) continue;
                    if (maxX > L.maxBX) maxX = (float)L.maxBX;
                    const int x0 = (int)iniX + 3, x1 = (int)maxX - 3;
                    if (x1 <= x0) continue;
                    row.push_back({x0, x1});
                    L.lastCol = std::max(L.lastCol, j);
                    L.lastRow = std::max(L.lastRow, i);
                }
                // k_fast works on strips of consecutive cells of one cell row
                const int per = std::max(1, std::min(std::min(VIORB_FAST_STRIP_CELLS, VIORB_FAST_MAXRW / L.wCell), VIORB_FAST_MAXAREA / (L.wCell * (y1 - y0))));
                for (size_t j0 = 0; j0 < row.size(); j0 += per) {
                    const size_t j1 = std::min(row.size(), j0 + per);
                    CellDesc c;
                    c.level = (short)l; c.ncell = (short)(j1 - j0);
                    c.x0 = (short)row[j0].first; c.x1 = (short)row[j1 - 1].second;
                    c.y0 = (short)y0; c.y1 = (short)y1;
                    {   // the constants k_fast derives from the strip (same expressions as in the kernel)
                        const int rw = c.x1 - c.x0, ox = c.x0 - 3, ax = ((ox & ~3) - 4) & ~15, shift = ox - ax, col0 = shift + 3;
                        const int nB = ((((shift + 3 + rw + 2) >> 2) + 2) << 2);
                        const int ng = ((col0 + rw - 1) >> 2) - (col0 >> 2) + 1;
                        c.nC = (short)((nB + 15) >> 4); c.ng = (short)ng; c.pad2[0] = c.pad2[1] = 0;
                        c.invNC = (1u << 20) / (unsigned)c.nC + 1;
                        c.ginv = (1u << 22) / (unsigned)ng + 1;
                    }
                    ex->cells.push_back(c);
                    maxRW = std::max(maxRW, c.x1 - c.x0);
                    maxRH = std::max(maxRH, c.y1 - c.y0);
                    maxArea = std::max(maxArea, (c.x1 - c.x0) * (c.y1 - c.y0));
                }
            }
        } else {
            L.wCell = L.hCell = 1;
            L.cinv = (1u << 20) + 1;
        }
        L.nCells = (int)ex->cells.size() - L.cellBase;
        // NMS leaves at most one corner per 2x2 block
        L.candOff = candOff;
        L.candCap = std::max(0, (L.maxBX - VIORB_MINB)) * std::max(0, (L.maxBY - VIORB_MINB)) / 4 + 64;
        candOff += round_up(L.candCap, 8);
        // DistributeOctTree roots (fextractor.cpp:533-536)
        L.nWant = ex->featPerLevel[l];
        const int dx = L.maxBX - VIORB_MINB, dy = L.maxBY - VIORB_MINB;
        L.nIni = dy > 0 ? (int)std::round((float)dx / dy) : 1;
        if (L.nIni < 1) L.nIni = 1;
        if (L.nIni > 64) return fail(VIORB_E_ARG, "aspect ratio above 64:1 is not supported");
        L.hX = (float)dx / L.nIni;
        L.keptOff = keptOff;
        L.keptCap = std::max(L.nWant + 3, 4 * L.nIni) + 1;
        keptOff += L.keptCap;
        maxNodes = std::max(maxNodes, L.keptCap + 4);
        ex->blurTiles.base[l] = blurBase;
        ex->blurTiles.tilesX[l] = (L.w + VIORB_BLUR_TW - 1) / VIORB_BLUR_TW;
        blurBase += ex->blurTiles.tilesX[l] * ((L.h + VIORB_BLUR_TH - 1) / VIORB_BLUR_TH);
    }
    ex->blurTiles.base[nl] = blurBase;
    ex->totalBlurTiles = blurBase;
    if (maxNodes > 65535) return fail(VIORB_E_ARG, "nfeatures too large for 16-bit node ids");
    g.totalCells = (int)ex->cells.size();
    g.candPerFrame = candOff;
    g.keptPerFrame = keptOff;
    g.outCap = keptOff;
    g.tileW = maxRW + 6;
    g.tileH = maxRH + 6;
    g.maxNodes = maxNodes;
    g.arenaStride = off;
    ex->fastSmem = fast_smem_bytes(g.tileH, maxArea);
    const int lutInts = octree_lut_ints(g.lv[0].maxBX - VIORB_MINB, g.lv[0].maxBY - VIORB_MINB);     // level 0 is the widest and tallest
    ex->octSmem = octree_smem_bytes(maxNodes, VIORB_OCT_LIGHT_PYR_CELLS, lutInts);
    ex->octSmemDense = octree_smem_bytes(maxNodes, VIORB_OCT_PYR_CELLS, lutInts);
    if (ex->octSmemDense > 220 * 1024) return fail(VIORB_E_ARG, "nfeatures too large for the quadtree's shared memory");
    if (maxRW > VIORB_FAST_MAXRW || maxRH > 127) return fail(VIORB_E_ARG, "FAST cell larger than supported");
    return VIORB_OK;
}

int upload_geometry(viorb_extractor* ex, const ResizeTabs& tabs) {
    if (ex->cells.size() > ex->capCells || tabs.xofs.size() > ex->capXTab || tabs.yofs.size() > ex->capYTab)
        return fail(VIORB_E_CAPACITY, "image size exceeds the max_width/max_height the handle was created with");
    CU(cudaMemcpyAsync(ex->dCells.p, ex->cells.data(), ex->cells.size() * sizeof(CellDesc), cudaMemcpyHostToDevice, ex->stream));
    CU(cudaMemcpyAsync(ex->dXofs.p, tabs.xofs.data(), tabs.xofs.size() * sizeof(int), cudaMemcpyHostToDevice, ex->stream));
    CU(cudaMemcpyAsync(ex->dXalpha.p, tabs.xalpha.data(), tabs.xalpha.size() * sizeof(short2), cudaMemcpyHostToDevice, ex->stream));
    CU(cudaMemcpyAsync(ex->dYofs.p, tabs.yofs.data(), tabs.yofs.size() * sizeof(int), cudaMemcpyHostToDevice, ex->stream));
    CU(cudaMemcpyAsync(ex->dYbeta.p, tabs.ybeta.data(), tabs.ybeta.size() * sizeof(short2), cudaMemcpyHostToDevice, ex->stream));
    std::vector<uint8_t> slotLevel((size_t)ex->g.keptPerFrame, 0);
    for (int l = 0; l < ex->g.nlevels; ++l)
        for (int i = 0; i < ex->g.lv[l].keptCap; ++i) slotLevel[ex->g.lv[l].keptOff + i] = (uint8_t)l;
    CU(cudaMemcpyAsync(ex->dSlotLevel.p, slotLevel.data(), slotLevel.size(), cudaMemcpyHostToDevice, ex->stream));
    OrientTab ot;
    build_orient_tab(ex->umaxv, ot);
    CU(cudaMemcpyAsync(ex->dOrientTab.p, &ot, sizeof ot, cudaMemcpyHostToDevice, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));      // the host vectors go out of scope
    return VIORB_OK;
}

int set_size(viorb_extractor* ex, int W, int H) {
    if (ex->g.W == W && ex->g.H == H) return VIORB_OK;
    if (W > ex->prm.max_width || H > ex->prm.max_height)
        return fail(VIORB_E_CAPACITY, "image larger than max_width x max_height of this handle");
    ResizeTabs tabs;
    int rc = build_geometry(ex, W, H, tabs);
    if (rc) { ex->g.W = ex->g.H = 0; return rc; }
    if (ex->g.arenaStride > ex->capArena || (size_t)ex->g.candPerFrame > ex->capCand ||
        (size_t)ex->g.keptPerFrame > ex->capKept || ex->g.outCap > ex->capOut) {
        ex->g.W = ex->g.H = 0;
        return fail(VIORB_E_CAPACITY, "image size exceeds the arenas sized at create time");
    }
    // per-frame stride of the keypoint / descriptor / match arrays = the handle's capacity, whatever the image
    // size: host buffers of viorb_extractor_capacity() rows then have the device layout (linear DMA)
    ex->g.outCap = ex->capOut;
    rc = upload_geometry(ex, tabs);
    if (rc) { ex->g.W = ex->g.H = 0; return rc; }
    return VIORB_OK;
}

// ---- once per device ----------------------------------------------------------------------------
// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to (function, device), not to a handle: it is raised to the
// opt-in maximum once, so handles with different needs (the monocular initialiser with 5 x nFeatures next to the
// tracking extractor, tracking.cpp:1087-1093) never lower each other's limit.  The kNN split scratch comes from a
// pool this library owns (the host application's default pool is left alone).
struct DeviceState {
    std::once_flag once;
    cudaError_t err = cudaSuccess;
    int sms = 148;
    int maxSmem = 0;
#ifndef VIORB_EMU
    cudaMemPool_t pool = nullptr;
#endif
    // ONE host -> device copy stream per device, shared by every handle's viorb_extract_batch: uploads of concurrent
    // calls (pipeline workers, left / right cameras) run in issue order at full link rate each, so the first caller's
    // kernels start after ITS images arrived - with one copy stream per handle the copy engines interleave all
    // uploads, every call's images arrive at the same late moment and the workers fall into lockstep
    // (all copy, then all compute: measured 13 ms cycles of three workers, 4.4 ms per step instead of 3.6).
    cudaStream_t h2d = nullptr;
    std::mutex h2dMu;                         // copy + event record of one chunk stay adjacent in the stream
};
DeviceState g_dev[64];
std::atomic<int> g_forceKnnSplit{0};      // viorb_debug_knn_split

int device_state(int dev, DeviceState** out) {
    if (dev < 0 || dev >= 64) return fail(VIORB_E_ARG, "device index out of range");
    DeviceState& d = g_dev[dev];
    std::call_once(d.once, [&d, dev] {
#ifndef VIORB_EMU
        auto ok = [&d](cudaError_t r) { if (d.err == cudaSuccess) d.err = r; };
        ok(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev));
        ok(cudaDeviceGetAttribute(&d.maxSmem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        // dynamic limit = opt-in maximum minus the kernel's static shared memory
        auto raise = [&](const void* fn) {
            cudaFuncAttributes fa{};
            ok(cudaFuncGetAttributes(&fa, fn));
            ok(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, d.maxSmem - (int)fa.sharedSizeBytes));
        };
        raise((const void*)k_fast);
        raise((const void*)k_octree);
        raise((const void*)k_octree_dense);
        raise((const void*)k_resolve_ini);
        raise((const void*)k_resolve_proj);
        // several ~40 KB CTAs of k_octree per SM: without this hint the driver picks a carveout that holds one
        ok(cudaFuncSetAttribute(k_octree, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
#if VIORB_KNN_TENSOR
        ok(cudaFuncSetAttribute(VIORB_KNN_TC_KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, tc5::SMEM_BYTES));
#endif
        {
            int cur = 0;
            ok(cudaGetDevice(&cur));
            ok(cudaSetDevice(dev));
            ok(cudaStreamCreateWithFlags(&d.h2d, cudaStreamNonBlocking));
            ok(cudaSetDevice(cur));
        }
        cudaMemPoolProps props{};
        props.allocType = cudaMemAllocationTypePinned;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        ok(cudaMemPoolCreate(&d.pool, &props));
        if (d.pool) {
            uint64_t keep = UINT64_MAX;                 // freed scratch rows stay in OUR pool for the next call
            ok(cudaMemPoolSetAttribute(d.pool, cudaMemPoolAttrReleaseThreshold, &keep));
        }
#else
        (void)dev;
        d.maxSmem = 227 * 1024;
#endif
    });
    if (d.err != cudaSuccess) return fail(VIORB_E_CUDA, std::string("per-device initialisation: ") + cudaGetErrorString(d.err));
    *out = &d;
    return VIORB_OK;
}

// ---- the extraction pipeline on n frames already on the device ------------------------------
// Frames [f0, f0 + n) of the current batch (ex->src); the kernels add g.frame0 to their frame index.
int run_pipeline(viorb_extractor* ex, int f0, int n, int lap0, int lap1) {
    ExtractGeom g = ex->g;
    g.frame0 = f0;
    const FrameSrc& fs = ex->src;
    cudaStream_t st = ex->stream;
    CU(cudaMemsetAsync(ex->dCandCount.p + (size_t)f0 * g.nlevels, 0, sizeof(int) * (size_t)n * g.nlevels, st));
    const bool tm = ex->timing;
    if (tm) CU(cudaEventRecord(ex->evX[0], st));
    for (int l = 1; l < g.nlevels; ++l) {
        const LevelGeom& L = g.lv[l];
        dim3 grid((L.w + 255) / 256, (L.h + 4 * VIORB_RS_ROWS - 1) / (4 * VIORB_RS_ROWS), n), block(64, 4);
        VIORB_LAUNCH(k_resize, grid, block, 0, st, g, fs, l, ex->dXofs.p, ex->dXalpha.p, ex->dYofs.p, ex->dYbeta.p);
    }
    if (tm) CU(cudaEventRecord(ex->evX[1], st));
    if (g.totalCells > 0) {
        VIORB_LAUNCH(k_fast, dim3(g.totalCells, n), dim3(VIORB_FAST_THREADS), ex->fastSmem, st, g, fs, ex->dCells.p,
                     ex->dCand.p, ex->dCandCount.p, ex->dStatus.p);
    }
    if (tm) CU(cudaEventRecord(ex->evX[2], st));
    // the dense jobs (few, long) run beside the light ones on the handle's side stream; with the stage timers on, in line
    {
        cudaStream_t sd = st;
        if (!tm && ex->sideOct) {
            CU(cudaEventRecord(ex->evOctFork, st));
            CU(cudaStreamWaitEvent(ex->sideOct, ex->evOctFork, 0));
            sd = ex->sideOct;
        }
        VIORB_LAUNCH(k_octree_dense, dim3(n, g.nlevels), dim3(VIORB_OCT_THREADS), ex->octSmemDense, sd, g, ex->dCand.p, ex->dCandCount.p,
                     ex->dKeyNode.p, ex->dKept.p, ex->dKeptCount.p, ex->dStatus.p);
        VIORB_LAUNCH(k_octree, dim3(n, g.nlevels), dim3(VIORB_OCT_LIGHT_THREADS), ex->octSmem, st, g, ex->dCand.p, ex->dCandCount.p,
                     ex->dKeyNode.p, ex->dKept.p, ex->dKeptCount.p, ex->dStatus.p);
        if (sd != st) {
            CU(cudaEventRecord(ex->evOctJoin, sd));
            CU(cudaStreamWaitEvent(st, ex->evOctJoin, 0));
        }
    }
    if (tm) CU(cudaEventRecord(ex->evX[3], st));
    VIORB_LAUNCH(k_orient, dim3((g.keptPerFrame + 8 * VIORB_ORIENT_KPW - 1) / (8 * VIORB_ORIENT_KPW), n), dim3(256), 0, st, g, fs, ex->dOrientTab.p, ex->dSlotLevel.p, ex->dKept.p,
                 ex->dKeptCount.p);
    if (tm) CU(cudaEventRecord(ex->evX[4], st));
    VIORB_LAUNCH(k_emit, dim3(n), dim3(256), 0, st, g, ex->dKept.p, ex->dKeptCount.p, lap0, lap1, ex->dKps.p,
                 ex->dSlot2Kept.p, ex->dNOut.p, ex->dMono.p, ex->cam, ex->dKpsUn.p);
    if (tm) CU(cudaEventRecord(ex->evX[5], st));
    VIORB_LAUNCH(k_blur, dim3(ex->totalBlurTiles, n), dim3(64), 0, st, g, fs, ex->blurTiles);
    if (tm) CU(cudaEventRecord(ex->evX[6], st));
    VIORB_LAUNCH(k_describe, dim3((g.outCap + 8 * VIORB_DESC_KPW - 1) / (8 * VIORB_DESC_KPW), n), dim3(256), 0, st, g, fs, ex->dKept.p, ex->dSlot2Kept.p,
                 ex->dNOut.p, ex->dDesc.p);
    if (tm) { CU(cudaEventRecord(ex->evX[7], st)); ex->haveX = true; }
    CU(cudaGetLastError());
    ex->lastN = f0 + n;
    return VIORB_OK;
}

// Host wait for a stream.  Single frames spin in the driver (lowest latency); batch calls, whose waits last
// milliseconds, sleep on a blocking-sync event so that several pipeline workers / ranks do not burn one host core each
// (on a 32-core host with 8 GPUs the spinning waits, not the GPUs, bounded the end-to-end rate).
#define VIORB_BLOCKING_BATCH 8
cudaError_t wait_stream(viorb_extractor* ex, cudaStream_t st) {
    if (ex->lastN < VIORB_BLOCKING_BATCH || !ex->evBlock) return cudaStreamSynchronize(st);
    cudaError_t e = cudaEventRecord(ex->evBlock, st);
    return e == cudaSuccess ? cudaEventSynchronize(ex->evBlock) : e;
}

int check_status(viorb_extractor* ex) {
    int st = 0;
    CU(cudaMemcpyAsync(&st, ex->dStatus.p, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(wait_stream(ex, ex->stream));
    if (st) {
        CU(cudaMemsetAsync(ex->dStatus.p, 0, sizeof(int), ex->stream));
        char buf[96];
        std::snprintf(buf, sizeof buf, "device arena overflow (status bits 0x%x)", st);
        return fail(VIORB_E_CAPACITY, buf);
    }
    return VIORB_OK;
}

int ensure_match_buffers(viorb_extractor* ex) {
    const size_t need = (size_t)ex->prm.max_batch * ex->capOut;
    if (ex->matchReady) return VIORB_OK;
    CU(ex->dBestIdx.alloc(need)); CU(ex->dBestDist.alloc(need)); CU(ex->dSad.alloc(need));
    CU(ex->dURight.alloc(need)); CU(ex->dDepth.alloc(need)); CU(ex->dLrMatch.alloc(need));
    CU(ex->dKnnIdx.alloc(need * 2)); CU(ex->dKnnDist.alloc(need * 2));
    CU(ex->dMatches.alloc(need));
    CU(ex->dNStereo.alloc(ex->prm.max_batch)); CU(ex->dNRatio.alloc(ex->prm.max_batch));
    // row table of the rectified stereo matcher: a keypoint of the coarsest level spans 2*ceil(2*scale)+3 rows at most
    const int span = 2 * (int)std::ceil(2.0f * ex->scale[ex->prm.nlevels - 1]) + 3;
    ex->rowListCap = ex->capOut * span;
    CU(ex->dRowStart.alloc((size_t)ex->prm.max_batch * (ex->prm.max_height + 2)));
    CU(ex->dRowList.alloc((size_t)ex->prm.max_batch * ex->rowListCap));
    ex->matchReady = true;           // only once EVERY buffer exists: a failed call is retried from the start
    return VIORB_OK;
}

bool is_pinned(const void* p) {
#ifdef VIORB_EMU
    (void)p;
    return false;
#else
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
#endif
}

bool same_geometry(const viorb_extractor* a, const viorb_extractor* b) {
    return a->g.W == b->g.W && a->g.H == b->g.H && a->g.nlevels == b->g.nlevels && a->g.outCap == b->g.outCap &&
           a->prm.scale_factor == b->prm.scale_factor && a->g.W > 0;
}

StereoParams make_stereo_params(const viorb_extractor* ex, float bf, float b) {
    StereoParams sp;
    std::memset(&sp, 0, sizeof sp);
    sp.bf = bf;
    sp.maxD = bf / b;                       // maxD = mbf / minZ, minZ = mb (frame.cpp:853-855)
    sp.nlevels = ex->g.nlevels;
    sp.outCap = ex->g.outCap;
    for (int l = 0; l < ex->g.nlevels; ++l) { sp.scale[l] = ex->scale[l]; sp.invScale[l] = ex->invScale[l]; }
    return sp;
}

int launch_stereo(viorb_extractor* L, viorb_extractor* R, int n, float bf, float b) {
    const StereoParams sp = make_stereo_params(L, bf, b);
    cudaStream_t st = L->stream;
    const int gx = (L->g.outCap + 7) / 8;
    const bool tm = L->timing;
    if (tm) CU(cudaEventRecord(L->evS[0], st));
    const int rows = L->g.H;
    VIORB_LAUNCH(k_stereo_rows, dim3(n), dim3(256), sizeof(int) * 2 * (rows + 1), st, sp, rows, L->rowListCap, R->dKps.p, R->dNOut.p,
                 L->dRowStart.p, L->dRowList.p, L->dStatus.p);
    VIORB_LAUNCH(k_stereo_coarse, dim3(gx, n), dim3(256), 0, st, sp, rows, L->rowListCap, L->dKps.p, L->dDesc.p, L->dNOut.p,
                 R->dDesc.p, L->dRowStart.p, L->dRowList.p, L->dBestIdx.p, L->dBestDist.p);
    if (tm) CU(cudaEventRecord(L->evS[1], st));
    VIORB_LAUNCH(k_stereo_refine, dim3(gx, n), dim3(256), 0, st, sp, L->g, L->src, R->src, L->dKps.p, L->dNOut.p, R->dKps.p,
                 L->dBestIdx.p, L->dBestDist.p, L->dURight.p, L->dDepth.p, L->dSad.p);
    if (tm) CU(cudaEventRecord(L->evS[2], st));
    VIORB_LAUNCH(k_stereo_cut, dim3(n), dim3(256), 0, st, sp.outCap, L->dNOut.p, L->dSad.p, L->dURight.p, L->dDepth.p,
                 L->dNStereo.p);
    if (tm) CU(cudaEventRecord(L->evS[3], st));
    CU(cudaGetLastError());
    return VIORB_OK;
}

// Work queued on L's stream from here on sees everything queued on R's stream so far (device-side
// ordering, the host does not wait) ...
int join_streams(viorb_extractor* L, viorb_extractor* R) {
    if (L->stream == R->stream) return VIORB_OK;
    CU(cudaEventRecord(R->evJoin, R->stream));
    CU(cudaStreamWaitEvent(L->stream, R->evJoin, 0));
    return VIORB_OK;
}
// ... and R's next work (its next extract overwrites the arenas the matcher reads) waits for what L queued.
int release_partner(viorb_extractor* L, viorb_extractor* R) {
    if (L->stream == R->stream) return VIORB_OK;
    CU(cudaEventRecord(L->evJoin, L->stream));
    CU(cudaStreamWaitEvent(R->stream, L->evJoin, 0));
    return VIORB_OK;
}

}  // namespace

// Brute-force kNN-2 of a batch of descriptor sets: k_knn2_f4 (tcgen05.mma kind::mxf4.block_scale, accumulators in tensor
// memory, kernels_knn_f4.cuh).  The product library holds this one kernel.  The earlier kernels (kind::i8 k_knn2_tc5,
// warp-level IMMA k_knn2_mma, LOP3/POPC k_knn2; identical results) are BUILD-time variants for
// experiments (-DVIORB_KNN_VARIANT=3 / 1 / 2, tools/variants.sh); the CPU emulation build of the tests uses the IMMA one.
static int launch_knn2(const DescSet& Q, const DescSet& T, int maxNq, int maxNt, int nSets, int32_t* dIdx, int32_t* dDist,
                       size_t outStride, cudaStream_t st) {
#if VIORB_KNN_TENSOR
    int dev = 0;
    CU(cudaGetDevice(&dev));
    DeviceState* ds = nullptr;
    if (int rc = device_state(dev, &ds)) return rc;
    const int qTiles = (maxNq + tc5::QT - 1) / tc5::QT, nBase = qTiles * nSets;
    if (nBase > 0) {
        const int nSm = ds->sms;
        // Few query tiles and a long train set (one stereo pair, the row-sharded sweep) leave SMs idle or a last
        // wave nearly empty: cut every query tile's train range over nSplit CTAs and merge the partial pairs.
        // Cost model in train tiles: waves x (tiles per item + ~2 tiles of per-item set-up).
        const int maxTiles = std::max(1, (maxNt + tc5::TT - 1) / tc5::TT);
        int nSplit = 1;
        long best = 0;
        for (int sp = 1; sp <= std::min(maxTiles, 32); ++sp) {
            const long cost = (long)((nBase * sp + nSm - 1) / nSm) * ((maxTiles + sp - 1) / sp + 2);
            if (sp == 1 || cost * 20 < best * 19) { best = cost; nSplit = sp; }      // a split must win by > 5 %
            if (sp == 1 && cost < 24) break;      // under ~20 us of kernel the merge launch costs more than it saves
        }
        if (const int forced = g_forceKnnSplit.load()) nSplit = std::max(1, std::min(forced, maxTiles));     // tests only
        const int nItems = nBase * nSplit;
        uint2* partial = nullptr;
        if (nSplit > 1) CU(cudaMallocFromPoolAsync((void**)&partial, (size_t)nItems * tc5::QT * sizeof(uint2), ds->pool, st));
        VIORB_LAUNCH(VIORB_KNN_TC_KERNEL, dim3(std::min(nItems, nSm)), dim3(tc5::THREADS), tc5::SMEM_BYTES, st, Q, T, dIdx, dDist,
                     outStride, qTiles, nSets, nItems, nSplit, partial);
        if (nSplit > 1) {
            VIORB_LAUNCH(tc5::k_knn2_tc5_merge, dim3((qTiles * tc5::QT + 255) / 256, nSets), dim3(256), 0, st, Q, partial, dIdx,
                         dDist, outStride, qTiles, nSplit);
            CU(cudaFreeAsync(partial, st));
        }
    }
#else
    const dim3 grid((maxNq + VIORB_KNN_THREADS - 1) / VIORB_KNN_THREADS, nSets), block(VIORB_KNN_THREADS);
    (void)maxNt;
#if VIORB_KNN_VARIANT == 2
    VIORB_LAUNCH(k_knn2, grid, block, 0, st, Q, T, dIdx, dDist, outStride);
#else
    VIORB_LAUNCH(k_knn2_mma, grid, block, 0, st, Q, T, dIdx, dDist, outStride);
#endif
#endif
    CU(cudaGetLastError());
    return VIORB_OK;
}

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

const char* viorb_last_error(void) { return g_err.c_str(); }
void viorb_debug_knn_split(int n_split) { g_forceKnnSplit.store(n_split < 0 ? 0 : n_split); }
const char* viorb_version(void) {
#ifdef VIORB_EMU
    return "viorb 0.1 (cpu emulation build: tests only)";
#else
    return "viorb 0.1 (sm_100a)";
#endif
}

int viorb_extractor_create(const viorb_params* params, viorb_extractor** out) {
    if (!params || !out) return fail(VIORB_E_ARG, "null argument");
    *out = nullptr;
    const viorb_params& p = *params;
    if (p.nlevels < 1 || p.nlevels > VIORB_MAX_LEVELS || p.nfeatures < 1 || !(p.scale_factor > 1.0f) || p.max_batch < 1 ||
        p.max_width < 1 || p.max_height < 1)
        return fail(VIORB_E_ARG, "invalid viorb_params");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= p.device)
        return fail(VIORB_E_NO_DEVICE, "no CUDA device: libviorb has no CPU path");
    CU(cudaSetDevice(p.device));
    DeviceState* ds = nullptr;
    if (int rc = device_state(p.device, &ds)) return rc;
    viorb_extractor* ex = new viorb_extractor();
    ex->prm = p;
    std::memset(&ex->g, 0, sizeof(ex->g));
    build_tables(ex);
    ResizeTabs tabs;
    int rc = build_geometry(ex, p.max_width, p.max_height, tabs);
    if (rc) { delete ex; return rc; }
    const ExtractGeom g = ex->g;
    const size_t B = (size_t)p.max_batch;
    ex->capArena = g.arenaStride + 4096;
    ex->capCand = (size_t)g.candPerFrame + 1024;
    ex->capKept = (size_t)g.keptPerFrame + 64;
    ex->capOut = g.outCap + 64;
    ex->capCells = ex->cells.size() + 256;
    ex->capXTab = tabs.xofs.size() + 64 * VIORB_MAX_LEVELS;
    ex->capYTab = tabs.yofs.size() + 64 * VIORB_MAX_LEVELS;
    ex->inPitch = round_up(p.max_width, 64);
    cudaError_t e = cudaStreamCreateWithFlags(&ex->stream, cudaStreamNonBlocking);
    ex->ownStream = true;
    auto ok = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    ok(ex->dIn.alloc(B * ex->inPitch * p.max_height));
    ok(ex->dPyr.alloc(B * ex->capArena));
    ok(ex->dBlur.alloc(B * ex->capArena));
    ok(ex->dCand.alloc(B * ex->capCand));
    ok(ex->dKeyNode.alloc(B * ex->capCand));
    ok(ex->dCandCount.alloc(B * VIORB_MAX_LEVELS));
    ok(ex->dKeptCount.alloc(B * VIORB_MAX_LEVELS));
    ok(ex->dKept.alloc(B * ex->capKept));
    ok(ex->dKps.alloc(B * ex->capOut));
    ok(ex->dDesc.alloc(B * ex->capOut * 32));
    ok(ex->dSlot2Kept.alloc(B * ex->capOut));
    ok(ex->dNOut.alloc(B));
    ok(ex->dMono.alloc(B));
    ok(ex->dStatus.alloc(1));
    ok(ex->dCells.alloc(ex->capCells));
    ok(ex->dOrientTab.alloc(256));
    ok(ex->dSlotLevel.alloc(ex->capKept));
    ok(ex->dXofs.alloc(ex->capXTab));
    ok(ex->dXalpha.alloc(ex->capXTab));
    ok(ex->dYofs.alloc(ex->capYTab));
    ok(ex->dYbeta.alloc(ex->capYTab));
    ok(ex->hIn.alloc(B * ex->inPitch * p.max_height));
    ok(ex->hKps.alloc(B * ex->capOut));
    ok(ex->hDesc.alloc(B * ex->capOut * 32));
    ok(ex->hCounts.alloc(B * 2 + 8));
    ok(cudaStreamCreateWithFlags(&ex->cpStream, cudaStreamNonBlocking));
    ok(cudaStreamCreateWithFlags(&ex->sideOct, cudaStreamNonBlocking));
    ok(cudaEventCreateWithFlags(&ex->evOctFork, cudaEventDisableTiming));
    ok(cudaEventCreateWithFlags(&ex->evOctJoin, cudaEventDisableTiming));
    ok(cudaStreamCreateWithFlags(&ex->outStream, cudaStreamNonBlocking));
    for (int i = 0; i < VIORB_MAX_CHUNKS; ++i) {
        ok(cudaEventCreateWithFlags(&ex->evIn[i], cudaEventDisableTiming));
        ok(cudaEventCreateWithFlags(&ex->evDone[i], cudaEventDisableTiming));
        ok(cudaEventCreateWithFlags(&ex->evCnt[i], cudaEventDisableTiming | cudaEventBlockingSync));
    }
    ok(cudaEventCreateWithFlags(&ex->evEntry, cudaEventDisableTiming));
    ok(cudaEventCreateWithFlags(&ex->evJoin, cudaEventDisableTiming));
    ok(cudaEventCreateWithFlags(&ex->evBlock, cudaEventDisableTiming | cudaEventBlockingSync));
    for (int i = 0; i < 8; ++i) ok(cudaEventCreate(&ex->evX[i]));
    for (int i = 0; i < 6; ++i) ok(cudaEventCreate(&ex->evS[i]));
    if (e == cudaSuccess) e = cudaMemset(ex->dStatus.p, 0, sizeof(int));
    if (e != cudaSuccess) {
        std::string msg = std::string("allocation failed: ") + cudaGetErrorString(e);
        viorb_extractor_destroy(ex);
        return fail(VIORB_E_CUDA, msg);
    }
    ex->g.W = ex->g.H = 0;      // force the geometry upload on first use
    *out = ex;
    return VIORB_OK;
}

void viorb_extractor_destroy(viorb_extractor* ex) {
    if (!ex) return;
    cudaSetDevice(ex->prm.device);
    if (ex->stream || ex->ownStream) cudaStreamSynchronize(ex->stream);
    ex->dIn.release(); ex->dPyr.release(); ex->dBlur.release(); ex->dDesc.release(); ex->dCand.release();
    ex->dKeyNode.release(); ex->dCandCount.release(); ex->dKeptCount.release(); ex->dSlot2Kept.release();
    ex->dNOut.release(); ex->dMono.release(); ex->dStatus.release(); ex->dXofs.release(); ex->dYofs.release();
    ex->dXalpha.release(); ex->dYbeta.release(); ex->dKept.release(); ex->dKps.release(); ex->dKpsUn.release(); ex->bow.release(); ex->dCells.release();
    ex->dOrientTab.release(); ex->dSlotLevel.release();
    ex->dBestIdx.release(); ex->dBestDist.release(); ex->dSad.release(); ex->dKnnIdx.release(); ex->dKnnDist.release();
    ex->dLrMatch.release(); ex->dURight.release(); ex->dDepth.release(); ex->dNStereo.release(); ex->dNRatio.release();
    ex->dMatches.release(); ex->dRowStart.release(); ex->dRowList.release();
    ex->hIn.release(); ex->hDesc.release(); ex->hKps.release(); ex->hCounts.release();
    if (ex->f2f) { ex->f2f->release(); delete ex->f2f; }
    for (int i = 0; i < 8; ++i) if (ex->evX[i]) cudaEventDestroy(ex->evX[i]);
    for (int i = 0; i < 6; ++i) if (ex->evS[i]) cudaEventDestroy(ex->evS[i]);
    for (int i = 0; i < VIORB_MAX_CHUNKS; ++i) {
        if (ex->evIn[i]) cudaEventDestroy(ex->evIn[i]);
        if (ex->evDone[i]) cudaEventDestroy(ex->evDone[i]);
        if (ex->evCnt[i]) cudaEventDestroy(ex->evCnt[i]);
    }
    if (ex->evEntry) cudaEventDestroy(ex->evEntry);
    if (ex->evJoin) cudaEventDestroy(ex->evJoin);
    if (ex->evBlock) cudaEventDestroy(ex->evBlock);
    if (ex->cpStream) { cudaStreamSynchronize(ex->cpStream); cudaStreamDestroy(ex->cpStream); }
    if (ex->sideOct) { cudaStreamSynchronize(ex->sideOct); cudaStreamDestroy(ex->sideOct); }
    if (ex->evOctFork) cudaEventDestroy(ex->evOctFork);
    if (ex->evOctJoin) cudaEventDestroy(ex->evOctJoin);
    if (ex->outStream) { cudaStreamSynchronize(ex->outStream); cudaStreamDestroy(ex->outStream); }
    if (ex->ownStream && ex->stream) cudaStreamDestroy(ex->stream);
    delete ex;
}

int viorb_extractor_tables(const viorb_extractor* ex, int32_t* nlevels, float* scale, float* inv_scale, float* sigma2,
                           float* inv_sigma2, int32_t* features_per_level) {
    if (!ex) return fail(VIORB_E_ARG, "null handle");
    const int nl = ex->prm.nlevels;
    if (nlevels) *nlevels = nl;
    for (int l = 0; l < nl; ++l) {
        if (scale) scale[l] = ex->scale[l];
        if (inv_scale) inv_scale[l] = ex->invScale[l];
        if (sigma2) sigma2[l] = ex->sigma2[l];
        if (inv_sigma2) inv_sigma2[l] = ex->invSigma2[l];
        if (features_per_level) features_per_level[l] = ex->featPerLevel[l];
    }
    return VIORB_OK;
}

int viorb_extractor_capacity(const viorb_extractor* ex) { return ex ? ex->capOut : 0; }

/* Page-locked host memory for images / results (what a camera driver or a caller that wants zero staging
 * copies should hand to viorb_extract*).  Thin wrappers over cudaMallocHost / cudaFreeHost. */
void* viorb_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { fail(VIORB_E_CUDA, "cudaMallocHost failed"); return nullptr; }
    return p;
}
void viorb_host_free(void* p) { if (p) cudaFreeHost(p); }

void* viorb_extractor_stream(viorb_extractor* ex) { return ex ? (void*)ex->stream : nullptr; }

/* Run this handle's work on a caller-owned stream (e.g. the framework's current stream). */
int viorb_extractor_set_stream(viorb_extractor* ex, void* stream) {
    if (!ex) return fail(VIORB_E_ARG, "null handle");
    CU(cudaSetDevice(ex->prm.device));
    CU(cudaStreamSynchronize(ex->stream));
    if (ex->ownStream && ex->stream) cudaStreamDestroy(ex->stream);
    ex->stream = (cudaStream_t)stream;
    ex->ownStream = false;
    return VIORB_OK;
}

int viorb_extractor_set_timing(viorb_extractor* ex, int enable) {
    if (!ex) return fail(VIORB_E_ARG, "null handle");
    ex->timing = enable != 0;
    ex->haveX = ex->haveS = false;
    return VIORB_OK;
}

/* ms[7]: resize, fast, octree, orient, emit, blur, describe of the last extract (waits for it). */
int viorb_extractor_stage_ms(viorb_extractor* ex, float* ms) {
    if (!ex || !ms) return fail(VIORB_E_ARG, "null argument");
    if (!ex->haveX) return fail(VIORB_E_ARG, "no timed extract: call viorb_extractor_set_timing(ex, 1) first");
    CU(cudaSetDevice(ex->prm.device));
    CU(cudaEventSynchronize(ex->evX[7]));
    for (int i = 0; i < 7; ++i) CU(cudaEventElapsedTime(&ms[i], ex->evX[i], ex->evX[i + 1]));
    return VIORB_OK;
}

/* ms[5]: stereo coarse, refine, cut, knn2, ratio of the last viorb_stereo_batch on this (left) handle. */
int viorb_stereo_stage_ms(viorb_extractor* ex, float* ms) {
    if (!ex || !ms) return fail(VIORB_E_ARG, "null argument");
    if (!ex->haveS) return fail(VIORB_E_ARG, "no timed stereo batch");
    CU(cudaSetDevice(ex->prm.device));
    CU(cudaEventSynchronize(ex->evS[5]));
    for (int i = 0; i < 5; ++i) CU(cudaEventElapsedTime(&ms[i], ex->evS[i], ex->evS[i + 1]));
    return VIORB_OK;
}

int viorb_extractor_sync(viorb_extractor* ex) {
    if (!ex) return fail(VIORB_E_ARG, "null handle");
    CU(cudaSetDevice(ex->prm.device));
    CU(cudaStreamSynchronize(ex->stream));
    return VIORB_OK;
}

int viorb_extract_device(viorb_extractor* ex, const uint8_t* d_imgs, int n, int width, int height, int step,
                         size_t frame_stride, int lap0, int lap1) {
    if (!ex || !d_imgs) return fail(VIORB_E_ARG, "null argument");
    if (width <= 0 || height <= 0 || n <= 0) return fail(VIORB_E_EMPTY, "empty image");
    if (n > ex->prm.max_batch) return fail(VIORB_E_CAPACITY, "batch larger than max_batch");
    if (step % 16 != 0 || ((uintptr_t)d_imgs & 15) || frame_stride % 16 != 0 || step < width)
        return fail(VIORB_E_ARG, "device images need a 16-byte aligned base, step and frame stride");
    CU(cudaSetDevice(ex->prm.device));
    int rc = set_size(ex, width, height);
    if (rc) return rc;
    ex->src.l0 = d_imgs;
    ex->src.l0Pitch = step;
    ex->src.l0Stride = frame_stride;
    ex->src.pyr = ex->dPyr.p;
    ex->src.blur = ex->dBlur.p;
    ex->lastMaxCount = 0;            // counts stay on the device
    return run_pipeline(ex, 0, n, lap0, lap1);
}

static std::atomic<int> g_batchCalls{0};      // viorb_extract_batch calls in flight in this process

int viorb_extract_batch(viorb_extractor* ex, const uint8_t* const* imgs, int n, int width, int height, int step, int lap0,
                        int lap1, viorb_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index_out) {
    if (!ex) return fail(VIORB_E_ARG, "null handle");
    if (!imgs || n <= 0 || width <= 0 || height <= 0) {
        if (mono_index_out) for (int i = 0; i < std::max(n, 1); ++i) mono_index_out[i] = -1;
        if (n_out) for (int i = 0; i < std::max(n, 1); ++i) n_out[i] = 0;
        return fail(VIORB_E_EMPTY, "empty image");
    }
    if (n > ex->prm.max_batch) return fail(VIORB_E_CAPACITY, "batch larger than max_batch");
    if (width > ex->prm.max_width || height > ex->prm.max_height)
        return fail(VIORB_E_CAPACITY, "image larger than max_width x max_height of this handle");
    CU(cudaSetDevice(ex->prm.device));
    const int pitch = ex->inPitch;
    const size_t fstride = (size_t)pitch * height;
    for (int i = 0; i < n; ++i)
        if (!imgs[i]) return fail(VIORB_E_EMPTY, "empty image");
    int rc = set_size(ex, width, height);
    if (rc) return rc;
    ex->src.l0 = ex->dIn.p;
    ex->src.l0Pitch = pitch;
    ex->src.l0Stride = fstride;
    ex->src.pyr = ex->dPyr.p;
    ex->src.blur = ex->dBlur.p;
    // Page-locked inputs (viorb_host_alloc / cudaHostRegister'ed camera buffers) are DMA'd straight from
    // the caller's memory; pageable inputs are first packed into the handle's pinned staging area.
    bool pinnedIn = true;
    for (int i = 0; i < n && pinnedIn; i += std::max(1, n - 1)) pinnedIn = is_pinned(imgs[i]) && is_pinned(imgs[i] + (size_t)(height - 1) * step);
    const size_t oc = ex->g.outCap;
    const size_t rowKp = std::min<size_t>(oc, cap);
    const bool pinnedKps = kps && is_pinned(kps) && is_pinned(kps + (size_t)n * cap - 1);
    const bool pinnedDesc = desc && is_pinned(desc) && is_pinned(desc + (size_t)n * cap * 32 - 1);
    // The batch is cut into chunks: while the kernels of chunk c run on the handle's stream, the images of
    // chunk c+1 travel host -> device on cpStream and the results of chunk c-1 device -> host on outStream
    // (the three engines of the PCIe / copy path work concurrently).  With stage timing on, one chunk.
    // 80 frames per chunk: measured with three pipeline workers (tools/e2e_chunks_cmp.sh, 160 frames per call):
    // 1 / 2 / 3 / 4 / 8 chunks = 39.2k / 40.2k / 37.6k / 37.4k / 34.6k stereo frames/s - smaller chunks leave the
    // level-synchronous kernels (k_octree, upper pyramid levels) with less than one wave of CTAs.
    // With several batch calls in flight (pipeline workers, each with its own handles: more than one left / right pair)
    // the calls' copies already run under each other's kernels and whole-batch launches are the better use of the GPU:
    // 1 / 2 / 4 chunks = 39.8k / 38.3k / 34.6k with three workers and the pointer tables prebuilt.
    struct InFlight {
        InFlight() { now = ++g_batchCalls; }
        ~InFlight() { --g_batchCalls; }
        int now;
    } inflight;
    int nChunks = (ex->timing || inflight.now > 2) ? 1 : std::min(VIORB_MAX_CHUNKS, std::max(1, (n + 40) / 80));
    if (const char* e = std::getenv("VIORB_CHUNKS")) nChunks = std::min(VIORB_MAX_CHUNKS, std::max(1, std::atoi(e)));   // tuning knob
    nChunks = std::min(nChunks, n);
    const int per = (n + nChunks - 1) / nChunks;
    nChunks = (n + per - 1) / per;
    // Uploads go through the device's shared copy stream (DeviceState::h2d) when nothing is pending on this handle -
    // the normal case of the synchronous calls; otherwise through the handle's own copy stream behind the pending work,
    // so that one busy handle never blocks the other handles' uploads.
    DeviceState* ds = nullptr;
    if (int drc = device_state(ex->prm.device, &ds)) return drc;
    const bool sharedUp = ds->h2d && !ex->timing && cudaStreamQuery(ex->stream) == cudaSuccess;
    cudaStream_t up = sharedUp ? ds->h2d : ex->cpStream;
    CU(cudaEventRecord(ex->evEntry, ex->stream));              // earlier work on the handle may still read the arenas
    if (!sharedUp) CU(cudaStreamWaitEvent(ex->cpStream, ex->evEntry, 0));
    CU(cudaStreamWaitEvent(ex->outStream, ex->evEntry, 0));
    for (int c = 0; c < nChunks; ++c) {
        const int f0 = c * per, m = std::min(per, n - f0);
        if (!pinnedIn)
            for (int i = f0; i < f0 + m; ++i)
                for (int y = 0; y < height; ++y)
                    std::memcpy(ex->hIn.p + i * fstride + (size_t)y * pitch, imgs[i] + (size_t)y * step, width);
        std::unique_lock<std::mutex> upLock(ds->h2dMu, std::defer_lock);
        if (sharedUp) upLock.lock();
        if (pinnedIn) {
            bool contiguous = true;
            for (int i = f0 + 1; i < f0 + m && contiguous; ++i) contiguous = imgs[i] == imgs[i - 1] + (size_t)step * height;
            if (contiguous && step == pitch) {
                // same row stride on both sides: one linear DMA (the bytes between width and step of a row belong
                // to the caller's rows; the copy stops after the last pixel of the last row)
                CU(cudaMemcpyAsync(ex->dIn.p + f0 * fstride, imgs[f0], ((size_t)height * m - 1) * step + width, cudaMemcpyHostToDevice, up));
            } else if (contiguous) {
                CU(cudaMemcpy2DAsync(ex->dIn.p + f0 * fstride, pitch, imgs[f0], step, width, (size_t)height * m, cudaMemcpyHostToDevice, up));
            } else {
                for (int i = f0; i < f0 + m; ++i)
                    CU(cudaMemcpy2DAsync(ex->dIn.p + i * fstride, pitch, imgs[i], step, width, height, cudaMemcpyHostToDevice, up));
            }
        } else {
            CU(cudaMemcpyAsync(ex->dIn.p + f0 * fstride, ex->hIn.p + f0 * fstride, fstride * m, cudaMemcpyHostToDevice, up));
        }
        CU(cudaEventRecord(ex->evIn[c], up));
        if (sharedUp) upLock.unlock();
        CU(cudaStreamWaitEvent(ex->stream, ex->evIn[c], 0));
        rc = run_pipeline(ex, f0, m, lap0, lap1);
        if (rc) return rc;
        CU(cudaEventRecord(ex->evDone[c], ex->stream));
        CU(cudaStreamWaitEvent(ex->outStream, ex->evDone[c], 0));
        // the chunk's counts come down first: the result copies below move only the rows that exist
        CU(cudaMemcpyAsync(ex->hCounts.p + f0, ex->dNOut.p + f0, sizeof(int) * m, cudaMemcpyDeviceToHost, ex->outStream));
        CU(cudaMemcpyAsync(ex->hCounts.p + n + f0, ex->dMono.p + f0, sizeof(int) * m, cudaMemcpyDeviceToHost, ex->outStream));
        CU(cudaEventRecord(ex->evCnt[c], ex->outStream));
    }
    // Device -> host of the results, chunk by chunk as the counts arrive (every chunk's uploads and kernels are already
    // queued, so the device stays busy while the host waits here).  Rows past the largest count of a chunk are never
    // copied: the arenas hold viorb_extractor_capacity() slots per frame, a frame fills nfeatures .. nfeatures + ~30.
    for (int c = 0; c < nChunks; ++c) {
        const int f0 = c * per, m = std::min(per, n - f0);
        if (n >= VIORB_BLOCKING_BATCH) CU(cudaEventSynchronize(ex->evCnt[c]));
        else { cudaError_t q; while ((q = cudaEventQuery(ex->evCnt[c])) == cudaErrorNotReady) {} CU(q); }
        int maxCnt = 0;
        for (int i = f0; i < f0 + m; ++i) maxCnt = std::max(maxCnt, ex->hCounts.p[i]);
        const size_t rows = std::min<size_t>(std::min<size_t>((size_t)maxCnt, oc), rowKp);
        if (rows == 0) continue;
        cudaStream_t os = ex->outStream;
        if (kps) {
            viorb_keypoint* dst = pinnedKps ? kps + (size_t)f0 * cap : ex->hKps.p + f0 * oc;
            const size_t dpitch = sizeof(viorb_keypoint) * (pinnedKps ? (size_t)cap : oc);
            CU(cudaMemcpy2DAsync(dst, dpitch, ex->dKps.p + f0 * oc, sizeof(viorb_keypoint) * oc, sizeof(viorb_keypoint) * rows, m,
                                 cudaMemcpyDeviceToHost, os));
        }
        if (desc) {
            uint8_t* dst = pinnedDesc ? desc + (size_t)f0 * cap * 32 : ex->hDesc.p + f0 * oc * 32;
            const size_t dpitch = 32 * (pinnedDesc ? (size_t)cap : oc);
            CU(cudaMemcpy2DAsync(dst, dpitch, ex->dDesc.p + f0 * oc * 32, 32 * oc, 32 * rows, m, cudaMemcpyDeviceToHost, os));
        }
    }
    CU(wait_stream(ex, ex->outStream));
    rc = check_status(ex);          // synchronises the handle's stream
    if (rc) return rc;
    int worst = VIORB_OK;
    ex->lastMaxCount = 0;
    for (int i = 0; i < n; ++i) {
        const int cnt = ex->hCounts.p[i];
        ex->lastMaxCount = std::max(ex->lastMaxCount, cnt);
        if (n_out) n_out[i] = cnt;
        if (mono_index_out) mono_index_out[i] = ex->hCounts.p[n + i];
        const int m = std::min(cnt, cap);
        if (cnt > cap) worst = VIORB_E_CAPACITY;
        if (kps && !pinnedKps && m > 0) std::memcpy(kps + (size_t)i * cap, ex->hKps.p + i * oc, sizeof(viorb_keypoint) * m);
        if (desc && !pinnedDesc && m > 0) std::memcpy(desc + (size_t)i * cap * 32, ex->hDesc.p + i * oc * 32, (size_t)32 * m);
    }
    if (worst) return fail(worst, "output buffer smaller than the number of keypoints (n_out holds the count)");
    return VIORB_OK;
}

int viorb_extract(viorb_extractor* ex, const uint8_t* img, int width, int height, int step, int lap0, int lap1,
                  viorb_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index_out) {
    const uint8_t* one[1] = {img};
    if (!img) {
        if (n_out) *n_out = 0;
        if (mono_index_out) *mono_index_out = -1;
        return fail(VIORB_E_EMPTY, "empty image");
    }
    return viorb_extract_batch(ex, one, 1, width, height, step, lap0, lap1, kps, desc, cap, n_out, mono_index_out);
}

int viorb_fetch(viorb_extractor* ex, int frame, viorb_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index_out) {
    if (!ex || frame < 0 || frame >= ex->lastN) return fail(VIORB_E_ARG, "bad handle or frame index");
    CU(cudaSetDevice(ex->prm.device));
    int rc = check_status(ex);
    if (rc) return rc;
    int cnt[2];
    CU(cudaMemcpyAsync(&cnt[0], ex->dNOut.p + frame, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaMemcpyAsync(&cnt[1], ex->dMono.p + frame, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    if (n_out) *n_out = cnt[0];
    if (mono_index_out) *mono_index_out = cnt[1];
    const int m = std::min(cnt[0], cap);
    const size_t oc = ex->g.outCap;
    if (kps && m > 0) CU(cudaMemcpyAsync(kps, ex->dKps.p + frame * oc, sizeof(viorb_keypoint) * m, cudaMemcpyDeviceToHost, ex->stream));
    if (desc && m > 0) CU(cudaMemcpyAsync(desc, ex->dDesc.p + frame * oc * 32, (size_t)32 * m, cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    if (cnt[0] > cap && (kps || desc)) return fail(VIORB_E_CAPACITY, "output buffer smaller than the number of keypoints");
    return VIORB_OK;
}

int viorb_pyramid_level(viorb_extractor* ex, int frame, int level, uint8_t* dst, int dst_step, int* width, int* height) {
    if (!ex || frame < 0 || frame >= ex->lastN || level < 0 || level >= ex->g.nlevels)
        return fail(VIORB_E_ARG, "bad handle, frame or level");
    CU(cudaSetDevice(ex->prm.device));
    const LevelGeom& L = ex->g.lv[level];
    if (width) *width = L.w;
    if (height) *height = L.h;
    if (!dst) return VIORB_OK;
    const uint8_t* p;
    int pitch;
    if (level == 0) { p = ex->src.l0 + (size_t)frame * ex->src.l0Stride; pitch = ex->src.l0Pitch; }
    else { p = ex->src.pyr + (size_t)frame * ex->g.arenaStride + L.off; pitch = L.pitch; }
    CU(cudaMemcpy2DAsync(dst, dst_step, p, pitch, L.w, L.h, cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    return VIORB_OK;
}

int viorb_debug_blurred(viorb_extractor* ex, int frame, int level, uint8_t* dst, int dst_step) {
    if (!ex || frame < 0 || frame >= ex->lastN || level < 0 || level >= ex->g.nlevels || !dst)
        return fail(VIORB_E_ARG, "bad handle, frame or level");
    CU(cudaSetDevice(ex->prm.device));
    const LevelGeom& L = ex->g.lv[level];
    CU(cudaMemcpy2DAsync(dst, dst_step, ex->src.blur + (size_t)frame * ex->g.arenaStride + L.off, L.pitch, L.w, L.h,
                         cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    return VIORB_OK;
}

int viorb_debug_candidates(viorb_extractor* ex, int frame, int level, int32_t* xys, int cap, int* n_out) {
    if (!ex || frame < 0 || frame >= ex->lastN || level < 0 || level >= ex->g.nlevels)
        return fail(VIORB_E_ARG, "bad handle, frame or level");
    CU(cudaSetDevice(ex->prm.device));
    const LevelGeom& L = ex->g.lv[level];
    int n = 0;
    CU(cudaMemcpyAsync(&n, ex->dCandCount.p + frame * ex->g.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    n = std::min(n, L.candCap);
    if (n_out) *n_out = n;
    if (!xys || n == 0) return VIORB_OK;
    std::vector<uint32_t> raw(n);
    CU(cudaMemcpyAsync(raw.data(), ex->dCand.p + (size_t)frame * ex->g.candPerFrame + L.candOff, sizeof(uint32_t) * n,
                       cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    // presentation only: put the (unordered) device list into the reference's cell-major raster order
    struct E { long long key; uint32_t e; };
    std::vector<E> v(n);
    for (int i = 0; i < n; ++i) {
        const int x = raw[i] & 0xFFF, y = (raw[i] >> 12) & 0xFFF;
        const long long cj = std::min(std::max(x - 3, 0) / L.wCell, L.lastCol), ci = std::min(std::max(y - 3, 0) / L.hCell, L.lastRow);
        v[i] = {(ci << 32) | (cj << 24) | ((long long)y << 12) | x, raw[i]};
    }
    std::sort(v.begin(), v.end(), [](const E& a, const E& b) { return a.key < b.key; });
    for (int i = 0; i < std::min(n, cap); ++i) {
        xys[3 * i] = v[i].e & 0xFFF; xys[3 * i + 1] = (v[i].e >> 12) & 0xFFF; xys[3 * i + 2] = v[i].e >> 24;
    }
    return VIORB_OK;
}

int viorb_debug_kept(viorb_extractor* ex, int frame, int level, int32_t* xys, float* angle, int cap, int* n_out) {
    if (!ex || frame < 0 || frame >= ex->lastN || level < 0 || level >= ex->g.nlevels)
        return fail(VIORB_E_ARG, "bad handle, frame or level");
    CU(cudaSetDevice(ex->prm.device));
    const LevelGeom& L = ex->g.lv[level];
    int n = 0;
    CU(cudaMemcpyAsync(&n, ex->dKeptCount.p + frame * ex->g.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    if (n_out) *n_out = n;
    if (n == 0) return VIORB_OK;
    std::vector<KeptKp> raw(n);
    CU(cudaMemcpyAsync(raw.data(), ex->dKept.p + (size_t)frame * ex->g.keptPerFrame + L.keptOff, sizeof(KeptKp) * n,
                       cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    for (int i = 0; i < std::min(n, cap); ++i) {
        if (xys) { xys[3 * i] = raw[i].x; xys[3 * i + 1] = raw[i].y; xys[3 * i + 2] = raw[i].score & 0xFFFF; }
        if (angle) angle[i] = raw[i].angle;
    }
    return VIORB_OK;
}

// ---- matching ------------------------------------------------------------------------------------

int viorb_knn2_device(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int32_t* d_idx, int32_t* d_dist, void* stream) {
    if (nq < 0 || nt < 0 || (nq > 0 && (!d_q || !d_idx || !d_dist))) return fail(VIORB_E_ARG, "bad argument");
    if (nt >= (1 << VIORB_KNN_IDX_BITS)) return fail(VIORB_E_ARG, "train set larger than 2^22 rows");
    if (((uintptr_t)d_q & 15) || ((uintptr_t)d_t & 15)) return fail(VIORB_E_ARG, "descriptor buffers must be 16-byte aligned");
    if (nq == 0) return VIORB_OK;
    DescSet Q{d_q, 0, nullptr, nq}, T{d_t, 0, nullptr, nt};
    return launch_knn2(Q, T, nq, nt, 1, d_idx, d_dist, (size_t)nq, (cudaStream_t)stream);
}

int viorb_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist, int device) {
    if (nq < 0 || nt < 0 || (nq > 0 && (!q || !idx || !dist)) || (nt > 0 && !t)) return fail(VIORB_E_ARG, "bad argument");
    if (nq == 0) return VIORB_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device) return fail(VIORB_E_NO_DEVICE, "no CUDA device: libviorb has no CPU path");
    CU(cudaSetDevice(device));
    DevBuf<uint8_t> dq, dt;
    DevBuf<int32_t> di, dd;
    int rc = VIORB_OK;
    cudaError_t e = dq.alloc((size_t)nq * 32);
    if (e == cudaSuccess) e = dt.alloc((size_t)std::max(nt, 1) * 32);
    if (e == cudaSuccess) e = di.alloc((size_t)nq * 2);
    if (e == cudaSuccess) e = dd.alloc((size_t)nq * 2);
    if (e == cudaSuccess) e = cudaMemcpy(dq.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && nt > 0) e = cudaMemcpy(dt.p, t, (size_t)nt * 32, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc) rc = viorb_knn2_device(dq.p, nq, dt.p, nt, di.p, dd.p, nullptr);
    if (!rc) {
        e = cudaMemcpy(idx, di.p, sizeof(int32_t) * 2 * nq, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(dist, dd.p, sizeof(int32_t) * 2 * nq, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
    }
    dq.release(); dt.release(); di.release(); dd.release();
    return rc;
}

int viorb_descriptor_distance(const uint8_t* a, const uint8_t* b, int n, int32_t* dist, int device) {
    if (n < 0 || (n > 0 && (!a || !b || !dist))) return fail(VIORB_E_ARG, "bad argument");
    if (n == 0) return VIORB_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device) return fail(VIORB_E_NO_DEVICE, "no CUDA device: libviorb has no CPU path");
    CU(cudaSetDevice(device));
    DevBuf<uint8_t> da, db;
    DevBuf<int32_t> dd;
    cudaError_t e = da.alloc((size_t)n * 32);
    if (e == cudaSuccess) e = db.alloc((size_t)n * 32);
    if (e == cudaSuccess) e = dd.alloc(n);
    if (e == cudaSuccess) e = cudaMemcpy(da.p, a, (size_t)n * 32, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(db.p, b, (size_t)n * 32, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        VIORB_LAUNCH(k_desc_dist, dim3((n + 255) / 256), dim3(256), 0, (cudaStream_t)0, da.p, db.p, n, dd.p);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(dist, dd.p, sizeof(int32_t) * n, cudaMemcpyDeviceToHost);
    da.release(); db.release(); dd.release();
    if (e != cudaSuccess) return fail(VIORB_E_CUDA, cudaGetErrorString(e));
    return VIORB_OK;
}

// kNN-2 + ratio (+ dedup) on device buffers of one set; result rows to host.
static int ratio_on_device(const uint8_t* dq, int nq, const uint8_t* dt, int nt, double ratio, int rule, int dedup,
                           viorb_match* out, int cap, int* n_out, cudaStream_t st) {
    if (n_out) *n_out = 0;
    if (nq == 0) return VIORB_OK;
    DevBuf<int32_t> di, dd;
    DevBuf<viorb_match> dm, dm2;
    DevBuf<int> dn, dFirst;
    cudaError_t e = di.alloc((size_t)nq * 2);
    if (e == cudaSuccess) e = dd.alloc((size_t)nq * 2);
    if (e == cudaSuccess) e = dm.alloc(nq);
    if (e == cudaSuccess) e = dn.alloc(2);
    int rc = VIORB_OK;
    if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc) rc = viorb_knn2_device(dq, nq, dt, nt, di.p, dd.p, (void*)st);
    int cnt = 0;
    if (!rc) {
        DescSet Q{dq, 0, nullptr, nq};
        VIORB_LAUNCH(k_ratio, dim3(1), dim3(256), 0, st, Q, di.p, dd.p, (size_t)nq, ratio, rule, dm.p, (size_t)nq, dn.p,
                     (int32_t*)nullptr);
        e = cudaMemcpyAsync(&cnt, dn.p, sizeof(int), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
    }
    const viorb_match* res = dm.p;
    if (!rc && dedup && cnt > 0) {
        e = dFirst.alloc(std::max(nt, 1));
        if (e == cudaSuccess) e = dm2.alloc(cnt);
        if (e == cudaSuccess) e = cudaMemsetAsync(dFirst.p, 0x7f, sizeof(int) * (size_t)nt, st);   // 0x7f7f7f7f > any index
        if (e == cudaSuccess) {
            // sentinel must be exactly INT_MAX for k_dedup_emit: overwrite with a fill kernel-free trick
            std::vector<int> init(nt, 0x7fffffff);
            e = cudaMemcpyAsync(dFirst.p, init.data(), sizeof(int) * (size_t)nt, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
        if (e == cudaSuccess) {
            VIORB_LAUNCH(k_dedup_mark, dim3((cnt + 255) / 256), dim3(256), 0, st, dm.p, cnt, dFirst.p);
            VIORB_LAUNCH(k_dedup_emit, dim3(1), dim3(256), 0, st, dFirst.p, nt, dd.p, dm2.p, dn.p + 1);
            e = cudaMemcpyAsync(&cnt, dn.p + 1, sizeof(int), cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        res = dm2.p;
    }
    if (!rc) {
        if (n_out) *n_out = cnt;
        const int m = std::min(cnt, cap);
        if (out && m > 0) {
            e = cudaMemcpyAsync(out, res, sizeof(viorb_match) * m, cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        }
        if (!rc && cnt > cap && out) rc = fail(VIORB_E_CAPACITY, "match buffer too small (n_out holds the count)");
    }
    di.release(); dd.release(); dm.release(); dm2.release(); dn.release(); dFirst.release();
    return rc;
}

int viorb_match_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio, int rule, int dedup, viorb_match* out,
                      int cap, int* n_out, int device) {
    if (nq < 0 || nt < 0 || (nq > 0 && !q) || (nt > 0 && !t)) return fail(VIORB_E_ARG, "bad argument");
    if (n_out) *n_out = 0;
    if (nq == 0) return VIORB_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device) return fail(VIORB_E_NO_DEVICE, "no CUDA device: libviorb has no CPU path");
    CU(cudaSetDevice(device));
    DevBuf<uint8_t> dq, dt;
    cudaError_t e = dq.alloc((size_t)nq * 32);
    if (e == cudaSuccess) e = dt.alloc((size_t)std::max(nt, 1) * 32);
    if (e == cudaSuccess) e = cudaMemcpy(dq.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && nt > 0) e = cudaMemcpy(dt.p, t, (size_t)nt * 32, cudaMemcpyHostToDevice);
    int rc = e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc) rc = ratio_on_device(dq.p, nq, dt.p, nt, ratio, rule, dedup, out, cap, n_out, (cudaStream_t)0);
    dq.release(); dt.release();
    return rc;
}

int viorb_match_resident(viorb_extractor* cur, int frame_q, viorb_extractor* prev, int frame_t, double ratio, int rule,
                         viorb_match* out, int cap, int* n_out) {
    if (!cur || !prev || frame_q < 0 || frame_q >= cur->lastN || frame_t < 0 || frame_t >= prev->lastN)
        return fail(VIORB_E_ARG, "bad handle or frame index");
    CU(cudaSetDevice(cur->prm.device));
    int rc = join_streams(cur, prev);
    if (rc) return rc;
    int nq = 0, nt = 0;
    CU(cudaMemcpyAsync(&nq, cur->dNOut.p + frame_q, sizeof(int), cudaMemcpyDeviceToHost, cur->stream));
    CU(cudaMemcpyAsync(&nt, prev->dNOut.p + frame_t, sizeof(int), cudaMemcpyDeviceToHost, cur->stream));
    CU(cudaStreamSynchronize(cur->stream));
    return ratio_on_device(cur->dDesc.p + (size_t)frame_q * cur->g.outCap * 32, nq,
                           prev->dDesc.p + (size_t)frame_t * prev->g.outCap * 32, nt, ratio, rule, 0, out, cap, n_out,
                           cur->stream);
}

int viorb_stereo_batch(viorb_extractor* L, viorb_extractor* R, int n, float bf, float b, double ratio) {
    if (!L || !R || n <= 0 || n > L->lastN || n > R->lastN) return fail(VIORB_E_ARG, "bad handle or frame count");
    if (!same_geometry(L, R)) return fail(VIORB_E_ARG, "left and right extractor must share parameters and image size");
    CU(cudaSetDevice(L->prm.device));
    int rc = ensure_match_buffers(L);
    if (rc) return rc;
    rc = join_streams(L, R);
    if (rc) return rc;
    // fisheye-style brute force kNN-2 + ratio, left -> right (frame.cpp:1167-1170).  It needs nothing from the rectified
    // matcher, whose kernels are bound by load latency: outside stage timing it runs beside them on the handle's side stream.
    const size_t oc = L->g.outCap;
    DescSet Q{L->dDesc.p, oc * 32, L->dNOut.p, 0}, T{R->dDesc.p, oc * 32, R->dNOut.p, 0};
    const bool fork = !L->timing && L->sideOct != nullptr;
    if (fork) {
        CU(cudaEventRecord(L->evOctFork, L->stream));
        CU(cudaStreamWaitEvent(L->sideOct, L->evOctFork, 0));
        rc = launch_knn2(Q, T, L->g.outCap, L->g.outCap, n, L->dKnnIdx.p, L->dKnnDist.p, oc, L->sideOct);
        if (rc) return rc;
        VIORB_LAUNCH(k_ratio, dim3(n), dim3(256), 0, L->sideOct, Q, L->dKnnIdx.p, L->dKnnDist.p, oc, ratio, 0, L->dMatches.p, oc,
                     L->dNRatio.p, L->dLrMatch.p);
        CU(cudaEventRecord(L->evOctJoin, L->sideOct));
    }
    rc = launch_stereo(L, R, n, bf, b);
    if (rc) return rc;
    if (fork) {
        CU(cudaStreamWaitEvent(L->stream, L->evOctJoin, 0));
    } else {
        rc = launch_knn2(Q, T, L->g.outCap, L->g.outCap, n, L->dKnnIdx.p, L->dKnnDist.p, oc, L->stream);
        if (rc) return rc;
        if (L->timing) CU(cudaEventRecord(L->evS[4], L->stream));
        VIORB_LAUNCH(k_ratio, dim3(n), dim3(256), 0, L->stream, Q, L->dKnnIdx.p, L->dKnnDist.p, oc, ratio, 0, L->dMatches.p, oc,
                     L->dNRatio.p, L->dLrMatch.p);
        if (L->timing) { CU(cudaEventRecord(L->evS[5], L->stream)); L->haveS = true; }
    }
    CU(cudaGetLastError());
    return release_partner(L, R);
}

int viorb_stereo_fetch(viorb_extractor* L, int frame, float* u_right, float* depth, int32_t* lr_match, int cap, int* n_left,
                       int* n_stereo, int* n_ratio) {
    if (!L || frame < 0 || frame >= L->lastN || !L->dURight.p) return fail(VIORB_E_ARG, "bad handle or frame index");
    CU(cudaSetDevice(L->prm.device));
    int c[3];
    CU(cudaMemcpyAsync(&c[0], L->dNOut.p + frame, sizeof(int), cudaMemcpyDeviceToHost, L->stream));
    CU(cudaMemcpyAsync(&c[1], L->dNStereo.p + frame, sizeof(int), cudaMemcpyDeviceToHost, L->stream));
    CU(cudaMemcpyAsync(&c[2], L->dNRatio.p + frame, sizeof(int), cudaMemcpyDeviceToHost, L->stream));
    CU(cudaStreamSynchronize(L->stream));
    if (n_left) *n_left = c[0];
    if (n_stereo) *n_stereo = c[1];
    if (n_ratio) *n_ratio = c[2];
    const int m = std::min(c[0], cap);
    const size_t o = (size_t)frame * L->g.outCap;
    if (m > 0) {
        if (u_right) CU(cudaMemcpyAsync(u_right, L->dURight.p + o, sizeof(float) * m, cudaMemcpyDeviceToHost, L->stream));
        if (depth) CU(cudaMemcpyAsync(depth, L->dDepth.p + o, sizeof(float) * m, cudaMemcpyDeviceToHost, L->stream));
        if (lr_match) CU(cudaMemcpyAsync(lr_match, L->dLrMatch.p + o, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, L->stream));
        CU(cudaStreamSynchronize(L->stream));
    }
    return VIORB_OK;
}

/* Whole-batch read-back of the stereo results: arrays are [n][cap] / [n]. One pinned-free D2H per array. */
int viorb_stereo_fetch_all(viorb_extractor* L, int n, float* u_right, float* depth, int32_t* lr_match, int cap, int* n_left,
                           int* n_stereo, int* n_ratio) {
    if (!L || n <= 0 || n > L->lastN || !L->dURight.p) return fail(VIORB_E_ARG, "bad handle or frame count");
    if (cap < L->g.outCap) return fail(VIORB_E_CAPACITY, "cap must be at least viorb_extractor_capacity()");
    CU(cudaSetDevice(L->prm.device));
    const size_t oc = L->g.outCap;
    cudaStream_t st = L->stream;
    if (n_left) CU(cudaMemcpyAsync(n_left, L->dNOut.p, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    if (n_stereo) CU(cudaMemcpyAsync(n_stereo, L->dNStereo.p, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    if (n_ratio) CU(cudaMemcpyAsync(n_ratio, L->dNRatio.p, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    // rows past the largest keypoint count of the batch do not exist: copy them only when the count is not known here
    const size_t rows = L->lastMaxCount > 0 ? std::min<size_t>((size_t)L->lastMaxCount, oc) : oc;
    if (u_right) CU(cudaMemcpy2DAsync(u_right, sizeof(float) * cap, L->dURight.p, sizeof(float) * oc, sizeof(float) * rows, n, cudaMemcpyDeviceToHost, st));
    if (depth) CU(cudaMemcpy2DAsync(depth, sizeof(float) * cap, L->dDepth.p, sizeof(float) * oc, sizeof(float) * rows, n, cudaMemcpyDeviceToHost, st));
    if (lr_match) CU(cudaMemcpy2DAsync(lr_match, sizeof(int32_t) * cap, L->dLrMatch.p, sizeof(int32_t) * oc, sizeof(int32_t) * rows, n, cudaMemcpyDeviceToHost, st));
    CU(wait_stream(L, st));
    return VIORB_OK;
}

int viorb_stereo_match(viorb_extractor* L, viorb_extractor* R, int frame, float bf, float b, float* u_right, float* depth,
                       int32_t* best_idx, int32_t* best_dist, int cap, int* n_left) {
    if (!L || !R || frame < 0 || frame >= L->lastN || frame >= R->lastN) return fail(VIORB_E_ARG, "bad handle or frame index");
    if (!same_geometry(L, R)) return fail(VIORB_E_ARG, "left and right extractor must share parameters and image size");
    CU(cudaSetDevice(L->prm.device));
    int rc = ensure_match_buffers(L);
    if (rc) return rc;
    rc = join_streams(L, R);
    if (rc) return rc;
    // the kernels are batched over frames [0, n): run up to and including `frame`
    rc = launch_stereo(L, R, frame + 1, bf, b);
    if (rc) return rc;
    rc = release_partner(L, R);
    if (rc) return rc;
    int nl = 0;
    CU(cudaMemcpyAsync(&nl, L->dNOut.p + frame, sizeof(int), cudaMemcpyDeviceToHost, L->stream));
    CU(cudaStreamSynchronize(L->stream));
    if (n_left) *n_left = nl;
    const int m = std::min(nl, cap);
    const size_t o = (size_t)frame * L->g.outCap;
    if (m > 0) {
        if (u_right) CU(cudaMemcpyAsync(u_right, L->dURight.p + o, sizeof(float) * m, cudaMemcpyDeviceToHost, L->stream));
        if (depth) CU(cudaMemcpyAsync(depth, L->dDepth.p + o, sizeof(float) * m, cudaMemcpyDeviceToHost, L->stream));
        if (best_idx) CU(cudaMemcpyAsync(best_idx, L->dBestIdx.p + o, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, L->stream));
        if (best_dist) CU(cudaMemcpyAsync(best_dist, L->dBestDist.p + o, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, L->stream));
        CU(cudaStreamSynchronize(L->stream));
    }
    if (nl > cap) return fail(VIORB_E_CAPACITY, "output buffer smaller than the number of left keypoints");
    return VIORB_OK;
}


// ---- windowed (grid) matchers ---------------------------------------------------------------------
}  // extern "C"

namespace {

template <class T>
cudaError_t upload(DevBuf<T>& d, const T* h, size_t n) {
    cudaError_t e = d.alloc(n);
    if (e == cudaSuccess && n > 0 && h) e = cudaMemcpy(d.p, h, n * sizeof(T), cudaMemcpyHostToDevice);
    return e;
}

GridBounds make_bounds(const viorb_grid_bounds& b) {
    GridBounds g;
    g.minX = b.min_x; g.maxX = b.max_x; g.minY = b.min_y; g.maxY = b.max_y;
    g.invW = (float)VIORB_GRID_COLS / (b.max_x - b.min_x);      // frame.cpp:235-236
    g.invH = (float)VIORB_GRID_ROWS / (b.max_y - b.min_y);
    return g;
}

// Grid build + window lists for a batch whose sources (dSrc) are already on the device.  Leaves listStart /
// list filled.  desc1 may be null (distance 0: plain GetFeaturesInArea).
struct WindowWork {
    DevBuf<int> cellStart, cellList, listStart, totals, pairBase, n1, n2;
    DevBuf<float4> cellKp;
    DevBuf<uint32_t> list;
    WinBatch wb{};
    void release() {
        cellStart.release(); cellList.release(); cellKp.release(); listStart.release(); totals.release(); pairBase.release();
        n1.release(); n2.release(); list.release();
    }
};

int window_lists(WindowWork& w, int npairs, int cap1, int cap2, const int32_t* n1, const int32_t* n2, const GridBounds& gb,
                 const WinSrc* dSrc, const uint8_t* dDesc1, const viorb_keypoint* dKps2, const uint8_t* dDesc2,
                 const float* dURight2, const int* dOff2 = nullptr) {
    CU(upload(w.n1, n1, npairs));
    CU(upload(w.n2, n2, npairs));
    w.wb.cap1 = cap1; w.wb.cap2 = cap2; w.wb.n1 = w.n1.p; w.wb.n2 = w.n2.p; w.wb.n1fill = nullptr; w.wb.off2 = dOff2;
    CU(w.cellStart.alloc((size_t)npairs * (VIORB_GRID_CELLS + 1)));
    CU(w.cellList.alloc((size_t)npairs * cap2));
    CU(w.cellKp.alloc((size_t)npairs * std::max(cap2, 1)));
    CU(w.listStart.alloc((size_t)npairs * (cap1 + 1)));
    CU(w.totals.alloc(npairs));
    CU(w.pairBase.alloc(npairs));
    cudaStream_t st = 0;
    VIORB_LAUNCH(k_grid_build, dim3(npairs), dim3(256), 0, st, w.wb, gb, dKps2, w.cellStart.p, w.cellList.p, w.cellKp.p);
    const dim3 gl((cap1 + 7) / 8, npairs);
    VIORB_LAUNCH(k_window_list, gl, dim3(256), 0, st, w.wb, gb, dSrc, dDesc1, dKps2, dDesc2, dURight2, w.cellStart.p, w.cellKp.p,
                 w.listStart.p, (uint32_t*)nullptr, 0);
    VIORB_LAUNCH(k_window_scan, dim3(npairs), dim3(256), 0, st, w.wb, w.listStart.p, w.totals.p);
    std::vector<int> tot(npairs), base(npairs);
    CU(cudaMemcpy(tot.data(), w.totals.p, sizeof(int) * npairs, cudaMemcpyDeviceToHost));
    size_t sum = 0;
    for (int p = 0; p < npairs; ++p) { base[p] = (int)sum; sum += (size_t)tot[p]; }
    if (sum >= 0x7fffffffull) return fail(VIORB_E_CAPACITY, "candidate lists exceed 2^31 entries");
    CU(cudaMemcpy(w.pairBase.p, base.data(), sizeof(int) * npairs, cudaMemcpyHostToDevice));
    VIORB_LAUNCH(k_window_rebase, dim3((cap1 + 256) / 256, npairs), dim3(256), 0, st, w.wb, w.listStart.p, w.pairBase.p);
    CU(w.list.alloc(sum));
    VIORB_LAUNCH(k_window_list, gl, dim3(256), 0, st, w.wb, gb, dSrc, dDesc1, dKps2, dDesc2, dURight2, w.cellStart.p, w.cellKp.p,
                 w.listStart.p, w.list.p, 1);
    CU(cudaGetLastError());
    return VIORB_OK;
}

// dynamic shared memory of k_resolve_ini: vMatchedDistance + vnMatches21 of the target frame (0 = keep them in global memory)
size_t resolve_ini_smem(int cap2) { return (size_t)cap2 * 8 <= 200 * 1024 ? (size_t)cap2 * 8 : 0; }

int need_device(int device) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device) return fail(VIORB_E_NO_DEVICE, "no CUDA device: libviorb has no CPU path");
    CU(cudaSetDevice(device));
    DeviceState* ds = nullptr;
    return device_state(device, &ds);          // kernel attributes (dynamic shared-memory limits) are set once per device
}

}  // namespace

extern "C" {

int viorb_features_in_area(const viorb_keypoint* kps, int n, const viorb_grid_bounds* bounds, float x, float y, float r,
                           int min_level, int max_level, int32_t* out, int cap, int* n_out, int device) {
    if (n < 0 || (n > 0 && !kps) || !bounds || !n_out) return fail(VIORB_E_ARG, "bad argument");
    if (n >= 65536) return fail(VIORB_E_ARG, "more than 65535 keypoints per frame are not supported");
    *n_out = 0;
    if (n == 0) return VIORB_OK;
    int rc = need_device(device);
    if (rc) return rc;
    const GridBounds gb = make_bounds(*bounds);
    DevBuf<viorb_keypoint> dK;
    DevBuf<WinSrc> dS;
    WindowWork w;
    WinSrc s{x, y, r, 0.f, min_level, max_level};
    const int32_t one = 1, nn = n;
    cudaError_t e = upload(dK, kps, n);
    if (e == cudaSuccess) e = upload(dS, &s, 1);
    rc = e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc) rc = window_lists(w, 1, 1, n, &one, &nn, gb, dS.p, nullptr, dK.p, nullptr, nullptr);
    if (!rc) {
        std::vector<uint32_t> l(w.list.n);
        if (!l.empty()) e = cudaMemcpy(l.data(), w.list.p, sizeof(uint32_t) * l.size(), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        else {
            *n_out = (int)l.size();
            for (int i = 0; i < (int)l.size() && i < cap && out; ++i) out[i] = (int32_t)(l[i] & 0xFFFF);
            if ((int)l.size() > cap && out) rc = fail(VIORB_E_CAPACITY, "output buffer too small (n_out holds the count)");
        }
    }
    dK.release(); dS.release(); w.release();
    return rc;
}

int viorb_search_for_initialization(int npairs, int cap1, int cap2, const viorb_keypoint* kps1, const uint8_t* desc1,
                                    const int32_t* n1, const viorb_keypoint* kps2, const uint8_t* desc2, const int32_t* n2,
                                    const viorb_grid_bounds* bounds, float* prev_matched, int window_size, float nn_ratio,
                                    int check_orientation, int32_t* matches12, int32_t* nmatches, int device) {
    if (npairs < 1 || cap1 < 0 || cap2 < 0 || !n1 || !n2 || !bounds || !matches12 || !prev_matched)
        return fail(VIORB_E_ARG, "bad argument");
    if (cap2 >= 65536) return fail(VIORB_E_ARG, "more than 65535 keypoints per frame are not supported");
    for (int p = 0; p < npairs; ++p)
        if (n1[p] < 0 || n1[p] > cap1 || n2[p] < 0 || n2[p] > cap2) return fail(VIORB_E_ARG, "n1 / n2 outside [0, cap]");
    if ((cap1 > 0 && (!kps1 || !desc1)) || (cap2 > 0 && (!kps2 || !desc2))) return fail(VIORB_E_ARG, "null keypoints or descriptors");
    if (cap1 == 0) { if (nmatches) for (int p = 0; p < npairs; ++p) nmatches[p] = 0; return VIORB_OK; }
    int rc = need_device(device);
    if (rc) return rc;
    const GridBounds gb = make_bounds(*bounds);
    const size_t N1 = (size_t)npairs * cap1, N2 = (size_t)npairs * std::max(cap2, 1);
    DevBuf<viorb_keypoint> dK1, dK2;
    DevBuf<uint8_t> dD1, dD2;
    DevBuf<float> dPM;
    DevBuf<WinSrc> dS;
    DevBuf<int> dMD, dM21, dM12, dNM, dAcc;
    DevBuf<int8_t> dBin;
    WindowWork w;
    cudaError_t e = upload(dK1, kps1, N1);
    if (e == cudaSuccess) e = dAcc.alloc(N1);
    if (e == cudaSuccess) e = upload(dD1, desc1, N1 * 32);
    if (e == cudaSuccess) e = upload(dK2, cap2 ? kps2 : nullptr, N2);
    if (e == cudaSuccess) e = upload(dD2, cap2 ? desc2 : nullptr, N2 * 32);
    if (e == cudaSuccess) e = upload(dPM, prev_matched, N1 * 2);
    if (e == cudaSuccess) e = dS.alloc(N1);
    if (e == cudaSuccess) e = dMD.alloc(N2);
    if (e == cudaSuccess) e = dM21.alloc(N2);
    if (e == cudaSuccess) e = dM12.alloc(N1);
    if (e == cudaSuccess) e = dBin.alloc(N1);
    if (e == cudaSuccess) e = dNM.alloc(npairs);
    rc = e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc) {
        DevBuf<int> tn1;                     // the source kernel needs n1 on the device before window_lists uploads it
        e = upload(tn1, n1, npairs);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        else {
            WinBatch wb{cap1, cap2, tn1.p, nullptr, nullptr};
            VIORB_LAUNCH(k_window_src_ini, dim3((cap1 + 255) / 256, npairs), dim3(256), 0, (cudaStream_t)0, wb, dK1.p, dPM.p,
                         (float)window_size, dS.p, 0);
            e = cudaDeviceSynchronize();
            if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        }
        tn1.release();
    }
    if (!rc) rc = window_lists(w, npairs, cap1, cap2, n1, n2, gb, dS.p, dD1.p, dK2.p, dD2.p, nullptr);
    if (!rc) {
        const int useSmem = resolve_ini_smem(cap2) ? 1 : 0;
        VIORB_LAUNCH(k_resolve_ini, dim3(npairs), dim3(32), resolve_ini_smem(cap2), (cudaStream_t)0, w.wb, dK1.p, dK2.p, w.listStart.p,
                     w.list.p, nn_ratio, check_orientation, dMD.p, dM21.p, dBin.p, dAcc.p, dM12.p, dPM.p, dNM.p, useSmem);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpy(matches12, dM12.p, sizeof(int32_t) * N1, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(prev_matched, dPM.p, sizeof(float) * 2 * N1, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && nmatches) e = cudaMemcpy(nmatches, dNM.p, sizeof(int32_t) * npairs, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
    }
    dK1.release(); dK2.release(); dD1.release(); dD2.release(); dPM.release(); dS.release(); dMD.release(); dM21.release();
    dM12.release(); dBin.release(); dNM.release(); dAcc.release(); w.release();
    return rc;
}

int viorb_search_by_projection(int npairs, int cap1, int cap2, const float* uv, const float* invz, const int32_t* last_octave,
                               const float* last_angle, const uint8_t* desc_mp, const uint8_t* valid, const uint8_t* observed,
                               const int32_t* n1, const viorb_keypoint* kps2, const uint8_t* desc2, const float* u_right2,
                               const uint8_t* blocked2, const int32_t* n2, const viorb_grid_bounds* bounds,
                               const float* scale_factors, int nlevels, float th, float bf, int direction,
                               int check_orientation, const float* uv_right, const int32_t* nleft2, int32_t* assigned,
                               int32_t* nmatches, int device) {
    if (npairs < 1 || cap1 < 0 || cap2 < 0 || !n1 || !n2 || !bounds || !assigned || !scale_factors || nlevels < 1 ||
        nlevels > VIORB_MAX_LEVELS || direction < 0 || direction > 2)
        return fail(VIORB_E_ARG, "bad argument");
    const bool rig = uv_right != nullptr;           // CurrentFrame.Nleft != -1 (fmatcher.cpp:2593-2660)
    if (rig && !nleft2) return fail(VIORB_E_ARG, "uv_right needs nleft2 (CurrentFrame.Nleft per pair)");
    if (rig && u_right2) return fail(VIORB_E_ARG, "two-camera frames have no mvuRight (the predicate is guarded by Nleft == -1)");
    if (rig) for (int p = 0; p < npairs; ++p)
        if (nleft2[p] < 0 || nleft2[p] > n2[p]) return fail(VIORB_E_ARG, "nleft2 outside [0, n2]");
    if (cap2 >= 65536) return fail(VIORB_E_ARG, "more than 65535 keypoints per frame are not supported");
    for (int p = 0; p < npairs; ++p)
        if (n1[p] < 0 || n1[p] > cap1 || n2[p] < 0 || n2[p] > cap2) return fail(VIORB_E_ARG, "n1 / n2 outside [0, cap]");
    if (cap1 > 0 && (!uv || !invz || !last_octave || !last_angle || !desc_mp || !valid || !observed)) return fail(VIORB_E_ARG, "null source array");
    if (cap2 > 0 && (!kps2 || !desc2)) return fail(VIORB_E_ARG, "null keypoints or descriptors");
    if (cap2 == 0) { if (nmatches) for (int p = 0; p < npairs; ++p) nmatches[p] = 0; return VIORB_OK; }
    int rc = need_device(device);
    if (rc) return rc;
    const GridBounds gb = make_bounds(*bounds);
    ProjParams pp;
    std::memset(&pp, 0, sizeof pp);
    pp.th = th; pp.bf = bf; pp.direction = direction; pp.nlevels = nlevels;
    for (int l = 0; l < nlevels; ++l) pp.scale[l] = scale_factors[l];
    const size_t N1 = (size_t)npairs * std::max(cap1, 1), N2 = (size_t)npairs * cap2;
    DevBuf<float> dUV, dIZ, dLA, dUR;
    DevBuf<int32_t> dLO, dAS, dNM, dPI;
    DevBuf<uint8_t> dDM, dVA, dOB, dD2, dBL;
    DevBuf<int8_t> dPB;
    DevBuf<viorb_keypoint> dK2;
    DevBuf<WinSrc> dS, dSR;
    DevBuf<float> dUVR;
    DevBuf<int> dNLeft, dN2;
    WindowWork w, wR;
    const bool have1 = cap1 > 0;
    std::vector<int32_t> nL(npairs), nR(npairs);
    for (int p = 0; p < npairs; ++p) { nL[p] = rig ? nleft2[p] : n2[p]; nR[p] = n2[p] - nL[p]; }
    cudaError_t e = upload(dUV, have1 ? uv : nullptr, N1 * 2);
    if (e == cudaSuccess && rig) e = upload(dUVR, have1 ? uv_right : nullptr, N1 * 2);
    if (e == cudaSuccess && rig) e = upload(dNLeft, nL.data(), npairs);
    if (e == cudaSuccess && rig) e = dSR.alloc(N1);
    if (e == cudaSuccess) e = upload(dN2, n2, npairs);
    if (e == cudaSuccess) e = upload(dIZ, have1 ? invz : nullptr, N1);
    if (e == cudaSuccess) e = upload(dLA, have1 ? last_angle : nullptr, N1);
    if (e == cudaSuccess) e = upload(dLO, have1 ? last_octave : nullptr, N1);
    if (e == cudaSuccess) e = upload(dDM, have1 ? desc_mp : nullptr, N1 * 32);
    if (e == cudaSuccess) e = upload(dVA, have1 ? valid : nullptr, N1);
    if (e == cudaSuccess) e = upload(dOB, have1 ? observed : nullptr, N1);
    if (e == cudaSuccess) e = upload(dK2, kps2, N2);
    if (e == cudaSuccess) e = upload(dD2, desc2, N2 * 32);
    if (e == cudaSuccess && u_right2) e = upload(dUR, u_right2, N2);
    if (e == cudaSuccess) e = upload(dBL, blocked2, N2);
    if (e == cudaSuccess && !blocked2) e = cudaMemset(dBL.p, 0, N2);
    if (e == cudaSuccess) e = dS.alloc(N1);
    if (e == cudaSuccess) e = dAS.alloc(N2);
    if (e == cudaSuccess) e = dPI.alloc(N1 * 4);
    if (e == cudaSuccess) e = dPB.alloc(N1 * 2);
    if (e == cudaSuccess) e = dNM.alloc(npairs);
    rc = e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc && have1) {
        DevBuf<int> tn1;
        e = upload(tn1, n1, npairs);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        else {
            WinBatch wb{cap1, cap2, tn1.p, nullptr};
            VIORB_LAUNCH(k_window_src_proj, dim3((cap1 + 255) / 256, npairs), dim3(256), 0, (cudaStream_t)0, wb, pp, dUV.p, dIZ.p, dLO.p,
                         dVA.p, dS.p);
            // the same map points projected into the right camera: same radius and octave window (:2596-2611)
            if (rig) VIORB_LAUNCH(k_window_src_proj, dim3((cap1 + 255) / 256, npairs), dim3(256), 0, (cudaStream_t)0, wb, pp, dUVR.p,
                                  dIZ.p, dLO.p, dVA.p, dSR.p);
            e = cudaDeviceSynchronize();
            if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        }
        tn1.release();
    }
    if (!rc) rc = window_lists(w, npairs, cap1, cap2, n1, nL.data(), gb, dS.p, dDM.p, dK2.p, dD2.p, u_right2 ? dUR.p : nullptr);
    if (!rc && rig) rc = window_lists(wR, npairs, cap1, cap2, n1, nR.data(), gb, dSR.p, dDM.p, dK2.p, dD2.p, nullptr, dNLeft.p);
    if (!rc) {
        VIORB_LAUNCH(k_resolve_proj, dim3(npairs), dim3(32), (size_t)cap2, (cudaStream_t)0, w.wb, dLA.p, dOB.p, dK2.p, w.listStart.p, w.list.p,
                     rig ? wR.listStart.p : (const int*)nullptr, rig ? wR.list.p : (const uint32_t*)nullptr,
                     rig ? dNLeft.p : (const int*)nullptr, cap2, check_orientation, dBL.p, dPI.p, dPB.p, dAS.p, dNM.p, dN2.p);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpy(assigned, dAS.p, sizeof(int32_t) * N2, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && nmatches) e = cudaMemcpy(nmatches, dNM.p, sizeof(int32_t) * npairs, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
    }
    dUV.release(); dIZ.release(); dLA.release(); dUR.release(); dLO.release(); dAS.release(); dNM.release(); dPI.release();
    dDM.release(); dVA.release(); dOB.release(); dD2.release(); dBL.release(); dPB.release(); dK2.release(); dS.release(); w.release();
    dSR.release(); dUVR.release(); dNLeft.release(); dN2.release(); wR.release();
    return rc;
}

// ---- keypoint undistortion (row f-2) ----------------------------------------------------------------
}  // extern "C"

namespace {

// mK / mDistCoef (CV_32F) -> the double model cvUndistortPointsInternal works on.
int make_cam(const float* cam, const float* dist, int ndist, CamModel* out) {
    CamModel c{};
    if (!dist) { *out = c; return VIORB_OK; }
    if (!cam || (ndist != 4 && ndist != 5)) return fail(VIORB_E_ARG, "camera needs fx fy cx cy and 4 or 5 radtan coefficients");
    c.fx = cam[0]; c.fy = cam[1]; c.cx = cam[2]; c.cy = cam[3];
    c.ifx = 1.0 / c.fx; c.ify = 1.0 / c.fy;
    for (int i = 0; i < ndist; ++i) c.k[i] = dist[i];
    c.mode = dist[0] == 0.0f ? 1 : 2;
    *out = c;
    return VIORB_OK;
}

}  // namespace

extern "C" {

int viorb_extractor_set_camera(viorb_extractor* ex, const float* cam, const float* dist, int ndist) {
    if (!ex) return fail(VIORB_E_ARG, "null handle");
    CamModel c;
    int rc = make_cam(cam, dist, ndist, &c);
    if (rc) return rc;
    CU(cudaSetDevice(ex->prm.device));
    if (c.mode && !ex->dKpsUn.p) CU(ex->dKpsUn.alloc((size_t)ex->prm.max_batch * ex->capOut));
    ex->cam = c;
    return VIORB_OK;
}

int viorb_fetch_undistorted(viorb_extractor* ex, int frame, viorb_keypoint* kps_un, int cap, int* n_out) {
    if (!ex || frame < 0 || frame >= ex->lastN) return fail(VIORB_E_ARG, "bad handle or frame index");
    if (!ex->cam.mode) return fail(VIORB_E_ARG, "no camera set: call viorb_extractor_set_camera before extracting");
    CU(cudaSetDevice(ex->prm.device));
    int rc = check_status(ex);
    if (rc) return rc;
    int cnt = 0;
    CU(cudaMemcpyAsync(&cnt, ex->dNOut.p + frame, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    if (n_out) *n_out = cnt;
    const int m = std::min(cnt, cap);
    if (kps_un && m > 0) {
        CU(cudaMemcpyAsync(kps_un, ex->dKpsUn.p + (size_t)frame * ex->g.outCap, sizeof(viorb_keypoint) * m, cudaMemcpyDeviceToHost,
                           ex->stream));
        CU(cudaStreamSynchronize(ex->stream));
    }
    if (cnt > cap && kps_un) return fail(VIORB_E_CAPACITY, "output buffer smaller than the number of keypoints");
    return VIORB_OK;
}

int viorb_undistort_keypoints(const viorb_keypoint* kps, int n, const float* cam, const float* dist, int ndist,
                              viorb_keypoint* out, int device) {
    if (n < 0 || (n > 0 && (!kps || !out)) || !dist) return fail(VIORB_E_ARG, "bad argument");
    CamModel c;
    int rc = make_cam(cam, dist, ndist, &c);
    if (rc) return rc;
    if (n == 0) return VIORB_OK;
    rc = need_device(device);
    if (rc) return rc;
    DevBuf<viorb_keypoint> dIn, dOut;
    cudaError_t e = upload(dIn, kps, n);
    if (e == cudaSuccess) e = dOut.alloc(n);
    if (e == cudaSuccess) {
        VIORB_LAUNCH(k_undistort_keypoints, dim3((n + 255) / 256, 1), dim3(256), 0, (cudaStream_t)0, dIn.p, (const int*)nullptr, n, c, dOut.p);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out, dOut.p, sizeof(viorb_keypoint) * n, cudaMemcpyDeviceToHost);
    dIn.release(); dOut.release();
    return e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
}

int viorb_undistort_points(const float* xy, int n, const float* cam, const float* dist, int ndist, float* out, int device) {
    if (n < 0 || (n > 0 && (!xy || !out)) || !dist) return fail(VIORB_E_ARG, "bad argument");
    CamModel c;
    int rc = make_cam(cam, dist, ndist, &c);
    if (rc) return rc;
    if (n == 0) return VIORB_OK;
    rc = need_device(device);
    if (rc) return rc;
    DevBuf<float2> dIn, dOut;
    cudaError_t e = upload(dIn, reinterpret_cast<const float2*>(xy), n);
    if (e == cudaSuccess) e = dOut.alloc(n);
    if (e == cudaSuccess) {
        // bare points are always undistorted (the k1 == 0 shortcut belongs to the two Frame functions)
        c.mode = 2;
        VIORB_LAUNCH(k_undistort_points, dim3((n + 255) / 256), dim3(256), 0, (cudaStream_t)0, dIn.p, n, c, dOut.p);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out, dOut.p, sizeof(float2) * n, cudaMemcpyDeviceToHost);
    dIn.release(); dOut.release();
    return e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
}

int viorb_image_bounds(int cols, int rows, const float* cam, const float* dist, int ndist, viorb_grid_bounds* bounds, int device) {
    if (!bounds || !dist || cols <= 0 || rows <= 0) return fail(VIORB_E_ARG, "bad argument");
    if (dist[0] == 0.0f) {
        bounds->min_x = 0.f; bounds->max_x = (float)cols; bounds->min_y = 0.f; bounds->max_y = (float)rows;
        return VIORB_OK;
    }
    const float c[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
    float u[8];
    int rc = viorb_undistort_points(c, 4, cam, dist, ndist, u, device);
    if (rc) return rc;
    bounds->min_x = std::min(u[0], u[4]);
    bounds->max_x = std::max(u[2], u[6]);
    bounds->min_y = std::min(u[1], u[3]);
    bounds->max_y = std::max(u[5], u[7]);
    return VIORB_OK;
}

// ---- bag of words (row f-3) ----------------------------------------------------------------------------
}  // extern "C"

namespace {

int bow_run(const viorb_vocabulary* voc, const uint8_t* dDesc, const int* dCounts, int nframes, int cap, int levelsup, BowBufs& b,
            cudaStream_t st) {
    if (cap > 16384) return fail(VIORB_E_CAPACITY, "bag-of-words transform supports at most 16384 descriptors per frame");
    CU(b.ensure(nframes, cap));
    int P = 256;                       // k_bow_build walks the sorted slots blockDim.x at a time: P >= 256
    while (P < cap) P <<= 1;
    const size_t smem = (size_t)P * sizeof(unsigned long long);
    CU(cudaFuncSetAttribute(k_bow_build, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    VIORB_LAUNCH(k_bow_walk, dim3((cap + 7) / 8, nframes), dim3(256), 0, st, voc->d, dDesc, dCounts, cap, levelsup, b.leaf.p, b.nid.p);
    VIORB_LAUNCH(k_bow_build, dim3(nframes), dim3(256), smem, st, voc->d, dCounts, cap, P, b.leaf.p, b.nid.p, b.bowWord.p,
                 b.bowValue.p, b.nBow.p, b.fvNode.p, b.fvStart.p, b.fvFeature.p, b.nFv.p);
    CU(cudaGetLastError());
    return VIORB_OK;
}

}  // namespace

extern "C" {

int viorb_vocabulary_create(const viorb_vocab_desc* desc, int device, viorb_vocabulary** out) {
    if (!desc || !out || desc->nnodes < 1 || !desc->child_start || !desc->descriptors || !desc->weight || !desc->word_id)
        return fail(VIORB_E_ARG, "bad vocabulary description");
    if (desc->weighting < 0 || desc->weighting > 3 || desc->scoring < 0 || desc->scoring > 5)
        return fail(VIORB_E_ARG, "unknown weighting or scoring type");
    const int nn = desc->nnodes;
    if (desc->child_start[0] != 0) return fail(VIORB_E_ARG, "child_start[0] must be 0");
    bool anyWord = false;
    for (int i = 0; i < nn; ++i) {
        const int c = desc->child_start[i + 1] - desc->child_start[i];
        if (c < 0 || c >= (1 << 22)) return fail(VIORB_E_ARG, "child_start must be non-decreasing (fewer than 2^22 children per node)");
        if (c == 0 && i > 0) anyWord = true;
    }
    const int nc = desc->child_start[nn];
    if (nc > 0 && !desc->children) return fail(VIORB_E_ARG, "children missing");
    for (int i = 0; i < nc; ++i)
        if (desc->children[i] <= 0 || desc->children[i] >= nn) return fail(VIORB_E_ARG, "child id out of range");
    int rc = need_device(device);
    if (rc) return rc;
    viorb_vocabulary* v = new viorb_vocabulary;
    v->device = device;
    cudaError_t e = upload(v->childStart, desc->child_start, (size_t)nn + 1);
    if (e == cudaSuccess) e = upload(v->children, desc->children, (size_t)nc);
    if (e == cudaSuccess) e = upload(v->wordId, desc->word_id, (size_t)nn);
    if (e == cudaSuccess) e = upload(v->weight, desc->weight, (size_t)nn);
    if (e == cudaSuccess) e = upload(v->desc, reinterpret_cast<const uint4*>(desc->descriptors), (size_t)nn * 2);
    if (e != cudaSuccess) {
        viorb_vocabulary_destroy(v);
        return fail(VIORB_E_CUDA, cudaGetErrorString(e));
    }
    v->d.nnodes = nn; v->d.levels = desc->levels; v->d.weighting = desc->weighting; v->d.scoring = desc->scoring;
    v->d.anyWord = anyWord ? 1 : 0;
    v->d.childStart = v->childStart.p; v->d.children = v->children.p; v->d.desc = v->desc.p; v->d.weight = v->weight.p;
    v->d.wordId = v->wordId.p;
    *out = v;
    return VIORB_OK;
}

void viorb_vocabulary_destroy(viorb_vocabulary* v) {
    if (!v) return;
    cudaSetDevice(v->device);
    v->childStart.release(); v->children.release(); v->wordId.release(); v->desc.release(); v->weight.release();
    delete v;
}

int viorb_bow_transform(const viorb_vocabulary* voc, int nframes, int cap, const uint8_t* desc, const int32_t* n, int levelsup,
                        uint32_t* bow_word, double* bow_value, int32_t* nbow, uint32_t* fv_node, int32_t* fv_start,
                        uint32_t* fv_feature, int32_t* nfv) {
    if (!voc || nframes < 0 || cap < 0 || (nframes > 0 && (!n || !nbow || !nfv))) return fail(VIORB_E_ARG, "bad argument");
    if (nframes == 0) return VIORB_OK;
    for (int f = 0; f < nframes; ++f)
        if (n[f] < 0 || n[f] > cap) return fail(VIORB_E_ARG, "n[f] outside [0, cap]");
    if (cap == 0) {
        for (int f = 0; f < nframes; ++f) { nbow[f] = 0; nfv[f] = 0; if (fv_start) fv_start[f] = 0; }
        return VIORB_OK;
    }
    if (!desc) return fail(VIORB_E_ARG, "descriptors missing");
    int rc = need_device(voc->device);
    if (rc) return rc;
    const size_t N = (size_t)nframes * cap;
    DevBuf<uint8_t> dD;
    DevBuf<int> dN;
    BowBufs b;
    cudaError_t e = upload(dD, desc, N * 32);
    if (e == cudaSuccess) e = upload(dN, n, (size_t)nframes);
    rc = e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc) rc = bow_run(voc, dD.p, dN.p, nframes, cap, levelsup, b, (cudaStream_t)0);
    if (!rc) {
        auto down = [&](void* h, const void* d, size_t bytes) { if (e == cudaSuccess && h) e = cudaMemcpy(h, d, bytes, cudaMemcpyDeviceToHost); };
        down(nbow, b.nBow.p, sizeof(int) * nframes);
        down(nfv, b.nFv.p, sizeof(int) * nframes);
        down(bow_word, b.bowWord.p, sizeof(uint32_t) * N);
        down(bow_value, b.bowValue.p, sizeof(double) * N);
        down(fv_node, b.fvNode.p, sizeof(uint32_t) * N);
        down(fv_start, b.fvStart.p, sizeof(int) * (size_t)nframes * (cap + 1));
        down(fv_feature, b.fvFeature.p, sizeof(uint32_t) * N);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
    }
    dD.release(); dN.release(); b.release();
    return rc;
}

int viorb_bow_resident(viorb_extractor* ex, const viorb_vocabulary* voc, int levelsup) {
    if (!ex || !voc || ex->lastN <= 0) return fail(VIORB_E_ARG, "bad handle, or nothing extracted yet");
    if (voc->device != ex->prm.device) return fail(VIORB_E_ARG, "vocabulary and extractor live on different devices");
    CU(cudaSetDevice(ex->prm.device));
    int rc = bow_run(voc, ex->dDesc.p, ex->dNOut.p, ex->lastN, ex->g.outCap, levelsup, ex->bow, ex->stream);
    if (rc) return rc;
    ex->bowN = ex->lastN;
    return VIORB_OK;
}

int viorb_bow_fetch(viorb_extractor* ex, int frame, uint32_t* bow_word, double* bow_value, int32_t* nbow, uint32_t* fv_node,
                    int32_t* fv_start, uint32_t* fv_feature, int32_t* nfv) {
    if (!ex || frame < 0 || frame >= ex->bowN) return fail(VIORB_E_ARG, "bad handle or frame index (call viorb_bow_resident first)");
    CU(cudaSetDevice(ex->prm.device));
    const BowBufs& b = ex->bow;
    const size_t cap = b.cap, o = (size_t)frame * cap;
    int cnt[2];
    CU(cudaMemcpyAsync(&cnt[0], b.nBow.p + frame, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaMemcpyAsync(&cnt[1], b.nFv.p + frame, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    if (nbow) *nbow = cnt[0];
    if (nfv) *nfv = cnt[1];
    int nlisted = 0;
    CU(cudaMemcpyAsync(&nlisted, b.fvStart.p + (size_t)frame * (cap + 1) + cnt[1], sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    if (bow_word && cnt[0]) CU(cudaMemcpyAsync(bow_word, b.bowWord.p + o, sizeof(uint32_t) * cnt[0], cudaMemcpyDeviceToHost, ex->stream));
    if (bow_value && cnt[0]) CU(cudaMemcpyAsync(bow_value, b.bowValue.p + o, sizeof(double) * cnt[0], cudaMemcpyDeviceToHost, ex->stream));
    if (fv_node && cnt[1]) CU(cudaMemcpyAsync(fv_node, b.fvNode.p + o, sizeof(uint32_t) * cnt[1], cudaMemcpyDeviceToHost, ex->stream));
    if (fv_start) CU(cudaMemcpyAsync(fv_start, b.fvStart.p + (size_t)frame * (cap + 1), sizeof(int) * (cnt[1] + 1), cudaMemcpyDeviceToHost, ex->stream));
    CU(cudaStreamSynchronize(ex->stream));
    if (fv_feature && nlisted) {
        CU(cudaMemcpyAsync(fv_feature, b.fvFeature.p + o, sizeof(uint32_t) * nlisted, cudaMemcpyDeviceToHost, ex->stream));
        CU(cudaStreamSynchronize(ex->stream));
    }
    return VIORB_OK;
}

// ---- fisheye stereo epilogue (row f-4) ------------------------------------------------------------------
static_assert(sizeof(Kb8Rig) == sizeof(viorb_kb8_rig), "rig layouts must match");

int viorb_kb8_triangulate_matches(const viorb_kb8_rig* rig, const float* pt1, const float* pt2, const float* sigma1,
                                  const float* sigma2, int n, float* depth, float* p3d, int device) {
    if (!rig || n < 0 || (n > 0 && (!pt1 || !pt2 || !sigma1 || !sigma2 || !depth || !p3d))) return fail(VIORB_E_ARG, "bad argument");
    if (n == 0) return VIORB_OK;
    int rc = need_device(device);
    if (rc) return rc;
    Kb8Rig g;
    std::memcpy(&g, rig, sizeof(g));
    DevBuf<float2> d1, d2;
    DevBuf<float> s1, s2, dz, dp;
    cudaError_t e = upload(d1, reinterpret_cast<const float2*>(pt1), n);
    if (e == cudaSuccess) e = upload(d2, reinterpret_cast<const float2*>(pt2), n);
    if (e == cudaSuccess) e = upload(s1, sigma1, n);
    if (e == cudaSuccess) e = upload(s2, sigma2, n);
    if (e == cudaSuccess) e = dz.alloc(n);
    if (e == cudaSuccess) e = dp.alloc((size_t)n * 3);
    if (e == cudaSuccess) {
        VIORB_LAUNCH(k_kb8_triangulate, dim3((n + 127) / 128), dim3(128), 0, (cudaStream_t)0, g, d1.p, d2.p, s1.p, s2.p, n, dz.p, dp.p);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(depth, dz.p, sizeof(float) * n, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(p3d, dp.p, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost);
    d1.release(); d2.release(); s1.release(); s2.release(); dz.release(); dp.release();
    return e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
}

int viorb_fisheye_stereo_matches(const viorb_kb8_rig* rig, const viorb_keypoint* kps_left, const uint8_t* desc_left, int n_left,
                                 int mono_left, const viorb_keypoint* kps_right, const uint8_t* desc_right, int n_right,
                                 int mono_right, const float* level_sigma2, int nlevels, int32_t* left_to_right,
                                 int32_t* right_to_left, float* depth, float* p3d, int32_t* n_desc_matches, int32_t* n_matches,
                                 int device) {
    if (!rig || n_left < 0 || n_right < 0 || mono_left < 0 || mono_left > n_left || mono_right < 0 || mono_right > n_right ||
        !level_sigma2 || nlevels <= 0 || (n_left > 0 && (!kps_left || !desc_left || !left_to_right || !depth || !p3d)) ||
        (n_right > 0 && (!kps_right || !desc_right || !right_to_left)))
        return fail(VIORB_E_ARG, "bad argument");
    for (int i = 0; i < n_left; ++i) { left_to_right[i] = -1; depth[i] = -1.0f; p3d[3 * i] = p3d[3 * i + 1] = p3d[3 * i + 2] = 0.f; }
    for (int i = 0; i < n_right; ++i) right_to_left[i] = -1;
    if (n_desc_matches) *n_desc_matches = 0;
    if (n_matches) *n_matches = 0;
    const int nq = n_left - mono_left, nt = n_right - mono_right;
    if (nq == 0 || nt < 2) return VIORB_OK;                  // knnMatch returns fewer than 2 neighbours: nothing passes
    for (int i = mono_left; i < n_left; ++i)
        if (kps_left[i].octave < 0 || kps_left[i].octave >= nlevels) return fail(VIORB_E_ARG, "keypoint octave outside level_sigma2");
    for (int i = mono_right; i < n_right; ++i)
        if (kps_right[i].octave < 0 || kps_right[i].octave >= nlevels) return fail(VIORB_E_ARG, "keypoint octave outside level_sigma2");
    int rc = need_device(device);
    if (rc) return rc;
    Kb8Rig g;
    std::memcpy(&g, rig, sizeof(g));
    DevBuf<viorb_keypoint> dKL, dKR;
    DevBuf<uint8_t> dDL, dDR;
    DevBuf<int32_t> dIdx, dDist, dL2R, dR2L;
    DevBuf<float> dSig, dZ, dP;
    DevBuf<int> dCnt;
    cudaError_t e = upload(dKL, kps_left, n_left);
    if (e == cudaSuccess) e = upload(dKR, kps_right, n_right);
    if (e == cudaSuccess) e = upload(dDL, desc_left + (size_t)mono_left * 32, (size_t)nq * 32);
    if (e == cudaSuccess) e = upload(dDR, desc_right + (size_t)mono_right * 32, (size_t)nt * 32);
    if (e == cudaSuccess) e = upload(dSig, level_sigma2, nlevels);
    if (e == cudaSuccess) e = upload(dL2R, left_to_right, n_left);
    if (e == cudaSuccess) e = upload(dR2L, right_to_left, n_right);
    if (e == cudaSuccess) e = upload(dZ, depth, n_left);
    if (e == cudaSuccess) e = upload(dP, p3d, (size_t)n_left * 3);
    if (e == cudaSuccess) e = dIdx.alloc((size_t)nq * 2);
    if (e == cudaSuccess) e = dDist.alloc((size_t)nq * 2);
    if (e == cudaSuccess) e = dCnt.alloc(2);
    if (e == cudaSuccess) e = cudaMemset(dCnt.p, 0, sizeof(int) * 2);
    rc = e == cudaSuccess ? VIORB_OK : fail(VIORB_E_CUDA, cudaGetErrorString(e));
    if (!rc) rc = viorb_knn2_device(dDL.p, nq, dDR.p, nt, dIdx.p, dDist.p, nullptr);
    if (!rc) {
        VIORB_LAUNCH(k_fisheye_epilogue, dim3((nq + 127) / 128), dim3(128), 0, (cudaStream_t)0, g, dKL.p, mono_left, dKR.p, mono_right, nq,
                     dIdx.p, dDist.p, dSig.p, dL2R.p, dR2L.p, dZ.p, dP.p, dCnt.p);
        e = cudaGetLastError();
        int cnt[2] = {0, 0};
        if (e == cudaSuccess) e = cudaMemcpy(left_to_right, dL2R.p, sizeof(int32_t) * n_left, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(right_to_left, dR2L.p, sizeof(int32_t) * n_right, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(depth, dZ.p, sizeof(float) * n_left, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(p3d, dP.p, sizeof(float) * 3 * n_left, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(cnt, dCnt.p, sizeof(int) * 2, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(VIORB_E_CUDA, cudaGetErrorString(e));
        if (n_desc_matches) *n_desc_matches = cnt[0];
        if (n_matches) *n_matches = cnt[1];
    }
    dKL.release(); dKR.release(); dDL.release(); dDR.release(); dIdx.release(); dDist.release(); dL2R.release(); dR2L.release();
    dSig.release(); dZ.release(); dP.release(); dCnt.release();
    return rc;
}

#include "host_sharded.inc"
#include "host_f2f.inc"

}  // extern "C"

// =================================================================================================
// viorb_knn2_device_static: kNN-2 on device buffers WITHOUT any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops,
// harnesses that forbid it).  The only scratch the kNN-2 path uses - the partial (best, second) key rows of a query tile
// whose train range is cut over several CTAs (k_knn2_tc5 -> k_knn2_tc5_merge) - lives in the file-scope device array
// below instead of the stream-ordered pool of launch_knn2.  Every GPU that loads the library holds these 4 MiB until the
// process exits.  Nothing in it has to start at zero: every row k_knn2_tc5_merge reads was written by k_knn2_tc5 in the
// same call.  All calls share the array, so two calls that overlap on the GPU (different streams) race.
// =================================================================================================
#define VIORB_STATIC_KNN_MAX_QUERIES VIORB_KNN2_STATIC_MAX_QUERIES   /* query rows per call (include/viorb.h) */
#define VIORB_STATIC_KNN_MAX_SPLIT 32                        /* CTAs per query tile, as in launch_knn2 */
#if VIORB_KNN_TENSOR
#define VIORB_STATIC_KNN_ROWS ((VIORB_STATIC_KNN_MAX_QUERIES / 128) * VIORB_STATIC_KNN_MAX_SPLIT * 128)     /* 524 288 uint2 = 4 MiB */
static __device__ __align__(256) uint2 g_knnStaticScratch[VIORB_STATIC_KNN_ROWS];

namespace {
struct StaticKnnState {
    std::once_flag once;
    cudaError_t err = cudaSuccess;
    int sms = 148;
    uint2* scratch = nullptr;         // DEVICE address of g_knnStaticScratch on this device (the name itself is a host-side stand-in)
};
StaticKnnState g_staticKnn[64];
}  // namespace
#endif

extern "C" int viorb_knn2_device_static(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int32_t* d_idx, int32_t* d_dist,
                                        void* stream) {
    if (nq < 0 || nt < 0 || (nq > 0 && (!d_q || !d_idx || !d_dist))) return fail(VIORB_E_ARG, "bad argument");
    if (nq > VIORB_STATIC_KNN_MAX_QUERIES)
        return fail(VIORB_E_CAPACITY, "viorb_knn2_device_static: nq = " + std::to_string(nq) + " exceeds the static scratch limit of " +
                                          std::to_string(VIORB_STATIC_KNN_MAX_QUERIES) + " query rows per call (use viorb_knn2_device, or split the queries)");
    if (nt >= (1 << VIORB_KNN_IDX_BITS)) return fail(VIORB_E_ARG, "train set larger than 2^22 rows");
    if (((uintptr_t)d_q & 15) || ((uintptr_t)d_t & 15)) return fail(VIORB_E_ARG, "descriptor buffers must be 16-byte aligned");
    if (nq == 0) return VIORB_OK;
#if VIORB_KNN_TENSOR
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(VIORB_E_ARG, "device index out of range");
    StaticKnnState& S = g_staticKnn[dev];
    std::call_once(S.once, [&S, dev] {                       // set-up, once per device, with that device current
        auto ok = [&S](cudaError_t r) { if (S.err == cudaSuccess) S.err = r; };
        ok(cudaDeviceGetAttribute(&S.sms, cudaDevAttrMultiProcessorCount, dev));
        ok(cudaFuncSetAttribute(VIORB_KNN_TC_KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, tc5::SMEM_BYTES));
        ok(cudaGetSymbolAddress((void**)&S.scratch, g_knnStaticScratch));
    });
    if (S.err != cudaSuccess || !S.scratch)
        return fail(VIORB_E_CUDA, std::string("viorb_knn2_device_static set-up: ") + cudaGetErrorString(S.err));
    cudaStream_t st = (cudaStream_t)stream;
    DescSet Q{d_q, 0, nullptr, nq}, T{d_t, 0, nullptr, nt};
    const int qTiles = (nq + tc5::QT - 1) / tc5::QT;
    // the train-range split of launch_knn2 (same cost model), bounded by what the static rows hold
    const int maxTiles = std::max(1, (nt + tc5::TT - 1) / tc5::TT);
    int nSplit = 1;
    long best = 0;
    for (int sp = 1; sp <= std::min(maxTiles, VIORB_STATIC_KNN_MAX_SPLIT); ++sp) {
        const long cost = (long)((qTiles * sp + S.sms - 1) / S.sms) * ((maxTiles + sp - 1) / sp + 2);
        if (sp == 1 || cost * 20 < best * 19) { best = cost; nSplit = sp; }
        if (sp == 1 && cost < 24) break;
    }
    if (const int forced = g_forceKnnSplit.load()) nSplit = std::max(1, std::min(std::min(forced, maxTiles), VIORB_STATIC_KNN_MAX_SPLIT));
    const int nItems = qTiles * nSplit;
    if ((size_t)nItems * tc5::QT > (size_t)VIORB_STATIC_KNN_ROWS) return fail(VIORB_E_CAPACITY, "viorb_knn2_device_static: scratch rows exceeded");
    uint2* partial = nSplit > 1 ? S.scratch : nullptr;
    VIORB_LAUNCH(VIORB_KNN_TC_KERNEL, dim3(std::min(nItems, S.sms)), dim3(tc5::THREADS), tc5::SMEM_BYTES, st, Q, T, d_idx, d_dist, (size_t)nq,
                 qTiles, 1, nItems, nSplit, partial);
    if (nSplit > 1)
        VIORB_LAUNCH(tc5::k_knn2_tc5_merge, dim3((qTiles * tc5::QT + 255) / 256, 1), dim3(256), 0, st, Q, partial, d_idx, d_dist, (size_t)nq,
                     qTiles, nSplit);
    CU(cudaGetLastError());
    return VIORB_OK;
#else
    (void)stream;
    return fail(VIORB_E_NO_DEVICE, "viorb_knn2_device_static exists only in the sm_100a build with the tcgen05 kNN-2 kernel");
#endif
}
