// viorb_common.cuh — shared declarations for the sm_100a ORB front-end kernels.
#pragma once

#ifdef VIORB_EMU
// g++ build of the same sources for CPU logic tests (tests/cuda_emu); never shipped.
#include "cuda_emu.h"
#define VIORB_LAUNCH(kernel, grid, block, smem, stream, ...) \
    cuda_emu::launch((grid), (block), (smem), [&]() { kernel(__VA_ARGS__); })
#define VIORB_DYN_SMEM(name) unsigned char* name = (unsigned char*)EMU_DYN_SMEM_PTR
#else
#include <cuda_runtime.h>
#include <cuda_pipeline.h>
#define VIORB_LAUNCH(kernel, grid, block, smem, stream, ...) \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define VIORB_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#endif

#include <stdint.h>

#include "../../include/viorb.h"

#define VIORB_MAX_LEVELS 16
#define VIORB_EDGE 19          // EDGE_THRESHOLD (fextractor.cpp:66)
#define VIORB_HALF_PATCH 15    // HALF_PATCH_SIZE (fextractor.cpp:65)
#define VIORB_MINB 16          // EDGE_THRESHOLD - 3 (fextractor.cpp:764)

// Geometry of one pyramid level for a given input size (host-computed once per size).
struct LevelGeom {
    int w, h, pitch;       // level image; pitch is a multiple of 16
    int off;               // byte offset inside one frame's pyramid / blur arena
    int maxBX, maxBY;      // w - 16, h - 16 (fextractor.cpp:766-767)
    int nCols, nRows, wCell, hCell, lastCol, lastRow;   // FAST cell grid (fextractor.cpp:775-778)
    unsigned cinv;         // (1 << 20) / wCell + 1: strip pixel -> cell of the strip
    int cellBase, nCells;  // slice of the cell table
    int candOff, candCap;  // slice of one frame's candidate arena (entries)
    int nWant, nIni;       // DistributeOctTree N and root count
    float hX;              // root width
    int keptOff, keptCap;  // slice of one frame's kept-keypoint arena
    int xTab, yTab;        // offsets into the resize coefficient tables (entries)
    int areaHalf;          // 1: exact 2x decimation (INTER_AREA fast path)
    int rsWords;           // 1: every 4-pixel output group reads <= 12 source bytes from its first aligned word (k_resize word path)
    float scale;           // mvScaleFactor[l]
    float kpSize;          // (float)(int)(31 * scale)
};

struct ExtractGeom {
    int nlevels, W, H;
    int iniTh, minTh;
    int totalCells;
    int candPerFrame, keptPerFrame;
    int outCap;            // per-frame output capacity (keypoints)
    int tileW, tileH;      // widest / tallest FAST strip region + 6 (k_fast's tiles have tileH rows of VIORB_FAST_TP bytes)
    int maxNodes;          // octree node capacity in shared memory
    int frame0;            // first frame of the batch this launch works on (chunked host batches)
    size_t arenaStride;    // bytes per frame in the pyramid and blur arenas
    LevelGeom lv[VIORB_MAX_LEVELS];
};

// One FAST strip: `ncell` consecutive cells of one cell row of level `level`.  [x0,x1) x [y0,y1) are the
// pixels on which cv::FAST would evaluate the corner test for those cells' ROIs (each ROI inset by 3),
// fextractor.cpp:780-806; cell k of the strip covers x0 + k*wCell .. x0 + (k+1)*wCell (the last one may
// be narrower).
#ifndef VIORB_FAST_STRIP_CELLS
#define VIORB_FAST_STRIP_CELLS 8
#endif
struct __align__(16) CellDesc {
    short level, ncell;
    short x0, y0, x1, y1;
    short nC, ng;          // 16-byte chunks per staged tile row; aligned 4-pixel words per region row
    unsigned invNC;        // (1 << 20) / nC + 1: chunk index -> tile row
    unsigned ginv;         // (1 << 22) / (words per region row) + 1: word index -> region row
    unsigned pad2[2];
};

// Where the images of a batch live.
struct FrameSrc {
    const uint8_t* l0;     // level 0 = the input images
    int l0Pitch;
    size_t l0Stride;
    uint8_t* pyr;          // levels >= 1 (arena; level 0 slot unused)
    uint8_t* blur;         // blurred levels (arena, all levels)
};

// Keypoint retained by the quadtree, before output ordering.
struct KeptKp {
    short x, y;            // level pixels
    int score;
    float angle;           // degrees
    float c, s;            // cos / sin of the angle (float32-rounded)
};

__device__ __forceinline__ const uint8_t* level_ptr(const ExtractGeom& g, const FrameSrc& fs, int l, int f, int& pitch) {
    if (l == 0) { pitch = fs.l0Pitch; return fs.l0 + (size_t)f * fs.l0Stride; }
    pitch = g.lv[l].pitch;
    return fs.pyr + (size_t)f * g.arenaStride + g.lv[l].off;
}

// ---- warp-level integer MMA (IMMA.16832.U8.U8) --------------------------------------------------------
// D(16x8, s32) += A(16x32, u8, row) * B(32x8, u8, col); PTX fragment layout of mma.m16n8k32 (g = lane >> 2, t = lane & 3):
// a0 (g, 4t..4t+3)  a1 (g+8, 4t..)  a2 (g, 16+4t..)  a3 (g+8, 16+4t..);  b0 (k 4t.., n g)  b1 (k 16+4t.., n g);
// c0 (g, 2t)  c1 (g, 2t+1)  c2 (g+8, 2t)  c3 (g+8, 2t+1).
__device__ __forceinline__ void mma_u8_16832(int c[4], const unsigned a[4], unsigned b0, unsigned b1) {
#ifdef VIORB_EMU
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    auto dot4 = [](unsigned x, unsigned y) {
        int s = 0;
        for (int i = 0; i < 4; ++i) s += (int)((x >> (8 * i)) & 0xFF) * (int)((y >> (8 * i)) & 0xFF);
        return s;
    };
    for (int tp = 0; tp < 4; ++tp) {
        const unsigned A0 = __shfl_sync(0xffffffffu, a[0], 4 * g + tp), A1 = __shfl_sync(0xffffffffu, a[1], 4 * g + tp);
        const unsigned A2 = __shfl_sync(0xffffffffu, a[2], 4 * g + tp), A3 = __shfl_sync(0xffffffffu, a[3], 4 * g + tp);
        for (int j = 0; j < 2; ++j) {
            const unsigned B0 = __shfl_sync(0xffffffffu, b0, 4 * (2 * t + j) + tp), B1 = __shfl_sync(0xffffffffu, b1, 4 * (2 * t + j) + tp);
            c[j] += dot4(A0, B0) + dot4(A2, B1);
            c[2 + j] += dot4(A1, B0) + dot4(A3, B1);
        }
    }
#else
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
#endif
}

