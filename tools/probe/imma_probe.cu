// Micro-benchmark: throughput of legacy warp-level integer MMA (mma.sync m16n8k32 s8/u8 -> s32) on sm_100a,
// to decide whether a {0,1}-expanded Hamming kNN on tensor cores can beat the LOP3/POPC kernel.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int iters, int* out) {
    int c[8][4] = {};
    unsigned a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    int s = 0;
    for (int j = 0; j < 8; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    int* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int warps = 4; warps <= 16; warps *= 2) {
        const int iters = 20000, blocks = 148 * 4;
        k<<<blocks, warps * 32>>>(100, d);
        cudaEventRecord(e0);
        k<<<blocks, warps * 32>>>(iters, d);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double mmas = (double)blocks * warps * iters * 8;
        double macs = mmas * 16 * 8 * 32;
        printf("warps/block %d: %.3f ms, %.1f T int8 MAC/s (%.1f TOPS), %.1f MMA/clk/SM @1.965GHz\n", warps, ms, macs / ms / 1e9,
               2 * macs / ms / 1e9, mmas / (ms * 1e-3) / 148 / 1.965e9);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
