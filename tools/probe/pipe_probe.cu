// pipe_probe.cu - which issue pipe do the packed 16-bit min/max forms use on sm_100a?
// Throughput of long independent chains, alone and interleaved: two instruction kinds on DIFFERENT pipes overlap
// (time of the mix ~ max of the two), on the SAME pipe they add up.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe pipe_probe.cu && ./pipe_probe
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>

#define CHAINS 8
__device__ __forceinline__ unsigned hmin2u(unsigned a, unsigned b) {
    __half2 r = __hmin2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
    return *reinterpret_cast<unsigned*>(&r);
}
__device__ __forceinline__ unsigned hmax2u(unsigned a, unsigned b) {
    __half2 r = __hmax2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
    return *reinterpret_cast<unsigned*>(&r);
}
__device__ __forceinline__ unsigned hsub2u(unsigned a, unsigned b) {
    __half2 r = __hsub2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
    return *reinterpret_cast<unsigned*>(&r);
}
// A: packed integer op, B: second op kind.  MODE bit 0: run A, bit 1: run B.  KA / KB choose the instruction.
template <int KA, int KB, int MODE>
__global__ void __launch_bounds__(256) k(unsigned* out, int iters, unsigned seed) {
    unsigned x[CHAINS], y[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { x[c] = seed * (threadIdx.x + 1) + c * 0x01010101u; y[c] = (seed ^ (c * 0x00370013u + threadIdx.x)) & 0x03ff03ffu; }
    const unsigned m = seed | 0x00010001u, m2 = (seed * 3 + 1) & 0x03ff03ffu;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (MODE & 1) {
                if (KA == 0) x[c] = __vmins2(x[c], m + c);                       // VIMNMX.S16x2
                if (KA == 1) x[c] = __vimin3_s16x2(x[c], m + c, m);              // VIMNMX3.S16x2
                if (KA == 2) x[c] = __byte_perm(x[c], m, 0x4140 + c);            // PRMT
                if (KA == 3) x[c] = (x[c] & m) ^ (m + c);                        // LOP3
            }
            if (MODE & 2) {
                if (KB == 0) y[c] = hmin2u(y[c], m2 + c);                        // HMNMX2
                if (KB == 1) y[c] = y[c] * m + c;                                // IMAD
                if (KB == 2) y[c] = hsub2u(y[c], m2);                            // HADD2
                if (KB == 3) y[c] = hmax2u(hmin2u(y[c], m2 + c), m2);            // 2 x HMNMX2
            }
        }
    }
    unsigned acc = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc ^= x[c] ^ y[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int KA, int KB, int MODE>
static float run(unsigned* d, int iters) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<KA, KB, MODE><<<148 * 8, 256>>>(d, 64, 12345u);
    cudaEventRecord(a);
    k<KA, KB, MODE><<<148 * 8, 256>>>(d, iters, 12345u);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
#define TRIPLE(KA, KB, name) { float ta = run<KA, KB, 1>(d, it), tb = run<KA, KB, 2>(d, it), tab = run<KA, KB, 3>(d, it); \
    printf("%-28s A %.3f ms  B %.3f ms  A+B interleaved %.3f ms  -> %s\n", name, ta, tb, tab, tab < 0.8f * (ta + tb) ? "different pipes" : "same pipe"); }
int main() {
    unsigned* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    const int it = 4096;
    TRIPLE(0, 0, "VIMNMX.S16x2 | HMNMX2");
    TRIPLE(1, 0, "VIMNMX3.S16x2 | HMNMX2");
    TRIPLE(0, 1, "VIMNMX.S16x2 | IMAD");
    TRIPLE(0, 2, "VIMNMX.S16x2 | HADD2");
    TRIPLE(2, 0, "PRMT | HMNMX2");
    TRIPLE(3, 0, "LOP3 | HMNMX2");
    TRIPLE(1, 3, "VIMNMX3.S16x2 | 2xHMNMX2");
    { float tb = run<0, 1, 2>(d, it), tc = run<0, 0, 2>(d, it); printf("IMAD alone %.3f ms, HMNMX2 alone %.3f ms\n", tb, tc); }
    // ops per clock per SM: 148*8 CTAs * 256 thr * CHAINS * it ops
    printf("ops per launch per kind: %.3e\n", 148.0 * 8 * 256 * CHAINS * it);
    cudaError_t e = cudaDeviceSynchronize(); printf("%s\n", cudaGetErrorString(e));
    return 0;
}
