// cuda_emu.h — a tiny single-process CUDA execution shim for LOGIC TESTS on a
// box without a GPU.  TEST INFRASTRUCTURE ONLY (lives under tests/): the product
// library is built by nvcc from the same .cu sources; this header lets g++ build
// those sources into tests/cuda_emu/_build/libviorb_emu.so so that index math,
// barriers and the host-side orchestration can be checked against the oracle
// before spending GPU time.  Nothing in vi_slam_b200/, bench.py or
// __graft_entry__.py loads the emulated library.
//
// Model: a kernel launch runs CTAs one after another; inside a CTA every CUDA
// thread is a ucontext fiber.  __syncthreads*() and the *_sync warp collectives
// are cooperative yield points with the usual "all participating threads arrive"
// release rule, so divergence / barrier bugs show up as a deadlock report.
// Atomics are plain read-modify-writes (fibers never run concurrently).
#pragma once
#include <ucontext.h>

#include <algorithm>
#include <cmath>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ __attribute__((noinline))
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static
#define __align__(n) __attribute__((aligned(n)))

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct uchar4 { unsigned char x, y, z, w; };
struct char4 { signed char x, y, z, w; };
struct uchar2 { unsigned char x, y; };
struct ushort2 { unsigned short x, y; };
struct short2 { short x, y; };
struct int2 { int x, y; };
struct uint2 { unsigned x, y; };
struct int4 { int x, y, z, w; };
struct __attribute__((aligned(16))) uint4 { unsigned x, y, z, w; };
struct float2 { float x, y; };
struct __attribute__((aligned(16))) float4 { float x, y, z, w; };
static inline float4 make_float4(float a, float b, float c, float d) { return float4{a, b, c, d}; }
static inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { return uint4{a, b, c, d}; }
static inline uint2 make_uint2(unsigned a, unsigned b) { return uint2{a, b}; }
static inline int2 make_int2(int a, int b) { return int2{a, b}; }
static inline int4 make_int4(int a, int b, int c, int d) { return int4{a, b, c, d}; }
static inline uchar4 make_uchar4(unsigned char a, unsigned char b, unsigned char c, unsigned char d) { return uchar4{a, b, c, d}; }
static inline short2 make_short2(short a, short b) { return short2{a, b}; }
static inline float2 make_float2(float a, float b) { return float2{a, b}; }

typedef int cudaError_t;
typedef struct emu_stream_s* cudaStream_t;
typedef struct emu_event_s { double t; }* cudaEvent_t;
enum { cudaSuccess = 0, cudaErrorEmu = 1, cudaErrorNotReady = 600 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum { cudaStreamNonBlocking = 1, cudaHostAllocDefault = 0, cudaEventDefault = 0, cudaEventDisableTiming = 2, cudaEventBlockingSync = 1 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
struct cudaDeviceProp { int multiProcessorCount; char name[64]; size_t sharedMemPerBlockOptin; int major, minor; };

static inline const char* cudaGetErrorString(cudaError_t e) { return e ? "emu error" : "no error"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
static inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return cudaSuccess; }
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int) {
    std::memset(p, 0, sizeof(*p)); p->multiProcessorCount = 148; std::strcpy(p->name, "emu"); p->sharedMemPerBlockOptin = 227 * 1024;
    p->major = 10; return cudaSuccess;
}
static inline cudaError_t cudaMalloc(void** p, size_t n) { *p = std::calloc(1, n ? n : 1); return *p ? cudaSuccess : cudaErrorEmu; }
template <class T> static inline cudaError_t cudaMalloc(T** p, size_t n) { return cudaMalloc((void**)p, n); }
static inline cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
template <class T> static inline cudaError_t cudaMallocHost(T** p, size_t n) { return cudaMalloc((void**)p, n); }
static inline cudaError_t cudaHostAlloc(void** p, size_t n, unsigned) { return cudaMalloc(p, n); }
static inline cudaError_t cudaFreeHost(void* p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { std::memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = 0) { std::memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpy2DAsync(void* d, size_t dp, const void* s, size_t sp, size_t w, size_t h, cudaMemcpyKind, cudaStream_t = 0) {
    for (size_t y = 0; y < h; ++y) std::memmove((char*)d + y * dp, (const char*)s + y * sp, w);
    return cudaSuccess;
}
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = 0) { std::memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemset(void* d, int v, size_t n) { std::memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = nullptr; return cudaSuccess; }
static inline cudaError_t cudaStreamQuery(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = nullptr; return cudaSuccess; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new emu_event_s{0}; return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { return cudaEventCreate(e); }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = 0) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventQuery(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return cudaSuccess; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return cudaSuccess; }
#define cudaMemcpyToSymbol(sym, src, n) (std::memcpy((void*)&(sym), (src), (n)), cudaSuccess)

namespace cuda_emu {

enum WaitKind { RUN = 0, WAIT_BLOCK = 1, WAIT_WARP = 2, DONE = 3 };

struct Fiber {
    ucontext_t ctx;
    char* stack = nullptr;
    int state = DONE;
    unsigned warp_mask = 0;
    dim3 tid;
    int linear = 0;
};

struct WarpX {                 // exchange area of one warp
    unsigned long long val[32];
    int pred[32];
    unsigned long long res_val[32];
    int res_pred[32];
};

struct Ctx {
    ucontext_t sched;
    std::vector<Fiber> fibers;
    std::vector<WarpX> warps;
    int cur = -1;
    int nthreads = 0;
    int block_pred_or = 0, block_pred_and = 1, block_pred_cnt = 0;
    int res_or = 0, res_and = 0, res_cnt = 0;
    dim3 gridDim_, blockDim_, blockIdx_;
    void (*entry)(void*) = nullptr;
    void* arg = nullptr;
    unsigned char* dyn_smem = nullptr;
    size_t dyn_cap = 0;
};
inline Ctx& C() { static Ctx c; return c; }
inline std::mutex& launch_mutex() { static std::mutex m; return m; }
static const size_t kStack = 96 * 1024;

inline void yield_to_sched() {
    Ctx& c = C();
    swapcontext(&c.fibers[c.cur].ctx, &c.sched);
}
inline void trampoline() {
    Ctx& c = C();
    c.entry(c.arg);
    c.fibers[c.cur].state = DONE;
    swapcontext(&c.fibers[c.cur].ctx, &c.sched);
}

inline void run_block() {
    Ctx& c = C();
    const int n = c.nthreads;
    if ((int)c.fibers.size() < n) {
        size_t old = c.fibers.size();
        c.fibers.resize(n);
        for (size_t i = old; i < (size_t)n; ++i) c.fibers[i].stack = (char*)std::malloc(kStack);
    }
    c.warps.resize((n + 31) / 32);
    for (int i = 0; i < n; ++i) {
        Fiber& f = c.fibers[i];
        getcontext(&f.ctx);
        f.ctx.uc_stack.ss_sp = f.stack;
        f.ctx.uc_stack.ss_size = kStack;
        f.ctx.uc_link = nullptr;
        makecontext(&f.ctx, (void (*)())trampoline, 0);
        f.state = RUN;
        f.linear = i;
        f.tid = dim3(i % c.blockDim_.x, (i / c.blockDim_.x) % c.blockDim_.y, i / (c.blockDim_.x * c.blockDim_.y));
    }
    c.block_pred_or = 0; c.block_pred_and = 1; c.block_pred_cnt = 0;
    for (;;) {
        bool progressed = false;
        int alive = 0;
        for (int i = 0; i < n; ++i) {
            if (c.fibers[i].state == RUN) {
                c.cur = i;
                swapcontext(&c.sched, &c.fibers[i].ctx);
                progressed = true;
            }
            if (c.fibers[i].state != DONE) ++alive;
        }
        if (!alive) break;
        // block barrier: every live thread waits at it
        int wb = 0;
        for (int i = 0; i < n; ++i) wb += c.fibers[i].state == WAIT_BLOCK;
        if (wb == alive) {
            c.res_or = c.block_pred_or; c.res_and = c.block_pred_and; c.res_cnt = c.block_pred_cnt;
            c.block_pred_or = 0; c.block_pred_and = 1; c.block_pred_cnt = 0;
            for (int i = 0; i < n; ++i) if (c.fibers[i].state == WAIT_BLOCK) c.fibers[i].state = RUN;
            continue;
        }
        // warp collectives: all live lanes named in the mask have arrived
        bool released = false;
        for (int w = 0; w < (n + 31) / 32; ++w) {
            unsigned waiting = 0, live = 0, want = 0;
            for (int l = 0; l < 32 && w * 32 + l < n; ++l) {
                Fiber& f = c.fibers[w * 32 + l];
                if (f.state != DONE) live |= 1u << l;
                if (f.state == WAIT_WARP) { waiting |= 1u << l; want |= f.warp_mask; }
            }
            if (waiting && (want & live & ~waiting) == 0) {
                WarpX& x = c.warps[w];
                std::memcpy(x.res_val, x.val, sizeof(x.val));
                std::memcpy(x.res_pred, x.pred, sizeof(x.pred));
                for (int l = 0; l < 32 && w * 32 + l < n; ++l)
                    if (c.fibers[w * 32 + l].state == WAIT_WARP) c.fibers[w * 32 + l].state = RUN;
                released = true;
            }
        }
        if (released) continue;
        if (!progressed) {
            std::fprintf(stderr, "cuda_emu: deadlock in block (%u,%u,%u): %d live, %d at __syncthreads\n",
                         c.blockIdx_.x, c.blockIdx_.y, c.blockIdx_.z, alive, wb);
            std::abort();
        }
    }
}

template <class F>
inline void launch(dim3 grid, dim3 block, size_t smem, F&& body) {
    std::lock_guard<std::mutex> g(launch_mutex());
    Ctx& c = C();
    c.gridDim_ = grid; c.blockDim_ = block;
    c.nthreads = (int)(block.x * block.y * block.z);
    if (smem > c.dyn_cap) { std::free(c.dyn_smem); c.dyn_smem = (unsigned char*)std::aligned_alloc(128, (smem + 127) / 128 * 128); c.dyn_cap = smem; }
    auto fn = [](void* p) { (*(F*)p)(); };
    c.entry = fn; c.arg = (void*)&body;
    for (unsigned z = 0; z < grid.z; ++z)
        for (unsigned y = 0; y < grid.y; ++y)
            for (unsigned x = 0; x < grid.x; ++x) {
                c.blockIdx_ = dim3(x, y, z);
                if (smem) std::memset(c.dyn_smem, 0xCD, smem);
                run_block();
            }
}

// ---- warp collective plumbing ----
inline int lane_id() { return C().fibers[C().cur].linear & 31; }
inline WarpX& my_warp() { return C().warps[C().fibers[C().cur].linear >> 5]; }
inline void warp_arrive(unsigned mask, unsigned long long v, int p) {
    Ctx& c = C();
    Fiber& f = c.fibers[c.cur];
    WarpX& x = my_warp();
    x.val[lane_id()] = v; x.pred[lane_id()] = p;
    f.warp_mask = mask; f.state = WAIT_WARP;
    yield_to_sched();
}
}  // namespace cuda_emu

#define threadIdx (cuda_emu::C().fibers[cuda_emu::C().cur].tid)
#define blockIdx (cuda_emu::C().blockIdx_)
#define blockDim (cuda_emu::C().blockDim_)
#define gridDim (cuda_emu::C().gridDim_)
static const int warpSize = 32;

static inline void __syncthreads() {
    auto& c = cuda_emu::C();
    c.fibers[c.cur].state = cuda_emu::WAIT_BLOCK;
    cuda_emu::yield_to_sched();
}
static inline int __syncthreads_or(int p) {
    auto& c = cuda_emu::C(); c.block_pred_or |= (p != 0); __syncthreads(); return c.res_or;
}
static inline int __syncthreads_and(int p) {
    auto& c = cuda_emu::C(); c.block_pred_and &= (p != 0); __syncthreads(); return c.res_and;
}
static inline int __syncthreads_count(int p) {
    auto& c = cuda_emu::C(); c.block_pred_cnt += (p != 0); __syncthreads(); return c.res_cnt;
}
static inline void __syncwarp(unsigned mask = 0xffffffffu) { cuda_emu::warp_arrive(mask, 0, 0); }
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline unsigned __activemask() { return 0xffffffffu; }   // only used with full warps in this code base

static inline unsigned __ballot_sync(unsigned mask, int pred) {
    cuda_emu::warp_arrive(mask, 0, pred != 0);
    auto& x = cuda_emu::my_warp();
    auto& c = cuda_emu::C();
    int base = (c.fibers[c.cur].linear >> 5) * 32;
    unsigned r = 0;
    for (int l = 0; l < 32; ++l)
        if ((mask >> l & 1) && base + l < c.nthreads && x.res_pred[l]) r |= 1u << l;
    return r;
}
static inline int __any_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) != 0; }
static inline int __all_sync(unsigned mask, int pred) { return (__ballot_sync(mask, !pred) == 0); }
template <class T> static inline T __shfl_sync(unsigned mask, T v, int src, int width = 32) {
    unsigned long long raw = 0; std::memcpy(&raw, &v, sizeof(T));
    cuda_emu::warp_arrive(mask, raw, 0);
    int lane = cuda_emu::lane_id();
    int s = (lane / width) * width + (src % width + width) % width;
    T out; unsigned long long r = cuda_emu::my_warp().res_val[s]; std::memcpy(&out, &r, sizeof(T)); return out;
}
template <class T> static inline T __shfl_xor_sync(unsigned mask, T v, int lm, int width = 32) {
    unsigned long long raw = 0; std::memcpy(&raw, &v, sizeof(T));
    cuda_emu::warp_arrive(mask, raw, 0);
    int lane = cuda_emu::lane_id();
    int s = lane ^ lm; if (s / width != lane / width) s = lane;
    T out; unsigned long long r = cuda_emu::my_warp().res_val[s]; std::memcpy(&out, &r, sizeof(T)); return out;
}
template <class T> static inline T __shfl_down_sync(unsigned mask, T v, unsigned d, int width = 32) {
    unsigned long long raw = 0; std::memcpy(&raw, &v, sizeof(T));
    cuda_emu::warp_arrive(mask, raw, 0);
    int lane = cuda_emu::lane_id();
    int s = lane + (int)d; if (s / width != lane / width) s = lane;
    T out; unsigned long long r = cuda_emu::my_warp().res_val[s]; std::memcpy(&out, &r, sizeof(T)); return out;
}
template <class T> static inline T __shfl_up_sync(unsigned mask, T v, unsigned d, int width = 32) {
    unsigned long long raw = 0; std::memcpy(&raw, &v, sizeof(T));
    cuda_emu::warp_arrive(mask, raw, 0);
    int lane = cuda_emu::lane_id();
    int s = lane - (int)d; if (s < (lane / width) * width) s = lane;
    T out; unsigned long long r = cuda_emu::my_warp().res_val[s]; std::memcpy(&out, &r, sizeof(T)); return out;
}
template <class T> static inline unsigned __match_any_sync(unsigned mask, T v) {
    unsigned long long raw = 0; std::memcpy(&raw, &v, sizeof(T));
    cuda_emu::warp_arrive(mask, raw, 1);
    auto& x = cuda_emu::my_warp();
    unsigned r = 0;
    for (int l = 0; l < 32; ++l) if ((mask >> l & 1) && x.res_val[l] == raw) r |= 1u << l;
    return r;
}
static inline int __reduce_add_sync(unsigned mask, int v) {
    cuda_emu::warp_arrive(mask, (unsigned long long)(long long)v, 0);
    auto& x = cuda_emu::my_warp(); long long s = 0;
    for (int l = 0; l < 32; ++l) if (mask >> l & 1) s += (long long)x.res_val[l];
    return (int)s;
}
static inline unsigned __reduce_or_sync(unsigned mask, unsigned v) {
    cuda_emu::warp_arrive(mask, v, 0);
    auto& x = cuda_emu::my_warp(); unsigned s = 0;
    for (int l = 0; l < 32; ++l) if (mask >> l & 1) s |= (unsigned)x.res_val[l];
    return s;
}
static inline int __reduce_max_sync(unsigned mask, int v) {
    cuda_emu::warp_arrive(mask, (unsigned long long)(long long)v, 0);
    auto& x = cuda_emu::my_warp(); long long s = LLONG_MIN;
    for (int l = 0; l < 32; ++l) if (mask >> l & 1) s = std::max(s, (long long)x.res_val[l]);
    return (int)s;
}
static inline int __reduce_min_sync(unsigned mask, int v) {
    cuda_emu::warp_arrive(mask, (unsigned long long)(long long)v, 0);
    auto& x = cuda_emu::my_warp(); long long s = LLONG_MAX;
    for (int l = 0; l < 32; ++l) if (mask >> l & 1) s = std::min(s, (long long)x.res_val[l]);
    return (int)s;
}

static inline unsigned __reduce_min_sync(unsigned mask, unsigned v) {
    cuda_emu::warp_arrive(mask, (unsigned long long)v, 0);
    auto& x = cuda_emu::my_warp(); unsigned long long s = ~0ull;
    for (int l = 0; l < 32; ++l) if (mask >> l & 1) s = std::min(s, (unsigned long long)x.res_val[l]);
    return (unsigned)s;
}

// ---- atomics (fibers are cooperative: plain RMW) ----
template <class T> static inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
template <class T> static inline T atomicSub(T* p, T v) { T o = *p; *p = o - v; return o; }
template <class T> static inline T atomicMax(T* p, T v) { T o = *p; *p = std::max(o, v); return o; }
template <class T> static inline T atomicMin(T* p, T v) { T o = *p; *p = std::min(o, v); return o; }
template <class T> static inline T atomicOr(T* p, T v) { T o = *p; *p = o | v; return o; }
template <class T> static inline T atomicAnd(T* p, T v) { T o = *p; *p = o & v; return o; }
template <class T> static inline T atomicExch(T* p, T v) { T o = *p; *p = v; return o; }
template <class T> static inline T atomicCAS(T* p, T c, T v) { T o = *p; if (o == c) *p = v; return o; }

// ---- intrinsics ----
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline unsigned __brev(unsigned v) { unsigned r = 0; for (int i = 0; i < 32; ++i) r |= (v >> i & 1) << (31 - i); return r; }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s) {
    unsigned long long v = ((unsigned long long)hi << 32) | lo; return (unsigned)(v >> (s & 31));
}
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned s) {
    unsigned long long v = ((unsigned long long)hi << 32) | lo; return (unsigned)((v << (s & 31)) >> 32);
}
static inline unsigned __byte_perm(unsigned a, unsigned b, unsigned sel) {
    unsigned long long v = ((unsigned long long)b << 32) | a; unsigned r = 0;
    for (int i = 0; i < 4; ++i) {
        unsigned s = (sel >> (4 * i)) & 0xF; unsigned byte = (unsigned)(v >> (8 * (s & 7))) & 0xFF;
        if (s & 8) byte = (byte & 0x80) ? 0xFF : 0x00;
        r |= byte << (8 * i);
    }
    return r;
}
#pragma GCC push_options
#pragma GCC optimize("fp-contract=off")
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __ddiv_rn(double a, double b) { return a / b; }
static inline float __double2float_rn(double a) { return (float)a; }
static inline double __dsqrt_rn(double a) { return std::sqrt(a); }
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }
#pragma GCC pop_options
static inline float __int_as_float(int v) { float f; std::memcpy(&f, &v, 4); return f; }
static inline int __float_as_int(float f) { int v; std::memcpy(&v, &f, 4); return v; }
static inline int __float2int_rn(float v) { return (int)lrintf(v); }
static inline int __float2int_rz(float v) { return (int)v; }
static inline int __float2int_rd(float v) { return (int)std::floor(v); }
static inline int __float2int_ru(float v) { return (int)std::ceil(v); }
static inline float __int2float_rn(int v) { return (float)v; }
static inline int __double2int_rn(double v) { return (int)lrint(v); }
static inline int __vimin3_s32(int a, int b, int c) { return std::min(a, std::min(b, c)); }
static inline int __vimax3_s32(int a, int b, int c) { return std::max(a, std::max(b, c)); }
static inline unsigned __vimin3_u32(unsigned a, unsigned b, unsigned c) { return std::min(a, std::min(b, c)); }
static inline unsigned __vimax3_u32(unsigned a, unsigned b, unsigned c) { return std::max(a, std::max(b, c)); }
// packed 16x2 (DPX-style) intrinsics
static inline unsigned emu_pack16(int lo, int hi) { return ((unsigned)lo & 0xFFFFu) | ((unsigned)hi << 16); }
static inline int emu_lo16(unsigned v) { return (short)(v & 0xFFFF); }
static inline int emu_hi16(unsigned v) { return (short)(v >> 16); }
static inline unsigned __vadd2(unsigned a, unsigned b) { return emu_pack16(emu_lo16(a) + emu_lo16(b), emu_hi16(a) + emu_hi16(b)); }
static inline unsigned __vsub2(unsigned a, unsigned b) { return emu_pack16(emu_lo16(a) - emu_lo16(b), emu_hi16(a) - emu_hi16(b)); }
static inline unsigned __vmaxs2(unsigned a, unsigned b) { return emu_pack16(std::max(emu_lo16(a), emu_lo16(b)), std::max(emu_hi16(a), emu_hi16(b))); }
static inline unsigned __vmins2(unsigned a, unsigned b) { return emu_pack16(std::min(emu_lo16(a), emu_lo16(b)), std::min(emu_hi16(a), emu_hi16(b))); }
static inline unsigned __vimax3_s16x2(unsigned a, unsigned b, unsigned c) { return __vmaxs2(__vmaxs2(a, b), c); }
static inline unsigned __vimin3_s16x2(unsigned a, unsigned b, unsigned c) { return __vmins2(__vmins2(a, b), c); }
static inline unsigned __dp2a_lo(unsigned a, unsigned b, unsigned c) { return c + (a & 0xFFFFu) * (b & 0xFFu) + (a >> 16) * ((b >> 8) & 0xFFu); }
static inline unsigned __dp4a(unsigned a, unsigned b, unsigned c) {
    for (int i = 0; i < 4; ++i) c += ((a >> (8 * i)) & 0xFF) * ((b >> (8 * i)) & 0xFF);
    return c;
}
static inline int __sad(int a, int b, unsigned c) { return (int)(c + (unsigned)std::abs(a - b)); }
using std::max;
using std::min;

static inline void __pipeline_memcpy_async(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
static inline void __pipeline_commit() {}
static inline void __pipeline_wait_prior(int) {}
#define EMU_DYN_SMEM_PTR ((void*)cuda_emu::C().dyn_smem)
