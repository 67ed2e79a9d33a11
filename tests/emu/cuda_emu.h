// Minimal CUDA execution-model emulator for CPU-side kernel tests.
//
// TEST INFRASTRUCTURE ONLY.  It lets the *unchanged* kernel sources in
// pfb-clean_b200/csrc be compiled with g++ (-DPFB_EMU -include cuda_emu.h) and
// run block by block on the host, so index arithmetic, barriers and shuffles
// can be checked against the oracle in the GPU-less build container.  The
// product library never includes this header and has no CPU path.
//
// Model: blocks run sequentially; the threads of a block are ucontext fibers
// scheduled round-robin on one OS thread.  __syncthreads() and the warp
// primitives are generation barriers that yield until every participant has
// arrived.  Shared memory is `static` storage (one block at a time).
#pragma once
#include <ucontext.h>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __constant__ static
#define __shared__ static
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct float2 { float x, y; };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) double2 { double x, y; };
struct int2 { int x, y; };
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
static inline double2 make_double2(double x, double y) { return double2{x, y}; }

typedef int cudaError_t;
typedef void *cudaStream_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2, cudaErrorInvalidValue = 1 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost,
                      cudaMemcpyDeviceToDevice, cudaMemcpyDefault };

namespace emu {
struct State {
    dim3 tIdx, bIdx, bDim, gDim;
    unsigned char *dyn_smem = nullptr;
    // scheduler
    ucontext_t main_ctx;
    int cur = 0;              // linear thread id running
    int nthreads = 0;
    int block_gen = 0, block_arrived = 0, alive = 0;
    std::vector<int> warp_gen, warp_arrived;
    std::vector<unsigned char> warp_slots;  // 32 lanes x 16 bytes per warp
    std::vector<unsigned> warp_ballot;
};
extern State S;
void yield();
void block_barrier();
void warp_barrier();
void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body);
}  // namespace emu

#define threadIdx (emu::S.tIdx)
#define blockIdx (emu::S.bIdx)
#define blockDim (emu::S.bDim)
#define gridDim (emu::S.gDim)
#define warpSize 32

static inline void __syncthreads() { emu::block_barrier(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::warp_barrier(); }
static inline void __threadfence() {}

template <class T> static inline T __ldg(const T *p) { return *p; }
template <class T> static inline T __ldcs(const T *p) { return *p; }
template <class T> static inline void __stcs(T *p, T v) { *p = v; }

template <class T> static inline T emu_shfl(T v, int src) {
    static_assert(sizeof(T) <= 16, "shfl payload");
    int w = emu::S.cur / 32, lane = emu::S.cur % 32;
    unsigned char *slots = emu::S.warp_slots.data() + (size_t)w * 32 * 16;
    std::memcpy(slots + lane * 16, &v, sizeof(T));
    emu::warp_barrier();
    int lanes = std::min(32, emu::S.nthreads - w * 32);
    T r = v;
    if (src >= 0 && src < lanes) std::memcpy(&r, slots + src * 16, sizeof(T));
    emu::warp_barrier();
    return r;
}
template <class T> static inline T __shfl_sync(unsigned, T v, int src, int width = 32) {
    int lane = emu::S.cur % 32;
    return emu_shfl(v, (lane / width) * width + (src % width));
}
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m, int width = 32) {
    int lane = emu::S.cur % 32;
    (void)width;
    return emu_shfl(v, lane ^ m);
}
template <class T> static inline T __shfl_down_sync(unsigned, T v, unsigned d, int width = 32) {
    int lane = emu::S.cur % 32;
    int src = lane + (int)d;
    if (src / width != lane / width) src = lane;
    return emu_shfl(v, src);
}
template <class T> static inline T __shfl_up_sync(unsigned, T v, unsigned d, int width = 32) {
    int lane = emu::S.cur % 32;
    int src = lane - (int)d;
    if (src < 0 || src / width != lane / width) src = lane;
    return emu_shfl(v, src);
}

static inline int __any_sync(unsigned, int pred) {
    int w = emu::S.cur / 32, lane = emu::S.cur % 32;
    unsigned char *slots = emu::S.warp_slots.data() + (size_t)w * 32 * 16;
    int p = pred ? 1 : 0;
    std::memcpy(slots + lane * 16, &p, sizeof(int));
    emu::warp_barrier();
    int lanes = std::min(32, emu::S.nthreads - w * 32), r = 0;
    for (int l = 0; l < lanes; ++l) { int v; std::memcpy(&v, slots + l * 16, sizeof(int)); r |= v; }
    emu::warp_barrier();
    return r;
}

template <class T> static inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; return o; }
static inline unsigned atomicOr(unsigned *p, unsigned v) { unsigned o = *p; *p = o | v; return o; }
static inline int atomicMax(int *p, int v) { int o = *p; *p = std::max(o, v); return o; }

static inline float __fmaf_rn(float a, float b, float c) { return std::fma(a, b, c); }
static inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
static inline void sincospi(double x, double *s, double *c) { *s = std::sin(M_PI * x); *c = std::cos(M_PI * x); }
static inline void sincospif(float x, float *s, float *c) { *s = (float)std::sin(M_PI * (double)x); *c = (float)std::cos(M_PI * (double)x); }
static inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }

// ---- host runtime subset (device memory == host memory) ----------------
static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
static inline cudaError_t cudaFree(void *p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void *p, int v, size_t n, cudaStream_t = 0) { std::memset(p, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = 0) { std::memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { std::memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
typedef void *cudaEvent_t;
enum { cudaEventDisableTiming = 2 };
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { *e = (void *)1; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = 0) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaMallocHost(void **p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
static inline cudaError_t cudaFreeHost(void *p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
template <class S> static inline cudaError_t cudaGetSymbolAddress(void **p, S &sym) { *p = (void *)&sym; return cudaSuccess; }
static inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
static inline const char *cudaGetErrorString(cudaError_t) { return "emu"; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return cudaSuccess; }
enum { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
#define cudaMemcpyToSymbol(sym, src, n) (std::memcpy((void *)&(sym), (src), (n)), cudaSuccess)
#define cudaMemcpyToSymbolAsync(sym, src, n, off, kind, st) (std::memcpy(((char *)&(sym)) + (off), (src), (n)), cudaSuccess)

template <class... KA, class... A>
static inline void pfb_launch(void (*k)(KA...), dim3 g, dim3 b, size_t smem, cudaStream_t, A... a) {
    emu::launch(g, b, smem, [&]() { k(a...); });
}
#define PFB_DYN_SMEM(name) unsigned char *name = emu::S.dyn_smem
