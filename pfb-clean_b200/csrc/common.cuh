// Shared helpers for the pfb_b200 CUDA library (sm_100a only).
#pragma once
#include "../../include/pfb_b200.h"

#ifndef PFB_EMU
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#endif
#include <string>

namespace pfb {

extern thread_local std::string g_last_error;
extern long long g_launch_count;

int fail(int status, const char *fmt, ...);

#ifndef PFB_EMU
template <class... KA, class... A>
static inline void pfb_launch(void (*k)(KA...), dim3 g, dim3 b, size_t smem,
                              cudaStream_t st, A... a) {
    k<<<g, b, smem, st>>>(a...);
}
#define PFB_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#else
using ::pfb_launch;
#endif

// Optional per-kernel timing (pfb_profile_*): when enabled, every launch is
// bracketed by two CUDA events recorded on the launching stream; the report
// groups them by kernel symbol.  Off by default (zero overhead).
extern int g_profile_on;
void profile_begin(const void *func, cudaStream_t st);
void profile_end(cudaStream_t st);

// every kernel launch of the library goes through here (counted)
template <class... KA, class... A>
static inline void launch(void (*k)(KA...), dim3 g, dim3 b, size_t smem,
                          cudaStream_t st, A... a) {
    ++g_launch_count;
#ifndef PFB_EMU
    if (g_profile_on) profile_begin((const void *)k, st);
#endif
    pfb_launch(k, g, b, smem, st, a...);
#ifndef PFB_EMU
    if (g_profile_on) profile_end(st);
#endif
}

#define PFB_CUDA_OK(expr)                                                      \
    do {                                                                       \
        cudaError_t e__ = (expr);                                              \
        if (e__ != cudaSuccess)                                                \
            return pfb::fail(PFB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,     \
                             cudaGetErrorString(e__), __FILE__, __LINE__);     \
    } while (0)

#define PFB_CHECK_LAUNCH()                                                     \
    do {                                                                       \
        cudaError_t e__ = cudaPeekAtLastError();                               \
        if (e__ != cudaSuccess)                                                \
            return pfb::fail(PFB_ERR_CUDA, "kernel launch failed: %s (%s:%d)", \
                             cudaGetErrorString(e__), __FILE__, __LINE__);     \
    } while (0)

static inline cudaStream_t as_stream(void *s) { return (cudaStream_t)s; }

template <class T> struct vec2;
template <> struct vec2<float> { typedef float2 type; };
template <> struct vec2<double> { typedef double2 type; };

static inline size_t dtype_size(int dtype) { return dtype == PFB_F64 ? 8 : 4; }

static inline int cdiv(long a, long b) { return (int)((a + b - 1) / b); }

// B200: 148 SMs.  Grid-stride kernels use a multiple of the SM count.
#ifdef PFB_EMU
static const int kNumSM = 2;  // CPU emulator (tests only): keep grids tiny
#else
static const int kNumSM = 148;
#endif

// ---------------------------------------------------------------------------
// cp.async (LDGSTS) helpers: global -> shared copies that tie up no registers
// while in flight; bytes beyond `src_bytes` are zero-filled by the copy itself.
template <int BYTES>
__device__ __forceinline__ void cp_async_zfill(void *smem_dst, const void *gsrc, int src_bytes) {
#ifdef PFB_EMU
    unsigned char tmp[BYTES];
    for (int i = 0; i < BYTES; ++i) tmp[i] = i < src_bytes ? ((const unsigned char *)gsrc)[i] : 0;
    for (int i = 0; i < BYTES; ++i) ((unsigned char *)smem_dst)[i] = tmp[i];
#else
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
    if (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sa), "l"(gsrc), "r"(src_bytes) : "memory");
    else if (BYTES == 4)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(sa), "l"(gsrc), "r"(src_bytes) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(sa), "l"(gsrc), "r"(src_bytes) : "memory");
#endif
}
// L2 prefetch of the 128-byte line holding `p` (no register, no scoreboard)
__device__ __forceinline__ void prefetch_l2(const void *p) {
#ifndef PFB_EMU
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    (void)p;
#endif
}
__device__ __forceinline__ void cp_async_commit() {
#ifndef PFB_EMU
    asm volatile("cp.async.commit_group;" ::: "memory");
#endif
}
template <int N> __device__ __forceinline__ void cp_async_wait() {
#ifndef PFB_EMU
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
#endif
}

}  // namespace pfb
