// Compile-time Stockham FFT core (power-of-two N >= 512, 16 points/thread).
//
// Same algorithm and register mapping as fft.cuh (thread t of the NT = N/16
// threads of a transform holds x[t + e*NT], radix-16 stages first, a 2/4/8
// remainder last) but with N, the stage plan and every shared-memory offset
// known at compile time, so that
//   * the exchange addresses are one per-thread base + immediates,
//   * twiddles come from a small per-stage table [(r-1)*Ns + k] that the
//     calling kernel keeps in SHARED memory (k = t mod Ns is per-thread
//     constant for all stages with Ns <= NT: one LDS per twiddle, no index
//     arithmetic) instead of a strided gather from a length-N global table,
//   * the barrier of an exchange is a NAMED barrier over the NT threads of the
//     transform only, so several transforms in one CTA do not wait on each
//     other.
// The first ncu profile of the generic kernels (profiles/r01a_*) showed ~40 %
// of the issued instructions were integer address arithmetic and the top stall
// was the twiddle gather; this core removes both.
#pragma once
#include "fft.cuh"

namespace pfb {

// named barrier over `n` threads (n % 32 == 0); id 1..15 (0 = __syncthreads)
struct GroupBar {
    int id, n;
    __device__ __forceinline__ void operator()() const {
#ifdef PFB_EMU
        __syncthreads();   // emulator: groups execute symmetric barrier sequences
#else
        asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
#endif
    }
};

__host__ __device__ constexpr int ct_radix(int N, int Ns) { return (N / Ns) >= 16 ? 16 : (N / Ns); }
// number of twiddle-table entries for length N: sum over stages with Ns > 1 of (R-1)*Ns
__host__ __device__ constexpr int ct_tab_size(int N) {
    int s = 0;
    for (int Ns = 1; Ns < N; Ns *= ct_radix(N, Ns))
        if (Ns > 1) s += (ct_radix(N, Ns) - 1) * Ns;
    return s;
}
// shared-memory elements of one transform line: padded (i + i/16) plus a skew
// of 4 so that lines of neighbouring rows start 4 complex banks apart (makes
// the row-interleaved post-processing of the real transforms conflict free)
__host__ __device__ constexpr int ct_stride(int N) { return N + N / 16 + 4; }

__device__ __forceinline__ int ct_pad(int i) { return i + (i >> 4); }

template <class T, int N, int Ns, int DIR, class Sync>
__device__ __forceinline__ void ct_stages(Cx<T> (&v)[16], const int t, Cx<T> *buf,
                                          const Cx<T> *tab, const Sync &sync) {
    constexpr int NT = N / 16;
    constexpr int R = ct_radix(N, Ns);
    constexpr int B = 16 / R;
    constexpr bool last = (Ns * R == N);
    // ---- twiddles
    if constexpr (Ns > 1) {
        if constexpr (Ns <= NT) {            // k = t mod Ns for every butterfly of the thread
            const int k = t & (Ns - 1);
#pragma unroll
            for (int r = 1; r < R; ++r) {
                const Cx<T> w = twd<T, DIR>(tab[(r - 1) * Ns + k]);
#pragma unroll
                for (int b = 0; b < B; ++b) v[b + r * B] = cmul(v[b + r * B], w);
            }
        } else {
#pragma unroll
            for (int b = 0; b < B; ++b) {
                const int k = (t + b * NT) & (Ns - 1);
#pragma unroll
                for (int r = 1; r < R; ++r)
                    v[b + r * B] = cmul(v[b + r * B], twd<T, DIR>(tab[(r - 1) * Ns + k]));
            }
        }
    }
    // ---- butterflies
#pragma unroll
    for (int b = 0; b < B; ++b) {
        Cx<T> u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = v[b + r * B];
        Dft<T, R, DIR>::run(u);
#pragma unroll
        for (int r = 0; r < R; ++r) v[b + r * B] = u[r];
    }
    if constexpr (!last) {
        sync();                               // previous readers of buf are done
#pragma unroll
        for (int b = 0; b < B; ++b) {
            const int j = t + b * NT;
            if constexpr (Ns == 1) {          // base = 16 j: pad -> 17 j + r
                Cx<T> *p = buf + 17 * j;
#pragma unroll
                for (int r = 0; r < R; ++r) p[r] = v[b + r * B];
            } else {                          // Ns >= 16: pad(base + r Ns) = pad(base) + r (Ns + Ns/16)
                const int k = j & (Ns - 1);
                Cx<T> *p = buf + ct_pad((j - k) * R + k);
#pragma unroll
                for (int r = 0; r < R; ++r) p[r * (Ns + Ns / 16)] = v[b + r * B];
            }
        }
        sync();
        const Cx<T> *q = buf + ct_pad(t);
#pragma unroll
        for (int e = 0; e < 16; ++e) v[e] = q[e * (NT + NT / 16)];
        ct_stages<T, N, Ns * R, DIR, Sync>(v, t, buf, tab + (Ns > 1 ? (R - 1) * Ns : 0), sync);
    }
}

// in: v[e] = x[t + e*NT]; out: v[e] = X[t + e*NT] (unnormalised).  All NT
// threads of the transform call it; `buf` = the transform's own line of
// ct_stride(N) elements; `tab` = table of ct_tab_size(N) entries (shared mem).
template <class T, int N, int DIR, class Sync>
__device__ __forceinline__ void ct_fft(Cx<T> (&v)[16], const int t, Cx<T> *buf,
                                       const Cx<T> *tab, const Sync &sync) {
    static_assert(N >= 512 && (N & (N - 1)) == 0, "ct_fft: N must be a power of two >= 512");
    ct_stages<T, N, 1, DIR, Sync>(v, t, buf, tab, sync);
}

// host: fill the stage tables, entry [(r-1)*Ns + k] = exp(-2 pi i r k / (Ns R))
template <class F>   // F(index, cos, sin)
static inline void ct_fill_table(int N, F put) {
    int off = 0;
    for (int Ns = 1; Ns < N; Ns *= ct_radix(N, Ns)) {
        int R = ct_radix(N, Ns);
        if (Ns > 1) {
            for (int r = 1; r < R; ++r)
                for (int k = 0; k < Ns; ++k) {
                    long double a = -2.0L * 3.14159265358979323846264338327950288L *
                                    (long double)r * (long double)k / ((long double)Ns * R);
                    put(off + (r - 1) * Ns + k, (double)cosl(a), (double)sinl(a));
                }
            off += (R - 1) * Ns;
        }
    }
}

}  // namespace pfb
