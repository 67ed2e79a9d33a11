// Padded sizes N = m * P2 with one small odd factor m (3, 5, 7) and a power of two
// P2 in [512, 4096] -- e.g. 20480 = 2 * (5 * 2048) for the 10240^2 cube of BASELINE
// config 5, 12288, 14336, 6144 ... (ducc0 good_size lengths, pfb/workers/grid.py:276-284).
//
// The run-time radix-list transform of hess_mixed.cuh walks the whole line through
// shared memory once per stage (5 stages for 10240) with generic index arithmetic and
// two table look-ups per twiddle; it runs at ~190 GB/s.  Here the odd factor is peeled
// off (Cooley-Tukey, n = m j + q, K = k + P2 s):
//     X[k + P2 s] = sum_q w_m^(q s) * [ w_N^(q k) * FFT_P2( x[m j + q] )[k] ]
// so the bulk of the work is m register-resident compile-time transforms of length P2
// (fft_ct.cuh: 16 points per thread, radix-16 stages, two exchanges), each done by its
// own group of P2/16 threads with a named barrier, followed by ONE in-place pass of
// m-point butterflies.  The inverse runs the mirror image (m-point butterflies first,
// conjugate twiddles, then the length-P2 transforms, output index m j + q).
//
// Buffers (shared memory, complex elements):
//     R1  natural order      L * N                       time side (input of the forward
//                                                        transform, output of the inverse)
//     R2  "segment layout"   L * m * ct_stride(P2)       spectrum side: X[k + P2 s] of line l
//                                                        at (l m + s) ct_stride(P2) + k
// Segment q of R2 doubles as the exchange buffer of sub-transform q, so nothing else is
// needed.  Same three passes, same pruned transposed scratch S[band][k][row] and the same
// even / odd split of the column pass as hess_mixed.cuh.
#pragma once
#include "hess_mixed.cuh"

namespace pfb {

struct PeelDesc {
    int N, m, P2;          // N = m * P2; m = 0: not applicable
};

// host: N = m * P2 with m in {3, 5, 7} and P2 in [512, 4096]?
static inline bool peel_make_desc(int N, PeelDesc &d) {
    d.N = N; d.m = 0; d.P2 = 0;
    const int ms[3] = {3, 5, 7};
    for (int i = 0; i < 3; ++i) {
        const int m = ms[i];
        if (N % m) continue;
        const int p = N / m;
        if (p >= 512 && p <= 4096 && (p & (p - 1)) == 0) { d.m = m; d.P2 = p; return true; }
    }
    return false;
}

template <int P2> __host__ __device__ constexpr int peel_log2() { return P2 == 512 ? 9 : P2 == 1024 ? 10 : P2 == 2048 ? 11 : 12; }
// index of spectrum element K inside a line's R2 block
template <int P2> __device__ __forceinline__ int peel_xidx(int K) {
    return (K >> peel_log2<P2>()) * ct_stride(P2) + (K & (P2 - 1));
}
// group barriers executed by one sub-transform (emulator: idle groups must keep in step)
template <int P2> __host__ __device__ constexpr int peel_nbar() {
    int n = 0;
    for (int Ns = 1; Ns < P2; Ns *= ct_radix(P2, Ns))
        if (Ns * ct_radix(P2, Ns) < P2) n += 2;
    return n + 1;
}

// shared memory: [roots 40][two-level twiddles of N: kMxTwMax][ct table of P2][R1: L N][R2: L m ct_stride(P2)]
static inline size_t peel_smem_elems(int N, int m, int P2, int L) {
    return (size_t)kMxHead + ct_tab_size(P2) + (size_t)L * N + (size_t)L * m * ct_stride(P2);
}
// threads per CTA: G groups of P2/16 threads, G <= number of sub-transforms, <= 15 named
// barriers, about 640 threads (register budget 102 per thread)
static inline int peel_threads(int m, int P2, int L) {
    const int NT = P2 / 16;
    int G = 640 / NT;
    if (G > 15) G = 15;
    if (G > m * L) G = m * L;
    if (G < 1) G = 1;
    return G * NT;
}

// one m-point butterfly over the segments of a line, in place.  DIR < 0: plain DFT (inputs
// already twiddled by the sub-transforms); DIR > 0: inverse DFT, then the conjugate
// twiddle w_N^(-q k)
template <class T, int R, int DIR>
__device__ __forceinline__ void peel_bfly(Cx<T> *base, int cts, int k, const Cx<T> *tws, const Cx<T> *rt) {
    Cx<T> u[R];
#pragma unroll
    for (int r = 0; r < R; ++r) u[r] = base[r * cts];
    MxDft<T, R, DIR>::run(u, rt);
    if (DIR > 0) {
#pragma unroll
        for (int r = 1; r < R; ++r) u[r] = cmul(u[r], cconj(mx_twiddle(tws, r * k)));
    }
#pragma unroll
    for (int r = 0; r < R; ++r) base[r * cts] = u[r];
}

template <class T, int DIR>
__device__ __forceinline__ void peel_butterflies(Cx<T> *R2, int L, int m, int P2, int cts,
                                                 const Cx<T> *tws, const Cx<T> *rt) {
    for (int i = threadIdx.x; i < P2 * L; i += blockDim.x) {
        const int l = i / P2, k = i - l * P2;
        Cx<T> *base = R2 + (long)l * m * cts + k;
        if (m == 3) peel_bfly<T, 3, DIR>(base, cts, k, tws, rt);
        else if (m == 5) peel_bfly<T, 5, DIR>(base, cts, k, tws, rt);
        else peel_bfly<T, 7, DIR>(base, cts, k, tws, rt);
    }
}

// forward: R1 (natural, L lines of N) -> R2 (segment layout).  Synchronised on exit.
template <class T, int P2>
__device__ __forceinline__ void peel_fwd(const Cx<T> *R1, Cx<T> *R2, int L, int m,
                                         const Cx<T> *tws, const Cx<T> *tab, const Cx<T> *rt) {
    constexpr int NT = P2 / 16, cts = ct_stride(P2);
    const int N = m * P2, G = blockDim.x / NT;
    const int g = threadIdx.x / NT, t = threadIdx.x - g * NT;
    const GroupBar bar{1 + g, NT};
    const int nsub = m * L;
    for (int s0 = 0; s0 < nsub; s0 += G) {
        const int sub = s0 + g;
        if (sub < nsub) {
            const int l = sub / m, q = sub - l * m;
            const Cx<T> *src = R1 + (long)l * N + q;
            Cx<T> *X = R2 + (long)sub * cts;
            Cx<T> v[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) v[e] = src[(t + e * NT) * m];
            ct_fft<T, P2, -1>(v, t, X, tab, bar);
            bar();                                // the last exchange of X has been read
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                const int k = t + e * NT;
                X[k] = q ? cmul(v[e], mx_twiddle(tws, q * k)) : v[e];
            }
        } else {
#ifdef PFB_EMU
            for (int i = 0; i < peel_nbar<P2>(); ++i) __syncthreads();
#endif
        }
    }
    __syncthreads();
    peel_butterflies<T, -1>(R2, L, m, P2, cts, tws, rt);
    __syncthreads();
}

// inverse: R2 (segment layout) -> R1 (natural), unnormalised.  Synchronised on exit.
template <class T, int P2>
__device__ __forceinline__ void peel_inv(Cx<T> *R2, Cx<T> *R1, int L, int m,
                                         const Cx<T> *tws, const Cx<T> *tab, const Cx<T> *rt) {
    constexpr int NT = P2 / 16, cts = ct_stride(P2);
    const int N = m * P2, G = blockDim.x / NT;
    const int g = threadIdx.x / NT, t = threadIdx.x - g * NT;
    const GroupBar bar{1 + g, NT};
    peel_butterflies<T, +1>(R2, L, m, P2, cts, tws, rt);
    __syncthreads();
    const int nsub = m * L;
    for (int s0 = 0; s0 < nsub; s0 += G) {
        const int sub = s0 + g;
        if (sub < nsub) {
            const int l = sub / m, q = sub - l * m;
            Cx<T> *X = R2 + (long)sub * cts;
            Cx<T> *dst = R1 + (long)l * N + q;
            Cx<T> v[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) v[e] = X[t + e * NT];
            ct_fft<T, P2, +1>(v, t, X, tab, bar);
#pragma unroll
            for (int e = 0; e < 16; ++e) dst[(t + e * NT) * m] = v[e];
#ifdef PFB_EMU
            __syncthreads();                      // keeps the count of peel_nbar (no bar() here)
#endif
        } else {
#ifdef PFB_EMU
            for (int i = 0; i < peel_nbar<P2>(); ++i) __syncthreads();
#endif
        }
    }
    __syncthreads();
}

template <class T, int P2>
__device__ __forceinline__ void peel_load_tables(Cx<T> *rt, const Cx<T> *__restrict__ roots,
                                                 const Cx<T> *__restrict__ tw, const Cx<T> *__restrict__ tabg,
                                                 int N) {
    mx_load_tables(rt, roots, tw, N);
    Cx<T> *tab = rt + kMxHead;
    for (int i = threadIdx.x; i < ct_tab_size(P2); i += blockDim.x) tab[i] = tabg[i];
}

// ------------------------------------------------------------------ pass A
// L rows per CTA: real row packed as M = nyp/2 complex, transform, real-FFT post-processing,
// transposed store (same arithmetic as k_hessm_rows_fwd).
template <class T, int P2>
__global__ void __launch_bounds__(640)
k_hessp_rows_fwd(const T *__restrict__ x, const T *__restrict__ beam, Cx<T> *__restrict__ S,
                 const Cx<T> *__restrict__ tw, const Cx<T> *__restrict__ twr,
                 const Cx<T> *__restrict__ roots, const Cx<T> *__restrict__ tabg, int m, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int M = h.M, L = h.R;
    constexpr int cts = ct_stride(P2);
    Cx<T> *tab = rt + kMxHead;
    Cx<T> *R1 = tab + ct_tab_size(P2), *R2 = R1 + (long)L * M;
    const int band = blockIdx.y, row0 = blockIdx.x * L;
    peel_load_tables<T, P2>(rt, roots, tw, tabg, M);
    for (int i = threadIdx.x; i < L * M; i += blockDim.x) {
        const int l = i / M, mm = i - l * M, row = row0 + l, c0 = 2 * mm;
        T a = (T)0, b = (T)0;
        if (row < h.nx) {
            const long ro = ((long)band * h.nx + row) * h.ny;
            if (c0 < h.ny) { a = x[ro + c0]; if (beam) a *= beam[ro + c0]; }
            if (c0 + 1 < h.ny) { b = x[ro + c0 + 1]; if (beam) b *= beam[ro + c0 + 1]; }
        }
        R1[i] = cx<T>(a, b);
    }
    __syncthreads();
    peel_fwd<T, P2>(R1, R2, L, m, rt + 40, tab, rt);
    // X[k] = E + w^k O,  E = (Z[k] + conj Z[M-k])/2,  O = -i (Z[k] - conj Z[M-k])/2
    for (int i = threadIdx.x; i < (M + 1) * L; i += blockDim.x) {
        const int k = i / L, l = i - k * L;
        const Cx<T> *z = R2 + (long)l * m * cts;
        const Cx<T> zk = z[peel_xidx<P2>(k == M ? 0 : k)], zm = cconj(z[peel_xidx<P2>(k == 0 ? 0 : M - k)]);
        const Cx<T> E = cx<T>((T)0.5 * (zk.x + zm.x), (T)0.5 * (zk.y + zm.y));
        const Cx<T> D = cx<T>((T)0.5 * (zk.x - zm.x), (T)0.5 * (zk.y - zm.y));
        const Cx<T> O = cx<T>(D.y, -D.x);
        const Cx<T> X = E + cmul(twr[k], O);
        if (row0 + l < h.nx) S[((long)band * h.nyo2 + k) * h.nx + row0 + l] = X;
    }
}

// ------------------------------------------------------------------ pass B
// One column per CTA.  SPLIT (nxp = 2H, nx <= H): two length-H pairs on the even / odd
// spectrum (psfT de-interleaved [even | odd], modg[n] = exp(-2 pi i n / nxp)); the even
// result is parked in the S line by the thread that later combines it.
template <class T, int P2, bool SPLIT>
__global__ void __launch_bounds__(640)
k_hessp_cols(Cx<T> *S, const Cx<T> *__restrict__ psfT, const Cx<T> *__restrict__ tw,
             const Cx<T> *__restrict__ modg, const Cx<T> *__restrict__ roots,
             const Cx<T> *__restrict__ tabg, int m, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int N = m * P2;
    Cx<T> *tab = rt + kMxHead;
    Cx<T> *R1 = tab + ct_tab_size(P2), *R2 = R1 + N;
    const int band = blockIdx.y, col = blockIdx.x;
    Cx<T> *line = S + ((long)band * h.nyo2 + col) * h.nx;
    const Cx<T> *pl = psfT + ((long)band * h.nyo2 + col) * h.nxp;
    peel_load_tables<T, P2>(rt, roots, tw, tabg, N);
    const Cx<T> *tws = rt + 40;
    for (int n = threadIdx.x; n < N; n += blockDim.x) R1[n] = n < h.nx ? line[n] : cx<T>((T)0, (T)0);
    __syncthreads();
    peel_fwd<T, P2>(R1, R2, 1, m, tws, tab, rt);
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        Cx<T> *c = R2 + peel_xidx<P2>(n);
        *c = cmul(*c, pl[n]);
    }
    __syncthreads();
    peel_inv<T, P2>(R2, R1, 1, m, tws, tab, rt);
    if (!SPLIT) {
        for (int n = threadIdx.x; n < h.nx; n += blockDim.x) line[n] = R1[n];
    } else {
    // odd spectrum: input a[n] w^n; element n of the line and of R1 is swapped by one thread
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        Cx<T> in = cx<T>((T)0, (T)0);
        if (n < h.nx) {
            in = cmul(line[n], modg[n]);
            line[n] = R1[n];
        }
        R1[n] = in;
    }
    __syncthreads();
    peel_fwd<T, P2>(R1, R2, 1, m, tws, tab, rt);
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        Cx<T> *c = R2 + peel_xidx<P2>(n);
        *c = cmul(*c, pl[N + n]);
    }
    __syncthreads();
    peel_inv<T, P2>(R2, R1, 1, m, tws, tab, rt);
    // the same thread wrote line[n] above (same index mapping): no fence needed
    for (int n = threadIdx.x; n < h.nx; n += blockDim.x)
        line[n] = line[n] + cmul(R1[n], cconj(modg[n]));
    }
}

// ------------------------------------------------------------------ pass C
template <class T, int P2>
__global__ void __launch_bounds__(640)
k_hessp_rows_inv(const Cx<T> *__restrict__ S, T *__restrict__ out, const Cx<T> *__restrict__ tw,
                 const Cx<T> *__restrict__ twr, const Cx<T> *__restrict__ roots,
                 const Cx<T> *__restrict__ tabg, int m, HessDims h, Epilogue<T> ep) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int M = h.M, L = h.R;
    constexpr int cts = ct_stride(P2);
    Cx<T> *tab = rt + kMxHead;
    Cx<T> *R1 = tab + ct_tab_size(P2), *R2 = R1 + (long)L * M;
    const int band = blockIdx.y, row0 = blockIdx.x * L;
    peel_load_tables<T, P2>(rt, roots, tw, tabg, M);
    // Z'[m] = (X[m] + conj X[M-m]) + i w^-m (X[m] - conj X[M-m])   (= 2 Z[m]), m < M
    for (int i = threadIdx.x; i < M * L; i += blockDim.x) {
        const int mm = i / L, l = i - mm * L;
        Cx<T> xm = cx<T>((T)0, (T)0), xr = xm;
        if (row0 + l < h.nx) {
            const Cx<T> *col = S + (long)band * h.nyo2 * h.nx + row0 + l;
            xm = col[(long)mm * h.nx];
            xr = cconj(col[(long)(M - mm) * h.nx]);
        }
        const Cx<T> P = xm + xr, Q = xm - xr;
        const Cx<T> wQ = cmul(cconj(twr[mm]), Q);
        R2[(long)l * m * cts + peel_xidx<P2>(mm)] = cx<T>(P.x - wQ.y, P.y + wQ.x);
    }
    __syncthreads();
    peel_inv<T, P2>(R2, R1, L, m, rt + 40, tab, rt);
    for (int i = threadIdx.x; i < L * M; i += blockDim.x) {
        const int l = i / M, mm = i - l * M, row = row0 + l, c0 = 2 * mm;
        if (row >= h.nx) continue;
        const long ro = ((long)band * h.nx + row) * h.ny;
        const Cx<T> z = R1[i];
        const T val[2] = {z.x, z.y};
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int c = c0 + q;
            if (c < h.ny) {
                T o = val[q];
                if (ep.beam) o *= ep.beam[ro + c];
                if (ep.has_wsum) o /= ep.wsum;
                if (ep.x) o += ep.x[ro + c] * ep.sigmainv;
                out[ro + c] = o;
            }
        }
    }
}

}  // namespace pfb
