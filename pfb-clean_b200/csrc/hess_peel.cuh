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
#include <type_traits>

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
// inpl (row passes): no R1, the natural-order line lives in its own R2 block
static inline size_t peel_smem_elems(int N, int m, int P2, int L, bool inpl = false) {
    return (size_t)kMxHead + ct_tab_size(P2) + 4 + (inpl ? 0 : (size_t)L * N) + (size_t)L * m * ct_stride(P2);
}
// second form of the column pass: segments (+ the parked even half when split)
static inline size_t peel_smem_elems_cols2(int N, int m, int P2, bool split) {
    return (size_t)kMxHead + ct_tab_size(P2) + 4 + (size_t)m * ct_stride(P2) + (split ? (size_t)N : 0);
}
// in-place row passes need exactly one thread group per segment
static inline bool peel_can_inplace(int m, int P2) { return m * (P2 / 16) <= 640; }
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

template <class T, int P2>
__device__ __forceinline__ void peel_load_tables(Cx<T> *rt, const Cx<T> *__restrict__ roots,
                                                 const Cx<T> *__restrict__ tw, const Cx<T> *__restrict__ tabg,
                                                 int N) {
    mx_load_tables(rt, roots, tw, N);
    Cx<T> *tab = rt + kMxHead;
    for (int i = threadIdx.x; i < ct_tab_size(P2); i += blockDim.x) tab[i] = tabg[i];
}

// Column pass, spectrum side: forward butterflies, multiplication with the PSF spectrum and
// inverse butterflies (+ conjugate twiddles) in one walk over the segments -- the spectrum
// never goes back to shared memory in between.  pl: this column's (half of the) PSF spectrum.
template <class T, int R>
__device__ __forceinline__ void peel_mid_r(Cx<T> *R2, int P2, int cts, const Cx<T> *__restrict__ pl,
                                           const Cx<T> *tws, const Cx<T> *rt) {
    constexpr bool AHEAD = sizeof(T) == 4;           // fp64: no registers left for the look-ahead
    Cx<T> pn[R];
    int k = threadIdx.x;
    if (AHEAD && k < P2) {
#pragma unroll
        for (int r = 0; r < R; ++r) pn[r] = pl[k + r * P2];
    }
    for (; k < P2; k += blockDim.x) {
        Cx<T> pc[R];
        if (AHEAD) {
#pragma unroll
            for (int r = 0; r < R; ++r) pc[r] = pn[r];
            const int kn = k + blockDim.x;
            if (kn < P2) {                           // next butterfly's PSF values, one trip ahead
#pragma unroll
                for (int r = 0; r < R; ++r) pn[r] = pl[kn + r * P2];
            }
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) pc[r] = pl[k + r * P2];
        }
        Cx<T> *base = R2 + k;
        Cx<T> u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = base[r * cts];
        MxDft<T, R, -1>::run(u, rt);
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = cmul(u[r], pc[r]);
        MxDft<T, R, +1>::run(u, rt);
#pragma unroll
        for (int r = 1; r < R; ++r) u[r] = cmul(u[r], cconj(mx_twiddle(tws, r * k)));
#pragma unroll
        for (int r = 0; r < R; ++r) base[r * cts] = u[r];
    }
}

// forward sub-transforms only (R1 -> twiddled segments of R2), synchronised on exit
template <class T, int P2>
__device__ __forceinline__ void peel_fwd_subs(const Cx<T> *R1, Cx<T> *R2, int L, int m,
                                              const Cx<T> *tws, const Cx<T> *tab,
                                              int ld1 = 0, bool inpl = false) {
    // inpl: R1 line l IS the R2 block of line l (ld1 = m ct_stride(P2)); requires one group per
    // segment (blockDim = m P2/16), so that a round is exactly one line and the block-wide
    // barrier below separates its last natural-order read from the first exchange write
    constexpr int NT = P2 / 16, cts = ct_stride(P2);
    const int N = ld1 ? ld1 : m * P2, G = blockDim.x / NT;
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
            if (inpl) __syncthreads();
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
}

// inverse sub-transforms only (segments of R2, already butterflied and twiddled -> R1)
template <class T, int P2>
__device__ __forceinline__ void peel_inv_subs(Cx<T> *R2, Cx<T> *R1, int L, int m, const Cx<T> *tab,
                                              int ld1 = 0, bool inpl = false) {
    constexpr int NT = P2 / 16, cts = ct_stride(P2);
    const int N = ld1 ? ld1 : m * P2, G = blockDim.x / NT;
    const int g = threadIdx.x / NT, t = threadIdx.x - g * NT;
    const GroupBar bar{1 + g, NT};
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
            if (inpl) __syncthreads();            // every segment of the line is done with its exchanges
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

// forward: R1 (natural, L lines of N) -> R2 (segment layout).  Synchronised on exit.
template <class T, int P2>
__device__ __forceinline__ void peel_fwd(const Cx<T> *R1, Cx<T> *R2, int L, int m,
                                         const Cx<T> *tws, const Cx<T> *tab, const Cx<T> *rt,
                                         int ld1 = 0, bool inpl = false) {
    peel_fwd_subs<T, P2>(R1, R2, L, m, tws, tab, ld1, inpl);
    peel_butterflies<T, -1>(R2, L, m, P2, ct_stride(P2), tws, rt);
    __syncthreads();
}

// inverse: R2 (segment layout) -> R1 (natural), unnormalised.  Synchronised on exit.
template <class T, int P2>
__device__ __forceinline__ void peel_inv(Cx<T> *R2, Cx<T> *R1, int L, int m,
                                         const Cx<T> *tws, const Cx<T> *tab, const Cx<T> *rt,
                                         int ld1 = 0, bool inpl = false) {
    peel_butterflies<T, +1>(R2, L, m, P2, ct_stride(P2), tws, rt);
    __syncthreads();
    peel_inv_subs<T, P2>(R2, R1, L, m, tab, ld1, inpl);
}

// forward pair - PSF - inverse pair of one column (half): R1 -> R1
template <class T, int P2>
__device__ __forceinline__ void peel_conv(Cx<T> *R1, Cx<T> *R2, int m, const Cx<T> *__restrict__ pl,
                                          const Cx<T> *tws, const Cx<T> *tab, const Cx<T> *rt) {
    constexpr int cts = ct_stride(P2);
    peel_fwd_subs<T, P2>(R1, R2, 1, m, tws, tab);
    if (m == 3) peel_mid_r<T, 3>(R2, P2, cts, pl, tws, rt);
    else if (m == 5) peel_mid_r<T, 5>(R2, P2, cts, pl, tws, rt);
    else peel_mid_r<T, 7>(R2, P2, cts, pl, tws, rt);
    __syncthreads();
    peel_inv_subs<T, P2>(R2, R1, 1, m, tab);
}

// ------------------------------------------------------------------ pass A
// L rows per CTA: real row packed as M = nyp/2 complex, transform, real-FFT post-processing,
// transposed store (same arithmetic as k_hessm_rows_fwd).
template <class T, int P2>
__global__ void __launch_bounds__(640)
k_hessp_rows_fwd(const T *__restrict__ x, const T *__restrict__ beam, Cx<T> *__restrict__ S,
                 const Cx<T> *__restrict__ tw, const Cx<T> *__restrict__ twr,
                 const Cx<T> *__restrict__ roots, const Cx<T> *__restrict__ tabg, int m, int inpl,
                 HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int M = h.M, L = h.R;                  // L = 1, 2 or 4
    const int ls = L == 4 ? 2 : L == 2 ? 1 : 0;
    constexpr int cts = ct_stride(P2);
    Cx<T> *tab = rt + kMxHead;
    Cx<T> *R1 = tab + ct_tab_size(P2) + 4, *R2 = inpl ? R1 : R1 + (long)L * M;
    const int ld1 = inpl ? m * cts : M;
    const int band = blockIdx.y, row0 = blockIdx.x * L;
    peel_load_tables<T, P2>(rt, roots, tw, tabg, M);
    for (int l = 0; l < L; ++l) {
        const int row = row0 + l;
        const long ro = ((long)band * h.nx + row) * h.ny;
        T a[8], b[8];
        peel_batched<8>(M,
            [&](int u, int mm) {
                const int c0 = 2 * mm;
                a[u] = b[u] = (T)0;
                if (row < h.nx) {
                    if (c0 < h.ny) { a[u] = x[ro + c0]; if (beam) a[u] *= beam[ro + c0]; }
                    if (c0 + 1 < h.ny) { b[u] = x[ro + c0 + 1]; if (beam) b[u] *= beam[ro + c0 + 1]; }
                }
            },
            [&](int u, int mm) { R1[(long)l * ld1 + mm] = cx<T>(a[u], b[u]); });
    }
    __syncthreads();
    peel_fwd<T, P2>(R1, R2, L, m, rt + 40, tab, rt, ld1, inpl != 0);
    // X[k] = E + w^k O,  E = (Z[k] + conj Z[M-k])/2,  O = -i (Z[k] - conj Z[M-k])/2
    Cx<T> w[4];
    peel_batched<4>((M + 1) * L,
        [&](int u, int i) { w[u] = twr[i >> ls]; },
        [&](int u, int i) {
            const int k = i >> ls, l = i & (L - 1);
            const Cx<T> *z = R2 + (long)l * m * cts;
            const Cx<T> zk = z[peel_xidx<P2>(k == M ? 0 : k)], zm = cconj(z[peel_xidx<P2>(k == 0 ? 0 : M - k)]);
            const Cx<T> E = cx<T>((T)0.5 * (zk.x + zm.x), (T)0.5 * (zk.y + zm.y));
            const Cx<T> D = cx<T>((T)0.5 * (zk.x - zm.x), (T)0.5 * (zk.y - zm.y));
            const Cx<T> O = cx<T>(D.y, -D.x);
            const Cx<T> X = E + cmul(w[u], O);
            if (row0 + l < h.nx) S[((long)band * h.nyo2 + k) * h.nx + row0 + l] = X;
        });
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
    Cx<T> *R1 = tab + ct_tab_size(P2) + 4, *R2 = R1 + N;
    const int band = blockIdx.y, col = blockIdx.x;
    Cx<T> *line = S + ((long)band * h.nyo2 + col) * h.nx;
    const Cx<T> *pl = psfT + ((long)band * h.nyo2 + col) * h.nxp;
    peel_load_tables<T, P2>(rt, roots, tw, tabg, N);
    const Cx<T> *tws = rt + 40;
    constexpr int U1 = sizeof(T) == 4 ? 8 : 4, U2 = U1 / 2;   // batch depths by register budget
    Cx<T> v[U1], g[U2];
    peel_batched<U1>(N,
        [&](int u, int n) { v[u] = n < h.nx ? line[n] : cx<T>((T)0, (T)0); },
        [&](int u, int n) { R1[n] = v[u]; });
    __syncthreads();
    peel_conv<T, P2>(R1, R2, m, pl, tws, tab, rt);
    if (!SPLIT) {
        for (int n = threadIdx.x; n < h.nx; n += blockDim.x) line[n] = R1[n];
    } else {
        // odd spectrum: input a[n] w^n; element n of the line and of R1 is swapped by one thread
        peel_batched<U2>(N,
            [&](int u, int n) {
                v[u] = cx<T>((T)0, (T)0);
                if (n < h.nx) { v[u] = line[n]; g[u] = modg[n]; }
            },
            [&](int u, int n) {
                Cx<T> in = cx<T>((T)0, (T)0);
                if (n < h.nx) {
                    in = cmul(v[u], g[u]);
                    line[n] = R1[n];
                }
                R1[n] = in;
            });
        __syncthreads();
        peel_conv<T, P2>(R1, R2, m, pl + N, tws, tab, rt);
        // the same thread wrote line[n] above (same index mapping): no fence needed
        peel_batched<U2>(h.nx,
            [&](int u, int n) { v[u] = line[n]; g[u] = modg[n]; },
            [&](int u, int n) { line[n] = v[u] + cmul(R1[n], cconj(g[u])); });
    }
}

// ------------------------------------------------------------------ pass B, second form
// Decimation in frequency on the way in, in time on the way out (n = j + P2 s, K = m k + q):
//     X[m k + q] = FFT_P2( b_q )[k],   b_q[j] = w_N^(j q) sum_s w_m^(s q) a[j + P2 s]
//     y[j + P2 s] = sum_q w_m^(-s q) w_N^(-j q) IFFT_P2( Y[m . + q] )[j]
// The m-point butterflies now sit at the GLOBAL-memory ends of the pipeline: a butterfly thread
// loads a[j + P2 s] straight from the S line (coalesced over j) and stores y[j + P2 s] straight
// back, so no natural-order staging line exists; a sub-transform's forward output is, element
// for element, the inverse sub-transform's input, so the PSF product (psfT stored permuted,
// [q][k], by k_psfhat_prepare) and the turn-around happen in registers like in the
// power-of-two kernel.  12 shared-memory accesses per point and half instead of 16, three
// block-wide barriers per half instead of seven.  Needs one thread group per segment
// (peel_can_inplace).  SPLIT: the even half's result waits in a shared-memory line (PARK).
template <class T, int P2, bool SPLIT>
__global__ void __launch_bounds__(640)
k_hessp_cols2(Cx<T> *S, const Cx<T> *__restrict__ psfP, const Cx<T> *__restrict__ tw,
              const Cx<T> *__restrict__ modg, const Cx<T> *__restrict__ roots,
              const Cx<T> *__restrict__ tabg, int m, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    constexpr int NT = P2 / 16, cts = ct_stride(P2);
    const int N = m * P2;
    Cx<T> *tab = rt + kMxHead;
    Cx<T> *SEG = tab + ct_tab_size(P2) + 4, *PARK = SEG + (long)m * cts;
    const int band = blockIdx.y, col = blockIdx.x;
    Cx<T> *line = S + ((long)band * h.nyo2 + col) * h.nx;
    const Cx<T> *pl = psfP + ((long)band * h.nyo2 + col) * h.nxp;
    peel_load_tables<T, P2>(rt, roots, tw, tabg, N);
    const Cx<T> *tws = rt + 40;
    const int g = threadIdx.x / NT, t = threadIdx.x - g * NT;      // group g = segment q
    const GroupBar bar{1 + g, NT};
    __syncthreads();
#pragma unroll 1
    for (int half = 0; half < (SPLIT ? 2 : 1); ++half) {
        const Cx<T> *ph = pl + (long)half * N;
        {   // this half's PSF values: DRAM -> L2 now, read after the forward sub-transform
            constexpr int PL = 128 / (int)sizeof(Cx<T>);
            for (int i = threadIdx.x * PL; i < N; i += (int)blockDim.x * PL) prefetch_l2(ph + i);
        }
        // ---- in: m-point butterflies on a[j + P2 s] (x w_2N^n for the odd half), twiddle, -> SEG
        auto fwd_bfly = [&](auto tag) {
            constexpr int R = decltype(tag)::value;
            for (int j = threadIdx.x; j < P2; j += blockDim.x) {
                Cx<T> u[R];
#pragma unroll
                for (int s = 0; s < R; ++s) {
                    const int n = j + s * P2;
                    u[s] = n < h.nx ? line[n] : cx<T>((T)0, (T)0);
                }
                if (SPLIT && half == 1) {
#pragma unroll
                    for (int s = 0; s < R; ++s) {
                        const int n = j + s * P2;
                        if (n < h.nx) u[s] = cmul(u[s], modg[n]);
                    }
                }
                MxDft<T, R, -1>::run(u, rt);
#pragma unroll
                for (int q = 1; q < R; ++q) u[q] = cmul(u[q], mx_twiddle(tws, q * j));
#pragma unroll
                for (int q = 0; q < R; ++q) SEG[q * cts + j] = u[q];
            }
        };
        if (m == 3) fwd_bfly(std::integral_constant<int, 3>());
        else if (m == 5) fwd_bfly(std::integral_constant<int, 5>());
        else fwd_bfly(std::integral_constant<int, 7>());
        __syncthreads();
        // ---- sub-transform q forward, x PSF, inverse: all in registers
        {
            Cx<T> *X = SEG + (long)g * cts;
            Cx<T> v[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) v[e] = X[t + e * NT];
            ct_fft<T, P2, -1>(v, t, X, tab, bar);
            const Cx<T> *pq = ph + (long)g * P2;
#pragma unroll
            for (int e = 0; e < 16; ++e) v[e] = cmul(v[e], pq[t + e * NT]);
            ct_fft<T, P2, +1>(v, t, X, tab, bar);
            bar();                                // the last exchange of X has been read
#pragma unroll
            for (int e = 0; e < 16; ++e) X[t + e * NT] = v[e];
        }
        __syncthreads();
        // ---- out: conjugate twiddle, inverse m-point butterflies, y[j + P2 s] -> PARK / line
        auto inv_bfly = [&](auto tag) {
            constexpr int R = decltype(tag)::value;
            for (int j = threadIdx.x; j < P2; j += blockDim.x) {
                Cx<T> u[R];
#pragma unroll
                for (int q = 0; q < R; ++q) u[q] = SEG[q * cts + j];
#pragma unroll
                for (int q = 1; q < R; ++q) u[q] = cmul(u[q], cconj(mx_twiddle(tws, q * j)));
                MxDft<T, R, +1>::run(u, rt);
                if (SPLIT && half == 0) {
#pragma unroll
                    for (int s = 0; s < R; ++s) PARK[j + s * P2] = u[s];
                } else if (SPLIT) {
#pragma unroll
                    for (int s = 0; s < R; ++s) {
                        const int n = j + s * P2;
                        if (n < h.nx) line[n] = PARK[n] + cmul(u[s], cconj(modg[n]));
                    }
                } else {
#pragma unroll
                    for (int s = 0; s < R; ++s) {
                        const int n = j + s * P2;
                        if (n < h.nx) line[n] = u[s];
                    }
                }
            }
        };
        if (m == 3) inv_bfly(std::integral_constant<int, 3>());
        else if (m == 5) inv_bfly(std::integral_constant<int, 5>());
        else inv_bfly(std::integral_constant<int, 7>());
        // no barrier: the next half's butterflies touch SEG[.][j], PARK[j + P2 s] and line[j + P2 s]
        // of the SAME j in the same thread
    }
}

// ------------------------------------------------------------------ pass C
// Mirror of pass A with the epilogue fused.  Every element of the S tile is fetched ONCE into
// the line's R2 block; X[m] and X[M-m] are then combined pairwise in place.
#ifdef PFB_EMU
struct PeelPrimal {};                            // the fused primal epilogue is not built for the emulator
#else
typedef PrimalArgs PeelPrimal;
#endif
// PRIMAL (fp32): the primal step of primal_dual_optimised is the epilogue, as in
// k_hessf_rows_inv_tma<.., PRIMAL> (same arithmetic and order); pp.part holds 2 partial eps
// sums per CTA (index blockIdx.y * gridDim.x + blockIdx.x).
template <class T, int P2, bool PRIMAL>
__global__ void __launch_bounds__(640)
k_hessp_rows_inv(const Cx<T> *__restrict__ S, T *__restrict__ out, const Cx<T> *__restrict__ tw,
                 const Cx<T> *__restrict__ twr, const Cx<T> *__restrict__ roots,
                 const Cx<T> *__restrict__ tabg, int m, int inpl, HessDims h, Epilogue<T> ep,
                 PeelPrimal pp) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int M = h.M, L = h.R;                  // L = 1, 2 or 4
    const int ls = L == 4 ? 2 : L == 2 ? 1 : 0;
    constexpr int cts = ct_stride(P2);
    Cx<T> *tab = rt + kMxHead;
    Cx<T> *xtop = tab + ct_tab_size(P2);         // X[M] of each line
    Cx<T> *R1 = xtop + 4, *R2 = inpl ? R1 : R1 + (long)L * M;
    const int ld1 = inpl ? m * cts : M;
    const int band = blockIdx.y, row0 = blockIdx.x * L;
    peel_load_tables<T, P2>(rt, roots, tw, tabg, M);
    constexpr int U3 = sizeof(T) == 4 ? 8 : 4;
    {
        Cx<T> xv[U3];
        const Cx<T> *tile = S + (long)band * h.nyo2 * h.nx + row0;
        peel_batched<U3>((M + 1) * L,
            [&](int u, int i) {
                const int k = i >> ls, l = i & (L - 1);
                xv[u] = row0 + l < h.nx ? tile[(long)k * h.nx + l] : cx<T>((T)0, (T)0);
            },
            [&](int u, int i) {
                const int k = i >> ls, l = i & (L - 1);
                if (k == M) xtop[l] = xv[u];
                else R2[(long)l * m * cts + peel_xidx<P2>(k)] = xv[u];
            });
    }
    __syncthreads();
    // Z'[j] = (X[j] + conj X[M-j]) + i w^-j (X[j] - conj X[M-j])   (= 2 Z[j]), j < M; the pair
    // j, M-j shares its inputs and w^(M-j) = -conj w^j
    {
        Cx<T> w[4];
        const int np = M / 2 + 1;                // pairs (j, M - j), j = 0 .. M/2
        peel_batched<4>(np * L,
            [&](int u, int i) { w[u] = twr[i >> ls]; },
            [&](int u, int i) {
                const int j = i >> ls, l = i & (L - 1);
                Cx<T> *z = R2 + (long)l * m * cts;
                const int ia = peel_xidx<P2>(j), ib = peel_xidx<P2>(j == 0 ? 0 : M - j);
                const Cx<T> a = z[ia], b = j == 0 ? xtop[l] : z[ib];
                {
                    const Cx<T> c = cconj(b);
                    const Cx<T> P = a + c, Q = a - c;
                    const Cx<T> wQ = cmul(cconj(w[u]), Q);
                    z[ia] = cx<T>(P.x - wQ.y, P.y + wQ.x);
                }
                if (j != 0 && 2 * j != M) {
                    const Cx<T> c = cconj(a);
                    const Cx<T> P = b + c, Q = b - c;
                    const Cx<T> wQ = cmul(cx<T>(-w[u].x, -w[u].y), Q);   // conj(w^(M-j)) = -w^j
                    z[ib] = cx<T>(P.x - wQ.y, P.y + wQ.x);
                }
            });
    }
    __syncthreads();
    peel_inv<T, P2>(R2, R1, L, m, rt + 40, tab, rt, ld1, inpl != 0);
#ifndef PFB_EMU
    constexpr bool DO_PRIMAL = PRIMAL && sizeof(T) == 4;
#else
    constexpr bool DO_PRIMAL = false;
#endif
    if constexpr (DO_PRIMAL) {
#ifndef PFB_EMU
        double acc0 = 0.0, acc1 = 0.0;
        for (int l = 0; l < L; ++l) {
            const int row = row0 + l;
            if (row >= h.nx) break;
            const long ro = ((long)band * h.nx + row) * h.ny;
            const int npair = (h.ny + 1) / 2 < M ? (h.ny + 1) / 2 : M;
            float pv[4][2], wv[4][2], qv[4][2], dv[4][2];
            peel_batched<4>(npair,
                [&](int u, int mm) {
#pragma unroll
                    for (int z = 0; z < 2; ++z) {
                        const int c = 2 * mm + z;
                        const bool ok = c < h.ny;
                        pv[u][z] = ok ? pp.xp[ro + c] : 0.0f;
                        wv[u][z] = ok ? pp.w[ro + c] : 0.0f;
                        qv[u][z] = ok ? pp.wprev[ro + c] : 0.0f;
                        dv[u][z] = ok ? pp.data[ro + c] : 0.0f;
                    }
                },
                [&](int u, int mm) {
                    const Cx<T> zc = R1[(long)l * ld1 + mm];
                    const float cv[2] = {(float)zc.x, (float)zc.y};
#pragma unroll
                    for (int z = 0; z < 2; ++z) {
                        const int c = 2 * mm + z;
                        if (c < h.ny) {
                            float ge = 2.0f * wv[u][z] - qv[u][z];
                            const float gr = cv[z] - dv[u][z];
                            ge = ge + gr;
                            float xi = pv[u][z] - pp.tau * ge;
                            if (pp.positivity == 1 && xi < 0.0f) xi = 0.0f;
                            pp.x[ro + c] = xi;
                            const float df = xi - pv[u][z];
                            acc0 += (double)(df * df);
                            acc1 += (double)(xi * xi);
                        }
                    }
                });
        }
        __shared__ double s_red[2][32];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            acc0 += __shfl_down_sync(0xffffffffu, acc0, o);
            acc1 += __shfl_down_sync(0xffffffffu, acc1, o);
        }
        const int wid = threadIdx.x >> 5, ln = threadIdx.x & 31;
        if (ln == 0) { s_red[0][wid] = acc0; s_red[1][wid] = acc1; }
        __syncthreads();
        if (wid == 0) {
            const int nw = blockDim.x >> 5;
            double a0 = ln < nw ? s_red[0][ln] : 0.0, a1 = ln < nw ? s_red[1][ln] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a0 += __shfl_down_sync(0xffffffffu, a0, o);
                a1 += __shfl_down_sync(0xffffffffu, a1, o);
            }
            if (ln == 0) {
                const long cta = (long)blockIdx.y * gridDim.x + blockIdx.x;
                pp.part[2 * cta] = a0;
                pp.part[2 * cta + 1] = a1;
            }
        }
#endif
    } else {
    for (int l = 0; l < L; ++l) {
        const int row = row0 + l;
        if (row >= h.nx) break;
        const long ro = ((long)band * h.nx + row) * h.ny;
#pragma unroll 4
        for (int mm = threadIdx.x; mm < M; mm += blockDim.x) {
            const Cx<T> z = R1[(long)l * ld1 + mm];
            const T val[2] = {z.x, z.y};
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int c = 2 * mm + q;
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
    }
}

}  // namespace pfb
