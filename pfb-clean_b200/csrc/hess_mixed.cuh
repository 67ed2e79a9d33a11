// Mixed-radix path of the PSF Hessian: padded sizes that are not powers of two.
//
// The reference pads to ducc0 `good_size` lengths made even
// (pfb/workers/grid.py:276-284): 2^a 3^b 5^c 7^d 11^e, e.g. 20480 = 2^12 * 5 for
// the 10240^2 cube of BASELINE config 5, or 12000 for a 6000^2 image.  Same
// three passes and the same pruned, transposed scratch S[band][k][row] as
// hess.cu; the transform itself is a shared-memory Stockham FFT with a run-time
// radix list (8, 4, 2, 3, 5, 7, 11, 13), ping-ponging between two line buffers:
//     butterfly j in [0, N/R):  k = j mod Ns
//     in : u[r] = src[j + r*N/R] * w^(r*k),   w = exp(-/+ 2 pi i / (Ns*R))
//     out: dst[(j-k)*R + k + r*Ns] = DFT_R(u)[r]
// Twiddles are the product of two small shared-memory tables (fp64 on the host,
// rounded once); the odd radices use a direct O(R^2) DFT on a table of R-th
// roots held in shared memory.
//
//  pass A  k_hessm_rows_fwd : L rows per CTA, real row packed as M = nyp/2
//          complex, FFT, real-FFT post-processing, transposed store.
//  pass B  k_hessm_cols     : one column per CTA.  SPLIT (nxp = 2H, nx <= H): the
//          length-2H pair is replaced by two length-H pairs on the even / odd
//          spectrum exactly as in hess_fast.cuh, run one after the other on the
//          same two buffers (the even result is parked in the S line, which is
//          dead once the odd input has been formed).  FULL otherwise.
//  pass C  k_hessm_rows_inv : mirror of pass A with the epilogue fused.
//
// Limits: two lines of the transform length must fit in shared memory
// (fp32: nyp, nxp(split) <= 28672; fp64: 14336).
#pragma once
#include "hess_fast.cuh"

namespace pfb {

struct MxDesc {
    int N, nstage;
    int radix[24];
};

// offsets of the R-th root tables inside the `roots` array
__host__ __device__ constexpr int mx_root_off(int R) {
    return R == 3 ? 0 : R == 5 ? 3 : R == 7 ? 8 : R == 11 ? 15 : 26;   // 13 -> 26
}
static const int kMxRoots = 39;

// host: radix list for N (true if N is 13-smooth)
static inline bool mx_make_desc(int N, MxDesc &d) {
    d.N = N;
    d.nstage = 0;
    int rem = N;
    if (N < 1) return false;
    while (rem % 8 == 0 && d.nstage < 24) { d.radix[d.nstage++] = 8; rem /= 8; }
    if (rem % 4 == 0) { d.radix[d.nstage++] = 4; rem /= 4; }
    if (rem % 2 == 0) { d.radix[d.nstage++] = 2; rem /= 2; }
    const int odd[5] = {3, 5, 7, 11, 13};
    for (int q = 0; q < 5; ++q)
        while (rem % odd[q] == 0 && d.nstage < 24) { d.radix[d.nstage++] = odd[q]; rem /= odd[q]; }
    return rem == 1;
}

template <class T, int R, int DIR> struct MxDft {
    // direct DFT of odd prime length R; rt[m] = exp(-2 pi i m / R)
    static __device__ __forceinline__ void run(Cx<T> (&u)[R], const Cx<T> *rt) {
        Cx<T> w[R];
#pragma unroll
        for (int m = 0; m < R; ++m) w[m] = twd<T, DIR>(rt[mx_root_off(R) + m]);
        Cx<T> o[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
            Cx<T> acc = u[0];
#pragma unroll
            for (int r = 1; r < R; ++r) acc = acc + cmul(u[r], w[(q * r) % R]);
            o[q] = acc;
        }
#pragma unroll
        for (int q = 0; q < R; ++q) u[q] = o[q];
    }
};
// 3- and 5-point butterflies in their symmetric form (pairs x[n], x[R-n] share their cosine /
// sine products): 16 and 56 real operations instead of 36 and 136 for the direct sums above --
// these two radices carry most of the arithmetic of the good_size paddings (2^a 3^b 5^c)
template <class T, int DIR> struct MxDft<T, 3, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[3], const Cx<T> *) {
        const T s = (T)0.86602540378443864676;                 // sin(2 pi / 3)
        const Cx<T> t1 = u[1] + u[2], d = u[1] - u[2];
        const Cx<T> m1 = cx<T>(u[0].x - (T)0.5 * t1.x, u[0].y - (T)0.5 * t1.y);
        // forward: -i s d = (s d.y, -s d.x); inverse: +i s d
        const Cx<T> m2 = DIR < 0 ? cx<T>(s * d.y, -s * d.x) : cx<T>(-s * d.y, s * d.x);
        u[0] = u[0] + t1;
        u[1] = m1 + m2;
        u[2] = m1 - m2;
    }
};
template <class T, int DIR> struct MxDft<T, 5, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[5], const Cx<T> *) {
        const T c1 = (T)0.30901699437494742410, c2 = (T)-0.80901699437494742410;   // cos(2 pi/5), cos(4 pi/5)
        const T s1 = (T)0.95105651629515357212, s2 = (T)0.58778525229247312917;    // sin(2 pi/5), sin(4 pi/5)
        const Cx<T> t1 = u[1] + u[4], t2 = u[2] + u[3], t3 = u[1] - u[4], t4 = u[2] - u[3];
        const Cx<T> a1 = cx<T>(u[0].x + c1 * t1.x + c2 * t2.x, u[0].y + c1 * t1.y + c2 * t2.y);
        const Cx<T> a2 = cx<T>(u[0].x + c2 * t1.x + c1 * t2.x, u[0].y + c2 * t1.y + c1 * t2.y);
        const Cx<T> b1 = cx<T>(s1 * t3.x + s2 * t4.x, s1 * t3.y + s2 * t4.y);
        const Cx<T> b2 = cx<T>(s2 * t3.x - s1 * t4.x, s2 * t3.y - s1 * t4.y);
        // forward: X1 = a1 - i b1, X4 = a1 + i b1, X2 = a2 - i b2, X3 = a2 + i b2; inverse: conjugate roles
        const Cx<T> ib1 = DIR < 0 ? cx<T>(b1.y, -b1.x) : cx<T>(-b1.y, b1.x);      // -/+ i b1
        const Cx<T> ib2 = DIR < 0 ? cx<T>(b2.y, -b2.x) : cx<T>(-b2.y, b2.x);
        u[0] = u[0] + t1 + t2;
        u[1] = a1 + ib1;
        u[4] = a1 - ib1;
        u[2] = a2 + ib2;
        u[3] = a2 - ib2;
    }
};
template <class T, int DIR> struct MxDft<T, 2, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[2], const Cx<T> *) { Dft<T, 2, DIR>::run(u); }
};
template <class T, int DIR> struct MxDft<T, 4, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[4], const Cx<T> *) { Dft<T, 4, DIR>::run(u); }
};
template <class T, int DIR> struct MxDft<T, 8, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[8], const Cx<T> *) { Dft<T, 8, DIR>::run(u); }
};

// Twiddles: w_N^m = T1[m >> 6] * T0[m & 63] from two small tables held in shared
// memory (T0[i] = w_N^i, i < 64; T1[i] = w_N^(64 i)): one extra complex multiply
// instead of a gather from a length-N table in global memory, whose L2 latency
// was exposed once per butterfly (first measurement of this path: 111 us per
// 10240-point column pair, 7x its shared-memory traffic bound).
static const int kMxTw0 = 64, kMxTwMax = 64 + 256;       // N <= 16384
static inline int mx_tw_entries(int N) { return kMxTw0 + (N + kMxTw0 - 1) / kMxTw0; }
// host: angle index (in units of 2 pi / N) of table entry i
static inline long mx_tw_index(int i) { return i < kMxTw0 ? i : (long)(i - kMxTw0) * kMxTw0; }

template <class T>
__device__ __forceinline__ Cx<T> mx_twiddle(const Cx<T> *tws, int m) {
    return cmul(tws[kMxTw0 + (m >> 6)], tws[m & 63]);
}

// one radix-R stage over L lines of length N (line l at src + l*N)
template <class T, int R, int DIR>
__device__ __forceinline__ void mx_stage(const Cx<T> *src, Cx<T> *dst, int L, int N, int Ns,
                                         const Cx<T> *tws, const Cx<T> *rt) {
    const int nb = N / R;
    const int twm = N / (Ns * R);
    const bool pow2 = (Ns & (Ns - 1)) == 0;       // every stage up to the first odd radix
    for (int i = threadIdx.x; i < nb * L; i += blockDim.x) {
        const int l = L == 1 ? 0 : i / nb, j = i - l * nb;
        const int k = pow2 ? (j & (Ns - 1)) : j % Ns;
        const Cx<T> *s = src + (long)l * N + j;
        Cx<T> u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = s[r * nb];
        if (Ns > 1) {
#pragma unroll
            for (int r = 1; r < R; ++r) u[r] = cmul(u[r], twd<T, DIR>(mx_twiddle(tws, r * k * twm)));
        }
        MxDft<T, R, DIR>::run(u, rt);
        Cx<T> *o = dst + (long)l * N + (j - k) * R + k;
#pragma unroll
        for (int r = 0; r < R; ++r) o[r * Ns] = u[r];
    }
}

// L transforms of length d.N: data in `a` (visible to all threads on entry: the
// caller has synchronised), scratch `b`.  Returns the buffer holding the result
// (synchronised).  Every thread of the block must call it.
template <class T, int DIR>
__device__ __forceinline__ Cx<T> *mx_fft(Cx<T> *a, Cx<T> *b, int L, const MxDesc &d,
                                         const Cx<T> *tw, const Cx<T> *rt) {
    int Ns = 1;
    for (int s = 0; s < d.nstage; ++s) {
        const int R = d.radix[s];
        switch (R) {
        case 8: mx_stage<T, 8, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        case 4: mx_stage<T, 4, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        case 2: mx_stage<T, 2, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        case 3: mx_stage<T, 3, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        case 5: mx_stage<T, 5, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        case 7: mx_stage<T, 7, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        case 11: mx_stage<T, 11, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        default: mx_stage<T, 13, DIR>(a, b, L, d.N, Ns, tw, rt); break;
        }
        __syncthreads();
        Cx<T> *t = a; a = b; b = t;
        Ns *= R;
    }
    return a;
}

// shared memory: [roots (kMxRoots, padded to 40)] [twiddle tables (kMxTwMax)] [A: L*N] [B: L*N]
static const int kMxHead = 40 + kMxTwMax;
static inline size_t mx_smem_bytes(int N, int L, size_t cs) { return cs * (kMxHead + (size_t)2 * L * N); }

// roots -> rt[0..40), two-level twiddle tables of length N -> rt[40..)
template <class T>
__device__ __forceinline__ void mx_load_tables(Cx<T> *rt, const Cx<T> *__restrict__ roots,
                                               const Cx<T> *__restrict__ tw, int N) {
    for (int i = threadIdx.x; i < kMxRoots; i += blockDim.x) rt[i] = roots[i];
    const int n = kMxTw0 + (N + kMxTw0 - 1) / kMxTw0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) rt[40 + i] = tw[i];
}

// Element loops of the passes: thread t takes elements t, t + nthreads, ...; U of them are
// fetched (ld) before the first is used (st), so U global / L2 latencies overlap instead of
// queueing one behind the other (ncu r02q / r02zf: 34-48 % of the stall samples of the
// one-at-a-time loops were long-scoreboard waits on single loads).
template <int U, class LD, class ST>
__device__ __forceinline__ void peel_batched(int n, LD ld, ST st) {
    for (int base = threadIdx.x; base < n; base += U * (int)blockDim.x) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * (int)blockDim.x;
            if (i < n) ld(u, i);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * (int)blockDim.x;
            if (i < n) st(u, i);
        }
    }
}

// ------------------------------------------------------------------ pass A
template <class T>
__global__ void __launch_bounds__(512)
k_hessm_rows_fwd(const T *__restrict__ x, const T *__restrict__ beam, Cx<T> *__restrict__ S,
                 const Cx<T> *__restrict__ tw, const Cx<T> *__restrict__ twr,
                 const Cx<T> *__restrict__ roots, MxDesc d, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int M = h.M, L = h.R;
    Cx<T> *A = rt + kMxHead, *B = A + (long)L * M;
    const int band = blockIdx.y, row0 = blockIdx.x * L;
    mx_load_tables(rt, roots, tw, M);
    {
        T va[8], vb[8];
        peel_batched<8>(L * M,
            [&](int u, int i) {
                const int l = i / M, m = i - l * M, row = row0 + l, c0 = 2 * m;
                va[u] = vb[u] = (T)0;
                if (row < h.nx) {
                    const long ro = ((long)band * h.nx + row) * h.ny;
                    if (c0 < h.ny) { va[u] = x[ro + c0]; if (beam) va[u] *= beam[ro + c0]; }
                    if (c0 + 1 < h.ny) { vb[u] = x[ro + c0 + 1]; if (beam) vb[u] *= beam[ro + c0 + 1]; }
                }
            },
            [&](int u, int i) { A[i] = cx<T>(va[u], vb[u]); });
    }
    __syncthreads();
    const Cx<T> *Z = mx_fft<T, -1>(A, B, L, d, rt + 40, rt);
    // X[k] = E + w^k O,  E = (Z[k] + conj Z[M-k])/2,  O = -i (Z[k] - conj Z[M-k])/2
    Cx<T> w[4];
    peel_batched<4>((M + 1) * L,
        [&](int u, int i) { w[u] = twr[i / L]; },
        [&](int u, int i) {
            const int k = i / L, l = i - k * L;
            const Cx<T> *z = Z + (long)l * M;
            const Cx<T> zk = z[k == M ? 0 : k], zm = cconj(z[k == 0 ? 0 : M - k]);
            const Cx<T> E = cx<T>((T)0.5 * (zk.x + zm.x), (T)0.5 * (zk.y + zm.y));
            const Cx<T> D = cx<T>((T)0.5 * (zk.x - zm.x), (T)0.5 * (zk.y - zm.y));
            const Cx<T> O = cx<T>(D.y, -D.x);
            const Cx<T> X = E + cmul(w[u], O);
            if (row0 + l < h.nx) S[((long)band * h.nyo2 + k) * h.nx + row0 + l] = X;
        });
}

// ------------------------------------------------------------------ pass B
// SPLIT: psfT de-interleaved [even | odd] per column, modg[n] = exp(-2 pi i n / nxp)
template <class T, bool SPLIT>
__global__ void __launch_bounds__(512)
k_hessm_cols(Cx<T> *S, const Cx<T> *__restrict__ psfT, const Cx<T> *__restrict__ tw,
             const Cx<T> *__restrict__ modg, const Cx<T> *__restrict__ roots, MxDesc d, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int N = d.N;
    Cx<T> *A = rt + kMxHead, *B = A + N;
    const int band = blockIdx.y, col = blockIdx.x;
    Cx<T> *line = S + ((long)band * h.nyo2 + col) * h.nx;
    const Cx<T> *pl = psfT + ((long)band * h.nyo2 + col) * h.nxp;
    mx_load_tables(rt, roots, tw, N);
    const Cx<T> *tws = rt + 40;
    constexpr int U1 = sizeof(T) == 4 ? 8 : 4, U2 = U1 / 2;   // batch depths by register budget
    Cx<T> v[U1], g[U2];
    peel_batched<U1>(N,
        [&](int u, int n) { v[u] = n < h.nx ? line[n] : cx<T>((T)0, (T)0); },
        [&](int u, int n) { A[n] = v[u]; });
    __syncthreads();
    Cx<T> *c = mx_fft<T, -1>(A, B, 1, d, tws, rt);
    Cx<T> *o = c == A ? B : A;
    peel_batched<U1>(N,
        [&](int u, int n) { v[u] = pl[n]; },
        [&](int u, int n) { c[n] = cmul(c[n], v[u]); });
    __syncthreads();
    Cx<T> *r = mx_fft<T, +1>(c, o, 1, d, tws, rt);
    if (!SPLIT) {
        for (int n = threadIdx.x; n < h.nx; n += blockDim.x) line[n] = r[n];
        return;
    }
    // odd spectrum: input a[n] w^n; park the even result in the (now dead) S line
    o = r == A ? B : A;
    peel_batched<U2>(N,
        [&](int u, int n) {
            v[u] = cx<T>((T)0, (T)0);
            if (n < h.nx) { v[u] = line[n]; g[u] = modg[n]; }
        },
        [&](int u, int n) { o[n] = n < h.nx ? cmul(v[u], g[u]) : cx<T>((T)0, (T)0); });
    __syncthreads();                              // every read of the line is done
    for (int n = threadIdx.x; n < h.nx; n += blockDim.x) line[n] = r[n];
    __syncthreads();
    c = mx_fft<T, -1>(o, r, 1, d, tws, rt);
    o = c == A ? B : A;
    peel_batched<U1>(N,
        [&](int u, int n) { v[u] = pl[N + n]; },
        [&](int u, int n) { c[n] = cmul(c[n], v[u]); });
    __syncthreads();
    r = mx_fft<T, +1>(c, o, 1, d, tws, rt);
    // same thread wrote line[n] above (same index mapping): no fence needed
    peel_batched<U2>(h.nx,
        [&](int u, int n) { v[u] = line[n]; g[u] = modg[n]; },
        [&](int u, int n) { line[n] = v[u] + cmul(r[n], cconj(g[u])); });
}

// ------------------------------------------------------------------ pass C
template <class T>
__global__ void __launch_bounds__(512)
k_hessm_rows_inv(const Cx<T> *__restrict__ S, T *__restrict__ out, const Cx<T> *__restrict__ tw,
                 const Cx<T> *__restrict__ twr, const Cx<T> *__restrict__ roots, MxDesc d,
                 HessDims h, Epilogue<T> ep) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int M = h.M, L = h.R;
    Cx<T> *A = rt + kMxHead, *B = A + (long)L * M;
    const int band = blockIdx.y, row0 = blockIdx.x * L;
    mx_load_tables(rt, roots, tw, M);
    // Z'[m] = (X[m] + conj X[M-m]) + i w^-m (X[m] - conj X[M-m])   (= 2 Z[m]), m < M
    {
        constexpr int U3 = sizeof(T) == 4 ? 8 : 2;
        Cx<T> xm[U3], xr[U3], w[U3];
        peel_batched<U3>(M * L,
            [&](int u, int i) {
                const int m = i / L, l = i - m * L;
                xm[u] = xr[u] = cx<T>((T)0, (T)0);
                if (row0 + l < h.nx) {
                    const Cx<T> *col = S + (long)band * h.nyo2 * h.nx + row0 + l;
                    xm[u] = col[(long)m * h.nx];
                    xr[u] = col[(long)(M - m) * h.nx];
                }
                w[u] = twr[m];
            },
            [&](int u, int i) {
                const int m = i / L, l = i - m * L;
                const Cx<T> c = cconj(xr[u]);
                const Cx<T> P = xm[u] + c, Q = xm[u] - c;
                const Cx<T> wQ = cmul(cconj(w[u]), Q);
                A[(long)l * M + m] = cx<T>(P.x - wQ.y, P.y + wQ.x);
            });
    }
    __syncthreads();
    const Cx<T> *Z = mx_fft<T, +1>(A, B, L, d, rt + 40, rt);
    for (int i = threadIdx.x; i < L * M; i += blockDim.x) {
        const int l = i / M, m = i - l * M, row = row0 + l, c0 = 2 * m;
        if (row >= h.nx) continue;
        const long ro = ((long)band * h.nx + row) * h.ny;
        const Cx<T> z = Z[i];
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
