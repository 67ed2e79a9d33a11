// Shared-memory Stockham FFT building blocks (power-of-two lengths, sm_100a).
//
// One transform of length N is done by NT = N/16 threads that each keep 16
// complex points in registers.  A stage of radix R (16, 8, 4 or 2) does 16/R
// register butterflies per thread.  With the Stockham auto-sort indexing
//     butterfly j in [0, N/R):  k = j mod Ns,
//     in : v[r] = buf[j + r*N/R] * w^(r*k),  w = exp(-/+ 2 pi i / (Ns*R))
//     out: buf[(j-k)*R + k + r*Ns] = DFT_R(v)[r]
// thread t (butterflies j = t + b*NT) always READS the 16 points {t + e*NT}:
// conflict-free in shared memory and coalesced when the first stage reads from
// global memory.  The LAST stage writes {t + e*NT} again, so its outputs can go
// straight to global memory (coalesced) or be consumed in registers by a
// following inverse transform without any exchange.  Only the S-1 exchanges
// between stages go through shared memory (index padded by i + i/16 so the
// stride-16 scatter of the first exchange is bank-conflict free).
// Twiddles come from a table of exp(-2 pi i m / twN) computed in fp64 on the
// host and rounded once, so fp32 accuracy is that of an exact-twiddle FFT.
#pragma once
#include "common.cuh"

namespace pfb {

template <class T> struct alignas(2 * sizeof(T)) Cx { T x, y; };

template <class T> __device__ __forceinline__ Cx<T> cx(T a, T b) { Cx<T> r; r.x = a; r.y = b; return r; }
template <class T> __device__ __forceinline__ Cx<T> operator+(Cx<T> a, Cx<T> b) { return cx<T>(a.x + b.x, a.y + b.y); }
template <class T> __device__ __forceinline__ Cx<T> operator-(Cx<T> a, Cx<T> b) { return cx<T>(a.x - b.x, a.y - b.y); }
template <class T> __device__ __forceinline__ Cx<T> cmul(Cx<T> a, Cx<T> b) {
    return cx<T>(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
template <class T> __device__ __forceinline__ Cx<T> cconj(Cx<T> a) { return cx<T>(a.x, -a.y); }
// multiply by -i (forward, DIR = -1) or +i (inverse, DIR = +1)
template <class T, int DIR> __device__ __forceinline__ Cx<T> mul_i(Cx<T> a) {
    return DIR < 0 ? cx<T>(a.y, -a.x) : cx<T>(-a.y, a.x);
}
// twiddle for direction: table holds exp(-2 pi i m/twN); inverse uses the conjugate
template <class T, int DIR> __device__ __forceinline__ Cx<T> twd(Cx<T> w) {
    return DIR < 0 ? w : cconj(w);
}

template <class T, int DIR> __device__ __forceinline__ void dft2(Cx<T> &a, Cx<T> &b) {
    Cx<T> t = a - b;
    a = a + b;
    b = t;
}

template <class T, int DIR>
__device__ __forceinline__ void dft4(Cx<T> &a0, Cx<T> &a1, Cx<T> &a2, Cx<T> &a3) {
    Cx<T> s02 = a0 + a2, d02 = a0 - a2, s13 = a1 + a3, d13 = mul_i<T, DIR>(a1 - a3);
    a0 = s02 + s13;
    a1 = d02 + d13;
    a2 = s02 - s13;
    a3 = d02 - d13;
}

// constants exp(-2 pi i m/16), m = 1..3 (cos, sin) ; w8 = w16^2
#define PFB_C1 0.92387953251128673848
#define PFB_S1 0.38268343236508978178
#define PFB_R2 0.70710678118654752440

// DFT_R on u[0..R), natural order in and out (Cooley-Tukey n = 4*n1 + n2 style
// decompositions written out; w_R^m = exp(DIR * 2 pi i m / R))
template <class T, int R, int DIR> struct Dft;

template <class T, int DIR> struct Dft<T, 2, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[2]) { dft2<T, DIR>(u[0], u[1]); }
};
template <class T, int DIR> struct Dft<T, 4, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[4]) { dft4<T, DIR>(u[0], u[1], u[2], u[3]); }
};
template <class T, int DIR> struct Dft<T, 8, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[8]) {
        // n = 2*n1 + n2 (n1 < 4, n2 < 2); X[k1 + 4*k2] = sum_n2 w8^(n2 k1) Y[n2][k1] (-1)^(n2 k2)
        dft4<T, DIR>(u[0], u[2], u[4], u[6]);
        dft4<T, DIR>(u[1], u[3], u[5], u[7]);
        const T r = (T)PFB_R2;
        Cx<T> w1 = cx<T>(r, DIR < 0 ? -r : r);
        Cx<T> w3 = cx<T>(-r, DIR < 0 ? -r : r);
        u[3] = cmul(u[3], w1);
        u[5] = mul_i<T, DIR>(u[5]);
        u[7] = cmul(u[7], w3);
        // now u[2*k1 + n2] holds Y[n2][k1] w8^(n2 k1); combine over n2
        Cx<T> o[8];
#pragma unroll
        for (int k1 = 0; k1 < 4; ++k1) {
            o[k1] = u[2 * k1] + u[2 * k1 + 1];
            o[k1 + 4] = u[2 * k1] - u[2 * k1 + 1];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) u[i] = o[i];
    }
};
template <class T, int DIR> struct Dft<T, 16, DIR> {
    static __device__ __forceinline__ void run(Cx<T> (&u)[16]) {
        // n = 4*n1 + n2; Y[n2][k1] = DFT4_{n1} u[4 n1 + n2]; X[k1 + 4 k2] = DFT4_{n2} (w16^(n2 k1) Y[n2][k1])
#pragma unroll
        for (int n2 = 0; n2 < 4; ++n2) dft4<T, DIR>(u[n2], u[4 + n2], u[8 + n2], u[12 + n2]);
        // after dft4 in place: u[4*k1 + n2] = Y[n2][k1]
        const T c1 = (T)PFB_C1, s1 = (T)PFB_S1, r = (T)PFB_R2;
        const T sg = DIR < 0 ? (T)-1 : (T)1;
        // w16^m = (cos, sg*sin): m=1 (c1,s1) m=2 (r,r) m=3 (s1,c1) m=4 (0,1) m=6 (-r,r) m=9 (-c1,-s1)
        u[4 * 1 + 1] = cmul(u[4 * 1 + 1], cx<T>(c1, sg * s1));     // m = 1
        u[4 * 1 + 2] = cmul(u[4 * 1 + 2], cx<T>(r, sg * r));       // m = 2
        u[4 * 1 + 3] = cmul(u[4 * 1 + 3], cx<T>(s1, sg * c1));     // m = 3
        u[4 * 2 + 1] = cmul(u[4 * 2 + 1], cx<T>(r, sg * r));       // m = 2
        u[4 * 2 + 2] = mul_i<T, DIR>(u[4 * 2 + 2]);                // m = 4
        u[4 * 2 + 3] = cmul(u[4 * 2 + 3], cx<T>(-r, sg * r));      // m = 6
        u[4 * 3 + 1] = cmul(u[4 * 3 + 1], cx<T>(s1, sg * c1));     // m = 3
        u[4 * 3 + 2] = cmul(u[4 * 3 + 2], cx<T>(-r, sg * r));      // m = 6
        u[4 * 3 + 3] = cmul(u[4 * 3 + 3], cx<T>(-c1, -sg * s1));   // m = 9
        Cx<T> o[16];
#pragma unroll
        for (int k1 = 0; k1 < 4; ++k1) {
            Cx<T> a0 = u[4 * k1], a1 = u[4 * k1 + 1], a2 = u[4 * k1 + 2], a3 = u[4 * k1 + 3];
            dft4<T, DIR>(a0, a1, a2, a3);
            o[k1] = a0; o[k1 + 4] = a1; o[k1 + 8] = a2; o[k1 + 12] = a3;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) u[i] = o[i];
    }
};

// stage plan of one transform length (host fills, passed by value)
struct FftDesc {
    int N;          // transform length (power of two, >= 16)
    int NT;         // threads per transform = N/16
    int nstage;
    int radix[8];
    int tw_step;    // twN / N : table index stride for this length
};

__device__ __forceinline__ int fft_pad(int i) { return i + (i >> 4); }
__host__ __device__ static inline int fft_smem_elems(int N) { return N + (N >> 4) + 8; }

// One radix-R stage on the 16 register points of thread t.  Element slot
// e = b + r*(16/R) is point r of butterfly j = t + b*NT.  If `exchange`, the
// outputs are scattered to shared memory at their Stockham positions and the
// next stage's inputs {t + e*NT} are gathered back; the last stage leaves
// v[e] = X[t + e*NT] in registers.
template <class T, int R, int DIR>
__device__ __forceinline__ void fft_stage(Cx<T> (&v)[16], int Ns, int NT, int t,
                                          Cx<T> *buf, const Cx<T> *__restrict__ tw,
                                          int tw_mul, bool exchange) {
    constexpr int B = 16 / R;
#pragma unroll
    for (int b = 0; b < B; ++b) {
        int j = t + b * NT;
        int k = j & (Ns - 1);
        Cx<T> u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = v[b + r * B];
        if (Ns > 1) {
#pragma unroll
            for (int r = 1; r < R; ++r) u[r] = cmul(u[r], twd<T, DIR>(tw[(long)r * k * tw_mul]));
        }
        Dft<T, R, DIR>::run(u);
#pragma unroll
        for (int r = 0; r < R; ++r) v[b + r * B] = u[r];
    }
    if (exchange) {
        __syncthreads();   // previous readers of buf are done
#pragma unroll
        for (int b = 0; b < B; ++b) {
            int j = t + b * NT;
            int k = j & (Ns - 1);
            int base = (j - k) * R + k;
#pragma unroll
            for (int r = 0; r < R; ++r) buf[fft_pad(base + r * Ns)] = v[b + r * B];
        }
        __syncthreads();
#pragma unroll
        for (int e = 0; e < 16; ++e) v[e] = buf[fft_pad(t + e * NT)];
    }
}

// Full transform: in v[e] = x[t + e*NT], out v[e] = X[t + e*NT] (unnormalised).
// All threads of the block must call it (block-wide barriers); `buf` is this
// transform's own shared-memory line of fft_smem_elems(N) points.
template <class T, int DIR>
__device__ __forceinline__ void fft_run(Cx<T> (&v)[16], const FftDesc &d, int t,
                                        Cx<T> *buf, const Cx<T> *__restrict__ tw) {
    int Ns = 1;
    for (int s = 0; s < d.nstage; ++s) {
        int R = d.radix[s];
        bool ex = s + 1 < d.nstage;
        int tw_mul = d.tw_step * (d.N / (Ns * R));
        switch (R) {
        case 16: fft_stage<T, 16, DIR>(v, Ns, d.NT, t, buf, tw, tw_mul, ex); break;
        case 8: fft_stage<T, 8, DIR>(v, Ns, d.NT, t, buf, tw, tw_mul, ex); break;
        case 4: fft_stage<T, 4, DIR>(v, Ns, d.NT, t, buf, tw, tw_mul, ex); break;
        default: fft_stage<T, 2, DIR>(v, Ns, d.NT, t, buf, tw, tw_mul, ex); break;
        }
        Ns *= R;
    }
}

// host: stage plan for length N (power of two >= 16): radix-16 stages first,
// the 2/4/8 remainder last.
static inline bool fft_make_desc(int N, int twN, FftDesc &d) {
    if (N < 16 || (N & (N - 1)) || twN % N) return false;
    d.N = N;
    d.NT = N / 16;
    d.tw_step = twN / N;
    d.nstage = 0;
    int rem = N;
    while (rem >= 16) { d.radix[d.nstage++] = 16; rem /= 16; }
    if (rem > 1) d.radix[d.nstage++] = rem;
    return d.nstage <= 8;
}

}  // namespace pfb
