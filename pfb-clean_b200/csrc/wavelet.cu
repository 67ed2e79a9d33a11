// Multi-basis Daubechies wavelet dictionary Psi (analysis `dot`, synthesis
// `hdot`) in the reference's exact coefficient layout (sm_100a).
//
// Reference: pfb/wavelets/wavelets.py (dwt2d :174, idwt2d :260, mode 'zero',
// transposed output) and pfb/operators/psi.py (psi_band.dot :187, hdot :221,
// pyramid bookkeeping :49-94).  Layout recap: per band and
// basis a (Nymax, Nxmax) slab, rows = y (image axis 1) index, cols = x (image
// axis 0) index; level l owns the block rows [lowy,highy) x cols [lowx,highx)
// with quadrants TL = LL (only kept for the last level), TR = x-hi/y-lo,
// BL = x-lo/y-hi, BR = hi/hi.  Cells outside the blocks are never written.
//
// Kernel design: one CTA = one 32x32 tile of coefficient positions (64x64
// pixels) of one (band, basis).  The tile is staged in shared memory, the row
// and column passes run out of shared memory with register sliding windows
// (each thread produces 4 consecutive outputs), and the 2-D transpose implied
// by the layout is folded into the shared-memory staging so that every global
// load and store is a coalesced row segment.  One launch per level covers all
// bands and bases; the final synthesis level accumulates all bases in shared
// memory and writes the image once.  HBM-bound: per level-0 tile 16 KB in,
// 16 KB out (fp32), ~64 FMA per 32 B.
#include "common.cuh"
#include "filters.h"

namespace pfb {

static const int kMaxBasis = 8;
static const int kMaxLevel = 10;
static const int TX = 32, TY = 32;   // tile of coefficient positions
static const int NS = 4;             // outputs per thread along the filter axis

__constant__ double c_taps[5][4][10] = PFB_TAPS_INIT;
__constant__ float c_tapsf[5][4][10] = PFB_TAPS_INIT;   // same taps rounded once to fp32

// tap as an operand of type T (compile-time indices fold into a constant-bank operand)
template <class T> __device__ __forceinline__ T tapc(int n, int which, int k);
template <> __device__ __forceinline__ float tapc<float>(int n, int which, int k) { return c_tapsf[n][which][k]; }
template <> __device__ __forceinline__ double tapc<double>(int n, int which, int k) { return c_taps[n][which][k]; }

// per-basis description of one level (device-visible, passed by value)
struct LevelDesc {
    int F;                 // filter length, 0 = 'self' basis (identity)
    int n0, n1;            // analysis input / synthesis output size (axis0, axis1)
    int C0, C1;            // coefficient sizes (x, y)
    int lowx, lowy;        // block origin in the slab (col, row)
    int ld_in;             // row stride of the natural-orientation scratch read
    int ld_out;            // row stride of the natural-orientation scratch written
    int ldn, qs;           // natural-orientation coefficient layout (wavelet_syn_stream.cuh):
    long off1;             //   row stride, elements per quadrant, offset of quadrant 1
};

struct LevelArgs {
    LevelDesc d[kMaxBasis];
    int nbasis, nband;
    int ntile0, ntile1;            // tiles along axis 0 / axis 1
    int Nxmax, Nymax;              // slab is (Nymax, Nxmax)
    long img_band_stride;          // nx*ny
    long scr_stride;               // elements per (band,basis) scratch image
    long cn_stride;                // elements per (band,basis) block of the natural layout
    int first, last;               // level is the finest / the coarsest
    int nx, ny;
    // fused dual update of the natural-layout analysis (pfb_psi_dot_nat_dual): when fz_vp is set
    // a coefficient u is not stored; v = (vp + sigma u) * f is, with vp laid out like the
    // output and f like ONE band of it (fz_dvp / fz_df: element offsets from the output cube to
    // vp and to f's band-0 position; the kernels subtract the band part)
    const void *fz_vp;
    long fz_dvp, fz_df;
    double fz_sigma;
};

template <class T, int F>
__device__ __forceinline__ void load_taps(int which, T (&f)[F]) {
#pragma unroll
    for (int k = 0; k < F; ++k) f[k] = (T)c_taps[F / 2 - 1][which][k];
}

// ---------------------------------------------------------------------------
// Analysis tile.  in: natural orientation (n0 x n1, ld).  Shared memory:
//   s_in  [IX][PIN]        input tile incl. halo (zero outside the image)
//   s_mid [2][TY][PM]      row-pass output, transposed: [lo/hi][t][r]
//   s_out [4][TY][TX+1]    the four sub-bands [q][t][s]   (aliases s_in)
template <class T, int F>
__device__ void ana_tile(const T *__restrict__ in, int ld, const LevelDesc &d,
                         T *slab, int ldslab, T *ll_out, bool last, int tile0,
                         int tile1, T *smem) {
    constexpr int IX = 2 * TX + F - 2, IY = 2 * TY + F - 2;
    constexpr int PIN = (IY % 2 == 0) ? IY + 1 : IY;      // odd pitch
    constexpr int PM = (IX % 2 == 0) ? IX + 1 : IX;
    constexpr int W = 2 * NS + F - 2;                      // window per thread
    T *s_in = smem;
    T *s_mid = smem + 72 * 73;     // sized for F = 10 (IX = 72, PIN = 73)
    T *s_out = smem;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int jx0 = tile0 * TX, jy0 = tile1 * TY;
    if (jx0 >= d.C0 || jy0 >= d.C1) return;
    const int gr0 = 2 * jx0 - F + 2, gc0 = 2 * jy0 - F + 2;

    for (int idx = tid; idx < IX * IY; idx += nthr) {
        int r = idx / IY, c = idx - r * IY;
        int gr = gr0 + r, gc = gc0 + c;
        T v = (T)0;
        if (gr >= 0 && gr < d.n0 && gc >= 0 && gc < d.n1) v = in[(long)gr * ld + gc];
        s_in[r * PIN + c] = v;
    }
    __syncthreads();

    T dlo[F], dhi[F];
    load_taps<T, F>(0, dlo);
    load_taps<T, F>(1, dhi);

    // row pass (filter along axis 1): item = (row r, chunk of NS outputs)
    for (int item = tid; item < IX * (TY / NS); item += nthr) {
        int ch = item / IX, r = item - ch * IX;
        int t0 = ch * NS;
        T w[W];
#pragma unroll
        for (int j = 0; j < W; ++j) w[j] = s_in[r * PIN + 2 * t0 + j];
#pragma unroll
        for (int t = 0; t < NS; ++t) {
            T lo = (T)0, hi = (T)0;
#pragma unroll
            for (int k = 0; k < F; ++k) {
                T x = w[2 * t + F - 1 - k];
                lo += dlo[k] * x;
                hi += dhi[k] * x;
            }
            s_mid[(0 * TY + t0 + t) * PM + r] = lo;
            s_mid[(1 * TY + t0 + t) * PM + r] = hi;
        }
    }
    __syncthreads();

    // column pass (filter along axis 0): item = (t, chunk of NS outputs s)
    for (int item = tid; item < TY * (TX / NS); item += nthr) {
        int sc = item / TY, t = item - sc * TY;
        int s0 = sc * NS;
        T wl[W], wh[W];
#pragma unroll
        for (int j = 0; j < W; ++j) {
            wl[j] = s_mid[(0 * TY + t) * PM + 2 * s0 + j];
            wh[j] = s_mid[(1 * TY + t) * PM + 2 * s0 + j];
        }
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            T ll = (T)0, lh = (T)0, hl = (T)0, hh = (T)0;
#pragma unroll
            for (int k = 0; k < F; ++k) {
                T a = wl[2 * s + F - 1 - k], b = wh[2 * s + F - 1 - k];
                ll += dlo[k] * a;   // x-lo of y-lo  -> TL
                lh += dhi[k] * a;   // x-hi of y-lo  -> TR
                hl += dlo[k] * b;   // x-lo of y-hi  -> BL
                hh += dhi[k] * b;   // x-hi of y-hi  -> BR
            }
            int o = t * (TX + 1) + s0 + s;
            s_out[0 * TY * (TX + 1) + o] = ll;
            s_out[1 * TY * (TX + 1) + o] = lh;
            s_out[2 * TY * (TX + 1) + o] = hl;
            s_out[3 * TY * (TX + 1) + o] = hh;
        }
    }
    __syncthreads();

    // store: slab rows = y index, cols = x index (lanes along s -> coalesced)
    for (int idx = tid; idx < 4 * TY * TX; idx += nthr) {
        int q = idx / (TY * TX), rem = idx - q * (TY * TX);
        int t = rem / TX, s = rem - t * TX;
        int jx = jx0 + s, jy = jy0 + t;
        if (jx >= d.C0 || jy >= d.C1) continue;
        if (q == 0 && !last) continue;   // LL of a non-final level: scratch only
        int qx = q & 1, qy = q >> 1;
        slab[(long)(d.lowy + qy * d.C1 + jy) * ldslab + d.lowx + qx * d.C0 + jx] =
            s_out[q * TY * (TX + 1) + t * (TX + 1) + s];
    }
    if (!last) {   // LL in natural orientation for the next level (lanes along t)
        for (int idx = tid; idx < TY * TX; idx += nthr) {
            int s = idx / TY, t = idx - s * TY;
            int jx = jx0 + s, jy = jy0 + t;
            if (jx < d.C0 && jy < d.C1)
                ll_out[(long)jx * d.ld_out + jy] = s_out[t * (TX + 1) + s];
        }
    }
}

// 'self' basis: alphao[i, 0:ny, 0:nx] = x.T  (pfb/operators/psi.py:196-199)
template <class T>
__device__ void self_tile(const T *__restrict__ x, int nx, int ny, T *slab,
                          int ldslab, int tile0, int tile1, T *smem) {
    const int B = 64;
    int r0 = tile0 * B, c0 = tile1 * B;
    if (r0 >= nx || c0 >= ny) return;
    for (int idx = threadIdx.x; idx < B * B; idx += blockDim.x) {
        int r = idx / B, c = idx - r * B;
        T v = (T)0;
        if (r0 + r < nx && c0 + c < ny) v = x[(long)(r0 + r) * ny + c0 + c];
        smem[r * (B + 1) + c] = v;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < B * B; idx += blockDim.x) {
        int c = idx / B, r = idx - c * B;
        if (r0 + r < nx && c0 + c < ny)
            slab[(long)(c0 + c) * ldslab + r0 + r] = smem[r * (B + 1) + c];
    }
}

// grid: x = tile1 * nbasis + basis, y = tile0, z = band
template <class T>
__global__ void __launch_bounds__(256)
k_dwt_level(const T *__restrict__ img, const T *__restrict__ scr_in, T *scr_out,
            T *alpha, LevelArgs a) {
    PFB_DYN_SMEM(raw);
    T *smem = reinterpret_cast<T *>(raw);
    int basis = blockIdx.x % a.nbasis, tile1 = blockIdx.x / a.nbasis;
    int tile0 = blockIdx.y, band = blockIdx.z;
    const LevelDesc &d = a.d[basis];
    long slabsz = (long)a.Nymax * a.Nxmax;
    T *slab = alpha + ((long)band * a.nbasis + basis) * slabsz;
    if (d.F == 0) {
        if (a.first) self_tile<T>(img + band * a.img_band_stride, a.nx, a.ny, slab, a.Nxmax, tile0, tile1, smem);
        return;
    }
    long so = ((long)band * a.nbasis + basis) * a.scr_stride;
    const T *in = a.first ? img + band * a.img_band_stride : scr_in + so;
    int ld = a.first ? a.ny : d.ld_in;
    T *ll = scr_out + so;
    switch (d.F) {
    case 2: ana_tile<T, 2>(in, ld, d, slab, a.Nxmax, ll, a.last, tile0, tile1, smem); break;
    case 4: ana_tile<T, 4>(in, ld, d, slab, a.Nxmax, ll, a.last, tile0, tile1, smem); break;
    case 6: ana_tile<T, 6>(in, ld, d, slab, a.Nxmax, ll, a.last, tile0, tile1, smem); break;
    case 8: ana_tile<T, 8>(in, ld, d, slab, a.Nxmax, ll, a.last, tile0, tile1, smem); break;
    case 10: ana_tile<T, 10>(in, ld, d, slab, a.Nxmax, ll, a.last, tile0, tile1, smem); break;
    }
}

}  // namespace pfb
#include "wavelet_stream.cuh"
#include "wavelet_syn_stream.cuh"
namespace pfb {

// shared memory (elements of T) needed by k_dwt_level
static const int kAnaSmemElems = 72 * 73 + 2 * TY * 73;

// ---------------------------------------------------------------------------
// Synthesis tile: adds the reconstruction of one (band, basis, level) tile of
// 64x64 output pixels into s_acc.  Shared memory:
//   s_c [4][TYH][PC]   coefficient tiles [q][iy][ix], TYH = TY + 4, PC = 37
//   s_t [2][TYH][PT]   x-synthesised rows [y-lo/y-hi][iy][X], PT = 65
//   s_acc[64][65]      output tile [X][Y] (accumulated over bases)
// coefficient source: slab (transposed orientation); the LL quadrant comes
// from `ll_in` (natural orientation scratch = previous level's output) unless
// this is the coarsest level.  If slab2 != null the coefficient used is
// 2*slab - slab2 (fuses `vp = 2v - vp` of pfb/opt/primal_dual.py:137).
static const int TYH = TY + 4, TXH = TX + 4, PC = TXH + 1, PT = 2 * TX + 1;
static const int kSynSmemElems = 4 * TYH * PC + 2 * TYH * PT + 64 * 65;

template <class T, int F>
__device__ void syn_tile(const T *__restrict__ slab, const T *__restrict__ slab2,
                         int ldslab, const T *__restrict__ ll_in, bool last,
                         const LevelDesc &d, int tile0, int tile1, T *smem) {
    constexpr int H = F / 2 - 1;
    constexpr int W = NS + H;
    T *s_c = smem;
    T *s_t = smem + 4 * TYH * PC;
    T *s_acc = s_t + 2 * TYH * PT;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int mx0 = tile0 * TX, my0 = tile1 * TY;
    const int nix = TX + H, niy = TY + H;

    // load the four quadrant tiles
    for (int q = 0; q < 4; ++q) {
        int qx = q & 1, qy = q >> 1;
        if (q == 0 && !last) {   // LL from the previous level's output [cx][cy]
            for (int idx = tid; idx < nix * niy; idx += nthr) {
                int ix = idx / niy, iy = idx - ix * niy;
                int cx = mx0 + ix, cy = my0 + iy;
                T v = (T)0;
                if (cx < d.C0 && cy < d.C1) v = ll_in[(long)cx * d.ld_in + cy];
                s_c[iy * PC + ix] = v;
            }
        } else {
            for (int idx = tid; idx < nix * niy; idx += nthr) {
                int iy = idx / nix, ix = idx - iy * nix;
                int cx = mx0 + ix, cy = my0 + iy;
                T v = (T)0;
                if (cx < d.C0 && cy < d.C1) {
                    long o = (long)(d.lowy + qy * d.C1 + cy) * ldslab + d.lowx + qx * d.C0 + cx;
                    v = slab[o];
                    if (slab2) v = (T)2 * v - slab2[o];
                }
                s_c[(q * TYH + iy) * PC + ix] = v;
            }
        }
    }
    __syncthreads();

    T rlo[F], rhi[F];
    load_taps<T, F>(2, rlo);
    load_taps<T, F>(3, rhi);

    // x direction: t_hy[iy][2mx+p] = sum_j rlo[2j+p] c_{hy,lo}[iy][mx+H-j] + rhi[2j+p] c_{hy,hi}[..]
    for (int item = tid; item < 2 * niy * (TX / NS); item += nthr) {
        int iy = item % niy, rest = item / niy;
        int hy = rest & 1, mc = rest >> 1;
        int m0 = mc * NS;
        T wl[W], wh[W];
#pragma unroll
        for (int j = 0; j < W; ++j) {
            wl[j] = s_c[((2 * hy + 0) * TYH + iy) * PC + m0 + j];
            wh[j] = s_c[((2 * hy + 1) * TYH + iy) * PC + m0 + j];
        }
#pragma unroll
        for (int m = 0; m < NS; ++m) {
            T e = (T)0, o = (T)0;
#pragma unroll
            for (int j = 0; j <= H; ++j) {
                T a = wl[m + H - j], b = wh[m + H - j];
                e += rlo[2 * j] * a + rhi[2 * j] * b;
                o += rlo[2 * j + 1] * a + rhi[2 * j + 1] * b;
            }
            s_t[(hy * TYH + iy) * PT + 2 * (m0 + m)] = e;
            s_t[(hy * TYH + iy) * PT + 2 * (m0 + m) + 1] = o;
        }
    }
    __syncthreads();

    // y direction: out[X][2my+p] += sum_j rlo[2j+p] t_lo[my+H-j][X] + rhi[2j+p] t_hi[..][X]
    for (int item = tid; item < 2 * TX * (TY / NS); item += nthr) {
        int X = item % (2 * TX), mc = item / (2 * TX);
        int m0 = mc * NS;
        T wl[W], wh[W];
#pragma unroll
        for (int j = 0; j < W; ++j) {
            wl[j] = s_t[(0 * TYH + m0 + j) * PT + X];
            wh[j] = s_t[(1 * TYH + m0 + j) * PT + X];
        }
#pragma unroll
        for (int m = 0; m < NS; ++m) {
            T e = (T)0, o = (T)0;
#pragma unroll
            for (int j = 0; j <= H; ++j) {
                T a = wl[m + H - j], b = wh[m + H - j];
                e += rlo[2 * j] * a + rhi[2 * j] * b;
                o += rlo[2 * j + 1] * a + rhi[2 * j + 1] * b;
            }
            s_acc[X * 65 + 2 * (m0 + m)] += e;
            s_acc[X * 65 + 2 * (m0 + m) + 1] += o;
        }
    }
    __syncthreads();
}

// grid: x = tile1, y = tile0, z = band * nb + j  where the CTA loops over
// bases [j*per, (j+1)*per) (per = nbasis for the final, accumulating level;
// per = 1 for the coarser levels, which write one scratch image per basis).
template <class T>
__global__ void __launch_bounds__(256)
k_idwt_level(const T *__restrict__ alpha, const T *__restrict__ alpha2,
             const T *__restrict__ scr_in, T *out, LevelArgs a, int per) {
    PFB_DYN_SMEM(raw);
    T *smem = reinterpret_cast<T *>(raw);
    T *s_acc = smem + 4 * TYH * PC + 2 * TYH * PT;
    int tile1 = blockIdx.x, tile0 = blockIdx.y;
    int groups = a.nbasis / per;
    int band = blockIdx.z / groups, grp = blockIdx.z - band * groups;
    long slabsz = (long)a.Nymax * a.Nxmax;
    for (int idx = threadIdx.x; idx < 64 * 65; idx += blockDim.x) s_acc[idx] = (T)0;
    __syncthreads();
    int n0 = 0, n1 = 0, ldo = 0;
    bool any = false;
    for (int basis = grp * per; basis < (grp + 1) * per; ++basis) {
        const LevelDesc &d = a.d[basis];
        long so = ((long)band * a.nbasis + basis) * slabsz;
        const T *slab = alpha + so;
        const T *slab2 = alpha2 ? alpha2 + so : nullptr;
        if (d.F == 0) {
            if (!a.first) continue;
            // xo += alpha[i, 0:ny, 0:nx].T  (pfb/operators/psi.py:229-232)
            int X0 = tile0 * 64, Y0 = tile1 * 64;
            for (int idx = threadIdx.x; idx < 64 * 64; idx += blockDim.x) {
                int Y = idx / 64, X = idx - Y * 64;
                if (X0 + X < a.nx && Y0 + Y < a.ny) {
                    long o = (long)(Y0 + Y) * a.Nxmax + X0 + X;
                    T v = slab[o];
                    if (slab2) v = (T)2 * v - slab2[o];
                    s_acc[X * 65 + Y] += v;
                }
            }
            __syncthreads();
            n0 = a.nx; n1 = a.ny; ldo = a.ny; any = true;
            continue;
        }
        if (2 * tile0 * TX >= d.n0 || 2 * tile1 * TY >= d.n1) continue;
        const T *ll = scr_in + ((long)band * a.nbasis + basis) * a.scr_stride;
        switch (d.F) {
        case 2: syn_tile<T, 2>(slab, slab2, a.Nxmax, ll, a.last, d, tile0, tile1, smem); break;
        case 4: syn_tile<T, 4>(slab, slab2, a.Nxmax, ll, a.last, d, tile0, tile1, smem); break;
        case 6: syn_tile<T, 6>(slab, slab2, a.Nxmax, ll, a.last, d, tile0, tile1, smem); break;
        case 8: syn_tile<T, 8>(slab, slab2, a.Nxmax, ll, a.last, d, tile0, tile1, smem); break;
        case 10: syn_tile<T, 10>(slab, slab2, a.Nxmax, ll, a.last, d, tile0, tile1, smem); break;
        }
        n0 = d.n0; n1 = d.n1; ldo = a.first ? a.ny : d.ld_out; any = true;
    }
    if (!any) return;
    // store (lanes along Y -> coalesced).  Final level: the image (nx, ny);
    // coarser levels: scratch image of this (band, basis).
    T *dst = a.first ? out + band * a.img_band_stride
                     : out + ((long)band * a.nbasis + grp * per) * a.scr_stride;
    if (a.first) { n0 = a.nx; n1 = a.ny; ldo = a.ny; }
    int X0 = tile0 * 64, Y0 = tile1 * 64;
    for (int idx = threadIdx.x; idx < 64 * 64; idx += blockDim.x) {
        int X = idx / 64, Y = idx - X * 64;
        if (X0 + X < n0 && Y0 + Y < n1) dst[(long)(X0 + X) * ldo + Y0 + Y] = s_acc[X * 65 + Y];
    }
}

// ---------------------------------------------------------------------------
// host side: plan = pyramid bookkeeping of pfb/operators/psi.py:49-94
struct BasisPlan {
    int F;
    int sx[kMaxLevel], sy[kMaxLevel], spx[kMaxLevel], spy[kMaxLevel];
    int highx[kMaxLevel], highy[kMaxLevel];
    int Ntotx, Ntoty;
};

}  // namespace pfb

struct pfb_psi_plan {
    int nband, nx, ny, nbasis, nlevel, dtype;
    int Nxmax, Nymax;
    int basis_id[pfb::kMaxBasis];
    pfb::BasisPlan bp[pfb::kMaxBasis];
    long scr_stride;
    void *scrA, *scrB;
    // natural-orientation coefficient scratch of the streaming synthesis
    // (wavelet_syn_stream.cuh): per (band, basis) a block of cn_stride elements
    int syn_ok;                                          // bases are distinct db1..db4 (+ one self)
    long cn_stride;
    long cn_off1[pfb::kMaxBasis][pfb::kMaxLevel];        // quadrant 1 of level l
    int cn_ld[pfb::kMaxBasis][pfb::kMaxLevel];
    void *cn;
};

namespace pfb {

static inline int round4(int n) { return (n + 3) & ~3; }
static inline int round32(int n) { return (n + 31) & ~31; }
static int coeff_size(int n, int F) { return (n + F - 1) / 2; }
static int signal_size(int c, int F) { return 2 * c - F + 2; }

// pywt.dwt_max_level(n, w) = floor(log2(n / (F - 1)))
static int dwt_max_level(int n, int F) {
    if (F < 2 || n < F - 1) return 0;
    int lvl = 0;
    while ((long)(F - 1) << (lvl + 1) <= n) ++lvl;
    return lvl;
}

static void fill_common(const pfb_psi_plan *p, LevelArgs &a, int level) {
    a.nbasis = p->nbasis;
    a.nband = p->nband;
    a.Nxmax = p->Nxmax;
    a.Nymax = p->Nymax;
    a.img_band_stride = (long)p->nx * p->ny;
    a.scr_stride = p->scr_stride;
    a.cn_stride = p->cn_stride;
    a.first = level == 0;
    a.last = level == p->nlevel - 1;
    a.nx = p->nx;
    a.ny = p->ny;
}

// level ping-pong scratch, allocated lazily so that a plan used only for size
// queries (or only in the other precision) costs nothing
static int ensure_scratch(const pfb_psi_plan *cp) {
    pfb_psi_plan *p = const_cast<pfb_psi_plan *>(cp);
    if (p->nlevel < 2 || p->scrA) return PFB_OK;
    size_t bytes = dtype_size(p->dtype) * (size_t)p->scr_stride * p->nband * p->nbasis;
    if (cudaMalloc(&p->scrA, bytes) != cudaSuccess) {
        p->scrA = nullptr;
        return fail(PFB_ERR_ALLOC, "psi plan scratch allocation failed (%zu bytes x2)", bytes);
    }
    if (cudaMalloc(&p->scrB, bytes) != cudaSuccess) {
        cudaFree(p->scrA);
        p->scrA = p->scrB = nullptr;
        return fail(PFB_ERR_ALLOC, "psi plan scratch allocation failed (%zu bytes x2)", bytes);
    }
    return PFB_OK;
}

// Rows of coefficients per chunk of a streaming walk: long chunks amortise the warm-up rows of
// the register history, but with few bands (one band per GPU when the cube is sharded 8 ways,
// or the single band-sum image of pfb_dual_update_rec) they leave less than a wave of CTAs;
// shorten them until there are about two waves.
static int pick_nj(int c0, long ctas_per_chunk) {
    int NJ = c0 >= 1024 ? 128 : (c0 >= 256 ? 64 : 32);
    while (NJ > 32 && ctas_per_chunk * cdiv(c0, NJ) < 8L * kNumSM) NJ >>= 1;
    return NJ;
}

// nat = false: alpha is the reference's slab cube; nat = true: alpha is the
// natural-orientation layout (streaming kernels only).
template <class T>
static int psi_dot_t(const pfb_psi_plan *p, const T *x, T *alpha, bool nat, cudaStream_t st,
                     const T *fz_vp = nullptr, const T *fz_f = nullptr, double fz_sigma = 0.0) {
    if (int rc = ensure_scratch(p)) return rc;
    size_t smem = sizeof(T) * kAnaSmemElems;
    PFB_CUDA_OK(cudaFuncSetAttribute(&k_dwt_level<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    T *A = (T *)p->scrA, *B = (T *)p->scrB;
    // streaming kernel: 4-element vector loads of image rows, F <= 8
    bool stream = (p->ny % 2 == 0) && (((uintptr_t)x & 15) == 0) && !getenv("PFB_DWT_TILE");
    for (int b = 0; b < p->nbasis; ++b)
        if (p->bp[b].F > 8) stream = false;
    if (nat && !(stream && p->syn_ok))
        return fail(PFB_ERR_BAD_ARG, "natural-orientation layout needs distinct db1..db4 (+self) bases and a 16-byte aligned image");
    for (int l = 0; l < p->nlevel; ++l) {
        LevelArgs a;
        memset(&a, 0, sizeof a);
        fill_common(p, a, l);
        if (fz_vp) {
            a.fz_vp = fz_vp;
            a.fz_dvp = fz_vp - alpha;
            a.fz_df = fz_f - alpha;
            a.fz_sigma = fz_sigma;
        }
        int maxC0 = 0, maxC1 = 0, wC0 = 0, wC1 = 0;
        for (int b = 0; b < p->nbasis; ++b) {
            const BasisPlan &bp = p->bp[b];
            LevelDesc &d = a.d[b];
            d.F = bp.F;
            if (bp.F == 0) {
                if (l == 0) {   // tiles of 64x64 pixels
                    maxC0 = std::max(maxC0, (p->nx + 1) / 2);
                    maxC1 = std::max(maxC1, (p->ny + 1) / 2);
                }
                continue;
            }
            d.n0 = l == 0 ? p->nx : bp.sx[l - 1];
            d.n1 = l == 0 ? p->ny : bp.sy[l - 1];
            d.C0 = bp.sx[l];
            d.C1 = bp.sy[l];
            d.lowx = bp.highx[l] - 2 * bp.sx[l];
            d.lowy = bp.highy[l] - 2 * bp.sy[l];
            d.ld_in = round4(d.n1);      // scratch rows padded to 16 bytes (fp32)
            d.ld_out = round4(d.C1);
            d.ldn = p->cn_ld[b][l];
            d.qs = d.C0 * d.ldn;
            d.off1 = p->cn_off1[b][l];
            maxC0 = std::max(maxC0, d.C0);
            maxC1 = std::max(maxC1, d.C1);
            wC0 = std::max(wC0, d.C0);
            wC1 = std::max(wC1, d.C1);
        }
        if (maxC0 == 0) continue;
        const T *in = (l % 2 == 1) ? A : B;   // level l reads what level l-1 wrote
        T *out = (l % 2 == 0) ? A : B;
        if (stream) {
            if (l == 0 && nat) {
                // finest level: every bank (and 'self') in one walk over the image
                DwtAllArgs g;
                memset(&g, 0, sizeof g);
                for (int s = 0; s < 4; ++s) g.s[s].basis = -1;
                g.self_basis = -1;
                g.ld_self = round32(p->ny);
                for (int b = 0; b < p->nbasis; ++b) {
                    if (p->bp[b].F == 0) { g.self_basis = b; continue; }
                    DwtAllSlot &d = g.s[p->bp[b].F / 2 - 1];
                    d.basis = b;
                    d.C0 = a.d[b].C0; d.C1 = a.d[b].C1;
                    d.ldn = a.d[b].ldn; d.qs = a.d[b].qs; d.ld_out = a.d[b].ld_out;
                    d.off1 = a.d[b].off1;
                }
                int SW = getenv("PFB_DWT_SW") ? atoi(getenv("PFB_DWT_SW")) : kStripNat;   // experiments only
                int NJ = pick_nj(wC0, (long)p->nband * cdiv(cdiv(wC1, SW), kStreamWarps));
                int dbg = getenv("PFB_DWT_DBG") ? atoi(getenv("PFB_DWT_DBG")) : 0;
                int nstrip = cdiv(wC1, SW), nchunk = cdiv(wC0, NJ);
                size_t sm2 = sizeof(T) * stream_ring_elems() * kStreamWarps;
                launch(&k_dwt_stream_all<T>, dim3((unsigned)((long)p->nband * nchunk * cdiv(nstrip, kStreamWarps))),
                       dim3(32 * kStreamWarps), sm2, st, x, out, alpha, a, g, nstrip, nchunk, NJ, SW, dbg);
                PFB_CHECK_LAUNCH();
                continue;
            }
            if (l == 0) {
                for (int b = 0; b < p->nbasis; ++b)
                    if (p->bp[b].F == 0) {
                        dim3 grid(cdiv(p->ny, 64), cdiv(p->nx, 64), p->nband);
                        launch(&k_self_dot<T>, grid, dim3(256), sizeof(T) * 64 * 65, st, x, alpha, a, b);
                        PFB_CHECK_LAUNCH();
                    }
            }
            if (wC0 == 0) continue;
            const int SW = nat ? kStripNat : kStripOut;
            int NJ = pick_nj(wC0, (long)p->nbasis * p->nband * cdiv(cdiv(wC1, SW), kStreamWarps));
            int nstrip = cdiv(wC1, SW), nchunk = cdiv(wC0, NJ);
            long ntask = (long)p->nband * nstrip * nchunk;      // per basis
            int dbg = getenv("PFB_DWT_DBG") ? atoi(getenv("PFB_DWT_DBG")) : 0;   // experiments only
            dim3 sgrid((unsigned)((long)p->nbasis * p->nband * nchunk * cdiv(nstrip, kStreamWarps)));
            if (nat) {
                size_t sm2 = sizeof(T) * stream_ring_elems() * kStreamWarps;
                launch(&k_dwt_stream<T, true>, sgrid, dim3(32 * kStreamWarps), sm2, st, x, in, out, alpha, a,
                       nstrip, nchunk, NJ, ntask, dbg, SW);
            } else {
                size_t sm2 = sizeof(T) * (stream_smem_warp(a.last ? 4 : 3) + stream_ring_elems()) * kStreamWarps;
                PFB_CUDA_OK(cudaFuncSetAttribute(&k_dwt_stream<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
                launch(&k_dwt_stream<T, false>, sgrid, dim3(32 * kStreamWarps), sm2, st, x, in, out, alpha, a,
                       nstrip, nchunk, NJ, ntask, dbg, SW);
            }
            PFB_CHECK_LAUNCH();
            continue;
        }
        a.ntile0 = cdiv(maxC0, TX);
        a.ntile1 = cdiv(maxC1, TY);
        dim3 grid(a.ntile1 * p->nbasis, a.ntile0, p->nband);
        launch(&k_dwt_level<T>, grid, dim3(256), smem, st, x, in, out, alpha, a);
        PFB_CHECK_LAUNCH();
    }
    return PFB_OK;
}

static int ensure_cn(const pfb_psi_plan *cp) {
    pfb_psi_plan *p = const_cast<pfb_psi_plan *>(cp);
    if (p->cn) return PFB_OK;
    size_t bytes = dtype_size(p->dtype) * (size_t)p->cn_stride * p->nband * p->nbasis;
    if (cudaMalloc(&p->cn, bytes) != cudaSuccess) {
        p->cn = nullptr;
        return fail(PFB_ERR_ALLOC, "psi plan coefficient scratch allocation failed (%zu bytes)", bytes);
    }
    return PFB_OK;
}

// slab <-> natural orientation for `nband` (band) cubes: 'self' on its own grid,
// then one launch per level.  to_nat fuses 2*alpha - alpha2 when alpha2 != NULL.
template <class T>
static int nat_transposes(const pfb_psi_plan *p, int nband, bool to_nat, const T *alpha,
                          const T *alpha2, T *alpha_out, T *cn, cudaStream_t st) {
    const int L = p->nlevel;
    for (int pass = -1; pass < L; ++pass) {
        NatArgs na;
        memset(&na, 0, sizeof na);
        na.nbasis = p->nbasis; na.Nxmax = p->Nxmax; na.Nymax = p->Nymax; na.cn_stride = p->cn_stride;
        int mr = 0, mc = 0;
        for (int b = 0; b < p->nbasis; ++b) {
            const BasisPlan &bp = p->bp[b];
            if (pass < 0) {
                if (bp.F != 0) continue;
                NatBlk &k = na.b[na.nblk++];
                k.basis = b; k.row0 = k.col0 = 0; k.nrow = p->ny; k.ncol = p->nx;
                k.ldn = round32(p->ny); k.off = 0;
                mr = std::max(mr, k.nrow); mc = std::max(mc, k.ncol);
                continue;
            }
            if (bp.F == 0) continue;
            const int l = pass, C0 = bp.sx[l], C1 = bp.sy[l], ldn = p->cn_ld[b][l];
            const int lowx = bp.highx[l] - 2 * C0, lowy = bp.highy[l] - 2 * C1;
            const long qsz = (long)C0 * ldn;
            for (int q = (l == L - 1 ? 0 : 1); q < 4; ++q) {
                NatBlk &k = na.b[na.nblk++];
                k.basis = b;
                k.row0 = lowy + (q >> 1) * C1; k.col0 = lowx + (q & 1) * C0;
                k.nrow = C1; k.ncol = C0; k.ldn = ldn;
                k.off = p->cn_off1[b][l] + (q - 1) * qsz;
            }
            mr = std::max(mr, C1); mc = std::max(mc, C0);
        }
        if (na.nblk == 0) continue;
        dim3 grid(cdiv(mc, 32), cdiv(mr, 32), nband * na.nblk);
        if (to_nat) launch(&k_slab_to_nat<T>, grid, dim3(256), 0, st, alpha, alpha2, cn, na);
        else launch(&k_nat_to_slab<T>, grid, dim3(256), 0, st, (const T *)cn, alpha_out, na);
        PFB_CHECK_LAUNCH();
    }
    return PFB_OK;
}

// Streaming synthesis from the natural-orientation layout (optionally of
// 2*cn - cn2): one k_idwt_stream launch per level (wavelet_syn_stream.cuh).
template <class T>
static int psi_hdot_nat_t(const pfb_psi_plan *p, const T *cn, const T *cn2, T *xo, cudaStream_t st) {
    if (int rc = ensure_scratch(p)) return rc;
    T *A = (T *)p->scrA, *B = (T *)p->scrB;
    const int L = p->nlevel;
    for (int l = L - 1; l >= 0; --l) {
        SynArgs a;
        memset(&a, 0, sizeof a);
        a.nbasis = p->nbasis; a.nband = p->nband;
        a.first = l == 0; a.last = l == L - 1;
        a.n0 = p->nx; a.n1 = p->ny; a.ld_out = p->ny;
        a.self_basis = -1; a.ld_self = round32(p->ny);
        a.cn_stride = p->cn_stride; a.scr_stride = p->scr_stride;
        a.img_band_stride = (long)p->nx * p->ny;
        for (int s = 0; s < kSynSlots; ++s) a.s[s].basis = -1;
        int mn0 = 0, mn1 = 0;
        for (int b = 0; b < p->nbasis; ++b) {
            const BasisPlan &bp = p->bp[b];
            if (bp.F == 0) { if (l == 0) a.self_basis = b; continue; }
            SynSlot &d = a.s[bp.F / 2 - 1];
            d.basis = b;
            d.C0 = bp.sx[l]; d.C1 = bp.sy[l];
            d.ldn = p->cn_ld[b][l];
            d.ld_ll = l + 1 < L ? round4(bp.spy[l + 1]) : 0;
            d.n0 = bp.spx[l]; d.n1 = bp.spy[l];
            d.ld_out = round4(bp.spy[l]);
            d.qs = d.C0 * d.ldn;
            d.off1 = p->cn_off1[b][l];
            mn0 = std::max(mn0, d.n0); mn1 = std::max(mn1, d.n1);
        }
        if (l == 0) { mn0 = p->nx; mn1 = p->ny; }
        const int h0 = mn0 / 2, h1 = mn1 / 2;
        const int per_basis = l > 0;
        const int NJ = pick_nj(h0, (long)(per_basis ? kSynSlots : 1) * p->nband * cdiv(cdiv(h1, kSynStrip), kSynWarps));
        const int nstrip = cdiv(h1, kSynStrip), nchunk = cdiv(h0, NJ);
        const long nblk = (long)(per_basis ? kSynSlots : 1) * p->nband * nchunk * cdiv(nstrip, kSynWarps);
        const T *in = ((L - 1 - l) % 2 == 1) ? A : B;
        T *out = l == 0 ? xo : (((L - 1 - l) % 2 == 0) ? A : B);
        const size_t sm1 = sizeof(T) * kSynWarps * kSynPD * syn_stage_elems(false);
        const size_t sm2 = sizeof(T) * kSynWarps * kSynPD * syn_stage_elems(true);
        if (cn2) {
            PFB_CUDA_OK(cudaFuncSetAttribute(&k_idwt_stream<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
            launch(&k_idwt_stream<T, true>, dim3((unsigned)nblk), dim3(32 * kSynWarps), sm2, st,
                   cn, cn2, in, out, a, nstrip, nchunk, NJ, per_basis);
        } else {
            PFB_CUDA_OK(cudaFuncSetAttribute(&k_idwt_stream<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
            launch(&k_idwt_stream<T, false>, dim3((unsigned)nblk), dim3(32 * kSynWarps), sm1, st,
                   cn, cn, in, out, a, nstrip, nchunk, NJ, per_basis);
        }
        PFB_CHECK_LAUNCH();
    }
    return PFB_OK;
}

// Reference-layout entry: slab -> plan scratch (fusing 2*alpha - alpha2), then
// the streaming synthesis.
template <class T>
static int psi_hdot_stream(const pfb_psi_plan *p, const T *alpha, const T *alpha2, T *xo,
                           cudaStream_t st) {
    if (int rc = ensure_cn(p)) return rc;
    T *cn = (T *)p->cn;
    if (int rc = nat_transposes<T>(p, p->nband, true, alpha, alpha2, nullptr, cn, st)) return rc;
    return psi_hdot_nat_t<T>(p, cn, (const T *)nullptr, xo, st);
}

template <class T>
static int psi_hdot_t(const pfb_psi_plan *p, const T *alpha, const T *alpha2, T *xo,
                      cudaStream_t st) {
    if (int rc = ensure_scratch(p)) return rc;
    if (p->syn_ok && ((uintptr_t)xo % (2 * sizeof(T))) == 0 && !getenv("PFB_IDWT_TILE"))
        return psi_hdot_stream<T>(p, alpha, alpha2, xo, st);
    size_t smem = sizeof(T) * kSynSmemElems;
    PFB_CUDA_OK(cudaFuncSetAttribute(&k_idwt_level<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    T *A = (T *)p->scrA, *B = (T *)p->scrB;
    for (int l = p->nlevel - 1; l >= 0; --l) {
        LevelArgs a;
        memset(&a, 0, sizeof a);
        fill_common(p, a, l);
        int maxn0 = 0, maxn1 = 0;
        for (int b = 0; b < p->nbasis; ++b) {
            const BasisPlan &bp = p->bp[b];
            LevelDesc &d = a.d[b];
            d.F = bp.F;
            if (bp.F == 0) continue;
            d.n0 = bp.spx[l];
            d.n1 = bp.spy[l];
            d.C0 = bp.sx[l];
            d.C1 = bp.sy[l];
            d.lowx = bp.highx[l] - 2 * bp.sx[l];
            d.lowy = bp.highy[l] - 2 * bp.sy[l];
            d.ld_in = l + 1 < p->nlevel ? bp.spy[l + 1] : 0;
            d.ld_out = bp.spy[l];
            maxn0 = std::max(maxn0, d.n0);
            maxn1 = std::max(maxn1, d.n1);
        }
        if (l == 0) { maxn0 = p->nx; maxn1 = p->ny; }
        if (maxn0 == 0) continue;
        a.ntile0 = cdiv(maxn0, 2 * TX);
        a.ntile1 = cdiv(maxn1, 2 * TY);
        int per = l == 0 ? p->nbasis : 1;
        dim3 grid(a.ntile1, a.ntile0, p->nband * (p->nbasis / per));
        // level l reads the scratch written by level l+1
        const T *in = ((p->nlevel - 1 - l) % 2 == 1) ? A : B;
        T *out = l == 0 ? xo : (((p->nlevel - 1 - l) % 2 == 0) ? A : B);
        launch(&k_idwt_level<T>, grid, dim3(256), smem, st, alpha, alpha2, in, out, a, per);
        PFB_CHECK_LAUNCH();
    }
    return PFB_OK;
}

}  // namespace pfb

using namespace pfb;

extern "C" {

int pfb_psi_plan_create(pfb_psi_plan **out, int nband, int nx, int ny, int nbasis,
                        const int *basis_ids, int nlevel, int dtype) {
    if (!out || !basis_ids || nband < 1 || nx < 1 || ny < 1 || nbasis < 1 ||
        nbasis > kMaxBasis || nlevel < 1 || nlevel > kMaxLevel ||
        (dtype != PFB_F32 && dtype != PFB_F64))
        return fail(PFB_ERR_BAD_ARG, "pfb_psi_plan_create: bad argument");
    if ((nx & 1) || (ny & 1))
        return fail(PFB_ERR_ODD_SIZE, "image size (%d,%d) must be even", nx, ny);
    pfb_psi_plan *p = new pfb_psi_plan();
    memset(p, 0, sizeof *p);
    p->nband = nband; p->nx = nx; p->ny = ny; p->nbasis = nbasis;
    p->nlevel = nlevel; p->dtype = dtype;
    long scr = 0;
    for (int b = 0; b < nbasis; ++b) {
        int id = basis_ids[b];
        if (id < 0 || id > 5) { delete p; return fail(PFB_ERR_BAD_ARG, "basis id %d not in 0..5", id); }
        p->basis_id[b] = id;
        BasisPlan &bp = p->bp[b];
        bp.F = 2 * id;
        if (id == 0) continue;
        int F = bp.F;
        if (nlevel > dwt_max_level(std::min(nx, ny), F)) {
            delete p;
            return fail(PFB_ERR_NLEVEL, "The requested decomposition level %d is not possible", nlevel);
        }
        int Nx = nx, Ny = ny, Nxm = 0, Nym = 0;
        for (int k = 0; k < nlevel; ++k) {
            int Cx = coeff_size(Nx, F), Cy = coeff_size(Ny, F);
            Nxm += Cx; Nym += Cy;
            bp.sx[k] = Cx; bp.sy[k] = Cy;
            Nx = Cx + Cx % 2; Ny = Cy + Cy % 2;
            bp.spx[k] = signal_size(Cx, F);
            bp.spy[k] = signal_size(Cy, F);
        }
        Nxm += bp.sx[nlevel - 1]; Nym += bp.sy[nlevel - 1];
        bp.Ntotx = Nxm; bp.Ntoty = Nym;
        p->Nxmax = std::max(p->Nxmax, Nxm);
        p->Nymax = std::max(p->Nymax, Nym);
        int lowx = 2 * bp.sx[nlevel - 1], lowy = 2 * bp.sy[nlevel - 1];
        bp.highx[nlevel - 1] = lowx; bp.highy[nlevel - 1] = lowy;
        for (int k = nlevel - 2; k >= 0; --k) {
            bp.highx[k] = lowx + bp.sx[k]; bp.highy[k] = lowy + bp.sy[k];
            lowx += bp.sx[k]; lowy += bp.sy[k];
        }
        scr = std::max(scr, (((long)(bp.sx[0] + 2) * (bp.sy[0] + 6)) + 3) & ~3L);
    }
    if (p->Nxmax == 0) {   // bases == ['self'] only: the reference cannot run this either
        delete p;
        return fail(PFB_ERR_BAD_ARG, "at least one wavelet basis is required");
    }
    p->scr_stride = scr;   // scratch itself is allocated on first use
    // natural-orientation layout for the streaming synthesis
    p->syn_ok = 1;
    int seen[6] = {0, 0, 0, 0, 0, 0};
    for (int b = 0; b < nbasis; ++b) {
        int id = p->basis_id[b];
        if (id == 5 || seen[id]++) p->syn_ok = 0;
        long tot = 0;
        if (id == 0) {
            tot = (long)nx * round32(ny);
        } else {
            const BasisPlan &bp = p->bp[b];
            for (int l = 0; l < nlevel; ++l) {
                int ldn = round32(bp.sy[l]);
                long qsz = (long)bp.sx[l] * ldn;
                if (qsz > (1L << 30)) p->syn_ok = 0;
                if (l == nlevel - 1) tot += qsz;      // LL in front of quadrant 1
                p->cn_off1[b][l] = tot;
                p->cn_ld[b][l] = ldn;
                tot += 3 * qsz;
            }
        }
        p->cn_stride = std::max(p->cn_stride, tot);
    }
    *out = p;
    return PFB_OK;
}

int pfb_psi_plan_destroy(pfb_psi_plan *p) {
    if (!p) return PFB_OK;
    if (p->scrA) cudaFree(p->scrA);
    if (p->scrB) cudaFree(p->scrB);
    if (p->cn) cudaFree(p->cn);
    delete p;
    return PFB_OK;
}

int pfb_filter_bank(int order, double *taps) {
    static const double h_taps[5][4][10] = PFB_TAPS_INIT;
    if (order < 1 || order > 5 || !taps) return fail(PFB_ERR_BAD_ARG, "pfb_filter_bank: order must be 1..5");
    for (int w = 0; w < 4; ++w)
        for (int k = 0; k < 10; ++k) taps[w * 10 + k] = h_taps[order - 1][w][k];
    return PFB_OK;
}

int pfb_psi_plan_query(const pfb_psi_plan *p, int *Nxmax, int *Nymax) {
    if (!p) return fail(PFB_ERR_BAD_ARG, "null plan");
    if (Nxmax) *Nxmax = p->Nxmax;
    if (Nymax) *Nymax = p->Nymax;
    return PFB_OK;
}

int pfb_psi_plan_tables(const pfb_psi_plan *p, int basis, int *tabs, int *Ntotx, int *Ntoty) {
    if (!p || basis < 0 || basis >= p->nbasis || !tabs) return fail(PFB_ERR_BAD_ARG, "bad argument");
    const BasisPlan &bp = p->bp[basis];
    int L = p->nlevel;
    for (int l = 0; l < L; ++l) {
        tabs[0 * L + l] = bp.highx[l]; tabs[1 * L + l] = bp.highy[l];
        tabs[2 * L + l] = bp.sx[l];    tabs[3 * L + l] = bp.sy[l];
        tabs[4 * L + l] = bp.spx[l];   tabs[5 * L + l] = bp.spy[l];
    }
    if (Ntotx) *Ntotx = bp.Ntotx;
    if (Ntoty) *Ntoty = bp.Ntoty;
    return PFB_OK;
}

int pfb_psi_dot(const pfb_psi_plan *p, const void *x, void *alpha, void *stream) {
    if (!p || !x || !alpha) return fail(PFB_ERR_BAD_ARG, "pfb_psi_dot: null argument");
    if (p->dtype == PFB_F32) return psi_dot_t<float>(p, (const float *)x, (float *)alpha, false, as_stream(stream));
    return psi_dot_t<double>(p, (const double *)x, (double *)alpha, false, as_stream(stream));
}

int pfb_psi_nat_stride(const pfb_psi_plan *p, long long *stride) {
    if (!p || !stride) return fail(PFB_ERR_BAD_ARG, "pfb_psi_nat_stride: null argument");
    *stride = p->syn_ok ? p->cn_stride : 0;
    return PFB_OK;
}

static int nat_check(const pfb_psi_plan *p, const void *a, const void *b, int nband, const char *who) {
    if (!p || !a || !b || nband < 1) return fail(PFB_ERR_BAD_ARG, "%s: bad argument", who);
    if (!p->syn_ok) return fail(PFB_ERR_BAD_ARG, "%s: natural-orientation layout needs distinct db1..db4 (+self) bases", who);
    return PFB_OK;
}

int pfb_psi_to_nat(const pfb_psi_plan *p, int nband, const void *alpha, void *nat, void *stream) {
    if (int rc = nat_check(p, alpha, nat, nband, "pfb_psi_to_nat")) return rc;
    cudaStream_t st = as_stream(stream);
    size_t bytes = dtype_size(p->dtype) * (size_t)p->cn_stride * nband * p->nbasis;
    PFB_CUDA_OK(cudaMemsetAsync(nat, 0, bytes, st));        // row / block padding stays zero
    if (p->dtype == PFB_F32)
        return nat_transposes<float>(p, nband, true, (const float *)alpha, nullptr, nullptr, (float *)nat, st);
    return nat_transposes<double>(p, nband, true, (const double *)alpha, nullptr, nullptr, (double *)nat, st);
}

int pfb_psi_from_nat(const pfb_psi_plan *p, int nband, const void *nat, void *alpha, void *stream) {
    if (int rc = nat_check(p, nat, alpha, nband, "pfb_psi_from_nat")) return rc;
    cudaStream_t st = as_stream(stream);
    if (p->dtype == PFB_F32)
        return nat_transposes<float>(p, nband, false, nullptr, nullptr, (float *)alpha, (float *)nat, st);
    return nat_transposes<double>(p, nband, false, nullptr, nullptr, (double *)alpha, (double *)nat, st);
}

int pfb_psi_dot_nat(const pfb_psi_plan *p, const void *x, void *nat, void *stream) {
    if (int rc = nat_check(p, x, nat, 1, "pfb_psi_dot_nat")) return rc;
    if (p->dtype == PFB_F32) return psi_dot_t<float>(p, (const float *)x, (float *)nat, true, as_stream(stream));
    return psi_dot_t<double>(p, (const double *)x, (double *)nat, true, as_stream(stream));
}

int pfb_psi_dot_nat_dual(const pfb_psi_plan *p, const void *x, const void *vp_nat, const void *f_nat,
                         double sigma, void *v_nat, void *stream) {
    if (int rc = nat_check(p, x, v_nat, 1, "pfb_psi_dot_nat_dual")) return rc;
    if (!vp_nat || !f_nat || vp_nat == v_nat)
        return fail(PFB_ERR_BAD_ARG, "pfb_psi_dot_nat_dual: vp and f must be given and vp must not alias the output");
    if (p->dtype == PFB_F32)
        return psi_dot_t<float>(p, (const float *)x, (float *)v_nat, true, as_stream(stream), (const float *)vp_nat,
                                (const float *)f_nat, sigma);
    return psi_dot_t<double>(p, (const double *)x, (double *)v_nat, true, as_stream(stream), (const double *)vp_nat,
                             (const double *)f_nat, sigma);
}

int pfb_psi_hdot_nat(const pfb_psi_plan *p, const void *nat, const void *nat2, void *xo, void *stream) {
    if (int rc = nat_check(p, nat, xo, 1, "pfb_psi_hdot_nat")) return rc;
    if ((uintptr_t)xo % (2 * dtype_size(p->dtype)) || (uintptr_t)nat % 16 || (uintptr_t)nat2 % 16)
        return fail(PFB_ERR_BAD_ARG, "pfb_psi_hdot_nat: coefficients must be 16-byte aligned, the image to two elements");
    if (p->dtype == PFB_F32)
        return psi_hdot_nat_t<float>(p, (const float *)nat, (const float *)nat2, (float *)xo, as_stream(stream));
    return psi_hdot_nat_t<double>(p, (const double *)nat, (const double *)nat2, (double *)xo, as_stream(stream));
}

int pfb_psi_hdot(const pfb_psi_plan *p, const void *alpha, const void *alpha2, void *xo,
                 void *stream) {
    if (!p || !alpha || !xo) return fail(PFB_ERR_BAD_ARG, "pfb_psi_hdot: null argument");
    if (p->dtype == PFB_F32) return psi_hdot_t<float>(p, (const float *)alpha, (const float *)alpha2, (float *)xo, as_stream(stream));
    return psi_hdot_t<double>(p, (const double *)alpha, (const double *)alpha2, (double *)xo, as_stream(stream));
}

}  // extern "C"
