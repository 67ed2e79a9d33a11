// Device-resident preconditioned conjugate gradients (pfb/opt/pcg.py:86-122 and the
// per-band form of _pcg_psf_impl, pfb/opt/pcg.py:243-291).
//
// The line-search scalars never visit the host: every kernel works on `nseg` independent
// segments of `seglen` elements (one segment = the whole cube for pcg, one band for
// pcg_psf) and reads / writes a per-segment state block of PFB_PCG_NSTATE doubles
//   [0] <r,y>  [1] <p,Ap>  [2] alpha  [3] <r,y>_next  [4] sum (x-xp)^2  [5] sum x^2
//   [6] beta   [7] count of non-zero p  [8] backtrack flag  [9] active
// in device memory.  Per iteration the host enqueues
//   A(p) -> pfb_pcg_seg_dot -> pfb_pcg_seg_scalar(ALPHA) -> pfb_pcg_seg_update ->
//   pfb_pcg_seg_scalar(BETA) -> pfb_pcg_seg_direction
// and reads ONE copy of the state (for eps, termination and the rare backtrack) while the
// next A(p) is already running.  Sums accumulate in fp64 in a fixed order; between a sum and
// the scalar step that consumes it the caller may all-reduce the state (band-sharded ranks).
#include "common.cuh"
#include <initializer_list>

namespace pfb {

#ifdef PFB_EMU
static const int kPcgBlocks = 3;
#else
static const int kPcgBlocks = kNumSM * 4;     // CTAs over all segments
#endif
static const int NS = PFB_PCG_NSTATE;

__device__ __forceinline__ double p_warp_sum(double x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    return x;
}

// block-reduce K values; thread 0 stores them to part[((seg * nblk) + blk) * K + k]
template <int K>
__device__ __forceinline__ void p_block_store(double (&acc)[K], double *part) {
    __shared__ double sh[4][32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double s = p_warp_sum(acc[k]);
        if (lane == 0) sh[k][wid] = s;
    }
    __syncthreads();
    if (wid == 0) {
        const int nw = (blockDim.x + 31) >> 5;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double s = lane < nw ? sh[k][lane] : 0.0;
            s = p_warp_sum(s);
            if (lane == 0) part[((long)blockIdx.y * gridDim.x + blockIdx.x) * K + k] = s;
        }
    }
}

// st[seg][slot0 + k] = sum over the segment's CTAs of part[...][k], fixed order (one warp per segment)
template <int K>
__global__ void k_pcg_final(const double *__restrict__ part, int nblk, double *st, int slot0) {
    const int seg = blockIdx.x;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double a = 0.0;
        for (int c = threadIdx.x; c < nblk; c += 32) a += part[((long)seg * nblk + c) * K + k];
        a = p_warp_sum(a);
        if (threadIdx.x == 0) st[(long)seg * NS + slot0 + k] = a;
    }
}

template <class T, int V> struct alignas(sizeof(T) * V) PPack { T v[V]; };
template <class T, int V> __device__ __forceinline__ PPack<T, V> pld(const T *p) {
    return *reinterpret_cast<const PPack<T, V> *>(p);
}
template <class T, int V> __device__ __forceinline__ void pst(T *p, const PPack<T, V> &x) {
    *reinterpret_cast<PPack<T, V> *>(p) = x;
}

// <a, b> per segment.  V = elements per 128-bit access (1 when sizes / pointers are unaligned).
template <class T, int V>
__global__ void __launch_bounds__(256)
k_pcg_seg_dot(const T *__restrict__ a, const T *__restrict__ b, long seglen, double *part) {
    const long o = (long)blockIdx.y * seglen;
    double acc[1] = {0.0};
    const long nv = seglen / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
        const PPack<T, V> x = pld<T, V>(a + o + i * V), y = pld<T, V>(b + o + i * V);
#pragma unroll
        for (int e = 0; e < V; ++e) acc[0] += (double)x.v[e] * (double)y.v[e];
    }
    p_block_store<1>(acc, part);
}

// x = xp + alpha p; r = rp + alpha Ap; y = minv r  (pfb/opt/pcg.py:92-95); sums {<r,y>, (x-xp)^2, x^2}.
// Inactive segments are carried over unchanged (x = xp, r = rp).
template <class T, int V>
__global__ void __launch_bounds__(256)
k_pcg_seg_update(T *__restrict__ x, T *__restrict__ r, const T *__restrict__ xp, const T *__restrict__ rp,
                 const T *__restrict__ p, const T *__restrict__ Ap, const double *__restrict__ st, T minv,
                 long seglen, double *part) {
    const int seg = blockIdx.y;
    const long o = (long)seg * seglen;
    const bool active = st[(long)seg * NS + 9] != 0.0;
    const T alpha = (T)st[(long)seg * NS + 2];
    double acc[3] = {0.0, 0.0, 0.0};
    const long nv = seglen / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
        const long k = o + i * V;
        const PPack<T, V> xo = pld<T, V>(xp + k), ro = pld<T, V>(rp + k);
        PPack<T, V> xi = xo, ri = ro;
        if (active) {
            const PPack<T, V> pp = pld<T, V>(p + k), ap = pld<T, V>(Ap + k);
#pragma unroll
            for (int e = 0; e < V; ++e) {
                xi.v[e] = xo.v[e] + alpha * pp.v[e];
                ri.v[e] = ro.v[e] + alpha * ap.v[e];
            }
        }
        pst<T, V>(x + k, xi);
        pst<T, V>(r + k, ri);
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const T yi = ri.v[e] * minv;
            const T d = xi.v[e] - xo.v[e];
            acc[0] += (double)ri.v[e] * (double)yi;
            acc[1] += (double)(d * d);
            acc[2] += (double)(xi.v[e] * xi.v[e]);
        }
    }
    p_block_store<3>(acc, part);
}

// pn = beta p - minv r; count of non-zero pn  (pfb/opt/pcg.py:103-107).  Inactive: pn = p.
template <class T, int V>
__global__ void __launch_bounds__(256)
k_pcg_seg_dir(T *__restrict__ pn, const T *__restrict__ p, const T *__restrict__ r,
              const double *__restrict__ st, T minv, long seglen, double *part) {
    const int seg = blockIdx.y;
    const long o = (long)seg * seglen;
    const bool active = st[(long)seg * NS + 9] != 0.0;
    const T beta = (T)st[(long)seg * NS + 6];
    double acc[1] = {0.0};
    const long nv = seglen / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
        const long k = o + i * V;
        PPack<T, V> pi = pld<T, V>(p + k);
        if (active) {
            const PPack<T, V> rr = pld<T, V>(r + k);
#pragma unroll
            for (int e = 0; e < V; ++e) pi.v[e] = beta * pi.v[e] - minv * rr.v[e];
        }
        pst<T, V>(pn + k, pi);
#pragma unroll
        for (int e = 0; e < V; ++e)
            if (pi.v[e] != (T)0) acc[0] += 1.0;
    }
    p_block_store<1>(acc, part);
}

// the scalar arithmetic of the line search, one thread per segment
__global__ void k_pcg_scalar(double *st, int nseg, int step) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nseg) return;
    double *q = st + (long)s * NS;
    if (step == PFB_PCG_ALPHA) {                // alpha = <r,y> / <p,Ap>           (pcg.py:90-91)
        q[2] = q[0] / q[1];
    } else if (step == PFB_PCG_BETA) {          // backtrack test, beta            (pcg.py:96,103)
        q[8] = (q[9] != 0.0 && q[3] > q[0]) ? 1.0 : 0.0;
        q[6] = q[3] / q[0];
    } else if (step == PFB_PCG_SHRINK) {        // alpha *= 0.75 where flagged     (pcg.py:97)
        if (q[8] != 0.0) q[2] *= 0.75;
    } else if (step == PFB_PCG_COMMIT) {        // rnorm = rnorm_next              (pcg.py:108)
        if (q[9] != 0.0) q[0] = q[3];
    }
}

static double *g_ppart = nullptr;
static long g_ppart_n = 0;

static int ensure_ppart(long n) {
    if (g_ppart_n < n) {
        if (g_ppart) cudaFree(g_ppart);
        g_ppart = nullptr;
        g_ppart_n = 0;
        if (cudaMalloc((void **)&g_ppart, sizeof(double) * n) != cudaSuccess)
            return fail(PFB_ERR_ALLOC, "cudaMalloc of the PCG reduction workspace failed");
        g_ppart_n = n;
    }
    return PFB_OK;
}

template <class T> static inline bool vec_ok(long seglen, std::initializer_list<const void *> ptrs) {
    const int V = 16 / (int)sizeof(T);
    if (seglen % V) return false;
    for (const void *q : ptrs)
        if ((uintptr_t)q & 15) return false;
    return true;
}

static inline int seg_blocks(int nseg, long seglen) {
    long per = kPcgBlocks / nseg;
    long need = (seglen / 4 + 255) / 256;
    if (per > need) per = need;
    return (int)(per < 1 ? 1 : per);
}

}  // namespace pfb

using namespace pfb;

extern "C" {

int pfb_pcg_seg_dot(const void *p, const void *Ap, int nseg, long long seglen, int dtype, double *state,
                    int slot, void *stream) {
    if (!p || !Ap || !state || nseg < 1 || seglen < 1 || slot < 0 || slot >= PFB_PCG_NSTATE)
        return fail(PFB_ERR_BAD_ARG, "pfb_pcg_seg_dot: bad argument");
    const int nblk = seg_blocks(nseg, seglen);
    if (int rc = ensure_ppart((long)nseg * nblk * 3)) return rc;
    cudaStream_t st = as_stream(stream);
    const dim3 g(nblk, nseg), b(256);
    if (dtype == PFB_F32) {
        if (vec_ok<float>(seglen, {p, Ap}))
            launch(&k_pcg_seg_dot<float, 4>, g, b, 0, st, (const float *)p, (const float *)Ap, (long)seglen, g_ppart);
        else
            launch(&k_pcg_seg_dot<float, 1>, g, b, 0, st, (const float *)p, (const float *)Ap, (long)seglen, g_ppart);
    } else if (dtype == PFB_F64) {
        if (vec_ok<double>(seglen, {p, Ap}))
            launch(&k_pcg_seg_dot<double, 2>, g, b, 0, st, (const double *)p, (const double *)Ap, (long)seglen, g_ppart);
        else
            launch(&k_pcg_seg_dot<double, 1>, g, b, 0, st, (const double *)p, (const double *)Ap, (long)seglen, g_ppart);
    } else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    launch(&k_pcg_final<1>, dim3(nseg), dim3(32), 0, st, (const double *)g_ppart, nblk, state, slot);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

int pfb_pcg_seg_scalar(double *state, int nseg, int step, void *stream) {
    if (!state || nseg < 1 || step < PFB_PCG_ALPHA || step > PFB_PCG_COMMIT)
        return fail(PFB_ERR_BAD_ARG, "pfb_pcg_seg_scalar: bad argument");
    launch(&k_pcg_scalar, dim3((nseg + 63) / 64), dim3(64), 0, as_stream(stream), state, nseg, step);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

int pfb_pcg_seg_update(void *x, void *r, const void *xp, const void *rp, const void *p, const void *Ap,
                       double minv, int nseg, long long seglen, int dtype, double *state, void *stream) {
    if (!x || !r || !xp || !rp || !p || !Ap || !state || nseg < 1 || seglen < 1)
        return fail(PFB_ERR_BAD_ARG, "pfb_pcg_seg_update: bad argument");
    const int nblk = seg_blocks(nseg, seglen);
    if (int rc = ensure_ppart((long)nseg * nblk * 3)) return rc;
    cudaStream_t st = as_stream(stream);
    const dim3 g(nblk, nseg), b(256);
#define PFB_UPD(T, V, MINV)                                                                                      \
    launch(&k_pcg_seg_update<T, V>, g, b, 0, st, (T *)x, (T *)r, (const T *)xp, (const T *)rp, (const T *)p,      \
           (const T *)Ap, (const double *)state, (T)(MINV), (long)seglen, g_ppart)
    if (dtype == PFB_F32) {
        if (vec_ok<float>(seglen, {x, r, xp, rp, p, Ap})) PFB_UPD(float, 4, minv);
        else PFB_UPD(float, 1, minv);
    } else if (dtype == PFB_F64) {
        if (vec_ok<double>(seglen, {x, r, xp, rp, p, Ap})) PFB_UPD(double, 2, minv);
        else PFB_UPD(double, 1, minv);
    } else
#undef PFB_UPD
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    launch(&k_pcg_final<3>, dim3(nseg), dim3(32), 0, st, (const double *)g_ppart, nblk, state, 3);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

int pfb_pcg_seg_direction(void *pn, const void *p, const void *r, double minv, int nseg, long long seglen,
                          int dtype, double *state, void *stream) {
    if (!pn || !p || !r || !state || nseg < 1 || seglen < 1)
        return fail(PFB_ERR_BAD_ARG, "pfb_pcg_seg_direction: bad argument");
    const int nblk = seg_blocks(nseg, seglen);
    if (int rc = ensure_ppart((long)nseg * nblk * 3)) return rc;
    cudaStream_t st = as_stream(stream);
    const dim3 g(nblk, nseg), b(256);
#define PFB_DIR(T, V, MINV)                                                                                      \
    launch(&k_pcg_seg_dir<T, V>, g, b, 0, st, (T *)pn, (const T *)p, (const T *)r, (const double *)state,         \
           (T)(MINV), (long)seglen, g_ppart)
    if (dtype == PFB_F32) {
        if (vec_ok<float>(seglen, {pn, p, r})) PFB_DIR(float, 4, minv);
        else PFB_DIR(float, 1, minv);
    } else if (dtype == PFB_F64) {
        if (vec_ok<double>(seglen, {pn, p, r})) PFB_DIR(double, 2, minv);
        else PFB_DIR(double, 1, minv);
    } else
#undef PFB_DIR
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    launch(&k_pcg_final<1>, dim3(nseg), dim3(32), 0, st, (const double *)g_ppart, nblk, state, 7);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

}  // extern "C"
