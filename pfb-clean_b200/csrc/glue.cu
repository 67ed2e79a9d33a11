// Worker glue of the spotless major cycle on the device:
// the pieces of pfb/workers/spotless.py that sit between two minor cycles and
// would otherwise bounce the cube through the host.
//
//   pfb_l1reweight : w = (1 + rmsfactor) / (1 + |S|^alpha / rms_b^alpha)
//                    pfb/utils/misc.py:1070-1080 (S = band-summed coefficients)
//   pfb_seg_stats  : per segment mean, population std and max |x| in fp64
//                    (np.std(..., axis=(-1,-2)) of rms_comps, spotless.py:219-223;
//                    np.std / np.abs().max() of residual_mfs, spotless.py:229-230,343-344)
//
// Streaming reductions (two passes for the std, like numpy): grid = multiple of
// the SM count, fp64 accumulation in a fixed order (deterministic).
#include "common.cuh"

namespace pfb {

#ifdef PFB_EMU
static const int kSegBlocks = 2;
#else
static const int kSegBlocks = kNumSM * 2;     // CTAs per segment
#endif

__device__ __forceinline__ double g_warp_sum(double x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    return x;
}
__device__ __forceinline__ double g_warp_max(double x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x = fmax(x, __shfl_down_sync(0xffffffffu, x, o));
    return x;
}

// pass 1 (PASS = 0): part[seg][blk] = {sum x, max |x|}
// pass 2 (PASS = 1): part[seg][blk] = {sum (x - mean_seg)^2, unused}; mean from stats[3 seg]
template <class T, int PASS>
__global__ void __launch_bounds__(256)
k_seg_stats(const T *__restrict__ x, long m, const double *__restrict__ stats, double *part) {
    const int seg = blockIdx.y;
    const T *xs = x + (long)seg * m;
    const double mean = PASS ? stats[3 * seg] : 0.0;
    double a = 0.0, b = 0.0;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long)gridDim.x * blockDim.x) {
        const double v = (double)xs[i];
        if (PASS == 0) { a += v; b = fmax(b, fabs(v)); }
        else { const double d = v - mean; a += d * d; }
    }
    __shared__ double sh[2][32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    a = g_warp_sum(a);
    b = g_warp_max(b);
    if (lane == 0) { sh[0][wid] = a; sh[1][wid] = b; }
    __syncthreads();
    if (wid == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        a = lane < nw ? sh[0][lane] : 0.0;
        b = lane < nw ? sh[1][lane] : 0.0;
        a = g_warp_sum(a);
        b = g_warp_max(b);
        if (lane == 0) {
            part[2 * ((long)seg * gridDim.x + blockIdx.x)] = a;
            part[2 * ((long)seg * gridDim.x + blockIdx.x) + 1] = b;
        }
    }
}

// one warp per segment, fixed order.  PASS 0: stats[3s] = mean, stats[3s+2] = absmax;
// PASS 1: stats[3s+1] = sqrt(sum / m)
template <int PASS>
__global__ void k_seg_final(const double *__restrict__ part, int nblk, long m, double *stats) {
    const int seg = blockIdx.x;
    double a = 0.0, b = 0.0;
    for (int c = threadIdx.x; c < nblk; c += 32) {
        a += part[2 * ((long)seg * nblk + c)];
        b = fmax(b, part[2 * ((long)seg * nblk + c) + 1]);
    }
    a = g_warp_sum(a);
    b = g_warp_max(b);
    if (threadIdx.x == 0) {
        if (PASS == 0) { stats[3 * seg] = a / (double)m; stats[3 * seg + 2] = b; }
        else stats[3 * seg + 1] = sqrt(a / (double)m);
    }
}

__device__ __forceinline__ float tpow(float a, float b) { return powf(a, b); }
__device__ __forceinline__ double tpow(double a, double b) { return pow(a, b); }

// w[b][i] = (1 + rmsfactor) / (1 + |s[b][i]|^alpha / rms[b]^alpha); arithmetic in the array
// precision like numpy (`mcomps**alpha / rms_comps**alpha` on fp32 arrays stays fp32)
template <class T>
__global__ void __launch_bounds__(256)
k_l1reweight(const T *__restrict__ s, const T *__restrict__ rms, long nslab, T rmsfactor, T alpha,
             T *__restrict__ w) {
    const int basis = blockIdx.y;
    const T ra = tpow(rms[basis], alpha);
    const T *sb = s + (long)basis * nslab;
    T *wb = w + (long)basis * nslab;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nslab; i += (long)gridDim.x * blockDim.x) {
        const T mc = fabs(sb[i]);
        wb[i] = ((T)1 + rmsfactor) / ((T)1 + tpow(mc, alpha) / ra);
    }
}

// out[i] += scale * x[i] * y[i]  (y may be null: out += scale * x) -- the per-dataset accumulation
// of dds2cubes, pfb/utils/misc.py:698-714 (dirty[b] += DIRTY * BEAM, mean_beam[b] += BEAM * wsum, ...)
template <class T>
__global__ void __launch_bounds__(256)
k_accum_mul(T *__restrict__ out, const T *__restrict__ x, const T *__restrict__ y, T scale, long n) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        T v = x[i];
        if (y) v = v * y[i];
        out[i] = out[i] + scale * v;
    }
}

static double *g_spart = nullptr;
static long g_spart_n = 0;

static int ensure_spart(long n) {
    if (g_spart_n < n) {
        if (g_spart) cudaFree(g_spart);
        g_spart = nullptr;
        g_spart_n = 0;
        if (cudaMalloc((void **)&g_spart, sizeof(double) * n) != cudaSuccess)
            return fail(PFB_ERR_ALLOC, "cudaMalloc of the statistics workspace failed");
        g_spart_n = n;
    }
    return PFB_OK;
}

template <class T>
static int seg_stats_t(const T *x, int nseg, long m, double *stats, cudaStream_t st) {
    long nb = (m + 255) / 256;
    const int blk = (int)(nb < kSegBlocks ? (nb < 1 ? 1 : nb) : kSegBlocks);
    if (int rc = ensure_spart(2L * nseg * blk)) return rc;
    launch(&k_seg_stats<T, 0>, dim3(blk, nseg), dim3(256), 0, st, x, m, (const double *)stats, g_spart);
    PFB_CHECK_LAUNCH();
    launch(&k_seg_final<0>, dim3(nseg), dim3(32), 0, st, (const double *)g_spart, blk, m, stats);
    PFB_CHECK_LAUNCH();
    launch(&k_seg_stats<T, 1>, dim3(blk, nseg), dim3(256), 0, st, x, m, (const double *)stats, g_spart);
    PFB_CHECK_LAUNCH();
    launch(&k_seg_final<1>, dim3(nseg), dim3(32), 0, st, (const double *)g_spart, blk, m, stats);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int l1reweight_t(const T *s, const T *rms, int nbasis, long nslab, double rmsfactor, double alpha,
                        T *w, cudaStream_t st) {
    long nb = (nslab + 255) / 256;
    const int blk = (int)(nb < kNumSM * 8 ? (nb < 1 ? 1 : nb) : kNumSM * 8);
    launch(&k_l1reweight<T>, dim3(blk, nbasis), dim3(256), 0, st, s, rms, nslab, (T)rmsfactor, (T)alpha, w);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

// ---------------------------------------------------------------------------
// No-allocation variant of pfb_seg_stats (for callers that cannot call cudaMalloc: CUDA graph
// capture, tight serving loops).  The partial sums live in a file-scope device array instead
// of the cudaMalloc'ed workspace: 2 doubles x kSegBlocks CTAs x kStaticMaxSeg segments
// = 303 104 bytes, held by every GPU that loads the library until the process exits.
// In host code `g_static_part` is only a stand-in for the symbol, NOT a device address: the
// address is resolved once per device with cudaGetSymbolAddress (that device current) and
// cached.  All calls share the array and it keeps the last call's contents; nothing in it
// has to start at zero (every slot a pass reads was written by that same pass), but two calls
// that overlap on the GPU (different streams) race on it.
static const int kStaticMaxSeg = 64;
static const int kStaticMaxDev = 64;
static __device__ __align__(256) double g_static_part[2 * kStaticMaxSeg * kSegBlocks];
static double *g_static_ptr[kStaticMaxDev] = {nullptr};

static int static_part_ptr(double **out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kStaticMaxDev)
        return fail(PFB_ERR_CUDA, "pfb_seg_stats_static: cannot identify the current device");
    if (!g_static_ptr[dev]) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_part);
        if (e != cudaSuccess || !ptr)
            return fail(PFB_ERR_CUDA, "pfb_seg_stats_static: cudaGetSymbolAddress failed: %s",
                        cudaGetErrorString(e));
        g_static_ptr[dev] = (double *)ptr;
    }
    *out = g_static_ptr[dev];
    return PFB_OK;
}

// same launches as seg_stats_t, on a caller-supplied workspace of 2 * nseg * kSegBlocks doubles
template <class T>
static int seg_stats_ws_t(const T *x, int nseg, long m, double *stats, double *part, cudaStream_t st) {
    long nb = (m + 255) / 256;
    const int blk = (int)(nb < kSegBlocks ? (nb < 1 ? 1 : nb) : kSegBlocks);
    launch(&k_seg_stats<T, 0>, dim3(blk, nseg), dim3(256), 0, st, x, m, (const double *)stats, part);
    PFB_CHECK_LAUNCH();
    launch(&k_seg_final<0>, dim3(nseg), dim3(32), 0, st, (const double *)part, blk, m, stats);
    PFB_CHECK_LAUNCH();
    launch(&k_seg_stats<T, 1>, dim3(blk, nseg), dim3(256), 0, st, x, m, (const double *)stats, part);
    PFB_CHECK_LAUNCH();
    launch(&k_seg_final<1>, dim3(nseg), dim3(32), 0, st, (const double *)part, blk, m, stats);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

}  // namespace pfb

using namespace pfb;

// Gaussian2D (pfb/utils/misc.py:109-138): out = exp(-c (a00 x^2 + a01 x y + a11 y^2)) where
// x^2 + y^2 <= extent, 0 elsewhere; coordinates and arithmetic fp64 like the reference
template <class T>
__global__ void __launch_bounds__(256)
k_gauss2d(const double *__restrict__ xx, const double *__restrict__ yy, long n, double a00, double a01,
          double a11, double c, double extent, T *out) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const double x = xx[i], y = yy[i];
        double g = 0.0;
        if (x * x + y * y <= extent) g = exp(-c * (a00 * x * x + a01 * x * y + a11 * y * y));
        out[i] = (T)g;
    }
}

// convkernhat = gausskernhat / thiskernhat where |thiskernhat| > 0, else 0 (misc.py:224-226);
// num has nnum elements and is broadcast over the nden / nnum blocks of den
template <class T>
__global__ void __launch_bounds__(256)
k_spec_ratio(const T *__restrict__ num, const T *__restrict__ den, T *out, long nnum, long nden) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nden; i += (long)gridDim.x * blockDim.x) {
        const long j = i % nnum;
        const T a = num[2 * j], b = num[2 * j + 1], c = den[2 * i], d = den[2 * i + 1];
        const T m = c * c + d * d;
        T re = (T)0, im = (T)0;
        if (c != (T)0 || d != (T)0) {
            re = (a * c + b * d) / m;
            im = (b * c - a * d) / m;
        }
        out[2 * i] = re;
        out[2 * i + 1] = im;
    }
}

extern "C" {

int pfb_seg_stats(const void *x, int nseg, long long m, double *stats3, int dtype, void *stream) {
    if (!x || !stats3 || nseg < 1 || m < 1) return fail(PFB_ERR_BAD_ARG, "pfb_seg_stats: bad argument");
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32) return seg_stats_t<float>((const float *)x, nseg, (long)m, stats3, st);
    if (dtype == PFB_F64) return seg_stats_t<double>((const double *)x, nseg, (long)m, stats3, st);
    return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
}

int pfb_accum_mul(void *out, const void *x, const void *y, double scale, long long n, int dtype, void *stream) {
    if (!out || !x || n < 0) return fail(PFB_ERR_BAD_ARG, "pfb_accum_mul: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    long nb = (n + 255) / 256;
    const int blk = (int)(nb < kNumSM * 8 ? nb : kNumSM * 8);
    if (dtype == PFB_F32)
        launch(&k_accum_mul<float>, dim3(blk), dim3(256), 0, st, (float *)out, (const float *)x, (const float *)y, (float)scale, (long)n);
    else if (dtype == PFB_F64)
        launch(&k_accum_mul<double>, dim3(blk), dim3(256), 0, st, (double *)out, (const double *)x, (const double *)y, scale, (long)n);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

int pfb_seg_stats_static(const void *x, int nseg, long long m, double *stats3, int dtype, void *stream) {
    if (!x || !stats3 || nseg < 1 || m < 1)
        return fail(PFB_ERR_BAD_ARG, "pfb_seg_stats_static: bad argument");
    if (nseg > kStaticMaxSeg)
        return fail(PFB_ERR_BAD_ARG, "pfb_seg_stats_static: %d segments exceed the static workspace "
                                     "(at most %d); use pfb_seg_stats", nseg, kStaticMaxSeg);
    if (dtype != PFB_F32 && dtype != PFB_F64) return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    double *part = nullptr;
    if (int rc = static_part_ptr(&part)) return rc;
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32) return seg_stats_ws_t<float>((const float *)x, nseg, (long)m, stats3, part, st);
    return seg_stats_ws_t<double>((const double *)x, nseg, (long)m, stats3, part, st);
}

int pfb_l1reweight(const void *bsum, const void *rms_comps, int nbasis, long long nslab, double rmsfactor,
                   double alpha, void *weight, int dtype, void *stream) {
    if (!bsum || !rms_comps || !weight || nbasis < 1 || nslab < 1)
        return fail(PFB_ERR_BAD_ARG, "pfb_l1reweight: bad argument");
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32)
        return l1reweight_t<float>((const float *)bsum, (const float *)rms_comps, nbasis, (long)nslab, rmsfactor,
                                   alpha, (float *)weight, st);
    if (dtype == PFB_F64)
        return l1reweight_t<double>((const double *)bsum, (const double *)rms_comps, nbasis, (long)nslab,
                                    rmsfactor, alpha, (double *)weight, st);
    return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
}


int pfb_gauss2d(const double *xx, const double *yy, long long n, double a00, double a01, double a11,
                double c, double extent, void *out, int dtype, void *stream) {
    if (!xx || !yy || !out || n < 0) return fail(PFB_ERR_BAD_ARG, "pfb_gauss2d: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    long nb = (n + 255) / 256;
    const int blk = (int)(nb < kNumSM * 8 ? nb : kNumSM * 8);
    if (dtype == PFB_F32)
        launch(&k_gauss2d<float>, dim3(blk), dim3(256), 0, st, xx, yy, (long)n, a00, a01, a11, c, extent, (float *)out);
    else if (dtype == PFB_F64)
        launch(&k_gauss2d<double>, dim3(blk), dim3(256), 0, st, xx, yy, (long)n, a00, a01, a11, c, extent, (double *)out);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

int pfb_spec_ratio(const void *num, const void *den, void *out, long long nnum, long long nden, int dtype,
                   void *stream) {
    if (!num || !den || !out || nnum < 1 || nden < 0 || nden % nnum)
        return fail(PFB_ERR_BAD_ARG, "pfb_spec_ratio: bad argument");
    if (nden == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    long nb = (nden + 255) / 256;
    const int blk = (int)(nb < kNumSM * 8 ? nb : kNumSM * 8);
    if (dtype == PFB_F32)
        launch(&k_spec_ratio<float>, dim3(blk), dim3(256), 0, st, (const float *)num, (const float *)den, (float *)out, (long)nnum, (long)nden);
    else if (dtype == PFB_F64)
        launch(&k_spec_ratio<double>, dim3(blk), dim3(256), 0, st, (const double *)num, (const double *)den, (double *)out, (long)nnum, (long)nden);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

}  // extern "C"
