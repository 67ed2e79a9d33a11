// CLEAN minor loops on the device: pfb/deconv/clark.py (clark :84-177,
// subminor :29-82, subtract :12-26) and pfb/deconv/hogbom.py (hogbom :8-72) -- the second
// consumer of the PSF-convolution kernels (clark calls psf_convolve_cube once per minor
// iteration, clark.py:136).
//
//   pfb_mfs_peak        IRsearch = (sum_b IR)^2, its argmax (first index on ties, like
//                       numpy) and value                                  clark.py:116-120,145-150
//   pfb_active_set      ordered list of the pixels with IRsearch > thr     clark.py:127
//   pfb_clark_subminor  the whole sub-minor loop in ONE launch: the active set lives in a
//                       scratch array, one CTA finds the peak, adds the component to the
//                       model and subtracts the shifted PSF, up to maxit times, with no host
//                       round trip                                        clark.py:29-82
//   pfb_hogbom_*        Hogbom iterations driven by device-side state (peak, gain-scaled
//                       flux, done flag): the host enqueues batches of iterations and polls
//                       the flag                                          hogbom.py:36-58
//
// The searches are discontinuous (argmax), so the arithmetic follows the reference's order
// and precision exactly and multiply-subtract pairs are NOT contracted into FMAs (numba and
// numpy do not contract): __fmul_rn / __fsub_rn and their fp64 forms.
#include "common.cuh"

namespace pfb {

#ifdef PFB_EMU
static const int kPeakBlocks = 3;
static inline float mul_rn(float a, float b) { volatile float r = a * b; return r; }
static inline double mul_rn(double a, double b) { volatile double r = a * b; return r; }
static inline float sub_rn(float a, float b) { return a - b; }
static inline double sub_rn(double a, double b) { return a - b; }
#else
static const int kPeakBlocks = kNumSM * 4;
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
#endif
__device__ __forceinline__ float tsqrt(float a) { return sqrtf(a); }
__device__ __forceinline__ double tsqrt(double a) { return sqrt(a); }

// (value, index) pair ordered like numpy's argmax: larger value wins, equal values -> lower index
template <class T> struct Peak { T v; long i; };
template <class T> __device__ __forceinline__ bool better(const Peak<T> &a, const Peak<T> &b) {
    return a.i >= 0 && (b.i < 0 || a.v > b.v || (a.v == b.v && a.i < b.i));
}

template <class T>
__device__ __forceinline__ Peak<T> block_peak(Peak<T> me) {
    __shared__ __align__(16) unsigned char raw[32 * sizeof(Peak<double>)];
    Peak<T> *sh = reinterpret_cast<Peak<T> *>(raw);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Peak<T> ot;
        ot.v = __shfl_down_sync(0xffffffffu, me.v, o);
        ot.i = __shfl_down_sync(0xffffffffu, me.i, o);
        if (better(ot, me)) me = ot;
    }
    __syncthreads();                          // sh may still be read from a previous call
    if (lane == 0) sh[wid] = me;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    Peak<T> r;
    r.v = (T)0; r.i = -1;
    if (wid == 0) {
        if (lane < nw) r = sh[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            Peak<T> ot;
            ot.v = __shfl_down_sync(0xffffffffu, r.v, o);
            ot.i = __shfl_down_sync(0xffffffffu, r.i, o);
            if (better(ot, r)) r = ot;
        }
        if (lane == 0) sh[0] = r;
    }
    __syncthreads();
    return sh[0];                              // every thread gets the block's peak
}

// search[i] = (sum_b IR[b][i])^2 (optional store); per-CTA peak -> part
template <class T>
__global__ void __launch_bounds__(256)
k_mfs_search(const T *__restrict__ IR, int nband, long npix, T *search, Peak<T> *part) {
    Peak<T> me;
    me.v = (T)0; me.i = -1;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (long)gridDim.x * blockDim.x) {
        T s = IR[i];
        for (int b = 1; b < nband; ++b) s = s + IR[(long)b * npix + i];
        const T v = mul_rn(s, s);
        if (search) search[i] = v;
        Peak<T> c;
        c.v = v; c.i = i;
        if (better(c, me)) me = c;
    }
    me = block_peak(me);
    if (threadIdx.x == 0) part[blockIdx.x] = me;
}

// out[0] = index of the peak (as double, exact below 2^53), out[1] = its value
template <class T>
__global__ void __launch_bounds__(256)
k_peak_final(const Peak<T> *__restrict__ part, int n, double *out) {
    Peak<T> me;
    me.v = (T)0; me.i = -1;
    for (int c = threadIdx.x; c < n; c += blockDim.x)
        if (better(part[c], me)) me = part[c];
    me = block_peak(me);
    if (threadIdx.x == 0) { out[0] = (double)me.i; out[1] = (double)me.v; }
}

// ---- ordered compaction: idx[] = ascending pixel indices with search > thr ------------------
const int kScanTile = 2048;                     // pixels per CTA (256 threads x 8)
template <class T>
__global__ void __launch_bounds__(256)
k_active_count(const T *__restrict__ search, long npix, T thr, int *counts) {
    const long base = (long)blockIdx.x * kScanTile;
    int c = 0;
    for (int j = 0; j < 8; ++j) {
        const long i = base + (long)threadIdx.x * 8 + j;
        if (i < npix && search[i] > thr) ++c;
    }
    __shared__ int sh[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if (lane == 0) sh[wid] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        counts[blockIdx.x] = t;
    }
}
// exclusive scan of counts[0..n) in place (single CTA, chunks of 1024), total -> *total
__global__ void __launch_bounds__(1024) k_active_scan(int *counts, int n, long long *total) {
    __shared__ int sh[1024];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < n ? counts[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {       // Hillis-Steele inclusive scan
            const int t = (int)threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < n) counts[i] = carry + sh[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += sh[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}
template <class T>
__global__ void __launch_bounds__(256)
k_active_write(const T *__restrict__ search, long npix, T thr, const int *__restrict__ offs, long cap,
               long long *idx) {
    const long base = (long)blockIdx.x * kScanTile;
    int flag[8], c = 0;
    for (int j = 0; j < 8; ++j) {
        const long i = base + (long)threadIdx.x * 8 + j;
        flag[j] = (i < npix && search[i] > thr) ? 1 : 0;
        c += flag[j];
    }
    // exclusive scan of c over the CTA's 256 threads
    __shared__ int sh[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) sh[wid] = inc;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < wid; ++w) woff += sh[w];
    long pos = (long)offs[blockIdx.x] + woff + inc - c;
    for (int j = 0; j < 8; ++j)
        if (flag[j]) {
            if (pos < cap) idx[pos] = base + (long)threadIdx.x * 8 + j;
            ++pos;
        }
}

// ---- Clark sub-minor loop, one CTA ------------------------------------------------------------
// A: scratch (nband, nact).  info[0] = iterations done, info[1] = last peak value Amax.
template <class T>
__global__ void __launch_bounds__(1024)
k_clark_subminor(const T *__restrict__ IR, const T *__restrict__ psf, const long long *__restrict__ idx,
                 long nact, T *model, const double *__restrict__ wsums, int nband, int nx, int ny, int nxp,
                 int nyp, double gamma, T th, int maxit, T *A, double *info) {
    const long npix = (long)nx * ny;
    const int nxo2 = nxp / 2, nyo2 = nyp / 2;
    __shared__ double s_xhat[64];
    // gather + first search
    Peak<T> me;
    me.v = (T)0; me.i = -1;
    for (long i = threadIdx.x; i < nact; i += blockDim.x) {
        const long pix = idx[i];
        T s = (T)0;
        for (int b = 0; b < nband; ++b) {
            const T a = IR[(long)b * npix + pix];
            A[(long)b * nact + i] = a;
            s = b ? s + a : a;
        }
        Peak<T> c;
        c.v = mul_rn(s, s); c.i = i;
        if (better(c, me)) me = c;
    }
    Peak<T> pk = block_peak(me);
    int k = 0;
    T Amax = pk.i >= 0 ? tsqrt(pk.v) : (T)0;
    while (pk.i >= 0 && Amax > th && k < maxit) {
        const long pix = idx[pk.i];
        const int p = (int)(pix / ny), q = (int)(pix - (long)p * ny);
        if (threadIdx.x < nband) {
            const int b = threadIdx.x;
            const T xh = A[(long)b * nact + pk.i];
            s_xhat[b] = (double)xh;
            if (wsums[b] > 0.0) {               // model[fsel, p, q] += gamma * xhat[fsel] / wsums[fsel]
                const double add = (gamma * (double)xh) / wsums[b];
                model[(long)b * npix + pix] = (T)((double)model[(long)b * npix + pix] + add);
            }
        }
        __syncthreads();
        me.v = (T)0; me.i = -1;
        for (long i = threadIdx.x; i < nact; i += blockDim.x) {
            const long pi = idx[i];
            const int ip = (int)(pi / ny), iq = (int)(pi - (long)ip * ny);
            const int pp = nxo2 - (p - ip), qq = nyo2 - (q - iq);
            T s = (T)0;
            for (int b = 0; b < nband; ++b) {
                const T ps = psf[((long)b * nxp + pp) * nyp + qq];
                const T a = sub_rn(A[(long)b * nact + i], mul_rn((T)s_xhat[b], ps));
                A[(long)b * nact + i] = a;
                s = b ? s + a : a;
            }
            Peak<T> c;
            c.v = mul_rn(s, s); c.i = i;
            if (better(c, me)) me = c;
        }
        pk = block_peak(me);                    // (its barriers also order the s_xhat reuse)
        Amax = tsqrt(pk.v);
        ++k;
    }
    if (threadIdx.x == 0) { info[0] = (double)k; info[1] = (double)Amax; }
}

// ---- Hogbom -----------------------------------------------------------------------------------
// state (doubles): [0] peak pixel index, [1] IRmax, [2] done flag, [3] iterations k, [4] tol,
//                  [5] maxit, [6] stall count (kept 0 like the reference), [8 + b] gamma-less xhat[b]
static const int kHogState = 8;

// One thread block: from the peak found by the last search decide whether to go on
// (hogbom.py:36), and if so take the component: xhat = IR[:, p, q] / wsums, x[:, p, q] += gamma xhat.
template <class T>
__global__ void __launch_bounds__(256)
k_hogbom_decide(const Peak<T> *__restrict__ part, int npart, const T *__restrict__ IR, T *x,
                const T *__restrict__ wsums, int nband, long npix, T gamma, double *st, int first) {
    if (!first && st[2] != 0.0) return;         // finished earlier: the queued tail of a batch is a no-op
    Peak<T> me;
    me.v = (T)0; me.i = -1;
    for (int c = threadIdx.x; c < npart; c += blockDim.x)
        if (better(part[c], me)) me = part[c];
    me = block_peak(me);
    __shared__ int go;
    if (threadIdx.x == 0) {
        const T IRmax = me.i >= 0 ? tsqrt(me.v) : (T)0;
        if (first) {                            // tol = max(pf * IRmax, threshold): st[4] holds pf, st[7] threshold
            const T tol = mul_rn((T)st[4], IRmax);
            st[4] = (double)(tol > (T)st[7] ? tol : (T)st[7]);
            st[3] = 0.0;
        } else {
            st[3] += 1.0;
        }
        st[0] = (double)me.i;
        st[1] = (double)IRmax;
        go = (me.i >= 0 && IRmax > (T)st[4] && st[3] < st[5]) ? 1 : 0;
        st[2] = go ? 0.0 : 1.0;
    }
    __syncthreads();
    if (go && threadIdx.x < nband) {
        const int b = threadIdx.x;
        const long pix = (long)st[0];
        const T xh = IR[(long)b * npix + pix] / wsums[b];
        st[kHogState + b] = (double)xh;
        x[(long)b * npix + pix] = x[(long)b * npix + pix] + mul_rn(gamma, xh);
    }
}

// IR -= gamma * xhat * psf shifted to the peak; new search + per-CTA peaks (hogbom.py:39-47)
template <class T>
__global__ void __launch_bounds__(256)
k_hogbom_subtract(T *IR, const T *__restrict__ psf, int nband, int nx, int ny, int nxp, int nyp, T gamma,
                  const double *__restrict__ st, Peak<T> *part) {
    if (st[2] != 0.0) return;                   // finished: the queued tail of the batch is a no-op
    const long npix = (long)nx * ny;
    const long pix0 = (long)st[0];
    const int p = (int)(pix0 / ny), q = (int)(pix0 - (long)p * ny);
    const int nx0 = nxp / 2, ny0 = nyp / 2;
    Peak<T> me;
    me.v = (T)0; me.i = -1;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (long)gridDim.x * blockDim.x) {
        const int r = (int)(i / ny), c = (int)(i - (long)r * ny);
        T s = (T)0;
        for (int b = 0; b < nband; ++b) {
            const T gx = mul_rn(gamma, (T)st[kHogState + b]);
            const T ps = psf[((long)b * nxp + (nx0 - p + r)) * nyp + (ny0 - q + c)];
            const T a = sub_rn(IR[(long)b * npix + i], mul_rn(gx, ps));
            IR[(long)b * npix + i] = a;
            s = b ? s + a : a;
        }
        Peak<T> cnd;
        cnd.v = mul_rn(s, s); cnd.i = i;
        if (better(cnd, me)) me = cnd;
    }
    me = block_peak(me);
    if (threadIdx.x == 0) part[blockIdx.x] = me;
}

static void *g_cpart = nullptr;
static int ensure_cpart() {
    if (!g_cpart && cudaMalloc(&g_cpart, sizeof(Peak<double>) * kPeakBlocks) != cudaSuccess)
        return fail(PFB_ERR_ALLOC, "cudaMalloc of the peak-search workspace failed");
    return PFB_OK;
}
static inline int peak_grid(long npix) {
    long b = (npix + 255) / 256;
    return (int)(b < 1 ? 1 : (b > kPeakBlocks ? kPeakBlocks : b));
}

template <class T>
static int mfs_peak_t(const T *IR, int nband, long npix, T *search, double *out2, cudaStream_t st) {
    if (int rc = ensure_cpart()) return rc;
    const int g = peak_grid(npix);
    launch(&k_mfs_search<T>, dim3(g), dim3(256), 0, st, IR, nband, npix, search, (Peak<T> *)g_cpart);
    PFB_CHECK_LAUNCH();
    launch(&k_peak_final<T>, dim3(1), dim3(256), 0, st, (const Peak<T> *)g_cpart, g, out2);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int active_set_t(const T *search, long npix, double thr, long long *idx, long cap, long long *count,
                        int *counts, cudaStream_t st) {
    const int nb = cdiv(npix, kScanTile);
    launch(&k_active_count<T>, dim3(nb), dim3(256), 0, st, search, npix, (T)thr, counts);
    PFB_CHECK_LAUNCH();
    launch(&k_active_scan, dim3(1), dim3(1024), 0, st, counts, nb, count);
    PFB_CHECK_LAUNCH();
    launch(&k_active_write<T>, dim3(nb), dim3(256), 0, st, search, npix, (T)thr, (const int *)counts, cap, idx);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

}  // namespace pfb

using namespace pfb;

extern "C" {

int pfb_mfs_peak(const void *IR, int nband, long long npix, void *search, double *out2, int dtype,
                 void *stream) {
    if (!IR || !out2 || nband < 1 || npix < 1) return fail(PFB_ERR_BAD_ARG, "pfb_mfs_peak: bad argument");
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32) return mfs_peak_t<float>((const float *)IR, nband, (long)npix, (float *)search, out2, st);
    if (dtype == PFB_F64) return mfs_peak_t<double>((const double *)IR, nband, (long)npix, (double *)search, out2, st);
    return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
}

int pfb_active_set(const void *search, long long npix, double thr, long long *idx, long long cap,
                   long long *count, int *scratch_counts, int dtype, void *stream) {
    if (!search || !idx || !count || !scratch_counts || npix < 1 || cap < 0)
        return fail(PFB_ERR_BAD_ARG, "pfb_active_set: bad argument");
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32) return active_set_t<float>((const float *)search, (long)npix, thr, idx, (long)cap, count, scratch_counts, st);
    if (dtype == PFB_F64) return active_set_t<double>((const double *)search, (long)npix, thr, idx, (long)cap, count, scratch_counts, st);
    return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
}

int pfb_clark_subminor(const void *IR, const void *psf, const long long *idx, long long nact, void *model,
                       const double *wsums, int nband, int nx, int ny, int nxp, int nyp, double gamma,
                       double th, int maxit, void *scratchA, double *info2, int dtype, void *stream) {
    if (!IR || !psf || !idx || !model || !wsums || !scratchA || !info2 || nact < 1 || nband < 1 || nband > 64)
        return fail(PFB_ERR_BAD_ARG, "pfb_clark_subminor: bad argument (1 <= nband <= 64, nact >= 1)");
    if (nxp < 2 * nx - 1 || nyp < 2 * ny - 1)
        return fail(PFB_ERR_BAD_ARG, "pfb_clark_subminor: the PSF must cover every shift of the image "
                                     "(nxp >= 2 nx - 1, nyp >= 2 ny - 1), as the reference's active-set "
                                     "masking assumes");
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32)
        launch(&k_clark_subminor<float>, dim3(1), dim3(1024), 0, st, (const float *)IR, (const float *)psf, idx,
               (long)nact, (float *)model, wsums, nband, nx, ny, nxp, nyp, gamma, (float)th, maxit, (float *)scratchA, info2);
    else if (dtype == PFB_F64)
        launch(&k_clark_subminor<double>, dim3(1), dim3(1024), 0, st, (const double *)IR, (const double *)psf, idx,
               (long)nact, (double *)model, wsums, nband, nx, ny, nxp, nyp, gamma, th, maxit, (double *)scratchA, info2);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

int pfb_hogbom_start(const void *IR, void *x, const void *wsums, int nband, long long npix, double gamma,
                     double pf, double threshold, int maxit, double *state, int dtype, void *stream) {
    if (!IR || !x || !wsums || !state || nband < 1 || nband > 64 || npix < 1)
        return fail(PFB_ERR_BAD_ARG, "pfb_hogbom_start: bad argument");
    if (int rc = ensure_cpart()) return rc;
    cudaStream_t st = as_stream(stream);
    double h[kHogState] = {-1.0, 0.0, 0.0, 0.0, pf, (double)maxit, 0.0, threshold};
    PFB_CUDA_OK(cudaMemcpyAsync(state, h, sizeof h, cudaMemcpyHostToDevice, st));
    const int g = peak_grid(npix);
    if (dtype == PFB_F32) {
        launch(&k_mfs_search<float>, dim3(g), dim3(256), 0, st, (const float *)IR, nband, (long)npix, (float *)nullptr, (Peak<float> *)g_cpart);
        launch(&k_hogbom_decide<float>, dim3(1), dim3(256), 0, st, (const Peak<float> *)g_cpart, g, (const float *)IR, (float *)x,
               (const float *)wsums, nband, (long)npix, (float)gamma, state, 1);
    } else if (dtype == PFB_F64) {
        launch(&k_mfs_search<double>, dim3(g), dim3(256), 0, st, (const double *)IR, nband, (long)npix, (double *)nullptr, (Peak<double> *)g_cpart);
        launch(&k_hogbom_decide<double>, dim3(1), dim3(256), 0, st, (const Peak<double> *)g_cpart, g, (const double *)IR, (double *)x,
               (const double *)wsums, nband, (long)npix, gamma, state, 1);
    } else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

int pfb_hogbom_iterate(void *IR, const void *psf, void *x, const void *wsums, int nband, int nx, int ny,
                       int nxp, int nyp, double gamma, int niter, double *state, int dtype, void *stream) {
    if (!IR || !psf || !x || !wsums || !state || nband < 1 || nband > 64 || niter < 1)
        return fail(PFB_ERR_BAD_ARG, "pfb_hogbom_iterate: bad argument");
    if (nxp < 2 * nx || nyp < 2 * ny)
        return fail(PFB_ERR_BAD_ARG, "pfb_hogbom_iterate: the PSF must be at least twice the image");
    if (int rc = ensure_cpart()) return rc;
    cudaStream_t st = as_stream(stream);
    const long npix = (long)nx * ny;
    const int g = peak_grid(npix);
    for (int it = 0; it < niter; ++it) {
        if (dtype == PFB_F32) {
            launch(&k_hogbom_subtract<float>, dim3(g), dim3(256), 0, st, (float *)IR, (const float *)psf, nband, nx, ny, nxp, nyp,
                   (float)gamma, (const double *)state, (Peak<float> *)g_cpart);
            launch(&k_hogbom_decide<float>, dim3(1), dim3(256), 0, st, (const Peak<float> *)g_cpart, g, (const float *)IR,
                   (float *)x, (const float *)wsums, nband, npix, (float)gamma, state, 0);
        } else if (dtype == PFB_F64) {
            launch(&k_hogbom_subtract<double>, dim3(g), dim3(256), 0, st, (double *)IR, (const double *)psf, nband, nx, ny, nxp, nyp,
                   gamma, (const double *)state, (Peak<double> *)g_cpart);
            launch(&k_hogbom_decide<double>, dim3(1), dim3(256), 0, st, (const Peak<double> *)g_cpart, g, (const double *)IR,
                   (double *)x, (const double *)wsums, nband, npix, gamma, state, 0);
        } else
            return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
        PFB_CHECK_LAUNCH();
    }
    return PFB_OK;
}

}  // extern "C"
