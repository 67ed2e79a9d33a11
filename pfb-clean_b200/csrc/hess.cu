// Image-space PSF Hessian: zero-pad -> r2c -> x psfhat -> c2r -> crop, fused
// into three shared-memory FFT passes (sm_100a).
//
// Reference: psf_convolve_cube pfb/operators/psf.py:32-56 (xpad = 0;
// xpad[:, :nx, :ny] = x; r2c(axes=(1,2)); xhat *= psfhat; c2r(inorm=2,
// lastsize=nyp); xout = xpad[:, :nx, :ny]) and hessian_psf_cube
// pfb/operators/hessian.py:254-281 (beam prologue/epilogue, /wsum,
// + sigmainv*x).  The padded image and the full half-spectrum are never
// materialised:
//   pass A  rows : the nx non-zero image rows only.  Row of ny reals (beam
//                  fused) packed as nyp/2 complex, complex FFT, real-FFT
//                  post-processing, stored TRANSPOSED: S[band][k][row].
//   pass B  cols : for each of the nyo2 = nyp/2+1 columns: nx contiguous
//                  points zero-extended to nxp, forward FFT, x psfhat^T
//                  (pre-transposed, pre-scaled by 1/(nxp*nyp)), inverse FFT in
//                  the same registers, first nx outputs stored in place.
//   pass C  rows : the nx output rows only.  Transposed gather, real-FFT
//                  pre-processing, inverse complex FFT, crop to ny, epilogue
//                  (beam, /wsum, + sigmainv*x) fused into the store.
// HBM traffic per band: x + out + psfhat^T + 4 x (nx x nyo2 complex) instead
// of the reference's ~5 passes over the 4x larger padded arrays.
#include "hess_plan.cuh"
#include <cmath>
#include <vector>

namespace pfb {

// ------------------------------------------------------------------ pass A
template <class T, int MAXT>
__global__ void __launch_bounds__(MAXT) k_hess_rows_fwd(const T *__restrict__ x, const T *__restrict__ beam,
                                Cx<T> *__restrict__ S, const Cx<T> *__restrict__ tw,
                                FftDesc d, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *smem = reinterpret_cast<Cx<T> *>(raw);
    const int NT = d.NT, M = h.M;
    const int stride = fft_smem_elems(M);
    const int lr = threadIdx.x / NT, t = threadIdx.x - lr * NT;
    const int band = blockIdx.y;
    const int row = blockIdx.x * h.R + lr;
    Cx<T> *buf = smem + (long)lr * stride;

    Cx<T> v[16];
    const long ro = ((long)band * h.nx + row) * h.ny;
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        int c0 = 2 * (t + e * NT);
        T a = (T)0, b = (T)0;
        if (row < h.nx) {
            if (c0 < h.ny) { a = x[ro + c0]; if (beam) a *= beam[ro + c0]; }
            if (c0 + 1 < h.ny) { b = x[ro + c0 + 1]; if (beam) b *= beam[ro + c0 + 1]; }
        }
        v[e] = cx<T>(a, b);
    }
    fft_run<T, -1>(v, d, t, buf, tw);
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 16; ++e) buf[fft_pad(t + e * NT)] = v[e];
    __syncthreads();

    // X[k] = E + w^k O,  E = (Z[k] + conj Z[M-k])/2,  O = -i (Z[k] - conj Z[M-k])/2
    const int twm = h.twN / h.nyp;
    for (int i = threadIdx.x; i < (M + 1) * h.R; i += blockDim.x) {
        int k = i / h.R, l2 = i - k * h.R;
        int row2 = blockIdx.x * h.R + l2;
        const Cx<T> *b2 = smem + (long)l2 * stride;
        Cx<T> zk = b2[fft_pad(k & (M - 1))], zm = cconj(b2[fft_pad((M - k) & (M - 1))]);
        Cx<T> E = cx<T>((T)0.5 * (zk.x + zm.x), (T)0.5 * (zk.y + zm.y));
        Cx<T> D = cx<T>((T)0.5 * (zk.x - zm.x), (T)0.5 * (zk.y - zm.y));
        Cx<T> O = cx<T>(D.y, -D.x);
        Cx<T> X = E + cmul(tw[(long)k * twm], O);
        if (row2 < h.nx) S[((long)band * h.nyo2 + k) * h.nx + row2] = X;
    }
}

// ------------------------------------------------------------------ pass B
template <class T, int MAXT>
__global__ void __launch_bounds__(MAXT) k_hess_cols(Cx<T> *S, const Cx<T> *__restrict__ psfT,
                            const Cx<T> *__restrict__ tw, FftDesc d, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *smem = reinterpret_cast<Cx<T> *>(raw);
    const int NT = d.NT;
    const int lc = threadIdx.x / NT, t = threadIdx.x - lc * NT;
    const int band = blockIdx.y;
    const int col = blockIdx.x * h.R + lc;
    const bool live = col < h.nyo2;
    Cx<T> *buf = smem + (long)lc * fft_smem_elems(h.nxp);
    Cx<T> *line = S + ((long)band * h.nyo2 + (live ? col : 0)) * h.nx;
    const Cx<T> *pl = psfT + ((long)band * h.nyo2 + (live ? col : 0)) * h.nxp;

    Cx<T> v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        int i = t + e * NT;
        v[e] = (live && i < h.nx) ? line[i] : cx<T>((T)0, (T)0);
    }
    fft_run<T, -1>(v, d, t, buf, tw);
#pragma unroll
    for (int e = 0; e < 16; ++e) v[e] = cmul(v[e], pl[t + e * NT]);
    fft_run<T, +1>(v, d, t, buf, tw);
    if (live) {
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            int i = t + e * NT;
            if (i < h.nx) line[i] = v[e];
        }
    }
}

// ------------------------------------------------------------------ pass C
template <class T, int MAXT>
__global__ void __launch_bounds__(MAXT) k_hess_rows_inv(const Cx<T> *__restrict__ S, T *__restrict__ out,
                                const Cx<T> *__restrict__ tw, FftDesc d, HessDims h,
                                Epilogue<T> ep) {
    PFB_DYN_SMEM(raw);
    Cx<T> *smem = reinterpret_cast<Cx<T> *>(raw);
    const int NT = d.NT, M = h.M;
    const int stride = fft_smem_elems(M);
    const int lr = threadIdx.x / NT, t = threadIdx.x - lr * NT;
    const int band = blockIdx.y;
    const int row = blockIdx.x * h.R + lr;
    Cx<T> *buf = smem + (long)lr * stride;

    for (int i = threadIdx.x; i < (M + 1) * h.R; i += blockDim.x) {
        int k = i / h.R, l2 = i - k * h.R;
        int row2 = blockIdx.x * h.R + l2;
        Cx<T> X = cx<T>((T)0, (T)0);
        if (row2 < h.nx) X = S[((long)band * h.nyo2 + k) * h.nx + row2];
        smem[(long)l2 * stride + fft_pad(k)] = X;
    }
    __syncthreads();

    // Z'[m] = (X[m] + conj X[M-m]) + i w^-m (X[m] - conj X[M-m])   (= 2 Z[m])
    const int twm = h.twN / h.nyp;
    Cx<T> v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        int m = t + e * NT;
        Cx<T> xm = buf[fft_pad(m)], xr = cconj(buf[fft_pad(M - m)]);
        Cx<T> A = xm + xr, B = xm - xr;
        Cx<T> wB = cmul(cconj(tw[(long)m * twm]), B);
        v[e] = cx<T>(A.x - wB.y, A.y + wB.x);
    }
    fft_run<T, +1>(v, d, t, buf, tw);

    if (row < h.nx) {
        const long ro = ((long)band * h.nx + row) * h.ny;
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            int c0 = 2 * (t + e * NT);
            T val[2] = {v[e].x, v[e].y};
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                int c = c0 + q;
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

// PSFHAT creation (pfb/operators/gridder.py:712-714, psfhat = r2c(iFs(psf))): column pass,
// forward only.  S holds the row transforms of the UNSHIFTED psf (pass A over all nxp rows);
// the ifftshift by (nxp/2, nyp/2) is the sign (-1)^(kx + ky) on the spectrum (both sizes even),
// which is exact.  In place: line[kx] = sign * FFT_nxp(line)[kx].
template <class T, int MAXT>
__global__ void __launch_bounds__(MAXT) k_psf_cols_fwd(Cx<T> *S, const Cx<T> *__restrict__ tw,
                                                       FftDesc d, HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *smem = reinterpret_cast<Cx<T> *>(raw);
    const int NT = d.NT;
    const int lc = threadIdx.x / NT, t = threadIdx.x - lc * NT;
    const int band = blockIdx.y;
    const int col = blockIdx.x * h.R + lc;
    const bool live = col < h.nyo2;
    Cx<T> *buf = smem + (long)lc * fft_smem_elems(h.nxp);
    Cx<T> *line = S + ((long)band * h.nyo2 + (live ? col : 0)) * h.nx;
    Cx<T> v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        int i = t + e * NT;
        v[e] = (live && i < h.nx) ? line[i] : cx<T>((T)0, (T)0);
    }
    fft_run<T, -1>(v, d, t, buf, tw);
    if (live) {
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            int i = t + e * NT;
            const T sg = ((i + col) & 1) ? (T)-1 : (T)1;
            if (i < h.nx) line[i] = cx<T>(sg * v[e].x, sg * v[e].y);
        }
    }
}

// same for padded sizes that are not powers of two (hess_mixed.cuh machinery, N = nxp)
template <class T>
__global__ void __launch_bounds__(512)
k_psfm_cols_fwd(Cx<T> *S, const Cx<T> *__restrict__ tw, const Cx<T> *__restrict__ roots, MxDesc d,
                HessDims h) {
    PFB_DYN_SMEM(raw);
    Cx<T> *rt = reinterpret_cast<Cx<T> *>(raw);
    const int N = d.N;
    Cx<T> *A = rt + kMxHead, *B = A + N;
    const int band = blockIdx.y, col = blockIdx.x;
    Cx<T> *line = S + ((long)band * h.nyo2 + col) * h.nx;
    mx_load_tables(rt, roots, tw, N);
    for (int n = threadIdx.x; n < N; n += blockDim.x) A[n] = n < h.nx ? line[n] : cx<T>((T)0, (T)0);
    __syncthreads();
    const Cx<T> *c = mx_fft<T, -1>(A, B, 1, d, rt + 40, rt);
    for (int n = threadIdx.x; n < h.nx; n += blockDim.x) {
        const T sg = ((n + col) & 1) ? (T)-1 : (T)1;
        line[n] = cx<T>(sg * c[n].x, sg * c[n].y);
    }
}

// psfhat (nband, nxp, nyo2) -> psfT (nband, nyo2, nxp) * scale
template <class T>
__global__ void k_psfhat_prepare(const Cx<T> *__restrict__ in, Cx<T> *__restrict__ out,
                                 int nxp, int nyo2, T scale, int deint, int pm, int pP2) {
    __shared__ Cx<T> tile[32][33];
    int band = blockIdx.z;
    const Cx<T> *src = in + (long)band * nxp * nyo2;
    Cx<T> *dst = out + (long)band * nxp * nyo2;
    int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int r = r0 + i, c = c0 + threadIdx.x;
        if (r < nxp && c < nyo2) tile[i][threadIdx.x] = src[(long)r * nyo2 + c];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int c = c0 + i, r = r0 + threadIdx.x;
        if (r < nxp && c < nyo2) {
            Cx<T> v = tile[threadIdx.x][i];
            // fast path: rows de-interleaved per column, [even rows | odd rows]
            int rr = deint ? ((r & 1) * (nxp >> 1) + (r >> 1)) : r;
            if (pm) {    // k_hessp_cols2: element K of a (half) spectrum at [K mod m][K / m]
                const int N = pm * pP2, hf = rr / N, K = rr - hf * N;
                rr = hf * N + (K % pm) * pP2 + K / pm;
            }
            dst[(long)c * nxp + rr] = cx<T>(v.x * scale, v.y * scale);
        }
    }
}

}  // namespace pfb


namespace pfb {


// kernels are compiled for several block-size caps so that the register
// allocation always fits (1024 threads -> 64 regs, 512 -> 128, <= 256 -> 255)
#define PFB_LAUNCH_T(MAXT, KERN, smem, grid, nthr, st, ...)                                   \
    do {                                                                                       \
        PFB_CUDA_OK(cudaFuncSetAttribute(&KERN<T, MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem))); \
        launch(&KERN<T, MAXT>, grid, dim3(nthr), smem, st, __VA_ARGS__);                       \
        PFB_CHECK_LAUNCH();                                                                    \
    } while (0)
#define PFB_BY_THREADS(nthr, KERN, smem, grid, st, ...)                                       \
    do {                                                                                       \
        if ((nthr) <= 256) PFB_LAUNCH_T(256, KERN, smem, grid, nthr, st, __VA_ARGS__);         \
        else if ((nthr) <= 512) PFB_LAUNCH_T(512, KERN, smem, grid, nthr, st, __VA_ARGS__);    \
        else PFB_LAUNCH_T(1024, KERN, smem, grid, nthr, st, __VA_ARGS__);                      \
    } while (0)

template <class T>
static int hess_set_psf_t(pfb_hess_plan *p, const void *psfhat, cudaStream_t st) {
    const HessDims &h = p->h;
    dim3 grid(cdiv(h.nyo2, 32), cdiv(h.nxp, 32), h.nband), block(32, 8);
    T scale = (T)(1.0 / ((double)h.nxp * (double)h.nyp));
    if (!p->psfT) {      // allocated on first use: a plan made only for pfb_psfhat_from_psf never needs it
        const size_t sP = 2 * sizeof(T) * (size_t)h.nband * h.nyo2 * h.nxp;
        if (cudaMalloc(&p->psfT, sP) != cudaSuccess) {
            cudaGetLastError();
            p->psfT = nullptr;
            return fail(PFB_ERR_ALLOC, "hessian plan: allocation of the transposed psfhat failed (%zu bytes)", sP);
        }
    }
    launch(&k_psfhat_prepare<T>, grid, block, 0, st, (const Cx<T> *)psfhat, (Cx<T> *)p->psfT,
           h.nxp, h.nyo2, scale, p->deint, p->pcol2 ? p->pcol.m : 0, p->pcol.P2);
    PFB_CHECK_LAUNCH();
    p->have_psf = 1;
    return PFB_OK;
}


// ---------------------------------------------------------------- fast path

template <class T, int M, int RB, bool TWS>
static int fast_rows_fwd_k(pfb_hess_plan *p, const T *x, const T *beam, cudaStream_t st, size_t smem) {
    const HessDims &h = p->h;
    auto kern = p->dyn ? &k_hessf_rows_fwd<T, M, RB, TWS, true> : &k_hessf_rows_fwd<T, M, RB, TWS, false>;
    PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int nbb = cdiv(h.nx, RB);
    int grid = persistent_grid((const void *)kern, M / 8, smem, nbb * h.nband, p->nsm);
    launch(kern, dim3(grid), dim3(M / 8), smem, st, x, beam, (Cx<T> *)p->S, (const Cx<T> *)p->tabM,
           (const Cx<T> *)p->twr, h, nbb, p->ctr + 0);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T, int M, int RB>
static int fast_rows_fwd(pfb_hess_plan *p, const T *x, const T *beam, cudaStream_t st) {
    const size_t smem = sizeof(Cx<T>) * ((size_t)RB * ct_stride(M) + ct_tab_size(M));
    const size_t smem_tw = smem + sizeof(Cx<T>) * (M + 1);
    if (smem_tw <= kMaxSmemFast && !getenv("PFB_HESS_NOTWS"))
        return fast_rows_fwd_k<T, M, RB, true>(p, x, beam, st, smem_tw);
    return fast_rows_fwd_k<T, M, RB, false>(p, x, beam, st, smem);
}

#ifndef PFB_EMU
// S described to the TMA unit: dims (row, k, band), 8-byte elements, box (2, 256, 1)
static bool make_tma_S(pfb_hess_plan *p) {
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError();
        return false;
    }
    const HessDims &h = p->h;
    const cuuint64_t dims[3] = {(cuuint64_t)h.nx, (cuuint64_t)h.nyo2, (cuuint64_t)h.nband};
    const cuuint64_t strides[2] = {(cuuint64_t)h.nx * 8, (cuuint64_t)h.nx * 8 * (cuuint64_t)h.nyo2};
    const cuuint32_t box[3] = {2, (cuuint32_t)kTmaBoxK, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult rc = ((encode_fn)fn)(&p->tmS, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, p->S, dims, strides, box, estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return rc == CUDA_SUCCESS;
}

#endif

template <class T, int H>
static int fast_cols1(pfb_hess_plan *p, cudaStream_t st) {
    const HessDims &h = p->h;
    const size_t smem = sizeof(Cx<T>) * ((size_t)2 * ct_stride(H) + ct_tab_size(H));
    auto kern = p->dyn ? &k_hessf_cols1<T, H, true> : &k_hessf_cols1<T, H, false>;
    PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = persistent_grid((const void *)kern, H / 16, smem, h.nyo2 * h.nband, p->nsm);
    static const int stagger = getenv("PFB_HESS_STAGGER") ? atoi(getenv("PFB_HESS_STAGGER")) : 0;
    launch(kern, dim3(grid), dim3(H / 16), smem, st, (Cx<T> *)p->S, (const Cx<T> *)p->psfT,
           (const Cx<T> *)p->tabH, (const Cx<T> *)p->modg, h, p->ctr + 1, H <= 4096 ? stagger : 0);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T, int H>
static int fast_cols(pfb_hess_plan *p, cudaStream_t st) {
    const HessDims &h = p->h;
    // one-group CTAs, two per SM (see k_hessf_cols1) when both fit in shared memory
    if (2 * sizeof(Cx<T>) * ((size_t)2 * ct_stride(H) + ct_tab_size(H)) + 2048 <= 227 * 1024 &&
        !getenv("PFB_HESS_COLS2G"))
        return fast_cols1<T, H>(p, st);
    size_t smem = sizeof(Cx<T>) * ((size_t)2 * ct_stride(H) + ct_tab_size(H));
    // prefetching variant: 3H more complex values; S columns must be 16-byte aligned
    const size_t smem_pf = smem + sizeof(Cx<T>) * 3 * (size_t)H;
    const bool pf = smem_pf <= kMaxSmemFast && (h.nx * sizeof(Cx<T>)) % 16 == 0 && !getenv("PFB_HESS_NOPF");
    if (pf) {
        auto kern = p->dyn ? &k_hessf_cols<T, H, true, true> : &k_hessf_cols<T, H, true, false>;
        PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pf));
        int grid = persistent_grid((const void *)kern, H / 8, smem_pf, h.nyo2 * h.nband, p->nsm);
        launch(kern, dim3(grid), dim3(H / 8), smem_pf, st, (Cx<T> *)p->S, (const Cx<T> *)p->psfT,
               (const Cx<T> *)p->tabH, (const Cx<T> *)p->modg, h, p->ctr + 1);
    } else {
        auto kern = p->dyn ? &k_hessf_cols<T, H, false, true> : &k_hessf_cols<T, H, false, false>;
        PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int grid = persistent_grid((const void *)kern, H / 8, smem, h.nyo2 * h.nband, p->nsm);
        launch(kern, dim3(grid), dim3(H / 8), smem, st, (Cx<T> *)p->S, (const Cx<T> *)p->psfT,
               (const Cx<T> *)p->tabH, (const Cx<T> *)p->modg, h, p->ctr + 1);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

// M = 8192 (fp32): one 512-thread transform group per CTA, two rows per batch
static int fast_rows_fwd_8k(pfb_hess_plan *p, const float *x, const float *beam, cudaStream_t st) {
    constexpr int M = 8192, RB = 2;
    const HessDims &h = p->h;
    const size_t smem = sizeof(Cx<float>) * ((size_t)RB * ct_stride(M) + ct_tab_size(M));
    auto kern = p->dyn ? &k_hessf_rows_fwd<float, M, RB, false, true, 1> : &k_hessf_rows_fwd<float, M, RB, false, false, 1>;
    PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int nbb = cdiv(h.nx, RB);
    int grid = persistent_grid((const void *)kern, M / 16, smem, nbb * h.nband, p->nsm);
    launch(kern, dim3(grid), dim3(M / 16), smem, st, x, beam, (Cx<float> *)p->S, (const Cx<float> *)p->tabM,
           (const Cx<float> *)p->twr, h, nbb, p->ctr + 0);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}


template <class T> static int fast_rows_fwd_d(pfb_hess_plan *p, const T *x, const T *beam, cudaStream_t st) {
    if (p->RB == 4) {
#define PFB_C(N) fast_rows_fwd<T, N, 4>(p, x, beam, st)
        PFB_SIZE_SWITCH(p->h.M, PFB_C)
#undef PFB_C
    }
#define PFB_C(N) fast_rows_fwd<T, N, 2>(p, x, beam, st)
    PFB_SIZE_SWITCH(p->h.M, PFB_C)
#undef PFB_C
}
template <class T> static int fast_cols_d(pfb_hess_plan *p, cudaStream_t st) {
#define PFB_C(N) fast_cols<T, N>(p, st)
    PFB_SIZE_SWITCH(p->h.nxp / 2, PFB_C)
#undef PFB_C
}

static inline int mixed_threads(long work) { return work >= 2048 ? 512 : work >= 512 ? 256 : 128; }

// passes A and B of a mixed-radix plan: peeled kernels where a pass's length allows, the run-time
// radix list otherwise
template <class T>
static int mixed_fwd_cols(pfb_hess_plan *p, const T *x, const T *beam_in, cudaStream_t st) {
    HessDims h = p->h;
    h.R = p->Lrow;
    Cx<T> *S = (Cx<T> *)p->S;
    const size_t cs = sizeof(Cx<T>);
    const size_t smr = mx_smem_bytes(h.M, h.R, cs), smc = mx_smem_bytes(p->mcol.N, 1, cs);
    const int tr = mixed_threads((long)h.M * h.R / 4), tc = mixed_threads(p->mcol.N / 4);
    const dim3 grow(cdiv(h.nx, h.R), h.nband), gcol(h.nyo2, h.nband);
    if (p->prow.m) {
        if (int rc = peel_rows_fwd_d<T>(p, x, beam_in, st)) return rc;
    } else {
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessm_rows_fwd<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smr));
        launch(&k_hessm_rows_fwd<T>, grow, dim3(tr), smr, st, x, beam_in, S, (const Cx<T> *)p->mtwM,
               (const Cx<T> *)p->twr, (const Cx<T> *)p->mroots, p->mrow, h);
    }
    PFB_CHECK_LAUNCH();
    if (p->pcol.m) {
        if (int rc = peel_cols_d<T>(p, st)) return rc;
    } else if (p->split) {
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessm_cols<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smc));
        launch(&k_hessm_cols<T, true>, gcol, dim3(tc), smc, st, S, (const Cx<T> *)p->psfT,
               (const Cx<T> *)p->mtwC, (const Cx<T> *)p->modg, (const Cx<T> *)p->mroots, p->mcol, h);
    } else {
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessm_cols<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smc));
        launch(&k_hessm_cols<T, false>, gcol, dim3(tc), smc, st, S, (const Cx<T> *)p->psfT,
               (const Cx<T> *)p->mtwC, (const Cx<T> *)p->modg, (const Cx<T> *)p->mroots, p->mcol, h);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

// Passes A and B (everything that needs x only).  `alt` = every other array the row passes will
// touch with 2-element vector accesses (their alignment decides between the fast and the generic
// row kernels, and both row passes must take the same decision).
template <class T>
static int hess_fwd_cols_t(pfb_hess_plan *p, const T *x, const T *beam_in, uintptr_t alt, cudaStream_t st) {
    const HessDims &h0 = p->h;
    const Cx<T> *tw = (const Cx<T> *)p->tw;
    Cx<T> *S = (Cx<T> *)p->S;
    if (p->mixed) return mixed_fwd_cols<T>(p, x, beam_in, st);
    const uintptr_t al = 2 * sizeof(T) - 1;
    const bool aligned = !(((uintptr_t)x | (uintptr_t)beam_in | alt) & al);
    const bool fast_rows = p->fast && aligned;
    const bool rows8k = p->fastrows && sizeof(T) == 4 && aligned;
    if ((p->fast || p->fastcols || p->fastrows) && p->dyn) PFB_CUDA_OK(cudaMemsetAsync(p->ctr, 0, 3 * sizeof(int), st));
    if (fast_rows) {
        if (int rc = fast_rows_fwd_d<T>(p, x, beam_in, st)) return rc;
    } else if (rows8k) {
        if constexpr (sizeof(T) == 4) {
            if (int rc = fast_rows_fwd_8k(p, x, beam_in, st)) return rc;
        }
    } else {
        HessDims h = h0;
        h.R = p->Rrow;
        size_t smem = sizeof(Cx<T>) * (size_t)fft_smem_elems(h.M) * h.R;
        int nthr = p->drow.NT * h.R;
        PFB_BY_THREADS(nthr, k_hess_rows_fwd, smem, dim3(cdiv(h.nx, h.R), h.nband), st,
                       x, beam_in, S, tw, p->drow, h);
    }
    if (p->fast) {
        if (int rc = fast_cols_d<T>(p, st)) return rc;
    } else if (p->fastcols) {
        if constexpr (sizeof(T) == 4) {
            if (int rc = fast_cols1<T, 8192>(p, st)) return rc;
        }
    } else {
        HessDims h = h0;
        h.R = p->Rcol;
        size_t smem = sizeof(Cx<T>) * (size_t)fft_smem_elems(h.nxp) * h.R;
        int nthr = p->dcol.NT * h.R;
        PFB_BY_THREADS(nthr, k_hess_cols, smem, dim3(cdiv(h.nyo2, h.R), h.nband), st,
                       S, (const Cx<T> *)p->psfT, tw, p->dcol, h);
    }
    return PFB_OK;
}

// Pass C.  pa != nullptr: the primal step is its epilogue (only where hess_can_fuse says so).
template <class T>
static int hess_rows_inv_t(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, bool aligned, cudaStream_t st,
                           const PrimalArgsT<T> *pa = nullptr, double *out2 = nullptr) {
    const HessDims &h0 = p->h;
    const Cx<T> *tw = (const Cx<T> *)p->tw;
    Cx<T> *S = (Cx<T> *)p->S;
    if (p->mixed) {
        HessDims h = h0;
        h.R = p->Lrow;
        if (p->prow.m) {
#ifndef PFB_EMU
            if constexpr (sizeof(T) == 4) {
                if (pa) return peel_rows_inv_primal_d(p, *pa, out2, st);
            }
#endif
            if (pa) return fail(PFB_ERR_BAD_ARG, "fused primal pass: not available for this plan");
            return peel_rows_inv_d<T>(p, out, ep, st);
        }
        if (pa) return fail(PFB_ERR_BAD_ARG, "fused primal pass: not available for this plan");
        const size_t smr = mx_smem_bytes(h.M, h.R, sizeof(Cx<T>));
        const int tr = mixed_threads((long)h.M * h.R / 4);
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessm_rows_inv<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smr));
        launch(&k_hessm_rows_inv<T>, dim3(cdiv(h.nx, h.R), h.nband), dim3(tr), smr, st, (const Cx<T> *)S, out,
               (const Cx<T> *)p->mtwM, (const Cx<T> *)p->twr, (const Cx<T> *)p->mroots, p->mrow, h, ep);
        PFB_CHECK_LAUNCH();
        return PFB_OK;
    }
    if (p->fast && aligned) return fast_rows_inv_d<T>(p, out, ep, st, pa, out2);
    if (p->fastrows && sizeof(T) == 4 && aligned) {
        if constexpr (sizeof(T) == 4) return fast_rows_inv_8k(p, out, ep, st, pa, out2);
    }
    if (pa) return fail(PFB_ERR_BAD_ARG, "fused primal pass: not available for this plan");
    HessDims h = h0;
    h.R = p->Rrow;
    size_t smem = sizeof(Cx<T>) * (size_t)fft_smem_elems(h.M) * h.R;
    int nthr = p->drow.NT * h.R;
    PFB_BY_THREADS(nthr, k_hess_rows_inv, smem, dim3(cdiv(h.nx, h.R), h.nband), st,
                   (const Cx<T> *)S, out, tw, p->drow, h, ep);
    return PFB_OK;
}

template <class T>
static int hess_apply_t(pfb_hess_plan *p, const T *x, const T *beam_in, const T *beam_out,
                        const T *addx, double sigmainv, int has_wsum, double wsum, T *out,
                        cudaStream_t st) {
    // the fast row passes use 2-element vector accesses on x / out / beam
    const uintptr_t al = 2 * sizeof(T) - 1;
    const uintptr_t alt = (uintptr_t)out | (uintptr_t)beam_out | (uintptr_t)addx;
    const bool aligned = !(((uintptr_t)x | (uintptr_t)beam_in | alt) & al);
    Epilogue<T> ep;
    ep.beam = beam_out;
    ep.x = addx;
    ep.wsum = (T)wsum;
    ep.sigmainv = (T)sigmainv;
    ep.has_wsum = has_wsum;
    if (int rc = hess_fwd_cols_t<T>(p, x, beam_in, alt, st)) return rc;
    return hess_rows_inv_t<T>(p, out, ep, aligned, st);
}

template <class T>
static int primal_end_t(pfb_hess_plan *p, const void *xp, const void *w, const void *wprev, const void *data,
                        double tau, int positivity, void *x, double *out2, cudaStream_t st) {
    PrimalArgsT<T> pa;
    pa.xp = (const T *)xp; pa.w = (const T *)w; pa.wprev = (const T *)wprev;
    pa.data = (const T *)data; pa.x = (T *)x; pa.tau = (T)tau; pa.positivity = positivity;
    pa.part = nullptr;
    return hess_rows_inv_t<T>(p, (T *)nullptr, Epilogue<T>(), true, st, &pa, out2);
}

// lines per CTA: fill up to 1024 threads / kMaxSmem, capped at `want`
static int lines_per_cta(int NT, int N, size_t csize, int want) {
    int r = want;
    while (r > 1 && (NT * r > 1024 || csize * fft_smem_elems(N) * r > kMaxSmem)) r >>= 1;
    return r;
}

// Plan of the mixed-radix path (hess_mixed.cuh) for padded sizes the power-of-two
// kernels do not take.
static int hess_plan_create_mixed(pfb_hess_plan **out, int nband, int nx, int ny, int nxp, int nyp,
                                  int dtype) {
    const size_t cs = 2 * dtype_size(dtype);
    const int M = nyp / 2, H = nxp / 2;
    const bool split = (nxp % 2 == 0) && nx <= H;
    const int Nc = split ? H : nxp;
    MxDesc mrow, mcol;
    if (nyp % 2 || M < 1)
        return fail(PFB_ERR_FFT_SIZE, "padded size (%d,%d) unsupported: nyp (lastsize) must be even", nxp, nyp);
    if (!mx_make_desc(M, mrow) || !mx_make_desc(Nc, mcol))
        return fail(PFB_ERR_FFT_SIZE,
                    "padded size (%d,%d) unsupported: nyp/2 and %s must factor into 2, 3, 5, 7, 11, 13",
                    nxp, nyp, split ? "nxp/2" : "nxp");
    if (mx_smem_bytes(M, 1, cs) > kMaxSmemFast || mx_smem_bytes(Nc, 1, cs) > kMaxSmemFast ||
        mx_tw_entries(M) > kMxTwMax || mx_tw_entries(Nc) > kMxTwMax)
        return fail(PFB_ERR_FFT_SIZE,
                    "padded size (%d,%d) unsupported: two transform lines (%d / %d complex) exceed %zu KB "
                    "of shared memory", nxp, nyp, M, Nc, kMaxSmemFast / 1024);
    pfb_hess_plan *p = new pfb_hess_plan();
    memset(p, 0, sizeof *p);
    p->dtype = dtype;
    HessDims &h = p->h;
    h.nband = nband; h.nx = nx; h.ny = ny; h.nxp = nxp; h.nyp = nyp;
    h.nyo2 = nyp / 2 + 1; h.M = M; h.twN = 0;
    p->mixed = 1;
    p->split = split ? 1 : 0;
    p->deint = p->split;
    p->mrow = mrow;
    p->mcol = mcol;
    p->nsm = kNumSM;
    // rows per CTA: transposed scratch written in 32-byte runs when shared memory allows
    p->Lrow = dtype == PFB_F32 ? 4 : 2;
    while (p->Lrow > 1 && mx_smem_bytes(M, p->Lrow, cs) > kMaxSmemFast) p->Lrow >>= 1;
    auto upload = [&](void **dst, int n, auto gen) -> bool {
        std::vector<char> hbuf(cs * (size_t)n);
        for (int i = 0; i < n; ++i) {
            const long double a = gen(i);
            const double c = (double)cosl(a), sn = (double)sinl(a);
            if (dtype == PFB_F32) { ((float *)hbuf.data())[2 * i] = (float)c; ((float *)hbuf.data())[2 * i + 1] = (float)sn; }
            else { ((double *)hbuf.data())[2 * i] = c; ((double *)hbuf.data())[2 * i + 1] = sn; }
        }
        return cudaMalloc(dst, hbuf.size()) == cudaSuccess &&
               cudaMemcpy(*dst, hbuf.data(), hbuf.size(), cudaMemcpyHostToDevice) == cudaSuccess;
    };
    const long double tau = -2.0L * 3.14159265358979323846264338327950288L;
    auto lin = [&](int N) { return [=](int i) { return tau * (long double)i / (long double)N; }; };
    auto root = [&](int i) {
        const int R = i < 3 ? 3 : i < 8 ? 5 : i < 15 ? 7 : i < 26 ? 11 : 13;
        return tau * (long double)(i - mx_root_off(R)) / (long double)R;
    };
    const size_t sS = cs * (size_t)nband * h.nyo2 * nx, sP = cs * (size_t)nband * h.nyo2 * nxp;
    auto two = [&](int N) { return [=](int i) { return tau * (long double)mx_tw_index(i) / (long double)N; }; };
    bool good = upload(&p->mtwM, mx_tw_entries(M), two(M)) && upload(&p->mtwC, mx_tw_entries(Nc), two(Nc)) &&
                upload(&p->twr, M + 1, lin(nyp)) && upload(&p->modg, std::max(H, 1), lin(nxp)) &&
                upload(&p->mroots, kMxRoots, root) &&
                cudaMalloc(&p->S, sS) == cudaSuccess;
    // passes whose length is (3, 5 or 7) x a power of two take the peeled kernels
    if (good && !getenv("PFB_HESS_NOPEEL")) {
        auto table = [&](void **dst, int P2) -> bool {
            std::vector<char> hbuf(cs * (size_t)ct_tab_size(P2));
            ct_fill_table(P2, [&](int i, double c, double sn) {
                if (dtype == PFB_F32) { ((float *)hbuf.data())[2 * i] = (float)c; ((float *)hbuf.data())[2 * i + 1] = (float)sn; }
                else { ((double *)hbuf.data())[2 * i] = c; ((double *)hbuf.data())[2 * i + 1] = sn; }
            });
            return cudaMalloc(dst, hbuf.size()) == cudaSuccess &&
                   cudaMemcpy(*dst, hbuf.data(), hbuf.size(), cudaMemcpyHostToDevice) == cudaSuccess;
        };
        PeelDesc pr, pc;
        const bool inpl = peel_make_desc(M, pr) && peel_can_inplace(pr.m, pr.P2) && !getenv("PFB_HESS_PEEL_NOINPL");
        if (pr.m && cs * peel_smem_elems(M, pr.m, pr.P2, 1, inpl) <= kMaxSmemFast) {
            p->inplR = inpl ? 1 : 0;
            p->LrowP = dtype == PFB_F32 ? 4 : 2;
            while (p->LrowP > 1 && cs * peel_smem_elems(M, pr.m, pr.P2, p->LrowP, inpl) > kMaxSmemFast) p->LrowP >>= 1;
            p->prow = pr;
            good = good && table(&p->ptabR, pr.P2);
            if (dtype == PFB_F32)    // eps partials of the fused primal epilogue, 2 per CTA of pass C
                good = good && cudaMalloc((void **)&p->pd_partP, sizeof(double) * 2 * (size_t)cdiv(nx, p->LrowP) * nband) == cudaSuccess;
        }
        const bool c2 = peel_make_desc(Nc, pc) && peel_can_inplace(pc.m, pc.P2) && !getenv("PFB_HESS_PEEL_COLS1") &&
                        cs * peel_smem_elems_cols2(Nc, pc.m, pc.P2, split) <= kMaxSmemFast;
        if (pc.m && (c2 || cs * peel_smem_elems(Nc, pc.m, pc.P2, 1) <= kMaxSmemFast)) {
            p->pcol = pc;
            p->pcol2 = c2 ? 1 : 0;
            good = good && table(&p->ptabC, pc.P2);
        }
    }
    if (!good) {
        pfb_hess_plan_destroy(p);
        return fail(PFB_ERR_ALLOC, "hessian plan (mixed radix) allocation failed (%zu + %zu bytes)", sS, sP);
    }
    *out = p;
    return PFB_OK;
}

// psfhat = r2c(ifftshift(psf)) on a plan created with nx = nxp, ny = nyp (all rows / columns live)
template <class T>
static int psfhat_from_psf_t(pfb_hess_plan *p, const T *psf, Cx<T> *out, cudaStream_t st) {
    const HessDims &h0 = p->h;
    Cx<T> *S = (Cx<T> *)p->S;
    if (p->mixed) {
        HessDims h = h0;
        h.R = p->Lrow;
        const size_t cs = sizeof(Cx<T>);
        const size_t smr = mx_smem_bytes(h.M, h.R, cs), smc = mx_smem_bytes(p->mcol.N, 1, cs);
        auto nthr = [](long work) { return work >= 2048 ? 512 : work >= 512 ? 256 : 128; };
        const int tr = nthr((long)h.M * h.R / 4), tc = nthr(p->mcol.N / 4);
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessm_rows_fwd<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smr));
        launch(&k_hessm_rows_fwd<T>, dim3(cdiv(h.nx, h.R), h.nband), dim3(tr), smr, st, psf, (const T *)nullptr, S,
               (const Cx<T> *)p->mtwM, (const Cx<T> *)p->twr, (const Cx<T> *)p->mroots, p->mrow, h);
        PFB_CHECK_LAUNCH();
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_psfm_cols_fwd<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smc));
        launch(&k_psfm_cols_fwd<T>, dim3(h.nyo2, h.nband), dim3(tc), smc, st, S, (const Cx<T> *)p->mtwC,
               (const Cx<T> *)p->mroots, p->mcol, h);
        PFB_CHECK_LAUNCH();
    } else {
        const Cx<T> *tw = (const Cx<T> *)p->tw;
        {
            HessDims h = h0;
            h.R = p->Rrow;
            size_t smem = sizeof(Cx<T>) * (size_t)fft_smem_elems(h.M) * h.R;
            int nthr = p->drow.NT * h.R;
            PFB_BY_THREADS(nthr, k_hess_rows_fwd, smem, dim3(cdiv(h.nx, h.R), h.nband), st,
                           psf, (const T *)nullptr, S, tw, p->drow, h);
        }
        {
            HessDims h = h0;
            h.R = p->Rcol;
            size_t smem = sizeof(Cx<T>) * (size_t)fft_smem_elems(h.nxp) * h.R;
            int nthr = p->dcol.NT * h.R;
            PFB_BY_THREADS(nthr, k_psf_cols_fwd, smem, dim3(cdiv(h.nyo2, h.R), h.nband), st, S, tw, p->dcol, h);
        }
    }
    // S (nband, nyo2, nxp) -> psfhat (nband, nxp, nyo2): the transposing kernel with the roles swapped
    dim3 grid(cdiv(h0.nxp, 32), cdiv(h0.nyo2, 32), h0.nband), block(32, 8);
    launch(&k_psfhat_prepare<T>, grid, block, 0, st, (const Cx<T> *)S, out, h0.nyo2, h0.nxp, (T)1, 0, 0, 0);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

}  // namespace pfb

using namespace pfb;

extern "C" {

int pfb_hess_plan_create(pfb_hess_plan **out, int nband, int nx, int ny, int nxp, int nyp,
                         int dtype) {
    if (!out || nband < 1 || nx < 1 || ny < 1 || nxp < nx || nyp < ny ||
        (dtype != PFB_F32 && dtype != PFB_F64))
        return fail(PFB_ERR_BAD_ARG, "pfb_hess_plan_create: bad argument");
    size_t cs = 2 * dtype_size(dtype);
    int M = nyp / 2;
    bool ok = (nyp % 2 == 0) && M >= 16 && !(M & (M - 1)) && nxp >= 16 && !(nxp & (nxp - 1)) &&
              M / 16 <= 1024 && nxp / 16 <= 1024 &&
              cs * fft_smem_elems(M) <= kMaxSmem && cs * fft_smem_elems(nxp) <= kMaxSmem;
    if (!ok) return hess_plan_create_mixed(out, nband, nx, ny, nxp, nyp, dtype);
    pfb_hess_plan *p = new pfb_hess_plan();
    memset(p, 0, sizeof *p);
    p->dtype = dtype;
    HessDims &h = p->h;
    h.nband = nband; h.nx = nx; h.ny = ny; h.nxp = nxp; h.nyp = nyp;
    h.nyo2 = nyp / 2 + 1; h.M = M;
    h.twN = std::max(nxp, nyp);
    fft_make_desc(M, h.twN, p->drow);
    fft_make_desc(nxp, h.twN, p->dcol);
    // rows: R lines so that the transposed scratch is written in >= 32 B runs
    p->Rrow = lines_per_cta(p->drow.NT, M, cs, dtype == PFB_F32 ? 4 : 2);
    if (p->drow.NT * p->Rrow < 128) p->Rrow = lines_per_cta(p->drow.NT, M, cs, 128 / p->drow.NT);
    p->Rcol = lines_per_cta(p->dcol.NT, nxp, cs, std::max(1, 256 / p->dcol.NT));
    std::vector<char> twh(cs * h.twN);
    for (int m = 0; m < h.twN; ++m) {
        long double a = -2.0L * 3.14159265358979323846264338327950288L * (long double)m / (long double)h.twN;
        double c = (double)cosl(a), s = (double)sinl(a);
        if (dtype == PFB_F32) { ((float *)twh.data())[2 * m] = (float)c; ((float *)twh.data())[2 * m + 1] = (float)s; }
        else { ((double *)twh.data())[2 * m] = c; ((double *)twh.data())[2 * m + 1] = s; }
    }
    size_t sS = cs * (size_t)nband * h.nyo2 * nx, sP = cs * (size_t)nband * h.nyo2 * nxp;
    if (cudaMalloc(&p->tw, twh.size()) != cudaSuccess || cudaMalloc(&p->S, sS) != cudaSuccess) {
        if (p->tw) cudaFree(p->tw);
        if (p->S) cudaFree(p->S);
        delete p;
        return fail(PFB_ERR_ALLOC, "hessian plan allocation failed (%zu + %zu bytes)", sS, sP);
    }
    if (cudaMemcpy(p->tw, twh.data(), twh.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaFree(p->tw); cudaFree(p->S);
        delete p;
        return fail(PFB_ERR_CUDA, "twiddle upload failed");
    }
    // ---- fast path set-up (hess_fast.cuh)
    {
        int H = nxp / 2;
        auto okN = [](int n) { return n == 512 || n == 1024 || n == 2048 || n == 4096; };
        p->nsm = kNumSM;
#ifndef PFB_EMU
        int dev = 0, nsm = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && nsm > 0)
            p->nsm = nsm;
#endif
        bool fast = okN(M) && okN(H) && nx <= H && ny <= M && (ny % 2 == 0) && !getenv("PFB_HESS_GENERIC");
        if (fast) {
            p->RB = 4;
            while (p->RB > 2 && cs * ((size_t)p->RB * ct_stride(M) + ct_tab_size(M)) > kMaxSmemFast) p->RB >>= 1;
            if (cs * ((size_t)p->RB * ct_stride(M) + ct_tab_size(M)) > kMaxSmemFast ||
                cs * ((size_t)2 * ct_stride(H) + ct_tab_size(H)) > kMaxSmemFast)
                fast = false;
        }
        if (fast) {
            auto upload = [&](void **dst, int n, auto gen) -> bool {
                std::vector<char> hbuf(cs * (size_t)n);
                gen([&](int i, double c, double sn) {
                    if (dtype == PFB_F32) { ((float *)hbuf.data())[2 * i] = (float)c; ((float *)hbuf.data())[2 * i + 1] = (float)sn; }
                    else { ((double *)hbuf.data())[2 * i] = c; ((double *)hbuf.data())[2 * i + 1] = sn; }
                });
                return cudaMalloc(dst, hbuf.size()) == cudaSuccess &&
                       cudaMemcpy(*dst, hbuf.data(), hbuf.size(), cudaMemcpyHostToDevice) == cudaSuccess;
            };
            auto lin = [](int n, int N) {
                return [n, N](auto put) {
                    for (int i = 0; i < n; ++i) {
                        long double a = -2.0L * 3.14159265358979323846264338327950288L * (long double)i / (long double)N;
                        put(i, (double)cosl(a), (double)sinl(a));
                    }
                };
            };
            bool ok2 = upload(&p->tabM, ct_tab_size(M), [&](auto put) { ct_fill_table(M, put); }) &&
                       upload(&p->tabH, ct_tab_size(H), [&](auto put) { ct_fill_table(H, put); }) &&
                       upload(&p->twr, M + 1, lin(M + 1, nyp)) && upload(&p->modg, H, lin(H, nxp));
            ok2 = ok2 && cudaMalloc((void **)&p->ctr, 3 * sizeof(int)) == cudaSuccess;
            if (!ok2) {
                pfb_hess_plan_destroy(p);
                return fail(PFB_ERR_ALLOC, "hessian plan (fast path tables) allocation failed");
            }
        }
        p->fast = fast ? 1 : 0;
        // config 4 (16384^2 PSF, fp32): the column pass -- 57 % of the generic Hessian there --
        // runs on the one-group compile-time kernel (one 512-thread CTA per SM); the row passes
        // stay on the generic kernels (a 16384-point row pair does not fit the fast row kernels)
        bool fastcols = !fast && dtype == PFB_F32 && H == 8192 && nxp == 2 * H && nx <= H &&
                        !getenv("PFB_HESS_GENERIC") &&
                        cs * ((size_t)2 * ct_stride(8192) + ct_tab_size(8192)) <= kMaxSmemFast;
        if (fastcols) {
            std::vector<float> tb(2 * (size_t)ct_tab_size(8192)), mg(2 * (size_t)H);
            ct_fill_table(8192, [&](int i, double c, double sn) { tb[2 * i] = (float)c; tb[2 * i + 1] = (float)sn; });
            for (int i = 0; i < H; ++i) {
                long double a = -2.0L * 3.14159265358979323846264338327950288L * (long double)i / (long double)nxp;
                mg[2 * i] = (float)cosl(a); mg[2 * i + 1] = (float)sinl(a);
            }
            bool ok3 = cudaMalloc(&p->tabH, tb.size() * 4) == cudaSuccess &&
                       cudaMemcpy(p->tabH, tb.data(), tb.size() * 4, cudaMemcpyHostToDevice) == cudaSuccess &&
                       cudaMalloc(&p->modg, mg.size() * 4) == cudaSuccess &&
                       cudaMemcpy(p->modg, mg.data(), mg.size() * 4, cudaMemcpyHostToDevice) == cudaSuccess &&
                       cudaMalloc((void **)&p->ctr, 3 * sizeof(int)) == cudaSuccess;
            if (!ok3) {
                pfb_hess_plan_destroy(p);
                return fail(PFB_ERR_ALLOC, "hessian plan (column-pass tables) allocation failed");
            }
        }
        p->fastcols = fastcols ? 1 : 0;
        p->deint = p->fast | p->fastcols;
        // same for the row passes of a 16384-wide PSF: one-group row kernels, two rows per batch
        bool fastrows = !fast && dtype == PFB_F32 && M == 8192 && ny <= M && ny % 2 == 0 &&
                        !getenv("PFB_HESS_GENERIC") &&
                        cs * ((size_t)2 * ct_stride(8192) + ct_tab_size(8192)) <= kMaxSmemFast;
        if (fastrows) {
            std::vector<float> tb(2 * (size_t)ct_tab_size(8192)), tr(2 * (size_t)(M + 1));
            ct_fill_table(8192, [&](int i, double c, double sn) { tb[2 * i] = (float)c; tb[2 * i + 1] = (float)sn; });
            for (int i = 0; i <= M; ++i) {
                long double a = -2.0L * 3.14159265358979323846264338327950288L * (long double)i / (long double)nyp;
                tr[2 * i] = (float)cosl(a); tr[2 * i + 1] = (float)sinl(a);
            }
            bool ok4 = cudaMalloc(&p->tabM, tb.size() * 4) == cudaSuccess &&
                       cudaMemcpy(p->tabM, tb.data(), tb.size() * 4, cudaMemcpyHostToDevice) == cudaSuccess &&
                       cudaMalloc(&p->twr, tr.size() * 4) == cudaSuccess &&
                       cudaMemcpy(p->twr, tr.data(), tr.size() * 4, cudaMemcpyHostToDevice) == cudaSuccess &&
                       (p->ctr || cudaMalloc((void **)&p->ctr, 3 * sizeof(int)) == cudaSuccess);
            if (!ok4) {
                pfb_hess_plan_destroy(p);
                return fail(PFB_ERR_ALLOC, "hessian plan (row-pass tables) allocation failed");
            }
        }
        p->fastrows = fastrows ? 1 : 0;
#ifndef PFB_EMU
        p->have_tma = (fast && dtype == PFB_F32 && nx % 2 == 0 && make_tma_S(p)) ? 1 : 0;
#endif
        // eps partials of the fused primal epilogue (every fast pass C kernel takes it)
        if ((p->fast || p->fastrows) &&
            cudaMalloc((void **)&p->pd_part, sizeof(double) * 2 * 4 * p->nsm) != cudaSuccess) {
            cudaGetLastError();
            p->pd_part = nullptr;
        }
        // Dynamic distribution costs ~5 % when the GPU is otherwise idle (measured on the
        // column pass) and wins when a concurrent kernel holds SMs (NCCL all-reduce of
        // the band-sharded solver): off by default, the sharded solver turns it on.
        p->dyn = getenv("PFB_HESS_DYNAMIC") ? atoi(getenv("PFB_HESS_DYNAMIC")) : 0;
    }
    *out = p;
    return PFB_OK;
}

int pfb_hess_plan_destroy(pfb_hess_plan *p) {
    if (!p) return PFB_OK;
    if (p->pd_partP) cudaFree(p->pd_partP);
    if (p->ptabR) cudaFree(p->ptabR);
    if (p->ptabC) cudaFree(p->ptabC);
    if (p->tabM) cudaFree(p->tabM);
    if (p->tabH) cudaFree(p->tabH);
    if (p->twr) cudaFree(p->twr);
    if (p->modg) cudaFree(p->modg);
    if (p->ctr) cudaFree(p->ctr);
    if (p->pd_part) cudaFree(p->pd_part);
    if (p->mtwM) cudaFree(p->mtwM);
    if (p->mtwC) cudaFree(p->mtwC);
    if (p->mroots) cudaFree(p->mroots);
    if (p->tw) cudaFree(p->tw);
    if (p->S) cudaFree(p->S);
    if (p->psfT) cudaFree(p->psfT);
    delete p;
    return PFB_OK;
}

int pfb_hess_plan_set_dynamic(pfb_hess_plan *p, int on) {
    if (!p) return fail(PFB_ERR_BAD_ARG, "pfb_hess_plan_set_dynamic: null plan");
    p->dyn = on ? 1 : 0;
    return PFB_OK;
}

int pfb_hess_plan_set_psfhat(pfb_hess_plan *p, const void *psfhat, void *stream) {
    if (!p || !psfhat) return fail(PFB_ERR_BAD_ARG, "pfb_hess_plan_set_psfhat: null argument");
    if (p->dtype == PFB_F32) return hess_set_psf_t<float>(p, psfhat, as_stream(stream));
    return hess_set_psf_t<double>(p, psfhat, as_stream(stream));
}

int pfb_hess_apply(pfb_hess_plan *p, const void *x, const void *beam, int add_x,
                   double sigmainv, int has_wsum, double wsum, void *out, void *stream) {
    if (!p || !x || !out) return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply: null argument");
    if (!p->have_psf) return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply: psfhat not set");
    if (x == out) return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply: out must not alias x");
    cudaStream_t st = as_stream(stream);
    if (p->dtype == PFB_F32)
        return hess_apply_t<float>(p, (const float *)x, (const float *)beam, (const float *)beam,
                                   add_x ? (const float *)x : nullptr, sigmainv, has_wsum, wsum,
                                   (float *)out, st);
    return hess_apply_t<double>(p, (const double *)x, (const double *)beam, (const double *)beam,
                                add_x ? (const double *)x : nullptr, sigmainv, has_wsum, wsum,
                                (double *)out, st);
}

int pfb_psfhat_from_psf(const void *psf, int nband, int nxp, int nyp, int dtype, void *psfhat,
                        void *stream) {
    if (!psf || !psfhat || nband < 1 || nxp < 2 || nyp < 2 || (dtype != PFB_F32 && dtype != PFB_F64))
        return fail(PFB_ERR_BAD_ARG, "pfb_psfhat_from_psf: bad argument");
    if ((nxp | nyp) & 1)
        return fail(PFB_ERR_FFT_SIZE, "pfb_psfhat_from_psf: padded sizes (%d,%d) must be even", nxp, nyp);
    pfb_hess_plan *p = nullptr;
    if (int rc = pfb_hess_plan_create(&p, nband, nxp, nyp, nxp, nyp, dtype)) return rc;
    cudaStream_t st = as_stream(stream);
    int rc = dtype == PFB_F32 ? psfhat_from_psf_t<float>(p, (const float *)psf, (Cx<float> *)psfhat, st)
                              : psfhat_from_psf_t<double>(p, (const double *)psf, (Cx<double> *)psfhat, st);
    // the plan's scratch is released here, so the stream is drained first
    if (rc == PFB_OK && cudaStreamSynchronize(st) != cudaSuccess)
        rc = fail(PFB_ERR_CUDA, "pfb_psfhat_from_psf: stream synchronise failed");
    pfb_hess_plan_destroy(p);
    return rc;
}

int pfb_hess_plan_kernels(const pfb_hess_plan *p) {
    if (!p) return 0;
    return (p->fast ? 1 : 0) | (p->mixed ? 2 : 0) | (p->prow.m ? 4 : 0) | (p->pcol.m ? 8 : 0) |
           (p->fastcols ? 16 : 0) | (p->fastrows ? 32 : 0);
}

int pfb_hess_plan_can_fuse_primal(const pfb_hess_plan *p) {
    if (!p || getenv("PFB_HESS_NOFUSE")) return 0;
    if (p->mixed) {                              // peeled pass C (hess_peel.cuh), fp32
#ifndef PFB_EMU
        return (p->prow.m && p->pd_partP && p->dtype == PFB_F32) ? 1 : 0;
#else
        return 0;
#endif
    }
    // power-of-two fast paths: every pass C kernel of hess_fast.cuh takes the primal epilogue.  On the
    // 16384-point row kernel it measured 0.1 ms slower than the separate primal kernel (config 4 on
    // 8 GPUs: 1.94 vs 1.36 + 0.48 ms), so there it is opt-in.
    if (p->fastrows && !p->fast) return (p->pd_part && getenv("PFB_HESS_FUSE_ROWS16K")) ? 1 : 0;
    return (p->fast && p->pd_part) ? 1 : 0;
}

static int primal_args_ok(const pfb_hess_plan *p, const char *who) {
    if (!p) return fail(PFB_ERR_BAD_ARG, "%s: bad argument", who);
    if (!p->have_psf) return fail(PFB_ERR_BAD_ARG, "%s: psfhat not set", who);
    if (!pfb_hess_plan_can_fuse_primal(p))
        return fail(PFB_ERR_BAD_ARG, "%s: not available for this plan (see pfb_hess_plan_can_fuse_primal)", who);
    return PFB_OK;
}

int pfb_hess_apply_primal_begin(pfb_hess_plan *p, const void *xp, void *stream) {
    if (int rc = primal_args_ok(p, "pfb_hess_apply_primal_begin")) return rc;
    const size_t al = 2 * dtype_size(p->dtype) - 1;
    if (!xp || ((uintptr_t)xp & al)) return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply_primal_begin: bad argument");
    cudaStream_t st = as_stream(stream);
    if (p->dtype == PFB_F32) return hess_fwd_cols_t<float>(p, (const float *)xp, nullptr, 0, st);
    return hess_fwd_cols_t<double>(p, (const double *)xp, nullptr, 0, st);
}

int pfb_hess_apply_primal_end(pfb_hess_plan *p, const void *xp, const void *w, const void *wprev,
                              const void *data, double tau, int positivity, void *x, double *out2,
                              void *stream) {
    if (int rc = primal_args_ok(p, "pfb_hess_apply_primal_end")) return rc;
    if (!xp || !w || !wprev || !data || !x || !out2 || (positivity != 0 && positivity != 1))
        return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply_primal_end: bad argument");
    if (x == xp) return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply_primal_end: x must not alias xp");
    const size_t al = 2 * dtype_size(p->dtype) - 1;
    if (((uintptr_t)xp | (uintptr_t)w | (uintptr_t)wprev | (uintptr_t)data | (uintptr_t)x) & al)
        return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply_primal_end: arrays must be aligned to two elements");
    if (p->h.ny % 2) return fail(PFB_ERR_BAD_ARG, "pfb_hess_apply_primal_end: ny must be even");
    cudaStream_t st = as_stream(stream);
    if (p->dtype == PFB_F32) return primal_end_t<float>(p, xp, w, wprev, data, tau, positivity, x, out2, st);
    return primal_end_t<double>(p, xp, w, wprev, data, tau, positivity, x, out2, st);
}

int pfb_hess_apply_primal(pfb_hess_plan *p, const void *xp, const void *w, const void *wprev,
                          const void *data, double tau, int positivity, void *x, double *out2,
                          void *stream) {
    if (int rc = pfb_hess_apply_primal_begin(p, xp, stream)) return rc;
    return pfb_hess_apply_primal_end(p, xp, w, wprev, data, tau, positivity, x, out2, stream);
}

}  // extern "C"
