// The Hessian plan object and the launch interface of the peeled kernels.  hess.cu (plan
// life cycle, power-of-two and run-time radix launch sites, the C-ABI) and hess_peel.cu (launch
// sites of hess_peel.cuh) are separate translation units only so that their many template
// instantiations compile in parallel.
#pragma once
#include "hess_fast.cuh"
#include "hess_mixed.cuh"
#include "hess_peel.cuh"

struct pfb_hess_plan {
    int dtype;
    pfb::HessDims h;
    pfb::FftDesc drow, dcol;
    int Rrow, Rcol;          // lines per CTA in the row / column passes
    void *tw;                // twN complex twiddles exp(-2 pi i m / twN)
    void *S;                 // (nband, nyo2, nx) complex scratch
    void *psfT;              // (nband, nyo2, nxp) complex, scaled
    int have_psf;
    // fast path (hess_fast.cuh): nxp = 2H >= 2nx, nyp = 2M >= 2ny, H, M in [512, 4096]
    int fast, RB, nsm;
    int fastcols;            // fp32, nxp = 16384: pass B on k_hessf_cols1<8192>
    int fastrows;            // fp32, nyp = 16384: passes A / C on the one-group row kernels (M = 8192)
    void *tabM, *tabH;       // ct_fft stage tables for lengths M and H
    void *twr;               // exp(-2 pi i k / nyp), k = 0..M
    void *modg;              // exp(-2 pi i n / nxp), n = 0..H-1
    int *ctr;                // work counters of the three persistent kernels (ItemQueue)
    int dyn;                 // hand items out dynamically (see pfb_hess_set_dynamic)
    // mixed-radix path (hess_mixed.cuh): padded sizes that are not powers of two
    int mixed, split, Lrow;  // split: even/odd pruned column pass (nxp = 2H, nx <= H)
    pfb::MxDesc mrow, mcol;
    void *mtwM, *mtwC, *mroots;   // length-M / length-(H or nxp) twiddle tables, odd-radix roots
    int deint;               // psfT rows stored [even | odd] per column
    // odd factor x power of two (hess_peel.cuh): passes whose length is m * P2 leave the
    // run-time radix list for m compile-time transforms of length P2
    pfb::PeelDesc prow, pcol;
    int LrowP, inplR;        // rows per CTA of the peeled row passes; lines transformed in place
    void *ptabR, *ptabC;     // ct_fft stage tables of prow.P2 / pcol.P2
    double *pd_partP;        // partial eps sums of the peeled pass C's primal epilogue (2 per CTA)
    int pcol2;               // column pass on k_hessp_cols2 (psfT stored permuted [q][k] per half)
#ifndef PFB_EMU
    CUtensorMap tmS;         // S as (band, k, row) tensor of 8-byte elements, box 2 rows x 256 k (fp32 fast path)
#endif
    int have_tma;
    double *pd_part;         // partial eps sums of the fused primal epilogue (2 per CTA)
};

namespace pfb {

static const size_t kMaxSmem = 200 * 1024;
static const size_t kMaxSmemFast = 224 * 1024;

// grid of a persistent kernel: one wave of resident CTAs, at most one CTA per work item
static inline int persistent_grid(const void *func, int threads, size_t smem, int total, int nsm) {
#ifdef PFB_EMU
    (void)func; (void)threads; (void)smem; (void)nsm;
    return total < 2 ? total : 2;
#else
    int per = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, func, threads, smem) != cudaSuccess || per < 1)
        per = 1;
    long g = (long)per * nsm;
    return (int)(g < total ? g : total);
#endif
}

#define PFB_SIZE_SWITCH(n, CALL)                         \
    switch (n) {                                         \
    case 512: return CALL(512);                          \
    case 1024: return CALL(1024);                        \
    case 2048: return CALL(2048);                        \
    case 4096: return CALL(4096);                        \
    default: return fail(PFB_ERR_FFT_SIZE, "fast path: unsupported length %d", (int)(n)); \
    }

// launch sites of pass C of the power-of-two fast paths (hess_rows_inv.cu): TMA, pipelined and
// basic kernels and the 16384-point row kernel; pa != nullptr: the primal step is the epilogue
// (PrimalArgsT), its eps terms go to out2
template <class T> int fast_rows_inv_d(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, cudaStream_t st,
                                       const PrimalArgsT<T> *pa = nullptr, double *out2 = nullptr);
int fast_rows_inv_8k(pfb_hess_plan *p, float *out, const Epilogue<float> &ep, cudaStream_t st,
                     const PrimalArgsT<float> *pa = nullptr, double *out2 = nullptr);

// launch sites of the peeled kernels (hess_peel.cu); T = float, double
template <class T> int peel_rows_fwd_d(pfb_hess_plan *p, const T *x, const T *beam, cudaStream_t st);
template <class T> int peel_rows_inv_d(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, cudaStream_t st);
template <class T> int peel_cols_d(pfb_hess_plan *p, cudaStream_t st);
#ifndef PFB_EMU
int peel_rows_inv_primal_d(pfb_hess_plan *p, const PrimalArgs &pa, double *out2, cudaStream_t st);
#endif

}  // namespace pfb
