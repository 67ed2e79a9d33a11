// Launch sites of pass C (rows inverse + epilogue) of the power-of-two fast paths; see
// hess_plan.cuh for why this is its own translation unit.
#include "hess_plan.cuh"

namespace pfb {

// eps terms of a fused primal pass: fixed-order sum of the CTAs' partials
static int primal_finish(pfb_hess_plan *p, int grid, double *out2, cudaStream_t st) {
    launch(&k_hess_red_final, dim3(1), dim3(32), 0, st, (const double *)p->pd_part, grid, out2);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

// pa != nullptr: the primal step is the epilogue (PrimalArgsT), eps terms -> out2
template <class T, int M, bool TWS>
static int fast_rows_inv2(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, cudaStream_t st, size_t smem,
                          const PrimalArgsT<T> *pa = nullptr, double *out2 = nullptr) {
    const HessDims &h = p->h;
    auto kern = pa ? (p->dyn ? &k_hessf_rows_inv2<T, M, TWS, true, true> : &k_hessf_rows_inv2<T, M, TWS, false, true>)
                   : (p->dyn ? &k_hessf_rows_inv2<T, M, TWS, true, false> : &k_hessf_rows_inv2<T, M, TWS, false, false>);
    PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int nbb = cdiv(h.nx, 4);
    int grid = persistent_grid((const void *)kern, M / 8, smem, nbb * h.nband, p->nsm);
    if (grid > 4 * p->nsm) grid = 4 * p->nsm;      // size of pd_part
    PrimalArgsT<T> a = pa ? *pa : PrimalArgsT<T>();
    a.part = p->pd_part;
    launch(kern, dim3(grid), dim3(M / 8), smem, st, (const Cx<T> *)p->S, out, (const Cx<T> *)p->tabM,
           (const Cx<T> *)p->twr, h, nbb, ep, p->ctr + 2, a);
    PFB_CHECK_LAUNCH();
    return pa ? primal_finish(p, grid, out2, st) : PFB_OK;
}

#ifndef PFB_EMU
template <int M, bool PRIMAL>
static int fast_rows_inv_tma(pfb_hess_plan *p, float *out, const Epilogue<float> &ep, PrimalArgs pa,
                             double *out2, cudaStream_t st) {
    const HessDims &h = p->h;
    const size_t smem = sizeof(Cx<float>) * ((size_t)2 * tma_stage_elems<M>() + ct_tab_size(M) + M) + 16 + 128;
    auto kern = p->dyn ? &k_hessf_rows_inv_tma<M, true, PRIMAL> : &k_hessf_rows_inv_tma<M, false, PRIMAL>;
    PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int nbb = cdiv(h.nx, 4);
    int grid = persistent_grid((const void *)kern, M / 8, smem, nbb * h.nband, p->nsm);
    if (grid > 4 * p->nsm) grid = 4 * p->nsm;      // size of pd_part
    pa.part = p->pd_part;
    launch(kern, dim3(grid), dim3(M / 8), smem, st, p->tmS, out, (const Cx<float> *)p->tabM,
           (const Cx<float> *)p->twr, h, nbb, ep, pa, p->ctr + 2);
    PFB_CHECK_LAUNCH();
    if (PRIMAL) {
        launch(&k_hess_red_final, dim3(1), dim3(32), 0, st, (const double *)p->pd_part, grid, out2);
        PFB_CHECK_LAUNCH();
    }
    return PFB_OK;
}
template <class T, int M> struct RowsInvTma {
    static bool fits() { return false; }
    static int run(pfb_hess_plan *, T *, const Epilogue<T> &, cudaStream_t) { return PFB_ERR_BAD_ARG; }
    static int run_primal(pfb_hess_plan *, const PrimalArgs &, double *, cudaStream_t) { return PFB_ERR_BAD_ARG; }
};
template <int M> struct RowsInvTma<float, M> {
    static bool fits() {
        return sizeof(Cx<float>) * ((size_t)2 * tma_stage_elems<M>() + ct_tab_size(M) + M) + 144 <= kMaxSmemFast;
    }
    static int run(pfb_hess_plan *p, float *out, const Epilogue<float> &ep, cudaStream_t st) {
        return fast_rows_inv_tma<M, false>(p, out, ep, PrimalArgs(), nullptr, st);
    }
    static int run_primal(pfb_hess_plan *p, const PrimalArgs &pa, double *out2, cudaStream_t st) {
        return fast_rows_inv_tma<M, true>(p, nullptr, Epilogue<float>(), pa, out2, st);
    }
};

#endif

template <class T, int M, int RB>
static int fast_rows_inv(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, cudaStream_t st,
                         const PrimalArgsT<T> *pa = nullptr, double *out2 = nullptr) {
    const HessDims &h = p->h;
#ifndef PFB_EMU
    if (p->have_tma && h.nx % 2 == 0 && RowsInvTma<T, M>::fits() && !getenv("PFB_HESS_NOTMA") &&
        !getenv("PFB_HESS_NOPIPE")) {
        if constexpr (sizeof(T) == 4) {
            if (pa) return RowsInvTma<T, M>::run_primal(p, *pa, out2, st);
        }
        return RowsInvTma<T, M>::run(p, out, ep, st);
    }
#endif
    // pipelined kernel (four lines: gather of the next row pair under the transforms) when it fits
    const size_t smem4 = sizeof(Cx<T>) * ((size_t)4 * ct_stride(M) + ct_tab_size(M));
    if (!getenv("PFB_HESS_NOPIPE") && h.nx % 2 == 0) {
        if (smem4 + sizeof(Cx<T>) * M <= kMaxSmemFast)
            return fast_rows_inv2<T, M, true>(p, out, ep, st, smem4 + sizeof(Cx<T>) * M, pa, out2);
        if (smem4 <= kMaxSmemFast) return fast_rows_inv2<T, M, false>(p, out, ep, st, smem4, pa, out2);
    }
    size_t smem = sizeof(Cx<T>) * ((size_t)RB * ct_stride(M) + ct_tab_size(M));
    auto kern = pa ? (p->dyn ? &k_hessf_rows_inv<T, M, RB, true, 2, true> : &k_hessf_rows_inv<T, M, RB, false, 2, true>)
                   : (p->dyn ? &k_hessf_rows_inv<T, M, RB, true, 2, false> : &k_hessf_rows_inv<T, M, RB, false, 2, false>);
    PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int nbb = cdiv(h.nx, RB);
    int grid = persistent_grid((const void *)kern, M / 8, smem, nbb * h.nband, p->nsm);
    if (grid > 4 * p->nsm) grid = 4 * p->nsm;      // size of pd_part
    PrimalArgsT<T> a = pa ? *pa : PrimalArgsT<T>();
    a.part = p->pd_part;
    launch(kern, dim3(grid), dim3(M / 8), smem, st, (const Cx<T> *)p->S, out, (const Cx<T> *)p->tabM,
           (const Cx<T> *)p->twr, h, nbb, ep, p->ctr + 2, a);
    PFB_CHECK_LAUNCH();
    return pa ? primal_finish(p, grid, out2, st) : PFB_OK;
}


int fast_rows_inv_8k(pfb_hess_plan *p, float *out, const Epilogue<float> &ep, cudaStream_t st,
                     const PrimalArgsT<float> *pa, double *out2) {
    constexpr int M = 8192, RB = 2;
    const HessDims &h = p->h;
    const size_t smem = sizeof(Cx<float>) * ((size_t)RB * ct_stride(M) + ct_tab_size(M));
    auto kern = pa ? (p->dyn ? &k_hessf_rows_inv<float, M, RB, true, 1, true> : &k_hessf_rows_inv<float, M, RB, false, 1, true>)
                   : (p->dyn ? &k_hessf_rows_inv<float, M, RB, true, 1, false> : &k_hessf_rows_inv<float, M, RB, false, 1, false>);
    PFB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int nbb = cdiv(h.nx, RB);
    int grid = persistent_grid((const void *)kern, M / 16, smem, nbb * h.nband, p->nsm);
    if (grid > 4 * p->nsm) grid = 4 * p->nsm;      // size of pd_part
    PrimalArgsT<float> a = pa ? *pa : PrimalArgsT<float>();
    a.part = p->pd_part;
    launch(kern, dim3(grid), dim3(M / 16), smem, st, (const Cx<float> *)p->S, out, (const Cx<float> *)p->tabM,
           (const Cx<float> *)p->twr, h, nbb, ep, p->ctr + 2, a);
    PFB_CHECK_LAUNCH();
    return pa ? primal_finish(p, grid, out2, st) : PFB_OK;
}

template <class T> int fast_rows_inv_d(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, cudaStream_t st,
                                       const PrimalArgsT<T> *pa, double *out2) {
    if (p->RB == 4) {
#define PFB_C(N) fast_rows_inv<T, N, 4>(p, out, ep, st, pa, out2)
        PFB_SIZE_SWITCH(p->h.M, PFB_C)
#undef PFB_C
    }
#define PFB_C(N) fast_rows_inv<T, N, 2>(p, out, ep, st, pa, out2)
    PFB_SIZE_SWITCH(p->h.M, PFB_C)
#undef PFB_C
}

template int fast_rows_inv_d<float>(pfb_hess_plan *, float *, const Epilogue<float> &, cudaStream_t,
                                    const PrimalArgsT<float> *, double *);
template int fast_rows_inv_d<double>(pfb_hess_plan *, double *, const Epilogue<double> &, cudaStream_t,
                                     const PrimalArgsT<double> *, double *);

}  // namespace pfb
