// Launch sites of the peeled Hessian kernels (hess_peel.cuh): (3 | 5 | 7) x 2^k lines, see
// hess_plan.cuh for why this is its own translation unit.
#include "hess_plan.cuh"

namespace pfb {

// ---- odd factor x power of two (hess_peel.cuh)
#define PFB_PEEL_SWITCH(n, CALL)                         \
    switch (n) {                                         \
    case 512: return CALL(512);                          \
    case 1024: return CALL(1024);                        \
    case 2048: return CALL(2048);                        \
    case 4096: return CALL(4096);                        \
    default: return fail(PFB_ERR_FFT_SIZE, "peeled transform: unsupported power of two %d", (int)(n)); \
    }
template <class T, int P2>
static int peel_rows_fwd(pfb_hess_plan *p, const T *x, const T *beam, cudaStream_t st) {
    HessDims h = p->h;
    h.R = p->LrowP;
    const size_t smem = sizeof(Cx<T>) * peel_smem_elems(h.M, p->prow.m, P2, h.R, p->inplR != 0);
    const int nthr = p->inplR ? p->prow.m * (P2 / 16) : peel_threads(p->prow.m, P2, h.R);
    PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessp_rows_fwd<T, P2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch(&k_hessp_rows_fwd<T, P2>, dim3(cdiv(h.nx, h.R), h.nband), dim3(nthr), smem, st,
           x, beam, (Cx<T> *)p->S, (const Cx<T> *)p->mtwM, (const Cx<T> *)p->twr, (const Cx<T> *)p->mroots,
           (const Cx<T> *)p->ptabR, p->prow.m, p->inplR, h);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}
template <class T, int P2>
static int peel_rows_inv(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, cudaStream_t st) {
    HessDims h = p->h;
    h.R = p->LrowP;
    const size_t smem = sizeof(Cx<T>) * peel_smem_elems(h.M, p->prow.m, P2, h.R, p->inplR != 0);
    const int nthr = p->inplR ? p->prow.m * (P2 / 16) : peel_threads(p->prow.m, P2, h.R);
    PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessp_rows_inv<T, P2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch(&k_hessp_rows_inv<T, P2, false>, dim3(cdiv(h.nx, h.R), h.nband), dim3(nthr), smem, st,
           (const Cx<T> *)p->S, out, (const Cx<T> *)p->mtwM, (const Cx<T> *)p->twr, (const Cx<T> *)p->mroots,
           (const Cx<T> *)p->ptabR, p->prow.m, p->inplR, h, ep, PeelPrimal());
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}
#ifndef PFB_EMU
// pass C with the primal step as its epilogue (fp32), then the fixed-order sum of the CTAs' eps terms
template <int P2>
static int peel_rows_inv_primal(pfb_hess_plan *p, PrimalArgs pa, double *out2, cudaStream_t st) {
    HessDims h = p->h;
    h.R = p->LrowP;
    const size_t smem = sizeof(Cx<float>) * peel_smem_elems(h.M, p->prow.m, P2, h.R, p->inplR != 0);
    const int nthr = p->inplR ? p->prow.m * (P2 / 16) : peel_threads(p->prow.m, P2, h.R);
    const dim3 grid(cdiv(h.nx, h.R), h.nband);
    pa.part = p->pd_partP;
    PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessp_rows_inv<float, P2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch(&k_hessp_rows_inv<float, P2, true>, grid, dim3(nthr), smem, st,
           (const Cx<float> *)p->S, (float *)nullptr, (const Cx<float> *)p->mtwM, (const Cx<float> *)p->twr,
           (const Cx<float> *)p->mroots, (const Cx<float> *)p->ptabR, p->prow.m, p->inplR, h, Epilogue<float>(), pa);
    PFB_CHECK_LAUNCH();
    launch(&k_hess_red_final_wide, dim3(1), dim3(1024), 0, st, (const double *)p->pd_partP, (int)(grid.x * grid.y), out2);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}
int peel_rows_inv_primal_d(pfb_hess_plan *p, const PrimalArgs &pa, double *out2, cudaStream_t st) {
#define PFB_C(N) peel_rows_inv_primal<N>(p, pa, out2, st)
    PFB_PEEL_SWITCH(p->prow.P2, PFB_C)
#undef PFB_C
}
#endif
template <class T, int P2>
static int peel_cols(pfb_hess_plan *p, cudaStream_t st) {
    const HessDims &h = p->h;
    if (p->pcol2) {
        const size_t smem2 = sizeof(Cx<T>) * peel_smem_elems_cols2(p->pcol.N, p->pcol.m, P2, p->split != 0);
        const dim3 grid2(h.nyo2, h.nband), blk2(p->pcol.m * (P2 / 16));
        if (p->split) {
            PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessp_cols2<T, P2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
            launch(&k_hessp_cols2<T, P2, true>, grid2, blk2, smem2, st, (Cx<T> *)p->S, (const Cx<T> *)p->psfT,
                   (const Cx<T> *)p->mtwC, (const Cx<T> *)p->modg, (const Cx<T> *)p->mroots, (const Cx<T> *)p->ptabC,
                   p->pcol.m, h);
        } else {
            PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessp_cols2<T, P2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
            launch(&k_hessp_cols2<T, P2, false>, grid2, blk2, smem2, st, (Cx<T> *)p->S, (const Cx<T> *)p->psfT,
                   (const Cx<T> *)p->mtwC, (const Cx<T> *)p->modg, (const Cx<T> *)p->mroots, (const Cx<T> *)p->ptabC,
                   p->pcol.m, h);
        }
        PFB_CHECK_LAUNCH();
        return PFB_OK;
    }
    const size_t smem = sizeof(Cx<T>) * peel_smem_elems(p->pcol.N, p->pcol.m, P2, 1);
    const dim3 grid(h.nyo2, h.nband), blk(peel_threads(p->pcol.m, P2, 1));
    if (p->split) {
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessp_cols<T, P2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        launch(&k_hessp_cols<T, P2, true>, grid, blk, smem, st, (Cx<T> *)p->S, (const Cx<T> *)p->psfT,
               (const Cx<T> *)p->mtwC, (const Cx<T> *)p->modg, (const Cx<T> *)p->mroots, (const Cx<T> *)p->ptabC,
               p->pcol.m, h);
    } else {
        PFB_CUDA_OK(cudaFuncSetAttribute(&k_hessp_cols<T, P2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        launch(&k_hessp_cols<T, P2, false>, grid, blk, smem, st, (Cx<T> *)p->S, (const Cx<T> *)p->psfT,
               (const Cx<T> *)p->mtwC, (const Cx<T> *)p->modg, (const Cx<T> *)p->mroots, (const Cx<T> *)p->ptabC,
               p->pcol.m, h);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}
template <class T> int peel_rows_fwd_d(pfb_hess_plan *p, const T *x, const T *beam, cudaStream_t st) {
#define PFB_C(N) peel_rows_fwd<T, N>(p, x, beam, st)
    PFB_PEEL_SWITCH(p->prow.P2, PFB_C)
#undef PFB_C
}
template <class T> int peel_rows_inv_d(pfb_hess_plan *p, T *out, const Epilogue<T> &ep, cudaStream_t st) {
#define PFB_C(N) peel_rows_inv<T, N>(p, out, ep, st)
    PFB_PEEL_SWITCH(p->prow.P2, PFB_C)
#undef PFB_C
}
template <class T> int peel_cols_d(pfb_hess_plan *p, cudaStream_t st) {
#define PFB_C(N) peel_cols<T, N>(p, st)
    PFB_PEEL_SWITCH(p->pcol.P2, PFB_C)
#undef PFB_C
}


template int peel_rows_fwd_d<float>(pfb_hess_plan *, const float *, const float *, cudaStream_t);
template int peel_rows_fwd_d<double>(pfb_hess_plan *, const double *, const double *, cudaStream_t);
template int peel_rows_inv_d<float>(pfb_hess_plan *, float *, const Epilogue<float> &, cudaStream_t);
template int peel_rows_inv_d<double>(pfb_hess_plan *, double *, const Epilogue<double> &, cudaStream_t);
template int peel_cols_d<float>(pfb_hess_plan *, cudaStream_t);
template int peel_cols_d<double>(pfb_hess_plan *, cudaStream_t);

}  // namespace pfb
