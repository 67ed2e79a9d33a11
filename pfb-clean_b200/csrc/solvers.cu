// Solver-level entry points of the C-ABI: the iteration loops of
//   primal_dual_optimised   pfb/opt/primal_dual.py:90-171
//   pcg                     pfb/opt/pcg.py:58-131     (A = hessian_psf_cube, M = scalar)
//   power_method            pfb/opt/power_method.py:11-46   (A = psf_convolve_cube)
// as host code over the library's own kernels, for callers that are not Python (the Python
// package runs the same launch sequences itself, pfb_b200/opt/*.py, because its operators
// are callables).  One stream, device pointers owned by the caller; the loops own only
// scratch, allocated on entry and released on exit.  They synchronise the stream once per
// iteration on a 16-byte record in pinned memory (the convergence test), with the next
// iteration's first kernel already queued where that kernel does not depend on the test.
#include "common.cuh"

#include <cmath>
#include <cstring>
#include <vector>

namespace pfb {

namespace {

struct Scratch {                       // device / pinned allocations released on scope exit
    std::vector<void *> dev, host;
    cudaEvent_t ev = nullptr;
    ~Scratch() {
        for (void *p : dev) cudaFree(p);
        for (void *p : host) cudaFreeHost(p);
        if (ev) cudaEventDestroy(ev);
    }
    void *alloc(size_t n) {
        void *p = nullptr;
        if (cudaMalloc(&p, n ? n : 1) != cudaSuccess) return nullptr;
        dev.push_back(p);
        return p;
    }
    void *pinned(size_t n) {
        void *p = nullptr;
        if (cudaMallocHost(&p, n) != cudaSuccess) return nullptr;
        host.push_back(p);
        return p;
    }
};

inline size_t esize(int dtype) { return dtype == PFB_F32 ? 4 : 8; }

}  // namespace
}  // namespace pfb

using namespace pfb;

#define PFB_TRY(expr)                          \
    do {                                       \
        const int rc_ = (expr);                \
        if (rc_ != PFB_OK) return rc_;         \
    } while (0)

extern "C" {

int pfb_pd_run(const pfb_psi_plan *psi, pfb_hess_plan *hess, void *x, void *v, const void *data,
               const void *l1weight, int nband, int nx, int ny, int nbasis, double lam, double L,
               double nu, double sigma, double gamma, double tol, int maxit, int positivity,
               int dtype, int *niter, double *eps_out, void *stream) {
    if (!psi || !hess || !x || !v || !data || !l1weight || nband < 1 || nx < 1 || ny < 1 || nbasis < 1 ||
        maxit < 0 || (dtype != PFB_F32 && dtype != PFB_F64) || positivity < 0 || positivity > 2 ||
        !(L > 0.0) || !(nu > 0.0) || !(gamma > 0.0))
        return fail(PFB_ERR_BAD_ARG, "pfb_pd_run: bad argument");
    long long stride = 0;
    int Nxmax = 0, Nymax = 0;
    PFB_TRY(pfb_psi_nat_stride(psi, &stride));
    PFB_TRY(pfb_psi_plan_query(psi, &Nxmax, &Nymax));
    if (stride <= 0)
        return fail(PFB_ERR_BAD_ARG, "pfb_pd_run: the plan's bases have no natural-orientation layout "
                                     "(db5 or repeated bases)");
    cudaStream_t st = as_stream(stream);
    const size_t s = esize(dtype);
    const long long npix = (long long)nx * ny, nimg = npix * nband;
    const long long n = (long long)nbasis * stride;              // coefficients per band (natural layout)
    if (((uintptr_t)x & 15) || (npix * (long long)s) % 16)
        return fail(PFB_ERR_BAD_ARG, "pfb_pd_run: x must be 16-byte aligned with 16-byte bands");
    if (sigma <= 0.0) sigma = L / (2.0 * gamma) / nu;                       // primal_dual.py:110-111
    const double tau = 0.9 / (L / (2.0 * gamma) + sigma * nu * nu);         // :114

    Scratch sc;
    void *va = sc.alloc(s * n * nband), *vb = sc.alloc(s * n * nband), *w = sc.alloc(s * n);
    void *xb = sc.alloc(s * nimg), *wa = sc.alloc(s * nimg), *wb = sc.alloc(s * nimg);
    double *terms = (double *)sc.alloc(2 * sizeof(double));
    double *hterms = (double *)sc.pinned(2 * sizeof(double));
    const bool fuse = positivity != 2 && pfb_hess_plan_can_fuse_primal(hess) != 0;
    void *conv = fuse ? nullptr : sc.alloc(s * nimg);
    if (!va || !vb || !w || !xb || !wa || !wb || !terms || !hterms || (!fuse && !conv))
        return fail(PFB_ERR_ALLOC, "pfb_pd_run: scratch allocation failed");
    if (cudaEventCreateWithFlags(&sc.ev, cudaEventDisableTiming) != cudaSuccess)
        return fail(PFB_ERR_CUDA, "pfb_pd_run: event creation failed");

    // dual and weights into the natural layout (padding cells zero), xp = x, w_{-1} = Psi(v_0)
    PFB_CUDA_OK(cudaMemsetAsync(va, 0, s * n * nband, st));
    PFB_CUDA_OK(cudaMemsetAsync(w, 0, s * n, st));
    PFB_TRY(pfb_psi_to_nat(psi, nband, v, va, st));
    PFB_TRY(pfb_psi_to_nat(psi, 1, l1weight, w, st));
    PFB_CUDA_OK(cudaMemcpyAsync(vb, va, s * n * nband, cudaMemcpyDeviceToDevice, st));
    PFB_CUDA_OK(cudaMemcpyAsync(xb, x, s * nimg, cudaMemcpyDeviceToDevice, st));
    void *xcur = x, *xprev = xb;          // the new model is written to xcur
    void *vcur = va, *vprev = vb;
    void *wcur = wa, *wprev = wb;
    PFB_TRY(pfb_psi_hdot_nat(psi, vprev, nullptr, wprev, st));

    double eps = 1.0;
    int k = 0;
    bool have_dot = false, broke = false;
    for (k = 0; k < maxit; ++k) {
        if (!have_dot) PFB_TRY(pfb_psi_dot_nat(psi, xprev, vcur, st));                  // :135
        have_dot = false;
        PFB_TRY(pfb_dual_update(vprev, vcur, w, nullptr, nband, n, lam, sigma, PFB_BAND_SUM, dtype, st));  // :136
        PFB_TRY(pfb_psi_hdot_nat(psi, vcur, nullptr, wcur, st));                        // :137-138 by linearity
        if (fuse) {
            PFB_TRY(pfb_hess_apply_primal(hess, xprev, wcur, wprev, data, tau, positivity, xcur, terms, st));
        } else {
            PFB_TRY(pfb_hess_apply(hess, xprev, nullptr, 0, 0.0, 0, 0.0, conv, st));
            PFB_TRY(pfb_pd_primal_lin(xcur, xprev, wcur, wprev, conv, data, tau, positivity, nband, npix,
                                      dtype, terms, st));                               // :139-150
        }
        { void *t = wcur; wcur = wprev; wprev = t; }
        PFB_CUDA_OK(cudaMemcpyAsync(hterms, terms, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
        PFB_CUDA_OK(cudaEventRecord(sc.ev, st));
        if (k + 1 < maxit) {              // next iteration's analysis, valid whatever the test decides
            PFB_TRY(pfb_psi_dot_nat(psi, xcur, vprev, st));
            have_dot = true;
        }
        PFB_CUDA_OK(cudaEventSynchronize(sc.ev));
        if (hterms[1] == 0.0)
            return fail(PFB_ERR_NAN, "pfb_pd_run: model is identically zero (the reference drops into pdb here)");
        eps = std::sqrt(hterms[0] / (1e-12 + hterms[1]));                               // :152 norm_diff
        if (std::isnan(eps) || std::isinf(eps)) return fail(PFB_ERR_NAN, "pfb_pd_run: eps is NaN/Inf");
        if (eps < tol) { broke = true; break; }                                         // :153-162
        { void *t = xcur; xcur = xprev; xprev = t; }                                    // :164-165 as swaps
        { void *t = vcur; vcur = vprev; vprev = t; }
    }
    const void *xfin = broke ? xcur : xprev, *vfin = broke ? vcur : vprev;
    if (xfin != x) PFB_CUDA_OK(cudaMemcpyAsync(x, xfin, s * nimg, cudaMemcpyDeviceToDevice, st));
    PFB_TRY(pfb_psi_from_nat(psi, nband, vfin, v, st));
    PFB_CUDA_OK(cudaStreamSynchronize(st));
    if (niter) *niter = broke ? k + 1 : k;                        // iterations run
    if (eps_out) *eps_out = eps;
    return PFB_OK;
}


int pfb_power_method_run(pfb_hess_plan *hess, void *b, int nband, int nx, int ny, double tol, int maxit,
                         int dtype, double *beta_out, int *niter, void *stream) {
    if (!hess || !b || nband < 1 || nx < 1 || ny < 1 || maxit < 0 || (dtype != PFB_F32 && dtype != PFB_F64))
        return fail(PFB_ERR_BAD_ARG, "pfb_power_method_run: bad argument");
    cudaStream_t st = as_stream(stream);
    const size_t s = esize(dtype);
    const long long n = (long long)nband * nx * ny;
    Scratch sc;
    void *ab = sc.alloc(s * n);
    double *d2 = (double *)sc.alloc(2 * sizeof(double));
    double *h2 = (double *)sc.pinned(2 * sizeof(double));
    if (!ab || !d2 || !h2) return fail(PFB_ERR_ALLOC, "pfb_power_method_run: scratch allocation failed");
    auto read2 = [&]() -> int {
        PFB_CUDA_OK(cudaMemcpyAsync(h2, d2, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
        PFB_CUDA_OK(cudaStreamSynchronize(st));
        return PFB_OK;
    };
    PFB_TRY(pfb_vdot(b, b, nullptr, nullptr, n, dtype, d2, st));                 // b = b0 / ||b0||  (:20-21)
    PFB_TRY(read2());
    if (!(h2[0] > 0.0)) return fail(PFB_ERR_NAN, "pfb_power_method_run: zero start vector");
    PFB_TRY(pfb_axpby(b, 1.0 / std::sqrt(h2[0]), b, 0.0, b, n, dtype, st));
    double beta = 1.0, eps = 1.0;
    int k = 0;
    while (eps > tol && k < maxit) {                                              // :26
        PFB_TRY(pfb_hess_apply(hess, b, nullptr, 0, 0.0, 0, 0.0, ab, st));        // b = A(bp)
        PFB_TRY(pfb_vdot(ab, ab, nullptr, nullptr, n, dtype, d2, st));
        PFB_TRY(read2());
        const double bnorm = std::sqrt(h2[0]);                                    // :28
        PFB_TRY(pfb_vdot(b, ab, b, b, n, dtype, d2, st));                         // beta = <bp, b> / <bp, bp>  (:30)
        PFB_TRY(read2());
        const double betap = beta;
        beta = h2[0] / h2[1];
        if (std::isnan(beta) || !(bnorm > 0.0)) return fail(PFB_ERR_NAN, "pfb_power_method_run: NaN");
        PFB_TRY(pfb_axpby(b, 1.0 / bnorm, ab, 0.0, ab, n, dtype, st));            // b /= bnorm; bp = b  (:31,35)
        eps = std::fabs(beta - betap) / betap;                                    // :33
        ++k;
    }
    PFB_CUDA_OK(cudaStreamSynchronize(st));
    if (beta_out) *beta_out = beta;
    if (niter) *niter = k;
    return PFB_OK;
}

int pfb_pcg_run(pfb_hess_plan *hess, const void *beam, double sigmainv, int has_wsum, double wsum,
                const void *bvec, void *x, double minv, int nband, int nx, int ny, double tol, int maxit,
                int minit, int backtrack, int dtype, int *niter, double *eps_out, void *stream) {
    if (!hess || !bvec || !x || nband < 1 || nx < 1 || ny < 1 || maxit < 0 || (dtype != PFB_F32 && dtype != PFB_F64))
        return fail(PFB_ERR_BAD_ARG, "pfb_pcg_run: bad argument");
    cudaStream_t st = as_stream(stream);
    const size_t s = esize(dtype);
    const long long n = (long long)nband * nx * ny;
    const int NS = PFB_PCG_NSTATE;
    Scratch sc;
    void *r = sc.alloc(s * n), *p = sc.alloc(s * n), *Ap = sc.alloc(s * n);
    void *xn = sc.alloc(s * n), *rn = sc.alloc(s * n), *pn = sc.alloc(s * n);
    double *state = (double *)sc.alloc(NS * sizeof(double));
    double *h = (double *)sc.pinned(NS * sizeof(double));
    if (!r || !p || !Ap || !xn || !rn || !pn || !state || !h)
        return fail(PFB_ERR_ALLOC, "pfb_pcg_run: scratch allocation failed");
    if (cudaEventCreateWithFlags(&sc.ev, cudaEventDisableTiming) != cudaSuccess)
        return fail(PFB_ERR_CUDA, "pfb_pcg_run: event creation failed");
    void *xc = x;                                    // current iterate (ping-pong with xn)
    auto A = [&](const void *in, void *out) {        // hessian_psf_cube, pfb/operators/hessian.py:254-281
        return pfb_hess_apply(hess, in, beam, 1, sigmainv, has_wsum, wsum, out, st);
    };
    auto dot = [&](const void *a, const void *c, int slot) {
        return pfb_pcg_seg_dot(a, c, 1, n, dtype, state, slot, st);
    };
    auto copy_state = [&]() -> int {                 // state -> pinned, completion on the event
        PFB_CUDA_OK(cudaMemcpyAsync(h, state, NS * sizeof(double), cudaMemcpyDeviceToHost, st));
        PFB_CUDA_OK(cudaEventRecord(sc.ev, st));
        return PFB_OK;
    };
    auto wait_state = [&]() -> int { PFB_CUDA_OK(cudaEventSynchronize(sc.ev)); return PFB_OK; };
    auto read_state = [&]() -> int { PFB_TRY(copy_state()); return wait_state(); };
    auto update = [&]() -> int {
        PFB_TRY(pfb_pcg_seg_update(xn, rn, xc, r, p, Ap, minv, 1, n, dtype, state, st));
        return pfb_pcg_seg_scalar(state, 1, PFB_PCG_BETA, st);
    };
    auto direction = [&]() { return pfb_pcg_seg_direction(pn, p, rn, minv, 1, n, dtype, state, st); };
    auto set_active = [&](double on) -> int {
        PFB_CUDA_OK(cudaMemcpyAsync(state + 9, &on, sizeof(double), cudaMemcpyHostToDevice, st));
        PFB_CUDA_OK(cudaStreamSynchronize(st));      // `on` lives on this stack frame
        return PFB_OK;
    };

    PFB_CUDA_OK(cudaMemsetAsync(state, 0, NS * sizeof(double), st));
    PFB_TRY(A(xc, r));                                                       // r = A(x0) - b   (pcg.py:72)
    PFB_TRY(pfb_axpby(r, 1.0, r, -1.0, bvec, n, dtype, st));
    // y = M(r) = minv r; <r, y>; <y, y> (zero residual test); p = -y          (:73-79)
    PFB_TRY(pfb_axpby(pn, minv, r, 0.0, r, n, dtype, st));
    PFB_TRY(dot(r, pn, 0));
    PFB_TRY(dot(pn, pn, 7));
    PFB_TRY(pfb_axpby(p, -1.0, pn, 0.0, pn, n, dtype, st));
    PFB_TRY(read_state());
    int k = 0;
    double eps = 1.0;
    if (h[7] == 0.0) {                                                       // "Initial residual is zero"
        if (niter) *niter = 0;
        if (eps_out) *eps_out = 0.0;
        return PFB_OK;
    }
    PFB_TRY(set_active(1.0));
    PFB_TRY(A(p, Ap));
    PFB_TRY(dot(p, Ap, 1));
    PFB_TRY(pfb_pcg_seg_scalar(state, 1, PFB_PCG_ALPHA, st));
    bool pending_nnz = false, running = true;
    while (running) {
        // the state is copied right behind the line update; the direction kernel is queued
        // behind the copy, so A(pn) follows without the GPU draining (only a backtrack redoes it)
        PFB_TRY(update());
        PFB_TRY(copy_state());
        PFB_TRY(direction());
        PFB_TRY(wait_state());
        if (pending_nnz && h[7] == 0.0) break;       // `if not np.any(p): break` of the previous iteration (:106-107)
        if (backtrack && h[8] != 0.0) {                                       // :96-101
            while (h[8] != 0.0) {
                PFB_TRY(pfb_pcg_seg_scalar(state, 1, PFB_PCG_SHRINK, st));
                PFB_TRY(update());
                PFB_TRY(read_state());
            }
            PFB_TRY(direction());
        }
        pending_nnz = true;
        ++k;
        eps = std::sqrt(h[4] / (1e-12 + h[5]));                               // :109 norm_diff
        if (std::isnan(eps)) return fail(PFB_ERR_NAN, "pfb_pcg_run: eps is NaN");
        if (!((eps > tol || k < minit) && k < maxit)) running = false;        // :86
        PFB_TRY(pfb_pcg_seg_scalar(state, 1, PFB_PCG_COMMIT, st));
        { void *t = xc; xc = xn; xn = t; }
        { void *t = r; r = rn; rn = t; }
        { void *t = p; p = pn; pn = t; }
        if (!running) break;
        PFB_TRY(A(p, Ap));
        PFB_TRY(dot(p, Ap, 1));
        PFB_TRY(pfb_pcg_seg_scalar(state, 1, PFB_PCG_ALPHA, st));
    }
    if (xc != x) PFB_CUDA_OK(cudaMemcpyAsync(x, xc, s * n, cudaMemcpyDeviceToDevice, st));
    PFB_CUDA_OK(cudaStreamSynchronize(st));
    if (niter) *niter = k;
    if (eps_out) *eps_out = eps;
    return PFB_OK;
}

}  // extern "C"
