/* pfb_b200 -- C-ABI of the B200-native pfb-imaging minor-cycle hot path.
 *
 * The reference (pfb-imaging v0.0.4, pure Python) has no FFI layer of its own:
 * its hot path is Python callables over caller-owned numpy buffers whose
 * arithmetic runs in ducc0 (C++), numba and numexpr.  Each entry point below
 * replaces one of those callables; the reference interface it stands in for is
 * cited as file:line relative to the reference tree.  INTEGRATION.md shows the
 * ctypes binding a pfb-imaging maintainer would add.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller unless named *_host;
 *  - arrays are C-contiguous; `dtype` is PFB_F32 or PFB_F64 (complex arrays are
 *    interleaved re,im of that real type);
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream); calls
 *    enqueue work and return without synchronising unless stated otherwise;
 *  - return value is a pfb_status; pfb_last_error() gives the message of the
 *    last failure on the calling thread;
 *  - plans own only scratch memory; they are not re-entrant (like the
 *    reference's shared xpad/xhat/xout scratch).
 */
#ifndef PFB_B200_H
#define PFB_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PFB_OK = 0,
    PFB_ERR_BAD_ARG = 1,     /* null pointer, non-positive size, unknown enum */
    PFB_ERR_ODD_SIZE = 2,    /* image sizes must be even (tests/test_wavelets.py:9-10) */
    PFB_ERR_NLEVEL = 3,      /* ValueError of pfb/operators/psi.py:43-46 */
    PFB_ERR_FFT_SIZE = 4,    /* padded size not supported by the FFT kernels */
    PFB_ERR_CUDA = 5,        /* CUDA runtime error, see pfb_last_error() */
    PFB_ERR_ALLOC = 6,
    PFB_ERR_NBAND = 7,       /* more bands than a kernel supports */
    PFB_ERR_NAN = 8          /* a solver loop met NaN / Inf or an identically zero model */
} pfb_status;

typedef enum { PFB_F32 = 0, PFB_F64 = 1 } pfb_dtype;

/* cross-band reduction of the l21 prox: sum (prox_21m) or l2 norm (prox_21) */
typedef enum { PFB_BAND_SUM = 0, PFB_BAND_L2 = 1 } pfb_band_mode;

int pfb_version(void);
const char *pfb_last_error(void);
/* number of kernel launches issued by this library since load (bench.py's
 * gpu_launches is the difference around the timed region) */
long long pfb_launch_count(void);

/* Per-kernel device timing for bench.py's roofline: while enabled every kernel
 * launch of the library is bracketed by CUDA events on its own stream.
 * pfb_profile_report synchronises the device and writes one line per kernel
 * symbol, "mangled_name\tlaunches\ttotal_ms\n", into buf (host memory).
 * Enabling clears the previous records. */
int pfb_profile_enable(int on);
int pfb_profile_report(char *buf_host, long long buflen);

/* ------------------------------------------------------------------ prox --
 * Coefficient arrays are (nband, n) with n = nbasis*Nymax*Nxmax flattened;
 * weight is (n).  Scalar math is fp64 as in the reference's numba kernels.   */

/* prox_21m_numba(v, result, lam, sigma, weight)  pfb/prox/prox_21m.py:30-61
 * (mode PFB_BAND_SUM);  prox_21_numba  pfb/prox/prox_21.py:25-52 (PFB_BAND_L2) */
int pfb_prox21(const void *v, void *result, const void *weight, int nband,
               long long n, double lam, double sigma, int mode, int dtype,
               void *stream);

/* dual_update_numba(vp, v, lam, sigma, weight), in place on v
 * pfb/prox/prox_21m.py:75-103 (PFB_BAND_SUM), pfb/prox/prox_21.py:66-91 (L2).
 * If u != NULL additionally u = 2*v_new - vp, the numexpr line that follows
 * it in primal_dual_optimised (pfb/opt/primal_dual.py:137); u may alias vp. */
int pfb_dual_update(const void *vp, void *v, const void *weight, void *u,
                    int nband, long long n, double lam, double sigma, int mode,
                    int dtype, void *stream);

/* Band-sharded split of the same update (the reference's distributed analogue
 * is vtilde_update + get_ratio + update, pfb/opt/primal_dual.py:183-222):
 *   partial: bsum[k] = sum (or sum of squares) of vp + sigma*v over the LOCAL
 *            bands; v is not modified.  The caller all-reduces bsum (SUM).
 *   apply  : v <- (vp + sigma*v) * (1 - max(a - lam w/sigma, 0)/a), a from
 *            bsum; u as above
 * band_stride = elements between consecutive bands of vp / v (0 => n); a value
 * > n lets the caller run the split on one basis slab of a (nband, nbasis, ...)
 * array at a time, so that the all-reduce of basis b overlaps the kernels of
 * basis b+1. */
/* Band-sharded 21m dual update without a coefficient-sized exchange: by linearity
 * of Psi^T the cross-band sum is  S = vbar + sigma * ubar  with ubar = Psi^T(sum_b x_b)
 * (one all-reduced IMAGE, analysed by every rank) and vbar = sum_b v_b of the previous
 * iteration, which each rank carries itself because  sum_b v_b = f(S) * S.
 *   v_b <- (vp_b + sigma v_b) f(S) for the nband local bands;  vbar <- f(S) S  (in place).
 * vbar, ubar, weight: n elements; vp, v: (nband, n). */
int pfb_dual_update_rec(const void *vp, void *v, void *vbar, const void *ubar, const void *weight,
                        int nband, long long n, double lam, double sigma, int dtype, void *stream);
/* The same update with the per-band half moved into the analysis: pfb_dual_factor does the
 * band-independent half (S = vbar + sigma ubar; f = shrink factor; vbar <- f S; f stored) and
 * pfb_psi_dot_nat_dual (below) stores v_b = (vp_b + sigma Psi^T x_b) f straight from the analysis
 * kernels' registers, so the analysed coefficients u_b are never written or read back. */
int pfb_dual_factor(void *vbar, const void *ubar, const void *weight, void *f, long long n,
                    double lam, double sigma, int dtype, void *stream);
int pfb_dual_partial(const void *vp, const void *v, void *bsum, int nband, long long n,
                     long long band_stride, double sigma, int mode, int dtype, void *stream);
int pfb_dual_apply(const void *vp, void *v, const void *bsum, const void *weight,
                   void *u, int nband, long long n, long long band_stride, double lam,
                   double sigma, int mode, int dtype, void *stream);

/* ------------------------------------------------- fused vector updates --
 * Reductions accumulate in fp64 in a fixed order (bitwise reproducible) and
 * write K doubles to a DEVICE buffer outK; they use one library-owned
 * workspace, so issue them from one stream at a time.                        */

/* out = a*x + b*y ('2.0*v - vp', 'xp - tau*xout' pfb/opt/primal_dual.py:137,140;
 * axpys of pfb/opt/pcg.py:92-104) */
int pfb_axpby(void *out, double a, const void *x, double b, const void *y,
              long long n, int dtype, void *stream);
/* out2 = {sum a*b, sum c*d} (c,d may both be NULL)  np.vdot in pfb/opt/pcg.py:76,91 */
int pfb_vdot(const void *a, const void *b, const void *c, const void *d,
             long long n, int dtype, double *out2, void *stream);
/* out2 = {sum (x-xp)^2, sum x^2}; eps = sqrt(out2[0]/(1e-12+out2[1]))
 * norm_diff  pfb/utils/misc.py:1316-1351 */
int pfb_norm_diff(const void *x, const void *xp, long long n, int dtype,
                  double *out2, void *stream);
/* primal step of primal_dual_optimised, pfb/opt/primal_dual.py:139-150:
 * x = P+(xp - tau*(xout + conv - data)), out2 as pfb_norm_diff.  conv/data may
 * be NULL.  positivity: 0 none, 1 clamp, 2 zero pixel if any band <= 0. */
int pfb_pd_primal(void *x, const void *xp, const void *xout, const void *conv,
                  const void *data, double tau, int positivity, int nband,
                  long long npix, int dtype, double *out2, void *stream);
/* Same step with the synthesis taken by linearity: the solver keeps w_k = Psi(v_k)
 * of the last two iterations and forms xout = Psi(2 v_k - v_{k-1}) = 2 w - wprev
 * (primal_dual.py:137-138) inside this kernel, so the dual update does not have
 * to write 2v - vp (one pass over the coefficient cube less per iteration). */
int pfb_pd_primal_lin(void *x, const void *xp, const void *w, const void *wprev,
                      const void *conv, const void *data, double tau, int positivity,
                      int nband, long long npix, int dtype, double *out2, void *stream);
/* pcg line update, pfb/opt/pcg.py:92-101: x = xp + alpha p; r = rp + alpha Ap;
 * out3 = {<r, minv*r>, sum (x-xp)^2, sum x^2} */
int pfb_pcg_update(void *x, void *r, const void *xp, const void *rp, const void *p,
                   const void *Ap, double alpha, double minv, long long n,
                   int dtype, double *out3, void *stream);
/* p = beta*p - minv*r ; out1 = {number of non-zero p}  pfb/opt/pcg.py:103-107 */
int pfb_pcg_direction(void *p, const void *r, double beta, double minv,
                      long long n, int dtype, double *out1, void *stream);

/* Device-resident line search of pcg (pfb/opt/pcg.py:86-122) and of the per-band solves of
 * _pcg_psf_impl (pfb/opt/pcg.py:243-291, which the reference runs concurrently over bands
 * with da.blockwise): the vectors are nseg independent segments of seglen elements (one
 * segment = the whole cube for pcg, one band for pcg_psf) and the scalars live in a device
 * state block of PFB_PCG_NSTATE doubles per segment:
 *   [0] <r,y>  [1] <p,Ap>  [2] alpha  [3] <r,y>_next  [4] sum (x-xp)^2  [5] sum x^2
 *   [6] beta   [7] number of non-zero p  [8] backtrack flag  [9] active (caller-set)
 * so that an iteration is enqueued without a host round trip:
 *   A(p); seg_dot(p, Ap -> [1]); seg_scalar(ALPHA); seg_update; seg_scalar(BETA);
 *   seg_direction; ...host reads the state once, late...; seg_scalar(COMMIT)
 * Inactive segments ([9] == 0) pass through unchanged.  Between a sum and the scalar step
 * that uses it the caller may all-reduce the state (band-sharded global pcg). */
#define PFB_PCG_NSTATE 10
typedef enum {
    PFB_PCG_ALPHA = 0,   /* [2] = [0] / [1]                                   pcg.py:90-91 */
    PFB_PCG_BETA = 1,    /* [8] = active && [3] > [0];  [6] = [3] / [0]       pcg.py:96,103 */
    PFB_PCG_SHRINK = 2,  /* [2] *= 0.75 where [8] is set                      pcg.py:97 */
    PFB_PCG_COMMIT = 3   /* [0] = [3] where active                            pcg.py:108 */
} pfb_pcg_step;
/* state[seg][slot] = sum over the segment of a*b (fp64, fixed order) */
int pfb_pcg_seg_dot(const void *a, const void *b, int nseg, long long seglen, int dtype,
                    double *state, int slot, void *stream);
int pfb_pcg_seg_scalar(double *state, int nseg, int step, void *stream);
/* x = xp + alpha p; r = rp + alpha Ap (alpha = state[seg][2]); state[seg][3..5] =
 * {<r, minv r>, sum (x-xp)^2, sum x^2}                                       pcg.py:92-95,111 */
int pfb_pcg_seg_update(void *x, void *r, const void *xp, const void *rp, const void *p,
                       const void *Ap, double minv, int nseg, long long seglen, int dtype,
                       double *state, void *stream);
/* pn = beta p - minv r (beta = state[seg][6]); state[seg][7] = count of non-zero pn; pn may
 * not alias p (the old direction survives for a late backtrack)              pcg.py:103-107 */
int pfb_pcg_seg_direction(void *pn, const void *p, const void *r, double minv, int nseg,
                          long long seglen, int dtype, double *state, void *stream);

/* ------------------------------------------------------------------- Psi --
 * Multi-basis wavelet dictionary, pfb/operators/psi.py:269-310 (class Psi) over
 * pfb/wavelets/wavelets.py dwt2d :174 / idwt2d :260.  basis_ids: 0 = 'self',
 * N = 'dbN' (1..5).  Coefficients are (nband, nbasis, Nymax, Nxmax).         */
/* Daubechies db<order> (order 1..5) filter bank compiled into the kernels, as
 * pywt.Wavelet('db<order>').filter_bank (pfb/operators/psi.py:37-41):
 * taps[w*10 + k], w = 0 dec_lo, 1 dec_hi, 2 rec_lo, 3 rec_hi, k < 2*order (rest 0). */
int pfb_filter_bank(int order, double *taps);
typedef struct pfb_psi_plan pfb_psi_plan;
/* Psi(nband, nx, ny, bases, nlevel, nthreads) + psi_band_maker
 * pfb/operators/psi.py:17-123, :270-282.  PFB_ERR_NLEVEL mirrors the
 * ValueError of psi.py:43-46, PFB_ERR_ODD_SIZE the even-size restriction. */
int pfb_psi_plan_create(pfb_psi_plan **out, int nband, int nx, int ny, int nbasis,
                        const int *basis_ids, int nlevel, int dtype);
int pfb_psi_plan_destroy(pfb_psi_plan *plan);
/* Psi.Nxmax / Psi.Nymax  pfb/operators/psi.py:281-282 */
int pfb_psi_plan_query(const pfb_psi_plan *plan, int *Nxmax, int *Nymax);
/* per-basis pyramid tables (host ints, 6*nlevel: highx, highy, sx, sy, spx, spy)
 * = ix/iy/sx/sy/spx/spy of psi_band_maker, for tests and integration */
int pfb_psi_plan_tables(const pfb_psi_plan *plan, int basis, int *tabs, int *Ntotx,
                        int *Ntoty);
/* Psi.dot(x, alphao): image -> coeffs; cells outside the pyramid blocks are
 * left untouched exactly like the reference.  pfb/operators/psi.py:284-296 */
int pfb_psi_dot(const pfb_psi_plan *plan, const void *x, void *alpha, void *stream);
/* Psi.hdot(alpha, xo): coeffs -> image (sum over bases, xo overwritten)
 * pfb/operators/psi.py:298-310.  If alpha2 != NULL the coefficients used are
 * 2*alpha - alpha2 (fuses pfb/opt/primal_dual.py:137 into the loader). */
int pfb_psi_hdot(const pfb_psi_plan *plan, const void *alpha, const void *alpha2,
                 void *xo, void *stream);

/* Natural-orientation coefficient layout.  The reference stores every sub-band
 * transposed w.r.t. the image (pfb/wavelets/wavelets.py:143-144, psi.py:200-218),
 * which costs a transposition per Psi.dot and per Psi.hdot.  A solver that owns
 * its dual variable for many iterations (primal_dual_optimised,
 * pfb/opt/primal_dual.py:117-166) may keep it in this layout instead and convert
 * on entry/exit: per (band, basis) one block of `stride` elements holding, for
 * every level, the sub-bands as [jx][jy] arrays (rows padded to 32 elements,
 * padding = 0), 'self' as the image itself.  Element-wise operators (dual
 * update / prox with the weights converted the same way) are layout-agnostic.
 * stride = 0: the plan's bases do not support it (db5, repeated bases). */
int pfb_psi_nat_stride(const pfb_psi_plan *plan, long long *stride);
/* slab cube (nband, nbasis, Nymax, Nxmax) -> natural layout (nband, nbasis, stride);
 * nband is given explicitly so the (nbasis, Nymax, Nxmax) weights convert with 1 */
int pfb_psi_to_nat(const pfb_psi_plan *plan, int nband, const void *alpha, void *nat,
                   void *stream);
/* natural layout -> pyramid cells of the slab cube (other cells untouched) */
int pfb_psi_from_nat(const pfb_psi_plan *plan, int nband, const void *nat, void *alpha,
                     void *stream);
/* Psi.dot / Psi.hdot on the natural layout (same arithmetic, no transposition);
 * nat2 != NULL: synthesis of 2*nat - nat2.  x must be 16-byte aligned. */
int pfb_psi_dot_nat(const pfb_psi_plan *plan, const void *x, void *nat, void *stream);
/* Analysis fused with the per-band half of the l21 dual update (pfb/opt/primal_dual.py:135-136,
 * pfb/prox/prox_21m.py:75-103): v_nat[b] = (vp_nat[b] + sigma * Psi^T x[b]) * f_nat, f_nat being
 * ONE band-shaped block (nbasis, stride) from pfb_dual_factor.  vp_nat must not alias v_nat;
 * padding cells of v_nat are not written (keep them zero). */
int pfb_psi_dot_nat_dual(const pfb_psi_plan *plan, const void *x, const void *vp_nat,
                         const void *f_nat, double sigma, void *v_nat, void *stream);
int pfb_psi_hdot_nat(const pfb_psi_plan *plan, const void *nat, const void *nat2,
                     void *xo, void *stream);

/* --------------------------------------------------------------- Hessian --
 * psf_convolve_cube / psf_convolve_slice  pfb/operators/psf.py:32-56, :11-29
 * hessian_psf_cube / _hessian_psf_slice   pfb/operators/hessian.py:254-281, :129-158
 * x, out, beam: (nband, nx, ny) real; psfhat: (nband, nxp, nyp/2+1) complex,
 * = r2c(ifftshift(psf)) as built by pfb/operators/gridder.py:712-714.
 * The plan owns what the reference keeps in xpad/xhat (here: a pruned,
 * transposed half-spectrum) and a transposed, 1/(nxp*nyp)-scaled copy of
 * psfhat.  Padded sizes: powers of two take the register-resident kernels; any
 * other even nyp / nxp whose transform lengths factor into 2,3,5,7,11,13 (the
 * ducc0 good_size lengths of pfb/workers/grid.py:276-284) take the mixed-radix
 * kernels (hess_mixed.cuh); PFB_ERR_FFT_SIZE otherwise (odd nyp, a larger prime
 * factor, or two transform lines beyond shared memory).  */
typedef struct pfb_hess_plan pfb_hess_plan;
int pfb_hess_plan_create(pfb_hess_plan **out, int nband, int nx, int ny, int nxp,
                         int nyp, int dtype);
int pfb_hess_plan_destroy(pfb_hess_plan *plan);
/* (re)load the PSF transfer function; call again whenever psfhat changes */
int pfb_hess_plan_set_psfhat(pfb_hess_plan *plan, const void *psfhat, void *stream);
/* Work distribution of the persistent kernels: 0 (default) = static round-robin,
 * 1 = items handed out by an atomic counter.  Dynamic costs a few percent on an
 * idle GPU and wins when a concurrent kernel holds SMs (the NCCL all-reduce
 * that the band-sharded solver overlaps with this operator). */
int pfb_hess_plan_set_dynamic(pfb_hess_plan *plan, int on);
/* out = conv(x)                                    beam NULL, add_x 0, has_wsum 0
 * out = beam*conv(beam*x) [/wsum] + sigmainv*x     add_x 1
 * out must not alias x. */
int pfb_hess_apply(pfb_hess_plan *plan, const void *x, const void *beam, int add_x,
                   double sigmainv, int has_wsum, double wsum, void *out,
                   void *stream);

/* PSFHAT creation on the device, pfb/operators/gridder.py:712-714:
 *   psfhat = r2c(iFs(psf, axes=(0,1)), axes=(0,1), forward=True, inorm=0)   per band
 * psf: (nband, nxp, nyp) real, peak at (nxp/2, nyp/2); psfhat: (nband, nxp, nyp/2+1) complex
 * (reference layout, ready for pfb_hess_plan_set_psfhat).  nxp, nyp even (what
 * pfb/workers/grid.py:276-284 produces) with the same factorisation limits as the Hessian
 * plans.  The forward row / column passes of the Hessian kernels are reused; the ifftshift
 * is applied as the exact sign (-1)^(kx+ky) on the spectrum.  Allocates and releases its own
 * scratch (nband*nxp*(nyp/2+1) complex) and therefore synchronises the stream. */
int pfb_psfhat_from_psf(const void *psf, int nband, int nxp, int nyp, int dtype, void *psfhat,
                        void *stream);

/* Which kernel family the plan selected (no reference counterpart; for tests and logs):
 * bit 0 power-of-two fast path, bit 1 mixed-radix plan (pfb/workers/grid.py:276-284 good_size
 * paddings), bit 2 / bit 3 its row / column passes run as (3|5|7) x 2^k peeled transforms,
 * bit 4 / bit 5 the 16384-point column / row kernels. */
int pfb_hess_plan_kernels(const pfb_hess_plan *plan);

/* PSF convolution with the primal step of primal_dual_optimised
 * (pfb/opt/primal_dual.py:138-150) fused into its last pass:
 *   x = P+( xp - tau * (2 w - wprev + psf_convolve(xp) - data) ),  out2 as pfb_norm_diff
 * (positivity 0 or 1; w, wprev = Psi(v) of this and the previous iteration, see
 * pfb_pd_primal_lin).  The convolution result is never written.  Available when
 * pfb_hess_plan_can_fuse_primal() is 1: the power-of-two fast paths (fp32 and fp64; every pass C
 * kernel of hess_fast.cuh takes the epilogue) and fp32 plans whose rows are (3|5|7) x 2^k points
 * (hess_peel.cuh); otherwise use pfb_hess_apply + pfb_pd_primal_lin.  ny even, arrays aligned to
 * two elements. */
int pfb_hess_plan_can_fuse_primal(const pfb_hess_plan *plan);
int pfb_hess_apply_primal(pfb_hess_plan *plan, const void *xp, const void *w, const void *wprev,
                          const void *data, double tau, int positivity, void *x, double *out2,
                          void *stream);
/* The same call in two halves: begin = the passes that need xp only (rows forward, columns
 * forward x psfhat x inverse; the result stays in the plan's scratch), end = the last pass with
 * the primal step.  A band-sharded solver issues begin right behind the previous primal step,
 * where it hides the all-reduce of the band-sum image (pfb_dual_update_rec), and end once the
 * synthesis w = Psi(v) of the new dual is there.  No other call on the plan in between. */
int pfb_hess_apply_primal_begin(pfb_hess_plan *plan, const void *xp, void *stream);
int pfb_hess_apply_primal_end(pfb_hess_plan *plan, const void *xp, const void *w, const void *wprev,
                              const void *data, double tau, int positivity, void *x, double *out2,
                              void *stream);

/* ---------------------------------------------------- solver loops --
 * The iteration loops themselves as library calls, for hosts that are not Python (the Python
 * package issues the same launch sequences from pfb_b200/opt/).  Device pointers, one
 * stream; scratch is allocated on entry and freed on exit; every call returns after its
 * last kernel has finished.  Band-sharded runs use the Python loops (NCCL lives there). */

/* primal_dual_optimised(x, v, lam, psiH, psi, L, prox_21m, l1weight, None, grad21, nu, sigma,
 * tol, maxit, positivity, gamma)  pfb/opt/primal_dual.py:90-171 with
 * grad21(x) = psf_convolve(x) - data (pfb/workers/spotless.py:259-263) and no reweighter.
 * x (nband, nx, ny) and v (nband, nbasis, Nymax, Nxmax) are updated in place; l1weight is
 * (nbasis, Nymax, Nxmax).  sigma <= 0: L / (2 gamma) / nu.  niter = iterations run,
 * eps = the last ||x - xp|| / ||x||.  Same kernels in the same order as the Python loop
 * (natural-orientation dual, synthesis by linearity, primal step fused into the last
 * Hessian pass where the plan allows). */
int pfb_pd_run(const pfb_psi_plan *psi, pfb_hess_plan *hess, void *x, void *v, const void *data,
               const void *l1weight, int nband, int nx, int ny, int nbasis, double lam, double L,
               double nu, double sigma, double gamma, double tol, int maxit, int positivity,
               int dtype, int *niter, double *eps, void *stream);
/* power_method(psf_convolve, imsize, b0, tol, maxit)  pfb/opt/power_method.py:11-46;
 * b holds b0 on entry and the normalised vector on exit. */
int pfb_power_method_run(pfb_hess_plan *hess, void *b, int nband, int nx, int ny, double tol,
                         int maxit, int dtype, double *beta, int *niter, void *stream);
/* pcg(A, b, x0, M, tol, maxit, minit, backtrack)  pfb/opt/pcg.py:58-131 with
 * A = hessian_psf_cube(beam, psfhat, sigmainv, wsum) (pfb/operators/hessian.py:254-281) and
 * M(x) = minv * x; x holds x0 on entry and the solution on exit; the line-search scalars stay
 * on the device (pfb_pcg_seg_*). */
int pfb_pcg_run(pfb_hess_plan *hess, const void *beam, double sigmainv, int has_wsum, double wsum,
                const void *b, void *x, double minv, int nband, int nx, int ny, double tol, int maxit,
                int minit, int backtrack, int dtype, int *niter, double *eps, void *stream);

/* ------------------------------------------------- CLEAN minor loops --
 * pfb/deconv/clark.py (clark :84-177, subminor :29-82, subtract :12-26) and
 * pfb/deconv/hogbom.py:8-72.  The searches are argmax decisions, so these kernels keep the
 * reference's evaluation order and do not contract multiply-subtract into FMA.          */

/* search[i] = (sum_b IR[b][i])^2 (search may be NULL); out2 = {argmax (first index on ties,
 * like numpy), search[argmax]}                       clark.py:116-120, 145-150; hogbom.py:26-29 */
int pfb_mfs_peak(const void *IR, int nband, long long npix, void *search, double *out2, int dtype,
                 void *stream);
/* Ip, Iq = np.where(IRsearch > thr) as ascending linear pixel indices (clark.py:127):
 * idx[0..min(count, cap)), *count = number found.  scratch_counts: ceil(npix / 2048) ints. */
int pfb_active_set(const void *search, long long npix, double thr, long long *idx, long long cap,
                   long long *count, int *scratch_counts, int dtype, void *stream);
/* subminor(IR[:, Ip, Iq], psf, Ip, Iq, model, wsums, gamma, th, maxit) in one launch
 * (clark.py:29-82): model updated in place; scratchA: nband * nact reals; info2 = {iterations
 * done, last peak}.  wsums: nband doubles on the device.  Needs nxp >= 2 nx - 1, nyp >= 2 ny - 1
 * (every PSF shift in range -- what the reference's active-set masking silently assumes). */
int pfb_clark_subminor(const void *IR, const void *psf, const long long *idx, long long nact,
                       void *model, const double *wsums, int nband, int nx, int ny, int nxp, int nyp,
                       double gamma, double th, int maxit, void *scratchA, double *info2, int dtype,
                       void *stream);
/* hogbom (hogbom.py:8-72) with the loop state on the device.  state: 8 + nband doubles:
 * [0] peak pixel, [1] IRmax, [2] done, [3] iterations k, [4] tol, [5] maxit.  start: first
 * search, tol = max(pf IRmax, threshold) and the first component; iterate: niter x
 * {IR -= gamma xhat psf(shifted); search; decide + next component}, no-ops once done is set --
 * the host enqueues batches and polls state[2].  wsums: nband reals (np.amax(PSF, axis=(1,2))). */
int pfb_hogbom_start(const void *IR, void *x, const void *wsums, int nband, long long npix,
                     double gamma, double pf, double threshold, int maxit, double *state, int dtype,
                     void *stream);
int pfb_hogbom_iterate(void *IR, const void *psf, void *x, const void *wsums, int nband, int nx, int ny,
                       int nxp, int nyp, double gamma, int niter, double *state, int dtype, void *stream);

/* ------------------------------------------------ worker glue (device) --
 * Pieces of the spotless major cycle between two minor cycles
 * (pfb/workers/spotless.py:217-260, 343-371), kept on the device so that model,
 * residual and dual never bounce through the host.                           */

/* stats3[3 s + {0,1,2}] = {mean, population std, max |x|} of segment s of x,
 * x = (nseg, m); fp64 accumulation, fixed order, two passes for the std like
 * numpy.  np.std(np.sum(psiHoutvar, axis=0), axis=(-1,-2)) = rms_comps with one
 * segment per basis (spotless.py:219-223, 360-363); np.std / np.abs().max() of
 * residual_mfs with one segment (spotless.py:229-230, 343-344). */
int pfb_seg_stats(const void *x, int nseg, long long m, double *stats3, int dtype, void *stream);
/* out += scale * x * y (y NULL: out += scale * x), complex arrays as 2n reals with y NULL:
 * the per-dataset accumulation of dds2cubes, pfb/utils/misc.py:698-714 */
int pfb_accum_mul(void *out, const void *x, const void *y, double scale, long long n, int dtype,
                  void *stream);
/* Restoring ("cheap f5"): the pieces of convolve2gaussres, pfb/utils/misc.py:186-238,
 * that are not the PSF-convolution kernels themselves.
 * pfb_gauss2d: Gaussian2D (misc.py:109-138) on the caller's coordinate grids (fp64, n values each):
 *   out = exp(-c (a00 x^2 + a01 x y + a11 y^2)) where x^2 + y^2 <= extent, else 0 (the caller forms
 *   the 2 x 2 matrix, a01 = A[0,1] + A[1,0], c = 2 sqrt(2 ln 2), extent = (nsigma Smaj)^2).
 * pfb_spec_ratio: convkernhat = gausskernhat / thiskernhat where |thiskernhat| > 0, else 0
 *   (misc.py:224-226); complex arrays, num (nnum values) broadcast over the blocks of den (nden). */
int pfb_gauss2d(const double *xx, const double *yy, long long n, double a00, double a01, double a11,
                double c, double extent, void *out, int dtype, void *stream);
int pfb_spec_ratio(const void *num, const void *den, void *out, long long nnum, long long nden, int dtype,
                   void *stream);
/* pfb_seg_stats without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * pfb_seg_stats; the partial sums live in a static device array of the library (303 104
 * bytes per GPU that loads it, held until exit) instead of a cudaMalloc'ed workspace, which
 * limits nseg to 64: more segments return PFB_ERR_BAD_ARG with a message naming the limit.
 * The call itself allocates nothing (no cudaMalloc / cudaFree, no events).  All calls share
 * the array: calls that overlap on the GPU (different streams) race; calls on one stream,
 * or otherwise ordered, are safe, and nothing has to be cleared between them. */
int pfb_seg_stats_static(const void *x, int nseg, long long m, double *stats3, int dtype,
                         void *stream);
/* l1reweight_func, pfb/utils/misc.py:1070-1080, after the band sum:
 * weight[b][i] = (1 + rmsfactor) / (1 + |bsum[b][i]|^alpha / rms_comps[b]^alpha),
 * bsum, weight = (nbasis, nslab), rms_comps = (nbasis) in the array precision. */
int pfb_l1reweight(const void *bsum, const void *rms_comps, int nbasis, long long nslab,
                   double rmsfactor, double alpha, void *weight, int dtype, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PFB_B200_H */
