/* CPU oracle (plain C + OpenMP) for the pfb-imaging minor-cycle kernels.
 *
 * TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.  Never linked into or
 * called from the product library.  Each function restates the algorithm of
 * the reference function cited above it (paths relative to /root/reference).
 * Arithmetic is fp64 with the reference's summation order; compile with
 * -ffp-contract=off so no FMA contraction changes the rounding.
 *
 * round_f32 != 0 emulates the reference's single-precision mode, in which the
 * numba kernels still accumulate in fp64 (fp64 scratch/taps,
 * pfb/operators/psi.py:114-117) but every store into a caller-owned fp32
 * array rounds to fp32.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline double rnd(double v, int round_f32) {
    return round_f32 ? (double)(float)v : v;
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* pfb/wavelets/wavelets.py:20-27 */
int orc_coeff_size(int nsignal, int nfilter) { return (nsignal + nfilter - 1) / 2; }
int orc_signal_size(int ncoeff, int nfilter) { return 2 * ncoeff - nfilter + 2; }

/* pfb/wavelets/wavelets.py:30-95 downsampling_convolution, step = 2,
 * mode 'zero'.  The four boundary loops of the reference all evaluate
 *   out[o] = sum_{j ascending, 0<=i-j<N} filt[j]*in[i-j],  i = 2o+1,
 * for o in [0,(N+F-1)/2); restated as one loop with the same j order. */
static void down_conv(const double *in, long is, int N, const double *filt,
                      int F, double *out, long os) {
    int o = 0;
    for (int i = 1; i < N + F - 1; i += 2, ++o) {
        int jlo = i - N + 1;
        if (jlo < 0) jlo = 0;
        int jhi = i < F - 1 ? i : F - 1;
        double s = 0.0;
        for (int j = jlo; j <= jhi; ++j) s += filt[j] * in[(long)(i - j) * is];
        out[(long)o * os] = s;
    }
}

/* pfb/wavelets/wavelets.py:98-123 upsampling_convolution_valid_sf:
 * accumulates into out (length O). */
static void up_conv(const double *in, long is, int N, const double *filt, int F,
                    double *out, long os, int O) {
    int Fo2 = F / 2;
    int o = 0;
    for (int i = Fo2 - 1; i < N && o < O; ++i, o += 2) {
        double se = 0.0, so = 0.0;
        for (int j = 0; j < Fo2; ++j) {
            double v = in[(long)(i - j) * is];
            so += filt[2 * j + 1] * v;
            se += filt[2 * j] * v;
        }
        out[(long)o * os] += se;
        out[(long)(o + 1) * os] += so;
    }
}

/* blocked transpose out(m x n, ldo) = in(n x m, ldi)^T;
 * pfb/wavelets/wavelets.py:6-17 copyT */
static void transpose(const double *in, long ldi, int n, int m, double *out,
                      long ldo) {
    const int B = 32;
#pragma omp parallel for schedule(static)
    for (int ib = 0; ib < m; ib += B)
        for (int jb = 0; jb < n; jb += B) {
            int ie = ib + B < m ? ib + B : m, je = jb + B < n ? jb + B : n;
            for (int i = ib; i < ie; ++i)
                for (int j = jb; j < je; ++j)
                    out[(long)i * ldo + j] = in[(long)j * ldi + i];
        }
}

/* pfb/wavelets/wavelets.py:126-171 dwt2d_level.
 * image (n0 x n1, ldi) -> coeffs block (2*C1 x 2*C0, ldc), transposed output.
 * cbuff (n0 x 2*C1), cbuffT (2*C1 x n0) are scratch.  approx (C0 x C1) gets
 * coeffs[0:C1,0:C0]^T (read back AFTER rounding, as the reference does). */
static void dwt2d_level(const double *image, long ldi, int n0, int n1,
                        double *coeffs, long ldc, int C0, int C1,
                        double *cbuff, double *cbuffT, const double *dec_lo,
                        const double *dec_hi, int F, double *approx,
                        int round_f32) {
    long ldb = 2L * C1;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n0; ++i) {
        down_conv(image + (long)i * ldi, 1, n1, dec_lo, F, cbuff + i * ldb, 1);
        down_conv(image + (long)i * ldi, 1, n1, dec_hi, F, cbuff + i * ldb + C1, 1);
    }
    transpose(cbuff, ldb, n0, 2 * C1, cbuffT, n0);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < 2 * C1; ++i) {
        double *row = coeffs + (long)i * ldc;
        down_conv(cbuffT + (long)i * n0, 1, n0, dec_lo, F, row, 1);
        down_conv(cbuffT + (long)i * n0, 1, n0, dec_hi, F, row + C0, 1);
        if (round_f32)
            for (int j = 0; j < 2 * C0; ++j) row[j] = (double)(float)row[j];
    }
    transpose(coeffs, ldc, C1, C0, approx, C1);
}

/* pfb/wavelets/wavelets.py:174-213 dwt2d.  highx/highy = ix[i][1], iy[i][1]
 * of pfb/operators/psi.py:80-94; sx/sy = coefficient sizes per level. */
void orc_dwt2d(const double *image, int nx, int ny, double *coeffs, long ldc,
               const int *highx, const int *highy, const int *sx, const int *sy,
               const double *dec_lo, const double *dec_hi, int F, int nlevel,
               int round_f32) {
    long n = (long)(2 * sx[0] + 2) * (2 * sy[0] + 2);
    double *cbuff = (double *)malloc(sizeof(double) * n);
    double *cbuffT = (double *)malloc(sizeof(double) * n);
    double *apA = (double *)malloc(sizeof(double) * n);
    double *apB = (double *)malloc(sizeof(double) * n);
    const double *cur = image;
    long ldcur = ny;
    int n0 = nx, n1 = ny;
    for (int i = 0; i < nlevel; ++i) {
        int lowx = highx[i] - 2 * sx[i], lowy = highy[i] - 2 * sy[i];
        double *nxt = (i % 2 == 0) ? apA : apB;
        dwt2d_level(cur, ldcur, n0, n1, coeffs + (long)lowy * ldc + lowx, ldc,
                    sx[i], sy[i], cbuff, cbuffT, dec_lo, dec_hi, F, nxt,
                    round_f32);
        cur = nxt;
        n0 = sx[i];
        n1 = sy[i];
        ldcur = n1;
    }
    free(cbuff); free(cbuffT); free(apA); free(apB);
}

/* pfb/wavelets/wavelets.py:216-257 idwt2d_level.
 * coeffs block (2*C1 x 2*C0, ldc) -> image (n0o x n1o, ldi) (overwritten). */
static void idwt2d_level(const double *coeffs, long ldc, int C0, int C1,
                         double *image, long ldi, int n0o, int n1o,
                         double *cbuff, double *cbuffT, const double *rec_lo,
                         const double *rec_hi, int F) {
    long ldt = 2L * C0, ldb = 2L * C1;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < 2 * C1; ++i) {
        double *row = cbuffT + i * ldt;
        memset(row, 0, sizeof(double) * ldt);
        up_conv(coeffs + (long)i * ldc, 1, C0, rec_lo, F, row, 1, (int)ldt);
        up_conv(coeffs + (long)i * ldc + C0, 1, C0, rec_hi, F, row, 1, (int)ldt);
    }
    transpose(cbuffT, ldt, 2 * C1, 2 * C0, cbuff, ldb);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n0o; ++i) {
        double *row = image + (long)i * ldi;
        memset(row, 0, sizeof(double) * n1o);
        up_conv(cbuff + i * ldb, 1, C1, rec_lo, F, row, 1, n1o);
        up_conv(cbuff + i * ldb + C1, 1, C1, rec_hi, F, row, 1, n1o);
    }
}

/* pfb/wavelets/wavelets.py:260-315 idwt2d.  image is an (nx x ny) scratch
 * that holds the reconstruction on return; coeffs is not modified. */
void orc_idwt2d(const double *coeffs, long ldc, int Ntotx, int Ntoty,
                double *image, int nx, int ny, const int *highx,
                const int *highy, const int *sx, const int *sy, const int *spx,
                const int *spy, const double *rec_lo, const double *rec_hi,
                int F, int nlevel) {
    double *alpha = (double *)malloc(sizeof(double) * (long)Ntotx * Ntoty);
    long n = (long)(2 * sx[0] + 2) * (2 * sy[0] + 2);
    double *cbuff = (double *)malloc(sizeof(double) * n);
    double *cbuffT = (double *)malloc(sizeof(double) * n);
    long lda = Ntotx;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < Ntoty; ++i)
        memcpy(alpha + i * lda, coeffs + (long)i * ldc, sizeof(double) * Ntotx);
    for (int i = nlevel - 1; i >= 0; --i) {
        int nax = sx[i], nay = sy[i];
        int lowx = highx[i] - 2 * nax, lowy = highy[i] - 2 * nay;
        if (i < nlevel - 1) /* alpha[lowy:lowy+nay, lowx:lowx+nax] = image[0:nax,0:nay].T */
            transpose(image, ny, nax, nay, alpha + (long)lowy * lda + lowx, lda);
        idwt2d_level(alpha + (long)lowy * lda + lowx, lda, nax, nay, image, ny,
                     spx[i], spy[i], cbuff, cbuffT, rec_lo, rec_hi, F);
    }
    free(alpha); free(cbuff); free(cbuffT);
}

/* pfb/operators/psi.py:187-218 psi_band.dot for ONE band.
 * basis_F[b] = 0 for 'self', else the filter length (2..10); taps holds
 * 4*10 doubles per basis: dec_lo, dec_hi, rec_lo, rec_hi (each padded to 10).
 * tabs holds per basis 6*nlevel ints: highx, highy, sx, sy, spx, spy.
 * alphao (nbasis, Nymax, Nxmax); cells outside the pyramid are NOT written. */
void orc_psi_band_dot(const double *x, int nx, int ny, double *alphao,
                      int nbasis, int Nymax, int Nxmax, const int *basis_F,
                      const double *taps, const int *tabs, int nlevel,
                      int round_f32) {
    for (int b = 0; b < nbasis; ++b) {
        double *slab = alphao + (long)b * Nymax * Nxmax;
        if (basis_F[b] == 0) { /* copyT(x, alphao[i, 0:Ny, 0:Nx]) */
            transpose(x, ny, nx, ny, slab, Nxmax);
            continue;
        }
        const int *t = tabs + (long)b * 6 * nlevel;
        orc_dwt2d(x, nx, ny, slab, Nxmax, t, t + nlevel, t + 2 * nlevel,
                  t + 3 * nlevel, taps + b * 40, taps + b * 40 + 10, basis_F[b],
                  nlevel, round_f32);
    }
}

/* pfb/operators/psi.py:221-256 psi_band.hdot for ONE band: xo = sum over
 * bases of the reconstructions; in fp32 mode every `xo += image` rounds. */
void orc_psi_band_hdot(const double *alpha, double *xo, int nx, int ny,
                       int nbasis, int Nymax, int Nxmax, const int *basis_F,
                       const int *Ntotx, const int *Ntoty, const double *taps,
                       const int *tabs, int nlevel, int round_f32) {
    double *image = (double *)malloc(sizeof(double) * (long)nx * ny);
    for (long k = 0; k < (long)nx * ny; ++k) xo[k] = 0.0;
    for (int b = 0; b < nbasis; ++b) {
        const double *slab = alpha + (long)b * Nymax * Nxmax;
        if (basis_F[b] == 0) {
            transpose(slab, Nxmax, ny, nx, image, ny);
        } else {
            const int *t = tabs + (long)b * 6 * nlevel;
            orc_idwt2d(slab, Nxmax, Ntotx[b], Ntoty[b], image, nx, ny, t,
                       t + nlevel, t + 2 * nlevel, t + 3 * nlevel,
                       t + 4 * nlevel, t + 5 * nlevel, taps + b * 40 + 20,
                       taps + b * 40 + 30, basis_F[b], nlevel);
        }
#pragma omp parallel for schedule(static)
        for (long k = 0; k < (long)nx * ny; ++k)
            xo[k] = rnd(xo[k] + image[k], round_f32);
    }
    free(image);
}

/* pfb/prox/prox_21m.py:30-61 prox_21m_numba (mode 0, sum over band) and
 * pfb/prox/prox_21.py:25-52 prox_21_numba (mode 1, l2 norm over band).
 * v, result (nband, n) with n = nbasis*nymax*nxmax flattened; weight (n). */
void orc_prox_21(const double *v, double *result, const double *weight,
                 int nband, long n, double lam, double sigma, int mode,
                 int round_f32) {
#pragma omp parallel for schedule(static)
    for (long k = 0; k < n; ++k) {
        double s = 0.0;
        for (int b = 0; b < nband; ++b) {
            double vv = v[b * n + k];
            s += mode ? vv * vv : vv;
        }
        double a = (mode ? sqrt(s) : s) / sigma;
        if (a == 0.0) {
            for (int b = 0; b < nband; ++b) result[b * n + k] = 0.0;
            continue;
        }
        double absv = fabs(a);
        double soft = fmax(absv - lam * weight[k] / sigma, 0.0);
        for (int b = 0; b < nband; ++b)
            result[b * n + k] = rnd(v[b * n + k] * soft / absv / sigma, round_f32);
    }
}

/* pfb/prox/prox_21m.py:75-103 dual_update_numba (mode 0) and
 * pfb/prox/prox_21.py:66-91 (mode 1).  In place on v (pre-loaded with
 * psi.dot(xp)); vp read only.  In fp32 mode vtilde = vp + sigma*v is formed
 * in fp64 by the reference only if sigma is a Python float and numba promotes;
 * numba types float32 + float64*float32 as float64, the store into v rounds. */
void orc_dual_update(const double *vp, double *v, const double *weight,
                     int nband, long n, double lam, double sigma, int mode,
                     int round_f32) {
#pragma omp parallel for schedule(static)
    for (long k = 0; k < n; ++k) {
        double s = 0.0;
        double vt[64];
        for (int b = 0; b < nband; ++b) {
            vt[b] = vp[b * n + k] + sigma * v[b * n + k];
            s += mode ? vt[b] * vt[b] : vt[b];
        }
        double a = fabs((mode ? sqrt(s) : s) / sigma);
        if (a != 0.0) {
            double soft = fmax(a - lam * weight[k] / sigma, 0.0);
            double fac = 1.0 - soft / a;
            for (int b = 0; b < nband; ++b) {
                /* v[...] = vtilde (store rounds), then v[...] *= fac (rounds) */
                double t = rnd(vt[b], round_f32);
                v[b * n + k] = rnd(t * fac, round_f32);
            }
        } else {
            for (int b = 0; b < nband; ++b) v[b * n + k] = rnd(vt[b], round_f32);
        }
    }
}

/* pfb/utils/misc.py:1316-1351 norm_diff: serial fp64 accumulation, the
 * difference formed in the input dtype. */
double orc_norm_diff(const double *x, const double *xp, long n, int round_f32) {
    double num = 0.0, den = 1e-12;
    for (long k = 0; k < n; ++k) {
        double d = rnd(x[k] - xp[k], round_f32);
        d = rnd(d * d, round_f32);
        num += d;
        den += rnd(x[k] * x[k], round_f32);
    }
    return sqrt(num / den);
}
