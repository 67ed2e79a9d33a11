// l21 prox / dual update, fused axpy + reduction kernels (sm_100a).
//
// All of these are single-pass streaming kernels bounded by HBM bandwidth:
// 128-bit coalesced loads/stores, grid = multiple of the 148 SMs, no shared
// memory except the block reduction.  The cross-band reduction of the prox is
// held in registers (one thread owns a coefficient index and walks the bands),
// scalar math in fp64 exactly as the reference's numba kernels promote
// (pfb/prox/prox_21m.py:30-61, 75-103; pfb/prox/prox_21.py:25-91).
#include "common.cuh"
#include <cstdarg>
#include <map>
#include <vector>

namespace pfb {

thread_local std::string g_last_error;
long long g_launch_count = 0;

int g_profile_on = 0;

#ifndef PFB_EMU
// one record per launch while profiling is on
struct ProfRec { const void *func; cudaEvent_t e0, e1; };
static std::vector<ProfRec> g_prof;

void profile_begin(const void *func, cudaStream_t st) {
    ProfRec r;
    r.func = func;
    cudaEventCreate(&r.e0);
    cudaEventCreate(&r.e1);
    cudaEventRecord(r.e0, st);
    g_prof.push_back(r);
}
void profile_end(cudaStream_t st) { cudaEventRecord(g_prof.back().e1, st); }
#endif

int fail(int status, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return status;
}

template <class T, int V> struct alignas(sizeof(T) * V) Pack { T v[V]; };

template <class T, int V>
__device__ __forceinline__ Pack<T, V> ld(const T *p) {
    return *reinterpret_cast<const Pack<T, V> *>(p);
}
template <class T, int V>
__device__ __forceinline__ void st(T *p, const Pack<T, V> &x) {
    *reinterpret_cast<Pack<T, V> *>(p) = x;
}

// ---------------------------------------------------------------------------
// prox_21m_numba / prox_21_numba:  result = v * soft(|S|)/|S| / sigma
// NB > 0: band count known at compile time -> all NB loads of a thread are
// issued back to back (no per-band branch), which is what keeps enough bytes
// in flight to saturate HBM; NB == 0: generic loop for nband > 8.
template <class T, int V, int NB>
__global__ void __launch_bounds__(256)
k_prox21(const T *__restrict__ v, T *__restrict__ res, const T *__restrict__ w,
         int nband, long n, double lam, double sigma, int mode) {
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long)gridDim.x * blockDim.x) {
        long k = i * V;
        Pack<T, V> c[NB > 0 ? NB : 1];
        double s[V];
#pragma unroll
        for (int e = 0; e < V; ++e) s[e] = 0.0;
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) c[b] = ld<T, V>(v + b * n + k);
#pragma unroll
            for (int b = 0; b < NB; ++b)
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)c[b].v[e];
                    s[e] += mode ? x * x : x;
                }
        } else {
            for (int b = 0; b < nband; ++b) {
                Pack<T, V> t = ld<T, V>(v + b * n + k);
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)t.v[e];
                    s[e] += mode ? x * x : x;
                }
            }
        }
        Pack<T, V> wk = ld<T, V>(w + k);
        double f[V];
#pragma unroll
        for (int e = 0; e < V; ++e) {
            double a = (mode ? sqrt(s[e]) : s[e]) / sigma;
            if (a == 0.0) {
                f[e] = 0.0;
            } else {
                double absv = fabs(a);
                double soft = fmax(absv - lam * (double)wk.v[e] / sigma, 0.0);
                f[e] = soft / absv / sigma;
            }
        }
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                Pack<T, V> o;
#pragma unroll
                for (int e = 0; e < V; ++e) o.v[e] = (T)((double)c[b].v[e] * f[e]);
                st<T, V>(res + b * n + k, o);
            }
        } else {
            for (int b = 0; b < nband; ++b) {
                Pack<T, V> t = ld<T, V>(v + b * n + k), o;
#pragma unroll
                for (int e = 0; e < V; ++e) o.v[e] = (T)((double)t.v[e] * f[e]);
                st<T, V>(res + b * n + k, o);
            }
        }
    }
}

// Shrink factor of the dual update, 1 - max(a - lam w/sigma, 0)/a with
// a = |r|/sigma and r = band sum (21m) or l2 norm over bands (21).  Numerator and
// denominator are multiplied by sigma, which leaves ONE fp64 division per
// coefficient instead of three: with one band per GPU the three divisions made
// the (then fp64-pipe bound) apply kernel twice as slow as its memory traffic.
__device__ __forceinline__ double shrink_factor(double s, double lam_w, int mode) {
    const double r = mode ? sqrt(s) : fabs(s);
    return r != 0.0 ? 1.0 - fmax(r - lam_w, 0.0) / r : 1.0;
}

// ---------------------------------------------------------------------------
// dual_update_numba, optionally fused with the `vp = 2 v - vp` that follows it
// in primal_dual_optimised (pfb/opt/primal_dual.py:136-137).
//   vt = vp + sigma*v;  a = |red(vt)|/sigma;  v = vt * (1 - max(a - lam w/sigma,0)/a)
//   u  = 2 v - vp   (written to `u` when non-null; u may alias vp)
// The raw vp/v packs stay in registers (vt is recomputed for the store instead
// of being kept in fp64) so a thread has 2*NB 16-byte loads in flight.
template <class T, int V, int NB>
__global__ void __launch_bounds__(256, 2)      // <= 128 registers: two CTAs (128 KB of loads) in flight per SM
k_dual(const T *vp, T *v, const T *__restrict__ w, T *u, int nband, long n,
       double lam, double sigma, int mode) {
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long)gridDim.x * blockDim.x) {
        long k = i * V;
        Pack<T, V> pa[NB > 0 ? NB : 1], pc[NB > 0 ? NB : 1];
        double s[V];
#pragma unroll
        for (int e = 0; e < V; ++e) s[e] = 0.0;
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                pa[b] = ld<T, V>(vp + b * n + k);
                pc[b] = ld<T, V>(v + b * n + k);
            }
#pragma unroll
            for (int b = 0; b < NB; ++b)
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)pa[b].v[e] + sigma * (double)pc[b].v[e];
                    s[e] += mode ? x * x : x;
                }
        } else {
            for (int b = 0; b < nband; ++b) {
                Pack<T, V> a = ld<T, V>(vp + b * n + k), c = ld<T, V>(v + b * n + k);
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)a.v[e] + sigma * (double)c.v[e];
                    s[e] += mode ? x * x : x;
                }
            }
        }
        Pack<T, V> wk = ld<T, V>(w + k);
        double f[V];
#pragma unroll
        for (int e = 0; e < V; ++e) {
            f[e] = shrink_factor(s[e], lam * (double)wk.v[e], mode);
        }
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                Pack<T, V> o, uo;
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)pa[b].v[e] + sigma * (double)pc[b].v[e];
                    o.v[e] = (T)(x * f[e]);
                    uo.v[e] = (T)2 * o.v[e] - pa[b].v[e];
                }
                st<T, V>(v + b * n + k, o);
                if (u) st<T, V>(u + b * n + k, uo);
            }
        } else {
            for (int b = 0; b < nband; ++b) {
                Pack<T, V> a = ld<T, V>(vp + b * n + k), c = ld<T, V>(v + b * n + k), o, uo;
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    o.v[e] = (T)(((double)a.v[e] + sigma * (double)c.v[e]) * f[e]);
                    uo.v[e] = (T)2 * o.v[e] - a.v[e];
                }
                st<T, V>(v + b * n + k, o);
                if (u) st<T, V>(u + b * n + k, uo);
            }
        }
    }
}

// ---------------------------------------------------------------------------
// Band-sharded (multi-GPU) split of the dual update.  Each rank owns nband
// local bands:
//   k_dual_partial : bsum[k] = red_local(vp + sigma*v)        (v is NOT modified)
//   (NCCL all-reduce SUM of bsum over ranks, issued by the host)
//   k_dual_apply   : v <- (vp + sigma*v) * factor(bsum), u <- 2 v - vp
// vt is recomputed by apply instead of being stored by partial: 2*nband*C fewer
// bytes per iteration.  NB as in k_dual (0 = run-time band count).
template <class T, int V, int NB>
__global__ void __launch_bounds__(256, 2)
k_dual_partial(const T *__restrict__ vp, const T *__restrict__ v, T *__restrict__ bsum, int nband,
               long n, long bs, double sigma, int mode) {
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long)gridDim.x * blockDim.x) {
        long k = i * V;
        double s[V];
#pragma unroll
        for (int e = 0; e < V; ++e) s[e] = 0.0;
        if (NB > 0) {
            Pack<T, V> pa[NB > 0 ? NB : 1], pc[NB > 0 ? NB : 1];
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                pa[b] = ld<T, V>(vp + b * bs + k);
                pc[b] = ld<T, V>(v + b * bs + k);
            }
#pragma unroll
            for (int b = 0; b < NB; ++b)
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)pa[b].v[e] + sigma * (double)pc[b].v[e];
                    s[e] += mode ? x * x : x;
                }
        } else {
            for (int b = 0; b < nband; ++b) {
                Pack<T, V> a = ld<T, V>(vp + b * bs + k), c = ld<T, V>(v + b * bs + k);
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)a.v[e] + sigma * (double)c.v[e];
                    s[e] += mode ? x * x : x;
                }
            }
        }
        Pack<T, V> so;
#pragma unroll
        for (int e = 0; e < V; ++e) so.v[e] = (T)s[e];
        st<T, V>(bsum + k, so);
    }
}

template <class T, int V, int NB>
__global__ void __launch_bounds__(256, 2)
k_dual_apply(const T *vp, T *v, const T *__restrict__ bsum,
             const T *__restrict__ w, T *u, int nband, long n, long bs, double lam,
             double sigma, int mode) {
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long)gridDim.x * blockDim.x) {
        long k = i * V;
        Pack<T, V> pa[NB > 0 ? NB : 1], pc[NB > 0 ? NB : 1];
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                pa[b] = ld<T, V>(vp + b * bs + k);
                pc[b] = ld<T, V>(v + b * bs + k);
            }
        }
        Pack<T, V> sk = ld<T, V>(bsum + k), wk = ld<T, V>(w + k);
        double f[V];
#pragma unroll
        for (int e = 0; e < V; ++e) {
            f[e] = shrink_factor((double)sk.v[e], lam * (double)wk.v[e], mode);
        }
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                Pack<T, V> o, uo;
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    double x = (double)pa[b].v[e] + sigma * (double)pc[b].v[e];
                    o.v[e] = (T)(x * f[e]);
                    uo.v[e] = (T)2 * o.v[e] - pa[b].v[e];
                }
                st<T, V>(v + b * bs + k, o);
                if (u) st<T, V>(u + b * bs + k, uo);
            }
        } else {
            for (int b = 0; b < nband; ++b) {
                Pack<T, V> a = ld<T, V>(vp + b * bs + k), c = ld<T, V>(v + b * bs + k), o, uo;
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    o.v[e] = (T)(((double)a.v[e] + sigma * (double)c.v[e]) * f[e]);
                    uo.v[e] = (T)2 * o.v[e] - a.v[e];
                }
                st<T, V>(v + b * bs + k, o);
                if (u) st<T, V>(u + b * bs + k, uo);
            }
        }
    }
}

// ---------------------------------------------------------------------------
// Band-sharded dual update WITHOUT a coefficient-sized exchange (21m / band-sum mode).
// The analysis is linear and identical for every band, and the shrink factor f is
// the same for all bands of a coefficient, so the cross-band sum obeys a recursion
// every rank can run on its own:
//     S_k   = sum_b (v_{k-1,b} + sigma Psi^T x_b) = Vbar_{k-1} + sigma Psi^T (sum_b x_b)
//     Vbar_k = sum_b v_{k,b} = f_k S_k
// Only xbar = sum_b x_b -- ONE image, nbasis x smaller than the coefficient slab of
// pfb/prox/prox_21m.py:99 -- is all-reduced; ubar = Psi^T xbar is computed by every
// rank.  This kernel:  S = vbar + sigma ubar;  f = shrink(S);  v_b <- (vp_b + sigma v_b) f
// for the local bands;  vbar <- f S.  (The host re-synchronises vbar with an exact
// all-reduce of sum_b v_b every few dozen iterations to bound rounding drift.)
template <class T, int V, int NB>
__global__ void __launch_bounds__(256, 2)
k_dual_rec(const T *vp, T *v, T *vbar, const T *__restrict__ ubar, const T *__restrict__ w,
           int nband, long n, double lam, double sigma) {
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long)gridDim.x * blockDim.x) {
        long k = i * V;
        Pack<T, V> pa[NB > 0 ? NB : 1], pc[NB > 0 ? NB : 1];
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                pa[b] = ld<T, V>(vp + b * n + k);
                pc[b] = ld<T, V>(v + b * n + k);
            }
        }
        Pack<T, V> vb = ld<T, V>(vbar + k), ub = ld<T, V>(ubar + k), wk = ld<T, V>(w + k), so;
        double f[V];
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const double S = (double)vb.v[e] + sigma * (double)ub.v[e];
            f[e] = shrink_factor(S, lam * (double)wk.v[e], 0);
            so.v[e] = (T)(S * f[e]);
        }
        st<T, V>(vbar + k, so);
        if (NB > 0) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                Pack<T, V> o;
#pragma unroll
                for (int e = 0; e < V; ++e)
                    o.v[e] = (T)(((double)pa[b].v[e] + sigma * (double)pc[b].v[e]) * f[e]);
                st<T, V>(v + b * n + k, o);
            }
        } else {
            for (int b = 0; b < nband; ++b) {
                Pack<T, V> a = ld<T, V>(vp + b * n + k), c = ld<T, V>(v + b * n + k), o;
#pragma unroll
                for (int e = 0; e < V; ++e)
                    o.v[e] = (T)(((double)a.v[e] + sigma * (double)c.v[e]) * f[e]);
                st<T, V>(v + b * n + k, o);
            }
        }
    }
}

// The band-independent half of k_dual_rec on its own: S = vbar + sigma ubar; f = shrink(S);
// vbar <- f S; fout = f.  The per-band half, v_b = (vp_b + sigma u_b) f, is then the epilogue of
// the analysis kernels (pfb_psi_dot_nat_dual), so u_b never exists in memory.
template <class T, int V>
__global__ void __launch_bounds__(256)
k_dual_factor(T *vbar, const T *__restrict__ ubar, const T *__restrict__ w, T *__restrict__ fout, long n,
              double lam, double sigma) {
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
        long k = i * V;
        Pack<T, V> vb = ld<T, V>(vbar + k), ub = ld<T, V>(ubar + k), wk = ld<T, V>(w + k), so, fo;
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const double S = (double)vb.v[e] + sigma * (double)ub.v[e];
            const double f = shrink_factor(S, lam * (double)wk.v[e], 0);
            so.v[e] = (T)(S * f);
            fo.v[e] = (T)f;
        }
        st<T, V>(vbar + k, so);
        st<T, V>(fout + k, fo);
    }
}

// ---------------------------------------------------------------------------
// out = a*x + b*y  (numexpr '2.0*v - vp', 'xp - tau*xout' of
// pfb/opt/primal_dual.py:137,140 and the axpys of pfb/opt/pcg.py:92-104);
// arithmetic in the array dtype (NumPy weak-scalar promotion).
template <class T, int V>
__global__ void __launch_bounds__(256)
k_axpby(T *out, T a, const T *x, T b, const T *y, long n) {
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long)gridDim.x * blockDim.x) {
        Pack<T, V> p = ld<T, V>(x + i * V), q = ld<T, V>(y + i * V), o;
#pragma unroll
        for (int e = 0; e < V; ++e) o.v[e] = a * p.v[e] + b * q.v[e];
        st<T, V>(out + i * V, o);
    }
}

// ---------------------------------------------------------------------------
// Deterministic two-stage reductions (fp64 accumulation, fixed order).
#ifdef PFB_EMU
static const int kRedBlocks = 3;  // emulator: same code, tiny grid
#else
static const int kRedBlocks = kNumSM * 4;
#endif
static const int kRedMaxK = 4;

__device__ __forceinline__ double warp_sum(double x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    return x;
}

// block-reduce K values; thread 0 stores them to part[blockIdx.x*K + k]
template <int K>
__device__ __forceinline__ void block_store(double (&acc)[K], double *part) {
    __shared__ double sh[kRedMaxK][32];
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double s = warp_sum(acc[k]);
        if (lane == 0) sh[k][wid] = s;
    }
    __syncthreads();
    if (wid == 0) {
        int nw = (blockDim.x + 31) >> 5;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double s = lane < nw ? sh[k][lane] : 0.0;
            s = warp_sum(s);
            if (lane == 0) part[(long)blockIdx.x * K + k] = s;
        }
    }
}

// out[k] = sum_b part[b*K + k] in fixed order (single block of 256 threads)
__global__ void k_red_final(const double *__restrict__ part, int nblocks, int K,
                            double *out) {
    __shared__ double sh[256];
    for (int k = 0; k < K; ++k) {
        double s = 0.0;
        for (int b = threadIdx.x; b < nblocks; b += blockDim.x) s += part[(long)b * K + k];
        sh[threadIdx.x] = s;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) out[k] = sh[0];
        __syncthreads();
    }
}

// out[0] = sum a*b ; if c != null also out[1] = sum c*d
template <class T>
__global__ void __launch_bounds__(256)
k_dot(const T *__restrict__ a, const T *__restrict__ b, const T *__restrict__ c,
      const T *__restrict__ d, long n, double *part) {
    double acc[2] = {0.0, 0.0};
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long)gridDim.x * blockDim.x) {
        acc[0] += (double)a[i] * (double)b[i];
        if (c) acc[1] += (double)c[i] * (double)d[i];
    }
    block_store<2>(acc, part);
}

// norm_diff of pfb/utils/misc.py:1316-1351: part -> {sum (x-xp)^2, sum x^2},
// difference and squares formed in T, accumulated in fp64.
template <class T>
__global__ void __launch_bounds__(256)
k_norm_diff(const T *__restrict__ x, const T *__restrict__ xp, long n, double *part) {
    double acc[2] = {0.0, 0.0};
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long)gridDim.x * blockDim.x) {
        T xi = x[i], d = xi - xp[i];
        acc[0] += (double)(d * d);
        acc[1] += (double)(xi * xi);
    }
    block_store<2>(acc, part);
}

// ---------------------------------------------------------------------------
// Primal step of primal_dual_optimised (pfb/opt/primal_dual.py:139-150):
//   xout += grad(xp) with grad = conv - data;  x = xp - tau*xout;
//   positivity 1: x[x<0] = 0;  2: pixel zeroed in all bands if any band <= 0;
//   eps terms: sum (x-xp)^2, sum x^2 (fp64 accumulate, squares in T).
// One thread owns a pixel and walks the bands (needed for positivity 2).
// conv/data may be null (grad term omitted / data omitted).
template <class T>
__global__ void __launch_bounds__(256)
k_pd_primal(T *x, const T *__restrict__ xp, const T *__restrict__ xout,
            const T *__restrict__ wprev, const T *__restrict__ conv, const T *__restrict__ data,
            T tau, int positivity, int nband, long npix, double *part) {
    double acc[2] = {0.0, 0.0};
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < npix;
         i += (long)gridDim.x * blockDim.x) {
        bool kill = false;
        for (int pass = 0; pass < (positivity == 2 ? 2 : 1); ++pass) {
            for (int b = 0; b < nband; ++b) {
                long k = b * npix + i;
                T g = xout[k];
                if (wprev) g = (T)2 * g - wprev[k];
                if (conv) {
                    T gr = conv[k];
                    if (data) gr = gr - data[k];
                    g = g + gr;
                }
                T p = xp[k];
                T xi = p - tau * g;
                if (positivity == 1) {
                    if (xi < (T)0) xi = (T)0;
                } else if (positivity == 2) {
                    if (pass == 0) {
                        if (xi <= (T)0) kill = true;
                        continue;
                    }
                    if (kill) xi = (T)0;
                }
                x[k] = xi;
                T d = xi - p;
                acc[0] += (double)(d * d);
                acc[1] += (double)(xi * xi);
            }
        }
    }
    block_store<2>(acc, part);
}

// Flat, 128-bit vectorised form for positivity 0/1 (no cross-band coupling):
// one thread handles V consecutive elements of the (nband*npix) cube.
template <class T, int V>
__global__ void __launch_bounds__(256)
k_pd_primal_vec(T *x, const T *__restrict__ xp, const T *__restrict__ xout,
                const T *__restrict__ wprev, const T *__restrict__ conv,
                const T *__restrict__ data, T tau, int positivity, long n, double *part) {
    double acc[2] = {0.0, 0.0};
    long nv = n / V;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long)gridDim.x * blockDim.x) {
        long k = i * V;
        Pack<T, V> g = ld<T, V>(xout + k), p = ld<T, V>(xp + k), c, d, o, wq;
        if (wprev) wq = ld<T, V>(wprev + k);
        if (conv) c = ld<T, V>(conv + k);
        if (data) d = ld<T, V>(data + k);
#pragma unroll
        for (int e = 0; e < V; ++e) {
            T ge = g.v[e];
            if (wprev) ge = (T)2 * ge - wq.v[e];
            if (conv) {
                T gr = c.v[e];
                if (data) gr = gr - d.v[e];
                ge = ge + gr;
            }
            T xi = p.v[e] - tau * ge;
            if (positivity == 1 && xi < (T)0) xi = (T)0;
            o.v[e] = xi;
            T df = xi - p.v[e];
            acc[0] += (double)(df * df);
            acc[1] += (double)(xi * xi);
        }
        st<T, V>(x + k, o);
    }
    block_store<2>(acc, part);
}

// ---------------------------------------------------------------------------
// PCG line update (pfb/opt/pcg.py:92-95, 97-101):
//   x = xp + alpha*p;  r = rp + alpha*Ap;  y = M(r) with M(r) = minv*r
// part -> {sum r*y, sum (x-xp)^2, sum x^2}.  xp/rp are kept by the caller so a
// backtracking step can be re-issued with a smaller alpha (no xp/rp copies).
template <class T>
__global__ void __launch_bounds__(256)
k_pcg_update(T *x, T *r, const T *__restrict__ xp, const T *__restrict__ rp,
             const T *__restrict__ p, const T *__restrict__ Ap, T alpha, T minv,
             long n, double *part) {
    double acc[3] = {0.0, 0.0, 0.0};
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long)gridDim.x * blockDim.x) {
        T xo = xp[i];
        T xi = xo + alpha * p[i];
        T ri = rp[i] + alpha * Ap[i];
        x[i] = xi;
        r[i] = ri;
        T yi = ri * minv;
        T d = xi - xo;
        acc[0] += (double)ri * (double)yi;
        acc[1] += (double)(d * d);
        acc[2] += (double)(xi * xi);
    }
    block_store<3>(acc, part);
}

// p = beta*p - minv*r ; part -> {count of non-zero p}  (pfb/opt/pcg.py:103-107)
template <class T>
__global__ void __launch_bounds__(256)
k_pcg_dir(T *p, const T *__restrict__ r, T beta, T minv, long n, double *part) {
    double acc[1] = {0.0};
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long)gridDim.x * blockDim.x) {
        T pi = beta * p[i] - minv * r[i];
        p[i] = pi;
        if (pi != (T)0) acc[0] += 1.0;
    }
    block_store<1>(acc, part);
}

// ---------------------------------------------------------------------------
// host side
static double *g_part = nullptr;  // kRedBlocks * kRedMaxK doubles

static int ensure_part() {
    if (!g_part) {
        if (cudaMalloc((void **)&g_part, sizeof(double) * kRedBlocks * kRedMaxK) != cudaSuccess)
            return fail(PFB_ERR_ALLOC, "cudaMalloc of reduction workspace failed");
    }
    return PFB_OK;
}

static inline int stream_grid(long nitems) {
    long b = (nitems + 255) / 256;
    long cap = (long)kNumSM * 16;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

template <class T> static inline bool aligned16(const void *p) {
    return ((uintptr_t)p & 15) == 0;
}

// dispatch the band count to a compile-time constant (1..8), else generic
#define PFB_NB_SWITCH(nband, CALL)                                            \
    switch (nband) {                                                           \
    case 1: CALL(1); break; case 2: CALL(2); break; case 3: CALL(3); break;    \
    case 4: CALL(4); break; case 5: CALL(5); break; case 6: CALL(6); break;    \
    case 7: CALL(7); break; case 8: CALL(8); break; default: CALL(0); break;   \
    }

template <class T>
static int prox21_t(const T *v, T *res, const T *w, int nband, long n, double lam,
                    double sigma, int mode, cudaStream_t st) {
    const int V = 16 / sizeof(T);
    bool vec = (n % V == 0) && aligned16<T>(v) && aligned16<T>(res) && aligned16<T>(w);
    dim3 b(256);
    if (vec) {
        dim3 g(stream_grid(n / V));
#define PFB_CALL(NB) launch(&k_prox21<T, V, NB>, g, b, 0, st, v, res, w, nband, n, lam, sigma, mode)
        PFB_NB_SWITCH(nband, PFB_CALL)
#undef PFB_CALL
    } else {
        dim3 g(stream_grid(n));
        launch(&k_prox21<T, 1, 0>, g, b, 0, st, v, res, w, nband, n, lam, sigma, mode);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int dual_t(const T *vp, T *v, const T *w, T *u, int nband, long n, double lam,
                  double sigma, int mode, cudaStream_t st) {
    const int V = 16 / sizeof(T);
    bool vec = (n % V == 0) && aligned16<T>(v) && aligned16<T>(vp) && aligned16<T>(w) &&
               aligned16<T>(u);
    dim3 b(256);
    if (vec) {
        dim3 g(stream_grid(n / V));
#define PFB_CALL(NB) launch(&k_dual<T, V, NB>, g, b, 0, st, vp, v, w, u, nband, n, lam, sigma, mode)
        PFB_NB_SWITCH(nband, PFB_CALL)
#undef PFB_CALL
    } else {
        dim3 g(stream_grid(n));
        launch(&k_dual<T, 1, 0>, g, b, 0, st, vp, v, w, u, nband, n, lam, sigma, mode);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int dual_partial_t(const T *vp, const T *v, T *bsum, int nband, long n, long bs, double sigma,
                          int mode, cudaStream_t st) {
    const int V = 16 / sizeof(T);
    bool vec = (n % V == 0) && (bs % V == 0) && aligned16<T>(v) && aligned16<T>(vp) && aligned16<T>(bsum);
    if (vec) {
        dim3 g(stream_grid(n / V)), b(256);
#define PFB_CALL(NB) launch(&k_dual_partial<T, V, NB>, g, b, 0, st, vp, v, bsum, nband, n, bs, sigma, mode)
        PFB_NB_SWITCH(nband, PFB_CALL)
#undef PFB_CALL
    } else {
        launch(&k_dual_partial<T, 1, 0>, dim3(stream_grid(n)), dim3(256), 0, st, vp, v, bsum, nband, n, bs, sigma, mode);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int dual_apply_t(const T *vp, T *v, const T *bsum, const T *w, T *u, int nband,
                        long n, long bs, double lam, double sigma, int mode, cudaStream_t st) {
    const int V = 16 / sizeof(T);
    bool vec = (n % V == 0) && (bs % V == 0) && aligned16<T>(v) && aligned16<T>(vp) && aligned16<T>(bsum) &&
               aligned16<T>(w) && aligned16<T>(u);
    if (vec) {
        dim3 g(stream_grid(n / V)), b(256);
#define PFB_CALL(NB) launch(&k_dual_apply<T, V, NB>, g, b, 0, st, vp, v, bsum, w, u, nband, n, bs, lam, sigma, mode)
        PFB_NB_SWITCH(nband, PFB_CALL)
#undef PFB_CALL
    } else {
        launch(&k_dual_apply<T, 1, 0>, dim3(stream_grid(n)), dim3(256), 0, st, vp, v, bsum, w, u, nband, n, bs, lam, sigma, mode);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int dual_rec_t(const T *vp, T *v, T *vbar, const T *ubar, const T *w, int nband, long n,
                      double lam, double sigma, cudaStream_t st) {
    const int V = 16 / sizeof(T);
    bool vec = (n % V == 0) && aligned16<T>(v) && aligned16<T>(vp) && aligned16<T>(vbar) &&
               aligned16<T>(ubar) && aligned16<T>(w);
    if (vec) {
        dim3 g(stream_grid(n / V)), b(256);
#define PFB_CALL(NB) launch(&k_dual_rec<T, V, NB>, g, b, 0, st, vp, v, vbar, ubar, w, nband, n, lam, sigma)
        PFB_NB_SWITCH(nband, PFB_CALL)
#undef PFB_CALL
    } else {
        launch(&k_dual_rec<T, 1, 0>, dim3(stream_grid(n)), dim3(256), 0, st, vp, v, vbar, ubar, w, nband, n, lam, sigma);
    }
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int dual_factor_t(T *vbar, const T *ubar, const T *w, T *f, long n, double lam, double sigma,
                         cudaStream_t st) {
    const int V = 16 / sizeof(T);
    bool vec = (n % V == 0) && aligned16<T>(vbar) && aligned16<T>(ubar) && aligned16<T>(w) && aligned16<T>(f);
    if (vec) launch(&k_dual_factor<T, V>, dim3(stream_grid(n / V)), dim3(256), 0, st, vbar, ubar, w, f, n, lam, sigma);
    else launch(&k_dual_factor<T, 1>, dim3(stream_grid(n)), dim3(256), 0, st, vbar, ubar, w, f, n, lam, sigma);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

template <class T>
static int axpby_t(T *out, double a, const T *x, double b, const T *y, long n, cudaStream_t st) {
    const int V = 16 / sizeof(T);
    bool vec = (n % V == 0) && aligned16<T>(out) && aligned16<T>(x) && aligned16<T>(y);
    if (vec) launch(&k_axpby<T, V>, dim3(stream_grid(n / V)), dim3(256), 0, st, out, (T)a, x, (T)b, y, n);
    else launch(&k_axpby<T, 1>, dim3(stream_grid(n)), dim3(256), 0, st, out, (T)a, x, (T)b, y, n);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

static int finish_reduce(int K, double *out, cudaStream_t st) {
    launch(&k_red_final, dim3(1), dim3(256), 0, st, (const double *)g_part, kRedBlocks, K, out);
    PFB_CHECK_LAUNCH();
    return PFB_OK;
}

}  // namespace pfb

using namespace pfb;

#define PFB_DISPATCH(dtype, CALL32, CALL64)                                    \
    do {                                                                       \
        if ((dtype) == PFB_F32) return CALL32;                                 \
        if ((dtype) == PFB_F64) return CALL64;                                 \
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", (int)(dtype));        \
    } while (0)

extern "C" {

int pfb_version(void) { return 100; }
const char *pfb_last_error(void) { return g_last_error.c_str(); }
long long pfb_launch_count(void) { return g_launch_count; }

/* profiling: see include/pfb_b200.h */
int pfb_profile_enable(int on) {
#ifndef PFB_EMU
    if (on && !g_profile_on) {
        for (auto &r : g_prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
        g_prof.clear();
    }
#endif
    g_profile_on = on ? 1 : 0;
    return PFB_OK;
}

int pfb_profile_report(char *buf, long long buflen) {
    if (!buf || buflen < 2) return fail(PFB_ERR_BAD_ARG, "pfb_profile_report: bad buffer");
    std::string out;
#ifndef PFB_EMU
    if (cudaDeviceSynchronize() != cudaSuccess)
        return fail(PFB_ERR_CUDA, "pfb_profile_report: device synchronise failed");
    struct Acc { long n; double ms; };
    std::map<std::string, Acc> acc;
    std::vector<std::string> order;
    for (auto &r : g_prof) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.e0, r.e1);
        const char *nm = nullptr;
        std::string key = "?";
        if (cudaFuncGetName(&nm, r.func) == cudaSuccess && nm) key = nm;
        auto it = acc.find(key);
        if (it == acc.end()) { acc[key] = Acc{1, (double)ms}; order.push_back(key); }
        else { it->second.n += 1; it->second.ms += ms; }
    }
    char line[1024];
    for (auto &k : order) {
        snprintf(line, sizeof line, "%s\t%ld\t%.6f\n", k.c_str(), acc[k].n, acc[k].ms);
        out += line;
    }
#endif
    if ((long long)out.size() + 1 > buflen) return fail(PFB_ERR_BAD_ARG, "pfb_profile_report: buffer too small (%zu needed)", out.size() + 1);
    memcpy(buf, out.c_str(), out.size() + 1);
    return PFB_OK;
}

int pfb_prox21(const void *v, void *result, const void *weight, int nband, long long n,
               double lam, double sigma, int mode, int dtype, void *stream) {
    if (!v || !result || !weight || nband < 1 || n < 0)
        return fail(PFB_ERR_BAD_ARG, "pfb_prox21: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    PFB_DISPATCH(dtype,
                 prox21_t<float>((const float *)v, (float *)result, (const float *)weight, nband, n, lam, sigma, mode, st),
                 prox21_t<double>((const double *)v, (double *)result, (const double *)weight, nband, n, lam, sigma, mode, st));
}

int pfb_dual_update(const void *vp, void *v, const void *weight, void *u, int nband,
                    long long n, double lam, double sigma, int mode, int dtype, void *stream) {
    if (!vp || !v || !weight || nband < 1 || n < 0)
        return fail(PFB_ERR_BAD_ARG, "pfb_dual_update: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    PFB_DISPATCH(dtype,
                 dual_t<float>((const float *)vp, (float *)v, (const float *)weight, (float *)u, nband, n, lam, sigma, mode, st),
                 dual_t<double>((const double *)vp, (double *)v, (const double *)weight, (double *)u, nband, n, lam, sigma, mode, st));
}

int pfb_dual_partial(const void *vp, const void *v, void *bsum, int nband, long long n,
                     long long band_stride, double sigma, int mode, int dtype, void *stream) {
    if (band_stride <= 0) band_stride = n;
    if (!vp || !v || !bsum || nband < 1 || n < 0 || band_stride < n)
        return fail(PFB_ERR_BAD_ARG, "pfb_dual_partial: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    PFB_DISPATCH(dtype,
                 dual_partial_t<float>((const float *)vp, (const float *)v, (float *)bsum, nband, n, band_stride, sigma, mode, st),
                 dual_partial_t<double>((const double *)vp, (const double *)v, (double *)bsum, nband, n, band_stride, sigma, mode, st));
}

int pfb_dual_apply(const void *vp, void *v, const void *bsum, const void *weight, void *u,
                   int nband, long long n, long long band_stride, double lam, double sigma,
                   int mode, int dtype, void *stream) {
    if (band_stride <= 0) band_stride = n;
    if (!vp || !v || !bsum || !weight || nband < 1 || n < 0 || band_stride < n)
        return fail(PFB_ERR_BAD_ARG, "pfb_dual_apply: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    PFB_DISPATCH(dtype,
                 dual_apply_t<float>((const float *)vp, (float *)v, (const float *)bsum, (const float *)weight, (float *)u, nband, n, band_stride, lam, sigma, mode, st),
                 dual_apply_t<double>((const double *)vp, (double *)v, (const double *)bsum, (const double *)weight, (double *)u, nband, n, band_stride, lam, sigma, mode, st));
}

int pfb_dual_update_rec(const void *vp, void *v, void *vbar, const void *ubar, const void *weight,
                        int nband, long long n, double lam, double sigma, int dtype, void *stream) {
    if (!vp || !v || !vbar || !ubar || !weight || nband < 1 || n < 0)
        return fail(PFB_ERR_BAD_ARG, "pfb_dual_update_rec: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    PFB_DISPATCH(dtype,
                 dual_rec_t<float>((const float *)vp, (float *)v, (float *)vbar, (const float *)ubar, (const float *)weight, nband, n, lam, sigma, st),
                 dual_rec_t<double>((const double *)vp, (double *)v, (double *)vbar, (const double *)ubar, (const double *)weight, nband, n, lam, sigma, st));
}

int pfb_dual_factor(void *vbar, const void *ubar, const void *weight, void *f, long long n, double lam,
                    double sigma, int dtype, void *stream) {
    if (!vbar || !ubar || !weight || !f || n < 0) return fail(PFB_ERR_BAD_ARG, "pfb_dual_factor: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    PFB_DISPATCH(dtype,
                 dual_factor_t<float>((float *)vbar, (const float *)ubar, (const float *)weight, (float *)f, n, lam, sigma, st),
                 dual_factor_t<double>((double *)vbar, (const double *)ubar, (const double *)weight, (double *)f, n, lam, sigma, st));
}

int pfb_axpby(void *out, double a, const void *x, double b, const void *y, long long n,
              int dtype, void *stream) {
    if (!out || !x || !y || n < 0) return fail(PFB_ERR_BAD_ARG, "pfb_axpby: bad argument");
    if (n == 0) return PFB_OK;
    cudaStream_t st = as_stream(stream);
    PFB_DISPATCH(dtype, axpby_t<float>((float *)out, a, (const float *)x, b, (const float *)y, n, st),
                 axpby_t<double>((double *)out, a, (const double *)x, b, (const double *)y, n, st));
}

int pfb_vdot(const void *a, const void *b, const void *c, const void *d, long long n,
             int dtype, double *out2, void *stream) {
    if (!a || !b || !out2 || n < 0 || ((c == nullptr) != (d == nullptr)))
        return fail(PFB_ERR_BAD_ARG, "pfb_vdot: bad argument");
    int rc = ensure_part();
    if (rc) return rc;
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32)
        launch(&k_dot<float>, dim3(kRedBlocks), dim3(256), 0, st, (const float *)a, (const float *)b, (const float *)c, (const float *)d, (long)n, g_part);
    else if (dtype == PFB_F64)
        launch(&k_dot<double>, dim3(kRedBlocks), dim3(256), 0, st, (const double *)a, (const double *)b, (const double *)c, (const double *)d, (long)n, g_part);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return finish_reduce(2, out2, st);
}

int pfb_norm_diff(const void *x, const void *xp, long long n, int dtype, double *out2,
                  void *stream) {
    if (!x || !xp || !out2 || n < 0) return fail(PFB_ERR_BAD_ARG, "pfb_norm_diff: bad argument");
    int rc = ensure_part();
    if (rc) return rc;
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32)
        launch(&k_norm_diff<float>, dim3(kRedBlocks), dim3(256), 0, st, (const float *)x, (const float *)xp, (long)n, g_part);
    else if (dtype == PFB_F64)
        launch(&k_norm_diff<double>, dim3(kRedBlocks), dim3(256), 0, st, (const double *)x, (const double *)xp, (long)n, g_part);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return finish_reduce(2, out2, st);
}

static int pd_primal_impl(void *x, const void *xp, const void *xout, const void *wprev,
                          const void *conv, const void *data, double tau, int positivity,
                          int nband, long long npix, int dtype, double *out2, void *stream) {
    if (!x || !xp || !xout || !out2 || nband < 1 || npix < 0 || (data && !conv))
        return fail(PFB_ERR_BAD_ARG, "pfb_pd_primal: bad argument");
    int rc = ensure_part();
    if (rc) return rc;
    cudaStream_t st = as_stream(stream);
    long ntot = (long)nband * (long)npix;
    bool al = (((uintptr_t)x | (uintptr_t)xp | (uintptr_t)xout | (uintptr_t)wprev | (uintptr_t)conv |
                (uintptr_t)data) & 15) == 0;
    if (positivity != 2 && al && dtype == PFB_F32 && ntot % 4 == 0)
        launch(&k_pd_primal_vec<float, 4>, dim3(kRedBlocks), dim3(256), 0, st, (float *)x, (const float *)xp, (const float *)xout, (const float *)wprev, (const float *)conv, (const float *)data, (float)tau, positivity, ntot, g_part);
    else if (positivity != 2 && al && dtype == PFB_F64 && ntot % 2 == 0)
        launch(&k_pd_primal_vec<double, 2>, dim3(kRedBlocks), dim3(256), 0, st, (double *)x, (const double *)xp, (const double *)xout, (const double *)wprev, (const double *)conv, (const double *)data, tau, positivity, ntot, g_part);
    else if (dtype == PFB_F32)
        launch(&k_pd_primal<float>, dim3(kRedBlocks), dim3(256), 0, st, (float *)x, (const float *)xp, (const float *)xout, (const float *)wprev, (const float *)conv, (const float *)data, (float)tau, positivity, nband, (long)npix, g_part);
    else if (dtype == PFB_F64)
        launch(&k_pd_primal<double>, dim3(kRedBlocks), dim3(256), 0, st, (double *)x, (const double *)xp, (const double *)xout, (const double *)wprev, (const double *)conv, (const double *)data, tau, positivity, nband, (long)npix, g_part);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return finish_reduce(2, out2, st);
}

int pfb_pd_primal(void *x, const void *xp, const void *xout, const void *conv,
                  const void *data, double tau, int positivity, int nband, long long npix,
                  int dtype, double *out2, void *stream) {
    return pd_primal_impl(x, xp, xout, nullptr, conv, data, tau, positivity, nband, npix, dtype, out2, stream);
}

int pfb_pd_primal_lin(void *x, const void *xp, const void *w, const void *wprev, const void *conv,
                      const void *data, double tau, int positivity, int nband, long long npix,
                      int dtype, double *out2, void *stream) {
    if (!wprev) return fail(PFB_ERR_BAD_ARG, "pfb_pd_primal_lin: wprev is null");
    return pd_primal_impl(x, xp, w, wprev, conv, data, tau, positivity, nband, npix, dtype, out2, stream);
}

int pfb_pcg_update(void *x, void *r, const void *xp, const void *rp, const void *p,
                   const void *Ap, double alpha, double minv, long long n, int dtype,
                   double *out3, void *stream) {
    if (!x || !r || !xp || !rp || !p || !Ap || !out3 || n < 0)
        return fail(PFB_ERR_BAD_ARG, "pfb_pcg_update: bad argument");
    int rc = ensure_part();
    if (rc) return rc;
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32)
        launch(&k_pcg_update<float>, dim3(kRedBlocks), dim3(256), 0, st, (float *)x, (float *)r, (const float *)xp, (const float *)rp, (const float *)p, (const float *)Ap, (float)alpha, (float)minv, (long)n, g_part);
    else if (dtype == PFB_F64)
        launch(&k_pcg_update<double>, dim3(kRedBlocks), dim3(256), 0, st, (double *)x, (double *)r, (const double *)xp, (const double *)rp, (const double *)p, (const double *)Ap, alpha, minv, (long)n, g_part);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return finish_reduce(3, out3, st);
}

int pfb_pcg_direction(void *p, const void *r, double beta, double minv, long long n,
                      int dtype, double *out1, void *stream) {
    if (!p || !r || !out1 || n < 0) return fail(PFB_ERR_BAD_ARG, "pfb_pcg_direction: bad argument");
    int rc = ensure_part();
    if (rc) return rc;
    cudaStream_t st = as_stream(stream);
    if (dtype == PFB_F32)
        launch(&k_pcg_dir<float>, dim3(kRedBlocks), dim3(256), 0, st, (float *)p, (const float *)r, (float)beta, (float)minv, (long)n, g_part);
    else if (dtype == PFB_F64)
        launch(&k_pcg_dir<double>, dim3(kRedBlocks), dim3(256), 0, st, (double *)p, (const double *)r, beta, minv, (long)n, g_part);
    else
        return fail(PFB_ERR_BAD_ARG, "unknown dtype %d", dtype);
    PFB_CHECK_LAUNCH();
    return finish_reduce(1, out1, st);
}

}  // extern "C"
