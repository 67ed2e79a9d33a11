// Streaming, register-resident analysis (DWT) level: the "warp-shuffle stencil"
// form of dwt2d_level (pfb/wavelets/wavelets.py:126-171).
//
// One WARP owns a strip of 29 coefficient columns jy (y = contiguous image
// axis) and walks down the image rows (x axis) of a chunk of NJ coefficient
// rows jx:
//   * each lane loads one 2-element vector per image row (256 B coalesced per
//     warp and row, straight from global memory -- no shared-memory staging of
//     the input);
//   * the row pass (filter along y) takes its 6 halo values from the three lower
//     lanes with warp shuffles (lanes 0..2 only carry halo => 29 outputs/strip);
//   * the column pass (filter along x) keeps the last F-2 row-pass results in
//     registers (shifted by two rows per step); the inner loops are pure FFMA
//     with the taps as constant-bank operands and the loop body is small enough
//     for the L0 instruction cache;
//   * the reference layout is transposed w.r.t. the image (slab rows = y), so
//     the three detail sub-bands (and LL on the last level) are staged through a
//     per-warp shared-memory block of 32 jx and written as 128-byte runs;
//     LL of a non-final level goes straight from registers to the next level's
//     natural-orientation scratch (coalesced).
// The tile kernel k_dwt_level of wavelet.cu stays as the generic path (image
// rows not a multiple of 4 elements, or unaligned pointers).
#pragma once

namespace pfb {

// ---------------------------------------------------------------------------
// Geometry.  A lane owns ONE coefficient column: it loads the image pair
// in[2t], in[2t+1] of every row (256 B per warp and row) and takes the 6 halo
// values of the 8-tap window from lanes l-1..l-3 => lanes 0..2 carry halo only
// and a warp strip yields 29 columns.  The transposed write-out block is 32 jx
// wide, i.e. every slab store instruction writes one 128-byte run (measured on
// B200: 8-wide runs cost 2.7 ms per Psi.dot at 8x4096^2, 16-wide 1.4 ms,
// 32-wide 0.9 ms -- profiles/dbg_dwt.sh), and with one column per lane its
// shared-memory staging (3 x 32 x 29 values) is small enough for 16 warps/SM.
static const int kStripOut = 29;     // coefficient columns per warp strip (slab output)
// Natural-orientation output stores one row segment per warp and sub-band straight
// from registers: 24 columns = 96 bytes (fp32) keep every segment aligned to the
// 32-byte DRAM sector.  Measured at 8x4096^2 (profiles/dbg_dwt.sh): 29-wide
// (unaligned) segments 1.68 ms per Psi.dot, 28-wide 1.36 ms, 24-wide 1.16 ms, no
// stores at all 0.73 ms -- partial-sector writes, not bytes, were the cost.
static const int kStripNat = 24;
static const int kJB = 32;           // jx per transposed write-out block
static const int kPitch = 31;        // staging pitch (odd: conflict-free transposed reads)
static const int kStreamWarps = 4;   // warps (independent tasks) per CTA
__host__ __device__ constexpr int stream_smem_warp(int nq) { return nq * kJB * kPitch; }

static const int kPD = 8;            // prefetch distance (coefficient rows in flight per warp)
// ring of kPD x 2 image-row pairs per warp, filled with cp.async (LDGSTS): no
// registers are tied up by the loads in flight and rows/columns outside the
// image are zero-filled by the copy itself (src-size operand).
__host__ __device__ constexpr int stream_ring_elems() { return kPD * 2 * 32 * 2; }

template <class T> struct Vec2 { T v[2]; };

// Fused dual update (pfb_psi_dot_nat_dual): store v = (vp + sigma u) * f instead of the
// coefficient u.  dv, df: element offsets from the output location to vp and to f.
template <class T> struct DualFz {
    bool on;
    long dv, df;
    T sigma;
    // vp and f are read-only here and never alias the output: loads through the read-only path
    __device__ __forceinline__ T ldvp(const T *dst) const { return __ldg(dst + dv); }
    __device__ __forceinline__ T ldf(const T *dst) const { return __ldg(dst + df); }
    // a, f: vp and f at dst, fetched one coefficient row ahead (see the kernels)
    __device__ __forceinline__ void put(T *dst, const T u, const T a, const T f) const {
        *dst = on ? (a + sigma * u) * f : u;
    }
};
__device__ __forceinline__ Vec2<float> ldv2(const float *p) {
    float2 q = *reinterpret_cast<const float2 *>(p);
    Vec2<float> r; r.v[0] = q.x; r.v[1] = q.y; return r;
}
__device__ __forceinline__ Vec2<double> ldv2(const double *p) {
    double2 q = *reinterpret_cast<const double2 *>(p);
    Vec2<double> r; r.v[0] = q.x; r.v[1] = q.y; return r;
}

__device__ __forceinline__ void stv2(float *p, const Vec2<float> &r) {
    *reinterpret_cast<float2 *>(p) = make_float2(r.v[0], r.v[1]);
}
__device__ __forceinline__ void stv2(double *p, const Vec2<double> &r) {
    *reinterpret_cast<double2 *>(p) = make_double2(r.v[0], r.v[1]);
}

// row pass of one image row for the lane's column: m = {lo, hi}
template <class T, int F>
__device__ __forceinline__ void dwt_row_pass(const Vec2<T> &q, T (&m)[2]) {
    // window w[j] = in[2t - 6 + j], j = 0..7
    T w[8];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        w[2 * j] = __shfl_up_sync(0xffffffffu, q.v[0], 3 - j);
        w[2 * j + 1] = __shfl_up_sync(0xffffffffu, q.v[1], 3 - j);
    }
    w[6] = q.v[0];
    w[7] = q.v[1];
    m[0] = m[1] = (T)0;
#pragma unroll
    for (int k = 0; k < F; ++k) {
        m[0] += tapc<T>(F / 2 - 1, 0, k) * w[7 - k];
        m[1] += tapc<T>(F / 2 - 1, 1, k) * w[7 - k];
    }
}

// The loop body handles ONE coefficient row jx (two image rows) and is kept
// small on purpose: the warps of an SM run this loop independently of each
// other, so the body has to fit the ~6 KB L0 instruction cache (an 8-row
// unrolled version with a static register ring was bound by instruction fetch:
// icc hit rate 65 %, `no_instruction` the top stall in profiles/r01d).  The F-2
// previous row-pass results are shifted in registers.
// NAT: the sub-bands go to the natural-orientation coefficient layout of the
// streaming synthesis (rows = jx, 116-byte coalesced runs straight from
// registers; `slab` then is the (band, basis) block of that layout) instead of
// the reference's transposed slab.
template <class T, int F, bool NAT>
__device__ __forceinline__ void dwt_stream_task(const T *__restrict__ in, const int ld,
                                                const LevelDesc &d, T *slab, const int ldslab,
                                                T *ll_out, const bool last, const int strip,
                                                const int chunk, const int NJ, T *s_out,
                                                T *ring, const int dbg, const int SW,
                                                const DualFz<T> fz = DualFz<T>{false, 0, 0, (T)0}) {
    const int lane = threadIdx.x & 31;
    const int T0 = SW * strip;
    const int c = 2 * T0 - 6 + 2 * lane;             // first column of this lane's pair
    const bool cvalid = c >= 0 && c < ld;            // ld, c even
    const int tl = lane - 3;                         // local output column (lanes 0..2: halo only)
    const int ta = T0 + tl;
    const bool tval = lane >= 3 && lane < 3 + SW && ta < d.C1;
    const int JX0 = chunk * NJ;
    const int jx_end = JX0 + NJ < d.C0 ? JX0 + NJ : d.C0;
    const int s0 = last ? 0 : 1;                     // first staged sub-band
    constexpr int NH = F > 2 ? F - 2 : 1;
    T hs[NH][2];                                     // row-pass results of rows 2jx-1, 2jx-2, ...
#pragma unroll
    for (int i = 0; i < NH; ++i) hs[i][0] = hs[i][1] = (T)0;
    int tmax0 = d.C1 - T0;
    tmax0 = tmax0 < SW ? tmax0 : SW;
    // bytes of this lane's pair that lie inside the image row (0 => zero fill)
    const int cbytes = (!cvalid || (dbg & 4)) ? 0 : (c + 1 < d.n1 ? 2 : (c < d.n1 ? 1 : 0)) * (int)sizeof(T);
    T *myring = ring + 2 * lane;                     // slot layout [iter % kPD][row 0/1][lane][2]
    auto prefetch = [&](const int j) {               // rows 2j, 2j+1 -> ring slot j mod kPD
        T *dst = myring + (j & (kPD - 1)) * 128;
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            const int r = 2 * j + rr;
            const bool ok = (unsigned)r < (unsigned)d.n0 && j < jx_end;
            cp_async_zfill<2 * sizeof(T)>(dst + rr * 64, ok ? in + (long)r * ld + c : in, ok ? cbytes : 0);
        }
        cp_async_commit();
    };
    const int jx_first = JX0 - (F - 2) / 2;
#pragma unroll 1
    for (int j = jx_first; j < jx_first + kPD; ++j) prefetch(j);
    // fused dual update: vp / f of the NEXT coefficient row are fetched into registers while the
    // current row is computed, so their latency is never exposed
    T za[4] = {(T)0, (T)0, (T)0, (T)0}, zf[4] = {(T)0, (T)0, (T)0, (T)0};
    auto zfetch = [&](const int j) {
        if (NAT && fz.on && tval && j >= JX0 && j < jx_end) {
            const T *dst = slab + d.off1 + (long)j * d.ldn + ta;
            if (last) { za[0] = fz.ldvp(dst - (long)d.qs); zf[0] = fz.ldf(dst - (long)d.qs); }
            za[1] = fz.ldvp(dst); zf[1] = fz.ldf(dst);
            za[2] = fz.ldvp(dst + d.qs); zf[2] = fz.ldf(dst + d.qs);
            za[3] = fz.ldvp(dst + 2 * (long)d.qs); zf[3] = fz.ldf(dst + 2 * (long)d.qs);
        }
    };
    zfetch(jx_first);

#pragma unroll 1
    for (int jx = jx_first; jx < jx_end; ++jx) {
        cp_async_wait<kPD - 1>();                    // the oldest group (this jx) has landed
        const T *slot = myring + (jx & (kPD - 1)) * 128;
        Vec2<T> q0 = ldv2(slot), q1 = ldv2(slot + 64);
        prefetch(jx + kPD);                          // refill the slot just consumed
        T m0[2], m1[2];
        dwt_row_pass<T, F>(q0, m0);
        dwt_row_pass<T, F>(q1, m1);
        // column pass: out[jx] = sum_k f[k] mid[2jx+1-k]; k=0 -> m1, k=1 -> m0, k>=2 -> hs[k-2]
        T o[4] = {(T)0, (T)0, (T)0, (T)0};
#pragma unroll
        for (int k = 0; k < F; ++k) {
            const T lo = tapc<T>(F / 2 - 1, 0, k), hi = tapc<T>(F / 2 - 1, 1, k);
            const T *mm = k == 0 ? m1 : (k == 1 ? m0 : hs[k >= 2 ? k - 2 : 0]);
            o[0] += lo * mm[0];       // x-lo of y-lo -> TL (LL)
            o[1] += hi * mm[0];       // x-hi of y-lo -> TR
            o[2] += lo * mm[1];       // x-lo of y-hi -> BL
            o[3] += hi * mm[1];       // x-hi of y-hi -> BR
        }
        if (F > 2) {                                 // shift the history by two rows
#pragma unroll
            for (int i = NH - 1; i >= 2; --i) { hs[i][0] = hs[i - 2][0]; hs[i][1] = hs[i - 2][1]; }
            hs[1][0] = m0[0]; hs[1][1] = m0[1];
            hs[0][0] = m1[0]; hs[0][1] = m1[1];
        }
        if (NAT) {
            if (jx >= JX0 && tval) {
                T *dst = slab + d.off1 + (long)jx * d.ldn + ta;
                if (last) fz.put(dst - (long)d.qs, o[0], za[0], zf[0]);
                else ll_out[(long)jx * d.ld_out + ta] = o[0];
                fz.put(dst, o[1], za[1], zf[1]);
                fz.put(dst + d.qs, o[2], za[2], zf[2]);
                fz.put(dst + 2 * (long)d.qs, o[3], za[3], zf[3]);
            }
            zfetch(jx + 1);
            continue;
        }
        // staging (warm-up rows hit cells that are overwritten before the flush)
        const int jxl = (jx - JX0) & (kJB - 1);
        if (lane >= 3) {
#pragma unroll
            for (int s = 0; s < 4; ++s)
                if (s > 0 || last) s_out[((s - s0) * kJB + jxl) * kPitch + tl] = o[s];
        }
        if (!last && jx >= JX0 && tval && !(dbg & 2)) ll_out[(long)jx * d.ld_out + ta] = o[0];
        // flush a completed block of 32 jx: one slab row (128 B) per store instruction
        if (jx >= JX0 && (jxl == kJB - 1 || jx == jx_end - 1)) {
            __syncwarp();
            const int fj = jx - jxl + lane;
            const int tmax = (fj < jx_end && !(dbg & 1)) ? tmax0 : 0;
            for (int s = s0; s < 4; ++s) {
                const int qx = s & 1, qy = s >> 1;
                T *dst = slab + (long)(d.lowy + qy * d.C1 + T0) * ldslab + d.lowx + qx * d.C0 + fj;
                const T *src = s_out + ((s - s0) * kJB + lane) * kPitch;
#pragma unroll 4
                for (int t = 0; t < tmax; ++t) dst[(long)t * ldslab] = src[t];
            }
            __syncwarp();
        }
    }
    cp_async_wait<0>();
}

// grid.x = nbasis * nband * nchunk * ceil(nstrip / kStreamWarps)
template <class T, bool NAT>
__global__ void __launch_bounds__(32 * kStreamWarps)
k_dwt_stream(const T *__restrict__ img, const T *__restrict__ scr_in, T *scr_out, T *alpha,
             LevelArgs a, int nstrip, int nchunk, int NJ, long ntask, int dbg, int SW) {
    PFB_DYN_SMEM(raw);
    const int warp = threadIdx.x >> 5;
    T *ring = reinterpret_cast<T *>(raw) + warp * stream_ring_elems();       // 16-byte aligned
    T *s_out = reinterpret_cast<T *>(raw) + kStreamWarps * stream_ring_elems() +
               warp * stream_smem_warp(a.last ? 4 : 3);
    // CTA order: basis fastest, so the four filter banks of one image region run
    // at the same time on neighbouring SMs and the image is read from DRAM once
    // (then from L2); the warps of a CTA take ADJACENT strips of the same rows so
    // that a DRAM page / L2 line of an image row is consumed while it is open.
    const int basis = blockIdx.x % a.nbasis;
    long cta = blockIdx.x / a.nbasis;
    const int nsg = (nstrip + kStreamWarps - 1) / kStreamWarps;      // strip groups
    const int sg = (int)(cta % nsg); cta /= nsg;
    const int chunk = (int)(cta % nchunk);
    const int band = (int)(cta / nchunk);
    const int strip = sg * kStreamWarps + warp;
    if (band >= a.nband) return;
    const LevelDesc &d = a.d[basis];
    if (d.F == 0 || SW * strip >= d.C1 || chunk * NJ >= d.C0) return;
    const long slabsz = (long)a.Nymax * a.Nxmax;
    T *slab = alpha + ((long)band * a.nbasis + basis) * (NAT ? a.cn_stride : slabsz);
    const long so = ((long)band * a.nbasis + basis) * a.scr_stride;
    const T *in = a.first ? img + band * a.img_band_stride : scr_in + so;
    const int ld = a.first ? a.ny : d.ld_in;
    T *ll = scr_out + so;
    DualFz<T> fz;
    fz.on = NAT && a.fz_vp != nullptr;
    fz.dv = a.fz_dvp;
    fz.df = a.fz_df - (long)band * a.nbasis * a.cn_stride;      // f has no band axis
    fz.sigma = (T)a.fz_sigma;
    switch (d.F) {
    case 2: dwt_stream_task<T, 2, NAT>(in, ld, d, slab, a.Nxmax, ll, a.last, strip, chunk, NJ, s_out, ring, dbg, SW, fz); break;
    case 4: dwt_stream_task<T, 4, NAT>(in, ld, d, slab, a.Nxmax, ll, a.last, strip, chunk, NJ, s_out, ring, dbg, SW, fz); break;
    case 6: dwt_stream_task<T, 6, NAT>(in, ld, d, slab, a.Nxmax, ll, a.last, strip, chunk, NJ, s_out, ring, dbg, SW, fz); break;
    case 8: dwt_stream_task<T, 8, NAT>(in, ld, d, slab, a.Nxmax, ll, a.last, strip, chunk, NJ, s_out, ring, dbg, SW, fz); break;
    default: break;   // F = 10 (db5) is routed to the tile kernel by the host
    }
}

// ---------------------------------------------------------------------------
// Finest level, natural-orientation output, ALL filter banks in one walk: the
// four Daubechies banks (slot s <-> db(s+1)) read the same image rows, so one
// warp loads each row pair once, shares the 12 halo shuffles, and runs the four
// row/column passes back to back (sum F = 20 taps); 'self' (the image itself in
// that layout) is stored from the same registers.  Per coefficient row this is
// ~270 instructions instead of 4 x ~130 for four single-bank walks, which matters
// because these walks are instruction-issue bound, not HBM bound.
struct DwtAllSlot {        // indexed with compile-time slot numbers only: every field
    int basis;             // is then a constant-bank operand (a run-time index would
    int C0, C1;            // turn each use into an LDC with a long-scoreboard wait)
    int ldn, qs, ld_out;
    long off1;
};
struct DwtAllArgs {
    DwtAllSlot s[4];       // db1..db4, basis = -1: absent
    int self_basis;        // -1 = absent
    int ld_self;
};

template <class T, int F>
__device__ __forceinline__ void dwt_all_bank(const T (&w0)[8], const T (&w1)[8],
                                             T (&hs)[F > 2 ? F - 2 : 1][2], T (&o)[4]) {
    constexpr int N = F / 2 - 1;
    T m0[2] = {(T)0, (T)0}, m1[2] = {(T)0, (T)0};
#pragma unroll
    for (int k = 0; k < F; ++k) {
        const T lo = tapc<T>(N, 0, k), hi = tapc<T>(N, 1, k);
        m0[0] = fma(lo, w0[7 - k], m0[0]);
        m0[1] = fma(hi, w0[7 - k], m0[1]);
        m1[0] = fma(lo, w1[7 - k], m1[0]);
        m1[1] = fma(hi, w1[7 - k], m1[1]);
    }
    o[0] = o[1] = o[2] = o[3] = (T)0;
#pragma unroll
    for (int k = 0; k < F; ++k) {
        const T lo = tapc<T>(N, 0, k), hi = tapc<T>(N, 1, k);
        const T *mm = k == 0 ? m1 : (k == 1 ? m0 : hs[k >= 2 ? k - 2 : 0]);
        o[0] = fma(lo, mm[0], o[0]);
        o[1] = fma(hi, mm[0], o[1]);
        o[2] = fma(lo, mm[1], o[2]);
        o[3] = fma(hi, mm[1], o[3]);
    }
    if (F > 2) {
#pragma unroll
        for (int i = F - 3; i >= 2; --i) { hs[i][0] = hs[i - 2][0]; hs[i][1] = hs[i - 2][1]; }
        hs[1][0] = m0[0]; hs[1][1] = m0[1];
        hs[0][0] = m1[0]; hs[0][1] = m1[1];
    }
}

// grid.x = nband * nchunk * ceil(nstrip / kStreamWarps); smem = ring only
template <class T>
__global__ void __launch_bounds__(32 * kStreamWarps)
k_dwt_stream_all(const T *__restrict__ img, T *scr_out, T *cn, LevelArgs a, DwtAllArgs g,
                 int nstrip, int nchunk, int NJ, int SW, int dbg) {
    PFB_DYN_SMEM(raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T *ring = reinterpret_cast<T *>(raw) + warp * stream_ring_elems();
    long cta = blockIdx.x;
    const int nsg = (nstrip + kStreamWarps - 1) / kStreamWarps;
    const int sg = (int)(cta % nsg); cta /= nsg;
    const int chunk = (int)(cta % nchunk);
    const int band = (int)(cta / nchunk);
    const int strip = sg * kStreamWarps + warp;
    if (band >= a.nband) return;
    const int T0 = SW * strip, JX0 = chunk * NJ;
    const int n0 = a.nx, n1 = a.ny, ld = a.ny;
    // extents over all banks (db4 has the largest coefficient arrays)
    int C0max = 0, C1max = 0;
#pragma unroll
    for (int s = 0; s < 4; ++s)
        if (g.s[s].basis >= 0) {
            C0max = g.s[s].C0 > C0max ? g.s[s].C0 : C0max;
            C1max = g.s[s].C1 > C1max ? g.s[s].C1 : C1max;
        }
    if (T0 >= C1max || JX0 >= C0max) return;
    const int jx_end = JX0 + NJ < C0max ? JX0 + NJ : C0max;
    const T *in = img + band * a.img_band_stride;
    const int c = 2 * T0 - 6 + 2 * lane;             // first column of this lane's pair
    const bool cvalid = c >= 0 && c < ld;
    const int ta = T0 + lane - 3;                    // coefficient column (lanes 0..2: halo only)
    const int cbytes = !cvalid ? 0 : (c + 1 < n1 ? 2 : (c < n1 ? 1 : 0)) * (int)sizeof(T);

    // per-bank output row pointers (quadrant 1 of row jx, column ta) and LL scratch rows
    T *pq[4], *pll[4];
    bool tv[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const DwtAllSlot &d = g.s[s];
        const long bb = (long)band * a.nbasis + (d.basis >= 0 ? d.basis : 0);
        pq[s] = cn + bb * a.cn_stride + d.off1 + ta;
        pll[s] = a.last ? pq[s] - d.qs : scr_out + bb * a.scr_stride + ta;
        tv[s] = d.basis >= 0 && lane >= 3 && lane < 3 + SW && ta < d.C1 && !(dbg & 1);
    }
    DualFz<T> fz;
    fz.on = a.fz_vp != nullptr;
    fz.dv = a.fz_dvp;
    fz.df = a.fz_df - (long)band * a.nbasis * a.cn_stride;      // f has no band axis
    fz.sigma = (T)a.fz_sigma;
    T *pself = cn;
    const bool sv = g.self_basis >= 0 && lane >= 3 && lane < 3 + SW && c < n1 && !(dbg & 2);
    if (g.self_basis >= 0) pself = cn + ((long)band * a.nbasis + g.self_basis) * a.cn_stride + c;

    T *myring = ring + 2 * lane;                     // slot layout [iter % kPD][row 0/1][lane][2]
    auto prefetch = [&](const int j) {               // rows 2j, 2j+1 -> ring slot j mod kPD
        T *dst = myring + (j & (kPD - 1)) * 128;
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            const int r = 2 * j + rr;
            const bool ok = (unsigned)r < (unsigned)n0 && j < jx_end;
            cp_async_zfill<2 * sizeof(T)>(dst + rr * 64, ok ? in + (long)r * ld + c : in, ok ? cbytes : 0);
        }
        cp_async_commit();
    };
    T h1[2][2], h2[4][2], h3[6][2], h0[1][2];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        h3[i][0] = h3[i][1] = (T)0;
        if (i < 4) h2[i][0] = h2[i][1] = (T)0;
        if (i < 2) h1[i][0] = h1[i][1] = (T)0;
    }
    h0[0][0] = h0[0][1] = (T)0;
    const int jx_first = JX0 - 3;                    // db4 needs three coefficient rows of history
#pragma unroll 1
    for (int j = jx_first; j < jx_first + kPD; ++j) prefetch(j);
    // fused dual update: vp / f of the NEXT coefficient row in registers (see dwt_stream_task)
    T za[4][4], zf[4][4];
    Vec2<T> zsa[2], zsf[2];
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int i = 0; i < 4; ++i) za[s][i] = zf[s][i] = (T)0;
    zsa[0].v[0] = zsa[0].v[1] = zsa[1].v[0] = zsa[1].v[1] = (T)0;
    zsf[0] = zsa[0]; zsf[1] = zsa[0];
    auto zfetch = [&](const int j) {
        if (!fz.on || j < JX0 || j >= jx_end) return;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const DwtAllSlot &d = g.s[s];
            if (tv[s] && j < d.C0) {
                const T *dst = pq[s] + (long)j * d.ldn;
                if (a.last) { za[s][0] = fz.ldvp(dst - (long)d.qs); zf[s][0] = fz.ldf(dst - (long)d.qs); }
                za[s][1] = fz.ldvp(dst); zf[s][1] = fz.ldf(dst);
                za[s][2] = fz.ldvp(dst + d.qs); zf[s][2] = fz.ldf(dst + d.qs);
                za[s][3] = fz.ldvp(dst + 2 * (long)d.qs); zf[s][3] = fz.ldf(dst + 2 * (long)d.qs);
            }
        }
        if (sv) {
#pragma unroll
            for (int rr = 0; rr < 2; ++rr)
                if (2 * j + rr < n0) {
                    const T *dst = pself + (long)(2 * j + rr) * g.ld_self;
                    zsa[rr] = ldv2(dst + fz.dv);
                    zsf[rr] = ldv2(dst + fz.df);
                }
        }
    };
    zfetch(jx_first);

#pragma unroll 1
    for (int jx = jx_first; jx < jx_end; ++jx) {
        cp_async_wait<kPD - 1>();
        const T *slot = myring + (jx & (kPD - 1)) * 128;
        const Vec2<T> q0 = ldv2(slot), q1 = ldv2(slot + 64);
        prefetch(jx + kPD);
        // the 8-wide window of both rows: 6 halo values from lanes l-1..l-3
        T w0[8], w1[8];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            w0[2 * j] = __shfl_up_sync(0xffffffffu, q0.v[0], 3 - j);
            w0[2 * j + 1] = __shfl_up_sync(0xffffffffu, q0.v[1], 3 - j);
            w1[2 * j] = __shfl_up_sync(0xffffffffu, q1.v[0], 3 - j);
            w1[2 * j + 1] = __shfl_up_sync(0xffffffffu, q1.v[1], 3 - j);
        }
        w0[6] = q0.v[0]; w0[7] = q0.v[1];
        w1[6] = q1.v[0]; w1[7] = q1.v[1];
        T o[4][4];
        dwt_all_bank<T, 2>(w0, w1, h0, o[0]);
        dwt_all_bank<T, 4>(w0, w1, h1, o[1]);
        dwt_all_bank<T, 6>(w0, w1, h2, o[2]);
        dwt_all_bank<T, 8>(w0, w1, h3, o[3]);
        if (jx < JX0) {                              // warm-up rows of the history
            zfetch(jx + 1);
            continue;
        }
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const DwtAllSlot &d = g.s[s];
            if (tv[s] && jx < d.C0) {
                T *dst = pq[s] + (long)jx * d.ldn;
                if (a.last) fz.put(dst - (long)d.qs, o[s][0], za[s][0], zf[s][0]);
                else pll[s][(long)jx * d.ld_out] = o[s][0];
                fz.put(dst, o[s][1], za[s][1], zf[s][1]);
                fz.put(dst + d.qs, o[s][2], za[s][2], zf[s][2]);
                fz.put(dst + 2 * (long)d.qs, o[s][3], za[s][3], zf[s][3]);
            }
        }
        if (sv) {
#pragma unroll
            for (int rr = 0; rr < 2; ++rr)
                if (2 * jx + rr < n0) {
                    T *dst = pself + (long)(2 * jx + rr) * g.ld_self;
                    Vec2<T> q = rr ? q1 : q0;
                    if (fz.on) {
                        q.v[0] = (zsa[rr].v[0] + fz.sigma * q.v[0]) * zsf[rr].v[0];
                        q.v[1] = (zsa[rr].v[1] + fz.sigma * q.v[1]) * zsf[rr].v[1];
                    }
                    stv2(dst, q);
                }
        }
        zfetch(jx + 1);
    }
    cp_async_wait<0>();
}

// 'self' basis alone: alphao[i, 0:ny, 0:nx] = x.T (pfb/operators/psi.py:196-199)
template <class T>
__global__ void __launch_bounds__(256)
k_self_dot(const T *__restrict__ img, T *alpha, LevelArgs a, int basis) {
    PFB_DYN_SMEM(raw);
    T *smem = reinterpret_cast<T *>(raw);
    const int band = blockIdx.z;
    T *slab = alpha + ((long)band * a.nbasis + basis) * (long)a.Nymax * a.Nxmax;
    self_tile<T>(img + band * a.img_band_stride, a.nx, a.ny, slab, a.Nxmax, blockIdx.y, blockIdx.x, smem);
}


// 'self' basis in the natural-orientation layout: its block is the image itself
// (row stride ld_self).  One thread copies a pair of columns.
template <class T>
__global__ void __launch_bounds__(256)
k_self_nat(const T *__restrict__ img, T *cn, int nx, int ny, int ld_self, long cn_stride,
           int nbasis, int basis, long total) {
    const int hy = ny / 2;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % hy);
        const long r = i / hy;                 // band * nx + row
        const long band = r / nx, row = r - band * nx;
        stv2(cn + (band * nbasis + basis) * cn_stride + row * ld_self + 2 * c, ldv2(img + r * ny + 2 * c));
    }
}

}  // namespace pfb
