// Streaming synthesis (IDWT) for Psi.hdot: the mirror of wavelet_stream.cuh.
//
// The reference coefficient slab is TRANSPOSED w.r.t. the image (slab rows = y,
// columns = x; pfb/wavelets/wavelets.py:143-144).  Reading it in the order the
// synthesis needs (a lane per image column, walking down the image rows) would
// be a 4-byte gather, so Psi.hdot runs in two steps:
//
//  1. k_slab_to_nat  -- one tiled transpose of every pyramid quadrant of the slab
//     into a natural-orientation coefficient scratch Cn[jx][jy] (32x32 tiles,
//     128-byte runs on both sides), with `2*alpha - alpha2` of
//     pfb/opt/primal_dual.py:137 fused into the copy.  The 'self' basis becomes
//     its image-orientation copy.
//  2. k_idwt_stream  -- per level, a warp owns 29 image column pairs and walks
//     down the coefficient rows.  Each lane loads its four quadrant values of a
//     row with coalesced 128-byte warp loads (no shared memory at all), does the
//     y synthesis with warp shuffles (halo from lanes l+1..l+3), the x synthesis
//     from a register history of the last F/2-1 rows, and -- on the finest level
//     -- accumulates all bases in registers in a FIXED order (db1, db2, db3, db4,
//     self), so the image is written exactly once and the sum is deterministic.
//
// Reference: idwt2d / idwt2d_level / upsampling_convolution_valid_sf
// (pfb/wavelets/wavelets.py:98-123, 216-315), psi_band.hdot (psi.py:221-256).
// out[2m+p] = sum_{j<=H} filt[2j+p] * in[m+H-j],  H = F/2-1  (valid part only).
#pragma once

namespace pfb {

static const int kSynSlots = 4;       // slot s <-> Daubechies db(s+1), F = 2(s+1)

struct SynSlot {
    int basis;             // index into the bases list, -1 = absent
    int C0, C1;            // coefficient sizes (x, y) of this level
    int ldn;               // row stride of the natural-orientation quadrants
    int ld_ll;             // row stride of the LL scratch written by the coarser level
    int n0, n1;            // size of this level's output (2C - F + 2)
    int ld_out;            // row stride of the LL scratch this level writes (coarser levels)
    int qs;                // elements per quadrant (C0 * ldn); quadrants 1,2,3 are consecutive
    long off1;             // offset of quadrant 1 inside the (band,basis) Cn block; on the
                           // coarsest level quadrant 0 (LL) sits right in front of it
};

struct SynArgs {
    SynSlot s[kSynSlots];
    int nbasis, nband;
    int first, last;               // level is the finest / the coarsest
    int n0, n1;                    // finest level: image size (nx, ny)
    int ld_out;                    // finest level: row stride of the image
    int self_basis;                // index of 'self' (finest level only), -1 = absent
    int ld_self;                   // row stride of its image-orientation copy (offset 0 of its block)
    long cn_stride, scr_stride, img_band_stride;
};

// ---------------------------------------------------------------------------
// step 1: slab -> natural orientation.  grid: (tiles over cols, tiles over rows,
// band * nblk + blk); blk enumerates the (basis, quadrant) blocks of one level
// (or the 'self' block).  Block descriptor: source origin in the slab, sizes,
// destination offset / stride.
struct NatBlk {
    int basis;
    int row0, col0;        // slab origin (row = jy, col = jx)
    int nrow, ncol;        // extent in the slab (nrow = C1 or ny, ncol = C0 or nx)
    int ldn;               // destination row stride (dest is [col][row])
    long off;              // destination offset inside the (band,basis) Cn block
};
static const int kMaxNatBlk = 32;
struct NatArgs {
    NatBlk b[kMaxNatBlk];
    int nblk, nbasis;
    int Nxmax, Nymax;
    long cn_stride;
};

template <class T>
__global__ void __launch_bounds__(256)
k_slab_to_nat(const T *__restrict__ alpha, const T *__restrict__ alpha2, T *__restrict__ cn,
              NatArgs a) {
    __shared__ T tile[32][33];
    const int band = blockIdx.z / a.nblk, blk = blockIdx.z - band * a.nblk;
    const NatBlk &d = a.b[blk];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    if (c0 >= d.ncol || r0 >= d.nrow) return;
    const long so = ((long)band * a.nbasis + d.basis) * (long)a.Nymax * a.Nxmax;
    const T *src = alpha + so;
    const T *src2 = alpha2 ? alpha2 + so : nullptr;
    T *dst = cn + ((long)band * a.nbasis + d.basis) * a.cn_stride + d.off;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        const int r = r0 + i, c = c0 + tx;
        T v = (T)0;
        if (r < d.nrow && c < d.ncol) {
            const long o = (long)(d.row0 + r) * a.Nxmax + d.col0 + c;
            v = src[o];
            if (src2) v = (T)2 * v - src2[o];
        }
        tile[i][tx] = v;
    }
    __syncthreads();
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, r = r0 + tx;       // dest row = slab column, dest column = slab row
        if (c < d.ncol && r < d.nrow) dst[(long)c * d.ldn + r] = tile[tx][i];
    }
}

// the way back (solver exit): natural orientation -> the pyramid cells of the slab
template <class T>
__global__ void __launch_bounds__(256)
k_nat_to_slab(const T *__restrict__ cn, T *__restrict__ alpha, NatArgs a) {
    __shared__ T tile[32][33];
    const int band = blockIdx.z / a.nblk, blk = blockIdx.z - band * a.nblk;
    const NatBlk &d = a.b[blk];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    if (c0 >= d.ncol || r0 >= d.nrow) return;
    T *dst = alpha + ((long)band * a.nbasis + d.basis) * (long)a.Nymax * a.Nxmax;
    const T *src = cn + ((long)band * a.nbasis + d.basis) * a.cn_stride + d.off;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, r = r0 + tx;
        tile[i][tx] = (c < d.ncol && r < d.nrow) ? src[(long)c * d.ldn + r] : (T)0;
    }
    __syncthreads();
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        const int r = r0 + i, c = c0 + tx;
        if (r < d.nrow && c < d.ncol) dst[(long)(d.row0 + r) * a.Nxmax + d.col0 + c] = tile[tx][i];
    }
}

// ---------------------------------------------------------------------------
// step 2: streaming synthesis.
static const int kSynStrip = 28;      // output column PAIRS per warp (lanes 28..31: halo only);
                                      // a multiple of 4 keeps every 16-byte load aligned
static const int kSynWarps = 4;       // warps (adjacent strips) per CTA
static const int kSynHmax = 3;        // db4: F/2 - 1

template <class T> struct SynRow {    // everything one walk step consumes
    T c[kSynSlots][4];                // [slot][quadrant] at (ix = mx + H_slot, iy = MY0 + lane)
    Vec2<T> sf[2];                    // 'self' rows 2mx, 2mx+1 at columns 2my, 2my+1
};
// cp.async ring geometry (elements): one coefficient set of one walk step, and
// a stage (one set, or two when the input is 2*cn - cn2)
static const int kSynPD = 3;          // walk steps in flight per warp
__host__ __device__ constexpr int syn_set_elems() { return kSynSlots * 4 * 32 + 2 * 64; }
__host__ __device__ constexpr int syn_stage_elems(bool two) { return (two ? 2 : 1) * syn_set_elems(); }

// One coefficient row of one filter bank: y synthesis across lanes, x synthesis
// against the register history h[k-1] = U(row ix - k), accumulate into
// acc[2*px + py] (output row 2mx+px, column 2my+py).
template <class T, int F>
__device__ __forceinline__ void syn_slot_step(const T (&c)[4], T (&h)[F > 2 ? F / 2 - 1 : 1][4],
                                              T (&acc)[4]) {
    constexpr int H = F / 2 - 1, N = F / 2 - 1;      // N: index of the tap table
    T u[4];                                          // u[2*hx + py]
#pragma unroll
    for (int hx = 0; hx < 2; ++hx) {
        T e = (T)0, o = (T)0;
#pragma unroll
        for (int j = 0; j <= H; ++j) {
            const int d = H - j;                     // iy + H - j  <->  lane + d
            const T lo = d ? __shfl_down_sync(0xffffffffu, c[hx], d) : c[hx];
            const T hi = d ? __shfl_down_sync(0xffffffffu, c[2 + hx], d) : c[2 + hx];
            // one FMA per tap (a*lo + b*hi added to e would cost FMUL + FFMA + FADD)
            e = fma(tapc<T>(N, 2, 2 * j), lo, e);
            e = fma(tapc<T>(N, 3, 2 * j), hi, e);
            o = fma(tapc<T>(N, 2, 2 * j + 1), lo, o);
            o = fma(tapc<T>(N, 3, 2 * j + 1), hi, o);
        }
        u[2 * hx] = e;
        u[2 * hx + 1] = o;
    }
#pragma unroll
    for (int py = 0; py < 2; ++py) {
        T e = acc[py], o = acc[2 + py];
        e = fma(tapc<T>(N, 2, 0), u[py], e);
        e = fma(tapc<T>(N, 3, 0), u[2 + py], e);
        o = fma(tapc<T>(N, 2, 1), u[py], o);
        o = fma(tapc<T>(N, 3, 1), u[2 + py], o);
#pragma unroll
        for (int j = 1; j <= H; ++j) {
            e = fma(tapc<T>(N, 2, 2 * j), h[j - 1][py], e);
            e = fma(tapc<T>(N, 3, 2 * j), h[j - 1][2 + py], e);
            o = fma(tapc<T>(N, 2, 2 * j + 1), h[j - 1][py], o);
            o = fma(tapc<T>(N, 3, 2 * j + 1), h[j - 1][2 + py], o);
        }
        acc[py] = e;
        acc[2 + py] = o;
    }
    if (H > 0) {
#pragma unroll
        for (int k = H - 1; k >= 1; --k)
#pragma unroll
            for (int i = 0; i < 4; ++i) h[k][i] = h[k - 1][i];
#pragma unroll
        for (int i = 0; i < 4; ++i) h[0][i] = u[i];
    }
}

// SLOT < 0: finest level, every slot takes part (absent ones see zeros);
// SLOT >= 0: a coarser level of that one filter bank.  Compile-time on purpose:
// the warp shuffles must not sit under a run-time condition (the compiler then
// wraps each of them in a convergence-barrier sequence).
template <class T, bool TWO, int SLOT>
__device__ __forceinline__ void syn_walk(const T *__restrict__ cn, const T *__restrict__ cn2,
                                         const T *__restrict__ scr_in, T *out, const SynArgs &a,
                                         int nstrip, int nchunk, int NJ, long idx) {
    constexpr bool per_basis = SLOT >= 0;
    constexpr int sel = SLOT >= 0 ? SLOT : 0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nsg = (nstrip + kSynWarps - 1) / kSynWarps;
    const int sg = (int)(idx % nsg); idx /= nsg;
    const int chunk = (int)(idx % nchunk);
    const int band = (int)(idx / nchunk);
    const int strip = sg * kSynWarps + warp;
    if (band >= a.nband) return;
    if (per_basis && a.s[sel].basis < 0) return;
    const int n0 = per_basis ? a.s[sel].n0 : a.n0, n1 = per_basis ? a.s[sel].n1 : a.n1;
    const int ldo = per_basis ? a.s[sel].ld_out : a.ld_out;
    const int MY0 = kSynStrip * strip, MX0 = chunk * NJ;
    if (2 * MY0 >= n1 || 2 * MX0 >= n0) return;
    const int mx_end = (MX0 + NJ) * 2 < n0 ? MX0 + NJ : n0 / 2;
    const int iy = MY0 + lane;

    // Loader geometry: every copy instruction moves 16 bytes per lane, i.e. a
    // warp fetches RPI quadrant rows of 32 coefficients at once (fp32: all four
    // quadrants of a slot in ONE LDGSTS).  Strips start at multiples of 28
    // coefficients, rows are padded to 32 => every 16-byte piece is aligned.
    constexpr int EPV = 16 / (int)sizeof(T);          // elements per vector
    constexpr int SPR = 32 / EPV;                     // vectors per 32-coefficient row
    constexpr int RPI = 32 / SPR;                     // rows per copy instruction
    constexpr int NCI = 4 / RPI;                      // copy instructions per slot
    const int seg = lane % SPR, rsub = lane / SPR;
    const int vcol = MY0 + seg * EPV;                 // first coefficient column of this lane's vector

    unsigned mask = 0;
    const T *pv[kSynSlots][NCI];                      // per-lane source of row 0
    int ldv[kSynSlots][NCI], nbv[kSynSlots];
#pragma unroll
    for (int s = 0; s < kSynSlots; ++s) {
        const SynSlot &d = a.s[s];
        nbv[s] = 0;
#pragma unroll
        for (int i = 0; i < NCI; ++i) { pv[s][i] = cn; ldv[s][i] = 0; }
        if (d.basis < 0 || (per_basis && s != sel)) continue;
        mask |= 1u << s;
        const long bb = (long)band * a.nbasis + d.basis;
        int left = (d.C1 - vcol) * (int)sizeof(T);
        nbv[s] = left < 0 ? 0 : (left > 16 ? 16 : left);
#pragma unroll
        for (int i = 0; i < NCI; ++i) {
            const int q = i * RPI + rsub;
            if (q == 0 && !a.last) {                  // LL: what the coarser level wrote
                pv[s][i] = scr_in + bb * a.scr_stride + vcol;
                ldv[s][i] = d.ld_ll;
            } else {
                pv[s][i] = cn + bb * a.cn_stride + d.off1 + (long)(q - 1) * d.qs + vcol;
                ldv[s][i] = d.ldn;
            }
        }
    }
    const bool has_self = !per_basis && a.self_basis >= 0;
    const bool ocol = lane < kSynStrip && 2 * iy < n1;
    // self rows: 2 * kSynStrip elements each, SSPR vectors reserved per row
    constexpr int SSPR = 64 / EPV, SRPI = 32 / SSPR, SNCI = 2 / SRPI;
    const int sseg = lane % SSPR, srow = lane / SSPR;
    const int scol = 2 * MY0 + sseg * EPV;
    const T *pself = cn;
    int nbself = 0;
    if (has_self) {
        pself = cn + ((long)band * a.nbasis + a.self_basis) * a.cn_stride + scol;
        int left = (n1 - scol) * (int)sizeof(T);
        nbself = (sseg * EPV >= 2 * kSynStrip || left < 0) ? 0 : (left > 16 ? 16 : left);
    }

    // byte distance from a coefficient of cn to the same coefficient of cn2
    const long d2 = TWO ? (long)((const char *)cn2 - (const char *)cn) : 0;
    auto p2 = [&](const T *q) { return reinterpret_cast<const T *>(reinterpret_cast<const char *>(q) + d2); };

    // cp.async ring: kSynPD walk steps of this warp's inputs in flight, no
    // registers tied up, out-of-range rows / columns zero-filled by the copy.
    // Stage layout per coefficient set: [slot][quadrant][32], then self [row][64].
    PFB_DYN_SMEM(raw);
    constexpr int SET = syn_set_elems();
    T *ring = reinterpret_cast<T *>(raw) + warp * (kSynPD * syn_stage_elems(TWO));
    auto prefetch = [&](const int mx) {
        T *st = ring + ((mx - (MX0 - kSynHmax)) % kSynPD) * syn_stage_elems(TWO);
#pragma unroll
        for (int s = 0; s < kSynSlots; ++s) {
            if (!(mask >> s & 1)) continue;
            const SynSlot &d = a.s[s];
            const int ix = mx + s;                    // H of slot s is s
            const bool ok = ix >= MX0 && ix < d.C0 && mx < mx_end;
            const int nb = ok ? nbv[s] : 0;
            const int ixc = ok ? ix : 0;              // keep the (unused) address inside the array
#pragma unroll
            for (int i = 0; i < NCI; ++i) {
                const T *p = pv[s][i] + (long)ixc * ldv[s][i];
                T *dq = st + s * 128 + i * (RPI * 32) + lane * EPV;
                cp_async_zfill<16>(dq, p, nb);
                // coarser-level LL already is the synthesis of the combined coefficients
                if (TWO) cp_async_zfill<16>(dq + SET, p2(p), (i * RPI + rsub == 0 && !a.last) ? 0 : nb);
            }
        }
        if (has_self) {
            const bool ok = mx >= MX0 && mx < mx_end;
#pragma unroll
            for (int i = 0; i < SNCI; ++i) {
                const int k = i * SRPI + srow;
                const T *q = pself + (long)(ok ? 2 * mx + k : 0) * a.ld_self;
                T *dq = st + kSynSlots * 128 + k * 64 + sseg * EPV;
                cp_async_zfill<16>(dq, q, ok ? nbself : 0);
                if (TWO) cp_async_zfill<16>(dq + SET, p2(q), ok ? nbself : 0);
            }
        }
        cp_async_commit();
    };
    auto consume = [&](const int mx, SynRow<T> &c) {
        const T *st = ring + ((mx - (MX0 - kSynHmax)) % kSynPD) * syn_stage_elems(TWO) + lane;
        __syncwarp();                                  // vectors were fetched by other lanes
#pragma unroll
        for (int s = 0; s < kSynSlots; ++s) {
            if (!(mask >> s & 1)) continue;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                T v = st[s * 128 + q * 32];
                if (TWO && (q > 0 || a.last)) v = (T)2 * v - st[SET + s * 128 + q * 32];
                c.c[s][q] = v;
            }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            c.sf[k].v[0] = c.sf[k].v[1] = (T)0;
            if (has_self) {
                c.sf[k] = ldv2(st + lane + kSynSlots * 128 + k * 64);
                if (TWO) {
                    const Vec2<T> o = ldv2(st + lane + kSynSlots * 128 + k * 64 + SET);
                    c.sf[k].v[0] = (T)2 * c.sf[k].v[0] - o.v[0];
                    c.sf[k].v[1] = (T)2 * c.sf[k].v[1] - o.v[1];
                }
            }
        }
        __syncwarp();                                  // all lanes have read before the stage is refilled
    };

    T h0[1][4], h1[1][4], h2[2][4], h3[3][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        h0[0][i] = h1[0][i] = (T)0;
        h2[0][i] = h2[1][i] = (T)0;
        h3[0][i] = h3[1][i] = h3[2][i] = (T)0;
    }
    T *dst = out + (per_basis ? ((long)band * a.nbasis + a.s[sel].basis) * a.scr_stride
                              : band * a.img_band_stride) + 2 * iy;
    SynRow<T> cur;
#pragma unroll
    for (int s = 0; s < kSynSlots; ++s) cur.c[s][0] = cur.c[s][1] = cur.c[s][2] = cur.c[s][3] = (T)0;
#pragma unroll 1
    for (int j = 0; j < kSynPD; ++j) prefetch(MX0 - kSynHmax + j);
#pragma unroll 1
    for (int mx = MX0 - kSynHmax; mx < mx_end; ++mx) {
        cp_async_wait<kSynPD - 1>();                           // the oldest group (this mx) has landed
        consume(mx, cur);
        prefetch(mx + kSynPD);                                 // refill the stage just read
        T acc[4] = {cur.sf[0].v[0], cur.sf[0].v[1], cur.sf[1].v[0], cur.sf[1].v[1]};
        T part[4] = {(T)0, (T)0, (T)0, (T)0};
        if (SLOT < 0 || SLOT == 0) syn_slot_step<T, 2>(cur.c[0], h0, part);
        if (SLOT < 0 || SLOT == 1) syn_slot_step<T, 4>(cur.c[1], h1, part);
        if (SLOT < 0 || SLOT == 2) syn_slot_step<T, 6>(cur.c[2], h2, part);
        if (SLOT < 0 || SLOT == 3) syn_slot_step<T, 8>(cur.c[3], h3, part);
        if (mx >= MX0 && ocol) {
            Vec2<T> r0, r1;
            r0.v[0] = part[0] + acc[0]; r0.v[1] = part[1] + acc[1];
            r1.v[0] = part[2] + acc[2]; r1.v[1] = part[3] + acc[3];
            stv2(dst + (long)(2 * mx) * ldo, r0);
            stv2(dst + (long)(2 * mx + 1) * ldo, r1);
        }
    }
    cp_async_wait<0>();
}

// grid.x = [kSynSlots *] nband * nchunk * ceil(nstrip / kSynWarps).
// per_basis = 0 (finest level): a task sums every basis and writes the image;
// per_basis = 1 (coarser levels): the task index also selects the slot and the
// result is that basis' LL scratch for the next finer level.
template <class T, bool TWO>
__global__ void __launch_bounds__(32 * kSynWarps)
k_idwt_stream(const T *__restrict__ cn, const T *__restrict__ cn2, const T *__restrict__ scr_in,
              T *out, SynArgs a, int nstrip, int nchunk, int NJ, int per_basis) {
    long idx = blockIdx.x;
    if (!per_basis) { syn_walk<T, TWO, -1>(cn, cn2, scr_in, out, a, nstrip, nchunk, NJ, idx); return; }
    const int sel = (int)(idx % kSynSlots);
    idx /= kSynSlots;
    switch (sel) {
    case 0: syn_walk<T, TWO, 0>(cn, cn2, scr_in, out, a, nstrip, nchunk, NJ, idx); break;
    case 1: syn_walk<T, TWO, 1>(cn, cn2, scr_in, out, a, nstrip, nchunk, NJ, idx); break;
    case 2: syn_walk<T, TWO, 2>(cn, cn2, scr_in, out, a, nstrip, nchunk, NJ, idx); break;
    default: syn_walk<T, TWO, 3>(cn, cn2, scr_in, out, a, nstrip, nchunk, NJ, idx); break;
    }
}

}  // namespace pfb
