// pdn_cg2.cuh -- round-2 PCG iteration: three kernels, ONE reduction (and one peer exchange) per iteration.
//
//   K1  sell_xs_kernel<CG2>   w = A z with the INPUT VECTOR STAGED IN SHARED MEMORY (x-band staging), fused
//                             gamma = (r,z), delta = (w,z), rr = (r,r) and the restriction c_w = P^T w of the
//                             two-level preconditioner; the last CTA sums the partials (several GPUs: ONE
//                             peer-memory exchange carries the dots and c_w together), runs the
//                             Chronopoulos-Gear scalar recurrence and keeps the coarse residual by recurrence:
//                                 c_s = c_w + beta c_s,   r_c -= alpha c_s          (P^T applied to s and r)
//   K2  coarse_apply2_kernel  x_c = A_c^-1 r_c   (dense fp32 inverse, L2 resident; the rank-local level rides along)
//   K3  cg_update2_kernel     p = z + beta p; s = w + beta s; x += alpha p; r -= alpha s;
//                             z = D^-1 r + P x_c;  max|z| for the next restriction scale;
//                             several GPUs: its last CTA pushes the interface entries of z (halo exchange)
//
// Round 1 ran four kernels and three exchanges (dots after the SpMV, r_c after the update, halo after z) and
// restricted r in the update kernel (match.any / redux on its critical path: 12-15 us against 5 us for the
// plain update).  Here the restriction rides in the SpMV, whose issue slots are idle while it waits for HBM.
//
// Restriction arithmetic: 42-bit FIXED POINT per entry (2^38 steps per scale), 64-bit integer atomics per
// (warp, aggregate) into a table whose entries sit TAB_STRIDE words apart (atomics on one cache line serialise in
// L2).  Integer sums are order-independent, so the solve stays bit-reproducible run to run.  Scale: |w_i| <=
// 2 a_ii max|z| for a diagonally dominant row, so fix = 2^38 / (8 amax zmax) cannot clip; zmax is the maximum
// the previous K3 (or the start kernel) left in sc->zmax_bits.  r_c is re-restricted from r once per CUDA-graph
// launch (every check_every iterations) at the same resolution, so drift cannot accumulate
// (scripts/proto_v2_pcg.py: same iteration count as textbook PCG, drift of r_c <= 1e-3 of its norm).
//
// x-band staging (north star kernel 3): a CTA owns a contiguous range of SELL slices; in mesh-local numbering
// the columns of its rows sit within +-halo of the rows themselves (one grid row + the sort window), so the
// CTA bulk-copies x[row_lo - halo, row_hi + halo) into shared memory once (cp.async.bulk onto an mbarrier;
// long ranges in double-buffered chunks) and gathers from there; the few out-of-window columns (via partners on
// other layers, override rows) take the read-only global path.  With the gathers out of L2 the matrix needs no
// shared-memory ring: plain streaming loads (L1::no_allocate, L2 evict_first policy) keep enough bytes in flight
// at 32 resident warps per SM, and all shared memory goes to the window.
//
// Included by pdn_kernels.cuh (namespace pdnk) after pdn_sell.cuh and pdn_coarse.cuh.
#pragma once

constexpr int SX_WARPS = 16;
constexpr int SX_BLK = SX_WARPS * 32;
constexpr int SX_CAP = 13824;          // doubles of shared memory for x windows per CTA (108 KB -> 2 CTAs per SM)
constexpr int SX_U = 8;                // entries per row fetched per step
constexpr size_t SX_SMEM = (size_t)SX_CAP * 8;

// how the CTAs stage x (host: plan_xs in pdn_solve.cu)
struct XsPlan {
    int32_t halo;           // rows staged on either side of a chunk (even)
    int32_t chunk_slices;   // slices per chunk; >= slices per CTA means one chunk using the whole buffer
};

__device__ __forceinline__ double ld_hint(const double *p, uint64_t pol) {
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ int32_t ld_hint(const int32_t *p, uint64_t pol) {
    int32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    return v;
}

// ---- K1 ---------------------------------------------------------------------------------------------------
// MODE: PLAIN / CG / BICG_V / BICG_T / RESID as in pdn_spmv.cuh (drop-in for sell_tma_kernel), or CG2.
template <int MODE, bool DIST>
__global__ void __launch_bounds__(SX_BLK, 2) sell_xs_kernel(SellOp A, XsPlan P, const double *__restrict__ x,
                                                            double *__restrict__ y, const double *__restrict__ aux0,
                                                            Scal *sc, double *partials, const PeerLink *Lp, Cg2 cg) {
    extern __shared__ __align__(128) unsigned char sx_smem[];
    double *xs = reinterpret_cast<double *>(sx_smem);
    __shared__ uint64_t bars[2];
    __shared__ double red[3][SX_WARPS];
    constexpr int EMODE = MODE == CG2 ? CG : MODE;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_trigger();
    if (tid == 0) {
        tma::mbar_init(&bars[0], 1);
        tma::mbar_init(&bars[1], 1);
        tma::fence_barrier_init();
    }
    const int64_t s_begin = A.row_begin / SELL_C, s_end = (A.row_end + SELL_C - 1) / SELL_C;
    const int64_t s_cnt = s_end - s_begin;
    const int64_t s_lo = s_begin + (int64_t)blockIdx.x * s_cnt / gridDim.x;
    const int64_t s_hi = s_begin + (int64_t)(blockIdx.x + 1) * s_cnt / gridDim.x;
    const int n_chunk = s_hi > s_lo ? (int)((s_hi - s_lo + P.chunk_slices - 1) / P.chunk_slices) : 0;
    const uint64_t pol = tma::policy_evict_first();
    __syncthreads();
    pdl_wait();                       // everything above is independent of the kernel that wrote x
    if (MODE != PLAIN && MODE != RESID) {
        if (sc->done) return;
    }
    if (DIST && Lp != nullptr && cg.halo != nullptr) {     // the peers' interface entries of x (consumer-side wait)
        peer_halo_wait(*Lp, *cg.halo);
        __syncthreads();
    }
    const double fix = MODE == CG2 ? restrict_fix(sc) : 0.0;
    double acc[3] = {0.0, 0.0, 0.0};

    // window of chunk ch: rows [wlo, wlo + wlen) of x, into buffer (ch & 1) when there are several chunks
    auto window = [&](int ch, int64_t &c0, int64_t &c1, int64_t &wlo, int32_t &wlen) {
        c0 = s_lo + (int64_t)ch * P.chunk_slices;
        c1 = c0 + P.chunk_slices < s_hi ? c0 + P.chunk_slices : s_hi;
        wlo = c0 * SELL_C - P.halo;
        if (wlo < 0) wlo = 0;
        int64_t whi = c1 * SELL_C + P.halo;
        if (whi > A.n) whi = A.n;
        wlen = (int32_t)(whi - wlo);
    };
    auto issue = [&](int ch) {        // thread 0
        int64_t c0, c1, wlo;
        int32_t wlen;
        window(ch, c0, c1, wlo, wlen);
        const uint32_t bytes = (uint32_t)((wlen + 1) & ~1) * 8u;     // vectors carry slack beyond n
        double *dst = xs + (n_chunk > 1 ? (size_t)(ch & 1) * (SX_CAP / 2) : 0);
        tma::mbar_expect_tx(&bars[ch & 1], bytes);
        tma::bulk_g2s(dst, x + wlo, bytes, &bars[ch & 1]);
    };
    if (tid == 0) {
        if (n_chunk > 0) issue(0);
        if (n_chunk > 1) issue(1);
    }

    for (int ch = 0; ch < n_chunk; ch++) {
        int64_t c0, c1, wlo;
        int32_t wlen;
        window(ch, c0, c1, wlo, wlen);
        const double *xw = xs + (n_chunk > 1 ? (size_t)(ch & 1) * (SX_CAP / 2) : 0);
        int64_t s = c0 + warp;
        int32_t off = 0, off1 = 0;
        uint32_t mask = 0;
        if (s < c1) { off = __ldg(&A.soff[s]); off1 = __ldg(&A.soff[s + 1]); mask = __ldg(&A.smask[s]); }
        // first slice of this warp: its matrix loads are issued before the wait for the x window
        int32_t c[SX_U];
        double v[SX_U];
        int w = (off1 - off) >> 5;
        if (s < c1) {
#pragma unroll
            for (int j = 0; j < SX_U; j++) c[j] = j < w ? ld_hint(A.ci + off + j * SELL_C + lane, pol) : 0;
#pragma unroll
            for (int j = 0; j < SX_U; j++) v[j] = j < w ? ld_hint(A.va + off + j * SELL_C + lane, pol) : 0.0;
        }
        tma::mbar_wait(&bars[ch & 1], (uint32_t)(ch >> 1) & 1u);
        while (s < c1) {
            const int64_t sn = s + SX_WARPS;
            int32_t n_off = 0, n_off1 = 0;
            uint32_t n_mask = 0;
            if (sn < c1) { n_off = __ldg(&A.soff[sn]); n_off1 = __ldg(&A.soff[sn + 1]); n_mask = __ldg(&A.smask[sn]); }
            const int64_t row = s * SELL_C + lane;
            const bool live = row < A.row_end && !((mask >> lane) & 1u);
            double xi = 0.0, ai = 0.0;
            int32_t a1 = -1, a2 = -1;
            if (live) {
                if (EMODE == CG) xi = xw[row - wlo];
                if (EMODE != PLAIN) ai = aux0[row];
                if (MODE == CG2) {
                    a1 = __ldg(cg.agg + row);
                    if (cg.nc2 > 0) a2 = __ldg(cg.agg2 + row);
                }
            }
            double sum = 0.0;
#pragma unroll
            for (int j = 0; j < SX_U; j++) {
                const int32_t d = c[j] - (int32_t)wlo;
                const double xv = j < w ? ((uint32_t)d < (uint32_t)wlen ? xw[d] : __ldg(x + c[j])) : 0.0;
                sum = fma(v[j], xv, sum);
            }
            for (int k = SX_U; k < w; k += SX_U) {       // rows longer than SX_U entries (lumped elements, CGAL valence > 7)
#pragma unroll
                for (int j = 0; j < SX_U; j++) c[j] = k + j < w ? ld_hint(A.ci + off + (k + j) * SELL_C + lane, pol) : 0;
#pragma unroll
                for (int j = 0; j < SX_U; j++) v[j] = k + j < w ? ld_hint(A.va + off + (k + j) * SELL_C + lane, pol) : 0.0;
#pragma unroll
                for (int j = 0; j < SX_U; j++) {
                    const int32_t d = c[j] - (int32_t)wlo;
                    const double xv = k + j < w ? ((uint32_t)d < (uint32_t)wlen ? xw[d] : __ldg(x + c[j])) : 0.0;
                    sum = fma(v[j], xv, sum);
                }
            }
            // next slice of this warp: loads in flight while this one finishes
            const int wn = (n_off1 - n_off) >> 5;
            if (sn < c1) {
#pragma unroll
                for (int j = 0; j < SX_U; j++) c[j] = j < wn ? ld_hint(A.ci + n_off + j * SELL_C + lane, pol) : 0;
#pragma unroll
                for (int j = 0; j < SX_U; j++) v[j] = j < wn ? ld_hint(A.va + n_off + j * SELL_C + lane, pol) : 0.0;
            }
            if (live) row_epilogue<EMODE>(row, sum, xi, ai, y, acc);
            if (MODE == CG2) {
                restrict_q(sum, a1, fix, cg.tab);
                if (cg.nc2 > 0) restrict_q(sum, a2, fix, cg.tab2);
            }
            s = sn; off = n_off; off1 = n_off1; mask = n_mask; w = wn;
        }
        if (n_chunk > 1) {
            __syncthreads();          // every warp is done with this buffer
            if (tid == 0 && ch + 2 < n_chunk) issue(ch + 2);
        }
    }
    __syncthreads();
    const int32_t gwarp = (int32_t)((blockIdx.x * SX_BLK + tid) >> 5), n_gwarp = (int32_t)((gridDim.x * SX_BLK) >> 5);
    if (MODE == CG2) special_rows_cg2(A, x, y, aux0, acc, gwarp, n_gwarp, cg, fix);
    else special_rows<EMODE>(A, x, y, aux0, acc, gwarp, n_gwarp);
    if (MODE == PLAIN) return;
    if (MODE == CG2) __threadfence();      // this thread's table atomics before the CTA's ticket
    if (!grid_reduce<3, SX_WARPS>(acc, partials, &sc->ticket[0], red)) return;
    if (MODE == CG2) cg2_tail<DIST>(acc, sc, cg, Lp, fix);
    else dots_tail<EMODE, DIST>(acc, sc, Lp);
}

// band statistics of the reduced operator (setup): hist[b] counts the stored entries with |col - row| in
// [64 b, 64 b + 64), b < 63; hist[63] the rest.  The host picks the halo from it (plan_xs).
__global__ void __launch_bounds__(256) band_hist_kernel(SellOp A, int64_t s0, int64_t s1, unsigned long long *hist) {
    __shared__ unsigned int sh[64];
    if (threadIdx.x < 64) sh[threadIdx.x] = 0u;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int64_t n_warp = ((int64_t)gridDim.x * 256) >> 5;
    for (int64_t s = s0 + warp; s < s1; s += n_warp) {
        const int64_t r = s * SELL_C + lane;
        const int32_t off = A.soff[s], w = (A.soff[s + 1] - off) >> 5;
        for (int32_t k = 0; k < w; k++) {
            const int32_t c = A.ci[off + k * SELL_C + lane];
            if (A.va[off + k * SELL_C + lane] == 0.0) continue;     // padding
            int64_t d = c - r;
            d = d < 0 ? -d : d;
            const int b = d >= 63 * 64 ? 63 : (int)(d >> 6);
            atomicAdd(&sh[b], 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x < 64 && sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
}

// ---- K2 ---------------------------------------------------------------------------------------------------
// RECUR: first the coarse-residual recurrence with the c_w table K1 just filled,
//     c_s' = c_w + beta c_s,   r_c' = r_c - alpha c_s'          (P^T applied to s and r, pdn_cg2_tail.cuh)
// computed by EVERY CTA into shared memory (3 nc independent loads per CTA instead of nc; as a loop of K1's last
// CTA it cost 3-6 us); the first CTA of each level also stores c_s', r_c' into the slot of this iteration (the
// others still read the previous slot: no race).  !RECUR (start / restart): r_c as the restriction kernel left it.
// Then x_c = A_c^-1 r_c' from the fp32 copy of the inverse; CTA b owns CO_RB rows, the CTAs from grid1 on work on
// the rank-local level.  The matrix loads are issued before the programmatic-launch wait (setup data).
template <bool RECUR>
__global__ void __launch_bounds__(CO_BLK) coarse_apply2_kernel(const Scal *__restrict__ sc, Cg2 cg, const float *__restrict__ Ainv,
                                                               double *__restrict__ xc, int grid1,
                                                               const float *__restrict__ Ainv2, double *__restrict__ xc2,
                                                               int summed, const PeerLink *Lp) {
    __shared__ double rcd[CO_MAX];
    __shared__ double red[CO_RB][CO_BLK / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool lvl2 = (int)blockIdx.x >= grid1;
    const int row0 = (lvl2 ? (int)blockIdx.x - grid1 : (int)blockIdx.x) * CO_RB;
    const int32_t nc = lvl2 ? cg.nc2 : cg.nc;
    if (lvl2) { Ainv = Ainv2; xc = xc2; }
    float a[CO_RB][CO_JMAX];
#pragma unroll
    for (int q = 0; q < CO_RB; q++)
#pragma unroll
        for (int t = 0; t < CO_JMAX; t++) {
            const int j = tid + t * CO_BLK;
            a[q][t] = (row0 + q < nc && j < nc) ? __ldg(&Ainv[(size_t)(row0 + q) * nc + j]) : 0.f;
        }
    pdl_trigger();
    pdl_wait();
    if (sc->done) return;
    if (Lp != nullptr && blockIdx.x == 0 && tid == 0) peer_log(*Lp, 2);      // event log: K2 start (PADNE_B200_PEER_STATS)
    double *rc = lvl2 ? cg.rc2 : cg.rc, *cs = lvl2 ? cg.cs2 : cg.cs;
    const int out = sc->iter & 1, in = out ^ 1;
    if (RECUR) {
        const unsigned long long *tab = lvl2 ? cg.tab2 : cg.tab;
        const double alpha = sc->alpha, beta = sc->beta, invfix = sc->cw_invfix;
        const bool from_sum = summed && !lvl2;           // several GPUs: the global level's c_w was summed by K1's tail
        double cw[CO_JMAX], c0[CO_JMAX], r0[CO_JMAX];
#pragma unroll
        for (int t = 0; t < CO_JMAX; t++) {
            const int j = tid + t * CO_BLK;
            cw[t] = c0[t] = r0[t] = 0.0;
            if (j < nc) {
                cw[t] = from_sum ? __ldcg(cg.cw + j) : (double)(long long)__ldcg(tab + (size_t)j * TAB_STRIDE) * invfix;
                c0[t] = __ldcg(cs + (size_t)in * nc + j);
                r0[t] = __ldcg(rc + (size_t)in * nc + j);
            }
        }
#pragma unroll
        for (int t = 0; t < CO_JMAX; t++) {
            const int j = tid + t * CO_BLK;
            if (j < nc) {
                const double c = cw[t] + beta * c0[t];
                const double r = r0[t] - alpha * c;
                rcd[j] = r;
                if (row0 == 0) { cs[(size_t)out * nc + j] = c; rc[(size_t)out * nc + j] = r; }
            }
        }
    } else {
        for (int j = tid; j < nc; j += CO_BLK) rcd[j] = __ldcg(rc + (size_t)out * nc + j);
    }
    __syncthreads();
    double acc[CO_RB];
#pragma unroll
    for (int q = 0; q < CO_RB; q++) {
        acc[q] = 0.0;
#pragma unroll
        for (int t = 0; t < CO_JMAX; t++) {
            const int j = tid + t * CO_BLK;
            if (j < nc) acc[q] = fma((double)a[q][t], rcd[j], acc[q]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[q] += __shfl_down_sync(0xffffffffu, acc[q], o);
        if (lane == 0) red[q][warp] = acc[q];
    }
    __syncthreads();
    if (tid < CO_RB && row0 + tid < nc) {
        double sum = 0.0;
        for (int k = 0; k < CO_BLK / 32; k++) sum += red[tid][k];
        xc[row0 + tid] = sum;
    }
    if (Lp != nullptr && blockIdx.x == 0 && tid == 0) peer_log(*Lp, 3);      // K2's first CTA done
}

// ---- K3 ---------------------------------------------------------------------------------------------------
// ZONLY: start / restart -- only z = D^-1 r + P x_c (p = s = 0 were set by the start kernel).
// Two rows per thread.  max|z| goes to sc->zmax_bits[iter & 1] (bits of a non-negative double order like integers).
// Several GPUs with a small interface: the last CTA pushes this rank's interface entries of z into the peers'
// copies, raises the halo flags and waits for the peers' (as coarse_z_kernel, pdn_coarse.cuh).
template <int BLK, bool ZONLY>
__global__ void __launch_bounds__(BLK) cg_update2_kernel(int64_t r0, int64_t n, Scal *sc, const double *__restrict__ dinv,
                                                         const double *__restrict__ w, double *__restrict__ z,
                                                         double *__restrict__ p, double *__restrict__ s,
                                                         double *__restrict__ x, double *__restrict__ r,
                                                         const int32_t *__restrict__ agg, const double *__restrict__ xc,
                                                         const int32_t *__restrict__ agg2, const double *__restrict__ xc2,
                                                         const PeerLink *Lp, const HaloSend *hsp, size_t vec_off,
                                                         const int32_t *__restrict__ send_idx, unsigned int *ticket,
                                                         unsigned long long *tab, int32_t nc, unsigned long long *tab2,
                                                         int32_t nc2) {
    pdl_trigger();
    pdl_wait();
    if (sc->done) return;
    if (Lp != nullptr && blockIdx.x == 0 && threadIdx.x == 0) peer_log(*Lp, 7);
    // K2 has consumed the c_w tables of this iteration: clear them for the next K1 (spread over the grid)
    if (!ZONLY) {
        for (int j = blockIdx.x * BLK + threadIdx.x; j < nc; j += gridDim.x * BLK) tab[(size_t)j * TAB_STRIDE] = 0ull;
        for (int j = blockIdx.x * BLK + threadIdx.x; j < nc2; j += gridDim.x * BLK) tab2[(size_t)j * TAB_STRIDE] = 0ull;
    }
    const double alpha = sc->alpha, beta = sc->beta;
    const int64_t npair = (n - r0 + 1) >> 1;
    double zm = 0.0;
    for (int64_t t = (int64_t)blockIdx.x * BLK + threadIdx.x; t < npair; t += (int64_t)gridDim.x * BLK) {
        const int64_t i = r0 + 2 * t;
        if (i + 1 < n) {
            double ra, rb;
            if (ZONLY) {
                const double2 rr = ld2(r + i);
                ra = rr.x;
                rb = rr.y;
            } else {
                const double2 zz = ld2(z + i), pp = ld2(p + i), ww = ld2(w + i), ss = ld2(s + i), rr = ld2(r + i);
                const double2 xx = __ldcs(reinterpret_cast<const double2 *>(x + i));   // x is touched once per iteration
                const double pa = zz.x + beta * pp.x, pb = zz.y + beta * pp.y;
                const double sa = ww.x + beta * ss.x, sb = ww.y + beta * ss.y;
                ra = rr.x - alpha * sa;
                rb = rr.y - alpha * sb;
                st2(p + i, pa, pb);
                st2(s + i, sa, sb);
                __stcs(reinterpret_cast<double2 *>(x + i), make_double2(xx.x + alpha * pa, xx.y + alpha * pb));
                st2(r + i, ra, rb);
            }
            const double2 dd = ld2(dinv + i);
            double ca = 0.0, cb = 0.0;
            if (agg != nullptr) {
                const int2 aa = *reinterpret_cast<const int2 *>(agg + i);
                ca = aa.x >= 0 ? __ldg(&xc[aa.x]) : 0.0;
                cb = aa.y == aa.x ? ca : (aa.y >= 0 ? __ldg(&xc[aa.y]) : 0.0);
                if (agg2 != nullptr) {
                    const int2 bb = *reinterpret_cast<const int2 *>(agg2 + i);
                    const double la = bb.x >= 0 ? __ldg(&xc2[bb.x]) : 0.0;
                    ca += la;
                    cb += bb.y == bb.x ? la : (bb.y >= 0 ? __ldg(&xc2[bb.y]) : 0.0);
                }
            }
            const double za = dd.x * ra + ca, zb = dd.y * rb + cb;
            st2(z + i, za, zb);
            zm = fmax(zm, fmax(fabs(za), fabs(zb)));
        } else {
            double ra;
            if (ZONLY) {
                ra = r[i];
            } else {
                const double pi = z[i] + beta * p[i];
                const double si = w[i] + beta * s[i];
                ra = r[i] - alpha * si;
                p[i] = pi;
                s[i] = si;
                x[i] += alpha * pi;
                r[i] = ra;
            }
            double c = 0.0;
            if (agg != nullptr) {
                const int32_t a = agg[i];
                c = a >= 0 ? __ldg(&xc[a]) : 0.0;
                if (agg2 != nullptr) {
                    const int32_t b = agg2[i];
                    c += b >= 0 ? __ldg(&xc2[b]) : 0.0;
                }
            }
            const double za = dinv[i] * ra + c;
            z[i] = za;
            zm = fmax(zm, fabs(za));
        }
    }
    // max |z| of this rank's rows (order-independent: an integer max on the bit pattern)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) zm = fmax(zm, __shfl_xor_sync(0xffffffffu, zm, o));
    if ((threadIdx.x & 31) == 0 && zm > 0.0)
        atomicMax(&sc->zmax_bits[sc->iter & 1], (unsigned long long)__double_as_longlong(zm));
    if (Lp == nullptr) return;
    // ---- several GPUs: the last CTA is the halo exchange of z ----
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    if (threadIdx.x == 0) *ticket = 0u;
    __threadfence();
    const PeerLink &L = *Lp;
    const HaloSend &hs = *hsp;
    const unsigned long long *mine = L.arena[L.me] + vec_off;
    for (int k = 0; k < L.world; k++) {
        if (k == L.me || hs.cnt[k] == 0) continue;
        unsigned long long *dst = L.arena[k] + vec_off;
        const int32_t *idx = send_idx + hs.off[k];
        for (int32_t q0 = 0; q0 < hs.cnt[k]; q0 += 8 * BLK) {
            int32_t g[8];
            unsigned long long v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int32_t q = q0 + u * BLK + (int32_t)threadIdx.x;
                g[u] = q < hs.cnt[k] ? __ldg(&idx[q]) : -1;
            }
#pragma unroll
            for (int u = 0; u < 8; u++) v[u] = g[u] >= 0 ? __ldcg(&mine[g[u]]) : 0ull;
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (g[u] >= 0) dst[g[u]] = v[u];
        }
    }
    peer_halo_post(L, hs);      // the wait is K1's (peer_halo_wait, pdn_peer_prim.cuh)
}

// ---- start / resync -------------------------------------------------------------------------------------------
// restriction of r into the fixed-point accumulator rc_fix (slot 0), 2^38 steps per ||r||_2 (as precise as K1's:
// an error made here stays in r_c until the next re-restriction);
// START also clears p and s.  Followed (several GPUs) by the integer sum over the ranks and by cg2_convert_kernel.
template <bool START>
__global__ void __launch_bounds__(RPB) cg2_restrict_kernel(int64_t r0, int64_t n, const Scal *__restrict__ sc,
                                                           const double *__restrict__ r, double *__restrict__ p,
                                                           double *__restrict__ s, const int32_t *__restrict__ agg,
                                                           long long *rc_fix, const int32_t *__restrict__ agg2,
                                                           long long *rc_fix2) {
    if (!START && sc->done) return;
    const double fix = Q_STEPS / sqrt(fmax(sc->rr, 1e-300));
    unsigned long long *rc = reinterpret_cast<unsigned long long *>(rc_fix);
    unsigned long long *rc2 = reinterpret_cast<unsigned long long *>(rc_fix2);
    for (int64_t i0 = r0 + (int64_t)blockIdx.x * RPB; i0 < n; i0 += (int64_t)gridDim.x * RPB) {
        const int64_t i = i0 + threadIdx.x;
        const bool act = i < n;
        double ri = 0.0;
        if (act) {
            ri = r[i];
            if (START) { p[i] = 0.0; s[i] = 0.0; }
        }
        restrict_q<1>(ri, act ? agg[i] : -1, fix, rc);
        if (agg2 != nullptr) restrict_q<1>(ri, act ? agg2[i] : -1, fix, rc2);
    }
}

// r_c = rc_fix / fix (fp64), accumulators cleared; START: c_s = 0 and the restriction tables of K1 are cleared
template <bool START>
__global__ void __launch_bounds__(256) cg2_convert_kernel(Scal *sc, Cg2 cg, long long *rc_fix, long long *rc_fix2) {
    if (!START && sc->done) return;
    const double scale = sqrt(fmax(sc->rr, 1e-300)) / Q_STEPS;
    const size_t slot = (size_t)(sc->iter & 1);       // the slot the next K2 reads (iter advances in K1's tail)
    for (int j = blockIdx.x * 256 + threadIdx.x; j < cg.nc; j += gridDim.x * 256) {
        cg.rc[slot * cg.nc + j] = (double)rc_fix[j] * scale;
        rc_fix[j] = 0;
        if (START) { cg.cs[slot * cg.nc + j] = 0.0; cg.tab[(size_t)j * TAB_STRIDE] = 0ull; }
    }
    for (int j = blockIdx.x * 256 + threadIdx.x; j < cg.nc2; j += gridDim.x * 256) {
        cg.rc2[slot * cg.nc2 + j] = (double)rc_fix2[j] * scale;
        rc_fix2[j] = 0;
        if (START) { cg.cs2[slot * cg.nc2 + j] = 0.0; cg.tab2[(size_t)j * TAB_STRIDE] = 0ull; }
    }
    if (START && blockIdx.x == 0 && threadIdx.x == 0) { sc->zmax_bits[0] = 0ull; sc->zmax_bits[1] = 0ull; }
}
