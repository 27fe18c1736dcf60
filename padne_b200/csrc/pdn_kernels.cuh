// pdn_kernels.cuh -- kernels (3) and (4): the CSR-stream SpMV with fused dot products, the fused
// Krylov vector updates and the Jacobi preconditioner.   Included by pdn_solve.cu only.
//
// Replaces scipy.sparse.linalg.spsolve at padne/solver.py:772-775 (SuperLU) by an iterative solve;
// the operator is the reduced system  A = -(Z^T Q L Z)  described in pdn_solve.cu.
//
// The SpMV (pdn_spmv.cuh) streams the CSR arrays through a TMA bulk-copy pipeline.
//
// Dot products are fused into the kernel that produces their operands.  Every block writes one
// partial per accumulator; the LAST block to finish (ticket counter) adds the partials in a
// fixed order and runs the scalar recurrence, so reductions are deterministic and cost no launch.
#pragma once

#include "pdn_common.cuh"

namespace pdnk {

constexpr int RPB = 256;       // threads per block of the vector kernels
constexpr int MAX_PART = 2048; // upper bound on blocks that write reduction partials

// device-resident scalars of one Krylov solve
struct Scal {
    // Chronopoulos-Gear PCG
    double gamma, gamma_old, delta, alpha, beta, rr;
    // BiCGSTAB
    double rho, rho_old, omega, r0v, ts, tt;
    // stopping
    double tol2;       // squared absolute threshold on ||r||_2
    double bnorm2;
    int32_t iter, done, breakdown, first;
    int32_t max_iter, pad0, pad1, pad2;
    unsigned int ticket[4];
    double part[4];    // this rank's partial sums awaiting the all-reduce (multi-GPU)
    // round-2 PCG iteration (pdn_cg2.cuh)
    unsigned long long zmax_bits[2];   // bits of max|z_i| over this rank's rows, slot = iteration parity
    double amax;                       // largest diagonal of the reduced operator (restriction scale, residual floor)
    double cw_invfix;                  // 1 / fixed-point scale of the c_w table the last K1 filled
};

// CG2: CG with the restriction of w and the coarse-residual recurrence in the tail (pdn_cg2.cuh)
enum Mode { PLAIN = 0, CG = 1, BICG_V = 2, BICG_T = 3, RESID = 4, CG2 = 5 };

// The reduced operator: main CSR part (rows flagged by a negative row pointer are skipped) plus a
// tiny override CSR for the special rows (constraint terminals), handled by block 0.
struct Operator {
    int64_t n;                 // rows
    const int32_t *rp;         // [n+1] row pointers; rp[i] < 0 flags a special row, offset = ~rp[i]
    const int32_t *ci;         // column indices
    const double *va;          // values
    int32_t n_special;         // override rows
    const int32_t *sp_row;     // [n_special] row index of each override row
    const int32_t *sp_rp;      // [n_special+1]
    const int32_t *sp_ci;
    const double *sp_va;
};

__device__ __forceinline__ int32_t rp_decode(int32_t v) { return v < 0 ? ~v : v; }

template <int NACC, int NW = RPB / 32>
__device__ __forceinline__ void block_reduce(double (&acc)[NACC], double (*sm)[NW]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int a = 0; a < NACC; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[a] += __shfl_down_sync(0xffffffffu, acc[a], o);
        if (lane == 0) sm[a][warp] = acc[a];
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int a = 0; a < NACC; a++) {
            double v = lane < NW ? sm[a][lane] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            acc[a] = v;  // valid in thread 0
        }
    }
    __syncthreads();
}

// Write this block's partials; returns true (to all threads) in the last block to arrive, with
// acc[] holding the grid totals summed in block order.
template <int NACC, int NW = RPB / 32>
__device__ __forceinline__ bool grid_reduce(double (&acc)[NACC], double *partials, unsigned int *ticket,
                                            double (*sm)[NW]) {
    __shared__ bool is_last;
    block_reduce<NACC, NW>(acc, sm);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int a = 0; a < NACC; a++) partials[a * MAX_PART + blockIdx.x] = acc[a];
        __threadfence();
        unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return false;
    __threadfence();
#pragma unroll
    for (int a = 0; a < NACC; a++) {
        double v = 0.0;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += NW * 32) v += __ldcg(&partials[a * MAX_PART + b]);
        acc[a] = v;
    }
    block_reduce<NACC, NW>(acc, sm);
    if (threadIdx.x == 0) *ticket = 0u;
    return true;
}

// ------------------------------------------------------------------------------------------------
// SpMV with fused dots (TMA-pipelined CSR-stream kernel)
//   PLAIN : y = A x
//   CG    : y = A x (x = z);   gamma = r.z, delta = z.y, rr = r.r        aux0 = r
//   BICG_V: y = A x (x = ph);  r0v = rhat.y                               aux0 = rhat
//   BICG_T: y = A x (x = sh);  ts = y.s, tt = y.y                         aux0 = s
//   RESID : y = b - A x;       rr = y.y                                   aux0 = b
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double2 ld2(const double *p) { return *reinterpret_cast<const double2 *>(p); }
__device__ __forceinline__ void st2(double *p, double a, double b) { *reinterpret_cast<double2 *>(p) = make_double2(a, b); }
#include "pdn_peer_prim.cuh"
#include "pdn_spmv.cuh"
#include "pdn_cg2_tail.cuh"
#include "pdn_sell.cuh"
#include "pdn_coarse.cuh"
#include "pdn_cg2.cuh"

// scalar recurrences of the BiCGSTAB vector kernels (one thread, after the grid-wide reduction)
__device__ __forceinline__ void bicg_x_finalize(const double (&acc)[3], Scal *sc) {
    const double rho_new = acc[0], rr = acc[2];
    sc->rr = rr;
    sc->iter += 1;
    if (rr <= sc->tol2) { sc->done = 1; return; }
    if (sc->iter >= sc->max_iter) { sc->done = 2; return; }
    const double beta = (rho_new / sc->rho) * (sc->alpha / sc->omega);
    if (!(beta == beta) || !(fabs(beta) < 1e300) || sc->omega == 0.0) { sc->breakdown = 1; sc->done = 3; return; }
    sc->rho_old = sc->rho;
    sc->rho = rho_new;
    sc->beta = beta;
}
__device__ __forceinline__ void bicg_init_finalize(const double (&acc)[3], Scal *sc) {
    sc->rho = acc[0];
    sc->rho_old = acc[0];
    sc->rr = acc[0];
    sc->omega = 1.0;
    sc->alpha = 1.0;
    if (acc[0] <= sc->tol2) sc->done = 1;
}

// Multi-GPU: the recurrence that follows an all-reduce of sc->part[0..2] (one thread).
//   WHICH: 0..4 = spmv_finalize<Mode>, 5 = bicg_x_finalize, 6 = bicg_init_finalize, 7 = *out = part[0]
template <int WHICH>
__global__ void scalar_kernel(Scal *sc, double *out) {
    if (WHICH != RESID && WHICH != 7 && WHICH != 6 && sc->done) return;
    const double acc[3] = {sc->part[0], sc->part[1], sc->part[2]};
    if (WHICH == 5) bicg_x_finalize(acc, sc);
    else if (WHICH == 6) bicg_init_finalize(acc, sc);
    else if (WHICH == 7) *out = acc[0];
    else spmv_finalize<WHICH>(acc, sc);
}

// ------------------------------------------------------------------------------------------------
// fused vector kernels (grid-stride, one element per thread per step) over the rows [r0, n) this
// rank owns (r0 = 0 and n = all rows on a single GPU)
// ------------------------------------------------------------------------------------------------
// Chronopoulos-Gear update:  p = z + beta p;  s = w + beta s;  x += alpha p;  r -= alpha s;  z = dinv r
// TL (two-level preconditioner): z is produced later by coarse_z_kernel; here the new residual is
// restricted onto the aggregates, r_c[agg] += r (fixed point, pdn_coarse.cuh).
// Two rows per thread (128-bit loads and stores, half the address arithmetic and half the warp collectives
// of the restriction): the one-row version was instruction-issue bound (~250 instructions per 32 rows) once
// the vectors stayed in L2.  r0 is a multiple of the sort window, so row pairs are the same on every rank count.

template <bool TL>
__global__ void __launch_bounds__(RPB) cg_update_kernel(int64_t r0, int64_t n, const Scal *__restrict__ sc,
                                                        const double *__restrict__ dinv,
                                                        const double *__restrict__ w, double *__restrict__ z,
                                                        double *__restrict__ p, double *__restrict__ s,
                                                        double *__restrict__ x, double *__restrict__ r,
                                                        const int32_t *__restrict__ agg, long long *rc_fix, int32_t nc) {
    pdl_trigger();
    pdl_wait();
    if (sc->done) return;
    const double alpha = sc->alpha, beta = sc->beta;
    const double fix = TL ? CO_RFIX32 / sqrt(fmax(sc->rr, 1e-300)) : 0.0;
    unsigned long long *rc = TL ? reinterpret_cast<unsigned long long *>(rc_fix) + (size_t)(sc->iter & 1) * nc : nullptr;
    const int64_t npair = (n - r0 + 1) >> 1;
    for (int64_t t0 = (int64_t)blockIdx.x * RPB; t0 < npair; t0 += (int64_t)gridDim.x * RPB) {
        const int64_t t = t0 + threadIdx.x;
        const int64_t i = r0 + 2 * t;
        const bool act = t < npair, full = act && i + 1 < n;
        double ra = 0.0, rb = 0.0;
        if (full) {
            const double2 zz = ld2(z + i), pp = ld2(p + i), ww = ld2(w + i), ss = ld2(s + i), rr = ld2(r + i), xx = ld2(x + i);
            const double pa = zz.x + beta * pp.x, pb = zz.y + beta * pp.y;
            const double sa = ww.x + beta * ss.x, sb = ww.y + beta * ss.y;
            ra = rr.x - alpha * sa;
            rb = rr.y - alpha * sb;
            st2(p + i, pa, pb);
            st2(s + i, sa, sb);
            st2(x + i, xx.x + alpha * pa, xx.y + alpha * pb);
            st2(r + i, ra, rb);
            if (!TL) {
                const double2 dd = ld2(dinv + i);
                st2(z + i, dd.x * ra, dd.y * rb);
            }
        } else if (act) {
            const double pi = z[i] + beta * p[i];
            const double si = w[i] + beta * s[i];
            ra = r[i] - alpha * si;
            p[i] = pi;
            s[i] = si;
            x[i] += alpha * pi;
            r[i] = ra;
            if (!TL) z[i] = dinv[i] * ra;
        }
        if (TL) {
            int32_t a0 = -1, a1 = -1;
            if (full) {
                const int2 aa = *reinterpret_cast<const int2 *>(agg + i);
                a0 = aa.x;
                a1 = aa.y;
            } else if (act) {
                a0 = agg[i];
            }
            restrict_add_pair(ra, a0, rb, a1, fix, rc);
        }
    }
}

// The two-level update: as above without z, plus the restriction of the new residual.  One global atomic per
// (warp, aggregate) -- 62K of them onto ~90 cache lines at 1M rows -- cost more than the whole vector update,
// so every CTA owns a CONTIGUOUS range of row pairs (few aggregates), sums them in a shared-memory table and
// flushes its non-zero entries at the end: ~2K global atomics per launch.  1024 threads per CTA, two CTAs per
// SM, so that the table is cleared and flushed by two entries per thread.
constexpr int TLB = 1024;
// LOCAL: also restrict into the rank-local second level (partitioned solves from 4 ranks on)
template <bool LOCAL>
__global__ void __launch_bounds__(TLB, 1) cg_update_tl_kernel(int64_t r0, int64_t n, const Scal *__restrict__ sc,
                                                              const double *__restrict__ w, const double *__restrict__ z,
                                                              double *__restrict__ p, double *__restrict__ s,
                                                              double *__restrict__ x, double *__restrict__ r,
                                                              const int32_t *__restrict__ agg, long long *rc_fix, int32_t nc,
                                                              const PeerLink *Lp, unsigned int *ticket,
                                                              const int32_t *__restrict__ agg2, long long *rc_fix2, int32_t nc2,
                                                              const int32_t *arange) {
    // Restriction tables: the 26-bit fixed-point value of a row goes in as TWO 14-bit limbs with native 32-bit shared
    // atomics (ATOMS.ADD), straight from every lane -- no warp collective.  (Round 1 grouped the lanes of a warp by
    // aggregate with match.any + redux.sync and added 64-bit sums, which on sm_100a is a CAS loop,
    // ATOMS.CAST.SPIN.64; a 32-row warp holds 3.5 distinct aggregates on these meshes, so the collectives ran 3-4
    // passes: 12-15 us per launch against 5 us for the plain update.  The integer sums are the same, so is the solve.)
    __shared__ int tab[2][CO_MAX];
    __shared__ int tab2[2][LOCAL ? CO_MAX : 1];   // rank-local second level
    pdl_trigger();
    for (int a = threadIdx.x; a < nc; a += TLB) tab[0][a] = tab[1][a] = 0;
    if (LOCAL)
        for (int a = threadIdx.x; a < nc2; a += TLB) tab2[0][a] = tab2[1][a] = 0;
    pdl_wait();
    if (sc->done) return;
    if (Lp != nullptr && blockIdx.x == 0 && threadIdx.x == 0) peer_log(*Lp, 6);
    const double alpha = sc->alpha, beta = sc->beta;
    const double fix = CO_RFIX32 / sqrt(fmax(sc->rr, 1e-300));
    unsigned long long *rc = reinterpret_cast<unsigned long long *>(rc_fix) + (size_t)(sc->iter & 1) * nc;
    unsigned long long *rc2 = LOCAL ? reinterpret_cast<unsigned long long *>(rc_fix2) + (size_t)(sc->iter & 1) * nc2 : nullptr;
    const int64_t npair = (n - r0 + 1) >> 1;
    const int64_t per = ((npair + gridDim.x - 1) / gridDim.x + 31) & ~(int64_t)31;   // whole warps
    const int64_t t_lo = (int64_t)blockIdx.x * per, t_hi = t_lo + per < npair ? t_lo + per : npair;
    __syncthreads();    // table cleared
    for (int64_t t0 = t_lo; t0 < t_hi; t0 += TLB) {
        const int64_t t = t0 + threadIdx.x;
        const int64_t i = r0 + 2 * t;
        const bool act = t < t_hi, full = act && i + 1 < n;
        double ra = 0.0, rb = 0.0;
        int32_t a0 = -1, a1 = -1;
        if (full) {
            const double2 zz = ld2(z + i), pp = ld2(p + i), ww = ld2(w + i), ss = ld2(s + i), rr = ld2(r + i);
            const double2 xx = __ldcs(reinterpret_cast<const double2 *>(x + i));   // x is touched once per iteration
            const int2 aa = *reinterpret_cast<const int2 *>(agg + i);
            const double pa = zz.x + beta * pp.x, pb = zz.y + beta * pp.y;
            const double sa = ww.x + beta * ss.x, sb = ww.y + beta * ss.y;
            ra = rr.x - alpha * sa;
            rb = rr.y - alpha * sb;
            st2(p + i, pa, pb);
            st2(s + i, sa, sb);
            __stcs(reinterpret_cast<double2 *>(x + i), make_double2(xx.x + alpha * pa, xx.y + alpha * pb));
            st2(r + i, ra, rb);
            a0 = aa.x;
            a1 = aa.y;
        } else if (act) {
            const double pi = z[i] + beta * p[i];
            const double si = w[i] + beta * s[i];
            ra = r[i] - alpha * si;
            p[i] = pi;
            s[i] = si;
            x[i] += alpha * pi;
            r[i] = ra;
            a0 = agg[i];
        }
        const int q0 = max(min(__double2int_rn(ra * fix), 33554431), -33554431);
        const int q1 = max(min(__double2int_rn(rb * fix), 33554431), -33554431);
        {
            int u0 = q0, u1 = q1;
            int32_t b1 = a1;
            if (b1 == a0) { u0 += u1; b1 = -1; }
            if (a0 >= 0 && u0 != 0) { atomicAdd(&tab[0][a0], u0 & 0x3fff); atomicAdd(&tab[1][a0], u0 >> 14); }
            if (b1 >= 0 && u1 != 0) { atomicAdd(&tab[0][b1], u1 & 0x3fff); atomicAdd(&tab[1][b1], u1 >> 14); }
        }
        if (LOCAL) {   // the same entries into the rank-local level's table
            int32_t b0 = -1, b1 = -1;
            if (full) {
                const int2 bb = *reinterpret_cast<const int2 *>(agg2 + i);
                b0 = bb.x;
                b1 = bb.y;
            } else if (act) {
                b0 = agg2[i];
            }
            int u0 = q0, u1 = q1;
            if (b1 == b0) { u0 += u1; b1 = -1; }
            if (b0 >= 0 && u0 != 0) { atomicAdd(&tab2[0][b0], u0 & 0x3fff); atomicAdd(&tab2[1][b0], u0 >> 14); }
            if (b1 >= 0 && u1 != 0) { atomicAdd(&tab2[0][b1], u1 & 0x3fff); atomicAdd(&tab2[1][b1], u1 >> 14); }
        }
    }
    __syncthreads();
    for (int a = threadIdx.x; a < nc; a += TLB) {
        const long long v = ((long long)tab[1][a] << 14) + (long long)tab[0][a];
        if (v != 0) atomicAdd(&rc[a], (unsigned long long)v);
    }
    for (int a = threadIdx.x; LOCAL && a < nc2; a += TLB) {
        const long long v = ((long long)tab2[1][a] << 14) + (long long)tab2[0][a];
        if (v != 0) atomicAdd(&rc2[a], (unsigned long long)v);
    }
    if (Lp == nullptr) return;
    // several GPUs: the last CTA to finish sums r_c over the ranks (peer-memory stores, pdn_peer_prim.cuh)
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    if (threadIdx.x == 0) *ticket = 0u;
    __threadfence();
    peer_sum_i64_cta(*Lp, reinterpret_cast<long long *>(rc), nc, arange);
}

// start / restart from the residual r:  z = dinv r (or its restriction, TL);  p = s = 0
template <bool TL>
__global__ void __launch_bounds__(RPB) cg_init_kernel(int64_t r0, int64_t n, const Scal *__restrict__ sc,
                                                      const double *__restrict__ dinv,
                                                      const double *__restrict__ r, double *__restrict__ z,
                                                      double *__restrict__ p, double *__restrict__ s,
                                                      const int32_t *__restrict__ agg, long long *rc_fix, int32_t nc) {
    const double fix = TL ? CO_RFIX32 / sqrt(fmax(sc->rr, 1e-300)) : 0.0;
    unsigned long long *rc = TL ? reinterpret_cast<unsigned long long *>(rc_fix) + (size_t)(sc->iter & 1) * nc : nullptr;
    for (int64_t i0 = r0 + (int64_t)blockIdx.x * RPB; i0 < n; i0 += (int64_t)gridDim.x * RPB) {
        const int64_t i = i0 + threadIdx.x;
        const bool act = i < n;
        double ri = 0.0;
        if (act) {
            ri = r[i];
            if (!TL) z[i] = dinv[i] * ri;
            p[i] = 0.0;
            s[i] = 0.0;
        }
        if (TL) {
            if (PDN_RESTRICT_WIDE) restrict_add(ri, act ? agg[i] : -1, fix, rc);
            else restrict_add_narrow(ri, act ? agg[i] : -1, fix, rc);
        }
    }
}

// the vectors BiCGSTAB restricts (s, p) are not bounded by ||r||: the fixed-point scale leaves 2^10 more
// headroom than in PCG (resolution 2^-30 ||r|| per entry)
constexpr double BICG_BOOST = 1024.0;

// BiCGSTAB step 2:  s = r - alpha v;  sh = dinv s   (TL: sh comes from coarse_z_kernel, s is restricted)
template <bool TL>
__global__ void __launch_bounds__(RPB) bicg_s_kernel(int64_t r0, int64_t n, const Scal *__restrict__ sc,
                                                     const double *__restrict__ dinv,
                                                     const double *__restrict__ r, const double *__restrict__ v,
                                                     double *__restrict__ s, double *__restrict__ sh,
                                                     const int32_t *__restrict__ agg, long long *rc_fix, int32_t nc) {
    if (sc->done) return;
    const double alpha = sc->alpha;
    const double fix = TL ? CO_RFIX / (BICG_BOOST * sqrt(fmax(sc->rr, 1e-300))) : 0.0;
    unsigned long long *rc = reinterpret_cast<unsigned long long *>(rc_fix);   // slot 0
    for (int64_t i0 = r0 + (int64_t)blockIdx.x * RPB; i0 < n; i0 += (int64_t)gridDim.x * RPB) {
        const int64_t i = i0 + threadIdx.x;
        const bool act = i < n;
        double si = 0.0;
        if (act) {
            si = r[i] - alpha * v[i];
            s[i] = si;
            if (!TL) sh[i] = dinv[i] * si;
        }
        if (TL) restrict_add(si, act ? agg[i] : -1, fix, rc);
    }
}

// BiCGSTAB step 4:  x += alpha ph + omega sh;  r = s - omega t;  rho_new = rhat.r;  rr = r.r
template <bool DIST>
__global__ void __launch_bounds__(RPB) bicg_x_kernel(int64_t r0, int64_t n, Scal *sc, const double *__restrict__ ph,
                                                     const double *__restrict__ sh, const double *__restrict__ s,
                                                     const double *__restrict__ t, const double *__restrict__ rhat,
                                                     double *__restrict__ x, double *__restrict__ r,
                                                     double *partials) {
    __shared__ double red[3][RPB / 32];
    if (sc->done) return;
    const double alpha = sc->alpha, omega = sc->omega;
    double acc[3] = {0.0, 0.0, 0.0};
    for (int64_t i = r0 + (int64_t)blockIdx.x * RPB + threadIdx.x; i < n; i += (int64_t)gridDim.x * RPB) {
        x[i] += alpha * ph[i] + omega * sh[i];
        const double ri = s[i] - omega * t[i];
        r[i] = ri;
        acc[0] += rhat[i] * ri;
        acc[2] += ri * ri;
    }
    if (!grid_reduce<3>(acc, partials, &sc->ticket[1], red)) return;
    if (threadIdx.x != 0) return;
    if (DIST) { sc->part[0] = acc[0]; sc->part[1] = acc[1]; sc->part[2] = acc[2]; }
    else bicg_x_finalize(acc, sc);
}

// BiCGSTAB step 5:  p = r + beta (p - omega v);  ph = dinv p   (TL: p is restricted into slot 1)
template <bool TL>
__global__ void __launch_bounds__(RPB) bicg_p_kernel(int64_t r0, int64_t n, const Scal *__restrict__ sc,
                                                     const double *__restrict__ dinv,
                                                     const double *__restrict__ r, const double *__restrict__ v,
                                                     double *__restrict__ p, double *__restrict__ ph,
                                                     const int32_t *__restrict__ agg, long long *rc_fix, int32_t nc) {
    if (sc->done) return;
    const double beta = sc->beta, omega = sc->omega;
    const double fix = TL ? CO_RFIX / (BICG_BOOST * sqrt(fmax(sc->rr, 1e-300))) : 0.0;
    unsigned long long *rc = reinterpret_cast<unsigned long long *>(rc_fix) + (TL ? nc : 0);   // slot 1
    for (int64_t i0 = r0 + (int64_t)blockIdx.x * RPB; i0 < n; i0 += (int64_t)gridDim.x * RPB) {
        const int64_t i = i0 + threadIdx.x;
        const bool act = i < n;
        double pi = 0.0;
        if (act) {
            pi = r[i] + beta * (p[i] - omega * v[i]);
            p[i] = pi;
            if (!TL) ph[i] = dinv[i] * pi;
        }
        if (TL) restrict_add(pi, act ? agg[i] : -1, fix, rc);
    }
}

// BiCGSTAB start from residual r:  rhat = r; p = r; ph = dinv p; rho = r.r; rr = r.r
// (TL: p is restricted into slot 1 with the scale of the CURRENT sc->rr, which the host set to ||r||^2)
template <bool DIST, bool TL>
__global__ void __launch_bounds__(RPB) bicg_init_kernel(int64_t r0, int64_t n, Scal *sc, const double *__restrict__ dinv,
                                                        const double *__restrict__ r, double *__restrict__ rhat,
                                                        double *__restrict__ p, double *__restrict__ ph,
                                                        double *partials, const int32_t *__restrict__ agg,
                                                        long long *rc_fix, int32_t nc) {
    __shared__ double red[3][RPB / 32];
    double acc[3] = {0.0, 0.0, 0.0};
    const double fix = TL ? CO_RFIX / (BICG_BOOST * sqrt(fmax(sc->rr, 1e-300))) : 0.0;
    unsigned long long *rc = reinterpret_cast<unsigned long long *>(rc_fix) + (TL ? nc : 0);   // slot 1
    for (int64_t i0 = r0 + (int64_t)blockIdx.x * RPB; i0 < n; i0 += (int64_t)gridDim.x * RPB) {
        const int64_t i = i0 + threadIdx.x;
        const bool act = i < n;
        double ri = 0.0;
        if (act) {
            ri = r[i];
            rhat[i] = ri;
            p[i] = ri;
            if (!TL) ph[i] = dinv[i] * ri;
            acc[0] += ri * ri;
        }
        if (TL) restrict_add(ri, act ? agg[i] : -1, fix, rc);
    }
    if (!grid_reduce<3>(acc, partials, &sc->ticket[1], red)) return;
    if (threadIdx.x != 0) return;
    if (DIST) { sc->part[0] = acc[0]; sc->part[1] = acc[1]; sc->part[2] = acc[2]; }
    else bicg_init_finalize(acc, sc);
}

// sum of squares (for ||b||, ||L v - r||)
template <bool DIST>
__global__ void __launch_bounds__(RPB) sumsq_kernel(int64_t r0, int64_t n, const double *__restrict__ a,
                                                    const double *__restrict__ b, Scal *sc, double *out,
                                                    double *partials) {
    // out = sum (a_i - b_i)^2   (b may be null)
    __shared__ double red[3][RPB / 32];
    double acc[3] = {0.0, 0.0, 0.0};
    for (int64_t i = r0 + (int64_t)blockIdx.x * RPB + threadIdx.x; i < n; i += (int64_t)gridDim.x * RPB) {
        const double d = b ? a[i] - b[i] : a[i];
        acc[0] += d * d;
    }
    if (!grid_reduce<3>(acc, partials, &sc->ticket[2], red)) return;
    if (threadIdx.x != 0) return;
    if (DIST) sc->part[0] = acc[0];
    else *out = acc[0];
}

#include "pdn_peer.cuh"

}  // namespace pdnk
