// pdn_cg2_tail.cuh -- what the SpMV kernels of the round-2 PCG iteration share (see pdn_cg2.cuh for the iteration):
// the restriction c_w = P^T w in fixed point, and the tail the LAST CTA of K1 runs -- sum of the partials (several
// GPUs: one peer-memory exchange of dots + c_w), Chronopoulos-Gear scalars, c_s = c_w + beta c_s, r_c -= alpha c_s.
// Included by pdn_kernels.cuh (namespace pdnk) after pdn_spmv.cuh and before the SpMV kernels.
#pragma once

constexpr int TAB_STRIDE = 8;          // words between two accumulators of the restriction table (64 B)
// Fixed-point resolution of the restrictions.  r_c lives by recurrence, so a quantisation error PERSISTS in it
// until the next re-restriction instead of being redrawn every iteration, and r_c itself is a small, cancellation-
// prone quantity (the coarse correction removes the aggregate means of r).  Measured on the CPU model
// (scripts/proto_v2_pcg.py, 16-50 rows per aggregate): 2^25 steps per bound -> no convergence, 2^32 -> drift 0.1,
// 2^36 -> 7e-3, 2^40 -> 4e-4.  So: 2^38 steps, reduced over the warp as two 19-bit limbs (two redux.sync).
constexpr double Q_STEPS = 274877906944.0;     // 2^38
constexpr long long Q_LIM = 1LL << 41;         // clamp: 8x the scale; 2^22 rows per aggregate cannot overflow int64
constexpr double W_BOUND = 8.0;        // headroom over the diagonal-dominance bound 2 a_ii max|z|

// what K1 needs beyond the operator: aggregates, restriction tables, the coarse vectors kept by recurrence
struct Cg2 {
    const int32_t *agg;               // [n] global-level aggregate of every row (-1: none)
    const int32_t *agg2;              // rank-local level (null without one)
    unsigned long long *tab, *tab2;   // [nc * TAB_STRIDE] fixed-point accumulators of c_w (zero between launches)
    double *rc, *cs;                  // [2][nc] coarse residual P^T r and P^T s, two slots (iteration parity):
                                      // K2 reads slot (iter & 1) ^ 1 and writes slot iter & 1 (pdn_cg2.cuh)
    double *rc2, *cs2;                // the same for the rank-local level
    double *cw;                       // [nc] several GPUs: c_w summed over the ranks by K1's tail (else unused)
    int32_t nc, nc2;
    // per-CTA shared-memory windows of the restriction (sell_tma_kernel<CG2>): first aggregate of the window of
    // every CTA, per level (setup: agg_min_kernel with the SpMV's own grid and slice split)
    const int32_t *amin, *amin2;
    // several GPUs, halo fused into K3: the plan (recv_cnt) K1 waits on before it reads z; else null
    const HaloSend *halo;
    // several GPUs: [first, last] aggregate this rank's rows touch (agg_range_kernel; first > last: none).  Only that
    // range of c_w crosses NVLink: a rank that owns 2 of 16 layers touches 1/8 of the aggregates.
    const int32_t *arange;
};

__device__ __forceinline__ void red_add_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ long long quantise(double val, double fix) {
    const double q = val * fix;
    const double lim = (double)Q_LIM;
    return __double2ll_rn(q > lim ? lim : (q < -lim ? -lim : q));      // NaN -> 0 (cvt.rni.s64.f64 of NaN is 0)
}
// Per-CTA restriction table in shared memory.  A CTA's contiguous rows touch a short, contiguous range of
// aggregates (bins are numbered mesh by mesh, row by row), so it accumulates c_w for the TABW aggregates from
// amin[cta] on in shared memory and flushes the non-zero ones with one global atomic each at the end; an aggregate
// outside the window (unstructured numbering) goes straight to the global table.  64-bit shared atomics are a CAS
// loop on sm_100a (ATOMS.CAST.SPIN.64) and a 32-row slice holds 3.5 distinct aggregates on these meshes (rows of
// valence 4 and 8 alternate and the length sort separates them), so neither 64-bit atomics nor warp collectives
// per group pay; the 42-bit value goes in as THREE 14-bit limbs with native 32-bit ATOMS.ADD -- no collective, no
// loop, the bank-conflict hardware serialises equal addresses.  2^17 entries per CTA and aggregate cannot overflow.
constexpr int CO_MAX_TAIL = PDN_CO_MAX;   // pdn_common.cuh: largest coarse space
constexpr int TABW = 512;
constexpr size_t TAB_SMEM = (size_t)2 * 3 * TABW * sizeof(int);      // two levels

__device__ __forceinline__ void restrict_smem(double val, int32_t a, double fix, int32_t amin, int *t /*[3][TABW]*/,
                                              unsigned long long *gtab) {
    if (a < 0) return;
    const long long q = quantise(val, fix);
    if (q == 0) return;
    const unsigned d = (unsigned)(a - amin);
    if (d < (unsigned)TABW) {
        atomicAdd(&t[d], (int)(q & 0x3fff));
        atomicAdd(&t[TABW + d], (int)((q >> 14) & 0x3fff));
        atomicAdd(&t[2 * TABW + d], (int)(q >> 28));
    } else {
        red_add_u64(gtab + (size_t)a * TAB_STRIDE, (unsigned long long)q);
    }
}
// all threads of the CTA, after a barrier: non-zero window entries -> the global table
__device__ __forceinline__ void flush_smem_table(const int *t, int32_t amin, int32_t nc, unsigned long long *gtab) {
    for (int d = threadIdx.x; d < TABW && amin + d < nc; d += blockDim.x) {
        const long long sum = ((long long)t[2 * TABW + d] << 28) + ((long long)t[TABW + d] << 14) + (long long)t[d];
        if (sum != 0) red_add_u64(gtab + (size_t)(amin + d) * TAB_STRIDE, (unsigned long long)sum);
    }
}

// tab[a] += val over the lanes of a warp (all 32 lanes call, converged; a < 0 contributes nothing).  One pass per
// DISTINCT aggregate in the warp -- 32 consecutive rows of a mesh sit in one or two bins -- each pass: the leader's
// aggregate by shuffle, its members by ballot, the 42-bit fixed-point values of the members summed with two
// redux.sync over 21-bit limbs, one 64-bit integer atomic.  (match.any + masked redux, the round-1 form, showed up
// as 10 % of the SpMV's stall samples: WARPSYNC.COLLECTIVE.)
template <int STRIDE = TAB_STRIDE>
__device__ __forceinline__ void restrict_q(double val, int32_t a, double fix, unsigned long long *tab) {
    const long long q = quantise(val, fix);
    const int lo = (int)(q & 0x1fffff), hi = (int)(q >> 21);          // q = hi * 2^21 + lo, 0 <= lo < 2^21, |hi| <= 2^20
    const int lane = (int)(threadIdx.x & 31);
    unsigned todo = __ballot_sync(0xffffffffu, a >= 0);
    while (todo) {                                                    // warp-uniform
        const int leader = __ffs(todo) - 1;
        const int32_t al = __shfl_sync(0xffffffffu, a, leader);
        const bool mine = a == al;
        const unsigned grp = __ballot_sync(0xffffffffu, mine);
        const int slo = __reduce_add_sync(0xffffffffu, mine ? lo : 0), shi = __reduce_add_sync(0xffffffffu, mine ? hi : 0);
        if (lane == leader) {
            const long long sum = ((long long)shi << 21) + (long long)slo;
            if (sum != 0) red_add_u64(tab + (size_t)al * STRIDE, (unsigned long long)sum);
        }
        todo &= ~grp;
    }
}

__device__ __forceinline__ double restrict_fix(const Scal *sc) {
    const double zmax = __longlong_as_double((long long)sc->zmax_bits[sc->iter & 1]);
    const double bound = W_BOUND * sc->amax * zmax;
    return bound > 0.0 && bound < 1e300 ? Q_STEPS / bound : 0.0;
}

// ---- the tail of K1: all threads of the LAST CTA call; thread 0 holds the grid totals of the three dots -------
template <bool DIST>
__device__ __forceinline__ void cg2_tail(double (&acc)[3], Scal *sc, const Cg2 &cg, const PeerLink *Lp, double fix) {
    __shared__ double sh[4];
    __shared__ int sh_done, sh_ok;
    const int tid = threadIdx.x, nt = blockDim.x;
    const double invfix = fix > 0.0 ? 1.0 / fix : 0.0;
    const volatile long long *src = nullptr;       // several GPUs: the slots of all ranks in the local arena
    if (DIST && Lp != nullptr) {
        // ONE exchange: every rank stores [gamma, delta, rr, 1/fix, q[0..nc)] into its slot on every rank
        const PeerLink &L = *Lp;
        const unsigned long long e = L.epoch[7] + 1;
        const size_t slot = (size_t)(e & 1) * PEER_MAX;
        if (tid == 0) { peer_log(L, 0); sh[0] = acc[0]; sh[1] = acc[1]; sh[2] = acc[2]; sh[3] = invfix; sh_ok = 1; }
        __syncthreads();
        // slot layout: 8 header words [gamma, delta, rr, 1/fix, first aggregate, count, -, -], then q[first .. first + count)
        const int32_t alo = __ldg(cg.arange), ahi = __ldg(cg.arange + 1);
        const int32_t acnt = ahi >= alo ? ahi - alo + 1 : 0;
        {   // this rank's range of the table once into registers, then one store per peer (remote stores are fire-and-forget)
            constexpr int TU = (CO_MAX_TAIL + 383) / 384;
            long long q[TU];
#pragma unroll
            for (int u = 0; u < TU; u++) {
                const int j = tid + u * nt;
                q[u] = j < acnt ? (long long)__ldcg(cg.tab + (size_t)(alo + j) * TAB_STRIDE) : 0;
            }
            for (int k = 0; k < L.world; k++) {
                long long *dst = reinterpret_cast<long long *>(L.arena[k] + PA_CW_VAL) + (slot + L.me) * PA_RC_SLOT;
                if (tid < 4) reinterpret_cast<double *>(dst)[tid] = sh[tid];
                if (tid == 4) dst[4] = alo;
                if (tid == 5) dst[5] = acnt;
#pragma unroll
                for (int u = 0; u < TU; u++)
                    if (tid + u * nt < acnt) dst[8 + tid + u * nt] = q[u];
            }
        }
        __threadfence_system();
        __syncthreads();
        if (tid < L.world) st_release_sys(L.arena[tid] + PA_CW_FLAG + slot + L.me, e);
        if (tid < L.world && !peer_wait(L.arena[L.me] + PA_CW_FLAG + slot + tid, e, L.status, L.epoch + 5)) sh_ok = 0;
        __syncthreads();
        src = reinterpret_cast<const volatile long long *>(L.arena[L.me] + PA_CW_VAL) + slot * PA_RC_SLOT;
        {   // c_w summed over the ranks in rank order, for K2.  The acquire of the waiting threads and the barrier order
            // these loads after the peers' stores; ld.cg reads L2, where the remote stores land (no L1 copy can be
            // stale).  All loads of a thread are independent: as volatile loads in a loop this took ~10 us on 2 GPUs.
            const long long *sl = const_cast<const long long *>(src);
            double inv[PEER_MAX];
            int lo[PEER_MAX], cnt[PEER_MAX];
#pragma unroll
            for (int r = 0; r < PEER_MAX; r++) {
                const long long *h = sl + (size_t)r * PA_RC_SLOT;
                inv[r] = r < L.world ? __ldcg(reinterpret_cast<const double *>(h) + 3) : 0.0;
                lo[r] = r < L.world ? (int)__ldcg(h + 4) : 0;
                cnt[r] = r < L.world ? (int)__ldcg(h + 5) : 0;
            }
#pragma unroll 1
            for (int j = tid; j < cg.nc; j += 2 * nt) {          // two entries x PEER_MAX independent loads in flight
                const int j2 = j + nt;
                long long va[PEER_MAX], vb[PEER_MAX];
#pragma unroll
                for (int r = 0; r < PEER_MAX; r++) {
                    const unsigned da = (unsigned)(j - lo[r]), db = (unsigned)(j2 - lo[r]);
                    va[r] = da < (unsigned)cnt[r] ? __ldcg(sl + (size_t)r * PA_RC_SLOT + 8 + da) : 0;
                    vb[r] = (db < (unsigned)cnt[r] && j2 < cg.nc) ? __ldcg(sl + (size_t)r * PA_RC_SLOT + 8 + db) : 0;
                }
                double ca = 0.0, cb = 0.0;
#pragma unroll
                for (int r = 0; r < PEER_MAX; r++) { ca += (double)va[r] * inv[r]; cb += (double)vb[r] * inv[r]; }
                cg.cw[j] = ca;
                if (j2 < cg.nc) cg.cw[j2] = cb;
            }
        }
        if (tid == 0) {
            L.epoch[7] = e;
            peer_log(L, 1);
            acc[0] = acc[1] = acc[2] = 0.0;
            for (int r = 0; r < L.world; r++) {       // rank order: identical bits on every rank
                const volatile double *h = reinterpret_cast<const volatile double *>(src + (size_t)r * PA_RC_SLOT);
                acc[0] += h[0]; acc[1] += h[1]; acc[2] += h[2];
            }
        }
    }
    if (tid == 0) {
        if (DIST && Lp != nullptr && !sh_ok) { sc->done = 3; sc->breakdown = 1; }
        else spmv_finalize<CG>(acc, sc);
        sh[0] = sc->alpha;
        sh[1] = sc->beta;
        sh_done = sc->done;
        // the slot the coming K3 maximises into (iter has advanced).  Not when `done` latched: iter did not
        // advance then, the slot is the one THIS kernel read, and the host may resume the very same iteration.
        if (!sc->done) sc->zmax_bits[sc->iter & 1] = 0ull;
    }
    __syncthreads();
    if (sh_done) return;                        // identical on every rank; the host clears the tables before a restart
    // The recurrence c_s = c_w + beta c_s, r_c -= alpha c_s itself runs in K2 (coarse_apply2_kernel), every CTA
    // redundantly and in parallel: as a loop of this ONE CTA it made the tail 3-6 us long (measured: SpMV 33.4 us
    // with it, 30.5 us without, 27.3 us with the plain dots tail).  Here only what K2 cannot know: 1/fix of this
    // launch and, on several GPUs, c_w summed over the ranks in rank order.
    if (tid == 0) sc->cw_invfix = invfix;
}

