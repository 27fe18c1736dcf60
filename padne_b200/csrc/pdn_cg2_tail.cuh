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
    double *rc, *cs;                  // [nc] coarse residual P^T r and P^T s
    double *rc2, *cs2;                // the same for the rank-local level
    int32_t nc, nc2;
};

__device__ __forceinline__ void red_add_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ long long quantise(double val, double fix) {
    const double q = val * fix;
    const double lim = (double)Q_LIM;
    return __double2ll_rn(q > lim ? lim : (q < -lim ? -lim : q));      // NaN -> 0 (cvt.rni.s64.f64 of NaN is 0)
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
        for (int k = 0; k < L.world; k++) {
            long long *dst = reinterpret_cast<long long *>(L.arena[k] + PA_CW_VAL) + (slot + L.me) * PA_RC_SLOT;
            if (tid < 4) reinterpret_cast<double *>(dst)[tid] = sh[tid];
            for (int j = tid; j < cg.nc; j += nt) dst[4 + j] = (long long)__ldcg(cg.tab + (size_t)j * TAB_STRIDE);
        }
        __threadfence_system();
        __syncthreads();
        if (tid < L.world) st_release_sys(L.arena[tid] + PA_CW_FLAG + slot + L.me, e);
        if (tid < L.world && !peer_wait(L.arena[L.me] + PA_CW_FLAG + slot + tid, e, L.status, L.epoch + 5)) sh_ok = 0;
        __syncthreads();
        src = reinterpret_cast<const volatile long long *>(L.arena[L.me] + PA_CW_VAL) + slot * PA_RC_SLOT;
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
    const double alpha = sh[0], beta = sh[1];
    for (int j = tid; j < cg.nc; j += nt) {
        double cw;
        if (DIST && Lp != nullptr) {
            cw = 0.0;
            for (int r = 0; r < Lp->world; r++) {
                const volatile long long *s = src + (size_t)r * PA_RC_SLOT;
                cw += (double)s[4 + j] * reinterpret_cast<const volatile double *>(s)[3];
            }
        } else {
            cw = (double)(long long)__ldcg(cg.tab + (size_t)j * TAB_STRIDE) * invfix;
        }
        cg.tab[(size_t)j * TAB_STRIDE] = 0ull;
        const double c = cw + beta * cg.cs[j];
        cg.cs[j] = c;
        cg.rc[j] -= alpha * c;
    }
    for (int j = tid; j < cg.nc2; j += nt) {    // rank-local level: block diagonal, nothing to exchange
        const double cw = (double)(long long)__ldcg(cg.tab2 + (size_t)j * TAB_STRIDE) * invfix;
        cg.tab2[(size_t)j * TAB_STRIDE] = 0ull;
        const double c = cw + beta * cg.cs2[j];
        cg.cs2[j] = c;
        cg.rc2[j] -= alpha * c;
    }
}

