// pdn_sell.cuh -- kernel (3a'): the Krylov loop's SpMV on a SELL-32-sigma copy of the reduced operator.
//
// Layout (built once per solve by the setup kernels at the bottom of this file):
//   * rows are sorted by stored length (descending, stable) inside windows of SELL_WIN rows and the
//     whole reduced system is permuted symmetrically with that order, so every Krylov vector lives
//     in the permuted numbering and all per-row accesses stay unit-stride;
//   * 32 consecutive rows form a slice, stored column-major and padded to the slice's longest row
//     (after the sort the padding is ~0: mesh rows of equal length sit together);
//   * lane l of a warp owns row 32*s + l: every load of the value / column arrays is one fully
//     coalesced 256 B / 128 B request, x is gathered through L1/L2 (mesh-local numbering keeps the
//     neighbours of a window inside a few KB), one DFMA per stored entry, no shared memory and no
//     barrier in the inner loop, full occupancy hides the gather latency.
// HBM traffic per row: 12 B per stored entry + 1/8 B of slice metadata + the vectors -- the
// algorithmic figure of SURVEY.md 8d minus the 4 B row pointer.
//
// Work is split statically (contiguous slice ranges per CTA, CTA count = SMs x resident CTAs) so the
// fused dot products are summed in a fixed order: results are bit-identical run to run.
//
// Included by pdn_kernels.cuh (namespace pdnk).
#pragma once

constexpr int SELL_C = 32;        // rows per slice = warp width
constexpr int SELL_WIN = 1024;    // sorting window (rows); slices never straddle a window
#ifndef PDN_SELL_BLK
#define PDN_SELL_BLK 256
#endif
constexpr int SELL_BLK = PDN_SELL_BLK;     // threads per CTA of the SpMV
constexpr int SELL_WARPS = SELL_BLK / 32;
constexpr int SELL_MAXLEN = 1 << 20;
#ifndef PDN_SELL_U
#define PDN_SELL_U 5
#endif
#ifndef PDN_SELL_PF
#define PDN_SELL_PF 0
#endif
#ifndef PDN_SELL_MINB
#define PDN_SELL_MINB 4
#endif
constexpr int SELL_U = PDN_SELL_U;     // entries per row fetched per step
constexpr int SELL_PF = PDN_SELL_PF;   // L2 prefetch distance in entries (12 B each) per CTA; 0 = off

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// matrix stream: read-only, do not allocate in L1 (L1 is for the gathered x)
#ifndef PDN_SELL_NOALLOC
#define PDN_SELL_NOALLOC 1
#endif
__device__ __forceinline__ double ld_stream(const double *p) {
#if PDN_SELL_NOALLOC
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}
__device__ __forceinline__ int32_t ld_stream(const int32_t *p) {
#if PDN_SELL_NOALLOC
    int32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}
#ifndef PDN_SELL_FAKEX
#define PDN_SELL_FAKEX 0    // timing experiment only: 1 replaces the gather x[col] by x[row] (wrong results)
#endif

struct SellOp {
    int64_t n;                 // rows of the reduced system (permuted numbering)
    int64_t row_begin, row_end;// rows this rank owns (multiples of SELL_WIN except the global end)
    const int32_t *soff;       // [n_slices+1] entry offset of every slice (multiple of 32)
    const uint32_t *smask;     // [n_slices] bit l: row 32*s+l is an override row (skipped here)
    const int32_t *ci;         // column indices (permuted numbering), column-major inside a slice
    const double *va;
    int32_t n_special;         // override rows (constraint terminals), handled by CTA 0 of the owner
    const int32_t *sp_row;     // [n_special] permuted row ids, ascending
    const int32_t *sp_rp;      // [n_special+1]
    const int32_t *sp_ci;      // permuted column ids
    const double *sp_va;
};

// DIST = false: the last CTA to finish runs the scalar recurrence (single GPU).
// DIST = true : the last CTA stores this rank's three partial sums to sc->part[] and the recurrence
//               runs in scalar_kernel after the all-reduce (pdn_multi.cuh).
// What the last CTA does with the grid totals of the fused dot products (thread 0 holds acc):
//   one GPU:                 the scalar recurrence (spmv_finalize)
//   several, peer memory:    warp 0 all-reduces them over the ranks in place (pdn_peer_prim.cuh), then the recurrence
//   several, NCCL fallback:  parked in sc->part for the all-reduce + scalar_kernel that follow
template <int MODE, bool DIST>
__device__ __forceinline__ void dots_tail(double (&acc)[3], Scal *sc, const PeerLink *Lp) {
    if (DIST && Lp != nullptr) {
        if (threadIdx.x >= 32) return;
        const bool ok = peer_allreduce3(*Lp, acc);
        if (threadIdx.x != 0) return;
        if (!ok) { sc->done = 3; sc->breakdown = 1; return; }
        sc->part[0] = acc[0]; sc->part[1] = acc[1]; sc->part[2] = acc[2];
        spmv_finalize<MODE>(acc, sc);
        return;
    }
    if (threadIdx.x != 0) return;
    if (DIST) {
        sc->part[0] = acc[0]; sc->part[1] = acc[1]; sc->part[2] = acc[2];
    } else {
        spmv_finalize<MODE>(acc, sc);
    }
}

// Override rows (constraint terminals): a terminal row sums the rows of every vertex its pad covers and can
// hold thousands of entries, so one WARP per row (lanes stride over the entries, fixed shuffle tree), the rows
// dealt round-robin to all warps of the grid.  (One thread per row in CTA 0 left a 30-80 us tail on the rank
// that owns the terminals.)
template <int MODE>
__device__ __forceinline__ void special_rows(const SellOp &A, const double *__restrict__ x, double *__restrict__ y,
                                             const double *__restrict__ aux0, double (&acc)[3], int32_t gwarp,
                                             int32_t n_gwarp) {
    const int lane = threadIdx.x & 31;
    for (int32_t t = gwarp; t < A.n_special; t += n_gwarp) {
        const int32_t row = A.sp_row[t];
        if (row < A.row_begin || row >= A.row_end) continue;
        double s2 = 0.0;
        for (int32_t q = A.sp_rp[t] + lane; q < A.sp_rp[t + 1]; q += 32) s2 = fma(A.sp_va[q], x[A.sp_ci[q]], s2);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        if (lane == 0) row_epilogue<MODE>(row, s2, MODE == CG ? x[row] : 0.0, MODE != PLAIN ? aux0[row] : 0.0, y, acc);
    }
}

// override rows (constraint terminals) for CG2: as special_rows, plus the restriction of their w entries
__device__ __forceinline__ void special_rows_cg2(const SellOp &A, const double *__restrict__ x, double *__restrict__ y,
                                                 const double *__restrict__ aux0, double (&acc)[3], int32_t gwarp,
                                                 int32_t n_gwarp, const Cg2 &cg, double fix) {
    const int lane = threadIdx.x & 31;
    for (int32_t t = gwarp; t < A.n_special; t += n_gwarp) {
        const int32_t row = A.sp_row[t];
        if (row < A.row_begin || row >= A.row_end) continue;
        double s2 = 0.0;
        for (int32_t q = A.sp_rp[t] + lane; q < A.sp_rp[t + 1]; q += 32) s2 = fma(A.sp_va[q], x[A.sp_ci[q]], s2);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        if (lane == 0) {
            row_epilogue<CG>(row, s2, x[row], aux0[row], y, acc);
            const long long q = quantise(s2, fix);
            const int32_t a = cg.agg[row];
            if (a >= 0 && q != 0) red_add_u64(cg.tab + (size_t)a * TAB_STRIDE, (unsigned long long)(long long)q);
            if (cg.nc2 > 0) {
                const int32_t a2 = cg.agg2[row];
                if (a2 >= 0 && q != 0) red_add_u64(cg.tab2 + (size_t)a2 * TAB_STRIDE, (unsigned long long)(long long)q);
            }
        }
    }
}

template <int MODE, bool DIST>
__global__ void __launch_bounds__(SELL_BLK, PDN_SELL_MINB) sell_spmv_kernel(SellOp A, const double *__restrict__ x,
                                                               double *__restrict__ y,
                                                               const double *__restrict__ aux0, Scal *sc,
                                                               double *partials, const PeerLink *Lp) {
    __shared__ double red[3][SELL_WARPS];
    if (MODE != PLAIN && MODE != RESID) {
        if (sc->done) return;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t s_begin = A.row_begin / SELL_C, s_end = (A.row_end + SELL_C - 1) / SELL_C;
    const int64_t s_cnt = s_end - s_begin;
    const int64_t s_lo = s_begin + (int64_t)blockIdx.x * s_cnt / gridDim.x;
    const int64_t s_hi = s_begin + (int64_t)(blockIdx.x + 1) * s_cnt / gridDim.x;
    double acc[3] = {0.0, 0.0, 0.0};

    // The CTA's slices are one contiguous range [e_lo, e_hi) of the value / column arrays.  Every warp
    // asks L2 for the lines SELL_PF entries ahead of the slice it is working on (same byte count as
    // the slice, so the union of all requests is the CTA's range shifted by SELL_PF, each line once):
    // the DRAM latency is paid by the prefetch, the loads below see L2 latency.
    // (Measured on B200: any distance > 0 is SLOWER than none -- 10M DOF: 198 us without, 221 us at 1 Ki,
    // 259 us at 4 Ki, 302 us at 16 Ki entries -- so it is compiled out by default, SELL_PF = 0.)
    const int32_t e_lo = SELL_PF > 0 && s_lo < s_hi ? __ldg(&A.soff[s_lo]) : 0,
                  e_hi = SELL_PF > 0 && s_lo < s_hi ? __ldg(&A.soff[s_hi]) : 0;
    if (SELL_PF > 0) {
        const int32_t e_pf = e_lo + SELL_PF < e_hi ? e_lo + SELL_PF : e_hi;
        for (int32_t e = e_lo + threadIdx.x * 16; e < e_pf; e += SELL_BLK * 16) prefetch_l2(A.va + e);
        for (int32_t e = e_lo + threadIdx.x * 32; e < e_pf; e += SELL_BLK * 32) prefetch_l2(A.ci + e);
    }

    int64_t s = s_lo + warp;
    int32_t off = 0, off1 = 0;
    uint32_t mask = 0;
    if (s < s_hi) { off = __ldg(&A.soff[s]); off1 = __ldg(&A.soff[s + 1]); mask = __ldg(&A.smask[s]); }
    while (s < s_hi) {
        // metadata of this warp's next slice is fetched before the entries of the current one
        const int64_t sn = s + SELL_WARPS;
        int32_t n_off = 0, n_off1 = 0;
        uint32_t n_mask = 0;
        if (sn < s_hi) { n_off = __ldg(&A.soff[sn]); n_off1 = __ldg(&A.soff[sn + 1]); n_mask = __ldg(&A.smask[sn]); }
        if (SELL_PF > 0) {
            const int32_t p0 = off + SELL_PF, p1 = off1 + SELL_PF < e_hi ? off1 + SELL_PF : e_hi;
            for (int32_t e = p0 + lane * 16; e < p1; e += 32 * 16) prefetch_l2(A.va + e);
            for (int32_t e = p0 + lane * 32; e < p1; e += 32 * 32) prefetch_l2(A.ci + e);
        }

        const int64_t row = s * SELL_C + lane;
        const bool live = row < A.row_end && !((mask >> lane) & 1u);
        double xi = 0.0, ai = 0.0;
        if (live) {
            if (MODE == CG) xi = x[row];
            if (MODE != PLAIN) ai = aux0[row];
        }
        const int w = (off1 - off) >> 5;
        const int32_t *__restrict__ cp = A.ci + off + lane;
        const double *__restrict__ vp = A.va + off + lane;
        double sum = 0.0;
        // SELL_U entries per step, all loads of a step issued before the first use; w is
        // warp-uniform, so the guards of the last step do not diverge
        for (int k = 0; k < w; k += SELL_U) {
            int32_t c[SELL_U];
            double v[SELL_U], xx[SELL_U];
#pragma unroll
            for (int j = 0; j < SELL_U; j++) c[j] = k + j < w ? ld_stream(cp + (k + j) * SELL_C) : 0;
#pragma unroll
            for (int j = 0; j < SELL_U; j++) v[j] = k + j < w ? ld_stream(vp + (k + j) * SELL_C) : 0.0;
#pragma unroll
            for (int j = 0; j < SELL_U; j++) xx[j] = k + j < w ? __ldg(x + (PDN_SELL_FAKEX ? (int32_t)row : c[j])) : 0.0;
#pragma unroll
            for (int j = 0; j < SELL_U; j++) sum = fma(v[j], xx[j], sum);
        }
        if (live) row_epilogue<MODE>(row, sum, xi, ai, y, acc);
        s = sn; off = n_off; off1 = n_off1; mask = n_mask;
    }
    special_rows<MODE>(A, x, y, aux0, acc, (int32_t)((blockIdx.x * SELL_BLK + threadIdx.x) >> 5),
                       (int32_t)((gridDim.x * SELL_BLK) >> 5));
    if (MODE == PLAIN) return;
    if (!grid_reduce<3, SELL_WARPS>(acc, partials, &sc->ticket[0], red)) return;
    dots_tail<MODE, DIST>(acc, sc, Lp);
}

// ------------------------------------------------------------------------------------------------
// TMA-ring variant (the default in the Krylov loop).
//
// The direct kernel above is bound by memory-level parallelism: a warp has one batch of matrix loads
// in flight, then waits, then gathers (ncu: 45 % of DRAM peak, the stalls sit on the first use of the
// column indices and of the gathered x).  Here every warp owns a private ring of SELL_NSLOT slice
// buffers in shared memory.  A slice is ONE contiguous run of the value array and one of the column
// array, so lane 0 streams the slices the warp will need SELL_NSLOT iterations from now with two
// cp.async.bulk copies each (the TMA engine, UBLKCP in SASS; completion on a per-slot mbarrier) --
// no registers are tied up and SELL_NSLOT slices per warp are always in flight.  The warp then reads its
// column indices / values from shared memory (lane-contiguous: conflict-free), issues ALL gathers of
// its row at once and runs the DFMA chain.  Warps never synchronise with each other.
// Slices wider than SELL_WCAP (a node with many lumped elements) fall back to direct global loads.
// Measured on B200 (in-situ SpMV, 2layer_1M / 8layer_10M, us): direct kernel with L1-allocating matrix
// loads 33.4 / 202, direct with L1::no_allocate 28.6 / 194, this kernel 12 warps x 2 slots 29.3 / 185,
// 8 warps x 3 slots 32.8 / 199, 4 warps x 6 slots 43.6 / 311: a DEEPER ring is slower (so is an L2
// prefetch), more resident warps are faster -- the loop is bound by the x gathers per warp, not by the
// matrix stream.
// ------------------------------------------------------------------------------------------------
#ifndef PDN_SELL_NSLOT
#define PDN_SELL_NSLOT 2
#endif
#ifndef PDN_SELL_WCAP
#define PDN_SELL_WCAP 10
#endif
#ifndef PDN_SELL_TW
#define PDN_SELL_TW 12
#endif
#ifndef PDN_SELL_EVICT
#define PDN_SELL_EVICT 1
#endif
#ifndef PDN_SELL_TMINB
#define PDN_SELL_TMINB 2
#endif
constexpr int ST_NSLOT = PDN_SELL_NSLOT;
constexpr int ST_WCAP = PDN_SELL_WCAP;
constexpr int ST_WARPS = PDN_SELL_TW;
constexpr int ST_BLK = ST_WARPS * 32;
constexpr int ST_SLOT_ENT = ST_WCAP * SELL_C;
constexpr int ST_SLOT_BYTES = ST_SLOT_ENT * 12;
constexpr size_t ST_SMEM = (size_t)ST_WARPS * ST_NSLOT * ST_SLOT_BYTES + (size_t)ST_WARPS * ST_NSLOT * 8 + 3 * ST_WARPS * 8;

// MODE == CG2 (round-2 PCG iteration, pdn_cg2.cuh): CG plus the restriction of w into cg.tab / cg.tab2 and the
// recurrence tail (pdn_cg2_tail.cuh); the other modes ignore cg.
template <int MODE, bool DIST>
__global__ void __launch_bounds__(ST_BLK, PDN_SELL_TMINB) sell_tma_kernel(SellOp A, const double *__restrict__ x,
                                                                     double *__restrict__ y,
                                                                     const double *__restrict__ aux0, Scal *sc,
                                                                     double *partials, const PeerLink *Lp, Cg2 cg) {
    extern __shared__ __align__(128) unsigned char st_smem[];
    constexpr int EMODE = MODE == CG2 ? CG : MODE;
    pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // CG2: restriction windows behind the ring (pdn_cg2_tail.cuh); cleared before the programmatic-launch wait
    int *rtab = reinterpret_cast<int *>(st_smem + ST_SMEM);
    int32_t amin1 = 0, amin2 = 0;
    if (MODE == CG2) {
        for (int e = threadIdx.x; e < (int)(TAB_SMEM / sizeof(int)); e += ST_BLK) rtab[e] = 0;
        amin1 = __ldg(cg.amin + blockIdx.x);
        if (cg.nc2 > 0) amin2 = __ldg(cg.amin2 + blockIdx.x);
    }
    pdl_wait();
    if (MODE != PLAIN && MODE != RESID) {
        if (sc->done) return;
    }
    // several GPUs, halo pushed by the tail of the kernel that wrote x (coarse_z_kernel / cg_update2_kernel): wait
    // here for the peers' interface entries (consumer-side wait, pdn_peer_prim.cuh)
    const bool halo_wait = DIST && Lp != nullptr && cg.halo != nullptr;
    if (halo_wait) peer_halo_wait(*Lp, *cg.halo);
    if (MODE == CG2 || halo_wait) __syncthreads();      // restriction windows cleared, halo entries in place
    const double fix = MODE == CG2 ? restrict_fix(sc) : 0.0;
#if PDN_SELL_EVICT
    const uint64_t l2pol = tma::policy_evict_first();
#endif
    unsigned char *ring = st_smem + (size_t)warp * ST_NSLOT * ST_SLOT_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(st_smem + (size_t)ST_WARPS * ST_NSLOT * ST_SLOT_BYTES) + warp * ST_NSLOT;
    double(*red)[ST_WARPS] = reinterpret_cast<double(*)[ST_WARPS]>(
        st_smem + (size_t)ST_WARPS * ST_NSLOT * ST_SLOT_BYTES + (size_t)ST_WARPS * ST_NSLOT * 8);
    if (lane == 0) {
        for (int q = 0; q < ST_NSLOT; q++) tma::mbar_init(&bars[q], 1);
        tma::fence_barrier_init();
    }
    __syncwarp();

    const int64_t s_begin = A.row_begin / SELL_C, s_end = (A.row_end + SELL_C - 1) / SELL_C;
    const int64_t s_cnt = s_end - s_begin;
    const int64_t s_lo = s_begin + (int64_t)blockIdx.x * s_cnt / gridDim.x;
    const int64_t s_hi = s_begin + (int64_t)(blockIdx.x + 1) * s_cnt / gridDim.x;
    // this warp's slices: s_lo + warp + j * ST_WARPS, j in [0, nj)
    const int64_t nj = s_hi - s_lo > warp ? (s_hi - s_lo - warp + ST_WARPS - 1) / ST_WARPS : 0;
    double acc[3] = {0.0, 0.0, 0.0};

    // slice metadata arrives 32 slices at a time, one slice per lane (separate copies for the issue
    // and the consume position, which sit SELL_NSLOT slices apart)
    int32_t io0 = 0, io1 = 0, co0 = 0, co1 = 0;
    uint32_t cmask = 0;
    auto load_batch = [&](int64_t jb, int32_t &o0, int32_t &o1, uint32_t *m) {
        const int64_t j = jb + lane;
        o0 = o1 = 0;
        if (m) *m = 0;
        if (j < nj) {
            const int64_t sl = s_lo + warp + j * ST_WARPS;
            o0 = __ldg(&A.soff[sl]);
            o1 = __ldg(&A.soff[sl + 1]);
            if (m) *m = __ldg(&A.smask[sl]);
        }
    };
    auto issue = [&](int64_t j) {   // all lanes call; lane 0 starts the copies of slice j
        const int32_t o0 = __shfl_sync(0xffffffffu, io0, (int)(j & 31)), o1 = __shfl_sync(0xffffffffu, io1, (int)(j & 31));
        const int32_t cnt = o1 - o0;
        if (lane == 0 && cnt > 0 && cnt <= ST_SLOT_ENT) {
            const int q = (int)(j % ST_NSLOT);
            unsigned char *slot = ring + (size_t)q * ST_SLOT_BYTES;
            tma::mbar_expect_tx(&bars[q], (uint32_t)cnt * 12u);
#if PDN_SELL_EVICT
            tma::bulk_g2s_hint(slot, A.va + o0, (uint32_t)cnt * 8u, &bars[q], l2pol);
            tma::bulk_g2s_hint(slot + (size_t)ST_SLOT_ENT * 8, A.ci + o0, (uint32_t)cnt * 4u, &bars[q], l2pol);
#else
            tma::bulk_g2s(slot, A.va + o0, (uint32_t)cnt * 8u, &bars[q]);
            tma::bulk_g2s(slot + (size_t)ST_SLOT_ENT * 8, A.ci + o0, (uint32_t)cnt * 4u, &bars[q]);
#endif
        }
    };

    load_batch(0, io0, io1, nullptr);
    load_batch(0, co0, co1, &cmask);
    int64_t jn = 0;
    for (; jn < ST_NSLOT && jn < nj; jn++) issue(jn);
    uint32_t phases = 0;   // bit q: parity of the next completion of slot q
    for (int64_t j = 0; j < nj; j++) {
        if (j && (j & 31) == 0) load_batch(j, co0, co1, &cmask);
        const int32_t off = __shfl_sync(0xffffffffu, co0, (int)(j & 31)), off1 = __shfl_sync(0xffffffffu, co1, (int)(j & 31));
        const uint32_t mask = __shfl_sync(0xffffffffu, cmask, (int)(j & 31));
        const int64_t sl = s_lo + warp + j * ST_WARPS;
        const int64_t row = sl * SELL_C + lane;
        const bool live = row < A.row_end && !((mask >> lane) & 1u);
        double xi = 0.0, ai = 0.0;
        int32_t a1 = -1, a2 = -1;
        if (live) {
            if (EMODE == CG) xi = x[row];
            if (EMODE != PLAIN) ai = aux0[row];
            if (MODE == CG2) {
                a1 = __ldg(cg.agg + row);
                if (cg.nc2 > 0) a2 = __ldg(cg.agg2 + row);
            }
        }
        const int w = (off1 - off) >> 5;
        double sum = 0.0;
        if (w > 0 && w <= ST_WCAP) {
            const int q = (int)(j % ST_NSLOT);
            const unsigned char *slot = ring + (size_t)q * ST_SLOT_BYTES;
            const double *vs = reinterpret_cast<const double *>(slot) + lane;
            const int32_t *cs = reinterpret_cast<const int32_t *>(slot + (size_t)ST_SLOT_ENT * 8) + lane;
            tma::mbar_wait(&bars[q], (phases >> q) & 1u);
            phases ^= 1u << q;
            int32_t c[ST_WCAP];
            double v[ST_WCAP], xx[ST_WCAP];
#pragma unroll
            for (int k = 0; k < ST_WCAP; k++) c[k] = k < w ? cs[k * SELL_C] : 0;
#pragma unroll
            for (int k = 0; k < ST_WCAP; k++) xx[k] = k < w ? __ldg(x + (PDN_SELL_FAKEX ? (int32_t)row : c[k])) : 0.0;
#pragma unroll
            for (int k = 0; k < ST_WCAP; k++) v[k] = k < w ? vs[k * SELL_C] : 0.0;
#pragma unroll
            for (int k = 0; k < ST_WCAP; k++) sum = fma(v[k], xx[k], sum);
        } else if (w > 0) {   // wide slice: straight from global memory
            const int32_t *__restrict__ cp = A.ci + off + lane;
            const double *__restrict__ vp = A.va + off + lane;
            for (int k = 0; k < w; k += SELL_U) {
                int32_t c[SELL_U];
                double v[SELL_U], xx[SELL_U];
#pragma unroll
                for (int i = 0; i < SELL_U; i++) c[i] = k + i < w ? __ldg(cp + (k + i) * SELL_C) : 0;
#pragma unroll
                for (int i = 0; i < SELL_U; i++) v[i] = k + i < w ? __ldg(vp + (k + i) * SELL_C) : 0.0;
#pragma unroll
                for (int i = 0; i < SELL_U; i++) xx[i] = k + i < w ? __ldg(x + c[i]) : 0.0;
#pragma unroll
                for (int i = 0; i < SELL_U; i++) sum = fma(v[i], xx[i], sum);
            }
        }
        if (live) row_epilogue<EMODE>(row, sum, xi, ai, y, acc);
        // every lane has consumed its words of slot j % NSLOT: refill it with slice j + NSLOT
        __syncwarp();
        if (jn < nj) {
            if ((jn & 31) == 0) load_batch(jn, io0, io1, nullptr);
            issue(jn);
            jn++;
        }
        if (MODE == CG2) {        // after the refill: the restriction runs while the next copies fly
            restrict_smem(sum, a1, fix, amin1, rtab, cg.tab);
            if (cg.nc2 > 0) restrict_smem(sum, a2, fix, amin2, rtab + 3 * TABW, cg.tab2);
        }
    }
    __syncthreads();
    if (MODE == CG2) {
        flush_smem_table(rtab, amin1, cg.nc, cg.tab);
        if (cg.nc2 > 0) flush_smem_table(rtab + 3 * TABW, amin2, cg.nc2, cg.tab2);
    }
    const int32_t gwarp = (int32_t)((blockIdx.x * ST_BLK + threadIdx.x) >> 5), n_gwarp = (int32_t)((gridDim.x * ST_BLK) >> 5);
    if (MODE == CG2) special_rows_cg2(A, x, y, aux0, acc, gwarp, n_gwarp, cg, fix);
    else special_rows<EMODE>(A, x, y, aux0, acc, gwarp, n_gwarp);
    if (MODE == PLAIN) return;
    if (MODE == CG2) __threadfence();      // this thread's table atomics before the CTA's ticket
    if (!grid_reduce<3, ST_WARPS>(acc, partials, &sc->ticket[0], red)) return;
    if (MODE == CG2) cg2_tail<DIST>(acc, sc, cg, Lp, fix);
    else dots_tail<EMODE, DIST>(acc, sc, Lp);
}

// [first, last] aggregate among the rows [row_begin, row_end): range[0] starts at INT_MAX, range[1] at -1
__global__ void __launch_bounds__(256) agg_range_kernel(int64_t row_begin, int64_t row_end, const int32_t *__restrict__ agg,
                                                        int32_t *range) {
    int lo = 0x7fffffff, hi = -1;
    for (int64_t r = row_begin + (int64_t)blockIdx.x * 256 + threadIdx.x; r < row_end; r += (int64_t)gridDim.x * 256) {
        const int32_t a = agg[r];
        if (a >= 0) { lo = a < lo ? a : lo; hi = a > hi ? a : hi; }
    }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    if ((threadIdx.x & 31) == 0) {
        if (lo != 0x7fffffff) atomicMin(&range[0], lo);
        if (hi >= 0) atomicMax(&range[1], hi);
    }
}

// first aggregate of the restriction window of every CTA of sell_tma_kernel<CG2>: launched with THAT kernel's grid,
// same static slice split; the smallest aggregate id among the CTA's rows (0 if it has none)
__global__ void __launch_bounds__(256) agg_min_kernel(int64_t row_begin, int64_t row_end, const int32_t *__restrict__ agg,
                                                      int32_t *__restrict__ amin) {
    __shared__ int sm;
    if (threadIdx.x == 0) sm = 0x7fffffff;
    __syncthreads();
    const int64_t s_begin = row_begin / SELL_C, s_end = (row_end + SELL_C - 1) / SELL_C;
    const int64_t s_cnt = s_end - s_begin;
    const int64_t s_lo = s_begin + (int64_t)blockIdx.x * s_cnt / gridDim.x;
    const int64_t s_hi = s_begin + (int64_t)(blockIdx.x + 1) * s_cnt / gridDim.x;
    int m = 0x7fffffff;
    for (int64_t r = s_lo * SELL_C + threadIdx.x; r < s_hi * SELL_C && r < row_end; r += 256) {
        const int32_t a = agg[r];
        if (a >= 0 && a < m) m = a;
    }
    m = __reduce_min_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0) atomicMin(&sm, m);
    __syncthreads();
    if (threadIdx.x == 0) amin[blockIdx.x] = sm == 0x7fffffff ? 0 : sm;
}

// ------------------------------------------------------------------------------------------------
// setup: reduced operator  A = -(Z^T L_nn Z)  (main part)  ->  SELL-32-sigma
// ------------------------------------------------------------------------------------------------
// stored length of every row of the reduced operator (-1: padding row beyond n, 0: override row)
__global__ void __launch_bounds__(256) sell_len_kernel(int64_t n, int64_t n_pad, const int32_t *__restrict__ rp,
                                                       const int32_t *__restrict__ ci,
                                                       const int32_t *__restrict__ map,
                                                       const uint8_t *__restrict__ special,
                                                       int32_t *__restrict__ len) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * 256) {
        int32_t l = -1;
        if (i < n) {
            l = 0;
            if (!special[i])
                for (int32_t k = rp[i]; k < rp[i + 1]; k++) {
                    const int32_t c = ci[k];
                    l += (c < n && map[c] >= 0) ? 1 : 0;
                }
            if (l > SELL_MAXLEN) l = SELL_MAXLEN;
        }
        len[i] = l;
    }
}

// one CTA per window: stable sort by descending length (bitonic network on unique keys)
// (win0: first window of the launch -- a partitioned setup sorts this rank's windows only)
__global__ void __launch_bounds__(SELL_WIN) sell_sort_kernel(const int32_t *__restrict__ len,
                                                            int32_t *__restrict__ perm,
                                                            int32_t *__restrict__ inv, int64_t win0) {
    __shared__ uint32_t key[SELL_WIN];
    const int t = threadIdx.x;
    const int64_t base = (win0 + (int64_t)blockIdx.x) * SELL_WIN;
    key[t] = ((uint32_t)(SELL_MAXLEN - len[base + t]) << 10) | (uint32_t)t;
    for (int k = 2; k <= SELL_WIN; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            const int p = t ^ j;
            if (p > t) {
                const uint32_t a = key[t], b = key[p];
                const bool up = (t & k) == 0;
                if ((a > b) == up) { key[t] = b; key[p] = a; }
            }
        }
    __syncthreads();
    const int o = (int)(key[t] & (SELL_WIN - 1));
    perm[base + t] = (int32_t)(base + o);
    inv[base + o] = (int32_t)(base + t);
}

// entries per slice = 32 x the length of its first (longest) row
__global__ void __launch_bounds__(256) sell_width_kernel(int64_t n_slices, const int32_t *__restrict__ len,
                                                         const int32_t *__restrict__ perm,
                                                         int32_t *__restrict__ cnt) {
    for (int64_t s = (int64_t)blockIdx.x * 256 + threadIdx.x; s < n_slices; s += (int64_t)gridDim.x * 256) {
        const int32_t l = len[perm[s * SELL_C]];
        cnt[s] = l > 0 ? l * SELL_C : 0;
    }
}

// one warp per slice: copy the rows in (negated, columns mapped through Z and the permutation),
// pad, build the Jacobi diagonal and the override mask
__global__ void __launch_bounds__(256) sell_fill_kernel(int64_t n, int64_t n_slices, const int32_t *__restrict__ rp,
                                                        const int32_t *__restrict__ ci,
                                                        const double *__restrict__ va,
                                                        const int32_t *__restrict__ map,
                                                        const uint8_t *__restrict__ special,
                                                        const int32_t *__restrict__ perm,
                                                        const int32_t *__restrict__ inv,
                                                        const int32_t *__restrict__ soff,
                                                        int32_t *__restrict__ o_ci, double *__restrict__ o_va,
                                                        uint32_t *__restrict__ smask, double *__restrict__ dinv,
                                                        int64_t s_first) {
    // slices [s_first, n_slices): a partitioned setup fills this rank's slices only
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int64_t n_warp = ((int64_t)gridDim.x * 256) >> 5;
    for (int64_t s = s_first + warp; s < n_slices; s += n_warp) {
        const int64_t r = s * SELL_C + lane;
        const int32_t off = soff[s], w = (soff[s + 1] - off) >> 5;
        const int64_t old = perm[r];
        const bool real = old < n;
        const bool sp = real && special[old];
        int32_t k = 0;
        double d = 0.0;
        if (real && !sp) {
            for (int32_t q = rp[old]; q < rp[old + 1] && k < w; q++) {
                const int32_t c = ci[q];
                const int32_t mc = c < n ? map[c] : -1;
                if (mc < 0) continue;
                const int32_t nc = inv[mc];
                const double v = -va[q];
                o_ci[off + k * SELL_C + lane] = nc;
                o_va[off + k * SELL_C + lane] = v;
                if (nc == r) d += v;
                k++;
            }
        }
        for (; k < w; k++) {
            o_ci[off + k * SELL_C + lane] = real ? (int32_t)r : 0;
            o_va[off + k * SELL_C + lane] = 0.0;
        }
        const uint32_t m = __ballot_sync(0xffffffffu, sp);
        if (lane == 0) smask[s] = m;
        if (real) dinv[r] = sp ? 1.0 : 1.0 / d;   // override rows are set afterwards
    }
}

__global__ void __launch_bounds__(256) remap_i32_kernel(int32_t k, const int32_t *__restrict__ inv, int32_t *idx) {
    for (int32_t i = blockIdx.x * 256 + threadIdx.x; i < k; i += gridDim.x * 256) idx[i] = inv[idx[i]];
}
__global__ void __launch_bounds__(256) flag_u8_kernel(int32_t k, const int32_t *__restrict__ idx, uint8_t *flag) {
    for (int32_t i = blockIdx.x * 256 + threadIdx.x; i < k; i += gridDim.x * 256) flag[idx[i]] = 1;
}
// b_new[r] = -g[perm[r]]
__global__ void __launch_bounds__(256) permute_negate_kernel(int64_t n, const int32_t *__restrict__ perm,
                                                             const double *__restrict__ g, double *__restrict__ b) {
    for (int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x; r < n; r += (int64_t)gridDim.x * 256) b[r] = -g[perm[r]];
}
