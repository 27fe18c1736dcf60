// pdn_spmv.cuh -- kernel (3a): CSR-stream SpMV with a TMA bulk-copy pipeline and fused dot products.
//
// The stored entries of SROWS consecutive rows are one contiguous range of the CSR value / column
// arrays.  A producer warp streams that range into shared memory with cp.async.bulk (the TMA
// engine; UBLKCP in SASS) STAGES work items ahead of the consumers, completion tracked by
// mbarriers, so the HBM stream never waits on the x gather.  The SROWS consumer threads then
//   (1) multiply every staged value by x[column] (column indices read from shared memory, x
//       gathered through the read-only path: L1/L2 resident for mesh-local numbering), ITEMS
//       independent gathers per thread in flight, products written back in place, and
//   (2) sum their own row out of shared memory in column order (fixed order: bit-reproducible).
// HBM traffic is the algorithmic 12 B per stored entry + 4 B per row + the vectors: no padding,
// no permutation.  Work is split statically (contiguous row-block ranges per CTA, grid = a
// multiple of the SM count), which keeps the fused reductions deterministic.
//
// Included by pdn_kernels.cuh (namespace pdnk).
#pragma once
#ifndef PDN_EXP
#define PDN_EXP 0
#endif

namespace tma {

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_addr(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy (bytes: multiple of 16, both addresses 16 B aligned)
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}
// the same with an L2 eviction policy (createpolicy): the matrix is streamed once per SpMV, evict_first keeps
// it from pushing the Krylov vectors out of the 126 MB L2
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_addr(dst)),
        "l"(src), "r"(bytes), "r"(smem_addr(bar)), "l"(pol)
        : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace tma

// consumer-only barrier (named barrier 1): the producer warp never joins it
template <int NTHREADS>
__device__ __forceinline__ void consumer_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(NTHREADS) : "memory");
}

template <int MODE>
__device__ __forceinline__ void row_epilogue(int64_t row, double sum, double xi, double ai, double *__restrict__ y,
                                             double (&acc)[3]) {
    // xi = x[row] (CG only), ai = aux0[row]
    if (MODE == PLAIN) {
        y[row] = sum;
    } else if (MODE == CG) {
        y[row] = sum;
        acc[0] += ai * xi; acc[1] += xi * sum; acc[2] += ai * ai;
    } else if (MODE == BICG_V) {
        y[row] = sum;
        acc[0] += ai * sum;
    } else if (MODE == BICG_T) {
        y[row] = sum;
        acc[0] += sum * ai; acc[1] += sum * sum;
    } else {  // RESID
        const double ri = ai - sum;
        y[row] = ri;
        acc[2] += ri * ri;
    }
}

// scalar recurrences, run by one thread of the last block to finish
template <int MODE>
__device__ __forceinline__ void spmv_finalize(const double (&acc)[3], Scal *sc) {
    if (MODE == CG) {
        const double gamma = acc[0], delta = acc[1], rr = acc[2];
        sc->rr = rr;
        if (sc->max_iter < 0) return;   // benchmark mode (pdn_bench_spmv): never latch `done`
        if (rr <= sc->tol2) { sc->done = 1; return; }
        if (sc->iter >= sc->max_iter) { sc->done = 2; return; }
        double beta, alpha;
        if (sc->first) {
            beta = 0.0;
            alpha = gamma / delta;
            sc->first = 0;
        } else {
            beta = gamma / sc->gamma_old;
            alpha = gamma / (delta - beta * gamma / sc->alpha);
        }
        if (!(alpha == alpha) || !(fabs(alpha) < 1e300) || delta <= 0.0) { sc->breakdown = 1; sc->done = 3; return; }
        sc->gamma_old = gamma;
        sc->gamma = gamma; sc->delta = delta;
        sc->alpha = alpha; sc->beta = beta;
        sc->iter += 1;
    } else if (MODE == BICG_V) {
        sc->r0v = acc[0];
        const double alpha = sc->rho / acc[0];
        if (!(alpha == alpha) || !(fabs(alpha) < 1e300)) { sc->breakdown = 1; sc->done = 3; return; }
        sc->alpha = alpha;
    } else if (MODE == BICG_T) {
        sc->ts = acc[0]; sc->tt = acc[1];
        sc->omega = acc[1] > 0.0 ? acc[0] / acc[1] : 0.0;
    } else if (MODE == RESID) {
        sc->rr = acc[2];
    }
}

// geometry of the pipelined kernel
#ifndef PDN_SROWS
#define PDN_SROWS 128
#endif
#ifndef PDN_STAGES
#define PDN_STAGES 4
#endif
#ifndef PDN_SITEMS
#define PDN_SITEMS 8
#endif
#ifndef PDN_MINB
#define PDN_MINB 3
#endif
constexpr int SROWS = PDN_SROWS;           // rows per work item = consumer threads per CTA
constexpr int SITEMS = PDN_SITEMS;         // gathers in flight per consumer thread
constexpr int SCAP = SROWS * SITEMS;       // stored entries per stage (multiple of 4)
constexpr int STAGES = PDN_STAGES;
constexpr int STHREADS = SROWS + 32;       // + one producer warp
constexpr int SWARPS = STHREADS / 32;
constexpr size_t SPMV_SMEM = (size_t)(STAGES * 12 + 16) * SCAP + 2 * STAGES * 8 + 3 * SWARPS * 8 + 64;

// number of stage-sized chunks of the aligned entry range [ka, k1)
__device__ __forceinline__ int32_t chunk_count(int32_t ka, int32_t k1) { return k1 > ka ? (k1 - ka + SCAP - 1) / SCAP : 0; }

// Raw prefetched metadata of one work item.  NO arithmetic is done on the loaded words until the
// item becomes current, so the loads stay in flight for a whole item instead of stalling at issue.
struct ItemRaw {
    int32_t v, v1, vk0, vk1;  // rp[row], rp[row+1], rp[r0], rp[r1]  (undecoded)
    double xi, ai;            // x[row], aux0[row]
    bool in_range;
};
struct ItemMeta {
    int32_t lo, hi, k0, k1;   // own row range, entry range of the whole work item
    double xi, ai;
    bool live;                // row exists and is not an override row
};

template <int MODE>
__device__ __forceinline__ ItemRaw load_meta(const Operator &A, const double *__restrict__ x,
                                             const double *__restrict__ aux0, int64_t blk, int tid) {
    ItemRaw m;
    m.v = m.v1 = m.vk0 = m.vk1 = 0;
    m.xi = m.ai = 0.0;
    m.in_range = false;
    if (blk < 0) return m;
    const int64_t r0 = blk * SROWS;
    const int64_t r1 = r0 + SROWS < A.n ? r0 + SROWS : A.n;
    const int64_t row = r0 + tid;
    m.vk0 = __ldg(&A.rp[r0]);
    m.vk1 = __ldg(&A.rp[r1]);
    if (row < r1) {
        m.in_range = true;
        m.v = __ldg(&A.rp[row]);
        m.v1 = __ldg(&A.rp[row + 1]);
        if (MODE == CG) m.xi = __ldg(&x[row]);
        if (MODE != PLAIN) m.ai = __ldg(&aux0[row]);
    }
    return m;
}

__device__ __forceinline__ ItemMeta decode_meta(const ItemRaw &r) {
    ItemMeta m;
    m.k0 = rp_decode(r.vk0);
    m.k1 = rp_decode(r.vk1);
    m.lo = rp_decode(r.v);
    m.live = r.in_range && r.v >= 0;
    m.hi = m.live ? rp_decode(r.v1) : m.lo;
    m.xi = r.xi;
    m.ai = r.ai;
    return m;
}

// sum ps[a-s0 .. b-s0) in order; rows of a P1 mesh hold ~7 entries, so the first 8 are predicated
__device__ __forceinline__ double row_sum(const double *__restrict__ ps, int32_t a, int32_t b, int32_t s0, double sum) {
#if PDN_EXP == 5
    for (int32_t k = a; k < b; k++) sum += ps[k - s0];
    return sum;
#endif
    const double *q = ps + (a - s0);
    const int n = b - a;
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (k < n) sum += q[k];
    for (int k = 8; k < n; k++) sum += q[k];
    return sum;
}

// a chunk whose products sit in prod[buf] and still has to be summed into its rows
struct Pending {
    int32_t lo, hi, s0, cnt;
    int buf;
    bool valid, last, live;
    int64_t row;
    double xi, ai;
};

template <int MODE>
__global__ void __launch_bounds__(STHREADS, PDN_MINB) spmv_kernel(Operator A, const double *__restrict__ x,
                                                        double *__restrict__ y, const double *__restrict__ aux0,
                                                        Scal *sc, double *partials) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *va_s = reinterpret_cast<double *>(smem_raw);                                  // [STAGES][SCAP]
    int32_t *ci_s = reinterpret_cast<int32_t *>(smem_raw + (size_t)STAGES * SCAP * 8);     // [STAGES][SCAP]
    double *prod = reinterpret_cast<double *>(smem_raw + (size_t)STAGES * SCAP * 12);      // [2][SCAP]
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + (size_t)(STAGES * 12 + 16) * SCAP);  // [STAGES]
    uint64_t *empty = full + STAGES;                                                      // [STAGES]
    double(*red)[SWARPS] = reinterpret_cast<double(*)[SWARPS]>(empty + STAGES);           // [3][SWARPS]

    if (MODE != PLAIN && MODE != RESID) {
        if (sc->done) return;
    }
    const int tid = threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < STAGES; s++) {
            tma::mbar_init(&full[s], 1);
            tma::mbar_init(&empty[s], SROWS / 32);
        }
        tma::fence_barrier_init();
    }
    __syncthreads();

    // static split: CTA b owns row-blocks [b * n_blk / grid, (b + 1) * n_blk / grid)
    const int64_t n_blk = (A.n + SROWS - 1) / SROWS;
    const int64_t blk_lo = (int64_t)blockIdx.x * n_blk / gridDim.x;
    const int64_t blk_hi = (int64_t)(blockIdx.x + 1) * n_blk / gridDim.x;
    double acc[3] = {0.0, 0.0, 0.0};

    if (tid >= SROWS) {
        // ---------------- producer warp ---------------------------------------------------------
        // The 32 lanes fetch the entry offsets of the next 31 work items with one coalesced load
        // (off the critical path); lane 0 then keeps every free stage refilled.
        const int lane = tid - SROWS;
        int stage = 0;
        uint32_t phase = 0;
        for (int64_t base = blk_lo; base < blk_hi; base += 31) {
            int64_t rb = (base + lane) * SROWS;
            if (rb > A.n) rb = A.n;
            const int32_t kb = rp_decode(__ldg(&A.rp[rb]));
            const int n_items = blk_hi - base < 31 ? (int)(blk_hi - base) : 31;
            for (int j = 0; j < n_items; j++) {
                const int32_t k0 = __shfl_sync(0xffffffffu, kb, j), k1 = __shfl_sync(0xffffffffu, kb, j + 1);
                if (lane == 0) {
                    const int32_t ka = k0 & ~3;   // 16 B alignment of both arrays
                    const int32_t nch = chunk_count(ka, k1);
                    for (int32_t ch = 0; ch < nch; ch++) {
                        const int32_t s0 = ka + ch * SCAP;
                        int32_t cnt = k1 - s0 < SCAP ? k1 - s0 : SCAP;
                        cnt = (cnt + 3) & ~3;     // the arrays carry >= 64 entries of slack
                        tma::mbar_wait(&empty[stage], phase ^ 1u);
                        tma::mbar_expect_tx(&full[stage], (uint32_t)cnt * 12u);
                        tma::bulk_g2s(va_s + (size_t)stage * SCAP, A.va + s0, (uint32_t)cnt * 8u, &full[stage]);
                        tma::bulk_g2s(ci_s + (size_t)stage * SCAP, A.ci + s0, (uint32_t)cnt * 4u, &full[stage]);
                        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------- consumers: one thread per row of the work item ----------------------------
        // Software pipeline over chunks: (1) read the staged column indices and issue the x gathers,
        // (2) while they fly, sum the rows of the PREVIOUS chunk out of its product buffer,
        // (3) multiply the staged values by the gathered x into this chunk's product buffer and
        // hand the stage back to the producer.
        int stage = 0;
        uint32_t phase = 0;
        int buf = 0;
        double sum = 0.0;
        Pending pend;
        pend.valid = false;
        ItemRaw nxt = load_meta<MODE>(A, x, aux0, blk_lo < blk_hi ? blk_lo : -1, tid);
        for (int64_t blk = blk_lo; blk < blk_hi; blk++) {
            const ItemMeta cur = decode_meta(nxt);
            nxt = load_meta<MODE>(A, x, aux0, blk + 1 < blk_hi ? blk + 1 : -1, tid);
            const int32_t ka = cur.k0 & ~3;
            const int32_t nch = chunk_count(ka, cur.k1);
            if (nch == 0 && cur.live) {   // rows without stored entries (cannot happen for live rows)
                row_epilogue<MODE>(blk * SROWS + tid, 0.0, cur.xi, cur.ai, y, acc);
            }
            for (int32_t ch = 0; ch < nch; ch++) {
                const int32_t s0 = ka + ch * SCAP;
                const int32_t cnt = cur.k1 - s0 < SCAP ? cur.k1 - s0 : SCAP;   // valid entries
                const double *vs = va_s + (size_t)stage * SCAP;
                const int32_t *cs = ci_s + (size_t)stage * SCAP;
                int32_t cc[SITEMS];
                double xx[SITEMS];
                tma::mbar_wait(&full[stage], phase);
#pragma unroll
                for (int j = 0; j < SITEMS; j++) cc[j] = tid + j * SROWS < cnt ? cs[tid + j * SROWS] : 0;
#pragma unroll
                for (int j = 0; j < SITEMS; j++) xx[j] = __ldg(&x[cc[j]]);
                // previous chunk: its products were stored before this barrier
                consumer_sync<SROWS>();
                if (pend.valid) {
                    const double *ps = prod + (size_t)pend.buf * SCAP;
                    const int32_t a = pend.lo > pend.s0 ? pend.lo : pend.s0;
                    const int32_t b = pend.hi < pend.s0 + pend.cnt ? pend.hi : pend.s0 + pend.cnt;
                    sum = row_sum(ps, a, b, pend.s0, sum);
                    if (pend.last) {
                        if (pend.live) row_epilogue<MODE>(pend.row, sum, pend.xi, pend.ai, y, acc);
                        sum = 0.0;
                    }
                }
                double *pd = prod + (size_t)buf * SCAP;
#pragma unroll
                for (int j = 0; j < SITEMS; j++) pd[tid + j * SROWS] = vs[tid + j * SROWS] * xx[j];
                // Hand the stage back only now: the multiplies above consumed every staged word, so no
                // shared-memory read of this stage can still be queued when the next bulk copy lands.
                __syncwarp();
                if ((tid & 31) == 0) tma::mbar_arrive(&empty[stage]);
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                pend.valid = true;
                pend.lo = cur.lo; pend.hi = cur.hi; pend.s0 = s0; pend.cnt = cnt; pend.buf = buf;
                pend.last = ch == nch - 1; pend.live = cur.live; pend.row = blk * SROWS + tid;
                pend.xi = cur.xi; pend.ai = cur.ai;
                buf ^= 1;
            }
        }
        consumer_sync<SROWS>();
        if (pend.valid) {   // drain
            const double *ps = prod + (size_t)pend.buf * SCAP;
            const int32_t a = pend.lo > pend.s0 ? pend.lo : pend.s0;
            const int32_t b = pend.hi < pend.s0 + pend.cnt ? pend.hi : pend.s0 + pend.cnt;
            sum = row_sum(ps, a, b, pend.s0, sum);
            if (pend.live) row_epilogue<MODE>(pend.row, sum, pend.xi, pend.ai, y, acc);
        }
        // override rows (constraint terminals): CTA 0, one consumer thread per row
        if (blockIdx.x == 0) {
            for (int32_t t = tid; t < A.n_special; t += SROWS) {
                const int32_t row = A.sp_row[t];
                double s2 = 0.0;
                for (int32_t k = A.sp_rp[t]; k < A.sp_rp[t + 1]; k++) s2 += A.sp_va[k] * x[A.sp_ci[k]];
                row_epilogue<MODE>(row, s2, MODE == CG ? x[row] : 0.0, MODE != PLAIN ? aux0[row] : 0.0, y, acc);
            }
        }
    }
    if (MODE == PLAIN) return;
    if (!grid_reduce<3, SWARPS>(acc, partials, &sc->ticket[0], red)) return;
    if (tid == 0) spmv_finalize<MODE>(acc, sc);
}
