// pdn_coarse.cuh -- kernel (4'): two-level preconditioner  M^-1 = D^-1 + P (P^T A P)^-1 P^T.
//
// Jacobi alone needs O(sqrt(N)) iterations on the plane Laplacians (4 204 at 1M DOF).  The second
// level is a coarse space of piecewise constants over AGGREGATES: the vertices of every mesh are
// binned on a uniform grid of their coordinates (pdn_assemble keeps xy resident), internal network
// nodes get aggregates of their own, rows eliminated by the constraints get none.  About a thousand
// aggregates remove the long-wavelength error the Jacobi sweep cannot see; the iteration count then
// depends on the bin size in mesh cells, not on N (measured on the CPU model: 1 474 -> 381 / 211 / 130
// at 32 / 16 / 8 cells per bin).
//
//   setup   aggregate ids (bins -> compact ids by a scan)
//           A_c = P^T A P accumulated in 64-bit FIXED POINT: integer atomics are associative, so the
//           coarse matrix -- and with it the whole solve -- stays bit-identical run to run
//           A_c^-1 by a cooperative blocked Gauss-Jordan: every CTA keeps its rows of the matrix in
//           shared memory for the whole inversion, one grid barrier per block of pivots
//   apply   r_c = P^T r  is fused into the PCG vector update (fixed point again, scaled by ||r||)
//           x_c = A_c^-1 r_c   dense mat-vec, one warp per row
//           z   = D^-1 r + P x_c
//
// Included by pdn_kernels.cuh (namespace pdnk).
#pragma once

constexpr int CO_MAX = PDN_CO_MAX;         // largest coarse space: 12 rows x 1776 doubles per CTA = 170 KB
constexpr int CO_BLK = 256;
constexpr double CO_RFIX = 1099511627776.0;   // 2^40: fixed-point steps per ||r|| in the restriction
// PCG restricts with 32-bit fixed point (one redux.sync instead of three, a fraction of the integer work:
// cg_update_kernel<TL> was instruction-issue bound, ~250 instructions per 32 rows).  2^22 steps per ||r_k||,
// entries clamped so that 32 lanes (of one or two rows each) cannot overflow: 8-16x headroom for a residual that grows in
// one step.  The quantisation (2.4e-7 ||r|| per entry) only perturbs the preconditioner.
#ifndef PDN_RESTRICT_WIDE
#define PDN_RESTRICT_WIDE 0
#endif
constexpr double CO_RFIX32 = PDN_RESTRICT_WIDE ? CO_RFIX : 4194304.0;
constexpr double CO_BOOST32 = CO_RFIX / CO_RFIX32;   // what coarse_apply_kernel multiplies back

struct CoarseGeom {
    int n_mesh;
    int64_t n_v;               // mesh vertices (old numbering: [0, n_v) have coordinates)
    double x0, y0, inv_h;      // bin = floor((x - x0) * inv_h)
    int32_t nbx, nby;          // bins per mesh
    int32_t n_bins;            // n_mesh * nbx * nby
    int32_t node_group;        // internal nodes per aggregate
};

// per-block bounding box partials: out[4*b + {0,1,2,3}] = min x, min y, max x, max y
__global__ void __launch_bounds__(CO_BLK) bbox_kernel(int64_t n_v, const double2 *__restrict__ xy, double *__restrict__ out) {
    __shared__ double sm[4][CO_BLK / 32];
    double v[4] = {1e300, 1e300, -1e300, -1e300};
    for (int64_t i = (int64_t)blockIdx.x * CO_BLK + threadIdx.x; i < n_v; i += (int64_t)gridDim.x * CO_BLK) {
        const double2 p = xy[i];
        v[0] = fmin(v[0], p.x); v[1] = fmin(v[1], p.y); v[2] = fmax(v[2], p.x); v[3] = fmax(v[3], p.y);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int a = 0; a < 4; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double t = __shfl_down_sync(0xffffffffu, v[a], o);
            v[a] = a < 2 ? fmin(v[a], t) : fmax(v[a], t);
        }
        if (lane == 0) sm[a][warp] = v[a];
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        const int a = threadIdx.x;
        double r = sm[a][0];
        for (int k = 1; k < CO_BLK / 32; k++) r = a < 2 ? fmin(r, sm[a][k]) : fmax(r, sm[a][k]);
        out[4 * blockIdx.x + a] = r;
    }
}

// raw aggregate id of every reduced row (permuted numbering); -1 for rows the constraints eliminated and for
// rows outside [row_lo, row_hi) (the rank-local level of a partitioned solve only aggregates its own rows)
__global__ void __launch_bounds__(CO_BLK) agg_raw_kernel(int64_t n, int64_t row_lo, int64_t row_hi, CoarseGeom g,
                                                         const int64_t *__restrict__ v_off,
                                                         const double2 *__restrict__ xy,
                                                         const int32_t *__restrict__ perm,
                                                         const int32_t *__restrict__ map, int32_t *__restrict__ raw,
                                                         int32_t *__restrict__ used) {
    for (int64_t r = (int64_t)blockIdx.x * CO_BLK + threadIdx.x; r < n; r += (int64_t)gridDim.x * CO_BLK) {
        const int32_t i = perm[r];
        int32_t a = -1;
        if (map[i] == i && r >= row_lo && r < row_hi) {
            if (i < g.n_v) {
                int lo = 0, hi = g.n_mesh;          // largest m with v_off[m] <= i
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (v_off[mid] <= i) lo = mid; else hi = mid;
                }
                const double2 p = xy[i];
                int bx = (int)floor((p.x - g.x0) * g.inv_h), by = (int)floor((p.y - g.y0) * g.inv_h);
                bx = bx < 0 ? 0 : (bx >= g.nbx ? g.nbx - 1 : bx);
                by = by < 0 ? 0 : (by >= g.nby ? g.nby - 1 : by);
                a = (lo * g.nby + by) * g.nbx + bx;
            } else {
                a = g.n_bins + (int32_t)((i - g.n_v) / g.node_group);
            }
            used[a] = 1;
        }
        raw[r] = a;
    }
}

// occupancy of the raw bins from an (all-gathered) raw id array
__global__ void __launch_bounds__(CO_BLK) agg_used_kernel(int64_t n, const int32_t *__restrict__ raw, int32_t *__restrict__ used) {
    for (int64_t r = (int64_t)blockIdx.x * CO_BLK + threadIdx.x; r < n; r += (int64_t)gridDim.x * CO_BLK) {
        const int32_t a = raw[r];
        if (a >= 0) used[a] = 1;
    }
}

__global__ void __launch_bounds__(CO_BLK) agg_final_kernel(int64_t n, const int32_t *__restrict__ cid, int32_t *__restrict__ agg) {
    for (int64_t r = (int64_t)blockIdx.x * CO_BLK + threadIdx.x; r < n; r += (int64_t)gridDim.x * CO_BLK) {
        const int32_t a = agg[r];
        agg[r] = a < 0 ? -1 : cid[a];
    }
}

// largest Jacobi diagonal (= 1 / smallest dinv): the fixed-point scale of A_c
__global__ void __launch_bounds__(CO_BLK) min_kernel(int64_t n, const double *__restrict__ v, double *__restrict__ out) {
    __shared__ double sm[CO_BLK / 32];
    double m = 1e300;
    for (int64_t i = (int64_t)blockIdx.x * CO_BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * CO_BLK) m = fmin(m, v[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmin(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < CO_BLK / 32; k++) m = fmin(m, sm[k]);
        out[blockIdx.x] = m;
    }
}

// one 64-bit integer atomic per distinct key in the warp (lanes with key < 0 contribute nothing)
__device__ __forceinline__ void warp_add_by_key(long long key, long long val, unsigned long long *table) {
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(peers) - 1;
    long long sum = 0;
    for (int l = 0; l < 32; l++) {
        const long long v = __shfl_sync(0xffffffffu, val, l);
        if ((peers >> l) & 1u) sum += v;
    }
    if (lane == leader && key >= 0 && sum != 0) atomicAdd(&table[key], (unsigned long long)sum);
}

// A_c[I][J] += a_ij (fixed point) over the SELL part (one warp per slice) ...
__global__ void __launch_bounds__(CO_BLK) coarse_accum_kernel(SellOp A, int64_t n_slices, const int32_t *__restrict__ agg,
                                                              int32_t nc, double fix, unsigned long long *__restrict__ acc) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * CO_BLK + threadIdx.x) >> 5;
    const int64_t n_warp = ((int64_t)gridDim.x * CO_BLK) >> 5;
    for (int64_t s = warp; s < n_slices; s += n_warp) {
        const int64_t r = s * SELL_C + lane;
        const int32_t off = A.soff[s], w = (A.soff[s + 1] - off) >> 5;
        const int32_t I = r < A.n ? agg[r] : -1;
        for (int32_t k = 0; k < w; k++) {
            const int32_t c = A.ci[off + k * SELL_C + lane];
            const double v = A.va[off + k * SELL_C + lane];
            const int32_t J = I >= 0 && v != 0.0 ? agg[c] : -1;
            const long long key = J >= 0 ? (long long)I * nc + J : -1;
            warp_add_by_key(key, llrint(v * fix), acc);
        }
    }
}
// ... and over the override rows (one thread per row, plain atomics: there are few)
__global__ void __launch_bounds__(CO_BLK) coarse_accum_special_kernel(SellOp A, const int32_t *__restrict__ agg, int32_t nc,
                                                                      double fix, unsigned long long *__restrict__ acc) {
    for (int32_t t = blockIdx.x * CO_BLK + threadIdx.x; t < A.n_special; t += gridDim.x * CO_BLK) {
        const int32_t I = agg[A.sp_row[t]];
        if (I < 0) continue;
        for (int32_t q = A.sp_rp[t]; q < A.sp_rp[t + 1]; q++) {
            const int32_t J = agg[A.sp_ci[q]];
            if (J >= 0) atomicAdd(&acc[(long long)I * nc + J], (unsigned long long)llrint(A.sp_va[q] * fix));
        }
    }
}

// fixed point -> double (symm: averaged with the transpose -- a no-op for a symmetric operator)
__global__ void __launch_bounds__(CO_BLK) coarse_convert_kernel(int32_t nc, double inv_fix, const long long *__restrict__ acc,
                                                                double *__restrict__ Ac, int symm) {
    const int64_t tot = (int64_t)nc * nc;
    for (int64_t e = (int64_t)blockIdx.x * CO_BLK + threadIdx.x; e < tot; e += (int64_t)gridDim.x * CO_BLK) {
        const int32_t i = (int32_t)(e / nc), j = (int32_t)(e - (int64_t)i * nc);
        Ac[e] = symm ? 0.5 * ((double)acc[e] + (double)acc[(int64_t)j * nc + i]) * inv_fix : (double)acc[e] * inv_fix;
    }
}

// ------------------------------------------------------------------------------------------------
// in-place inverse of the SPD coarse matrix: cooperative blocked Gauss-Jordan (no pivoting).
// CTA b owns rows [b*R, (b+1)*R) and keeps them in shared memory.  Block step b (the sweep operator over
// the pivot block K = rows of CTA b, D = A_KK):
//   owner:   A_K,j <- D^-1 A_K,j (j not in K),  A_KK <- D^-1; D^-1 is an R x R Gauss-Jordan by the whole
//            CTA, the product is one pass over the rows; the new rows are published to piv[b] and
//            flag[b] is released.  (v1 swept the R pivots one after the other over all nc columns: 2.4 us
//            per pivot, 2.3 ms at nc = 968, and taller blocks made it slower.)
//   others:  wait for flag[b] only -- no grid-wide barrier, piv[b] is never reused -- then
//            A_i,j <- (j in K ? 0 : A_i,j) - sum_q A_i,K[q] * piv_q,j.
// Launched cooperatively (all CTAs co-resident); status[0] != 0 on a non-positive pivot / timeout.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool flag_wait(const unsigned int *flag, int32_t *status) {
    __shared__ int ok;
    __syncthreads();
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        int good = 1;
        while (true) {
            unsigned int v;
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
            if (v != 0u) break;
            if (clock64() - t0 > 4000000000LL) { good = 0; atomicExch(&status[0], PDN_ECUDA); break; }   // ~2 s
        }
        ok = good;
    }
    __syncthreads();
    return ok != 0;
}

// out_i,j = sum_q c[i][q] * P_q,j for four columns per thread at a time: the 16 x 4 operand block sits
// in registers, every multiplier is read from shared memory once per four columns (one LDS per FMA made
// both phases LDS-bound).
//   OWNER:  P = my old rows, c = D^-1:        A_K,j <- out (j not in K) | c[i][j-k0] (j in K)
//   else:   P = pb (the owner's OLD rows), c = A_iK D^-1:  A_i,j <- A_i,j - out (j not in K) | -c[i][j-k0] (j in K)
template <bool OWNER>
__device__ __forceinline__ void co_rank_update(double *rows, const double (*cs)[16], const double *pb, int nc, int ni,
                                               int nk, int k0) {
    const int tid = threadIdx.x;
    for (int j0 = tid; j0 < nc; j0 += 4 * CO_BLK) {
        double p[4][16];
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const int j = j0 + t * CO_BLK;
#pragma unroll
            for (int q = 0; q < 16; q++)
                p[t][q] = (j < nc && q < nk) ? (OWNER ? rows[q * nc + j] : __ldcg(&pb[q * nc + j])) : 0.0;
        }
        for (int i = 0; i < ni; i++) {
            double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
            for (int q = 0; q < 16; q += 2) {
                const double2 c = *reinterpret_cast<const double2 *>(&cs[i][q]);
#pragma unroll
                for (int t = 0; t < 4; t++) acc[t] += c.x * p[t][q] + c.y * p[t][q + 1];
            }
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const int j = j0 + t * CO_BLK;
                if (j >= nc) continue;
                const bool in_k = j >= k0 && j < k0 + nk;
                if (OWNER) rows[i * nc + j] = in_k ? cs[i][j - k0] : acc[t];
                else rows[i * nc + j] = in_k ? -cs[i][j - k0] : rows[i * nc + j] - acc[t];
            }
        }
    }
}

// piv[kb]: the owner's rows BEFORE the multiplication by D^-1 (R * nc doubles), then D^-1 (16 x 16).
// Publishing the old rows lets every other CTA start its update while the owner still multiplies its own
// rows: the serial chain per block step is (update my rows, invert D, publish), not (..., multiply, publish).
__global__ void __launch_bounds__(CO_BLK, 1) coarse_invert_kernel(int32_t nc, int32_t R, double *__restrict__ A,
                                                                  double *__restrict__ piv /*[n_blk][R * nc + 256]*/,
                                                                  unsigned int *flags /*[n_blk], zeroed*/, int32_t *status,
                                                                  float *__restrict__ Af /* the inverse in fp32 */) {
    extern __shared__ double rows[];                 // [R][nc]
    __shared__ __align__(16) double cs[16][16];      // D^-1 (owner) / A_iK D^-1 (others); R <= 16
    __shared__ __align__(16) double dv[16][16], ak[16][16];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int r0 = b * R;
    const int nr = r0 >= nc ? 0 : (r0 + R <= nc ? R : nc - r0);
    for (int e = tid; e < nr * nc; e += CO_BLK) rows[e] = A[(size_t)r0 * nc + e];
    __syncthreads();
    const int n_blk = (nc + R - 1) / R;
    const size_t pstride = (size_t)R * nc + 256;
    const int di = tid >> 4, dj = tid & 15;          // CO_BLK == 256: one thread per entry of a 16 x 16 block
    for (int kb = 0; kb < n_blk; kb++) {
        const int k0 = kb * R, nk = k0 + R <= nc ? R : nc - k0;
        double *pb = piv + (size_t)kb * pstride;
        if (kb == b) {
            // D^-1 in place (identity padding beyond nk)
            cs[di][dj] = (di < nk && dj < nk) ? rows[di * nc + k0 + dj] : (di == dj ? 1.0 : 0.0);
            for (int q = 0; q < nk; q++) {
                __syncthreads();
                const double p = cs[q][q], rq = cs[q][dj], cq = cs[di][q], cur = cs[di][dj];
                __syncthreads();
                if (!(p > 0.0) && tid == 0) atomicExch(&status[0], PDN_ESINGULAR);
                const double ip = 1.0 / p;
                const double pr = (dj == q ? 1.0 : rq) * ip;
                cs[di][dj] = di == q ? pr : (dj == q ? 0.0 : cur) - cq * pr;
            }
            __syncthreads();
            for (int e = tid; e < nk * nc; e += CO_BLK) pb[e] = rows[e];
            pb[(size_t)R * nc + tid] = cs[di][dj];
            __threadfence();
            __syncthreads();
            if (tid == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flags + kb), "r"(1u) : "memory");
            co_rank_update<true>(rows, cs, nullptr, nc, nk, nk, k0);
            __syncthreads();
        } else {
            if (!flag_wait(flags + kb, status)) return;
            dv[di][dj] = __ldcg(&pb[(size_t)R * nc + tid]);
            ak[di][dj] = (di < nr && dj < nk) ? rows[di * nc + k0 + dj] : 0.0;
            __syncthreads();
            double c = 0.0;
#pragma unroll
            for (int q = 0; q < 16; q++) c += ak[di][q] * dv[q][dj];
            cs[di][dj] = c;
            __syncthreads();
            co_rank_update<false>(rows, cs, pb, nc, nr, nk, k0);
            __syncthreads();
        }
    }
    // the iteration applies the inverse from an fp32 copy (half the L2 traffic of the latency-bound
    // coarse_apply_kernel; a 6e-8 relative perturbation of a preconditioner)
    for (int e = tid; e < nr * nc; e += CO_BLK) {
        A[(size_t)r0 * nc + e] = rows[e];
        Af[(size_t)r0 * nc + e] = (float)rows[e];
    }
}

// ------------------------------------------------------------------------------------------------
// Round-2 form of the same sweep (default; PADNE_B200_INVERT=v1 keeps the kernel above).  Measured with per-phase
// cycle counters (-DPDN_INV_TIMING, nc = 1458, R = 16, 92 block steps), a step of the kernel above is a serial chain
//   rank-16 update 24.2k cycles -> owner publishes its 16 rows 8.5k -> owner inverts D 7.1k -> c = A_iK D^-1 1.3k
// = 41k cycles = 22 us, and the update itself runs at 1.2k cycles per row where the same loop in isolation
// (scripts/micro/upd_loop.cu) needs 540: with 16 pivot rows x 3 columns of the published block in registers (96)
// the loop sits on the 128-register ceiling and every broadcast LDS of c is issued only after the DFMAs of the
// previous one (the form `acc += c.x * p + c.y * p'` also compiled to DMUL + DFMA + DADD).  Here:
//   * RT = 12 rows per CTA (compile time): 72 registers of published rows, so the c loads run ahead of the DFMAs;
//     per step the update is (12/16)^2 of the work, for 4/3 of the steps;
//   * the update is a pure DFMA chain per entry, acc = fma(c_q, p_q, acc) starting from the entry itself;
//   * the NEXT owner publishes its rows with one bulk copy shared -> global by the TMA engine as soon as its 15
//     column warps are done (plain stores: 8.5k cycles as a separate copy, +6k when issued from the update loop);
//   * warp 15 prepares the NEXT owner's D^-1: while the others update the rows, it forms the new diagonal block from
//     the published rows with the same fma chain (same bits as the update will store), inverts it in shared memory
//     with warp-level barriers only and publishes it; an owner therefore raises ONE flag, right after its fence, for
//     rows and D^-1 together.  Warp 15 owns only the LAST 32 * NCOL columns of a pass (18 of 1458: one column per
//     lane instead of three), so the inversion hides behind the other warps' update;
//   * the others pull the published rows from L2 into registers before they form c = -A_iK D^-1.
// Same sweep and pivot order as above; an entry differs in the last bit (fma chain instead of mul/fma/add).
// ------------------------------------------------------------------------------------------------
constexpr int INV_BLK = 512;               // columns per pass and per column-per-thread
constexpr int INV_THREADS = 512;           // 16 warps (a 17th would be allocated as 20: 96 registers per thread)
constexpr int INV_CW = INV_THREADS - 32;   // warps 0..14 own the first 480 * NCOL columns of a pass; warp 15, the
                                           // inverter, the last 32 * NCOL (none at all when nc <= 480 * NCOL + ...)

// In-place Gauss-Jordan of the n x n block in d (16 x 17 storage -- a 16-double row stride puts a whole column on one
// pair of banks: 1000 cycles per pivot --, identity beyond n) by ONE warp:
// lane = (row r = lane & 15, half h = lane >> 4) holds 8 entries.  Returns false on a non-positive pivot.
__device__ __forceinline__ bool warp_invert_block(double (*d)[17], int n, int lane) {
    const int r = lane & 15, h = lane >> 4;
    bool ok = true;
#pragma unroll 1
    for (int q = 0; q < n; q++) {
        __syncwarp();
        const double p = d[q][q], cq = d[r][q];
        double rq[8], cur[8];
#pragma unroll
        for (int c = 0; c < 8; c++) { rq[c] = d[q][8 * h + c]; cur[c] = d[r][8 * h + c]; }
        __syncwarp();
        ok = ok && (p > 0.0);
        const double ip = 1.0 / p;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const int col = 8 * h + c;
            const double pr = (col == q ? 1.0 : rq[c]) * ip;
            d[r][col] = r == q ? pr : (col == q ? 0.0 : cur[c]) - cq * pr;
        }
    }
    __syncwarp();
    return ok;
}

__device__ __forceinline__ void inv_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void inv_bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
// shared -> global bulk copy by the TMA engine (bytes: multiple of 16, both addresses 16 B aligned), and its completion
__device__ __forceinline__ void inv_bulk_s2g(void *dst, const void *src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(tma::smem_addr(src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void inv_bulk_commit_wait() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");
}
__device__ __forceinline__ void inv_release(unsigned int *flag) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(1u) : "memory");
}

template <int NCOL, int RT>
__global__ void __launch_bounds__(INV_THREADS, 1) coarse_invert2_kernel(int32_t nc, double *__restrict__ A,
                                                                        double *__restrict__ piv /*[n_blk + 1][RT * nc + 256]*/,
                                                                        unsigned int *flags /*[2 * n_blk], zeroed*/,
                                                                        int32_t *status, float *__restrict__ Af) {
    extern __shared__ __align__(128) double rows[];  // [RT][nc]
    __shared__ __align__(16) double cs[16][16];      // this step: D^-1 (owner) / -A_iK D^-1 (others); RT <= 16
    __shared__ __align__(16) double cst[16][17];     // its transpose, for the inverter warp (conflict-free by row)
    __shared__ __align__(16) double dv[16][16], ak[16][16];
    __shared__ __align__(16) double dn[16][17];      // inverter warp: my diagonal block -> its inverse, one step ahead
    __shared__ __align__(16) double pd[16][17];      // ... and the columns of that block in the published rows
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const bool inverter = tid >= INV_CW;             // warp 15
    const int ir = lane & 15, ih = lane >> 4;        // inverter lane: row, half of the columns
    const int r0 = b * RT;
    const int nr = r0 >= nc ? 0 : (r0 + RT <= nc ? RT : nc - r0);
    const int n_blk = (nc + RT - 1) / RT;
    const size_t pstride = (size_t)RT * nc + 256;
    for (int e = tid; e < nr * nc; e += INV_THREADS) {
        const double v = A[(size_t)r0 * nc + e];
        rows[e] = v;
        if (b == 0) piv[e] = v;                      // the first owner's rows are published as they are
    }
    __syncthreads();
    if (b == 0) {                                    // ... and its D^-1 has nobody to prepare it
        if (inverter) {
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const int col = 8 * ih + c;
                dn[ir][col] = (ir < nr && col < nr) ? rows[ir * nc + col] : (ir == col ? 1.0 : 0.0);
            }
            if (!warp_invert_block(dn, nr, lane) && lane == 0) atomicExch(&status[0], PDN_ESINGULAR);
#pragma unroll
            for (int c = 0; c < 8; c++) piv[(size_t)RT * nc + ir * 16 + 8 * ih + c] = dn[ir][8 * ih + c];
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) { inv_release(flags + 0); inv_release(flags + 1); }
    }
    const int di = (tid >> 4) & 15, dj = tid & 15;   // threads 0..255: one per entry of a 16 x 16 block
    const bool blk16 = tid < 256;
#ifdef PDN_INV_TIMING
    long long tw_b = 0, tw_a = 0, t_cs = 0, t_upd = 0, t_own_upd = 0, t_next_upd = 0, tc = clock64();
    __shared__ long long t_invw[2];
#define INV_TICK(acc) do { if (tid == 0) { const long long n_ = clock64(); acc += n_ - tc; tc = n_; } } while (0)
#else
#define INV_TICK(acc) do { } while (0)
#endif
    for (int kb = 0; kb < n_blk; kb++) {
        const int k0 = kb * RT, nk = k0 + RT <= nc ? RT : nc - k0;
        double *pb = piv + (size_t)kb * pstride;
        double *pbn = piv + (size_t)(kb + 1) * pstride;      // publication slot of the next owner (used by it alone)
        const bool owner = kb == b, next_owner = kb + 1 == b;
        // flags[2 kb] = the owner's rows are in pb, flags[2 kb + 1] = its D^-1 is; both were raised during step kb - 1
        if (owner) {
            if (blk16) cs[di][dj] = dn[di][dj];      // (the step ended with a block-wide barrier: dn is complete)
        } else {
            if (!flag_wait(flags + 2 * kb, status)) return;
            INV_TICK(tw_b);
        }
        // one pass per NCOL * INV_BLK columns (uniform trip count: the passes contain barriers)
        for (int base = 0, pass = 0; base < nc; base += NCOL * INV_BLK, pass++) {
            // columns of this thread in this pass
            const int jbase = base + (inverter ? INV_CW * NCOL + lane : tid), jstep = inverter ? 32 : INV_CW;
            double p[NCOL][RT];
#pragma unroll
            for (int t = 0; t < NCOL; t++) {
                const int j = jbase + t * jstep;
#pragma unroll
                for (int q = 0; q < RT; q++)
                    p[t][q] = (j < nc && q < nk) ? (owner ? rows[q * nc + j] : __ldcg(&pb[q * nc + j])) : 0.0;
            }
            if (pass == 0) {
                if (!owner) {
                    // the published rows are on their way into registers; now D^-1 and c = -A_iK D^-1
                    if (!flag_wait(flags + 2 * kb + 1, status)) return;
                    INV_TICK(tw_a);
                    if (blk16) {
                        dv[di][dj] = __ldcg(&pb[(size_t)RT * nc + tid]);
                        ak[di][dj] = (di < nr && dj < nk) ? rows[di * nc + k0 + dj] : 0.0;
                    }
                    if (next_owner && inverter) {
                        // my diagonal block BEFORE this step's update, and its columns of the published rows
#pragma unroll
                        for (int c = 0; c < 8; c++) {
                            const int col = 8 * ih + c;
                            dn[ir][col] = (ir < nr && col < nr) ? rows[ir * nc + r0 + col] : (ir == col ? 1.0 : 0.0);
                            pd[ir][col] = (ir < nk && col < nr) ? __ldcg(&pb[ir * nc + r0 + col]) : 0.0;
                        }
                    }
                    __syncthreads();
                    if (blk16) {
                        double c = 0.0;
#pragma unroll
                        for (int q = 0; q < 16; q++) c += ak[di][q] * dv[q][dj];
                        cs[di][dj] = -c;
                        cst[dj][di] = -c;
                    }
                }
                __syncthreads();
                INV_TICK(t_cs);
            }
            const int ni = owner ? nk : nr;
            for (int i = 0; i < ni; i++) {
                double acc[NCOL];
#pragma unroll
                for (int t = 0; t < NCOL; t++) {
                    const int j = jbase + t * jstep;
                    acc[t] = (owner || j >= nc) ? 0.0 : rows[i * nc + j];
                }
#pragma unroll
                for (int q = 0; q < RT; q += 2) {
                    const double2 c = *reinterpret_cast<const double2 *>(&cs[i][q]);
#pragma unroll
                    for (int t = 0; t < NCOL; t++) {
                        acc[t] = fma(c.x, p[t][q], acc[t]);
                        acc[t] = fma(c.y, p[t][q + 1], acc[t]);
                    }
                }
#pragma unroll
                for (int t = 0; t < NCOL; t++) {
                    const int j = jbase + t * jstep;
                    if (j >= nc) continue;
                    rows[i * nc + j] = acc[t];               // (the pivot columns are overwritten below)
                }
            }
        }
        // Columns done.  In the next owner's CTA warp 15 only ARRIVES and goes on to its D^-1, the other 15 warps
        // finish the rows and raise the rows flag without waiting for it; everywhere else this is a plain barrier.
        if (next_owner && inverter) {
            tma::fence_proxy_async();                // my columns of the rows, for the bulk copy below
            inv_bar_arrive(1, INV_THREADS);
#ifdef PDN_INV_TIMING
            const long long ti0 = clock64();
#endif
            // this step's update of my diagonal block with the update's own fma chain (identical bits) ...
            double val[8];
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const int col = 8 * ih + c;
                double acc = dn[ir][col];
#pragma unroll
                for (int q = 0; q < RT; q += 2) {
                    acc = fma(cst[q][ir], pd[q][col], acc);
                    acc = fma(cst[q + 1][ir], pd[q + 1][col], acc);
                }
                val[c] = acc;
            }
#pragma unroll
            for (int c = 0; c < 8; c++) dn[ir][8 * ih + c] = val[c];
            // ... and its inverse, published under its own flag
            if (!warp_invert_block(dn, nr, lane) && lane == 0) atomicExch(&status[0], PDN_ESINGULAR);
#pragma unroll
            for (int c = 0; c < 8; c++) pbn[(size_t)RT * nc + ir * 16 + 8 * ih + c] = dn[ir][8 * ih + c];
            __threadfence();
            __syncwarp();
            if (lane == 0) inv_release(flags + 2 * (kb + 1) + 1);
#ifdef PDN_INV_TIMING
            if (lane == 0) t_invw[0] = clock64() - ti0;
#endif
        } else {
            inv_bar_sync(1, INV_THREADS);
            // the pivot columns of every row: D^-1 (owner) / -A_iK D^-1 (others), i.e. c itself
            if (blk16 && di < (owner ? nk : nr) && dj < nk) rows[di * nc + k0 + dj] = cs[di][dj];
            if (next_owner) {
                // publish my rows: one bulk copy shared -> global by the TMA engine (the same stores issued from
                // the update loop cost the next owner 6k cycles per step), then the rows flag
                tma::fence_proxy_async();
                inv_bar_sync(2, INV_CW);
                if (tid == 0) {
                    const uint32_t n_ent = (uint32_t)nr * (uint32_t)nc, n_even = n_ent & ~1u;
                    for (uint32_t o = 0; o < n_even; o += 4096u)
                        inv_bulk_s2g(pbn + o, rows + o, (n_even - o < 4096u ? n_even - o : 4096u) * 8u);
                    if (n_ent & 1u) pbn[n_ent - 1] = rows[n_ent - 1];
                    inv_bulk_commit_wait();
                    __threadfence();
                    inv_release(flags + 2 * (kb + 1));
                }
            }
        }
        __syncthreads();
#ifdef PDN_INV_TIMING
        if (owner) INV_TICK(t_own_upd); else if (next_owner) INV_TICK(t_next_upd); else INV_TICK(t_upd);
#endif
    }
#ifdef PDN_INV_TIMING
    if (tid == 0 && (b % 17 == 0 || b == n_blk - 1))
        printf("[inv b=%d of %d] per other-step cycles: wait_rows %lld wait_dinv %lld cs %lld update %lld | as next owner: update+publish %lld (warp 15: block + inverse %lld) | as owner: update %lld\n",
               b, n_blk, tw_b / (n_blk - 1), tw_a / (n_blk - 1), t_cs / (n_blk - 1), t_upd / (n_blk - 2), t_next_upd, t_invw[0], t_own_upd);
#endif
    for (int e = tid; e < nr * nc; e += INV_THREADS) {
        A[(size_t)r0 * nc + e] = rows[e];
        Af[(size_t)r0 * nc + e] = (float)rows[e];
    }
}

// ------------------------------------------------------------------------------------------------
// apply
// ------------------------------------------------------------------------------------------------
// x_c = A_c^-1 r_c  (r_c arrives in fixed point, in slot `slot`, or iteration & 1 if slot < 0, scaled by
// 2^40 / (boost ||r||)); the other slot's
// accumulator is cleared for the next restriction.  CTA b owns CO_RB consecutive rows: every thread
// issues the loads of ALL its matrix entries before anything else (the matrix is L2-resident, the
// kernel is latency-bound: 8.7 us with 8 rows per CTA and loads after the r_c conversion), then one
// block reduction of CO_RB sums in a fixed order.
constexpr int CO_RB = 4;
constexpr int CO_JMAX = (CO_MAX + CO_BLK - 1) / CO_BLK;   // columns per thread
// A second level (the rank-local one of a partitioned solve: nc2, Ainv2, rc_fix2, xc2) rides in the same
// launch: the CTAs from grid1 on work on it.
__global__ void __launch_bounds__(CO_BLK) coarse_apply_kernel(int32_t nc, const float *__restrict__ Ainv, const Scal *__restrict__ sc,
                                                              long long *__restrict__ rc_fix, double *__restrict__ xc,
                                                              int slot, double boost, int grid1, int32_t nc2,
                                                              const float *__restrict__ Ainv2, long long *__restrict__ rc_fix2,
                                                              double *__restrict__ xc2) {
    __shared__ double rcd[CO_MAX];
    __shared__ double red[CO_RB][CO_BLK / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int row0 = blockIdx.x * CO_RB;
    if ((int)blockIdx.x >= grid1) {
        row0 = ((int)blockIdx.x - grid1) * CO_RB;
        nc = nc2; Ainv = Ainv2; rc_fix = rc_fix2; xc = xc2;
    }
    // the matrix entries do not depend on anything this iteration produced: request them first
    float a[CO_RB][CO_JMAX];
#pragma unroll
    for (int q = 0; q < CO_RB; q++)
#pragma unroll
        for (int t = 0; t < CO_JMAX; t++) {
            const int j = tid + t * CO_BLK;
            a[q][t] = (row0 + q < nc && j < nc) ? __ldg(&Ainv[(size_t)(row0 + q) * nc + j]) : 0.f;
        }
    pdl_trigger();
    pdl_wait();                 // the inverse is setup data; everything below depends on the update kernel
    if (sc->done) return;
    const int par = slot < 0 ? (sc->iter & 1) : slot;
    const long long *__restrict__ rc = rc_fix + (size_t)par * nc;
    long long *__restrict__ other = rc_fix + (size_t)(par ^ 1) * nc;
    const double scale = boost * sqrt(fmax(sc->rr, 1e-300)) / CO_RFIX;
    for (int j = tid; j < nc; j += CO_BLK) rcd[j] = (double)rc[j] * scale;
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
        other[row0 + tid] = 0;
    }
}

// z = D^-1 r + P x_c  over the rows [r0, n); two rows per thread (r0 is even, the vectors are 16 B aligned).
// Several GPUs with a small interface (L.world > 1): z lives in the peer arena at word offset vec_off and the
// kernel IS the halo exchange -- the last CTA to finish stores this rank's interface entries (send_idx, peer after
// peer) into the peers' copies, fences ONCE, raises the halo flags and waits for the peers' entries.  (Pushing
// from every CTA cost a system-scope fence per CTA: 18 us instead of 5 for this kernel.)
// BLK: 256 on one GPU; 1024 (two CTAs per SM) when the tail runs, because the ticket is one atomic per CTA on
// one address: 1184 CTAs spent 12 us queueing on it.
template <int BLK>
__global__ void __launch_bounds__(BLK) coarse_z_kernel(int64_t r0, int64_t n, const Scal *__restrict__ sc,
                                                       const double *__restrict__ dinv, const double *__restrict__ r,
                                                       const int32_t *__restrict__ agg, const double *__restrict__ xc,
                                                       double *__restrict__ z, const PeerLink *Lp, const HaloSend *hsp,
                                                       size_t vec_off, const int32_t *__restrict__ send_idx,
                                                       unsigned int *ticket, const int32_t *__restrict__ agg2,
                                                       const double *__restrict__ xc2) {
    pdl_trigger();
    pdl_wait();
    if (sc->done) return;
    if (Lp != nullptr && blockIdx.x == 0 && threadIdx.x == 0) peer_log(*Lp, 7);
    const int64_t npair = (n - r0 + 1) >> 1;
    for (int64_t t = (int64_t)blockIdx.x * BLK + threadIdx.x; t < npair; t += (int64_t)gridDim.x * BLK) {
        const int64_t i = r0 + 2 * t;
        if (i + 1 < n) {
            const double2 dd = *reinterpret_cast<const double2 *>(dinv + i), rr = *reinterpret_cast<const double2 *>(r + i);
            const int2 aa = *reinterpret_cast<const int2 *>(agg + i);
            const double ca = aa.x >= 0 ? __ldg(&xc[aa.x]) : 0.0;
            double cb = aa.y == aa.x ? ca : (aa.y >= 0 ? __ldg(&xc[aa.y]) : 0.0);
            double ca2 = ca;
            if (agg2 != nullptr) {          // rank-local second level (multi-GPU instantiation only)
                const int2 bb = *reinterpret_cast<const int2 *>(agg2 + i);
                const double la = bb.x >= 0 ? __ldg(&xc2[bb.x]) : 0.0;
                ca2 += la;
                cb += bb.y == bb.x ? la : (bb.y >= 0 ? __ldg(&xc2[bb.y]) : 0.0);
            }
            *reinterpret_cast<double2 *>(z + i) = make_double2(dd.x * rr.x + ca2, dd.y * rr.y + cb);
        } else {
            const int32_t a = agg[i];
            double c = a >= 0 ? __ldg(&xc[a]) : 0.0;
            if (agg2 != nullptr) {
                const int32_t b = agg2[i];
                c += b >= 0 ? __ldg(&xc2[b]) : 0.0;
            }
            z[i] = dinv[i] * r[i] + c;
        }
    }
    if (BLK != 1024 || Lp == nullptr) return;    // the tail only exists in the multi-GPU instantiation
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
        // eight entries per thread and round: all indices, then all values, then all remote stores (three
        // latencies per round instead of three per entry)
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
    peer_halo_post(L, hs);      // the reader waits (peer_halo_wait at the start of the SpMV, pdn_peer_prim.cuh)
}

// restriction of one residual entry into the fixed-point accumulator (called by the update kernels,
// whole warps converged).  Lanes with the same aggregate are summed with three redux.sync over 19-bit
// limbs of the biased value (|value| < 2^56), one integer atomic per distinct aggregate in the warp.
__device__ __forceinline__ void restrict_add_narrow(double ri, int32_t a, double fix, unsigned long long *rc) {
    int q = __double2int_rn(ri * fix);             // saturating
    q = max(min(q, 67108863), -67108863);
    const unsigned peers = __match_any_sync(0xffffffffu, a);
    const int sum = __reduce_add_sync(peers, q);
    if ((int)(threadIdx.x & 31) == __ffs(peers) - 1 && a >= 0 && sum != 0)
        atomicAdd(&rc[a], (unsigned long long)(long long)sum);
}
__device__ __forceinline__ void restrict_add(double ri, int32_t a, double fix, unsigned long long *rc);
// two rows per lane (cg_update_kernel): entries clamped to +-(2^25 - 1), equal aggregates are summed in the
// lane, the second collective round only runs when some lane straddles two aggregates
__device__ __forceinline__ void restrict_add_pair(double ra, int32_t a0, double rb, int32_t a1, double fix,
                                                  unsigned long long *rc) {
#if PDN_RESTRICT_WIDE
    restrict_add(ra, a0, fix, rc);
    restrict_add(rb, a1, fix, rc);
#else
    int q0 = max(min(__double2int_rn(ra * fix), 33554431), -33554431);
    int q1 = max(min(__double2int_rn(rb * fix), 33554431), -33554431);
    if (a1 == a0) { q0 += q1; a1 = -1; }
    const int lane = threadIdx.x & 31;
    {
        const unsigned peers = __match_any_sync(0xffffffffu, a0);
        const int sum = __reduce_add_sync(peers, q0);
        if (lane == __ffs(peers) - 1 && a0 >= 0 && sum != 0) atomicAdd(&rc[a0], (unsigned long long)(long long)sum);
    }
    if (__any_sync(0xffffffffu, a1 >= 0)) {
        const unsigned peers = __match_any_sync(0xffffffffu, a1);
        const int sum = __reduce_add_sync(peers, q1);
        if (lane == __ffs(peers) - 1 && a1 >= 0 && sum != 0) atomicAdd(&rc[a1], (unsigned long long)(long long)sum);
    }
#endif
}
__device__ __forceinline__ void restrict_add(double ri, int32_t a, double fix, unsigned long long *rc) {
    const double lim = 72057594037927935.0;   // 2^56 - 1
    double q = ri * fix;
    q = q > lim ? lim : (q < -lim ? -lim : q);
    const unsigned long long u = (unsigned long long)(llrint(q) + (1LL << 56));
    const unsigned peers = __match_any_sync(0xffffffffu, a);
    const unsigned s0 = __reduce_add_sync(peers, (unsigned)(u & 0x7ffffu));
    const unsigned s1 = __reduce_add_sync(peers, (unsigned)((u >> 19) & 0x7ffffu));
    const unsigned s2 = __reduce_add_sync(peers, (unsigned)(u >> 38));
    if ((int)(threadIdx.x & 31) == __ffs(peers) - 1 && a >= 0) {
        const long long sum = (long long)s0 + ((long long)s1 << 19) + ((long long)s2 << 38) - ((long long)__popc(peers) << 56);
        if (sum != 0) atomicAdd(&rc[a], (unsigned long long)sum);
    }
}
