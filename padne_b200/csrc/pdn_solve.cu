// pdn_solve.cu -- the solve: constraint elimination on the host, Krylov iterations on the device.
//
// Replaces solve_system (padne/solver.py:767-780).  The reference hands the whole saddle-point
// system to SuperLU.  Here the full matrix (padne/solver.py:757-760 unknown order) is split as
//
//        [ L_nn  B ] [ v ]   [ r_n ]        n  node voltages (mesh vertices + internal nodes)
//        [ C     0 ] [ x ] = [ e   ]        m  extra unknowns (source currents, ground current)
//
// C holds the voltage-source / ground rows (solver.py:503-511, 526-529, 558-560): each row is
// "v_p - v_n = V" or "v_g = 0".  They are eliminated with a union-find that carries voltage
// offsets: v = Z u + o, where Z maps every node to the representative of its source-connected
// group (nodes tied to ground are fixed and drop out).  B = C^T for plain sources; a
// VoltageRegulator adds +-gain entries (solver.py:537-538), B = C^T + G, which makes the system
// non-symmetric.  With M = C B (m x m, dense, tiny) and Q = I - B M^-1 C the currents are
// x = M^-1 C (r_n - L_nn v) and the reduced system is
//
//        A u = b,   A = -(Z^T Q L_nn Z),   b = -Z^T Q (r_n - L_nn o).
//
// Z^T Q differs from the identity only on the "special" nodes T (constraint terminals and
// regulator sense nodes), so A is stored as L_nn with remapped columns, converted on the device to
// the SELL-32-sigma layout of pdn_sell.cuh (rows sorted by length inside 1024-row windows, the whole
// reduced system permuted symmetrically), plus a tiny override CSR for the rows in T (built on the
// host from the |T| fetched rows).  With several GPUs (pdn_comm_init) every rank builds the same
// operator and owns a contiguous range of sort windows in the Krylov loop: halo exchange of the
// interface entries of the SpMV input, all-reduce of the fused dot products (pdn_multi.cu).  A is SPD when
// B = C^T (Jacobi-PCG, Chronopoulos-Gear single-reduction form) and non-symmetric otherwise
// (Jacobi-BiCGSTAB).  After convergence v = Z u + o, x is recovered, and ||L v - r||_2 of the
// FULL system is reported exactly as solver.py:775.
#include <math.h>

#include <chrono>
#include <string>
#include <algorithm>
#include <map>
#include <unordered_map>

#include "pdn_kernels.cuh"

using namespace pdnk;

namespace {

constexpr int BLK = RPB;

// ------------------------------------------------------------------------------------------------
// setup kernels
// ------------------------------------------------------------------------------------------------
// entries of node rows that sit in extra columns (the B block); columns are sorted, so they are
// at the end of each row
__global__ void __launch_bounds__(BLK) border_kernel(int64_t n, const int32_t *__restrict__ rp,
                                                     const int32_t *__restrict__ ci, const double *__restrict__ va,
                                                     int32_t cap, int32_t *count, int32_t *o_row,
                                                     int32_t *o_col, double *o_val) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
        const int32_t lo = rp[i];
        for (int32_t k = rp[i + 1] - 1; k >= lo && ci[k] >= n; k--) {
            int32_t slot = atomicAdd(count, 1);
            if (slot < cap) { o_row[slot] = (int32_t)i; o_col[slot] = ci[k]; o_val[slot] = va[k]; }
        }
    }
}

__global__ void __launch_bounds__(BLK) iota_kernel(int64_t n, int32_t *map) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) map[i] = (int32_t)i;
}

__global__ void __launch_bounds__(BLK) scatter_i32_kernel(int32_t k, const int32_t *__restrict__ idx,
                                                          const int32_t *__restrict__ val, int32_t *dst) {
    for (int32_t i = blockIdx.x * BLK + threadIdx.x; i < k; i += gridDim.x * BLK) dst[idx[i]] = val[i];
}
__global__ void __launch_bounds__(BLK) scatter_f64_kernel(int32_t k, const int32_t *__restrict__ idx,
                                                          const double *__restrict__ val, double *dst) {
    for (int32_t i = blockIdx.x * BLK + threadIdx.x; i < k; i += gridDim.x * BLK) dst[idx[i]] = val[i];
}
__global__ void __launch_bounds__(BLK) gather_f64_kernel(int32_t k, const int32_t *__restrict__ idx,
                                                         const double *__restrict__ src, double *out) {
    for (int32_t i = blockIdx.x * BLK + threadIdx.x; i < k; i += gridDim.x * BLK) out[i] = src[idx[i]];
}
// row lengths (node columns only) of the rows in T, then their entries
__global__ void __launch_bounds__(BLK) rowlen_kernel(int32_t k, int64_t n, const int32_t *__restrict__ rows,
                                                     const int32_t *__restrict__ rp, const int32_t *__restrict__ ci,
                                                     int32_t *len) {
    for (int32_t t = blockIdx.x * BLK + threadIdx.x; t < k; t += gridDim.x * BLK) {
        const int32_t i = rows[t];
        int32_t c = 0;
        for (int32_t q = rp[i]; q < rp[i + 1]; q++) c += ci[q] < n;
        len[t] = c;
    }
}
__global__ void __launch_bounds__(BLK) rowget_kernel(int32_t k, int64_t n, const int32_t *__restrict__ rows,
                                                     const int32_t *__restrict__ rp, const int32_t *__restrict__ ci,
                                                     const double *__restrict__ va, const int32_t *__restrict__ optr,
                                                     int32_t *o_ci, double *o_va) {
    for (int32_t t = blockIdx.x * BLK + threadIdx.x; t < k; t += gridDim.x * BLK) {
        const int32_t i = rows[t];
        int32_t o = optr[t];
        for (int32_t q = rp[i]; q < rp[i + 1]; q++)
            if (ci[q] < n) { o_ci[o] = ci[q]; o_va[o] = va[q]; o++; }
    }
}

__global__ void __launch_bounds__(BLK) check_dinv_kernel(int64_t n, const double *__restrict__ dinv, int32_t *status) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
        const double d = dinv[i];
        if (!(d > 0.0) || !(d < 1e300)) {
            if (atomicCAS(&status[0], 0, PDN_ESINGULAR) == 0) status[1] = (int32_t)i;
        }
    }
}

// y[i] += x[i] (first == 0) or y[i] = x[i] (first != 0) over [r0, r1)
__global__ void __launch_bounds__(BLK) accumulate_kernel(int64_t r0, int64_t r1, const double *__restrict__ x,
                                                         double *__restrict__ y, int first) {
    for (int64_t i = r0 + (int64_t)blockIdx.x * BLK + threadIdx.x; i < r1; i += (int64_t)gridDim.x * BLK)
        y[i] = first ? x[i] : y[i] + x[i];
}

__global__ void __launch_bounds__(BLK) ones_kernel(int64_t n, double *__restrict__ d) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) d[i] = 1.0;
}

// v_j = u[inv[map_j]] + o_j for nodes (u lives in the permuted numbering); extras start at 0
__global__ void __launch_bounds__(BLK) expand_kernel(int64_t n, int64_t N, const int32_t *__restrict__ map,
                                                     const int32_t *__restrict__ inv,
                                                     const double *__restrict__ o, const double *__restrict__ u,
                                                     double *__restrict__ v) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < N; i += (int64_t)gridDim.x * BLK) {
        if (i < n) {
            const int32_t m = map[i];
            v[i] = (m >= 0 ? u[inv[m]] : 0.0) + o[i];
        } else {
            v[i] = 0.0;
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// per-context solve workspace
// ------------------------------------------------------------------------------------------------
// halo exchange plan of one rank (multi-GPU): which entries of a distributed vector it sends to /
// receives from every peer, in ascending index order on both sides
struct HaloPlan {
    int world = 1, me = 0;
    std::vector<int64_t> begin;                 // [world+1] first owned row of every rank
    std::vector<int32_t> send_off, send_cnt, recv_off, recv_cnt;   // [world]
    int32_t n_send = 0, n_recv = 0;
    int32_t n_interface = 0;                    // rows read by another rank, over ALL ranks (identical everywhere)
    DevBuf<int32_t> send_idx, recv_idx;
    DevBuf<double> sendbuf, recvbuf;
    HaloSend peer{};                            // the same plan for the peer-memory path (pdn_peer.cuh)
    DevBuf<HaloSend> peer_d;                    // ... in device memory (coarse_z_kernel's tail)
};

struct SolveWork {
    // reduced operator (SELL-32-sigma, permuted numbering)
    int64_t n = 0, N = 0, n_pad = 0, n_slices = 0;
    DevBuf<int32_t> map, len, perm, inv, scnt, soff, s_ci, sp_row, sp_row_old, sp_rp, sp_ci, tmp_i32a, tmp_i32b, tmp_i32c;
    DevBuf<int32_t> t_ptr_d, t_ci_d, halo_col, halo_mask;
    DevBuf<double> t_va_d;
    DevBuf<uint32_t> smask, needmask;
    DevBuf<uint8_t> special;
    DevBuf<double> s_va, dinv, sp_va, off, tmp_f64a, tmp_f64b;
    int32_t n_special = 0;
    // two-level preconditioner (pdn_coarse.cuh)
    bool two_level = false;
    bool pdl = false;          // programmatic dependent launches inside the captured PCG iterations (one GPU)
    int32_t nc = 0;
    float coarse_ms = 0.f;
    DevBuf<int32_t> agg, co_used, co_cid;
    DevBuf<long long> co_acc, rc_fix;
    DevBuf<double> co_A, co_piv, xc, co_tmp;
    DevBuf<float> co_Af;
    // rank-local second coarse level of a partitioned PCG solve (aggregates over this rank's rows only, finer
    // bins, block-diagonal: couplings to other ranks' rows are dropped; no exchange)
    DevBuf<int32_t> agg2;
    DevBuf<long long> rc_fix2;
    DevBuf<double> co_A2, xc2;
    DevBuf<float> co_Af2;
    int32_t nc2 = 0;
    DevBuf<unsigned int> co_counter;
    std::string coarse_note;   // why the last build_coarse returned false ("" when it built)
    int local_level = 0;       // 0: not requested, 1: built, < 0: requested but not built (-PDN_E* / -100: see build_coarse)
    // round-2 PCG iteration (pdn_cg2.cuh): restriction tables of K1, coarse vectors kept by recurrence
    bool v2 = false;           // this solve runs the three-kernel iteration
    bool part = false;         // several GPUs, partitioned assembly: the resident L holds this rank's rows (+ terminals, extras)
    bool halo_fused = false;   // v2, several GPUs: K3's last CTA pushes the halo of z and K1 waits for the peers'
    bool xs = true;            // SpMV with x-band staging (else the round-1 TMA-ring kernel)
    XsPlan xsp{0, 1 << 28};
    int xs_per_sm = 0;
    double xs_coverage = 0.0;  // fraction of this rank's stored entries whose column falls inside the staged window
    DevBuf<unsigned long long> tab, tab2, band_hist;
    DevBuf<int32_t> agg_range;             // [2] first / last aggregate of this rank's rows (exchange payload)
    DevBuf<int32_t> cta_amin, cta_amin2;   // first aggregate of every SpMV CTA's restriction window, per level
    DevBuf<double> rc_d, cs_d, rc2_d, cs2_d, cw_d;   // [2][nc] slots by iteration parity; cw: rank-summed c_w
    double amax = 0.0;         // largest diagonal of the reduced operator
    // vectors
    DevBuf<double> b, x, r, z, p, s, w, rhat, vv, sh, ph, tt, full_a, full_b, rhs_dev;
    DevBuf<double> x_acc, b_corr;   // refinement-style restarts: solution so far, right-hand side of the correction
    DevBuf<double> partials;
    Scal *d_sc = nullptr;
    Scal *h_sc = nullptr;  // pinned
    double *d_out = nullptr;
    bool have_operator = false;
    SellOp A{};
    HaloPlan halo;
    int64_t row_begin = 0, row_end = 0;   // rows of the Krylov loop this rank owns
    ~SolveWork() {
        if (d_sc) cudaFree(d_sc);
        if (h_sc) cudaFreeHost(h_sc);
        if (d_out) cudaFree(d_out);
    }
};

void pdn_solve_release(pdn_ctx *ctx) {
    delete ctx->solve;
    ctx->solve = nullptr;
}

static SolveWork *get_work(pdn_ctx *ctx) {
    if (!ctx->solve) {
        SolveWork *w = new SolveWork();
        ctx->solve = w;
        PDN_CUDA(cudaMalloc((void **)&w->d_sc, sizeof(Scal)));
        PDN_CUDA(cudaMallocHost((void **)&w->h_sc, sizeof(Scal)));
        PDN_CUDA(cudaMalloc((void **)&w->d_out, 8 * sizeof(double)));
        w->partials.reserve(3 * MAX_PART);
        PDN_CUDA(cudaMemsetAsync(w->d_sc, 0, sizeof(Scal), ctx->stream));
    }
    return ctx->solve;
}

// ---- CSR-stream kernel (pdn_spmv.cuh): the FULL system (setup, recovery, diagnostics) -----------------
static int spmv_grid(const pdn_ctx *ctx, int64_t n) {
    int64_t nb = (n + SROWS - 1) / SROWS;
    int per_sm = (int)((227 * 1024) / (SPMV_SMEM + 1024));
    if (per_sm > 8) per_sm = 8;
    int64_t full = (int64_t)ctx->n_sm * per_sm;
    if (full > MAX_PART) full = MAX_PART;
    if (nb < 1) nb = 1;
    return (int)(nb < full ? nb : full);
}

static Operator full_operator(const pdn_ctx *ctx, int64_t rows) {
    Operator A{};
    A.n = rows;
    A.rp = ctx->L.indptr.p;
    A.ci = ctx->L.indices.p;
    A.va = ctx->L.data.p;
    A.n_special = 0;
    return A;
}

template <int MODE>
static void launch_spmv(pdn_ctx *ctx, const Operator &A, const double *x, double *y, const double *aux,
                        SolveWork *w) {
    static bool configured = false;
    if (!configured) {
        PDN_CUDA(cudaFuncSetAttribute(spmv_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SPMV_SMEM));
        configured = true;
    }
    PDN_LAUNCH(ctx, spmv_kernel<MODE>, spmv_grid(ctx, A.n), STHREADS, SPMV_SMEM, A, x, y, aux, w->d_sc, w->partials.p);
}

// ---- SELL kernels (pdn_sell.cuh): the reduced operator of the Krylov loop ------------------------------
#ifndef PDN_SELL_TMA
#define PDN_SELL_TMA 1     // 1: per-warp TMA ring kernel, 0: direct global-load kernel
#endif

static bool is_dist(const pdn_ctx *ctx) { return pdn_comm_world(ctx) > 1; }

// the peer-memory link of this rank in device memory, or null (single GPU / NCCL fallback)
static const PeerLink *peer_link_or_none(pdn_ctx *ctx) {
    if (is_dist(ctx) && pdn_peer_enabled(ctx)) return pdn_peer_link_dev(ctx);
    return nullptr;
}

// CTAs of the TMA-ring kernel: resident CTAs of THIS instantiation (CG2 carries its restriction windows in shared
// memory), capped by the work
template <int MODE, bool DIST>
static int ring_grid(pdn_ctx *ctx, int64_t slices) {
#if PDN_SELL_TMA
    constexpr size_t smem = ST_SMEM + (MODE == CG2 ? TAB_SMEM : 0);
    static int per_sm = 0;
    if (!per_sm) {
        PDN_CUDA(cudaFuncSetAttribute(sell_tma_kernel<MODE, DIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PDN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sell_tma_kernel<MODE, DIST>, ST_BLK, smem));
        if (per_sm < 1) per_sm = 1;
    }
    int64_t grid = std::min<int64_t>((int64_t)ctx->n_sm * per_sm, MAX_PART);
    grid = std::min<int64_t>(grid, std::max<int64_t>(1, (slices + ST_WARPS - 1) / ST_WARPS));
    return (int)grid;
#else
    return 1;
#endif
}

template <int MODE, bool DIST>
static void launch_sell_kernel(pdn_ctx *ctx, SolveWork *w, const double *x, double *y, const double *aux, const Cg2 &cg = Cg2{}) {
    const int64_t slices = (w->A.row_end - w->A.row_begin + SELL_C - 1) / SELL_C;
#if PDN_SELL_TMA
    constexpr size_t smem = ST_SMEM + (MODE == CG2 ? TAB_SMEM : 0);
    const int grid = ring_grid<MODE, DIST>(ctx, slices);
    PDN_LAUNCH_PDL(ctx, w->pdl, (sell_tma_kernel<MODE, DIST>), grid, ST_BLK, smem, w->A, x, y, aux,
                   w->d_sc, w->partials.p, peer_link_or_none(ctx), cg);
#else
    static_assert(MODE != CG2, "the direct-load kernel has no CG2 mode");
    auto kern = sell_spmv_kernel<MODE, DIST>;
    static int per_sm = 0;
    if (!per_sm) {
        PDN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, SELL_BLK, 0));
        if (per_sm < 1) per_sm = 1;
    }
    int64_t grid = std::min<int64_t>((int64_t)ctx->n_sm * per_sm, MAX_PART);
    grid = std::min<int64_t>(grid, std::max<int64_t>(1, (slices + SELL_WARPS - 1) / SELL_WARPS));
    PDN_LAUNCH(ctx, kern, (int)grid, SELL_BLK, 0, w->A, x, y, aux, w->d_sc, w->partials.p, peer_link_or_none(ctx));
#endif
}

// ---- SELL SpMV with x-band staging (pdn_cg2.cuh) -------------------------------------------------------------
static Cg2 make_cg2(SolveWork *w) {
    Cg2 cg{};
    cg.agg = w->agg.p;
    cg.agg2 = w->nc2 > 0 ? w->agg2.p : nullptr;
    cg.tab = w->tab.p;
    cg.tab2 = w->tab2.p;
    cg.rc = w->rc_d.p;
    cg.cs = w->cs_d.p;
    cg.rc2 = w->rc2_d.p;
    cg.cs2 = w->cs2_d.p;
    cg.cw = w->cw_d.p;
    cg.nc = w->nc;
    cg.nc2 = w->nc2;
    cg.amin = w->cta_amin.p;
    cg.amin2 = w->cta_amin2.p;
    cg.halo = w->halo_fused ? w->halo.peer_d.p : nullptr;
    cg.arange = w->agg_range.p;
    return cg;
}

// CTAs of the x-staging kernel over `slices` slices: resident CTAs of the CG2 instantiation (every mode is launched
// with the same grid, so that the staging plan fits them all)
static int xs_grid(pdn_ctx *ctx, SolveWork *w, int64_t slices) {
    if (!w->xs_per_sm) {
        PDN_CUDA(cudaFuncSetAttribute(sell_xs_kernel<CG2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SX_SMEM));
        int per_sm = 0;
        PDN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sell_xs_kernel<CG2, false>, SX_BLK, SX_SMEM));
        w->xs_per_sm = per_sm < 1 ? 1 : per_sm;
    }
    int64_t grid = std::min<int64_t>((int64_t)ctx->n_sm * w->xs_per_sm, MAX_PART);
    grid = std::min<int64_t>(grid, std::max<int64_t>(1, (slices + SX_WARPS - 1) / SX_WARPS));
    return (int)grid;
}

template <int MODE, bool DIST>
static void launch_xs_kernel(pdn_ctx *ctx, SolveWork *w, const double *x, double *y, const double *aux, const Cg2 &cg) {
    const int64_t slices = (w->A.row_end - w->A.row_begin + SELL_C - 1) / SELL_C;
    auto kern = sell_xs_kernel<MODE, DIST>;
    static bool configured = false;
    if (!configured) {
        PDN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SX_SMEM));
        configured = true;
    }
    PDN_LAUNCH_PDL(ctx, w->pdl, kern, xs_grid(ctx, w, slices), SX_BLK, SX_SMEM, w->A, w->xsp, x, y, aux, w->d_sc, w->partials.p,
                   peer_link_or_none(ctx), cg);
}

// The staging plan of this rank's rows: band histogram of |col - row| on the device, then the smallest halo
// (a multiple of 64 rows) whose window holds >= 99.5 % of the stored entries, the range of a CTA in ONE chunk
// when rows + 2 halo fit the buffer, else in double-buffered chunks.  Out-of-window columns take the global path,
// so any plan is correct; a mesh without locality (coverage low at every halo) just gathers through L2 as before.
static void plan_xs(pdn_ctx *ctx, SolveWork *w) {
    cudaStream_t st = ctx->stream;
    const int64_t s0 = w->A.row_begin / SELL_C, s1 = (w->A.row_end + SELL_C - 1) / SELL_C, slices = s1 - s0;
    w->xsp = XsPlan{0, 1 << 28};
    w->xs_coverage = 0.0;
    if (slices <= 0) return;
    const int grid = xs_grid(ctx, w, slices);
    const int64_t spc = (slices + grid - 1) / grid;          // slices of the longest CTA range
    w->band_hist.reserve(64);
    PDN_CUDA(cudaMemsetAsync(w->band_hist.p, 0, 64 * sizeof(unsigned long long), st));
    PDN_LAUNCH(ctx, band_hist_kernel, pdn_grid(ctx, slices * 32, 256, 8), 256, 0, w->A, s0, s1, w->band_hist.p);
    unsigned long long hist[64];
    PDN_CUDA(cudaMemcpyAsync(hist, w->band_hist.p, sizeof hist, cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    double total = 0.0;
    for (int b = 0; b < 64; b++) total += (double)hist[b];
    if (!(total > 0.0)) return;
    // largest halo the double-buffered layout can hold next to a useful chunk (>= 1024 rows)
    const int halo_cap = std::min(63 * 64, ((SX_CAP / 2 - 1024) / 2) & ~63);
    int halo = halo_cap;
    double cum = 0.0, cov = 0.0;
    for (int b = 0; b < 63; b++) {
        cum += (double)hist[b];
        const int h = 64 * (b + 1);
        if (h > halo_cap) break;
        cov = cum / total;
        if (cov >= 0.995) { halo = h; break; }
    }
    if (const char *e = getenv("PADNE_B200_XS_HALO")) {       // tuning: force the halo (rows, rounded to 64)
        halo = std::max(0, std::min(halo_cap, atoi(e) / 64 * 64));
        cum = 0.0;
        for (int b = 0; b < halo / 64; b++) cum += (double)hist[b];
        cov = cum / total;
    }
    w->xs_coverage = cov;
    w->xsp.halo = halo;
    if (spc * SELL_C + 2 * halo <= SX_CAP) w->xsp.chunk_slices = (int32_t)std::max<int64_t>(spc, 1);
    else w->xsp.chunk_slices = (SX_CAP / 2 - 2 * halo) / SELL_C;
}

template <int WHICH>
static void dist_reduce(pdn_ctx *ctx, SolveWork *w);

// y = A x (+ fused dots) over the rows this rank owns.  Multi-GPU: the partial sums are all-reduced and
// the scalar recurrence runs in scalar_kernel; x must already hold the halo entries.
template <int MODE>
static void launch_sell(pdn_ctx *ctx, SolveWork *w, const double *x, double *y, const double *aux) {
    Cg2 no_cg{};
    // several GPUs: the vector's interface entries were pushed by the tail of the kernel that wrote it (or by the
    // explicit exchange kernels, whose wait has then already happened): the SpMV waits for the peers' flags
    no_cg.halo = w->halo_fused ? w->halo.peer_d.p : nullptr;
    if (is_dist(ctx)) {
        if (w->xs) launch_xs_kernel<MODE, true>(ctx, w, x, y, aux, no_cg);
        else launch_sell_kernel<MODE, true>(ctx, w, x, y, aux);
        // peer memory: the kernel's last CTA has all-reduced the dots and run the recurrence itself
        if (MODE != PLAIN && !pdn_peer_enabled(ctx)) dist_reduce<MODE>(ctx, w);
    } else {
        if (w->xs) launch_xs_kernel<MODE, false>(ctx, w, x, y, aux, no_cg);
        else launch_sell_kernel<MODE, false>(ctx, w, x, y, aux);
    }
}

// refresh the halo entries of a distributed vector (no-op on a single GPU)
__global__ void __launch_bounds__(256) halo_pack_kernel(int32_t k, const int32_t *__restrict__ idx,
                                                        const double *__restrict__ vec, double *__restrict__ buf) {
    for (int32_t i = blockIdx.x * 256 + threadIdx.x; i < k; i += gridDim.x * 256) buf[i] = vec[idx[i]];
}
__global__ void __launch_bounds__(256) halo_unpack_kernel(int32_t k, const int32_t *__restrict__ idx,
                                                          const double *__restrict__ buf, double *__restrict__ vec) {
    for (int32_t i = blockIdx.x * 256 + threadIdx.x; i < k; i += gridDim.x * 256) vec[idx[i]] = buf[i];
}

static void halo_exchange(pdn_ctx *ctx, SolveWork *w, double *vec) {
    if (!is_dist(ctx)) return;
    HaloPlan &h = w->halo;
    if (h.n_send == 0 && h.n_recv == 0) return;
    if (pdn_peer_enabled(ctx)) {
        // the vector lives in the arena: owners store their interface entries into the peers' copies
        int which = -1;
        for (int v = 0; v < PDN_PA_NVEC; v++)
            if (vec == pdn_peer_vec(ctx, v)) which = v;
        if (which < 0) PDN_FAIL(PDN_ECUDA, "halo_exchange: vector is not inside the peer arena");
        const PeerLink L = pdn_peer_link(ctx);
        if (h.n_send > 0)
            PDN_LAUNCH(ctx, halo_push_kernel, std::min(ctx->n_sm, pdn_blocks(h.n_send, 256)), 256, 0, L, h.peer,
                       pdn_peer_vec_offset(ctx, which), h.send_idx.p, pdn_peer_ticket(ctx));
        if (h.n_recv > 0) PDN_LAUNCH(ctx, halo_wait_kernel, 1, 32, 0, L, h.peer);
        return;
    }
    if (h.n_send > 0)
        PDN_LAUNCH(ctx, halo_pack_kernel, pdn_grid(ctx, h.n_send, 256, 4), 256, 0, h.n_send, h.send_idx.p, vec, h.sendbuf.p);
    pdn_comm_exchange(ctx, h.sendbuf.p, h.send_off.data(), h.send_cnt.data(), h.recvbuf.p, h.recv_off.data(),
                      h.recv_cnt.data());
    if (h.n_recv > 0)
        PDN_LAUNCH(ctx, halo_unpack_kernel, pdn_grid(ctx, h.n_recv, 256, 4), 256, 0, h.n_recv, h.recv_idx.p, h.recvbuf.p, vec);
}

// sum sc->part[0..2] over the ranks and run recurrence WHICH (scalar_kernel numbering)
template <int WHICH>
static void dist_reduce(pdn_ctx *ctx, SolveWork *w) {
    if (pdn_peer_enabled(ctx)) {
        PDN_LAUNCH(ctx, peer_reduce_kernel<WHICH>, 1, 32, 0, pdn_peer_link(ctx), w->d_sc, w->d_out);
    } else {
        pdn_comm_allreduce_sum(ctx, w->d_sc->part, 3);
        PDN_LAUNCH(ctx, scalar_kernel<WHICH>, 1, 1, 0, w->d_sc, w->d_out);
    }
}

static void fetch_scal(pdn_ctx *ctx, SolveWork *w) {
    PDN_CUDA(cudaMemcpyAsync(w->h_sc, w->d_sc, sizeof(Scal), cudaMemcpyDeviceToHost, ctx->stream));
    PDN_CUDA(cudaStreamSynchronize(ctx->stream));
}

// ------------------------------------------------------------------------------------------------
// host side: constraint elimination
// ------------------------------------------------------------------------------------------------
namespace {

struct Term { int loc; double coef; };

struct Elim {
    int64_t n = 0, m = 0;
    std::vector<int32_t> T;                    // special node ids (sorted)
    std::unordered_map<int32_t, int> tloc;     // node id -> local index
    std::vector<std::vector<Term>> C;          // per constraint row: terms over local indices
    std::vector<double> e;                     // constraint right-hand sides
    std::vector<std::vector<std::pair<int, double>>> Bn;  // per local node: (extra index, value)
    std::vector<int> par;                      // union-find over locals + datum (index nT)
    std::vector<double> pot;                   // v_x - v_parent
    std::vector<int32_t> rep;                  // per local: representative NODE id, -1 if fixed
    std::vector<double> off;                   // per local: offset to representative / fixed value
    std::vector<double> Minv;                  // m x m
    std::vector<double> F;                     // nT x nT  (Z^T Q restricted to T)
    bool symmetric = true;

    int local(int32_t node) {
        auto it = tloc.find(node);
        if (it != tloc.end()) return it->second;
        int l = (int)T.size();
        tloc.emplace(node, l);
        T.push_back(node);
        return l;
    }
    int find(int x) {
        // iterative find with potential accumulation
        std::vector<int> path;
        while (par[x] != x) { path.push_back(x); x = par[x]; }
        // compress: walk back from the node nearest the root
        for (int k = (int)path.size() - 1; k >= 0; k--) {
            int y = path[k];
            if (par[y] != x) { pot[y] += pot[par[y]]; par[y] = x; }
        }
        return x;
    }
    // v_p - v_q = d ; false on a cycle
    bool unite(int p, int q, double d) {
        int rp = find(p), rq = find(q);
        if (rp == rq) return false;
        par[rp] = rq;
        pot[rp] = d + pot[q] - pot[p];
        return true;
    }
};

// dense inverse by Gauss-Jordan with partial pivoting; false if singular
bool invert_dense(std::vector<double> &a, int m, std::vector<double> &inv) {
    inv.assign((size_t)m * m, 0.0);
    for (int i = 0; i < m; i++) inv[(size_t)i * m + i] = 1.0;
    for (int c = 0; c < m; c++) {
        int piv = c;
        double best = fabs(a[(size_t)c * m + c]);
        for (int r = c + 1; r < m; r++)
            if (fabs(a[(size_t)r * m + c]) > best) { best = fabs(a[(size_t)r * m + c]); piv = r; }
        if (!(best > 1e-300)) return false;
        if (piv != c)
            for (int k = 0; k < m; k++) {
                std::swap(a[(size_t)c * m + k], a[(size_t)piv * m + k]);
                std::swap(inv[(size_t)c * m + k], inv[(size_t)piv * m + k]);
            }
        const double d = 1.0 / a[(size_t)c * m + c];
        for (int k = 0; k < m; k++) { a[(size_t)c * m + k] *= d; inv[(size_t)c * m + k] *= d; }
        for (int r = 0; r < m; r++) {
            if (r == c) continue;
            const double f = a[(size_t)r * m + c];
            if (f == 0.0) continue;
            for (int k = 0; k < m; k++) {
                a[(size_t)r * m + k] -= f * a[(size_t)c * m + k];
                inv[(size_t)r * m + k] -= f * inv[(size_t)c * m + k];
            }
        }
    }
    return true;
}

// Build the elimination from the border blocks.
//   c_ptr/c_col/c_val : CSR of the extra rows n..N-1 (all columns)
//   b_row/b_col/b_val : entries of node rows in extra columns
void build_elim(Elim &E, int64_t n, int64_t N, const std::vector<int32_t> &c_ptr,
                const std::vector<int32_t> &c_col, const std::vector<double> &c_val,
                const std::vector<int32_t> &b_row, const std::vector<int32_t> &b_col,
                const std::vector<double> &b_val, const double *rhs_tail) {
    const int m = (int)(N - n);
    E.n = n;
    E.m = m;
    E.C.assign(m, {});
    E.e.assign(rhs_tail, rhs_tail + m);
    for (int r = 0; r < m; r++) {
        for (int32_t k = c_ptr[r] - c_ptr[0]; k < c_ptr[r + 1] - c_ptr[0]; k++) {
            if (c_col[k] >= n) PDN_FAIL(PDN_EUNSUPPORTED, "solve: extra-extra coupling at row %lld", (long long)(n + r));
            E.C[r].push_back({E.local(c_col[k]), c_val[k]});
        }
        if (E.C[r].empty()) PDN_FAIL(PDN_ESINGULAR, "solve: empty constraint row %lld", (long long)(n + r));
        if (E.C[r].size() > 2 || (E.C[r].size() == 2 && E.C[r][0].coef != -E.C[r][1].coef))
            PDN_FAIL(PDN_EUNSUPPORTED, "solve: constraint row %lld is not a voltage difference", (long long)(n + r));
    }
    // B entries sorted by (row, col) for determinism
    std::vector<size_t> order(b_row.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](size_t a, size_t b) {
        return b_row[a] != b_row[b] ? b_row[a] < b_row[b] : b_col[a] < b_col[b];
    });
    for (size_t i : order) E.local(b_row[i]);
    const int nT = (int)E.T.size();
    E.Bn.assign(nT, {});
    for (size_t i : order) E.Bn[E.tloc[b_row[i]]].push_back({(int)(b_col[i] - n), b_val[i]});

    // union-find with potentials; the datum (0 V) is local index nT
    E.par.resize(nT + 1);
    E.pot.assign(nT + 1, 0.0);
    for (int i = 0; i <= nT; i++) E.par[i] = i;
    for (int r = 0; r < m; r++) {
        const auto &row = E.C[r];
        bool ok = row.size() == 1 ? E.unite(row[0].loc, nT, E.e[r] / row[0].coef)
                                  : E.unite(row[0].loc, row[1].loc, E.e[r] / row[0].coef);
        if (!ok) PDN_FAIL(PDN_ESINGULAR, "solve: voltage-source loop at constraint row %lld", (long long)(n + r));
    }
    E.rep.assign(nT, -1);
    E.off.assign(nT, 0.0);
    const int root_datum = E.find(nT);
    const double pot_datum = E.pot[nT];  // relative to root_datum (0 if the datum is the root)
    for (int l = 0; l < nT; l++) {
        const int r = E.find(l);
        if (r == root_datum) {
            E.rep[l] = -1;
            E.off[l] = E.pot[l] - (nT == root_datum ? 0.0 : pot_datum);
        } else {
            E.rep[l] = E.T[r];
            E.off[l] = (l == r) ? 0.0 : E.pot[l];
        }
    }

    // symmetric <=> B == C^T
    size_t n_c = 0;
    E.symmetric = true;
    for (int r = 0; r < m && E.symmetric; r++)
        for (const Term &t : E.C[r]) {
            n_c++;
            bool found = false;
            for (auto &bv : E.Bn[t.loc])
                if (bv.first == r && bv.second == t.coef) { found = true; break; }
            if (!found) { E.symmetric = false; break; }
        }
    if (E.symmetric && n_c != b_row.size()) E.symmetric = false;

    // M = C B and its inverse
    if (m > 4096) PDN_FAIL(PDN_EUNSUPPORTED, "solve: %d extra unknowns exceed the dense elimination limit", m);
    std::vector<double> M((size_t)m * m, 0.0);
    for (int r = 0; r < m; r++)
        for (const Term &t : E.C[r])
            for (auto &bv : E.Bn[t.loc]) M[(size_t)r * m + bv.first] += t.coef * bv.second;
    if (!invert_dense(M, m, E.Minv)) PDN_FAIL(PDN_ESINGULAR, "solve: constraint block C*B is singular");

    // F = (Z^T Q)[T, T]
    if (nT > 8192) PDN_FAIL(PDN_EUNSUPPORTED, "solve: %d constraint terminals exceed the dense limit", nT);
    E.F.assign((size_t)nT * nT, 0.0);
    std::vector<double> QT;
    if (!E.symmetric) {
        QT.assign((size_t)nT * nT, 0.0);
        std::vector<double> BM(m);
        for (int j = 0; j < nT; j++) {
            QT[(size_t)j * nT + j] = 1.0;
            if (E.Bn[j].empty()) continue;
            std::fill(BM.begin(), BM.end(), 0.0);
            for (auto &bv : E.Bn[j])
                for (int c = 0; c < m; c++) BM[c] += bv.second * E.Minv[(size_t)bv.first * m + c];
            for (int r = 0; r < m; r++)
                for (const Term &t : E.C[r]) QT[(size_t)j * nT + t.loc] -= BM[r] * t.coef;
        }
    }
    for (int j = 0; j < nT; j++) {
        if (E.rep[j] < 0) continue;
        const int lt = E.tloc[E.rep[j]];
        if (E.symmetric) E.F[(size_t)lt * nT + j] += 1.0;
        else
            for (int k = 0; k < nT; k++) E.F[(size_t)lt * nT + k] += QT[(size_t)j * nT + k];
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// Krylov drivers
// ------------------------------------------------------------------------------------------------
namespace {

struct KrylovResult {
    int iterations = 0, restarts = 0, stagnated = 0;
    double rel_residual = 0.0;
    double bnorm = 0.0;          // ||b||_2 of the reduced system
    double abs_target = 0.0;     // the absolute bar the loop finally ran to
    double floor_abs = 0.0;      // 4 u amax ||x||: what one fp64 residual evaluation can resolve
    float ms = 0.f;
};

void reset_scal(pdn_ctx *ctx, SolveWork *w, double tol2, double bnorm2, int max_iter) {
    Scal h{};
    h.tol2 = tol2;
    h.bnorm2 = bnorm2;
    h.rr = bnorm2;     // r = b at the start: the scale of the first fixed-point restriction
    h.first = 1;
    h.max_iter = max_iter;
    h.omega = 1.0;
    h.alpha = 1.0;
    h.gamma_old = 1.0;
    h.amax = w->amax;
    *w->h_sc = h;
    PDN_CUDA(cudaMemcpyAsync(w->d_sc, w->h_sc, sizeof(Scal), cudaMemcpyHostToDevice, ctx->stream));
}

// keep the iteration counter, restart the recurrences (tol2: the threshold from now on)
void restart_scal(pdn_ctx *ctx, SolveWork *w, double tol2) {
    Scal h = *w->h_sc;
    h.tol2 = tol2;
    h.first = 1;
    h.done = 0;
    h.breakdown = 0;
    h.omega = 1.0;
    h.alpha = 1.0;
    h.gamma_old = 1.0;
    memset(h.ticket, 0, sizeof h.ticket);
    *w->h_sc = h;
    PDN_CUDA(cudaMemcpyAsync(w->d_sc, w->h_sc, sizeof(Scal), cudaMemcpyHostToDevice, ctx->stream));
}

// one iteration = SpMV with fused dots (+ halo exchange and all-reduce on several GPUs) + vector update
// two-level: x_c = A_c^-1 r_c, z = D^-1 r + P x_c  (r_c was accumulated by the update / init kernel)
// slot: which half of rc_fix holds the restriction (-1: iteration parity, PCG); src / dst: the vector that
// was restricted and the preconditioned one
// true when coarse_z_kernel also performs the halo exchange of the vector it writes: peer memory and an
// interface small enough for one CTA to push (a via-heavy board keeps the multi-CTA push kernel)
static bool halo_in_coarse_z(pdn_ctx *ctx, SolveWork *w) {
    const char *e = getenv("PADNE_B200_HALO_FUSED");   // 0: keep the separate push / wait kernels (same on every rank)
    if (e && e[0] == '0') return false;
    // n_interface counts the interface rows of ALL ranks (every rank sees the same number: same decision)
    return w->two_level && is_dist(ctx) && pdn_peer_enabled(ctx) && w->halo.n_interface <= 16384 * pdn_comm_world(ctx);
}

void enqueue_coarse_correction(pdn_ctx *ctx, SolveWork *w, int vgrid, int slot, double boost, const double *src,
                               double *dst, bool rc_summed = false, bool with_local = false) {
    // several GPUs: every rank restricted its own rows -> exact integer sum, identical everywhere.  NCCL
    // fallback: both slots (the idle one is all zeros).  Peer memory: cg_update_tl_kernel's last CTA has already
    // done it (rc_summed); the other restricting kernels are followed by the exchange kernel.
    if (is_dist(ctx)) {
        if (!pdn_peer_enabled(ctx)) pdn_comm_allreduce_sum_i64(ctx, w->rc_fix.p, 2 * w->nc);
        else if (!rc_summed)
            PDN_LAUNCH(ctx, peer_sum_i64_kernel, 1, 256, 0, pdn_peer_link(ctx), w->d_sc, w->rc_fix.p, w->nc, slot,
                       (const int32_t *)w->agg_range.p);
    }
    const int cgrid = (w->nc + CO_RB - 1) / CO_RB;
    const int32_t nc2 = with_local ? w->nc2 : 0;
    const int cgrid2 = (nc2 + CO_RB - 1) / CO_RB;
    PDN_LAUNCH_PDL(ctx, w->pdl, coarse_apply_kernel, cgrid + cgrid2, CO_BLK, 0, w->nc, w->co_Af.p, w->d_sc, w->rc_fix.p, w->xc.p,
                   slot, boost, cgrid, nc2, w->co_Af2.p, w->rc_fix2.p, w->xc2.p);
    const int32_t *agg2 = nc2 > 0 ? w->agg2.p : nullptr;
    const double *xc2 = nc2 > 0 ? w->xc2.p : nullptr;
    const PeerLink *Lp = nullptr;
    size_t vec_off = 0;
    unsigned int *ticket = nullptr;
    if (halo_in_coarse_z(ctx, w) && (w->halo.n_send > 0 || w->halo.n_recv > 0)) {
        int which = -1;
        for (int v = 0; v < PDN_PA_NVEC; v++)
            if (dst == pdn_peer_vec(ctx, v)) which = v;
        if (which < 0) PDN_FAIL(PDN_ECUDA, "coarse correction: the vector is not inside the peer arena");
        Lp = pdn_peer_link_dev(ctx);
        vec_off = pdn_peer_vec_offset(ctx, which);
        ticket = pdn_peer_ticket(ctx);
    }
    if (Lp != nullptr) {
        const int64_t npair = (w->row_end - w->row_begin + 1) / 2;
        const int zgrid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)ctx->n_sm, (npair + 1023) / 1024));   // one CTA per SM (60 registers)
        PDN_LAUNCH(ctx, coarse_z_kernel<1024>, zgrid, 1024, 0, w->row_begin, w->row_end, w->d_sc, w->dinv.p, src, w->agg.p,
                   w->xc.p, dst, Lp, w->halo.peer_d.p, vec_off, w->halo.send_idx.p, ticket, agg2, xc2);
    } else {
        PDN_LAUNCH_PDL(ctx, w->pdl, coarse_z_kernel<CO_BLK>, vgrid, CO_BLK, 0, w->row_begin, w->row_end, w->d_sc, w->dinv.p,
                       src, w->agg.p, w->xc.p, dst, Lp, (const HaloSend *)nullptr, vec_off, w->halo.send_idx.p, ticket, agg2, xc2);
    }
}

void enqueue_pcg_iteration(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    const int64_t r0 = w->row_begin, r1 = w->row_end;
    if (!halo_in_coarse_z(ctx, w)) halo_exchange(ctx, w, w->z.p);
    launch_sell<CG>(ctx, w, w->z.p, w->w.p, w->r.p);
    if (w->two_level) {
        const int64_t npair = (r1 - r0 + 1) / 2;
        auto kern = w->nc2 > 0 ? cg_update_tl_kernel<true> : cg_update_tl_kernel<false>;
        int tl_per_sm = 1;
        PDN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&tl_per_sm, kern, TLB, 0));
        if (tl_per_sm < 1) tl_per_sm = 1;
        const int tgrid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)ctx->n_sm * tl_per_sm, (npair + TLB - 1) / TLB));
        PDN_LAUNCH_PDL(ctx, w->pdl, kern, tgrid, TLB, 0, r0, r1, w->d_sc, w->w.p, w->z.p, w->p.p, w->s.p, w->x.p,
                       w->r.p, w->agg.p, w->rc_fix.p, w->nc, peer_link_or_none(ctx), &w->d_sc->ticket[3],
                       w->nc2 > 0 ? w->agg2.p : (const int32_t *)nullptr, w->rc_fix2.p, w->nc2, (const int32_t *)w->agg_range.p);
        enqueue_coarse_correction(ctx, w, vgrid, -1, CO_BOOST32, w->r.p, w->z.p, /*rc_summed=*/pdn_peer_enabled(ctx),
                                  /*with_local=*/true);
    } else {
        PDN_LAUNCH_PDL(ctx, w->pdl, cg_update_kernel<false>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->w.p, w->z.p, w->p.p, w->s.p,
                   w->x.p, w->r.p, (const int32_t *)nullptr, (long long *)nullptr, 0);
    }
}

// ---- round-2 PCG iteration (pdn_cg2.cuh): K1 SpMV + restriction + recurrence, K2 coarse apply, K3 update + z ----
template <bool RECUR>
static void launch_coarse_apply2(pdn_ctx *ctx, SolveWork *w, bool summed) {
    const int cgrid = (w->nc + CO_RB - 1) / CO_RB, cgrid2 = (w->nc2 + CO_RB - 1) / CO_RB;
    PDN_LAUNCH_PDL(ctx, w->pdl, coarse_apply2_kernel<RECUR>, cgrid + cgrid2, CO_BLK, 0, w->d_sc, make_cg2(w), w->co_Af.p, w->xc.p,
                   cgrid, w->co_Af2.p, w->xc2.p, summed ? 1 : 0, summed ? peer_link_or_none(ctx) : (const PeerLink *)nullptr);
}

template <bool ZONLY>
static void launch_update2(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    const int64_t r0 = w->row_begin, r1 = w->row_end;
    const int32_t *agg2 = w->nc2 > 0 ? w->agg2.p : nullptr;
    const double *xc2 = w->nc2 > 0 ? w->xc2.p : nullptr;
    if (halo_in_coarse_z(ctx, w) && (w->halo.n_send > 0 || w->halo.n_recv > 0)) {
        int which = -1;
        for (int v = 0; v < PDN_PA_NVEC; v++)
            if (w->z.p == pdn_peer_vec(ctx, v)) which = v;
        if (which < 0) PDN_FAIL(PDN_ECUDA, "update: z is not inside the peer arena");
        const int64_t npair = (r1 - r0 + 1) / 2;
        // 512 threads, two CTAs per SM: the body runs like the single-GPU kernel, and 296 tickets do not queue
        // (PADNE_B200_K3=1024: one 1024-thread CTA per SM, the round-1 shape of this tail)
        const char *e = getenv("PADNE_B200_K3");
        if (e && strcmp(e, "1024") == 0) {
            const int zgrid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)ctx->n_sm, (npair + 1023) / 1024));
            PDN_LAUNCH_PDL(ctx, w->pdl, (cg_update2_kernel<1024, ZONLY>), zgrid, 1024, 0, r0, r1, w->d_sc, w->dinv.p, w->w.p, w->z.p,
                           w->p.p, w->s.p, w->x.p, w->r.p, w->agg.p, w->xc.p, agg2, xc2, pdn_peer_link_dev(ctx), w->halo.peer_d.p,
                           pdn_peer_vec_offset(ctx, which), w->halo.send_idx.p, pdn_peer_ticket(ctx), w->tab.p, w->nc, w->tab2.p,
                           w->nc2);
        } else {
            const int zgrid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)2 * ctx->n_sm, (npair + 511) / 512));
            PDN_LAUNCH_PDL(ctx, w->pdl, (cg_update2_kernel<512, ZONLY>), zgrid, 512, 0, r0, r1, w->d_sc, w->dinv.p, w->w.p, w->z.p,
                           w->p.p, w->s.p, w->x.p, w->r.p, w->agg.p, w->xc.p, agg2, xc2, pdn_peer_link_dev(ctx), w->halo.peer_d.p,
                           pdn_peer_vec_offset(ctx, which), w->halo.send_idx.p, pdn_peer_ticket(ctx), w->tab.p, w->nc, w->tab2.p,
                           w->nc2);
        }
    } else {
        PDN_LAUNCH_PDL(ctx, w->pdl, (cg_update2_kernel<RPB, ZONLY>), vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->w.p, w->z.p,
                       w->p.p, w->s.p, w->x.p, w->r.p, w->agg.p, w->xc.p, agg2, xc2, (const PeerLink *)nullptr,
                       (const HaloSend *)nullptr, (size_t)0, (const int32_t *)nullptr, (unsigned int *)nullptr, w->tab.p, w->nc,
                       w->tab2.p, w->nc2);
    }
}

// exact restriction of the current r (start: also p = s = 0, c_s = 0, tables cleared); several GPUs: integer sum
template <bool START>
static void enqueue_pcg2_restrict(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    const int64_t r0 = w->row_begin, r1 = w->row_end;
    const bool saved_pdl = w->pdl;
    w->pdl = false;     // plain launches: these kernels do not call griddepcontrol.wait
    PDN_LAUNCH(ctx, cg2_restrict_kernel<START>, vgrid, RPB, 0, r0, r1, w->d_sc, w->r.p, w->p.p, w->s.p, w->agg.p, w->rc_fix.p,
               w->nc2 > 0 ? w->agg2.p : (const int32_t *)nullptr, w->rc_fix2.p);
    if (is_dist(ctx))
        PDN_LAUNCH(ctx, peer_sum_i64_kernel, 1, 256, 0, pdn_peer_link(ctx), w->d_sc, w->rc_fix.p, w->nc, 0,
                   (const int32_t *)w->agg_range.p);
    PDN_LAUNCH(ctx, cg2_convert_kernel<START>, std::max(1, (std::max(w->nc, w->nc2) + 255) / 256), 256, 0, w->d_sc, make_cg2(w),
               w->rc_fix.p, w->rc_fix2.p);
    w->pdl = saved_pdl;
}

void enqueue_pcg2_iteration(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    if (!halo_in_coarse_z(ctx, w)) halo_exchange(ctx, w, w->z.p);
    const Cg2 cg = make_cg2(w);
    if (w->xs) {
        if (is_dist(ctx)) launch_xs_kernel<CG2, true>(ctx, w, w->z.p, w->w.p, w->r.p, cg);
        else launch_xs_kernel<CG2, false>(ctx, w, w->z.p, w->w.p, w->r.p, cg);
    } else {
        if (is_dist(ctx)) launch_sell_kernel<CG2, true>(ctx, w, w->z.p, w->w.p, w->r.p, cg);
        else launch_sell_kernel<CG2, false>(ctx, w, w->z.p, w->w.p, w->r.p, cg);
    }
    launch_coarse_apply2<true>(ctx, w, is_dist(ctx));
    launch_update2<false>(ctx, w, vgrid);
}

// start / restart from the residual r
void enqueue_pcg2_start(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    cudaStream_t st = ctx->stream;
    PDN_CUDA(cudaMemsetAsync(w->rc_fix.p, 0, (size_t)2 * w->nc * sizeof(long long), st));
    if (w->nc2 > 0) PDN_CUDA(cudaMemsetAsync(w->rc_fix2.p, 0, (size_t)2 * w->nc2 * sizeof(long long), st));
    enqueue_pcg2_restrict<true>(ctx, w, vgrid);
    launch_coarse_apply2<false>(ctx, w, false);
    launch_update2<true>(ctx, w, vgrid);
}

void enqueue_bicg_iteration(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    const int64_t r0 = w->row_begin, r1 = w->row_end;
    const bool dist = is_dist(ctx);
    const int32_t *no_agg = nullptr;
    long long *no_rc = nullptr;
    if (!halo_in_coarse_z(ctx, w)) halo_exchange(ctx, w, w->ph.p);
    launch_sell<BICG_V>(ctx, w, w->ph.p, w->vv.p, w->rhat.p);
    if (w->two_level) {
        PDN_LAUNCH(ctx, bicg_s_kernel<true>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->vv.p, w->s.p, w->sh.p,
                   w->agg.p, w->rc_fix.p, w->nc);
        enqueue_coarse_correction(ctx, w, vgrid, 0, BICG_BOOST, w->s.p, w->sh.p);
    } else {
        PDN_LAUNCH(ctx, bicg_s_kernel<false>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->vv.p, w->s.p, w->sh.p,
                   no_agg, no_rc, 0);
    }
    if (!halo_in_coarse_z(ctx, w)) halo_exchange(ctx, w, w->sh.p);
    launch_sell<BICG_T>(ctx, w, w->sh.p, w->tt.p, w->s.p);
    if (dist) {
        PDN_LAUNCH(ctx, bicg_x_kernel<true>, vgrid, RPB, 0, r0, r1, w->d_sc, w->ph.p, w->sh.p, w->s.p, w->tt.p,
                   w->rhat.p, w->x.p, w->r.p, w->partials.p);
        dist_reduce<5>(ctx, w);
    } else {
        PDN_LAUNCH(ctx, bicg_x_kernel<false>, vgrid, RPB, 0, r0, r1, w->d_sc, w->ph.p, w->sh.p, w->s.p, w->tt.p,
                   w->rhat.p, w->x.p, w->r.p, w->partials.p);
    }
    if (w->two_level) {
        PDN_LAUNCH(ctx, bicg_p_kernel<true>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->vv.p, w->p.p, w->ph.p,
                   w->agg.p, w->rc_fix.p, w->nc);
        enqueue_coarse_correction(ctx, w, vgrid, 1, BICG_BOOST, w->p.p, w->ph.p);
    } else {
        PDN_LAUNCH(ctx, bicg_p_kernel<false>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->vv.p, w->p.p, w->ph.p,
                   no_agg, no_rc, 0);
    }
}

// Solve A x = b (x starts at 0).  method: PDN_METHOD_PCG or PDN_METHOD_BICGSTAB.
// Multi-GPU: b, dinv and the operator are replicated; every rank iterates on its own rows and all
// ranks see identical scalars (the all-reduce returns the same bits everywhere), so they take the
// same decisions; x is all-gathered at the end.
//
// Stopping (SolverInfo parity, padne/solver.py:775 and tests/test_solver.py:2083-2089): the reference's direct
// solve leaves ||L v - r||_2 ~ 1e-10 and its own test bar is 1e-9 ABSOLUTE, while rtol = 1e-10 on a right-hand
// side of norm ~2e3 stops at 2e-7.  So the loop stops when the TRUE residual of the reduced system is below
//     min( rtol ||b||,  max( atol / 4,  4 u amax ||x||_2 ) )
// -- the relative bar of the north star AND the reference's absolute one, the latter no lower than what fp64 can
// resolve (u amax ||x|| is the rounding of one residual evaluation; below it the recurrence keeps shrinking but
// b - A x does not: scripts/proto_v2_pcg.py).  The absolute target needs ||x||, so the loop first runs to
// rtol ||b||, then tightens the threshold ONCE and simply continues (PCG: the recurrences are intact when `done`
// latches; BiCGSTAB: restart from the true residual).  Stagnation above the target but below rtol ||b|| is
// accepted and reported (stagnated = 1); above rtol ||b|| it is PDN_ENOTCONVERGED.
struct KrylovTarget {
    double rtol, atol_full;
};

KrylovResult run_krylov(pdn_ctx *ctx, SolveWork *w, int method, KrylovTarget tgt, int max_iter, int check_every) {
    KrylovResult res;
    const int64_t n = w->n;
    const int64_t r0 = w->row_begin, r1 = w->row_end;
    const bool dist = is_dist(ctx);
    const bool v2 = w->v2 && method == PDN_METHOD_PCG;
    cudaStream_t st = ctx->stream;
    const int vgrid = pdn_grid(ctx, r1 - r0, RPB, 8);
    cudaEvent_t ev0, ev1;
    PDN_CUDA(cudaEventCreate(&ev0));
    PDN_CUDA(cudaEventCreate(&ev1));
    PDN_CUDA(cudaEventRecord(ev0, st));

    // ||b|| (b is replicated: no communication -- unless the assembly was partitioned: own rows, summed over the ranks)
    reset_scal(ctx, w, 0.0, 0.0, max_iter);
    if (w->part) {
        PDN_LAUNCH(ctx, sumsq_kernel<true>, vgrid, RPB, 0, r0, r1, w->b.p, (const double *)nullptr, w->d_sc, w->d_out,
                   w->partials.p);
        dist_reduce<7>(ctx, w);
    } else
    PDN_LAUNCH(ctx, sumsq_kernel<false>, pdn_grid(ctx, n, RPB, 8), RPB, 0, (int64_t)0, n, w->b.p,
               (const double *)nullptr, w->d_sc, w->d_out, w->partials.p);
    double bnorm2 = 0.0;
    PDN_CUDA(cudaMemcpyAsync(&bnorm2, w->d_out, sizeof(double), cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    PDN_CUDA(cudaMemsetAsync(w->x.p, 0, n * sizeof(double), st));
    res.bnorm = sqrt(bnorm2 > 0.0 ? bnorm2 : 0.0);
    if (!(bnorm2 > 0.0)) {  // b == 0 -> u == 0
        PDN_CUDA(cudaEventRecord(ev1, st));
        PDN_CUDA(cudaEventSynchronize(ev1));
        cudaEventDestroy(ev0);
        cudaEventDestroy(ev1);
        if (bnorm2 != bnorm2) PDN_FAIL(PDN_ESINGULAR, "solve: right-hand side is not finite");
        return res;
    }
    const double tol_req2 = tgt.rtol * tgt.rtol * bnorm2;      // the required bar: rtol ||b||
    double tol2 = tol_req2;                                    // the bar in force (tightened once, below)
    bool refined = !(tgt.atol_full > 0.0);
    reset_scal(ctx, w, tol2, bnorm2, max_iter);
    PDN_CUDA(cudaMemcpyAsync(w->r.p, w->b.p, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    if (dist) {
        // the halo entries of the SpMV inputs are written by the exchange only; start them defined.
        // One exchange outside the capture also lets NCCL set up its point-to-point channels.
        PDN_CUDA(cudaMemsetAsync(w->z.p, 0, n * sizeof(double), st));
        if (method == PDN_METHOD_BICGSTAB) {
            PDN_CUDA(cudaMemsetAsync(w->ph.p, 0, n * sizeof(double), st));
            PDN_CUDA(cudaMemsetAsync(w->sh.p, 0, n * sizeof(double), st));
        }
        halo_exchange(ctx, w, w->z.p);
        PDN_CUDA(cudaStreamSynchronize(st));
    }

    // one CUDA graph = check_every iterations; kernels turn into no-ops once sc->done is set
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int per_iter = 0;
    auto enqueue_iteration = [&]() {
        if (v2) enqueue_pcg2_iteration(ctx, w, vgrid);
        else if (method == PDN_METHOD_PCG) enqueue_pcg_iteration(ctx, w, vgrid);
        else enqueue_bicg_iteration(ctx, w, vgrid);
    };
    auto build_graph = [&]() {
        const int64_t launches_before = ctx->launches;
        // PCG: the kernels of an iteration are chained by programmatic dependent launches (round-1 iteration:
        // one GPU only; round-2 iteration: any number of GPUs, every kernel of the chain waits on its predecessor)
        const char *pdl_env = getenv("PADNE_B200_PDL");
        w->pdl = method == PDN_METHOD_PCG && (v2 || !dist) && !(pdl_env && pdl_env[0] == '0');
        PDN_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        try {
            if (v2) enqueue_pcg2_restrict<false>(ctx, w, vgrid);      // r_c re-restricted once per graph launch
            for (int k = 0; k < check_every; k++) enqueue_iteration();
        } catch (...) {
            w->pdl = false;
            cudaGraph_t g2 = nullptr;
            cudaStreamEndCapture(st, &g2);
            if (g2) cudaGraphDestroy(g2);
            throw;
        }
        w->pdl = false;
        PDN_CUDA(cudaStreamEndCapture(st, &graph));
        PDN_CUDA(cudaGraphInstantiate(&exec, graph, 0));
        per_iter = (int)((ctx->launches - launches_before) / check_every);   // kernels of one captured iteration
        ctx->launches = launches_before;  // capture is not execution
    };
    {
        const auto tb0 = std::chrono::steady_clock::now();
        build_graph();
        if (getenv("PADNE_B200_TRACE"))
            fprintf(stderr, "[padne_b200 trace rank %d] graph capture + instantiate: %.3f ms (%d kernels per iteration x %d)\n",
                    pdn_comm_rank(ctx), std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tb0).count(),
                    per_iter, check_every);
    }

    auto cleanup = [&]() {
        if (exec) cudaGraphExecDestroy(exec);
        if (graph) cudaGraphDestroy(graph);
        cudaEventDestroy(ev0);
        cudaEventDestroy(ev1);
    };
    // ||x||_2 of the current iterate (this rank's rows, summed over the ranks)
    auto x_norm = [&]() -> double {
        double xx = 0.0;
        if (dist) {
            PDN_LAUNCH(ctx, sumsq_kernel<true>, vgrid, RPB, 0, r0, r1, w->x.p, (const double *)nullptr, w->d_sc, w->d_out,
                       w->partials.p);
            dist_reduce<7>(ctx, w);
        } else {
            PDN_LAUNCH(ctx, sumsq_kernel<false>, vgrid, RPB, 0, r0, r1, w->x.p, (const double *)nullptr, w->d_sc, w->d_out,
                       w->partials.p);
        }
        PDN_CUDA(cudaMemcpyAsync(&xx, w->d_out, sizeof(double), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        return sqrt(xx > 0.0 ? xx : 0.0);
    };
    // what ONE fp64 evaluation of b - A x can resolve: u amax ||x||_2 (computed once, on the first pass's x).
    // Measured (scripts/gpu_diag.py, round 2): SuperLU lands at ~1x this, a PCG pass that updates x in place at
    // 3-5x (its x += alpha p roundings add up over hundreds of iterations), a refinement-style restart at 1-1.5x.
    double floor1 = -1.0;
    auto get_floor = [&]() -> double {
        if (floor1 < 0.0) {
            const Scal keep = *w->h_sc;                  // the reduction borrows the scalar block
            floor1 = 1.1102230246251565e-16 * w->amax * x_norm();
            *w->h_sc = keep;
            res.floor_abs = floor1;
        }
        return floor1;
    };
    // The absolute target once ||x|| is known; returns true when it tightened the bar in force.  When the
    // reference's bar is within reach of the arithmetic (floor <= 0.7 atol) the loop aims below it (0.7 atol, or
    // 4 floor / atol / 4 if those are lower still); on boards so large that even one residual evaluation is
    // coarser than atol it aims at 4 floor, the level a PCG pass reaches without stagnating.
    double accept_abs = 0.0;                             // true residuals up to here are taken (>= the target)
    auto refine = [&]() -> bool {
        if (refined) return false;
        refined = true;
        const double f = get_floor(), a = tgt.atol_full;
        double target;
        if (f <= 0.7 * a) {
            target = std::max(0.25 * a, std::min(4.0 * f, 0.7 * a));
            accept_abs = std::max(target, std::min(1.5 * target, 0.9 * a));
        } else {
            target = 4.0 * f;
            accept_abs = 1.5 * target;
        }
        if (target * target < tol2 * 0.999) { tol2 = target * target; return true; }
        accept_abs = 0.0;
        return false;
    };
    // true residual good enough?  The bar in force, or the acceptance level of the absolute target (never above
    // the required rtol ||b||)
    auto accepted = [&](double rr_true) {
        return rr_true <= std::max(tol2 * 1.0000001, std::min(accept_abs * accept_abs, tol_req2));
    };
    // stagnation: the required bar, but never below 8x the evaluation floor (a request the arithmetic cannot meet)
    auto acceptable_floor = [&](double rr_true) {
        const double f = 8.0 * get_floor();
        return rr_true <= std::max(tol_req2 * 1.0000001, f * f);
    };
    // Restarts are refinement steps: the solution so far moves to x_acc, the true residual becomes the right-hand
    // side of a CORRECTION system solved from zero, so the roundings of the new pass scale with the correction, not
    // with x; the pieces are added once at the end.
    const double *b_cur = w->b.p;
    bool have_acc = false;
    auto clear_tables = [&]() {
        if (!v2) return;
        PDN_CUDA(cudaMemsetAsync(w->tab.p, 0, (size_t)w->nc * TAB_STRIDE * sizeof(unsigned long long), st));
        if (w->nc2 > 0) PDN_CUDA(cudaMemsetAsync(w->tab2.p, 0, (size_t)w->nc2 * TAB_STRIDE * sizeof(unsigned long long), st));
    };

    try {
        const int max_restarts = 6;
        double best_rr = -1.0;
        for (int attempt = 0;; attempt++) {
            // start / restart from the current residual r
            if (v2) {
                enqueue_pcg2_start(ctx, w, vgrid);
            } else if (method == PDN_METHOD_PCG && w->two_level) {
                PDN_CUDA(cudaMemsetAsync(w->rc_fix.p, 0, (size_t)2 * w->nc * sizeof(long long), st));
                PDN_LAUNCH(ctx, cg_init_kernel<true>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->z.p, w->p.p,
                           w->s.p, w->agg.p, w->rc_fix.p, w->nc);
                if (w->nc2 > 0) {   // the same restriction into the rank-local level
                    PDN_CUDA(cudaMemsetAsync(w->rc_fix2.p, 0, (size_t)2 * w->nc2 * sizeof(long long), st));
                    PDN_LAUNCH(ctx, cg_init_kernel<true>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->z.p, w->p.p,
                               w->s.p, w->agg2.p, w->rc_fix2.p, w->nc2);
                }
                enqueue_coarse_correction(ctx, w, vgrid, -1, CO_BOOST32, w->r.p, w->z.p, false, /*with_local=*/true);
            } else if (method == PDN_METHOD_PCG) {
                PDN_LAUNCH(ctx, cg_init_kernel<false>, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->z.p, w->p.p,
                           w->s.p, (const int32_t *)nullptr, (long long *)nullptr, 0);
            } else {
                const bool tl = w->two_level;
                if (tl) PDN_CUDA(cudaMemsetAsync(w->rc_fix.p, 0, (size_t)2 * w->nc * sizeof(long long), st));
                const int32_t *agg = tl ? w->agg.p : nullptr;
                long long *rc = tl ? w->rc_fix.p : nullptr;
                auto kern = dist ? (tl ? bicg_init_kernel<true, true> : bicg_init_kernel<true, false>)
                                 : (tl ? bicg_init_kernel<false, true> : bicg_init_kernel<false, false>);
                PDN_LAUNCH(ctx, kern, vgrid, RPB, 0, r0, r1, w->d_sc, w->dinv.p, w->r.p, w->rhat.p, w->p.p, w->ph.p,
                           w->partials.p, agg, rc, w->nc);
                if (dist) dist_reduce<6>(ctx, w);
                if (tl) enqueue_coarse_correction(ctx, w, vgrid, 1, BICG_BOOST, w->p.p, w->ph.p);
            }
            if (attempt == 0 && getenv("PADNE_B200_KTIME")) {
                KernelTimes kt;
                ctx->ktime = &kt;
                kt.active = true;
                kt.mark("start", st);
                for (int k = 0; k < check_every; k++) enqueue_iteration();
                kt.active = false;
                ctx->ktime = nullptr;
                kt.report(pdn_comm_rank(ctx));
            }
            while (true) {
                PDN_CUDA(cudaGraphLaunch(exec, st));
                fetch_scal(ctx, w);
                res.iterations = w->h_sc->iter;
                ctx->launches += (int64_t)per_iter * check_every + (v2 ? (dist ? 3 : 2) : 0);  // kernels inside the graph
                if (!w->h_sc->done) continue;
                // converged to the bar in force: the first time, bring in the absolute target (needs ||x||).
                // PCG keeps going with the tighter bar -- `done` latched before anything was updated, so clearing
                // it resumes the very same iteration; BiCGSTAB restarts from the true residual below.
                if (w->h_sc->done == 1 && method == PDN_METHOD_PCG && w->h_sc->iter < max_iter && !refined) {
                    const bool tighter = refine();
                    Scal keep = *w->h_sc;
                    memset(keep.ticket, 0, sizeof keep.ticket);
                    if (tighter) { keep.done = 0; keep.tol2 = tol2; }
                    *w->h_sc = keep;
                    PDN_CUDA(cudaMemcpyAsync(w->d_sc, w->h_sc, sizeof(Scal), cudaMemcpyHostToDevice, st));
                    if (tighter) {
                        clear_tables();                 // K1's aborted launch left its restriction in them
                        continue;
                    }
                }
                break;
            }
            const int done = w->h_sc->done;
            // true residual of the current x
            halo_exchange(ctx, w, w->x.p);
            launch_sell<RESID>(ctx, w, w->x.p, w->r.p, b_cur);
            fetch_scal(ctx, w);
            const double rr_true = w->h_sc->rr;
            res.rel_residual = sqrt(rr_true / bnorm2);
            if (rr_true != rr_true) PDN_FAIL(PDN_ESINGULAR, "solve: residual is not finite (singular system?)");
            if (attempt == 0) get_floor();                         // from the first pass's x, before it moves to x_acc
            if (done == 1 && method != PDN_METHOD_PCG) refine();   // BiCGSTAB: the restart below runs to the tightened bar
            if (accepted(rr_true)) break;
            if (w->h_sc->iter >= max_iter) {
                if (rr_true <= tol_req2 * 1.0000001) { res.stagnated = 1; break; }    // the required bar holds
                PDN_FAIL(PDN_ENOTCONVERGED, "solve: %d iterations, relative residual %.3e > %.3e", w->h_sc->iter,
                         res.rel_residual, tgt.rtol);
            }
            // converged by the recurrence but not truly (drift, or the fp64 floor of b - A x), or breakdown:
            // restart from the true residual, unless that stopped paying
            if (attempt >= max_restarts || (done == 1 && best_rr >= 0.0 && rr_true >= 0.25 * best_rr)) {
                if (acceptable_floor(rr_true)) { res.stagnated = 1; break; }    // the floor of the arithmetic
                PDN_FAIL(PDN_ENOTCONVERGED, "solve: stagnated at relative residual %.3e (required %.3e)",
                         res.rel_residual, tgt.rtol);
            }
            if (best_rr < 0.0 || rr_true < best_rr) best_rr = rr_true;
            res.restarts++;
            // x_acc += x;  right-hand side of the correction = the true residual just computed;  x = 0
            w->x_acc.reserve((size_t)n);
            w->b_corr.reserve((size_t)n);
            PDN_LAUNCH(ctx, accumulate_kernel, vgrid, BLK, 0, r0, r1, w->x.p, w->x_acc.p, have_acc ? 0 : 1);
            have_acc = true;
            PDN_CUDA(cudaMemcpyAsync(w->b_corr.p + r0, w->r.p + r0, (size_t)(r1 - r0) * sizeof(double), cudaMemcpyDeviceToDevice, st));
            b_cur = w->b_corr.p;
            PDN_CUDA(cudaMemsetAsync(w->x.p, 0, n * sizeof(double), st));
            restart_scal(ctx, w, tol2);
            clear_tables();
        }
        if (have_acc) PDN_LAUNCH(ctx, accumulate_kernel, vgrid, BLK, 0, r0, r1, w->x_acc.p, w->x.p, 0);
    } catch (...) {
        cleanup();
        throw;
    }
    res.abs_target = sqrt(tol2);
    if (dist) {   // every rank needs the whole u for the (replicated) recovery step
        try {
            pdn_comm_allgatherv(ctx, w->x.p, w->halo.begin.data());
        } catch (...) {
            cleanup();
            throw;
        }
    }
    PDN_CUDA(cudaEventRecord(ev1, st));
    PDN_CUDA(cudaEventSynchronize(ev1));
    PDN_CUDA(cudaEventElapsedTime(&res.ms, ev0, ev1));
    cleanup();
    return res;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// pdn_solve
// ------------------------------------------------------------------------------------------------
namespace {

__global__ void __launch_bounds__(BLK) last_diag_kernel(int64_t N, const int32_t *__restrict__ rp,
                                                        const int32_t *__restrict__ ci, int32_t *out) {
    // out[0] = 1 + largest row index that stores a diagonal entry;
    // out[1] = 1 + largest row index WITHOUT a diagonal that has an entry right of the diagonal position: an internal
    // node touched only by voltage sources (its row holds nothing but +-1 in source-current columns, which lie after
    // it), whereas a constraint row only references node columns, all of them before it.  n_nodes = max of the two.
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < N; i += (int64_t)gridDim.x * BLK) {
        bool has = false, upper = false;
        for (int32_t k = rp[i]; k < rp[i + 1]; k++) { has = has || ci[k] == i; upper = upper || ci[k] > i; }
        if (has) atomicMax(out, (int32_t)i + 1);
        else if (upper) atomicMax(out + 1, (int32_t)i + 1);
    }
}

template <typename T>
void upload(pdn_ctx *ctx, DevBuf<T> &buf, const std::vector<T> &h) {
    buf.reserve(h.size() + 1);
    if (!h.empty())
        PDN_CUDA(cudaMemcpyAsync(buf.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// multi-GPU: halo plan.  Every rank holds the same operator, so every rank can derive both what it
// needs from each peer and what each peer needs from it, with no negotiation: needmask[c] gets bit k
// when a row owned by rank k references column c owned by somebody else.
// ------------------------------------------------------------------------------------------------
struct RowPart {
    int world;
    int64_t begin[33];
    __host__ __device__ int owner(int64_t r) const {
        int k = 0;
        while (k + 1 < world && r >= begin[k + 1]) k++;
        return k;
    }
};

// what: bit 0 the SELL part, bit 1 the override rows.  (A partitioned setup marks its own SELL rows, sums the marks
// over the ranks -- every rank sets only its own bit, so the sum is the OR -- and then adds the override rows,
// which every rank holds.)
__global__ void __launch_bounds__(256) halo_mark_kernel(SellOp A, int64_t n_slices, RowPart part,
                                                        uint32_t *__restrict__ needmask, int what) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int64_t n_warp = ((int64_t)gridDim.x * 256) >> 5;
    for (int64_t s = warp; (what & 1) && s < n_slices; s += n_warp) {
        const int64_t r = s * SELL_C + lane;
        if (r >= A.n) continue;
        const int32_t off = A.soff[s], w = (A.soff[s + 1] - off) >> 5;
        const int ko = part.owner(r);
        const int64_t lo = part.begin[ko], hi = part.begin[ko + 1];
        for (int32_t k = 0; k < w; k++) {
            const int32_t c = A.ci[off + k * SELL_C + lane];
            if (c < lo || c >= hi) atomicOr(&needmask[c], 1u << ko);
        }
    }
    // override rows
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; (what & 2) && t < A.n_special; t += (int64_t)gridDim.x * 256) {
        const int ko = part.owner(A.sp_row[t]);
        const int64_t lo = part.begin[ko], hi = part.begin[ko + 1];
        for (int32_t q = A.sp_rp[t]; q < A.sp_rp[t + 1]; q++) {
            const int32_t c = A.sp_ci[q];
            if (c < lo || c >= hi) atomicOr(&needmask[c], 1u << ko);
        }
    }
}

// The send / receive lists are ordered compactions of needmask (flag -> exclusive scan -> scatter), built on
// the device: a board with many vias has ~1M interface entries and sorting them on the host took 15 ms.
//   list 0 (send):  item v = (k, c) = (v / len_me, begin[me] + v % len_me);  kept if k != me and rank k reads c
//   list 1 (recv):  item v = c, k = owner(c);                                  kept if k != me and I read c
// Both come out ordered by (peer, column), the same order on the sending and on the receiving rank.
__device__ __forceinline__ bool halo_item(int list, const RowPart &part, int me, int64_t v, const uint32_t *needmask,
                                          int32_t &c) {
    if (list == 0) {
        const int64_t len = part.begin[me + 1] - part.begin[me];
        const int k = (int)(v / len);
        c = (int32_t)(part.begin[me] + (v - (int64_t)k * len));
        return k != me && ((needmask[c] >> k) & 1u);
    }
    c = (int32_t)v;
    return part.owner(v) != me && ((needmask[c] >> me) & 1u);
}
__global__ void __launch_bounds__(256) halo_flag_kernel(int list, RowPart part, int me, int64_t n_items,
                                                        const uint32_t *__restrict__ needmask, int32_t *__restrict__ flag) {
    for (int64_t v = (int64_t)blockIdx.x * 256 + threadIdx.x; v < n_items; v += (int64_t)gridDim.x * 256) {
        int32_t c;
        flag[v] = halo_item(list, part, me, v, needmask, c) ? 1 : 0;
    }
}
__global__ void __launch_bounds__(256) halo_scatter_kernel(int list, RowPart part, int me, int64_t n_items,
                                                           const uint32_t *__restrict__ needmask,
                                                           const int32_t *__restrict__ offs, int32_t *__restrict__ out) {
    for (int64_t v = (int64_t)blockIdx.x * 256 + threadIdx.x; v < n_items; v += (int64_t)gridDim.x * 256) {
        int32_t c;
        if (halo_item(list, part, me, v, needmask, c)) out[offs[v]] = c;
    }
}
// number of rows some other rank reads (the same on every rank: the marks are replicated)
__global__ void __launch_bounds__(256) halo_count_kernel(int64_t n, const uint32_t *__restrict__ needmask, int32_t *count) {
    int32_t c = 0;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) c += needmask[i] != 0u;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}
// first list position of every peer: out[k] (send), out[world + 1 + k] (recv), k = 0 .. world
__global__ void halo_offsets_kernel(RowPart part, int me, const int32_t *__restrict__ offs_send,
                                    const int32_t *__restrict__ offs_recv, int32_t *__restrict__ out) {
    const int k = threadIdx.x;
    if (k > part.world) return;
    const int64_t len = part.begin[me + 1] - part.begin[me];
    out[k] = offs_send[(int64_t)k * len];
    out[part.world + 1 + k] = offs_recv[part.begin[k]];
}

static void build_halo_plan(pdn_ctx *ctx, SolveWork *w, int32_t /*n_sp_entries*/) {
    cudaStream_t st = ctx->stream;
    HaloPlan &h = w->halo;
    const int world = h.world, me = h.me;
    const int64_t n = w->n;
    RowPart part{};
    part.world = world;
    for (int k = 0; k <= world; k++) part.begin[k] = h.begin[k];
    w->needmask.reserve((size_t)n);
    PDN_CUDA(cudaMemsetAsync(w->needmask.p, 0, (size_t)n * sizeof(uint32_t), st));
    if (w->part) {
        PDN_LAUNCH(ctx, halo_mark_kernel, pdn_grid(ctx, w->n_slices * 32, 256, 8), 256, 0, w->A, w->n_slices, part,
                   w->needmask.p, 1);
        pdn_comm_allreduce_sum_i32(ctx, reinterpret_cast<int32_t *>(w->needmask.p), (size_t)n);
        PDN_LAUNCH(ctx, halo_mark_kernel, pdn_grid(ctx, w->n_slices * 32, 256, 8), 256, 0, w->A, w->n_slices, part,
                   w->needmask.p, 2);
    } else {
        PDN_LAUNCH(ctx, halo_mark_kernel, pdn_grid(ctx, w->n_slices * 32, 256, 8), 256, 0, w->A, w->n_slices, part,
                   w->needmask.p, 3);
    }
    const int64_t n_send_items = (int64_t)world * (h.begin[me + 1] - h.begin[me]), n_recv_items = n;
    const int64_t n_items = std::max(n_send_items, n_recv_items);
    w->tmp_i32b.reserve((size_t)n_items + 1);       // flags
    w->halo_col.reserve((size_t)n_send_items + 1);  // offsets of the send items
    w->halo_mask.reserve((size_t)n_recv_items + 1); // offsets of the receive items
    w->tmp_i32a.reserve((size_t)2 * (world + 1) + 1);
    std::vector<int32_t> offs((size_t)2 * (world + 1) + 1, 0);
    PDN_CUDA(cudaMemsetAsync(w->tmp_i32a.p + 2 * (world + 1), 0, sizeof(int32_t), st));
    PDN_LAUNCH(ctx, halo_count_kernel, pdn_grid(ctx, n, 256, 8), 256, 0, n, w->needmask.p, w->tmp_i32a.p + 2 * (world + 1));
    if (n_send_items > 0) {
        PDN_LAUNCH(ctx, halo_flag_kernel, pdn_grid(ctx, n_send_items, 256, 8), 256, 0, 0, part, me, n_send_items,
                   w->needmask.p, w->tmp_i32b.p);
        pdn_exclusive_scan_i32(ctx, w->tmp_i32b.p, n_send_items, w->halo_col.p);
    } else {
        PDN_CUDA(cudaMemsetAsync(w->halo_col.p, 0, sizeof(int32_t), st));
    }
    PDN_LAUNCH(ctx, halo_flag_kernel, pdn_grid(ctx, n_recv_items, 256, 8), 256, 0, 1, part, me, n_recv_items,
               w->needmask.p, w->tmp_i32b.p);
    pdn_exclusive_scan_i32(ctx, w->tmp_i32b.p, n_recv_items, w->halo_mask.p);
    PDN_LAUNCH(ctx, halo_offsets_kernel, 1, 32, 0, part, me, w->halo_col.p, w->halo_mask.p, w->tmp_i32a.p);
    PDN_CUDA(cudaMemcpyAsync(offs.data(), w->tmp_i32a.p, offs.size() * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    h.send_off.assign(world, 0); h.send_cnt.assign(world, 0);
    h.recv_off.assign(world, 0); h.recv_cnt.assign(world, 0);
    for (int k = 0; k < world; k++) {
        h.send_off[k] = offs[k];
        h.send_cnt[k] = offs[k + 1] - offs[k];
        h.recv_off[k] = offs[world + 1 + k];
        h.recv_cnt[k] = offs[world + 2 + k] - offs[world + 1 + k];
    }
    h.n_send = offs[world];
    h.n_recv = offs[2 * world + 1];
    h.n_interface = offs[2 * (world + 1)];
    h.peer = HaloSend{};
    for (int k = 0; k < world && k < PEER_MAX; k++) {
        h.peer.off[k] = h.send_off[k];
        h.peer.cnt[k] = h.send_cnt[k];
        h.peer.recv_cnt[k] = h.recv_cnt[k];
    }
    h.send_idx.reserve((size_t)h.n_send + 1);
    h.recv_idx.reserve((size_t)h.n_recv + 1);
    h.peer_d.reserve(1);
    PDN_CUDA(cudaMemcpyAsync(h.peer_d.p, &h.peer, sizeof(HaloSend), cudaMemcpyHostToDevice, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    if (h.n_send > 0)
        PDN_LAUNCH(ctx, halo_scatter_kernel, pdn_grid(ctx, n_send_items, 256, 8), 256, 0, 0, part, me, n_send_items,
                   w->needmask.p, w->halo_col.p, h.send_idx.p);
    if (h.n_recv > 0)
        PDN_LAUNCH(ctx, halo_scatter_kernel, pdn_grid(ctx, n_recv_items, 256, 8), 256, 0, 1, part, me, n_recv_items,
                   w->needmask.p, w->halo_mask.p, h.recv_idx.p);
    h.sendbuf.reserve((size_t)h.n_send + 1);
    h.recvbuf.reserve((size_t)h.n_recv + 1);
}

// ------------------------------------------------------------------------------------------------
// two-level preconditioner: aggregates, coarse matrix, its inverse (kernels in pdn_coarse.cuh).
// Returns false (and leaves Jacobi in place) when no useful coarse space can be built.
// ------------------------------------------------------------------------------------------------
// local == false: the global level over all rows (replicated on every rank).
// local == true:  the rank-local level: bins `world` times finer, only the rows [row_begin, row_end) get an
//                 aggregate, so A_loc = P^T A P keeps the couplings inside this rank only.
static bool build_coarse(pdn_ctx *ctx, SolveWork *w, int32_t target, int64_t n_entries, bool symmetric, bool local = false) {
    cudaStream_t st = ctx->stream;
    const int64_t n = w->n;
    DevBuf<int32_t> &agg = local ? w->agg2 : w->agg;
    DevBuf<long long> &rc_fix = local ? w->rc_fix2 : w->rc_fix;
    DevBuf<double> &co_A = local ? w->co_A2 : w->co_A;
    DevBuf<double> &xc = local ? w->xc2 : w->xc;
    DevBuf<float> &co_Af = local ? w->co_Af2 : w->co_Af;
    const int64_t row_lo = local ? w->row_begin : 0, row_hi = local ? w->row_end : n;
    const int world = local ? std::max(1, w->halo.world) : 1;
    const int world_all = std::max(1, w->halo.world);
    if (local) w->nc2 = 0;
    else { w->two_level = false; w->nc = 0; }
    // every exit that leaves the level unbuilt says why (pdn_info / the bench line / the caller's error text)
    auto fail = [&](const char *why) {
        char buf[200];
        snprintf(buf, sizeof buf, "%s coarse level not built: %s", local ? "rank-local" : "global", why);
        w->coarse_note = buf;
        return false;
    };
    if (!ctx->geo_valid || ctx->geo_n_v <= 0 || ctx->geo_n_v > n) return fail("no mesh geometry resident");
    // default: the inverse costs ~2.5 ms at 1600 aggregates and buys ~20% fewer iterations than 1024 at 1M rows
    if (target <= 0) target = n >= 200000 ? 1600 : 1024;
    if (target > CO_MAX) target = CO_MAX;
    if (target > (row_hi - row_lo) / 16) target = (int32_t)((row_hi - row_lo) / 16);     // tiny systems: keep >= 16 rows per aggregate
    if (target < 4) return fail("fewer than 64 rows");
    // bins are laid over ALL meshes; the local level keeps the ~1/world of them that hold its rows
    const int64_t raw_target = (int64_t)target * world;

    // bounding box of the vertices
    // (partitioned assembly: only this rank's vertices are sure to be resident -- box of those, then the union over the ranks)
    const int64_t bv0 = w->part ? std::min<int64_t>(w->row_begin, ctx->geo_n_v) : 0;
    const int64_t bv1 = w->part ? std::min<int64_t>(w->row_end, ctx->geo_n_v) : ctx->geo_n_v;
    const int bgrid = pdn_grid(ctx, std::max<int64_t>(bv1 - bv0, 1), CO_BLK, 4);
    w->co_tmp.reserve((size_t)4 * bgrid + 2048);
    PDN_LAUNCH(ctx, bbox_kernel, bgrid, CO_BLK, 0, bv1 - bv0, (const double2 *)ctx->geo_xy + bv0, w->co_tmp.p);
    std::vector<double> hb((size_t)4 * bgrid);
    PDN_CUDA(cudaMemcpyAsync(hb.data(), w->co_tmp.p, hb.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    double x0 = 1e300, y0 = 1e300, x1 = -1e300, y1 = -1e300;
    for (int b = 0; b < bgrid; b++) {
        x0 = std::min(x0, hb[4 * b]); y0 = std::min(y0, hb[4 * b + 1]);
        x1 = std::max(x1, hb[4 * b + 2]); y1 = std::max(y1, hb[4 * b + 3]);
    }
    double n_entries_d = (double)n_entries;
    if (w->part) {
        double box[5] = {-x0, -y0, x1, y1, (double)n_entries};
        w->tmp_f64a.reserve(8);
        PDN_CUDA(cudaMemcpyAsync(w->tmp_f64a.p, box, sizeof box, cudaMemcpyHostToDevice, st));
        pdn_comm_allreduce_max(ctx, w->tmp_f64a.p, 5);
        PDN_CUDA(cudaMemcpyAsync(box, w->tmp_f64a.p, sizeof box, cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        x0 = -box[0]; y0 = -box[1]; x1 = box[2]; y1 = box[3];
        n_entries_d = box[4] * world_all;       // bound on the stored entries of the whole operator (same on every rank)
    }
    if (!(x1 >= x0) || !(y1 >= y0)) return fail("degenerate bounding box");

    // bins: equal squares, about `target` aggregates in total
    CoarseGeom g{};
    g.n_mesh = ctx->geo_n_mesh;
    g.n_v = ctx->geo_n_v;
    const int64_t n_internal = n - ctx->geo_n_v;
    g.node_group = n_internal > 256 ? (int32_t)((n_internal + 255) / 256) : 1;
    const int32_t node_aggs = (int32_t)((n_internal + g.node_group - 1) / g.node_group);
    const double W = std::max(x1 - x0, 1e-12), Ht = std::max(y1 - y0, 1e-12);
    double bins_per_mesh = std::max(1.0, (double)(raw_target - node_aggs) / g.n_mesh);
    double H = sqrt(W * Ht / bins_per_mesh);
    int32_t nc = 0;
    agg.reserve((size_t)n);
    for (int attempt = 0;; attempt++) {
        for (int guard = 0; guard < 200; guard++) {
            g.nbx = std::max(1, (int)ceil(W / H));
            g.nby = std::max(1, (int)ceil(Ht / H));
            if ((int64_t)g.n_mesh * g.nbx * g.nby + node_aggs <= std::max<int64_t>(raw_target, 4)) break;
            H *= 1.05;
        }
        if ((int64_t)g.n_mesh * g.nbx * g.nby + node_aggs > (int64_t)CO_MAX * world) return fail("more meshes + internal-node aggregates than CO_MAX");
        g.x0 = x0; g.y0 = y0; g.inv_h = 1.0 / H;
        g.n_bins = g.n_mesh * g.nbx * g.nby;
        const int32_t n_raw = g.n_bins + node_aggs;

        // aggregate ids: raw bin -> compact id over the non-empty bins
        w->co_used.reserve((size_t)n_raw + 1);
        w->co_cid.reserve((size_t)n_raw + 1);
        PDN_CUDA(cudaMemsetAsync(w->co_used.p, 0, ((size_t)n_raw + 1) * sizeof(int32_t), st));
        if (w->part && !local) {
            // every rank bins its own rows (their coordinates are resident), the raw ids are all-gathered, and the
            // occupancy of the bins is read off the gathered array: identical compact ids everywhere
            PDN_LAUNCH(ctx, agg_raw_kernel, pdn_grid(ctx, n, CO_BLK, 8), CO_BLK, 0, n, w->row_begin, w->row_end, g,
                       ctx->geo_voff.p, (const double2 *)ctx->geo_xy, w->perm.p, w->map.p, agg.p, w->co_used.p);
            pdn_comm_allgatherv_i32(ctx, agg.p, w->halo.begin.data());
            PDN_LAUNCH(ctx, agg_used_kernel, pdn_grid(ctx, n, CO_BLK, 8), CO_BLK, 0, n, agg.p, w->co_used.p);
        } else {
            PDN_LAUNCH(ctx, agg_raw_kernel, pdn_grid(ctx, n, CO_BLK, 8), CO_BLK, 0, n, row_lo, row_hi, g, ctx->geo_voff.p,
                       (const double2 *)ctx->geo_xy, w->perm.p, w->map.p, agg.p, w->co_used.p);
        }
        pdn_exclusive_scan_i32(ctx, w->co_used.p, n_raw, w->co_cid.p);
        PDN_CUDA(cudaMemcpyAsync(&nc, w->co_cid.p + n_raw, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        if (nc < 2) return fail("fewer than 2 non-empty aggregates");
        if (nc <= CO_MAX) break;
        if (!local || attempt >= 12) return fail("more non-empty bins than CO_MAX after 12 coarsenings");
        H *= 1.15;              // an uneven partition put more bins on this rank than the table holds: coarser bins
    }
    PDN_LAUNCH(ctx, agg_final_kernel, pdn_grid(ctx, n, CO_BLK, 8), CO_BLK, 0, n, w->co_cid.p, agg.p);

    // fixed-point scale of A_c: |sum| <= entries * largest diagonal must stay below 2^62 (amax: the solve's, the
    // same on every rank)
    const double amax = w->amax;
    if (!(amax > 0.0) || !(amax < 1e300)) return fail("non-positive Jacobi diagonal");
    const double fix = exp2(floor(log2(4.0e18 / (std::max(n_entries_d, 1.0) * amax))));

    const size_t nc2 = (size_t)nc * nc;
    w->co_acc.reserve(nc2);
    co_A.reserve(nc2);
    PDN_CUDA(cudaMemsetAsync(w->co_acc.p, 0, nc2 * sizeof(long long), st));
    PDN_LAUNCH(ctx, coarse_accum_kernel, pdn_grid(ctx, w->n_slices * 32, CO_BLK, 8), CO_BLK, 0, w->A, w->n_slices, agg.p,
               nc, fix, (unsigned long long *)w->co_acc.p);
    // partitioned: the SELL part above held this rank's rows only -- exact integer sum over the ranks; the override
    // rows (every rank holds them all) are added afterwards, once
    if (w->part && !local) pdn_comm_allreduce_sum_i64(ctx, w->co_acc.p, (int)nc2);
    if (w->A.n_special > 0)
        PDN_LAUNCH(ctx, coarse_accum_special_kernel, pdn_blocks(w->A.n_special, CO_BLK), CO_BLK, 0, w->A, agg.p, nc, fix,
                   (unsigned long long *)w->co_acc.p);
    PDN_LAUNCH(ctx, coarse_convert_kernel, pdn_grid(ctx, (int64_t)nc2, CO_BLK, 8), CO_BLK, 0, nc, 1.0 / fix, w->co_acc.p,
               co_A.p, symmetric ? 1 : 0);

    // inverse: cooperative blocked Gauss-Jordan, rows of the matrix resident in shared memory (pdn_coarse.cuh).
    // Round-2 kernel: 12 rows per CTA (PADNE_B200_CO_R=16 for the taller block where it fits), 16 column warps + an
    // inverter warp; PADNE_B200_INVERT=v1 keeps the round-1 kernel (16 rows, 256 threads).
    static const bool invert_v1 = [] { const char *e = getenv("PADNE_B200_INVERT"); return e && strcmp(e, "v1") == 0; }();
    int32_t R;
    if (invert_v1) {
        R = (int32_t)std::min<int64_t>(16, (int64_t)(200 * 1024) / ((int64_t)nc * 8));
        if (const char *e = getenv("PADNE_B200_CO_R")) R = std::min<int32_t>(R, std::max(1, atoi(e)));   // tuning
        R = std::max(R, (nc + ctx->n_sm - 1) / ctx->n_sm);
    } else {
        R = 12;
        if (const char *e = getenv("PADNE_B200_CO_R")) R = atoi(e) == 16 ? 16 : 12;
        if (R == 16 && (int64_t)nc * 16 * 8 > 205 * 1024) R = 12;
    }
    if (R < 1 || R > 16) return fail("inverse: rows per CTA out of range");
    const int grid = (nc + R - 1) / R;
    if (grid > ctx->n_sm) return fail("inverse: more row blocks than SMs");
    const size_t smem = (size_t)R * nc * sizeof(double);
    if (smem > 205 * 1024) return fail("inverse: rows do not fit shared memory");
    // columns per thread: one pass covers the matrix when nc <= 1536, else two passes of two
    const int ncol = nc <= INV_BLK ? 1 : (nc <= 2 * INV_BLK ? 2 : (nc <= 3 * INV_BLK ? 3 : 2));
    const void *kern = (const void *)coarse_invert_kernel;
    if (!invert_v1) {
        if (R == 12)
            kern = ncol == 1 ? (const void *)coarse_invert2_kernel<1, 12>
                 : ncol == 2 ? (const void *)coarse_invert2_kernel<2, 12> : (const void *)coarse_invert2_kernel<3, 12>;
        else
            kern = ncol == 1 ? (const void *)coarse_invert2_kernel<1, 16>
                 : ncol == 2 ? (const void *)coarse_invert2_kernel<2, 16> : (const void *)coarse_invert2_kernel<3, 16>;
    }
    {
        static const void *configured[8] = {};
        const int slot = invert_v1 ? 0 : 1 + (R == 16 ? 3 : 0) + (ncol - 1);
        if (configured[slot] != kern) {
            PDN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(205 * 1024)));
            configured[slot] = kern;
        }
    }
    w->co_piv.reserve(((size_t)grid + 1) * ((size_t)R * nc + 256));
    w->co_counter.reserve(2 * (size_t)grid + 4);
    PDN_CUDA(cudaMemsetAsync(w->co_counter.p, 0, (2 * (size_t)grid + 4) * sizeof(unsigned int), st));
    pdn_status_reset(ctx);
    {
        int32_t nc_arg = nc, R_arg = R;
        double *A_arg = co_A.p, *piv_arg = w->co_piv.p;
        unsigned int *cnt_arg = w->co_counter.p;
        int32_t *st_arg = ctx->d_status;
        co_Af.reserve(nc2);
        float *af_arg = co_Af.p;
        void *args_v1[] = {&nc_arg, &R_arg, &A_arg, &piv_arg, &cnt_arg, &st_arg, &af_arg};
        void *args_v2[] = {&nc_arg, &A_arg, &piv_arg, &cnt_arg, &st_arg, &af_arg};
        PDN_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(invert_v1 ? CO_BLK : INV_THREADS),
                                             invert_v1 ? args_v1 : args_v2, smem, st));
        ctx->launches++;
    }
    int64_t detail = 0;
    const int stt = pdn_status_fetch(ctx, &detail);
    if (stt != 0) {   // not positive definite (or a stuck barrier): fall back to Jacobi, loudly in last_error
        pdn_status_reset(ctx);
        ctx->last_error = std::string("coarse matrix inversion failed (") + pdn_strerror(stt) + "): Jacobi only";
        return fail(stt == PDN_ESINGULAR ? "inverse: non-positive pivot" : "inverse: a CTA missed its deadline");
    }
    rc_fix.reserve((size_t)2 * nc);
    xc.reserve((size_t)nc);
    if (local) {
        w->nc2 = nc;
    } else {
        w->nc = nc;
        w->two_level = true;
    }
    return true;
}

void KernelTimes::report(int rank) {
    if (marks.size() < 2) return;
    cudaEventSynchronize(marks.back().second);
    std::map<std::string, std::pair<double, int>> agg;
    for (size_t i = 1; i < marks.size(); i++) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, marks[i - 1].second, marks[i].second);
        auto &a = agg[marks[i].first];
        a.first += ms;
        a.second++;
    }
    std::string out;
    char buf[160];
    for (auto &kv : agg) {
        snprintf(buf, sizeof buf, "  %-44s n=%4d  mean %.1f us\n", kv.first.c_str(), kv.second.second,
                 1e3 * kv.second.first / kv.second.second);
        out += buf;
    }
    fprintf(stderr, "[padne_b200 ktime rank %d] event-to-event, eager launches:\n%s", rank, out.c_str());
    for (auto &m : marks) cudaEventDestroy(m.second);
    marks.clear();
}

// PADNE_B200_TRACE=1: wall-clock per phase of one solve (each mark drains the stream, so the phases add
// up to more than an untraced solve); printed to stderr by every rank
struct PhaseTrace {
    bool on;
    cudaStream_t st;
    int rank;
    std::chrono::steady_clock::time_point t0;
    std::string out;
    PhaseTrace(cudaStream_t s, int r) : on(getenv("PADNE_B200_TRACE") != nullptr), st(s), rank(r) {
        if (on) { cudaStreamSynchronize(st); t0 = std::chrono::steady_clock::now(); }
    }
    void mark(const char *name) {
        if (!on) return;
        cudaStreamSynchronize(st);
        const auto t1 = std::chrono::steady_clock::now();
        char buf[96];
        snprintf(buf, sizeof buf, " %s=%.3f", name, std::chrono::duration<double, std::milli>(t1 - t0).count());
        out += buf;
        t0 = t1;
    }
    void note(const char *name, long long v) {
        if (!on) return;
        char buf[96];
        snprintf(buf, sizeof buf, " %s=%lld", name, v);
        out += buf;
    }
    ~PhaseTrace() {
        if (on) fprintf(stderr, "[padne_b200 trace rank %d]%s (ms)\n", rank, out.c_str());
    }
};

void pdn_solve_impl(pdn_ctx *ctx, const double *rhs, const pdn_opts *opts_in, double *v_out, pdn_info *info,
                    bool on_device) {
    cudaStream_t st = ctx->stream;
    PhaseTrace trace(st, pdn_comm_rank(ctx));
    SolveWork *w = get_work(ctx);
    const CsrDev &L = ctx->L;
    const int64_t N = L.n;
    pdn_opts opts{};
    if (opts_in) opts = *opts_in;
    else opts.n_nodes = -1;
    const double rtol = opts.rtol > 0.0 ? opts.rtol : 1e-10;
    // absolute bar on ||L v - r||_2 (the reference's own: tests/test_solver.py:2083-2089); < 0 switches it off
    const double atol = opts.atol > 0.0 ? opts.atol : (opts.atol < 0.0 ? 0.0 : 1e-9);
    const int max_iter = opts.max_iter > 0 ? opts.max_iter : (int)(20.0 * sqrt((double)N)) + 2000;
    const int check_every = opts.check_every > 0 ? opts.check_every : 32;
    if (opts.precond != PDN_PRECOND_AUTO && opts.precond != PDN_PRECOND_NONE && opts.precond != PDN_PRECOND_TWOLEVEL &&
        opts.precond != PDN_PRECOND_JACOBI)
        PDN_FAIL(PDN_EINVAL, "solve: unknown preconditioner %d", opts.precond);

    cudaEvent_t ev_start, ev_setup, ev_end;
    PDN_CUDA(cudaEventCreate(&ev_start));
    PDN_CUDA(cudaEventCreate(&ev_setup));
    PDN_CUDA(cudaEventCreate(&ev_end));
    struct EvGuard {
        cudaEvent_t a, b, c;
        ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); cudaEventDestroy(c); }
    } evg{ev_start, ev_setup, ev_end};
    PDN_CUDA(cudaEventRecord(ev_start, st));
    pdn_status_reset(ctx);
    w->have_operator = false;

    // right-hand side on the device
    const double *d_rhs = rhs;
    if (!on_device) {
        w->rhs_dev.reserve((size_t)N);
        PDN_CUDA(cudaMemcpyAsync(w->rhs_dev.p, rhs, N * sizeof(double), cudaMemcpyHostToDevice, st));
        d_rhs = w->rhs_dev.p;
    }

    // n_nodes
    int64_t n = opts.n_nodes;
    if (n < 0) {
        w->tmp_i32a.reserve(4);
        PDN_CUDA(cudaMemsetAsync(w->tmp_i32a.p, 0, 2 * sizeof(int32_t), st));
        PDN_LAUNCH(ctx, last_diag_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, N, L.indptr.p, L.indices.p, w->tmp_i32a.p);
        int32_t n32[2] = {0, 0};
        PDN_CUDA(cudaMemcpyAsync(n32, w->tmp_i32a.p, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        n = std::max(n32[0], n32[1]);
    }
    if (n <= 0 || n > N) PDN_FAIL(PDN_EINVAL, "solve: n_nodes = %lld out of range (N = %lld)", (long long)n, (long long)N);
    const int m = (int)(N - n);
    w->n = n;
    w->N = N;

    // ---- border blocks to the host --------------------------------------------------------------
    std::vector<int32_t> c_ptr(m + 1, 0), c_col, b_row, b_col;
    std::vector<double> c_val, b_val, rhs_tail(m > 0 ? m : 1, 0.0);
    PDN_CUDA(cudaMemcpyAsync(c_ptr.data(), L.indptr.p + n, (m + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    if (m > 0) {
        if (on_device) PDN_CUDA(cudaMemcpyAsync(rhs_tail.data(), rhs + n, m * sizeof(double), cudaMemcpyDeviceToHost, st));
        else memcpy(rhs_tail.data(), rhs + n, m * sizeof(double));
    }
    PDN_CUDA(cudaStreamSynchronize(st));
    const int32_t n_c = c_ptr[m] - c_ptr[0];
    c_col.resize(n_c);
    c_val.resize(n_c);
    if (n_c > 0) {
        PDN_CUDA(cudaMemcpyAsync(c_col.data(), L.indices.p + c_ptr[0], n_c * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaMemcpyAsync(c_val.data(), L.data.p + c_ptr[0], n_c * sizeof(double), cudaMemcpyDeviceToHost, st));
    }
    {
        int32_t cap = std::max(4096, 8 * n_c);
        for (int pass = 0; pass < 2; pass++) {
            w->tmp_i32a.reserve(4);
            w->tmp_i32b.reserve(cap);
            w->tmp_i32c.reserve(cap);
            w->tmp_f64a.reserve(cap);
            PDN_CUDA(cudaMemsetAsync(w->tmp_i32a.p, 0, sizeof(int32_t), st));
            PDN_LAUNCH(ctx, border_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, L.indptr.p, L.indices.p, L.data.p, cap,
                       w->tmp_i32a.p, w->tmp_i32b.p, w->tmp_i32c.p, w->tmp_f64a.p);
            int32_t cnt = 0;
            PDN_CUDA(cudaMemcpyAsync(&cnt, w->tmp_i32a.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            PDN_CUDA(cudaStreamSynchronize(st));
            if (cnt <= cap) {
                b_row.resize(cnt); b_col.resize(cnt); b_val.resize(cnt);
                if (cnt > 0) {
                    PDN_CUDA(cudaMemcpyAsync(b_row.data(), w->tmp_i32b.p, cnt * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
                    PDN_CUDA(cudaMemcpyAsync(b_col.data(), w->tmp_i32c.p, cnt * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
                    PDN_CUDA(cudaMemcpyAsync(b_val.data(), w->tmp_f64a.p, cnt * sizeof(double), cudaMemcpyDeviceToHost, st));
                }
                PDN_CUDA(cudaStreamSynchronize(st));
                break;
            }
            if (pass == 1) PDN_FAIL(PDN_ECUDA, "solve: border extraction overflow");
            cap = cnt;
        }
    }

    trace.mark("border");
    // ---- elimination ------------------------------------------------------------------------------
    Elim E;
    build_elim(E, n, N, c_ptr, c_col, c_val, b_row, b_col, b_val, rhs_tail.data());
    const int nT = (int)E.T.size();
    int method = opts.method;
    if (method == PDN_METHOD_AUTO) method = E.symmetric ? PDN_METHOD_PCG : PDN_METHOD_BICGSTAB;
    if (method == PDN_METHOD_PCG && !E.symmetric)
        PDN_FAIL(PDN_EINVAL, "solve: PCG requested but the regulator coupling makes the system non-symmetric");

    trace.mark("elim_host");
    // map / offsets on the device
    w->map.reserve((size_t)n);
    w->off.reserve((size_t)N);
    PDN_LAUNCH(ctx, iota_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, w->map.p);
    PDN_CUDA(cudaMemsetAsync(w->off.p, 0, N * sizeof(double), st));
    std::vector<int32_t> h_T(E.T.begin(), E.T.end()), h_rep(E.rep.begin(), E.rep.end());
    upload(ctx, w->sp_row_old, h_T);    // original numbering (setup, recovery)
    upload(ctx, w->sp_row, h_T);        // mapped to the permuted numbering once the permutation exists
    upload(ctx, w->tmp_i32b, h_rep);
    upload(ctx, w->tmp_f64a, E.off);
    if (nT > 0) {
        PDN_LAUNCH(ctx, scatter_i32_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row_old.p, w->tmp_i32b.p, w->map.p);
        PDN_LAUNCH(ctx, scatter_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row_old.p, w->tmp_f64a.p, w->off.p);
    }

    // main part of the reduced operator -> SELL-32-sigma (pdn_sell.cuh) + Jacobi diagonal
    const int64_t n_pad = (n + SELL_WIN - 1) / SELL_WIN * SELL_WIN, n_slices = n_pad / SELL_C;
    w->n_pad = n_pad;
    w->n_slices = n_slices;
    w->special.reserve((size_t)n_pad);
    w->len.reserve((size_t)n_pad);
    w->perm.reserve((size_t)n_pad);
    w->inv.reserve((size_t)n_pad);
    w->scnt.reserve((size_t)n_slices);
    w->soff.reserve((size_t)n_slices + 1);
    w->smask.reserve((size_t)n_slices);
    w->dinv.reserve((size_t)n_pad);
    PDN_CUDA(cudaMemsetAsync(w->special.p, 0, (size_t)n_pad, st));
    if (nT > 0) PDN_LAUNCH(ctx, flag_u8_kernel, pdn_blocks(nT, 256), 256, 0, nT, w->sp_row_old.p, w->special.p);
    PDN_LAUNCH(ctx, sell_len_kernel, pdn_grid(ctx, n_pad, 256, 8), 256, 0, n, n_pad, L.indptr.p, L.indices.p, w->map.p,
               w->special.p, w->len.p);
    // partitioned setup: the row ranges of the ranks (whole sort windows), known before the operator exists
    const int world = pdn_comm_world(ctx), me = pdn_comm_rank(ctx);
    std::vector<int64_t> part_begin(world + 1, 0);
    pdn_row_partition(n, world, part_begin.data());
    w->part = world > 1 && ctx->part_valid;
    if (w->part && (ctx->part_n_nodes != n || ctx->part_lo != part_begin[me] || ctx->part_hi != part_begin[me + 1]))
        PDN_FAIL(PDN_EINVAL, "solve: the system was assembled for another partition (n_nodes %lld, rows %lld..%lld)",
                 (long long)ctx->part_n_nodes, (long long)ctx->part_lo, (long long)ctx->part_hi);
    if (world > 1 && !w->part && ctx->part_valid) PDN_FAIL(PDN_EINVAL, "solve: partitioned system without a communicator");
    const int64_t own_win0 = w->part ? part_begin[me] / SELL_WIN : 0;
    const int64_t own_win1 = w->part ? (me == world - 1 ? n_pad / SELL_WIN : part_begin[me + 1] / SELL_WIN) : n_pad / SELL_WIN;
    if (own_win1 > own_win0)
        PDN_LAUNCH(ctx, sell_sort_kernel, (int)(own_win1 - own_win0), SELL_WIN, 0, w->len.p, w->perm.p, w->inv.p, own_win0);
    if (w->part) {
        // every rank sorted its own windows (it alone knows its rows' lengths): all-gather the permutation
        std::vector<int64_t> bp(part_begin);
        bp[world] = n_pad;
        pdn_comm_allgatherv_i32(ctx, w->perm.p, bp.data());
        pdn_comm_allgatherv_i32(ctx, w->inv.p, bp.data());
    }
    PDN_LAUNCH(ctx, sell_width_kernel, pdn_grid(ctx, n_slices, 256, 8), 256, 0, n_slices, w->len.p, w->perm.p, w->scnt.p);
    pdn_exclusive_scan_i32(ctx, w->scnt.p, n_slices, w->soff.p);
    int32_t sell_total = 0;
    PDN_CUDA(cudaMemcpyAsync(&sell_total, w->soff.p + n_slices, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    if (sell_total < 0) PDN_FAIL(PDN_EUNSUPPORTED, "solve: padded operator exceeds int32 indexing");
    w->s_ci.reserve((size_t)sell_total + 64);
    w->s_va.reserve((size_t)sell_total + 64);
    {
        const int64_t fs0 = own_win0 * (SELL_WIN / SELL_C), fs1 = own_win1 * (SELL_WIN / SELL_C);
        if (w->part) {     // the slices of other ranks: no override mask, unit diagonal
            PDN_CUDA(cudaMemsetAsync(w->smask.p, 0, (size_t)n_slices * sizeof(uint32_t), st));
            PDN_LAUNCH(ctx, ones_kernel, pdn_grid(ctx, n_pad, BLK, 8), BLK, 0, n_pad, w->dinv.p);
        }
        if (fs1 > fs0)
            PDN_LAUNCH(ctx, sell_fill_kernel, pdn_grid(ctx, (fs1 - fs0) * 32, 256, 8), 256, 0, n, fs1, L.indptr.p, L.indices.p,
                       L.data.p, w->map.p, w->special.p, w->perm.p, w->inv.p, w->soff.p, w->s_ci.p, w->s_va.p, w->smask.p,
                       w->dinv.p, fs0);
    }

    trace.mark("sell");
    // g = r_n - L_nn o
    w->full_a.reserve((size_t)N);
    w->full_b.reserve((size_t)N);
    reset_scal(ctx, w, 0.0, 0.0, 0);
    launch_spmv<RESID>(ctx, full_operator(ctx, n), w->off.p, w->full_a.p, d_rhs, w);

    // rows of T and g_T to the host
    std::vector<int32_t> t_ptr(nT + 1, 0), t_ci;
    std::vector<double> t_va, g_T(nT > 0 ? nT : 1, 0.0);
    if (nT > 0) {
        w->tmp_i32c.reserve(nT + 1);
        PDN_LAUNCH(ctx, rowlen_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, n, w->sp_row_old.p, L.indptr.p, L.indices.p, w->tmp_i32c.p);
        std::vector<int32_t> len(nT);
        PDN_CUDA(cudaMemcpyAsync(len.data(), w->tmp_i32c.p, nT * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        w->tmp_f64b.reserve(nT);
        PDN_LAUNCH(ctx, gather_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row_old.p, w->full_a.p, w->tmp_f64b.p);
        PDN_CUDA(cudaMemcpyAsync(g_T.data(), w->tmp_f64b.p, nT * sizeof(double), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        for (int t = 0; t < nT; t++) t_ptr[t + 1] = t_ptr[t] + len[t];
        const int32_t tot = t_ptr[nT];
        t_ci.resize(tot);
        t_va.resize(tot);
        DevBuf<int32_t> &d_tptr = w->t_ptr_d, &d_tci = w->t_ci_d;
        DevBuf<double> &d_tva = w->t_va_d;
        upload(ctx, d_tptr, t_ptr);
        d_tci.reserve(tot + 1);
        d_tva.reserve(tot + 1);
        PDN_LAUNCH(ctx, rowget_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, n, w->sp_row_old.p, L.indptr.p, L.indices.p, L.data.p,
                   d_tptr.p, d_tci.p, d_tva.p);
        if (tot > 0) {
            PDN_CUDA(cudaMemcpyAsync(t_ci.data(), d_tci.p, tot * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            PDN_CUDA(cudaMemcpyAsync(t_va.data(), d_tva.p, tot * sizeof(double), cudaMemcpyDeviceToHost, st));
        }
        PDN_CUDA(cudaStreamSynchronize(st));
    }

    // override rows R = -F L'[T,:], b_T = -F g_T, their Jacobi diagonal
    std::vector<int32_t> r_ptr(nT + 1, 0), r_ci;
    std::vector<double> r_va, b_T(nT > 0 ? nT : 1, 0.0), dinv_T(nT > 0 ? nT : 1, 1.0);
    for (int lt = 0; lt < nT; lt++) {
        const int32_t row = E.T[lt];
        std::map<int32_t, double> acc;
        if (E.rep[lt] == row) {
            double bt = 0.0;
            for (int lj = 0; lj < nT; lj++) {
                const double f = E.F[(size_t)lt * nT + lj];
                if (f == 0.0) continue;
                bt -= f * g_T[lj];
                for (int32_t k = t_ptr[lj]; k < t_ptr[lj + 1]; k++) {
                    int32_t c = t_ci[k];
                    auto it = E.tloc.find(c);
                    if (it != E.tloc.end()) c = E.rep[it->second];
                    if (c < 0) continue;  // fixed column: already in g
                    acc[c] -= f * t_va[k];
                }
            }
            b_T[lt] = bt;
            auto d = acc.find(row);
            if (d == acc.end() || !(d->second != 0.0))
                PDN_FAIL(PDN_ESINGULAR, "solve: zero diagonal in the reduced row of node %d", row);
            dinv_T[lt] = 1.0 / d->second;
        } else {
            acc[row] = 1.0;  // slave or fixed node: u = 0
            b_T[lt] = 0.0;
            dinv_T[lt] = 1.0;
        }
        for (auto &kv : acc) { r_ci.push_back(kv.first); r_va.push_back(kv.second); }
        r_ptr[lt + 1] = (int32_t)r_ci.size();
    }
    upload(ctx, w->sp_rp, r_ptr);
    upload(ctx, w->sp_ci, r_ci);
    upload(ctx, w->sp_va, r_va);
    upload(ctx, w->tmp_f64a, b_T);
    upload(ctx, w->tmp_f64b, dinv_T);
    w->b.reserve((size_t)n);
    PDN_LAUNCH(ctx, permute_negate_kernel, pdn_grid(ctx, n, 256, 8), 256, 0, n, w->perm.p, w->full_a.p, w->b.p);
    if (nT > 0) {
        // override rows move to the permuted numbering
        PDN_LAUNCH(ctx, remap_i32_kernel, pdn_blocks(nT, 256), 256, 0, nT, w->inv.p, w->sp_row.p);
        if (!r_ci.empty())
            PDN_LAUNCH(ctx, remap_i32_kernel, pdn_blocks((int64_t)r_ci.size(), 256), 256, 0, (int32_t)r_ci.size(),
                       w->inv.p, w->sp_ci.p);
        PDN_LAUNCH(ctx, scatter_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->tmp_f64a.p, w->b.p);
        PDN_LAUNCH(ctx, scatter_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->tmp_f64b.p, w->dinv.p);
    }
    PDN_LAUNCH(ctx, check_dinv_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, w->dinv.p, ctx->d_status);
    // largest diagonal of the reduced operator: scale of the K1 restriction and of the residual floor
    std::vector<double> h_dmin;
    {
        const int mgrid = pdn_grid(ctx, n, CO_BLK, 4);
        w->co_tmp.reserve((size_t)mgrid + 8);
        PDN_LAUNCH(ctx, min_kernel, mgrid, CO_BLK, 0, n, w->dinv.p, w->co_tmp.p);
        h_dmin.resize((size_t)mgrid);
        PDN_CUDA(cudaMemcpyAsync(h_dmin.data(), w->co_tmp.p, h_dmin.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    }
    if (opts.precond == PDN_PRECOND_NONE) PDN_LAUNCH(ctx, ones_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, w->dinv.p);
    {
        int64_t detail = 0;
        int stt = pdn_status_fetch(ctx, &detail);
        if (stt != 0) {
            pdn_status_reset(ctx);
            int32_t node = -1;
            if (detail >= 0 && detail < n)
                PDN_CUDA(cudaMemcpy(&node, w->perm.p + detail, sizeof(int32_t), cudaMemcpyDeviceToHost));
            PDN_FAIL(stt, "solve: node %d has no conductance to any other node (%s)", node, pdn_strerror(stt));
        }
        double dmin = 1e300;
        for (double v : h_dmin) dmin = std::min(dmin, v);
        w->amax = dmin > 0.0 && dmin < 1e300 ? 1.0 / dmin : 0.0;
        if (w->part) {     // every rank saw its own rows' diagonals only
            w->tmp_f64a.reserve(4);
            PDN_CUDA(cudaMemcpyAsync(w->tmp_f64a.p, &w->amax, sizeof(double), cudaMemcpyHostToDevice, st));
            pdn_comm_allreduce_max(ctx, w->tmp_f64a.p, 1);
            PDN_CUDA(cudaMemcpyAsync(&w->amax, w->tmp_f64a.p, sizeof(double), cudaMemcpyDeviceToHost, st));
            PDN_CUDA(cudaStreamSynchronize(st));
        }
    }
    w->n_special = nT;
    w->A.n = n;
    w->A.soff = w->soff.p;
    w->A.smask = w->smask.p;
    w->A.ci = w->s_ci.p;
    w->A.va = w->s_va.p;
    w->A.n_special = nT;
    w->A.sp_row = w->sp_row.p;
    w->A.sp_rp = w->sp_rp.p;
    w->A.sp_ci = w->sp_ci.p;
    w->A.sp_va = w->sp_va.p;
    trace.mark("override_rows");
    trace.note("n_special", (long long)nT);
    trace.note("special_nnz", (long long)r_ci.size());
    // rows of the Krylov loop this rank owns: whole sort windows, split evenly
    {
        const int world = pdn_comm_world(ctx), me = pdn_comm_rank(ctx);
        const int64_t n_win = n_pad / SELL_WIN;
        HaloPlan &h = w->halo;
        h.world = world;
        h.me = me;
        h.begin.assign(world + 1, 0);
        for (int k = 0; k <= world; k++) h.begin[k] = std::min<int64_t>(n, (int64_t)k * n_win / world * SELL_WIN);
        h.begin[world] = n;
        w->row_begin = w->A.row_begin = h.begin[me];
        w->row_end = w->A.row_end = h.begin[me + 1];
        if (world > 1) build_halo_plan(ctx, w, (int32_t)r_ci.size());
    }
    w->have_operator = true;
    trace.mark("halo_plan");
    // SpMV kernel of the Krylov loop: the per-warp TMA ring (pdn_sell.cuh).  PADNE_B200_SPMV=xs selects the
    // x-band staging kernel (pdn_cg2.cuh), measured SLOWER at 1M rows in round 2 (31.9 against 26.0 us in situ:
    // the CTA waits for its whole 50 KB window before the first gather, and the per-entry window test costs issue
    // slots); kept for A/B runs.
    {
        const char *e = getenv("PADNE_B200_SPMV");
        w->xs = e && strcmp(e, "xs") == 0;
        if (w->xs) plan_xs(ctx, w);
        trace.mark("xs_plan");
    }

    // second level of the preconditioner (systems assembled from meshes: the aggregates are coordinate bins)
    w->two_level = false;
    w->nc = 0;
    w->nc2 = 0;
    w->coarse_ms = 0.f;
    {
        const bool want = opts.precond == PDN_PRECOND_TWOLEVEL || (opts.precond == PDN_PRECOND_AUTO && n >= 1024);
        if (opts.precond == PDN_PRECOND_TWOLEVEL && !ctx->geo_valid)
            PDN_FAIL(PDN_EUNSUPPORTED, "solve: the two-level preconditioner needs a system assembled by pdn_assemble");
        if (want && ctx->geo_valid) {
            cudaEvent_t c0, c1;
            PDN_CUDA(cudaEventCreate(&c0));
            PDN_CUDA(cudaEventCreate(&c1));
            PDN_CUDA(cudaEventRecord(c0, st));
            bool ok = false;
            try {
                ok = build_coarse(ctx, w, opts.n_coarse, (int64_t)sell_total + (int64_t)r_ci.size(), E.symmetric);
            } catch (...) {
                cudaEventDestroy(c0);
                cudaEventDestroy(c1);
                throw;
            }
            PDN_CUDA(cudaEventRecord(c1, st));
            PDN_CUDA(cudaEventSynchronize(c1));
            cudaEventElapsedTime(&w->coarse_ms, c0, c1);
            cudaEventDestroy(c0);
            cudaEventDestroy(c1);
            if (!ok && opts.precond == PDN_PRECOND_TWOLEVEL)
                PDN_FAIL(PDN_EUNSUPPORTED, "solve: no coarse space could be built (%s)", ctx->last_error.c_str());
            // several GPUs, PCG: a second, rank-local level keeps the bins as fine as on one GPU (the global level
            // is capped at CO_MAX aggregates for the whole board).  PADNE_B200_LOCAL_LEVEL: 0 off, 1 on, "try" on but
            // optional; default: on from 4 ranks.  Every rank must take the same decision, so the outcome is summed
            // over the ranks: unless optional, ONE rank that cannot build its level fails the solve on ALL ranks,
            // with the reason -- round 1 fell back silently and lost 1.8x in iterations at 8 GPUs.
            w->nc2 = 0;
            w->local_level = 0;
            const char *le = getenv("PADNE_B200_LOCAL_LEVEL");
            const bool optional = le && strcmp(le, "try") == 0;
            const bool want_local = le ? le[0] != '0' : pdn_comm_world(ctx) >= 4;
            if (ok && is_dist(ctx) && method == PDN_METHOD_PCG && want_local) {
                const auto t_loc0 = std::chrono::steady_clock::now();
                const bool built = build_coarse(ctx, w, opts.n_coarse, (int64_t)sell_total + (int64_t)r_ci.size(), E.symmetric,
                                                /*local=*/true);
                const std::string note = built ? std::string() : w->coarse_note;
                w->tmp_f64a.reserve(4);
                const double mine = built ? 0.0 : 1.0;
                PDN_CUDA(cudaMemcpyAsync(w->tmp_f64a.p, &mine, sizeof(double), cudaMemcpyHostToDevice, st));
                pdn_comm_allreduce_sum(ctx, w->tmp_f64a.p, 1);
                double failed = 0.0;
                PDN_CUDA(cudaMemcpyAsync(&failed, w->tmp_f64a.p, sizeof(double), cudaMemcpyDeviceToHost, st));
                PDN_CUDA(cudaStreamSynchronize(st));
                if (failed > 0.0) {
                    fprintf(stderr, "[padne_b200 rank %d] rank-local coarse level: %d of %d ranks could not build it%s%s\n",
                            pdn_comm_rank(ctx), (int)failed, pdn_comm_world(ctx), built ? "" : " -- this rank: ", note.c_str());
                    if (!optional) {
                        PDN_FAIL(PDN_EUNSUPPORTED, "solve: the rank-local coarse level failed on %d of %d ranks (%s); "
                                 "PADNE_B200_LOCAL_LEVEL=try makes it optional, =0 switches it off", (int)failed,
                                 pdn_comm_world(ctx), built ? "built here" : note.c_str());
                    }
                    w->nc2 = 0;               // optional: every rank drops it
                    w->local_level = -1;
                } else {
                    w->local_level = 1;
                }
                // build_coarse ends with a stream synchronisation, so wall clock is the device time here
                w->coarse_ms += (float)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_loc0).count();
            }
        }
    }
    // Round-2 iteration (three kernels, ONE exchange of dots + coarse sums per iteration, pdn_cg2.cuh): opt-in with
    // PADNE_B200_PCG=v2.  Measured (round 2, B200): once the round-1 update kernel restricted with two-limb shared
    // atomics instead of warp collectives, the four-kernel iteration is the faster one at every GPU count -- 43.4
    // against 54 us per iteration on one GPU (1M DOF), 74.8 against 76.3 on two, 90.8 against 100.1 on four: the
    // restriction + recurrence cost the SpMV 6 us, the fused update moves the same ~100 B per row through L2 as
    // the two kernels it replaces, and the exchange saved (one of three) is worth 6-10 us.
    {
        const char *e = getenv("PADNE_B200_PCG");
        const bool can_v2 = method == PDN_METHOD_PCG && w->two_level && (!is_dist(ctx) || pdn_peer_enabled(ctx));
        const bool want_v2 = e && strcmp(e, "v2") == 0;
        w->v2 = can_v2 && want_v2;
        if (is_dist(ctx) && w->two_level) {
            // the range of aggregates this rank's rows touch: all its coarse-sum exchanges carry (both iterations)
            const int32_t init[2] = {0x7fffffff, -1};
            w->agg_range.reserve(8);
            PDN_CUDA(cudaMemcpyAsync(w->agg_range.p, init, sizeof init, cudaMemcpyHostToDevice, st));
            PDN_LAUNCH(ctx, agg_range_kernel, pdn_grid(ctx, std::max<int64_t>(w->row_end - w->row_begin, 1), 256, 4), 256, 0,
                       w->row_begin, w->row_end, w->agg.p, w->agg_range.p);
        }
        w->halo_fused = halo_in_coarse_z(ctx, w) && (w->halo.n_send > 0 || w->halo.n_recv > 0);
        if (w->v2) {
            w->tab.reserve((size_t)w->nc * TAB_STRIDE + 8);
            w->tab2.reserve((size_t)std::max(w->nc2, 1) * TAB_STRIDE + 8);
            const size_t c1 = (size_t)2 * w->nc + 8, c2 = (size_t)2 * std::max(w->nc2, 1) + 8;
            w->rc_d.reserve(c1); w->cs_d.reserve(c1); w->cw_d.reserve(c1);
            w->rc2_d.reserve(c2); w->cs2_d.reserve(c2);
            PDN_CUDA(cudaMemsetAsync(w->rc_d.p, 0, c1 * sizeof(double), st));
            PDN_CUDA(cudaMemsetAsync(w->cs_d.p, 0, c1 * sizeof(double), st));
            PDN_CUDA(cudaMemsetAsync(w->cw_d.p, 0, c1 * sizeof(double), st));
            PDN_CUDA(cudaMemsetAsync(w->rc2_d.p, 0, c2 * sizeof(double), st));
            PDN_CUDA(cudaMemsetAsync(w->cs2_d.p, 0, c2 * sizeof(double), st));
            if (w->nc2 == 0) { w->rc_fix2.reserve(8); w->xc2.reserve(8); w->co_Af2.reserve(8); }
            // restriction windows of the ring kernel's CTAs (same grid, same slice split)
            if (!w->xs) {
                const int64_t slices = (w->row_end - w->row_begin + SELL_C - 1) / SELL_C;
                const int rgrid = is_dist(ctx) ? ring_grid<CG2, true>(ctx, slices) : ring_grid<CG2, false>(ctx, slices);
                w->cta_amin.reserve((size_t)rgrid + 8);
                w->cta_amin2.reserve((size_t)rgrid + 8);
                PDN_LAUNCH(ctx, agg_min_kernel, rgrid, 256, 0, w->row_begin, w->row_end, w->agg.p, w->cta_amin.p);
                if (w->nc2 > 0)
                    PDN_LAUNCH(ctx, agg_min_kernel, rgrid, 256, 0, w->row_begin, w->row_end, w->agg2.p, w->cta_amin2.p);
            }
            PDN_CUDA(cudaMemsetAsync(w->tab.p, 0, ((size_t)w->nc * TAB_STRIDE + 8) * sizeof(unsigned long long), st));
            PDN_CUDA(cudaMemsetAsync(w->tab2.p, 0, ((size_t)std::max(w->nc2, 1) * TAB_STRIDE + 8) * sizeof(unsigned long long), st));
        }
    }

    trace.mark("coarse");
    // vectors.  The SpMV inputs (z, ph, sh, x) are the halo-exchanged ones: with peer-memory exchanges
    // they live in the arena every peer has mapped
    {
        DevBuf<double> *ex[PDN_PA_NVEC] = {&w->z, &w->ph, &w->sh, &w->x};
        if (pdn_peer_enabled(ctx)) {
            pdn_peer_ensure(ctx, n);     // collective (same n on every rank)
            for (int v = 0; v < PDN_PA_NVEC; v++) ex[v]->attach(pdn_peer_vec(ctx, v), (size_t)n);
        } else {
            for (int v = 0; v < PDN_PA_NVEC; v++)
                if (!ex[v]->owned) ex[v]->release();
        }
    }
    w->x.reserve((size_t)n); w->r.reserve((size_t)n); w->z.reserve((size_t)n); w->p.reserve((size_t)n);
    w->s.reserve((size_t)n); w->w.reserve((size_t)n);
    if (method == PDN_METHOD_BICGSTAB) {
        w->rhat.reserve((size_t)n); w->vv.reserve((size_t)n); w->sh.reserve((size_t)n);
        w->ph.reserve((size_t)n); w->tt.reserve((size_t)n);
    }
    PDN_CUDA(cudaEventRecord(ev_setup, st));
    trace.mark("vectors");

    // ---- Krylov -------------------------------------------------------------------------------------
    KrylovResult kr = run_krylov(ctx, w, method, KrylovTarget{rtol, atol}, max_iter, check_every);

    trace.mark("krylov");
    if (is_dist(ctx)) {   // a peer that never answered (pdn_peer.cuh deadline)
        int64_t detail = 0;
        const int stt = pdn_status_fetch(ctx, &detail);
        if (stt != 0) {
            pdn_status_reset(ctx);
            PDN_FAIL(stt, "solve: a peer rank did not answer within the deadline (%s)", pdn_strerror(stt));
        }
    }

    // ---- recovery: v = Z u + o, x = M^-1 C (r_n - L_nn v), full residual -------------------------------
    double *d_v = on_device ? v_out : w->full_b.p;
    PDN_LAUNCH(ctx, expand_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, n, N, w->map.p, w->inv.p, w->off.p, w->x.p, d_v);
    std::vector<double> xcur(m > 0 ? m : 1, 0.0);
    if (m > 0) {
        reset_scal(ctx, w, 0.0, 0.0, 0);
        launch_spmv<RESID>(ctx, full_operator(ctx, n), d_v, w->full_a.p, d_rhs, w);
        std::vector<double> h_T(nT > 0 ? nT : 1, 0.0);
        if (nT > 0) {
            w->tmp_f64b.reserve(nT);
            PDN_LAUNCH(ctx, gather_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row_old.p, w->full_a.p, w->tmp_f64b.p);
            PDN_CUDA(cudaMemcpyAsync(h_T.data(), w->tmp_f64b.p, nT * sizeof(double), cudaMemcpyDeviceToHost, st));
            PDN_CUDA(cudaStreamSynchronize(st));
        }
        std::vector<double> ch(m, 0.0);
        for (int r = 0; r < m; r++)
            for (const Term &t : E.C[r]) ch[r] += t.coef * h_T[t.loc];
        for (int r = 0; r < m; r++) {
            double acc = 0.0;
            for (int c = 0; c < m; c++) acc += E.Minv[(size_t)r * m + c] * ch[c];
            xcur[r] = acc;
        }
        PDN_CUDA(cudaMemcpyAsync(d_v + n, xcur.data(), m * sizeof(double), cudaMemcpyHostToDevice, st));
    }
    reset_scal(ctx, w, 0.0, 0.0, 0);
    launch_spmv<RESID>(ctx, full_operator(ctx, N), d_v, w->full_a.p, d_rhs, w);
    fetch_scal(ctx, w);
    double residual_norm = sqrt(w->h_sc->rr);
    if (w->part) {
        // this rank's L holds its own node rows, the terminals and the extra rows: ||L v - r||^2 = the own node rows
        // summed over the ranks + the extra rows (identical on every rank, counted once)
        double parts[2] = {0.0, 0.0};
        const int rgrid = pdn_grid(ctx, std::max<int64_t>(w->row_end - w->row_begin, 1), RPB, 8);
        PDN_LAUNCH(ctx, sumsq_kernel<true>, rgrid, RPB, 0, w->row_begin, w->row_end, w->full_a.p, (const double *)nullptr,
                   w->d_sc, w->d_out, w->partials.p);
        dist_reduce<7>(ctx, w);
        PDN_CUDA(cudaMemcpyAsync(&parts[0], w->d_out, sizeof(double), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        if (N > n) {
            PDN_LAUNCH(ctx, sumsq_kernel<false>, pdn_grid(ctx, N - n, RPB, 8), RPB, 0, n, N, w->full_a.p, (const double *)nullptr,
                       w->d_sc, w->d_out, w->partials.p);
            PDN_CUDA(cudaMemcpyAsync(&parts[1], w->d_out, sizeof(double), cudaMemcpyDeviceToHost, st));
            PDN_CUDA(cudaStreamSynchronize(st));
        }
        residual_norm = sqrt(parts[0] + parts[1]);
    }
    if (!on_device) PDN_CUDA(cudaMemcpyAsync(v_out, d_v, N * sizeof(double), cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaEventRecord(ev_end, st));
    PDN_CUDA(cudaEventSynchronize(ev_end));
    if (info) {
        float t_setup = 0.f, t_total = 0.f;
        cudaEventElapsedTime(&t_setup, ev_start, ev_setup);
        cudaEventElapsedTime(&t_total, ev_start, ev_end);
        memset(info, 0, sizeof *info);
        info->method = method;
        info->iterations = kr.iterations;
        info->restarts = kr.restarts;
        info->n_constraints = m;
        info->rel_residual = kr.rel_residual;
        info->residual_norm = residual_norm;
        info->ground_node_current = m > 0 ? xcur[m - 1] : 0.0;
        info->t_setup_ms = t_setup;
        info->t_krylov_ms = kr.ms;
        info->t_total_ms = t_total;
        info->spmv_ms = 0.0;
        info->precond = w->two_level ? PDN_PRECOND_TWOLEVEL
                                     : (opts.precond == PDN_PRECOND_NONE ? PDN_PRECOND_NONE : PDN_PRECOND_JACOBI);
        info->n_coarse = w->two_level ? w->nc : 0;
        info->t_coarse_ms = w->coarse_ms;
        info->n_coarse_local = w->nc2;
        info->local_level = w->local_level;
        info->stagnated = kr.stagnated;
        info->iteration_kind = w->v2 && method == PDN_METHOD_PCG ? 2 : 1;
        info->abs_target = kr.abs_target;
        info->floor_abs = kr.floor_abs;
        info->rhs_norm = kr.bnorm;
        info->spmv_kind = w->xs ? 1 : 0;
        info->spmv_halo = w->xs ? w->xsp.halo : 0;
        info->spmv_window_coverage = w->xs ? w->xs_coverage : 0.0;
    }
}

// ------------------------------------------------------------------------------------------------
// diagnostics
// ------------------------------------------------------------------------------------------------
void pdn_spmv_impl(pdn_ctx *ctx, const double *x, double *y, bool on_device) {
    SolveWork *w = get_work(ctx);
    const int64_t N = ctx->L.n;
    DevIn<double> dx;
    dx.set(x, (size_t)N, on_device, ctx->stream);
    DevBuf<double> dy;
    double *yp = on_device ? y : dy.reserve((size_t)N);
    launch_spmv<PLAIN>(ctx, full_operator(ctx, N), dx.p, yp, nullptr, w);
    if (!on_device) PDN_CUDA(cudaMemcpyAsync(y, yp, N * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    PDN_CUDA(cudaStreamSynchronize(ctx->stream));
}

// Times the kernel that dominates the solve: the CG-mode SELL SpMV (fused r.z, z.w, r.r) on the reduced
// operator of the last pdn_solve.  (Before any solve there is no reduced operator: the CSR-stream
// kernel on the full system is timed instead.)
//   mode 0: back-to-back launches (warm L2)
//   mode 1: a >L2 buffer is written before every launch (cold L2)
//   mode 2: in situ -- every SpMV is followed by the PCG vector update, exactly like one solver
//           iteration, and only the SpMV is bracketed by events (the cache state of the real solve)
void pdn_bench_spmv_impl(pdn_ctx *ctx, int32_t reps, int mode, double *ms) {
    SolveWork *w = get_work(ctx);
    cudaStream_t st = ctx->stream;
    const bool sell = w->have_operator;   // multi-GPU: this rank's rows only, no collectives
    Operator Afull = full_operator(ctx, ctx->L.n);
    const int64_t n = sell ? w->n : Afull.n;
    DevBuf<double> x, y, aux, v3, v4, v5, v6;
    x.reserve((size_t)n); y.reserve((size_t)n); aux.reserve((size_t)n);
    const int vgrid = pdn_grid(ctx, n, BLK, 8);
    PDN_LAUNCH(ctx, ones_kernel, vgrid, BLK, 0, n, x.p);
    PDN_LAUNCH(ctx, ones_kernel, vgrid, BLK, 0, n, aux.p);
    if (mode == 2) {
        v3.reserve((size_t)n); v4.reserve((size_t)n); v5.reserve((size_t)n); v6.reserve((size_t)n);
        PDN_LAUNCH(ctx, ones_kernel, vgrid, BLK, 0, n, v3.p);
        PDN_LAUNCH(ctx, ones_kernel, vgrid, BLK, 0, n, v4.p);
        PDN_LAUNCH(ctx, ones_kernel, vgrid, BLK, 0, n, v5.p);
        PDN_LAUNCH(ctx, ones_kernel, vgrid, BLK, 0, n, v6.p);
    }
    // max_iter < 0 tells the kernel's scalar epilogue to leave `done` alone (benchmark mode), so all
    // launches can be queued back to back; every SpMV gets its own event pair and the host
    // synchronises once at the end, like the kernels inside the solver's CUDA graph.
    reset_scal(ctx, w, 0.0, 0.0, -1);
    const int64_t r0 = sell ? w->row_begin : 0, r1 = sell ? w->row_end : n;
    // the kernel exactly as the last solve ran it: CG2 (restriction + recurrence tail) for the round-2 iteration.
    // Peer exchanges are left out (null link): this rank's launch only.
    const bool v2 = sell && w->v2;
    const Cg2 cg = v2 ? make_cg2(w) : Cg2{};
    if (v2) {
        w->h_sc->amax = w->amax;
        w->h_sc->zmax_bits[0] = w->h_sc->zmax_bits[1] = (unsigned long long)0x3ff0000000000000ull;   // max|z| = 1.0
        PDN_CUDA(cudaMemcpyAsync(w->d_sc, w->h_sc, sizeof(Scal), cudaMemcpyHostToDevice, st));
    }
    auto launch = [&]() {
        if (sell && w->xs) {
            const int64_t slices = (w->A.row_end - w->A.row_begin + SELL_C - 1) / SELL_C;
            if (v2) {
                PDN_CUDA(cudaFuncSetAttribute(sell_xs_kernel<CG2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SX_SMEM));
                PDN_LAUNCH(ctx, (sell_xs_kernel<CG2, false>), xs_grid(ctx, w, slices), SX_BLK, SX_SMEM, w->A, w->xsp, x.p, y.p, aux.p,
                           w->d_sc, w->partials.p, (const PeerLink *)nullptr, cg);
            } else {
                PDN_CUDA(cudaFuncSetAttribute(sell_xs_kernel<CG, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SX_SMEM));
                PDN_LAUNCH(ctx, (sell_xs_kernel<CG, false>), xs_grid(ctx, w, slices), SX_BLK, SX_SMEM, w->A, w->xsp, x.p, y.p, aux.p,
                           w->d_sc, w->partials.p, (const PeerLink *)nullptr, cg);
            }
        } else if (sell && v2) {
            launch_sell_kernel<CG2, false>(ctx, w, x.p, y.p, aux.p, cg);
        } else if (sell) {
            launch_sell_kernel<CG, false>(ctx, w, x.p, y.p, aux.p);
        } else {
            launch_spmv<CG>(ctx, Afull, x.p, y.p, aux.p, w);
        }
    };
    for (int k = 0; k < 3; k++) launch();
    std::vector<cudaEvent_t> ev(2 * (size_t)reps);
    for (auto &e : ev) PDN_CUDA(cudaEventCreate(&e));
    for (int k = 0; k < reps; k++) {
        if (mode == 1) pdn_flush_l2(ctx);
        PDN_CUDA(cudaEventRecord(ev[2 * k], st));
        launch();
        PDN_CUDA(cudaEventRecord(ev[2 * k + 1], st));
        if (mode == 2 && v2) {   // the rest of a round-2 iteration: K2, K3 (dinv = v6, w = y, z = x, p = v3, s = v4, x = v5, r = aux)
            launch_coarse_apply2<true>(ctx, w, false);
            PDN_LAUNCH(ctx, (cg_update2_kernel<RPB, false>), pdn_grid(ctx, r1 - r0, RPB, 8), RPB, 0, r0, r1, w->d_sc, v6.p, y.p, x.p,
                       v3.p, v4.p, v5.p, aux.p, w->agg.p, w->xc.p, w->nc2 > 0 ? w->agg2.p : (const int32_t *)nullptr,
                       w->nc2 > 0 ? w->xc2.p : (const double *)nullptr, (const PeerLink *)nullptr, (const HaloSend *)nullptr,
                       (size_t)0, (const int32_t *)nullptr, (unsigned int *)nullptr, w->tab.p, w->nc, w->tab2.p, w->nc2);
        } else if (mode == 2)   // dinv = v6, w = y, z = x, p = v3, s = v4, x = v5, r = aux
            PDN_LAUNCH(ctx, cg_update_kernel<false>, pdn_grid(ctx, r1 - r0, RPB, 8), RPB, 0, r0, r1, w->d_sc, v6.p, y.p, x.p,
                       v3.p, v4.p, v5.p, aux.p, (const int32_t *)nullptr, (long long *)nullptr, 0);
    }
    PDN_CUDA(cudaStreamSynchronize(st));
    double total = 0.0;
    for (int k = 0; k < reps; k++) {
        float t = 0.f;
        cudaEventElapsedTime(&t, ev[2 * k], ev[2 * k + 1]);
        total += t;
    }
    for (auto &e : ev) cudaEventDestroy(e);
    *ms = total / reps;
}
