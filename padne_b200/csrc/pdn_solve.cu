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
// regulator sense nodes), so A is stored as L_nn with remapped columns (device kernel) plus a tiny
// override CSR for the rows in T (built on the host from the |T| fetched rows).  A is SPD when
// B = C^T (Jacobi-PCG, Chronopoulos-Gear single-reduction form) and non-symmetric otherwise
// (Jacobi-BiCGSTAB).  After convergence v = Z u + o, x is recovered, and ||L v - r||_2 of the
// FULL system is reported exactly as solver.py:775.
#include <math.h>

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
__global__ void __launch_bounds__(BLK) flag_rows_kernel(int32_t k, const int32_t *__restrict__ idx, int32_t *rp) {
    for (int32_t i = blockIdx.x * BLK + threadIdx.x; i < k; i += gridDim.x * BLK) rp[idx[i]] = ~rp[idx[i]];
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

// A_main = -L_nn with columns mapped through Z; fixed / extra columns are neutralised in place
// (value 0, column = own row) so the pattern and the row pointers are shared with L.
__global__ void __launch_bounds__(BLK) reduce_kernel(int64_t n, const int32_t *__restrict__ rp,
                                                     const int32_t *__restrict__ ci, const double *__restrict__ va,
                                                     const int32_t *__restrict__ map, int32_t *__restrict__ rp2,
                                                     int32_t *__restrict__ ci2, double *__restrict__ va2,
                                                     double *__restrict__ dinv) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i <= n; i += (int64_t)gridDim.x * BLK) {
        rp2[i] = rp[i];
        if (i == n) break;
        double d = 0.0;
        for (int32_t k = rp[i]; k < rp[i + 1]; k++) {
            const int32_t c = ci[k];
            const int32_t mc = c < n ? map[c] : -1;
            if (mc < 0) { ci2[k] = (int32_t)i; va2[k] = 0.0; }
            else {
                const double v = -va[k];
                ci2[k] = mc; va2[k] = v;
                if (mc == i) d += v;
            }
        }
        dinv[i] = 1.0 / d;   // special rows are overwritten afterwards
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

__global__ void __launch_bounds__(BLK) negate_kernel(int64_t n, const double *__restrict__ g, double *__restrict__ b) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) b[i] = -g[i];
}

__global__ void __launch_bounds__(BLK) ones_kernel(int64_t n, double *__restrict__ d) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) d[i] = 1.0;
}

// v_j = u[map_j] + o_j for nodes; extras start at 0
__global__ void __launch_bounds__(BLK) expand_kernel(int64_t n, int64_t N, const int32_t *__restrict__ map,
                                                     const double *__restrict__ o, const double *__restrict__ u,
                                                     double *__restrict__ v) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < N; i += (int64_t)gridDim.x * BLK) {
        if (i < n) {
            const int32_t m = map[i];
            v[i] = (m >= 0 ? u[m] : 0.0) + o[i];
        } else {
            v[i] = 0.0;
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// per-context solve workspace
// ------------------------------------------------------------------------------------------------
struct SolveWork {
    // reduced operator
    int64_t n = 0, N = 0;
    DevBuf<int32_t> rp2, ci2, map, sp_row, sp_rp, sp_ci, tmp_i32a, tmp_i32b, tmp_i32c;
    DevBuf<double> va2, dinv, sp_va, off, tmp_f64a, tmp_f64b;
    int32_t n_special = 0;
    // vectors
    DevBuf<double> b, x, r, z, p, s, w, rhat, vv, sh, ph, tt, full_a, full_b, rhs_dev;
    DevBuf<double> partials;
    Scal *d_sc = nullptr;
    Scal *h_sc = nullptr;  // pinned
    double *d_out = nullptr;
    bool have_operator = false;
    Operator A{};
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
    int iterations = 0, restarts = 0;
    double rel_residual = 0.0;
    float ms = 0.f;
};

void reset_scal(pdn_ctx *ctx, SolveWork *w, double tol2, double bnorm2, int max_iter) {
    Scal h{};
    h.tol2 = tol2;
    h.bnorm2 = bnorm2;
    h.first = 1;
    h.max_iter = max_iter;
    h.omega = 1.0;
    h.alpha = 1.0;
    h.gamma_old = 1.0;
    *w->h_sc = h;
    PDN_CUDA(cudaMemcpyAsync(w->d_sc, w->h_sc, sizeof(Scal), cudaMemcpyHostToDevice, ctx->stream));
}

// keep the iteration counter / thresholds, restart the recurrences
void restart_scal(pdn_ctx *ctx, SolveWork *w) {
    Scal h = *w->h_sc;
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

void enqueue_pcg_iteration(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    const int64_t n = w->n;
    launch_spmv<CG>(ctx, w->A, w->z.p, w->w.p, w->r.p, w);
    PDN_LAUNCH(ctx, cg_update_kernel, vgrid, RPB, 0, n, w->d_sc, w->dinv.p, w->w.p, w->z.p, w->p.p, w->s.p,
               w->x.p, w->r.p);
}

void enqueue_bicg_iteration(pdn_ctx *ctx, SolveWork *w, int vgrid) {
    const int64_t n = w->n;
    launch_spmv<BICG_V>(ctx, w->A, w->ph.p, w->vv.p, w->rhat.p, w);
    PDN_LAUNCH(ctx, bicg_s_kernel, vgrid, RPB, 0, n, w->d_sc, w->dinv.p, w->r.p, w->vv.p, w->s.p, w->sh.p);
    launch_spmv<BICG_T>(ctx, w->A, w->sh.p, w->tt.p, w->s.p, w);
    PDN_LAUNCH(ctx, bicg_x_kernel, vgrid, RPB, 0, n, w->d_sc, w->ph.p, w->sh.p, w->s.p, w->tt.p, w->rhat.p,
               w->x.p, w->r.p, w->partials.p);
    PDN_LAUNCH(ctx, bicg_p_kernel, vgrid, RPB, 0, n, w->d_sc, w->dinv.p, w->r.p, w->vv.p, w->p.p, w->ph.p);
}

// Solve A x = b (x starts at 0).  method: PDN_METHOD_PCG or PDN_METHOD_BICGSTAB.
KrylovResult run_krylov(pdn_ctx *ctx, SolveWork *w, int method, double rtol, double atol, int max_iter,
                        int check_every) {
    KrylovResult res;
    const int64_t n = w->n;
    cudaStream_t st = ctx->stream;
    const int vgrid = pdn_grid(ctx, n, RPB, 8);
    cudaEvent_t ev0, ev1;
    PDN_CUDA(cudaEventCreate(&ev0));
    PDN_CUDA(cudaEventCreate(&ev1));
    PDN_CUDA(cudaEventRecord(ev0, st));

    // ||b||
    reset_scal(ctx, w, 0.0, 0.0, max_iter);
    PDN_LAUNCH(ctx, sumsq_kernel, vgrid, RPB, 0, n, w->b.p, (const double *)nullptr, w->d_sc, w->d_out,
               w->partials.p);
    double bnorm2 = 0.0;
    PDN_CUDA(cudaMemcpyAsync(&bnorm2, w->d_out, sizeof(double), cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    PDN_CUDA(cudaMemsetAsync(w->x.p, 0, n * sizeof(double), st));
    if (!(bnorm2 > 0.0)) {  // b == 0 -> u == 0
        PDN_CUDA(cudaEventRecord(ev1, st));
        PDN_CUDA(cudaEventSynchronize(ev1));
        cudaEventDestroy(ev0);
        cudaEventDestroy(ev1);
        if (bnorm2 != bnorm2) PDN_FAIL(PDN_ESINGULAR, "solve: right-hand side is not finite");
        return res;
    }
    double tol2 = rtol * rtol * bnorm2;
    if (atol * atol > tol2) tol2 = atol * atol;
    reset_scal(ctx, w, tol2, bnorm2, max_iter);
    PDN_CUDA(cudaMemcpyAsync(w->r.p, w->b.p, n * sizeof(double), cudaMemcpyDeviceToDevice, st));

    // one CUDA graph = check_every iterations; kernels turn into no-ops once sc->done is set
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    auto build_graph = [&]() {
        const int64_t launches_before = ctx->launches;
        PDN_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        try {
            for (int k = 0; k < check_every; k++) {
                if (method == PDN_METHOD_PCG) enqueue_pcg_iteration(ctx, w, vgrid);
                else enqueue_bicg_iteration(ctx, w, vgrid);
            }
        } catch (...) {
            cudaGraph_t g2 = nullptr;
            cudaStreamEndCapture(st, &g2);
            if (g2) cudaGraphDestroy(g2);
            throw;
        }
        PDN_CUDA(cudaStreamEndCapture(st, &graph));
        PDN_CUDA(cudaGraphInstantiate(&exec, graph, 0));
        ctx->launches = launches_before;  // capture is not execution
    };
    const int per_iter = method == PDN_METHOD_PCG ? 2 : 5;
    build_graph();

    auto cleanup = [&]() {
        if (exec) cudaGraphExecDestroy(exec);
        if (graph) cudaGraphDestroy(graph);
        cudaEventDestroy(ev0);
        cudaEventDestroy(ev1);
    };

    try {
        const int max_restarts = 6;
        double best_rr = -1.0;
        for (int attempt = 0;; attempt++) {
            // start / restart from the current residual r
            if (method == PDN_METHOD_PCG)
                PDN_LAUNCH(ctx, cg_init_kernel, vgrid, RPB, 0, n, w->dinv.p, w->r.p, w->z.p, w->p.p, w->s.p);
            else
                PDN_LAUNCH(ctx, bicg_init_kernel, vgrid, RPB, 0, n, w->d_sc, w->dinv.p, w->r.p, w->rhat.p, w->p.p,
                           w->ph.p, w->partials.p);
            while (true) {
                PDN_CUDA(cudaGraphLaunch(exec, st));
                fetch_scal(ctx, w);
                res.iterations = w->h_sc->iter;
                ctx->launches += (int64_t)per_iter * check_every;  // kernels inside the graph
                if (w->h_sc->done) break;
            }
            // true residual of the current x
            launch_spmv<RESID>(ctx, w->A, w->x.p, w->r.p, w->b.p, w);
            fetch_scal(ctx, w);
            const double rr_true = w->h_sc->rr;
            res.rel_residual = sqrt(rr_true / bnorm2);
            const int done = w->h_sc->done;
            if (rr_true != rr_true) PDN_FAIL(PDN_ESINGULAR, "solve: residual is not finite (singular system?)");
            if (rr_true <= tol2 * 1.0000001) break;
            if (w->h_sc->iter >= max_iter)
                PDN_FAIL(PDN_ENOTCONVERGED, "solve: %d iterations, relative residual %.3e > %.3e", w->h_sc->iter,
                         res.rel_residual, rtol);
            // converged by the recurrence but not truly (drift), or breakdown: restart from the true residual
            if (attempt >= max_restarts || (done == 1 && best_rr >= 0.0 && rr_true >= best_rr)) {
                // stagnation at the attainable accuracy: accept if within 100x of the target, else fail
                if (rr_true <= tol2 * 1e4) break;
                PDN_FAIL(PDN_ENOTCONVERGED, "solve: stagnated at relative residual %.3e (target %.3e)",
                         res.rel_residual, rtol);
            }
            if (best_rr < 0.0 || rr_true < best_rr) best_rr = rr_true;
            res.restarts++;
            restart_scal(ctx, w);
        }
    } catch (...) {
        cleanup();
        throw;
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
    // out = 1 + largest row index that stores a diagonal entry
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < N; i += (int64_t)gridDim.x * BLK) {
        bool has = false;
        for (int32_t k = rp[i]; k < rp[i + 1]; k++) has = has || ci[k] == i;
        if (has) atomicMax(out, (int32_t)i + 1);
    }
}

template <typename T>
void upload(pdn_ctx *ctx, DevBuf<T> &buf, const std::vector<T> &h) {
    buf.reserve(h.size() + 1);
    if (!h.empty())
        PDN_CUDA(cudaMemcpyAsync(buf.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
}

}  // namespace

void pdn_solve_impl(pdn_ctx *ctx, const double *rhs, const pdn_opts *opts_in, double *v_out, pdn_info *info,
                    bool on_device) {
    cudaStream_t st = ctx->stream;
    SolveWork *w = get_work(ctx);
    const CsrDev &L = ctx->L;
    const int64_t N = L.n;
    pdn_opts opts{};
    if (opts_in) opts = *opts_in;
    else opts.n_nodes = -1;
    const double rtol = opts.rtol > 0.0 ? opts.rtol : 1e-10;
    const double atol = opts.atol > 0.0 ? opts.atol : 0.0;
    const int max_iter = opts.max_iter > 0 ? opts.max_iter : (int)(20.0 * sqrt((double)N)) + 2000;
    const int check_every = opts.check_every > 0 ? opts.check_every : 32;
    if (opts.precond == PDN_PRECOND_AMG) PDN_FAIL(PDN_EUNSUPPORTED, "solve: AMG preconditioner not built in this version");

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
        PDN_CUDA(cudaMemsetAsync(w->tmp_i32a.p, 0, sizeof(int32_t), st));
        PDN_LAUNCH(ctx, last_diag_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, N, L.indptr.p, L.indices.p, w->tmp_i32a.p);
        int32_t n32 = 0;
        PDN_CUDA(cudaMemcpyAsync(&n32, w->tmp_i32a.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        n = n32;
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

    // ---- elimination ------------------------------------------------------------------------------
    Elim E;
    build_elim(E, n, N, c_ptr, c_col, c_val, b_row, b_col, b_val, rhs_tail.data());
    const int nT = (int)E.T.size();
    int method = opts.method;
    if (method == PDN_METHOD_AUTO) method = E.symmetric ? PDN_METHOD_PCG : PDN_METHOD_BICGSTAB;
    if (method == PDN_METHOD_PCG && !E.symmetric)
        PDN_FAIL(PDN_EINVAL, "solve: PCG requested but the regulator coupling makes the system non-symmetric");

    // map / offsets on the device
    w->map.reserve((size_t)n);
    w->off.reserve((size_t)N);
    PDN_LAUNCH(ctx, iota_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, w->map.p);
    PDN_CUDA(cudaMemsetAsync(w->off.p, 0, N * sizeof(double), st));
    std::vector<int32_t> h_T(E.T.begin(), E.T.end()), h_rep(E.rep.begin(), E.rep.end());
    upload(ctx, w->sp_row, h_T);
    upload(ctx, w->tmp_i32b, h_rep);
    upload(ctx, w->tmp_f64a, E.off);
    if (nT > 0) {
        PDN_LAUNCH(ctx, scatter_i32_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->tmp_i32b.p, w->map.p);
        PDN_LAUNCH(ctx, scatter_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->tmp_f64a.p, w->off.p);
    }

    // main part of the reduced operator + Jacobi diagonal
    const int64_t nnz_n = (int64_t)c_ptr[0];  // entries of the node rows
    w->rp2.reserve((size_t)n + 1);
    w->ci2.reserve((size_t)nnz_n + 1);
    w->va2.reserve((size_t)nnz_n + 1);
    w->dinv.reserve((size_t)n);
    PDN_LAUNCH(ctx, reduce_kernel, pdn_grid(ctx, n + 1, BLK, 8), BLK, 0, n, L.indptr.p, L.indices.p, L.data.p, w->map.p,
               w->rp2.p, w->ci2.p, w->va2.p, w->dinv.p);

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
        PDN_LAUNCH(ctx, rowlen_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, n, w->sp_row.p, L.indptr.p, L.indices.p, w->tmp_i32c.p);
        std::vector<int32_t> len(nT);
        PDN_CUDA(cudaMemcpyAsync(len.data(), w->tmp_i32c.p, nT * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        w->tmp_f64b.reserve(nT);
        PDN_LAUNCH(ctx, gather_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->full_a.p, w->tmp_f64b.p);
        PDN_CUDA(cudaMemcpyAsync(g_T.data(), w->tmp_f64b.p, nT * sizeof(double), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        for (int t = 0; t < nT; t++) t_ptr[t + 1] = t_ptr[t] + len[t];
        const int32_t tot = t_ptr[nT];
        t_ci.resize(tot);
        t_va.resize(tot);
        DevBuf<int32_t> d_tptr, d_tci;
        DevBuf<double> d_tva;
        upload(ctx, d_tptr, t_ptr);
        d_tci.reserve(tot + 1);
        d_tva.reserve(tot + 1);
        PDN_LAUNCH(ctx, rowget_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, n, w->sp_row.p, L.indptr.p, L.indices.p, L.data.p,
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
    PDN_LAUNCH(ctx, negate_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, w->full_a.p, w->b.p);
    if (nT > 0) {
        PDN_LAUNCH(ctx, flag_rows_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->rp2.p);
        PDN_LAUNCH(ctx, scatter_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->tmp_f64a.p, w->b.p);
        PDN_LAUNCH(ctx, scatter_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->tmp_f64b.p, w->dinv.p);
    }
    PDN_LAUNCH(ctx, check_dinv_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, w->dinv.p, ctx->d_status);
    if (opts.precond == PDN_PRECOND_NONE) PDN_LAUNCH(ctx, ones_kernel, pdn_grid(ctx, n, BLK, 8), BLK, 0, n, w->dinv.p);
    {
        int64_t detail = 0;
        int stt = pdn_status_fetch(ctx, &detail);
        if (stt != 0) {
            pdn_status_reset(ctx);
            PDN_FAIL(stt, "solve: node %lld has no conductance to any other node (%s)", (long long)detail, pdn_strerror(stt));
        }
    }
    w->n_special = nT;
    w->A.n = n;
    w->A.rp = w->rp2.p;
    w->A.ci = w->ci2.p;
    w->A.va = w->va2.p;
    w->A.n_special = nT;
    w->A.sp_row = w->sp_row.p;
    w->A.sp_rp = w->sp_rp.p;
    w->A.sp_ci = w->sp_ci.p;
    w->A.sp_va = w->sp_va.p;
    w->have_operator = true;

    // vectors
    w->x.reserve((size_t)n); w->r.reserve((size_t)n); w->z.reserve((size_t)n); w->p.reserve((size_t)n);
    w->s.reserve((size_t)n); w->w.reserve((size_t)n);
    if (method == PDN_METHOD_BICGSTAB) {
        w->rhat.reserve((size_t)n); w->vv.reserve((size_t)n); w->sh.reserve((size_t)n);
        w->ph.reserve((size_t)n); w->tt.reserve((size_t)n);
    }
    PDN_CUDA(cudaEventRecord(ev_setup, st));

    // ---- Krylov -------------------------------------------------------------------------------------
    KrylovResult kr = run_krylov(ctx, w, method, rtol, atol, max_iter, check_every);

    // ---- recovery: v = Z u + o, x = M^-1 C (r_n - L_nn v), full residual -------------------------------
    double *d_v = on_device ? v_out : w->full_b.p;
    PDN_LAUNCH(ctx, expand_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, n, N, w->map.p, w->off.p, w->x.p, d_v);
    std::vector<double> xcur(m > 0 ? m : 1, 0.0);
    if (m > 0) {
        reset_scal(ctx, w, 0.0, 0.0, 0);
        launch_spmv<RESID>(ctx, full_operator(ctx, n), d_v, w->full_a.p, d_rhs, w);
        std::vector<double> h_T(nT > 0 ? nT : 1, 0.0);
        if (nT > 0) {
            w->tmp_f64b.reserve(nT);
            PDN_LAUNCH(ctx, gather_f64_kernel, pdn_blocks(nT, BLK), BLK, 0, nT, w->sp_row.p, w->full_a.p, w->tmp_f64b.p);
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
    const double residual_norm = sqrt(w->h_sc->rr);
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

// Times the kernel that dominates the solve: the CG-mode SpMV (fused r.z, z.w, r.r) on the reduced
// operator of the last pdn_solve; falls back to the full system when no solve has run yet.
//   mode 0: back-to-back launches (warm L2)
//   mode 1: a >L2 buffer is written before every launch (cold L2)
//   mode 2: in situ -- every SpMV is followed by the PCG vector update, exactly like one solver
//           iteration, and only the SpMV is bracketed by events (the cache state of the real solve)
void pdn_bench_spmv_impl(pdn_ctx *ctx, int32_t reps, int mode, double *ms) {
    SolveWork *w = get_work(ctx);
    cudaStream_t st = ctx->stream;
    Operator A = w->have_operator ? w->A : full_operator(ctx, ctx->L.n);
    const int64_t n = A.n;
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
    for (int k = 0; k < 3; k++) launch_spmv<CG>(ctx, A, x.p, y.p, aux.p, w);
    std::vector<cudaEvent_t> ev(2 * (size_t)reps);
    for (auto &e : ev) PDN_CUDA(cudaEventCreate(&e));
    for (int k = 0; k < reps; k++) {
        if (mode == 1) pdn_flush_l2(ctx);
        PDN_CUDA(cudaEventRecord(ev[2 * k], st));
        launch_spmv<CG>(ctx, A, x.p, y.p, aux.p, w);
        PDN_CUDA(cudaEventRecord(ev[2 * k + 1], st));
        if (mode == 2)   // dinv = v6, w = y, z = x, p = v3, s = v4, x = v5, r = aux
            PDN_LAUNCH(ctx, cg_update_kernel, vgrid, RPB, 0, n, w->d_sc, v6.p, y.p, x.p, v3.p, v4.p, v5.p, aux.p);
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
