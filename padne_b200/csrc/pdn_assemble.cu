// pdn_assemble.cu -- kernels (1) and (2): element weights and deterministic CSR assembly.
//
// Reference path being replaced (paths relative to atx/padne):
//   HalfEdge.cotan                       padne/mesh.py:124-139
//   Vertex.orbit                         padne/mesh.py:78-84   (fixes the fp summation order)
//   laplace_operator                     padne/solver.py:171-213
//   process_mesh_laplace_operators       padne/solver.py:563-575
//   stamp_network_into_system / ground   padne/solver.py:469-560  (arrive as an ordered op list)
//   scipy lil_matrix get/set semantics   an entry that becomes exactly 0.0 leaves the pattern
//
// Pipeline (all on the context stream, no host round trip until the final nnz read):
//   element_kernel      one thread per triangle (a warp = a batch of 32 triangles): three corner
//                       weights abs(dot/cross)/2 in fp64 with separately rounded operations,
//                       globalised vertex ids, per-vertex incidence counts
//   stamp_count_kernel  per-row stamp counts
//   scans               incidence / stamp / row-capacity offsets   (counting sort by ROW)
//   fill_kernel         incidence and stamp lists per row (order inside a row is fixed later)
//   row_kernel          one thread per row: walks the vertex orbit exactly like the reference
//                       (start = half-edge of the highest-numbered incident triangle, then
//                       twin.next), reduces the two faces of every edge, accumulates the
//                       diagonal in orbit order, replays the row's stamps in their original
//                       order with LIL semantics, sorts the row by column       (segmented
//                       sort + reduce inside the row's segment)
//   scan + compact      final indptr (int32) and the packed CSR
// The result is bit-identical run to run: the only atomics are integer counters whose arrival
// order is erased by the per-row orderings above.
#include "pdn_common.cuh"

namespace {

constexpr int BLK = 256;

__device__ __forceinline__ void raise_status(int32_t *status, int code, int64_t detail) {
    if (atomicCAS(&status[0], 0, code) == 0) status[1] = (int32_t)detail;
}

// abs(dot/cross)/2 at corner o for the edge (i,k) -- padne/mesh.py:136-138.  Each operation is
// rounded separately (CPython floats): explicit _rn intrinsics never contract into FMA.
__device__ __forceinline__ bool corner_weight(double ox, double oy, double ix, double iy, double kx,
                                              double ky, double *w) {
    double vix = __dsub_rn(ix, ox), viy = __dsub_rn(iy, oy);
    double vkx = __dsub_rn(kx, ox), vky = __dsub_rn(ky, oy);
    double dot = __dadd_rn(__dmul_rn(vix, vkx), __dmul_rn(viy, vky));
    double cross = __dsub_rn(__dmul_rn(vix, vky), __dmul_rn(viy, vkx));
    if (cross == 0.0) return false;
    *w = __ddiv_rn(fabs(__ddiv_rn(dot, cross)), 2.0);
    return true;
}

__device__ __forceinline__ int find_segment(const int64_t *__restrict__ off, int n_seg, int64_t i) {
    // largest m with off[m] <= i   (off has n_seg+1 entries, off[0] == 0)
    int lo = 0, hi = n_seg;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (off[mid] <= i) lo = mid; else hi = mid;
    }
    return lo;
}

// ------------------------------------------------------------------------------------------------
// kernel (1)
// ------------------------------------------------------------------------------------------------
// Standalone form: tri indexes xy directly (pdn_element_weights).
__global__ void __launch_bounds__(BLK) element_weights_kernel(int64_t n_v, const double2 *__restrict__ xy,
                                                              int64_t n_t, const int32_t *__restrict__ tri,
                                                              double *__restrict__ w, int32_t *status) {
    for (int64_t t = (int64_t)blockIdx.x * BLK + threadIdx.x; t < n_t; t += (int64_t)gridDim.x * BLK) {
        int32_t v[3] = {tri[3 * t], tri[3 * t + 1], tri[3 * t + 2]};
        bool ok = true;
#pragma unroll
        for (int c = 0; c < 3; c++) ok = ok && v[c] >= 0 && v[c] < n_v;
        if (!ok) { raise_status(status, PDN_EINVAL, t); continue; }
        double2 p[3] = {xy[v[0]], xy[v[1]], xy[v[2]]};
#pragma unroll
        for (int c = 0; c < 3; c++) {
            double wc = 0.0;
            const double2 o = p[c], a = p[(c + 1) % 3], b = p[(c + 2) % 3];
            if (!corner_weight(o.x, o.y, a.x, a.y, b.x, b.y, &wc)) raise_status(status, PDN_EDEGENERATE, t);
            w[3 * t + c] = wc;
        }
    }
}

// Assembly form: mesh-local triangles -> weights, globalised triangles, incidence counts.
__global__ void __launch_bounds__(BLK) element_kernel(int n_mesh, const int64_t *__restrict__ v_off,
                                                      const int64_t *__restrict__ t_off,
                                                      const double2 *__restrict__ xy,
                                                      const int32_t *__restrict__ tri, int64_t n_t,
                                                      double *__restrict__ w, int32_t *__restrict__ gtri,
                                                      int32_t *__restrict__ cnt, int32_t *status,
                                                      const int32_t *__restrict__ mesh_needed) {
    for (int64_t t = (int64_t)blockIdx.x * BLK + threadIdx.x; t < n_t; t += (int64_t)gridDim.x * BLK) {
        const int m = n_mesh > 1 ? find_segment(t_off, n_mesh, t) : 0;
        if (mesh_needed != nullptr && !mesh_needed[m]) {     // partitioned assembly: this mesh is not resident here
#pragma unroll
            for (int c = 0; c < 3; c++) gtri[3 * t + c] = -1;
            continue;
        }
        const int64_t vo = v_off[m], nv = v_off[m + 1] - vo;
        int32_t v[3] = {tri[3 * t], tri[3 * t + 1], tri[3 * t + 2]};
        bool ok = true;
#pragma unroll
        for (int c = 0; c < 3; c++) ok = ok && v[c] >= 0 && v[c] < nv;
        if (!ok) {
            raise_status(status, PDN_EINVAL, t);
#pragma unroll
            for (int c = 0; c < 3; c++) { gtri[3 * t + c] = -1; w[3 * t + c] = 0.0; }
            continue;
        }
        double2 p[3];
#pragma unroll
        for (int c = 0; c < 3; c++) { v[c] += (int32_t)vo; p[c] = xy[v[c]]; }
#pragma unroll
        for (int c = 0; c < 3; c++) {
            double wc = 0.0;
            const double2 o = p[c], a = p[(c + 1) % 3], b = p[(c + 2) % 3];
            if (!corner_weight(o.x, o.y, a.x, a.y, b.x, b.y, &wc)) raise_status(status, PDN_EDEGENERATE, t);
            w[3 * t + c] = wc;
            gtri[3 * t + c] = v[c];
            atomicAdd(&cnt[v[c]], 1);
        }
    }
}

__global__ void __launch_bounds__(BLK) stamp_count_kernel(int64_t n_stamp, int64_t N,
                                                          const int32_t *__restrict__ s_row,
                                                          const int32_t *__restrict__ s_col,
                                                          int32_t *__restrict__ cnt_s, int32_t *status) {
    for (int64_t s = (int64_t)blockIdx.x * BLK + threadIdx.x; s < n_stamp; s += (int64_t)gridDim.x * BLK) {
        int32_t r = s_row[s], c = s_col[s];
        if (r < 0 || r >= N || c < 0 || c >= N) { raise_status(status, PDN_EINVAL, s); continue; }
        atomicAdd(&cnt_s[r], 1);
    }
}

// row capacity = neighbours (<= incident triangles + 1) + diagonal + stamps
__global__ void __launch_bounds__(BLK) capacity_kernel(int64_t N, const int32_t *__restrict__ cnt,
                                                       const int32_t *__restrict__ cnt_s,
                                                       int32_t *__restrict__ ub, const uint8_t *__restrict__ want) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < N; i += (int64_t)gridDim.x * BLK) {
        int32_t c = cnt[i];
        ub[i] = (want == nullptr || want[i]) ? (c > 0 ? c + 2 : 0) + cnt_s[i] : 0;
    }
}

// ---- partitioned assembly (several GPUs): which rows this rank builds -------------------------------------------
// want[i] = 1 for the rows of this rank's range, for the extra rows (constraints: i >= n_nodes), and -- second kernel --
// for every constraint TERMINAL: both ends of any stamp that touches an extra row or column (voltage-source /
// regulator / ground terminals and regulator sense nodes, padne/solver.py:503-538, 558-560).  The solve eliminates
// the constraints on the host of EVERY rank, so those few rows are built everywhere; need[m] flags their meshes.
__global__ void __launch_bounds__(BLK) want_range_kernel(int64_t N, int64_t lo, int64_t hi, int64_t n_nodes,
                                                         uint8_t *__restrict__ want) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < N; i += (int64_t)gridDim.x * BLK)
        want[i] = ((i >= lo && i < hi) || i >= n_nodes) ? 1 : 0;
}
__global__ void __launch_bounds__(BLK) want_terminal_kernel(int64_t n_stamp, int64_t N, int64_t n_nodes, int64_t n_v,
                                                            int n_mesh, const int64_t *__restrict__ v_off,
                                                            const int32_t *__restrict__ s_row,
                                                            const int32_t *__restrict__ s_col, uint8_t *__restrict__ want,
                                                            int32_t *__restrict__ need) {
    for (int64_t s = (int64_t)blockIdx.x * BLK + threadIdx.x; s < n_stamp; s += (int64_t)gridDim.x * BLK) {
        const int32_t r = s_row[s], c = s_col[s];
        if (r < 0 || r >= N || c < 0 || c >= N) continue;           // reported by stamp_count_kernel
        if (r < n_nodes && c < n_nodes) continue;
        const int32_t e[2] = {r, c};
        for (int k = 0; k < 2; k++) {
            want[e[k]] = 1;
            if (e[k] < n_v) need[n_mesh > 1 ? find_segment(v_off, n_mesh, e[k]) : 0] = 1;
        }
    }
}

__global__ void __launch_bounds__(BLK) fill_kernel(int64_t n_t, const int32_t *__restrict__ gtri,
                                                   const int32_t *__restrict__ inc_ptr,
                                                   int32_t *__restrict__ cur, int32_t *__restrict__ inc,
                                                   int64_t n_stamp, int64_t N, const int32_t *__restrict__ s_row,
                                                   const int32_t *__restrict__ s_col,
                                                   const int32_t *__restrict__ sinc_ptr,
                                                   int32_t *__restrict__ cur_s, int32_t *__restrict__ sinc) {
    const int64_t n_corner = 3 * n_t, total = n_corner + n_stamp;
    for (int64_t k = (int64_t)blockIdx.x * BLK + threadIdx.x; k < total; k += (int64_t)gridDim.x * BLK) {
        if (k < n_corner) {
            int32_t g = gtri[k];
            if (g < 0) continue;
            int32_t slot = inc_ptr[g] + atomicAdd(&cur[g], 1);
            inc[slot] = (int32_t)k;  // packed t*3+c
        } else {
            int64_t s = k - n_corner;
            int32_t r = s_row[s], c = s_col[s];
            if (r < 0 || r >= N || c < 0 || c >= N) continue;
            int32_t slot = sinc_ptr[r] + atomicAdd(&cur_s[r], 1);
            sinc[slot] = (int32_t)s;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// kernel (2): one thread per row
// ------------------------------------------------------------------------------------------------
struct RowBuf {  // a row under construction in its scratch segment, unsorted, lil semantics
    int32_t *col;
    double *val;
    int len;
    __device__ __forceinline__ int find(int32_t j) const {
        for (int p = 0; p < len; p++)
            if (col[p] == j) return p;
        return -1;
    }
    __device__ __forceinline__ double get(int32_t j) const {
        int p = find(j);
        return p >= 0 ? val[p] : 0.0;
    }
    // L[i, j] = x  (x == 0.0 deletes -- scipy lil_insert)
    __device__ __forceinline__ void set(int32_t j, double x) {
        int p = find(j);
        if (x == 0.0) {
            if (p >= 0) { len--; col[p] = col[len]; val[p] = val[len]; }
            return;
        }
        if (p >= 0) { val[p] = x; return; }
        col[len] = j; val[len] = x; len++;
    }
    __device__ __forceinline__ void add(int32_t j, double x) { set(j, __dadd_rn(get(j), x)); }
};

struct Incident {
    int32_t nxt, prv;  // vertex after / before this one in the triangle
    double w_nxt;      // corner weight of the edge (self, nxt) = weight at the corner prv
    double w_prv;      // corner weight of the edge (self, prv) = weight at the corner nxt
};

__device__ __forceinline__ Incident load_incident(int32_t pk, const int32_t *__restrict__ gtri,
                                                  const double *__restrict__ w) {
    const int32_t t3 = (pk / 3) * 3, c = pk - t3;
    const int c1 = c == 2 ? 0 : c + 1, c2 = c == 0 ? 2 : c - 1;
    Incident r;
    r.nxt = gtri[t3 + c1];
    r.prv = gtri[t3 + c2];
    r.w_nxt = w[t3 + c2];
    r.w_prv = w[t3 + c1];
    return r;
}
__device__ __forceinline__ int32_t incident_nxt(int32_t pk, const int32_t *__restrict__ gtri) {
    const int32_t t3 = (pk / 3) * 3, c = pk - t3;
    return gtri[t3 + (c == 2 ? 0 : c + 1)];
}
__device__ __forceinline__ int32_t incident_prv(int32_t pk, const int32_t *__restrict__ gtri) {
    const int32_t t3 = (pk / 3) * 3, c = pk - t3;
    return gtri[t3 + (c == 0 ? 2 : c - 1)];
}

__global__ void __launch_bounds__(BLK) row_kernel(
    int64_t N, int64_t n_v, int n_mesh, const int64_t *__restrict__ v_off,
    const double *__restrict__ conductance, const int32_t *__restrict__ gtri, const double *__restrict__ w,
    const int32_t *__restrict__ inc_ptr, const int32_t *__restrict__ inc,
    const int32_t *__restrict__ sinc_ptr, int32_t *__restrict__ sinc, const int32_t *__restrict__ s_col,
    const double *__restrict__ s_val, const int32_t *__restrict__ s_op, const int32_t *__restrict__ ub_ptr,
    int32_t *__restrict__ tcol, double *__restrict__ tval, int32_t *__restrict__ row_len, int32_t *status,
    const uint8_t *__restrict__ want) {
    for (int64_t i = (int64_t)blockIdx.x * BLK + threadIdx.x; i < N; i += (int64_t)gridDim.x * BLK) {
        if (want != nullptr && !want[i]) { row_len[i] = 0; continue; }      // another rank's row
        RowBuf row;
        row.col = tcol + ub_ptr[i];
        row.val = tval + ub_ptr[i];
        row.len = 0;

        if (i < n_v) {
            const int32_t base = inc_ptr[i];
            const int ninc = inc_ptr[i + 1] - base;
            if (ninc == 0) {
                raise_status(status, PDN_EINVAL, i);  // vertex in no triangle: Vertex.orbit() fails
            } else {
                const double cond = conductance[n_mesh > 1 ? find_segment(v_off, n_mesh, i) : 0];
                // Vertex.out = the half-edge created from this vertex by the LAST triangle that
                // contains it (mesh.py:322): the incident entry with the largest triangle id.
                int j0 = 0;
                int32_t pk0 = inc[base];
                for (int j = 1; j < ninc; j++) {
                    int32_t pk = inc[base + j];
                    if (pk > pk0) { pk0 = pk; j0 = j; }
                }
                double diag = 0.0;
                int cur = j0, visited = 0;
                bool bad = false;
                while (true) {
                    const Incident a = load_incident(inc[base + cur], gtri, w);
                    // half-edge self -> a.nxt lives in triangle `cur`; its twin lives in the
                    // triangle whose predecessor of self is a.nxt
                    int j1 = -1, n_match = 0;
                    for (int j = 0; j < ninc; j++)
                        if (incident_prv(inc[base + j], gtri) == a.nxt) { j1 = j; n_match++; }
                    if (n_match > 1 || ++visited > ninc) { bad = true; break; }
                    if (n_match == 1) {
                        const Incident b = load_incident(inc[base + j1], gtri, w);
                        const double ratio = __dadd_rn(a.w_nxt, b.w_prv);   // own face, then twin face
                        if (ratio != 0.0) {
                            row.add((int32_t)a.nxt, __dmul_rn(cond, ratio));
                            diag = __dsub_rn(diag, ratio);
                        }
                        cur = j1;
                    } else {
                        // twin is a boundary half-edge: only the own face contributes
                        if (a.w_nxt != 0.0) {
                            row.add((int32_t)a.nxt, __dmul_rn(cond, a.w_nxt));
                            diag = __dsub_rn(diag, a.w_nxt);
                        }
                        // twin.next = the boundary half-edge leaving self: self -> y where (y, self)
                        // is interior and (self, y) is not: the incident triangle whose prv is
                        // nobody's nxt
                        int jb = -1, n_open = 0;
                        for (int j = 0; j < ninc; j++) {
                            const int32_t y = incident_prv(inc[base + j], gtri);
                            bool closed = false;
                            for (int k = 0; k < ninc; k++)
                                if (incident_nxt(inc[base + k], gtri) == y) { closed = true; break; }
                            if (!closed) { jb = j; n_open++; }
                        }
                        if (n_open != 1) { bad = true; break; }
                        const Incident b = load_incident(inc[base + jb], gtri, w);
                        if (b.w_prv != 0.0) {   // boundary half-edge self -> y: twin face only
                            row.add((int32_t)b.prv, __dmul_rn(cond, b.w_prv));
                            diag = __dsub_rn(diag, b.w_prv);
                        }
                        cur = jb;
                    }
                    if (cur == j0) break;
                }
                if (bad || visited != ninc) raise_status(status, PDN_ENONMANIFOLD, i);
                row.add((int32_t)i, __dmul_rn(cond, diag));
            }
        }

        // stamps of this row, replayed in their original order
        const int32_t sb = sinc_ptr[i];
        const int ns = sinc_ptr[i + 1] - sb;
        for (int a = 1; a < ns; a++) {  // insertion sort by stamp index
            int32_t key = sinc[sb + a];
            int b = a - 1;
            while (b >= 0 && sinc[sb + b] > key) { sinc[sb + b + 1] = sinc[sb + b]; b--; }
            sinc[sb + b + 1] = key;
        }
        for (int a = 0; a < ns; a++) {
            const int32_t s = sinc[sb + a];
            const int32_t j = s_col[s];
            if (s_op[s] == PDN_OP_SET) row.set(j, s_val[s]);
            else row.add(j, s_val[s]);
        }

        // sort the row by column
        for (int a = 1; a < row.len; a++) {
            int32_t kc = row.col[a];
            double kv = row.val[a];
            int b = a - 1;
            while (b >= 0 && row.col[b] > kc) { row.col[b + 1] = row.col[b]; row.val[b + 1] = row.val[b]; b--; }
            row.col[b + 1] = kc;
            row.val[b + 1] = kv;
        }
        row_len[i] = row.len;
    }
}

__global__ void __launch_bounds__(BLK) compact_kernel(int64_t N, const int32_t *__restrict__ ub_ptr,
                                                      const int32_t *__restrict__ indptr,
                                                      const int32_t *__restrict__ tcol,
                                                      const double *__restrict__ tval,
                                                      int32_t *__restrict__ indices, double *__restrict__ data) {
    // one warp per 32 consecutive rows; lanes stride over each row's entries
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * BLK + threadIdx.x) >> 5;
    const int64_t n_warp = ((int64_t)gridDim.x * BLK) >> 5;
    for (int64_t r0 = warp * 32; r0 < N; r0 += n_warp * 32) {
        const int64_t r1 = r0 + 32 < N ? r0 + 32 : N;
        for (int64_t i = r0; i < r1; i++) {
            const int32_t src = ub_ptr[i], dst = indptr[i], len = indptr[i + 1] - dst;
            for (int k = lane; k < len; k += 32) {
                indices[dst + k] = tcol[src + k];
                data[dst + k] = tval[src + k];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// next row F1: per-face power density (padne/solver.py:689-745)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLK) power_density_kernel(int64_t n_v, const double2 *__restrict__ xy,
                                                            const double *__restrict__ val, int64_t n_t,
                                                            const int32_t *__restrict__ tri, double conductivity,
                                                            double *__restrict__ p_out, int32_t *status) {
    for (int64_t t = (int64_t)blockIdx.x * BLK + threadIdx.x; t < n_t; t += (int64_t)gridDim.x * BLK) {
        const int32_t a = tri[3 * t], b = tri[3 * t + 1], c = tri[3 * t + 2];
        if (a < 0 || a >= n_v || b < 0 || b >= n_v || c < 0 || c >= n_v) {
            raise_status(status, PDN_EINVAL, t);
            continue;
        }
        const double2 p1 = xy[a], p2 = xy[b], p3 = xy[c];
        const double x1 = p1.x, y1 = p1.y, x2 = p2.x, y2 = p2.y, x3 = p3.x, y3 = p3.y;
        const double f1 = val[a], f2 = val[b], f3 = val[c];
        double g[2];
#pragma unroll
        for (int k = 0; k < 2; k++) {
            // barycentric interpolant evaluated at (x1+1, y1) and (x1, y1+1), minus f1
            const double x = k == 0 ? __dadd_rn(x1, 1.0) : x1, y = k == 0 ? y1 : __dadd_rn(y1, 1.0);
            const double y23 = __dsub_rn(y2, y3), x32 = __dsub_rn(x3, x2);
            const double D = __dadd_rn(__dmul_rn(y23, __dsub_rn(x1, x3)), __dmul_rn(x32, __dsub_rn(y1, y3)));
            const double xx3 = __dsub_rn(x, x3), yy3 = __dsub_rn(y, y3);
            const double l1 = __ddiv_rn(__dadd_rn(__dmul_rn(y23, xx3), __dmul_rn(x32, yy3)), D);
            const double l2 = __ddiv_rn(
                __dadd_rn(__dmul_rn(__dsub_rn(y3, y1), xx3), __dmul_rn(__dsub_rn(x1, x3), yy3)), D);
            const double l3 = __dsub_rn(__dsub_rn(1.0, l1), l2);
            const double f = __dadd_rn(__dadd_rn(__dmul_rn(l1, f1), __dmul_rn(l2, f2)), __dmul_rn(l3, f3));
            g[k] = __dsub_rn(f, f1);
        }
        const double jx = __dmul_rn(g[0], conductivity), jy = __dmul_rn(g[1], conductivity);
        p_out[t] = __dadd_rn(__dmul_rn(jx, g[0]), __dmul_rn(jy, g[1]));
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// host drivers
// ------------------------------------------------------------------------------------------------
static void raise_device_status(pdn_ctx *ctx, const char *where) {
    int64_t detail = 0;
    int st = pdn_status_fetch(ctx, &detail);
    if (st != 0) {
        pdn_status_reset(ctx);
        PDN_FAIL(st, "%s: %s (at index %lld)", where, pdn_strerror(st), (long long)detail);
    }
}

void pdn_element_weights_impl(pdn_ctx *ctx, int64_t n_v, const double *xy, int64_t n_t, const int32_t *tri,
                              double *w, bool on_device) {
    if (n_t == 0) return;
    DevIn<double> d_xy;
    DevIn<int32_t> d_tri;
    DevBuf<double> d_w;
    d_xy.set(xy, (size_t)n_v * 2, on_device, ctx->stream);
    d_tri.set(tri, (size_t)n_t * 3, on_device, ctx->stream);
    double *wp = on_device ? w : d_w.reserve((size_t)n_t * 3);
    pdn_status_reset(ctx);
    PDN_LAUNCH(ctx, element_weights_kernel, pdn_grid(ctx, n_t, BLK, 8), BLK, 0, n_v, (const double2 *)d_xy.p,
               n_t, d_tri.p, wp, ctx->d_status);
    if (!on_device)
        PDN_CUDA(cudaMemcpyAsync(w, wp, (size_t)n_t * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    raise_device_status(ctx, "element_weights");
}

void pdn_power_density_impl(pdn_ctx *ctx, int64_t n_v, const double *xy, const double *val, int64_t n_t,
                            const int32_t *tri, double conductivity, double *p, bool on_device) {
    if (n_t == 0) return;
    DevIn<double> d_xy, d_val;
    DevIn<int32_t> d_tri;
    DevBuf<double> d_p;
    d_xy.set(xy, (size_t)n_v * 2, on_device, ctx->stream);
    d_val.set(val, (size_t)n_v, on_device, ctx->stream);
    d_tri.set(tri, (size_t)n_t * 3, on_device, ctx->stream);
    double *pp = on_device ? p : d_p.reserve((size_t)n_t);
    pdn_status_reset(ctx);
    PDN_LAUNCH(ctx, power_density_kernel, pdn_grid(ctx, n_t, BLK, 8), BLK, 0, n_v, (const double2 *)d_xy.p,
               d_val.p, n_t, d_tri.p, conductivity, pp, ctx->d_status);
    if (!on_device)
        PDN_CUDA(cudaMemcpyAsync(p, pp, (size_t)n_t * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    raise_device_status(ctx, "power_density");
}

// ------------------------------------------------------------------------------------------------
// next row F3: nearest mesh vertex of every connection point (NodeIndexer, padne/solver.py:356-396, 421-433).
// The reference builds one scipy KDTree per layer over all vertices of all meshes of that layer and
// queries it once per connection.  Here: a CTA takes NQ queries, walks the meshes whose layer matches and
// keeps per-thread minima of the squared distance dx*dx + dy*dy (separately rounded: this unit is compiled
// with -fmad=false), then reduces (distance, global index) with the LOWEST index winning a tie.  The
// vertices are read once per NQ queries; everything is L2-resident at board sizes.
// ------------------------------------------------------------------------------------------------
constexpr int NQ = 8;
__global__ void __launch_bounds__(256) nearest_vertex_kernel(int32_t n_mesh, const int64_t *__restrict__ v_off,
                                                             const int32_t *__restrict__ mesh_layer,
                                                             const double2 *__restrict__ xy, int64_t n_q,
                                                             const double2 *__restrict__ q_xy,
                                                             const int32_t *__restrict__ q_layer,
                                                             int64_t *__restrict__ out_index, double *__restrict__ out_d2) {
    __shared__ double s_d[NQ][8];
    __shared__ long long s_i[NQ][8];
    const int64_t q0 = (int64_t)blockIdx.x * NQ;
    double qx[NQ], qy[NQ], best[NQ];
    long long bi[NQ];
    int32_t ql[NQ];
#pragma unroll
    for (int k = 0; k < NQ; k++) {
        const bool live = q0 + k < n_q;
        const double2 q = live ? q_xy[q0 + k] : make_double2(0.0, 0.0);
        qx[k] = q.x; qy[k] = q.y;
        ql[k] = live ? q_layer[q0 + k] : -2147483647;
        best[k] = INFINITY;
        bi[k] = -1;
    }
    for (int32_t m = 0; m < n_mesh; m++) {
        const int32_t layer = mesh_layer[m];
        unsigned mask = 0;
#pragma unroll
        for (int k = 0; k < NQ; k++) mask |= (ql[k] == layer ? 1u : 0u) << k;
        if (!mask) continue;
        for (int64_t i = v_off[m] + threadIdx.x; i < v_off[m + 1]; i += 256) {   // ascending index per thread
            const double2 p = xy[i];
#pragma unroll
            for (int k = 0; k < NQ; k++) {
                const double dx = p.x - qx[k], dy = p.y - qy[k];
                const double d = dx * dx + dy * dy;
                if (((mask >> k) & 1u) && d < best[k]) { best[k] = d; bi[k] = i; }
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NQ; k++) {
        double d = best[k];
        long long i = bi[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double d2 = __shfl_xor_sync(0xffffffffu, d, o);
            const long long i2 = __shfl_xor_sync(0xffffffffu, i, o);
            if (i2 >= 0 && (i < 0 || d2 < d || (d2 == d && i2 < i))) { d = d2; i = i2; }
        }
        if (lane == 0) { s_d[k][warp] = d; s_i[k][warp] = i; }
    }
    __syncthreads();
    if (threadIdx.x < NQ && q0 + threadIdx.x < n_q) {
        const int k = threadIdx.x;
        double d = s_d[k][0];
        long long i = s_i[k][0];
        for (int w = 1; w < 8; w++) {
            const double d2 = s_d[k][w];
            const long long i2 = s_i[k][w];
            if (i2 >= 0 && (i < 0 || d2 < d || (d2 == d && i2 < i))) { d = d2; i = i2; }
        }
        out_index[q0 + k] = i;                       // -1: no vertex on that layer
        if (out_d2) out_d2[q0 + k] = d;
    }
}

void pdn_nearest_vertex_impl(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const int32_t *mesh_layer,
                             const double *xy, int64_t n_q, const double *q_xy, const int32_t *q_layer,
                             int64_t *out_index, double *out_d2, bool on_device) {
    if (n_q == 0) return;
    cudaStream_t st = ctx->stream;
    int64_t n_v = 0;
    if (on_device) {
        PDN_CUDA(cudaMemcpyAsync(&n_v, v_off + n_mesh, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
    } else {
        n_v = v_off[n_mesh];
    }
    DevIn<int64_t> d_voff;
    DevIn<int32_t> d_ml, d_ql;
    DevIn<double> d_xy, d_q;
    DevBuf<int64_t> d_oi;
    DevBuf<double> d_od;
    d_voff.set(v_off, (size_t)n_mesh + 1, on_device, st);
    d_ml.set(mesh_layer, (size_t)n_mesh, on_device, st);
    d_xy.set(xy, (size_t)n_v * 2, on_device, st);
    d_q.set(q_xy, (size_t)n_q * 2, on_device, st);
    d_ql.set(q_layer, (size_t)n_q, on_device, st);
    int64_t *oi = on_device ? out_index : d_oi.reserve((size_t)n_q);
    double *od = out_d2 ? (on_device ? out_d2 : d_od.reserve((size_t)n_q)) : nullptr;
    PDN_LAUNCH(ctx, nearest_vertex_kernel, pdn_blocks(n_q, NQ), 256, 0, n_mesh, d_voff.p, d_ml.p, (const double2 *)d_xy.p, n_q,
               (const double2 *)d_q.p, d_ql.p, oi, od);
    if (!on_device) {
        PDN_CUDA(cudaMemcpyAsync(out_index, oi, (size_t)n_q * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        if (out_d2) PDN_CUDA(cudaMemcpyAsync(out_d2, od, (size_t)n_q * sizeof(double), cudaMemcpyDeviceToHost, st));
    }
    PDN_CUDA(cudaStreamSynchronize(st));
}

struct AssembleWork {
    DevIn<int64_t> d_voff, d_toff;
    DevIn<double> d_xy, d_cond, d_sval;
    DevIn<int32_t> d_tri, d_srow, d_scol, d_sop;
    DevBuf<double> w, tval;
    DevBuf<int32_t> gtri, cnt, cnt_s, ub, inc_ptr, sinc_ptr, ub_ptr, inc, sinc, tcol, row_len;
    DevBuf<uint8_t> want;          // partitioned assembly: rows this rank builds
    DevBuf<int32_t> mesh_needed;   // ... and the meshes it needs resident
};

void pdn_assemble_release(pdn_ctx *ctx) {
    delete ctx->asmw;
    ctx->asmw = nullptr;
}

void pdn_assemble_impl(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const double *xy,
                       const int64_t *t_off, const int32_t *tri, const double *conductance, int64_t N,
                       int64_t n_stamp, const int32_t *s_row, const int32_t *s_col, const double *s_val,
                       const int32_t *s_op, bool on_device, int64_t *nnz_out, int64_t n_nodes_part) {
    cudaStream_t st = ctx->stream;
    // Partitioned assembly (n_nodes_part >= 0 on a context with a communicator): this rank builds the rows of its
    // range of the row partition (pdn_row_partition: the one the solve uses), the extra rows and the constraint
    // terminals; from HOST inputs only the meshes those rows live in are uploaded.  Everything else of the resident
    // system stays empty -- the matrix is block diagonal per mesh (padne/solver.py:216-229, 563-575) and the
    // coupling is the stamp list, so no rank needs another rank's rows.
    const int world = pdn_comm_world(ctx), me = pdn_comm_rank(ctx);
    const bool part = n_nodes_part >= 0 && world > 1;
    int64_t part_lo = 0, part_hi = N;
    if (part) {
        if (n_nodes_part > N) PDN_FAIL(PDN_EINVAL, "assemble: n_nodes larger than N");
        std::vector<int64_t> begin(world + 1);
        pdn_row_partition(n_nodes_part, world, begin.data());
        part_lo = begin[me];
        part_hi = begin[me + 1];
    }
    ctx->part_valid = false;
    // offsets are needed on both sides; they are tiny
    std::vector<int64_t> h_voff(n_mesh + 1), h_toff(n_mesh + 1);
    if (on_device) {
        PDN_CUDA(cudaMemcpyAsync(h_voff.data(), v_off, (n_mesh + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaMemcpyAsync(h_toff.data(), t_off, (n_mesh + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        PDN_CUDA(cudaStreamSynchronize(st));
    } else {
        memcpy(h_voff.data(), v_off, (n_mesh + 1) * sizeof(int64_t));
        memcpy(h_toff.data(), t_off, (n_mesh + 1) * sizeof(int64_t));
    }
    if (h_voff[0] != 0 || h_toff[0] != 0) PDN_FAIL(PDN_EINVAL, "assemble: offsets must start at 0");
    for (int m = 0; m < n_mesh; m++)
        if (h_voff[m + 1] < h_voff[m] || h_toff[m + 1] < h_toff[m])
            PDN_FAIL(PDN_EINVAL, "assemble: offsets must be non-decreasing");
    const int64_t n_v = h_voff[n_mesh], n_t = h_toff[n_mesh];
    if (n_v > N) PDN_FAIL(PDN_EINVAL, "assemble: N smaller than the vertex count");
    if (n_t > 0 && (!xy || !tri || !conductance)) PDN_FAIL(PDN_EINVAL, "assemble: mesh arrays missing");
    if (N >= 0x7fffffffLL || 3 * n_t + 2 * n_v + n_stamp >= 0x7fffffffLL)
        PDN_FAIL(PDN_EUNSUPPORTED, "assemble: system exceeds int32 indexing");

    // scratch and staging buffers live in the context and only ever grow: repeated assemblies of
    // same-sized boards do no cudaMalloc / cudaFree at all
    if (!ctx->asmw) ctx->asmw = new AssembleWork();
    AssembleWork &aw = *ctx->asmw;
    DevIn<int64_t> &d_voff = aw.d_voff, &d_toff = aw.d_toff;
    DevIn<double> &d_xy = aw.d_xy, &d_cond = aw.d_cond, &d_sval = aw.d_sval;
    DevIn<int32_t> &d_tri = aw.d_tri, &d_srow = aw.d_srow, &d_scol = aw.d_scol, &d_sop = aw.d_sop;
    d_voff.set(v_off, n_mesh + 1, on_device, st);
    d_toff.set(t_off, n_mesh + 1, on_device, st);
    d_cond.set(conductance, n_mesh, on_device, st);
    d_srow.set(s_row, n_stamp, on_device, st);
    d_scol.set(s_col, n_stamp, on_device, st);
    d_sval.set(s_val, n_stamp, on_device, st);
    d_sop.set(s_op, n_stamp, on_device, st);
    const uint8_t *want = nullptr;
    const int32_t *mesh_needed = nullptr;
    std::vector<int32_t> h_need(n_mesh > 0 ? n_mesh : 1, 1);
    if (part) {
        aw.want.reserve((size_t)N);
        aw.mesh_needed.reserve((size_t)n_mesh + 1);
        PDN_CUDA(cudaMemsetAsync(aw.mesh_needed.p, 0, ((size_t)n_mesh + 1) * sizeof(int32_t), st));
        PDN_LAUNCH(ctx, want_range_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, N, part_lo, part_hi, n_nodes_part, aw.want.p);
        if (n_stamp > 0)
            PDN_LAUNCH(ctx, want_terminal_kernel, pdn_grid(ctx, n_stamp, BLK, 8), BLK, 0, n_stamp, N, n_nodes_part, n_v,
                       (int)n_mesh, d_voff.p, d_srow.p, d_scol.p, aw.want.p, aw.mesh_needed.p);
        if (n_mesh > 0) {
            PDN_CUDA(cudaMemcpyAsync(h_need.data(), aw.mesh_needed.p, n_mesh * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            PDN_CUDA(cudaStreamSynchronize(st));
            for (int m = 0; m < n_mesh; m++)
                if (h_voff[m] < part_hi && h_voff[m + 1] > part_lo) h_need[m] = 1;      // overlaps this rank's rows
            PDN_CUDA(cudaMemcpyAsync(aw.mesh_needed.p, h_need.data(), n_mesh * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        }
        want = aw.want.p;
        mesh_needed = aw.mesh_needed.p;
    }
    if (on_device || !part) {
        d_xy.set(xy, (size_t)n_v * 2, on_device, st);
        d_tri.set(tri, (size_t)n_t * 3, on_device, st);
    } else {
        // host inputs, partitioned: full-size device arrays (the kernels index them globally), only the needed
        // meshes copied in -- consecutive needed meshes as one copy
        d_xy.own.reserve((size_t)n_v * 2);
        d_tri.own.reserve((size_t)n_t * 3);
        d_xy.p = d_xy.own.p;
        d_tri.p = d_tri.own.p;
        for (int m = 0; m < n_mesh;) {
            if (!h_need[m]) { m++; continue; }
            int m1 = m;
            while (m1 + 1 < n_mesh && h_need[m1 + 1]) m1++;
            const int64_t v0 = h_voff[m], v1 = h_voff[m1 + 1], t0 = h_toff[m], t1 = h_toff[m1 + 1];
            if (v1 > v0)
                PDN_CUDA(cudaMemcpyAsync(d_xy.own.p + 2 * v0, xy + 2 * v0, (size_t)(v1 - v0) * 2 * sizeof(double), cudaMemcpyHostToDevice, st));
            if (t1 > t0)
                PDN_CUDA(cudaMemcpyAsync(d_tri.own.p + 3 * t0, tri + 3 * t0, (size_t)(t1 - t0) * 3 * sizeof(int32_t), cudaMemcpyHostToDevice, st));
            m = m1 + 1;
        }
    }

    DevBuf<double> &w = aw.w, &tval = aw.tval;
    DevBuf<int32_t> &gtri = aw.gtri, &cnt = aw.cnt, &cnt_s = aw.cnt_s, &ub = aw.ub, &inc_ptr = aw.inc_ptr,
                    &sinc_ptr = aw.sinc_ptr, &ub_ptr = aw.ub_ptr, &inc = aw.inc, &sinc = aw.sinc, &tcol = aw.tcol,
                    &row_len = aw.row_len;
    w.reserve((size_t)n_t * 3 + 1);
    gtri.reserve((size_t)n_t * 3 + 1);
    cnt.reserve((size_t)N);
    cnt_s.reserve((size_t)N);
    ub.reserve((size_t)N);
    inc_ptr.reserve((size_t)N + 1);
    sinc_ptr.reserve((size_t)N + 1);
    ub_ptr.reserve((size_t)N + 1);
    inc.reserve((size_t)n_t * 3 + 1);
    sinc.reserve((size_t)n_stamp + 1);
    row_len.reserve((size_t)N);
    const size_t cap_total = (size_t)(3 * n_t + 2 * n_v + n_stamp) + 1;
    tcol.reserve(cap_total);
    tval.reserve(cap_total);

    pdn_status_reset(ctx);
    PDN_CUDA(cudaMemsetAsync(cnt.p, 0, (size_t)N * sizeof(int32_t), st));
    PDN_CUDA(cudaMemsetAsync(cnt_s.p, 0, (size_t)N * sizeof(int32_t), st));
    if (n_t > 0)
        PDN_LAUNCH(ctx, element_kernel, pdn_grid(ctx, n_t, BLK, 8), BLK, 0, (int)n_mesh, d_voff.p, d_toff.p,
                   (const double2 *)d_xy.p, d_tri.p, n_t, w.p, gtri.p, cnt.p, ctx->d_status, mesh_needed);
    if (n_stamp > 0)
        PDN_LAUNCH(ctx, stamp_count_kernel, pdn_grid(ctx, n_stamp, BLK, 8), BLK, 0, n_stamp, N, d_srow.p,
                   d_scol.p, cnt_s.p, ctx->d_status);
    PDN_LAUNCH(ctx, capacity_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, N, cnt.p, cnt_s.p, ub.p, want);
    pdn_exclusive_scan_i32(ctx, cnt.p, N, inc_ptr.p);
    pdn_exclusive_scan_i32(ctx, cnt_s.p, N, sinc_ptr.p);
    pdn_exclusive_scan_i32(ctx, ub.p, N, ub_ptr.p);
    // the count arrays become the fill cursors
    PDN_CUDA(cudaMemsetAsync(cnt.p, 0, (size_t)N * sizeof(int32_t), st));
    PDN_CUDA(cudaMemsetAsync(cnt_s.p, 0, (size_t)N * sizeof(int32_t), st));
    if (3 * n_t + n_stamp > 0)
        PDN_LAUNCH(ctx, fill_kernel, pdn_grid(ctx, 3 * n_t + n_stamp, BLK, 8), BLK, 0, n_t, gtri.p, inc_ptr.p,
                   cnt.p, inc.p, n_stamp, N, d_srow.p, d_scol.p, sinc_ptr.p, cnt_s.p, sinc.p);
    PDN_LAUNCH(ctx, row_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, N, n_v, (int)n_mesh, d_voff.p, d_cond.p,
               gtri.p, w.p, inc_ptr.p, inc.p, sinc_ptr.p, sinc.p, d_scol.p, d_sval.p, d_sop.p, ub_ptr.p,
               tcol.p, tval.p, row_len.p, ctx->d_status, want);

    CsrDev &L = ctx->L;
    L.indptr.reserve((size_t)N + 1);
    pdn_exclusive_scan_i32(ctx, row_len.p, N, L.indptr.p);
    int32_t nnz32 = 0;
    PDN_CUDA(cudaMemcpyAsync(&nnz32, L.indptr.p + N, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    raise_device_status(ctx, "assemble");  // synchronises
    const int64_t nnz = nnz32;
    L.indices.reserve((size_t)nnz + 1);
    L.data.reserve((size_t)nnz + 1);
    PDN_LAUNCH(ctx, compact_kernel, pdn_grid(ctx, N, BLK, 8), BLK, 0, N, ub_ptr.p, L.indptr.p, tcol.p, tval.p,
               L.indices.p, L.data.p);
    PDN_CUDA(cudaStreamSynchronize(st));
    L.n = N;
    L.nnz = nnz;
    L.valid = true;
    if (part) {
        ctx->part_valid = true;
        ctx->part_lo = part_lo;
        ctx->part_hi = part_hi;
        ctx->part_n_nodes = n_nodes_part;
    }
    // keep the geometry for the solve's coarse space (pdn_coarse.cuh)
    if (n_v > 0 && n_mesh > 0) {
        if (on_device) {   // borrowed pointers: take copies
            ctx->geo_xy_buf.reserve((size_t)n_v * 2);
            PDN_CUDA(cudaMemcpyAsync(ctx->geo_xy_buf.p, xy, (size_t)n_v * 2 * sizeof(double), cudaMemcpyDeviceToDevice, st));
            ctx->geo_xy = ctx->geo_xy_buf.p;
        } else {
            ctx->geo_xy = d_xy.p;   // the staging copy lives in the context until the next assembly
        }
        ctx->geo_voff.reserve((size_t)n_mesh + 1);
        PDN_CUDA(cudaMemcpyAsync(ctx->geo_voff.p, h_voff.data(), (n_mesh + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
        PDN_CUDA(cudaStreamSynchronize(st));
        ctx->geo_n_mesh = n_mesh;
        ctx->geo_n_v = n_v;
        ctx->geo_valid = true;
    }
    if (nnz_out) *nnz_out = nnz;
}
