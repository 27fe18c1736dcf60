/*
 * padne_oracle.c -- CPU restatement of the reference's assembly path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library, and only as the checker or the timed CPU baseline.  The product
 * (padne_b200/) never links, loads or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against (a) the golden vectors of
 * the reference's own tests (unit square Laplacian, tests/test_solver.py:776-852; lumped stamps,
 * :65-147) and (b) fixtures under tests/golden/ produced by running the UNMODIFIED reference
 * (padne/mesh.py, padne/solver.py imported from /root/reference through oracle/refload.py;
 * generator: tests/golden/make_golden.py).
 *
 * What is restated, with the reference lines each function follows
 * (paths relative to /root/reference):
 *   he_build()          Mesh.from_triangle_soup           padne/mesh.py:302-378
 *   he_cotan()          HalfEdge.cotan                    padne/mesh.py:124-139
 *                       (Point.__sub__ :57-60, Vector.dot :24-25, Vector.__xor__ :41-43)
 *   laplace_into_lil()  Vertex.orbit + laplace_operator   padne/mesh.py:78-84, padne/solver.py:171-213
 *                       + process_mesh_laplace_operators  padne/solver.py:563-575
 *   lil_get/lil_set     scipy.sparse.lil_matrix item get/set: an entry set to exactly 0.0 is
 *                       deleted (scipy/sparse/_csparsetools lil_insert; probed)
 *   oracle_assemble()   assemble_system                   padne/solver.py:783-812 (stamps arrive as
 *                       the ordered op list of padne_b200/stamps.py = solver.py:469-560)
 *   oracle_power_density()  compute_triangle_gradient / compute_power_density  solver.py:689-745
 *
 * Every floating-point expression keeps the reference's operation order; build with
 * -ffp-contract=off so that no FMA is formed (CPython rounds every * + - / separately).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_OK 0
#define ORACLE_EINVAL 1
#define ORACLE_EDEGENERATE 2   /* ZeroDivisionError in cotan (mesh.py:138) */
#define ORACLE_ENONMANIFOLD 7  /* ValueError("Non-manifold mesh") (mesh.py:343) */
#define ORACLE_ENOMEM 8

/* ------------------------------------------------------------------------------------------ */
/* half-edge mesh                                                                             */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    int64_t n_v, n_he;
    int32_t *origin, *twin, *next, *face; /* face: >=0 interior face id, -1 none yet, -2 boundary */
    int32_t *out;                         /* per vertex: Vertex.out */
    /* open-addressing edge map (mesh._edge_map) */
    uint64_t *keys;
    int32_t *vals;
    uint64_t cap;
} hemesh;

static uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
static int32_t map_get(const hemesh *m, uint64_t key) {
    uint64_t h = mix64(key) & (m->cap - 1);
    while (m->keys[h] != UINT64_MAX) {
        if (m->keys[h] == key) return m->vals[h];
        h = (h + 1) & (m->cap - 1);
    }
    return -1;
}
static void map_put(hemesh *m, uint64_t key, int32_t v) {
    uint64_t h = mix64(key) & (m->cap - 1);
    while (m->keys[h] != UINT64_MAX) h = (h + 1) & (m->cap - 1);
    m->keys[h] = key; m->vals[h] = v;
}
static void he_free(hemesh *m) {
    free(m->origin); free(m->twin); free(m->next); free(m->face); free(m->out);
    free(m->keys); free(m->vals);
    memset(m, 0, sizeof *m);
}

/* Mesh.connect_vertices (mesh.py:266-297) */
static int32_t he_connect(hemesh *m, int32_t v1, int32_t v2) {
    uint64_t k12 = ((uint64_t)(uint32_t)v1 << 32) | (uint32_t)v2;
    int32_t e = map_get(m, k12);
    if (e >= 0) return e;
    int32_t e12 = (int32_t)m->n_he, e21 = e12 + 1;
    m->n_he += 2;
    m->origin[e12] = v1; m->origin[e21] = v2;
    m->twin[e12] = e21; m->twin[e21] = e12;
    m->next[e12] = m->next[e21] = -1;
    m->face[e12] = m->face[e21] = -1;
    map_put(m, k12, e12);
    map_put(m, ((uint64_t)(uint32_t)v2 << 32) | (uint32_t)v1, e21);
    if (m->out[v1] < 0) m->out[v1] = e12;
    if (m->out[v2] < 0) m->out[v2] = e21;
    return e12;
}

/* Mesh.from_triangle_soup (mesh.py:302-378) */
static int he_build(hemesh *m, int64_t n_v, int64_t n_t, const int32_t *tri) {
    memset(m, 0, sizeof *m);
    m->n_v = n_v;
    int64_t max_he = 6 * n_t + 2;
    m->cap = 64;
    while (m->cap < (uint64_t)max_he * 2) m->cap <<= 1;
    m->origin = malloc(sizeof(int32_t) * max_he);
    m->twin = malloc(sizeof(int32_t) * max_he);
    m->next = malloc(sizeof(int32_t) * max_he);
    m->face = malloc(sizeof(int32_t) * max_he);
    m->out = malloc(sizeof(int32_t) * (n_v > 0 ? n_v : 1));
    m->keys = malloc(sizeof(uint64_t) * m->cap);
    m->vals = malloc(sizeof(int32_t) * m->cap);
    if (!m->origin || !m->twin || !m->next || !m->face || !m->out || !m->keys || !m->vals) return ORACLE_ENOMEM;
    memset(m->keys, 0xff, sizeof(uint64_t) * m->cap);
    for (int64_t v = 0; v < n_v; v++) m->out[v] = -1;

    for (int64_t t = 0; t < n_t; t++) {
        int32_t v[3] = {tri[3 * t], tri[3 * t + 1], tri[3 * t + 2]};
        int32_t h[3];
        for (int k = 0; k < 3; k++) {
            if (v[k] < 0 || v[k] >= n_v) return ORACLE_EINVAL;
        }
        for (int k = 0; k < 3; k++) {          /* (v1,v2), (v2,v3), (v3,v1) */
            int32_t u = v[k], w = v[(k + 1) % 3];
            h[k] = he_connect(m, u, w);
            m->out[u] = h[k];                  /* u.out = hedge  (mesh.py:322) */
            m->face[h[k]] = (int32_t)t;
        }
        for (int k = 0; k < 3; k++) m->next[h[k]] = h[(k + 1) % 3];
    }
    /* boundary loops: next of a boundary half-edge = the boundary half-edge leaving its head */
    int32_t *v2b = malloc(sizeof(int32_t) * (n_v > 0 ? n_v : 1));
    if (!v2b) return ORACLE_ENOMEM;
    for (int64_t v = 0; v < n_v; v++) v2b[v] = -1;
    for (int64_t e = 0; e < m->n_he; e++) {
        if (m->face[e] != -1) continue;
        if (v2b[m->origin[e]] >= 0) { free(v2b); return ORACLE_ENONMANIFOLD; }
        v2b[m->origin[e]] = (int32_t)e;
    }
    for (int64_t e = 0; e < m->n_he; e++) {
        if (m->face[e] != -1) continue;
        int32_t head = m->origin[m->twin[e]];
        m->next[e] = v2b[head];                /* HalfEdge.connect(hedge_prev, hedge_next) */
    }
    for (int64_t e = 0; e < m->n_he; e++)
        if (m->face[e] == -1) m->face[e] = -2; /* Face(is_boundary=True) */
    free(v2b);
    return ORACLE_OK;
}

/* HalfEdge.cotan (mesh.py:124-139).  xy is this mesh's coordinate block. */
static int he_cotan(const hemesh *m, const double *xy, int32_t self, double *out) {
    int32_t vi = m->origin[self];
    int32_t vk = m->origin[m->twin[self]];
    double ratio = 0.;
    int32_t cand[2];
    cand[0] = m->next[m->next[self]];
    cand[1] = m->next[m->next[m->twin[self]]];
    for (int c = 0; c < 2; c++) {
        int32_t other = cand[c];
        if (other < 0 || m->next[other] < 0) return ORACLE_EINVAL;
        if (m->face[m->next[other]] == -2) continue;          /* boundary face: skip */
        int32_t vo = m->origin[other];
        double vix = xy[2 * vi] - xy[2 * vo], viy = xy[2 * vi + 1] - xy[2 * vo + 1];
        double vkx = xy[2 * vk] - xy[2 * vo], vky = xy[2 * vk + 1] - xy[2 * vo + 1];
        double dot = vix * vkx + viy * vky;
        double cross = vix * vky - viy * vkx;
        if (cross == 0.0) return ORACLE_EDEGENERATE;          /* ZeroDivisionError */
        ratio += fabs(dot / cross) / 2;
    }
    *out = ratio;
    return ORACLE_OK;
}

/* ------------------------------------------------------------------------------------------ */
/* lil_matrix: per-row sorted (col, val) lists; a value of exactly 0 deletes the entry          */
/* ------------------------------------------------------------------------------------------ */
typedef struct { int32_t *col; double *val; int32_t len, cap; } lilrow;
typedef struct { int64_t n; lilrow *rows; } lilmat;

static int lil_find(const lilrow *r, int32_t j) { /* lower bound */
    int lo = 0, hi = r->len;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (r->col[mid] < j) lo = mid + 1; else hi = mid; }
    return lo;
}
static double lil_get(const lilmat *L, int64_t i, int32_t j) {
    const lilrow *r = &L->rows[i];
    int p = lil_find(r, j);
    return (p < r->len && r->col[p] == j) ? r->val[p] : 0.0;
}
static int lil_set(lilmat *L, int64_t i, int32_t j, double x) {
    lilrow *r = &L->rows[i];
    int p = lil_find(r, j);
    int present = (p < r->len && r->col[p] == j);
    if (x == 0.0) {                                /* lil_insert: zero -> delete */
        if (present) {
            memmove(r->col + p, r->col + p + 1, sizeof(int32_t) * (r->len - p - 1));
            memmove(r->val + p, r->val + p + 1, sizeof(double) * (r->len - p - 1));
            r->len--;
        }
        return ORACLE_OK;
    }
    if (present) { r->val[p] = x; return ORACLE_OK; }
    if (r->len == r->cap) {
        int32_t nc = r->cap ? r->cap * 2 : 8;
        int32_t *c2 = realloc(r->col, sizeof(int32_t) * nc);
        if (!c2) return ORACLE_ENOMEM;
        r->col = c2;
        double *v2 = realloc(r->val, sizeof(double) * nc);
        if (!v2) return ORACLE_ENOMEM;
        r->val = v2; r->cap = nc;
    }
    memmove(r->col + p + 1, r->col + p, sizeof(int32_t) * (r->len - p));
    memmove(r->val + p + 1, r->val + p, sizeof(double) * (r->len - p));
    r->col[p] = j; r->val[p] = x; r->len++;
    return ORACLE_OK;
}
static void lil_free(lilmat *L) {
    if (!L->rows) return;
    for (int64_t i = 0; i < L->n; i++) { free(L->rows[i].col); free(L->rows[i].val); }
    free(L->rows); L->rows = NULL;
}

/* laplace_operator (solver.py:171-213) scaled by the conductance and scattered with
 * L[gi, gj] += v (solver.py:568-575).  The COO is emitted and consumed in the reference's order:
 * all off-diagonals in vertex/orbit order, then the diagonals. */
static int laplace_into_lil(const hemesh *m, const double *xy, double conductance,
                            int64_t goff, lilmat *L) {
    int64_t n = m->n_v;
    double *diag = calloc(n > 0 ? n : 1, sizeof(double));
    if (!diag) return ORACLE_ENOMEM;
    int rc = ORACLE_OK;
    for (int64_t i = 0; i < n && rc == ORACLE_OK; i++) {
        int32_t start = m->out[i];
        if (start < 0) { rc = ORACLE_EINVAL; break; }   /* vertex in no triangle: orbit() fails */
        int32_t e = start;
        int64_t guard = 0;
        do {                                            /* Vertex.orbit (mesh.py:78-84) */
            double ratio;
            rc = he_cotan(m, xy, e, &ratio);
            if (rc != ORACLE_OK) break;
            if (ratio != 0) {
                int32_t k = m->origin[m->twin[e]];
                double v = conductance * ratio;           /* conductance * laplace_operator(msh) */
                rc = lil_set(L, goff + i, (int32_t)(goff + k), lil_get(L, goff + i, (int32_t)(goff + k)) + v);
                if (rc != ORACLE_OK) break;
                diag[i] -= ratio;
            }
            e = m->next[m->twin[e]];
            if (e < 0 || ++guard > m->n_he) { rc = ORACLE_EINVAL; break; }
        } while (e != start);
    }
    for (int64_t i = 0; i < n && rc == ORACLE_OK; i++) {
        double v = conductance * diag[i];
        rc = lil_set(L, goff + i, (int32_t)(goff + i), lil_get(L, goff + i, (int32_t)(goff + i)) + v);
    }
    free(diag);
    return rc;
}

/* ------------------------------------------------------------------------------------------ */
/* public entry points (ctypes)                                                                 */
/* ------------------------------------------------------------------------------------------ */

/* assemble_system (solver.py:783-812).  Two-call protocol: pass indices == NULL to get nnz.
 * The handle keeps the LIL between the two calls. */
typedef struct { lilmat L; int64_t nnz; } oracle_system;

int oracle_assemble(int32_t n_mesh, const int64_t *v_off, const double *xy,
                    const int64_t *t_off, const int32_t *tri, const double *conductance,
                    int64_t N, int64_t n_stamp, const int32_t *s_row, const int32_t *s_col,
                    const double *s_val, const int32_t *s_op,
                    oracle_system **out, int64_t *nnz_out) {
    if (!out || N <= 0) return ORACLE_EINVAL;
    oracle_system *S = calloc(1, sizeof *S);
    if (!S) return ORACLE_ENOMEM;
    S->L.n = N;
    S->L.rows = calloc(N, sizeof(lilrow));
    if (!S->L.rows) { free(S); return ORACLE_ENOMEM; }
    int rc = ORACLE_OK;
    for (int32_t m = 0; m < n_mesh && rc == ORACLE_OK; m++) {
        hemesh hm;
        int64_t nv = v_off[m + 1] - v_off[m], nt = t_off[m + 1] - t_off[m];
        rc = he_build(&hm, nv, nt, tri + 3 * t_off[m]);
        if (rc == ORACLE_OK) rc = laplace_into_lil(&hm, xy + 2 * v_off[m], conductance[m], v_off[m], &S->L);
        he_free(&hm);
    }
    for (int64_t s = 0; s < n_stamp && rc == ORACLE_OK; s++) {
        int64_t i = s_row[s]; int32_t j = s_col[s];
        if (i < 0 || i >= N || j < 0 || j >= N) { rc = ORACLE_EINVAL; break; }
        if (s_op[s] == 1) rc = lil_set(&S->L, i, j, s_val[s]);                         /* L[i,j] = x  */
        else rc = lil_set(&S->L, i, j, lil_get(&S->L, i, j) + s_val[s]);              /* L[i,j] += x */
    }
    if (rc != ORACLE_OK) { lil_free(&S->L); free(S); return rc; }
    int64_t nnz = 0;
    for (int64_t i = 0; i < N; i++) nnz += S->L.rows[i].len;
    S->nnz = nnz;
    *out = S;
    if (nnz_out) *nnz_out = nnz;
    return ORACLE_OK;
}

/* L.tocsr(): sorted column indices, int32, no explicit zeros */
int oracle_system_csr(const oracle_system *S, int32_t *indptr, int32_t *indices, double *data) {
    if (!S) return ORACLE_EINVAL;
    int64_t p = 0;
    for (int64_t i = 0; i < S->L.n; i++) {
        indptr[i] = (int32_t)p;
        const lilrow *r = &S->L.rows[i];
        memcpy(indices + p, r->col, sizeof(int32_t) * r->len);
        memcpy(data + p, r->val, sizeof(double) * r->len);
        p += r->len;
    }
    indptr[S->L.n] = (int32_t)p;
    return ORACLE_OK;
}

void oracle_system_free(oracle_system *S) {
    if (!S) return;
    lil_free(&S->L);
    free(S);
}

/* Per-corner element weights: w[t][c] = abs(dot/cross)/2 at corner c of triangle t, i.e. the
 * contribution of face t to the cotan() of the edge opposite corner c (mesh.py:136-138).
 * tri holds indices into xy (mesh-local + offset already applied by the caller, or one mesh). */
int oracle_corner_weights(int64_t n_v, const double *xy, int64_t n_t, const int32_t *tri, double *w) {
    for (int64_t t = 0; t < n_t; t++) {
        for (int c = 0; c < 3; c++) {
            int32_t vo = tri[3 * t + c], vi = tri[3 * t + (c + 1) % 3], vk = tri[3 * t + (c + 2) % 3];
            if (vo < 0 || vo >= n_v || vi < 0 || vi >= n_v || vk < 0 || vk >= n_v) return ORACLE_EINVAL;
            double vix = xy[2 * vi] - xy[2 * vo], viy = xy[2 * vi + 1] - xy[2 * vo + 1];
            double vkx = xy[2 * vk] - xy[2 * vo], vky = xy[2 * vk + 1] - xy[2 * vo + 1];
            double dot = vix * vkx + viy * vky;
            double cross = vix * vky - viy * vkx;
            if (cross == 0.0) return ORACLE_EDEGENERATE;
            w[3 * t + c] = fabs(dot / cross) / 2;
        }
    }
    return ORACLE_OK;
}

/* compute_triangle_gradient + compute_power_density (solver.py:689-745), per face:
 * p = (E * conductivity) . E with E from the barycentric interpolant evaluated at
 * (x1+1, y1) and (x1, y1+1).  val holds the vertex voltages indexed like xy. */
int oracle_power_density(int64_t n_v, const double *xy, const double *val, int64_t n_t,
                         const int32_t *tri, double conductivity, double *p_out) {
    for (int64_t t = 0; t < n_t; t++) {
        int32_t a = tri[3 * t], b = tri[3 * t + 1], c = tri[3 * t + 2];
        if (a < 0 || a >= n_v || b < 0 || b >= n_v || c < 0 || c >= n_v) return ORACLE_EINVAL;
        double x1 = xy[2 * a], y1 = xy[2 * a + 1], x2 = xy[2 * b], y2 = xy[2 * b + 1];
        double x3 = xy[2 * c], y3 = xy[2 * c + 1];
        double f1 = val[a], f2 = val[b], f3 = val[c];
        double g[2];
        for (int k = 0; k < 2; k++) {
            double x = k == 0 ? x1 + 1 : x1, y = k == 0 ? y1 : y1 + 1;
            double D = (y2 - y3) * (x1 - x3) + (x3 - x2) * (y1 - y3);
            double l1 = ((y2 - y3) * (x - x3) + (x3 - x2) * (y - y3)) / D;
            double l2 = ((y3 - y1) * (x - x3) + (x1 - x3) * (y - y3)) / D;
            double l3 = 1 - l1 - l2;
            g[k] = (l1 * f1 + l2 * f2 + l3 * f3) - f1;
        }
        double jx = g[0] * conductivity, jy = g[1] * conductivity;  /* J = E * conductivity */
        p_out[t] = jx * g[0] + jy * g[1];                            /* J.dot(E) */
    }
    return ORACLE_OK;
}
