/*
 * padne_b200.h -- C ABI of the B200-native padne FEM hot path (libpadne_b200.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch or
 * Python types.  Every entry point names the reference interface it replaces; paths are relative
 * to the reference repository (atx/padne).
 *
 * Conventions
 *   - every function returns an int status, PDN_OK == 0 (pdn_strerror() gives the text);
 *   - all array arguments are caller-owned and only borrowed for the duration of the call;
 *   - `on_device != 0` means every ARRAY pointer of that call is a device pointer on the context's
 *     GPU (e.g. a torch tensor's data_ptr()); otherwise they are host pointers and the call does
 *     the host<->device copies itself;
 *   - one in-flight call per pdn_ctx (the reference is single-threaded, solver.py:783-812);
 *   - the assembled system stays RESIDENT on the device inside the context between
 *     pdn_assemble / pdn_set_csr and pdn_solve, so the solve needs no second upload.
 *
 * Unknown order (padne/solver.py:757-760):
 *   [ mesh vertices | internal network nodes | one current per VoltageSource/VoltageRegulator |
 *     ground current ]            n_nodes = vertices + internal nodes;  N = n_nodes + extras + 1
 */
#ifndef PADNE_B200_H
#define PADNE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes -> the exception the Python shim raises (reference behaviour in brackets) */
#define PDN_OK 0
#define PDN_EINVAL 1        /* ValueError                                                         */
#define PDN_EDEGENERATE 2   /* ZeroDivisionError  [HalfEdge.cotan, padne/mesh.py:138]             */
#define PDN_ESINGULAR 3     /* singular / inconsistent constraints [SuperLU: "Matrix is exactly   */
                            /* singular" warning and an all-NaN v]                                */
#define PDN_ENOTCONVERGED 4 /* RuntimeError: Krylov solver hit max_iter (never silent)            */
#define PDN_ECUDA 5         /* RuntimeError: CUDA / NCCL failure                                  */
#define PDN_EUNSUPPORTED 6  /* NotImplementedError [padne/solver.py:454-456, 540-541]             */
#define PDN_ENONMANIFOLD 7  /* ValueError("Non-manifold mesh") [padne/mesh.py:343]                */
#define PDN_ENOMEM 8        /* MemoryError                                                        */
#define PDN_ENOSYSTEM 9     /* RuntimeError: no resident system (call pdn_assemble/pdn_set_csr)   */

/* stamp operations: scipy lil_matrix `+=` versus `=` (padne/solver.py:480-484 vs :503-511) */
#define PDN_OP_ADD 0
#define PDN_OP_SET 1

/* Krylov methods (pdn_opts.method / pdn_info.method) */
#define PDN_METHOD_AUTO 0     /* PCG when the reduced system is symmetric, else BiCGSTAB          */
#define PDN_METHOD_PCG 1
#define PDN_METHOD_BICGSTAB 2

/* preconditioners */
#define PDN_PRECOND_AUTO 0      /* two-level when the system came from pdn_assemble (coordinates are
                                   resident) and has >= 1024 rows, else Jacobi                          */
#define PDN_PRECOND_NONE 1
/* (2 is unassigned) */
#define PDN_PRECOND_TWOLEVEL 3  /* Jacobi + coarse space of piecewise constants over coordinate bins (PCG
                                   and BiCGSTAB); partitioned PCG adds a rank-local level from 4 ranks  */
#define PDN_PRECOND_JACOBI 4

typedef struct pdn_ctx pdn_ctx;

typedef struct pdn_opts {
    double rtol;        /* required: ||b - A u||_2 <= rtol * ||b||_2 on the reduced system; 0 -> 1e-10 */
    double atol;        /* AND the reference's absolute bar on ||L v - r||_2 (its own test asserts < 1e-9,
                           tests/test_solver.py:2083-2089): the loop runs on to max(atol / 4, the fp64 floor
                           4 u amax ||u||_2); 0 -> 1e-9, < 0 -> disabled (rtol only)                     */
    int32_t max_iter;   /* 0 -> 20 * sqrt(N) + 2000                                                 */
    int32_t method;     /* PDN_METHOD_*                                                            */
    int32_t precond;    /* PDN_PRECOND_*                                                           */
    int32_t check_every;/* iterations per CUDA-graph launch between host convergence polls; 0 -> 32 */
    int64_t n_nodes;    /* voltage unknowns; <0 -> detect: the trailing rows without a diagonal entry that
                           reference only earlier columns are the constraint rows                     */
    int32_t n_coarse;   /* target number of aggregates of the two-level preconditioner; 0 -> 1024 below
                           200 000 rows, 1600 above; capped at 1776                                    */
    int32_t reserved;
} pdn_opts;

typedef struct pdn_info {
    int32_t method;             /* PDN_METHOD_PCG or PDN_METHOD_BICGSTAB actually used             */
    int32_t iterations;
    int32_t restarts;           /* true-residual restarts                                          */
    int32_t n_constraints;      /* extra unknowns eliminated                                       */
    double rel_residual;        /* ||b - A u|| / ||b|| of the reduced system (true residual)       */
    double residual_norm;       /* ||L v - r||_2 of the FULL system, as padne/solver.py:775        */
    double ground_node_current; /* v[N-1], padne/solver.py:777                                      */
    double t_setup_ms;          /* constraint elimination + reduced operator                        */
    double t_krylov_ms;         /* Krylov iterations (CUDA events)                                  */
    double t_total_ms;
    double spmv_ms;             /* mean duration of one reduced-operator SpMV launch, measured by   */
                                /* pdn_bench_spmv-style events during the solve (0 if not measured) */
    int32_t precond;            /* PDN_PRECOND_* actually used                                      */
    int32_t n_coarse;           /* aggregates of the coarse space (0 without one)                   */
    double t_coarse_ms;         /* part of t_setup_ms spent building and inverting the coarse matrix */
    int32_t n_coarse_local;     /* aggregates of THIS rank's local level (partitioned PCG; 0 without one) */
    int32_t local_level;        /* 0 not requested, 1 built on every rank, -1 requested but dropped         */
    int32_t stagnated;          /* 1: the true residual stopped improving above the absolute target (fp64
                                   floor) -- rtol still holds, or the call would have failed               */
    int32_t iteration_kind;     /* 2: three-kernel PCG iteration with one reduction; 1: round-1 iteration   */
    double abs_target;          /* absolute bar on the reduced residual the loop finally ran to             */
    double floor_abs;           /* 4 u amax ||u||_2 (0 if it was never needed)                              */
    double rhs_norm;            /* ||b||_2 of the reduced system                                            */
    int32_t spmv_kind;          /* 1: x-band staging SpMV, 0: TMA-ring SpMV                                 */
    int32_t spmv_halo;          /* rows of x staged either side of a CTA's rows                             */
    double spmv_window_coverage;/* fraction of stored entries gathered from the shared-memory window       */
} pdn_info;

/* ---- context ------------------------------------------------------------------------------- */
/* One context = one GPU, one stream, one resident system.  `device` is the CUDA ordinal.
 * Scratch buffers are kept inside the context and only grow, so repeated calls on same-sized
 * boards allocate nothing. */
int pdn_create(pdn_ctx **out, int device);
int pdn_destroy(pdn_ctx *ctx);
const char *pdn_strerror(int status);
/* Text of the last CUDA/NCCL failure or argument complaint on this context ("" if none). */
const char *pdn_last_error(pdn_ctx *ctx);
/* Number of kernel launches issued by this context since creation (bench.py's gpu_launches). */
int64_t pdn_launch_count(pdn_ctx *ctx);
/* cudaStream_t of the context as an integer, so callers can record CUDA events on it. */
uint64_t pdn_stream(pdn_ctx *ctx);

/* ---- kernel (1): element weights --------------------------------------------------------------
 * Replaces HalfEdge.cotan (padne/mesh.py:124-139): w[t][c] = abs(dot/cross)/2 at corner c of
 * triangle t, i.e. the contribution of face t to the weight of the edge opposite corner c.
 * fp64, every * + - / rounded separately (no FMA), like CPython.
 * tri indexes xy directly.  A zero cross product -> PDN_EDEGENERATE. */
int pdn_element_weights(pdn_ctx *ctx, int64_t n_v, const double *xy /*[n_v,2]*/, int64_t n_t,
                        const int32_t *tri /*[n_t,3]*/, double *w /*[n_t,3]*/, int on_device);

/* ---- kernels (1)+(2): global assembly -----------------------------------------------------------
 * Replaces assemble_system (padne/solver.py:783-812) = laplace_operator per mesh (:171-213) scaled
 * by the mesh conductance and scattered at the VertexIndexer offset (:216-229, :563-575), then the
 * ordered lumped-element / ground stamps (:469-560) replayed with lil_matrix semantics
 * (an entry that becomes exactly 0.0 leaves the pattern).
 *   v_off/t_off [n_mesh+1]: vertex / triangle offsets of each mesh (back to back);
 *   tri: MESH-LOCAL vertex indices;  conductance [n_mesh];
 *   s_row/s_col/s_val/s_op [n_stamp]: stamp operations in the reference's execution order.
 * The result stays resident; *nnz_out receives the number of stored entries. */
int pdn_assemble(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const double *xy,
                 const int64_t *t_off, const int32_t *tri, const double *conductance, int64_t N,
                 int64_t n_stamp, const int32_t *s_row, const int32_t *s_col, const double *s_val,
                 const int32_t *s_op, int on_device, int64_t *nnz_out);

/* The same on a context with a communicator (pdn_comm_init), PARTITIONED: the matrix is block diagonal per mesh
 * (padne/solver.py:216-229, 563-575) and the coupling is the stamp list, so every rank builds only the rows of its
 * range of the solve's row partition (contiguous 1024-row windows of the n_nodes voltage rows, split evenly), the
 * extra rows (>= n_nodes) and the few constraint-terminal rows, and -- from host inputs -- uploads only the meshes
 * those rows live in.  All ranks pass the SAME full arrays; the resident system holds this rank's rows only
 * (pdn_get_csr returns them, the other rows empty; *nnz_out is the local count) and pdn_solve works from that,
 * exchanging what the setup needs (row permutation, halo marks, coarse matrix).  Without a communicator it is
 * pdn_assemble. */
int pdn_assemble_part(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const double *xy,
                      const int64_t *t_off, const int32_t *tri, const double *conductance, int64_t N,
                      int64_t n_nodes, int64_t n_stamp, const int32_t *s_row, const int32_t *s_col,
                      const double *s_val, const int32_t *s_op, int on_device, int64_t *nnz_out);

/* Copy the resident CSR out (L.tocsr(): int32 sorted indices, no explicit zeros).
 * indptr [N+1], indices [nnz], data [nnz].  Any pointer may be NULL to skip that array. */
int pdn_get_csr(pdn_ctx *ctx, int32_t *indptr, int32_t *indices, double *data, int on_device);
/* Shape of the resident system. */
int pdn_get_shape(pdn_ctx *ctx, int64_t *N, int64_t *nnz);
/* Make a caller-provided CSR the resident system (the solve_system(L, r) seam, solver.py:767). */
int pdn_set_csr(pdn_ctx *ctx, int64_t N, int64_t nnz, const int32_t *indptr, const int32_t *indices,
                const double *data, int on_device);

/* ---- kernels (3)+(4): the solve ------------------------------------------------------------------
 * Replaces solve_system (padne/solver.py:767-780): solves L v = rhs for the resident L.
 * The constraint rows/columns (indices >= n_nodes) are eliminated on the host (union-find with
 * voltage offsets), the reduced operator is solved with preconditioned CG (symmetric) or BiCGSTAB
 * (regulator gains present), the eliminated currents are recovered, and the residual of the
 * FULL system is reported exactly as solver.py:775.   rhs [N], v [N]. */
int pdn_solve(pdn_ctx *ctx, const double *rhs, const pdn_opts *opts, double *v, pdn_info *info,
              int on_device);

/* y = L x with the resident full system (diagnostics, tests, SpMV benchmarking). x,y [N]. */
int pdn_spmv(pdn_ctx *ctx, const double *x, double *y, int on_device);
/* Time `reps` launches of the solver's dominant kernel (the SELL-32-sigma SpMV with fused dot
 * products, on the reduced operator of the last pdn_solve) with CUDA events on the context stream; returns the mean
 * milliseconds per launch.  flush_l2: 0 = back-to-back (warm L2), 1 = a >L2 buffer is written
 * before every launch (cold), 2 = in situ: each SpMV is followed by the PCG vector update like a
 * solver iteration and only the SpMV is bracketed by the events. */
int pdn_bench_spmv(pdn_ctx *ctx, int32_t reps, int flush_l2, double *ms_per_launch);

/* ---- r = b - A x and ||r||_2^2 WITHOUT any device allocation -----------------------------------------
 * No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * A variant of the residual form of the CSR SpMV (the kernel behind pdn_spmv and pdn_info.residual_norm)
 * whose reduction scratch is a fixed file-scope device array instead of context-owned cudaMalloc'd memory:
 * the call makes no cudaMalloc / cudaFree / allocating library call and needs no pdn_ctx.
 *   device: CUDA ordinal; stream: a cudaStream_t of that device cast to an integer (0 = default stream);
 *   indptr [N+1], indices / data [nnz], x, b, r [N], rr [1]: DEVICE pointers, caller-owned;
 *   r[i] = b[i] - sum_k data[k] * x[indices[k]] (stored order, every operation rounded separately: the
 *   same bits as b - pdn_spmv(x)), rr[0] = sum_i r[i]^2 (fixed summation order: reproducible);
 *   err (may be NULL) receives a message of up to err_cap bytes when the status is not PDN_OK.
 * ASYNCHRONOUS: the work is queued on `stream`; r and rr are valid once the stream reaches that point.
 * Limit: N <= PDN_NOALLOC_MAX_ROWS, larger N -> PDN_EUNSUPPORTED.  The scratch (131 328 bytes: one partial
 * sum per 1024 rows) is held by every GPU that loads the library until the process exits, and is SHARED by
 * all calls on a GPU: calls that overlap on one GPU (different streams or threads) race on it. */
#define PDN_NOALLOC_MAX_ROWS 16777216 /* 2^24 */
int pdn_residual_noalloc(int device, uint64_t stream, int64_t N, const int32_t *indptr, const int32_t *indices,
                         const double *data, const double *x, const double *b, double *r, double *rr,
                         char *err, int64_t err_cap);

/* ---- multi-GPU: one process per GPU, NCCL over NVLink / NVSwitch -----------------------------------
 * No reference equivalent (the reference is single-threaded CPU code); this is the north star's
 * "large boards are partitioned across the GPUs of one box".  After pdn_comm_init every pdn_solve on
 * this context is COLLECTIVE: all ranks must call it with the same resident system and the same
 * right-hand side.  The reduced system's rows are split into contiguous ranges (whole 1024-row sort
 * windows), each rank iterates on its own rows, exchanges the interface entries of the SpMV input
 * with the ranks that need them (grouped ncclSend/ncclRecv) and all-reduces the fused dot products;
 * every rank returns the full v.
 *   pdn_comm_unique_id: call on rank 0, ship the 128 bytes to the other ranks (any transport);
 *   pdn_comm_init:      collective; world == 1 leaves the context in single-GPU mode. */
int pdn_comm_unique_id(void *out128);
int pdn_comm_init(pdn_ctx *ctx, int rank, int world, const void *id128);
int pdn_comm_destroy(pdn_ctx *ctx);
int pdn_comm_info(pdn_ctx *ctx, int *rank, int *world);

/* ---- next row F1: per-face power density ----------------------------------------------------------
 * Replaces compute_power_density / compute_triangle_gradient (padne/solver.py:689-745):
 * p[t] = conductivity * |grad V|^2 with the reference's exact barycentric expression order.
 * tri indexes xy / val directly. */
int pdn_power_density(pdn_ctx *ctx, int64_t n_v, const double *xy, const double *val, int64_t n_t,
                      const int32_t *tri, double conductivity, double *p /*[n_t]*/, int on_device);

/* ---- next row F3: nearest mesh vertex of every connection point ------------------------------------
 * Replaces the per-layer scipy KDTree of NodeIndexer (padne/solver.py:356-396 builds it, :421-433 queries
 * it once per Connection): out_index[q] = global index (VertexIndexer order, padne/solver.py:216-229) of the
 * vertex nearest to q_xy[q] among ALL meshes m with mesh_layer[m] == q_layer[q]; -1 when the layer has no
 * vertex.  Squared Euclidean distance dx*dx + dy*dy in binary64, the lowest index wins an exact tie
 * (a KD-tree's tie order is unspecified).  out_d2 (squared distances) may be NULL. */
int pdn_nearest_vertex(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off /*[n_mesh+1]*/,
                       const int32_t *mesh_layer /*[n_mesh]*/, const double *xy /*[n_v,2]*/, int64_t n_q,
                       const double *q_xy /*[n_q,2]*/, const int32_t *q_layer /*[n_q]*/,
                       int64_t *out_index /*[n_q]*/, double *out_d2 /*[n_q] or NULL*/, int on_device);

/* ---- next row F4: the text of a ParaView .vtu piece from flat arrays (host only: the output is a file) ----
 * Replaces the per-vertex / per-face object walks and string joins of padne/paraview.py:109-231
 * (create_point_data, create_points, _extract_triangle_connectivity, create_cells, create_piece).
 * Unlike the functions above these return a BYTE COUNT (>= 0) or a negated status code (< 0), need no
 * pdn_ctx and no GPU, and take host pointers only.
 *
 * pdn_format_f64: " ".join(str(v) for v in x) (padne/paraview.py:91) -- CPython's float repr, byte for byte.
 *   cap must be >= 25 * n.
 * pdn_vtu_piece: the <Piece> element of one mesh exactly as lxml pretty-prints it inside
 *   VTKFile/UnstructuredGrid (padne/paraview.py:276-287): PointData "voltage" = voltage[0..n_v), Points =
 *   (x, -y, 0.0) per vertex, Cells = tri in mesh.faces order / offsets 3(i+1) / types 5.
 *   cap must be >= pdn_vtu_piece_bound(n_v, n_t); a triangle index outside [0, n_v) gives -PDN_EINVAL. */
int64_t pdn_format_f64(const double *x, int64_t n, char *out, int64_t cap);
int64_t pdn_vtu_piece_bound(int64_t n_v, int64_t n_t);
int64_t pdn_vtu_piece(int64_t n_v, const double *xy /*[n_v,2]*/, const double *voltage /*[n_v]*/, int64_t n_t,
                      const int32_t *tri /*[n_t,3]*/, char *out, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif /* PADNE_B200_H */
