// pdn_common.cuh -- context, device buffers, launch bookkeeping and small device-wide primitives
// shared by the kernels of libpadne_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <utility>

#include <string>
#include <vector>

#include "../../include/padne_b200.h"

// ------------------------------------------------------------------------------------------------
// error handling
// ------------------------------------------------------------------------------------------------
struct pdn_error {
    int code;
    std::string msg;
};

#define PDN_CUDA(expr)                                                                          \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            char _b[512];                                                                       \
            snprintf(_b, sizeof _b, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr,               \
                     cudaGetErrorString(_e));                                                   \
            throw pdn_error{_e == cudaErrorMemoryAllocation ? PDN_ENOMEM : PDN_ECUDA, _b};      \
        }                                                                                       \
    } while (0)

#define PDN_FAIL(code_, ...)                                                                    \
    do {                                                                                        \
        char _b[512];                                                                           \
        snprintf(_b, sizeof _b, __VA_ARGS__);                                                   \
        throw pdn_error{(code_), _b};                                                           \
    } while (0)

// ------------------------------------------------------------------------------------------------
// device buffer (RAII, grows only)
// ------------------------------------------------------------------------------------------------
template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t cap = 0;  // elements
    bool owned = true;   // false: p points into memory owned by somebody else (attach)
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    void release() {
        if (p && owned) cudaFree(p);
        p = nullptr;
        cap = 0;
        owned = true;
    }
    // borrow n elements at ptr (never freed here)
    void attach(T *ptr, size_t n) {
        release();
        p = ptr;
        cap = n;
        owned = false;
    }
    // contents are NOT preserved
    T *reserve(size_t n) {
        if (n > cap) {
            release();
            size_t want = n + n / 16 + 64;
            PDN_CUDA(cudaMalloc((void **)&p, want * sizeof(T)));
            cap = want;
        }
        return p;
    }
    void swap(DevBuf &o) {
        T *tp = p; p = o.p; o.p = tp;
        size_t tc = cap; cap = o.cap; o.cap = tc;
        bool to = owned; owned = o.owned; o.owned = to;
    }
};

// A borrowed-or-staged input array: device pointer either given by the caller or uploaded.
template <typename T>
struct DevIn {
    const T *p = nullptr;
    DevBuf<T> own;
    void set(const T *src, size_t n, bool on_device, cudaStream_t s) {
        if (on_device || n == 0) { p = src; return; }
        own.reserve(n);
        PDN_CUDA(cudaMemcpyAsync(own.p, src, n * sizeof(T), cudaMemcpyHostToDevice, s));
        p = own.p;
    }
};

// ------------------------------------------------------------------------------------------------
// resident systems
// ------------------------------------------------------------------------------------------------
struct CsrDev {
    int64_t n = 0, nnz = 0;
    DevBuf<int32_t> indptr, indices;
    DevBuf<double> data;
    bool valid = false;
};

struct SolveWork;     // pdn_solve.cu
struct AssembleWork;  // pdn_assemble.cu (scratch buffers kept between calls)
struct PdnComm;       // pdn_multi.cu

// PADNE_B200_KTIME=1: the Krylov loop runs its first graph-worth of iterations eagerly with an event after
// every launch and prints mean event-to-event times per kernel and rank (multi-rank runs cannot use ncu).
// Short kernels are launch-bound in eager mode: the numbers locate a slow kernel, they are not a profile.
struct KernelTimes {
    std::vector<std::pair<const char *, cudaEvent_t>> marks;
    bool active = false;
    void mark(const char *name, cudaStream_t st) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        cudaEventRecord(e, st);
        marks.emplace_back(name, e);
    }
    void report(int rank);
};

struct pdn_ctx {
    KernelTimes *ktime = nullptr;
    int device = 0;
    cudaStream_t stream = nullptr;
    int n_sm = 148;
    int64_t launches = 0;
    std::string last_error;
    CsrDev L;                  // the full system of padne/solver.py:783-812
    // Partitioned assembly (pdn_assemble_part on a context with a communicator): only the rows [part_lo, part_hi),
    // the constraint terminals and the extra rows of L are assembled on this rank; the other rows are empty.
    bool part_valid = false;
    int64_t part_lo = 0, part_hi = 0, part_n_nodes = 0;
    int32_t *d_status = nullptr;  // [4] device status words (error code, detail)
    int32_t *h_status = nullptr;  // pinned mirror
    DevBuf<char> l2_flush;
    SolveWork *solve = nullptr;
    AssembleWork *asmw = nullptr;
    PdnComm *comm = nullptr;      // set by pdn_comm_init (one rank per GPU); null = single GPU
    // geometry of the resident system (kept by pdn_assemble for the coarse space of the solve)
    bool geo_valid = false;
    const double *geo_xy = nullptr;   // [geo_n_v, 2] device pointer (into asmw or geo_xy_buf)
    DevBuf<double> geo_xy_buf;
    DevBuf<int64_t> geo_voff;         // [geo_n_mesh + 1] device copy
    int geo_n_mesh = 0;
    int64_t geo_n_v = 0;
    DevBuf<int32_t> scan_tmp32;   // tile totals of the device-wide scans
    DevBuf<int64_t> scan_tmp64;
};

#define PDN_LAUNCH(ctx, kernel, grid, block, smem, ...)                                         \
    do {                                                                                        \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                        \
        (ctx)->launches++;                                                                      \
        PDN_CUDA(cudaGetLastError());                                                           \
        if ((ctx)->ktime && (ctx)->ktime->active) (ctx)->ktime->mark(#kernel, (ctx)->stream);   \
    } while (0)

// Programmatic dependent launch (PDL): the kernel may be scheduled while its predecessor in the stream is
// still draining, which hides most of the ~2 us kernel boundary inside the Krylov graph.  Such a kernel calls
// pdl_wait() before it touches anything the predecessor wrote (everything before the wait must be
// independent of it) and pdl_trigger() to let its own successor be scheduled early.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
static inline cudaError_t pdn_launch_ex(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                        bool pdl, Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
#define PDN_LAUNCH_PDL(ctx, pdl, kernel, grid, block, smem, ...)                                              \
    do {                                                                                                      \
        PDN_CUDA(pdn_launch_ex(kernel, dim3(grid), dim3(block), (size_t)(smem), (ctx)->stream, (pdl), __VA_ARGS__)); \
        (ctx)->launches++;                                                                                    \
        if ((ctx)->ktime && (ctx)->ktime->active) (ctx)->ktime->mark(#kernel, (ctx)->stream);                 \
    } while (0)
#endif

static inline int pdn_blocks(int64_t n, int block) { return (int)((n + block - 1) / block); }

// grid for a grid-stride kernel: a multiple of the SM count, capped by the work available
static inline int pdn_grid(const pdn_ctx *ctx, int64_t n, int block, int per_sm) {
    int64_t need = (n + block - 1) / block;
    int64_t full = (int64_t)ctx->n_sm * per_sm;
    if (need < 1) need = 1;
    return (int)(need < full ? need : full);
}

// wrap a C-ABI body: set device, translate exceptions to status codes
template <typename F>
static inline int pdn_guard(pdn_ctx *ctx, F &&body) {
    if (!ctx) return PDN_EINVAL;
    try {
        PDN_CUDA(cudaSetDevice(ctx->device));
        body();
        return PDN_OK;
    } catch (const pdn_error &e) {
        ctx->last_error = e.msg;
        cudaGetLastError();  // clear non-sticky errors
        return e.code;
    } catch (const std::bad_alloc &) {
        ctx->last_error = "host allocation failed";
        return PDN_ENOMEM;
    }
}

// device status helpers (pdn_api.cu)
void pdn_status_reset(pdn_ctx *ctx);
// returns the device status code (0 if none) after synchronising the stream; detail in *detail
int pdn_status_fetch(pdn_ctx *ctx, int64_t *detail);

// device-wide exclusive scan of int32 counts into int64 or int32 offsets (pdn_scan.cu).
// out has n+1 entries; out[n] = total.  Returns nothing; total can be read from out[n].
void pdn_exclusive_scan_i32(pdn_ctx *ctx, const int32_t *in, int64_t n, int32_t *out);
void pdn_exclusive_scan_i64(pdn_ctx *ctx, const int32_t *in, int64_t n, int64_t *out);

// flush L2 by writing a buffer larger than it
void pdn_flush_l2(pdn_ctx *ctx);

// entry points implemented across translation units
void pdn_assemble_impl(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const double *xy,
                       const int64_t *t_off, const int32_t *tri, const double *conductance,
                       int64_t N, int64_t n_stamp, const int32_t *s_row, const int32_t *s_col,
                       const double *s_val, const int32_t *s_op, bool on_device, int64_t *nnz_out,
                       int64_t n_nodes_part = -1);
void pdn_element_weights_impl(pdn_ctx *ctx, int64_t n_v, const double *xy, int64_t n_t,
                              const int32_t *tri, double *w, bool on_device);
void pdn_nearest_vertex_impl(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const int32_t *mesh_layer,
                             const double *xy, int64_t n_q, const double *q_xy, const int32_t *q_layer,
                             int64_t *out_index, double *out_d2, bool on_device);
void pdn_power_density_impl(pdn_ctx *ctx, int64_t n_v, const double *xy, const double *val,
                            int64_t n_t, const int32_t *tri, double conductivity, double *p,
                            bool on_device);
void pdn_solve_impl(pdn_ctx *ctx, const double *rhs, const pdn_opts *opts, double *v,
                    pdn_info *info, bool on_device);
void pdn_spmv_impl(pdn_ctx *ctx, const double *x, double *y, bool on_device);
void pdn_bench_spmv_impl(pdn_ctx *ctx, int32_t reps, int mode, double *ms);
void pdn_solve_release(pdn_ctx *ctx);
void pdn_assemble_release(pdn_ctx *ctx);

// ---- peer-memory exchanges (pdn_peer.cuh / pdn_multi.cu) ---------------------------------------------
constexpr int PDN_CO_MAX = 1776;          // largest coarse space (pdn_coarse.cuh)
constexpr int PDN_PEER_MAX = 8;
// arena layout, in 8-byte words
constexpr size_t PDN_PA_HALO_FLAG = 0;                              // [PEER_MAX]
constexpr size_t PDN_PA_RED_FLAG = 16;                              // [2][PEER_MAX]
constexpr size_t PDN_PA_RC_FLAG = 32;                               // [2][PEER_MAX]
constexpr size_t PDN_PA_CW_FLAG = 48;                               // [2][PEER_MAX] merged dots + c_w exchange (pdn_cg2.cuh)
constexpr size_t PDN_PA_RED_VAL = 64;                               // [2][PEER_MAX][8] flag-in-word pairs of 3 doubles
constexpr size_t PDN_PA_RC_SLOT = 2 * PDN_CO_MAX + 8;               // header + two flag-in-word words per 64-bit entry
constexpr size_t PDN_PA_RC_VAL = 192;                               // [2][PEER_MAX][RC_SLOT]
constexpr size_t PDN_PA_CW_VAL = PDN_PA_RC_VAL + (size_t)2 * PDN_PEER_MAX * PDN_PA_RC_SLOT;   // [2][PEER_MAX][RC_SLOT]: 4 header words + c_w
constexpr size_t PDN_PA_VEC = PDN_PA_CW_VAL + (size_t)2 * PDN_PEER_MAX * PDN_PA_RC_SLOT;   // [4][n_cap]
constexpr int PDN_PA_NVEC = 4;            // halo-exchanged vectors living in the arena: z, ph, sh, x

constexpr int PDN_PEER_LOG = 4096;          // entries of the exchange event log behind the 8 epoch words
struct PdnPeerLink {
    int world, me;
    unsigned long long *arena[PDN_PEER_MAX];   // arena[k]: rank k's arena as mapped in this process
    unsigned long long *epoch;                 // local: [0] halo, [1] dot sums, [2] coarse residual, [7] merged dots + c_w
    int32_t *status;                           // local device status words
};
// true when the communicator exchanges through mapped peer memory (else NCCL collectives)
bool pdn_peer_enabled(const pdn_ctx *ctx);
// collective: (re)allocate the arena for vectors of n_cap doubles and map every peer's
void pdn_peer_ensure(pdn_ctx *ctx, int64_t n_cap);
PdnPeerLink pdn_peer_link(const pdn_ctx *ctx);
// the same in device memory (valid after pdn_peer_ensure): hot kernels take this pointer, a by-value struct
// indexed by rank would be copied to local memory by every thread
const PdnPeerLink *pdn_peer_link_dev(const pdn_ctx *ctx);
double *pdn_peer_vec(const pdn_ctx *ctx, int which);   // vector `which` (0..3) inside the local arena
size_t pdn_peer_vec_offset(const pdn_ctx *ctx, int which);   // its word offset inside every arena
unsigned int *pdn_peer_ticket(const pdn_ctx *ctx);

// multi-GPU plumbing (pdn_multi.cu); all collectives are enqueued on the context stream
int pdn_comm_rank(const pdn_ctx *ctx);
int pdn_comm_world(const pdn_ctx *ctx);
void pdn_comm_release(pdn_ctx *ctx);
void pdn_comm_allreduce_sum(pdn_ctx *ctx, double *buf, int count);
void pdn_comm_allreduce_sum_i64(pdn_ctx *ctx, long long *buf, int count);
void pdn_comm_allreduce_max(pdn_ctx *ctx, double *buf, int count);
void pdn_comm_allreduce_sum_i32(pdn_ctx *ctx, int32_t *buf, size_t count);
void pdn_comm_allgatherv_i32(pdn_ctx *ctx, int32_t *vec, const int64_t *begin);

// Row partition of a system with n node rows over `world` ranks: contiguous ranges of whole 1024-row sort windows
// (pdn_sell.cuh), split evenly; begin has world + 1 entries.  Used by the partitioned assembly and by the solve.
constexpr int64_t PDN_PART_WIN = 1024;
static inline void pdn_row_partition(int64_t n, int world, int64_t *begin) {
    const int64_t n_win = (n + PDN_PART_WIN - 1) / PDN_PART_WIN;
    for (int k = 0; k <= world; k++) {
        const int64_t b = (int64_t)k * n_win / world * PDN_PART_WIN;
        begin[k] = b < n ? b : n;
    }
    begin[world] = n;
}
void pdn_comm_allgatherv(pdn_ctx *ctx, double *vec, const int64_t *begin);
void pdn_comm_exchange(pdn_ctx *ctx, const double *sendbuf, const int32_t *send_off, const int32_t *send_cnt,
                       double *recvbuf, const int32_t *recv_off, const int32_t *recv_cnt);

