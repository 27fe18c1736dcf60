// pdn_api.cu -- the extern "C" surface of libpadne_b200.so (include/padne_b200.h) plus the small
// device-wide primitives (scan, L2 flush, status words) the kernel files share.
#include "pdn_common.cuh"

// ------------------------------------------------------------------------------------------------
// status words
// ------------------------------------------------------------------------------------------------
void pdn_status_reset(pdn_ctx *ctx) {
    PDN_CUDA(cudaMemsetAsync(ctx->d_status, 0, 4 * sizeof(int32_t), ctx->stream));
}

int pdn_status_fetch(pdn_ctx *ctx, int64_t *detail) {
    PDN_CUDA(cudaMemcpyAsync(ctx->h_status, ctx->d_status, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost,
                             ctx->stream));
    PDN_CUDA(cudaStreamSynchronize(ctx->stream));
    if (detail) *detail = ctx->h_status[1];
    return ctx->h_status[0];
}

// ------------------------------------------------------------------------------------------------
// exclusive scan: three phases (block totals, scan of totals by one block, add back)
// ------------------------------------------------------------------------------------------------
static constexpr int SCAN_BLOCK = 256;
static constexpr int SCAN_ITEMS = 8;  // per thread
static constexpr int SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;

template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T *total, T *smem /*[SCAN_BLOCK/32]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) smem[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        T w = lane < (SCAN_BLOCK / 32) ? smem[lane] : T(0);
        T winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            T t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        if (lane < (SCAN_BLOCK / 32)) smem[lane] = winc - w;  // exclusive warp offsets
        if (lane == (SCAN_BLOCK / 32) - 1) smem[SCAN_BLOCK / 32] = winc;
    }
    __syncthreads();
    T res = smem[warp] + inc - v;
    *total = smem[SCAN_BLOCK / 32];
    __syncthreads();
    return res;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_BLOCK) scan_tile_totals(const int32_t *__restrict__ in, int64_t n,
                                                               T *__restrict__ tile_tot) {
    __shared__ T sm[SCAN_BLOCK / 32 + 1];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    T s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (base + k < n) s += (T)in[base + k];
    T tot;
    block_exclusive_scan<T>(s, &tot, sm);
    if (threadIdx.x == 0) tile_tot[blockIdx.x] = tot;
}

// one block scans the tile totals in place (exclusive); tile count is small (n / 2048)
template <typename T>
__global__ void __launch_bounds__(SCAN_BLOCK) scan_totals_inplace(T *tile_tot, int64_t n_tiles, T *grand) {
    __shared__ T sm[SCAN_BLOCK / 32 + 1];
    T carry = 0;
    for (int64_t base = 0; base < n_tiles; base += SCAN_BLOCK) {
        int64_t i = base + threadIdx.x;
        T v = i < n_tiles ? tile_tot[i] : T(0);
        T tot;
        T ex = block_exclusive_scan<T>(v, &tot, sm);
        if (i < n_tiles) tile_tot[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) *grand = carry;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_BLOCK) scan_apply(const int32_t *__restrict__ in, int64_t n,
                                                         const T *__restrict__ tile_off, T *__restrict__ out) {
    __shared__ T sm[SCAN_BLOCK / 32 + 1];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    T v[SCAN_ITEMS];
    T s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = base + k < n ? (T)in[base + k] : T(0);
        s += v[k];
    }
    T tot;
    T ex = block_exclusive_scan<T>(s, &tot, sm) + tile_off[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) out[base + k] = ex;
        ex += v[k];
    }
}

template <typename T>
static void exclusive_scan_impl(pdn_ctx *ctx, const int32_t *in, int64_t n, T *out, DevBuf<T> &tmp) {
    int64_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (n_tiles < 1) n_tiles = 1;
    if ((size_t)n_tiles + 1 > tmp.cap) {
        PDN_CUDA(cudaStreamSynchronize(ctx->stream));   // an earlier scan may still read the old buffer
        tmp.reserve((size_t)n_tiles + 1);
    }
    PDN_LAUNCH(ctx, scan_tile_totals<T>, (int)n_tiles, SCAN_BLOCK, 0, in, n, tmp.p);
    PDN_LAUNCH(ctx, scan_totals_inplace<T>, 1, SCAN_BLOCK, 0, tmp.p, n_tiles, out + n);
    PDN_LAUNCH(ctx, scan_apply<T>, (int)n_tiles, SCAN_BLOCK, 0, in, n, tmp.p, out);
}

// the tile-total scratch lives in the context (stream-ordered reuse: no synchronisation needed)
void pdn_exclusive_scan_i32(pdn_ctx *ctx, const int32_t *in, int64_t n, int32_t *out) {
    exclusive_scan_impl<int32_t>(ctx, in, n, out, ctx->scan_tmp32);
}
void pdn_exclusive_scan_i64(pdn_ctx *ctx, const int32_t *in, int64_t n, int64_t *out) {
    exclusive_scan_impl<int64_t>(ctx, in, n, out, ctx->scan_tmp64);
}

// ------------------------------------------------------------------------------------------------
// L2 flush
// ------------------------------------------------------------------------------------------------
void pdn_flush_l2(pdn_ctx *ctx) {
    const size_t bytes = (size_t)256 << 20;  // 256 MiB > 126 MB L2
    ctx->l2_flush.reserve(bytes);
    PDN_CUDA(cudaMemsetAsync(ctx->l2_flush.p, 0, bytes, ctx->stream));
}

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

const char *pdn_strerror(int status) {
    switch (status) {
        case PDN_OK: return "ok";
        case PDN_EINVAL: return "invalid argument";
        case PDN_EDEGENERATE: return "degenerate (zero-area) triangle: float division by zero";
        case PDN_ESINGULAR: return "Matrix is exactly singular";
        case PDN_ENOTCONVERGED: return "Krylov solver did not converge";
        case PDN_ECUDA: return "CUDA failure";
        case PDN_EUNSUPPORTED: return "unsupported system structure";
        case PDN_ENONMANIFOLD: return "Non-manifold mesh";
        case PDN_ENOMEM: return "out of memory";
        case PDN_ENOSYSTEM: return "no resident system";
        default: return "unknown status";
    }
}

int pdn_create(pdn_ctx **out, int device) {
    if (!out) return PDN_EINVAL;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) {
        cudaGetLastError();
        return PDN_ECUDA;
    }
    pdn_ctx *ctx = new pdn_ctx();
    ctx->device = device;
    try {
        PDN_CUDA(cudaSetDevice(device));
        PDN_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        PDN_CUDA(cudaDeviceGetAttribute(&ctx->n_sm, cudaDevAttrMultiProcessorCount, device));
        PDN_CUDA(cudaMalloc((void **)&ctx->d_status, 4 * sizeof(int32_t)));
        PDN_CUDA(cudaMallocHost((void **)&ctx->h_status, 4 * sizeof(int32_t)));
        PDN_CUDA(cudaMemsetAsync(ctx->d_status, 0, 4 * sizeof(int32_t), ctx->stream));
    } catch (const pdn_error &) {
        delete ctx;
        return PDN_ECUDA;
    }
    *out = ctx;
    return PDN_OK;
}

int pdn_destroy(pdn_ctx *ctx) {
    if (!ctx) return PDN_EINVAL;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    pdn_solve_release(ctx);
    pdn_assemble_release(ctx);
    pdn_comm_release(ctx);
    ctx->scan_tmp32.release();
    ctx->scan_tmp64.release();
    ctx->geo_xy_buf.release();
    ctx->geo_voff.release();
    if (ctx->d_status) cudaFree(ctx->d_status);
    if (ctx->h_status) cudaFreeHost(ctx->h_status);
    ctx->L.indptr.release();
    ctx->L.indices.release();
    ctx->L.data.release();
    ctx->l2_flush.release();
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return PDN_OK;
}

const char *pdn_last_error(pdn_ctx *ctx) { return ctx ? ctx->last_error.c_str() : ""; }
int64_t pdn_launch_count(pdn_ctx *ctx) { return ctx ? ctx->launches : 0; }
uint64_t pdn_stream(pdn_ctx *ctx) { return ctx ? (uint64_t)(uintptr_t)ctx->stream : 0; }

int pdn_element_weights(pdn_ctx *ctx, int64_t n_v, const double *xy, int64_t n_t, const int32_t *tri,
                        double *w, int on_device) {
    return pdn_guard(ctx, [&] {
        if (n_v < 0 || n_t < 0 || (n_t > 0 && (!xy || !tri || !w))) PDN_FAIL(PDN_EINVAL, "element_weights: bad arguments");
        pdn_element_weights_impl(ctx, n_v, xy, n_t, tri, w, on_device != 0);
    });
}

int pdn_assemble(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const double *xy, const int64_t *t_off,
                 const int32_t *tri, const double *conductance, int64_t N, int64_t n_stamp,
                 const int32_t *s_row, const int32_t *s_col, const double *s_val, const int32_t *s_op,
                 int on_device, int64_t *nnz_out) {
    return pdn_guard(ctx, [&] {
        if (n_mesh < 0 || N <= 0 || n_stamp < 0 || !v_off || !t_off)
            PDN_FAIL(PDN_EINVAL, "assemble: bad sizes");
        if (n_stamp > 0 && (!s_row || !s_col || !s_val || !s_op)) PDN_FAIL(PDN_EINVAL, "assemble: stamp arrays missing");
        ctx->L.valid = false;
        ctx->geo_valid = false;
        pdn_assemble_impl(ctx, n_mesh, v_off, xy, t_off, tri, conductance, N, n_stamp, s_row, s_col, s_val,
                          s_op, on_device != 0, nnz_out);
    });
}

int pdn_assemble_part(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const double *xy, const int64_t *t_off,
                      const int32_t *tri, const double *conductance, int64_t N, int64_t n_nodes, int64_t n_stamp,
                      const int32_t *s_row, const int32_t *s_col, const double *s_val, const int32_t *s_op,
                      int on_device, int64_t *nnz_out) {
    return pdn_guard(ctx, [&] {
        if (n_mesh < 0 || N <= 0 || n_stamp < 0 || !v_off || !t_off || n_nodes <= 0 || n_nodes > N)
            PDN_FAIL(PDN_EINVAL, "assemble_part: bad sizes");
        if (n_stamp > 0 && (!s_row || !s_col || !s_val || !s_op)) PDN_FAIL(PDN_EINVAL, "assemble_part: stamp arrays missing");
        ctx->L.valid = false;
        ctx->geo_valid = false;
        pdn_assemble_impl(ctx, n_mesh, v_off, xy, t_off, tri, conductance, N, n_stamp, s_row, s_col, s_val,
                          s_op, on_device != 0, nnz_out, n_nodes);
    });
}

int pdn_get_shape(pdn_ctx *ctx, int64_t *N, int64_t *nnz) {
    if (!ctx) return PDN_EINVAL;
    if (!ctx->L.valid) return PDN_ENOSYSTEM;
    if (N) *N = ctx->L.n;
    if (nnz) *nnz = ctx->L.nnz;
    return PDN_OK;
}

int pdn_get_csr(pdn_ctx *ctx, int32_t *indptr, int32_t *indices, double *data, int on_device) {
    return pdn_guard(ctx, [&] {
        if (!ctx->L.valid) PDN_FAIL(PDN_ENOSYSTEM, "get_csr: no resident system");
        cudaMemcpyKind k = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
        const CsrDev &L = ctx->L;
        if (indptr) PDN_CUDA(cudaMemcpyAsync(indptr, L.indptr.p, (L.n + 1) * sizeof(int32_t), k, ctx->stream));
        if (indices && L.nnz) PDN_CUDA(cudaMemcpyAsync(indices, L.indices.p, L.nnz * sizeof(int32_t), k, ctx->stream));
        if (data && L.nnz) PDN_CUDA(cudaMemcpyAsync(data, L.data.p, L.nnz * sizeof(double), k, ctx->stream));
        PDN_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

int pdn_set_csr(pdn_ctx *ctx, int64_t N, int64_t nnz, const int32_t *indptr, const int32_t *indices,
                const double *data, int on_device) {
    return pdn_guard(ctx, [&] {
        if (N <= 0 || nnz < 0 || !indptr || (nnz > 0 && (!indices || !data)) || nnz > 0x7fffffffLL)
            PDN_FAIL(PDN_EINVAL, "set_csr: bad arguments");
        cudaMemcpyKind k = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
        CsrDev &L = ctx->L;
        L.valid = false;
        ctx->geo_valid = false;   // a foreign matrix carries no coordinates: Jacobi only
        L.indptr.reserve((size_t)N + 1);
        L.indices.reserve((size_t)nnz + 1);
        L.data.reserve((size_t)nnz + 1);
        PDN_CUDA(cudaMemcpyAsync(L.indptr.p, indptr, (N + 1) * sizeof(int32_t), k, ctx->stream));
        if (nnz) {
            PDN_CUDA(cudaMemcpyAsync(L.indices.p, indices, nnz * sizeof(int32_t), k, ctx->stream));
            PDN_CUDA(cudaMemcpyAsync(L.data.p, data, nnz * sizeof(double), k, ctx->stream));
        }
        PDN_CUDA(cudaStreamSynchronize(ctx->stream));
        L.n = N;
        L.nnz = nnz;
        L.valid = true;
    });
}

int pdn_solve(pdn_ctx *ctx, const double *rhs, const pdn_opts *opts, double *v, pdn_info *info, int on_device) {
    return pdn_guard(ctx, [&] {
        if (!ctx->L.valid) PDN_FAIL(PDN_ENOSYSTEM, "solve: no resident system");
        if (!rhs || !v) PDN_FAIL(PDN_EINVAL, "solve: rhs / v missing");
        pdn_solve_impl(ctx, rhs, opts, v, info, on_device != 0);
    });
}

int pdn_spmv(pdn_ctx *ctx, const double *x, double *y, int on_device) {
    return pdn_guard(ctx, [&] {
        if (!ctx->L.valid) PDN_FAIL(PDN_ENOSYSTEM, "spmv: no resident system");
        if (!x || !y) PDN_FAIL(PDN_EINVAL, "spmv: x / y missing");
        pdn_spmv_impl(ctx, x, y, on_device != 0);
    });
}

int pdn_bench_spmv(pdn_ctx *ctx, int32_t reps, int flush_l2, double *ms_per_launch) {
    return pdn_guard(ctx, [&] {
        if (!ctx->L.valid) PDN_FAIL(PDN_ENOSYSTEM, "bench_spmv: no resident system");
        if (reps <= 0 || !ms_per_launch) PDN_FAIL(PDN_EINVAL, "bench_spmv: bad arguments");
        pdn_bench_spmv_impl(ctx, reps, flush_l2, ms_per_launch);
    });
}

int pdn_power_density(pdn_ctx *ctx, int64_t n_v, const double *xy, const double *val, int64_t n_t,
                      const int32_t *tri, double conductivity, double *p, int on_device) {
    return pdn_guard(ctx, [&] {
        if (n_v < 0 || n_t < 0 || (n_t > 0 && (!xy || !val || !tri || !p)))
            PDN_FAIL(PDN_EINVAL, "power_density: bad arguments");
        pdn_power_density_impl(ctx, n_v, xy, val, n_t, tri, conductivity, p, on_device != 0);
    });
}

int pdn_nearest_vertex(pdn_ctx *ctx, int32_t n_mesh, const int64_t *v_off, const int32_t *mesh_layer, const double *xy,
                       int64_t n_q, const double *q_xy, const int32_t *q_layer, int64_t *out_index, double *out_d2,
                       int on_device) {
    return pdn_guard(ctx, [&] {
        if (n_mesh < 0 || n_q < 0 || !v_off || (n_mesh > 0 && !mesh_layer) || (n_q > 0 && (!q_xy || !q_layer || !out_index)))
            PDN_FAIL(PDN_EINVAL, "nearest_vertex: bad arguments");
        if (!on_device) {
            for (int32_t m = 0; m < n_mesh; m++)
                if (v_off[m + 1] < v_off[m]) PDN_FAIL(PDN_EINVAL, "nearest_vertex: v_off must be non-decreasing");
            if (n_mesh > 0 && v_off[n_mesh] > v_off[0] && !xy) PDN_FAIL(PDN_EINVAL, "nearest_vertex: xy is null");
        }
        pdn_nearest_vertex_impl(ctx, n_mesh, v_off, mesh_layer, xy, n_q, q_xy, q_layer, out_index, out_d2, on_device != 0);
    });
}

}  // extern "C"
