// pdn_noalloc.cu -- pdn_residual_noalloc: r = b - A x and ||r||^2 with NO device allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It is a variant of the residual form of the CSR SpMV (spmv_kernel<RESID> in pdn_spmv.cuh: y = aux - A x with
// rr = y.y through grid_reduce, whose per-CTA partial sums and arrival ticket live in context-owned, cudaMalloc'd
// scratch).  Here that scratch is a file-scope `static __device__` array of a fixed size, so the entry point
// allocates nothing -- at the price of a hard row limit and of every call on a GPU sharing one array:
//
//     [0, 256)        header: word 0 = arrival ticket of the running launch (cleared before every launch)
//     [256, ...)      one fp64 partial sum of r_i^2 per CTA of NA_ROWS rows, NA_MAX_BLOCKS of them
//
// In host code `pdn_na_scratch` is only a host-side stand-in for the symbol: the DEVICE address is fetched once
// per device with cudaGetSymbolAddress (that device current) and handed to the kernel as an ordinary pointer.
//
// Arithmetic: one thread per row, entries summed in stored (column) order, every product and sum rounded
// separately (__dmul_rn / __dadd_rn: no FMA contraction), r_i = b_i - sum -- the same bits as pdn_spmv's
// y followed by b - y.  The last CTA to arrive adds the partials in CTA order, so rr is reproducible run to run.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "padne_b200.h"

namespace {

constexpr int NA_ROWS = 1024;                                    // rows (= threads) per CTA
constexpr int64_t NA_MAX_ROWS = PDN_NOALLOC_MAX_ROWS;            // 2^24
constexpr int NA_MAX_BLOCKS = (int)(NA_MAX_ROWS / NA_ROWS);      // 16384 partials
constexpr size_t NA_HEADER = 256;
constexpr size_t NA_SCRATCH_BYTES = NA_HEADER + (size_t)NA_MAX_BLOCKS * sizeof(double);   // 131 328 B per GPU
static_assert(NA_MAX_ROWS % NA_ROWS == 0, "row limit must be a whole number of CTAs");
constexpr int NA_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char pdn_na_scratch[NA_SCRATCH_BYTES];

__global__ void __launch_bounds__(NA_ROWS) residual_noalloc_kernel(int64_t n, const int32_t *__restrict__ rp,
                                                                   const int32_t *__restrict__ ci,
                                                                   const double *__restrict__ va,
                                                                   const double *__restrict__ x,
                                                                   const double *__restrict__ b, double *__restrict__ r,
                                                                   double *__restrict__ rr, unsigned char *scratch) {
    __shared__ double red[NA_ROWS / 32];
    __shared__ bool is_last;
    unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch);
    double *partials = reinterpret_cast<double *>(scratch + NA_HEADER);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t row = (int64_t)blockIdx.x * NA_ROWS + tid;
    double q = 0.0;
    if (row < n) {
        const int32_t k0 = __ldg(&rp[row]), k1 = __ldg(&rp[row + 1]);
        double sum = 0.0;
        for (int32_t k = k0; k < k1; k++) sum = __dadd_rn(sum, __dmul_rn(__ldg(&va[k]), __ldg(&x[__ldg(&ci[k])])));
        const double ri = __dadd_rn(__ldg(&b[row]), -sum);
        r[row] = ri;
        q = __dmul_rn(ri, ri);
    }
    // CTA sum in a fixed order: lanes by butterfly, warps in index order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_down_sync(0xffffffffu, q, o);
    if (lane == 0) red[warp] = q;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < NA_ROWS / 32; w++) s += red[w];
        partials[blockIdx.x] = s;
        __threadfence();
        is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // last CTA: the partials in CTA order (thread t takes t, t + 1024, ...; then the same fixed tree)
    double v = 0.0;
    for (int i = tid; i < (int)gridDim.x; i += NA_ROWS) v += __ldcg(&partials[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp] = v;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < NA_ROWS / 32; w++) s += red[w];
        *rr = s;
        *ticket = 0u;
    }
}

std::mutex na_mu;
unsigned char *na_ptr[NA_MAX_DEVICES];     // device address of pdn_na_scratch on each device (null: not set up yet)

}  // namespace

// restores the caller's current device on every way out
struct NaDeviceGuard {
    int prev = -1;
    bool switched = false;
    ~NaDeviceGuard() {
        if (switched) cudaSetDevice(prev);
    }
};

#define NA_FAIL(code_, ...)                                                                  \
    do {                                                                                     \
        if (err != nullptr && err_cap > 0) snprintf(err, (size_t)err_cap, __VA_ARGS__);      \
        return (code_);                                                                      \
    } while (0)
#define NA_CUDA(expr)                                                                        \
    do {                                                                                     \
        cudaError_t e_ = (expr);                                                             \
        if (e_ != cudaSuccess)                                                               \
            NA_FAIL(PDN_ECUDA, "pdn_residual_noalloc: %s failed: %s", #expr, cudaGetErrorString(e_)); \
    } while (0)

extern "C" int pdn_residual_noalloc(int device, uint64_t stream, int64_t N, const int32_t *indptr, const int32_t *indices,
                                    const double *data, const double *x, const double *b, double *r, double *rr,
                                    char *err, int64_t err_cap) {
    if (err != nullptr && err_cap > 0) err[0] = 0;
    if (N < 0) NA_FAIL(PDN_EINVAL, "pdn_residual_noalloc: N = %lld is negative", (long long)N);
    if (N > NA_MAX_ROWS)
        NA_FAIL(PDN_EUNSUPPORTED,
                "pdn_residual_noalloc: N = %lld rows exceeds PDN_NOALLOC_MAX_ROWS = %lld (the scratch is a fixed "
                "%zu-byte device array, one partial sum per %d rows); use pdn_set_csr + pdn_spmv for larger systems",
                (long long)N, (long long)NA_MAX_ROWS, NA_SCRATCH_BYTES, NA_ROWS);
    if (rr == nullptr) NA_FAIL(PDN_EINVAL, "pdn_residual_noalloc: rr is NULL");
    if (N > 0 && (!indptr || !indices || !data || !x || !b || !r))
        NA_FAIL(PDN_EINVAL, "pdn_residual_noalloc: NULL array argument");
    if (device < 0 || device >= NA_MAX_DEVICES)
        NA_FAIL(PDN_EINVAL, "pdn_residual_noalloc: device %d outside [0, %d)", device, NA_MAX_DEVICES);
    int n_dev = 0;
    NA_CUDA(cudaGetDeviceCount(&n_dev));
    if (device >= n_dev) NA_FAIL(PDN_EINVAL, "pdn_residual_noalloc: device %d, but %d visible", device, n_dev);

    NaDeviceGuard guard;
    NA_CUDA(cudaGetDevice(&guard.prev));
    if (guard.prev != device) {
        NA_CUDA(cudaSetDevice(device));
        guard.switched = true;
    }
    // set-up, once per device, with that device current: the device address of the scratch array
    unsigned char *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(na_mu);
        if (na_ptr[device] == nullptr) {
            void *p = nullptr;
            NA_CUDA(cudaGetSymbolAddress(&p, pdn_na_scratch));
            if (p == nullptr) NA_FAIL(PDN_ECUDA, "pdn_residual_noalloc: cudaGetSymbolAddress returned NULL");
            na_ptr[device] = static_cast<unsigned char *>(p);
        }
        scratch = na_ptr[device];
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (N == 0) {                                    // empty system: ||r||^2 = 0 (all-zero bits are 0.0)
        NA_CUDA(cudaMemsetAsync(rr, 0, sizeof(double), st));
        return PDN_OK;
    }
    // the ticket must start at zero whatever an earlier (possibly aborted) call left behind; the partials need no
    // clearing: every CTA of this launch writes its own before the last one reads them
    NA_CUDA(cudaMemsetAsync(scratch, 0, NA_HEADER, st));
    const int grid = (int)((N + NA_ROWS - 1) / NA_ROWS);
    residual_noalloc_kernel<<<grid, NA_ROWS, 0, st>>>(N, indptr, indices, data, x, b, r, rr, scratch);
    NA_CUDA(cudaGetLastError());
    return PDN_OK;
}
