// the rank-16 update C[16 x NC] += A[16 x 16] B[16 x NC] with FP64 tensor-core MMAs (m8n8k4), C in shared memory,
// B prefetched into registers (48 doubles per thread), A fragments in registers: cycles per 16-row update.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int BLK = 512, NC = 1458, R = 16, WARPS = BLK / 32, TPW = ((NC + 7) / 8 + WARPS - 1) / WARPS;
__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(BLK, 1) k(const double *__restrict__ P, double *out, int reps, int nc) {
    extern __shared__ double rows[];
    __shared__ __align__(16) double cs[16][16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int e = tid; e < R * NC; e += BLK) rows[e] = 1.0 + e * 1e-9;
    if (tid < 256) cs[tid >> 4][tid & 15] = 1e-3 * (tid + 1);
    __syncthreads();
    const int n_tiles = (nc + 7) / 8;
    long long t0 = clock64();
    for (int rep = 0; rep < reps; rep++) {
        double b[TPW][4];
#pragma unroll
        for (int t = 0; t < TPW; t++) {
            const int n = warp + WARPS * t, col = 8 * n + (lane >> 2);
#pragma unroll
            for (int ks = 0; ks < 4; ks++) b[t][ks] = (n < n_tiles && col < nc) ? __ldcg(&P[(4 * ks + (lane & 3)) * nc + col]) : 0.0;
        }
        double a[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
            for (int ks = 0; ks < 4; ks++) a[mt][ks] = cs[8 * mt + (lane >> 2)][4 * ks + (lane & 3)];
#pragma unroll
        for (int t = 0; t < TPW; t++) {
            const int n = warp + WARPS * t;
            if (n >= n_tiles) break;
            const int col = 8 * n + 2 * (lane & 3);
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
                double *cp = &rows[(8 * mt + (lane >> 2)) * nc + col];
                double c0 = col < nc ? cp[0] : 0.0, c1 = col + 1 < nc ? cp[1] : 0.0;
#pragma unroll
                for (int ks = 0; ks < 4; ks++) dmma(c0, c1, a[mt][ks], b[t][ks]);
                if (col < nc) cp[0] = c0;
                if (col + 1 < nc) cp[1] = c1;
            }
        }
        __syncthreads();
    }
    long long t1 = clock64();
    out[blockIdx.x * BLK + tid] = rows[tid];
    if (tid == 0 && blockIdx.x == 0) printf("DMMA m8n8k4: %.0f cycles per 16-row update of %d columns (incl. B loads from L2)\n", (double)(t1 - t0) / reps, nc);
}
int main() {
    double *out, *P; cudaMalloc(&out, 148 * BLK * 8); cudaMalloc(&P, 16 * NC * 8 + 1024); cudaMemset(P, 0, 16 * NC * 8 + 1024);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, R * NC * 8);
    k<<<148, BLK, R * NC * 8>>>(P, out, 50, NC); cudaDeviceSynchronize();
    k<<<92, BLK, R * NC * 8>>>(P, out, 50, NC); cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
