// the rank-16 update loop of coarse_invert2_kernel in isolation (one CTA per SM): where do 1200 cycles per row go?
#include <cstdio>
#include <cuda_runtime.h>
constexpr int BLK = 512, NC = 1458, R = 16;
// V 0: as in the kernel   1: no cs loads (registers)   2: no rows read-modify-write   3: two rows per iteration
// V 4: cs read as 16 LDS.64 instead of 8 LDS.128   5: rows RMW kept, cs via __shfl from a per-lane register copy
template <int V, int NCOL>
__global__ void __launch_bounds__(BLK, 1) k(double *out, int reps) {
    extern __shared__ double rows[];
    __shared__ __align__(16) double cs[16][16];
    const int tid = threadIdx.x;
    for (int e = tid; e < R * NC; e += BLK) rows[e] = 1.0 + e * 1e-9;
    if (tid < 256) cs[tid >> 4][tid & 15] = 1e-3 * (tid + 1);
    __syncthreads();
    double p[NCOL][16];
    for (int t = 0; t < NCOL; t++)
        for (int q = 0; q < 16; q++) p[t][q] = 1.0 + 1e-6 * (tid + q + t);
    const double mylane = cs[0][tid & 15];
    long long t0 = clock64();
    for (int rep = 0; rep < reps; rep++) {
        if (V == 3) {
            for (int i = 0; i < R; i += 2) {
                double acc[2][NCOL];
#pragma unroll
                for (int t = 0; t < NCOL; t++) {
                    const int j = tid + t * BLK;
                    acc[0][t] = j < NC ? rows[i * NC + j] : 0.0;
                    acc[1][t] = j < NC ? rows[(i + 1) * NC + j] : 0.0;
                }
#pragma unroll
                for (int q = 0; q < 16; q += 2) {
                    const double2 c0 = *reinterpret_cast<const double2 *>(&cs[i][q]);
                    const double2 c1 = *reinterpret_cast<const double2 *>(&cs[i + 1][q]);
#pragma unroll
                    for (int t = 0; t < NCOL; t++) {
                        acc[0][t] = fma(c0.x, p[t][q], acc[0][t]);
                        acc[1][t] = fma(c1.x, p[t][q], acc[1][t]);
                        acc[0][t] = fma(c0.y, p[t][q + 1], acc[0][t]);
                        acc[1][t] = fma(c1.y, p[t][q + 1], acc[1][t]);
                    }
                }
#pragma unroll
                for (int t = 0; t < NCOL; t++) {
                    const int j = tid + t * BLK;
                    if (j < NC) { rows[i * NC + j] = acc[0][t]; rows[(i + 1) * NC + j] = acc[1][t]; }
                }
            }
        } else {
            for (int i = 0; i < R; i++) {
                double acc[NCOL];
#pragma unroll
                for (int t = 0; t < NCOL; t++) {
                    const int j = tid + t * BLK;
                    acc[t] = (V == 2 || j >= NC) ? 1e-3 * i : rows[i * NC + j];
                }
#pragma unroll
                for (int q = 0; q < 16; q += 2) {
                    double2 c;
                    if (V == 1) c = make_double2(1e-3 * (i + q), 1e-3 * (i - q));
                    else if (V == 4) { c.x = cs[i][q]; c.y = cs[i][q + 1]; }
                    else if (V == 5) { c.x = __shfl_sync(0xffffffffu, mylane, q) + i; c.y = __shfl_sync(0xffffffffu, mylane, q + 1) + i; }
                    else c = *reinterpret_cast<const double2 *>(&cs[i][q]);
#pragma unroll
                    for (int t = 0; t < NCOL; t++) {
                        acc[t] = fma(c.x, p[t][q], acc[t]);
                        acc[t] = fma(c.y, p[t][q + 1], acc[t]);
                    }
                }
#pragma unroll
                for (int t = 0; t < NCOL; t++) {
                    const int j = tid + t * BLK;
                    if (j < NC) { if (V == 2) p[t][i] += acc[t] * 1e-30; else rows[i * NC + j] = acc[t]; }
                }
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
    for (int t = 0; t < NCOL; t++) for (int q = 0; q < 16; q++) s += p[t][q];
    __syncthreads();
    out[blockIdx.x * BLK + tid] = s + rows[tid];
    if (tid == 0 && blockIdx.x == 0) printf("V%d NCOL %d: %.0f cycles per row (16 warps, %d columns)\n", V, NCOL, (double)(t1 - t0) / (reps * R), NC);
}
template <int V, int NCOL> void run(double *out) {
    cudaFuncSetAttribute(k<V, NCOL>, cudaFuncAttributeMaxDynamicSharedMemorySize, R * NC * 8);
    k<V, NCOL><<<148, BLK, R * NC * 8>>>(out, 50);
    cudaDeviceSynchronize();
}
int main() {
    double *out; cudaMalloc(&out, 148 * BLK * 8);
    run<0, 3>(out); run<1, 3>(out); run<2, 3>(out); run<3, 3>(out); run<4, 3>(out); run<5, 3>(out);
    run<0, 2>(out); run<3, 2>(out);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
