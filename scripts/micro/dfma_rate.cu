// DFMA issue rate and latency on one SM-full of warps: ILP chains per thread x warps per SM (scripts/micro, diagnostics only)
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void k(double *out, int iters, double a, double b) {
    double acc[ILP];
#pragma unroll
    for (int c = 0; c < ILP; c++) acc[c] = threadIdx.x + c;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < ILP; c++) acc[c] = fma(acc[c], a, b);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < ILP; c++) s += acc[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0)
        printf("ILP %2d threads %4d: %.2f cycles per warp-DFMA per SMSP, %.1f DFMA lanes/clk/SM\n", ILP, blockDim.x,
               (double)(t1 - t0) / ((double)iters * ILP * (blockDim.x / 128.0)), 32.0 * iters * ILP * (blockDim.x / 32.0) / (double)(t1 - t0));
}
int main() {
    double *out; cudaMalloc(&out, 148 * 1024 * 8);
    for (int th : {128, 512, 1024}) {
        k<1><<<148, th>>>(out, 4096, 1.0000001, 1e-9); cudaDeviceSynchronize();
        k<3><<<148, th>>>(out, 4096, 1.0000001, 1e-9); cudaDeviceSynchronize();
        k<8><<<148, th>>>(out, 4096, 1.0000001, 1e-9); cudaDeviceSynchronize();
        k<16><<<148, th>>>(out, 4096, 1.0000001, 1e-9); cudaDeviceSynchronize();
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
