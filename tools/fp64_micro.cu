// FP64 latency / single-warp throughput probe for sm_100a (informs the shape of the Riccati sweep kernel).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp64_micro tools/fp64_micro.cu && build/fp64_micro
#include <cstdio>
#include <cuda_runtime.h>
template <int CHAINS>
__global__ void chain_kernel(double *out, long long *cyc, int iters, double m, double c)
{
    double a[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) a[k] = threadIdx.x + k;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) a[k] = fma(a[k], m, c);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += a[k];
    if (s == 1234.5) out[0] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
// same, with only the first `active` lanes of the warp executing: does a partly filled warp issue FP64 faster?
template <int CHAINS>
__global__ void partial_kernel(double *out, long long *cyc, int iters, double m, double c, int active)
{
    double a[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) a[k] = threadIdx.x + k;
    long long t0 = clock64();
    if ((threadIdx.x & 31) < active) {
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < CHAINS; ++k) a[k] = fma(a[k], m, c);
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += a[k];
    if (s == 1234.5) out[0] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void sincos_kernel(double *out, long long *cyc, int iters, double x0)
{
    double x = x0 + threadIdx.x * 1e-3, s, c;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) { sincos(x, &s, &c); x = x + 1e-3 * s * c; }
    long long t1 = clock64();
    if (x == 1234.5) out[0] = x;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void div_kernel(double *out, long long *cyc, int iters, double x0)
{
    double x = x0 + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) x = 1.0 / (x + 1.5);
    long long t1 = clock64();
    if (x == 1234.5) out[0] = x;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main()
{
    double *o; long long *c, h[4];
    cudaMalloc(&o, 8); cudaMalloc(&c, 64);
    const int it = 4096;
#define RUN(K, W) chain_kernel<K><<<1, 32 * W>>>(o, c, it, 0.999, 1e-9); cudaMemcpy(h, c, 8, cudaMemcpyDeviceToHost); \
    printf("DFMA chains=%d warps=%d: %.2f cycles per DFMA-per-chain step, %.2f cycles per warp-instruction\n", K, W, (double)h[0] / it, (double)h[0] / it / K);
    RUN(1, 1) RUN(2, 1) RUN(4, 1) RUN(8, 1) RUN(16, 1) RUN(1, 4) RUN(8, 4) RUN(8, 8) RUN(8, 16)
    for (int act : {32, 16, 8, 1}) {
        partial_kernel<8><<<1, 32>>>(o, c, it, 0.999, 1e-9, act); cudaMemcpy(h, c, 8, cudaMemcpyDeviceToHost);
        printf("DFMA chains=8 warps=1 active lanes=%d: %.2f cycles per warp-instruction\n", act, (double)h[0] / it / 8);
    }
    sincos_kernel<<<1, 32>>>(o, c, 1024, 0.3); cudaMemcpy(h, c, 8, cudaMemcpyDeviceToHost);
    printf("dependent sincos: %.1f cycles each\n", (double)h[0] / 1024);
    div_kernel<<<1, 32>>>(o, c, 1024, 0.3); cudaMemcpy(h, c, 8, cudaMemcpyDeviceToHost);
    printf("dependent 1/x (+add): %.1f cycles each\n", (double)h[0] / 1024);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
