// fp64 FMA issue rate and dependent-issue latency on this GPU:  dfma_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int CHAINS>
__global__ void k(double* out, double a, double b, int iters) {
    double acc[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc[c] = threadIdx.x + c;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) acc[c] = fma(acc[c], a, b);
    double s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += acc[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int CHAINS>
void run(int threads, int blocks) {
    double* out; cudaMalloc(&out, (size_t)threads * blocks * 8);
    const int iters = 20000;
    k<CHAINS><<<blocks, threads>>>(out, 1.0000001, 1e-9, 100);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<CHAINS><<<blocks, threads>>>(out, 1.0000001, 1e-9, iters);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fma = (double)threads * blocks * iters * CHAINS;
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("chains %d threads/SM %4d: %.2f TFMA/s, %.1f lanes/clk/SM at %.2f GHz nominal, %.1f cycles per dependent step per warp\n", CHAINS, threads,
           fma / ms / 1e9, fma / ms / 1e-3 / blocks / (clk * 1e3), clk / 1e6, (double)ms * 1e-3 * clk * 1e3 / iters);
    cudaFree(out);
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    run<1>(32, sms); run<1>(128, sms); run<1>(512, sms); run<4>(512, sms); run<8>(512, sms); run<8>(1024, sms);
    return 0;
}
