// Phase timers of the per-anchor tridiagonalisation kernel (tridiag_fused_small_kernel, -DTF_PROFILE):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTF_PROFILE -Iiddiff_b200/csrc tools/tridiag_small_profile.cu \
//        iddiff_b200/csrc/idb_tridiag.cu -o tools/tridiag_small_profile;   tools/tridiag_small_profile [n] [anchors]
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "idb_kernels.h"
namespace idb {
#ifdef TF_PROFILE
extern __device__ double tf_prof[8];
#endif
size_t tridiag_workspace_bytes(int n_anchors, int n);
}
__global__ void fill_spd(double* G, int n, int A) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)A * n * n) return;
    const int a = idx / ((size_t)n * n), i = (idx / n) % n, j = idx % n;
    const unsigned lo = min(i, j), hi = max(i, j);
    unsigned h = lo * 2654435761u ^ (hi * 40503u + 0x9E3779B9u) ^ (a * 97u);
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13; h *= 3266489917u; h ^= h >> 16;
    G[idx] = (double)h / 4294967296.0 - 0.5 + (i == j ? 0.1 * n : 0.0);
}
int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 200, A = argc > 2 ? atoi(argv[2]) : 148;
    double* G; float* sv; void* ws;
    cudaMalloc(&G, (size_t)A * n * n * 8); cudaMalloc(&sv, (size_t)A * n * 4);
    cudaMalloc(&ws, idb::tridiag_workspace_bytes(A, n));
    fill_spd<<<(unsigned)(((size_t)A * n * n + 255) / 256), 256>>>(G, n, A);
    for (int rep = 0; rep < 3; ++rep) {
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        cudaEventRecord(a);
        cudaError_t e = idb::launch_spectrum_tridiag(G, A, n, n, sv, nullptr, ws, 0);
        cudaEventRecord(b);
        cudaError_t e2 = cudaDeviceSynchronize();
        float ms = 0; cudaEventElapsedTime(&ms, a, b);
        printf("n %d anchors %d rep %d: tridiag + bisect %.3f ms (%s / %s)\n", n, A, rep, ms, cudaGetErrorString(e), cudaGetErrorString(e2));
    }
#ifdef TF_PROFILE
    double acc[8];
    cudaMemcpyFromSymbol(acc, idb::tf_prof, sizeof(acc));
    const char* names[8] = {"(1) column + reflector", "(2) wait at the barrier after the pass", "(3) p, w", "(2b0) batch set-up (row sides, pointers)", "(2a) column-side registers", "(2b) batch: loads + fma", "(2c) batch: row reduction", "(2d) column partials out"};
    double tot = 0; for (int i = 0; i < 8; ++i) tot += acc[i];
    for (int i = 0; i < 8; ++i) printf("  %-40s %6.1f %%  %9.0f cycles / column\n", names[i], 100 * acc[i] / tot, acc[i] / n);
#endif
    return 0;
}
