// Phase timers of the panel tridiagonalisation kernel (build with -DTP_PROFILE, see tools/panel_profile.sh):
//   panel_profile n  ->  per-phase share of CTA 0's cycles + the kernel's wall time
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
namespace idb {
size_t tridiag_panel_workspace_doubles(int n, int grid);
cudaError_t launch_tridiag_panel(double* G, int n, double* d, double* e, double* ws, int num_sms, cudaStream_t stream);
}
__global__ void fill_sym(double* A, int n) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)n * n) return;
    const int i = idx / n, j = idx % n;
    const unsigned lo = min(i, j), hi = max(i, j);
    unsigned h = lo * 2654435761u ^ (hi * 40503u + 0x9E3779B9u);
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13; h *= 3266489917u; h ^= h >> 16;
    A[idx] = (double)h / 4294967296.0 - 0.5 + (i == j ? 4.0 : 0.0);
}
int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 4096;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double *A, *d, *e, *ws;
    const size_t wsn = idb::tridiag_panel_workspace_doubles(n, sms);
    cudaMalloc(&A, (size_t)n * n * 8); cudaMalloc(&d, n * 8); cudaMalloc(&e, n * 8); cudaMalloc(&ws, wsn * 8);
    const char* names[7] = {"phase A", "sync 1", "phase B", "read sweeps", "reduce", "sync 2", "flush sweeps"};
    for (int rep = 0; rep < 2; ++rep) {
        fill_sym<<<(unsigned)(((size_t)n * n + 255) / 256), 256>>>(A, n);
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        cudaEventRecord(a);
        cudaError_t err = idb::launch_tridiag_panel(A, n, d, e, ws, sms, 0);
        cudaEventRecord(b);
        cudaError_t err2 = cudaDeviceSynchronize();
        float ms = 0; cudaEventElapsedTime(&ms, a, b);
        printf("n %d  rep %d  %.2f ms  (%s / %s)\n", n, rep, ms, cudaGetErrorString(err), cudaGetErrorString(err2));
    }
    std::vector<double> acc(7);
    const size_t off = (size_t)(2 * 8 + 2) * n + (size_t)sms * 17;
    cudaMemcpy(acc.data(), ws + off, 56, cudaMemcpyDeviceToHost);
    double tot = 0; for (double v : acc) tot += v;
    for (int i = 0; i < 7; ++i) printf("  %-12s %6.1f %%  %8.2f us / column\n", names[i], 100 * acc[i] / tot, acc[i] / 1.9e3 / n);
    return 0;
}
