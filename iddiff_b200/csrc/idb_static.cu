// An entry point that allocates NOTHING: the whole-GPU eigen-solver of idb_eig_large.cu with its workspace in a file-scope
// device array instead of a caller-supplied buffer.  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  It always takes the one-class fused tridiagonalisation kernel (the panel kernel's per-CTA partial-sum buffers
// are megabytes).
//
// The array costs IDB_STATIC_SCRATCH_BYTES (768 KB) on every GPU that loads the library, until the process exits; that
// bounds the matrix at n <= IDB_SPECTRUM_STATIC_MAX_N (4096).  Every call on a device shares the one array: calls that
// overlap on the GPU (different streams, or different host threads) race on it -- serialise them.  The solver writes every
// workspace word before it reads it (d, e, the reflector / partial-product vectors and the per-CTA partial sums are all
// produced inside each launch), so nothing has to be cleared between calls and the previous call's contents are harmless.
//
// In host code the array's NAME is only a host-side stand-in for the symbol, not a device address: the device address is
// looked up once per device with cudaGetSymbolAddress (that device current) and passed to the launcher.
#include "idb_kernels.h"
#include <mutex>

namespace {
constexpr int STATIC_MAX_N = 4096;
constexpr size_t STATIC_DOUBLES = (size_t)24 * STATIC_MAX_N;      // 768 KB >= the fused kernel's workspace at n = 4096 (256 KB)
constexpr int STATIC_MAX_DEVICES = 64;

static __device__ __align__(256) double g_spectrum_scratch[STATIC_DOUBLES];

std::mutex g_mu;
void* g_scratch_addr[STATIC_MAX_DEVICES] = {};

// device address of the array on the CURRENT device (looked up on first use there, cached afterwards)
cudaError_t scratch_address(void** out) {
    int dev = -1;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return err;
    if (dev < 0 || dev >= STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        err = cudaGetSymbolAddress(&p, g_spectrum_scratch);
        if (err != cudaSuccess) return err;
        if (!p) return cudaErrorInvalidSymbol;
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return cudaSuccess;
}
}  // namespace

extern "C" {

int idb_spectrum_large_static_max_n(void) { return STATIC_MAX_N; }

int idb_spectrum_large_static(double* G, int n, int n_out, float* sv, void* stream) {
    if (!G || !sv || n <= 0 || n_out <= 0 || n_out > n) return idb::api_fail("idb_spectrum_large_static: bad argument");
    if (n > STATIC_MAX_N || idb::spectrum_large_fused_workspace_bytes(n) > sizeof(double) * STATIC_DOUBLES)
        return idb::api_fail("idb_spectrum_large_static: n exceeds IDB_SPECTRUM_STATIC_MAX_N = 4096 (the built-in 768 KB "
                             "scratch array); use idb_spectrum_large with a caller-supplied workspace");
    void* ws = nullptr;
    cudaError_t err = scratch_address(&ws);
    if (err != cudaSuccess) return idb::api_check(err, "idb_spectrum_large_static: cudaGetSymbolAddress");
    return idb::api_check(idb::launch_spectrum_large(G, n, n_out, sv, ws, idb::api_num_sms(), static_cast<cudaStream_t>(stream), 1),
                          "idb_spectrum_large_static");
}

}  // extern "C"
