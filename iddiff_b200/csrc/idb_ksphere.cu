// Analytic k-sphere score, the device counterpart of models/ksphere_gt.py:14-26 (reference):
//   score = (x / r) * (ive(p/2, k) / ive(p/2 - 1, k) - r) / sigma^2,  r = |x|,  k = r / sigma^2.
// The reference evaluates this per row in NumPy with float32 scalars (NumPy-2 promotion: every
// intermediate is rounded to float32, the two Bessel values included); the roundings are
// replicated here, the Bessel functions themselves are evaluated in fp64.
#include "idb_common.cuh"
#include "idb_kernels.h"
#include <math.h>

namespace idb {

// exponentially scaled modified Bessel function of the first kind, exp(-x) I_nu(x), x > 0, nu >= 0
__device__ double ive_f64(double nu, double x) {
    const double mu = 4.0 * nu * nu;
    if (x >= 30.0 && x >= 2.0 * nu * nu + 30.0) {
        // Hankel asymptotic expansion: sum_k (-1)^k a_k(nu) / x^k, a_k = prod_{j<=k} (mu - (2j-1)^2) / (k! 8^k)
        double term = 1.0, sum = 1.0, prev = 1.0;
        for (int k = 1; k < 200; ++k) {
            const double o = 2.0 * k - 1.0;
            term *= -(mu - o * o) / (8.0 * k * x);
            if (fabs(term) > prev) break;  // asymptotic series started to diverge
            sum += term;
            prev = fabs(term);
            if (prev < 1e-18 * fabs(sum)) break;
        }
        return sum * rsqrt(2.0 * 3.14159265358979323846 * x);
    }
    // ascending series (x/2)^nu sum_k (x^2/4)^k / (k! Gamma(nu+k+1)), rescaled to stay in range
    const double q = 0.25 * x * x;
    double term = 1.0, sum = 1.0, logscale = 0.0;
    for (int k = 1; k < 100000; ++k) {
        term *= q / ((double)k * (nu + (double)k));
        sum += term;
        if (term < 1e-18 * sum) break;
        if (sum > 1e200) {
            sum *= 1e-200;
            term *= 1e-200;
            logscale += 200.0 * 2.302585092994045684;
        }
    }
    return exp(nu * log(0.5 * x) - lgamma(nu + 1.0) - x + logscale + log(sum));
}

// one warp per row
__global__ void __launch_bounds__(256) ksphere_score_kernel(const float* __restrict__ x,
                                                            const float* __restrict__ sigma_rows, float sigma,
                                                            int rows, int D, int manifold_dim,
                                                            float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int row = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5);
    if (row >= rows) return;
    const float* xr = x + (size_t)row * D;
    double acc = 0.0;
    for (int c = lane; c < D; c += 32) {
        const double v = (double)xr[c];
        acc = fma(v, v, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    const float r = (float)sqrt(acc);
    const float sg = sigma_rows ? sigma_rows[row] : sigma;
    const float s2 = __fmul_rn(sg, sg);
    const float kf = __fdiv_rn(r, s2);
    float coef = 0.f;
    if (lane == 0) {
        const double nu = 0.5 * (double)manifold_dim;
        const float i_hi = (float)ive_f64(nu, (double)kf);
        const float i_lo = (float)ive_f64(nu - 1.0, (double)kf);
        const float ratio = __fdiv_rn(i_hi, i_lo);  // = -vm in the reference
        coef = __fsub_rn(ratio, r);
    }
    coef = __shfl_sync(0xffffffffu, coef, 0);
    float* orow = out + (size_t)row * D;
    for (int c = lane; c < D; c += 32) {
        const float u = __fdiv_rn(xr[c], r);
        orow[c] = __fdiv_rn(__fmul_rn(u, coef), s2);
    }
}

cudaError_t launch_ksphere_score(const float* x, const float* sigma_or_null, float sigma, int rows, int D,
                                 int manifold_dim, float* out, cudaStream_t stream) {
    if (rows <= 0) return cudaSuccess;
    const int warps_per_block = 8;
    ksphere_score_kernel<<<(rows + warps_per_block - 1) / warps_per_block, 256, 0, stream>>>(
        x, sigma_or_null, sigma, rows, D, manifold_dim, out);
    return cudaGetLastError();
}

}  // namespace idb
