// Spectrum stage: centre -> Gram (fp64) -> Cholesky (fp64) -> one-sided Jacobi on the triangular
// factor (fp32, shared memory, one anchor per CTA, warp-shuffle rotations) -> singular values ->
// spectral-gap read-out.  Replaces dim_reduction.py:190-199 (column centring + full LAPACK SVD
// on the host) and plot_utils.py:224-229 (gap rule) of the reference.
//
// Why not an fp32 Gram + eigensolver: the spectra this path exists for have a cliff of 1e3 or more
// between the tangent and normal singular values; squaring that in fp32 loses the tail.  The Gram
// is therefore accumulated in fp64 (B200 has full-rate fp64 FMA; the stage is < 5 % of a step),
// its Cholesky factor L (G = L L^T, so sv(L) = sv(centred scores)) is taken in fp64, and only
// then rounded to fp32, where a one-sided Jacobi keeps high RELATIVE accuracy on graded columns.
#include "idb_common.cuh"
#include "idb_kernels.h"
#include <math.h>

namespace idb {

// ---------------------------------- column means --------------------------------------------
__global__ void __launch_bounds__(256) colmean_kernel(const float* __restrict__ X, long long anchor_stride, int ld,
                                                      int K, int D, double divisor, double* __restrict__ colmean) {
    __shared__ double s_part[8][33];
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    const int rg = threadIdx.x >> 5;
    const float* Xa = X + (size_t)blockIdx.y * anchor_stride;
    double acc = 0.0;
    if (c < D)
        for (int k = rg; k < K; k += 8) acc += (double)Xa[(size_t)k * ld + c];
    s_part[rg][threadIdx.x & 31] = acc;
    __syncthreads();
    if (rg == 0 && c < D) {
        double t = 0.0;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += s_part[i][threadIdx.x];
        colmean[(size_t)blockIdx.y * D + c] = t / divisor;
    }
}

// ---------------------------------- centred Gram (fp64) -------------------------------------
// One CTA = one 64 x 64 tile of G for one anchor (upper-triangular tile pairs only; mirrored on store).
// The contraction runs on the fp64 tensor-core path (mma.sync m8n8k4 f64, SASS DMMA): 8 warps, each a 16 x 32 block of the
// tile = 2 x 4 MMA tiles, operands converted fp32 -> centred fp64 while they are staged in shared memory.
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

constexpr int CG_LD = 72;  // padded row (doubles): the 4 k-rows of a fragment load fall into disjoint bank halves

__global__ void __launch_bounds__(256) center_gram_kernel(const float* __restrict__ X, long long anchor_stride, int ld,
                                                          int K, int D, double scale,
                                                          const double* __restrict__ colmean, double* __restrict__ G) {
    __shared__ double sA[16][CG_LD];
    __shared__ double sB[16][CG_LD];
    // decode (ti <= tj) from the linear pair index
    const int T = (D + 63) >> 6;
    int ti = 0, rem = blockIdx.x;
    while (rem >= T - ti) { rem -= T - ti; ++ti; }
    const int tj = ti + rem;
    const int a = blockIdx.y;
    const float* Xa = X + (size_t)a * anchor_stride;
    const double* mu = colmean + (size_t)a * D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i0 = (warp >> 1) * 16, j0 = (warp & 1) * 32;
    const int fr = lane >> 2, fk = lane & 3;   // fragment row / k index of this lane
    double acc[2][4][2];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int n = 0; n < 4; ++n) acc[m][n][0] = acc[m][n][1] = 0.0;

    // staging: thread -> (k row kk = t / 16, 4 consecutive columns c4 = (t % 16) * 4) of both operand tiles
    const int skk = threadIdx.x >> 4, sc = (threadIdx.x & 15) << 2;
    double mua[4], mub[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int ca = ti * 64 + sc + q, cb = tj * 64 + sc + q;
        mua[q] = ca < D ? mu[ca] : 0.0;
        mub[q] = cb < D ? mu[cb] : 0.0;
    }
    for (int k0 = 0; k0 < K; k0 += 16) {
        const int k = k0 + skk;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int ca = ti * 64 + sc + q, cb = tj * 64 + sc + q;
            double va = 0.0, vb = 0.0;
            if (k < K) {
                if (ca < D) va = ((double)__ldg(Xa + (size_t)k * ld + ca) - mua[q]) * scale;
                if (cb < D) vb = ((double)__ldg(Xa + (size_t)k * ld + cb) - mub[q]) * scale;
            }
            sA[skk][sc + q] = va;
            sB[skk][sc + q] = vb;
        }
        __syncthreads();
#pragma unroll
        for (int k4 = 0; k4 < 16; k4 += 4) {
            double af[2], bf[4];
#pragma unroll
            for (int m = 0; m < 2; ++m) af[m] = sA[k4 + fk][i0 + m * 8 + fr];   // A[i][k] = X[k][i]
#pragma unroll
            for (int n = 0; n < 4; ++n) bf[n] = sB[k4 + fk][j0 + n * 8 + fr];   // B[k][j] = X[k][j]
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int n = 0; n < 4; ++n) dmma884(acc[m][n][0], acc[m][n][1], af[m], bf[n]);
        }
        __syncthreads();
    }
    double* Ga = G + (size_t)a * D * D;
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        const int r = ti * 64 + i0 + m * 8 + fr;
        if (r >= D) continue;
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int c = tj * 64 + j0 + n * 8 + 2 * fk + q;
                if (c >= D) continue;
                Ga[(size_t)r * D + c] = acc[m][n][q];
                if (ti != tj) Ga[(size_t)c * D + r] = acc[m][n][q];
            }
    }
}

// 40-wide tiles (5 warps, each one 8-row strip of 5 MMA tiles) for column counts that pad badly to 64: D = 200 is
// exactly 5 x 40, so the upper triangle costs 15 x 40^2 = 24 000 outputs instead of 10 x 64^2 = 40 960 -- the stage is
// bound by the fp64 tensor pipe, so that ratio is its time.
constexpr int CG40 = 40;
constexpr int CG40_LD = 52;   // row stride = 4 (mod 16) doubles: the 4 k-rows of a half-warp's fragment load hit disjoint banks

__global__ void __launch_bounds__(160) center_gram40_kernel(const float* __restrict__ X, long long anchor_stride, int ld,
                                                            int K, int D, double scale,
                                                            const double* __restrict__ colmean, double* __restrict__ G) {
    __shared__ __align__(16) double sA[16][CG40_LD];
    __shared__ __align__(16) double sB[16][CG40_LD];
    const int T = (D + CG40 - 1) / CG40;
    int ti = 0, rem = blockIdx.x;
    while (rem >= T - ti) { rem -= T - ti; ++ti; }
    const int tj = ti + rem;
    const int a = blockIdx.y;
    const float* Xa = X + (size_t)a * anchor_stride;
    const double* mu = colmean + (size_t)a * D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i0 = warp * 8;
    const int fr = lane >> 2, fk = lane & 3;
    double acc[5][2];
#pragma unroll
    for (int n = 0; n < 5; ++n) acc[n][0] = acc[n][1] = 0.0;
    // staging: 20 threads per k-row, two columns each (one conflict-free 16-byte shared store per operand), 8 rows per pass,
    // two passes per 16-row slab.  ncu: the kernel is bound by shared-memory wavefronts (LSU data pipe 96 % busy), and the
    // former mapping (4 scalar stores per thread, 32 bytes apart across lanes) spent 41 % of them on bank-conflicted stores.
    const int skk = threadIdx.x / 20, sc = (threadIdx.x % 20) << 1;
    const int ca = ti * CG40 + sc, cb = tj * CG40 + sc;
    const double mua0 = ca < D ? mu[ca] : 0.0, mua1 = ca + 1 < D ? mu[ca + 1] : 0.0;
    const double mub0 = cb < D ? mu[cb] : 0.0, mub1 = cb + 1 < D ? mu[cb + 1] : 0.0;
    const bool vec2 = ((ld & 1) == 0) && ((reinterpret_cast<uintptr_t>(Xa) & 7) == 0) && ((D & 1) == 0);
    // the next 16-row slab is fetched into registers while the tensor cores work on the current one
    float2 fa[2], fb[2];
    auto fetch = [&](int k0) {
#pragma unroll
        for (int ps = 0; ps < 2; ++ps) {
            const int k = k0 + ps * 8 + skk;
            fa[ps] = make_float2(0.f, 0.f);
            fb[ps] = make_float2(0.f, 0.f);
            if (k < K) {
                const float* row = Xa + (size_t)k * ld;
                if (vec2) {
                    if (ca < D) fa[ps] = __ldg(reinterpret_cast<const float2*>(row + ca));
                    if (cb < D) fb[ps] = __ldg(reinterpret_cast<const float2*>(row + cb));
                } else {
                    if (ca < D) fa[ps].x = __ldg(row + ca);
                    if (ca + 1 < D) fa[ps].y = __ldg(row + ca + 1);
                    if (cb < D) fb[ps].x = __ldg(row + cb);
                    if (cb + 1 < D) fb[ps].y = __ldg(row + cb + 1);
                }
            }
        }
    };
    fetch(0);
    for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
        for (int ps = 0; ps < 2; ++ps) {
            const int kr = ps * 8 + skk;
            const bool kin = k0 + kr < K;
            double2 va, vb;
            va.x = (kin && ca < D) ? ((double)fa[ps].x - mua0) * scale : 0.0;
            va.y = (kin && ca + 1 < D) ? ((double)fa[ps].y - mua1) * scale : 0.0;
            vb.x = (kin && cb < D) ? ((double)fb[ps].x - mub0) * scale : 0.0;
            vb.y = (kin && cb + 1 < D) ? ((double)fb[ps].y - mub1) * scale : 0.0;
            *reinterpret_cast<double2*>(&sA[kr][sc]) = va;
            *reinterpret_cast<double2*>(&sB[kr][sc]) = vb;
        }
        __syncthreads();
        if (k0 + 16 < K) fetch(k0 + 16);
#pragma unroll
        for (int k4 = 0; k4 < 16; k4 += 4) {
            const double af = sA[k4 + fk][i0 + fr];
            double bf[5];
#pragma unroll
            for (int n = 0; n < 5; ++n) bf[n] = sB[k4 + fk][n * 8 + fr];
#pragma unroll
            for (int n = 0; n < 5; ++n) dmma884(acc[n][0], acc[n][1], af, bf[n]);
        }
        __syncthreads();
    }
    double* Ga = G + (size_t)a * D * D;
    const int r = ti * CG40 + i0 + fr;
    if (r >= D) return;
#pragma unroll
    for (int n = 0; n < 5; ++n)
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int c = tj * CG40 + n * 8 + 2 * fk + q;
            if (c >= D) continue;
            Ga[(size_t)r * D + c] = acc[n][q];
            if (ti != tj) Ga[(size_t)c * D + r] = acc[n][q];
        }
}

cudaError_t launch_column_sums(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D,
                               double divisor, double* out, cudaStream_t stream) {
    if (n_anchors <= 0 || D <= 0) return cudaSuccess;
    colmean_kernel<<<dim3((D + 31) / 32, n_anchors), 256, 0, stream>>>(X, anchor_stride, ld, K, D, divisor, out);
    return cudaGetLastError();
}

cudaError_t launch_center_gram(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D,
                               double scale, double* colmean, double* G, cudaStream_t stream) {
    cudaError_t e = launch_column_sums(X, anchor_stride, ld, n_anchors, K, D, (double)K, colmean, stream);
    if (e != cudaSuccess) return e;
    return launch_gram_about(X, anchor_stride, ld, n_anchors, K, D, scale, colmean, G, stream);
}

// Gram of the rows about a GIVEN centre vector (the K-sharded mode passes the all-reduced global column mean)
cudaError_t launch_gram_about(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D,
                              double scale, const double* colmean, double* G, cudaStream_t stream) {
    if (n_anchors <= 0 || D <= 0) return cudaSuccess;
    const int T = (D + 63) / 64, T40 = (D + CG40 - 1) / CG40;
    const long long cost64 = (long long)T * (T + 1) / 2 * 64 * 64, cost40 = (long long)T40 * (T40 + 1) / 2 * CG40 * CG40;
    // padded outputs computed; the narrow tiles have a worse shared-load : MMA ratio (6 : 5 against 6 : 8), so they only pay
    // where they save >= 40 % of the outputs (measured: D = 200 0.82 -> 0.71 ms per 128 anchors, D = 100 0.31 -> 0.33)
    if (cost40 * 100 < cost64 * 60) {
        center_gram40_kernel<<<dim3(T40 * (T40 + 1) / 2, n_anchors), 160, 0, stream>>>(X, anchor_stride, ld, K, D, scale,
                                                                                      colmean, G);
    } else {
        center_gram_kernel<<<dim3(T * (T + 1) / 2, n_anchors), 256, 0, stream>>>(X, anchor_stride, ld, K, D, scale,
                                                                                 colmean, G);
    }
    return cudaGetLastError();
}

// ---------------------------------- Cholesky + one-sided Jacobi -----------------------------
constexpr int SPEC_THREADS = 512;
constexpr int SPEC_WARPS = SPEC_THREADS / 32;
constexpr size_t SPEC_SMEM_MAX = 224 * 1024;

__host__ __device__ inline size_t spec_ldb(int n) { return (size_t)((n + 1) & ~1) | 1; }  // odd stride >= n (+pad col room)
__host__ __device__ inline size_t spec_packed_bytes(int n) { return (size_t)n * (n + 1) / 2 * sizeof(double); }
__host__ __device__ inline size_t spec_square_bytes(int n) { return spec_ldb(n) * (size_t)((n + 1) & ~1) * sizeof(float); }

size_t spectrum_workspace_bytes(int n_anchors, int n) {
    // per anchor: packed fp64 factor (only used when it does not fit shared memory) + fp32 square factor
    return (size_t)n_anchors * (spec_packed_bytes(n) + spec_square_bytes(n)) + 256;
}

// Tournament (circle-method) ordering: ne players (even), round r in [0, ne-1), slot i in [0, ne/2).
__device__ __forceinline__ void rr_pair(int ne, int r, int i, int& p, int& q) {
    const int m = ne - 1;
    if (i == 0) {
        p = m;
        q = r;
    } else {
        p = (r + i) % m;
        q = (r - i + m) % m;
    }
    if (p > q) { const int t = p; p = q; q = t; }
}

// L: packed lower triangle (row-major: L[i][j] at i(i+1)/2 + j), in shared or global memory.
// B: fp32 column-major [ne columns][ldb], in shared or global memory.
__global__ void __launch_bounds__(SPEC_THREADS, 1) spectrum_kernel(const double* __restrict__ G, int n, int n_out,
                                                                   float* __restrict__ sv, int* __restrict__ sweeps,
                                                                   uint8_t* __restrict__ workspace, int l_in_smem,
                                                                   int b_in_smem) {
    extern __shared__ __align__(16) uint8_t spec_smem[];
    __shared__ double s_red[SPEC_WARPS];
    __shared__ double s_pivot;
    __shared__ int s_flag;
    const int a = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t per_anchor = spec_packed_bytes(n) + spec_square_bytes(n);
    uint8_t* ws = workspace + (size_t)a * per_anchor;
    double* L = l_in_smem ? reinterpret_cast<double*>(spec_smem) : reinterpret_cast<double*>(ws);
    float* Bg = reinterpret_cast<float*>(ws + spec_packed_bytes(n));
    const double* Ga = G + (size_t)a * n * n;
    const int ne = (n + 1) & ~1;
    const size_t ldb = spec_ldb(n);

    // ---- load the lower triangle, find max diagonal
    double dmax = 0.0;
    for (int idx = tid; idx < n * n; idx += SPEC_THREADS) {
        const int i = idx / n, j = idx - i * n;
        if (j <= i) {
            const double v = Ga[idx];
            L[(size_t)i * (i + 1) / 2 + j] = v;
            if (i == j) dmax = fmax(dmax, v);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    if (lane == 0) s_red[warp] = dmax;
    __syncthreads();
    dmax = 0.0;
    for (int w = 0; w < SPEC_WARPS; ++w) dmax = fmax(dmax, s_red[w]);
    const double thresh = dmax * (double)n * 2.3e-16;

    // ---- right-looking Cholesky, positive semi-definite safe (a vanished pivot zeroes its column)
    for (int j = 0; j < n; ++j) {
        if (tid == 0) {
            const double pv = L[(size_t)j * (j + 1) / 2 + j];
            s_pivot = (pv > thresh && pv > 0.0) ? sqrt(pv) : 0.0;
        }
        __syncthreads();
        const double d = s_pivot;
        const double inv = d > 0.0 ? 1.0 / d : 0.0;
        for (int i = j + tid; i < n; i += SPEC_THREADS) {
            double* e = &L[(size_t)i * (i + 1) / 2 + j];
            *e = (i == j) ? d : (*e) * inv;
        }
        __syncthreads();
        if (d > 0.0) {
            for (int i = j + 1 + warp; i < n; i += SPEC_WARPS) {
                double* Li = &L[(size_t)i * (i + 1) / 2];
                const double lij = Li[j];
                for (int k = j + 1 + lane; k <= i; k += 32) Li[k] = fma(-lij, L[(size_t)k * (k + 1) / 2 + j], Li[k]);
            }
        }
        __syncthreads();
    }

    // ---- B = L rounded to fp32, column-major with a zero pad column when n is odd
    for (size_t idx = tid; idx < (size_t)ne * ldb; idx += SPEC_THREADS) {
        const int j = (int)(idx / ldb), i = (int)(idx - (size_t)j * ldb);
        float v = 0.f;
        if (j < n && i < n && i >= j) v = (float)L[(size_t)i * (i + 1) / 2 + j];
        Bg[idx] = v;
    }
    __syncthreads();
    float* B = Bg;
    if (b_in_smem) {
        B = reinterpret_cast<float*>(spec_smem);
        for (size_t idx = tid; idx < (size_t)ne * ldb; idx += SPEC_THREADS) B[idx] = Bg[idx];
        __syncthreads();
    }

    // ---- one-sided Jacobi sweeps: orthogonalise the columns of B, one column pair per warp
    const float tol = 2e-6f;
    int sweep = 0;
    for (; sweep < 40; ++sweep) {
        int rotated = 0;
        for (int r = 0; r < ne - 1; ++r) {
            for (int i = warp; i < ne / 2; i += SPEC_WARPS) {
                int p, q;
                rr_pair(ne, r, i, p, q);
                float* bp = B + (size_t)p * ldb;
                float* bq = B + (size_t)q * ldb;
                float al = 0.f, be = 0.f, ga = 0.f;
                // columns of a lower-triangular start are zero above the diagonal only before the first
                // rotation touches them, so always run the full length
                for (int k = lane; k < n; k += 32) {
                    const float x = bp[k], y = bq[k];
                    al = fmaf(x, x, al);
                    be = fmaf(y, y, be);
                    ga = fmaf(x, y, ga);
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    al += __shfl_xor_sync(0xffffffffu, al, o);
                    be += __shfl_xor_sync(0xffffffffu, be, o);
                    ga += __shfl_xor_sync(0xffffffffu, ga, o);
                }
                if (fabsf(ga) > tol * sqrtf(al * be) && al > 0.f && be > 0.f) {
                    rotated = 1;
                    const float zeta = (be - al) / (2.f * ga);
                    const float tt = copysignf(1.f, zeta) / (fabsf(zeta) + sqrtf(fmaf(zeta, zeta, 1.f)));
                    const float cs = rsqrtf(fmaf(tt, tt, 1.f));
                    const float sn = cs * tt;
                    for (int k = lane; k < n; k += 32) {
                        const float x = bp[k], y = bq[k];
                        bp[k] = fmaf(cs, x, -sn * y);
                        bq[k] = fmaf(sn, x, cs * y);
                    }
                }
            }
            __syncthreads();
        }
        if (__syncthreads_or(rotated) == 0) break;
    }

    // ---- singular values = column norms (fp64 accumulate), sorted descending by rank
    double* nrm = reinterpret_cast<double*>(ws);  // reuse the packed-L area of the workspace (n doubles)
    __syncthreads();
    for (int j = warp; j < n; j += SPEC_WARPS) {
        const float* bj = B + (size_t)j * ldb;
        double s = 0.0;
        for (int k = lane; k < n; k += 32) s = fma((double)bj[k], (double)bj[k], s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) nrm[j] = sqrt(s);
    }
    __threadfence_block();
    __syncthreads();
    for (int j = tid; j < n; j += SPEC_THREADS) {
        const double v = nrm[j];
        int rank = 0;
        for (int i = 0; i < n; ++i) {
            const double u = nrm[i];
            rank += (u > v) || (u == v && i < j);
        }
        if (rank < n_out) sv[(size_t)a * n_out + rank] = (float)v;
    }
    if (tid == 0 && sweeps) sweeps[a] = sweep;
    (void)s_flag;
}

cudaError_t launch_spectrum(const double* G, int n_anchors, int n, int n_out, float* sv, int* sweeps, void* workspace,
                            cudaStream_t stream) {
    if (n_anchors <= 0 || n <= 0) return cudaSuccess;
    const int l_in_smem = spec_packed_bytes(n) <= SPEC_SMEM_MAX;
    const int b_in_smem = spec_square_bytes(n) <= SPEC_SMEM_MAX;
    size_t smem = 16;
    if (l_in_smem) smem = spec_packed_bytes(n);
    if (b_in_smem && spec_square_bytes(n) > smem) smem = spec_square_bytes(n);
    cudaError_t e = cudaFuncSetAttribute(spectrum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    spectrum_kernel<<<n_anchors, SPEC_THREADS, smem, stream>>>(G, n, n_out, sv, sweeps,
                                                               reinterpret_cast<uint8_t*>(workspace), l_in_smem,
                                                               b_in_smem);
    return cudaGetLastError();
}

// ---------------------------------- spectral-gap read-out -----------------------------------
// plot_utils.py:224-229 in the reference's own arithmetic: Python floats (fp64) on the fp32 singular
// values widened by s.tolist(); softmax as in plot_utils.py:211-214 (x - max, exp, / sum), first-occurrence
// argmax with NumPy's NaN rule (the first NaN wins).  dims = -1 flags the reference's ZeroDivisionError.
// One warp per anchor (lanes stride over the gaps; the three passes -- max, sum, first-occurrence argmax -- are warp
// reductions).
__global__ void __launch_bounds__(128) id_readout_kernel(const float* __restrict__ sv, int A, int n, int ld,
                                                         int* __restrict__ dims) {
    const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (a >= A) return;
    const float* s = sv + (size_t)a * ld;
    const double nf = (double)s[1] - (double)s[2];
    if (nf == 0.0) {
        if (lane == 0) dims[a] = -1;
        return;
    }
    const int m = n - 2;  // len(diff); diff[q] = (s[q+1] - s[q+2]) / nf, q = 0 .. m-1
    const unsigned full = 0xffffffffu;
    double mx = -INFINITY;
    bool has_nan = false;
    for (int q = lane; q < m; q += 32) {
        const double d = ((double)s[q + 1] - (double)s[q + 2]) / nf;
        if (d != d) has_nan = true;
        else if (d > mx) mx = d;
    }
    if (__any_sync(full, has_nan)) {
        // np.max propagates NaN -> every soft entry is NaN -> argmax = 0
        if (lane == 0) dims[a] = m;
        return;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(full, mx, o));
    double sum = 0.0;
    for (int q = lane; q < m; q += 32) sum += exp(((double)s[q + 1] - (double)s[q + 2]) / nf - mx);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(full, sum, o);
    int best = 0x7fffffff, first_nan = 0x7fffffff;
    double bv = -1.0;
    for (int q = lane; q < m; q += 32) {
        const double v = exp(((double)s[q + 1] - (double)s[q + 2]) / nf - mx) / sum;
        if (v != v) first_nan = min(first_nan, q);
        else if (v > bv) { bv = v; best = q; }       // strict: the lane keeps its first occurrence
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(full, bv, o);
        const int ob = __shfl_xor_sync(full, best, o);
        const int on = __shfl_xor_sync(full, first_nan, o);
        if (ov > bv || (ov == bv && ob < best)) { bv = ov; best = ob; }
        first_nan = min(first_nan, on);
    }
    if (best == 0x7fffffff) best = 0;
    if (lane == 0) dims[a] = m - (first_nan != 0x7fffffff ? first_nan : best);
}

cudaError_t launch_id_readout(const float* sv, int n_anchors, int n, int ld, int* dims, cudaStream_t stream) {
    if (n_anchors <= 0) return cudaSuccess;
    id_readout_kernel<<<(n_anchors + 3) / 4, 128, 0, stream>>>(sv, n_anchors, n, ld, dims);
    return cudaGetLastError();
}

}  // namespace idb
