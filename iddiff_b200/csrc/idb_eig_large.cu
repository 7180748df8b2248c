// Spectrum stage for IMAGE-sized score matrices (BASELINE configs 4-5: K = 4096 ... 16768 rows, D = 12288 columns,
// so n = min(K, D) is 4096 ... 12288 and the Gram matrix no longer fits one SM):
//
//   row_gram_kernel        G = (s (X - mean))(s (X - mean))^T, K x K, fp64 -- the row side, used when K < D
//                          (north_star: "S.S^T or S^T.S, whichever side is smaller")
//   tridiag_coop_kernel    Householder tridiagonalisation of ONE n x n fp64 matrix by the whole GPU: a cooperative
//                          (grid-synchronised) kernel, trailing block streamed from L2/HBM by every SM
//   bisect_large_kernel    Sturm multisection, one warp per eigenvalue (32 probes per round), d / e^2 in shared memory
//
// Replaces torch.linalg.svd at dim_reduction.py:197-198 (LAPACK gesdd of the K x D matrix on the host: 16 s at
// 4096 x 12288).  Roofline: memory-bound -- per eliminated column the trailing m x m block is read once
// for the symmetric matrix-vector product and read + written once for the rank-2 update: 24 m^2 bytes, 8 n^3 in total.
#include "idb_common.cuh"
#include "idb_kernels.h"
#include <cooperative_groups.h>
#include <math.h>

namespace cg = cooperative_groups;

namespace idb {

// ---------------------------------- row-side centred Gram -----------------------------------
// One CTA = one 64 x 64 tile (ti <= tj) of G; contraction over the D columns in 16-wide slabs through shared memory on
// the fp64 tensor-core path (mma.sync m8n8k4 f64, SASS DMMA): 8 warps x (2 x 4) MMA tiles.  Rows are staged row-major
// ([64][20] doubles: float4 global loads along the contiguous column axis, conflict-free fragment reads).
__device__ __forceinline__ void rg_dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
constexpr int RG_LD = 20;

__global__ void __launch_bounds__(256) row_gram_kernel(const float* __restrict__ X, int ld, int K, int D, double scale,
                                                       const double* __restrict__ colmean, double* __restrict__ G) {
    __shared__ __align__(16) double sA[64][RG_LD];
    __shared__ __align__(16) double sB[64][RG_LD];
    const int T = (K + 63) >> 6;
    int ti = 0, rem = blockIdx.x;
    while (rem >= T - ti) { rem -= T - ti; ++ti; }
    const int tj = ti + rem;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i0 = (warp >> 1) * 16, j0 = (warp & 1) * 32;
    const int fr = lane >> 2, fk = lane & 3;
    double acc[2][4][2];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int n = 0; n < 4; ++n) acc[m][n][0] = acc[m][n][1] = 0.0;
    const int sr = threadIdx.x >> 2, sq = (threadIdx.x & 3) << 2;   // staging: row sr, columns sq .. sq+3 of the slab
    const int ra = ti * 64 + sr, rb = tj * 64 + sr;
    const bool vec = (ld & 3) == 0 && (D & 3) == 0 && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    for (int d0 = 0; d0 < D; d0 += 16) {
        const int d = d0 + sq;
        double va[4] = {0.0, 0.0, 0.0, 0.0}, vb[4] = {0.0, 0.0, 0.0, 0.0};
        if (vec && d < D) {
            const double m0 = colmean[d], m1 = colmean[d + 1], m2 = colmean[d + 2], m3 = colmean[d + 3];
            if (ra < K) {
                const float4 f = __ldg(reinterpret_cast<const float4*>(X + (size_t)ra * ld + d));
                va[0] = ((double)f.x - m0) * scale; va[1] = ((double)f.y - m1) * scale;
                va[2] = ((double)f.z - m2) * scale; va[3] = ((double)f.w - m3) * scale;
            }
            if (rb < K) {
                const float4 f = __ldg(reinterpret_cast<const float4*>(X + (size_t)rb * ld + d));
                vb[0] = ((double)f.x - m0) * scale; vb[1] = ((double)f.y - m1) * scale;
                vb[2] = ((double)f.z - m2) * scale; vb[3] = ((double)f.w - m3) * scale;
            }
        } else if (!vec) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (d + q < D) {
                    const double mu = colmean[d + q];
                    if (ra < K) va[q] = ((double)__ldg(X + (size_t)ra * ld + d + q) - mu) * scale;
                    if (rb < K) vb[q] = ((double)__ldg(X + (size_t)rb * ld + d + q) - mu) * scale;
                }
            }
        }
        *reinterpret_cast<double2*>(&sA[sr][sq]) = make_double2(va[0], va[1]);
        *reinterpret_cast<double2*>(&sA[sr][sq + 2]) = make_double2(va[2], va[3]);
        *reinterpret_cast<double2*>(&sB[sr][sq]) = make_double2(vb[0], vb[1]);
        *reinterpret_cast<double2*>(&sB[sr][sq + 2]) = make_double2(vb[2], vb[3]);
        __syncthreads();
#pragma unroll
        for (int k4 = 0; k4 < 16; k4 += 4) {
            double af[2], bf[4];
#pragma unroll
            for (int m = 0; m < 2; ++m) af[m] = sA[i0 + m * 8 + fr][k4 + fk];   // A[i][k] = row i, column k
#pragma unroll
            for (int n = 0; n < 4; ++n) bf[n] = sB[j0 + n * 8 + fr][k4 + fk];   // B[k][j] = row j, column k
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int n = 0; n < 4; ++n) rg_dmma884(acc[m][n][0], acc[m][n][1], af[m], bf[n]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        const int r = ti * 64 + i0 + m * 8 + fr;
        if (r >= K) continue;
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int c = tj * 64 + j0 + n * 8 + 2 * fk + q;
                if (c >= K) continue;
                G[(size_t)r * K + c] = acc[m][n][q];
                if (ti != tj) G[(size_t)c * K + r] = acc[m][n][q];
            }
    }
}

// column means of X[K][D] (fp64), one thread per column, rows strided over blockIdx.y slices with atomics
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ X, int ld, int K, int D, int rows_per,
                                                     double* __restrict__ colsum) {
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c >= D) return;
    const int k0 = blockIdx.y * rows_per, k1 = min(K, k0 + rows_per);
    double a = 0.0;
    for (int k = k0; k < k1; ++k) a += (double)__ldg(X + (size_t)k * ld + c);
    atomicAdd(colsum + c, a);
}
__global__ void scale_vec_kernel(double* v, int n, double s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] *= s;
}

cudaError_t launch_row_gram(const float* X, int ld, int K, int D, double scale, double* colmean, double* G,
                            cudaStream_t stream) {
    if (K <= 0 || D <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(colmean, 0, (size_t)D * sizeof(double), stream);
    if (e != cudaSuccess) return e;
    const int slices = K >= 2048 ? 16 : 1;
    const int rows_per = (K + slices - 1) / slices;
    colsum_kernel<<<dim3((D + 255) / 256, slices), 256, 0, stream>>>(X, ld, K, D, rows_per, colmean);
    scale_vec_kernel<<<(D + 255) / 256, 256, 0, stream>>>(colmean, D, 1.0 / (double)K);
    const int T = (K + 63) / 64;
    row_gram_kernel<<<T * (T + 1) / 2, 256, 0, stream>>>(X, ld, K, D, scale, colmean, G);
    return cudaGetLastError();
}

// ---------------------------------- cooperative tridiagonalisation --------------------------
constexpr int TC_THREADS = 512;
constexpr int TC_WARPS = TC_THREADS / 32;

__device__ __forceinline__ double tc_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double tc_block_sum(double v, double* red) {
    v = tc_warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < TC_WARPS; ++w) t += red[w];
    return t;
}

// A: full symmetric n x n, row-major, overwritten.  d[n], e[n] out (e[n-1] = 0).  p: n doubles of scratch.
// LAPACK dsytd2 / dlarfg arithmetic: H = I - tau v v^T, p = tau A v, w = p - (tau/2)(p.v) v, A <- A - v w^T - w v^T.
__global__ void __launch_bounds__(TC_THREADS, 1) tridiag_coop_kernel(double* __restrict__ A, int n,
                                                                     double* __restrict__ d, double* __restrict__ e,
                                                                     double* __restrict__ p) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) double tc_smem[];
    double* v = tc_smem;        // [n] Householder vector, indexed by absolute row (valid for rows > k)
    double* w = tc_smem + n;    // [n]
    __shared__ double red[TC_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gw = blockIdx.x * TC_WARPS + warp, n_gw = gridDim.x * TC_WARPS;

    for (int k = 0; k + 2 < n; ++k) {
        // (a) every CTA builds the reflector of column k from row k (symmetric copy, contiguous)
        const double* rowk = A + (size_t)k * n;
        double ss = 0.0;
        for (int j = k + 2 + tid; j < n; j += TC_THREADS) {
            const double x = rowk[j];
            v[j] = x;
            ss = fma(x, x, ss);
        }
        ss = tc_block_sum(ss, red);           // ||x[1:]||^2
        const double alpha = rowk[k + 1];
        double tau = 0.0, beta = alpha;
        if (ss > 0.0) {
            const double nrm = sqrt(fma(alpha, alpha, ss));
            beta = alpha >= 0.0 ? -nrm : nrm;
            tau = (beta - alpha) / beta;
            const double inv = 1.0 / (alpha - beta);
            for (int j = k + 2 + tid; j < n; j += TC_THREADS) v[j] *= inv;
        }
        if (tid == 0) v[k + 1] = 1.0;
        if (blockIdx.x == 0 && tid == 0) {
            d[k] = rowk[k];
            e[k] = beta;
        }
        __syncthreads();
        if (tau != 0.0) {   // uniform over the grid (same data, same arithmetic)
            // (b) p = tau * A[k+1:, k+1:] v : one warp per row, rows dealt cyclically over the grid's warps
            for (int i = k + 1 + gw; i < n; i += n_gw) {
                const double* row = A + (size_t)i * n;
                double s = 0.0;
                for (int j = k + 1 + lane; j < n; j += 32) s = fma(row[j], v[j], s);
                s = tc_warp_sum(s);
                if (lane == 0) p[i] = tau * s;
            }
            grid.sync();
            // (c) w = p - (tau/2)(p.v) v  (redundantly per CTA), then the rank-2 update of the CTA's rows
            double pv = 0.0;
            for (int j = k + 1 + tid; j < n; j += TC_THREADS) {
                const double pj = p[j];
                w[j] = pj;
                pv = fma(pj, v[j], pv);
            }
            pv = tc_block_sum(pv, red);
            const double kk = -0.5 * tau * pv;
            for (int j = k + 1 + tid; j < n; j += TC_THREADS) w[j] = fma(kk, v[j], w[j]);
            __syncthreads();
            for (int i = k + 1 + gw; i < n; i += n_gw) {
                double* row = A + (size_t)i * n;
                const double vi = v[i], wi = w[i];
                for (int j = k + 1 + lane; j < n; j += 32) row[j] = row[j] - vi * w[j] - wi * v[j];
            }
            grid.sync();
        }
    }
    if (blockIdx.x == 0 && tid == 0) {
        if (n >= 2) {
            d[n - 2] = A[(size_t)(n - 2) * n + (n - 2)];
            e[n - 2] = A[(size_t)(n - 1) * n + (n - 2)];
        }
        d[n - 1] = A[(size_t)(n - 1) * n + (n - 1)];
        e[n - 1] = 0.0;
    }
}

// Fused variant: the rank-2 update of step k-1 and the symmetric matrix-vector product of step k are ONE pass over the
// trailing block -- every row is owned by a fixed warp (row i -> warp i mod #warps), which reads it, applies the
// pending pair (v, w) of the previous step, writes it back and dots it with the new reflector u in the same sweep.
// 16 m^2 bytes and one grid synchronisation per column instead of 24 m^2 and two.  Row k itself gets the pending pair
// applied on the fly when its reflector is formed.  (v, w, u) must sit in shared memory for the sweep: three
// n-vectors fit one SM up to n = TCF_MAX_N; above that
constexpr int TCF_MAX_N = 9000;
constexpr int PANEL_MIN_N = 3584;   // the automatic choice takes the panel kernel (variant 3, idb_eig_panel.cu) from here on; measured
                                    // fused / panel: n = 3072 59 / 63 ms, 4096 106 / 92, 6144 298 / 181, 8192 645 / 302 (tridiag. only),
                                    // 12288 1916 / 778 ms
// the 64-wide column chunks of the trailing block are dealt round-robin over C classes of CTAs, a CTA of class c keeps
// only ITS chunks of (v, w, u) in shared memory (3 n / C doubles) and produces the partial products p_c[i]; the full
// reflector of every step lives in a double-buffered global copy (ubuf, written by CTA 0, L2-resident) and w is
// re-derived where a row index needs it: w[i] = kk u[i] + sum_c p_c[i].  Same 16 m^2 bytes and one grid
// synchronisation per column as the C = 1 kernel; n = 12288 (K = 16768 rows of a 64x64x3 image) runs with C = 2.
constexpr int TCF_UNROLL = 8;

template <int C>
__global__ void __launch_bounds__(TC_THREADS, 1) tridiag_fused_cls_kernel(double* __restrict__ A, int n,
                                                                          double* __restrict__ d, double* __restrict__ e,
                                                                          double* __restrict__ pbuf,
                                                                          double* __restrict__ ubuf) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) double tc_smem[];
    const int Ls = ((((n + 63) >> 6) + C - 1) / C) << 6;
    double* v = tc_smem;
    double* w = tc_smem + Ls;
    double* u = tc_smem + 2 * Ls;
    __shared__ double red[TC_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cls = blockIdx.x % C, rg = blockIdx.x / C, n_rg = gridDim.x / C;
    const int gw = rg * TC_WARPS + warp, n_gw = n_rg * TC_WARPS;
    const bool lead = blockIdx.x == 0;
    for (int j = tid; j < 3 * Ls; j += TC_THREADS) tc_smem[j] = 0.0;
    __syncthreads();
    bool pend = false;
    double kk_prev = 0.0;
    auto own = [&](int j) { return ((j >> 6) % C) == cls; };
    auto loc = [&](int j) { return (((j >> 6) / C) << 6) | (j & 63); };
    auto psum = [&](const double* p, int j) {
        double t = p[j];
#pragma unroll
        for (int c = 1; c < C; ++c) t += p[(size_t)c * n + j];
        return t;
    };

    for (int k = 0; k + 2 < n; ++k) {
        const double* pp = pbuf + (size_t)((k + 1) & 1) * C * n;   // partial products of step k-1
        double* pc = pbuf + (size_t)(k & 1) * C * n;               // ... of this step
        const double* up = ubuf + (size_t)((k + 1) & 1) * n;       // reflector of step k-1 (= pending v)
        double* uc = ubuf + (size_t)(k & 1) * n;                   // reflector of this step
        auto wprev = [&](int j) { return fma(kk_prev, up[j], psum(pp, j)); };
        // (1) row k with the pending pair applied -> d[k], reflector u of column k
        const double* rowk = A + (size_t)k * n;
        double vk = 0.0, wk = 0.0;
        if (pend) { vk = up[k]; wk = wprev(k); }
        double ss = 0.0;
        for (int j = k + 2 + tid; j < n; j += TC_THREADS) {
            double x = rowk[j];
            if (pend) x = x - vk * wprev(j) - wk * up[j];
            if (own(j)) u[loc(j)] = x;
            if (lead) uc[j] = x;
            ss = fma(x, x, ss);
        }
        ss = tc_block_sum(ss, red);
        double alpha = rowk[k + 1], dk = rowk[k];
        if (pend) {
            alpha = alpha - vk * wprev(k + 1) - wk * up[k + 1];
            dk = dk - 2.0 * vk * wk;
        }
        double tau = 0.0, beta = alpha;
        if (ss > 0.0) {
            const double nrm = sqrt(fma(alpha, alpha, ss));
            beta = alpha >= 0.0 ? -nrm : nrm;
            tau = (beta - alpha) / beta;
            const double inv = 1.0 / (alpha - beta);
            for (int j = k + 2 + tid; j < n; j += TC_THREADS) {
                if (own(j)) u[loc(j)] *= inv;
                if (lead) uc[j] *= inv;
            }
        }
        if (tid == 0) {
            if (own(k + 1)) u[loc(k + 1)] = 1.0;
            if (lead) {
                uc[k + 1] = 1.0;
                d[k] = dk;
                e[k] = beta;
            }
        }
        __syncthreads();
        // (2) one pass over this warp's rows, this CTA's column chunks
        if (pend || tau != 0.0) {
            int i = gw + ((k + 1 - gw + n_gw - 1) / n_gw) * n_gw;
            if (gw >= k + 1) i = gw;
            const int q0 = (k + 1) >> 6;
            const int qs = q0 + ((cls - q0 % C) + C) % C;   // first chunk of this class holding a column >= k+1
            for (; i < n; i += n_gw) {
                double* row = A + (size_t)i * n;
                double vi = 0.0, wi = 0.0;
                if (pend) { vi = up[i]; wi = wprev(i); }
                double s = 0.0;
                if ((n & 1) == 0) {
                    // TCF_UNROLL independent 16-byte loads per lane in flight before the first dependent store
                    for (int q = qs; (q << 6) < n; q += TCF_UNROLL * C) {
                        double2 a[TCF_UNROLL];
#pragma unroll
                        for (int t = 0; t < TCF_UNROLL; ++t) {
                            const int j = ((q + t * C) << 6) + 2 * lane;
                            if (j < n) a[t] = *reinterpret_cast<const double2*>(row + j);
                        }
#pragma unroll
                        for (int t = 0; t < TCF_UNROLL; ++t) {
                            const int qq = q + t * C;
                            const int j = (qq << 6) + 2 * lane;
                            if (j >= n) break;
                            const int lj = ((qq / C) << 6) + 2 * lane;
                            if (pend) {
                                const double2 wj = *reinterpret_cast<const double2*>(w + lj);
                                const double2 vj = *reinterpret_cast<const double2*>(v + lj);
                                a[t].x = a[t].x - vi * wj.x - wi * vj.x;
                                a[t].y = a[t].y - vi * wj.y - wi * vj.y;
                                *reinterpret_cast<double2*>(row + j) = a[t];   // columns <= k are dead data
                            }
                            const double2 uj = *reinterpret_cast<const double2*>(u + lj);
                            if (j >= k + 1) s = fma(a[t].x, uj.x, s);
                            if (j + 1 >= k + 1) s = fma(a[t].y, uj.y, s);
                        }
                    }
                } else {
                    for (int q = qs; (q << 6) < n; q += C) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const int j = (q << 6) + lane + 32 * h;
                            if (j >= n || j < k + 1) continue;
                            const int lj = ((q / C) << 6) + lane + 32 * h;
                            double a = row[j];
                            if (pend) {
                                a = a - vi * w[lj] - wi * v[lj];
                                row[j] = a;
                            }
                            s = fma(a, u[lj], s);
                        }
                    }
                }
                s = tc_warp_sum(s);
                if (lane == 0) pc[(size_t)cls * n + i] = tau * s;
            }
        }
        grid.sync();
        // (3) w_new = p - (tau/2)(p.u) u ;  (v, w) <- (u, w_new) for this CTA's chunks
        if (tau != 0.0) {
            double pv = 0.0;
            for (int j = k + 1 + tid; j < n; j += TC_THREADS) pv = fma(psum(pc, j), uc[j], pv);
            pv = tc_block_sum(pv, red);
            const double kk = -0.5 * tau * pv;
            for (int j = k + 1 + tid; j < n; j += TC_THREADS) {
                if (!own(j)) continue;
                const int lj = loc(j);
                const double uj = u[lj];
                w[lj] = fma(kk, uj, psum(pc, j));
                v[lj] = uj;
            }
            kk_prev = kk;
            pend = true;
        } else {
            pend = false;
        }
        __syncthreads();
    }
    if (lead && tid == 0) {
        const int a = n - 2, b = n - 1;
        const int kl = n - 3;   // last eliminated column (only meaningful when pend)
        const double* pp = pbuf + (size_t)(kl & 1) * C * n;
        const double* up = ubuf + (size_t)(kl & 1) * n;
        double va = 0.0, vb = 0.0, wa = 0.0, wb = 0.0;
        if (pend) {
            va = up[a]; vb = up[b];
            wa = fma(kk_prev, va, psum(pp, a));
            wb = fma(kk_prev, vb, psum(pp, b));
        }
        if (n >= 2) {
            d[a] = A[(size_t)a * n + a] - 2.0 * va * wa;
            e[a] = A[(size_t)a * n + b] - (va * wb + wa * vb);
        }
        d[b] = A[(size_t)b * n + b] - 2.0 * vb * wb;
        e[b] = 0.0;
    }
}

// #eigenvalues of the symmetric tridiagonal (d, e) below x, from the sign changes of the leading principal minors
//   p_0 = 1, p_1 = d_0 - x, p_i = (d_{i-1} - x) p_{i-1} - e_{i-2}^2 p_{i-2}      (Sturm sequence, product form).
// The quotient form q_i = p_i / p_{i-1} used by LAPACK's dlaebz costs an fp64 division (~10 fp64 instructions) per
// step and fp64 vector issue is the scarce resource here (measured: the quotient form ran at the fp64 issue limit);
// this form is 3 fp64 instructions per step, signs and magnitudes are examined on the integer pipe, and the pair is
// rescaled by an exact power of two whenever it leaves [2^-400, 2^400].
__device__ __forceinline__ int sturm_count_l(const double* __restrict__ d, const double* __restrict__ e2, int n, double x) {
    double p0 = 1.0, p1 = d[0] - x;
    if (p1 == 0.0) p1 = -1e-300;
    int cnt = __double2hiint(p1) < 0;
    // branch-free step (the rare events -- an exact zero, a rescale -- are selects): the warps of a CTA stay converged and
    // the compiler can overlap the loads and the d[i] - x of the next steps with the dependent fma chain
#pragma unroll 4
    for (int i = 1; i < n; ++i) {
        double p2 = fma(d[i] - x, p1, -(e2[i - 1] * p0));
        const int h1 = __double2hiint(p1);
        p2 = (p2 == 0.0) ? __hiloint2double((h1 ^ 0x80000000) - (60 << 20), __double2loint(p1)) : p2;   // -p1 2^-60: a sign change
        const int hi = __double2hiint(p2);
        cnt += (unsigned)(hi ^ h1) >> 31;
        const int ex = (hi >> 20) & 0x7ff;                   // biased exponent
        const int adj = ex > 1023 + 400 ? -(400 << 20) : (ex < 1023 - 400 ? (400 << 20) : 0);   // exact power-of-two rescale of the pair
        p0 = __hiloint2double(h1 + ((h1 & 0x7ff00000) ? adj : 0), __double2loint(p1));
        p1 = __hiloint2double(hi + adj, __double2loint(p2));
    }
    return cnt;
}

// ---------------------------------- multisection over many CTAs -----------------------------
// One WARP per eigenvalue: every round the 32 lanes evaluate the Sturm count at 32 interior points of the current interval
// and the interval shrinks 33-fold (5 bits per round instead of the 1 bit of a bisection step: ~8 rounds of n dependent
// steps from the Gershgorin interval to fp32 resolution instead of ~40-60).  The recurrence is sequential in i, so the time
// of this stage is rounds x n x (latency of one step); the extra evaluations ride on parallelism the one-thread-per-
// eigenvalue version left idle (n = 4096: 16 CTAs on 148 SMs).  d / e^2 in shared memory, 32 eigenvalues per CTA.
constexpr int BL_THREADS = 1024;
constexpr int BL_WARPS = BL_THREADS / 32;

__global__ void __launch_bounds__(BL_THREADS, 1) bisect_large_kernel(const double* __restrict__ dg, const double* __restrict__ eg,
                                                                     int n, int n_out, float* __restrict__ sv) {
    extern __shared__ __align__(16) double bl_smem[];
    double* d = bl_smem;
    double* e2 = bl_smem + n;
    __shared__ double s_lo[BL_WARPS], s_hi[BL_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double lo = 1e300, hi = -1e300;
    for (int i = threadIdx.x; i < n; i += BL_THREADS) {
        const double di = dg[i], ei = eg[i];
        d[i] = di;
        e2[i] = ei * ei;
        const double el = i > 0 ? fabs(eg[i - 1]) : 0.0, er = i + 1 < n ? fabs(ei) : 0.0;
        lo = fmin(lo, di - el - er);
        hi = fmax(hi, di + el + er);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) { s_lo[warp] = lo; s_hi[warp] = hi; }
    __syncthreads();
    for (int q = 0; q < BL_WARPS; ++q) { lo = fmin(lo, s_lo[q]); hi = fmax(hi, s_hi[q]); }
    const double span = fmax(fabs(lo), fabs(hi));
    lo = lo - 1e-14 * span - 1e-300;
    hi = hi + 1e-14 * span + 1e-300;
    const int k = n - 1 - (blockIdx.x * BL_WARPS + warp);   // this warp's eigenvalue: the k-th smallest (largest ones first)
    if (k < 0 || n - 1 - k >= n_out) return;
    const double atol = 1e-17 * fmax(fabs(lo), fabs(hi));
    for (int it = 0; it < 64; ++it) {
        // the result is rounded to fp32 after a square root: 4e-10 relative on lambda is 3e-3 ulp of sqrt(lambda)
        if (hi - lo <= fmax(atol, 4e-10 * fmax(fabs(lo), fabs(hi)))) break;
        const double x = fma((hi - lo) * (1.0 / 33.0), (double)(lane + 1), lo);   // lo < x_0 < ... < x_31 < hi (up to rounding)
        const bool above = sturm_count_l(d, e2, n, x) > k;   // x is above the k-th eigenvalue
        const unsigned m = __ballot_sync(0xffffffffu, above);
        const int f = m ? __ffs(m) - 1 : 32;                  // first probe above it
        const double x_hi = __shfl_sync(0xffffffffu, x, f < 32 ? f : 31);
        const double x_lo = __shfl_sync(0xffffffffu, x, f > 0 ? f - 1 : 0);
        const double nlo = f > 0 ? fmax(x_lo, lo) : lo, nhi = f < 32 ? fmin(x_hi, hi) : hi;
        if (!(nhi - nlo < hi - lo)) break;                    // no progress: the interval is at the resolution of fp64
        lo = nlo;
        hi = nhi;
    }
    if (lane == 0) sv[n - 1 - k] = (float)sqrt(fmax(0.5 * (lo + hi), 0.0));
}

// d[n], e[n], partial products [2][classes <= 2][n], reflectors [2][n]
size_t tridiag_panel_workspace_doubles(int n, int grid);
cudaError_t launch_tridiag_panel(double* G, int n, double* d, double* e, double* ws, int num_sms, cudaStream_t stream);
bool tridiag_panel_supported(int n, int num_sms);

// d, e, then the larger of the fused kernels' 6 n doubles and the panel kernel's buffers (sized for up to 192 CTAs)
// what the fused kernels alone need (classes 1, 2, -1): d, e and 6 n doubles
size_t spectrum_large_fused_workspace_bytes(int n) { return (size_t)8 * n * sizeof(double) + 256; }

size_t spectrum_large_workspace_bytes(int n) {
    const size_t panel = tridiag_panel_workspace_doubles(n, 192), fused = (size_t)6 * n;
    return ((size_t)2 * n + (panel > fused ? panel : fused)) * sizeof(double) + 256;
}

template <int C>
static cudaError_t launch_tridiag_cls(double* G, int n, double* d, double* e, double* p, double* u, int num_sms,
                                      cudaStream_t stream) {
    const int Ls = ((((n + 63) >> 6) + C - 1) / C) << 6;
    const size_t tsmem = (size_t)3 * Ls * sizeof(double);
    if (tsmem > 220 * 1024) return cudaErrorInvalidValue;
    cudaError_t err = cudaFuncSetAttribute(tridiag_fused_cls_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)tsmem);
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tridiag_fused_cls_kernel<C>, TC_THREADS, tsmem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) return cudaErrorInvalidValue;
    int grid = num_sms;
    const int max_useful = C * ((n + TC_WARPS - 1) / TC_WARPS);
    if (grid > max_useful) grid = max_useful;
    grid = (grid / C) * C;
    if (grid < C) return cudaErrorInvalidValue;
    void* args[] = {&G, &n, &d, &e, &p, &u};
    return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(tridiag_fused_cls_kernel<C>), dim3(grid),
                                       dim3(TC_THREADS), args, tsmem, stream);
}

// classes: 0 = choose, 1 = one-class fused kernel, 2 = two column classes, 3 = panel kernel with deferred updates
// (idb_eig_panel.cu; even n), -1 = the unfused two-pass kernel (kept as the cross-check of the others)
cudaError_t launch_spectrum_large(double* G, int n, int n_out, float* sv, void* workspace, int num_sms,
                                  cudaStream_t stream, int classes) {
    if (n <= 0) return cudaSuccess;
    double* d = reinterpret_cast<double*>(workspace);
    double* e = d + n;
    double* p = e + n;
    double* u = p + (size_t)4 * n;
    if (classes == 0) classes = (n >= PANEL_MIN_N && tridiag_panel_supported(n, num_sms)) ? 3 : (n <= TCF_MAX_N ? 1 : 2);
    cudaError_t err;
    if (classes == 3) {
        err = launch_tridiag_panel(G, n, d, e, p, num_sms, stream);
        if (err != cudaSuccess) return err;
    } else if (classes == 2) {
        err = launch_tridiag_cls<2>(G, n, d, e, p, u, num_sms, stream);
        if (err != cudaSuccess) return err;
    } else if (classes == 1) {
        err = launch_tridiag_cls<1>(G, n, d, e, p, u, num_sms, stream);
        if (err != cudaSuccess) return err;
    } else {
        void* kernel = reinterpret_cast<void*>(tridiag_coop_kernel);
        const size_t tsmem = (size_t)2 * n * sizeof(double);
        if (tsmem > 220 * 1024) return cudaErrorInvalidValue;
        err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem);
        if (err != cudaSuccess) return err;
        int per_sm = 0;
        err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tridiag_coop_kernel, TC_THREADS, tsmem);
        if (err != cudaSuccess) return err;
        if (per_sm < 1) return cudaErrorInvalidValue;
        int grid = num_sms;   // one CTA per SM
        const int max_useful = (n + TC_WARPS - 1) / TC_WARPS;
        if (grid > max_useful) grid = max_useful;
        void* args[] = {&G, &n, &d, &e, &p};
        err = cudaLaunchCooperativeKernel(kernel, dim3(grid), dim3(TC_THREADS), args, tsmem, stream);
        if (err != cudaSuccess) return err;
    }
    const size_t smem = (size_t)2 * n * sizeof(double);
    err = cudaFuncSetAttribute(bisect_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    bisect_large_kernel<<<(n_out + BL_WARPS - 1) / BL_WARPS, BL_THREADS, smem, stream>>>(d, e, n, n_out, sv);
    return cudaGetLastError();
}

}  // namespace idb
