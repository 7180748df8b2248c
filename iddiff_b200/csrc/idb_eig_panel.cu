// Panel (deferred-update) Householder tridiagonalisation of ONE large symmetric fp64 matrix by the whole GPU.
//
// The fused kernel of idb_eig_large.cu reads AND writes the trailing block once per eliminated column (16 m^2 bytes): it is
// a level-2 algorithm at its HBM roofline (5.1 TB/s at n = 12288).  This kernel defers the rank-2 updates the way LAPACK's
// dlatrd does: the reflector pairs (v_p, w_p) of up to NB columns stay PENDING while the trailing block in HBM is stale,
//     A_true = A_stale - sum_p (v_p w_p^T + w_p v_p^T),
// a column's symmetric product is taken with the stale block and corrected with 2 NB dot products,
//     A_true u = A_stale u - sum_p (v_p (w_p.u) + w_p (v_p.u)),
// and only every NB-th column's sweep applies all pending pairs and writes the block back: (8 + 8 / NB) m^2 bytes per column
// instead of 16 -- 9 at NB = 8.  Replaces torch.linalg.svd at dim_reduction.py:197-198 for image-sized score matrices
// (same role as tridiag_fused_cls_kernel; selected by launch_spectrum_large for even n >= 2048).
//
// Work split per column k (cooperative launch, one CTA per SM, three grid synchronisations):
//   phase A  every CTA owns a cyclic SLICE of the indices j > k: finalises the previous column's w there
//            (w = tau (A_stale u - corrections) + kk u), forms its slice of row k of A_true and the partial norm;
//   phase B  norm -> (tau, beta), reflector slice u -> global;
//   sweep    rows are owned by fixed warps (row i -> warp i mod #warps); columns are walked in blocks of LB whose vectors
//            (u, and on a flush all pending v_p, w_p) sit in shared memory, each loaded vector element serves all of the
//            warp's rows; the sweep also accumulates the dot products the next phase A needs (pc.u, w_p.u, v_p.u).
#include "idb_common.cuh"
#include "idb_kernels.h"
#include <cooperative_groups.h>
#include <math.h>

namespace cg = cooperative_groups;

namespace idb {

namespace {
constexpr int TP_THREADS = 512;
constexpr int TP_WARPS = TP_THREADS / 32;
constexpr int TP_NB = 8;       // pending reflector pairs before a flush
constexpr int TP_LB = 1024;    // columns per shared-memory block
constexpr int TP_RMAX = 8;     // rows of the trailing block per warp (n <= TP_RMAX * gridDim.x * TP_WARPS)
constexpr int TP_NDOT = 2 * TP_NB + 1;

__device__ __forceinline__ double tp_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double tp_block_sum(double v, double* red) {
    v = tp_warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < TP_WARPS; ++w) t += red[w];
    return t;
}
}  // namespace

size_t tridiag_panel_workspace_doubles(int n, int grid) {
    return (size_t)(2 * TP_NB + 2) * n + (size_t)grid * (TP_NDOT + 1) + 64;
}

__global__ void __launch_bounds__(TP_THREADS, 1) tridiag_panel_kernel(double* __restrict__ A, int n, double* __restrict__ d,
                                                                      double* __restrict__ e, double* __restrict__ ws) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) double tp_smem[];
    double* ub = tp_smem;                                   // [LB]
    double* vb = ub + TP_LB;                                // [NB][LB]
    double* wb = vb + TP_NB * TP_LB;                        // [NB][LB]
    double* rowsc = wb + TP_NB * TP_LB;                     // [WARPS][RMAX][2 NB]: v_p[i], w_p[i] of the warp's rows
    __shared__ double red[TP_WARPS];
    __shared__ double s_dot[TP_NDOT];                       // pc.u, (w_p.u)_p, (v_p.u)_p of the previous sweep
    __shared__ double s_vk[TP_NB + 1], s_wk[TP_NB + 1];     // v_p[k], w_p[k]
    __shared__ double s_sc[4];

    const int G = gridDim.x, b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_gw = G * TP_WARPS, gw = b * TP_WARPS + warp;
    double* Vbuf = ws;                                      // [NB][n]
    double* Wbuf = Vbuf + (size_t)TP_NB * n;                // [NB][n]
    double* uc = Wbuf + (size_t)TP_NB * n;                  // [n] reflector of the current column
    double* pc = uc + n;                                    // [n] A u of the current column (stale or true, unscaled)
    double* part_ss = pc + n;                               // [G]
    double* part_dot = part_ss + G;                         // [G][NDOT]
    double* alpha_g = part_dot + (size_t)G * TP_NDOT;       // [1]

    // replicated scalar state (identical in every CTA)
    int r = 0;                 // pending pairs whose w is final (slots 0 .. r-1)
    bool have_prev = false;    // the previous column produced a reflector whose w still has to be finalised in slot r
    bool prev_flushed = false; // ... and its sweep applied the pending pairs (its A u is already the true one)
    int r_prev = 0;            // pending pairs that were outstanding during the previous sweep
    double tau_prev = 0.0;

    for (int k = 0; k + 1 < n; ++k) {
        const bool last = (k + 2 >= n);                     // k = n - 2: only finalise + read the last 2 x 2 block
        // ---------------- phase A ----------------
        // (a) scalars of the previous column from the sweep's partial dot products
        double kk = 0.0;
        if (have_prev) {
            for (int c = warp; c < TP_NDOT; c += TP_WARPS) {
                double t = 0.0;
                for (int q = lane; q < G; q += 32) t += part_dot[(size_t)q * TP_NDOT + c];
                t = tp_warp_sum(t);
                if (lane == 0) s_dot[c] = t;
            }
            __syncthreads();
            double corr = 0.0;
            if (!prev_flushed)
                for (int p = 0; p < r_prev; ++p) corr += s_dot[1 + p] * s_dot[1 + TP_NB + p];
            const double pv = tau_prev * (s_dot[0] - 2.0 * corr);
            kk = -0.5 * tau_prev * pv;
        }
        const int slot = r;                                 // where the previous column's pair goes
        // w_new[j] of the previous column (only meaningful when have_prev)
        auto w_new_at = [&](int j) {
            double base = pc[j];
            if (!prev_flushed)
                for (int p = 0; p < r_prev; ++p)
                    base -= Vbuf[(size_t)p * n + j] * s_dot[1 + p] + Wbuf[(size_t)p * n + j] * s_dot[1 + TP_NB + p];
            return fma(kk, uc[j], tau_prev * base);
        };
        // (b) v_p[k], w_p[k] of every pending pair (the newest: v[k] = u_prev[k] = 1, w[k] computed here)
        if (tid <= r) {
            if (tid < r) { s_vk[tid] = Vbuf[(size_t)tid * n + k]; s_wk[tid] = Wbuf[(size_t)tid * n + k]; }
            else if (have_prev) { s_vk[tid] = uc[k]; s_wk[tid] = w_new_at(k); }
        }
        __syncthreads();
        const int r_now = r + (have_prev ? 1 : 0);
        // (c) this CTA's slice of j in (k, n): finalise w_new there, then row k of A_true
        const int j = k + 1 + tid * G + b;
        double x = 0.0;
        if (j < n) {
            double vj_new = 0.0, wj_new = 0.0;
            if (have_prev) {
                vj_new = uc[j];
                wj_new = w_new_at(j);
                Vbuf[(size_t)slot * n + j] = vj_new;
                Wbuf[(size_t)slot * n + j] = wj_new;
            }
            x = A[(size_t)k * n + j];
            for (int p = 0; p < r; ++p) x -= s_vk[p] * Wbuf[(size_t)p * n + j] + s_wk[p] * Vbuf[(size_t)p * n + j];
            if (have_prev) x -= s_vk[r] * wj_new + s_wk[r] * vj_new;
        }
        if (b == 0 && tid == 0) {
            double dk = A[(size_t)k * n + k];
            for (int p = 0; p < r_now; ++p) dk -= 2.0 * s_vk[p] * s_wk[p];
            d[k] = dk;
        }
        if (last) {
            // x (thread 0 of the CTA owning j = n - 1) is A_true[n-2][n-1]; the last diagonal entry needs the pairs too
            if (j == n - 1) {
                e[n - 2] = x;
                e[n - 1] = 0.0;
                double dl = A[(size_t)(n - 1) * n + (n - 1)];
                for (int p = 0; p < r; ++p) dl -= 2.0 * Vbuf[(size_t)p * n + j] * Wbuf[(size_t)p * n + j];
                if (have_prev) dl -= 2.0 * uc[j] * w_new_at(j);
                d[n - 1] = dl;
            }
            break;
        }
        if (j == k + 1) *alpha_g = x;
        const double ss_part = tp_block_sum((j < n && j >= k + 2) ? x * x : 0.0, red);
        if (tid == 0) part_ss[b] = ss_part;
        grid.sync();
        // ---------------- phase B ----------------
        if (warp == 0) {
            double t = 0.0;
            for (int q = lane; q < G; q += 32) t += part_ss[q];
            t = tp_warp_sum(t);
            if (lane == 0) { s_sc[0] = t; s_sc[1] = *alpha_g; }
        }
        __syncthreads();
        const double ss = s_sc[0], alpha = s_sc[1];
        double tau = 0.0, beta = alpha;
        if (ss > 0.0) {
            const double nrm = sqrt(fma(alpha, alpha, ss));
            beta = alpha >= 0.0 ? -nrm : nrm;
            tau = (beta - alpha) / beta;
            const double inv = 1.0 / (alpha - beta);
            if (j < n) uc[j] = (j == k + 1) ? 1.0 : x * inv;
        }
        if (b == 0 && tid == 0) e[k] = beta;
        r = r_now;                                          // the previous pair is final now (slots 0 .. r-1)
        grid.sync();
        // ---------------- sweep ----------------
        const bool flush = (r == TP_NB);
        if (tau != 0.0) {
            // rows of this warp: i_t = first + t n_gw >= k + 1
            int first = gw;
            if (first < k + 1) first += ((k + 1 - first + n_gw - 1) / n_gw) * n_gw;
            int nrows = 0;
            if (first < n) nrows = (n - 1 - first) / n_gw + 1;
            double* rs = rowsc + (size_t)warp * TP_RMAX * 2 * TP_NB;
            for (int q = lane; q < nrows * 2 * r; q += 32) {
                const int t = q / (2 * r), c = q % (2 * r);
                const int i = first + t * n_gw;
                if (c < r) rs[t * 2 * TP_NB + c] = Vbuf[(size_t)c * n + i];                         // v_p[i] at [p]
                else rs[t * 2 * TP_NB + TP_NB + (c - r)] = Wbuf[(size_t)(c - r) * n + i];          // w_p[i] at [NB + p]
            }
            double s[TP_RMAX];
#pragma unroll
            for (int t = 0; t < TP_RMAX; ++t) s[t] = 0.0;
            const int cb0 = ((k + 1) / TP_LB) * TP_LB;
            for (int c0 = cb0; c0 < n; c0 += TP_LB) {
                __syncthreads();
                for (int q = tid; q < TP_LB; q += TP_THREADS) {
                    const int jj = c0 + q;
                    ub[q] = (jj >= k + 1 && jj < n) ? uc[jj] : 0.0;
                }
                if (flush) {
                    for (int q = tid; q < TP_NB * TP_LB; q += TP_THREADS) {
                        const int p = q / TP_LB, jj = c0 + q % TP_LB;
                        const bool ok = jj < n;
                        vb[q] = ok ? Vbuf[(size_t)p * n + jj] : 0.0;
                        wb[q] = ok ? Wbuf[(size_t)p * n + jj] : 0.0;
                    }
                }
                __syncthreads();
                const int cend = min(TP_LB, n - c0);
                for (int q0 = 2 * lane; q0 < cend; q0 += 64) {
                    double2 a[TP_RMAX];
#pragma unroll
                    for (int t = 0; t < TP_RMAX; ++t)
                        if (t < nrows) a[t] = *reinterpret_cast<const double2*>(A + (size_t)(first + t * n_gw) * n + c0 + q0);
                    if (flush) {
#pragma unroll
                        for (int p = 0; p < TP_NB; ++p) {
                            const double2 vj = *reinterpret_cast<const double2*>(vb + p * TP_LB + q0);
                            const double2 wj = *reinterpret_cast<const double2*>(wb + p * TP_LB + q0);
#pragma unroll
                            for (int t = 0; t < TP_RMAX; ++t) {
                                if (t < nrows) {
                                    const double vi = rs[t * 2 * TP_NB + p], wi = rs[t * 2 * TP_NB + TP_NB + p];
                                    a[t].x = a[t].x - vi * wj.x - wi * vj.x;
                                    a[t].y = a[t].y - vi * wj.y - wi * vj.y;
                                }
                            }
                        }
#pragma unroll
                        for (int t = 0; t < TP_RMAX; ++t)
                            if (t < nrows) *reinterpret_cast<double2*>(A + (size_t)(first + t * n_gw) * n + c0 + q0) = a[t];
                    }
                    const double2 uj = *reinterpret_cast<const double2*>(ub + q0);
#pragma unroll
                    for (int t = 0; t < TP_RMAX; ++t)
                        if (t < nrows) s[t] = fma(a[t].y, uj.y, fma(a[t].x, uj.x, s[t]));
                }
            }
            // row results + the dot products of the next phase A
            double dots[TP_NDOT];
#pragma unroll
            for (int c = 0; c < TP_NDOT; ++c) dots[c] = 0.0;
#pragma unroll
            for (int t = 0; t < TP_RMAX; ++t) {
                if (t < nrows) {
                    const double st = tp_warp_sum(s[t]);
                    if (lane == 0) {
                        const int i = first + t * n_gw;
                        const double ui = uc[i];
                        pc[i] = st;
                        dots[0] = fma(st, ui, dots[0]);
                        if (!flush) {
                            for (int p = 0; p < r; ++p) {
                                dots[1 + p] = fma(rs[t * 2 * TP_NB + TP_NB + p], ui, dots[1 + p]);      // w_p[i] u[i]
                                dots[1 + TP_NB + p] = fma(rs[t * 2 * TP_NB + p], ui, dots[1 + TP_NB + p]);  // v_p[i] u[i]
                            }
                        }
                    }
                }
            }
            // CTA partial sums (lane 0 of every warp holds its share)
            __syncthreads();
            double* scratch = ub;                              // [WARPS][NDOT] reuse
            if (lane == 0)
                for (int c = 0; c < TP_NDOT; ++c) scratch[warp * TP_NDOT + c] = dots[c];
            __syncthreads();
            if (tid < TP_NDOT) {
                double t = 0.0;
                for (int w2 = 0; w2 < TP_WARPS; ++w2) t += scratch[w2 * TP_NDOT + tid];
                part_dot[(size_t)b * TP_NDOT + tid] = t;
            }
            have_prev = true;
            prev_flushed = flush;
            r_prev = r;
            tau_prev = tau;
            if (flush) r = 0;
        } else {
            have_prev = false;
        }
        grid.sync();
    }
}

cudaError_t launch_tridiag_panel(double* G, int n, double* d, double* e, double* ws, int num_sms, cudaStream_t stream) {
    const size_t smem = ((size_t)(1 + 2 * TP_NB) * TP_LB + (size_t)TP_WARPS * TP_RMAX * 2 * TP_NB) * sizeof(double);
    cudaError_t err = cudaFuncSetAttribute(tridiag_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tridiag_panel_kernel, TP_THREADS, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) return cudaErrorInvalidValue;
    int grid = num_sms;
    if ((n & 1) || (long long)n > (long long)TP_RMAX * grid * TP_WARPS || (n + grid - 1) / grid > TP_THREADS)
        return cudaErrorInvalidValue;
    void* args[] = {&G, &n, &d, &e, &ws};
    return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(tridiag_panel_kernel), dim3(grid), dim3(TP_THREADS), args, smem,
                                       stream);
}

}  // namespace idb
