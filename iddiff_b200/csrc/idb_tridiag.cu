// Spectrum of the centred score matrix from its fp64 Gram matrix, method 1 (default):
//   Householder tridiagonalisation of G (fp64, packed lower triangle resident in shared memory, one
//   anchor per CTA)  ->  Sturm-sequence bisection of the tridiagonal (fp64, one thread per eigenvalue)
//   ->  sigma_i = sqrt(lambda_i), descending.
// Replaces torch.linalg.svd at dim_reduction.py:197-198 of the reference (LAPACK gesdd on the host).
//
// Why next to the one-sided Jacobi of idb_spectrum.cu (method 0): Jacobi on a 200 x 200 factor re-reads
// both columns of every pair from shared memory in every rotation -- 12 n^3 bytes of smem traffic per
// sweep, ~9 sweeps -- and measured 63 ms per 512 anchors; the two-stage solver touches the packed matrix
// ~3 times per eliminated column (n^3 bytes in total) and has no convergence loop.  Backward error is
// n * eps64 * |G|, i.e. relative 1e-16 * n * (s_max / s_i)^2 on singular value s_i: 2e-8 at a 1e3 cliff.
#include "idb_common.cuh"
#include "idb_kernels.h"
#include <math.h>

namespace idb {

constexpr int TRI_THREADS = 1024;
constexpr int TRI_WARPS = TRI_THREADS / 32;
constexpr int TRI_GROUPS = TRI_THREADS / 256;
constexpr size_t TRI_SMEM_MAX = 226 * 1024;

__host__ __device__ inline size_t tri_packed_bytes(int n) { return (size_t)n * (n + 1) / 2 * sizeof(double); }
// v, p and TRI_GROUPS partial-sum vectors
__host__ __device__ inline size_t tri_vec_bytes(int n) { return (size_t)(2 + TRI_GROUPS) * ((n + 1) & ~1) * sizeof(double); }

size_t tridiag_workspace_bytes(int n_anchors, int n) {
    // per anchor: packed fp64 matrix (used only when it does not fit shared memory) + d[n] + e[n]
    return (size_t)n_anchors * (tri_packed_bytes(n) + 2 * (size_t)n * sizeof(double)) + 256;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sum; every thread gets the result.  red: TRI_WARPS doubles of shared memory.
__device__ __forceinline__ double block_sum(double v, double* red) {
    v = warp_sum(v);
    __syncthreads();  // red may still be read from the previous reduction
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < TRI_WARPS; ++w) t += red[w];
    return t;
}

// A: packed lower triangle, row-major (A[i][j], j <= i, at i(i+1)/2 + j).
__global__ void __launch_bounds__(TRI_THREADS, 1) tridiag_kernel(const double* __restrict__ G, int n,
                                                                 uint8_t* __restrict__ workspace, int a_in_smem) {
    extern __shared__ __align__(16) uint8_t tri_smem[];
    __shared__ double red[TRI_WARPS];
    const int a = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t per_anchor = tri_packed_bytes(n) + 2 * (size_t)n * sizeof(double);
    uint8_t* ws = workspace + (size_t)a * per_anchor;
    double* vec = reinterpret_cast<double*>(tri_smem);
    double* v = vec;                      // Householder vector (indexed by absolute row)
    const int nv = (n + 1) & ~1;
    double* p = vec + nv;                 // p = tau A v, then w
    double* cpart = vec + 2 * nv;         // [TRI_GROUPS][nv]
    double* A = a_in_smem ? reinterpret_cast<double*>(tri_smem + tri_vec_bytes(n)) : reinterpret_cast<double*>(ws);
    double* d_out = reinterpret_cast<double*>(ws + tri_packed_bytes(n));
    double* e_out = d_out + n;
    const double* Ga = G + (size_t)a * n * n;

    for (int idx = tid; idx < n * n; idx += TRI_THREADS) {
        const int i = idx / n, j = idx - i * n;
        if (j <= i) A[(size_t)i * (i + 1) / 2 + j] = Ga[idx];
    }
    __syncthreads();

    for (int k = 0; k + 2 < n; ++k) {
        const int r0 = k + 1;  // first row of the trailing block
        // ---- Householder reflector for x = A[r0:n, k]  (LAPACK dlarfg)
        double part = 0.0;
        for (int r = r0 + 1 + tid; r < n; r += TRI_THREADS) {
            const double x = A[r * (r + 1) / 2 + k];
            part = fma(x, x, part);
        }
        const double xnorm2 = block_sum(part, red);
        const double alpha = A[(size_t)r0 * (r0 + 1) / 2 + k];
        if (tid == 0) d_out[k] = A[(size_t)k * (k + 1) / 2 + k];
        if (xnorm2 == 0.0) {  // already tridiagonal in this column
            if (tid == 0) e_out[k] = alpha;
            continue;
        }
        const double beta = -copysign(sqrt(alpha * alpha + xnorm2), alpha);
        const double tau = (beta - alpha) / beta;
        const double inv = 1.0 / (alpha - beta);
        if (tid == 0) e_out[k] = beta;
        for (int r = r0 + tid; r < n; r += TRI_THREADS) v[r] = (r == r0) ? 1.0 : A[r * (r + 1) / 2 + k] * inv;
        __syncthreads();
        // ---- p = tau * A22 v: row part (warp per row, lanes along the contiguous row) ...
        for (int i = r0 + warp; i < n; i += TRI_WARPS) {
            const double* Ai = A + i * (i + 1) / 2;
            double s = 0.0;
            for (int j = r0 + lane; j <= i; j += 32) s = fma(Ai[j], v[j], s);
            s = warp_sum(s);
            if (lane == 0) p[i] = s;
        }
        __syncthreads();
        // ... column part: p_j += sum_{i > j} A[i][j] v_i.  Lanes own consecutive columns j and walk the SAME row i
        // together (consecutive addresses, conflict-free); the rows are dealt round-robin to TRI_GROUPS groups of
        // 256 threads whose partial sums meet in `cpart`.
        {
            const int m = n - r0;
            const int jj0 = tid & 255, grp = tid >> 8;
            for (int jj = jj0; jj < m; jj += 256) {
                const int j = r0 + jj;
                double s = 0.0;
                for (int i = r0 + 1 + grp; i < n; i += TRI_GROUPS)
                    if (i > j) s = fma(A[i * (i + 1) / 2 + j], v[i], s);
                cpart[grp * nv + j] = s;
            }
        }
        __syncthreads();
        for (int r = r0 + tid; r < n; r += TRI_THREADS) {
            double s = p[r];
#pragma unroll
            for (int g = 0; g < TRI_GROUPS; ++g) s += cpart[g * nv + r];
            p[r] = s;
        }
        __syncthreads();
        // ---- w = tau p - (tau^2/2)(p.v) v
        double pv = 0.0;
        for (int r = r0 + tid; r < n; r += TRI_THREADS) pv = fma(p[r], v[r], pv);
        pv = block_sum(pv, red);
        const double c = 0.5 * tau * tau * pv;
        for (int r = r0 + tid; r < n; r += TRI_THREADS) p[r] = tau * p[r] - c * v[r];
        __syncthreads();
        // ---- A22 -= v w^T + w v^T  (lower triangle)
        for (int i = r0 + warp; i < n; i += TRI_WARPS) {
            double* Ai = A + i * (i + 1) / 2;
            const double vi = v[i], wi = p[i];
            for (int j = r0 + lane; j <= i; j += 32) Ai[j] -= vi * p[j] + wi * v[j];
        }
        __syncthreads();
    }
    if (tid == 0) {
        if (n >= 2) {
            d_out[n - 2] = A[(size_t)(n - 2) * (n - 1) / 2 + (n - 2)];
            e_out[n - 2] = A[(size_t)(n - 1) * n / 2 + (n - 2)];
        }
        d_out[n - 1] = A[(size_t)(n - 1) * n / 2 + (n - 1)];
        e_out[n - 1] = 0.0;
    }
}

// Fused variant (default): the rank-2 update of step k-1 and the symmetric matrix-vector product of step k in ONE pass
// over the packed trailing block.  A warp owns batches of four consecutive rows (r0 + 4 warp + 64 q ...) and walks each along its
// contiguous columns (conflict-free); an element A[i][j] is read once, gets the pending pair (v, w) applied, is
// written back, and feeds BOTH halves of the symmetric product: the row sum p_i (a warp reduction) and the column sum
// p_j, which a lane accumulates in registers for its own columns j = r0 + lane + 32 t over all rows of the warp.  The
// per-warp column sums meet in shared memory.  2 accesses of the matrix per eliminated column instead of 4, and 5
// block-wide synchronisations instead of 10 -- the stage is latency-bound (one anchor per SM), so that is its time.
#ifdef TF_EXP_THREADS
constexpr int TF_THREADS = TF_EXP_THREADS;
#else
constexpr int TF_THREADS = 256;   // measured (tools/tridiag_small_profile.cu, 512 anchors of n = 200, tridiag + bisect): 512 threads x 4-row
#endif                              // batches 5.07 ms (128-register cap: spills), 384 x 4 3.92, 256 x 4 3.87, 256 x 8 3.79, 192 x 8 4.14, 128 x 8 4.23
constexpr int TF_WARPS = TF_THREADS / 32;
#ifdef TF_EXP_RB
constexpr int TF_RB = TF_EXP_RB;
#else
constexpr int TF_RB = 8;
#endif

__host__ __device__ inline size_t tf_vec_bytes(int n) { return (size_t)(4 + TF_WARPS) * ((n + 1) & ~1) * sizeof(double); }

__device__ __forceinline__ double tf_block_sum(double v, double* red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < TF_WARPS; ++w) t += red[w];
    return t;
}

#ifdef TF_PROFILE   // phase timers of CTA 0 (tools/tridiag_small_profile.cu)
__device__ double tf_prof[8];
#define TF_TICK_START() long long tf_t = clock64()
#define TF_TICK(i) do { long long tf_n = clock64(); tf_acc[i] += tf_n - tf_t; tf_t = tf_n; } while (0)
#else
#define TF_TICK_START() do {} while (0)
#define TF_TICK(i) do {} while (0)
#endif

template <int T, bool CACHE, bool SMEM_A>
__global__ void __launch_bounds__(TF_THREADS, 1) tridiag_fused_small_kernel(const double* __restrict__ G, int n,
                                                                            uint8_t* __restrict__ workspace) {
    extern __shared__ __align__(16) uint8_t tri_smem[];
    __shared__ double red[TF_WARPS];
    const int a = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31;
    // warp-uniform by construction AND known to be so to the compiler: the shuffles of the pass below then need no
    // re-convergence sequences (WARPSYNC.COLLECTIVE / ENDCOLLECTIVE)
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const size_t per_anchor = tri_packed_bytes(n) + 2 * (size_t)n * sizeof(double);
    uint8_t* ws = workspace + (size_t)a * per_anchor;
    const int nv = (n + 1) & ~1;
    double* vec = reinterpret_cast<double*>(tri_smem);
    double* v = vec;                 // pending pair of the previous step (rows >= k)
    double* w = vec + nv;
    double* u = vec + 2 * nv;        // reflector of this step (rows >= k+1)
    double* rowp = vec + 3 * nv;     // row half of A u, then p
    double* cpart = vec + 4 * nv;    // [TF_WARPS][nv] column half of A u, per warp
    double* A;   // compile-time address space: shared-memory loads/stores with 32-bit addressing when it fits
    if (SMEM_A) A = vec + (4 + TF_WARPS) * nv;
    else A = reinterpret_cast<double*>(ws);
    double* d_out = reinterpret_cast<double*>(ws + tri_packed_bytes(n));
    double* e_out = d_out + n;
    const double* Ga = G + (size_t)a * n * n;

    for (int idx = tid; idx < n * n; idx += TF_THREADS) {
        const int i = idx / n, j = idx - i * n;
        if (j <= i) A[(size_t)i * (i + 1) / 2 + j] = Ga[idx];
    }
    __syncthreads();

#ifdef TF_PROFILE
    long long tf_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
    TF_TICK_START();
    bool pend = false;
    for (int k = 0; k + 2 < n; ++k) {
        const int r0 = k + 1;
        // (1) column k of the trailing rows with the pending pair applied -> d[k], reflector u (LAPACK dlarfg)
        const double vk = pend ? v[k] : 0.0, wk = pend ? w[k] : 0.0;
        double part = 0.0;
        for (int r = r0 + 1 + tid; r < n; r += TF_THREADS) {
            double x = A[r * (r + 1) / 2 + k];
            if (pend) x = x - v[r] * wk - w[r] * vk;
            u[r] = x;
            part = fma(x, x, part);
        }
        const double xnorm2 = tf_block_sum(part, red);
        double alpha = A[(size_t)r0 * (r0 + 1) / 2 + k], dk = A[(size_t)k * (k + 1) / 2 + k];
        if (pend) {
            alpha = alpha - v[r0] * wk - w[r0] * vk;
            dk = dk - 2.0 * vk * wk;
        }
        double tau = 0.0, beta = alpha;
        if (xnorm2 > 0.0) {
            beta = -copysign(sqrt(alpha * alpha + xnorm2), alpha);
            tau = (beta - alpha) / beta;
            const double inv = 1.0 / (alpha - beta);
            for (int r = r0 + 1 + tid; r < n; r += TF_THREADS) u[r] *= inv;   // same thread wrote u[r]
        }
        if (tid == 0) {
            u[r0] = 1.0;
            d_out[k] = dk;
            e_out[k] = beta;
        }
        __syncthreads();
        TF_TICK(0);
        // (2) one pass over the trailing block
        if (__any_sync(0xffffffffu, pend || tau != 0.0)) {      // (uniform over the CTA; the vote tells the compiler)
            double cacc[T], uj[CACHE ? T : 1], wj[CACHE ? T : 1], vj[CACHE ? T : 1];
#pragma unroll
            for (int t = 0; t < T; ++t) {
                cacc[t] = 0.0;
                if (CACHE) {
                    const int j = r0 + lane + 32 * t;
                    const bool in = j < n;
                    uj[t] = in ? u[j] : 0.0;
                    wj[t] = in && pend ? w[j] : 0.0;
                    vj[t] = in && pend ? v[j] : 0.0;
                }
            }
            TF_TICK(4);
            // TF_RB rows of this warp in flight together: independent dependency chains and interleaved reductions
            // a batch = TF_RB CONSECUTIVE rows (lengths differ by at most TF_RB - 1: no column group is walked for a row
            // that does not reach it; with rows dealt one by one over the warps a batch spanned 48 rows of length)
            for (int ib = r0 + warp * TF_RB; ib < n; ib += TF_WARPS * TF_RB) {
                double* Ai[TF_RB];
                double vi[TF_RB], wi[TF_RB], ui[TF_RB], sr[TF_RB];
                int ir[TF_RB];
#pragma unroll
                for (int b = 0; b < TF_RB; ++b) {
                    const int i = ib + b;
                    const bool on = i < n;
                    ir[b] = on ? i : -1;                       // -1: no column satisfies j <= i
                    const int ic = on ? i : r0;
                    Ai[b] = A + ((ic * (ic + 1)) >> 1);        // 32-bit: n <= 225 on this path
                    vi[b] = on && pend ? v[ic] : 0.0;
                    wi[b] = on && pend ? w[ic] : 0.0;
                    ui[b] = on ? u[ic] : 0.0;
                    sr[b] = 0.0;
                }
                const int imax = min(ib + TF_RB - 1, n - 1);
                TF_TICK(3);
#pragma unroll
                for (int t = 0; t < T; ++t) {
                    if (r0 + 32 * t > imax) break;             // warp-uniform: no row of this batch reaches these columns
                    const int j = r0 + lane + 32 * t;
                    const double ujt = CACHE ? uj[t] : (j < n ? u[j] : 0.0);
                    const double wjt = CACHE ? wj[t] : (pend && j < n ? w[j] : 0.0);
                    const double vjt = CACHE ? vj[t] : (pend && j < n ? v[j] : 0.0);
                    double x[TF_RB];
                    if (r0 + 32 * t + 31 < ib && ib + TF_RB <= n) {
                        // warp-uniform: the column group lies strictly inside EVERY row of the batch (consecutive rows) --
                        // no masks, no address selects; most of the elements of the block go through here
#pragma unroll
                        for (int b = 0; b < TF_RB; ++b) x[b] = Ai[b][j];
#pragma unroll
                        for (int b = 0; b < TF_RB; ++b) {
                            const double xn = x[b] - vi[b] * wjt - wi[b] * vjt;  // vi = wi = 0 unless pending
                            if (pend) Ai[b][j] = xn;
                            sr[b] = fma(xn, ujt, sr[b]);
                            cacc[t] = fma(xn, ui[b], cacc[t]);
                        }
                    } else {
                        // the groups holding the diagonal: every load is issued up front from a safe address, masked lanes
                        // contribute zeros; straight-line per-row code so that the rows' fp64 chains interleave
#pragma unroll
                        for (int b = 0; b < TF_RB; ++b) x[b] = Ai[b][j <= ir[b] ? j : r0];
#pragma unroll
                        for (int b = 0; b < TF_RB; ++b) {
                            const bool in = j <= ir[b];
                            const double xn = x[b] - vi[b] * wjt - wi[b] * vjt;
                            if (in && pend) Ai[b][j] = xn;
                            sr[b] = fma(xn, in ? ujt : 0.0, sr[b]);
                            cacc[t] = fma(xn, j < ir[b] ? ui[b] : 0.0, cacc[t]);
                        }
                    }
                }
                TF_TICK(5);
                if constexpr (TF_RB == 4) {   // transposing butterfly: four row sums over 32 lanes in 2 + 1 + 3 shuffles; lane 8 g ends with row g
                    const bool h16 = lane & 16, h8 = lane & 8;
                    double t2[2];
#pragma unroll
                    for (int e2 = 0; e2 < 2; ++e2) {
                        const double keep = h16 ? sr[e2 + 2] : sr[e2], send = h16 ? sr[e2] : sr[e2 + 2];
                        t2[e2] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                    }
                    const double keep = h8 ? t2[1] : t2[0], send = h8 ? t2[0] : t2[1];
                    double t1 = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 4);
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 2);
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 1);
                    const int irg = h16 ? (h8 ? ir[3] : ir[2]) : (h8 ? ir[1] : ir[0]);
                    if ((lane & 7) == 0 && irg >= 0) rowp[irg] = t1;
                } else {                      // eight rows: 4 + 2 + 1 + 2 shuffles; lane 4 g ends with row g
                    static_assert(TF_RB == 4 || TF_RB == 8, "butterfly written for four or eight rows");
                    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
                    double t4[4], t2[2];
#pragma unroll
                    for (int e2 = 0; e2 < 4; ++e2) {
                        const double keep = h16 ? sr[(e2 + 4) % TF_RB] : sr[e2], send = h16 ? sr[e2] : sr[(e2 + 4) % TF_RB];
                        t4[e2] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                    }
#pragma unroll
                    for (int e2 = 0; e2 < 2; ++e2) {
                        const double keep = h8 ? t4[e2 + 2] : t4[e2], send = h8 ? t4[e2] : t4[e2 + 2];
                        t2[e2] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                    }
                    const double keep = h4 ? t2[1] : t2[0], send = h4 ? t2[0] : t2[1];
                    double t1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 2);
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 1);
                    const int rg = (h16 ? 4 : 0) + (h8 ? 2 : 0) + (h4 ? 1 : 0);
                    int irg = ir[0];
#pragma unroll
                    for (int e2 = 1; e2 < TF_RB; ++e2) irg = rg == e2 ? ir[e2] : irg;
                    if ((lane & 3) == 0 && irg >= 0) rowp[irg] = t1;
                }
                TF_TICK(6);
            }
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const int j = r0 + lane + 32 * t;
                if (j < n) cpart[warp * nv + j] = cacc[t];
            }
            TF_TICK(7);
        }
        __syncthreads();
        TF_TICK(1);
        // (3) p = tau A u,  w_new = p - (tau/2)(p.u) u,  (v, w) <- (u, w_new)
        if (tau != 0.0) {
            double pv = 0.0;
            for (int r = r0 + tid; r < n; r += TF_THREADS) {
                double sr = rowp[r];
#pragma unroll
                for (int q = 0; q < TF_WARPS; ++q) sr += cpart[q * nv + r];
                sr *= tau;
                rowp[r] = sr;
                pv = fma(sr, u[r], pv);
            }
            pv = tf_block_sum(pv, red);
            const double kk = -0.5 * tau * pv;
            for (int r = r0 + tid; r < n; r += TF_THREADS) w[r] = fma(kk, u[r], rowp[r]);   // same thread wrote rowp[r]
            double* tmp = v;
            v = u;
            u = tmp;
            pend = true;
        } else {
            pend = false;
        }
        __syncthreads();
        TF_TICK(2);
    }
#ifdef TF_PROFILE
    if (blockIdx.x == 0 && tid == 0)
        for (int i = 0; i < 8; ++i) tf_prof[i] = (double)tf_acc[i];
#endif
    if (tid == 0) {
        const int ia = n - 2, ib = n - 1;
        double va = 0.0, vb = 0.0, wa = 0.0, wb = 0.0;
        if (pend) {
            vb = v[ib]; wb = w[ib];
            if (n >= 2) { va = v[ia]; wa = w[ia]; }
        }
        if (n >= 2) {
            d_out[ia] = A[(size_t)ia * (ia + 1) / 2 + ia] - 2.0 * va * wa;
            e_out[ia] = A[(size_t)ib * (ib + 1) / 2 + ia] - (vb * wa + wb * va);
        }
        d_out[ib] = A[(size_t)ib * (ib + 1) / 2 + ib] - 2.0 * vb * wb;
        e_out[ib] = 0.0;
    }
}

template <int T, bool CACHE>
static cudaError_t launch_tridiag_fused_small(const double* G, int n_anchors, int n, void* workspace, cudaStream_t stream) {
    const size_t need = tf_vec_bytes(n) + tri_packed_bytes(n);
    const bool a_in_smem = need <= TRI_SMEM_MAX;
    const size_t smem = a_in_smem ? need : tf_vec_bytes(n);
    if (smem > TRI_SMEM_MAX) return cudaErrorInvalidValue;
    uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
    cudaError_t e;
    if (a_in_smem) {
        e = cudaFuncSetAttribute(tridiag_fused_small_kernel<T, CACHE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem);
        if (e != cudaSuccess) return e;
        tridiag_fused_small_kernel<T, CACHE, true><<<n_anchors, TF_THREADS, smem, stream>>>(G, n, ws);
    } else {
        e = cudaFuncSetAttribute(tridiag_fused_small_kernel<T, CACHE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem);
        if (e != cudaSuccess) return e;
        tridiag_fused_small_kernel<T, CACHE, false><<<n_anchors, TF_THREADS, smem, stream>>>(G, n, ws);
    }
    return cudaGetLastError();
}

// #eigenvalues of the symmetric tridiagonal (d, e) below x, from the sign changes of the leading principal minors
//   p_0 = 1, p_1 = d_0 - x, p_i = (d_{i-1} - x) p_{i-1} - e_{i-2}^2 p_{i-2}      (Sturm sequence, product form).
// The quotient form q_i = p_i / p_{i-1} used by LAPACK's dlaebz costs an fp64 division (~10 fp64 instructions) per
// step and fp64 vector issue is the scarce resource here (measured: the quotient form ran at the fp64 issue limit);
// this form is 3 fp64 instructions per step, signs and magnitudes are examined on the integer pipe, and the pair is
// rescaled by an exact power of two whenever it leaves [2^-400, 2^400].
__device__ __forceinline__ int sturm_count(const double* __restrict__ d, const double* __restrict__ e2, int n, double x) {
    double p0 = 1.0, p1 = d[0] - x;
    if (p1 == 0.0) p1 = -1e-300;
    int cnt = __double2hiint(p1) < 0;
    // branch-free step, as in idb_eig_large.cu: the rare events (an exact zero, a rescale) are selects and exponent arithmetic
#pragma unroll 4
    for (int i = 1; i < n; ++i) {
        double p2 = fma(d[i] - x, p1, -(e2[i - 1] * p0));
        const int h1 = __double2hiint(p1);
        p2 = (p2 == 0.0) ? __hiloint2double((h1 ^ 0x80000000) - (60 << 20), __double2loint(p1)) : p2;   // -p1 2^-60: a sign change
        const int hi = __double2hiint(p2);
        cnt += (unsigned)(hi ^ h1) >> 31;
        const int ex = (hi >> 20) & 0x7ff;                   // biased exponent
        const int adj = ex > 1023 + 400 ? -(400 << 20) : (ex < 1023 - 400 ? (400 << 20) : 0);   // exact power-of-two rescale of the pair
        p0 = __hiloint2double(h1 + adj, __double2loint(p1));
        p1 = __hiloint2double(hi + adj, __double2loint(p2));
    }
    return cnt;
}

// ---------------------------------- Sturm bisection -----------------------------------------
// One CTA per anchor, one thread per eigenvalue (index k = k-th smallest).  d, e^2 live in shared memory and
// are read as warp broadcasts.  count(x) = #eigenvalues < x by the LAPACK dlaebz recurrence
//   q_0 = d_0 - x,  q_i = d_i - x - e_{i-1}^2 / q_{i-1}   (|q| < pivmin -> -pivmin),   count = #{q_i < 0}.
__global__ void __launch_bounds__(256) bisect_kernel(const uint8_t* __restrict__ workspace, int n, int n_out,
                                                     float* __restrict__ sv, int* __restrict__ iters_out) {
    extern __shared__ __align__(16) uint8_t bis_smem[];
    double* d = reinterpret_cast<double*>(bis_smem);
    double* e2 = d + n;
    __shared__ double s_lo, s_hi;
    const int a = blockIdx.x;
    const size_t per_anchor = tri_packed_bytes(n) + 2 * (size_t)n * sizeof(double);
    const double* dg = reinterpret_cast<const double*>(workspace + (size_t)a * per_anchor + tri_packed_bytes(n));
    const double* eg = dg + n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        d[i] = dg[i];
        const double e = eg[i];
        e2[i] = e * e;
    }
    __syncthreads();
    if (threadIdx.x == 0) {  // Gershgorin interval and pivmin
        double lo = d[0], hi = d[0], emax = 0.0;
        for (int i = 0; i < n; ++i) {
            const double el = i > 0 ? sqrt(e2[i - 1]) : 0.0, er = i + 1 < n ? sqrt(e2[i]) : 0.0;
            lo = fmin(lo, d[i] - el - er);
            hi = fmax(hi, d[i] + el + er);
            emax = fmax(emax, e2[i]);
        }
        const double span = fmax(fabs(lo), fabs(hi));
        s_lo = lo - 1e-14 * span - 1e-300;
        s_hi = hi + 1e-14 * span + 1e-300;
    }
    __syncthreads();
    // Shared first level: the CTA's threads evaluate the count at blockDim.x points spaced geometrically over
    // [2^-30, 1] * (upper Gershgorin bound) -- one Sturm sequence each, the cost of one bisection step -- and every
    // eigenvalue then starts from the table cell that brackets it (~10 % wide) instead of the whole interval.
    // The bracket invariant count(lo) <= k < count(hi) is taken from the evaluated counts themselves.
    __shared__ double s_x[256];
    __shared__ int s_c[256];
    const int S = blockDim.x;
    {
        const double top = s_hi;
        const double x = top > 0.0 ? top * exp2(-30.0 * (double)(S - 1 - (int)threadIdx.x) / (double)max(S - 1, 1)) : top;
        s_x[threadIdx.x] = x;
        s_c[threadIdx.x] = top > 0.0 ? sturm_count(d, e2, n, x) : n;
    }
    __syncthreads();
    int it = 0;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        double lo = s_lo, hi = s_hi;
        {
            int a0 = 0, b0 = S - 1;                     // s_c[S-1] = count(upper bound) = n > k
            if (s_c[b0] > k) {
                while (a0 < b0) {                        // first index with count > k
                    const int m = (a0 + b0) >> 1;
                    if (s_c[m] > k) b0 = m; else a0 = m + 1;
                }
                while (b0 > 0 && s_c[b0 - 1] > k) --b0;  // rounding may break monotonicity: keep the invariant exact
                hi = s_x[b0];
                if (b0 > 0) lo = s_x[b0 - 1];
            }
        }
        const double atol = 1e-17 * fmax(fabs(s_lo), fabs(s_hi));
        for (it = 0; it < 200; ++it) {
            const double mid = 0.5 * (lo + hi);
            // the result is rounded to fp32 after a square root: 4e-10 relative on lambda is 3e-3 ulp of sqrt(lambda)
            if (hi - lo <= fmax(atol, 4e-10 * fmax(fabs(lo), fabs(hi))) || mid <= lo || mid >= hi) break;
            const int cnt = sturm_count(d, e2, n, mid);
            if (cnt > k) hi = mid; else lo = mid;  // the k-th smallest (0-based) is < mid iff count(mid) > k
        }
        const double lam = 0.5 * (lo + hi);
        const int pos = n - 1 - k;  // descending order
        if (pos < n_out) sv[(size_t)a * n_out + pos] = (float)sqrt(fmax(lam, 0.0));
    }
    if (threadIdx.x == 0 && iters_out) iters_out[a] = it;
}

cudaError_t launch_spectrum_tridiag(const double* G, int n_anchors, int n, int n_out, float* sv, int* iters,
                                    void* workspace, cudaStream_t stream) {
    if (n_anchors <= 0 || n <= 0) return cudaSuccess;
    cudaError_t e;
    if (n <= 225) {   // fused one-pass kernel (7 column slots per lane); measured slower than the four-pass kernel once
                      // the packed matrix no longer fits shared memory (n = 500: 23 vs 21 ms, n = 1024: 174 vs 129 ms)
        e = launch_tridiag_fused_small<7, true>(G, n_anchors, n, workspace, stream);
        if (e != cudaSuccess) return e;
    } else {
        const size_t need = tri_vec_bytes(n) + tri_packed_bytes(n);
        const int a_in_smem = need <= TRI_SMEM_MAX;
        const size_t smem = a_in_smem ? need : tri_vec_bytes(n);
        if (smem > TRI_SMEM_MAX) return cudaErrorInvalidValue;
        e = cudaFuncSetAttribute(tridiag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        tridiag_kernel<<<n_anchors, TRI_THREADS, smem, stream>>>(G, n, reinterpret_cast<uint8_t*>(workspace), a_in_smem);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    const size_t bsmem = 2 * (size_t)n * sizeof(double);
    if (bsmem > 48 * 1024) {
        e = cudaFuncSetAttribute(bisect_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bsmem);
        if (e != cudaSuccess) return e;
    }
    int threads = ((n + 31) / 32) * 32;
    if (threads > 256) threads = 256;
    bisect_kernel<<<n_anchors, threads, bsmem, stream>>>(reinterpret_cast<const uint8_t*>(workspace), n, n_out, sv,
                                                        iters);
    return cudaGetLastError();
}

}  // namespace idb
