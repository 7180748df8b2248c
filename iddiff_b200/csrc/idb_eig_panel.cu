// Panel (deferred-update) Householder tridiagonalisation of ONE large symmetric fp64 matrix by the whole GPU.
//
// The fused kernel of idb_eig_large.cu reads AND writes the whole trailing block once per eliminated column (16 m^2 bytes): a
// level-2 algorithm at its HBM roofline (5.1 TB/s at n = 12288).  This kernel cuts the traffic to 4.5 m^2 bytes per column:
//   * deferred updates, the way LAPACK's dlatrd does: the reflector pairs (v_p, w_p) of up to NB = 8 columns stay PENDING while
//     the trailing block in HBM is stale,  A_true = A_stale - sum_p (v_p w_p^T + w_p v_p^T);  a column's symmetric product is
//     taken with the stale block and corrected with 2 NB dot products,
//         A_true u = A_stale u - sum_p (v_p (w_p.u) + w_p (v_p.u)),
//     and only every NB-th column's sweep applies the pending pairs and writes the block back;
//   * symmetry: only the LOWER TRIANGLE is ever read or kept up to date.  Every element a_ij, i > j, serves both
//     (A u)[i] += a_ij u[j] (row side) and (A u)[j] += a_ij u[i] (column side) from one load.
// Per column: 7/8 x 4 m^2 (read-only sweeps) + 1/8 x 8 m^2 (flush sweeps: read + write) = 4.5 m^2 bytes.  Replaces
// torch.linalg.svd at dim_reduction.py:197-198 for image-sized score matrices (same role as tridiag_fused_cls_kernel; taken
// by launch_spectrum_large for even n >= 3584).  Measured (B200): n = 12288 0.69 s (fused kernel 1.83 s), 8192 0.30 s, 4096 77 ms.
//
// Work split per column k (cooperative launch, one CTA per SM, 8 warps, TWO grid synchronisations):
//   phase A  every CTA owns a contiguous SLICE of the indices j > k: sums the CTAs' column-side partials of the previous
//            sweep, finalises the previous column's w there (w = tau (A_stale u - corrections) + kk u) and writes its slice
//            of row k of A_true (read as column k of the lower triangle);                                -> grid sync
//   phase B  every CTA pulls the whole row into shared memory with one bulk copy, forms the norm, (tau, beta) and the
//            reflector u redundantly (u lives in shared memory only: 8 n bytes);
//   sweep    the CTA owns the rows i = b (mod G); its warps own column chunks of those rows (dealt in snake order so every
//            warp gets the same share of the triangle) and stream them as 4 KB tiles through private rings of 1-D bulk copies
//            (cp.async.bulk + mbarrier transaction counts, 3 stages per warp = 96 KB in flight per SM at no register cost).
//            Column-side sums stay in registers over all the CTA's rows of a chunk and go to part_col[b][.]; row-side sums
//            are reduced over the lanes eight rows at a time by a transposing butterfly.  A flush sweep additionally holds
//            the pending pairs' column sides of its chunk in registers (row sides in shared memory), brings every tile up
//            to date before using it and writes it back from registers.  Every sweep also produces the dot products the
//            next phase A needs (u.(A u), w_p.u, v_p.u).                                                 -> grid sync
// tools/panel_profile.sh builds the kernel with per-phase clock64 timers (-DTP_PROFILE).
#include "idb_common.cuh"
#include "idb_kernels.h"
#include <cooperative_groups.h>
#include <math.h>
#include <type_traits>

namespace cg = cooperative_groups;

namespace idb {

namespace {
constexpr int TP_THREADS = 256;
constexpr int TP_WARPS = TP_THREADS / 32;
constexpr int TP_NB = 8;       // pending reflector pairs before a flush
constexpr int TP_RMAX = 11;    // rows of the trailing block per warp (n <= TP_RMAX * gridDim.x * TP_WARPS = 13024 on 148 SMs)
constexpr int TP_CW = 256;     // columns per chunk of a read-only sweep
constexpr int TP_TR = 2;       // rows per tile there (one ring stage = TR rows x CW columns = 4 KB, two 2 KB bulk copies)
constexpr int TP_S = 3;        // ring stages per warp: 8 warps x 3 x 4 KB = 96 KB in flight per SM
constexpr int TP_TILE = TP_TR * TP_CW;
constexpr int TP_RPC = TP_RMAX * TP_WARPS;   // rows of the trailing block per CTA
constexpr int TP_NDOT = 2 * TP_NB + 1;
constexpr int TP_DC = (TP_NDOT + 1 + TP_WARPS - 1) / TP_WARPS;   // dot-product columns per warp in phase A (+ 1: column sum at k)
constexpr int TP_GL = 5;       // CTAs per lane there (grid <= 160)

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
constexpr size_t tp_smem_doubles(int n) {
    return (size_t)((n + 63) & ~63) + (size_t)TP_WARPS * TP_S * TP_TILE + (size_t)TP_WARPS * TP_RMAX * 2 * TP_NB +
           (size_t)TP_WARPS * TP_RPC;
}
}  // namespace

size_t tridiag_panel_workspace_doubles(int n, int grid) {
    return (size_t)(2 * TP_NB + 2) * n + (size_t)grid * TP_NDOT + 64 + (size_t)grid * ((n + 63) & ~63);
}

#ifdef TP_PROFILE   // phase timers of CTA 0 (tools/panel_profile.sh), written behind part_dot
#define TP_TICK_START() long long tp_t = clock64()
#define TP_TICK(i) do { long long tp_n = clock64(); tp_acc[i] += tp_n - tp_t; tp_t = tp_n; } while (0)
#else
#define TP_TICK_START() do {} while (0)
#define TP_TICK(i) do {} while (0)
#endif

__global__ void __launch_bounds__(TP_THREADS, 1) tridiag_panel_kernel(double* __restrict__ A, int n, double* __restrict__ d,
                                                                      double* __restrict__ e, double* __restrict__ ws) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(128) double tp_smem[];
    const int nu = (n + 63) & ~63;
    double* us = tp_smem;                                   // [nu] the WHOLE reflector of the current column (0 for j <= k)
    double* ring = us + nu;                                 // [WARPS][S][CH] bulk-copy ring of the read-only sweeps ...
                                                            // ... aliased by 3 stages of [2 NB][LB] in the flush sweeps
    double* rowsc = ring + TP_WARPS * TP_S * TP_TILE;       // [WARPS][RMAX][2 NB]: v_p[i], w_p[i] of the warp's rows
    double* rowacc = rowsc + TP_WARPS * TP_RMAX * 2 * TP_NB; // [WARPS][RPC] row sums of a read-only sweep, per warp (= per column set)
    __shared__ __align__(8) uint64_t full[TP_WARPS * TP_S];
    __shared__ __align__(8) uint64_t xbar;
    __shared__ double red[TP_WARPS];
    __shared__ double s_dot[TP_NDOT + 1];                   // pc.u, (w_p.u)_p, (v_p.u)_p of the previous sweep; [NDOT]: the column
                                                            // side of (A u)[k]
    __shared__ double s_vk[TP_NB + 1], s_wk[TP_NB + 1];     // v_p[k], w_p[k]
    __shared__ double s_part[TP_WARPS * TP_NDOT];

    const int G = gridDim.x, b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_gw = G * TP_WARPS, gw = warp * G + b;       // rows cyclic over the CTAs first: a short tail still uses every SM
    double* Vbuf = ws;                                      // [NB][n]
    double* Wbuf = Vbuf + (size_t)TP_NB * n;                // [NB][n]
    double* xbuf = Wbuf + (size_t)TP_NB * n;                // [n] row k of A_true, written in slices
    double* pc = xbuf + n;                                  // [n] A u of the current column (stale or true, unscaled)
    double* part_dot = pc + n;                              // [NDOT][G] (+ 64 spare)
    double* part_col = part_dot + (((size_t)G * TP_NDOT + 64 + 1) & ~(size_t)1);   // [G][nu] column-side partial products of a
                                                                                    // read-only sweep, one row per CTA

    for (int q = tid; q < nu; q += TP_THREADS) us[q] = 0.0;
    if (tid == 0) {
        for (int q = 0; q < TP_WARPS * TP_S; ++q) mbar_init(&full[q], 1);
        mbar_init(&xbar, 1);
        fence_barrier_init();
    }
    fence_proxy_async();
    __syncthreads();
    double* ring_w = ring + (size_t)warp * TP_S * TP_TILE;
    uint64_t* full_w = full + warp * TP_S;
    int stage = 0;             // the warp's ring position and phase bit run on across sweeps (strict round-robin)
    uint32_t phase = 0, xphase = 0;

    // replicated scalar state (identical in every CTA)
    int r = 0;                 // pending pairs whose w is final (slots 0 .. r-1)
    bool have_prev = false;    // the previous column produced a reflector (still in `us`) whose w goes to slot r now
    bool prev_flushed = false; // ... and its sweep applied the pending pairs (its A u is already the true one)
    double tau_prev = 0.0;

#ifdef TP_PROFILE
    long long tp_acc[7] = {0, 0, 0, 0, 0, 0, 0};
#endif
    for (int k = 0; k + 1 < n; ++k) {
        TP_TICK_START();
        const bool last = (k + 2 >= n);                     // k = n - 2: only finalise + read the last 2 x 2 block
        // ---------------- phase A: this CTA's slice of the indices j > k ----------------
        const int sl = (n - k - 1 + G - 1) / G;              // contiguous slices: coalesced reads of the G column partials
        const int j = k + 1 + b * sl + tid;
        const bool mine = tid < sl && j < n;
        const bool sym_prev = have_prev;                     // every sweep walks the lower triangle only: (A u)[j] =
                                                             // pc[j] (row side) + sum over the CTAs of part_col[.][j]
        // every global load of the phase is issued up front: ONE L2 round trip.  (The pairs outstanding during the previous
        // sweep are exactly the r final ones unless that sweep flushed them, so w_new and x share the same loads.)
        double pd[TP_DC][TP_GL];                             // the CTAs' partial dot products of the previous sweep
#pragma unroll
        for (int ci = 0; ci < TP_DC; ++ci)
#pragma unroll
            for (int i = 0; i < TP_GL; ++i) {
                const int c = warp + TP_WARPS * ci, q = lane + 32 * i;
                pd[ci][i] = 0.0;
                if (have_prev && q < G) {
                    if (c < TP_NDOT) pd[ci][i] = part_dot[(size_t)c * G + q];
                    else if (c == TP_NDOT && sym_prev) pd[ci][i] = part_col[(size_t)q * nu + k];
                }
            }
        double akj = 0.0, pcj = 0.0, upj = 0.0, pck = 0.0, vj[TP_NB], wj[TP_NB];
#pragma unroll
        for (int p = 0; p < TP_NB; ++p) { vj[p] = 0.0; wj[p] = 0.0; }
        if (mine) {
            akj = A[(size_t)j * n + k];                      // column k of the lower triangle (the only one kept up to date)
            if (have_prev) { pcj = pc[j]; upj = us[j]; }
#pragma unroll
            for (int p = 0; p < TP_NB; ++p)
                if (p < r) { vj[p] = Vbuf[(size_t)p * n + j]; wj[p] = Wbuf[(size_t)p * n + j]; }
        }
        if (have_prev) pck = pc[k];
        if (sym_prev) {
            // column side of (A u)[j] over the slice: warp w sums the partials of the CTAs w, w + 8, ... (all loads independent)
            constexpr int JL = (TP_RPC + 31) / 32;
            double acc[JL];
#pragma unroll
            for (int e = 0; e < JL; ++e) acc[e] = 0.0;
            const int j0 = k + 1 + b * sl;
            constexpr int QL = (32 * TP_GL) / TP_WARPS;      // CTAs per warp
            double pv[QL][JL];                               // every load is issued before the first add (clamped addresses,
#pragma unroll                                               // masked values): one L2 round trip
            for (int qi = 0; qi < QL; ++qi) {
                const int q = min(warp + TP_WARPS * qi, G - 1);
#pragma unroll
                for (int e = 0; e < JL; ++e) pv[qi][e] = part_col[(size_t)q * nu + min(j0 + lane + 32 * e, n - 1)];
            }
#pragma unroll
            for (int qi = 0; qi < QL; ++qi) {
                const bool qok = warp + TP_WARPS * qi < G;
#pragma unroll
                for (int e = 0; e < JL; ++e) acc[e] += (qok && lane + 32 * e < sl) ? pv[qi][e] : 0.0;
            }
#pragma unroll
            for (int e = 0; e < JL; ++e)
                if (lane + 32 * e < TP_RPC) rowacc[warp * TP_RPC + lane + 32 * e] = acc[e];
        }
        if (tid < r) { s_vk[tid] = Vbuf[(size_t)tid * n + k]; s_wk[tid] = Wbuf[(size_t)tid * n + k]; }
        if (have_prev) {
#pragma unroll
            for (int ci = 0; ci < TP_DC; ++ci) {
                double t = 0.0;
#pragma unroll
                for (int i = 0; i < TP_GL; ++i) t += pd[ci][i];
                t = tp_warp_sum(t);
                if (lane == 0 && warp + TP_WARPS * ci <= TP_NDOT) s_dot[warp + TP_WARPS * ci] = t;
            }
        }
        __syncthreads();
        double kk = 0.0, wk_new = 0.0;
        const double upk = have_prev ? us[k] : 0.0;          // = 1: the previous reflector's leading element
        if (have_prev) {
            double corr = 0.0;
            if (!prev_flushed)
                for (int p = 0; p < r; ++p) corr += s_dot[1 + p] * s_dot[1 + TP_NB + p];
            kk = -0.5 * tau_prev * tau_prev * (s_dot[0] - 2.0 * corr);
            // w_new[k] of the previous column: tau (A_true u)[k] + kk u[k], redundantly in every thread
            double base = pck + s_dot[TP_NDOT];
            if (!prev_flushed)
                for (int p = 0; p < r; ++p) base -= s_vk[p] * s_dot[1 + p] + s_wk[p] * s_dot[1 + TP_NB + p];
            wk_new = fma(kk, upk, tau_prev * base);
        }
        const int slot = r;
        const int r_now = r + (have_prev ? 1 : 0);
        double x = 0.0, wj_new = 0.0;
        if (mine) {
            if (sym_prev) {
#pragma unroll
                for (int w2 = 0; w2 < TP_WARPS; ++w2) pcj += rowacc[w2 * TP_RPC + tid];
            }
            if (have_prev) {
                double base = pcj;
                if (!prev_flushed) {
#pragma unroll
                    for (int p = 0; p < TP_NB; ++p)
                        if (p < r) base -= vj[p] * s_dot[1 + p] + wj[p] * s_dot[1 + TP_NB + p];
                }
                wj_new = fma(kk, upj, tau_prev * base);
                Vbuf[(size_t)slot * n + j] = upj;
                Wbuf[(size_t)slot * n + j] = wj_new;
            }
            x = akj;
#pragma unroll
            for (int p = 0; p < TP_NB; ++p)
                if (p < r) x -= s_vk[p] * wj[p] + s_wk[p] * vj[p];
            if (have_prev) x -= upk * wj_new + wk_new * upj;
            xbuf[j] = x;
            fence_proxy_async();                             // read by the other CTAs' bulk copies after the grid sync
        }
        if (b == 0 && tid == 0) {
            double dk = A[(size_t)k * n + k];
            for (int p = 0; p < r; ++p) dk -= 2.0 * s_vk[p] * s_wk[p];
            if (have_prev) dk -= 2.0 * upk * wk_new;
            d[k] = dk;
        }
        if (last) {
            if (mine && j == n - 1) {                        // thread 0 of CTA 0: x = A_true[n-2][n-1]
                e[n - 2] = x;
                e[n - 1] = 0.0;
                double dl = A[(size_t)(n - 1) * n + (n - 1)];
#pragma unroll
                for (int p = 0; p < TP_NB; ++p)
                    if (p < r) dl -= 2.0 * vj[p] * wj[p];
                if (have_prev) dl -= 2.0 * upj * wj_new;
                d[n - 1] = dl;
            }
            break;
        }
        r = r_now;                                          // the previous pair is final now (slots 0 .. r-1)
        TP_TICK(0);
        grid.sync();
        TP_TICK(1);
        // ---------------- phase B: the whole row -> norm -> reflector, redundantly in every CTA ----------------
        int first = gw;
        if (first < k + 1) first += ((k + 1 - first + n_gw - 1) / n_gw) * n_gw;
        int nrows = 0;
        if (first < n) nrows = (n - 1 - first) / n_gw + 1;
        double* rs = rowsc + (size_t)warp * TP_RMAX * 2 * TP_NB;
        const int rs_cnt = nrows * 2 * r;
        // the row x: ONE bulk copy straight into `us` (from the even index at or below k + 1)
        const int xs0 = (k + 1) & ~1;
        if (tid == 0) {
            const uint32_t bytes = (uint32_t)(n - xs0) * 8u;
            mbar_arrive_expect_tx(&xbar, bytes);
            tma_bulk_g2s(us + xs0, xbuf + xs0, bytes, &xbar);
        }
        // the row sides v_p[i], w_p[i] of this warp's rows, under the latency of that copy
        constexpr int RSL = (TP_RMAX * 2 * TP_NB + 31) / 32;
        double rs_val[RSL];
#pragma unroll
        for (int qq = 0; qq < RSL; ++qq) {
            const int q = lane + 32 * qq;
            rs_val[qq] = 0.0;
            if (q < rs_cnt) {
                const int t = q / (2 * r), c = q % (2 * r);
                const int i = first + t * n_gw;
                rs_val[qq] = c < r ? Vbuf[(size_t)c * n + i] : Wbuf[(size_t)(c - r) * n + i];
            }
        }
#pragma unroll
        for (int qq = 0; qq < RSL; ++qq) {
            const int q = lane + 32 * qq;
            if (q < rs_cnt) {
                const int t = q / (2 * r), c = q % (2 * r);
                rs[t * 2 * TP_NB + (c < r ? c : TP_NB + (c - r))] = rs_val[qq];      // v_p[i] at [p], w_p[i] at [NB + p]
            }
        }
        mbar_wait(&xbar, xphase);
        xphase ^= 1u;
        double ssp = 0.0;
        for (int q = k + 2 + tid; q < n; q += TP_THREADS) {
            const double xv = us[q];
            ssp = fma(xv, xv, ssp);
        }
        const double alpha = us[k + 1];
        const double ss = tp_block_sum(ssp, red);            // (its barriers also publish `rs`)
        if (tid == 0) us[k] = 0.0;                           // u is zero up to k
        double tau = 0.0, beta = alpha;
        if (ss > 0.0) {
            const double nrm = sqrt(fma(alpha, alpha, ss));
            beta = alpha >= 0.0 ? -nrm : nrm;
            tau = (beta - alpha) / beta;
            const double inv = 1.0 / (alpha - beta);
            for (int q = k + 1 + tid; q < n; q += TP_THREADS) us[q] = (q == k + 1) ? 1.0 : us[q] * inv;
        }
        if (b == 0 && tid == 0) e[k] = beta;
        fence_proxy_async();                                 // the next column's bulk copy overwrites these generic writes
        __syncthreads();
        TP_TICK(2);
        // ---------------- sweep ----------------
        const bool flush = (r == TP_NB);
        if (tau != 0.0) {
            double dot0 = 0.0;   // sum_i (A u)[i] u[i] over the warp's rows (identical in every lane)
            double dotl = 0.0;   // lane c < 2 NB: sum_i rs[.][c] u[i]  (v_p.u at c = p, w_p.u at c = NB + p)
            const bool dot_lane = lane < 2 * TP_NB && (lane & (TP_NB - 1)) < r;
            // One sweep over the LOWER TRIANGLE of the trailing block (the block is symmetric: every element a_ij, i > j,
            // serves (A u)[i] += a_ij u[j] and (A u)[j] += a_ij u[i]): 4 m^2 bytes read instead of 8 m^2, and only that
            // triangle is ever kept up to date.  The CTA owns the rows i = b (mod G); its warps own CW-column chunks of those
            // rows (dealt in snake order, so every warp gets the same share of the triangle) and stream them as 4 KB tiles
            // (TR rows x CW columns, one bulk copy per row) through their rings.  Column-side sums stay in registers over all
            // the CTA's rows of a chunk and go to part_col[b][.] (summed over the CTAs by the next phase A); row-side sums
            // are reduced over the lanes eight rows at a time by a transposing butterfly and collected per warp.
            //   read-only sweep (seven of eight): CW = 256, the stale block is only read;
            //   flush sweep: CW = 128, the pending pairs' column sides of the chunk sit in registers, their row sides in
            //   shared memory; every tile is brought up to date, written back from registers, and then used.
            auto sweep = [&](auto cw_tag, auto flush_tag) {
                constexpr int CW = decltype(cw_tag)::value;
                constexpr bool FL = decltype(flush_tag)::value;
                constexpr int TR = TP_TILE / CW, NE = CW / 64;
                const int Q = (n - 1 - b) / G + 1;           // the CTA's rows i_q = b + G q < n
                const int q0 = (b >= k + 1) ? 0 : (k + 1 - b + G - 1) / G;
                const int c_lo = (k + 1) / CW, c_hi = (n - 1) / CW;
                double* rowacc_w = rowacc + warp * TP_RPC;
                for (int q = lane; q < TP_RPC; q += 32) rowacc_w[q] = 0.0;
                auto chunk_of = [&](int m) {                 // m-th chunk of this warp: w, 15 - w, 16 + w, 31 - w, ...
                    return c_lo + 2 * TP_WARPS * (m >> 1) + ((m & 1) ? 2 * TP_WARPS - 1 - warp : warp);
                };
                auto first_row = [&](int c) {                // first row of the CTA that reaches chunk c (i >= CW c, i >= k + 1)
                    const int lo = CW * c;
                    const int qc = (lo <= b) ? 0 : (lo - b + G - 1) / G;
                    return max(qc, q0);
                };
                int p_m = 0, p_c = chunk_of(0), p_q = 0;     // producer cursor (chunk, first row of the tile), lane 0 only
                bool p_done = p_c > c_hi;
                if (!p_done) { p_q = first_row(p_c); p_done = p_q >= Q; }
                auto issue = [&](int st_i) {
                    const int nr = min(TR, Q - p_q);
                    const int col = CW * p_c;
                    const uint32_t bytes = (uint32_t)min(CW, n - col) * 8u;
                    mbar_arrive_expect_tx(&full_w[st_i], bytes * nr);
                    const double* src = A + (size_t)(b + G * p_q) * n + col;
                    for (int rr = 0; rr < nr; ++rr)
                        tma_bulk_g2s(ring_w + st_i * TP_TILE + rr * CW, src + (size_t)rr * G * n, bytes, &full_w[st_i]);
                    p_q += TR;
                    if (p_q >= Q) {                          // next chunk of this warp (later chunks reach fewer rows)
                        p_c = chunk_of(++p_m);
                        if (p_c > c_hi) p_done = true;
                        else { p_q = first_row(p_c); p_done = p_q >= Q; }
                    }
                };
                if (lane == 0) {
                    int st_i = stage;
                    for (int q = 0; q < TP_S && !p_done; ++q) {
                        issue(st_i);
                        if (++st_i == TP_S) st_i = 0;
                    }
                }
                __syncwarp();
                double dotc = 0.0;                           // sum_j colsum[j] u[j] over this lane's columns
                for (int m = 0;; ++m) {
                    const int c = chunk_of(m);
                    if (c > c_hi) break;
                    const int col0 = CW * c + 2 * lane;      // this lane's columns: col0 + 64 e, + 1
                    const int q_first = first_row(c);
                    double2 uj[NE], cs[NE];
                    double2 vj[FL ? TP_NB : 1][NE], wj[FL ? TP_NB : 1][NE];   // flush: the pairs' column sides of this chunk
#pragma unroll
                    for (int e = 0; e < NE; ++e) {
                        const bool in = col0 + 64 * e < nu;
                        uj[e] = in ? *reinterpret_cast<const double2*>(us + col0 + 64 * e) : make_double2(0.0, 0.0);   // 0 up to k, beyond n
                        cs[e] = make_double2(0.0, 0.0);
                        if (FL) {
#pragma unroll
                            for (int p = 0; p < TP_NB; ++p) {
                                const bool ld = q_first < Q && col0 + 64 * e < n;
                                vj[p][e] = ld ? *reinterpret_cast<const double2*>(Vbuf + (size_t)p * n + col0 + 64 * e) : make_double2(0.0, 0.0);
                                wj[p][e] = ld ? *reinterpret_cast<const double2*>(Wbuf + (size_t)p * n + col0 + 64 * e) : make_double2(0.0, 0.0);
                            }
                        }
                    }
                    for (int q = q_first; q < Q; q += 8) {
                        double pr[8];
#pragma unroll
                        for (int h = 0; h < 8 / TR; ++h) {
                            if (q + TR * h < Q) {            // (warp-uniform) one tile = rows q + TR h ... + TR - 1
                                mbar_wait(&full_w[stage], phase);
                                const double* src = ring_w + stage * TP_TILE + 2 * lane;
#pragma unroll
                                for (int r2 = 0; r2 < TR; ++r2) {
                                    const int rr = TR * h + r2;
                                    double acc = 0.0;
                                    if (q + rr < Q) {
                                        const int i = b + G * (q + rr);
                                        const double ui = us[i];
                                        double2 a[NE];
#pragma unroll
                                        for (int e = 0; e < NE; ++e) a[e] = *reinterpret_cast<const double2*>(src + r2 * CW + 64 * e);
                                        if (FL) {
                                            // row sides of CTA row q + rr: warp (q + rr) mod 8 loaded them in phase B at its slot (q + rr - q0) / 8
                                            const double* rsq = rowsc + (((q + rr) & (TP_WARPS - 1)) * TP_RMAX + ((q + rr - q0) >> 3)) * 2 * TP_NB;
#pragma unroll
                                            for (int p = 0; p < TP_NB; ++p) {
                                                const double vi = rsq[p], wi = rsq[TP_NB + p];
#pragma unroll
                                                for (int e = 0; e < NE; ++e) {
                                                    a[e].x = a[e].x - vi * wj[p][e].x - wi * vj[p][e].x;
                                                    a[e].y = a[e].y - vi * wj[p][e].y - wi * vj[p][e].y;
                                                }
                                            }
#pragma unroll
                                            for (int e = 0; e < NE; ++e)
                                                if (col0 + 64 * e < n) *reinterpret_cast<double2*>(A + (size_t)i * n + col0 + 64 * e) = a[e];
                                        }
#pragma unroll
                                        for (int e = 0; e < NE; ++e) {
                                            const int col = col0 + 64 * e;
                                            // row side: lower triangle + diagonal; column side: strictly lower
                                            acc = fma(col + 1 <= i ? a[e].y : 0.0, uj[e].y, fma(col <= i ? a[e].x : 0.0, uj[e].x, acc));
                                            cs[e].x = fma(col < i ? a[e].x : 0.0, ui, cs[e].x);
                                            cs[e].y = fma(col + 1 < i ? a[e].y : 0.0, ui, cs[e].y);
                                        }
                                    }
                                    pr[rr] = acc;
                                }
                                __syncwarp();
                                if (lane == 0 && !p_done) issue(stage);
                                if (++stage == TP_S) { stage = 0; phase ^= 1u; }
                            } else {
#pragma unroll
                                for (int r2 = 0; r2 < TR; ++r2) pr[TR * h + r2] = 0.0;
                            }
                        }
                        // transposing butterfly: 8 row sums over 32 lanes in 4 + 2 + 1 + 2 shuffles; lane 4 g holds row g
                        const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
                        double t4[4], t2[2], t1;
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const double keep = h16 ? pr[e + 4] : pr[e], send = h16 ? pr[e] : pr[e + 4];
                            t4[e] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                        }
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const double keep = h8 ? t4[e + 2] : t4[e], send = h8 ? t4[e] : t4[e + 2];
                            t2[e] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                        }
                        {
                            const double keep = h4 ? t2[1] : t2[0], send = h4 ? t2[0] : t2[1];
                            t1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                        }
                        t1 += __shfl_xor_sync(0xffffffffu, t1, 2);
                        t1 += __shfl_xor_sync(0xffffffffu, t1, 1);
                        const int rg = (h16 ? 4 : 0) + (h8 ? 2 : 0) + (h4 ? 1 : 0);
                        if ((lane & 3) == 0 && q + rg < Q) rowacc_w[q + rg] += t1;
                    }
#pragma unroll
                    for (int e = 0; e < NE; ++e) {
                        if (col0 + 64 * e < nu) *reinterpret_cast<double2*>(part_col + (size_t)b * nu + col0 + 64 * e) = cs[e];
                        dotc = fma(cs[e].x, uj[e].x, fma(cs[e].y, uj[e].y, dotc));
                    }
                }
                if (FL) fence_proxy_async();                 // the next sweeps' bulk copies read what this one stored
                __syncthreads();                             // every warp's row sums are complete
                double mine_dot = dotc;
                if (tid >= q0 && tid < Q) {                  // one CTA row per thread: row side of (A u)[i]
                    double rsum = 0.0;
#pragma unroll
                    for (int w2 = 0; w2 < TP_WARPS; ++w2) rsum += rowacc[w2 * TP_RPC + tid];
                    const int i = b + G * tid;
                    pc[i] = rsum;
                    mine_dot = fma(rsum, us[i], mine_dot);
                }
                dot0 = tp_warp_sum(mine_dot);
            };
            if (!flush) {
                sweep(std::integral_constant<int, 256>{}, std::false_type{});
                if (dot_lane)
                    for (int t = 0; t < nrows; ++t) dotl = fma(rs[t * 2 * TP_NB + lane], us[first + t * n_gw], dotl);
            } else {
                sweep(std::integral_constant<int, 128>{}, std::true_type{});
            }
            TP_TICK(flush ? 6 : 3);
            // the dot products of the next phase A: pc.u, (w_p.u)_p at [1 + p], (v_p.u)_p at [1 + NB + p]
            if (lane == 0) s_part[warp * TP_NDOT] = dot0;
            if (lane < 2 * TP_NB) s_part[warp * TP_NDOT + 1 + (lane < TP_NB ? TP_NB + lane : lane - TP_NB)] = dotl;
            __syncthreads();
            if (tid < TP_NDOT) {
                double t = 0.0;
                for (int w2 = 0; w2 < TP_WARPS; ++w2) t += s_part[w2 * TP_NDOT + tid];
                part_dot[(size_t)tid * G + b] = t;
            }
            have_prev = true;
            prev_flushed = flush;
            tau_prev = tau;
            if (flush) r = 0;
        } else {
            have_prev = false;
        }
        TP_TICK(4);
        grid.sync();
        TP_TICK(5);
    }
#ifdef TP_PROFILE
    if (b == 0 && tid == 0)
        for (int i = 0; i < 7; ++i) part_dot[(size_t)G * TP_NDOT + i] = (double)tp_acc[i];
#endif
}

// even n (16-byte bulk copies), every row owned (RMAX rows per warp), one slice element and <= GL partial sums per thread
bool tridiag_panel_supported(int n, int num_sms) {
    return n >= 4 && (n & 1) == 0 && (long long)n <= (long long)TP_RMAX * num_sms * TP_WARPS &&
           (n + num_sms - 1) / num_sms <= TP_THREADS && num_sms <= 32 * TP_GL && tp_smem_doubles(n) * sizeof(double) <= 225 * 1024;
}

cudaError_t launch_tridiag_panel(double* G, int n, double* d, double* e, double* ws, int num_sms, cudaStream_t stream) {
    if (!tridiag_panel_supported(n, num_sms)) return cudaErrorInvalidValue;
    const size_t smem = tp_smem_doubles(n) * sizeof(double);
    cudaError_t err = cudaFuncSetAttribute(tridiag_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tridiag_panel_kernel, TP_THREADS, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) return cudaErrorInvalidValue;
    void* args[] = {&G, &n, &d, &e, &ws};
    return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(tridiag_panel_kernel), dim3(num_sms), dim3(TP_THREADS), args, smem,
                                       stream);
}

}  // namespace idb
