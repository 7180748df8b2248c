// HBM-bound passes of the U-Net score networks (reference: models/ddpm.py, models/layers.py, models/ncsnpp.py,
// models/BeatGANsblocks.py).  Activations live in HBM as fp32 NHWC; every convolution operand is produced by ONE
// "operand preparation" pass that fuses everything the reference runs as separate elementwise / resampling kernels
// between two convolutions:
//
//   concat of the skip tensor (ddpm.py:198) -> GroupNorm apply (layers.py:638,648) -> SiLU (:660,666)
//   -> nearest x2 up-sampling (layers.py:600) | stride-2 parity split (layers.py:621-623) | 2x2 average pool
//   -> hi/lo fp16 split (x 2^k) in the NHWC layout the implicit-GEMM convolution reads through TMA.
//
// GroupNorm statistics are one extra read of the tensor (fp64 accumulation).  Attention at 16 x 16 (layers.py:567-590)
// is < 1 % of the network's FLOPs and is done in fp32 on the CUDA cores, scores resident in shared memory.
#include "idb_common.cuh"
#include "idb_kernels.h"

namespace idb {

// ---------------------------------- GroupNorm statistics -------------------------------------
// sums[n][g] = (sum x, sum x^2) over the group's channels and all pixels, x = concat(src0, src1) along C.
__global__ void __launch_bounds__(256) gn_stats_kernel(const float* __restrict__ s0, int C0,
                                                       const float* __restrict__ s1, int C1, int HW, int groups,
                                                       int CL, double* __restrict__ sums) {
    extern __shared__ double gs[];  // [groups][2]
    const int C = C0 + C1, cg = C / groups;
    const int n = blockIdx.x;
    for (int i = threadIdx.x; i < 2 * groups; i += blockDim.x) gs[i] = 0.0;
    __syncthreads();
    const int cl = threadIdx.x % CL, pl = threadIdx.x / CL, PL = blockDim.x / CL;
    const int per = (HW + gridDim.y - 1) / gridDim.y;
    const int p0 = blockIdx.y * per, p1 = min(HW, p0 + per);
    for (int c = cl; c < C; c += CL) {
        const float* src = c < C0 ? s0 + (size_t)n * HW * C0 + c : s1 + (size_t)n * HW * C1 + (c - C0);
        const int ld = c < C0 ? C0 : C1;
        double a = 0.0, b = 0.0;
        for (int px = p0 + pl; px < p1; px += PL) {
            const double v = (double)__ldg(src + (size_t)px * ld);
            a += v;
            b += v * v;
        }
        atomicAdd(&gs[2 * (c / cg)], a);
        atomicAdd(&gs[2 * (c / cg) + 1], b);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * groups; i += blockDim.x) atomicAdd(&sums[(size_t)n * 2 * groups + i], gs[i]);
}

// float4 variant (C0 % 4 == 0 and C1 % 4 == 0): one thread = 4 channels, pixels strided over the block
__global__ void __launch_bounds__(256) gn_stats_vec_kernel(const float* __restrict__ s0, int C0,
                                                           const float* __restrict__ s1, int C1, int HW, int groups,
                                                           int QL, double* __restrict__ sums) {
    extern __shared__ double gs[];  // [groups][2]
    const int C = C0 + C1, cg = C / groups;
    const int n = blockIdx.x;
    for (int i = threadIdx.x; i < 2 * groups; i += blockDim.x) gs[i] = 0.0;
    __syncthreads();
    const int ql = threadIdx.x % QL, pl = threadIdx.x / QL, PL = blockDim.x / QL;
    const int per = (HW + gridDim.y - 1) / gridDim.y;
    const int p0 = blockIdx.y * per, p1 = min(HW, p0 + per);
    for (int q = ql; q < (C >> 2); q += QL) {
        const int c = q << 2;
        const float* src = c < C0 ? s0 + (size_t)n * HW * C0 + c : s1 + (size_t)n * HW * C1 + (c - C0);
        const int ld = c < C0 ? C0 : C1;
        double a[4] = {0.0, 0.0, 0.0, 0.0}, b[4] = {0.0, 0.0, 0.0, 0.0};
        for (int px = p0 + pl; px < p1; px += PL) {
            const float4 f = __ldg(reinterpret_cast<const float4*>(src + (size_t)px * ld));
            const double v0 = f.x, v1 = f.y, v2 = f.z, v3 = f.w;
            a[0] += v0; b[0] = fma(v0, v0, b[0]);
            a[1] += v1; b[1] = fma(v1, v1, b[1]);
            a[2] += v2; b[2] = fma(v2, v2, b[2]);
            a[3] += v3; b[3] = fma(v3, v3, b[3]);
        }
        if ((cg & 3) == 0) {  // the four channels share a group
            atomicAdd(&gs[2 * (c / cg)], (a[0] + a[1]) + (a[2] + a[3]));
            atomicAdd(&gs[2 * (c / cg) + 1], (b[0] + b[1]) + (b[2] + b[3]));
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                atomicAdd(&gs[2 * ((c + j) / cg)], a[j]);
                atomicAdd(&gs[2 * ((c + j) / cg) + 1], b[j]);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * groups; i += blockDim.x) atomicAdd(&sums[(size_t)n * 2 * groups + i], gs[i]);
}

// (mean, rstd) in fp32 from the fp64 sums: var = E[x^2] - mean^2 (biased, as torch GroupNorm), rstd = 1/sqrt(var+eps)
__global__ void gn_finalize_kernel(const double* __restrict__ sums, int total, double count, double eps,
                                   float* __restrict__ mr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const double m = sums[2 * i] / count;
    double var = sums[2 * i + 1] / count - m * m;
    if (var < 0.0) var = 0.0;
    mr[2 * i] = (float)m;
    mr[2 * i + 1] = (float)(1.0 / sqrt(var + eps));
}

cudaError_t launch_gn_stats(const float* s0, int C0, const float* s1, int C1, int N, int HW, int groups, double eps,
                            double* sums, float* mean_rstd, int num_sms, cudaStream_t stream) {
    if (N <= 0) return cudaSuccess;
    const int C = C0 + C1;
    cudaError_t e = cudaMemsetAsync(sums, 0, (size_t)N * groups * 2 * sizeof(double), stream);
    if (e != cudaSuccess) return e;
    int CL = 32;
    while (CL < C && CL < 256) CL <<= 1;
    int slices = (4 * num_sms + N - 1) / N;
    const int max_slices = (HW + 15) / 16;
    if (slices > max_slices) slices = max_slices;
    if (slices < 1) slices = 1;
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if (C0 % 4 == 0 && C1 % 4 == 0 && al16(s0) && (!s1 || al16(s1))) {
        int QL = 8;
        while (QL < (C >> 2) && QL < 256) QL <<= 1;
        gn_stats_vec_kernel<<<dim3(N, slices), 256, 2 * groups * sizeof(double), stream>>>(s0, C0, s1, C1, HW, groups, QL,
                                                                                          sums);
    } else {
        gn_stats_kernel<<<dim3(N, slices), 256, 2 * groups * sizeof(double), stream>>>(s0, C0, s1, C1, HW, groups, CL,
                                                                                      sums);
    }
    const int total = N * groups;
    gn_finalize_kernel<<<(total + 255) / 256, 256, 0, stream>>>(sums, total, (double)(C / groups) * HW, eps, mean_rstd);
    return cudaGetLastError();
}

// GroupNorm statistics from the per-(image, slot, channel) partial sums the convolution epilogue wrote
// (idb_conv.cu, ConvParams::stats): one warp per (image, group) adds the (sum, sum of squares) pairs of the group's
// channels over all slots in fp64, in a fixed order (bit-reproducible), over the virtual concat [st0 | st1].
__global__ void __launch_bounds__(128) gn_from_partials_kernel(const float2* __restrict__ st0, int C0, int S0,
                                                               const float2* __restrict__ st1, int C1, int S1, int N,
                                                               int groups, double count, double eps,
                                                               float* __restrict__ mr) {
    const int wid = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (wid >= N * groups) return;
    const int img = wid / groups, g = wid % groups;
    const int cg = (C0 + C1) / groups;
    double s = 0.0, q = 0.0;
    for (int j = 0; j < cg; ++j) {
        const int c = g * cg + j;
        const float2* base;
        int S, C, cc;
        if (c < C0) { base = st0; S = S0; C = C0; cc = c; }
        else { base = st1; S = S1; C = C1; cc = c - C0; }
        const float2* ptr = base + (size_t)img * S * C + cc;
        for (int k = lane; k < S; k += 32) {
            const float2 v = __ldg(ptr + (size_t)k * C);
            s += (double)v.x;
            q += (double)v.y;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (lane == 0) {
        const double m = s / count;
        double var = q / count - m * m;
        if (var < 0.0) var = 0.0;
        mr[2 * wid] = (float)m;
        mr[2 * wid + 1] = (float)(1.0 / sqrt(var + eps));
    }
}

cudaError_t launch_gn_from_partials(const float* st0, int C0, int S0, const float* st1, int C1, int S1, int N, int HW,
                                    int groups, double eps, float* mean_rstd, cudaStream_t stream) {
    if (N <= 0) return cudaSuccess;
    const int warps = N * groups;
    gn_from_partials_kernel<<<(warps + 3) / 4, 128, 0, stream>>>(reinterpret_cast<const float2*>(st0), C0, S0,
                                                                 reinterpret_cast<const float2*>(st1), C1, S1, N, groups,
                                                                 (double)((C0 + C1) / groups) * HW, eps, mean_rstd);
    return cudaGetLastError();
}

// ---------------------------------- operand preparation --------------------------------------
// one thread = one output pixel x one 8-channel chunk (16 B of hi and 16 B of lo).  VEC: the chunk lies inside one NHWC
// source and is 32-byte aligned -> two float4 loads; otherwise a scalar gather (NCHW network input, odd channel counts).
struct PrepChunk {
    float g[8], b[8];   // GroupNorm / affine scale and shift of the 8 channels
    float m[8], r[8];   // mean and rstd of the 8 channels' groups
};

template <bool VEC>
__device__ __forceinline__ void prep_load8(const PrepParams& p, int n, int y, int x, int c0, int C, float (&v)[8]) {
    if (VEC) {
        const float* src = c0 < p.C0 ? p.src0 + (((size_t)n * p.Hs + y) * p.Ws + x) * p.C0 + c0
                                     : p.src1 + (((size_t)n * p.Hs + y) * p.Ws + x) * p.C1 + (c0 - p.C0);
        const float4 a = __ldg(reinterpret_cast<const float4*>(src));
        const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = c0 + j;
            float o = 0.f;
            if (c < C) {
                if (p.src_nchw) o = __ldg(p.src0 + (((size_t)n * p.C0 + c) * p.Hs + y) * p.Ws + x);
                else if (c < p.C0) o = __ldg(p.src0 + (((size_t)n * p.Hs + y) * p.Ws + x) * p.C0 + c);
                else o = __ldg(p.src1 + (((size_t)n * p.Hs + y) * p.Ws + x) * p.C1 + (c - p.C0));
            }
            v[j] = o;
        }
    }
}

// ((x - mean) * rstd) * gamma + beta, then SiLU: the reference's order of operations (F.group_norm, then act).
// The pass is issue-bound (ncu: 0.71 warp instructions per scheduler-cycle at 3.6 TB/s), so SiLU uses the 2-instruction
// ex2.approx exponential and the approximate divide: |error| <= ~3e-7 relative, below the fp32 accumulation error of
// the convolution that consumes the operand.
template <bool VEC>
__device__ __forceinline__ void prep_transform8(const PrepParams& p, const PrepChunk& k, int c0, int C, float (&v)[8]) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float o = v[j];
        if (p.mean_rstd) o = (o - k.m[j]) * k.r[j];
        if (p.gamma) o = fmaf(o, k.g[j], k.b[j]);
        if (p.act == 1) o = __fdividef(o, 1.f + __expf(-o));
        else if (p.act == 2) o = o / (1.f + expf(-o));   // IEEE divide + full-accuracy expf (accuracy experiments)
        v[j] = (VEC || c0 + j < C) ? o : 0.f;
    }
}

// two values -> packed fp16 (hi, hi) and (lo, lo) words with the paired conversion instructions
__device__ __forceinline__ void split2_fp16(float a, float b, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(a, b);
    const float2 back = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - back.x, b - back.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}

template <bool VEC, int KIND>   // KIND 0: modes 0-2, 1: 2x2 average pool, 2: FIR resampling
__global__ void __launch_bounds__(256, KIND ? 2 : 4) conv_prep_kernel(const PrepParams p) {
    constexpr bool POOL = KIND == 1;
    // The launcher sizes the grid so that (gridDim.x * 256) % chunks == 0: a thread keeps ONE 8-channel chunk for its
    // whole grid-stride loop (per-channel parameters live in registers) and only the pixel index advances; all index
    // arithmetic is 32-bit, spatial sizes that are powers of two use shifts.
    const unsigned chunks = (unsigned)(p.Cp >> 3);
    const int C = p.C0 + p.C1;
    const unsigned n_pix = (unsigned)p.planes * p.N * p.Ho * p.Wo;
    const unsigned gtid = blockIdx.x * 256u + threadIdx.x;
    const unsigned ch = gtid % chunks;
    const unsigned pix_stride = (gridDim.x * 256u) / chunks;
    const int c0 = (int)ch * 8;
    const bool live = c0 < C;
    float kg[8], kb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = min(c0 + j, C - 1);
        kg[j] = (live && p.gamma) ? __ldg(p.gamma + c) : 1.f;
        kb[j] = (live && p.gamma && p.beta) ? __ldg(p.beta + c) : 0.f;
    }
    const int g_first = live ? c0 / p.cg : 0;
    for (unsigned pix = gtid / chunks; pix < n_pix; pix += pix_stride) {
        unsigned r = pix;
        int xo, yo;
        if (p.wo_shift >= 0) { xo = (int)(r & (unsigned)(p.Wo - 1)); r >>= p.wo_shift; } else { xo = (int)(r % (unsigned)p.Wo); r /= (unsigned)p.Wo; }
        if (p.ho_shift >= 0) { yo = (int)(r & (unsigned)(p.Ho - 1)); r >>= p.ho_shift; } else { yo = (int)(r % (unsigned)p.Ho); r /= (unsigned)p.Ho; }
        int n = (int)r, plane = 0;
        if (p.planes > 1) { n = (int)(r % (unsigned)p.N); plane = (int)(r / (unsigned)p.N); }
        float v[8];
        if (!live) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.f;
        } else {
            PrepChunk k;
#pragma unroll
            for (int j = 0; j < 8; ++j) { k.g[j] = kg[j]; k.b[j] = kb[j]; k.m[j] = 0.f; k.r[j] = 1.f; }
            if (p.mean_rstd) {
                const float2* mrp = reinterpret_cast<const float2*>(p.mean_rstd) + (size_t)n * p.groups;
                if ((p.cg & 7) == 0) {            // the chunk lies in one group
                    const float2 mr = __ldg(mrp + g_first);
#pragma unroll
                    for (int j = 0; j < 8; ++j) { k.m[j] = mr.x; k.r[j] = mr.y; }
                } else if ((p.cg & 3) == 0) {     // two groups per chunk (cg = 4, 12, ...)
                    const float2 a = __ldg(mrp + g_first), b = __ldg(mrp + (c0 + 4) / p.cg);
#pragma unroll
                    for (int j = 0; j < 4; ++j) { k.m[j] = a.x; k.r[j] = a.y; k.m[j + 4] = b.x; k.r[j + 4] = b.y; }
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float2 mr = __ldg(mrp + min(c0 + j, C - 1) / p.cg);
                        k.m[j] = mr.x; k.r[j] = mr.y;
                    }
                }
            }
            if (POOL) {  // 2 x 2 average pool of the transformed values (BeatGANs_nn.py avg_pool_nd)
                float a[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = 0.f;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    prep_load8<VEC>(p, n, 2 * yo + (q >> 1), 2 * xo + (q & 1), c0, C, a);
                    prep_transform8<VEC>(p, k, c0, C, a);
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] += a[j];
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] *= 0.25f;
            } else if (p.mode == 0) {
                prep_load8<VEC>(p, n, yo, xo, c0, C, v);
                prep_transform8<VEC>(p, k, c0, C, v);
            } else if (p.mode == 1) {  // nearest-neighbour x2 (F.interpolate 'nearest', layers.py:600)
                prep_load8<VEC>(p, n, yo >> 1, xo >> 1, c0, C, v);
                prep_transform8<VEC>(p, k, c0, C, v);
            } else if (p.mode == 2) {  // parity plane (py, px) of a stride-2 read: src[2*yo + py][2*xo + px]
                const int y = 2 * yo + (plane >> 1), x = 2 * xo + (plane & 1);
                if (y < p.Hs && x < p.Ws) {
                    prep_load8<VEC>(p, n, y, x, c0, C, v);
                    prep_transform8<VEC>(p, k, c0, C, v);
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] = 0.f;
                }
            } else if (KIND == 2 && p.mode == 4) {
                // upfirdn2d of the TRANSFORMED tensor (layerspp.py:244-257: FIR up/down-sampling of act(GN(x)) and of x;
                // ncsnpp.py:297-304: FIR pad before the stride-2 convolution of the input pyramid), gathered per output
                // pixel: out[y][x] = sum_ab fir[a][b] * U[y*down + a - pad0][x*down + b - pad0], U = zero-stuffed (up)
                // transformed source, zero outside.  No intermediate tensor exists.  With out_raw the SAME gather also
                // filters the untransformed source (the res-block's skip branch, layerspp.py:246-248,253-255): one read
                // of the activation serves both branches.
                float raw[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) { v[j] = 0.f; raw[j] = 0.f; }
                const int yf = p.fir_planes ? 2 * yo + (plane >> 1) : yo, xf = p.fir_planes ? 2 * xo + (plane & 1) : xo;
                if (yf < p.fir_hf && xf < p.fir_wf) {
                    const int y0 = yf * p.fir_down - p.fir_pad0, x0 = xf * p.fir_down - p.fir_pad0;
                    for (int a = 0; a < p.fir_kh; ++a) {
                        int i = y0 + a;
                        if (i < 0 || (p.fir_up == 2 && (i & 1))) continue;
                        if (p.fir_up == 2) i >>= 1;
                        if (i >= p.Hs) continue;
                        for (int b = 0; b < p.fir_kw; ++b) {
                            int jx = x0 + b;
                            if (jx < 0 || (p.fir_up == 2 && (jx & 1))) continue;
                            if (p.fir_up == 2) jx >>= 1;
                            if (jx >= p.Ws) continue;
                            float t[8];
                            prep_load8<VEC>(p, n, i, jx, c0, C, t);
                            const float wgt = p.fir[a * 4 + b];
                            if (p.out_raw) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) raw[j] = fmaf(wgt, t[j], raw[j]);
                            }
                            prep_transform8<VEC>(p, k, c0, C, t);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] = fmaf(wgt, t[j], v[j]);
                        }
                    }
                }
                if (p.out_raw) {
                    if (VEC) {
                        float4* dst = reinterpret_cast<float4*>(p.out_raw + (size_t)pix * C + c0);
                        dst[0] = make_float4(raw[0], raw[1], raw[2], raw[3]);
                        dst[1] = make_float4(raw[4], raw[5], raw[6], raw[7]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            if (c0 + j < C) p.out_raw[(size_t)pix * C + c0 + j] = raw[j];
                    }
                }
            }
        }
        if (p.out_hi) {
            uint32_t h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) split2_fp16(v[2 * j] * p.scale, v[2 * j + 1] * p.scale, h[j], l[j]);
            *reinterpret_cast<uint4*>(p.out_hi + ((size_t)pix * p.Cp + c0) * 2) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(p.out_lo + ((size_t)pix * p.Cp + c0) * 2) = make_uint4(l[0], l[1], l[2], l[3]);
        }
        if (p.out_f32 && live) {
            if (VEC) {
                float4* dst = reinterpret_cast<float4*>(p.out_f32 + (size_t)pix * C + c0);
                dst[0] = make_float4(v[0], v[1], v[2], v[3]);
                dst[1] = make_float4(v[4], v[5], v[6], v[7]);
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (c0 + j < C) p.out_f32[(size_t)pix * C + c0 + j] = v[j];
            }
        }
    }
}

// Same-size fast path (mode 0, NHWC sources, 8-channel aligned, one plane) -- 70 % of the prep traffic of a U-Net.
// The source pixel IS the output pixel, so the hot loop has no spatial index arithmetic at all: n = pix >> log2(HW)
// (or one division) selects the GroupNorm statistics, every per-channel quantity sits in registers.
template <bool HW_POW2>
__global__ void __launch_bounds__(256, 4) conv_prep_same_kernel(const PrepParams p, int hw_shift, unsigned hw) {
    const unsigned chunks = (unsigned)(p.Cp >> 3);
    const int C = p.C0 + p.C1;
    const unsigned n_pix = (unsigned)p.N * hw;
    const unsigned gtid = blockIdx.x * 256u + threadIdx.x;
    const unsigned ch = gtid % chunks;
    const unsigned pix_stride = (gridDim.x * 256u) / chunks;
    const int c0 = (int)ch * 8;
    const bool live = c0 < C;
    PrepChunk k;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = live ? c0 + j : 0;
        k.g[j] = (live && p.gamma) ? __ldg(p.gamma + c) : 1.f;
        k.b[j] = (live && p.gamma && p.beta) ? __ldg(p.beta + c) : 0.f;
        k.m[j] = 0.f; k.r[j] = 1.f;
    }
    const float* src = p.src0;
    unsigned ldc = 0;
    if (live) {
        if (c0 < p.C0) { src = p.src0 + c0; ldc = (unsigned)p.C0; } else { src = p.src1 + (c0 - p.C0); ldc = (unsigned)p.C1; }
    }
    const int ga = live ? c0 / p.cg : 0, gb = live ? (c0 + 4) / p.cg : 0;
    const bool two = (p.cg & 3) == 0;     // at most two groups per chunk; otherwise per-channel lookups
    int gj[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) gj[j] = live ? (c0 + j) / p.cg : 0;
    const float2* mr_base = reinterpret_cast<const float2*>(p.mean_rstd);
    for (unsigned pix = gtid / chunks; pix < n_pix; pix += pix_stride) {
        float v[8];
        if (live) {
            const float4* q = reinterpret_cast<const float4*>(src + (size_t)pix * ldc);
            const float4 a0 = __ldg(q), a1 = __ldg(q + 1);
            v[0] = a0.x; v[1] = a0.y; v[2] = a0.z; v[3] = a0.w; v[4] = a1.x; v[5] = a1.y; v[6] = a1.z; v[7] = a1.w;
            if (mr_base) {
                const unsigned n = HW_POW2 ? pix >> hw_shift : pix / hw;
                const float2* mrp = mr_base + (size_t)n * p.groups;
                if (two) {
                    const float2 x = __ldg(mrp + ga), y = __ldg(mrp + gb);
#pragma unroll
                    for (int j = 0; j < 4; ++j) { k.m[j] = x.x; k.r[j] = x.y; k.m[j + 4] = y.x; k.r[j + 4] = y.y; }
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float2 mr = __ldg(mrp + gj[j]);
                        k.m[j] = mr.x; k.r[j] = mr.y;
                    }
                }
            }
            prep_transform8<true>(p, k, c0, C, v);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.f;
        }
        if (p.out_hi) {
            uint32_t h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) split2_fp16(v[2 * j] * p.scale, v[2 * j + 1] * p.scale, h[j], l[j]);
            *reinterpret_cast<uint4*>(p.out_hi + ((size_t)pix * p.Cp + c0) * 2) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(p.out_lo + ((size_t)pix * p.Cp + c0) * 2) = make_uint4(l[0], l[1], l[2], l[3]);
        }
        if (p.out_f32 && live) {
            float4* dst = reinterpret_cast<float4*>(p.out_f32 + (size_t)pix * C + c0);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
    }
}

static int ilog2_exact(int v) {
    if (v <= 0 || (v & (v - 1))) return -1;
    int s = 0;
    while ((1 << s) < v) ++s;
    return s;
}

cudaError_t launch_conv_prep(const PrepParams& p_in, int num_sms, cudaStream_t stream) {
    PrepParams p = p_in;
    const long long n_pix = (long long)p.planes * p.N * p.Ho * p.Wo;
    const long long chunks = p.Cp >> 3;
    const long long total = n_pix * chunks;
    if (total <= 0) return cudaSuccess;
    if (n_pix >= (1ll << 31)) return cudaErrorInvalidValue;
    p.wo_shift = ilog2_exact(p.Wo);
    p.ho_shift = ilog2_exact(p.Ho);
    // grid * 256 must be a multiple of chunks: grid = multiple of chunks / gcd(chunks, 256)
    long long g = chunks, b = 256;
    while (b) { const long long t = g % b; g = b; b = t; }
    const long long unit = chunks / g;
    long long grid = (total + 255) / 256;
    const long long cap = (long long)num_sms * 16;
    if (grid > cap) grid = cap;
    grid = (grid + unit - 1) / unit * unit;
    auto al32 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const bool vec = !p.src_nchw && p.C0 % 8 == 0 && p.C1 % 8 == 0 && al32(p.src0) && (!p.src1 || al32(p.src1)) &&
                     (!p.out_f32 || al32(p.out_f32));
    if (vec && p.mode == 0 && p.planes == 1) {
        const int hw_shift = ilog2_exact(p.Ho * p.Wo);
        if (hw_shift >= 0) conv_prep_same_kernel<true><<<(unsigned)grid, 256, 0, stream>>>(p, hw_shift, (unsigned)(p.Ho * p.Wo));
        else conv_prep_same_kernel<false><<<(unsigned)grid, 256, 0, stream>>>(p, 0, (unsigned)(p.Ho * p.Wo));
        return cudaGetLastError();
    }
    if (p.mode == 3) {
        if (vec) conv_prep_kernel<true, 1><<<(unsigned)grid, 256, 0, stream>>>(p);
        else conv_prep_kernel<false, 1><<<(unsigned)grid, 256, 0, stream>>>(p);
    } else if (p.mode == 4) {
        if (vec) conv_prep_kernel<true, 2><<<(unsigned)grid, 256, 0, stream>>>(p);
        else conv_prep_kernel<false, 2><<<(unsigned)grid, 256, 0, stream>>>(p);
    } else {
        if (vec) conv_prep_kernel<true, 0><<<(unsigned)grid, 256, 0, stream>>>(p);
        else conv_prep_kernel<false, 0><<<(unsigned)grid, 256, 0, stream>>>(p);
    }
    return cudaGetLastError();
}

// ---------------------------------- attention -------------------------------------------------
// out[n][i][h*dh + c] = sum_j softmax_j( scale * <q_i, k_j> ) v[j][c]     (layers.py:581-587; heads for BeatGANs)
// CTA = 32 queries of one (image, head); scores for all T keys live in shared memory.
constexpr int AT_Q = 32;

__global__ void __launch_bounds__(256) attention_kernel(const AttnParams p) {
    extern __shared__ float at_smem[];
    float* Sm = at_smem;                    // [T][32]   scores, then probabilities (transposed: key-major)
    float* Qc = Sm + (size_t)p.T * AT_Q;    // [32 c][32 i]
    float* Kc = Qc + 32 * AT_Q;             // [256 j][33]
    float* red = Kc + 256 * 33;             // [8][32] x 2
    const int t = threadIdx.x;
    const int i0 = blockIdx.x * AT_Q, head = blockIdx.y, n = blockIdx.z;
    const float* base = p.qkv + (size_t)n * p.T * p.ld;
    const int qo = p.q0 + head * p.head_stride, ko = p.k0 + head * p.head_stride, vo = p.v0 + head * p.head_stride;

    // phase 1: S[i][j] = scale * <q_i, k_j>
    for (int j0 = 0; j0 < p.T; j0 += 256) {
        float acc[AT_Q];
#pragma unroll
        for (int i = 0; i < AT_Q; ++i) acc[i] = 0.f;
        for (int c0 = 0; c0 < p.dh; c0 += 32) {
            __syncthreads();
            for (int e = t; e < 32 * AT_Q; e += 256) {  // Qc[c][i]
                const int i = e >> 5, c = e & 31;
                Qc[c * AT_Q + i] = (i0 + i < p.T && c0 + c < p.dh) ? __ldg(base + (size_t)(i0 + i) * p.ld + qo + c0 + c) : 0.f;
            }
            for (int e = t; e < 256 * 32; e += 256) {  // Kc[j][c]
                const int j = e >> 5, c = e & 31;
                Kc[j * 33 + c] = (j0 + j < p.T && c0 + c < p.dh) ? __ldg(base + (size_t)(j0 + j) * p.ld + ko + c0 + c) : 0.f;
            }
            __syncthreads();
#pragma unroll 4
            for (int c = 0; c < 32; ++c) {
                const float kv = Kc[t * 33 + c];
                const float4* q4 = reinterpret_cast<const float4*>(Qc + c * AT_Q);
#pragma unroll
                for (int i4 = 0; i4 < AT_Q / 4; ++i4) {
                    const float4 q = q4[i4];
                    acc[4 * i4] = fmaf(q.x, kv, acc[4 * i4]);
                    acc[4 * i4 + 1] = fmaf(q.y, kv, acc[4 * i4 + 1]);
                    acc[4 * i4 + 2] = fmaf(q.z, kv, acc[4 * i4 + 2]);
                    acc[4 * i4 + 3] = fmaf(q.w, kv, acc[4 * i4 + 3]);
                }
            }
        }
        if (j0 + t < p.T) {
            float4* dst = reinterpret_cast<float4*>(Sm + (size_t)(j0 + t) * AT_Q);
#pragma unroll
            for (int i4 = 0; i4 < AT_Q / 4; ++i4)
                dst[i4] = make_float4(acc[4 * i4] * p.scale, acc[4 * i4 + 1] * p.scale, acc[4 * i4 + 2] * p.scale,
                                      acc[4 * i4 + 3] * p.scale);
        }
    }
    __syncthreads();
    // phase 2: softmax over keys for each of the 32 queries (F.softmax, layers.py:584)
    {
        const int i = t & 31, part = t >> 5;
        float m = -INFINITY;
        for (int j = part; j < p.T; j += 8) m = fmaxf(m, Sm[(size_t)j * AT_Q + i]);
        red[part * 32 + i] = m;
        __syncthreads();
        m = red[i];
#pragma unroll
        for (int q = 1; q < 8; ++q) m = fmaxf(m, red[q * 32 + i]);
        float s = 0.f;
        for (int j = part; j < p.T; j += 8) {
            const float e = expf(Sm[(size_t)j * AT_Q + i] - m);
            Sm[(size_t)j * AT_Q + i] = e;
            s += e;
        }
        red[256 + part * 32 + i] = s;
        __syncthreads();
        s = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) s += red[256 + q * 32 + i];
        const float inv = 1.f / s;
        for (int j = part; j < p.T; j += 8) Sm[(size_t)j * AT_Q + i] *= inv;
    }
    __syncthreads();
    // phase 3: O[i][c] = sum_j P[i][j] v[j][c]
    for (int c = t; c < p.dh; c += 256) {
        float acc[AT_Q];
#pragma unroll
        for (int i = 0; i < AT_Q; ++i) acc[i] = 0.f;
        for (int j = 0; j < p.T; ++j) {
            const float vv = __ldg(base + (size_t)j * p.ld + vo + c);
            const float4* p4 = reinterpret_cast<const float4*>(Sm + (size_t)j * AT_Q);
#pragma unroll
            for (int i4 = 0; i4 < AT_Q / 4; ++i4) {
                const float4 q = p4[i4];
                acc[4 * i4] = fmaf(q.x, vv, acc[4 * i4]);
                acc[4 * i4 + 1] = fmaf(q.y, vv, acc[4 * i4 + 1]);
                acc[4 * i4 + 2] = fmaf(q.z, vv, acc[4 * i4 + 2]);
                acc[4 * i4 + 3] = fmaf(q.w, vv, acc[4 * i4 + 3]);
            }
        }
#pragma unroll
        for (int i = 0; i < AT_Q; ++i)
            if (i0 + i < p.T) p.out[((size_t)n * p.T + i0 + i) * p.ld_out + head * p.dh + c] = acc[i];
    }
}

cudaError_t launch_attention(const AttnParams& p, cudaStream_t stream) {
    if (p.N <= 0 || p.T <= 0) return cudaSuccess;
    const size_t smem = ((size_t)p.T * AT_Q + 32 * AT_Q + 256 * 33 + 512) * sizeof(float);
    if (smem > 220 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attention_kernel<<<dim3((p.T + AT_Q - 1) / AT_Q, p.heads, p.N), 256, smem, stream>>>(p);
    return cudaGetLastError();
}

// ---------------------------------- attention on the tensor cores ------------------------------
// layers.py:577-587 / BeatGANsblocks.py:480-491 for one head of T >= 128 tokens: S = scale Q K^T and O = softmax(S) V are
// run by conv_fp16x3_kernel as 1x1 convolutions whose weight matrix differs per image (ConvParams::w_img_stride): the
// "pixels" are the queries, the "output channels" the keys (then the value channels).  These two kernels build the
// operands: Q as the NHWC activation operand, K as a per-image weight slab [dh/64][T][128 B] (rows = keys), V^T as a
// per-image weight slab [T/64][dh][128 B] (rows = value channels, contraction = keys), and the softmax that turns the
// fp32 scores into the P operand.  All operands are fp16 hi/lo pairs of value * scale, like every other convolution.
__global__ void __launch_bounds__(256) attn_pack_kernel(const float* __restrict__ qkv, int N, int T, int ld, int dh, int q0,
                                                        int k0, int v0, float scale, uint8_t* __restrict__ q_hi,
                                                        uint8_t* __restrict__ q_lo, uint8_t* __restrict__ k_hi,
                                                        uint8_t* __restrict__ k_lo, uint8_t* __restrict__ vt_hi,
                                                        uint8_t* __restrict__ vt_lo) {
    const int chunks = dh >> 3;
    const long long plain = (long long)N * T * chunks;          // phase 1: (n, t, 8 channels) -> Q operand and K slab
    const long long trans = (long long)N * dh * (T >> 3);       // phase 2: (n, c, 8 tokens)   -> V^T slab
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < plain + trans; i += (long long)gridDim.x * 256) {
        float v[8];
        uint32_t h[4], l[4];
        if (i < plain) {
            const int ch = (int)(i % chunks), t = (int)((i / chunks) % T), n = (int)(i / ((long long)chunks * T));
            const float* row = qkv + ((size_t)n * T + t) * ld;
            const int c0 = ch << 3;
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __ldg(row + q0 + c0 + j);
#pragma unroll
            for (int j = 0; j < 4; ++j) split2_fp16(v[2 * j] * scale, v[2 * j + 1] * scale, h[j], l[j]);
            const size_t qo = (((size_t)n * T + t) * dh + c0) * 2;
            *reinterpret_cast<uint4*>(q_hi + qo) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(q_lo + qo) = make_uint4(l[0], l[1], l[2], l[3]);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __ldg(row + k0 + c0 + j);
#pragma unroll
            for (int j = 0; j < 4; ++j) split2_fp16(v[2 * j] * scale, v[2 * j + 1] * scale, h[j], l[j]);
            const size_t ko = (size_t)n * T * dh * 2 + slab_offset_bytes((size_t)T, (size_t)t, (size_t)c0);
            *reinterpret_cast<uint4*>(k_hi + ko) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(k_lo + ko) = make_uint4(l[0], l[1], l[2], l[3]);
        } else {
            const long long r = i - plain;
            const int c = (int)(r % dh), t0 = (int)((r / dh) % (T >> 3)) << 3, n = (int)(r / ((long long)dh * (T >> 3)));
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __ldg(qkv + ((size_t)n * T + t0 + j) * ld + v0 + c);
#pragma unroll
            for (int j = 0; j < 4; ++j) split2_fp16(v[2 * j] * scale, v[2 * j + 1] * scale, h[j], l[j]);
            const size_t vo = (size_t)n * T * dh * 2 + slab_offset_bytes((size_t)dh, (size_t)c, (size_t)t0);
            *reinterpret_cast<uint4*>(vt_hi + vo) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(vt_lo + vo) = make_uint4(l[0], l[1], l[2], l[3]);
        }
    }
}

cudaError_t launch_attn_pack(const float* qkv, int N, int T, int ld, int dh, int q0, int k0, int v0, float scale,
                             uint8_t* q_hi, uint8_t* q_lo, uint8_t* k_hi, uint8_t* k_lo, uint8_t* vt_hi, uint8_t* vt_lo,
                             cudaStream_t stream) {
    const long long total = (long long)N * T * (dh >> 3) * 2;
    if (total <= 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    attn_pack_kernel<<<(int)blocks, 256, 0, stream>>>(qkv, N, T, ld, dh, q0, k0, v0, scale, q_hi, q_lo, k_hi, k_lo, vt_hi, vt_lo);
    return cudaGetLastError();
}

// F.softmax over the keys of one query row (layers.py:584), one warp per row, the reference's order of operations
// (max, exp, sum, divide); writes the probabilities as the hi/lo operand [rows][T] of P * scale.
__global__ void __launch_bounds__(256) attn_softmax_kernel(const float* __restrict__ S, long long rows, int T, float scale,
                                                           uint8_t* __restrict__ p_hi, uint8_t* __restrict__ p_lo) {
    const long long row = blockIdx.x * 8LL + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* s = S + row * T;
    float m = -INFINITY;
    for (int c = lane * 8; c < T; c += 256) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(s + c)), b = __ldg(reinterpret_cast<const float4*>(s + c) + 1);
        m = fmaxf(m, fmaxf(fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w)), fmaxf(fmaxf(b.x, b.y), fmaxf(b.z, b.w))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float sum = 0.f;
    for (int c = lane * 8; c < T; c += 256) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(s + c)), b = __ldg(reinterpret_cast<const float4*>(s + c) + 1);
        sum += expf(a.x - m) + expf(a.y - m) + expf(a.z - m) + expf(a.w - m) + expf(b.x - m) + expf(b.y - m) + expf(b.z - m) +
               expf(b.w - m);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.f / sum;
    for (int c = lane * 8; c < T; c += 256) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(s + c)), b = __ldg(reinterpret_cast<const float4*>(s + c) + 1);
        const float v[8] = {expf(a.x - m) * inv, expf(a.y - m) * inv, expf(a.z - m) * inv, expf(a.w - m) * inv,
                            expf(b.x - m) * inv, expf(b.y - m) * inv, expf(b.z - m) * inv, expf(b.w - m) * inv};
        uint32_t h[4], l[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split2_fp16(v[2 * j] * scale, v[2 * j + 1] * scale, h[j], l[j]);
        const size_t off = ((size_t)row * T + c) * 2;
        *reinterpret_cast<uint4*>(p_hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(p_lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

cudaError_t launch_attn_softmax(const float* S, int rows, int T, float scale, uint8_t* p_hi, uint8_t* p_lo,
                                cudaStream_t stream) {
    if (rows <= 0) return cudaSuccess;
    attn_softmax_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(S, rows, T, scale, p_hi, p_lo);
    return cudaGetLastError();
}

}  // namespace idb
