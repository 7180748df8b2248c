// Denoising-score-matching training step of the FCN score network: the kernels AROUND the tensor-core
// GEMMs.  Reference: losses.py:164-188 (unconditional loss_fn: t ~ U(eps, T), x_t = mean + std z, score_fn, squared error),
// models/utils.py:254-266 (VE: score = -model / std, so the loss is mean (model(x_t, (N-1) t) - z)^2),
// losses.py:26-51 (Adam, linear warm-up, clip_grad_norm_), BaseSdeGenerativeModel.py:60-63 (training_step).
//
// The forward and backward matrix products run on the SAME tcgen05 kernel as the estimation path (idb_gemm.cu) in its
// bf16-pair format (16 significand bits, fp32 exponent range: gradients need no loss scaling):
//   forward   pre_l = H_{l-1} W_l^T + b_l,  H_l = ELU(pre_l)          idb_fcn_anchor_layer  (fp32 pre + next operand)
//   backward  dW_l  = dZ_l^T H_{l-1}                                  idb_fcn_output_layer on TRANSPOSED operand slabs
//             dH_{l-1} = dZ_l W_l                                     idb_fcn_output_layer on the transposed weight slab
//             dZ_{l-1} = dH_{l-1} * ELU'(pre_{l-1})                   fused into the operand packing below
// What lives here: sampling (Philox noise + marginal scaling, written in the network's input layout), the loss and its
// gradient, operand packing (plain and transposed, with the activation / its derivative fused in), bias gradients, the
// global gradient norm, and Adam with warm-up and clipping (no host synchronisation anywhere in a step).
#include "idb_common.cuh"
#include "idb_kernels.h"
#include <math.h>

namespace idb {

// ---------------------------------- sampling ------------------------------------------------
// One thread = one row of the batch x 8 consecutive input columns.  X[r] = (mean_coeff[r] * x0 + std[r] * z, label_scale * t[r], 0...)
__global__ void __launch_bounds__(256) dsm_sample_kernel(const float* __restrict__ data, const long long* __restrict__ idx,
                                                         const float* __restrict__ t, const float* __restrict__ mean_coeff,
                                                         const float* __restrict__ stdv, const float* __restrict__ z_ext,
                                                         int B, int D, int Mp, int ldx, float label_scale, uint64_t seed,
                                                         uint64_t step_host, const long long* __restrict__ step_dev,
                                                         float* __restrict__ X, float* __restrict__ z_out) {
    const uint64_t step = step_dev ? (uint64_t)*step_dev : step_host;   // device counter: the step replays as a CUDA graph
    const int chunks = ldx >> 3;
    const long long total = (long long)Mp * chunks;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int ch = (int)(i % chunks), r = (int)(i / chunks), c0 = ch << 3;
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = 0.f;
        if (r < B) {
            float zz[8];
            if (z_ext) {
#pragma unroll
                for (int j = 0; j < 8; ++j) zz[j] = (c0 + j < D) ? __ldg(z_ext + (size_t)r * D + c0 + j) : 0.f;
            } else {
                const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const uint4 q = philox4x32_10(make_uint4((uint32_t)(ch * 2 + h), (uint32_t)r, (uint32_t)step,
                                                             (uint32_t)(step >> 32) ^ 0x5eedu), key);
                    box_muller(q.x, q.y, zz[4 * h + 0], zz[4 * h + 1]);
                    box_muller(q.z, q.w, zz[4 * h + 2], zz[4 * h + 3]);
                }
            }
            const float* x0 = data + (size_t)(idx ? idx[r] : r) * D;
            const float mc = __ldg(mean_coeff + r), sd = __ldg(stdv + r);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int c = c0 + j;
                if (c < D) {
                    o[j] = __fadd_rn(__fmul_rn(mc, __ldg(x0 + c)), __fmul_rn(sd, zz[j]));   // mean + std * z, fp32 ops as the reference
                    z_out[(size_t)r * D + c] = zz[j];
                } else if (c == D) {
                    o[j] = __fmul_rn(__ldg(t + r), label_scale);                            // torch.cat([x, labels]) (fcn.py:35-38)
                }
            }
        }
        float4* dst = reinterpret_cast<float4*>(X + (size_t)r * ldx + c0);
        dst[0] = make_float4(o[0], o[1], o[2], o[3]);
        dst[1] = make_float4(o[4], o[5], o[6], o[7]);
    }
}

cudaError_t launch_dsm_sample(const float* data, const long long* idx, const float* t, const float* mean_coeff,
                              const float* stdv, const float* z_ext, int B, int D, int Mp, int ldx, float label_scale,
                              uint64_t seed, uint64_t step, const long long* step_dev, float* X, float* z_out,
                              cudaStream_t stream) {
    const long long total = (long long)Mp * (ldx >> 3);
    if (total <= 0) return cudaSuccess;
    dsm_sample_kernel<<<(int)((total + 255) / 256), 256, 0, stream>>>(data, idx, t, mean_coeff, stdv, z_ext, B, D, Mp, ldx,
                                                                     label_scale, seed, step, step_dev, X, z_out);
    return cudaGetLastError();
}

// ---------------------------------- operand packing -----------------------------------------
// value(r, c) = f(src[r][c], pre[r][c]) inside [R) x [C), 0 outside;  f: 0 identity | 1 ELU(src) | 2 src * ELU'(pre).
// Plain:      slab of the matrix [Rp][Cp]   (rows = batch, contraction = features)  + optional fp32 copy [Rp][ld_out]
// Transposed: slab of the matrix [CpT][RpT] (rows = features, contraction = batch)
__device__ __forceinline__ float train_value(const TrainPackParams& p, int r, int c) {
    if (r >= p.R || c >= p.C) return 0.f;
    const float s = __ldg(p.src + (size_t)r * p.ld + c);
    if (p.mode == 0) return s;
    if (p.mode == 1) return s > 0.f ? s : expm1f(s);
    const float z = __ldg(p.pre + (size_t)r * p.ld_pre + c);
    return z > 0.f ? s : s * expf(z);
}

__global__ void __launch_bounds__(256) train_pack_kernel(const TrainPackParams p) {
    if (!p.transpose) {
        const int chunks = p.Cp >> 3;
        const long long total = (long long)p.Rp * chunks;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
             i += (long long)gridDim.x * blockDim.x) {
            const int ch = (int)(i % chunks), r = (int)(i / chunks), c0 = ch << 3;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = train_value(p, r, c0 + j);
            if (p.hi) {
                uint32_t h[4], l[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) split_pair16(v[2 * j] * p.scale, v[2 * j + 1] * p.scale, p.fmt, h[j], l[j]);
                const size_t off = slab_offset_bytes((size_t)p.Rp, (size_t)r, (size_t)c0);
                *reinterpret_cast<uint4*>(p.hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
                *reinterpret_cast<uint4*>(p.lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
            }
            if (p.out_f32) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (c0 + j < p.ld_out) p.out_f32[(size_t)r * p.ld_out + c0 + j] = v[j];
            }
        }
    } else {
        // rows of the slab = features (index c of the source), contraction = batch (index r): consecutive threads take
        // consecutive features so that the strided source reads coalesce across the warp
        const int rchunks = p.Rp >> 3;
        const long long total = (long long)p.Cp * rchunks;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
             i += (long long)gridDim.x * blockDim.x) {
            const int c = (int)(i % p.Cp), r0 = (int)(i / p.Cp) << 3;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = train_value(p, r0 + j, c);
            uint32_t h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) split_pair16(v[2 * j] * p.scale, v[2 * j + 1] * p.scale, p.fmt, h[j], l[j]);
            const size_t off = slab_offset_bytes((size_t)p.Cp, (size_t)c, (size_t)r0);
            *reinterpret_cast<uint4*>(p.hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(p.lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
        }
    }
}

cudaError_t launch_train_pack(const TrainPackParams& p, cudaStream_t stream) {
    const long long total = p.transpose ? (long long)p.Cp * (p.Rp >> 3) : (long long)p.Rp * (p.Cp >> 3);
    if (total <= 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    train_pack_kernel<<<(int)blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

// ---------------------------------- loss and its gradient -----------------------------------
// loss = mean_{r < B, c < D} (out[r][c] - z[r][c])^2 (losses.py:173-185 with reduce_mean, VE, no likelihood weighting);
// dZ[r][c] = 2 (out - z) / (B D) inside, 0 in the padding.  One CTA: the sum is formed in a fixed order (reproducible).
constexpr int LOSS_BLOCKS = 64;
__global__ void __launch_bounds__(256) dsm_loss_kernel(const float* __restrict__ out, int ldo, const float* __restrict__ z,
                                                       int B, int D, int Mp, float* __restrict__ dZ,
                                                       double* __restrict__ partial) {
    __shared__ double red[8];
    const double inv = 1.0 / ((double)B * (double)D);
    double acc = 0.0;
    const long long total = (long long)Mp * ldo;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += (long long)LOSS_BLOCKS * 256) {
        const int r = (int)(i / ldo), c = (int)(i % ldo);
        float g = 0.f;
        if (r < B && c < D) {
            const float e = out[i] - __ldg(z + (size_t)r * D + c);
            acc += (double)e * (double)e;
            g = (float)(2.0 * (double)e * inv);
        }
        dZ[i] = g;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += red[w];
        partial[blockIdx.x] = s;
    }
}
// second stage: the block partials in a fixed order (bit-reproducible loss)
__global__ void dsm_loss_final_kernel(const double* __restrict__ partial, double inv, float* __restrict__ loss) {
    double s = 0.0;
    for (int i = 0; i < LOSS_BLOCKS; ++i) s += partial[i];
    *loss = (float)(s * inv);
}

cudaError_t launch_dsm_loss(const float* out, int ldo, const float* z, int B, int D, int Mp, float* dZ, float* loss,
                            double* partial, cudaStream_t stream) {
    dsm_loss_kernel<<<LOSS_BLOCKS, 256, 0, stream>>>(out, ldo, z, B, D, Mp, dZ, partial);
    dsm_loss_final_kernel<<<1, 1, 0, stream>>>(partial, 1.0 / ((double)B * (double)D), loss);
    return cudaGetLastError();
}

// ---------------------------------- bias gradient -------------------------------------------
// out[c] = sum_{r < R} src[r][c]: one thread per column, rows in order (reproducible); coalesced across the warp.
__global__ void __launch_bounds__(256) colsum_f32_kernel(const float* __restrict__ src, int ld, int R, int C,
                                                         float* __restrict__ out) {
    __shared__ float part[8][33];
    const int c = blockIdx.x * 32 + (threadIdx.x & 31), rg = threadIdx.x >> 5;
    float a = 0.f;
    if (c < C)
        for (int r = rg; r < R; r += 8) a += src[(size_t)r * ld + c];
    part[rg][threadIdx.x & 31] = a;
    __syncthreads();
    if (rg == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += part[i][threadIdx.x];
        out[c] = t;
    }
}

cudaError_t launch_colsum_f32(const float* src, int ld, int R, int C, float* out, cudaStream_t stream) {
    if (C <= 0) return cudaSuccess;
    colsum_f32_kernel<<<(C + 31) / 32, 256, 0, stream>>>(src, ld, R, C, out);
    return cudaGetLastError();
}

// ---------------------------------- gradient norm, Adam -------------------------------------
// accum += sum over [N) x [K) of g^2 (torch.nn.utils.clip_grad_norm_, losses.py:47-48)
__global__ void __launch_bounds__(256) grad_sumsq_kernel(const float* __restrict__ g, int ld, int N, int K,
                                                         double* __restrict__ accum) {
    __shared__ double red[8];
    double a = 0.0;
    for (int n = blockIdx.x; n < N; n += gridDim.x) {      // one row per block iteration: no per-element division
        const float* row = g + (size_t)n * ld;
        float f = 0.f;
        for (int k = threadIdx.x; k < K; k += 256) f = fmaf(row[k], row[k], f);
        a += (double)f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += red[w];
        atomicAdd(accum, s);
    }
}

cudaError_t launch_grad_sumsq(const float* g, int ld, int N, int K, double* accum, cudaStream_t stream) {
    if (N <= 0 || K <= 0) return cudaSuccess;
    const int blocks = N < 592 ? N : 592;
    grad_sumsq_kernel<<<blocks, 256, 0, stream>>>(g, ld, N, K, accum);
    return cudaGetLastError();
}

// torch.optim.Adam (no amsgrad, no weight decay) after clip_grad_norm_: the clip coefficient is formed on the device
// from the accumulated squared norm, so a step never synchronises with the host.
//   g *= min(1, max_norm / (||g|| + 1e-6));  m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;
//   p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ prm, int ldp, const float* __restrict__ g, int ldg,
                                                   float* __restrict__ m, float* __restrict__ v, int N, int K,
                                                   float lr_base, int warmup, float beta1, float beta2, float eps,
                                                   long long step_host, const long long* __restrict__ step_dev,
                                                   float max_norm, const double* __restrict__ normsq) {
    // step-dependent scalars on the device (the step counter lives there so that a captured step can be replayed):
    // lr = lr_base * min(step / warmup, 1) (losses.py:44-46), bias corrections 1 - beta^step in fp64 like torch
    const long long step = step_dev ? *step_dev : step_host;
    const double wl = warmup > 0 ? fmin((double)step / (double)warmup, 1.0) : 1.0;
    const float step_size = (float)((double)lr_base * wl / (1.0 - pow((double)beta1, (double)step)));
    const float inv_sqrt_bias2 = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)step)));
    float coef = 1.f;
    if (max_norm >= 0.f && normsq) {
        const float c = max_norm / ((float)sqrt(*normsq) + 1e-6f);
        coef = c < 1.f ? c : 1.f;
    }
    const long long total = (long long)N * K;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int n = (int)(i / K), k = (int)(i % K);
        const float gr = g[(size_t)n * ldg + k] * coef;
        const float mm = beta1 * m[i] + (1.f - beta1) * gr;
        const float vv = beta2 * v[i] + (1.f - beta2) * gr * gr;
        m[i] = mm;
        v[i] = vv;
        const float denom = sqrtf(vv) * inv_sqrt_bias2 + eps;
        prm[(size_t)n * ldp + k] -= step_size * (mm / denom);
    }
}

cudaError_t launch_adam(float* prm, int ldp, const float* g, int ldg, float* m, float* v, int N, int K, float lr_base,
                        int warmup, float beta1, float beta2, float eps, long long step_host, const long long* step_dev,
                        float max_norm, const double* normsq, cudaStream_t stream) {
    const long long total = (long long)N * K;
    if (total <= 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    adam_kernel<<<(int)blocks, 256, 0, stream>>>(prm, ldp, g, ldg, m, v, N, K, lr_base, warmup, beta1, beta2, eps, step_host,
                                                 step_dev, max_norm, normsq);
    return cudaGetLastError();
}

}  // namespace idb
