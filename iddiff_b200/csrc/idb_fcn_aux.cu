// HBM-bound helpers around the tensor-core chain: perturbation sampling (dim_reduction.py:178-182),
// weight packing, the fp32 anchor chain (models/fcn.py:32-40 evaluated at the un-perturbed anchor)
// and raw-score materialisation (models/utils.py:264-267).
#include "idb_common.cuh"
#include "idb_kernels.h"

namespace idb {

// One thread = one 16-byte output chunk = 8 consecutive columns of one perturbation row.
// IDX = uint32_t when the chunk count fits 31 bits (every case benched): 32-bit divisions instead of three 64-bit ones.
template <typename IDX>
__global__ void __launch_bounds__(256) perturb_kernel(const PerturbParams p) {
    const IDX chunks = (IDX)(p.Dp >> 3);
    // slab rows: anchor-major, Kp rows per anchor (Kp = K: densely packed), padded with zero rows to a multiple of 128
    const long long Mtot = p.d_hi ? (((long long)p.n_anchors * p.Kp + 127) / 128) * 128 : (long long)p.n_anchors * p.Kp;
    const IDX total = (IDX)(Mtot * (long long)chunks);
    for (IDX i = (IDX)blockIdx.x * (IDX)blockDim.x + threadIdx.x; i < total; i += (IDX)gridDim.x * (IDX)blockDim.x) {
        const int ch = (int)(i % chunks);
        const IDX m = i / chunks;
        const int k = (int)(m % (IDX)p.Kp);
        const int a = (int)(m / (IDX)p.Kp);
        const int c0 = ch << 3;
        float dev[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) dev[j] = 0.f;
        if (a < p.n_anchors && k < p.K && c0 < p.D) {
            float zz[8];
            if (p.z != nullptr) {
                const float* zr = p.z + ((size_t)a * p.K + k) * p.D;
#pragma unroll
                for (int j = 0; j < 8; ++j) zz[j] = (c0 + j < p.D) ? __ldg(zr + c0 + j) : 0.f;
            } else {
                const uint64_t aid = p.anchor_id0 + (uint64_t)a;
                const uint2 key = make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const uint4 r = philox4x32_10(make_uint4((uint32_t)(ch * 2 + h), (uint32_t)k + p.row0, (uint32_t)aid,
                                                             (uint32_t)(aid >> 32)), key);
                    box_muller(r.x, r.y, zz[4 * h + 0], zz[4 * h + 1]);
                    box_muller(r.z, r.w, zz[4 * h + 2], zz[4 * h + 3]);
                }
            }
            const float* x0 = p.x0 + (size_t)a * p.D;
            float* xo = p.x_out ? p.x_out + ((size_t)a * p.K + k) * p.D : nullptr;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (c0 + j < p.D) {
                    // reference arithmetic: batch = mean + std * z, every op rounded to fp32
                    const float mean = __fmul_rn(p.mean_coeff, __ldg(x0 + c0 + j));
                    const float x = __fadd_rn(mean, __fmul_rn(p.std, zz[j]));
                    if (xo) xo[c0 + j] = x;
                    dev[j] = __fsub_rn(x, mean);
                }
            }
        }
        if (p.d_hi) {
            uint32_t h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                split_pair16(dev[2 * j] * p.d_scale, dev[2 * j + 1] * p.d_scale, p.fmt, h[j], l[j]);
            const size_t off = slab_offset_bytes((size_t)Mtot, (size_t)m, (size_t)c0);
            *reinterpret_cast<uint4*>(p.d_hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(p.d_lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
        }
    }
}

cudaError_t launch_perturb(const PerturbParams& p, cudaStream_t stream) {
    const long long rows = p.d_hi ? (((long long)p.n_anchors * p.Kp + 127) / 128) * 128 : (long long)p.n_anchors * p.Kp;
    const long long total = rows * (p.Dp >> 3);
    if (total <= 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 64) blocks = 148LL * 64;
    if (total + blocks * 256 < 0x7fffffffLL) perturb_kernel<uint32_t><<<(int)blocks, 256, 0, stream>>>(p);
    else perturb_kernel<unsigned long long><<<(int)blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

// ---------------------------------- weight packing ------------------------------------------
__global__ void __launch_bounds__(256) pack_weight_kernel(const float* __restrict__ w, int N, int K, int ldw, int Np,
                                                          int Kc, int fp16, float scale, uint8_t* __restrict__ hi,
                                                          uint8_t* __restrict__ lo) {
    const int chunks = Kc >> 3;
    const long long total = (long long)Np * chunks;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int ch = (int)(i % chunks);
        const int n = (int)(i / chunks);
        const int k0 = ch << 3;
        uint32_t h[4], l[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v0 = 0.f, v1 = 0.f;
            if (n < N) {
                if (k0 + 2 * j < K) v0 = w[(size_t)n * ldw + k0 + 2 * j];
                if (k0 + 2 * j + 1 < K) v1 = w[(size_t)n * ldw + k0 + 2 * j + 1];
            }
            split_pair16(v0 * scale, v1 * scale, fp16, h[j], l[j]);  // scale is a power of two: exact
        }
        const size_t off = slab_offset_bytes((size_t)Np, (size_t)n, (size_t)k0);
        *reinterpret_cast<uint4*>(hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

cudaError_t launch_pack_weight(const float* w, int N, int K, int ldw, int Np, int Kc, int fp16, float scale,
                               uint8_t* hi, uint8_t* lo, cudaStream_t stream) {
    const long long total = (long long)Np * (Kc >> 3);
    if (total <= 0) return cudaSuccess;
    pack_weight_kernel<<<(int)((total + 255) / 256), 256, 0, stream>>>(w, N, K, ldw, Np, Kc, fp16, scale, hi, lo);
    return cudaGetLastError();
}

// ---------------------------------- fp32 anchor chain ---------------------------------------
// out[a][n] = sum_k w[n][k] * f(in_scale * in[a][k]) + bias[n] + extra_col[n * extra_stride] * extra_scalar
// 64 x 64 output tile per CTA (16 x 16 threads, 4 x 4 each), 16-wide k tiles through shared memory.
__global__ void __launch_bounds__(256) anchor_linear_kernel(const float* __restrict__ in, int ld_in, int apply_elu,
                                                            float in_scale, const float* __restrict__ w, int ldw,
                                                            const float* __restrict__ bias,
                                                            const float* __restrict__ extra_col, int extra_stride,
                                                            float extra_scalar, float* __restrict__ out, int ld_out,
                                                            int A, int N, int K) {
    __shared__ float s_in[16][64 + 1];
    __shared__ float s_w[16][64 + 1];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int a0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int k0 = 0; k0 < K; k0 += 16) {
        for (int i = threadIdx.x; i < 64 * 16; i += 256) {
            const int r = i >> 4, kk = i & 15;
            float v = 0.f;
            if (a0 + r < A && k0 + kk < K) {
                v = __fmul_rn(in_scale, in[(size_t)(a0 + r) * ld_in + k0 + kk]);
                if (apply_elu == 1) v = v > 0.f ? v : expm1f(v);
                else if (apply_elu == 2) v = v / (1.f + expf(-v));  // SiLU (time-embedding path of the U-Nets)
            }
            s_in[kk][r] = v;
            float wv = 0.f;
            if (n0 + r < N && k0 + kk < K) wv = w[(size_t)(n0 + r) * ldw + k0 + kk];
            s_w[kk][r] = wv;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float av[4], wv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = s_in[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) wv[j] = s_w[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int a = a0 + ty * 4 + i;
        if (a >= A) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= N) continue;
            float v = acc[i][j];
            if (extra_col) v = fmaf(extra_col[(size_t)n * extra_stride], extra_scalar, v);
            if (bias) v += bias[n];
            out[(size_t)a * ld_out + n] = v;
        }
    }
}

cudaError_t launch_anchor_linear(const float* in, int ld_in, int apply_elu, float in_scale, const float* w, int ldw,
                                 const float* bias, const float* extra_col, int extra_stride, float extra_scalar,
                                 float* out, int ld_out, int n_anchors, int N, int K, cudaStream_t stream) {
    if (n_anchors <= 0 || N <= 0) return cudaSuccess;
    dim3 grid((N + 63) / 64, (n_anchors + 63) / 64);
    anchor_linear_kernel<<<grid, 256, 0, stream>>>(in, ld_in, apply_elu, in_scale, w, ldw, bias, extra_col,
                                                   extra_stride, extra_scalar, out, ld_out, n_anchors, N, K);
    return cudaGetLastError();
}

// ---------------------------------- raw scores ----------------------------------------------
__global__ void __launch_bounds__(256) fcn_scores_kernel(const float* __restrict__ a_out, int ld_a,
                                                         const float* __restrict__ d_out, int ld_d, int A, int K,
                                                         int Kp, int D, float std, float* __restrict__ score) {
    const long long total = (long long)A * K * D;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % D);
        const long long r = i / D;
        const int k = (int)(r % K);
        const int a = (int)(r / K);
        const float o = a_out[(size_t)a * ld_a + c] + d_out[((size_t)a * Kp + k) * ld_d + c];
        score[i] = __fdiv_rn(-o, std);
    }
}

cudaError_t launch_fcn_scores(const float* a_out, int ld_a, const float* d_out, int ld_d, int n_anchors, int K, int Kp,
                              int D, float std, float* score, cudaStream_t stream) {
    const long long total = (long long)n_anchors * K * D;
    if (total <= 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    fcn_scores_kernel<<<(int)blocks, 256, 0, stream>>>(a_out, ld_a, d_out, ld_d, n_anchors, K, Kp, D, std, score);
    return cudaGetLastError();
}

}  // namespace idb
