// Stand-alone Blackwell versions of the reference's two StyleGAN-derived native ops, with the op
// semantics of the reference extensions (so `op.upfirdn2d` / `op.fused_leaky_relu` can be re-pointed here):
//
//   upfirdn2d       op/upfirdn2d.cpp:12-22, op/upfirdn2d_kernel.cu:49-369, op/upfirdn2d.py:145-200 (reference)
//   fused_bias_act  op/fused_bias_act.cpp:11-20, op/fused_bias_act_kernel.cu:19-99 (reference)
//
// Both are HBM-bound elementwise/stencil ops: one read and one write of every activation.  The
// kernels below keep that traffic minimal (FIR taps and the input patch in shared memory, float4 I/O
// on the elementwise op, grids sized from the SM count); inside the U-Net engine the same arithmetic
// is fused into the operand-preparation pass of the adjacent convolution instead (idb_unet.cu).
#include "idb_common.cuh"
#include "idb_kernels.h"

namespace idb {

// out[m][oy][ox][c] = sum_{ky,kx} U[m][oy*down_y + ky][ox*down_x + kx][c] * k[kh-1-ky][kw-1-kx]
// U = input zero-stuffed by (up_y, up_x) and padded / cropped by (pad_*): U[y][x] = in[(y-pad_y0)/up_y][(x-pad_x0)/up_x]
// when both divisions are exact and in range, else 0  (upfirdn2d_native, op/upfirdn2d.py:162-200).
// One CTA = one `major` slice x a TY x TX output tile (all minor); the input patch the tile touches is staged
// once in shared memory, taps are read from shared memory, accumulation order is (ky, kx) ascending in fp32.
constexpr int UF_TX = 32, UF_TY = 8;

__global__ void __launch_bounds__(UF_TX* UF_TY) upfirdn2d_kernel(const UpFirDnParams p) {
    extern __shared__ float uf_smem[];
    float* s_k = uf_smem;                  // [kh*kw] flipped taps
    float* s_in = uf_smem + p.kh * p.kw;   // [patch_h][patch_w][minor_tile]
    const int tiles_x = (p.out_w + UF_TX - 1) / UF_TX;
    const int tiles_y = (p.out_h + UF_TY - 1) / UF_TY;
    const int tid = threadIdx.y * UF_TX + threadIdx.x;
    for (int i = tid; i < p.kh * p.kw; i += UF_TX * UF_TY) {
        const int ky = i / p.kw, kx = i % p.kw;
        s_k[i] = p.kernel[(p.kh - 1 - ky) * p.kw + (p.kw - 1 - kx)];
    }
    const long long n_work = (long long)p.major * tiles_y * tiles_x;
    for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
        const int tx = (int)(w % tiles_x), ty = (int)((w / tiles_x) % tiles_y);
        const long long m = w / ((long long)tiles_x * tiles_y);
        const int ox0 = tx * UF_TX, oy0 = ty * UF_TY;
        // rows/cols of the input that the tile can touch
        const int uy0 = oy0 * p.down_y - p.pad_y0, ux0 = ox0 * p.down_x - p.pad_x0;  // U coords minus pad -> stuffed coords
        const int uy1 = (oy0 + UF_TY - 1) * p.down_y + p.kh - 1 - p.pad_y0;
        const int ux1 = (ox0 + UF_TX - 1) * p.down_x + p.kw - 1 - p.pad_x0;
        const int iy0 = uy0 <= 0 ? 0 : (uy0 + p.up_y - 1) / p.up_y;
        const int ix0 = ux0 <= 0 ? 0 : (ux0 + p.up_x - 1) / p.up_x;
        int iy1 = uy1 < 0 ? -1 : uy1 / p.up_y;
        int ix1 = ux1 < 0 ? -1 : ux1 / p.up_x;
        if (iy1 > p.in_h - 1) iy1 = p.in_h - 1;
        if (ix1 > p.in_w - 1) ix1 = p.in_w - 1;
        const int ph = iy1 - iy0 + 1, pw = ix1 - ix0 + 1;
        for (int c0 = 0; c0 < p.minor; c0 += p.minor_tile) {
            const int mc = min(p.minor_tile, p.minor - c0);
            __syncthreads();
            if (ph > 0 && pw > 0) {
                const int row_elems = pw * mc;
                for (int i = tid; i < ph * row_elems; i += UF_TX * UF_TY) {
                    const int r = i / row_elems, rem = i % row_elems;
                    const int x = rem / mc, c = rem % mc;
                    s_in[(r * pw + x) * mc + c] =
                        __ldg(p.in + (((m * p.in_h + iy0 + r) * p.in_w + ix0 + x) * (long long)p.minor + c0 + c));
                }
            }
            __syncthreads();
            const int oy = oy0 + threadIdx.y, ox = ox0 + threadIdx.x;
            if (oy < p.out_h && ox < p.out_w) {
                for (int c = 0; c < mc; ++c) {
                    float acc = 0.f;
                    for (int ky = 0; ky < p.kh; ++ky) {
                        const int sy = oy * p.down_y + ky - p.pad_y0;
                        if (sy < 0 || sy % p.up_y) continue;
                        const int iy = sy / p.up_y;
                        if (iy < iy0 || iy > iy1) continue;
                        for (int kx = 0; kx < p.kw; ++kx) {
                            const int sx = ox * p.down_x + kx - p.pad_x0;
                            if (sx < 0 || sx % p.up_x) continue;
                            const int ix = sx / p.up_x;
                            if (ix < ix0 || ix > ix1) continue;
                            acc = fmaf(s_in[((iy - iy0) * pw + (ix - ix0)) * mc + c], s_k[ky * p.kw + kx], acc);
                        }
                    }
                    p.out[((m * p.out_h + oy) * p.out_w + ox) * (long long)p.minor + c0 + c] = acc;
                }
            }
        }
    }
}

// Channel-minor fast path (minor % 4 == 0, the NHWC activations of the U-Net engine): one thread = one output pixel x
// 4 channels; taps are gathered straight from global memory as float4 (neighbouring outputs share them through L1/L2),
// consecutive threads walk the channel axis so every load and store is a coalesced 16-byte access.  Same (ky, kx)
// ascending fp32 accumulation order as the tiled kernel above.
__global__ void __launch_bounds__(256) upfirdn2d_nhwc_kernel(const UpFirDnParams p) {
    __shared__ float s_k[1024];
    for (int i = threadIdx.x; i < p.kh * p.kw; i += 256) {
        const int ky = i / p.kw, kx = i % p.kw;
        s_k[i] = p.kernel[(p.kh - 1 - ky) * p.kw + (p.kw - 1 - kx)];
    }
    __syncthreads();
    const unsigned c4n = (unsigned)p.minor >> 2;
    const unsigned total = (unsigned)p.major * p.out_h * p.out_w * c4n;
    for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
        const unsigned c4 = i % c4n;
        unsigned r = i / c4n;
        const int ox = (int)(r % (unsigned)p.out_w); r /= (unsigned)p.out_w;
        const int oy = (int)(r % (unsigned)p.out_h);
        const unsigned m = r / (unsigned)p.out_h;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int ky = 0; ky < p.kh; ++ky) {
            const int sy = oy * p.down_y + ky - p.pad_y0;
            if (sy < 0 || sy % p.up_y) continue;
            const int iy = sy / p.up_y;
            if (iy >= p.in_h) continue;
            for (int kx = 0; kx < p.kw; ++kx) {
                const int sx = ox * p.down_x + kx - p.pad_x0;
                if (sx < 0 || sx % p.up_x) continue;
                const int ix = sx / p.up_x;
                if (ix >= p.in_w) continue;
                const float4 v = __ldg(reinterpret_cast<const float4*>(
                    p.in + (((size_t)m * p.in_h + iy) * p.in_w + ix) * p.minor) + c4);
                const float k = s_k[ky * p.kw + kx];
                acc.x = fmaf(v.x, k, acc.x); acc.y = fmaf(v.y, k, acc.y);
                acc.z = fmaf(v.z, k, acc.z); acc.w = fmaf(v.w, k, acc.w);
            }
        }
        reinterpret_cast<float4*>(p.out + (((size_t)m * p.out_h + oy) * p.out_w + ox) * p.minor)[c4] = acc;
    }
}

cudaError_t launch_upfirdn2d(UpFirDnParams p, int num_sms, cudaStream_t stream) {
    p.out_h = (p.in_h * p.up_y + p.pad_y0 + p.pad_y1 - p.kh) / p.down_y + 1;
    p.out_w = (p.in_w * p.up_x + p.pad_x0 + p.pad_x1 - p.kw) / p.down_x + 1;
    if (p.out_h <= 0 || p.out_w <= 0 || p.major <= 0) return cudaSuccess;
    {
        const long long total4 = (long long)p.major * p.out_h * p.out_w * (p.minor >> 2);
        const bool al = ((reinterpret_cast<uintptr_t>(p.in) | reinterpret_cast<uintptr_t>(p.out)) & 15) == 0;
        if (p.minor % 4 == 0 && al && total4 < (1ll << 32)) {
            long long grid = (total4 + 255) / 256;
            if (grid > (long long)num_sms * 32) grid = (long long)num_sms * 32;
            upfirdn2d_nhwc_kernel<<<(unsigned)grid, 256, 0, stream>>>(p);
            return cudaGetLastError();
        }
    }
    const int patch_h = ((UF_TY - 1) * p.down_y + p.kh - 1) / p.up_y + 2;
    const int patch_w = ((UF_TX - 1) * p.down_x + p.kw - 1) / p.up_x + 2;
    const size_t budget = 96 * 1024;
    size_t per_minor = (size_t)patch_h * patch_w * sizeof(float);
    int mt = (int)((budget - (size_t)p.kh * p.kw * sizeof(float)) / per_minor);
    if (mt < 1) return cudaErrorInvalidValue;
    if (mt > p.minor) mt = p.minor;
    p.minor_tile = mt;
    const size_t smem = (size_t)p.kh * p.kw * sizeof(float) + per_minor * mt;
    cudaError_t e = cudaFuncSetAttribute(upfirdn2d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget);
    if (e != cudaSuccess) return e;
    const long long tiles = (long long)p.major * ((p.out_h + UF_TY - 1) / UF_TY) * ((p.out_w + UF_TX - 1) / UF_TX);
    long long grid = (long long)num_sms * 8;
    if (grid > tiles) grid = tiles;
    upfirdn2d_kernel<<<(unsigned)grid, dim3(UF_TX, UF_TY), smem, stream>>>(p);
    return cudaGetLastError();
}

// out[i] = act(x[i] + b[(i / step_b) % size_b]) * scale  with the reference's (act, grad) table:
//   act 1 = linear, 3 = leaky ReLU(alpha); grad 0 = forward, 1 = first derivative w.r.t. x times incoming
//   gradient x (sign taken from `ref`), 2 = second derivative (0).   op/fused_bias_act_kernel.cu:37-49.
__device__ __forceinline__ float bias_act_one(float x, float ref, int code, float alpha, float scale) {
    float y;
    switch (code) {
        default:
        case 10: case 11: y = x; break;
        case 12: case 32: y = 0.f; break;
        case 30: y = (x > 0.f) ? x : x * alpha; break;
        case 31: y = (ref > 0.f) ? x : x * alpha; break;
    }
    return y * scale;
}

__global__ void __launch_bounds__(256) fused_bias_act_kernel(float* __restrict__ out, const float* __restrict__ x,
                                                             const float* __restrict__ b, const float* __restrict__ ref,
                                                             int code, float alpha, float scale, long long size_x,
                                                             int step_b, int size_b, int vec) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {  // size_x % 4 == 0, step_b % 4 == 0 (or no bias), 16-byte aligned: one bias value per float4
        const long long n4 = size_x >> 2;
        for (; i < n4; i += stride) {
            float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
            const float bb = b ? __ldg(b + ((i * 4) / step_b) % size_b) : 0.f;
            float4 r = ref ? __ldg(reinterpret_cast<const float4*>(ref) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            v.x = bias_act_one(v.x + bb, r.x, code, alpha, scale);
            v.y = bias_act_one(v.y + bb, r.y, code, alpha, scale);
            v.z = bias_act_one(v.z + bb, r.z, code, alpha, scale);
            v.w = bias_act_one(v.w + bb, r.w, code, alpha, scale);
            reinterpret_cast<float4*>(out)[i] = v;
        }
    } else {
        for (; i < size_x; i += stride) {
            float v = __ldg(x + i);
            if (b) v += __ldg(b + (i / step_b) % size_b);
            out[i] = bias_act_one(v, ref ? __ldg(ref + i) : 0.f, code, alpha, scale);
        }
    }
}

cudaError_t launch_fused_bias_act(const float* x, const float* b, const float* ref, int act, int grad, float alpha,
                                  float scale, long long size_x, int step_b, int size_b, float* out, int num_sms,
                                  cudaStream_t stream) {
    if (size_x <= 0) return cudaSuccess;
    auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const int vec = (size_x % 4 == 0) && (!b || step_b % 4 == 0) && al(x) && al(out) && (!ref || al(ref));
    const long long items = vec ? size_x / 4 : size_x;
    long long grid = (items + 255) / 256;
    const long long cap = (long long)num_sms * 16;
    if (grid > cap) grid = cap;
    fused_bias_act_kernel<<<(unsigned)grid, 256, 0, stream>>>(out, x, b, ref, act * 10 + grad, alpha, scale, size_x,
                                                              step_b > 0 ? step_b : 1, size_b > 0 ? size_b : 1, vec);
    return cudaGetLastError();
}

}  // namespace idb
