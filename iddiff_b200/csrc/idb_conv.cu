// Implicit-GEMM convolution on the 5th-gen tensor cores for the U-Net score networks
// (reference: every nn.Conv2d 3x3 / 1x1 and NIN of models/ddpm.py, models/ncsnpp.py, models/BeatGANsUNET.py --
//  layers.py:93-100 ddpm_conv3x3, layers.py:555-564 NIN, layerspp.py conv3x3/conv1x1; cuDNN in the reference).
//
//   out[n,h,w,co] = out_scale * ( acc_scale * sum_{tap,ci} A[plane(tap)][n, h+dy(tap), w+dx(tap), ci] * Wt[co][tap][ci]
//                                 + bias[co] + residual[n,h,w,co] )
//
// A is the activation in NHWC as hi/lo fp16 pairs (written by the operand-preparation pass, idb_unet.cu, which has
// already applied GroupNorm + activation + up/down-sampling), so the M dimension of the GEMM is "output pixels" and
// each filter tap is the SAME tensor shifted by (dy, dx): one 5-D TMA tensor map (C, W, H, N, plane) with
// SWIZZLE_128B delivers the [128 pixels x 64 channels] operand tile of a tap straight into the K-major shared-memory
// image tcgen05.mma reads, and the hardware's out-of-bounds zero fill IS the convolution's zero padding.  No im2col
// buffer exists anywhere.  Stride-2 convolutions read four parity planes (space-to-depth done by the producer).
//
// Same split-precision scheme as the FCN chain (idb_gemm.cu): 3 tcgen05.mma (lo.hi + hi.lo + hi.hi) per product,
// fp32 accumulation in TMEM, power-of-two operand scales undone exactly on the accumulator.
//
// Kernel shape: persistent, one CTA per SM, 64 + 512 threads: warp 0 TMA producer, warp 1 MMA issuer, 16 epilogue
// warps; TMEM double-buffered (2 x BN fp32 columns) so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "idb_common.cuh"
#include "idb_kernels.h"

namespace idb {

namespace {
constexpr int CBM = 128;
constexpr int CBK = 64;
constexpr uint32_t CA_TILE = CBM * 128;  // 16 KB: 128 pixels x 64 fp16 channels
constexpr int C_EPI_WARPS = 16;
constexpr int C_EPI_THREADS = C_EPI_WARPS * 32;
constexpr int C_THREADS = 64 + C_EPI_THREADS;
constexpr int C_SEG = 8;   // k-blocks (512 product terms, 96 tcgen05.mma) accumulated in TMEM before a flush to registers

// BN output channels x (MT x 128) output pixels per CTA tile.  MT = 2 ("two M-subtiles share one weight tile") is the
// shape for Cout = 128 layers: a 128 x 128 tile moves 64 KB per k-block for 12 MMAs of N = 128 (85 B per tensor-pipe cycle,
// more than the L2 feeds: ncu 68 % pipe activity), a 256 x 128 tile moves 96 KB for 24 (62 B per cycle, like 128 x 256).
template <int BN, int MT> struct ConvCfg {
    static constexpr uint32_t B_TILE = BN * 128;
    static constexpr uint32_t A_BYTES = MT * 2 * CA_TILE;
    static constexpr uint32_t STAGE = A_BYTES + 2 * B_TILE;
    static constexpr int STAGES = STAGE >= 96 * 1024 ? 2 : 3;
    static constexpr int TCOLS = MT * BN;                      // fp32 accumulator columns of one TMEM buffer
    static constexpr size_t SMEM = 1024 + (size_t)STAGES * STAGE + 256;
};

__device__ __forceinline__ void tma_tensor5d_g2s(void* smem_dst, const CUtensorMap* tm, int c0, int c1, int c2, int c3,
                                                 int c4, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::
            "r"(smem_u32(smem_dst)),
        "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}
}  // namespace

__device__ __forceinline__ void st_global_256f(float* dst, const float (&o)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "f"(o[0]), "f"(o[1]), "f"(o[2]),
                 "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]), "f"(o[7])
                 : "memory");
}
__device__ __forceinline__ void ld_global_nc_256(const float* src, float (&q)[8]) {
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(q[0]), "=f"(q[1]), "=f"(q[2]), "=f"(q[3]), "=f"(q[4]), "=f"(q[5]), "=f"(q[6]), "=f"(q[7])
                 : "l"(src));
}

// Transposing butterfly over the 32 per-lane values of one quantity: afterwards lane l holds in a[0] the total of value
// index l over the warp's 32 lanes (full warp), or -- half_warp, two images of 16 pixels per warp -- in a[0], a[1] the
// totals of indices 2 (l & 15) + {0, 1} over the 16 lanes of its half.  31 (30) shuffles instead of 160.
__device__ __forceinline__ void stat_butterfly(float (&a)[32], int lane, bool half_warp) {
    if (!half_warp) {
        const bool up = lane & 16;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float keep = up ? a[i + 16] : a[i], send = up ? a[i] : a[i + 16];
            a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
#pragma unroll
        for (int m = 8, w = 16; m >= 1; m >>= 1, w >>= 1) {
            const bool u = lane & m;
#pragma unroll
            for (int i = 0; i < w / 2; ++i) {
                const float keep = u ? a[i + w / 2] : a[i], send = u ? a[i] : a[i + w / 2];
                a[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
            }
        }
    } else {
#pragma unroll
        for (int m = 8, w = 32; m >= 1; m >>= 1, w >>= 1) {
            const bool u = lane & m;
#pragma unroll
            for (int i = 0; i < w / 2; ++i) {
                const float keep = u ? a[i + w / 2] : a[i], send = u ? a[i] : a[i + w / 2];
                a[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
            }
        }
    }
}

template <int BN, int MT>
__global__ void __launch_bounds__(C_THREADS, 1) conv_fp16x3_kernel(const __grid_constant__ ConvParams p) {
    using Cfg = ConvCfg<BN, MT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)Cfg::STAGES * Cfg::STAGE);
    uint64_t* full = bars;
    uint64_t* empty = bars + Cfg::STAGES;
    uint64_t* tfull = bars + 2 * Cfg::STAGES;
    uint64_t* tempty = bars + 2 * Cfg::STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tiles_w = p.W / p.BW, tiles_h = p.H / p.BH;
    const int tiles_img = (p.N + p.BI - 1) / p.BI;
    const int m_tiles = tiles_w * tiles_h * tiles_img;
    const int n_tiles_n = p.Np / BN;
    const int n_tiles = ((m_tiles + MT - 1) / MT) * n_tiles_n;   // a CTA tile = MT consecutive 128-pixel tiles x BN channels
    const int kc = p.Cp / CBK;          // 64-channel chunks per tap
    const int n_kb = p.n_taps * kc;

    if (threadIdx.x == 0) {
        for (int s = 0; s < Cfg::STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tfull[s], 1);
            mbar_init(&tempty[s], C_EPI_THREADS);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 2 * Cfg::TCOLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int mt0 = (t / n_tiles_n) * MT;
                const size_t n0 = (size_t)(t % n_tiles_n) * BN;
                for (int kb = 0; kb < n_kb; ++kb) {
                    const int tap = kb / kc, c0 = (kb % kc) * CBK;
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* st = smem + (size_t)stage * Cfg::STAGE;
                    mbar_arrive_expect_tx(&full[stage], Cfg::STAGE);
                    int i_first = 0;
#pragma unroll
                    for (int sub = 0; sub < MT; ++sub) {
                        // a sub-tile past the last one (odd tile count) reads image indices >= N: TMA fills zeros
                        const int mt = mt0 + sub;
                        const int w0 = (mt % tiles_w) * p.BW;
                        const int h0 = ((mt / tiles_w) % tiles_h) * p.BH;
                        const int i0 = (mt / (tiles_w * tiles_h)) * p.BI;
                        if (sub == 0) i_first = i0;
                        const int cw = w0 + p.tap_dx[tap], ch = h0 + p.tap_dy[tap], cp = p.tap_plane[tap];
                        tma_tensor5d_g2s(st + sub * 2 * CA_TILE, &p.tm_hi, c0, cw, ch, i0, cp, &full[stage]);
                        tma_tensor5d_g2s(st + sub * 2 * CA_TILE + CA_TILE, &p.tm_lo, c0, cw, ch, i0, cp, &full[stage]);
                    }
                    const size_t b_off = ((size_t)kb * p.Np + n0) * 128 + (size_t)i_first * (size_t)p.w_img_stride;
                    tma_bulk_g2s(st + Cfg::A_BYTES, p.w_hi + b_off, Cfg::B_TILE, &full[stage]);
                    tma_bulk_g2s(st + Cfg::A_BYTES + Cfg::B_TILE, p.w_lo + b_off, Cfg::B_TILE, &full[stage]);
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_f16(CBM, BN, 0u, 0u);  // fp16 x fp16 -> fp32
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;   // counts (tile, K-segment) accumulation phases: TMEM buffer it & 1
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                for (int kb0 = 0; kb0 < n_kb; kb0 += C_SEG, ++it) {
                    const int as = it & 1;
                    const uint32_t aphase = (it >> 1) & 1;
                    mbar_wait(&tempty[as], aphase ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + as * Cfg::TCOLS;
                    const int kb1 = kb0 + C_SEG < n_kb ? kb0 + C_SEG : n_kb;
                    for (int kb = kb0; kb < kb1; ++kb) {
                        mbar_wait(&full[stage], phase);
                        tc_fence_after();
                        const uint32_t sa = smem_u32(smem + (size_t)stage * Cfg::STAGE);
                        const uint64_t bh = umma_desc_sw128(sa + Cfg::A_BYTES);
                        const uint64_t bl = umma_desc_sw128(sa + Cfg::A_BYTES + Cfg::B_TILE);
#pragma unroll
                        for (int sub = 0; sub < MT; ++sub) {     // the sub-tiles share the weight tile of this k-block
                            const uint64_t ah = umma_desc_sw128(sa + sub * 2 * CA_TILE);
                            const uint64_t al = umma_desc_sw128(sa + sub * 2 * CA_TILE + CA_TILE);
#pragma unroll
                            for (int k = 0; k < CBK / 16; ++k) {
                                const uint64_t adv = (uint64_t)((k * 32) >> 4);
                                umma_bf16(d_tmem + sub * BN, al + adv, bh + adv, idesc, ((kb - kb0) | k) != 0);
                                umma_bf16(d_tmem + sub * BN, ah + adv, bl + adv, idesc, 1);
                                umma_bf16(d_tmem + sub * BN, ah + adv, bh + adv, idesc, 1);
                            }
                        }
                        umma_commit(&empty[stage]);
                        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                    }
                    umma_commit(&tfull[as]);
                }
            }
        }
    } else {
        // the 32-column chunks of the CTA tile this warp owns: MT = 1: columns cc = cgroup + 4 ci of the one sub-tile;
        // MT = 2 (BN = 128): column chunk cgroup of sub-tile ci
        constexpr int NCH = MT * BN / 128;
        const int quarter = warp & 3;
        const int cgroup = (warp - 2) >> 2;
        const int r = quarter * 32 + lane;  // row of a sub-tile = TMEM lane = output pixel
        const bool ptrs32 = ((reinterpret_cast<uintptr_t>(p.out) | reinterpret_cast<uintptr_t>(p.residual)) & 31) == 0;
        const int stat_pimg = p.BW * p.BH;                      // pixels of one image inside a tile
        const int stat_ppw = stat_pimg < 32 ? stat_pimg : 32;   // pixels per warp that belong to one image (32 or 16)
        const int stat_slots = (p.H * p.W) / stat_ppw;
        int it = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
            const int mt0 = (t / n_tiles_n) * MT;
            const int n0 = (t % n_tiles_n) * BN;
            // The tensor core adds into its fp32 accumulator with truncation, and the error of a long chain grows
            // linearly with its length (measured: 3.6e-7 / 1.4e-6 / 6.0e-6 rms of the output scale at 576 / 2304 / 9216
            // terms against 3e-7 for an fp32 FMA chain).  So a TMEM accumulator only ever holds C_SEG k-blocks; the
            // partial sums are added here in registers with round-to-nearest, which keeps every convolution at the
            // short-chain error whatever its depth.  The TMEM buffer is released as soon as it is in registers.
            uint32_t acc[NCH][32];
            for (int kb0 = 0; kb0 < n_kb; kb0 += C_SEG, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(&tfull[as], aphase);
                tc_fence_after();
                const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * Cfg::TCOLS;
#pragma unroll
                for (int ci = 0; ci < NCH; ++ci) {
                    const int cc = MT == 2 ? cgroup + (BN / 32) * ci : cgroup + 4 * ci;   // column chunk inside the TMEM buffer
                    if (n0 + (MT == 2 ? cgroup : cc) * 32 >= p.Cout) continue;  // warp-uniform: padded output channels
                    if (kb0 == 0) {
                        tmem_ld32(t_row + cc * 32, acc[ci]);
                        tmem_ld_wait();
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {   // 8 columns at a time: the partial sums never need 32 more registers
                            uint32_t v[8];
                            tmem_ld8(t_row + cc * 32 + q * 8, v);
                            tmem_ld_wait();
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                acc[ci][8 * q + j] = __float_as_uint(__uint_as_float(acc[ci][8 * q + j]) + __uint_as_float(v[j]));
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&tempty[as]);
            }
#pragma unroll
            for (int ci = 0; ci < NCH; ++ci) {
                const int c0 = n0 + (MT == 2 ? cgroup : cgroup + 4 * ci) * 32;
                if (c0 >= p.Cout) continue;  // warp-uniform: padded output channels
                uint32_t (&v)[32] = acc[ci];
                // output pixel of this thread in sub-tile mt (MT = 2: sub-tile ci; one past the end holds images >= N)
                const int mt = mt0 + (MT == 2 ? ci : 0);
                const int w = (mt % tiles_w) * p.BW + r % p.BW;
                const int h = ((mt / tiles_w) % tiles_h) * p.BH + (r / p.BW) % p.BH;
                const int img = (mt / (tiles_w * tiles_h)) * p.BI + r / (p.BW * p.BH);
                const bool row_ok = img < p.N;
                const size_t pix = ((size_t)img * p.H + h) * p.W + w;
                // statistics slot of this warp's 32 pixels inside their image (see the stats block below)
                const int img_warp = (mt / (tiles_w * tiles_h)) * p.BI + (quarter * 32) / stat_pimg;
                const bool row_ok_warp = img_warp < p.N;
                const int stat_slot = (p.BI == 1 ? (mt % (tiles_w * tiles_h)) * 4 + quarter : ((quarter * 32) % stat_pimg) / 32);
                if (row_ok) {
                    float* dst = p.out + pix * p.ldo + c0;
                    const float* res = p.residual ? p.residual + pix * p.ld_res + c0 : nullptr;
                    if (c0 + 32 <= p.Cout && (p.ldo & 7) == 0 && (!res || (p.ld_res & 7) == 0) && ptrs32) {
                        // a thread owns 128 contiguous bytes of its pixel: 256-bit accesses move whole 32-byte sectors
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float o[8], q[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                const float b = p.bias ? __ldg(p.bias + c0 + 8 * j + e) : 0.f;
                                o[e] = __uint_as_float(v[8 * j + e]) * p.acc_scale + b;
                            }
                            if (res) {
                                ld_global_nc_256(res + 8 * j, q);
#pragma unroll
                                for (int e = 0; e < 8; ++e) o[e] += q[e];
                            }
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                o[e] *= p.out_scale;
                                v[8 * j + e] = __float_as_uint(o[e]);
                            }
                            st_global_256f(dst + 8 * j, o);
                        }
                    } else if (c0 + 32 <= p.Cout && (p.ldo & 3) == 0 && (!res || (p.ld_res & 3) == 0)) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float4 o;
                            float4 b = p.bias ? __ldg(reinterpret_cast<const float4*>(p.bias + c0) + j)
                                              : make_float4(0.f, 0.f, 0.f, 0.f);
                            o.x = __uint_as_float(v[4 * j]) * p.acc_scale + b.x;
                            o.y = __uint_as_float(v[4 * j + 1]) * p.acc_scale + b.y;
                            o.z = __uint_as_float(v[4 * j + 2]) * p.acc_scale + b.z;
                            o.w = __uint_as_float(v[4 * j + 3]) * p.acc_scale + b.w;
                            if (res) {
                                const float4 q = __ldg(reinterpret_cast<const float4*>(res) + j);
                                o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w;
                            }
                            o.x *= p.out_scale; o.y *= p.out_scale; o.z *= p.out_scale; o.w *= p.out_scale;
                            reinterpret_cast<float4*>(dst)[j] = o;
                            v[4 * j] = __float_as_uint(o.x); v[4 * j + 1] = __float_as_uint(o.y);
                            v[4 * j + 2] = __float_as_uint(o.z); v[4 * j + 3] = __float_as_uint(o.w);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            float o = 0.f;
                            if (c0 + j < p.Cout) {
                                o = __uint_as_float(v[j]) * p.acc_scale + (p.bias ? __ldg(p.bias + c0 + j) : 0.f);
                                if (res) o += __ldg(res + j);
                                o *= p.out_scale;
                                dst[j] = o;
                            }
                            v[j] = __float_as_uint(o);
                        }
                    }
                }
                if (p.stats) {
                    // GroupNorm statistics of the tensor this convolution writes (layers.py:638,648: the NEXT block's
                    // nn.GroupNorm), from the registers that already hold it: per (image, 32-pixel slot, channel) sums
                    // of x and x^2.  A transposing butterfly leaves lane l with the total of channel c0 + l over the
                    // warp's 32 pixels in 31 shuffles per quantity; fixed slots, no atomics -> bit-reproducible.
                    const bool half_warp = stat_ppw == 16;   // 4x4 images: a warp holds two images of 16 pixels
                    float s0[2], q0[2];
                    {
                        float a[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) a[j] = row_ok ? __uint_as_float(v[j]) : 0.f;
                        stat_butterfly(a, lane, half_warp);
                        s0[0] = a[0]; s0[1] = a[1];
                    }
                    {
                        float a[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float x = row_ok ? __uint_as_float(v[j]) : 0.f;
                            a[j] = x * x;
                        }
                        stat_butterfly(a, lane, half_warp);
                        q0[0] = a[0]; q0[1] = a[1];
                    }
                    if (!half_warp) {
                        if (row_ok_warp && c0 + lane < p.Cout) {
                            float2* dst2 = reinterpret_cast<float2*>(p.stats) +
                                           ((size_t)img_warp * stat_slots + stat_slot) * p.Cout + c0 + lane;
                            *dst2 = make_float2(s0[0], q0[0]);
                        }
                    } else {
                        const int ch = c0 + 2 * (lane & 15);
                        if (row_ok && ch < p.Cout) {   // lanes of one half share img
                            float2* dst2 = reinterpret_cast<float2*>(p.stats) + (size_t)img * p.Cout + ch;
                            dst2[0] = make_float2(s0[0], q0[0]);
                            if (ch + 1 < p.Cout) dst2[1] = make_float2(s0[1], q0[1]);
                        }
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 2 * Cfg::TCOLS);
}

template <int BN, int MT>
static cudaError_t launch_conv_bn(const ConvParams& p, int num_sms, cudaStream_t stream) {
    using Cfg = ConvCfg<BN, MT>;
    const int m_tiles = (p.W / p.BW) * (p.H / p.BH) * ((p.N + p.BI - 1) / p.BI);
    const int n_tiles = ((m_tiles + MT - 1) / MT) * (p.Np / BN);
    const int grid = n_tiles < num_sms ? n_tiles : num_sms;
    if (grid <= 0) return cudaSuccess;
    cudaError_t e =
        cudaFuncSetAttribute(conv_fp16x3_kernel<BN, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM);
    if (e != cudaSuccess) return e;
    conv_fp16x3_kernel<BN, MT><<<grid, C_THREADS, Cfg::SMEM, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_conv(const ConvParams& p, int num_sms, cudaStream_t stream) {
    if (p.Np % 256 == 0) return launch_conv_bn<256, 1>(p, num_sms, stream);
    // Cout = 128 (384) layers: pair two pixel tiles on one weight tile when there are enough tiles to keep every SM busy
    // (and the weights are shared by all images)
    const int m_tiles = (p.W / p.BW) * (p.H / p.BH) * ((p.N + p.BI - 1) / p.BI);
    if (p.w_img_stride == 0 && (long long)(m_tiles / 2) * (p.Np / 128) >= 2LL * num_sms) return launch_conv_bn<128, 2>(p, num_sms, stream);
    return launch_conv_bn<128, 1>(p, num_sms, stream);
}

}  // namespace idb
