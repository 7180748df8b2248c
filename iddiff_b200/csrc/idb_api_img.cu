// C ABI, image-network side (U-Net score models): see include/iddiff_b200.h.
#include "../../include/iddiff_b200.h"
#include "idb_kernels.h"

using idb::api_check;
using idb::api_fail;

extern "C" {

int idb_upfirdn2d(const float* input, const float* kernel, int major, int in_h, int in_w, int minor, int kernel_h,
                  int kernel_w, int up_x, int up_y, int down_x, int down_y, int pad_x0, int pad_x1, int pad_y0,
                  int pad_y1, float* out, void* stream) {
    if (!input || !kernel || !out) return api_fail("idb_upfirdn2d: null pointer");
    if (major < 0 || in_h <= 0 || in_w <= 0 || minor <= 0 || kernel_h <= 0 || kernel_w <= 0)
        return api_fail("idb_upfirdn2d: bad shape");
    if (up_x <= 0 || up_y <= 0 || down_x <= 0 || down_y <= 0) return api_fail("idb_upfirdn2d: up/down must be >= 1");
    if (kernel_h * kernel_w > 1024) return api_fail("idb_upfirdn2d: FIR kernel larger than 1024 taps");
    idb::UpFirDnParams p;
    p.in = input; p.kernel = kernel; p.out = out;
    p.major = major; p.in_h = in_h; p.in_w = in_w; p.minor = minor; p.kh = kernel_h; p.kw = kernel_w;
    p.up_x = up_x; p.up_y = up_y; p.down_x = down_x; p.down_y = down_y;
    p.pad_x0 = pad_x0; p.pad_x1 = pad_x1; p.pad_y0 = pad_y0; p.pad_y1 = pad_y1;
    p.out_h = p.out_w = p.minor_tile = 0;
    return api_check(idb::launch_upfirdn2d(p, idb::api_num_sms(), static_cast<cudaStream_t>(stream)), "idb_upfirdn2d");
}

int idb_fused_bias_act(const float* input, const float* bias, const float* refer, int act, int grad, float alpha,
                       float scale, long long size_x, int step_b, int size_b, float* out, void* stream) {
    if (!input || !out || size_x < 0) return api_fail("idb_fused_bias_act: bad argument");
    if (act != 1 && act != 3) return api_fail("idb_fused_bias_act: act must be 1 (linear) or 3 (leaky ReLU)");
    if (grad < 0 || grad > 2) return api_fail("idb_fused_bias_act: grad must be 0, 1 or 2");
    if (bias && (step_b <= 0 || size_b <= 0)) return api_fail("idb_fused_bias_act: bias needs step_b, size_b > 0");
    return api_check(idb::launch_fused_bias_act(input, bias, refer, act, grad, alpha, scale, size_x, step_b, size_b, out,
                                                idb::api_num_sms(), static_cast<cudaStream_t>(stream)),
                     "idb_fused_bias_act");
}

int idb_groupnorm_stats(const float* src0, int C0, const float* src1, int C1, int N, int HW, int groups, double eps,
                        double* sums, float* mean_rstd, void* stream) {
    if (!src0 || !sums || !mean_rstd || C0 <= 0 || C1 < 0 || (C1 > 0 && !src1) || N < 0 || HW <= 0 || groups <= 0)
        return api_fail("idb_groupnorm_stats: bad argument");
    if ((C0 + C1) % groups) return api_fail("idb_groupnorm_stats: channels must divide into groups");
    if (groups > 1024) return api_fail("idb_groupnorm_stats: at most 1024 groups");
    return api_check(idb::launch_gn_stats(src0, C0, src1, C1, N, HW, groups, eps, sums, mean_rstd, idb::api_num_sms(),
                                          static_cast<cudaStream_t>(stream)),
                     "idb_groupnorm_stats");
}

int idb_conv_prep(const float* src0, int C0, const float* src1, int C1, int N, int Hs, int Ws, int src_nchw, int mode,
                  const float* mean_rstd, int groups, const float* gamma, const float* beta, int act, float scale,
                  void* out_hi, void* out_lo, int Cp, float* out_f32, void* stream) {
    if (!src0 || C0 <= 0 || C1 < 0 || (C1 > 0 && !src1) || N < 0 || Hs <= 0 || Ws <= 0)
        return api_fail("idb_conv_prep: bad argument");
    if (mode < 0 || mode > 3) return api_fail("idb_conv_prep: mode must be 0..3");
    if (src_nchw && C1) return api_fail("idb_conv_prep: NCHW source cannot be concatenated");
    if ((out_hi == nullptr) != (out_lo == nullptr)) return api_fail("idb_conv_prep: out_hi/out_lo both or neither");
    if (!out_hi && !out_f32) return api_fail("idb_conv_prep: no output requested");
    const int C = C0 + C1;
    if (out_hi && (Cp % 64 || Cp < C)) return api_fail("idb_conv_prep: Cp must be a multiple of 64 and >= C");
    if (mean_rstd && (groups <= 0 || C % groups)) return api_fail("idb_conv_prep: bad groups");
    if (mode == 3 && ((Hs | Ws) & 1)) return api_fail("idb_conv_prep: average pool needs even H, W");
    if (act < 0 || act > 2) return api_fail("idb_conv_prep: act must be 0, 1 or 2");
    if (out_hi && !(scale > 0.f)) return api_fail("idb_conv_prep: scale must be positive");
    idb::PrepParams p;
    p.src0 = src0; p.src1 = src1; p.C0 = C0; p.C1 = C1; p.N = N; p.Hs = Hs; p.Ws = Ws; p.src_nchw = src_nchw;
    p.mode = mode;
    p.fir_kh = p.fir_kw = 0; p.fir_up = p.fir_down = 1; p.fir_pad0 = p.fir_planes = p.fir_hf = p.fir_wf = 0;
    p.planes = mode == 2 ? 4 : 1;
    p.Ho = mode == 1 ? 2 * Hs : (mode == 0 ? Hs : (mode == 2 ? (Hs + 1) / 2 : Hs / 2));
    p.Wo = mode == 1 ? 2 * Ws : (mode == 0 ? Ws : (mode == 2 ? (Ws + 1) / 2 : Ws / 2));
    p.mean_rstd = mean_rstd; p.groups = mean_rstd ? groups : 1; p.cg = mean_rstd ? C / groups : C;
    p.gamma = gamma; p.beta = beta; p.act = act; p.scale = scale;
    p.out_hi = static_cast<uint8_t*>(out_hi); p.out_lo = static_cast<uint8_t*>(out_lo);
    p.Cp = out_hi ? Cp : ((C + 7) & ~7);
    p.out_f32 = out_f32; p.out_raw = nullptr;
    return api_check(idb::launch_conv_prep(p, idb::api_num_sms(), static_cast<cudaStream_t>(stream)), "idb_conv_prep");
}

int idb_conv_prep_fir(const float* src0, int C0, int N, int Hs, int Ws, const float* mean_rstd, int groups,
                      const float* gamma, const float* beta, int act, const float* kernel_host, int kh, int kw, int up,
                      int down, int pad0, int pad1, int as_planes, float scale, void* out_hi, void* out_lo, int Cp,
                      float* out_f32, float* out_raw, void* stream) {
    if (!src0 || !kernel_host || C0 <= 0 || N < 0 || Hs <= 0 || Ws <= 0) return api_fail("idb_conv_prep_fir: bad argument");
    if (kh < 1 || kh > 4 || kw < 1 || kw > 4) return api_fail("idb_conv_prep_fir: FIR kernels up to 4 x 4");
    if ((up != 1 && up != 2) || (down != 1 && down != 2) || (as_planes && down != 1) || pad0 < 0 || pad1 < 0)
        return api_fail("idb_conv_prep_fir: up, down in {1, 2}, pads >= 0, as_planes needs down = 1");
    if ((out_hi == nullptr) != (out_lo == nullptr)) return api_fail("idb_conv_prep_fir: out_hi/out_lo both or neither");
    if (!out_hi && !out_f32 && !out_raw) return api_fail("idb_conv_prep_fir: no output requested");
    if (out_hi && (Cp % 64 || Cp < C0)) return api_fail("idb_conv_prep_fir: Cp must be a multiple of 64 and >= C");
    if (mean_rstd && (groups <= 0 || C0 % groups)) return api_fail("idb_conv_prep_fir: bad groups");
    if (act < 0 || act > 2) return api_fail("idb_conv_prep_fir: act must be 0, 1 or 2");
    if (out_hi && !(scale > 0.f)) return api_fail("idb_conv_prep_fir: scale must be positive");
    const int hf = (Hs * up + pad0 + pad1 - kh) / down + 1, wf = (Ws * up + pad0 + pad1 - kw) / down + 1;
    if (hf <= 0 || wf <= 0) return api_fail("idb_conv_prep_fir: empty output");
    idb::PrepParams p;
    p.src0 = src0; p.src1 = nullptr; p.C0 = C0; p.C1 = 0; p.N = N; p.Hs = Hs; p.Ws = Ws; p.src_nchw = 0;
    p.mode = 4;
    for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b)   // the reference correlates with the FLIPPED kernel (upfirdn2d_kernel.cu:137)
            p.fir[a * 4 + b] = (a < kh && b < kw) ? kernel_host[(kh - 1 - a) * kw + (kw - 1 - b)] : 0.f;
    p.fir_kh = kh; p.fir_kw = kw; p.fir_up = up; p.fir_down = down; p.fir_pad0 = pad0; p.fir_planes = as_planes ? 1 : 0;
    p.fir_hf = hf; p.fir_wf = wf;
    p.planes = as_planes ? 4 : 1;
    p.Ho = as_planes ? (hf + 1) / 2 : hf;
    p.Wo = as_planes ? (wf + 1) / 2 : wf;
    p.mean_rstd = mean_rstd; p.groups = mean_rstd ? groups : 1; p.cg = mean_rstd ? C0 / groups : C0;
    p.gamma = gamma; p.beta = beta; p.act = act; p.scale = scale;
    p.out_hi = static_cast<uint8_t*>(out_hi); p.out_lo = static_cast<uint8_t*>(out_lo);
    p.Cp = out_hi ? Cp : ((C0 + 7) & ~7);
    p.out_f32 = out_f32; p.out_raw = out_raw;
    return api_check(idb::launch_conv_prep(p, idb::api_num_sms(), static_cast<cudaStream_t>(stream)), "idb_conv_prep_fir");
}

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}
}  // namespace

int idb_conv2d(const void* a_hi, const void* a_lo, int planes, int N, int Ha, int Wa, int H, int W, int Cp, const void* w_hi,
               const void* w_lo, int n_taps, const int* taps_host, int Np, int Cout, float acc_scale, float out_scale,
               const float* bias, const float* residual, int ld_res, float* out, int ldo, void* stream) {
    return idb_conv2d_stats(a_hi, a_lo, planes, N, Ha, Wa, H, W, Cp, w_hi, w_lo, n_taps, taps_host, Np, Cout, acc_scale,
                            out_scale, bias, residual, ld_res, out, ldo, nullptr, stream);
}

int idb_conv_stats_slots(int H, int W) {
    if (H <= 0 || W <= 0) return 0;
    const int BW = W < 128 ? W : 128;
    const int BH = 128 / BW < H ? 128 / BW : H;
    const int pimg = BW * BH;
    if (pimg < 16 || (H * W) % (pimg < 32 ? pimg : 32)) return 0;
    return (H * W) / (pimg < 32 ? pimg : 32);
}

int idb_groupnorm_from_conv_stats(const float* stats0, int C0, int slots0, const float* stats1, int C1, int slots1, int N,
                                  int HW, int groups, double eps, float* mean_rstd, void* stream) {
    if (!stats0 || !mean_rstd || N < 0 || HW <= 0 || groups <= 0 || C0 <= 0 || slots0 <= 0 || C1 < 0 ||
        (C1 > 0 && (!stats1 || slots1 <= 0)) || (C0 + C1) % groups)
        return api_fail("idb_groupnorm_from_conv_stats: bad argument");
    return api_check(idb::launch_gn_from_partials(stats0, C0, slots0, C1 > 0 ? stats1 : nullptr, C1, slots1, N, HW,
                                                  groups, eps, mean_rstd, static_cast<cudaStream_t>(stream)),
                     "idb_groupnorm_from_conv_stats");
}

static int conv2d_impl(const void* a_hi, const void* a_lo, int planes, int N, int Ha, int Wa, int H, int W, int Cp,
                       const void* w_hi, const void* w_lo, int n_taps, const int* taps_host, int Np, int Cout,
                       float acc_scale, float out_scale, const float* bias, const float* residual, int ld_res, float* out,
                       int ldo, float* stats, long long w_img_stride, void* stream);

int idb_conv2d_stats(const void* a_hi, const void* a_lo, int planes, int N, int Ha, int Wa, int H, int W, int Cp,
                     const void* w_hi, const void* w_lo, int n_taps, const int* taps_host, int Np, int Cout,
                     float acc_scale, float out_scale, const float* bias, const float* residual, int ld_res, float* out,
                     int ldo, float* stats, void* stream) {
    return conv2d_impl(a_hi, a_lo, planes, N, Ha, Wa, H, W, Cp, w_hi, w_lo, n_taps, taps_host, Np, Cout, acc_scale, out_scale,
                       bias, residual, ld_res, out, ldo, stats, 0, stream);
}

size_t idb_attention_tc_workspace_bytes(int N, int T, int dh) {
    if (N <= 0 || T <= 0 || dh <= 0) return 256;
    const size_t op = (size_t)N * T * dh * 2;          // one fp16 plane of Q / K / V^T
    return 6 * op + (size_t)N * T * T * 4 + 2 * (size_t)N * T * T * 2 + 1024;
}

int idb_attention_tc(const float* qkv, int N, int H, int W, int ld, int dh, int q0, int k0, int v0, float scale,
                     void* workspace, float* out, int ld_out, void* stream) {
    const int T = H * W;
    if (!qkv || !workspace || !out || N <= 0 || H <= 0 || W <= 0 || ld_out < dh)
        return api_fail("idb_attention_tc: bad argument");
    if (T < 128 || T % 128 || dh % 64 || dh > 256 * 4 || T > 1024)
        return api_fail("idb_attention_tc: needs T = H*W a multiple of 128 (one image per 128-pixel tile) and dh % 64 == 0; "
                        "use idb_attention otherwise");
    if ((reinterpret_cast<uintptr_t>(workspace) & 1023) || (reinterpret_cast<uintptr_t>(qkv) & 15))
        return api_fail("idb_attention_tc: workspace must be 1024-byte aligned, qkv 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t op = (size_t)N * T * dh * 2;
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    uint8_t *q_hi = ws, *q_lo = ws + op, *k_hi = ws + 2 * op, *k_lo = ws + 3 * op, *vt_hi = ws + 4 * op, *vt_lo = ws + 5 * op;
    float* S = reinterpret_cast<float*>(ws + 6 * op);
    uint8_t* p_hi = reinterpret_cast<uint8_t*>(S) + (size_t)N * T * T * 4;
    uint8_t* p_lo = p_hi + (size_t)N * T * T * 2;
    const float OPS = 64.f, PS = 4096.f;     // operand scales: activations x 2^6 (as every convolution), probabilities x 2^12
    int rc = api_check(idb::launch_attn_pack(qkv, N, T, ld, dh, q0, k0, v0, OPS, q_hi, q_lo, k_hi, k_lo, vt_hi, vt_lo, st),
                       "idb_attention_tc(pack)");
    if (rc) return rc;
    const int tap[3] = {0, 0, 0};
    // S[n][query][key] = scale * sum_c Q[n][query][c] K[n][key][c]: "1x1 convolution" with the image's own K as weights
    rc = conv2d_impl(q_hi, q_lo, 1, N, H, W, H, W, dh, k_hi, k_lo, 1, tap, T, T, scale / (OPS * OPS), 1.f, nullptr, nullptr, 0,
                     S, T, nullptr, (long long)T * dh * 2, st);
    if (rc) return rc;
    rc = api_check(idb::launch_attn_softmax(S, N * T, T, PS, p_hi, p_lo, st), "idb_attention_tc(softmax)");
    if (rc) return rc;
    // O[n][query][c] = sum_key P[n][query][key] V[n][key][c]: weights = the image's own V^T
    const int Np = (dh + 127) / 128 * 128;
    if (Np != dh) return api_fail("idb_attention_tc: dh must be a multiple of 128");
    return conv2d_impl(p_hi, p_lo, 1, N, H, W, H, W, T, vt_hi, vt_lo, 1, tap, Np, dh, 1.f / (PS * OPS), 1.f, nullptr, nullptr, 0,
                       out, ld_out, nullptr, (long long)T * dh * 2, st);
}

static int conv2d_impl(const void* a_hi, const void* a_lo, int planes, int N, int Ha, int Wa, int H, int W, int Cp,
                       const void* w_hi, const void* w_lo, int n_taps, const int* taps_host, int Np, int Cout,
                       float acc_scale, float out_scale, const float* bias, const float* residual, int ld_res, float* out,
                       int ldo, float* stats, long long w_img_stride, void* stream) {
    if (!a_hi || !a_lo || !w_hi || !w_lo || !out || !taps_host) return api_fail("idb_conv2d: null pointer");
    if (planes <= 0 || N <= 0 || H <= 0 || W <= 0 || Cp <= 0 || Cp % 64) return api_fail("idb_conv2d: bad shape (Cp % 64)");
    if (Ha < H || Wa < W) return api_fail("idb_conv2d: operand planes must be at least as large as the output (Ha >= H, Wa >= W)");
    if (n_taps <= 0 || n_taps > 16) return api_fail("idb_conv2d: 1..16 taps");
    if (Np % 128 || Cout <= 0 || Cout > Np || ldo < Cout) return api_fail("idb_conv2d: need Np % 128 == 0, Cout <= Np, ldo >= Cout");
    if (residual && ld_res < Cout) return api_fail("idb_conv2d: ld_res < Cout");
    idb::ConvParams p;
    p.BW = W < 128 ? W : 128;
    if (128 % p.BW || W % p.BW) return api_fail("idb_conv2d: W must be a power of two <= 128 or a multiple of 128");
    p.BH = 128 / p.BW < H ? 128 / p.BW : H;
    if ((128 / p.BW) % p.BH || H % p.BH) return api_fail("idb_conv2d: H must be a power of two or a multiple of 128/W");
    p.BI = 128 / (p.BW * p.BH);
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return api_fail("idb_conv2d: cuTensorMapEncodeTiled not available from the driver");
    const cuuint64_t dims[5] = {(cuuint64_t)Cp, (cuuint64_t)Wa, (cuuint64_t)Ha, (cuuint64_t)N, (cuuint64_t)planes};
    const cuuint64_t strides[4] = {(cuuint64_t)Cp * 2, (cuuint64_t)Wa * Cp * 2, (cuuint64_t)Ha * Wa * Cp * 2,
                                   (cuuint64_t)N * Ha * Wa * Cp * 2};
    const cuuint32_t box[5] = {64, (cuuint32_t)p.BW, (cuuint32_t)p.BH, (cuuint32_t)p.BI, 1};
    const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    for (int which = 0; which < 2; ++which) {
        CUresult r = enc(which ? &p.tm_lo : &p.tm_hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5,
                         const_cast<void*>(which ? a_lo : a_hi), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return api_fail("idb_conv2d: cuTensorMapEncodeTiled rejected the activation tensor");
    }
    p.w_hi = static_cast<const uint8_t*>(w_hi); p.w_lo = static_cast<const uint8_t*>(w_lo);
    p.N = N; p.H = H; p.W = W; p.Cp = Cp; p.n_taps = n_taps;
    for (int t = 0; t < n_taps; ++t) {
        p.tap_plane[t] = taps_host[3 * t]; p.tap_dy[t] = taps_host[3 * t + 1]; p.tap_dx[t] = taps_host[3 * t + 2];
        if (p.tap_plane[t] < 0 || p.tap_plane[t] >= planes) return api_fail("idb_conv2d: tap plane out of range");
    }
    p.Np = Np; p.Cout = Cout; p.acc_scale = acc_scale; p.out_scale = out_scale;
    p.bias = bias; p.residual = residual; p.ld_res = ld_res; p.out = out; p.ldo = ldo;
    p.stats = stats;
    p.w_img_stride = w_img_stride;
    if (w_img_stride && p.BI != 1) return api_fail("idb_conv2d: per-image weights need one image per 128-pixel tile");
    if (stats && (idb_conv_stats_slots(H, W) == 0 || (reinterpret_cast<uintptr_t>(stats) & 15)))
        return api_fail("idb_conv2d_stats: statistics need >= 16 pixels of an image per tile and a 16-byte aligned buffer");
    return api_check(idb::launch_conv(p, idb::api_num_sms(), static_cast<cudaStream_t>(stream)), "idb_conv2d");
}

int idb_attention(const float* qkv, int N, int T, int ld, int heads, int dh, int q0, int k0, int v0, int head_stride,
                  float scale, float* out, int ld_out, void* stream) {
    if (!qkv || !out || N < 0 || T <= 0 || heads <= 0 || dh <= 0 || ld_out < heads * dh)
        return api_fail("idb_attention: bad argument");
    if (T > 1536) return api_fail("idb_attention: at most 1536 tokens (scores are kept in shared memory)");
    idb::AttnParams p;
    p.qkv = qkv; p.N = N; p.T = T; p.ld = ld; p.heads = heads; p.dh = dh; p.q0 = q0; p.k0 = k0; p.v0 = v0;
    p.head_stride = head_stride; p.scale = scale; p.out = out; p.ld_out = ld_out;
    return api_check(idb::launch_attention(p, static_cast<cudaStream_t>(stream)), "idb_attention");
}

int idb_row_gram(const float* X, int ld, int K, int D, double scale, double* colmean, double* G, void* stream) {
    if (!X || !colmean || !G || K <= 0 || D <= 0 || ld < D) return api_fail("idb_row_gram: bad argument");
    return api_check(idb::launch_row_gram(X, ld, K, D, scale, colmean, G, static_cast<cudaStream_t>(stream)),
                     "idb_row_gram");
}

size_t idb_spectrum_large_workspace_bytes(int n) { return idb::spectrum_large_workspace_bytes(n > 0 ? n : 1); }

int idb_spectrum_large_variant(double* G, int n, int n_out, float* sv, void* workspace, int variant, void* stream) {
    if (!G || !sv || !workspace || n <= 0 || n_out <= 0 || n_out > n) return api_fail("idb_spectrum_large: bad argument");
    if (n > 14000) return api_fail("idb_spectrum_large: n <= 14000 (d and e^2 of the bisection in shared memory)");
    if (variant < -1 || variant > 3) return api_fail("idb_spectrum_large_variant: variant in {-1, 0, 1, 2, 3}");
    return api_check(idb::launch_spectrum_large(G, n, n_out, sv, workspace, idb::api_num_sms(),
                                                static_cast<cudaStream_t>(stream), variant),
                     "idb_spectrum_large");
}

int idb_spectrum_large(double* G, int n, int n_out, float* sv, void* workspace, void* stream) {
    return idb_spectrum_large_variant(G, n, n_out, sv, workspace, 0, stream);
}

}  // extern "C"
