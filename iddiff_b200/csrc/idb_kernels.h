// Internal launcher declarations shared by the .cu files and the C-ABI (idb_api.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace idb {

struct GemmParams {
    const uint8_t* a_hi;  // slab [Kc/64][M][128 B]
    const uint8_t* a_lo;
    const uint8_t* b_hi;  // slab [Kc/64][Np][128 B]
    const uint8_t* b_lo;
    int M, Np, Kc;
    int fmt;          // operand format of A, B and the mode-0 output: 0 = bf16 hi/lo pairs, 1 = fp16 hi/lo pairs
    float acc_scale;  // accumulator * acc_scale = un-scaled product (1 / (a_scale * w_scale), exact power of two)
    float out_scale;  // mode 0: deviations are multiplied by this before the hi/lo split (a_scale of the next layer)
    // mode 0 (hidden layer): anchor pre-activations in, slab-layout deviations out
    const float* a0;  // [n_anchors][Np]
    int rows_per_anchor;  // rows of A per anchor, anchor-major and DENSE (>= 128; a 128-row tile spans at most two anchors)
    int n_anchors;        // rows past n_anchors * rows_per_anchor are padding (zero deviations)
    uint8_t* out_hi;  // slab [Np/64][M][128 B]
    uint8_t* out_lo;
    // mode 1 (last layer): fp32 row-major [M][Np]
    float* out_f32;
};

cudaError_t launch_gemm_bf16x3(int mode, const GemmParams& p, int num_sms, cudaStream_t stream);

// fp32 [N][K] (row stride ldw) -> hi/lo bf16 slabs [Kc/64][Np][128 B], zero padded
cudaError_t launch_pack_weight(const float* w, int N, int K, int ldw, int Np, int Kc, int fp16, float scale,
                               uint8_t* hi, uint8_t* lo, cudaStream_t stream);

struct PerturbParams {
    const float* x0;  // [n_anchors][D]
    const float* z;   // [n_anchors][K][D] or null (Philox)
    int n_anchors, K, D;
    int Kp, Dp;       // padded rows per anchor / padded columns (slab outputs)
    float mean_coeff, std;
    uint64_t seed;
    uint64_t anchor_id0;  // global id of anchor 0 of this call (Philox stream key)
    uint32_t row0;        // global index of perturbation row 0 of this call (K sharded over ranks)
    int fmt;          // operand format of d_hi/d_lo (0 = bf16 pair, 1 = fp16 pair)
    float d_scale;    // deviations are multiplied by this (power of two) before the split
    float* x_out;     // optional fp32 [n_anchors][K][D] (perturbed points, reference arithmetic)
    uint8_t* d_hi;    // optional slab [Dp/64][n_anchors*Kp][128 B]: deviation x - mean
    uint8_t* d_lo;
};
cudaError_t launch_perturb(const PerturbParams& p, cudaStream_t stream);

// out[a][n] = sum_k w[n][k] * act(in[a][k]) + bias[n] (+ extra_col[n] * extra_scalar);  act = identity or ELU
cudaError_t launch_anchor_linear(const float* in, int ld_in, int apply_elu, float in_scale, const float* w, int ldw,
                                 const float* bias, const float* extra_col, int extra_stride, float extra_scalar,
                                 float* out, int ld_out, int n_anchors, int N, int K, cudaStream_t stream);

// score[a][k][c] = -(a_out[a][c] + d_out[a*Kp + k][c]) / std   (models/utils.py:264-267)
cudaError_t launch_fcn_scores(const float* a_out, int ld_a, const float* d_out, int ld_d, int n_anchors, int K, int Kp,
                              int D, float std, float* score, cudaStream_t stream);

cudaError_t launch_ksphere_score(const float* x, const float* sigma_or_null, float sigma, int rows, int D,
                                 int manifold_dim, float* out, cudaStream_t stream);

// column means (fp64) and centred Gram G = (scale (X - mean))^T (scale (X - mean)) in fp64
cudaError_t launch_column_sums(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D,
                               double divisor, double* out, cudaStream_t stream);
cudaError_t launch_gram_about(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D,
                              double scale, const double* colmean, double* G, cudaStream_t stream);
cudaError_t launch_center_gram(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D,
                               double scale, double* colmean, double* G, cudaStream_t stream);

size_t spectrum_workspace_bytes(int n_anchors, int n);
cudaError_t launch_spectrum(const double* G, int n_anchors, int n, int n_out, float* sv, int* sweeps, void* workspace,
                            cudaStream_t stream);

size_t tridiag_workspace_bytes(int n_anchors, int n);
cudaError_t launch_spectrum_tridiag(const double* G, int n_anchors, int n, int n_out, float* sv, int* iters,
                                    void* workspace, cudaStream_t stream);

cudaError_t launch_id_readout(const float* sv, int n_anchors, int n, int ld, int* dims, cudaStream_t stream);

struct UpFirDnParams {
    const float* in;      // [major][in_h][in_w][minor]
    const float* kernel;  // [kh][kw]
    float* out;           // [major][out_h][out_w][minor]
    int major, in_h, in_w, minor, kh, kw;
    int up_x, up_y, down_x, down_y, pad_x0, pad_x1, pad_y0, pad_y1;
    int out_h, out_w, minor_tile;  // filled by the launcher
};
cudaError_t launch_upfirdn2d(UpFirDnParams p, int num_sms, cudaStream_t stream);
cudaError_t launch_fused_bias_act(const float* x, const float* b, const float* ref, int act, int grad, float alpha,
                                  float scale, long long size_x, int step_b, int size_b, float* out, int num_sms,
                                  cudaStream_t stream);

struct alignas(64) ConvParams {
    CUtensorMap tm_hi, tm_lo;  // A operand: 5-D (Cp, W, H, N, planes) fp16, box (64, BW, BH, BI, 1), SWIZZLE_128B
    const uint8_t* w_hi;       // slab [n_taps * Cp/64][Np][128 B]
    const uint8_t* w_lo;
    int N, H, W, Cp, n_taps;
    int tap_plane[16], tap_dy[16], tap_dx[16];
    int Np, Cout;
    int BW, BH, BI;            // BW * BH * BI == 128 output pixels per tile
    float acc_scale, out_scale;
    const float* bias;         // [Cout] or null
    const float* residual;     // [N][H][W][ld_res] or null
    int ld_res;
    float* out;                // [N][H][W][ldo]
    int ldo;
    float* stats;              // optional [N][H*W / min(32, BW*BH)][Cout][2]: per-slot (sum, sum of squares) of `out`
    long long w_img_stride;    // bytes between the weight slabs of consecutive images (0: one weight matrix for all) --
                               // attention's Q K^T and P V are convolutions whose "weights" are per-image (BI == 1 only)
};
cudaError_t launch_conv(const ConvParams& p, int num_sms, cudaStream_t stream);

struct PrepParams {
    const float* src0;  // NHWC [N][Hs][Ws][C0]  (NCHW [N][C0][Hs][Ws] when src_nchw)
    const float* src1;  // optional second tensor concatenated along C: [N][Hs][Ws][C1]
    int C0, C1, N, Hs, Ws, src_nchw;
    int mode;           // 0 same size, 1 nearest x2 up, 2 stride-2 parity planes, 3 2x2 average pool, 4 FIR resampling
    // mode 4: upfirdn2d(transformed src, fir, up, down, pad) -- op/upfirdn2d_kernel.cu:107-207 folded into this pass
    float fir[16];      // FLIPPED kernel, row-major [kh][4]
    int fir_kh, fir_kw, fir_up, fir_down, fir_pad0, fir_planes;  // fir_planes: emit the 4 stride-2 parity planes of the
    int fir_hf, fir_wf;                                          // (down = 1) result, whose size is fir_hf x fir_wf
    int planes, Ho, Wo;
    int wo_shift, ho_shift;  // log2 when Wo / Ho are powers of two, else -1 (filled by the launcher)
    const float* mean_rstd;  // [N][groups][2] or null
    int groups, cg;
    const float* gamma;      // [C] or null
    const float* beta;
    int act;                 // 0 none, 1 SiLU
    float scale;             // power of two applied before the fp16 hi/lo split
    uint8_t* out_hi;         // [planes][N][Ho][Wo][Cp] fp16 (may be null)
    uint8_t* out_lo;
    int Cp;
    float* out_f32;          // optional fp32 [planes][N][Ho][Wo][C] copy of the transformed tensor
    float* out_raw;          // mode 4 only, optional: the same FIR applied to the UNtransformed source, fp32 [planes][N][Ho][Wo][C]
};
cudaError_t launch_conv_prep(const PrepParams& p, int num_sms, cudaStream_t stream);

cudaError_t launch_gn_from_partials(const float* st0, int C0, int S0, const float* st1, int C1, int S1, int N, int HW,
                                    int groups, double eps, float* mean_rstd, cudaStream_t stream);
cudaError_t launch_gn_stats(const float* s0, int C0, const float* s1, int C1, int N, int HW, int groups, double eps,
                            double* sums, float* mean_rstd, int num_sms, cudaStream_t stream);

struct AttnParams {
    const float* qkv;  // [N][T][ld]
    int N, T, ld, heads, dh;
    int q0, k0, v0, head_stride;  // channel offset of q / k / v of head h = x0 + h * head_stride
    float scale;
    float* out;        // [N][T][ld_out], channel h * dh + c
    int ld_out;
};
cudaError_t launch_attention(const AttnParams& p, cudaStream_t stream);

cudaError_t launch_row_gram(const float* X, int ld, int K, int D, double scale, double* colmean, double* G,
                            cudaStream_t stream);
size_t spectrum_large_workspace_bytes(int n);
size_t spectrum_large_fused_workspace_bytes(int n);
cudaError_t launch_spectrum_large(double* G, int n, int n_out, float* sv, void* workspace, int num_sms,
                                  cudaStream_t stream, int classes = 0);

// helpers of the C-ABI translation units (thread-local error string lives in idb_api.cu)
int api_fail(const char* what);
int api_check(cudaError_t e, const char* where);
int api_num_sms();

// ---- DSM training step of the FCN (idb_train.cu) ----
struct TrainPackParams {
    const float* src;   // fp32 [R][ld]
    int ld, R, C;       // valid extent; everything outside packs as zero
    const float* pre;   // mode 2: pre-activations [R][ld_pre]
    int ld_pre;
    int mode;           // 0 identity, 1 ELU(src), 2 src * ELU'(pre)
    int transpose;      // 0: slab [Rp][Cp] (+ out_f32 [Rp][ld_out]); 1: slab [Cp][Rp]
    int fmt;
    float scale;
    int Rp, Cp;
    uint8_t* hi;
    uint8_t* lo;
    float* out_f32;
    int ld_out;
};
cudaError_t launch_dsm_sample(const float* data, const long long* idx, const float* t, const float* mean_coeff,
                              const float* stdv, const float* z_ext, int B, int D, int Mp, int ldx, float label_scale,
                              uint64_t seed, uint64_t step, const long long* step_dev, float* X, float* z_out,
                              cudaStream_t stream);
cudaError_t launch_train_pack(const TrainPackParams& p, cudaStream_t stream);
cudaError_t launch_dsm_loss(const float* out, int ldo, const float* z, int B, int D, int Mp, float* dZ, float* loss,
                            double* partial, cudaStream_t stream);
cudaError_t launch_colsum_f32(const float* src, int ld, int R, int C, float* out, cudaStream_t stream);
cudaError_t launch_grad_sumsq(const float* g, int ld, int N, int K, double* accum, cudaStream_t stream);
cudaError_t launch_adam(float* prm, int ldp, const float* g, int ldg, float* m, float* v, int N, int K, float lr_base,
                        int warmup, float beta1, float beta2, float eps, long long step_host, const long long* step_dev,
                        float max_norm, const double* normsq, cudaStream_t stream);

// ---- attention on the tensor cores: operand packing + softmax around two per-image-weight convolutions (idb_unet.cu) ----
cudaError_t launch_attn_pack(const float* qkv, int N, int T, int ld, int dh, int q0, int k0, int v0, float scale,
                             uint8_t* q_hi, uint8_t* q_lo, uint8_t* k_hi, uint8_t* k_lo, uint8_t* vt_hi, uint8_t* vt_lo,
                             cudaStream_t stream);
cudaError_t launch_attn_softmax(const float* S, int rows, int T, float scale, uint8_t* p_hi, uint8_t* p_lo,
                                cudaStream_t stream);

}  // namespace idb
