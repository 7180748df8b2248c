/* iddiff_b200 -- C ABI of the B200-native `manifold_dimension` hot path of GBATZOLIS/ID-diff.
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer into caller-owned memory (torch tensors on the Python side);
 *     nothing is allocated, freed or synchronised inside; the work is enqueued on `stream`
 *     (a cudaStream_t passed as void*), mirroring the reference ops which launch on the current
 *     ATen stream without syncing (op/upfirdn2d_kernel.cu:213-215 in the reference);
 *   - return value 0 = enqueued, negative = rejected (bad argument or CUDA launch error);
 *     idb_last_error() returns a thread-local description of the last failure;
 *   - no global mutable state: callable from any host thread with its own stream.
 *
 * "slab layout": bf16 matrix [R rows][C cols], C % 64 == 0, R % 128 == 0, stored as C/64 k-slabs of
 * R lines x 128 bytes, the eight 16-byte chunks of a line XOR-permuted by (row % 8): the
 * SWIZZLE_128B K-major shared-memory image of tcgen05.mma, so an operand tile is one contiguous
 * bulk-TMA copy.  byte offset of (r, c) = ((c/64)*R + r)*128 + ((((c%64)/8) ^ (r%8)) * 16) + (c%8)*2.
 *
 * Reference file:line citations are relative to the root of GBATZOLIS/ID-diff.
 */
#ifndef IDDIFF_B200_H
#define IDDIFF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

int idb_abi_version(void);
const char* idb_last_error(void);

/* Perturbation sampling: dim_reduction.py:178-182 (x.clone(); sde.marginal_prob; randn_like; mean + std*z)
 * with sde_lib.py:342-347 / 250-254 / 300-304 / 181-186 folded into (mean_coeff, std) by the host
 * (t is one constant for the whole run, dim_reduction.py:173-174).
 *   x0 [n_anchors][D] fp32;  z [n_anchors][K][D] fp32 or NULL -> fused Philox4x32-10 + Box-Muller keyed by
 *   (seed, anchor_id0 + a, row, column), so results do not depend on sharding or chunking.
 * Outputs (either may be NULL): x_out [n_anchors][K][D] fp32 = fl(fl(mean_coeff*x0) + fl(std*z));
 *   d_hi/d_lo slab layout [Dp/64][M][128B], M = ceil128(n_anchors*Kp) = hi/lo pair (format fmt) of d_scale*(x - mean):
 *   anchor a owns rows [a*Kp, a*Kp + K) (Kp >= K; Kp = K packs the anchors densely), all other rows and columns >= D zero.
 *   Dp % 64 == 0. */
int idb_perturb(const float* x0, const float* z, int n_anchors, int K, int D, int Kp, int Dp, float mean_coeff,
                float std, uint64_t seed, uint64_t anchor_id0, float* x_out, int fmt, float d_scale, void* d_hi,
                void* d_lo, void* stream);

/* Same, for a SLICE of the K perturbation rows of each anchor: the K rows computed by this call are rows
 * [row0, row0 + K) of the anchor's global perturbation set (Philox counter = global row), so one huge anchor can be
 * sharded by rows over the GPUs of a box ("one huge anchor") with results independent of the world size. */
int idb_perturb_rows(const float* x0, const float* z, int n_anchors, int K, int D, int Kp, int Dp, float mean_coeff,
                     float std, uint64_t seed, uint64_t anchor_id0, uint32_t row0, float* x_out, int fmt, float d_scale,
                     void* d_hi, void* d_lo, void* stream);

/* Operand formats of the tensor-core chain.  Every operand is a hi/lo PAIR of 16-bit numbers (v ~= hi + lo) so
 * that three tcgen05.mma (hi.hi + lo.hi + hi.lo) reproduce an fp32-accurate product:
 *   IDB_FMT_BF16  bf16 pairs, 16 significand bits, fp32 exponent range, no scaling needed (scales must be 1);
 *   IDB_FMT_FP16  fp16 pairs, 22 significand bits; values are pre-multiplied by a power-of-two scale so that hi
 *                 and lo are normal fp16 numbers, and the fp32 accumulator is multiplied by the exact inverse.
 * A and B of one layer must use the SAME format (kind::f16 faults on bf16 x fp16). */
#define IDB_FMT_BF16 0
#define IDB_FMT_FP16 1

/* nn.Linear weight [N][K] fp32 (row stride ldw) -> hi/lo slabs [Kc/64][Np][128B] of (w * scale), zero padded.
 * Np % 128 == 0 (% 256 when used as a W operand), Kc % 64 == 0; also packs anchor rows [A][D] into an A operand.
 * Done once per model (models/fcn.py:16-30 holds the weights).  For IDB_FMT_FP16
 * choose scale = 2^k with max|w|*scale in [2^13, 2^14). */
int idb_pack_weight(const float* w, int N, int K, int ldw, int Np, int Kc, int fmt, float scale, void* hi, void* lo,
                    void* stream);

/* One nn.Linear of models/fcn.py:20-30 evaluated in fp32 at the un-perturbed anchors:
 *   out[a][n] = sum_k w[n][k] * f(in_scale*in[a][k]) + bias[n] + extra_col[n*extra_stride]*extra_scalar,
 * f = ELU when apply_elu else identity; extra_col/extra_scalar carry the time column of
 * torch.cat([x, t]) (models/fcn.py:35-38).  bias / extra_col may be NULL. */
int idb_anchor_linear(const float* in, int ld_in, int apply_elu, float in_scale, const float* w, int ldw,
                      const float* bias, const float* extra_col, int extra_stride, float extra_scalar, float* out,
                      int ld_out, int n_anchors, int N, int K, void* stream);

/* Hidden nn.Linear + ELU on the deviations (tcgen05 kind::f16, 3 MMAs per product, fp32 accumulate in TMEM):
 *   acc = acc_scale * A (hi+lo) . W^T (hi+lo);  e = ELU(a0 + acc) - ELU(a0);  out = hi/lo pair of out_scale * e in
 *   slabs [Np/64][M][128B].  acc_scale = 1 / (scale of A * scale of W); out_scale = scale the next layer expects.
 * a0 [n_anchors][Np] fp32 = anchor pre-activations of this layer.  Rows of A are anchor-major and DENSE: anchor a owns rows
 * [a * rows_per_anchor, (a + 1) * rows_per_anchor) (rows_per_anchor = K >= 128: a 128-row tile spans at most two anchors and
 * every row looks up its own), rows from n_anchors * rows_per_anchor to M are zero padding.  M % 128 == 0, Np % 256 == 0,
 * Kc % 64 == 0. */
int idb_fcn_hidden_layer(const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, int fmt,
                         float acc_scale, float out_scale, int M, int Np, int Kc, const float* a0,
                         int rows_per_anchor, int n_anchors, void* out_hi, void* out_lo, void* stream);

/* Last nn.Linear on the deviations: out_f32 [M][Np] fp32 row-major = acc_scale * A (hi+lo) . W^T (hi+lo). */
int idb_fcn_output_layer(const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, int fmt,
                         float acc_scale, int M, int Np, int Kc, float* out_f32, void* stream);

/* One nn.Linear of models/fcn.py:20-30 evaluated at the un-perturbed anchors on the tensor cores (same kernel,
 * third epilogue): pre[m][n] = acc_scale * A (hi+lo) . W^T (hi+lo) + bias[n] -> out_f32 [M][Np] fp32, and, when
 * out_hi/out_lo are set, hi/lo pairs of out_scale * ELU(pre) in slabs [Np/64][M][128B] = the next layer's anchor
 * operand.  A rows are anchors (M % 128 == 0, rows past the last anchor are padding); bias [Np] may be NULL.
 * The time column of torch.cat([x, t]) (models/fcn.py:35-38) is folded into the first layer's bias by the host.
 * idb_anchor_linear above is the plain-fp32 SIMT version of the same op, kept as its unit-test cross-check. */
int idb_fcn_anchor_layer(const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, int fmt,
                         float acc_scale, float out_scale, int M, int Np, int Kc, const float* bias, float* out_f32,
                         void* out_hi, void* out_lo, void* stream);

/* Raw scores as the reference's score_fn returns them (models/utils.py:264-267):
 *   score[a][k][c] = -(a_out[a][c] + d_out[a*Kp + k][c]) / std,  score [n_anchors][K][D] fp32. */
int idb_fcn_scores(const float* a_out, int ld_a, const float* d_out, int ld_d, int n_anchors, int K, int Kp, int D,
                   float std, float* score, void* stream);

/* Analytic k-sphere score, models/ksphere_gt.py:14-26 (+ models/utils.py:169-175: std passed, output
 * returned unchanged).  x/out [rows][D] fp32; sigma_rows [rows] or NULL -> scalar sigma. */
int idb_ksphere_gt_score(const float* x, const float* sigma_rows, float sigma, int rows, int D, int manifold_dim,
                         float* out, void* stream);

/* dim_reduction.py:190-194 + the Gram side of the SVD (:197): per anchor a, over rows k < K of
 * X + a*anchor_stride (row stride ld, fp32): colmean[a][D] (fp64) and
 * G[a][D][D] = sum_k (scale (X[k]-mean))^T (scale (X[k]-mean)) accumulated in fp64. */
int idb_center_gram(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D, double scale,
                    double* colmean, double* G, void* stream);

/* The two halves of idb_center_gram, for ONE anchor whose K rows are sharded over the GPUs of a box ("one
 * huge anchor": partial Grams + one all-reduce).  Each rank calls idb_column_sums on its rows (colsum[a][D] fp64, plain
 * sums), the sums are all-reduced and divided by the global K, then idb_gram_about accumulates
 * G[a][D][D] = sum_{k local} (scale (X[k]-center))^T (scale (X[k]-center)) about that GLOBAL mean, so the partial Grams
 * add up exactly to the centred Gram of dim_reduction.py:193-197 and one all-reduce of G finishes it. */
int idb_column_sums(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D, double* colsum,
                    void* stream);
int idb_gram_about(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D, double scale,
                   const double* center, double* G, void* stream);

/* Singular values of the centred score matrix from its Gram matrix (replaces torch.linalg.svd at
 * dim_reduction.py:197-198), one anchor per CTA, matrix resident in shared memory while it fits:
 *   IDB_SPECTRUM_TRIDIAG  fp64 Householder tridiagonalisation + Sturm bisection (one thread per eigenvalue);
 *   IDB_SPECTRUM_JACOBI   fp64 Cholesky G = L L^T, then one-sided Jacobi on L in fp32 with warp-shuffle rotations.
 * sv [n_anchors][n_out] fp32 descending (n_out <= n: the reference returns min(K, D) values);
 * sweeps [n_anchors] (may be NULL) = Jacobi sweeps / bisection steps used.  workspace: idb_spectrum_workspace_bytes. */
#define IDB_SPECTRUM_JACOBI 0
#define IDB_SPECTRUM_TRIDIAG 1
size_t idb_spectrum_workspace_bytes(int n_anchors, int n);
int idb_spectrum(const double* G, int n_anchors, int n, int n_out, int method, float* sv, int* sweeps, void* workspace,
                 void* stream);

/* Spectral-gap read-out, plot_utils.py:224-229 (dup :173-182) with softmax :211-214:
 * dims[a] = (n-2) - argmax softmax((s[i]-s[i+1])/(s[1]-s[2]), i=1..n-2); -1 where the reference raises
 * ZeroDivisionError (s[1]==s[2]).  sv [n_anchors][ld] fp32, n >= 3. */
int idb_id_from_spectrum(const float* sv, int n_anchors, int n, int ld, int* dims, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Denoising-score-matching training step of the FCN (losses.py:164-188, :26-51,
 * lightning_modules/BaseSdeGenerativeModel.py:60-63).  The matrix products of the forward AND backward pass are
 * idb_fcn_anchor_layer / idb_fcn_output_layer in the bf16-pair format; these entry points are everything around them.
 * Host side: iddiff_b200/train.py (FcnTrainer).
 * ------------------------------------------------------------------------------------------------ */

/* losses.py:176-179 + models/fcn.py:35-38: X[r] = (mean_coeff[r] * data[idx[r]] + std[r] * z[r], label_scale * t[r], 0 ...)
 * for r < B, zero rows up to Mp; X fp32 [Mp][ldx] (ldx >= D + 1, % 8), z_out fp32 [B][D].  z_ext [B][D] or NULL (Philox
 * keyed by (seed, step, row, column)); idx [B] int64 or NULL (rows 0 .. B-1 of data).  step_dev: optional device counter that
 * replaces `step` (a whole training step is then one replayable CUDA graph). */
int idb_dsm_sample(const float* data, const long long* idx, const float* t, const float* mean_coeff, const float* stdv,
                   const float* z_ext, int B, int D, int Mp, int ldx, float label_scale, uint64_t seed, uint64_t step,
                   const long long* step_dev, float* X, float* z_out, void* stream);

/* fp32 [R][ld] -> tensor-core operand slabs with the activation (or its derivative) fused in:
 *   value(r, c) = src (mode 0) | ELU(src) (mode 1) | src * ELU'(pre) (mode 2: dZ = dH * ELU'(pre)) inside [R) x [C), else 0;
 *   transpose 0: slab of [Rp][Cp] (A operand: rows = batch) and / or the fp32 values out_f32 [Rp][ld_out];
 *   transpose 1: slab of [Cp][Rp] (rows = features, contraction = batch: the operands of dW = dZ^T H). */
int idb_train_pack(const float* src, int ld, int R, int C, const float* pre, int ld_pre, int mode, int transpose, int fmt,
                   float scale, int Rp, int Cp, void* hi, void* lo, float* out_f32, int ld_out, void* stream);

/* losses.py:181-187 (VE, reduce_mean, no likelihood weighting): loss = mean (out - z)^2 over [B) x [D);
 * dZ [Mp][ldo] = 2 (out - z) / (B D) inside, 0 in the padding.  loss: one float on the device; scratch: 64 doubles (block
 * partial sums, added in a fixed order: the loss is bit-reproducible). */
int idb_dsm_loss(const float* out, int ldo, const float* z, int B, int D, int Mp, float* dZ, float* loss, double* scratch,
                 void* stream);

/* out[c] = sum_{r < R} src[r][c] (bias gradients). */
int idb_column_sums_f32(const float* src, int ld, int R, int C, float* out, void* stream);

/* accum (one double on the device, zeroed by the caller once per step) += sum of squares of grad [N][K] (row stride ld):
 * the global norm of torch.nn.utils.clip_grad_norm_ (losses.py:47-48). */
int idb_grad_sumsq(const float* g, int ld, int N, int K, double* accum, void* stream);

/* One torch.optim.Adam step (no amsgrad / weight decay; losses.py:26-35) on param [N][K] (row stride ldp) with gradient
 * [N][K] (row stride ldg), after clipping by the global norm sqrt(*grad_normsq) to max_norm (max_norm < 0 or NULL: no
 * clipping; losses.py:47-48) and with the linear warm-up lr * min(step / warmup, 1) (losses.py:44-46; warmup <= 0: none).
 * step >= 1 is the optimiser step; when step_dev is set the counter is read from the device instead.  Clip coefficient,
 * learning rate and bias corrections are all formed on the device: a step never synchronises with the host. */
int idb_adam_step(float* param, int ldp, const float* grad, int ldg, float* m, float* v, int N, int K, float lr,
                  int warmup, float beta1, float beta2, float eps, long long step, const long long* step_dev,
                  float max_norm, const double* grad_normsq, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Image score networks.
 * ------------------------------------------------------------------------------------------------ */

/* Replaces the reference extension op `upfirdn2d(input[major,in_h,in_w,minor], kernel[kh,kw], up_x, up_y, down_x,
 * down_y, pad_x0, pad_x1, pad_y0, pad_y1) -> [major,out_h,out_w,minor]` (op/upfirdn2d.cpp:12-22; kernels
 * op/upfirdn2d_kernel.cu:49-369; semantics = upfirdn2d_native, op/upfirdn2d.py:162-200): zero-stuff by up, pad (negative
 * pad crops), correlate with the flipped FIR kernel, keep every down-th sample.  fp32, contiguous.
 * out_h = (in_h*up_y + pad_y0 + pad_y1 - kh)/down_y + 1 (same for w); the caller allocates `out`. */
int idb_upfirdn2d(const float* input, const float* kernel, int major, int in_h, int in_w, int minor, int kernel_h,
                  int kernel_w, int up_x, int up_y, int down_x, int down_y, int pad_x0, int pad_x1, int pad_y0,
                  int pad_y1, float* out, void* stream);

/* Replaces the reference extension op `fused_bias_act(input, bias, refer, act, grad, alpha, scale)`
 * (op/fused_bias_act.cpp:11-20, op/fused_bias_act_kernel.cu:19-99): out[i] = f(x[i] + bias[(i/step_b) % size_b]) * scale,
 * f by (act, grad): act 1 linear, 3 leaky ReLU(alpha); grad 0 forward, 1 gradient (sign from refer), 2 -> 0.
 * bias / refer may be NULL.  The reference passes step_b = prod(shape[2:]), size_b = shape[1] (fused_bias_act_kernel.cu:80-84). */
int idb_fused_bias_act(const float* input, const float* bias, const float* refer, int act, int grad, float alpha,
                       float scale, long long size_x, int step_b, int size_b, float* out, void* stream);

/* U-Net activations are fp32 NHWC [N][H][W][C] in HBM.  A convolution is three calls: idb_groupnorm_stats (when the
 * conv input is normalised), idb_conv_prep (writes the tensor-core operand), idb_conv2d. */

/* nn.GroupNorm statistics (layers.py:638,648; ddpm.py:143; BeatGANs_nn.py:23-25 GroupNorm32): per (image, group)
 * mean and 1/sqrt(biased var + eps) of x = concat(src0 [N][HW][C0], src1 [N][HW][C1]) along C (src1 NULL / C1 = 0:
 * no concat -- the concat of ddpm.py:198 is never materialised).  sums [N][groups][2] fp64 is scratch,
 * mean_rstd [N][groups][2] fp32 the result. */
int idb_groupnorm_stats(const float* src0, int C0, const float* src1, int C1, int N, int HW, int groups, double eps,
                        double* sums, float* mean_rstd, void* stream);

/* Operand preparation = everything between two convolutions in ONE pass over the activation:
 *   v = concat(src0, src1)[n, y, x, c];  v = (v - mean) * rstd (if mean_rstd);  v = v * gamma[c] + beta[c] (if gamma);
 *   v = SiLU(v) (act 1);  then by mode: 0 same size | 1 nearest x2 up (layers.py:600 F.interpolate) | 2 the four
 *   stride-2 parity planes src[2y+py][2x+px], plane = 2 py + px, zero beyond the edge (layers.py:621-623 pad (0,1,0,1) +
 *   stride-2 conv) | 3 2x2 average pool of the transformed values (BeatGANsblocks.py Downsample, avg_pool_nd).
 * Outputs: out_hi/out_lo fp16 [planes][N][Ho][Wo][Cp] = hi/lo pair of scale * v (channels >= C zero), Cp % 64 == 0;
 *          out_f32 [planes][N][Ho][Wo][C] (optional).  src_nchw: src0 is [N][C0][Hs][Ws] (the network input). */
#define IDB_PREP_SAME 0
#define IDB_PREP_UP2 1
#define IDB_PREP_STRIDE2 2
#define IDB_PREP_AVGPOOL2 3
int idb_conv_prep(const float* src0, int C0, const float* src1, int C1, int N, int Hs, int Ws, int src_nchw, int mode,
                  const float* mean_rstd, int groups, const float* gamma, const float* beta, int act, float scale,
                  void* out_hi, void* out_lo, int Cp, float* out_f32, void* stream);

/* Operand preparation with the FIR resampling FOLDED IN (BASELINE.json: "upfirdn2d ... fused into the adjacent conv"):
 *   v = upfirdn2d(act(GN(src0)) , kernel, up, down, (pad0, pad1))        [op/upfirdn2d.py:145 semantics, kernel <= 4 x 4]
 * written as the convolution operand (out_hi/out_lo) and / or fp32 (out_f32), gathered per output pixel so the
 * normalised, activated and the filtered tensors are never materialised.  Covers layerspp.py:244-257 (BigGAN res-block:
 * downsample_2d = up 1, down 2, pad (1, 1); upsample_2d = kernel * 4, up 2, down 1, pad (2, 1)) and, with as_planes,
 * up_or_down_sampling.py:144-178 conv_downsample_2d: FIR pad (2, 2) at down = 1 followed by a stride-2 convolution,
 * emitted directly as the four stride-2 parity planes idb_conv2d reads.  kernel_host: HOST array [kh][kw].
 * out_raw (optional, fp32): the same filter applied to the UNtransformed src0 in the same gather -- the res-block's skip
 * branch upfirdn2d(x) (layerspp.py:246-248,253-255) costs no second read of the activation. */
int idb_conv_prep_fir(const float* src0, int C0, int N, int Hs, int Ws, const float* mean_rstd, int groups,
                      const float* gamma, const float* beta, int act, const float* kernel_host, int kh, int kw, int up,
                      int down, int pad0, int pad1, int as_planes, float scale, void* out_hi, void* out_lo, int Cp,
                      float* out_f32, float* out_raw, void* stream);

/* nn.Conv2d / NIN as an implicit GEMM on tcgen05 (3 MMAs per product on fp16 hi/lo pairs, fp32 accumulate in TMEM):
 *   out[n,h,w,co] = out_scale * (acc_scale * sum_{t,ci} A[plane_t][n, h+dy_t, w+dx_t, ci] * Wt[co][t][ci]
 *                                + bias[co] + residual[n,h,w,co])
 * A = idb_conv_prep output [planes][N][Ha][Wa][Cp] (Ha >= H, Wa >= W: the planes of a stride-2 read of an odd-sized
 * tensor are one row/column larger than the output); reads outside the planes are zero (TMA out-of-bounds fill = the
 * convolution's zero padding).  Wt = idb_pack_weight of the matrix [Cout][n_taps*Cp] (tap-major, channel-minor).
 * taps_host: HOST array [n_taps][3] = (plane, dy, dx).  3x3 pad 1: plane 0, dy,dx in {-1,0,1};  3x3 stride 2 after
 * pad (0,1,0,1): tap (ky,kx) -> plane 2(ky&1)+(kx&1), dy = ky>>1, dx = kx>>1;  1x1 / NIN: one tap (0,0,0).
 * bias [Cout] / residual [N][H][W][ld_res] may be NULL.  H, W powers of two (<= 128) so that 128-pixel tiles are boxes.
 * Replaces cuDNN conv + bias add + temb add (layers.py:662-664; t is constant on this path so Dense_0(act(temb)) is
 * folded into `bias` by the host) + residual add (:675) + skip rescale (layerspp.py:271-274). */
int idb_conv2d(const void* a_hi, const void* a_lo, int planes, int N, int Ha, int Wa, int H, int W, int Cp, const void* w_hi,
               const void* w_lo, int n_taps, const int* taps_host, int Np, int Cout, float acc_scale, float out_scale,
               const float* bias, const float* residual, int ld_res, float* out, int ldo, void* stream);

/* idb_conv2d that ALSO emits the GroupNorm statistics of the tensor it writes, from the epilogue registers (the fusion
 * BASELINE.json asks for: "GroupNorm ... fused into their consumers" -- the next nn.GroupNorm, layers.py:638,648, no
 * longer re-reads the activation): stats [N][slots][Cout][2] fp32 = per (image, 32-pixel slot, channel) sum and sum of
 * squares of out, slots = idb_conv_stats_slots(H, W) (0: this shape cannot carry statistics, use idb_groupnorm_stats).
 * Fixed slots, no atomics: bit-reproducible.  idb_groupnorm_from_conv_stats folds the partials of the (virtually
 * concatenated) producers [stats0 | stats1] into mean_rstd [N][groups][2] in fp64, same result layout as
 * idb_groupnorm_stats. */
int idb_conv_stats_slots(int H, int W);
int idb_conv2d_stats(const void* a_hi, const void* a_lo, int planes, int N, int Ha, int Wa, int H, int W, int Cp,
                     const void* w_hi, const void* w_lo, int n_taps, const int* taps_host, int Np, int Cout,
                     float acc_scale, float out_scale, const float* bias, const float* residual, int ld_res, float* out,
                     int ldo, float* stats, void* stream);
int idb_groupnorm_from_conv_stats(const float* stats0, int C0, int slots0, const float* stats1, int C1, int slots1, int N,
                                  int HW, int groups, double eps, float* mean_rstd, void* stream);

/* Self-attention over the T = H*W pixels of each image (layers.py:567-590 AttnBlock: 1 head, scale C^-1/2;
 * BeatGANsblocks.py:434-491 QKVAttentionLegacy: heads, scale dh^-1/2): out[n][i][h*dh+c] =
 * sum_j softmax_j(scale <q_i,k_j>) v[j][c], q/k/v of head h at channels q0/k0/v0 + h*head_stride of qkv [N][T][ld].
 * fp32 on the CUDA cores (< 1 % of the network FLOPs), scores resident in shared memory. */
int idb_attention(const float* qkv, int N, int T, int ld, int heads, int dh, int q0, int k0, int v0, int head_stride,
                  float scale, float* out, int ld_out, void* stream);

/* The same attention on the TENSOR CORES for one head of T = H*W >= 128 tokens (T % 128 == 0, dh % 128 == 0: the 16x16
 * attention blocks of the 64x64 networks, 256 tokens x 256 channels): S = scale Q K^T and O = softmax(S) V run on
 * conv_fp16x3_kernel (tcgen05, 3 MMAs per product) as 1x1 convolutions whose weight matrix is the image's own K (then
 * V^T); Q / K / V^T are packed as fp16 hi/lo operands and the softmax (the reference's max / exp / sum / divide order)
 * writes P as the second convolution's operand.  workspace: idb_attention_tc_workspace_bytes(N, T, dh), 1024-byte aligned.
 * Other shapes (4x4 mid-block, multi-head reduced-width nets): idb_attention. */
size_t idb_attention_tc_workspace_bytes(int N, int T, int dh);
int idb_attention_tc(const float* qkv, int N, int H, int W, int ld, int dh, int q0, int k0, int v0, float scale,
                     void* workspace, float* out, int ld_out, void* stream);

/* Row-side centred Gram for one anchor when K < D (image shapes: K = 4096 perturbations, D = 12288 pixels):
 * colmean[D] (fp64) and G[K][K] = (scale (X - mean)) (scale (X - mean))^T in fp64; its eigenvalues are the squared
 * singular values of the centred K x D score matrix, the same ones the D-side Gram of idb_center_gram has
 * (dim_reduction.py:190-199).  X [K][ld] fp32. */
int idb_row_gram(const float* X, int ld, int K, int D, double scale, double* colmean, double* G, void* stream);

/* Singular values from ONE large Gram matrix (n up to 14000) by the whole GPU: cooperative Householder
 * tridiagonalisation (fp64, G is overwritten) + Sturm bisection.  sv [n_out] fp32 descending.
 * workspace: idb_spectrum_large_workspace_bytes(n).  Replaces torch.linalg.svd at dim_reduction.py:197-198 for
 * image-sized score matrices. */
size_t idb_spectrum_large_workspace_bytes(int n);
int idb_spectrum_large(double* G, int n, int n_out, float* sv, void* workspace, void* stream);
/* Same, with the tridiagonalisation kernel chosen by the caller (cross-checks and tuning): 0 = automatic (what
 * idb_spectrum_large runs), 1 = fused one-pass kernel (n <= 9000), 2 = fused kernel with two column classes
 * (n <= 14000; the automatic choice above 9000), 3 = panel kernel with deferred rank-2 updates (even n; (8 + 8/NB) m^2
 * bytes per column instead of 16), -1 = unfused two-pass kernel. */
int idb_spectrum_large_variant(double* G, int n, int n_out, float* sv, void* workspace, int variant, void* stream);

/* idb_spectrum_large WITHOUT a workspace argument and without any allocation: the workspace is a file-scope device array
 * inside the library (768 KB, held by every GPU that loads the library until the process exits), so the call is usable
 * where cudaMalloc is not (under CUDA graph capture, in tight serving loops, in harnesses that forbid it).  It has no
 * counterpart in the reference.  Limit: n <= 4096
 * (idb_spectrum_large_static_max_n()); a larger n is rejected with an error, use idb_spectrum_large there.  Same
 * arguments and output as idb_spectrum_large otherwise (G [n][n] fp64 is overwritten; sv [n_out] fp32 descending); it
 * always runs the one-class fused tridiagonalisation kernel (idb_spectrum_large_variant 1), whose workspace is 8 n
 * doubles -- the panel kernel idb_spectrum_large prefers from n = 3584 keeps megabytes of per-CTA partial sums.  Every call on a device shares the one array: calls that overlap on the GPU race -- serialise them.  Nothing
 * needs clearing between calls (every workspace word is written before it is read). */
#define IDB_SPECTRUM_STATIC_MAX_N 4096
int idb_spectrum_large_static_max_n(void);
int idb_spectrum_large_static(double* G, int n, int n_out, float* sv, void* stream);

#ifdef __cplusplus
}
#endif
#endif
