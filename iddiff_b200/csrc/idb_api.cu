// C ABI of libiddiff_b200.so -- see include/iddiff_b200.h for the contract of every entry point.
#include "../../include/iddiff_b200.h"
#include "idb_kernels.h"
#include <math.h>
#include <stdio.h>

namespace {
thread_local char g_err[512] = "";

int fail(const char* what) {
    snprintf(g_err, sizeof(g_err), "%s", what);
    return -1;
}
int check(cudaError_t e, const char* where) {
    if (e == cudaSuccess) return 0;
    snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
    return -2;
}
int num_sms() {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}
bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
}  // namespace

namespace idb {
int api_fail(const char* what) { return fail(what); }
int api_check(cudaError_t e, const char* where) { return check(e, where); }
int api_num_sms() { return num_sms(); }
}  // namespace idb

extern "C" {

int idb_abi_version(void) { return 1; }
const char* idb_last_error(void) { return g_err; }

int idb_perturb_rows(const float* x0, const float* z, int n_anchors, int K, int D, int Kp, int Dp, float mean_coeff,
                     float std, uint64_t seed, uint64_t anchor_id0, uint32_t row0, float* x_out, int fmt, float d_scale,
                     void* d_hi, void* d_lo, void* stream);

int idb_perturb(const float* x0, const float* z, int n_anchors, int K, int D, int Kp, int Dp, float mean_coeff,
                float std, uint64_t seed, uint64_t anchor_id0, float* x_out, int fmt, float d_scale, void* d_hi,
                void* d_lo, void* stream) {
    return idb_perturb_rows(x0, z, n_anchors, K, D, Kp, Dp, mean_coeff, std, seed, anchor_id0, 0u, x_out, fmt, d_scale,
                            d_hi, d_lo, stream);
}

int idb_perturb_rows(const float* x0, const float* z, int n_anchors, int K, int D, int Kp, int Dp, float mean_coeff,
                     float std, uint64_t seed, uint64_t anchor_id0, uint32_t row0, float* x_out, int fmt, float d_scale,
                     void* d_hi, void* d_lo, void* stream) {
    if (!x0 || n_anchors < 0 || K <= 0 || D <= 0) return fail("idb_perturb: bad shape");
    if (d_hi && fmt != IDB_FMT_BF16 && fmt != IDB_FMT_FP16) return fail("idb_perturb: unknown fmt");
    if (d_hi && !(d_scale > 0.f)) return fail("idb_perturb: d_scale must be positive");
    if ((d_hi == nullptr) != (d_lo == nullptr)) return fail("idb_perturb: d_hi and d_lo must both be set or both NULL");
    if (d_hi && (Dp % 64 || Kp < K || Dp < D || !aligned16(d_hi) || !aligned16(d_lo)))
        return fail("idb_perturb: need Dp % 64 == 0, Kp >= K, Dp >= D, 16-byte aligned slabs");
    if (!d_hi && !x_out) return fail("idb_perturb: no output requested");
    idb::PerturbParams p;
    p.x0 = x0; p.z = z; p.n_anchors = n_anchors; p.K = K; p.D = D;
    p.Kp = d_hi ? Kp : K;
    p.Dp = d_hi ? Dp : ((D + 7) & ~7);
    p.mean_coeff = mean_coeff; p.std = std; p.seed = seed; p.anchor_id0 = anchor_id0; p.row0 = row0;
    p.fmt = fmt; p.d_scale = d_scale;
    p.x_out = x_out; p.d_hi = static_cast<uint8_t*>(d_hi); p.d_lo = static_cast<uint8_t*>(d_lo);
    return check(idb::launch_perturb(p, static_cast<cudaStream_t>(stream)), "idb_perturb");
}

int idb_pack_weight(const float* w, int N, int K, int ldw, int Np, int Kc, int w_fmt, float scale, void* hi, void* lo,
                    void* stream) {
    const int IDB_W_BF16 = IDB_FMT_BF16, IDB_W_FP16 = IDB_FMT_FP16;
    if (!w || !hi || !lo || N <= 0 || K <= 0 || ldw < K) return fail("idb_pack_weight: bad argument");
    if (w_fmt != IDB_W_BF16 && w_fmt != IDB_W_FP16) return fail("idb_pack_weight: unknown w_fmt");
    if (!(scale > 0.f)) return fail("idb_pack_weight: scale must be positive");
    if (Np % 128 || Kc % 64 || Np < N || Kc < K || !aligned16(hi) || !aligned16(lo))
        return fail("idb_pack_weight: need Np % 128 == 0, Kc % 64 == 0, Np >= N, Kc >= K, 16-byte aligned slabs");
    return check(idb::launch_pack_weight(w, N, K, ldw, Np, Kc, w_fmt, scale, static_cast<uint8_t*>(hi), static_cast<uint8_t*>(lo),
                                         static_cast<cudaStream_t>(stream)),
                 "idb_pack_weight");
}

int idb_anchor_linear(const float* in, int ld_in, int apply_elu, float in_scale, const float* w, int ldw,
                      const float* bias, const float* extra_col, int extra_stride, float extra_scalar, float* out,
                      int ld_out, int n_anchors, int N, int K, void* stream) {
    if (!in || !w || !out || n_anchors < 0 || N <= 0 || K <= 0 || ld_in < K || ldw < K || ld_out < N)
        return fail("idb_anchor_linear: bad argument");
    return check(idb::launch_anchor_linear(in, ld_in, apply_elu, in_scale, w, ldw, bias, extra_col, extra_stride,
                                           extra_scalar, out, ld_out, n_anchors, N, K,
                                           static_cast<cudaStream_t>(stream)),
                 "idb_anchor_linear");
}

static int gemm_common(int mode, const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, int w_fmt,
                       float w_inv_scale, float out_scale, int M, int Np, int Kc, const float* a0, int rows_per_anchor,
                       int n_anchors, void* out_hi, void* out_lo, float* out_f32,
                       void* stream, const char* name) {
    if (!a_hi || !a_lo || !w_hi || !w_lo) return fail("idb_fcn_*_layer: NULL operand");
    if (w_fmt != IDB_FMT_BF16 && w_fmt != IDB_FMT_FP16) return fail("idb_fcn_*_layer: unknown fmt");
    if (!(w_inv_scale > 0.f) || !(out_scale > 0.f)) return fail("idb_fcn_*_layer: scales must be positive");
    if (M < 0 || M % 128 || Np <= 0 || Np % 256 || Kc <= 0 || Kc % 64)
        return fail("idb_fcn_*_layer: need M % 128 == 0, Np % 256 == 0, Kc % 64 == 0");
    if (!aligned16(a_hi) || !aligned16(a_lo) || !aligned16(w_hi) || !aligned16(w_lo))
        return fail("idb_fcn_*_layer: operands must be 16-byte aligned");
    idb::GemmParams p;
    p.a_hi = static_cast<const uint8_t*>(a_hi); p.a_lo = static_cast<const uint8_t*>(a_lo);
    p.b_hi = static_cast<const uint8_t*>(w_hi); p.b_lo = static_cast<const uint8_t*>(w_lo);
    p.M = M; p.Np = Np; p.Kc = Kc;
    p.fmt = w_fmt; p.acc_scale = w_inv_scale; p.out_scale = out_scale;
    p.a0 = a0; p.rows_per_anchor = rows_per_anchor; p.n_anchors = n_anchors;
    p.out_hi = static_cast<uint8_t*>(out_hi); p.out_lo = static_cast<uint8_t*>(out_lo);
    p.out_f32 = out_f32;
    return check(idb::launch_gemm_bf16x3(mode, p, num_sms(), static_cast<cudaStream_t>(stream)), name);
}

int idb_fcn_hidden_layer(const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, int w_fmt,
                         float w_inv_scale, float out_scale, int M, int Np, int Kc, const float* a0,
                         int rows_per_anchor, int n_anchors, void* out_hi, void* out_lo, void* stream) {
    if (!a0 || !out_hi || !out_lo || rows_per_anchor < 128 || n_anchors <= 0 ||
        (long long)n_anchors * rows_per_anchor > (long long)M)
        return fail("idb_fcn_hidden_layer: need a0, outputs, rows_per_anchor >= 128 and n_anchors * rows_per_anchor <= M");
    if (!aligned16(out_hi) || !aligned16(out_lo)) return fail("idb_fcn_hidden_layer: outputs must be 16-byte aligned");
    return gemm_common(0, a_hi, a_lo, w_hi, w_lo, w_fmt, w_inv_scale, out_scale, M, Np, Kc, a0, rows_per_anchor, n_anchors,
                       out_hi, out_lo, nullptr, stream,
                       "idb_fcn_hidden_layer");
}

int idb_fcn_output_layer(const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, int w_fmt,
                         float w_inv_scale, int M, int Np, int Kc, float* out_f32, void* stream) {
    if (!out_f32 || !aligned16(out_f32)) return fail("idb_fcn_output_layer: out_f32 must be set and 16-byte aligned");
    return gemm_common(1, a_hi, a_lo, w_hi, w_lo, w_fmt, w_inv_scale, 1.0f, M, Np, Kc, nullptr, 128, 1, nullptr, nullptr, out_f32, stream,
                       "idb_fcn_output_layer");
}

int idb_fcn_anchor_layer(const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, int fmt,
                         float acc_scale, float out_scale, int M, int Np, int Kc, const float* bias, float* out_f32,
                         void* out_hi, void* out_lo, void* stream) {
    if (!out_f32 || !aligned16(out_f32)) return fail("idb_fcn_anchor_layer: out_f32 must be set and 16-byte aligned");
    if ((out_hi == nullptr) != (out_lo == nullptr) || !aligned16(out_hi) || !aligned16(out_lo))
        return fail("idb_fcn_anchor_layer: out_hi/out_lo must both be set (16-byte aligned) or both NULL");
    return gemm_common(2, a_hi, a_lo, w_hi, w_lo, fmt, acc_scale, out_scale, M, Np, Kc, bias, 128, 1, out_hi, out_lo,
                       out_f32, stream, "idb_fcn_anchor_layer");
}

int idb_fcn_scores(const float* a_out, int ld_a, const float* d_out, int ld_d, int n_anchors, int K, int Kp, int D,
                   float std, float* score, void* stream) {
    if (!a_out || !d_out || !score || K <= 0 || Kp < K || D <= 0 || ld_a < D || ld_d < D || !(std > 0.f))
        return fail("idb_fcn_scores: bad argument");
    return check(idb::launch_fcn_scores(a_out, ld_a, d_out, ld_d, n_anchors, K, Kp, D, std, score,
                                        static_cast<cudaStream_t>(stream)),
                 "idb_fcn_scores");
}

int idb_ksphere_gt_score(const float* x, const float* sigma_rows, float sigma, int rows, int D, int manifold_dim,
                         float* out, void* stream) {
    if (!x || !out || rows < 0 || D <= 0 || manifold_dim < 2) return fail("idb_ksphere_gt_score: bad argument");
    if (!sigma_rows && !(sigma > 0.f)) return fail("idb_ksphere_gt_score: sigma must be positive");
    return check(idb::launch_ksphere_score(x, sigma_rows, sigma, rows, D, manifold_dim, out,
                                           static_cast<cudaStream_t>(stream)),
                 "idb_ksphere_gt_score");
}

int idb_center_gram(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D, double scale,
                    double* colmean, double* G, void* stream) {
    if (!X || !colmean || !G || n_anchors < 0 || K <= 0 || D <= 0 || ld < D)
        return fail("idb_center_gram: bad argument");
    return check(idb::launch_center_gram(X, anchor_stride, ld, n_anchors, K, D, scale, colmean, G,
                                         static_cast<cudaStream_t>(stream)),
                 "idb_center_gram");
}

int idb_column_sums(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D, double* colsum,
                    void* stream) {
    if (!X || !colsum || n_anchors < 0 || K <= 0 || D <= 0 || ld < D) return fail("idb_column_sums: bad argument");
    return check(idb::launch_column_sums(X, anchor_stride, ld, n_anchors, K, D, 1.0, colsum,
                                         static_cast<cudaStream_t>(stream)),
                 "idb_column_sums");
}

int idb_gram_about(const float* X, long long anchor_stride, int ld, int n_anchors, int K, int D, double scale,
                   const double* center, double* G, void* stream) {
    if (!X || !center || !G || n_anchors < 0 || K <= 0 || D <= 0 || ld < D) return fail("idb_gram_about: bad argument");
    return check(idb::launch_gram_about(X, anchor_stride, ld, n_anchors, K, D, scale, center, G,
                                        static_cast<cudaStream_t>(stream)),
                 "idb_gram_about");
}

/* ---- DSM training step ---- */
int idb_dsm_sample(const float* data, const long long* idx, const float* t, const float* mean_coeff, const float* stdv,
                   const float* z_ext, int B, int D, int Mp, int ldx, float label_scale, uint64_t seed, uint64_t step,
                   const long long* step_dev, float* X, float* z_out, void* stream) {
    if (!data || !t || !mean_coeff || !stdv || !X || !z_out || B <= 0 || D <= 0 || Mp < B || ldx < D + 1 || ldx % 8 ||
        !aligned16(X))
        return fail("idb_dsm_sample: bad argument (need Mp >= B, ldx >= D + 1, ldx % 8 == 0, 16-byte aligned X)");
    return check(idb::launch_dsm_sample(data, idx, t, mean_coeff, stdv, z_ext, B, D, Mp, ldx, label_scale, seed, step,
                                        step_dev, X, z_out, static_cast<cudaStream_t>(stream)),
                 "idb_dsm_sample");
}

int idb_train_pack(const float* src, int ld, int R, int C, const float* pre, int ld_pre, int mode, int transpose, int fmt,
                   float scale, int Rp, int Cp, void* hi, void* lo, float* out_f32, int ld_out, void* stream) {
    if (!src || R < 0 || C < 0 || ld < C || mode < 0 || mode > 2 || (mode == 2 && (!pre || ld_pre < C)))
        return fail("idb_train_pack: bad argument");
    if (fmt != IDB_FMT_BF16 && fmt != IDB_FMT_FP16) return fail("idb_train_pack: unknown fmt");
    if (!(scale > 0.f)) return fail("idb_train_pack: scale must be positive");
    if (transpose) {
        if (!hi || !lo || out_f32) return fail("idb_train_pack: the transposed form writes the operand slabs only");
        if (Cp % 128 || Rp % 64 || Cp < C || Rp < R) return fail("idb_train_pack: transposed slab needs Cp % 128 == 0, Rp % 64 == 0");
    } else {
        if ((hi == nullptr) != (lo == nullptr) || (!hi && !out_f32)) return fail("idb_train_pack: no output requested");
        if (Rp % 8 || Cp % 64 || Cp < C || Rp < R) return fail("idb_train_pack: slab needs Rp % 8 == 0, Cp % 64 == 0");
        if (out_f32 && ld_out < C) return fail("idb_train_pack: ld_out < C");
    }
    if (hi && (!aligned16(hi) || !aligned16(lo))) return fail("idb_train_pack: slabs must be 16-byte aligned");
    idb::TrainPackParams p;
    p.src = src; p.ld = ld; p.R = R; p.C = C; p.pre = pre; p.ld_pre = ld_pre; p.mode = mode; p.transpose = transpose;
    p.fmt = fmt; p.scale = scale; p.Rp = Rp; p.Cp = Cp; p.hi = static_cast<uint8_t*>(hi); p.lo = static_cast<uint8_t*>(lo);
    p.out_f32 = out_f32; p.ld_out = ld_out;
    return check(idb::launch_train_pack(p, static_cast<cudaStream_t>(stream)), "idb_train_pack");
}

int idb_dsm_loss(const float* out, int ldo, const float* z, int B, int D, int Mp, float* dZ, float* loss, double* scratch,
                 void* stream) {
    if (!out || !z || !dZ || !loss || !scratch || B <= 0 || D <= 0 || Mp < B || ldo < D)
        return fail("idb_dsm_loss: bad argument");
    return check(idb::launch_dsm_loss(out, ldo, z, B, D, Mp, dZ, loss, scratch, static_cast<cudaStream_t>(stream)),
                 "idb_dsm_loss");
}

int idb_column_sums_f32(const float* src, int ld, int R, int C, float* out, void* stream) {
    if (!src || !out || R < 0 || C < 0 || ld < C) return fail("idb_column_sums_f32: bad argument");
    return check(idb::launch_colsum_f32(src, ld, R, C, out, static_cast<cudaStream_t>(stream)), "idb_column_sums_f32");
}

int idb_grad_sumsq(const float* g, int ld, int N, int K, double* accum, void* stream) {
    if (!g || !accum || N < 0 || K < 0 || ld < K) return fail("idb_grad_sumsq: bad argument");
    return check(idb::launch_grad_sumsq(g, ld, N, K, accum, static_cast<cudaStream_t>(stream)), "idb_grad_sumsq");
}

int idb_adam_step(float* param, int ldp, const float* grad, int ldg, float* m, float* v, int N, int K, float lr,
                  int warmup, float beta1, float beta2, float eps, long long step, const long long* step_dev,
                  float max_norm, const double* grad_normsq, void* stream) {
    if (!param || !grad || !m || !v || N < 0 || K < 0 || ldp < K || ldg < K || (!step_dev && step < 1))
        return fail("idb_adam_step: bad argument");
    return check(idb::launch_adam(param, ldp, grad, ldg, m, v, N, K, lr, warmup, beta1, beta2, eps, step, step_dev, max_norm,
                                  grad_normsq, static_cast<cudaStream_t>(stream)),
                 "idb_adam_step");
}

size_t idb_spectrum_workspace_bytes(int n_anchors, int n) {
    if (n_anchors <= 0 || n <= 0) return 256;
    const size_t a = idb::spectrum_workspace_bytes(n_anchors, n), b = idb::tridiag_workspace_bytes(n_anchors, n);
    return a > b ? a : b;
}

int idb_spectrum(const double* G, int n_anchors, int n, int n_out, int method, float* sv, int* sweeps, void* workspace,
                 void* stream) {
    if (method == IDB_SPECTRUM_TRIDIAG) {
        if (!G || !sv || !workspace || n_anchors < 0 || n <= 0 || n_out <= 0 || n_out > n)
            return fail("idb_spectrum: bad argument");
        if (!aligned16(workspace)) return fail("idb_spectrum: workspace must be 16-byte aligned");
        return check(idb::launch_spectrum_tridiag(G, n_anchors, n, n_out, sv, sweeps, workspace,
                                                  static_cast<cudaStream_t>(stream)),
                     "idb_spectrum(tridiag)");
    }
    if (method != IDB_SPECTRUM_JACOBI) return fail("idb_spectrum: unknown method");
    if (!G || !sv || !workspace || n_anchors < 0 || n <= 0 || n_out <= 0 || n_out > n)
        return fail("idb_spectrum: bad argument");
    if (!aligned16(workspace)) return fail("idb_spectrum: workspace must be 16-byte aligned");
    return check(idb::launch_spectrum(G, n_anchors, n, n_out, sv, sweeps, workspace,
                                      static_cast<cudaStream_t>(stream)),
                 "idb_spectrum");
}

int idb_id_from_spectrum(const float* sv, int n_anchors, int n, int ld, int* dims, void* stream) {
    if (!sv || !dims || n_anchors < 0 || n < 3 || ld < n) return fail("idb_id_from_spectrum: need n >= 3, ld >= n");
    return check(idb::launch_id_readout(sv, n_anchors, n, ld, dims, static_cast<cudaStream_t>(stream)),
                 "idb_id_from_spectrum");
}

}  // extern "C"
