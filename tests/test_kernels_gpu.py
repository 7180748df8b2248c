"""GPU: every C-ABI kernel against a plain torch fp64 reference of the same op (unit level).
End-to-end parity against the oracle / golden fixtures lives in test_parity_gpu.py."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from iddiff_b200 import _lib, layout  # noqa: E402
from iddiff_b200._lib import call, ptr, stream_ptr  # noqa: E402


def dev():
    return torch.device("cuda:0")


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dev())


def test_pack_weight_matches_host_layout():
    N, K, Np, Kc = 200, 100, 256, 128
    w = _rand(N, K + 1, seed=1)  # ldw = K + 1 like the first FCN layer
    hi = torch.zeros(Np * Kc * 2, dtype=torch.uint8, device=dev())
    lo = torch.zeros_like(hi)
    call("idb_pack_weight", ptr(w), N, K, K + 1, Np, Kc, _lib.FMT_BF16, 1.0, ptr(hi), ptr(lo), stream_ptr())
    torch.cuda.synchronize()
    full = torch.zeros(Np, Kc, device=dev())
    full[:N, :K] = w[:, :K]
    h, l = layout.split_bf16(full)
    assert torch.equal(layout.from_slab(hi, Np, Kc).view(torch.int16), h.view(torch.int16))
    assert torch.equal(layout.from_slab(lo, Np, Kc).view(torch.int16), l.view(torch.int16))
    assert torch.equal(layout.to_slab(h), hi)


A_SCALE = 64.0


def _split16(x, fmt):
    dt = torch.float16 if fmt == _lib.FMT_FP16 else torch.bfloat16
    h = x.to(dt)
    return h, (x - h.float()).to(dt)


def _slabs(x, fmt=_lib.FMT_BF16):
    """Activation operand in format fmt: (hi slab, lo slab, fp64 value represented, scale)."""
    sc = A_SCALE if fmt == _lib.FMT_FP16 else 1.0
    h, l = _split16(x * sc, fmt)
    return layout.to_slab(h.view(torch.bfloat16)), layout.to_slab(l.view(torch.bfloat16)), (h.double() + l.double()) / sc, sc


def _w_slabs(w, fmt):
    """Device-packed weight slabs in either format + the fp64 value they represent + 1/scale."""
    Np, Kc = w.shape
    scale = 1.0
    if fmt == _lib.FMT_FP16:
        scale = 2.0 ** (14 - math.frexp(float(w.abs().max()))[1])
    hi = torch.zeros(Np * Kc * 2, dtype=torch.uint8, device=dev())
    lo = torch.zeros_like(hi)
    call("idb_pack_weight", ptr(w), Np, Kc, Kc, Np, Kc, fmt, scale, ptr(hi), ptr(lo), stream_ptr())
    torch.cuda.synchronize()
    if fmt == _lib.FMT_FP16:
        h = (w * scale).to(torch.float16)
        l = (w * scale - h.float()).to(torch.float16)
        assert torch.equal(layout.from_slab(hi, Np, Kc).view(torch.int16), h.view(torch.int16))
        assert torch.equal(layout.from_slab(lo, Np, Kc).view(torch.int16), l.view(torch.int16))
        return hi, lo, (h.double() + l.double()) / scale, 1.0 / scale
    h, l = layout.split_bf16(w)
    return hi, lo, h.double() + l.double(), 1.0


@pytest.mark.parametrize("fmt", [_lib.FMT_BF16, _lib.FMT_FP16])
@pytest.mark.parametrize("M,Np,Kc", [(128, 256, 64), (256, 512, 128), (384, 256, 2048), (19 * 128, 2048, 256)])
def test_output_layer_gemm(M, Np, Kc, fmt):
    a = _rand(M, Kc, seed=2, scale=0.01)
    w = _rand(Np, Kc, seed=3, scale=1 / math.sqrt(Kc))
    a_hi, a_lo, a64, asc = _slabs(a, fmt)
    w_hi, w_lo, w64, inv = _w_slabs(w, fmt)
    if fmt == _lib.FMT_FP16:  # 22 significand bits
        assert (w64 - w.double()).abs().max().item() < 2 ** -21 * w.abs().max().item()
    out = torch.full((M, Np), float("nan"), device=dev())
    call("idb_fcn_output_layer", ptr(a_hi), ptr(a_lo), ptr(w_hi), ptr(w_lo), fmt, inv / asc, M, Np, Kc, ptr(out),
         stream_ptr())
    torch.cuda.synchronize()
    ref = a64 @ w64.T
    err = (out.double() - ref).abs().max().item()
    scale = ref.abs().max().item()
    # dropped lo*lo term (2^-16 relative per product for bf16 pairs, 2^-22 for fp16 pairs), fp32 accumulation
    assert err < (3e-5 if fmt == _lib.FMT_BF16 else 2e-5) * scale, (err, scale)  # fp32 tensor-core accumulation over Kc
    # and the split operands reproduce the fp32 inputs to 2^-16 (bf16 pair) / 2^-22 + 2^-30 absolute (fp16 pair)
    tol = 2 ** -15 * a.abs().max().item() if fmt == _lib.FMT_BF16 else 2 ** -21 * a.abs().max().item()
    assert (a64 - a.double()).abs().max().item() < tol


def _elu(x):
    return torch.where(x > 0, x, torch.expm1(x))


@pytest.mark.parametrize("fmt", [_lib.FMT_BF16, _lib.FMT_FP16])
@pytest.mark.parametrize("A,Kp,Np,Kc", [(1, 128, 256, 64), (3, 256, 512, 192), (2, 128, 2048, 2048),
                                        (3, 151, 256, 64), (5, 333, 512, 128)])   # dense packing: tiles straddle anchors
def test_hidden_layer_gemm_elu_deviation(A, Kp, Np, Kc, fmt):
    M = (A * Kp + 127) // 128 * 128
    a = _rand(M, Kc, seed=4, scale=0.02)
    a[A * Kp:] = 0                                   # rows past the last anchor are padding
    w = _rand(Np, Kc, seed=5, scale=1 / math.sqrt(Kc))
    a0 = _rand(A, Np, seed=6, scale=0.5)
    a_hi, a_lo, a64, asc = _slabs(a, fmt)
    w_hi, w_lo, w64, inv = _w_slabs(w, fmt)
    o_hi = torch.zeros(M * Np * 2, dtype=torch.uint8, device=dev())
    o_lo = torch.zeros_like(o_hi)
    call("idb_fcn_hidden_layer", ptr(a_hi), ptr(a_lo), ptr(w_hi), ptr(w_lo), fmt, inv / asc, asc, M, Np, Kc, ptr(a0),
         Kp, A, ptr(o_hi), ptr(o_lo), stream_ptr())
    torch.cuda.synchronize()
    dt = torch.float16 if fmt == _lib.FMT_FP16 else torch.bfloat16
    got = (layout.from_slab(o_hi, M, Np).view(dt).double() + layout.from_slab(o_lo, M, Np).view(dt).double()) / asc
    d = a64 @ w64.T
    base = torch.cat([a0.double().repeat_interleave(Kp, dim=0), a0[-1:].double().expand(M - A * Kp, Np)])
    ref = _elu(base + d) - _elu(base)
    assert float(got[A * Kp:].abs().max() if M > A * Kp else 0.0) == 0.0
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err < (5e-5 if fmt == _lib.FMT_BF16 else 2e-5) * scale, (err, scale)


def test_perturb_external_noise_is_reference_arithmetic():
    A, K, D, Kp, Dp = 3, 151, 100, 256, 128
    x0 = _rand(A, D, seed=7)
    z = _rand(A, K, D, seed=8)
    mc, std = 0.9995, 0.0105
    x = torch.empty(A, K, D, device=dev())
    hi = torch.empty(A * Kp * Dp * 2, dtype=torch.uint8, device=dev())
    lo = torch.empty_like(hi)
    call("idb_perturb", ptr(x0), ptr(z), A, K, D, Kp, Dp, mc, std, 0, 0, ptr(x), _lib.FMT_BF16, 1.0, ptr(hi), ptr(lo),
         stream_ptr())
    torch.cuda.synchronize()
    mcf, stdf = torch.tensor(mc, dtype=torch.float32).item(), torch.tensor(std, dtype=torch.float32).item()
    mean = (torch.tensor(mcf, device=dev()) * x0)[:, None, :]
    ref = mean + torch.tensor(stdf, device=dev()) * z
    assert torch.equal(x, ref)  # bit-exact: same fp32 ops in the same order
    devn = torch.zeros(A, Kp, Dp, device=dev())
    devn[:, :K, :D] = ref - mean
    h, l = layout.split_bf16(devn.reshape(A * Kp, Dp))
    assert torch.equal(layout.from_slab(hi, A * Kp, Dp).view(torch.int16), h.view(torch.int16))
    assert torch.equal(layout.from_slab(lo, A * Kp, Dp).view(torch.int16), l.view(torch.int16))


def test_perturb_philox_statistics_and_sharding_invariance():
    A, K, D = 4, 1501, 200
    x0 = torch.zeros(A, D, device=dev())
    x = torch.empty(A, K, D, device=dev())
    call("idb_perturb", ptr(x0), None, A, K, D, 0, 0, 1.0, 1.0, 1234, 10, ptr(x), 0, 1.0, None, None, stream_ptr())
    torch.cuda.synchronize()
    assert abs(x.mean().item()) < 5e-3 and abs(x.var().item() - 1) < 5e-3
    assert abs((x ** 4).mean().item() - 3) < 0.05
    # columns / rows / anchors are uncorrelated
    assert abs((x[:, :, 0] * x[:, :, 1]).mean().item()) < 0.05
    assert abs((x[0] * x[1]).mean().item()) < 0.01
    # anchor with global id 12 is the same whether it is slot 2 of a chunk starting at 10 or slot 0 of one at 12
    y = torch.empty(1, K, D, device=dev())
    call("idb_perturb", ptr(x0[:1]), None, 1, K, D, 0, 0, 1.0, 1.0, 1234, 12, ptr(y), 0, 1.0, None, None, stream_ptr())
    torch.cuda.synchronize()
    assert torch.equal(y[0], x[2])
    call("idb_perturb", ptr(x0[:1]), None, 1, K, D, 0, 0, 1.0, 1.0, 1235, 12, ptr(y), 0, 1.0, None, None, stream_ptr())
    torch.cuda.synchronize()
    assert not torch.equal(y[0], x[2])


def test_anchor_linear_matches_fp64():
    A, N, K = 70, 130, 97
    x = _rand(A, K, seed=9)
    w = _rand(N, K + 1, seed=10)
    b = _rand(N, seed=11)
    out = torch.zeros(A, 256, device=dev())
    call("idb_anchor_linear", ptr(x), K, 0, 0.75, ptr(w), K + 1, ptr(b), w.data_ptr() + 4 * K, K + 1, 0.00999,
         ptr(out), 256, A, N, K, stream_ptr())
    torch.cuda.synchronize()
    ref = (0.75 * x.double()) @ w[:, :K].double().T + w[:, K].double() * 0.00999 + b.double()
    assert (out[:, :N].double() - ref).abs().max().item() < 1e-5 * ref.abs().max().item()
    assert out[:, N:].abs().max().item() == 0
    out2 = torch.zeros(A, 128, device=dev())
    w2 = _rand(100, N, seed=12)
    call("idb_anchor_linear", ptr(out), 256, 1, 1.0, ptr(w2), N, None, None, 0, 0.0, ptr(out2), 128, A, 100, N,
         stream_ptr())
    torch.cuda.synchronize()
    ref2 = _elu(out[:, :N].double()) @ w2.double().T
    assert (out2[:, :100].double() - ref2).abs().max().item() < 1e-5 * ref2.abs().max().item()


@pytest.mark.parametrize("A,K,D,ld,R", [(2, 151, 24, 24, 151), (3, 1501, 200, 256, 1536), (1, 400, 100, 100, 400)])
def test_center_gram_fp64(A, K, D, ld, R):
    X = _rand(A, R, ld, seed=13) + 5.0
    mean = torch.empty(A, D, dtype=torch.float64, device=dev())
    G = torch.empty(A, D, D, dtype=torch.float64, device=dev())
    call("idb_center_gram", ptr(X), R * ld, ld, A, K, D, 100.0, ptr(mean), ptr(G), stream_ptr())
    torch.cuda.synchronize()
    Xv = X[:, :K, :D].double()
    mref = Xv.mean(1)
    C = (Xv - mref[:, None, :]) * 100.0
    Gref = C.transpose(1, 2) @ C
    assert (mean - mref).abs().max().item() < 1e-12
    assert (G - Gref).abs().max().item() < 1e-11 * Gref.abs().max().item()


def _cliff_matrix(K, D, r, gap, seed):
    g = torch.Generator().manual_seed(seed)
    U, _ = torch.linalg.qr(torch.randn(K, D, generator=g, dtype=torch.float64))
    V, _ = torch.linalg.qr(torch.randn(D, D, generator=g, dtype=torch.float64))
    s = torch.cat([torch.linspace(6.4e3, 4.6e3, D - r, dtype=torch.float64),
                   gap * torch.linspace(5.0, 1.0, r, dtype=torch.float64)])
    return (U * s) @ V.T, s


@pytest.mark.parametrize("method", [_lib.SPECTRUM_JACOBI, _lib.SPECTRUM_TRIDIAG])
@pytest.mark.parametrize("K,D,r,gap", [(151, 24, 5, 1.0), (400, 100, 11, 1.0), (1501, 200, 51, 1.0),
                                       (1501, 201, 7, 1e-2), (700, 300, 30, 1.0), (8, 3, 1, 1.0), (64, 33, 2, 1e-1)])
def test_spectrum_from_gram_matches_fp64_svd(K, D, r, gap, method):
    A = 3
    mats, specs = zip(*[_cliff_matrix(K, D, r, gap, 20 + i) for i in range(A)])
    C = torch.stack(mats).to(dev())
    G = C.transpose(1, 2) @ C
    sref = torch.linalg.svdvals(C)
    ws = torch.empty(_lib.load().idb_spectrum_workspace_bytes(A, D), dtype=torch.uint8, device=dev())
    sv = torch.empty(A, D, device=dev())
    sweeps = torch.empty(A, dtype=torch.int32, device=dev())
    call("idb_spectrum", ptr(G), A, D, D, method, ptr(sv), ptr(sweeps), ptr(ws), stream_ptr())
    torch.cuda.synchronize()
    rel = ((sv.double() - sref).abs() / sref).max().item()
    # fp32 Jacobi on the Cholesky factor: 1e-4; fp64 tridiagonal + bisection: fp32 output rounding plus
    # n eps64 (s_max/s_i)^2 (the 1e5 cliff of the gap=1e-2 case costs 1e-16 * 200 * 1e10 = 2e-4 on its tail)
    bound = 1e-4 if method == _lib.SPECTRUM_JACOBI else (1e-6 if gap >= 1e-1 else 1e-3)
    assert rel < bound, (rel, sweeps.tolist())
    assert sweeps.max().item() < (40 if method == _lib.SPECTRUM_JACOBI else 200)
    assert torch.all(sv[:, :-1] >= sv[:, 1:])


@pytest.mark.parametrize("method", [_lib.SPECTRUM_JACOBI, _lib.SPECTRUM_TRIDIAG])
def test_spectrum_rank_deficient_and_zero(method):
    A, K, D = 2, 50, 16
    g = torch.Generator().manual_seed(3)
    C = torch.randn(A, K, D, generator=g, dtype=torch.float64)
    C[0, :, 8:] = C[0, :, :8]  # rank 8
    C[1] = 0
    C = C.to(dev())
    G = C.transpose(1, 2) @ C
    ws = torch.empty(_lib.load().idb_spectrum_workspace_bytes(A, D), dtype=torch.uint8, device=dev())
    sv = torch.empty(A, D, device=dev())
    call("idb_spectrum", ptr(G), A, D, D, method, ptr(sv), None, ptr(ws), stream_ptr())
    torch.cuda.synchronize()
    sref = torch.linalg.svdvals(C)
    assert (sv.double() - sref).abs().max().item() < 1e-6 * sref.max().item()
    assert sv[1].abs().max().item() == 0
    dims = torch.empty(A, dtype=torch.int32, device=dev())
    call("idb_id_from_spectrum", ptr(sv), A, D, D, ptr(dims), stream_ptr())
    torch.cuda.synchronize()
    assert dims[1].item() == -1  # s[1] == s[2]: the reference raises ZeroDivisionError


def test_id_readout_matches_oracle_on_golden_spectra():
    from oracle import fixtures as FX
    from oracle import id_oracle as O
    g = FX.golden("readout")
    for s, n, d in zip(g["spectra"], g["lens"], g["dims"]):
        s32 = torch.tensor(s[:n], dtype=torch.float32)
        sv = s32[None, :].contiguous().to(dev())
        dims = torch.empty(1, dtype=torch.int32, device=dev())
        call("idb_id_from_spectrum", ptr(sv), 1, int(n), int(n), ptr(dims), stream_ptr())
        torch.cuda.synchronize()
        try:
            want = O.id_from_spectrum(s32.double().tolist())
        except ZeroDivisionError:
            want = -1
        assert dims.item() == want
        if np.array_equal(np.float32(s[:n]).astype(np.float64), s[:n]):
            assert want == int(d) or want == -1


def test_ksphere_score_matches_oracle():
    from oracle import id_oracle as O
    rows, D, p = 64, 100, 10
    g = torch.Generator().manual_seed(5)
    x = torch.randn(rows, D, generator=g)
    x = x / x.norm(dim=1, keepdim=True) + 0.01 * torch.randn(rows, D, generator=g)
    for sigma in [0.01, 0.05, 0.5, 2.0]:
        ref = O.ksphere_gt_forward(x, torch.full((rows,), sigma), p)
        out = torch.empty(rows, D, device=dev())
        call("idb_ksphere_gt_score", ptr(x.to(dev())), None, sigma, rows, D, p, ptr(out), stream_ptr())
        torch.cuda.synchronize()
        err = (out.cpu().double() - ref.double()).abs().max().item()
        assert err < 2e-4 * ref.abs().max().item(), (sigma, err, ref.abs().max().item())


@pytest.mark.parametrize("fmt", [_lib.FMT_BF16, _lib.FMT_FP16])
def test_anchor_chain_tensor_core_matches_fp32_and_fp64(fmt):
    from iddiff_b200.engine import FcnEngine
    D, H, L, A = 100, 320, 2, 37
    g = torch.Generator().manual_seed(21)
    dims = [(H, D + 1)] + [(H, H)] * L + [(D, H)]
    ws = [((torch.rand(o, i, generator=g) * 2 - 1) / math.sqrt(i)).to(dev()) for o, i in dims]
    bs = [((torch.rand(o, generator=g) * 2 - 1) * 0.5).to(dev()) for o, _ in dims]
    eng = FcnEngine(ws, bs, fmt=fmt)
    x0 = _rand(A, D, seed=22)
    mc, label = 0.9995, 0.999
    P = eng.anchor_chain(x0, mc, label)
    Q = eng.anchor_chain_fp32(x0, mc, label)
    torch.cuda.synchronize()
    h = torch.cat([(torch.tensor(mc, device=dev()) * x0).double(), torch.full((A, 1), label, dtype=torch.float64, device=dev())], 1)
    for l in range(L + 2):
        ref = h @ ws[l].double().T + bs[l].double()
        n = ref.shape[1]
        tol = (2e-5 if fmt == _lib.FMT_FP16 else 1e-4) * ref.abs().max().item()
        assert (P[l][:A, :n].double() - ref).abs().max().item() < tol
        assert (Q[l][:A, :n].double() - ref).abs().max().item() < 2e-5 * ref.abs().max().item()
        assert P[l][:A, n:].abs().max().item() == 0 if P[l].shape[1] > n else True
        h = _elu(ref)


def test_bad_arguments_are_rejected_loudly():
    with pytest.raises(_lib.IdbError):
        call("idb_fcn_output_layer", None, None, None, None, 0, 1.0, 128, 256, 64, None, None)
    x = torch.zeros(4, 4)
    with pytest.raises(_lib.IdbError):
        ptr(x)  # CPU tensor: no CPU path
    a = torch.zeros(128 * 64 * 2, dtype=torch.uint8, device=dev())
    o = torch.zeros(100, 256, device=dev())
    with pytest.raises(_lib.IdbError):
        call("idb_fcn_output_layer", ptr(a), ptr(a), ptr(a), ptr(a), 0, 1.0, 100, 256, 64, ptr(o), stream_ptr())


def test_partial_grams_about_the_global_mean_add_up_to_the_centred_gram():
    """idb_column_sums + idb_gram_about (the K-sharded mode's per-rank pieces) on two row shards of one matrix: the
    partial Grams about the all-reduced mean sum to idb_center_gram of the whole matrix, and both match fp64."""
    from iddiff_b200.dim_reduction import _DeviceGramOps as Ops
    K, D = 700, 300
    X = (torch.randn(K, D, generator=torch.Generator().manual_seed(5)) * 0.01 + 1.5).cuda()
    halves = (X[:333].contiguous(), X[333:].contiguous())
    mean = sum(Ops.column_sums(h) for h in halves) / K
    G = sum(Ops.gram_about(h, mean, -100.0) for h in halves)
    whole = torch.empty(D, D, dtype=torch.float64, device="cuda")
    cm = torch.empty(D, dtype=torch.float64, device="cuda")
    _lib.call("idb_center_gram", X.data_ptr(), 0, D, 1, K, D, -100.0, cm.data_ptr(), whole.data_ptr(),
              torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    Y = (X.double() - X.double().mean(0, keepdim=True)) * -100.0
    want = Y.T @ Y
    torch.testing.assert_close(mean, X.double().mean(0), rtol=1e-13, atol=0)
    torch.testing.assert_close(whole, want, rtol=1e-11, atol=1e-11 * float(want.abs().max()))
    torch.testing.assert_close(G, want, rtol=1e-11, atol=1e-11 * float(want.abs().max()))
