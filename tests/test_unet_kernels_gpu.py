"""GPU numerics of the U-Net kernels (through the C ABI) against plain PyTorch fp64 references of the same ops.
Tolerances: the split-precision convolution carries 22-bit operands and fp32 accumulation, so it is compared with the
fp64 result at 2e-6 of the output scale (the fp32 cuDNN/ATen result of the reference has the same order of error)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def rel_err(a, b):
    return float((a.double().cpu() - b.double().cpu()).abs().max() / b.double().abs().max())


@pytest.mark.parametrize("n,c,h,w,groups", [(3, 32, 8, 8, 32), (2, 128, 16, 16, 32), (5, 96, 4, 4, 32), (2, 384, 8, 8, 32)])
def test_groupnorm_stats(n, c, h, w, groups):
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(0)
    x = torch.randn(n, c, h, w) * 3 + 1.5
    c0 = c // 4 * 3
    mr = U.gn_stats(nhwc(x[:, :c0]).cuda(), nhwc(x[:, c0:]).cuda(), groups, 1e-6).cpu()
    xd = x.double().reshape(n, groups, -1)
    mean, var = xd.mean(-1), xd.var(-1, unbiased=False)
    np.testing.assert_allclose(mr[..., 0].numpy(), mean.numpy(), rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(mr[..., 1].numpy(), (1 / torch.sqrt(var + 1e-6)).numpy(), rtol=1e-6)


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_prep_modes_f32(mode):
    """GroupNorm + SiLU + resampling of a concatenated pair, fp32 output and the hi+lo operand."""
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(1)
    n, c0, c1, h, w = 2, 40, 16, 8, 8
    x = torch.randn(n, c0 + c1, h, w)
    gamma, beta = torch.randn(c0 + c1), torch.randn(c0 + c1)
    a, b = nhwc(x[:, :c0]).cuda(), nhwc(x[:, c0:]).cuda()
    mr = U.gn_stats(a, b, 8, 1e-6)
    op, f32 = U.prep(a, b, mode=mode, mean_rstd=mr, groups=8, gamma=gamma.cuda(), beta=beta.cuda(), act=1, want_f32=True)
    ref = F.silu(F.group_norm(x.double(), 8, gamma.double(), beta.double(), eps=1e-6))
    if mode == 1:
        ref = F.interpolate(ref, scale_factor=2, mode="nearest")
    elif mode == 3:
        ref = F.avg_pool2d(ref, 2)
    if mode == 2:
        refp = F.pad(ref, (0, 1, 0, 1))
        planes = [nhwc(refp[:, :, py::2, px::2][:, :, :h // 2, :w // 2]) for py in (0, 1) for px in (0, 1)]
        ref_nhwc = torch.stack(planes)
    else:
        ref_nhwc = nhwc(ref)[None]
    got = f32 if mode == 2 else f32[None]
    np.testing.assert_allclose(got.cpu().numpy(), ref_nhwc.numpy(), rtol=2e-5, atol=2e-6)
    rec = (op.hi.double() + op.lo.double()) / U.ACT_SCALE
    assert rec.shape[-1] == 64 and float(rec[..., c0 + c1:].abs().max()) == 0.0
    np.testing.assert_allclose(rec[..., :c0 + c1].cpu().numpy(), got.double().cpu().numpy(), rtol=1e-6, atol=1e-8)


@pytest.mark.parametrize("n,cin,cout,h,w", [(2, 3, 128, 16, 16), (3, 128, 128, 16, 16), (5, 192, 256, 8, 8),
                                            (9, 64, 96, 4, 4), (1, 320, 3, 32, 32), (2, 128, 384, 64, 64)])
def test_conv3x3_vs_torch_fp64(n, cin, cout, h, w):
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(2)
    x = torch.randn(n, cin, h, w)
    wt = torch.randn(cout, cin, 3, 3) / np.sqrt(9 * cin)
    bias = torch.randn(cout)
    res = torch.randn(n, cout, h, w)
    pc = U.pack_conv(wt.cuda(), bias.cuda())
    op = U.prep(nhwc(x).cuda())
    out = U.conv(op, pc, residual=nhwc(res).cuda(), out_scale=0.5)
    ref = nhwc((F.conv2d(x.double(), wt.double(), bias.double(), padding=1) + res.double()) * 0.5)
    assert rel_err(out, ref) < 1e-5  # fp32 accumulation over up to 9*Cin products


def test_conv3x3_stride2_ddpm_downsample():
    """layers.py:607-629: F.pad(x, (0,1,0,1)) then 3x3 stride-2 conv, padding 0."""
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(3)
    n, c, h, w = 3, 128, 16, 16
    x, wt, bias = torch.randn(n, c, h, w), torch.randn(c, c, 3, 3) / 30, torch.randn(c)
    pc = U.pack_conv(wt.cuda(), bias.cuda(), taps=U.TAPS_3x3_S2, planes=4)
    out = U.conv(U.prep(nhwc(x).cuda(), mode=U.PREP_STRIDE2), pc)
    ref = nhwc(F.conv2d(F.pad(x.double(), (0, 1, 0, 1)), wt.double(), bias.double(), stride=2))
    assert tuple(out.shape) == (n, h // 2, w // 2, c)
    assert rel_err(out, ref) < 1e-5  # fp32 accumulation over up to 9*Cin products


def test_nin_1x1_and_nchw_input():
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(4)
    n, cin, cout, h, w = 4, 3, 64, 8, 8
    x, wt = torch.rand(n, cin, h, w), torch.randn(cout, cin)
    pc = U.pack_conv(wt.cuda(), None)
    two, m1 = torch.full((cin,), 2.0).cuda(), torch.full((cin,), -1.0).cuda()
    out = U.conv(U.prep(x.cuda(), nchw=True, gamma=two, beta=m1), pc)   # 2x - 1 of ddpm.py:168 folded into the prep
    ref = torch.einsum("nchw,oc->nhwo", 2 * x.double() - 1, wt.double())
    assert rel_err(out, ref) < 1e-5  # fp32 accumulation over up to 9*Cin products


@pytest.mark.parametrize("n,t_hw,c,heads", [(3, 8, 64, 1), (2, 16, 256, 1), (2, 16, 128, 4), (1, 4, 32, 2)])
def test_attention_vs_torch(n, t_hw, c, heads):
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(5)
    T, dh = t_hw * t_hw, c // heads
    qkv = torch.randn(n, T, 3 * c)
    # layout A (DDPM): q | k | v blocks of C channels, head h at h*dh inside each block
    out = U.attention(qkv.reshape(n, t_hw, t_hw, 3 * c).cuda(), heads, dh, 0, c, 2 * c, dh, dh ** -0.5)
    q, k, v = (qkv[..., i * c:(i + 1) * c].double().reshape(n, T, heads, dh).transpose(1, 2) for i in range(3))
    ref = (torch.softmax(q @ k.transpose(-1, -2) * dh ** -0.5, -1) @ v).transpose(1, 2).reshape(n, T, c)
    np.testing.assert_allclose(out.reshape(n, T, c).cpu().numpy(), ref.numpy(), rtol=1e-4, atol=2e-6)


def test_dense_silu():
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(6)
    x, w, b = torch.randn(1, 512), torch.randn(256, 512) / 20, torch.randn(256)
    out = U.dense(x.cuda(), w.cuda(), b.cuda(), act_in=2)
    ref = F.linear(F.silu(x.double()), w.double(), b.double())
    np.testing.assert_allclose(out.cpu().numpy(), ref.numpy(), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("n,cin,cout,h,w,groups", [(3, 64, 128, 64, 64, 32),    # BI = 1, 32 tiles per image, 128 slots
                                                   (2, 128, 256, 16, 16, 32),   # two tiles per image
                                                   (5, 64, 384, 8, 8, 32),      # 2 images per tile, ragged last tile, 12 channels per group
                                                   (11, 128, 128, 4, 4, 32),    # 16-pixel images: half-warp reduction, ragged
                                                   (2, 64, 96, 32, 32, 8)])     # Cout < Np: partial 32-channel chunk
def test_groupnorm_statistics_from_the_conv_epilogue(n, cin, cout, h, w, groups):
    """Row g1: the convolution's epilogue emits the per-(image, slot, channel) sums the NEXT GroupNorm needs; folded by
    idb_groupnorm_from_conv_stats they equal the statistics of the written tensor (the re-reading kernel and fp64 torch),
    alone and as the second half of a virtual concat; the result is bit-reproducible."""
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(n * 100 + cout)
    x = torch.randn(n, cin, h, w)
    wt = torch.randn(cout, cin, 3, 3) / np.sqrt(9 * cin)
    bias = torch.randn(cout)
    res = nhwc(torch.randn(n, cout, h, w)).cuda()
    pc = U.pack_conv(wt.cuda(), bias.cuda())
    a = U.prep(nhwc(x).cuda())
    y = U.conv(a, pc, residual=res, out_scale=0.7, want_stats=True)
    assert hasattr(y, "_idb_stats") and y._idb_stats.shape == (n, max(1, h * w // 32), cout, 2)
    y_plain = U.conv(a, pc, residual=res, out_scale=0.7)
    assert torch.equal(y, y_plain) and not hasattr(y_plain, "_idb_stats")
    mr_fused = U.gn_stats(y, None, groups, 1e-6)
    mr_reread = U.gn_stats(y_plain, None, groups, 1e-6)
    yd = y.double().permute(0, 3, 1, 2).reshape(n, groups, -1).cpu()
    mean, rstd = yd.mean(-1), 1 / torch.sqrt(yd.var(-1, unbiased=False) + 1e-6)
    np.testing.assert_allclose(mr_fused[..., 0].cpu().numpy(), mean.numpy(), rtol=2e-6, atol=2e-6)
    np.testing.assert_allclose(mr_fused[..., 1].cpu().numpy(), rstd.numpy(), rtol=2e-6)
    np.testing.assert_allclose(mr_fused.cpu().numpy(), mr_reread.cpu().numpy(), rtol=2e-6, atol=2e-6)
    # virtual concat [other | y]: other carries no partial sums -> falls back to the re-reading kernel; both carry -> fused
    y2 = U.conv(a, pc, want_stats=True)
    cat_fused = U.gn_stats(y2, y, groups, 1e-6)
    cd = torch.cat([y2, y], dim=3).double().permute(0, 3, 1, 2).reshape(n, groups, -1).cpu()
    np.testing.assert_allclose(cat_fused[..., 0].cpu().numpy(), cd.mean(-1).numpy(), rtol=2e-6, atol=2e-6)
    np.testing.assert_allclose(cat_fused[..., 1].cpu().numpy(), (1 / torch.sqrt(cd.var(-1, unbiased=False) + 1e-6)).numpy(), rtol=2e-6)
    again = U.conv(a, pc, residual=res, out_scale=0.7, want_stats=True)
    assert torch.equal(again._idb_stats, y._idb_stats)


@pytest.mark.parametrize("kind", ["down", "up", "pyramid"])
@pytest.mark.parametrize("norm", [False, True])
def test_fir_folded_into_operand_preparation(kind, norm):
    """Row g1: idb_conv_prep_fir = upfirdn2d(act(GN(x))) gathered per output pixel, against (a) the oracle's numpy
    upfirdn2d of the torch fp64 transform and (b) the unfused kernel sequence it replaces (prep -> idb_upfirdn2d -> prep):
    layerspp.py:244-257 down (1, 2, pad 1 1) / up (kernel x 4, 2, 1, pad 2 1), and the input pyramid's pad-only FIR
    (up_or_down_sampling.py:144-178) emitted as the four stride-2 parity planes."""
    from iddiff_b200 import unet_ops as U
    from oracle import img_oracle as IO
    torch.manual_seed(7)
    n, c, h, w, groups = 3, 48, 16, 16, 12
    x = torch.randn(n, c, h, w) * 2 + 0.5
    gamma, beta = torch.randn(c), torch.randn(c)
    xg = nhwc(x).cuda()
    gain = 4.0 if kind == "up" else 1.0
    kh = U.fir_kernel_host([1, 3, 3, 1], gain)
    kd = U.fir_kernel([1, 3, 3, 1], gain, "cuda")
    up, down, p0, p1 = {"down": (1, 2, 1, 1), "up": (2, 1, 2, 1), "pyramid": (1, 1, 2, 2)}[kind]
    kw = {}
    t = x.double()
    if norm:
        mr = U.gn_stats(xg, None, groups, 1e-6)
        kw = dict(mean_rstd=mr, groups=groups, gamma=gamma.cuda(), beta=beta.cuda(), act=1)
        t = F.silu(F.group_norm(t, groups, gamma.double(), beta.double(), eps=1e-6))
    op, f32, raw = U.prep_fir(xg, kh, up, down, p0, p1, as_planes=kind == "pyramid", want_f32=True, want_raw=True, **kw)
    # the skip branch from the same gather: the filter on the untransformed tensor
    plain = U.prep_fir(xg, kh, up, down, p0, p1, as_planes=kind == "pyramid", want_operand=False, want_f32=True)[1]
    assert torch.equal(raw, plain)
    if not norm:
        assert torch.equal(raw, f32)
    # (a) oracle: numpy upfirdn2d on [major = N*C, H, W, minor = 1]
    ref = IO.upfirdn2d_np(t.numpy().reshape(n * c, h, w, 1), kh.astype(np.float64), up, up, down, down, p0, p1, p0, p1)
    ref = torch.from_numpy(ref.reshape(n, c, ref.shape[1], ref.shape[2]))
    if kind == "pyramid":
        refp = F.pad(ref, (0, 1, 0, 1))
        ho = (ref.shape[2] + 1) // 2
        ref_nhwc = torch.stack([nhwc(refp[:, :, py::2, px::2][:, :, :ho, :ho]) for py in (0, 1) for px in (0, 1)])
        got = f32
    else:
        ref_nhwc, got = nhwc(ref)[None], f32[None]
    assert got.shape == ref_nhwc.shape
    np.testing.assert_allclose(got.cpu().numpy(), ref_nhwc.numpy(), rtol=2e-5, atol=2e-6)
    rec = (op.hi.double() + op.lo.double()) / U.ACT_SCALE
    np.testing.assert_allclose(rec[..., :c].cpu().numpy(), got.double().cpu().numpy(), rtol=1e-6, atol=1e-8)
    assert float(rec[..., c:].abs().max()) == 0.0
    # (b) the unfused sequence
    tf = U.prep(xg, want_operand=False, want_f32=True, **kw)[1]
    f = U.upfirdn2d_nhwc(tf, kd, up, down, p0, p1)
    if kind == "pyramid":
        old = U.prep(f, mode=U.PREP_STRIDE2, want_f32=True)[1]
    else:
        old = f[None]
    np.testing.assert_allclose(got.cpu().numpy(), old.cpu().numpy(), rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("n,hh,ww,c", [(3, 16, 16, 256), (2, 8, 16, 128), (5, 16, 16, 384)])
def test_tensor_core_attention_vs_cuda_core_and_fp64(n, hh, ww, c):
    """The 16x16 attention blocks of the 64x64 networks (layers.py:577-587: one head, 256 tokens x 256 channels) on the
    tensor cores -- Q K^T and P V as per-image-weight 1x1 convolutions on conv_fp16x3_kernel -- against the fp32 CUDA-core
    kernel and torch fp64."""
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(8)
    T = hh * ww
    qkv = (torch.randn(n, T, 3 * c) * 1.5).reshape(n, hh, ww, 3 * c).cuda()
    assert U.TC_ATTENTION
    got = U.attention(qkv, 1, c, 0, c, 2 * c, c, float(c) ** -0.5)
    U.TC_ATTENTION = False
    try:
        old = U.attention(qkv, 1, c, 0, c, 2 * c, c, float(c) ** -0.5)
    finally:
        U.TC_ATTENTION = True
    q, k, v = (qkv.reshape(n, T, 3 * c)[..., i * c:(i + 1) * c].double().cpu() for i in range(3))
    ref = torch.softmax(q @ k.transpose(-1, -2) * float(c) ** -0.5, -1) @ v
    sc = float(ref.abs().max())
    e_tc, e_cc = float((got.reshape(n, T, c).double().cpu() - ref).abs().max()), float((old.reshape(n, T, c).double().cpu() - ref).abs().max())
    print(f"T={T} C={c}: tensor-core error {e_tc / sc:.2e}, CUDA-core error {e_cc / sc:.2e} of the output scale")
    # the scores carry the 22-bit operand rounding times |S| ~ 10 into the softmax exponent: ~4e-6 of the output scale
    assert e_tc < 1e-5 * sc and e_cc < 3e-6 * sc


@pytest.mark.parametrize("n,cin,cout,h,w", [(20, 64, 128, 64, 64),      # 640 pixel tiles -> paired (256 x 128 CTA tiles)
                                            (1185, 64, 128, 8, 8),      # odd tile count (593): the last CTA tile has a phantom half
                                            (200, 128, 384, 16, 16)])   # Np = 384: three channel tiles per pixel-tile pair
def test_paired_pixel_tiles_share_one_weight_tile(n, cin, cout, h, w):
    """conv_fp16x3_kernel<128, 2>: the Cout = 128 / 384 layers run two 128-pixel sub-tiles per CTA on one weight tile.
    Against torch fp64, with bias + residual + GroupNorm statistics, bit-identical to the unpaired kernel (forced by a
    batch small enough to fall under the pairing threshold)."""
    from iddiff_b200 import unet_ops as U
    torch.manual_seed(n + cout)
    x = torch.randn(n, cin, h, w)
    wt = torch.randn(cout, cin, 3, 3) / np.sqrt(9 * cin)
    bias = torch.randn(cout)
    res = nhwc(torch.randn(n, cout, h, w)).cuda()
    pc = U.pack_conv(wt.cuda(), bias.cuda())
    a = U.prep(nhwc(x).cuda())
    y = U.conv(a, pc, residual=res, out_scale=0.5, want_stats=True)
    ref = 0.5 * (F.conv2d(x.double(), wt.double(), bias.double(), padding=1) + res.cpu().double().permute(0, 3, 1, 2))
    assert rel_err(y.permute(0, 3, 1, 2), ref) < 1e-5
    # the same images in small batches take the unpaired kernel: identical bits
    step = max(1, n // 16)
    for s in range(0, n, step):
        e = min(n, s + step)
        part = U.conv(U.prep(nhwc(x[s:e]).cuda()), pc, residual=res[s:e].contiguous(), out_scale=0.5, want_stats=True)
        assert torch.equal(part, y[s:e])
        assert torch.equal(part._idb_stats, y._idb_stats[s:e])
    mr = U.gn_stats(y, None, 32, 1e-6)
    yd = y.double().permute(0, 3, 1, 2).reshape(n, 32, -1).cpu()
    np.testing.assert_allclose(mr[..., 0].cpu().numpy(), yd.mean(-1).numpy(), rtol=2e-6, atol=2e-6)
