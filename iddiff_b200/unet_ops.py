"""Host-side handles of the U-Net kernels (C ABI: idb_groupnorm_stats, idb_conv_prep, idb_conv2d, idb_attention).

Activations are fp32 NHWC torch tensors on the device; torch only owns the memory and the stream.  Weight packing
(once per model) turns an nn.Conv2d / NIN / nn.Linear weight into the hi/lo fp16 slab operand of the tensor-core
kernel.  No CPU path: every function raises IdbError on a CPU tensor.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import torch

from . import _lib

ACT_SCALE = 64.0  # activations are multiplied by 2^6 before the fp16 hi/lo split (|v| < 1023 stays finite)
PREP_SAME, PREP_UP2, PREP_STRIDE2, PREP_AVGPOOL2 = 0, 1, 2, 3

TAPS_3x3 = [(0, ky - 1, kx - 1) for ky in range(3) for kx in range(3)]
TAPS_3x3_S2 = [(2 * (ky & 1) + (kx & 1), ky >> 1, kx >> 1) for ky in range(3) for kx in range(3)]
TAPS_1x1 = [(0, 0, 0)]


def _ceil(a: int, b: int) -> int:
    return (a + b - 1) // b * b


@dataclass
class PackedConv:
    """One convolution's weights in tensor-core operand form."""
    w_hi: torch.Tensor
    w_lo: torch.Tensor
    bias: Optional[torch.Tensor]
    cin: int
    cp: int
    cout: int
    np_: int
    taps: Sequence[Tuple[int, int, int]]
    w_scale: float
    planes: int

    @property
    def taps_c(self):
        flat = [v for t in self.taps for v in t]
        return (C.c_int * len(flat))(*flat)


def pack_conv(weight: torch.Tensor, bias: Optional[torch.Tensor], taps=None, planes: int = 1) -> PackedConv:
    """weight: nn.Conv2d layout [Cout][Cin][kh][kw] (or [Cout][Cin] for 1x1 / Linear / NIN^T), fp32, on the device.
    The GEMM B operand is the matrix [Cout][n_taps * Cp], tap-major (ky, kx row-major), channel-minor, zero padded."""
    if not weight.is_cuda:
        raise _lib.IdbError("pack_conv: CUDA tensor required")
    w = weight.detach().float()
    if w.dim() == 2:
        w = w[:, :, None, None]
    cout, cin, kh, kw = w.shape
    n_taps = kh * kw
    if taps is None:
        taps = TAPS_3x3 if n_taps == 9 else TAPS_1x1
    assert len(taps) == n_taps
    cp = _ceil(cin, 64)
    np_ = _ceil(cout, 128)
    mat = torch.zeros(cout, n_taps, cp, device=w.device, dtype=torch.float32)
    mat[:, :, :cin] = w.permute(0, 2, 3, 1).reshape(cout, n_taps, cin)
    mat = mat.reshape(cout, n_taps * cp).contiguous()
    amax = float(mat.abs().max())
    # power-of-two scale with max|w| * scale in [2^13, 2^14)
    w_scale = 2.0 ** (13 - math.floor(math.log2(amax))) if amax > 0 else 1.0
    kc = n_taps * cp
    hi = torch.empty(kc // 64 * np_ * 128, dtype=torch.uint8, device=w.device)
    lo = torch.empty_like(hi)
    _lib.call("idb_pack_weight", _lib.ptr(mat), cout, kc, kc, np_, kc, _lib.FMT_FP16, w_scale, _lib.ptr(hi), _lib.ptr(lo),
              _lib.stream_ptr())
    b = bias.detach().float().contiguous() if bias is not None else None
    return PackedConv(hi, lo, b, cin, cp, cout, np_, list(taps), w_scale, planes)


SILU_ACT = 1            # 1: ex2.approx + approximate divide (|error| ~3e-7 relative); 2: expf + IEEE divide
FUSED_GN_STATS = True   # GroupNorm statistics from the producing convolution's epilogue (False: re-read the tensor)


def gn_stats(x0: torch.Tensor, x1: Optional[torch.Tensor], groups: int, eps: float) -> torch.Tensor:
    """x0 [N,H,W,C0] (+ x1 [N,H,W,C1] concatenated along C) -> mean_rstd [N, groups, 2] fp32.  When every input
    carries the partial sums its convolution wrote (conv(..., want_stats=True)), they are folded by one small kernel;
    otherwise the tensor is read once more (idb_groupnorm_stats)."""
    n, h, w, c0 = x0.shape
    c1 = x1.shape[3] if x1 is not None else 0
    st0 = getattr(x0, "_idb_stats", None)
    st1 = getattr(x1, "_idb_stats", None) if x1 is not None else None
    if st0 is not None and (x1 is None or st1 is not None):
        mr = torch.empty(n, groups, 2, dtype=torch.float32, device=x0.device)
        _lib.call("idb_groupnorm_from_conv_stats", _lib.ptr(st0), c0, st0.shape[1], _lib.ptr(st1), c1,
                  st1.shape[1] if st1 is not None else 0, n, h * w, groups, float(eps), _lib.ptr(mr), _lib.stream_ptr())
        return mr
    sums = torch.empty(n, groups, 2, dtype=torch.float64, device=x0.device)
    mr = torch.empty(n, groups, 2, dtype=torch.float32, device=x0.device)
    _lib.call("idb_groupnorm_stats", _lib.ptr(x0), c0, _lib.ptr(x1), c1, n, h * w, groups, float(eps), _lib.ptr(sums),
              _lib.ptr(mr), _lib.stream_ptr())
    return mr


@dataclass
class Operand:
    hi: torch.Tensor
    lo: torch.Tensor
    planes: int
    n: int
    h: int
    w: int
    cp: int


def prep(x0: torch.Tensor, x1: Optional[torch.Tensor] = None, *, mode: int = PREP_SAME, mean_rstd=None, groups: int = 0,
         gamma=None, beta=None, act: int = 0, nchw: bool = False, want_operand: bool = True, want_f32: bool = False):
    """One pass: concat -> GroupNorm apply -> SiLU -> resample -> fp16 hi/lo operand (and / or fp32 NHWC copy)."""
    if nchw:
        n, c0, hs, ws = x0.shape
    else:
        n, hs, ws, c0 = x0.shape
    c1 = x1.shape[3] if x1 is not None else 0
    c = c0 + c1
    planes = 4 if mode == PREP_STRIDE2 else 1
    ho = {PREP_SAME: hs, PREP_UP2: 2 * hs, PREP_STRIDE2: (hs + 1) // 2, PREP_AVGPOOL2: hs // 2}[mode]
    wo = {PREP_SAME: ws, PREP_UP2: 2 * ws, PREP_STRIDE2: (ws + 1) // 2, PREP_AVGPOOL2: ws // 2}[mode]
    cp = _ceil(c, 64)
    dev = x0.device
    hi = lo = None
    if want_operand:
        hi = torch.empty(planes, n, ho, wo, cp, dtype=torch.float16, device=dev)
        lo = torch.empty_like(hi)
    f32 = torch.empty(planes, n, ho, wo, c, dtype=torch.float32, device=dev) if want_f32 else None
    _lib.call("idb_conv_prep", _lib.ptr(x0), c0, _lib.ptr(x1), c1, n, hs, ws, int(nchw), mode, _lib.ptr(mean_rstd),
              groups, _lib.ptr(gamma), _lib.ptr(beta), act and SILU_ACT, ACT_SCALE, _lib.ptr(hi), _lib.ptr(lo), cp, _lib.ptr(f32),
              _lib.stream_ptr())
    op = Operand(hi, lo, planes, n, ho, wo, cp) if want_operand else None
    if want_f32:
        return op, (f32[0] if planes == 1 else f32)
    return op


FUSED_FIR = True   # FIR resampling folded into the operand-preparation pass (False: stand-alone idb_upfirdn2d launches)


def fir_kernel_host(taps, gain: float):
    """up_or_down_sampling._setup_kernel (:165-172) * gain as a contiguous numpy float32 [kh][kw] (host)."""
    import numpy as np
    k = np.asarray(taps, dtype=np.float32)
    if k.ndim == 1:
        k = np.outer(k, k)
    k /= np.sum(k)
    return np.ascontiguousarray(k * np.float32(gain), dtype=np.float32)


def prep_fir(x0: torch.Tensor, k_host, up: int, down: int, pad0: int, pad1: int, *, as_planes: bool = False,
             mean_rstd=None, groups: int = 0, gamma=None, beta=None, act: int = 0, want_operand: bool = True,
             want_f32: bool = False, want_raw: bool = False):
    """upfirdn2d(act(GN(x0)), k, up, down, (pad0, pad1)) written straight as the convolution operand and / or fp32 NHWC
    (idb_conv_prep_fir); as_planes: the four stride-2 parity planes of the down = 1 result (conv_downsample_2d)."""
    n, hs, ws, c = x0.shape
    kh, kw = k_host.shape
    hf = (hs * up + pad0 + pad1 - kh) // down + 1
    wf = (ws * up + pad0 + pad1 - kw) // down + 1
    planes = 4 if as_planes else 1
    ho, wo = ((hf + 1) // 2, (wf + 1) // 2) if as_planes else (hf, wf)
    cp = _ceil(c, 64)
    dev = x0.device
    hi = lo = None
    if want_operand:
        hi = torch.empty(planes, n, ho, wo, cp, dtype=torch.float16, device=dev)
        lo = torch.empty_like(hi)
    f32 = torch.empty(planes, n, ho, wo, c, dtype=torch.float32, device=dev) if want_f32 else None
    raw = torch.empty(planes, n, ho, wo, c, dtype=torch.float32, device=dev) if want_raw else None
    kptr = k_host.ctypes.data_as(C.POINTER(C.c_float))
    _lib.call("idb_conv_prep_fir", _lib.ptr(x0), c, n, hs, ws, _lib.ptr(mean_rstd), groups, _lib.ptr(gamma), _lib.ptr(beta),
              act and SILU_ACT, kptr, kh, kw, up, down, pad0, pad1, int(as_planes), ACT_SCALE, _lib.ptr(hi), _lib.ptr(lo), cp,
              _lib.ptr(f32), _lib.ptr(raw), _lib.stream_ptr())
    op = Operand(hi, lo, planes, n, ho, wo, cp) if want_operand else None
    out = [op]
    if want_f32:
        out.append(f32[0] if planes == 1 else f32)
    if want_raw:
        out.append(raw[0] if planes == 1 else raw)
    return out[0] if len(out) == 1 else tuple(out)


def conv(a: Operand, pc: PackedConv, *, bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
         out_scale: float = 1.0, out: Optional[torch.Tensor] = None, out_hw=None, want_stats: bool = False) -> torch.Tensor:
    """out [N,H,W,Cout] fp32 = out_scale * (conv(a, W) + bias + residual).  `bias` overrides pc.bias (the host folds
    the time-embedding projection into it).  want_stats: the epilogue also writes the GroupNorm partial sums of `out`
    (attached as out._idb_stats [N, slots, Cout, 2]; gn_stats picks them up)."""
    if a.cp != pc.cp or a.planes != pc.planes:
        raise _lib.IdbError(f"conv: operand (Cp={a.cp}, planes={a.planes}) does not match weights (Cp={pc.cp}, planes={pc.planes})")
    oh, ow = out_hw if out_hw is not None else (a.h, a.w)
    if out is None:
        out = torch.empty(a.n, oh, ow, pc.cout, dtype=torch.float32, device=a.hi.device)
    b = bias if bias is not None else pc.bias
    ld_res = residual.shape[-1] if residual is not None else 0
    stats = None
    if want_stats and FUSED_GN_STATS:
        slots = _lib.load().idb_conv_stats_slots(oh, ow)
        if slots > 0:
            stats = torch.empty(a.n, slots, pc.cout, 2, dtype=torch.float32, device=a.hi.device)
    _lib.call("idb_conv2d_stats", _lib.ptr(a.hi), _lib.ptr(a.lo), a.planes, a.n, a.h, a.w, oh, ow, a.cp, _lib.ptr(pc.w_hi),
              _lib.ptr(pc.w_lo), len(pc.taps), pc.taps_c, pc.np_, pc.cout, 1.0 / (ACT_SCALE * pc.w_scale), float(out_scale),
              _lib.ptr(b), _lib.ptr(residual), ld_res, _lib.ptr(out), out.shape[-1], _lib.ptr(stats), _lib.stream_ptr())
    if stats is not None:
        out._idb_stats = stats
    return out


TC_ATTENTION = True   # one-head attention over >= 128 tokens on the tensor cores (False: the fp32 CUDA-core kernel)


def attention(qkv: torch.Tensor, heads: int, dh: int, q0: int, k0: int, v0: int, head_stride: int,
              scale: float) -> torch.Tensor:
    """qkv [N,H,W,ld] fp32 -> [N,H,W,heads*dh] fp32."""
    n, h, w, ld = qkv.shape
    out = torch.empty(n, h, w, heads * dh, dtype=torch.float32, device=qkv.device)
    t = h * w
    if TC_ATTENTION and heads == 1 and t >= 128 and t % 128 == 0 and t <= 1024 and dh % 128 == 0:
        ws = torch.empty(_lib.load().idb_attention_tc_workspace_bytes(n, t, dh) + 1024, dtype=torch.uint8, device=qkv.device)
        base = (ws.data_ptr() + 1023) // 1024 * 1024
        _lib.call("idb_attention_tc", _lib.ptr(qkv), n, h, w, ld, dh, q0, k0, v0, float(scale), base, _lib.ptr(out), dh,
                  _lib.stream_ptr())
        return out
    _lib.call("idb_attention", _lib.ptr(qkv), n, h * w, ld, heads, dh, q0, k0, v0, head_stride, float(scale),
              _lib.ptr(out), heads * dh, _lib.stream_ptr())
    return out


def dense(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], act_in: int = 0) -> torch.Tensor:
    """fp32 nn.Linear on a few rows (time-embedding path): out = W f(x) + b, f = identity (0) / ELU (1) / SiLU (2)."""
    a, k = x.shape
    n = weight.shape[0]
    out = torch.empty(a, n, dtype=torch.float32, device=x.device)
    _lib.call("idb_anchor_linear", _lib.ptr(x), k, act_in, 1.0, _lib.ptr(weight), weight.stride(0), _lib.ptr(bias), None,
              0, 0.0, _lib.ptr(out), n, a, n, k, _lib.stream_ptr())
    return out


def fir_kernel(taps, gain: float, device) -> torch.Tensor:
    """up_or_down_sampling._setup_kernel (:165-172) * gain, in numpy float32 like the reference."""
    import numpy as np
    k = np.asarray(taps, dtype=np.float32)
    if k.ndim == 1:
        k = np.outer(k, k)
    k /= np.sum(k)
    return torch.from_numpy(np.ascontiguousarray(k * gain)).to(device)


def upfirdn2d_nhwc(x: torch.Tensor, k: torch.Tensor, up: int, down: int, pad0: int, pad1: int) -> torch.Tensor:
    """The reference op on an NHWC activation: [major = N, H, W, minor = C] is exactly its native layout
    (op/upfirdn2d.cpp:12-22), so no NCHW reshape is needed."""
    n, h, w, c = x.shape
    kh, kw = k.shape
    oh = (h * up + pad0 + pad1 - kh) // down + 1
    ow = (w * up + pad0 + pad1 - kw) // down + 1
    out = torch.empty(n, oh, ow, c, dtype=torch.float32, device=x.device)
    _lib.call("idb_upfirdn2d", _lib.ptr(x), _lib.ptr(k), n, h, w, c, kh, kw, up, up, down, down, pad0, pad1, pad0, pad1,
              _lib.ptr(out), _lib.stream_ptr())
    return out
