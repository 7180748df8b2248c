"""Drop-in for the reference's `op` package (op/__init__.py:1-2): `upfirdn2d`, `fused_leaky_relu`,
`FusedLeakyReLU`, same call signatures, served by the sm_100a kernels behind the C ABI
(idb_upfirdn2d / idb_fused_bias_act).  CUDA tensors only: the reference's CPU branches
(op/upfirdn2d.py:146-149, op/fused_act.py:87-94) are the oracle, not a fallback here."""
from __future__ import annotations

import torch
from torch import nn

from . import _lib


def _as_f32_cuda(t: torch.Tensor, what: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.IdbError(f"{what}: CUDA tensor required (no CPU path)")
    if t.dtype != torch.float32:
        raise _lib.IdbError(f"{what}: float32 required, got {t.dtype}")
    return t.contiguous()


def upfirdn2d_op(input: torch.Tensor, kernel: torch.Tensor, up_x, up_y, down_x, down_y, pad_x0, pad_x1, pad_y0,
                 pad_y1) -> torch.Tensor:
    """The extension-level op (op/upfirdn2d.cpp:12-22): input [major, in_h, in_w, minor] -> [major, out_h, out_w, minor]."""
    x = _as_f32_cuda(input, "upfirdn2d")
    k = _as_f32_cuda(kernel, "upfirdn2d")
    if x.dim() != 4 or k.dim() != 2:
        raise _lib.IdbError("upfirdn2d: input must be 4-D [major,in_h,in_w,minor] and kernel 2-D")
    major, in_h, in_w, minor = x.shape
    kh, kw = k.shape
    out_h = (in_h * up_y + pad_y0 + pad_y1 - kh) // down_y + 1
    out_w = (in_w * up_x + pad_x0 + pad_x1 - kw) // down_x + 1
    if out_h <= 0 or out_w <= 0:
        raise _lib.IdbError("upfirdn2d: empty output")
    out = torch.empty(major, out_h, out_w, minor, device=x.device, dtype=torch.float32)
    _lib.call("idb_upfirdn2d", _lib.ptr(x), _lib.ptr(k), major, in_h, in_w, minor, kh, kw, up_x, up_y, down_x, down_y,
              pad_x0, pad_x1, pad_y0, pad_y1, _lib.ptr(out), _lib.stream_ptr())
    return out


def upfirdn2d(input: torch.Tensor, kernel: torch.Tensor, up=1, down=1, pad=(0, 0)) -> torch.Tensor:
    """op/upfirdn2d.py:145-159: NCHW in, NCHW out, same (up, down, pad) on both axes."""
    b, c, in_h, in_w = input.shape
    out = upfirdn2d_op(input.reshape(-1, in_h, in_w, 1), kernel, up, up, down, down, pad[0], pad[1], pad[0], pad[1])
    return out.view(b, c, out.shape[1], out.shape[2])


def fused_bias_act(input: torch.Tensor, bias, refer, act: int, grad: int, alpha: float, scale: float) -> torch.Tensor:
    """The extension-level op (op/fused_bias_act.cpp:11-20); empty bias / refer tensors mean "absent"."""
    x = _as_f32_cuda(input, "fused_bias_act")
    b = _as_f32_cuda(bias, "fused_bias_act") if bias is not None and bias.numel() else None
    r = _as_f32_cuda(refer, "fused_bias_act") if refer is not None and refer.numel() else None
    if r is not None and r.numel() != x.numel():
        raise _lib.IdbError("fused_bias_act: refer must have the shape of input")
    step_b = 1
    for d in x.shape[2:]:
        step_b *= d
    size_b = x.shape[1] if x.dim() > 1 else 1
    if b is not None and b.numel() != size_b:
        raise _lib.IdbError("fused_bias_act: bias must have input.shape[1] elements")
    out = torch.empty_like(x)
    _lib.call("idb_fused_bias_act", _lib.ptr(x), _lib.ptr(b), _lib.ptr(r), act, grad, float(alpha), float(scale),
              x.numel(), step_b, size_b, _lib.ptr(out), _lib.stream_ptr())
    return out


def fused_leaky_relu(input: torch.Tensor, bias: torch.Tensor, negative_slope=0.2, scale=2 ** 0.5) -> torch.Tensor:
    """op/fused_act.py:86-97 (CUDA branch: honours negative_slope, Appendix A.17). Forward only."""
    return fused_bias_act(input, bias, None, 3, 0, negative_slope, scale)


class FusedLeakyReLU(nn.Module):
    """op/fused_act.py:74-84."""

    def __init__(self, channel, negative_slope=0.2, scale=2 ** 0.5):
        super().__init__()
        self.bias = nn.Parameter(torch.zeros(channel))
        self.negative_slope = negative_slope
        self.scale = scale

    def forward(self, input):
        return fused_leaky_relu(input, self.bias, self.negative_slope, self.scale)
