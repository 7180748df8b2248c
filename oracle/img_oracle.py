"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy / torch-CPU) of the image-network side of the
`manifold_dimension` path: the two StyleGAN-derived native ops and the U-Net score networks.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product path never does.

Parity status: PINNED by tests/golden/ops_*.npz and tests/golden/unet_*.npz, which oracle/gen_golden_img.py
produced by running the real reference (/root/reference) in the build container.

Every function cites the reference file:line it follows.
"""
from __future__ import annotations

import numpy as np


def upfirdn2d_np(x, k, up_x, up_y, down_x, down_y, pad_x0, pad_x1, pad_y0, pad_y1):
    """op/upfirdn2d.py:162-200 (upfirdn2d_native) on [major, in_h, in_w, minor] -> [major, out_h, out_w, minor].
    Zero-stuff (:171-173), pad then crop negative pads (:175-183), correlate with the flipped kernel = true
    convolution (:189-190), decimate (:198).  Accumulates in the dtype of x, taps in (ky, kx) order."""
    x = np.asarray(x)
    k = np.asarray(k, dtype=x.dtype)
    major, in_h, in_w, minor = x.shape
    kh, kw = k.shape
    st = np.zeros((major, in_h * up_y, in_w * up_x, minor), x.dtype)
    st[:, ::up_y, ::up_x, :] = x
    st = np.pad(st, [(0, 0), (max(pad_y0, 0), max(pad_y1, 0)), (max(pad_x0, 0), max(pad_x1, 0)), (0, 0)])
    st = st[:, max(-pad_y0, 0):st.shape[1] - max(-pad_y1, 0), max(-pad_x0, 0):st.shape[2] - max(-pad_x1, 0), :]
    fh, fw = st.shape[1] - kh + 1, st.shape[2] - kw + 1
    out = np.zeros((major, fh, fw, minor), x.dtype)
    kf = k[::-1, ::-1]
    for ky in range(kh):
        for kx in range(kw):
            out += st[:, ky:ky + fh, kx:kx + fw, :] * kf[ky, kx]
    return out[:, ::down_y, ::down_x, :]


def fused_bias_act_np(x, b, ref, act, grad, alpha, scale):
    """op/fused_bias_act_kernel.cu:19-53: y = f(x + b[channel]) * scale, bias along dim 1 (:80-84)."""
    x = np.asarray(x)
    v = x.copy()
    if b is not None:
        shape = [1, -1] + [1] * (x.ndim - 2)
        v = v + np.asarray(b, x.dtype).reshape(shape)
    code = act * 10 + grad
    if code in (10, 11):
        y = v
    elif code == 30:
        y = np.where(v > 0, v, v * x.dtype.type(alpha))
    elif code == 31:
        y = np.where(np.asarray(ref) > 0, v, v * x.dtype.type(alpha))
    else:
        y = np.zeros_like(v)
    return (y * x.dtype.type(scale)).astype(x.dtype)


def fused_leaky_relu_cpu_np(x, b, scale):
    """op/fused_act.py:87-94, the reference's CPU branch: slope hard-coded to 0.2 whatever `negative_slope` is."""
    return fused_bias_act_np(x, b, None, 3, 0, 0.2, scale)
