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


# ------------------------------------------------------------------------------------------------
# DDPM U-Net (models/ddpm.py:149-213) as a function of a state_dict -- torch CPU, dtype follows the inputs
# (float32 = the reference's arithmetic; float64 = "truth" for measuring the reference's own rounding floor).
# ------------------------------------------------------------------------------------------------
def _silu(x):
    import torch
    return x * torch.sigmoid(x)


def ddpm_forward(state, cfg, x, labels):
    """state: {'all_modules.<i>....': tensor}; cfg: dict(nf, ch_mult, num_res_blocks, attn_resolutions, image_size,
    resamp_with_conv, centered); x [B,C,H,W]; labels [B].  Follows ddpm.py:149-213 block by block."""
    import math
    import torch
    import torch.nn.functional as F
    dt = x.dtype
    P = lambda i, name: state[f"all_modules.{i}.{name}"].to(dt)
    nf, ch_mult, nrb = cfg["nf"], cfg["ch_mult"], cfg["num_res_blocks"]
    nres = len(ch_mult)
    idx = [0]

    def nxt():
        idx[0] += 1
        return idx[0] - 1

    def gn(i, name, h):
        return F.group_norm(h, 32, P(i, name + ".weight"), P(i, name + ".bias"), eps=1e-6)

    def nin(i, name, h):  # layers.py:555-564
        return torch.einsum("bchw,co->bohw", h, P(i, name + ".W")) + P(i, name + ".b")[None, :, None, None]

    def res(h, temb):  # layers.py:654-675
        i = nxt()
        out = F.conv2d(_silu(gn(i, "GroupNorm_0", h)), P(i, "Conv_0.weight"), P(i, "Conv_0.bias"), padding=1)
        out = out + F.linear(_silu(temb), P(i, "Dense_0.weight"), P(i, "Dense_0.bias"))[:, :, None, None]
        out = F.conv2d(_silu(gn(i, "GroupNorm_1", out)), P(i, "Conv_1.weight"), P(i, "Conv_1.bias"), padding=1)
        if f"all_modules.{i}.NIN_0.W" in state:
            h = nin(i, "NIN_0", h)
        elif f"all_modules.{i}.Conv_2.weight" in state:
            h = F.conv2d(h, P(i, "Conv_2.weight"), P(i, "Conv_2.bias"), padding=1)
        return h + out

    def attn(h):  # layers.py:576-590
        i = nxt()
        B, C, H, W = h.shape
        g = gn(i, "GroupNorm_0", h)
        q, k, v = nin(i, "NIN_0", g), nin(i, "NIN_1", g), nin(i, "NIN_2", g)
        w = torch.einsum("bchw,bcij->bhwij", q, k) * (int(C) ** (-0.5))
        w = F.softmax(w.reshape(B, H, W, H * W), dim=-1).reshape(B, H, W, H, W)
        o = torch.einsum("bhwij,bcij->bchw", w, v)
        return h + nin(i, "NIN_3", o)

    # time embedding, layers.py:524-538 + ddpm.py:152-159
    half = nf // 2
    freqs = torch.exp(torch.arange(half, dtype=torch.float32) * -(math.log(10000) / (half - 1))).to(dt)
    arg = labels.to(dt)[:, None] * freqs[None, :]
    temb = torch.cat([torch.sin(arg), torch.cos(arg)], dim=1)
    i = nxt()
    temb = F.linear(temb, P(i, "weight"), P(i, "bias"))
    i = nxt()
    temb = F.linear(_silu(temb), P(i, "weight"), P(i, "bias"))

    h = x if cfg["centered"] else 2 * x - 1.0
    i = nxt()
    hs = [F.conv2d(h, P(i, "weight"), P(i, "bias"), padding=1)]
    for lvl in range(nres):
        for _ in range(nrb):
            h = res(hs[-1], temb)
            if h.shape[-1] in cfg["attn_resolutions"]:
                h = attn(h)
            hs.append(h)
        if lvl != nres - 1:
            i = nxt()
            if cfg["resamp_with_conv"]:  # layers.py:621-623
                hs.append(F.conv2d(F.pad(hs[-1], (0, 1, 0, 1)), P(i, "Conv_0.weight"), P(i, "Conv_0.bias"), stride=2))
            else:
                hs.append(F.avg_pool2d(hs[-1], 2))
    h = hs[-1]
    h = res(h, temb)
    h = attn(h)
    h = res(h, temb)
    for lvl in reversed(range(nres)):
        for _ in range(nrb + 1):
            h = res(torch.cat([h, hs.pop()], dim=1), temb)
        if h.shape[-1] in cfg["attn_resolutions"]:
            h = attn(h)
        if lvl != 0:
            i = nxt()
            h = F.interpolate(h, scale_factor=2, mode="nearest")  # layers.py:600
            if cfg["resamp_with_conv"]:
                h = F.conv2d(h, P(i, "Conv_0.weight"), P(i, "Conv_0.bias"), padding=1)
    i = nxt()
    h = _silu(F.group_norm(h, 32, P(i, "weight"), P(i, "bias"), eps=1e-6))
    i = nxt()
    return F.conv2d(h, P(i, "weight"), P(i, "bias"), padding=1)


def unet_score_matrix(forward, state, cfg, x0, z, mean_coeff, std, label, batch=64):
    """dim_reduction.py:177-194 + models/utils.py:264-267 for one anchor: x = mean_coeff*x0 + std*z, score = -net(x)/std,
    flatten NCHW, centre.  x0 [C,H,W], z [K,C,H,W]; dtype follows z."""
    import torch
    rows = []
    with torch.no_grad():
        for s in range(0, z.shape[0], batch):
            zz = z[s:s + batch]
            x = mean_coeff * x0[None].to(zz.dtype) + std * zz
            lab = torch.full((zz.shape[0],), label, dtype=zz.dtype)
            rows.append(-forward(state, cfg, x, lab) / std)
    S = torch.cat(rows).flatten(1)
    return S - S.mean(0, keepdim=True)


# ------------------------------------------------------------------------------------------------
# NCSN++ (models/ncsnpp.py:238-388) for the style_gan_base configuration with model.name='ncsnpp':
# Fourier embedding, BigGAN++ res-blocks with FIR resampling, skip rescale, input pyramid ('residual'), progressive='none'.
# ------------------------------------------------------------------------------------------------
def upfirdn2d_torch(x, k, up=1, down=1, pad=(0, 0)):
    """op/upfirdn2d.py:145-200 on NCHW tensors (same pads on both axes), any float dtype."""
    import torch
    import torch.nn.functional as F
    n, c, h, w = x.shape
    k = torch.as_tensor(k, dtype=x.dtype)
    st = x.new_zeros(n, c, h * up, w * up)
    st[:, :, ::up, ::up] = x
    p0, p1 = pad
    st = F.pad(st, [max(p0, 0), max(p1, 0), max(p0, 0), max(p1, 0)])
    st = st[:, :, max(-p0, 0):st.shape[2] - max(-p1, 0), max(-p0, 0):st.shape[3] - max(-p1, 0)]
    wt = torch.flip(k, [0, 1])[None, None].repeat(c, 1, 1, 1)
    out = F.conv2d(st, wt, groups=c)
    return out[:, :, ::down, ::down]


def _fir(taps, gain, dtype):
    import torch
    k = np.asarray(taps, dtype=np.float32)
    k = np.outer(k, k)
    k /= np.sum(k)
    return torch.from_numpy(k * np.float32(gain)).to(dtype)


def ncsnpp_forward(state, cfg, x, labels):
    """cfg: DDPM keys + fir_kernel, fourier_scale (unused here: W is in the state), skip_rescale=True."""
    import math
    import torch
    import torch.nn.functional as F
    dt = x.dtype
    P = lambda i, name: state[f"all_modules.{i}.{name}"].to(dt)
    has = lambda i, name: f"all_modules.{i}.{name}" in state
    nf, ch_mult, nrb = cfg["nf"], cfg["ch_mult"], cfg["num_res_blocks"]
    nres = len(ch_mult)
    taps = cfg["fir_kernel"]
    rs = math.sqrt(2.0)
    idx = [0]

    def nxt():
        idx[0] += 1
        return idx[0] - 1

    def gn(i, name, h):
        c = h.shape[1]
        return F.group_norm(h, min(c // 4, 32), P(i, name + ".weight"), P(i, name + ".bias"), eps=1e-6)

    def nin(i, name, h):
        return torch.einsum("bchw,co->bohw", h, P(i, name + ".W")) + P(i, name + ".b")[None, :, None, None]

    def res(h, temb, up=False, down=False):  # layerspp.py:241-274
        i = nxt()
        x_ = h
        a = _silu(gn(i, "GroupNorm_0", h))
        if up:  # upsample_2d: gain factor^2, pad (2, 1)
            k = _fir(taps, 4.0, dt)
            a, x_ = upfirdn2d_torch(a, k, up=2, pad=(2, 1)), upfirdn2d_torch(x_, k, up=2, pad=(2, 1))
        elif down:  # downsample_2d: pad (1, 1)
            k = _fir(taps, 1.0, dt)
            a, x_ = upfirdn2d_torch(a, k, down=2, pad=(1, 1)), upfirdn2d_torch(x_, k, down=2, pad=(1, 1))
        out = F.conv2d(a, P(i, "Conv_0.weight"), P(i, "Conv_0.bias"), padding=1)
        out = out + F.linear(_silu(temb), P(i, "Dense_0.weight"), P(i, "Dense_0.bias"))[:, :, None, None]
        out = F.conv2d(_silu(gn(i, "GroupNorm_1", out)), P(i, "Conv_1.weight"), P(i, "Conv_1.bias"), padding=1)
        if has(i, "Conv_2.weight"):
            x_ = F.conv2d(x_, P(i, "Conv_2.weight"), P(i, "Conv_2.bias"))
        return (x_ + out) / rs

    def attn(h):  # layerspp.py:75-91
        i = nxt()
        B, C, H, W = h.shape
        g = gn(i, "GroupNorm_0", h)
        q, k, v = nin(i, "NIN_0", g), nin(i, "NIN_1", g), nin(i, "NIN_2", g)
        w = torch.einsum("bchw,bcij->bhwij", q, k) * (int(C) ** (-0.5))
        w = F.softmax(w.reshape(B, H, W, H * W), dim=-1).reshape(B, H, W, H, W)
        o = nin(i, "NIN_3", torch.einsum("bhwij,bcij->bchw", w, v))
        return (h + o) / rs

    i = nxt()
    proj = labels.to(dt)[:, None] * P(i, "W")[None, :] * 2 * math.pi   # layerspp.py:39-41
    temb = torch.cat([torch.sin(proj), torch.cos(proj)], dim=-1)
    i = nxt()
    temb = F.linear(temb, P(i, "weight"), P(i, "bias"))
    i = nxt()
    temb = F.linear(_silu(temb), P(i, "weight"), P(i, "bias"))

    h = x if cfg["centered"] else 2 * x - 1.0
    pyramid = h
    i = nxt()
    hs = [F.conv2d(h, P(i, "weight"), P(i, "bias"), padding=1)]
    for lvl in range(nres):
        for _ in range(nrb):
            h = res(hs[-1], temb)
            if h.shape[-1] in cfg["attn_resolutions"]:
                h = attn(h)
            hs.append(h)
        if lvl != nres - 1:
            h = res(hs[-1], temb, down=True)
            i = nxt()  # conv_downsample_2d: FIR pad (2, 2), then 3x3 stride 2 (up_or_down_sampling.py:144-178)
            f = upfirdn2d_torch(pyramid, _fir(taps, 1.0, dt), pad=(2, 2))
            pyramid = F.conv2d(f, P(i, "Conv2d_0.weight"), stride=2) + P(i, "Conv2d_0.bias")[None, :, None, None]
            pyramid = (pyramid + h) / rs
            h = pyramid
            hs.append(h)
    h = hs[-1]
    h = res(h, temb)
    h = attn(h)
    h = res(h, temb)
    for lvl in reversed(range(nres)):
        for _ in range(nrb + 1):
            h = res(torch.cat([h, hs.pop()], dim=1), temb)
        if h.shape[-1] in cfg["attn_resolutions"]:
            h = attn(h)
        if lvl != 0:
            h = res(h, temb, up=True)
    i = nxt()
    h = _silu(F.group_norm(h, min(h.shape[1] // 4, 32), P(i, "weight"), P(i, "bias"), eps=1e-6))
    i = nxt()
    return F.conv2d(h, P(i, "weight"), P(i, "bias"), padding=1)


# ------------------------------------------------------------------------------------------------
# BeatGANs U-Net (models/BeatGANsUNET.py:235-284), state_dict keys as in the reference module tree.
# ------------------------------------------------------------------------------------------------
def beatgans_forward(state, cfg, x, t):
    import math
    import torch
    import torch.nn.functional as F
    dt = x.dtype
    P = lambda name: state[name].to(dt)
    mult, nrb, mc = cfg["channel_mult"], cfg["num_res_blocks"], cfg["model_channels"]
    heads = cfg["num_heads"]

    def gn(prefix, h):  # GroupNorm32(min(32, C), C), eps 1e-5 (BeatGANs_nn.py:23-25, 98-104)
        return F.group_norm(h, min(32, h.shape[1]), P(prefix + ".weight"), P(prefix + ".bias"), eps=1e-5)

    def res(prefix, h, emb, lateral=None, up=False, down=False):  # BeatGANsblocks.py:185-255
        if lateral is not None:
            h = torch.cat([h, lateral], dim=1)
        x_ = h
        a = _silu(gn(prefix + ".in_layers.0", h))
        if up:
            a, x_ = F.interpolate(a, scale_factor=2, mode="nearest"), F.interpolate(x_, scale_factor=2, mode="nearest")
        elif down:
            a, x_ = F.avg_pool2d(a, 2), F.avg_pool2d(x_, 2)
        a = F.conv2d(a, P(prefix + ".in_layers.2.weight"), P(prefix + ".in_layers.2.bias"), padding=1)
        e = F.linear(_silu(emb), P(prefix + ".emb_layers.1.weight"), P(prefix + ".emb_layers.1.bias"))
        scale, shift = torch.chunk(e[:, :, None, None], 2, dim=1)
        a = gn(prefix + ".out_layers.0", a) * (1 + scale) + shift          # apply_conditions :318-323
        a = F.conv2d(_silu(a), P(prefix + ".out_layers.3.weight"), P(prefix + ".out_layers.3.bias"), padding=1)
        if prefix + ".skip_connection.weight" in state:
            x_ = F.conv2d(x_, P(prefix + ".skip_connection.weight"), P(prefix + ".skip_connection.bias"))
        return x_ + a

    def attn(prefix, h):  # BeatGANsblocks.py:437-443, 475-491
        b, c, hh, ww = h.shape
        xf = h.reshape(b, c, -1)
        qkv = F.conv1d(gn(prefix + ".norm", xf), P(prefix + ".qkv.weight"), P(prefix + ".qkv.bias"))
        ch = c // heads
        q, k, v = qkv.reshape(b * heads, ch * 3, -1).split(ch, dim=1)
        s = 1 / math.sqrt(math.sqrt(ch))
        w = torch.softmax(torch.einsum("bct,bcs->bts", q * s, k * s), dim=-1)
        a = torch.einsum("bts,bcs->bct", w, v).reshape(b, -1, xf.shape[-1])
        a = F.conv1d(a, P(prefix + ".proj_out.weight"), P(prefix + ".proj_out.bias"))
        return (xf + a).reshape(b, c, hh, ww)

    half = mc // 2   # BeatGANs_nn.py:107-125
    freqs = torch.exp(-math.log(10000) * torch.arange(start=0, end=half, dtype=torch.float32) / half).to(dt)
    args = t.to(dt)[:, None] * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    emb = F.linear(emb, P("time_embed.0.weight"), P("time_embed.0.bias"))
    emb = F.linear(_silu(emb), P("time_embed.2.weight"), P("time_embed.2.bias"))

    h = F.conv2d(x, P("input_blocks.0.0.weight"), P("input_blocks.0.0.bias"), padding=1)
    hs, k, resolution = [h], 1, cfg["image_size"]
    for level in range(len(mult)):
        for _ in range(nrb):
            h = res(f"input_blocks.{k}.0", h, emb)
            if resolution in cfg["attention_resolutions"]:
                h = attn(f"input_blocks.{k}.1", h)
            hs.append(h)
            k += 1
        if level != len(mult) - 1:
            resolution //= 2
            h = res(f"input_blocks.{k}.0", h, emb, down=True)
            hs.append(h)
            k += 1
    h = res("middle_block.0", h, emb)
    h = attn("middle_block.1", h)
    h = res("middle_block.2", h, emb)
    k = 0
    for level in reversed(range(len(mult))):
        for i in range(nrb + 1):
            h = res(f"output_blocks.{k}.0", h, emb, lateral=hs.pop())
            j = 1
            if resolution in cfg["attention_resolutions"]:
                h = attn(f"output_blocks.{k}.{j}", h)
                j += 1
            if level and i == nrb:
                resolution *= 2
                h = res(f"output_blocks.{k}.{j}", h, emb, up=True)
            k += 1
    h = _silu(gn("out.0", h))
    return F.conv2d(h, P("out.2.weight"), P("out.2.bias"), padding=1)
