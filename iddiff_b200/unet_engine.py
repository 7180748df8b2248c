"""Device pipeline of the image score networks (DDPM U-Net) for the `manifold_dimension` path.

    anchors [A, C*H*W] --idb_perturb--> x [A, K, C, H, W] fp32
      --(per image chunk) U-Net: idb_conv_prep / idb_groupnorm_stats / idb_conv2d / idb_attention--> raw net output
      written straight into the score matrix S [A, K, H*W*C] (NHWC column order: a column permutation of the reference's
      NCHW flatten, dim_reduction.py:191, which leaves singular values unchanged)
      --centre + Gram + spectrum + gap read-out (engine.spectrum_from_rows, scale = -1/std)--> singular values, ID

Python only sequences C-ABI calls; torch owns memory and streams.  `t` is one constant on this path
(dim_reduction.py:173-174), so everything that depends only on t -- the time embedding MLP (ddpm.py:152-159) and each
block's Dense_0(act(temb)) (layers.py:662-664) -- is evaluated once per label on the device and folded into the
convolution biases.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch

from . import _lib, unet_ops as U
from .engine import SpectrumResult, _stage, spectrum_from_rows


def timestep_embedding(label: float, dim: int, max_positions: int = 10000) -> torch.Tensor:
    """layers.py:524-538 for one timestep, same float32 torch ops as the reference -> [1, dim] on the host."""
    half = dim // 2
    step = math.log(max_positions) / (half - 1)
    freqs = torch.exp(torch.arange(half, dtype=torch.float32) * -step)
    arg = torch.tensor([label], dtype=torch.float32)[:, None] * freqs[None, :]
    emb = torch.cat([torch.sin(arg), torch.cos(arg)], dim=1)
    if dim % 2 == 1:
        emb = torch.nn.functional.pad(emb, (0, 1))
    return emb


class _Res:
    """ResnetBlockDDPM weights in kernel form."""

    def __init__(self, m, dev):
        g = lambda t: t.detach().to(dev, torch.float32).contiguous()
        self.in_ch, self.out_ch = m.in_ch, m.out_ch
        self.gn0 = (g(m.GroupNorm_0.weight), g(m.GroupNorm_0.bias))
        self.gn1 = (g(m.GroupNorm_1.weight), g(m.GroupNorm_1.bias))
        self.conv0 = U.pack_conv(g(m.Conv_0.weight), g(m.Conv_0.bias))
        self.conv1 = U.pack_conv(g(m.Conv_1.weight), g(m.Conv_1.bias))
        self.dense = (g(m.Dense_0.weight), g(m.Dense_0.bias)) if hasattr(m, "Dense_0") else None
        self.short = None
        if m.in_ch != m.out_ch:
            if m.conv_shortcut:
                self.short = U.pack_conv(g(m.Conv_2.weight), g(m.Conv_2.bias))
            else:
                self.short = U.pack_conv(g(m.NIN_0.W).t().contiguous(), g(m.NIN_0.b))


class _Attn:
    def __init__(self, m, dev):
        g = lambda t: t.detach().to(dev, torch.float32).contiguous()
        self.c = m.channels
        self.gn = (g(m.GroupNorm_0.weight), g(m.GroupNorm_0.bias))
        wqkv = torch.cat([g(m.NIN_0.W).t(), g(m.NIN_1.W).t(), g(m.NIN_2.W).t()], dim=0).contiguous()  # [3C, C]
        bqkv = torch.cat([g(m.NIN_0.b), g(m.NIN_1.b), g(m.NIN_2.b)]).contiguous()
        self.qkv = U.pack_conv(wqkv, bqkv)   # the three NINs of layers.py:577-579 as one 1x1 convolution
        self.proj = U.pack_conv(g(m.NIN_3.W).t().contiguous(), g(m.NIN_3.b))


class UNetEngine:
    """DDPM U-Net (models/ddpm.py) on the sm_100a kernels.  Build with `UNetEngine.from_module(holder, device)`."""

    GROUPS, EPS = 32, 1e-6

    def __init__(self, holder, device):
        self.dev = torch.device(device)
        self.h = holder
        g = lambda t: t.detach().to(self.dev, torch.float32).contiguous()
        mods = holder.all_modules
        self.temb_w = [(g(mods[0].weight), g(mods[0].bias)), (g(mods[1].weight), g(mods[1].bias))]
        self.blocks: Dict[int, object] = {}
        for kind, idx, _ in holder.plan:
            m = mods[idx] if idx >= 0 else None
            if kind == "res":
                self.blocks[idx] = _Res(m, self.dev)
            elif kind == "attn":
                self.blocks[idx] = _Attn(m, self.dev)
            elif kind in ("conv_in", "conv_out"):
                self.blocks[idx] = U.pack_conv(g(m.weight), g(m.bias))
            elif kind == "down":
                self.blocks[idx] = U.pack_conv(g(m.Conv_0.weight), g(m.Conv_0.bias), taps=U.TAPS_3x3_S2, planes=4) \
                    if m.with_conv else None
            elif kind == "up":
                self.blocks[idx] = U.pack_conv(g(m.Conv_0.weight), g(m.Conv_0.bias)) if m.with_conv else None
            elif kind == "gn_out":
                self.blocks[idx] = (g(m.weight), g(m.bias))
        cin = mods[holder.plan[0][1]].weight.shape[1]
        self.in_affine = None
        if not holder.centered:  # ddpm.py:163-168: h = 2x - 1 for data in [0, 1]
            self.in_affine = (torch.full((cin,), 2.0, device=self.dev), torch.full((cin,), -1.0, device=self.dev))
        self.cin = cin
        self.cout = mods[holder.plan[-1][1]].weight.shape[0]
        self._label_cache: Dict[float, Dict[int, torch.Tensor]] = {}

    @classmethod
    def from_module(cls, holder, device="cuda"):
        return cls(holder, device)

    # ---- per-label constants ------------------------------------------------------------------
    def _biases_for(self, label: float) -> Dict[int, torch.Tensor]:
        """conv0 bias + Dense_0(SiLU(temb)) for every ResnetBlock (layers.py:662-664), evaluated on the device."""
        key = float(label)
        if key in self._label_cache:
            return self._label_cache[key]
        temb = timestep_embedding(key, self.h.nf).to(self.dev)
        temb = U.dense(temb, *self.temb_w[0], act_in=0)
        temb = U.dense(temb, *self.temb_w[1], act_in=2)
        out = {}
        for idx, blk in self.blocks.items():
            if isinstance(blk, _Res) and blk.dense is not None:
                proj = U.dense(temb, *blk.dense, act_in=2)            # [1, out_ch]
                eff = torch.empty_like(proj)
                # idb_fused_bias_act as a plain "add the per-channel vector" (act 1 = linear, scale 1)
                _lib.call("idb_fused_bias_act", _lib.ptr(proj), _lib.ptr(blk.conv0.bias), None, 1, 0, 0.0, 1.0,
                          proj.numel(), 1, proj.numel(), _lib.ptr(eff), _lib.stream_ptr())
                out[idx] = eff.reshape(-1)
        self._label_cache = {key: out}
        return out

    # ---- blocks --------------------------------------------------------------------------------
    def _res(self, blk: _Res, bias0, x0, x1=None):
        mr = U.gn_stats(x0, x1, self.GROUPS, self.EPS)
        a = U.prep(x0, x1, mean_rstd=mr, groups=self.GROUPS, gamma=blk.gn0[0], beta=blk.gn0[1], act=1)
        h = U.conv(a, blk.conv0, bias=bias0)
        mr = U.gn_stats(h, None, self.GROUPS, self.EPS)
        a = U.prep(h, mean_rstd=mr, groups=self.GROUPS, gamma=blk.gn1[0], beta=blk.gn1[1], act=1)
        if blk.short is not None:
            xs = U.conv(U.prep(x0, x1), blk.short)
        else:
            xs = x0
        return U.conv(a, blk.conv1, residual=xs)

    def _attn(self, blk: _Attn, x):
        mr = U.gn_stats(x, None, self.GROUPS, self.EPS)
        a = U.prep(x, mean_rstd=mr, groups=self.GROUPS, gamma=blk.gn[0], beta=blk.gn[1])
        qkv = U.conv(a, blk.qkv)
        c = blk.c
        o = U.attention(qkv, 1, c, 0, c, 2 * c, c, float(int(c) ** (-0.5)))
        return U.conv(U.prep(o), blk.proj, residual=x)

    def forward(self, x_nchw: torch.Tensor, label: float, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x [B, C, H, W] fp32 (device) -> raw network output [B, H, W, Cout] fp32 NHWC (ddpm.py:149-213)."""
        if not (x_nchw.is_cuda and x_nchw.dtype == torch.float32 and x_nchw.is_contiguous() and x_nchw.dim() == 4):
            raise _lib.IdbError("UNetEngine.forward: contiguous float32 CUDA [B, C, H, W] required")
        biases = self._biases_for(label)
        h, hs = None, []
        for kind, idx, extra in self.h.plan:
            blk = self.blocks.get(idx)
            if kind == "conv_in":
                ga, be = self.in_affine if self.in_affine else (None, None)
                h = U.conv(U.prep(x_nchw, nchw=True, gamma=ga, beta=be), blk)
            elif kind == "push":
                hs.append(h)
            elif kind == "res":
                h = self._res(blk, biases.get(idx), h, hs.pop() if extra.get("skip") else None)
            elif kind == "attn":
                h = self._attn(blk, h)
            elif kind == "down":
                if blk is not None:
                    h = U.conv(U.prep(h, mode=U.PREP_STRIDE2), blk)
                else:
                    h = U.prep(h, mode=U.PREP_AVGPOOL2, want_operand=False, want_f32=True)[1]
            elif kind == "up":
                if blk is not None:
                    h = U.conv(U.prep(h, mode=U.PREP_UP2), blk)
                else:
                    h = U.prep(h, mode=U.PREP_UP2, want_operand=False, want_f32=True)[1]
            elif kind == "gn_out":
                mr = U.gn_stats(h, None, self.GROUPS, self.EPS)
                h = U.prep(h, mean_rstd=mr, groups=self.GROUPS, gamma=blk[0], beta=blk[1], act=1)
            elif kind == "conv_out":
                h = U.conv(h, blk, out=out)
        assert not hs
        return h

    # ---- the manifold_dimension step ------------------------------------------------------------
    def bytes_per_anchor(self, K: int, D: int) -> int:
        return K * D * 4 * 2 + D * D * 8

    def estimate(self, x0: torch.Tensor, K: int, mean_coeff: float, std: float, label: float, *,
                 z: Optional[torch.Tensor] = None, seed: int = 0, anchor_id0: int = 0, image_shape=None,
                 image_chunk: int = 128, want_scores: bool = False) -> SpectrumResult:
        """x0 [A, C*H*W] (NCHW flatten) -> spectra + IDs of the K x D centred score matrices."""
        A, D = x0.shape
        c, hh, ww = image_shape if image_shape is not None else self.image_shape
        assert c * hh * ww == D
        dev = self.dev
        x = torch.empty(A, K, D, dtype=torch.float32, device=dev)
        with _stage("perturb", A * K * D * 4.0):
            _lib.call("idb_perturb", _lib.ptr(x0), _lib.ptr(z), A, K, D, 0, 0, float(mean_coeff), float(std), seed,
                      anchor_id0, _lib.ptr(x), 0, 1.0, None, None, _lib.stream_ptr())
        S = torch.empty(A, K, hh * ww * self.cout, dtype=torch.float32, device=dev)
        xi = x.view(A * K, c, hh, ww)
        So = S.view(A * K, hh, ww, self.cout)
        with _stage("score_net", 0.0):
            for s in range(0, A * K, image_chunk):
                e = min(A * K, s + image_chunk)
                self.forward(xi[s:e], label, out=So[s:e])
        del x
        res = spectrum_from_rows(S, K, hh * ww * self.cout, -1.0 / std)
        if want_scores:
            # models/utils.py:264-267 score = -model(x)/std, back in the reference's NCHW column order
            res.scores = (-(So.view(A, K, hh, ww, self.cout).permute(0, 1, 4, 2, 3)) / std).reshape(A, K, -1)
        return res
