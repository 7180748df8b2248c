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


def _f32(t, dev):
    return t.detach().to(dev, torch.float32).contiguous()


class _Res:
    """ResnetBlockDDPM (layers.py:632-675) / ResnetBlockBigGANpp (layerspp.py:212-274) weights in kernel form."""

    def __init__(self, m, dev):
        g = lambda t: _f32(t, dev)
        self.in_ch, self.out_ch = m.in_ch, m.out_ch
        self.up, self.down = bool(getattr(m, "up", False)), bool(getattr(m, "down", False))
        self.groups0, self.groups1 = m.GroupNorm_0.num_groups, m.GroupNorm_1.num_groups
        self.out_scale = float(1.0 / math.sqrt(2.0)) if getattr(m, "skip_rescale", False) else 1.0
        self.gn0 = (g(m.GroupNorm_0.weight), g(m.GroupNorm_0.bias))
        self.gn1 = (g(m.GroupNorm_1.weight), g(m.GroupNorm_1.bias))
        self.conv0 = U.pack_conv(g(m.Conv_0.weight), g(m.Conv_0.bias))
        self.conv1 = U.pack_conv(g(m.Conv_1.weight), g(m.Conv_1.bias))
        self.dense = (g(m.Dense_0.weight), g(m.Dense_0.bias)) if hasattr(m, "Dense_0") else None
        self.short = None
        if hasattr(m, "Conv_2"):      # 3x3 (DDPM conv_shortcut) or 1x1 (BigGAN++)
            self.short = U.pack_conv(g(m.Conv_2.weight), g(m.Conv_2.bias))
        elif hasattr(m, "NIN_0"):
            self.short = U.pack_conv(g(m.NIN_0.W).t().contiguous(), g(m.NIN_0.b))
        if self.up or self.down:
            if not m.fir:
                raise NotImplementedError("BigGAN res-block resampling without FIR (naive_up/downsample) is not built")
            # upsample_2d: k * factor^2, up 2, pad (2, 1); downsample_2d: k, down 2, pad (1, 1)  (up_or_down_sampling.py:181-257)
            self.fir = U.fir_kernel(m.fir_kernel, 4.0 if self.up else 1.0, dev)
            self.fir_host = U.fir_kernel_host(m.fir_kernel, 4.0 if self.up else 1.0)


class _Attn:
    """AttnBlock (layers.py:567-590) / AttnBlockpp (layerspp.py:62-91)."""

    def __init__(self, m, dev):
        g = lambda t: _f32(t, dev)
        self.c = m.channels
        self.groups = m.GroupNorm_0.num_groups
        self.out_scale = float(1.0 / math.sqrt(2.0)) if getattr(m, "skip_rescale", False) else 1.0
        self.gn = (g(m.GroupNorm_0.weight), g(m.GroupNorm_0.bias))
        wqkv = torch.cat([g(m.NIN_0.W).t(), g(m.NIN_1.W).t(), g(m.NIN_2.W).t()], dim=0).contiguous()  # [3C, C]
        bqkv = torch.cat([g(m.NIN_0.b), g(m.NIN_1.b), g(m.NIN_2.b)]).contiguous()
        self.qkv = U.pack_conv(wqkv, bqkv)   # the three NINs of layers.py:577-579 as one 1x1 convolution
        self.proj = U.pack_conv(g(m.NIN_3.W).t().contiguous(), g(m.NIN_3.b))


class _PyrDown:
    """layerspp.Downsample(fir, with_conv) on the input pyramid (ncsnpp.py:297-304): FIR pad (2, 2) then 3x3 stride 2."""

    def __init__(self, m, dev):
        self.conv = U.pack_conv(_f32(m.Conv2d_0.weight, dev), _f32(m.Conv2d_0.bias, dev), taps=U.TAPS_3x3_S2, planes=4)
        self.fir = U.fir_kernel(m.fir_kernel, 1.0, dev)
        self.fir_host = U.fir_kernel_host(m.fir_kernel, 1.0)


class _BRes:
    """BeatGANs ResBlock (BeatGANsblocks.py:83-255): GN32-SiLU-[avg-pool | nearest up]-conv, scale-shift conditioning
    (apply_conditions :258-332), GN32-SiLU-conv, 1x1 skip when the channel count changes."""

    def __init__(self, m, dev):
        g = lambda t: _f32(t, dev)
        self.in_ch, self.out_ch, self.up, self.down = m.in_ch, m.out_ch, m.up, m.down
        self.groups0, self.groups1 = m.in_layers[0].num_groups, m.out_layers[0].num_groups
        self.gn0 = (g(m.in_layers[0].weight), g(m.in_layers[0].bias))
        self.gn1 = (g(m.out_layers[0].weight), g(m.out_layers[0].bias))
        self.conv0 = U.pack_conv(g(m.in_layers[2].weight), g(m.in_layers[2].bias))
        self.conv1 = U.pack_conv(g(m.out_layers[3].weight), g(m.out_layers[3].bias))
        self.emb = (g(m.emb_layers[1].weight), g(m.emb_layers[1].bias))
        sc = m.skip_connection
        self.short = U.pack_conv(g(sc.weight), g(sc.bias)) if isinstance(sc, torch.nn.Conv2d) else None


class _BAttn:
    """BeatGANs AttentionBlock + QKVAttentionLegacy (BeatGANsblocks.py:399-491)."""

    def __init__(self, m, dev):
        g = lambda t: _f32(t, dev)
        self.c, self.heads = m.channels, m.num_heads
        self.groups = m.norm.num_groups
        self.gn = (g(m.norm.weight), g(m.norm.bias))
        self.qkv = U.pack_conv(g(m.qkv.weight)[:, :, 0].contiguous(), g(m.qkv.bias))
        self.proj = U.pack_conv(g(m.proj_out.weight)[:, :, 0].contiguous(), g(m.proj_out.bias))


class UNetEngine:
    """The U-Net score networks (models/ddpm.py, models/ncsnpp.py, models/BeatGANsUNET.py) on the sm_100a kernels.
    Build with `UNetEngine.from_module(holder, device)`."""

    def __init__(self, holder, device):
        self.dev = torch.device(device)
        self.h = holder
        g = lambda t: t.detach().to(self.dev, torch.float32).contiguous()
        self.EPS = float(holder.gn_eps)
        self.embedding = getattr(holder, "embedding_type", "positional")
        self.fourier = self.embedding == "fourier"
        if self.embedding == "beatgans":
            lin = [holder.time_embed[0], holder.time_embed[2]]
        else:
            t0 = 1 if self.fourier else 0
            lin = [holder.all_modules[t0], holder.all_modules[t0 + 1]]
        self.fourier_w = holder.all_modules[0].W.detach().float().cpu() if self.fourier else None
        self.temb_w = [(g(l.weight), g(l.bias)) for l in lin]
        self.rescale = float(1.0 / math.sqrt(2.0)) if getattr(holder, "skip_rescale", False) else 1.0
        self.blocks: Dict[int, object] = {}
        for kind, idx, _ in holder.plan:
            m = holder.plan_module(idx) if idx >= 0 else None
            if kind == "bres":
                self.blocks[idx] = _BRes(m, self.dev)
            elif kind == "battn":
                self.blocks[idx] = _BAttn(m, self.dev)
            elif kind == "res":
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
            elif kind == "pyr_down":
                self.blocks[idx] = _PyrDown(m, self.dev)
            elif kind == "gn_out":
                self.blocks[idx] = (g(m.weight), g(m.bias), m.num_groups)
        cin = holder.plan_module(holder.plan[0][1]).weight.shape[1]
        self.in_affine = None
        if not holder.centered:  # ddpm.py:163-168: h = 2x - 1 for data in [0, 1]
            self.in_affine = (torch.full((cin,), 2.0, device=self.dev), torch.full((cin,), -1.0, device=self.dev))
        self.cin = cin
        self.cout = holder.plan_module(holder.plan[-1][1]).weight.shape[0]
        self._label_cache: Dict[float, Dict[int, torch.Tensor]] = {}
        self._graphs: Dict[tuple, tuple] = {}

    @classmethod
    def from_module(cls, holder, device="cuda"):
        return cls(holder, device)

    # ---- per-label constants ------------------------------------------------------------------
    def _biases_for(self, label: float) -> Dict[int, torch.Tensor]:
        """conv0 bias + Dense_0(SiLU(temb)) for every ResnetBlock (layers.py:662-664), evaluated on the device."""
        key = float(label)
        if key in self._label_cache:
            return self._label_cache[key]
        if self.fourier:  # layerspp.py:39-41, same float32 torch expression as the reference
            xp = torch.tensor([key], dtype=torch.float32)[:, None] * self.fourier_w[None, :] * 2 * math.pi
            temb = torch.cat([torch.sin(xp), torch.cos(xp)], dim=-1).to(self.dev)
        elif self.embedding == "beatgans":  # BeatGANs_nn.py:107-125: cos first, frequencies exp(-ln(1e4) i / half)
            half = self.h.nf // 2
            freqs = torch.exp(-math.log(10000) * torch.arange(start=0, end=half, dtype=torch.float32) / half)
            args = torch.tensor([key], dtype=torch.float32)[:, None] * freqs[None]
            temb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
            if self.h.nf % 2:
                temb = torch.cat([temb, torch.zeros_like(temb[:, :1])], dim=-1)
            temb = temb.to(self.dev)
        else:
            temb = timestep_embedding(key, self.h.nf).to(self.dev)
        temb = U.dense(temb, *self.temb_w[0], act_in=0)
        temb = U.dense(temb, *self.temb_w[1], act_in=2)
        out = {}
        for idx, blk in self.blocks.items():
            if isinstance(blk, _BRes):
                # apply_conditions (BeatGANsblocks.py:258-332): GN(h) * (1 + scale) + shift with (scale, shift) =
                # chunk(Linear(SiLU(emb))) -- constants of the label, folded into the GroupNorm affine (parameter
                # preparation on [C]-sized vectors, done once per label like weight packing).
                ss = U.dense(temb, *blk.emb, act_in=2).reshape(-1)
                scale1 = 1.0 + ss[:blk.out_ch]
                out[idx] = ((blk.gn1[0] * scale1).contiguous(), (blk.gn1[1] * scale1 + ss[blk.out_ch:]).contiguous())
            elif isinstance(blk, _Res) and blk.dense is not None:
                proj = U.dense(temb, *blk.dense, act_in=2)            # [1, out_ch]
                eff = torch.empty_like(proj)
                # idb_fused_bias_act as a plain "add the per-channel vector" (act 1 = linear, scale 1)
                _lib.call("idb_fused_bias_act", _lib.ptr(proj), _lib.ptr(blk.conv0.bias), None, 1, 0, 0.0, 1.0,
                          proj.numel(), 1, proj.numel(), _lib.ptr(eff), _lib.stream_ptr())
                out[idx] = eff.reshape(-1)
        self._label_cache[key] = out   # captured graphs reference these vectors: an entry lives as long as a graph of its label
        # a t-sweep visits many labels (12 per call of get_conditional_manifold_dimension): keep the cache bounded, never
        # dropping a label a live graph was captured with
        if len(self._label_cache) > self.LABEL_CACHE:
            pinned = {k[1] for k in self._graphs}
            for old in [k for k in self._label_cache if k not in pinned and k != key][:len(self._label_cache) - self.LABEL_CACHE]:
                del self._label_cache[old]
        return out

    # ---- blocks --------------------------------------------------------------------------------
    def _res(self, blk: _Res, bias0, x0, x1=None):
        mr = U.gn_stats(x0, x1, blk.groups0, self.EPS)
        if blk.up or blk.down:
            # layerspp.py:244-257: FIR resampling of BOTH act(GN(x)) and x before the convolutions
            assert x1 is None
            if U.FUSED_FIR:
                # ONE gathering pass for both branches: act(GN(x)) -> FIR -> operand, and x -> FIR -> fp32; neither the
                # normalised nor the filtered activation is materialised (no upfirdn2d launch)
                updn = (2, 1, 2, 1) if blk.up else (1, 2, 1, 1)
                a, x0 = U.prep_fir(x0, blk.fir_host, *updn, mean_rstd=mr, groups=blk.groups0, gamma=blk.gn0[0],
                                   beta=blk.gn0[1], act=1, want_raw=True)
                return self._res_tail(blk, bias0, a, x0, None)
            _, hf = U.prep(x0, mean_rstd=mr, groups=blk.groups0, gamma=blk.gn0[0], beta=blk.gn0[1], act=1,
                           want_operand=False, want_f32=True)
            if blk.up:
                hf, xr = U.upfirdn2d_nhwc(hf, blk.fir, 2, 1, 2, 1), U.upfirdn2d_nhwc(x0, blk.fir, 2, 1, 2, 1)
            else:
                hf, xr = U.upfirdn2d_nhwc(hf, blk.fir, 1, 2, 1, 1), U.upfirdn2d_nhwc(x0, blk.fir, 1, 2, 1, 1)
            a = U.prep(hf)
            x0 = xr
        else:
            a = U.prep(x0, x1, mean_rstd=mr, groups=blk.groups0, gamma=blk.gn0[0], beta=blk.gn0[1], act=1)
        return self._res_tail(blk, bias0, a, x0, x1)

    def _res_tail(self, blk: _Res, bias0, a, x0, x1):
        h = U.conv(a, blk.conv0, bias=bias0, want_stats=True)
        mr = U.gn_stats(h, None, blk.groups1, self.EPS)
        a = U.prep(h, mean_rstd=mr, groups=blk.groups1, gamma=blk.gn1[0], beta=blk.gn1[1], act=1)
        if blk.short is not None:
            xs = U.conv(U.prep(x0, x1), blk.short)
        elif x1 is not None:
            raise _lib.IdbError("res-block with a concatenated input needs a shortcut projection")
        else:
            xs = x0
        return U.conv(a, blk.conv1, residual=xs, out_scale=blk.out_scale, want_stats=True)

    def _attn(self, blk: _Attn, x):
        mr = U.gn_stats(x, None, blk.groups, self.EPS)
        a = U.prep(x, mean_rstd=mr, groups=blk.groups, gamma=blk.gn[0], beta=blk.gn[1])
        qkv = U.conv(a, blk.qkv)
        c = blk.c
        o = U.attention(qkv, 1, c, 0, c, 2 * c, c, float(int(c) ** (-0.5)))
        return U.conv(U.prep(o), blk.proj, residual=x, out_scale=blk.out_scale, want_stats=True)

    def _bres(self, blk: _BRes, affine1, x0, x1=None):
        mode = U.PREP_UP2 if blk.up else (U.PREP_AVGPOOL2 if blk.down else U.PREP_SAME)
        mr = U.gn_stats(x0, x1, blk.groups0, self.EPS)
        a = U.prep(x0, x1, mode=mode, mean_rstd=mr, groups=blk.groups0, gamma=blk.gn0[0], beta=blk.gn0[1], act=1)
        h = U.conv(a, blk.conv0, want_stats=True)
        mr = U.gn_stats(h, None, blk.groups1, self.EPS)
        a = U.prep(h, mean_rstd=mr, groups=blk.groups1, gamma=affine1[0], beta=affine1[1], act=1)
        if blk.up or blk.down:   # x_upd on the skip branch (BeatGANsblocks.py:199-204); channel count unchanged
            assert x1 is None and blk.short is None
            xs = U.prep(x0, mode=mode, want_operand=False, want_f32=True)[1]
        elif blk.short is not None:
            xs = U.conv(U.prep(x0, x1), blk.short)
        elif x1 is not None:
            raise _lib.IdbError("res-block with a lateral input needs a skip projection")
        else:
            xs = x0
        return U.conv(a, blk.conv1, residual=xs, want_stats=True)

    def _battn(self, blk: _BAttn, x):
        mr = U.gn_stats(x, None, blk.groups, self.EPS)
        a = U.prep(x, mean_rstd=mr, groups=blk.groups, gamma=blk.gn[0], beta=blk.gn[1])
        qkv = U.conv(a, blk.qkv)
        dh = blk.c // blk.heads
        # legacy head layout: qkv.reshape(bs * heads, 3 * ch, T) -> head h owns channels [3 ch h, 3 ch (h + 1)) = q | k | v
        o = U.attention(qkv, blk.heads, dh, 0, dh, 2 * dh, 3 * dh, 1.0 / math.sqrt(dh))
        return U.conv(U.prep(o), blk.proj, residual=x, want_stats=True)

    def forward(self, x_nchw: torch.Tensor, label: float, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x [B, C, H, W] fp32 (device) -> raw network output [B, H, W, Cout] fp32 NHWC (ddpm.py:149-213)."""
        if not (x_nchw.is_cuda and x_nchw.dtype == torch.float32 and x_nchw.is_contiguous() and x_nchw.dim() == 4):
            raise _lib.IdbError("UNetEngine.forward: contiguous float32 CUDA [B, C, H, W] required")
        biases = self._biases_for(label)
        h, hs, pyramid = None, [], None
        for kind, idx, extra in self.h.plan:
            blk = self.blocks.get(idx)
            if kind == "conv_in":
                ga, be = self.in_affine if self.in_affine else (None, None)
                if extra.get("pyramid"):  # ncsnpp.py:268-271: the (rescaled) input also feeds the input pyramid
                    a, pyramid = U.prep(x_nchw, nchw=True, gamma=ga, beta=be, want_f32=True)
                else:
                    a = U.prep(x_nchw, nchw=True, gamma=ga, beta=be)
                h = U.conv(a, blk, want_stats=True)
            elif kind == "pyr_down":
                # conv_downsample_2d (up_or_down_sampling.py:144-178) then (pyramid + h) / sqrt(2) (ncsnpp.py:300-304)
                if U.FUSED_FIR and pyramid.shape[-1] % 8 == 0:   # (the 3-channel image itself takes the scalar path: unfused is faster)
                    a = U.prep_fir(pyramid, blk.fir_host, 1, 1, 2, 2, as_planes=True)
                else:
                    a = U.prep(U.upfirdn2d_nhwc(pyramid, blk.fir, 1, 1, 2, 2), mode=U.PREP_STRIDE2)
                h = U.conv(a, blk.conv, residual=h, out_scale=self.rescale,
                           out_hw=(h.shape[1], h.shape[2]), want_stats=True)
                pyramid = h
            elif kind == "push":
                hs.append(h)
            elif kind == "res":
                h = self._res(blk, biases.get(idx), h, hs.pop() if extra.get("skip") else None)
            elif kind == "bres":
                h = self._bres(blk, biases[idx], h, hs.pop() if extra.get("skip") else None)
            elif kind == "battn":
                h = self._battn(blk, h)
            elif kind == "attn":
                h = self._attn(blk, h)
            elif kind == "down":
                if blk is not None:
                    h = U.conv(U.prep(h, mode=U.PREP_STRIDE2), blk, want_stats=True)
                else:
                    h = U.prep(h, mode=U.PREP_AVGPOOL2, want_operand=False, want_f32=True)[1]
            elif kind == "up":
                if blk is not None:
                    h = U.conv(U.prep(h, mode=U.PREP_UP2), blk, want_stats=True)
                else:
                    h = U.prep(h, mode=U.PREP_UP2, want_operand=False, want_f32=True)[1]
            elif kind == "gn_out":
                mr = U.gn_stats(h, None, blk[2], self.EPS)
                h = U.prep(h, mean_rstd=mr, groups=blk[2], gamma=blk[0], beta=blk[1], act=1)
            elif kind == "conv_out":
                h = U.conv(h, blk, out=out)
        assert not hs
        return h

    # ---- the manifold_dimension step ------------------------------------------------------------
    def bytes_per_anchor(self, K: int, D: int) -> int:
        return K * D * 4 * 2 + D * D * 8

    LABEL_CACHE = 16    # per-label constant sets kept (time-embedding biases / GroupNorm affines), besides those graphs pin
    IMAGE_CHUNK = 1024  # images per forward: fills the 148 SMs at the 8x8 / 4x4 levels too (measured +7 % at 512 over 128, +2 % more at 1024)
    USE_GRAPH = True    # replay one captured CUDA graph per full chunk (494-594 launches per forward; measured +5 %)

    def _run_chunks(self, xi: torch.Tensor, So: torch.Tensor, label: float, image_chunk: Optional[int]):
        """xi [M, C, H, W] -> So [M, H, W, Cout], `image_chunk` images at a time.  Full chunks replay a CUDA graph of the
        whole forward (captured once per (chunk, label, shape); static input / output buffers, torch's graph memory
        pool); the ragged last chunk runs eagerly."""
        chunk = image_chunk or self.IMAGE_CHUNK
        M = xi.shape[0]
        n_full = M // chunk
        if self.USE_GRAPH and n_full >= 2:
            key = (chunk, float(label), tuple(xi.shape[1:]))
            g = self._graphs.get(key)
            if g is None:
                xs = torch.empty(chunk, *xi.shape[1:], dtype=torch.float32, device=self.dev)
                ys = torch.empty(chunk, *So.shape[1:], dtype=torch.float32, device=self.dev)
                xs.copy_(xi[:chunk])
                self._biases_for(label)                      # per-label constants are built outside the capture
                side = torch.cuda.Stream(device=self.dev)
                side.wait_stream(torch.cuda.current_stream(self.dev))
                with torch.cuda.stream(side):                # warm-up run (cudaFuncSetAttribute etc.)
                    self.forward(xs, label, out=ys)
                torch.cuda.current_stream(self.dev).wait_stream(side)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    self.forward(xs, label, out=ys)
                while len(self._graphs) >= 2:                # each graph owns a private activation pool: keep two
                    self._graphs.pop(next(iter(self._graphs)))
                g = self._graphs[key] = (graph, xs, ys)
            graph, xs, ys = g
            for i in range(n_full):
                xs.copy_(xi[i * chunk:(i + 1) * chunk])
                graph.replay()
                So[i * chunk:(i + 1) * chunk].copy_(ys)
            start = n_full * chunk
        else:
            start = 0
        for s in range(start, M, chunk):
            e = min(M, s + chunk)
            self.forward(xi[s:e], label, out=So[s:e])

    def score_rows(self, x0: torch.Tensor, k_rows: int, row0: int, mean_coeff: float, std: float, label: float, *,
                   seed: int = 0, anchor_id0: int = 0, image_shape=None, image_chunk: Optional[int] = None) -> torch.Tensor:
        """Raw network outputs for perturbation rows [row0, row0 + k_rows) of every anchor -> [A, k_rows, H*W*C]
        (Philox noise keyed by the GLOBAL row, so the rows do not depend on how K is sharded over ranks)."""
        A, D = x0.shape
        c, hh, ww = image_shape if image_shape is not None else self.image_shape
        x = torch.empty(A, k_rows, D, dtype=torch.float32, device=self.dev)
        with _stage("perturb", A * k_rows * D * 4.0):
            _lib.call("idb_perturb_rows", _lib.ptr(x0), None, A, k_rows, D, 0, 0, float(mean_coeff), float(std), seed,
                      anchor_id0, row0, _lib.ptr(x), 0, 1.0, None, None, _lib.stream_ptr())
        S = torch.empty(A, k_rows, hh * ww * self.cout, dtype=torch.float32, device=self.dev)
        xi, So = x.view(A * k_rows, c, hh, ww), S.view(A * k_rows, hh, ww, self.cout)
        with _stage("score_net", 0.0):
            self._run_chunks(xi, So, label, image_chunk)
        return S

    def estimate_row_sharded(self, x0: torch.Tensor, K: int, mean_coeff: float, std: float, label: float, *,
                             seed: int = 0, anchor_id0: int = 0, group=None, image_chunk: Optional[int] = None,
                             gather_rows: bool = False) -> SpectrumResult:
        """One (or a few) huge anchors over all ranks of the process group: rank r evaluates the score
        network on its K / world rows; the centred Gram is assembled from PARTIAL Grams with one all-reduce
        (dim_reduction.sharded_centred_gram; K < D: one all-to-all to column panels first); the eigen-solve and the
        read-out run replicated.  gather_rows=True is the older variant (all-gather of the rows, Gram replicated),
        kept as the cross-check."""
        import torch.distributed as dist
        from .dim_reduction import all_gather_rows, shard_range, sharded_centred_gram
        from .engine import spectrum_from_gram
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        lo, hi = shard_range(K, rank, world)
        S_loc = self.score_rows(x0, hi - lo, lo, mean_coeff, std, label, seed=seed, anchor_id0=anchor_id0,
                                image_chunk=image_chunk)
        if gather_rows:
            with _stage("all_gather_rows", 0.0):
                S = all_gather_rows(S_loc, K, group)
            return spectrum_from_rows(S, K, S.shape[2], -1.0 / std)
        A, _, D = S_loc.shape
        out = []
        for a in range(A):
            with _stage("sharded_gram", 0.0):
                G = sharded_centred_gram(S_loc[a], K, -1.0 / std, group)
            out.append(spectrum_from_gram(G[None], min(K, D)))
        return SpectrumResult(torch.cat([r.singular_values for r in out]), torch.cat([r.dims for r in out]),
                              torch.cat([r.sweeps for r in out]))

    def estimate(self, x0: torch.Tensor, K: int, mean_coeff: float, std: float, label: float, *,
                 z: Optional[torch.Tensor] = None, seed: int = 0, anchor_id0: int = 0, image_shape=None,
                 image_chunk: Optional[int] = None, want_scores: bool = False) -> SpectrumResult:
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
            self._run_chunks(xi, So, label, image_chunk)
        del x
        res = spectrum_from_rows(S, K, hh * ww * self.cout, -1.0 / std)
        if want_scores:
            # models/utils.py:264-267 score = -model(x)/std, back in the reference's NCHW column order
            res.scores = (-(So.view(A, K, hh, ww, self.cout).permute(0, 1, 4, 2, 3)) / std).reshape(A, K, -1)
        return res
