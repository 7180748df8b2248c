"""Weight holders for the image score networks, registered under the reference's model names.

Like models.FCN they only OWN parameters, in exactly the reference's state_dict layout (so its Lightning
checkpoints load with strict=True), plus a `plan` -- the ordered list of blocks the CUDA engine walks
(unet_engine.py).  The forward pass itself is not here: calling the module raises.

'ddpm' mirrors models/ddpm.py:82-147 (constructor order defines the `all_modules.<i>` indices) with the blocks of
models/layers.py: ResnetBlockDDPM :632-675, AttnBlock :567-590, NIN :555-564, Downsample :607-629, Upsample :593-604,
ddpm_conv3x3 :93-100, default_init :80-83.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .models import get_model, register_model  # noqa: F401


def fan_avg_uniform_(t: torch.Tensor, scale: float = 1.0, in_axis: int = 1, out_axis: int = 0) -> torch.Tensor:
    """layers.py:56-83 `default_init(scale)` = variance_scaling(scale, 'fan_avg', 'uniform'); scale 0 -> 1e-10."""
    scale = 1e-10 if scale == 0 else scale
    shape = t.shape
    receptive = t.numel() / shape[in_axis] / shape[out_axis]
    fan_in, fan_out = shape[in_axis] * receptive, shape[out_axis] * receptive
    bound = math.sqrt(3.0 * scale / ((fan_in + fan_out) / 2.0))
    with torch.no_grad():
        return t.uniform_(-bound, bound)


def _conv3x3(cin, cout, init_scale=1.0, stride=1, padding=1):
    conv = nn.Conv2d(cin, cout, 3, stride=stride, padding=padding)
    fan_avg_uniform_(conv.weight, init_scale)
    nn.init.zeros_(conv.bias)
    return conv


def _dense(cin, cout):
    lin = nn.Linear(cin, cout)
    fan_avg_uniform_(lin.weight)
    nn.init.zeros_(lin.bias)
    return lin


def _gn(ch):
    return nn.GroupNorm(num_groups=32, num_channels=ch, eps=1e-6)


class _Holder(nn.Module):
    gn_eps = 1e-6

    def plan_module(self, idx):
        return self.all_modules[idx]

    def forward(self, *a, **k):
        raise RuntimeError(f"{type(self).__name__} holds weights only; the forward pass runs in unet_engine.UNetEngine "
                           "(no PyTorch/CPU path for this network here)")


class NIN(_Holder):
    def __init__(self, cin, cout, init_scale=0.1):
        super().__init__()
        self.W = nn.Parameter(fan_avg_uniform_(torch.empty(cin, cout), init_scale))
        self.b = nn.Parameter(torch.zeros(cout))


class ResnetBlockDDPM(_Holder):
    def __init__(self, in_ch, out_ch=None, temb_dim=None, conv_shortcut=False, dropout=0.1):
        super().__init__()
        out_ch = out_ch or in_ch
        self.in_ch, self.out_ch, self.conv_shortcut = in_ch, out_ch, conv_shortcut
        self.GroupNorm_0 = _gn(in_ch)
        self.Conv_0 = _conv3x3(in_ch, out_ch)
        if temb_dim is not None:
            self.Dense_0 = _dense(temb_dim, out_ch)
        self.GroupNorm_1 = _gn(out_ch)
        self.Dropout_0 = nn.Dropout(dropout)
        self.Conv_1 = _conv3x3(out_ch, out_ch, init_scale=0.0)
        if in_ch != out_ch:
            if conv_shortcut:
                self.Conv_2 = _conv3x3(in_ch, out_ch)
            else:
                self.NIN_0 = NIN(in_ch, out_ch)


class AttnBlock(_Holder):
    def __init__(self, channels):
        super().__init__()
        self.channels = channels
        self.GroupNorm_0 = _gn(channels)
        self.NIN_0, self.NIN_1, self.NIN_2 = NIN(channels, channels), NIN(channels, channels), NIN(channels, channels)
        self.NIN_3 = NIN(channels, channels, init_scale=0.0)


class Downsample(_Holder):
    def __init__(self, channels, with_conv=False):
        super().__init__()
        self.with_conv = with_conv
        if with_conv:
            self.Conv_0 = _conv3x3(channels, channels, stride=2, padding=0)


class Upsample(_Holder):
    def __init__(self, channels, with_conv=False):
        super().__init__()
        self.with_conv = with_conv
        if with_conv:
            self.Conv_0 = _conv3x3(channels, channels)


@register_model("ddpm")
class DDPM(_Holder):
    """models/ddpm.py:82-147.  `plan` entries: (kind, module index, extra) in execution order (ddpm.py:149-213)."""

    def __init__(self, config):
        super().__init__()
        m = config.model
        if m.nonlinearity.lower() != "swish":
            raise NotImplementedError("iddiff_b200 DDPM engine: only nonlinearity='swish' (every shipped config) is built")
        self.nf = nf = m.nf
        ch_mult = tuple(m.ch_mult)
        self.num_res_blocks = nrb = m.num_res_blocks
        self.attn_resolutions = tuple(m.attn_resolutions)
        self.num_resolutions = nres = len(ch_mult)
        self.all_resolutions = [config.data.effective_image_size // (2 ** i) for i in range(nres)]
        self.conditional = m.conditional
        self.centered = config.data.centered
        if not self.conditional:
            raise NotImplementedError("DDPM with conditional=False (no time embedding) is not built")
        temb_dim = 4 * nf
        mods, plan = [], []

        def add(kind, mod, **extra):
            plan.append((kind, len(mods), extra))
            mods.append(mod)

        mods.append(_dense(nf, temb_dim))
        mods.append(_dense(temb_dim, temb_dim))
        add("conv_in", _conv3x3(m.input_channels, nf))
        plan.append(("push", -1, {}))
        hs_c, in_ch = [nf], nf
        for lvl in range(nres):
            for _ in range(nrb):
                out_ch = nf * ch_mult[lvl]
                add("res", ResnetBlockDDPM(in_ch, out_ch, temb_dim, dropout=m.dropout))
                in_ch = out_ch
                if self.all_resolutions[lvl] in self.attn_resolutions:
                    add("attn", AttnBlock(in_ch))
                plan.append(("push", -1, {}))
                hs_c.append(in_ch)
            if lvl != nres - 1:
                add("down", Downsample(in_ch, m.resamp_with_conv))
                plan.append(("push", -1, {}))
                hs_c.append(in_ch)
        in_ch = hs_c[-1]
        add("res", ResnetBlockDDPM(in_ch, None, temb_dim, dropout=m.dropout))
        add("attn", AttnBlock(in_ch))
        add("res", ResnetBlockDDPM(in_ch, None, temb_dim, dropout=m.dropout))
        for lvl in reversed(range(nres)):
            for _ in range(nrb + 1):
                out_ch = nf * ch_mult[lvl]
                add("res", ResnetBlockDDPM(in_ch + hs_c.pop(), out_ch, temb_dim, dropout=m.dropout), skip=True)
                in_ch = out_ch
            if self.all_resolutions[lvl] in self.attn_resolutions:
                add("attn", AttnBlock(in_ch))
            if lvl != 0:
                add("up", Upsample(in_ch, m.resamp_with_conv))
        assert not hs_c
        add("gn_out", _gn(in_ch))
        add("conv_out", _conv3x3(in_ch, m.output_channels, init_scale=0.0))
        self.all_modules = nn.ModuleList(mods)
        self.plan = plan


# ------------------------------------------------------------------------------------------------
# NCSN++ (models/ncsnpp.py:40-236) in the configuration BASELINE config 4 names: style_gan_base.py:106-129 with
# model.name='ncsnpp' -- Fourier embedding, BigGAN res-blocks with FIR up/down-sampling, skip rescale,
# progressive_input='residual' (input pyramid), progressive='none'.
# Blocks: layerspp.py GaussianFourierProjection :32-41, AttnBlockpp :62-91, Downsample (FIR conv) :129-163,
# ResnetBlockBigGANpp :212-274; up_or_down_sampling.py Conv2d :23-57.
# ------------------------------------------------------------------------------------------------
def _gnpp(ch):
    return nn.GroupNorm(num_groups=min(ch // 4, 32), num_channels=ch, eps=1e-6)


class GaussianFourierProjection(_Holder):
    def __init__(self, embedding_size=256, scale=1.0):
        super().__init__()
        self.W = nn.Parameter(torch.randn(embedding_size) * scale, requires_grad=False)


class ResnetBlockBigGANpp(_Holder):
    def __init__(self, in_ch, out_ch=None, temb_dim=None, up=False, down=False, dropout=0.1, fir=False,
                 fir_kernel=(1, 3, 3, 1), skip_rescale=True, init_scale=0.0):
        super().__init__()
        out_ch = out_ch or in_ch
        self.in_ch, self.out_ch, self.up, self.down = in_ch, out_ch, up, down
        self.fir, self.fir_kernel, self.skip_rescale = fir, tuple(fir_kernel), skip_rescale
        self.GroupNorm_0 = _gnpp(in_ch)
        self.Conv_0 = _conv3x3(in_ch, out_ch)
        if temb_dim is not None:
            self.Dense_0 = _dense(temb_dim, out_ch)
        self.GroupNorm_1 = _gnpp(out_ch)
        self.Dropout_0 = nn.Dropout(dropout)
        self.Conv_1 = _conv3x3(out_ch, out_ch, init_scale=init_scale)
        if in_ch != out_ch or up or down:
            self.Conv_2 = nn.Conv2d(in_ch, out_ch, 1)
            fan_avg_uniform_(self.Conv_2.weight)
            nn.init.zeros_(self.Conv_2.bias)


class AttnBlockpp(_Holder):
    def __init__(self, channels, skip_rescale=False, init_scale=0.0):
        super().__init__()
        self.channels, self.skip_rescale = channels, skip_rescale
        self.GroupNorm_0 = _gnpp(channels)
        self.NIN_0, self.NIN_1, self.NIN_2 = NIN(channels, channels), NIN(channels, channels), NIN(channels, channels)
        self.NIN_3 = NIN(channels, channels, init_scale=init_scale)


class _FirConv2d(_Holder):
    """up_or_down_sampling.Conv2d parameters (weight [out, in, k, k], bias)."""

    def __init__(self, in_ch, out_ch, kernel=3):
        super().__init__()
        self.weight = nn.Parameter(fan_avg_uniform_(torch.empty(out_ch, in_ch, kernel, kernel)))
        self.bias = nn.Parameter(torch.zeros(out_ch))


class DownsampleFirConv(_Holder):
    """layerspp.Downsample(fir=True, with_conv=True): FIR pre-filter + stride-2 3x3 conv (conv_downsample_2d)."""

    def __init__(self, in_ch, out_ch, fir_kernel=(1, 3, 3, 1)):
        super().__init__()
        self.fir_kernel = tuple(fir_kernel)
        self.Conv2d_0 = _FirConv2d(in_ch, out_ch)


@register_model("ncsnpp")
class NCSNpp(_Holder):
    def __init__(self, config):
        super().__init__()
        m = config.model
        want = dict(nonlinearity="swish", embedding_type="fourier", resblock_type="biggan", progressive="none",
                    progressive_input="residual")
        for k, v in want.items():
            if str(getattr(m, k)).lower() != v:
                raise NotImplementedError(f"iddiff_b200 NCSN++ engine is built for model.{k}={v!r} (style_gan_base.py), "
                                          f"got {getattr(m, k)!r}")
        if not (m.fir and m.conditional and config.training.continuous):
            raise NotImplementedError("NCSN++ engine: fir=True, conditional=True, continuous training only")
        self.nf = nf = m.nf
        ch_mult = tuple(m.ch_mult)
        nrb = m.num_res_blocks
        self.attn_resolutions = tuple(m.attn_resolutions)
        nres = len(ch_mult)
        res_all = [config.data.effective_image_size // (2 ** i) for i in range(nres)]
        self.centered = config.data.centered
        self.skip_rescale = m.skip_rescale
        self.embedding_type = "fourier"
        channels = config.data.num_channels
        temb_dim = 4 * nf
        mods, plan = [], []

        def add(kind, mod, **extra):
            plan.append((kind, len(mods), extra))
            mods.append(mod)

        def res(i, o=None, **kw):
            return ResnetBlockBigGANpp(i, o, temb_dim, dropout=m.dropout, fir=m.fir, fir_kernel=m.fir_kernel,
                                       skip_rescale=m.skip_rescale, init_scale=m.init_scale, **kw)

        attn = lambda c: AttnBlockpp(c, skip_rescale=m.skip_rescale, init_scale=m.init_scale)
        mods.append(GaussianFourierProjection(embedding_size=nf, scale=m.fourier_scale))
        mods.append(_dense(2 * nf, temb_dim))
        mods.append(_dense(temb_dim, temb_dim))
        add("conv_in", _conv3x3(channels, nf), pyramid=True)
        plan.append(("push", -1, {}))
        hs_c, in_ch, pyr_ch = [nf], nf, channels
        for lvl in range(nres):
            for _ in range(nrb):
                out_ch = nf * ch_mult[lvl]
                add("res", res(in_ch, out_ch))
                in_ch = out_ch
                if res_all[lvl] in self.attn_resolutions:
                    add("attn", attn(in_ch))
                plan.append(("push", -1, {}))
                hs_c.append(in_ch)
            if lvl != nres - 1:
                add("res", res(in_ch, down=True))
                add("pyr_down", DownsampleFirConv(pyr_ch, in_ch, m.fir_kernel))
                pyr_ch = in_ch
                plan.append(("push", -1, {}))
                hs_c.append(in_ch)
        in_ch = hs_c[-1]
        add("res", res(in_ch))
        add("attn", attn(in_ch))
        add("res", res(in_ch))
        for lvl in reversed(range(nres)):
            for _ in range(nrb + 1):
                out_ch = nf * ch_mult[lvl]
                add("res", res(in_ch + hs_c.pop(), out_ch), skip=True)
                in_ch = out_ch
            if res_all[lvl] in self.attn_resolutions:
                add("attn", attn(in_ch))
            if lvl != 0:
                add("res", res(in_ch, up=True))
        assert not hs_c
        add("gn_out", _gnpp(in_ch))
        add("conv_out", _conv3x3(in_ch, channels, init_scale=m.init_scale))
        self.all_modules = nn.ModuleList(mods)
        self.plan = plan


# ------------------------------------------------------------------------------------------------
# BeatGANs U-Net (models/BeatGANsUNET.py:18-284; blocks models/BeatGANsblocks.py: ResBlock :83-255,
# apply_conditions :258-332, Upsample/Downsample :335-396, AttentionBlock :399-451, QKVAttentionLegacy :468-491;
# models/BeatGANs_nn.py: GroupNorm32 :23-25, normalization :98-104, timestep_embedding :107-125, zero_module).
# Parameter names follow the reference's nn.Sequential / ModuleList nesting (input_blocks.<k>.<j>.in_layers.0.weight ...).
# ------------------------------------------------------------------------------------------------
def _gn32(ch):
    return nn.GroupNorm(min(32, ch), ch)  # GroupNorm32: default eps 1e-5


class BeatResBlock(_Holder):
    def __init__(self, channels, emb_channels, dropout, out_channels=None, up=False, down=False, has_lateral=False,
                 use_zero_module=True):
        super().__init__()
        out_channels = out_channels or channels
        self.in_ch, self.out_ch, self.up, self.down, self.has_lateral = channels, out_channels, up, down, has_lateral
        self.in_layers = nn.Sequential(_gn32(channels), nn.SiLU(), nn.Conv2d(channels, out_channels, 3, padding=1))
        self.emb_layers = nn.Sequential(nn.SiLU(), nn.Linear(emb_channels, 2 * out_channels))
        conv = nn.Conv2d(out_channels, out_channels, 3, padding=1)
        if use_zero_module:
            for p in conv.parameters():
                p.detach().zero_()
        self.out_layers = nn.Sequential(_gn32(out_channels), nn.SiLU(), nn.Dropout(p=dropout), conv)
        self.skip_connection = nn.Identity() if out_channels == channels else nn.Conv2d(channels, out_channels, 1)


class BeatAttentionBlock(_Holder):
    def __init__(self, channels, num_heads=1, num_head_channels=-1):
        super().__init__()
        self.channels = channels
        self.num_heads = num_heads if num_head_channels == -1 else channels // num_head_channels
        self.norm = _gn32(channels)
        self.qkv = nn.Conv1d(channels, channels * 3, 1)
        self.proj_out = nn.Conv1d(channels, channels, 1)
        for p in self.proj_out.parameters():
            p.detach().zero_()


class _Seq(nn.Sequential):
    def forward(self, *a, **k):
        raise RuntimeError("weight holder only; the forward pass runs in unet_engine.UNetEngine")


@register_model("BeatGANsUNetModel")
class BeatGANsUNetModel(_Holder):
    gn_eps = 1e-5

    def __init__(self, config):
        super().__init__()
        c = config.model
        if not c.resblock_updown or c.num_classes is not None or c.resnet_two_cond or c.use_new_attention_order:
            raise NotImplementedError("BeatGANs engine: resblock_updown=True, no classes, one condition, legacy attention "
                                      "order (style_gan_BeatGAN.py) only")
        if c.dims != 2 or c.input_channel_mult or c.num_input_res_blocks:
            raise NotImplementedError("BeatGANs engine: dims=2, symmetric channel_mult / num_res_blocks only")
        self.nf = self.time_emb_channels = c.time_embed_channels or c.model_channels
        E = c.embed_channels
        self.centered = True          # no 2x - 1 in BeatGANsUNET.py:257
        self.embedding_type = "beatgans"
        self.time_embed = _Seq(nn.Linear(self.time_emb_channels, E), nn.SiLU(), nn.Linear(E, E))
        mult = tuple(c.channel_mult)
        heads_up = c.num_heads if c.num_heads_upsample == -1 else c.num_heads_upsample
        mods, plan = [], []

        def add(kind, mod, **extra):
            plan.append((kind, len(mods), extra))
            mods.append(mod)

        def res(ch, out=None, **kw):
            return BeatResBlock(ch, E, c.dropout, out_channels=out, use_zero_module=c.resnet_use_zero_module, **kw)

        ch = input_ch = int(mult[0] * c.model_channels)
        conv_in = nn.Conv2d(c.in_channels, ch, 3, padding=1)
        self.input_blocks = nn.ModuleList([_Seq(conv_in)])
        add("conv_in", conv_in)
        plan.append(("push", -1, {}))
        chans = [ch]
        resolution = c.image_size
        for level, m_ in enumerate(mult):
            for _ in range(c.num_res_blocks):
                layers = [res(ch, int(m_ * c.model_channels))]
                add("bres", layers[0])
                ch = int(m_ * c.model_channels)
                if resolution in c.attention_resolutions:
                    layers.append(BeatAttentionBlock(ch, c.num_heads, c.num_head_channels))
                    add("battn", layers[-1])
                self.input_blocks.append(_Seq(*layers))
                plan.append(("push", -1, {}))
                chans.append(ch)
            if level != len(mult) - 1:
                resolution //= 2
                blk = res(ch, ch, down=True)
                self.input_blocks.append(_Seq(blk))
                add("bres", blk)
                plan.append(("push", -1, {}))
                chans.append(ch)
        mid = [res(ch), BeatAttentionBlock(ch, c.num_heads, c.num_head_channels), res(ch)]
        self.middle_block = _Seq(*mid)
        add("bres", mid[0])
        add("battn", mid[1])
        add("bres", mid[2])
        self.output_blocks = nn.ModuleList([])
        for level, m_ in list(enumerate(mult))[::-1]:
            for i in range(c.num_res_blocks + 1):
                ich = chans.pop()
                layers = [res(ch + ich, int(c.model_channels * m_), has_lateral=True)]
                add("bres", layers[0], skip=True)
                ch = int(c.model_channels * m_)
                if resolution in c.attention_resolutions:
                    layers.append(BeatAttentionBlock(ch, heads_up, c.num_head_channels))
                    add("battn", layers[-1])
                if level and i == c.num_res_blocks:
                    resolution *= 2
                    layers.append(res(ch, ch, up=True))
                    add("bres", layers[-1])
                self.output_blocks.append(_Seq(*layers))
        assert not chans
        conv_out = nn.Conv2d(input_ch, c.out_channels, 3, padding=1)
        if c.resnet_use_zero_module:
            for p in conv_out.parameters():
                p.detach().zero_()
        self.out = _Seq(_gn32(ch), nn.SiLU(), conv_out)
        add("gn_out", self.out[0])
        add("conv_out", conv_out)
        self._plan_modules = mods   # plain list: the modules are registered under the reference's names above
        self.plan = plan

    def plan_module(self, idx):
        return self._plan_modules[idx]
