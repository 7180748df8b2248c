"""Synthetic inputs and the full-size image-network specifications of BASELINE configs 4 and 5.

No checkpoint and no dataset exist offline (BASELINE.json: "all runs use synthetic data ... random-init weights"), so the
bench, the tools and the tests build their networks and anchors here.  Everything is regenerated bit-identically from
integer seeds: uniform draws from a seeded torch.Generator in float64, rounded once to float32.

Reference shapes: configs/dimension_estimation/extra_experiments/styleGAN/style_gan_base.py:108-130 (DDPM / NCSN++),
style_gan_BeatGAN.py:29-82 + style_gan_64d_BeatGAN.py:22 (BeatGANs).
"""
from __future__ import annotations

import math
from typing import Dict

import torch

# style_gan_base.py:108-119 (model.name = 'ddpm'): 116.1 M parameters, 79.7 GFLOP per 64x64x3 image
DDPM_FULL = dict(nf=128, ch_mult=(1, 2, 2, 3, 3), num_res_blocks=4, attn_resolutions=(16,), image_size=64,
                 resamp_with_conv=True, centered=False, input_channels=3, output_channels=3, dropout=0.0,
                 conditional=True, nonlinearity="swish")
# the same file with model.name = 'ncsnpp' (BASELINE config 4; style_gan_base.py:120-130): 126.3 M, 88.6 GFLOP
NCSNPP_FULL = dict(DDPM_FULL, fir=True, fir_kernel=[1, 3, 3, 1], skip_rescale=True, resblock_type="biggan",
                   progressive="none", progressive_input="residual", progressive_combine="sum",
                   embedding_type="fourier", init_scale=0.0, fourier_scale=16)
# style_gan_64d_BeatGAN.py (BASELINE config 5): 87.5 M parameters, 37.4 GFLOP per image
BEATGANS_FULL = dict(image_size=64, centered=True, input_channels=3, in_channels=3, model_channels=128, out_channels=3,
                     num_res_blocks=2, num_input_res_blocks=None, embed_channels=64, attention_resolutions=(16,),
                     time_embed_channels=None, dropout=0.3, channel_mult=(1, 1, 2, 3, 4), input_channel_mult=None,
                     conv_resample=True, dims=2, num_classes=None, use_checkpoint=False, num_heads=1,
                     num_head_channels=-1, num_heads_upsample=-1, resblock_updown=True, use_new_attention_order=False,
                     resnet_two_cond=False, resnet_cond_channels=None, resnet_use_zero_module=True,
                     attn_checkpoint=False)
# the paper's synthetic-image configs (squares / gaussian_blobs: 1x32x32, configs/.../squares/10.py:101-135)
SQUARES_FULL = dict(DDPM_FULL, ch_mult=(1, 2, 2, 2), image_size=32, input_channels=1, output_channels=1)
# name -> (registry name, spec, GFLOP per image forward measured on the reference with FlopCounterMode)
FULL_MODELS = {"squares": ("ddpm", SQUARES_FULL, None), "ddpm": ("ddpm", DDPM_FULL, 79.7),
               "ncsnpp": ("ncsnpp", NCSNPP_FULL, 88.6), "beatgans": ("BeatGANsUNetModel", BEATGANS_FULL, 37.4)}


def unet_config(spec: dict, name: str = "ddpm"):
    """A config carrying the keys models/ddpm.py:84-109 reads, for the reference and for iddiff_b200 alike."""
    from .config import ConfigDict
    model = {k: v for k, v in spec.items() if k not in ("image_size", "centered")}
    if name.startswith("BeatGANs"):
        model["image_size"] = spec["image_size"]
    return ConfigDict(model=dict(name=name, sigma_min=1e-2, sigma_max=4.0, beta_min=0.1, beta_max=20.0,
                                 num_scales=1000, checkpoint_path=None, **model),
                      data=dict(effective_image_size=spec["image_size"], image_size=spec["image_size"],
                                centered=spec["centered"], num_channels=spec["input_channels"],
                                shape=[spec["input_channels"], spec["image_size"], spec["image_size"]],
                                use_data_mean=False),
                      training=dict(sde="vesde", batch_size=64, continuous=True))


def make_unet_state(holder, seed: int) -> Dict[str, torch.Tensor]:
    """Deterministic non-degenerate weights for a U-Net holder (the reference's zero-initialised output convolutions,
    appendix A.11, are re-randomised): every tensor of holder.state_dict(), in order, from one seeded generator;
    float64 draws rounded once to float32.  Matrices/filters: U(-b, b), b = sqrt(3 / fan_in); biases U(-0.1, 0.1);
    GroupNorm scale 1 + U(-0.2, 0.2)."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k, v in holder.state_dict().items():
        u = torch.rand(v.shape, generator=g, dtype=torch.float64) * 2 - 1
        if v.dim() >= 2:
            fan_in = v.shape[0] if k.endswith(".W") else v[0].numel()
            if k.endswith("emb_layers.1.weight"):
                fan_in *= 16   # BeatGANs scale-shift projection: keep (1 + scale) well away from 0
            t = u * math.sqrt(3.0 / fan_in)
        elif v.dim() == 1 and _is_norm_weight(holder, k):
            t = 1.0 + 0.2 * u
        else:
            t = 0.1 * u
        out[k] = t.float()
    return out


def _is_norm_weight(holder, key: str) -> bool:
    mod = holder
    parts = key.split(".")[:-1]
    for p in parts:
        mod = getattr(mod, p) if not p.isdigit() else mod[int(p)]
    return isinstance(mod, torch.nn.GroupNorm) and key.endswith("weight")


def smooth_images(n: int, c: int, size: int, seed: int) -> torch.Tensor:
    """Synthetic image manifold in [0, 1]: a seeded 2-D latent -> one Gaussian blob per channel."""
    g = torch.Generator().manual_seed(seed)
    lat = torch.rand(n, 2, generator=g, dtype=torch.float64)
    yy, xx = torch.meshgrid(torch.linspace(0, 1, size, dtype=torch.float64),
                            torch.linspace(0, 1, size, dtype=torch.float64), indexing="ij")
    imgs = []
    for ch in range(c):
        cx = (lat[:, 0] + 0.2 * ch) % 1.0
        cy = (lat[:, 1] + 0.13 * ch) % 1.0
        imgs.append(torch.exp(-((xx[None] - cx[:, None, None]) ** 2 + (yy[None] - cy[:, None, None]) ** 2) / 0.05))
    return torch.stack(imgs, 1).float()


def build_full_holder(which: str, seed: int = 0):
    """Random-init weight holder of a full-size image score network ('ddpm' | 'ncsnpp' | 'beatgans' | 'squares') with
    the reference's zero-initialised output convolutions re-randomised (they make the score
    exactly / nearly zero); returns (holder, spec, GFLOP per image or None)."""
    from . import unet_models
    name, spec, gflop = FULL_MODELS[which]
    torch.manual_seed(seed)
    holder = unet_models.get_model(name)(unet_config(spec, name))
    with torch.no_grad():
        for _, p in holder.named_parameters():
            if p.dim() >= 2 and float(p.abs().max()) < 1e-3:
                unet_models.fan_avg_uniform_(p, 1.0)
    return holder, spec, gflop


def build_full_engine(which: str, device="cuda:0", seed: int = 0):
    """(UNetEngine, parameter count, GFLOP per image) for a full-size network on `device`."""
    from .unet_engine import UNetEngine
    holder, spec, gflop = build_full_holder(which, seed)
    eng = UNetEngine.from_module(holder, device)
    eng.image_shape = (spec["input_channels"], spec["image_size"], spec["image_size"])
    return eng, sum(p.numel() for p in holder.parameters()), gflop
