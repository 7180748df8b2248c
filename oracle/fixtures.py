"""TEST INFRASTRUCTURE ONLY -- deterministic builders shared by oracle/gen_golden.py (which runs
the real reference on them) and tests/ (which replay them through the CUDA path and the oracle).

Everything here is regenerated bit-identically from integer seeds: uniform draws come from a
seeded torch.Generator (integer Philox/MT streams), and every floating-point reduction is done
in float64 and rounded once to float32, so different host CPUs (AVX2 vs AVX-512 summation order)
produce the same float32 weights.
"""
from __future__ import annotations

import math
from typing import Dict, List

import numpy as np
import torch


def fcn_layer_shapes(D: int, H: int, L: int):
    """Linear layer (out, in) shapes of models/fcn.py:16-30 and their Sequential indices
    (Linear, Dropout, ELU triples -> Linear modules sit at mlp.0, mlp.3, ...)."""
    shapes = [(H, D + 1)] + [(H, H)] * L + [(D, H)]
    keys = [3 * i for i in range(L + 2)]
    return shapes, keys


def make_fcn_state(D: int, H: int, L: int, seed: int) -> Dict[str, torch.Tensor]:
    """nn.Linear-style init, U(-1/sqrt(fan_in), 1/sqrt(fan_in)), from a private generator."""
    g = torch.Generator().manual_seed(seed)
    shapes, keys = fcn_layer_shapes(D, H, L)
    state = {}
    for (o, i), k in zip(shapes, keys):
        bound = 1.0 / math.sqrt(i)
        state[f"mlp.{k}.weight"] = ((torch.rand(o, i, generator=g, dtype=torch.float64) * 2 - 1) * bound).float()
        state[f"mlp.{k}.bias"] = ((torch.rand(o, generator=g, dtype=torch.float64) * 2 - 1) * bound).float()
    return state


def _orthogonal(n: int, g: torch.Generator) -> torch.Tensor:
    a = torch.randn(n, n, generator=g, dtype=torch.float64)
    q, r = torch.linalg.qr(a)
    return q * torch.sign(torch.diagonal(r))[None, :]


def make_gap_fcn_state(D: int, H: int, L: int, seed: int, gap_rank: int, gap: float = 1e-3):
    """A dense FCN with a *decisive* spectral gap, built without training or stored checkpoints.

    Hidden layers are dense random orthogonal matrices R_i with biases in [-0.25, 0.75] (so roughly
    a third of the ELU units sit in their exponential branch), the input layer is gain*R_0[:, :D]
    plus a random time column, and the output layer is P R_0[:, :D]^T R_1^T ... R_L^T / gain with
    P = Q diag(d) Q^T, d = linspace(1, .7) on D-gap_rank directions and `gap`*linspace(1, .5) on the
    remaining gap_rank ones.  The Jacobian at any anchor is then ~ c * P: D-gap_rank nearly flat
    singular values, a cliff of 1/gap, then gap_rank small ones -- the shape a trained score network
    produces on a (gap_rank-1)-sphere (figures/paper/10_sphere_spectrum.png); the reference's
    read-out returns gap_rank.  All construction is float64 from a seeded generator, rounded once."""
    g = torch.Generator().manual_seed(seed)
    Rs = [_orthogonal(H, g) for _ in range(L + 1)]
    gain = 0.5 * math.sqrt(H)
    rb = lambda n: torch.rand(n, generator=g, dtype=torch.float64) - 0.25
    _, keys = fcn_layer_shapes(D, H, L)
    state = {}
    tcol = torch.rand(H, 1, generator=g, dtype=torch.float64) - 0.5
    state[f"mlp.{keys[0]}.weight"] = torch.cat([gain * Rs[0][:, :D], tcol], 1)
    state[f"mlp.{keys[0]}.bias"] = rb(H)
    for i in range(1, L + 1):
        state[f"mlp.{keys[i]}.weight"] = Rs[i]
        state[f"mlp.{keys[i]}.bias"] = rb(H)
    v = torch.rand(D, generator=g, dtype=torch.float64) - 0.5
    Q = torch.eye(D, dtype=torch.float64) - 2.0 * torch.outer(v, v) / torch.dot(v, v)
    d = torch.cat([torch.linspace(1.0, 0.7, D - gap_rank, dtype=torch.float64),
                   gap * torch.linspace(1.0, 0.5, gap_rank, dtype=torch.float64)])
    P = Q @ torch.diag(d) @ Q.T
    A = Rs[0][:, :D].T.clone()
    for i in range(1, L + 1):
        A = A @ Rs[i].T
    state[f"mlp.{keys[-1]}.weight"] = P @ A / gain
    state[f"mlp.{keys[-1]}.bias"] = torch.rand(D, generator=g, dtype=torch.float64) - 0.5
    return {k: v.float().contiguous() for k, v in state.items()}


def fcn_state_to_lists(state: Dict[str, torch.Tensor], L: int):
    _, keys = fcn_layer_shapes(0, 0, L)
    ws = [state[f"mlp.{k}.weight"] for k in keys]
    bs = [state[f"mlp.{k}.bias"] for k in keys]
    return ws, bs


def seeded_noise(seed: int, n_anchors: int, num_batches: int, B: int, shape) -> torch.Tensor:
    """The noise the reference run drew: torch.manual_seed(seed) immediately before the first
    randn_like of the hot loop (gen_golden.run_reference re-seeds there), then one [B, *shape]
    draw per perturbation batch, anchors in order."""
    torch.manual_seed(seed)
    proto = torch.empty(B, *shape)
    return torch.stack([torch.stack([torch.randn_like(proto) for _ in range(num_batches)])
                        for _ in range(n_anchors)])


def seeded_noise_checksum(z: torch.Tensor) -> float:
    return float(z.double().sum().item())


def golden(name: str):
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", name + ".npz")
    return np.load(path, allow_pickle=False)


def golden_fcn_state(gold) -> Dict[str, torch.Tensor]:
    """Weights of an FCN fixture: stored arrays if present, else regenerated from the seed."""
    keys = [k for k in gold.files if k.startswith("w::")]
    if keys:
        return {k[3:]: torch.from_numpy(gold[k]) for k in keys}
    return make_gap_fcn_state(int(gold["D"]), int(gold["H"]), int(gold["L"]), seed=int(gold["seed"]),
                              gap_rank=int(gold["gap_rank"]))


# ------------------------------------------------------------------------------------------------
# U-Net fixtures
# ------------------------------------------------------------------------------------------------
DDPM_SMALL = dict(nf=32, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=(8,), image_size=16, resamp_with_conv=True,
                  centered=False, input_channels=3, output_channels=3, dropout=0.0, conditional=True,
                  nonlinearity="swish")


# the paper's image configs (squares / gaussian_blobs / MNIST: one channel, 32x32, ch_mult (1,2,2,2), attention at 16;
# configs/dimension_estimation/paper/image_data/squares/10.py:101-135) at a reduced width
DDPM_GRAY_SMALL = dict(DDPM_SMALL, nf=32, ch_mult=(1, 2, 2, 2), num_res_blocks=2, attn_resolutions=(16,), image_size=32,
                       input_channels=1, output_channels=1)


NCSNPP_SMALL = dict(DDPM_SMALL, nf=16, fir=True, fir_kernel=[1, 3, 3, 1], skip_rescale=True, resblock_type="biggan",
                    progressive="none", progressive_input="residual", progressive_combine="sum",
                    embedding_type="fourier", init_scale=0.0, fourier_scale=16)


BEATGANS_SMALL = dict(image_size=16, centered=True, input_channels=3, in_channels=3, model_channels=32, out_channels=3,
                      num_res_blocks=1, num_input_res_blocks=None, embed_channels=8, attention_resolutions=(8,),
                      time_embed_channels=None, dropout=0.1, channel_mult=(1, 2), input_channel_mult=None,
                      conv_resample=True, dims=2, num_classes=None, use_checkpoint=False, num_heads=2,
                      num_head_channels=-1, num_heads_upsample=-1, resblock_updown=True, use_new_attention_order=False,
                      resnet_two_cond=False, resnet_cond_channels=None, resnet_use_zero_module=True,
                      attn_checkpoint=False)


from iddiff_b200.synthetic import unet_config, make_unet_state, smooth_images  # noqa: E402,F401  (shared with tools/ and bench.py)
