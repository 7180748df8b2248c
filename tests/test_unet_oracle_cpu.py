"""CPU: the torch restatement of the DDPM U-Net path (oracle/img_oracle.py) against the goldens the REAL reference
produced (tests/golden/unet_ddpm_small.npz via oracle/gen_golden_unet.py), and the weight holder's state_dict layout."""
import numpy as np
import torch

from oracle import fixtures as FX
from oracle import img_oracle as IO
from oracle import id_oracle as O


import pytest

MODELS = {"ddpm": ("unet_ddpm_small", FX.DDPM_SMALL, IO.ddpm_forward),
          "ddpm_gray": ("unet_ddpm_gray_small", FX.DDPM_GRAY_SMALL, IO.ddpm_forward),   # the paper's 1x32x32 configs
          "ncsnpp": ("unet_ncsnpp_small", FX.NCSNPP_SMALL, IO.ncsnpp_forward),
          "BeatGANsUNetModel": ("unet_beatgans_small", FX.BEATGANS_SMALL, IO.beatgans_forward)}


def _name(model):
    return "ddpm" if model == "ddpm_gray" else model


def _shape(spec):
    return (spec["input_channels"], spec["image_size"], spec["image_size"])


def _setup(model="ddpm"):
    from iddiff_b200 import unet_models
    gold, spec, _ = MODELS[model]
    g = FX.golden(gold)
    holder = unet_models.get_model(_name(model))(FX.unet_config(spec, _name(model)))
    state = FX.make_unet_state(holder, int(g["seed"]))
    wsum = sum(v.double().abs().sum().item() for v in state.values())
    assert abs(wsum - float(g["w_checksum"])) < 1e-6 * wsum, "fixture weights drifted from the ones the reference ran"
    return g, holder, state


def test_holder_state_dict_layout_and_plan():
    g, holder, state = _setup()
    holder.load_state_dict(state, strict=True)
    kinds = [k for k, _, _ in holder.plan]
    assert kinds.count("res") == 8 and kinds.count("attn") == 3 and kinds[0] == "conv_in" and kinds[-1] == "conv_out"
    try:
        holder(torch.zeros(1, 3, 16, 16), torch.zeros(1))
    except RuntimeError as e:
        assert "holds weights only" in str(e)
    else:
        raise AssertionError("the holder must not have a PyTorch forward path")


def test_ncsnpp_holder_state_dict_layout():
    g, holder, state = _setup("ncsnpp")
    holder.load_state_dict(state, strict=True)   # the REAL reference loaded the same dict strict=True (gen_golden_unet)
    kinds = [k for k, _, _ in holder.plan]
    assert kinds.count("pyr_down") == 1 and kinds.count("res") == 10 and "all_modules.0.W" in state


@pytest.mark.parametrize("model", ["ddpm", "ddpm_gray", "ncsnpp", "BeatGANsUNetModel"])
def test_unet_forward_oracle_matches_reference(model):
    g, _, state = _setup(model)
    _, spec, fwd = MODELS[model]
    x = torch.from_numpy(g["x"])
    lab = torch.full((x.shape[0],), float(g["label"]))
    with torch.no_grad():
        y = fwd(state, spec, x, lab)
    scale = np.abs(g["y"]).max()
    assert np.abs(y.numpy() - g["y"]).max() < 1e-5 * scale


@pytest.mark.parametrize("model", ["ddpm", "ddpm_gray", "ncsnpp", "BeatGANsUNetModel"])
def test_unet_path_oracle_matches_reference_spectrum(model):
    g, _, state = _setup(model)
    _, spec, fwd = MODELS[model]
    B, K, nb = int(g["B"]), int(g["K"]), int(g["nb"])
    x0 = torch.from_numpy(g["x0"])
    z = FX.seeded_noise(int(g["seed"]) + 3, x0.shape[0], nb, B, x0.shape[1:])
    assert abs(FX.seeded_noise_checksum(z) - float(g["z_checksum"])) < 1e-6 * abs(float(g["z_checksum"])) + 1e-3
    t = torch.ones(1) * 1e-5
    std = float(O.ve_marginal(torch.zeros(1, 1), t, 1e-2, 4.0)[1][0])
    a = 0
    za = z[a].reshape(nb * B, *x0.shape[1:])[:K]
    S = IO.unet_score_matrix(fwd, state, spec, x0[a], za, 1.0, std, float(g["label"]))
    s = torch.linalg.svdvals(S).double().numpy()
    np.testing.assert_allclose(s, g["singular_values"][a], rtol=1e-3)
    assert O.id_from_spectrum(s.tolist()) == int(g["dims"][a])


@pytest.mark.parametrize("which", ["beatgans", "ddpm", "ncsnpp"])
def test_oracle_forward_at_full_size_against_the_real_reference(which):
    """The oracle restatement at the size BASELINE configs 4-5 name (64x64x3, 87-126 M parameters) against the output of
    the REAL reference module stored in tests/golden/unet_*_full.npz (oracle/gen_golden_unet_full.py)."""
    import numpy as np
    import torch
    from oracle import fixtures as FX, img_oracle as IO
    from iddiff_b200 import synthetic as SY, unet_models
    g = FX.golden(f"unet_{which}_full")
    name, spec, _ = SY.FULL_MODELS[which]
    holder = unet_models.get_model(name)(SY.unet_config(spec, name))
    state = SY.make_unet_state(holder, int(g["seed"]))
    fwd = {"ddpm": IO.ddpm_forward, "ncsnpp": IO.ncsnpp_forward, "beatgans": IO.beatgans_forward}[which]
    x = torch.from_numpy(g["x"][:1])
    with torch.no_grad():
        y = fwd(state, spec, x, torch.full((1,), float(g["label"]))).numpy()
    scale = np.abs(g["y"]).max()
    assert np.abs(y - g["y"][:1]).max() < 1e-5 * scale
