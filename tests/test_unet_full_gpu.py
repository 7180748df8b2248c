"""GPU parity of the FULL-SIZE image score networks -- the sizes BASELINE configs 4 and 5 name (64x64x3, D = 12288,
nf / model_channels = 128: DDPM 116.1 M, NCSN++ 126.3 M, BeatGANs 87.5 M parameters) -- against goldens produced by the
REAL reference modules (oracle/gen_golden_unet_full.py -> tests/golden/unet_*_full.npz).

One tolerance rule: raw network output within 2e-5 of its largest magnitude and within 8x the fp32
reference's own distance to the fp64 oracle; singular values rtol 1e-3 per value plus ONE times the reference's rounding
floor max_i |s_ref - s_fp64|; integer ID exact; raw scores 2e-4 of the largest score.  The centred K x D matrix with
K < D has rank K - 1: its last singular value is structurally zero (the reference returns rounding noise there), so it
is compared absolutely and excluded from the floor."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fixtures as FX  # noqa: E402
from oracle import id_oracle as O  # noqa: E402
from iddiff_b200 import synthetic as SY  # noqa: E402

WHICH = pytest.mark.parametrize("which", ["beatgans", "ddpm", "ncsnpp"])
_CACHE = {}


def _engine(which):
    """One engine per model for the whole module (weights regenerated from the golden's seed, checksum-guarded)."""
    if which in _CACHE:
        return _CACHE[which]
    _CACHE.clear()
    torch.cuda.empty_cache()
    from iddiff_b200 import unet_models
    from iddiff_b200.unet_engine import UNetEngine
    name, spec, _ = SY.FULL_MODELS[which]
    g = FX.golden(f"unet_{which}_full")
    holder = unet_models.get_model(name)(SY.unet_config(spec, name))
    state = SY.make_unet_state(holder, int(g["seed"]))
    wsum = sum(v.double().abs().sum().item() for v in state.values())
    assert abs(wsum - float(g["w_checksum"])) <= 1e-9 * float(g["w_checksum"])
    assert sum(v.numel() for k, v in holder.state_dict().items()) >= int(g["n_params"])
    holder.load_state_dict(state, strict=True)
    eng = UNetEngine.from_module(holder, "cuda:0")
    eng.image_shape = (spec["input_channels"], spec["image_size"], spec["image_size"])
    _CACHE[which] = (g, eng)
    return g, eng


@WHICH
def test_full_size_forward_vs_reference(which):
    g, eng = _engine(which)
    x = torch.from_numpy(g["x"]).cuda()
    got = eng.forward(x, float(g["label"])).permute(0, 3, 1, 2).cpu().numpy()
    assert np.isfinite(got).all()
    scale = np.abs(g["y"]).max()
    err = np.abs(got - g["y"]).max()
    assert err < 2e-5 * scale, (err, scale)
    ref_err = np.abs(g["y"] - g["y64"]).max()
    eng_err = np.abs(got - g["y64"]).max()
    print(f"{which}: |engine - ref| {err:.3e}  |engine - fp64| {eng_err:.3e}  |ref - fp64| {ref_err:.3e}  scale {scale:.3f}")
    assert eng_err < max(8 * ref_err, 2e-6 * scale), (eng_err, ref_err)


@WHICH
def test_full_size_path_k256_vs_reference(which):
    """dim_reduction.py:177-199 on one anchor, K = 256 rows (row-side Gram, n = 256), the reference's own noise."""
    g, eng = _engine(which)
    K, B, seed = int(g["K"]), int(g["B"]), int(g["seed"])
    x0 = torch.from_numpy(g["x0"])
    torch.manual_seed(seed + 3)
    proto = torch.empty(B, *x0.shape)
    z = torch.cat([torch.randn_like(proto) for _ in range(K // B)])
    assert abs(float(z.double().sum()) - float(g["z_checksum"])) < 1e-6
    res = eng.estimate(x0.reshape(1, -1).cuda(), K, 1.0, float(g["std"]), float(g["label"]),
                       z=z.reshape(1, K, -1).cuda(), want_scores=True)
    torch.cuda.synchronize()
    sv = res.singular_values[0].cpu().numpy().astype(np.float64)
    want, s64 = g["singular_values"], g["s64"]
    m = K - 1
    floor = np.abs(want[:m] - s64[:m]).max()
    rel_ref = np.abs(sv[:m] / want[:m] - 1).max()
    rel_64 = np.abs(sv[:m] / s64[:m] - 1).max()
    print(f"{which}: spectra vs reference {rel_ref:.3e}, vs fp64 {rel_64:.3e}; reference vs fp64 "
          f"{np.abs(want[:m] / s64[:m] - 1).max():.3e}; floor {floor:.3e}")
    assert np.all(np.abs(sv[:m] - want[:m]) <= 1e-3 * want[:m] + floor)
    assert np.all(np.abs(sv[:m] - s64[:m]) <= 1e-3 * s64[:m] + floor)
    assert abs(sv[-1]) < 1e-4 * want[0]
    assert res.dims.cpu().tolist() == g["dims"].tolist() == [O.id_from_spectrum(want.tolist())]
    head = res.scores[0, :4].reshape(4, *x0.shape).cpu().numpy()
    scale = np.abs(g["scores_head"]).max()
    assert np.abs(head - g["scores_head"]).max() < 2e-4 * scale


def test_release_engines():
    _CACHE.clear()
    torch.cuda.empty_cache()


def test_config5_true_size_whole_path_vs_the_reference_driver():
    """BASELINE config 5 at its TRUE size through the reference's own driver: golden = the REAL
    dim_reduction.get_manifold_dimension (oracle/gen_golden_c5_full.py; full-size BeatGANs, loader batch 128 -> the
    reference's K schedule gives K = 16768, a 16768 x 12288 score matrix, LAPACK full SVD, ~40 min on 8 cores).  Here:
    the drop-in path on the reference's own noise -- perturb, 16768 forwards, D-side fp64 Gram, the n = 12288 whole-GPU
    eigen-solver (two column classes), read-out.  All 12288 singular values rtol 1e-3 (+ 1e-6 s_0: no fp64 oracle of this
    size exists to measure the reference's own floor), integer ID exact."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c5_beatgans_k16768.npz")
    if not os.path.exists(path):
        pytest.skip("tests/golden/c5_beatgans_k16768.npz not generated")
    from iddiff_b200 import unet_models
    from iddiff_b200.unet_engine import UNetEngine
    _CACHE.clear()
    torch.cuda.empty_cache()
    g = np.load(path)
    name, spec, _ = SY.FULL_MODELS["beatgans"]
    holder = unet_models.get_model(name)(SY.unet_config(spec, name))
    state = SY.make_unet_state(holder, int(g["seed"]))
    wsum = sum(v.double().abs().sum().item() for v in state.values())
    assert abs(wsum - float(g["w_checksum"])) <= 1e-9 * float(g["w_checksum"])
    holder.load_state_dict(state, strict=True)
    eng = UNetEngine.from_module(holder, "cuda:0")
    eng.image_shape = (3, 64, 64)
    K, B, draws = int(g["K"]), int(g["B"]), int(g["draws"])
    assert K == 16768 and draws == 132
    x0 = torch.from_numpy(g["x0"])
    torch.manual_seed(int(g["seed"]) + 3)
    proto = torch.empty(B, *x0.shape)
    z = torch.empty(K, x0.numel())
    for i in range(draws):                       # the last batch is drawn and discarded by the reference
        zi = torch.randn_like(proto)
        if (i + 1) * B <= K:
            z[i * B:(i + 1) * B] = zi.reshape(B, -1)
    from iddiff_b200 import sde_lib
    sde = sde_lib.VESDE(sigma_min=float(g["sigma_min"]), sigma_max=float(g["sigma_max"]), N=1000)
    mc, std = sde.coefficients(1e-5)
    label = float((torch.ones(1) * 1e-5 * 999)[0])
    res = eng.estimate(x0.reshape(1, -1).cuda(), K, mc, std, label, z=z.reshape(1, K, -1).cuda())
    torch.cuda.synchronize()
    sv = res.singular_values[0].cpu().numpy().astype(np.float64)
    want = g["singular_values"].astype(np.float64)
    rel = np.abs(sv / want - 1)
    print(f"config 5 true size: n = {sv.size}, max relative difference {rel.max():.3e} (median {np.median(rel):.2e}), "
          f"s0 {want[0]:.2f} s_last {want[-1]:.4f}, ID {res.dims.cpu().tolist()} vs {g['dims'].tolist()}")
    assert np.all(np.abs(sv - want) <= 1e-3 * want + 1e-6 * want[0])
    assert res.dims.cpu().tolist() == g["dims"].tolist()
    del eng
    torch.cuda.empty_cache()
