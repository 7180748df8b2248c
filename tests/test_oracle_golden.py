"""CPU: the oracle restatement (oracle/id_oracle.py) against the fixtures that oracle/gen_golden.py
produced by running the REAL reference (tests/golden/*.npz).  This is what pins the oracle."""
import numpy as np
import pytest
import torch

from oracle import id_oracle as O
from oracle import fixtures as FX

SDE_PARAMS = dict(sigma_min=1e-2, sigma_max=4.0, beta_min=0.1, beta_max=20.0)


def test_marginals_match_reference():
    g = FX.golden("sde_marginals")
    t = torch.from_numpy(g["t"])
    x = torch.from_numpy(g["x"])
    cases = {
        "ve_ksphere": lambda: O.ve_marginal(x, t, 1e-2, 4),
        "ve_image": lambda: O.ve_marginal(x, t, 0.009, 50),
        "vp": lambda: O.vp_marginal(x, t, 0.1, 20),
        "subvp": lambda: O.subvp_marginal(x, t, 0.1, 20),
        "snr": lambda: O.snr_marginal(x, t),
    }
    for name, fn in cases.items():
        mean, std = fn()
        # same torch ops in the same order on the same dtype: bit-exact
        assert np.array_equal(mean.numpy(), g[name + "_mean"]), name
        assert np.array_equal(std.numpy(), g[name + "_std"]), name


def test_k_schedule_matches_reference():
    g = FX.golden("k_schedule")
    for shape, nd, B, K, nb in zip(g["shape"], g["ndim"], g["B"], g["K"], g["num_batches"]):
        num_batches, extra, k = O.k_schedule(list(shape[:nd]), int(B))
        assert (num_batches, k) == (int(nb), int(K))
    # SURVEY Appendix A.2 values
    assert O.k_schedule([100], 500)[2] == 1501
    assert O.k_schedule([1, 32, 32], 64)[2] == 4288
    assert O.k_schedule([3, 64, 64], 128)[2] == 16768
    assert O.k_schedule([100], 133)[2] == 400


def test_readout_matches_reference():
    g = FX.golden("readout")
    for s, n, d in zip(g["spectra"], g["lens"], g["dims"]):
        assert O.id_from_spectrum(s[:n].tolist()) == int(d)
    sv = g["spectra"][0][: g["lens"][0]]
    m = O.extract_sing_vals({"singular_values": [sv.tolist(), sv[::-1].tolist()]}, "mean")[0]
    assert np.array_equal(np.asarray(m), g["mean_mode"])
    with pytest.raises(ZeroDivisionError):
        O.id_from_spectrum([3.0, 1.0, 1.0, 0.5])


@pytest.mark.parametrize("name", ["fcn_small_ve", "fcn_small_vp", "fcn_small_subvp", "fcn_small_snr"])
def test_fcn_small_matches_reference(name):
    g = FX.golden(name)
    D, L, B, K = int(g["D"]), int(g["L"]), int(g["B"]), int(g["K"])
    sde = str(g["sde"])
    ws, bs = FX.fcn_state_to_lists(FX.golden_fcn_state(g), L)
    score_fn = O.make_score_fn("fcn", sde, SDE_PARAMS, fcn_params=(ws, bs))
    x0 = torch.from_numpy(g["x0"])
    n = x0.shape[0]
    z = torch.from_numpy(g["z"])
    zs = iter(z.reshape(-1, B, D))
    batch = x0[torch.arange(B) % n]
    cap = {}
    info = O.manifold_dimension([batch], n + 1, score_fn, sde, SDE_PARAMS, randn_like=lambda b: next(zs), capture=cap)
    sv = np.asarray(info["singular_values"])
    assert sv.shape == g["singular_values"].shape and cap["scores"][0].shape == (K, D)
    # same ATen/LAPACK kernels on the same inputs; allow last-bit differences across host CPUs
    np.testing.assert_allclose(torch.stack(cap["scores"]).numpy(), g["scores"], rtol=2e-5, atol=2e-5)
    np.testing.assert_allclose(sv, g["singular_values"], rtol=1e-4, atol=1e-5 * sv.max())
    assert [O.id_from_spectrum(s) for s in sv.tolist()] == g["dims"].tolist()
    assert [O.id_from_spectrum(s) for s in g["singular_values"].tolist()] == g["dims"].tolist()
    assert g["dims"].tolist() == [int(g["gap_rank"])] * n


def test_ksphere_c1_matches_reference():
    g = FX.golden("ksphere_c1")
    D, B, K, p = int(g["D"]), int(g["B"]), int(g["K"]), int(g["manifold_dim"])
    x0 = torch.from_numpy(g["x0"])
    n = x0.shape[0]
    nb, _, K2 = O.k_schedule([D], B)
    assert K2 == K == 400
    z = FX.seeded_noise(int(g["seed"]) + 2, n, nb, B, [D])
    assert abs(FX.seeded_noise_checksum(z) - float(g["z_checksum"])) < 1e-6 * z.numel() ** 0.5
    zs = iter(z.reshape(-1, B, D))
    score_fn = O.make_score_fn("ksphere_gt", "vesde", SDE_PARAMS, manifold_dim=p)
    batch = torch.cat([x0, x0[:1].repeat(B - n, 1)])
    cap = {}
    info = O.manifold_dimension([batch], n + 1, score_fn, "vesde", SDE_PARAMS, randn_like=lambda b: next(zs), capture=cap)
    sc = torch.stack(cap["scores"]).numpy()
    np.testing.assert_allclose(sc[:, :16], g["scores_head"], rtol=1e-4, atol=1e-2)
    sv = np.asarray(info["singular_values"])
    np.testing.assert_allclose(sv, g["singular_values"], rtol=1e-3)
    # anchors of the fixture came from the reference's KSphereDataset with the same seed
    torch.manual_seed(int(g["seed"]))
    data = O.ksphere_data(B, 1, D, p, 0.0, "random_isometry")
    np.testing.assert_allclose(data[:8].numpy(), g["data_head"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", ["fcn_d100", "fcn_c2"])
def test_fcn_bench_shape_matches_reference(name):
    g = FX.golden(name)
    D, L, B, K = int(g["D"]), int(g["L"]), int(g["B"]), int(g["K"])
    state = FX.golden_fcn_state(g)
    wsum = sum(v.double().abs().sum().item() for v in state.values())
    assert abs(wsum - float(g["w_checksum"])) < 1e-6 * wsum
    ws, bs = FX.fcn_state_to_lists(state, L)
    x0 = torch.from_numpy(g["x0"])
    n = x0.shape[0]
    nb, _, K2 = O.k_schedule([D], B)
    assert K2 == K == 1501
    z = FX.seeded_noise(int(g["seed"]) + 2, n, nb, B, [D])
    assert abs(FX.seeded_noise_checksum(z) - float(g["z_checksum"])) < 1e-6 * z.numel() ** 0.5
    zs = iter(z.reshape(-1, B, D))
    score_fn = O.make_score_fn("fcn", "vesde", SDE_PARAMS, fcn_params=(ws, bs))
    batch = x0[torch.arange(B) % n]
    cap = {}
    info = O.manifold_dimension([batch], n + 1, score_fn, "vesde", SDE_PARAMS, randn_like=lambda b: next(zs), capture=cap)
    sv = np.asarray(info["singular_values"])
    sc = torch.stack(cap["scores"]).numpy()
    np.testing.assert_allclose(sc[:, :16], g["scores_head"], rtol=1e-4, atol=1e-3)
    np.testing.assert_allclose(sv, g["singular_values"], rtol=1e-3, atol=1e-5 * sv.max())
    assert [O.id_from_spectrum(s) for s in sv.tolist()] == g["dims"].tolist() == [int(g["gap_rank"])] * n
