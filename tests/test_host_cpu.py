"""CPU: the host-side mirror of the reference interface (no kernels are launched)."""
import os
import tempfile

import numpy as np
import pytest
import torch

from iddiff_b200 import data, models, plot_utils, sde_lib
from iddiff_b200 import dim_reduction as DR
from iddiff_b200.config import ConfigDict, lookup
from oracle import fixtures as FX
from oracle import id_oracle as O


def test_k_schedule_matches_oracle_and_reference_fixture():
    g = FX.golden("k_schedule")
    for shape, nd, B, K, nb in zip(g["shape"], g["ndim"], g["B"], g["K"], g["num_batches"]):
        assert DR.k_schedule(list(shape[:nd]), int(B)) == O.k_schedule(list(shape[:nd]), int(B))
        assert DR.k_schedule(list(shape[:nd]), int(B))[2] == int(K)
    assert DR.k_schedule([200], 500)[2] == 1501 and DR.k_schedule([3, 64, 64], 128)[2] == 16768


def test_gather_anchors_keeps_the_n_minus_one_quirk():
    batches = [torch.arange(12.).reshape(4, 3), torch.arange(12., 18.).reshape(2, 3)]
    got = DR.gather_anchors(batches, 6)
    assert len(got) == 5 and [b for _, b in got] == [4, 4, 4, 4, 2]       # last loader batch is smaller -> other K
    assert len(DR.gather_anchors(batches, 3)) == 2 and DR.gather_anchors(batches, 1) == []
    assert len(DR.gather_anchors(batches, 100)) == 6


def test_sde_marginals_match_reference_fixture():
    g = FX.golden("sde_marginals")
    t, x = torch.from_numpy(g["t"]), torch.from_numpy(g["x"])
    cases = {"ve_ksphere": sde_lib.VESDE(1e-2, 4, 1000), "ve_image": sde_lib.VESDE(0.009, 50, 1000),
             "vp": sde_lib.VPSDE(0.1, 20, 1000), "subvp": sde_lib.subVPSDE(0.1, 20, 1000), "snr": sde_lib.SNRSDE(1000)}
    for name, sde in cases.items():
        mean, std = sde.marginal_prob(x, t)
        assert np.array_equal(mean.numpy(), g[name + "_mean"]) and np.array_equal(std.numpy(), g[name + "_std"]), name
        for ti in (1e-5, 1e-3):   # the two scalars the kernels get are the per-row fp32 values of marginal_prob
            mc, sd = sde.coefficients(ti)
            m2, s2 = sde.marginal_prob(torch.full((2, 3), 0.7), torch.ones(2) * ti)
            assert np.float32(sd) == s2[0].item() and torch.equal(torch.tensor(mc) * torch.full((3,), 0.7), m2[0])
            assert DR.sde_coefficients(sde, ti) == (mc, sd)

    class RefLike:  # an object with only the reference's surface
        N = 1000
        marginal_prob = staticmethod(cases["vp"].marginal_prob)
    assert DR.sde_coefficients(RefLike(), 1e-3) == cases["vp"].coefficients(1e-3)


def test_configure_sde_picks_eps_like_the_reference():
    base = dict(sigma_min=0.01, sigma_max=4.0, beta_min=0.1, beta_max=20.0, num_scales=1000)
    for name, eps in [("vesde", 1e-5), ("VPSDE", 1e-3), ("subvpsde", 1e-3), ("snrsde", 1e-3)]:
        cfg = ConfigDict(training=dict(sde=name), model=base, data=dict(use_data_mean=False))
        sde, e = sde_lib.configure_sde(cfg)
        assert e == eps and sde.N == 1000
    with pytest.raises(NotImplementedError):
        sde_lib.configure_sde(ConfigDict(training=dict(sde="foo"), model=base, data={}))


def test_readout_matches_reference_fixture():
    g = FX.golden("readout")
    for s, n, d in zip(g["spectra"], g["lens"], g["dims"]):
        assert plot_utils.dim_from_spectrum(s[:n].tolist()) == int(d) == O.id_from_spectrum(s[:n].tolist())
    svd = {"singular_values": [g["spectra"][0][: g["lens"][0]].tolist(), g["spectra"][0][: g["lens"][0]][::-1].tolist()]}
    assert np.array_equal(np.asarray(plot_utils.extract_sing_vals(svd, "mean")[0]), g["mean_mode"])
    assert plot_utils.dims_from_spectra(svd, "first") == [int(g["dims"][0])]
    with pytest.raises(ZeroDivisionError):
        plot_utils.dim_from_spectrum([3.0, 1.0, 1.0, 0.5])


def test_ksphere_generator_matches_oracle():
    for kw in [dict(embedding_type="random_isometry"), dict(embedding_type="first", angle_std=0.5),
               dict(embedding_type="separating"), dict(embedding_type="along_axis", radii=[1.0, 2.5])]:
        torch.manual_seed(5)
        a = data.ksphere_points(20, 2, 40, [3, 6] if "radii" not in kw else 6, 0.01, **kw)
        torch.manual_seed(5)
        b = O.ksphere_data(20, 2, 40, [3, 6] if "radii" not in kw else 6, 0.01, **kw)
        assert a.shape == b.shape == (40, 40) and (a - b).abs().max().item() < 1e-6
    g = FX.golden("ksphere_c1")   # and against the reference's own KSphereDataset output
    torch.manual_seed(int(g["seed"]))
    d = data.ksphere_points(int(g["B"]), 1, int(g["D"]), int(g["manifold_dim"]), 0.0, "random_isometry")
    np.testing.assert_allclose(d[:8].numpy(), g["data_head"], rtol=1e-5, atol=1e-6)


def test_config_lookup_and_registries():
    cfg = ConfigDict(logging=dict(svd_points=5), dim_estimation={}, model=dict(name="fcn"))
    assert lookup(cfg, "logging.svd_points") == 5 and lookup(cfg, "dim_estimation.num_datapoints") is None
    assert cfg.logging.svd_points == 5 and cfg["model"]["name"] == "fcn"
    assert models.get_model("fcn") is models.FCN and models.get_model("ksphere_gt") is models.KSphereGT
    with pytest.raises(ValueError):
        models.register_model("fcn")(object)


def test_fcn_holder_has_reference_state_dict_layout_and_loads_lightning_ckpt():
    g = FX.golden("fcn_small_ve")
    D, H, L = int(g["D"]), int(g["H"]), int(g["L"])
    cfg = ConfigDict(model=dict(name="fcn", state_size=D, hidden_layers=L, hidden_nodes=H, dropout=0.0))
    m = models.create_model(cfg)
    state = FX.golden_fcn_state(g)
    assert set(m.state_dict()) == set(state)                      # mlp.0.weight, mlp.3.weight, ... like models/fcn.py
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "last.ckpt")
        torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}, "hyper_parameters": {}}, path)
        models.load_checkpoint_into(m, path)
    for k, v in state.items():
        assert torch.equal(m.state_dict()[k], v)
    with pytest.raises(RuntimeError, match="holds weights only"):
        m(torch.zeros(1, D), torch.zeros(1))


def test_drop_in_rejects_cpu_device_and_missing_num_datapoints():
    from iddiff_b200._lib import IdbError
    with tempfile.TemporaryDirectory() as td:
        cfg = ConfigDict(logging=dict(log_path=td, log_name="x", svd_points=3), device="cpu",
                         training=dict(sde="vesde", batch_size=4), data=dict(datamodule="Tensor", anchors=torch.zeros(4, 3),
                                                                            use_data_mean=False, manifold_dim=2),
                         model=dict(name="ksphere_gt", sigma_min=0.01, sigma_max=4.0, num_scales=1000))
        with pytest.raises(IdbError, match="CUDA devices only"):
            DR.get_manifold_dimension(cfg, return_svd=True)
        assert os.path.isdir(os.path.join(td, "x", "svd"))        # the reference creates the directory first


def test_read_config_loads_reference_style_config_files(tmp_path):
    """main.py:32-34 / configs/utils.py:6-10: a config FILE written the reference's way (`import ml_collections`,
    a `get_config()` building nested ConfigDicts, a base config imported as `configs.<...>`) loads unmodified, with or
    without ml_collections installed; a pickled config loads too."""
    import pickle
    from iddiff_b200 import main as M, config as C
    pkg = tmp_path / "repo" / "configs" / "sub"
    pkg.mkdir(parents=True)
    (tmp_path / "repo" / "configs" / "__init__.py").write_text("")
    (pkg / "__init__.py").write_text("")
    (pkg / "base.py").write_text(
        "import ml_collections\n"
        "def get_config():\n"
        "  config = ml_collections.ConfigDict()\n"
        "  config.logging = logging = ml_collections.ConfigDict()\n"
        "  logging.svd_points = 5\n"
        "  config.model = model = ml_collections.ConfigDict()\n"
        "  model.name = 'fcn'\n"
        "  model.sigma_min = 0.01\n"
        "  config.dim_estimation = ml_collections.ConfigDict()\n"
        "  return config\n")
    (pkg / "derived.py").write_text(
        "import ml_collections\n"
        "from configs.sub.base import get_config as get_base_config\n"
        "def get_config():\n"
        "  config = get_base_config()\n"
        "  config.model.sigma_max = 4\n"
        "  config.data = data = ml_collections.ConfigDict()\n"
        "  data.shape = [200]\n"
        "  return config\n")
    cfg = M.read_config(str(pkg / "derived.py"))
    assert cfg.model.name == "fcn" and cfg.model.sigma_max == 4 and cfg.data.shape == [200]
    assert C.lookup(cfg, "logging.svd_points") == 5 and C.lookup(cfg, "dim_estimation.num_datapoints") is None
    with open(tmp_path / "c.pkl", "wb") as f:
        pickle.dump(C.ConfigDict(model=dict(name="ddpm")), f)
    assert M.read_config(str(tmp_path / "c.pkl")).model.name == "ddpm"


def test_read_config_loads_the_reference_tree_when_present():
    """In the build container the reference checkout exists: its own dimension-estimation config files load."""
    import os
    import pytest
    from iddiff_b200 import main as M, dim_reduction as DR
    root = "/root/reference/configs/dimension_estimation/paper"
    if not os.path.isdir(root):
        pytest.skip("reference checkout not on this machine")
    cfg = M.read_config(os.path.join(root, "euclidean_data/ksphere/50dim.py"))
    assert cfg.model.name == "fcn" and cfg.training.sde == "vesde" and cfg.logging.svd_points > 0
    img = M.read_config(os.path.join(root, "image_data/squares/10.py"))
    assert img.model.name == "ddpm" and list(img.data.shape) == [1, 32, 32]
    assert DR.k_schedule(img.data.shape, img.training.batch_size)[2] == 4480


def test_synthetic_image_manifolds_match_the_reference_generators():
    """'Synthetic' data module (the paper's squares / gaussian_blobs configs): the vectorised generators reproduce the
    images the REAL reference's FixedSquaresManifold / FixedGaussiansManifold built from the same seed
    (tests/golden/synthetic_images.npz via oracle/gen_golden_data.py) bit for bit, and the module serves batches of
    the config's shape."""
    import numpy as np
    import torch
    from iddiff_b200 import data
    from iddiff_b200.config import ConfigDict
    from oracle import fixtures as FX
    g = FX.golden("synthetic_images")
    for seed in (42, 7):
        sq = data.fixed_squares_images(seed, 24, 10, [3, 5], 32)
        assert sq.dtype == torch.float32 and np.array_equal(sq.numpy(), g[f"squares_{seed}"])
        ga = data.fixed_gaussians_images(seed, 24, 10, [1, 5], 32)
        assert ga.dtype == torch.float32 and np.array_equal(ga.numpy(), g[f"gaussians_{seed}"])
    cfg = ConfigDict(seed=42, training=dict(batch_size=8, workers=0),
                     data=dict(datamodule="Synthetic", dataset_type="FixedSquaresManifold", data_samples=40,
                               split=[0.8, 0.1, 0.1], image_size=32, num_squares=10, square_range=[3, 5],
                               return_labels=False, shape=[1, 32, 32]))
    dm = data.create_lightning_datamodule(cfg)
    dm.setup()
    batch = next(iter(dm.train_dataloader()))
    assert tuple(batch.shape) == (8, 1, 32, 32) and len(dm.train_data) == 32 and len(dm.valid_data) == 4


def test_plot_distribution_tail_truncates_the_softmax_like_the_reference():
    """plot_utils.py:178-179: `tail` keeps the last entries of the SOFTMAX (not of the spectrum); golden from the real
    reference's plot_distribution(mode='all', tail=t) on the readout fixtures."""
    import numpy as np
    from iddiff_b200 import plot_utils
    from oracle import fixtures as FX
    g, gt = FX.golden("readout"), FX.golden("readout_tail")
    spectra = [g["spectra"][i, :int(n)].tolist() for i, n in enumerate(g["lens"])]
    for t in (1, 3, 8, 50):
        assert plot_utils.plot_distribution({"singular_values": spectra}, mode="all", tail=t) == gt[f"tail_{t}"].tolist()
    assert plot_utils.plot_distribution({"singular_values": spectra}, mode="all") == g["dims"].tolist()
    assert plot_utils.plot_dims({"singular_values": spectra})[1] == g["dims"].tolist()
