"""GPU parity of the DDPM U-Net path (UNetEngine, through the C ABI) against the REAL reference's goldens
(tests/golden/unet_ddpm_small.npz) and the fp64 oracle.

Tolerances (BASELINE.json north_star): integer ID bit-exact; singular values rtol 1e-3 per value (tail included)
against the fp32 reference plus the reference's own rounding floor max_i |s_ref - s_fp64| measured with the fp64
oracle; raw network output 2e-5 of its largest magnitude (fp32 reference arithmetic has the same order of error)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fixtures as FX  # noqa: E402
from oracle import id_oracle as O  # noqa: E402
from oracle import img_oracle as IO  # noqa: E402


MODELS = {"ddpm": ("unet_ddpm_small", FX.DDPM_SMALL, IO.ddpm_forward),
          "ddpm_gray": ("unet_ddpm_gray_small", FX.DDPM_GRAY_SMALL, IO.ddpm_forward),   # the paper's 1x32x32 configs
          "ncsnpp": ("unet_ncsnpp_small", FX.NCSNPP_SMALL, IO.ncsnpp_forward),
          "BeatGANsUNetModel": ("unet_beatgans_small", FX.BEATGANS_SMALL, IO.beatgans_forward)}
BOTH = pytest.mark.parametrize("model", ["ddpm", "ddpm_gray", "ncsnpp", "BeatGANsUNetModel"])


def _name(model):
    return "ddpm" if model == "ddpm_gray" else model


def _shape(spec):
    return (spec["input_channels"], spec["image_size"], spec["image_size"])


def _engine(model="ddpm"):
    from iddiff_b200 import unet_models
    from iddiff_b200.unet_engine import UNetEngine
    gold, spec, _ = MODELS[model]
    g = FX.golden(gold)
    holder = unet_models.get_model(_name(model))(FX.unet_config(spec, _name(model)))
    state = FX.make_unet_state(holder, int(g["seed"]))
    holder.load_state_dict(state, strict=True)
    eng = UNetEngine.from_module(holder, "cuda:0")
    eng.image_shape = _shape(spec)
    return g, state, eng


@BOTH
def test_unet_forward_vs_reference_golden(model):
    g, state, eng = _engine(model)
    _, spec, fwd = MODELS[model]
    x = torch.from_numpy(g["x"]).cuda()
    y = eng.forward(x, float(g["label"]))           # NHWC
    got = y.permute(0, 3, 1, 2).cpu().numpy()
    scale = np.abs(g["y"]).max()
    assert np.abs(got - g["y"]).max() < 2e-5 * scale
    # and against the fp64 evaluation of the oracle: the engine must stay within 8x of the fp32 reference's own
    # distance to the truth (measured: 2-3x on the 16x16 nets, 5.5x on the four-level 32x32 one -- the tensor cores'
    # fp32 accumulation truncates where the reference's FMA chain rounds)
    with torch.no_grad():
        y64 = fwd(state, spec, torch.from_numpy(g["x"]).double(),
                              torch.full((x.shape[0],), float(g["label"]), dtype=torch.float64)).numpy()
    ref_err = np.abs(g["y"] - y64).max()
    eng_err = np.abs(got - y64).max()
    assert eng_err < max(8 * ref_err, 2e-6 * scale), (eng_err, ref_err)


@BOTH
def test_unet_forward_batch_sizes_and_chunking(model):
    """Ragged image counts (1, 7, 130: partial 128-pixel tiles at the 8x8 level pack 2 images per tile)."""
    g, state, eng = _engine(model)
    gen = torch.Generator().manual_seed(3)
    x = torch.rand(130, *_shape(MODELS[model][1]), generator=gen)
    full = eng.forward(x.cuda(), float(g["label"])).cpu()
    for n in (1, 7):
        part = eng.forward(x[:n].contiguous().cuda(), float(g["label"])).cpu()
        assert torch.equal(part, full[:n])


@BOTH
def test_unet_manifold_dimension_vs_reference_golden(model):
    g, state, eng = _engine(model)
    _, spec, fwd = MODELS[model]
    from iddiff_b200 import sde_lib
    B, K, nb = int(g["B"]), int(g["K"]), int(g["nb"])
    x0 = torch.from_numpy(g["x0"])
    A = x0.shape[0]
    z = FX.seeded_noise(int(g["seed"]) + 3, A, nb, B, x0.shape[1:]).reshape(A, nb * B, -1)[:, :K].contiguous()
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    mc, std = sde.coefficients(1e-5)
    res = eng.estimate(x0.reshape(A, -1).cuda(), K, mc, std, float(g["label"]), z=z.cuda(), want_scores=True)
    torch.cuda.synchronize()
    sv = res.singular_values.cpu().numpy().astype(np.float64)
    want = g["singular_values"]
    # the reference's own rounding floor from the fp64 oracle on anchor 0
    za = z[0].reshape(K, *x0.shape[1:]).double()
    S64 = IO.unet_score_matrix(fwd, state, spec, x0[0].double(), za, mc, std, float(g["label"]))
    s64 = torch.linalg.svdvals(S64).numpy()
    floor = np.abs(want[0] - s64).max()
    assert np.all(np.abs(sv[0] - s64) <= 1e-3 * s64 + floor), np.abs(sv[0] / s64 - 1).max()
    for a in range(A):
        assert np.all(np.abs(sv[a] - want[a]) <= 1e-3 * want[a] + floor), (a, np.abs(sv[a] / want[a] - 1).max())
    assert res.dims.cpu().tolist() == g["dims"].tolist() == [O.id_from_spectrum(s) for s in want.tolist()]
    head = res.scores[:, :8].reshape(A, 8, *_shape(spec)).cpu().numpy()
    for a in range(A):
        scale = np.abs(g["scores_head"][a]).max()
        assert np.abs(head[a] - g["scores_head"][a]).max() < 2e-4 * scale


@BOTH
def test_unet_drop_in_get_manifold_dimension(tmp_path, model):
    """The reference entry point end to end for an image model: config (model.name='ddpm') -> Lightning-style
    checkpoint -> image anchors from a loader batch -> {'singular_values': ...}, against the REAL reference's output."""
    from iddiff_b200 import dim_reduction as DR, plot_utils
    from iddiff_b200 import unet_models
    gold, spec, _ = MODELS[model]
    g = FX.golden(gold)
    holder = unet_models.get_model(_name(model))(FX.unet_config(spec, _name(model)))
    state = FX.make_unet_state(holder, int(g["seed"]))
    ckpt = tmp_path / "last.ckpt"
    torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}}, ckpt)
    B, K, nb = int(g["B"]), int(g["K"]), int(g["nb"])
    x0 = torch.from_numpy(g["x0"])
    n_used, svd_points = x0.shape[0], int(g["svd_points"])
    imgs = FX.smooth_images(svd_points, spec["input_channels"], spec["image_size"], int(g["seed"]) + 2)
    assert torch.equal(imgs[:n_used], x0)
    cfg = FX.unet_config(spec, _name(model))
    cfg.logging = dict(log_path=str(tmp_path), log_name="run", svd_points=svd_points)
    cfg.dim_estimation = {}
    cfg.device = "cuda:0"
    cfg.training.batch_size = B
    cfg.training.lightning_module = "base"
    cfg.data.datamodule = "Tensor"
    cfg.data.anchors = imgs[torch.arange(B) % svd_points]
    cfg.model.checkpoint_path = str(ckpt)
    assert DR.k_schedule(x0.shape[1:], B)[2] == K
    z = FX.seeded_noise(int(g["seed"]) + 3, n_used, nb, B, x0.shape[1:]).reshape(n_used, nb * B, -1)[:, :K]
    info = DR.get_manifold_dimension(cfg, name="img", return_svd=True, noise=[z[i] for i in range(n_used)])
    sv = np.asarray(info["singular_values"])
    assert sv.shape == g["singular_values"].shape
    np.testing.assert_allclose(sv, g["singular_values"], rtol=1e-3)
    assert plot_utils.dims_from_spectra(info) == g["dims"].tolist() == DR.get_manifold_dimension.last_dims


def test_conditional_manifold_dimension_t_sweep(tmp_path):
    """dim_reduction.py:12-114: label filter (y == 1), N-1 anchors, one directory per time with the three pickles;
    one time slice is checked against the oracle's spectrum at that t."""
    import pickle
    from iddiff_b200 import dim_reduction as DR, unet_models, sde_lib
    model = "ddpm"
    gold, spec, fwd = MODELS[model]
    g = FX.golden(gold)
    holder = unet_models.get_model(_name(model))(FX.unet_config(spec, _name(model)))
    state = FX.make_unet_state(holder, int(g["seed"]))
    ckpt = tmp_path / "last.ckpt"
    torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}}, ckpt)
    B = 16                                              # K = (4 (256 // 16 + 1) - 1) * 16 + 0 = 1072 >= D = 768
    imgs = FX.smooth_images(B, 3, 16, 21)
    labels = torch.tensor([1, 0, 1, 1, 0, 1] + [0] * (B - 6))
    cfg = FX.unet_config(spec, _name(model))
    cfg.logging = dict(log_path=str(tmp_path), log_name="cond")
    cfg.dim_estimation = dict(num_datapoints=4)         # -> 3 anchors: images 0, 2, 3
    cfg.device = "cuda:0"
    cfg.training.batch_size = B
    cfg.data.datamodule = "Tensor"
    cfg.data.anchors, cfg.data.labels = imgs, labels
    cfg.model.checkpoint_path = str(ckpt)
    times = torch.linspace(1e-5, 0.3, 12)[:3]
    DR.get_conditional_manifold_dimension(cfg, times=times, seed=3)
    K = DR.k_schedule((3, 16, 16), B)[2]
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    by_t = {}
    for t in times:
        d = tmp_path / "cond" / "svd" / ("%.3f" % float(t))
        with open(d / "labels.pkl", "rb") as f:
            assert pickle.load(f)["labels"] == [1, 1, 1]
        with open(d / "images.pkl", "rb") as f:
            im = pickle.load(f)["images"]
        assert im.shape == (3, 16, 16, 3) and np.array_equal(im[1], imgs[2].permute(1, 2, 0).numpy())
        with open(d / "labels_svd.pkl", "rb") as f:
            sv = np.asarray(pickle.load(f)["singular_values"])
        assert sv.shape == (3, 768) and np.all(np.diff(sv, axis=1) <= 0)
        by_t[float(t)] = sv
    # oracle at the second time slice for anchor image 2 (statistical check: different noise, same distribution ->
    # leading singular values agree to a few per cent; exact-noise parity is covered by the golden tests above)
    t = float(times[1])
    mc, std = sde.coefficients(t)
    label = float((torch.ones(1) * t * 999)[0])
    z = torch.randn(K, 3, 16, 16, generator=torch.Generator().manual_seed(0))
    S = IO.unet_score_matrix(fwd, state, spec, imgs[2], z, mc, std, label)
    want = torch.linalg.svdvals(S).numpy()
    np.testing.assert_allclose(by_t[t][1, :8], want[:8], rtol=0.1)


def test_full_size_beatgans_properties():
    """BASELINE config 5 at its full network size (87.5 M parameters, 64x64x3, D = 12288), K = 256 rows, through
    size-independent properties: chunking invariance (bit-identical), finite sorted spectrum with K-1 non-trivial values,
    K-sharding invariance of the Philox rows (rows [128, 256) computed alone equal rows 128.. of the full run), and
    linearity of the read-out in 1/std."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import bench_unet
    from iddiff_b200 import sde_lib
    from iddiff_b200.engine import spectrum_from_rows
    eng, n_params, _ = bench_unet.build("beatgans")
    assert abs(n_params - 87.48e6) < 0.02e6          # 87.5 M (64d)
    sde = sde_lib.VESDE(sigma_min=0.01, sigma_max=50.0, N=1000)
    mc, std = sde.coefficients(1e-5)
    x0 = FX.smooth_images(1, 3, 64, 5).reshape(1, -1).cuda()
    K = 256
    S = eng.score_rows(x0, K, 0, mc, std, 0.00999, seed=4, image_chunk=128)
    S64 = eng.score_rows(x0, K, 0, mc, std, 0.00999, seed=4, image_chunk=64)
    assert torch.equal(S, S64)
    tail = eng.score_rows(x0, 128, 128, mc, std, 0.00999, seed=4)
    assert torch.equal(tail, S[:, 128:])
    r = spectrum_from_rows(S, K, 12288, -1.0 / std)
    sv = r.singular_values[0].cpu().numpy()
    assert sv.shape == (K,) and np.isfinite(sv).all() and np.all(np.diff(sv) <= 0)
    assert sv[-2] > 1e-3 * sv[0] and sv[-1] < 1e-4 * sv[0]     # centred K x D matrix with K < D has rank K - 1
    r2 = spectrum_from_rows(S, K, 12288, -2.0 / std)
    np.testing.assert_allclose(r2.singular_values[0, :-1].cpu().numpy(), 2 * sv[:-1], rtol=1e-5)


def test_drop_in_runs_the_papers_squares_config_shape_from_the_synthetic_datamodule(tmp_path):
    """configs/dimension_estimation/paper/image_data/squares/10.py end to end at a reduced width: data module
    'Synthetic' / FixedSquaresManifold (anchors from its shuffled train loader, as dim_reduction.py:150-163 takes
    them), model.name='ddpm' on 1x32x32, checkpoint, pickle out."""
    import pickle
    from iddiff_b200 import dim_reduction as DR, unet_models
    spec = FX.DDPM_GRAY_SMALL
    holder = unet_models.get_model("ddpm")(FX.unet_config(spec, "ddpm"))
    state = FX.make_unet_state(holder, 23)
    ckpt = tmp_path / "last.ckpt"
    torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}}, ckpt)
    cfg = FX.unet_config(spec, "ddpm")
    cfg.seed = 42
    cfg.logging = dict(log_path=str(tmp_path), log_name="squares", svd_points=3)
    cfg.dim_estimation = {}
    cfg.device = "cuda:0"
    cfg.training.batch_size = 16
    cfg.training.workers = 0
    cfg.training.lightning_module = "base"
    cfg.data.update(dict(datamodule="Synthetic", dataset_type="FixedSquaresManifold", data_samples=64,
                         split=[0.75, 0.125, 0.125], image_size=32, num_squares=10, square_range=[3, 5],
                         return_labels=False, shape=[1, 32, 32], use_data_mean=False))
    cfg.model.checkpoint_path = str(ckpt)
    torch.manual_seed(0)
    assert DR.get_manifold_dimension(cfg, name="squares") is None
    with open(tmp_path / "squares" / "svd" / "squares.pkl", "rb") as f:
        sv = np.asarray(pickle.load(f)["singular_values"])
    assert sv.shape == (2, 1024) and np.isfinite(sv).all() and np.all(np.diff(sv, axis=1) <= 0) and sv[:, 0].min() > 0


@pytest.mark.parametrize("gold", ["cond_ddpm_small_ve", "cond_ddpm_small_vp"])
def test_conditional_sweep_exact_noise_vs_reference_golden(tmp_path, gold):
    """Row a18: the REAL reference's get_conditional_manifold_dimension (dim_reduction.py:12-114; golden from
    oracle/gen_golden_cond.py) against the drop-in on the reference's own noise, at the first, a middle and the LAST of
    its 12 diffusion times (t = 0.3: VE sigma 0.06, VP mean coefficient 0.63 / std 0.78 -- the large-deviation end)."""
    import pickle
    from iddiff_b200 import dim_reduction as DR, plot_utils, sde_lib, unet_models
    g = FX.golden(gold)
    spec, fwd = FX.DDPM_SMALL, IO.ddpm_forward
    sde_name, seed, B, nb, K = str(g["sde"]), int(g["seed"]), int(g["B"]), int(g["nb"]), int(g["K"])
    holder = unet_models.get_model("ddpm")(FX.unet_config(spec, "ddpm"))
    state = FX.make_unet_state(holder, seed)
    ckpt = tmp_path / "last.ckpt"
    torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}}, ckpt)
    imgs = FX.smooth_images(6, 3, 16, seed + 2)
    batch = imgs[torch.arange(B) % 6]
    labels = torch.tensor([0, 1, 1, 0, 1, 1])[torch.arange(B) % 6]
    n_anchor = g["singular_values"].shape[1]
    assert np.array_equal(g["images"], np.stack([imgs[1].permute(1, 2, 0).numpy(), imgs[2].permute(1, 2, 0).numpy()]))
    # the reference's noise: one [B, C, H, W] draw per perturbation batch, anchors inside time slices, in order
    torch.manual_seed(seed + 3)
    proto = torch.empty(B, 3, 16, 16)
    z = torch.stack([torch.randn_like(proto) for _ in range(12 * n_anchor * nb)]).reshape(12, n_anchor, nb * B, -1)[:, :, :K]
    pick = [0, 5, 11]
    cfg = FX.unet_config(spec, "ddpm")
    cfg.training.sde = sde_name
    cfg.logging = dict(log_path=str(tmp_path), log_name="cond")
    cfg.dim_estimation = dict(num_datapoints=int(g["num_datapoints"]))
    cfg.device = "cuda:0"
    cfg.training.batch_size = B
    cfg.data.datamodule = "Tensor"
    cfg.data.anchors, cfg.data.labels = batch, labels
    cfg.model.checkpoint_path = str(ckpt)
    times = torch.from_numpy(g["times"])[pick]
    DR.get_conditional_manifold_dimension(cfg, times=times, noise=[[z[ti, j] for j in range(n_anchor)] for ti in pick])
    sde, _ = sde_lib.configure_sde(cfg)
    for ti in pick:
        d = tmp_path / "cond" / "svd" / str(g["dirs"][ti])
        with open(d / "labels_svd.pkl", "rb") as f:
            sv = np.asarray(pickle.load(f)["singular_values"])
        with open(d / "labels.pkl", "rb") as f:
            assert pickle.load(f)["labels"] == g["labels"].tolist()
        want = g["singular_values"][ti]
        t = float(g["times"][ti])
        mc, std = sde.coefficients(t)
        label = float((torch.ones(1) * t * (sde.N - 1))[0])
        S64 = IO.unet_score_matrix(fwd, state, spec, imgs[1].double(), z[ti, 0].reshape(K, 3, 16, 16).double(), mc, std, label)
        s64 = torch.linalg.svdvals(S64).numpy()
        floor = np.abs(want[0] - s64).max()
        rel = np.abs(sv / want - 1).max()
        print(f"{gold} t={t:.3f} std={std:.4f}: vs reference {rel:.3e}, vs fp64 {np.abs(sv[0] / s64 - 1).max():.3e}, "
              f"reference vs fp64 {np.abs(want[0] / s64 - 1).max():.3e}")
        assert np.isfinite(sv).all()
        assert np.all(np.abs(sv[0] - s64) <= 1e-3 * s64 + floor)
        assert np.all(np.abs(sv - want) <= 1e-3 * want + floor), rel
        assert plot_utils.dims_from_spectra({"singular_values": sv.tolist()}) == g["dims"][ti].tolist()
