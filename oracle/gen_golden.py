"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the REAL reference
(/root/reference, imported through oracle/ref_shims.py) on fixed seeds and injected anchors.

Run in the build container only:   python oracle/gen_golden.py
/root/reference is read, never written.  The fixtures are small; large inputs (noise, bench-size
weights) are regenerated from the recorded torch seeds and guarded by checksums.

What is executed from the reference, unmodified:
  dim_reduction.get_manifold_dimension   (dim_reduction.py:116-211)
  models.utils.get_score_fn              (models/utils.py:156-282)
  models.fcn.FCN, models.ksphere_gt.KSphereGT
  sde_lib.VESDE/VPSDE/subVPSDE/SNRSDE.marginal_prob
  lightning_modules.BaseSdeGenerativeModel (configure_sde), KSphereGroundTruthModel
  lightning_data_modules.KSphereDataset.KSphereDataset
  plot_utils.plot_distribution / plot_dims (gap read-out; matplotlib replaced by mocks)
"""
import os
import sys
import tempfile
import types
from unittest import mock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_shims  # noqa: E402

ref_shims.install()
# matplotlib is not installed: plot_utils / utils only need the names to exist.
for name in ["matplotlib", "matplotlib.pyplot", "matplotlib.cm", "matplotlib.colors", "lpips", "iunets",
             "iunets.layers", "iunets.networks"]:
    sys.modules.setdefault(name, mock.MagicMock())

import dim_reduction  # noqa: E402  (reference)
import sde_lib  # noqa: E402
from models import utils as mutils  # noqa: E402
import models.fcn  # noqa: E402,F401
import models.ksphere_gt  # noqa: E402,F401
import lightning_modules.BaseSdeGenerativeModel  # noqa: E402,F401
import lightning_modules.KSphereGroundTruthModel  # noqa: E402,F401
from lightning_data_modules import utils as dm_utils  # noqa: E402
from lightning_data_modules.KSphereDataset import KSphereDataset  # noqa: E402

from oracle.fixtures import make_gap_fcn_state, seeded_noise_checksum  # noqa: E402

CD = ref_shims.ConfigDict
GOLD = os.path.join(ROOT, "tests", "golden")

_INJECT = {}


@dm_utils.register_lightning_datamodule(name="IdbInjected")
class _Injected:
    """A loader that yields the injected anchor batches in order (no shuffle, no workers)."""

    def __init__(self, config):
        pass

    def setup(self, stage=None):
        pass

    def train_dataloader(self):
        return list(_INJECT["batches"])

    def val_dataloader(self):   # (batch, labels) pairs for get_conditional_manifold_dimension (dim_reduction.py:21,48)
        return list(_INJECT["val_batches"])


@mutils.register_model(name="idb_identity")
class _Identity(torch.nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dummy = torch.nn.Linear(1, 1)

    def forward(self, x, labels):
        return x


def base_config(sde, data_shape, batch_size, model_name, lightning_module, svd_points, **model_kw):
    c = CD()
    c.logging = CD(log_path=tempfile.mkdtemp(prefix="idb_gold_"), log_name="g", svd_points=svd_points)
    c.training = CD(lightning_module=lightning_module, batch_size=batch_size, sde=sde, workers=0,
                    continuous=True, gpus=0)
    c.validation = CD(batch_size=batch_size, workers=0)
    c.eval = CD(batch_size=batch_size, workers=0)
    c.data = CD(datamodule="IdbInjected", shape=list(data_shape), use_data_mean=False)
    c.model = CD(name=model_name, checkpoint_path=None, sigma_min=1e-2, sigma_max=4.0, beta_min=0.1,
                 beta_max=20.0, num_scales=1000, **model_kw)
    c.dim_estimation = CD()
    c.device = torch.device("cpu")
    return c


def run_reference(config, batches, seed, ckpt_state=None):
    """Calls the reference's get_manifold_dimension and records noise, raw scores and the SVD input."""
    _INJECT["batches"] = batches
    rec = {"z": [], "scores": [], "svd_in": []}
    if ckpt_state is not None:
        f = tempfile.NamedTemporaryFile(suffix=".ckpt", delete=False)
        torch.save({"state_dict": {"score_model." + k: v for k, v in ckpt_state.items()}}, f.name)
        config.model.checkpoint_path = f.name
    orig_randn_like, orig_svd, orig_get = torch.randn_like, torch.linalg.svd, mutils.get_score_fn

    def randn_like(x, *a, **k):
        if not rec["z"]:
            # module construction inside get_manifold_dimension consumed the global generator for its
            # random init; re-seed at the first draw of the hot loop so tests can regenerate the noise.
            torch.manual_seed(seed)
        z = orig_randn_like(x, *a, **k)
        rec["z"].append(z.clone())
        return z

    def svd(a, *ar, **k):
        rec["svd_in"].append(a.clone())
        return orig_svd(a, *ar, **k)

    def get_score_fn(*a, **k):
        fn = orig_get(*a, **k)

        def wrapped(x, t):
            out = fn(x, t)
            rec["scores"].append(out.detach().clone())
            return out
        return wrapped

    with mock.patch.object(torch, "randn_like", randn_like), \
            mock.patch.object(torch.linalg, "svd", svd), \
            mock.patch.object(mutils, "get_score_fn", get_score_fn):
        info = dim_reduction.get_manifold_dimension(config, name=None, return_svd=True)
    return info, rec


def reference_dims(svd):
    """plot_utils.plot_distribution(svd, mode='all') -> dims (plot_utils.py:158-195); also plot_dims."""
    import plot_utils
    plot_utils.plt = mock.MagicMock()
    with mock.patch("builtins.print"):
        dims = plot_utils.plot_distribution(svd, mode="all")
    patch = mock.MagicMock()
    patch.get_x.return_value = 0.0
    patch.get_width.return_value = 0.5
    plot_utils.plt.hist.return_value = (np.zeros(1), np.arange(2.0), [patch])
    with mock.patch("builtins.print"):
        _, dims2 = plot_utils.plot_dims(svd)
    assert list(dims) == list(dims2)
    return [int(d) for d in dims]


def gen_fcn(name, sde, D, H, L, B, svd_points, seed, store_weights, store_z, gap_rank):
    state = make_gap_fcn_state(D, H, L, seed=seed, gap_rank=gap_rank)
    cfg = base_config(sde, [D], B, "fcn", "base", svd_points, state_size=D, hidden_layers=L, hidden_nodes=H,
                      dropout=0.0)
    g = torch.Generator().manual_seed(seed + 1)
    n_anchor = svd_points  # one loader batch holds more anchors than are consumed
    x0 = torch.randn(n_anchor, D, generator=g)
    x0 = x0 / x0.norm(dim=1, keepdim=True)
    # the loader batch must have exactly B rows (K depends on the loader batch size): pad with copies
    batch = x0[torch.arange(B) % n_anchor]
    info, rec = run_reference(cfg, [batch], seed + 2, ckpt_state=state)
    sv = np.asarray(info["singular_values"], dtype=np.float64)
    n_used = sv.shape[0]
    nb = len(rec["z"]) // n_used
    z = torch.stack(rec["z"]).reshape(n_used, nb, B, D)
    K = rec["svd_in"][0].shape[0]
    scores = torch.stack([torch.cat(rec["scores"][a * nb:(a + 1) * nb], 0)[:K] for a in range(n_used)])
    out = dict(sde=sde, D=D, H=H, L=L, B=B, svd_points=svd_points, seed=seed, gap_rank=gap_rank, K=K,
               x0=x0[:n_used].numpy(), singular_values=sv, dims=np.asarray(reference_dims(info)),
               z_checksum=np.float64(z.double().sum().item()),
               w_checksum=np.float64(sum(v.double().abs().sum().item() for v in state.values())),
               sigma_min=1e-2, sigma_max=4.0, beta_min=0.1, beta_max=20.0, num_scales=1000)
    if store_weights:
        for k, v in state.items():
            out["w::" + k] = v.numpy()
    if store_z:
        out["z"] = z.numpy()
        out["scores"] = scores.numpy()
    else:
        out["scores_head"] = scores[:, :16].numpy()
        out["scores_tail"] = scores[:, -4:].numpy()
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print(name, "K", K, "dims", out["dims"], "s[:3]", sv[0, :3], "s[-3:]", sv[0, -3:])


def gen_ksphere(name, D, p, B, svd_points, seed):
    cfg = base_config("vesde", [D], B, "ksphere_gt", "ksphere_gt", svd_points)
    cfg.data.update(dict(data_samples=B, n_spheres=1, ambient_dim=D, manifold_dim=p, noise_std=0.0,
                         embedding_type="random_isometry"))
    torch.manual_seed(seed)
    data = KSphereDataset(cfg).data  # reference generator (KSphereDataset.py:21-95)
    info, rec = run_reference(cfg, [data[:B]], seed + 2)
    sv = np.asarray(info["singular_values"], dtype=np.float64)
    n_used = sv.shape[0]
    nb = len(rec["z"]) // n_used
    K = rec["svd_in"][0].shape[0]
    z = torch.stack(rec["z"]).reshape(n_used, nb, B, D)
    scores = torch.stack([torch.cat(rec["scores"][a * nb:(a + 1) * nb], 0)[:K] for a in range(n_used)])
    out = dict(D=D, manifold_dim=p, B=B, svd_points=svd_points, seed=seed, K=K, x0=data[:n_used].numpy(),
               data_head=data[:8].numpy(), singular_values=sv, dims=np.asarray(reference_dims(info)),
               z_checksum=np.float64(z.double().sum().item()), scores_head=scores[:, :16].numpy(),
               scores_tail=scores[:, -4:].numpy(), sigma_min=1e-2, sigma_max=4.0, num_scales=1000)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print(name, "K", K, "dims", out["dims"], "s[:3]", sv[0, :3], "s[-3:]", sv[0, -3:])


def gen_marginals():
    t = torch.tensor([1e-5, 1e-3, 0.01, 0.1, 0.3, 0.5, 1.0], dtype=torch.float32)
    x = torch.linspace(-1, 1, 7 * 3).reshape(7, 3)
    out = {"t": t.numpy(), "x": x.numpy()}
    for nm, sde in [("ve_ksphere", sde_lib.VESDE(sigma_min=1e-2, sigma_max=4, N=1000)),
                    ("ve_image", sde_lib.VESDE(sigma_min=0.009, sigma_max=50, N=1000)),
                    ("vp", sde_lib.VPSDE(beta_min=0.1, beta_max=20, N=1000)),
                    ("subvp", sde_lib.subVPSDE(beta_min=0.1, beta_max=20, N=1000)),
                    ("snr", sde_lib.SNRSDE(N=1000))]:
        mean, std = sde.marginal_prob(x, t)
        out[nm + "_mean"] = mean.numpy()
        out[nm + "_std"] = std.numpy()
    np.savez_compressed(os.path.join(GOLD, "sde_marginals.npz"), **out)
    print("sde_marginals ok")


def gen_kschedule():
    rows = []
    for shape, B in [([7], 5), ([1, 8, 8], 5), ([3, 4, 4], 16), ([1, 6, 6], 36), ([2, 3, 5], 4)]:
        cfg = base_config("vesde", shape, B, "idb_identity", "base", 2)
        batch = torch.rand(B, *shape)
        info, rec = run_reference(cfg, [batch], 0)
        rows.append((shape + [0] * (3 - len(shape)), len(shape), B, rec["svd_in"][0].shape[0],
                     len(rec["z"])))
    np.savez_compressed(os.path.join(GOLD, "k_schedule.npz"),
                        shape=np.asarray([r[0] for r in rows]), ndim=np.asarray([r[1] for r in rows]),
                        B=np.asarray([r[2] for r in rows]), K=np.asarray([r[3] for r in rows]),
                        num_batches=np.asarray([r[4] for r in rows]))
    print("k_schedule", [(r[0][:r[1]], r[2], r[3], r[4]) for r in rows])


def gen_readout():
    rng = np.random.default_rng(0)
    spectra = []
    # cliff spectra (trained-net like), smooth decays, plateaus with ties, tiny tails, fp32-valued
    for n, r in [(100, 11), (200, 51), (24, 5), (50, 1), (30, 28)]:
        s = np.concatenate([np.sort(rng.uniform(4e3, 6.5e3, n - r))[::-1], np.sort(rng.uniform(0, 5, r))[::-1]])
        spectra.append(np.float32(s).astype(np.float64))
    for n in [5, 16, 100, 500]:
        s = np.sort(rng.lognormal(0, 2, n))[::-1]
        spectra.append(np.float32(s).astype(np.float64))
    s = np.float32(np.linspace(10, 1, 12)).astype(np.float64)  # near-equal gaps
    spectra.append(s)
    s = np.asarray([9.0, 7.0, 5.0, 5.0, 5.0, 1.0, 0.5])  # exact ties
    spectra.append(s)
    s = np.asarray([3.0, 2.0, 1.0])  # minimum length
    spectra.append(s)
    dims = []
    for s in spectra:
        dims.append(reference_dims({"singular_values": [s.tolist()]})[0])
    L = max(len(s) for s in spectra)
    pad = np.full((len(spectra), L), np.nan)
    for i, s in enumerate(spectra):
        pad[i, :len(s)] = s
    # 'mean' aggregation mode of extract_sing_vals (plot_utils.py:197-205)
    import plot_utils
    with mock.patch("builtins.print"):
        mean_sv = plot_utils.extract_sing_vals({"singular_values": [spectra[0].tolist(), spectra[0][::-1].tolist()]},
                                               "mean")[0]
    np.savez_compressed(os.path.join(GOLD, "readout.npz"), spectra=pad, lens=np.asarray([len(s) for s in spectra]),
                        dims=np.asarray(dims), mean_mode=np.asarray(mean_sv))
    print("readout dims", dims)


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    gen_marginals()
    gen_kschedule()
    gen_readout()
    gen_fcn("fcn_small_ve", "vesde", D=24, H=64, L=2, B=50, svd_points=4, seed=10, store_weights=True,
            store_z=True, gap_rank=5)
    gen_fcn("fcn_small_vp", "vpsde", D=24, H=64, L=2, B=50, svd_points=3, seed=20, store_weights=True,
            store_z=True, gap_rank=7)
    gen_fcn("fcn_small_subvp", "subvpsde", D=24, H=64, L=2, B=50, svd_points=3, seed=30, store_weights=True,
            store_z=True, gap_rank=3)
    gen_fcn("fcn_small_snr", "snrsde", D=24, H=64, L=2, B=50, svd_points=3, seed=40, store_weights=True,
            store_z=True, gap_rank=9)
    gen_ksphere("ksphere_c1", D=100, p=10, B=133, svd_points=4, seed=50)
    gen_fcn("fcn_c2", "vesde", D=200, H=2048, L=5, B=500, svd_points=3, seed=60, store_weights=False,
            store_z=False, gap_rank=51)
    gen_fcn("fcn_d100", "vesde", D=100, H=2048, L=5, B=500, svd_points=3, seed=70, store_weights=False,
            store_z=False, gap_rank=11)
