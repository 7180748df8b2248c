"""TEST INFRASTRUCTURE ONLY -- golden of the conditional t-sweep from the REAL reference:
dim_reduction.get_conditional_manifold_dimension (dim_reduction.py:12-114), unmodified, on the reduced-width DDPM of
fixtures.DDPM_SMALL (3x16x16, B = 64 -> K = 1216, n = 768), labelled anchors from an injected validation loader
(label filter y == 1), all 12 diffusion times linspace(sampling_eps, 0.3, 12), for the VE and the VP SDE.

The noise is regenerated in the tests from the recorded seed (torch.manual_seed immediately before the first
randn_like of the run, then one [B, C, H, W] draw per perturbation batch, anchors inside time slices, in order).
  python oracle/gen_golden_cond.py
"""
import os
import pickle
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from unittest import mock  # noqa: E402
import torch.utils.cpp_extension as _cpp  # noqa: E402

_cpp.load = lambda *a, **k: mock.MagicMock()

from oracle import gen_golden as GG  # noqa: E402
from oracle import fixtures as FX  # noqa: E402


def gen_cond(name, sde, seed, B=64, num_datapoints=3):
    import models.ddpm  # noqa: F401
    import tempfile
    import dim_reduction
    from iddiff_b200 import unet_models
    spec = FX.DDPM_SMALL
    holder = unet_models.get_model("ddpm")(FX.unet_config(spec, "ddpm"))
    state = FX.make_unet_state(holder, seed)
    size, cin = spec["image_size"], spec["input_channels"]
    model_kw = {k: v for k, v in spec.items() if k not in ("image_size", "centered")}
    cfg = GG.base_config(sde, [cin, size, size], B, "ddpm", "base", num_datapoints, **model_kw)
    cfg.data.update(dict(centered=spec["centered"], effective_image_size=size, image_size=size, num_channels=cin))
    cfg.dim_estimation = GG.CD(num_datapoints=num_datapoints)
    f = tempfile.NamedTemporaryFile(suffix=".ckpt", delete=False)
    torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}}, f.name)
    cfg.model.checkpoint_path = f.name
    # one validation batch of exactly B rows; labels: anchors 0 and 3 are skipped by the y == 1 filter
    imgs = FX.smooth_images(6, cin, size, seed + 2)
    batch = imgs[torch.arange(B) % 6]
    labels = torch.tensor([0, 1, 1, 0, 1, 1])[torch.arange(B) % 6]
    GG._INJECT["val_batches"] = [(batch, labels)]
    rec = {"n": 0}
    orig_randn_like = torch.randn_like

    def randn_like(x, *a, **k):
        if rec["n"] == 0:
            torch.manual_seed(seed + 3)
        rec["n"] += 1
        return orig_randn_like(x, *a, **k)

    with mock.patch.object(torch, "randn_like", randn_like), torch.no_grad():
        dim_reduction.get_conditional_manifold_dimension(cfg, name=None)
    base = os.path.join(cfg.logging.log_path, cfg.logging.log_name, "svd")
    times = torch.linspace(1e-5 if sde == "vesde" else 1e-3, 0.3, 12)
    svs, labs = [], None
    for t in times:
        d = os.path.join(base, "%.3f" % t.item())
        svs.append(np.asarray(pickle.load(open(os.path.join(d, "labels_svd.pkl"), "rb"))["singular_values"]))
        labs = pickle.load(open(os.path.join(d, "labels.pkl"), "rb"))["labels"]
        images = pickle.load(open(os.path.join(d, "images.pkl"), "rb"))["images"]
    sv = np.stack(svs)            # [12, anchors, n]
    dims = np.asarray([GG.reference_dims({"singular_values": s.tolist()}) for s in sv])
    n_anchor = sv.shape[1]
    nb = rec["n"] // (12 * n_anchor)
    print(name, "sv", sv.shape, "draws", rec["n"], "nb", nb, "dims", dims[[0, 5, 11]].tolist(),
          "s[0]", sv[[0, 5, 11], 0, 0], "s[-1]", sv[[0, 5, 11], 0, -1])
    np.savez_compressed(os.path.join(GG.GOLD, name + ".npz"), sde=sde, seed=seed, B=B, nb=nb, K=(nb - 1) * B,
                        num_datapoints=num_datapoints, times=times.numpy(), singular_values=sv, dims=dims,
                        labels=np.asarray(labs), images=images, batch_index=np.asarray([1, 2]),
                        dirs=np.asarray(["%.3f" % t.item() for t in times]),
                        sigma_min=1e-2, sigma_max=4.0, beta_min=0.1, beta_max=20.0, num_scales=1000)


if __name__ == "__main__":
    gen_cond("cond_ddpm_small_ve", "vesde", seed=23)
    gen_cond("cond_ddpm_small_vp", "vpsde", seed=29)


def gen_readout_tail():
    """plot_utils.plot_distribution(svd, mode='all', tail=t) of the REAL reference on the readout.npz spectra."""
    import plot_utils
    plot_utils.plt = mock.MagicMock()
    g = np.load(os.path.join(GG.GOLD, "readout.npz"))
    spectra = [g["spectra"][i, :int(n)].tolist() for i, n in enumerate(g["lens"])]
    out = {}
    for t in (1, 3, 8, 50):
        with mock.patch("builtins.print"):
            out[f"tail_{t}"] = np.asarray(plot_utils.plot_distribution({"singular_values": spectra}, mode="all", tail=t))
    np.savez_compressed(os.path.join(GG.GOLD, "readout_tail.npz"), **out)
    print("readout_tail", {k: v.tolist() for k, v in out.items()})
