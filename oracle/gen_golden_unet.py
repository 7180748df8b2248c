"""TEST INFRASTRUCTURE ONLY -- U-Net goldens from the REAL reference (imported from /root/reference on CPU).

  unet_ddpm_small.npz : models.ddpm.DDPM (reference, unmodified) with the weights of fixtures.make_unet_state loaded
      strict=True (which also proves iddiff_b200.unet_models.DDPM reproduces the reference's state_dict layout):
      forward outputs on seeded inputs, and one dim_reduction.get_manifold_dimension run (2 anchors, K from the
      reference's own schedule) with its singular values and plot_utils dims.
Weights and noise are regenerated from seeds in the tests (checksums stored here).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from unittest import mock  # noqa: E402
import torch.utils.cpp_extension as _cpp  # noqa: E402

_cpp.load = lambda *a, **k: mock.MagicMock()  # the reference's `op` package: no nvcc JIT, its CPU branches run

from oracle import gen_golden as GG  # noqa: E402  (installs the shims, imports the reference's dim_reduction)
from oracle import fixtures as FX  # noqa: E402
from oracle import img_oracle as IO  # noqa: E402

GOLD = GG.GOLD


def gen_unet(name="unet_ddpm_small", spec=FX.DDPM_SMALL, seed=7, B=64, svd_points=3, model="ddpm"):
    import models.ddpm  # noqa: F401  (reference; registers 'ddpm')
    import models.ncsnpp  # noqa: F401  (reference; registers 'ncsnpp')
    import models.BeatGANsUNET  # noqa: F401  (reference; registers 'BeatGANsUNetModel')
    from models import utils as mutils
    from iddiff_b200 import unet_models
    holder = unet_models.get_model(model)(FX.unet_config(spec, model))
    oracle_forward = {"ddpm": IO.ddpm_forward, "ncsnpp": IO.ncsnpp_forward,
                      "BeatGANsUNetModel": IO.beatgans_forward}[model]
    state = FX.make_unet_state(holder, seed)
    size, cin = spec["image_size"], spec["input_channels"]
    model_kw = {k: v for k, v in spec.items() if k not in ("image_size", "centered")}
    if model.startswith("BeatGANs"):
        model_kw["image_size"] = size
    cfg = GG.base_config("vesde", [cin, size, size], B, model, "base", svd_points, **model_kw)
    cfg.data.update(dict(centered=spec["centered"], effective_image_size=size, image_size=size, num_channels=cin))
    ref = mutils.create_model(cfg)
    ref.load_state_dict(state, strict=True)
    ref.eval()
    # forward-level golden
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.rand(5, cin, size, size, generator=g)
    label = float((torch.ones(1) * 1e-5 * 999)[0])
    with torch.no_grad():
        y = ref(x, torch.full((5,), label))
        y_oracle = oracle_forward(state, spec, x, torch.full((5,), label))
    print("oracle vs reference forward: max abs diff", float((y - y_oracle).abs().max()), "scale", float(y.abs().max()))
    # path-level golden: anchors = smooth synthetic images, one loader batch of exactly B rows
    imgs = FX.smooth_images(svd_points, cin, size, seed + 2)
    batch = imgs[torch.arange(B) % svd_points]
    info, rec = GG.run_reference(cfg, [batch], seed + 3, ckpt_state=state)
    sv = np.asarray(info["singular_values"], dtype=np.float64)
    n_used = sv.shape[0]
    nb = len(rec["z"]) // n_used
    K = rec["svd_in"][0].shape[0]
    z = torch.stack(rec["z"]).reshape(n_used, nb, B, cin, size, size)
    scores = torch.stack([torch.cat(rec["scores"][a * nb:(a + 1) * nb], 0)[:K] for a in range(n_used)])
    out = dict(seed=seed, B=B, K=K, nb=nb, svd_points=svd_points, label=label, x=x.numpy(), y=y.numpy(),
               x0=imgs[:n_used].numpy(), singular_values=sv, dims=np.asarray(GG.reference_dims(info)),
               z_checksum=np.float64(FX.seeded_noise_checksum(z)),
               w_checksum=np.float64(sum(v.double().abs().sum().item() for v in state.values())),
               scores_head=scores[:, :8].numpy(), sigma_min=1e-2, sigma_max=4.0, num_scales=1000)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print(name, "K", K, "nb", nb, "dims", out["dims"], "s[:3]", sv[0, :3], "s[-3:]", sv[0, -3:])


def main(what):
    if "ddpm" in what:
        gen_unet()
    if "gray" in what:
        gen_unet("unet_ddpm_gray_small", FX.DDPM_GRAY_SMALL, seed=17, B=64, model="ddpm")
    if "beatgans" in what:
        gen_unet("unet_beatgans_small", FX.BEATGANS_SMALL, seed=13, model="BeatGANsUNetModel")
    if "ncsnpp" in what:
        gen_unet("unet_ncsnpp_small", FX.NCSNPP_SMALL, seed=11, model="ncsnpp")


if __name__ == "__main__":
    main(sys.argv[1:] or ["ddpm"])
