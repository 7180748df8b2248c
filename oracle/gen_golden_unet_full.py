"""TEST INFRASTRUCTURE ONLY -- goldens of the FULL-SIZE image score networks (BASELINE configs 4 and 5) from the REAL
reference, imported from /root/reference on CPU.  Run in the build container:  python oracle/gen_golden_unet_full.py [ddpm ncsnpp beatgans]

  unet_{ddpm,ncsnpp,beatgans}_full.npz
    forward : the reference module (models.ddpm.DDPM / models.ncsnpp.NCSNpp / models.BeatGANsUNET.BeatGANsUNetModel at
              nf / model_channels = 128, 64x64x3, 116.1 / 126.3 / 87.5 M parameters), weights of synthetic.make_unet_state
              loaded strict=True, evaluated on 2 seeded images -> y; the fp64 oracle's y64 beside it (rounding floor).
    path    : K = 256 rows of one anchor.  The reference's driver cannot produce K < 16387 on a 64x64 image (its K
              schedule, dim_reduction.py:166-171), so this golden executes the reference's own pieces in the
              driver's order (dim_reduction.py:177-199): sde_lib.VESDE.marginal_prob -> torch.randn_like ->
              models.utils.get_score_fn(sde, model, conditional=False, train=False, continuous=True) -> cat / flatten /
              centre -> torch.linalg.svd -> plot_utils dims.  The fp64 oracle's spectrum is stored beside it.
Weights and noise are regenerated from seeds in the tests (checksums stored here).
"""
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from unittest import mock  # noqa: E402
import torch.utils.cpp_extension as _cpp  # noqa: E402

_cpp.load = lambda *a, **k: mock.MagicMock()  # the reference's `op` package: no nvcc JIT, its CPU branches run

from oracle import gen_golden as GG  # noqa: E402  (installs the shims)
from oracle import img_oracle as IO  # noqa: E402
from iddiff_b200 import synthetic as SY  # noqa: E402

GOLD = GG.GOLD
SDE = {"ddpm": (0.009, 50.0), "ncsnpp": (0.009, 50.0), "beatgans": (0.01, 50.0)}  # style_gan_base.py:100-101, style_gan_BeatGAN.py:23-24
SEEDS = {"ddpm": 101, "ncsnpp": 103, "beatgans": 107}
K_PATH, B_PATH = 256, 128


def gen_full(which):
    import models.ddpm  # noqa: F401
    import models.ncsnpp  # noqa: F401
    import models.BeatGANsUNET  # noqa: F401
    import sde_lib
    from models import utils as mutils
    from iddiff_b200 import unet_models
    model, spec, _ = SY.FULL_MODELS[which]
    seed = SEEDS[which]
    smin, smax = SDE[which]
    oracle_forward = {"ddpm": IO.ddpm_forward, "ncsnpp": IO.ncsnpp_forward, "beatgans": IO.beatgans_forward}[which]
    holder = unet_models.get_model(model)(SY.unet_config(spec, model))
    state = SY.make_unet_state(holder, seed)
    size, cin = spec["image_size"], spec["input_channels"]
    model_kw = {k: v for k, v in spec.items() if k not in ("image_size", "centered")}
    if model.startswith("BeatGANs"):
        model_kw["image_size"] = size
    cfg = GG.base_config("vesde", [cin, size, size], B_PATH, model, "base", 2, **model_kw)
    cfg.model.update(dict(sigma_min=smin, sigma_max=smax))
    cfg.data.update(dict(centered=spec["centered"], effective_image_size=size, image_size=size, num_channels=cin))
    ref = mutils.create_model(cfg)
    ref.load_state_dict(state, strict=True)
    ref.eval()
    n_params = sum(p.numel() for p in ref.parameters())
    sde = sde_lib.VESDE(sigma_min=smin, sigma_max=smax, N=1000)
    t = 1e-5
    label = float((torch.ones(1) * t * (sde.N - 1))[0])
    # ---- forward golden
    imgs = SY.smooth_images(3, cin, size, seed + 2)
    g = torch.Generator().manual_seed(seed + 1)
    x = (imgs[:2] + 0.02 * torch.randn(2, cin, size, size, generator=g)).clamp(0, 1)
    t0 = time.time()
    with torch.no_grad():
        y = ref(x, torch.full((2,), label))
        y64 = oracle_forward(state, spec, x.double(), torch.full((2,), label, dtype=torch.float64))
        y_or = oracle_forward(state, spec, x, torch.full((2,), label))
    print(which, "params %.2f M" % (n_params / 1e6), "forward %.1f s" % (time.time() - t0),
          "| oracle32 vs reference", float((y - y_or).abs().max()), "| reference vs fp64", float((y - y64).abs().max()),
          "| scale", float(y.abs().max()), flush=True)
    # ---- path golden: dim_reduction.py:166-199 with the reference's own objects, K = 256
    score_fn = mutils.get_score_fn(sde, ref, conditional=False, train=False, continuous=True)
    x0 = imgs[2]
    xr = x0.repeat([B_PATH, 1, 1, 1])
    vec_t = torch.ones(B_PATH) * t
    torch.manual_seed(seed + 3)
    zs, rows = [], []
    t0 = time.time()
    with torch.no_grad():
        for _ in range(K_PATH // B_PATH):
            batch = xr.clone()
            mean, std = sde.marginal_prob(batch, vec_t)
            z = torch.randn_like(batch)
            zs.append(z)
            batch = mean + std[(...,) + (None,) * len(batch.shape[1:])] * z
            rows.append(score_fn(batch, vec_t).detach().cpu())
    scores = torch.flatten(torch.cat(rows, dim=0), start_dim=1)
    means = scores.mean(dim=0, keepdim=True)
    normalized = scores - means
    u, s, v = torch.linalg.svd(normalized)
    sv = np.asarray(s.tolist(), dtype=np.float64)
    info = {"singular_values": [s.tolist()]}
    dims = GG.reference_dims(info)
    print(which, "path %.1f s" % (time.time() - t0), "dims", dims, "s[:3]", sv[:3], "s[-3:]", sv[-3:], flush=True)
    # fp64 oracle spectrum on the same noise (the reference's rounding floor)
    mc, sd = 1.0, float(sde.marginal_prob(torch.zeros(1, 1), torch.ones(1) * t)[1][0])
    zz = torch.cat(zs).double()
    t0 = time.time()
    S64 = IO.unet_score_matrix(oracle_forward, state, spec, x0.double(), zz, mc, sd, label, batch=32)
    s64 = torch.linalg.svdvals(S64).numpy()
    print(which, "fp64 path %.1f s" % (time.time() - t0), "floor", np.abs(sv - s64).max(), "rel",
          np.abs(sv / s64 - 1).max(), flush=True)
    out = dict(seed=seed, B=B_PATH, K=K_PATH, label=label, sigma_min=smin, sigma_max=smax, num_scales=1000, std=sd,
               n_params=n_params, x=x.numpy(), y=y.numpy(), y64=y64.numpy(), x0=x0.numpy(), singular_values=sv,
               s64=s64, dims=np.asarray(dims), z_checksum=np.float64(SY_checksum(torch.cat(zs))),
               w_checksum=np.float64(sum(v.double().abs().sum().item() for v in state.values())),
               scores_head=scores[:4].reshape(4, cin, size, size).numpy())
    np.savez_compressed(os.path.join(GOLD, f"unet_{which}_full.npz"), **out)


def SY_checksum(z):
    return float(z.double().sum().item())


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    for w in (sys.argv[1:] or ["beatgans", "ddpm", "ncsnpp"]):
        gen_full(w)
