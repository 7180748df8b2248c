"""TEST INFRASTRUCTURE ONLY -- BASELINE config 5 at its TRUE size through the REAL reference's driver:
dim_reduction.get_manifold_dimension (unmodified, dim_reduction.py:116-211) on one 64x64x3 anchor with the full-size BeatGANs
U-Net (87.5 M parameters) and the loader batch size of the reference's config (B = 128) -> its own K schedule gives
K = 16768 perturbations (132 forward batches, the last one discarded), a 16768 x 12288 score matrix, full
LAPACK SVD -> 12288 singular values.  ~40 min on 8 cores.  Noise is regenerated in the test from the seed.
  python oracle/gen_golden_c5_full.py
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

_cpp.load = lambda *a, **k: mock.MagicMock()

from oracle import gen_golden as GG  # noqa: E402
from iddiff_b200 import synthetic as SY  # noqa: E402


def main():
    import models.BeatGANsUNET  # noqa: F401
    from iddiff_b200 import unet_models
    which, seed, B = "beatgans", 211, 128
    model, spec, _ = SY.FULL_MODELS[which]
    holder = unet_models.get_model(model)(SY.unet_config(spec, model))
    state = SY.make_unet_state(holder, seed)
    size, cin = spec["image_size"], spec["input_channels"]
    kw = {k: v for k, v in spec.items() if k not in ("image_size", "centered")}
    kw["image_size"] = size
    cfg = GG.base_config("vesde", [cin, size, size], B, model, "base", 2, **kw)     # svd_points 2 -> ONE anchor
    cfg.model.update(dict(sigma_min=0.01, sigma_max=50.0))                          # style_gan_BeatGAN.py:23-24
    cfg.data.update(dict(centered=spec["centered"], effective_image_size=size, image_size=size, num_channels=cin))
    imgs = SY.smooth_images(1, cin, size, seed + 2)
    batch = imgs[torch.zeros(B, dtype=torch.long)]
    t0 = time.time()
    with torch.no_grad():       # autograd graphs of 132 x 128-image forwards would not fit; values are unaffected
        GG._INJECT["batches"] = [batch]
        rec = {"n": 0, "svd_shape": None}
        orig_randn_like, orig_svd = torch.randn_like, torch.linalg.svd

        def randn_like(x, *a, **k):
            if rec["n"] == 0:
                torch.manual_seed(seed + 3)
            rec["n"] += 1
            if rec["n"] % 10 == 0:
                print("batch", rec["n"], "%.0f s" % (time.time() - t0), flush=True)
            return orig_randn_like(x, *a, **k)

        def svd(a, *ar, **k):
            rec["svd_shape"] = tuple(a.shape)
            print("svd of", a.shape, "%.0f s" % (time.time() - t0), flush=True)
            return orig_svd(a, *ar, **k)

        import tempfile
        f = tempfile.NamedTemporaryFile(suffix=".ckpt", delete=False)
        torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}}, f.name)
        cfg.model.checkpoint_path = f.name
        import dim_reduction
        with mock.patch.object(torch, "randn_like", randn_like), mock.patch.object(torch.linalg, "svd", svd):
            info = dim_reduction.get_manifold_dimension(cfg, name=None, return_svd=True)
    sv = np.asarray(info["singular_values"], dtype=np.float64)
    dims = GG.reference_dims(info)
    print("done %.0f s" % (time.time() - t0), "sv", sv.shape, "draws", rec["n"], "svd_in", rec["svd_shape"], "dims", dims,
          "s[:3]", sv[0, :3], "s[-3:]", sv[0, -3:], flush=True)
    np.savez_compressed(os.path.join(GG.GOLD, "c5_beatgans_k16768.npz"), seed=seed, B=B, K=rec["svd_shape"][0],
                        draws=rec["n"], x0=imgs[0].numpy(), singular_values=sv[0].astype(np.float32), dims=np.asarray(dims),
                        sigma_min=0.01, sigma_max=50.0, num_scales=1000,
                        w_checksum=np.float64(sum(v.double().abs().sum().item() for v in state.values())))


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    main()
