"""TEST INFRASTRUCTURE ONLY -- runs the UNMODIFIED reference (from /root/reference here, from the git-ignored copy
oracle/_ref on the GPU box, see oracle/build_ref.sh) on the host cores, for bench.py's `--impl reference` arm and its
`cpu_baseline` legs.  Never on the product path.

  fcn_get_manifold_dimension : the reference's own entry point dim_reduction.get_manifold_dimension (dim_reduction.py:116-211)
      with an injected loader and an in-memory Lightning-style checkpoint -- module construction, checkpoint load, score
      batches with their autograd graphs (no torch.no_grad there), .cpu() copies and the full LAPACK SVD
      included, exactly what a user of the reference pays.
  unet_forward : the reference's own nn.Module forward (models/ddpm.py, models/ncsnpp.py, models/BeatGANsUNET.py).
"""
import os
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402


def available() -> bool:
    return ref_shims.available()


_GG = None


def _gg():
    """oracle.gen_golden installs the shims and imports the reference's dim_reduction + registries once."""
    global _GG
    if _GG is None:
        from unittest import mock
        import torch.utils.cpp_extension as _cpp
        _cpp.load = lambda *a, **k: mock.MagicMock()   # the reference's `op` package: no nvcc JIT; CPU branches run
        from oracle import gen_golden
        _GG = gen_golden
    return _GG


class FcnReference:
    """BASELINE configs[1] through the reference's entry point; the checkpoint file is written once."""

    def __init__(self, state, D, H, L, B, sigma_min=1e-2, sigma_max=4.0):
        GG = _gg()
        self.GG, self.D, self.B = GG, D, B
        f = tempfile.NamedTemporaryFile(suffix=".ckpt", delete=False)
        torch.save({"state_dict": {"score_model." + k: v for k, v in state.items()}}, f.name)
        self.cfg = GG.base_config("vesde", [D], B, "fcn", "base", 2, state_size=D, hidden_layers=L, hidden_nodes=H,
                                  dropout=0.0)
        self.cfg.model.update(dict(sigma_min=sigma_min, sigma_max=sigma_max, checkpoint_path=f.name))

    def run(self, anchors):
        """anchors [n, D] -> spectra list; n <= B anchors ride in ONE loader batch of exactly B rows (K = 3B + 1)."""
        import dim_reduction  # the reference's
        n = anchors.shape[0]
        batch = anchors[torch.arange(self.B) % n]
        self.GG._INJECT["batches"] = [batch]
        self.cfg.logging.svd_points = n + 1           # N - 1 anchors are processed
        info = dim_reduction.get_manifold_dimension(self.cfg, name=None, return_svd=True)
        assert len(info["singular_values"]) == n
        return info["singular_values"]


def unet_reference_module(which, state):
    """The reference's own module for 'ddpm' | 'ncsnpp' | 'beatgans' at full size with `state` loaded strict."""
    GG = _gg()
    import models.ddpm  # noqa: F401
    import models.ncsnpp  # noqa: F401
    import models.BeatGANsUNET  # noqa: F401
    from models import utils as mutils
    from iddiff_b200 import synthetic as SY
    model, spec, _ = SY.FULL_MODELS[which]
    size, cin = spec["image_size"], spec["input_channels"]
    kw = {k: v for k, v in spec.items() if k not in ("image_size", "centered")}
    if model.startswith("BeatGANs"):
        kw["image_size"] = size
    cfg = GG.base_config("vesde", [cin, size, size], 128, model, "base", 2, **kw)
    cfg.data.update(dict(centered=spec["centered"], effective_image_size=size, image_size=size, num_channels=cin))
    ref = mutils.create_model(cfg)
    ref.load_state_dict(state, strict=True)
    return ref.eval()


def time_unet_forward(which, state, n_images, label=0.00999):
    """score-evals/s of the reference module's forward on the host cores (as dim_reduction.py:183 calls it: autograd
    graph built and dropped)."""
    torch.set_num_threads(os.cpu_count() or 1)
    ref = unet_reference_module(which, state)
    from iddiff_b200 import synthetic as SY
    _, spec, _ = SY.FULL_MODELS[which]
    x = torch.rand(n_images, spec["input_channels"], spec["image_size"], spec["image_size"])
    lab = torch.full((n_images,), label)
    ref(x[:1], lab[:1]).detach()
    t0 = time.perf_counter()
    y = ref(x, lab).detach()
    dt = time.perf_counter() - t0
    return n_images / dt, dt, torch.get_num_threads(), x, y
