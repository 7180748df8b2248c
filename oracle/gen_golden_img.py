"""TEST INFRASTRUCTURE ONLY -- golden vectors for the image-network side, produced by running the REAL reference
(/root/reference) on CPU in the build container:   python oracle/gen_golden_img.py [ops] [ddpm] [ncsnpp] [beatgans]

The reference's `op` package JIT-compiles its CUDA extensions at import (op/upfirdn2d.py:9-16, op/fused_act.py:9-16);
there is no GPU here, so torch.utils.cpp_extension.load is replaced by a stub and the reference's own pure-PyTorch CPU
branches run (op/upfirdn2d.py:146-149 -> upfirdn2d_native, op/fused_act.py:87-94) -- they are the FIR oracle.
/root/reference is read, never written.
"""
import os
import sys
from unittest import mock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_shims  # noqa: E402

ref_shims.install()
for name in ["matplotlib", "matplotlib.pyplot", "matplotlib.cm", "matplotlib.colors", "lpips", "iunets",
             "iunets.layers", "iunets.networks"]:
    sys.modules.setdefault(name, mock.MagicMock())
import torch.utils.cpp_extension as _cpp  # noqa: E402

_cpp.load = lambda *a, **k: mock.MagicMock()  # no nvcc JIT: CPU branches of the reference ops only

GOLD = os.path.join(ROOT, "tests", "golden")

# (name, N, C, H, W, kernel taps 1-D, gain, up, down, pad0, pad1)
OPS_CASES = [
    ("up2_fir", 2, 3, 8, 8, (1, 3, 3, 1), 4.0, 2, 1, 2, 1),        # up_or_down_sampling.py:181-214 upsample_2d
    ("down2_fir", 2, 3, 16, 16, (1, 3, 3, 1), 1.0, 1, 2, 1, 1),    # up_or_down_sampling.py:217-257 downsample_2d
    ("convdown_pad", 1, 5, 9, 11, (1, 3, 3, 1), 1.0, 1, 1, 2, 2),  # conv_downsample_2d pre-filter (:144-178)
    ("odd_k3", 1, 2, 7, 5, (1, 2, 1), 1.0, 3, 2, 1, 0),
    ("crop", 1, 2, 10, 10, (1, 3, 3, 1), 1.0, 1, 1, -1, -2),
    ("wide", 1, 1, 4, 70, (1, 3, 3, 1), 4.0, 2, 1, 2, 1),
]


def gen_ops():
    from op.upfirdn2d import upfirdn2d as ref_upfirdn2d
    from op.fused_act import fused_leaky_relu as ref_flr
    out = {}
    g = torch.Generator().manual_seed(11)
    for name, n, c, h, w, taps, gain, up, down, p0, p1 in OPS_CASES:
        x = torch.randn(n, c, h, w, generator=g)
        k1 = torch.tensor(taps, dtype=torch.float32)
        k = torch.outer(k1, k1)
        k = k / k.sum() * gain
        y = ref_upfirdn2d(x, k, up=up, down=down, pad=(p0, p1))
        out[f"{name}_x"], out[f"{name}_k"], out[f"{name}_y"] = x.numpy(), k.numpy(), y.numpy()
        out[f"{name}_args"] = np.array([up, down, p0, p1])
    x = torch.randn(3, 6, 5, 7, generator=g)
    b = torch.randn(6, generator=g)
    out["flr_x"], out["flr_b"] = x.numpy(), b.numpy()
    out["flr_y"] = ref_flr(x, b, 0.2, 2 ** 0.5).numpy()
    np.savez_compressed(os.path.join(GOLD, "ops_upfirdn_fused.npz"), **out)
    print("ops goldens written:", len(out), "arrays")


if __name__ == "__main__":
    what = sys.argv[1:] or ["ops"]
    if "ops" in what:
        gen_ops()
    if any(w in what for w in ("ddpm", "ncsnpp", "beatgans")):
        from oracle import gen_golden_unet
        gen_golden_unet.main(what)
