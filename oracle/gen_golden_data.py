"""TEST INFRASTRUCTURE ONLY -- goldens for the paper's synthetic image manifolds, from the REAL reference
(lightning_data_modules/SyntheticDataset.py imported from /root/reference on CPU): FixedSquaresManifold and
FixedGaussiansManifold at the shipped configs' settings (32x32, 10 shapes, seed 42; squares/10.py:96-111,158,
gaussian_blobs/10.py:96-111,158) with a small data_samples.  Stored: every generated image.

  python oracle/gen_golden_data.py      -> tests/golden/synthetic_images.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import gen_golden as GG  # noqa: E402  (installs the shims and the reference on sys.path)
from iddiff_b200.config import ConfigDict  # noqa: E402


def main():
    from lightning_data_modules.SyntheticDataset import FixedSquaresManifold, FixedGaussiansManifold  # reference
    out = {}
    for seed in (42, 7):
        cfg = ConfigDict(seed=seed, data=dict(return_labels=False, data_samples=24, num_squares=10, square_range=[3, 5],
                                              num_gaussians=10, std_range=[1, 5], image_size=32))
        out[f"squares_{seed}"] = FixedSquaresManifold(cfg).data.numpy()
        out[f"gaussians_{seed}"] = FixedGaussiansManifold(cfg).data.numpy()
    np.savez_compressed(os.path.join(GG.GOLD, "synthetic_images.npz"), **out)
    for k, v in out.items():
        print(k, v.shape, v.dtype, float(v.min()), float(v.max()))


if __name__ == "__main__":
    main()
