"""CPU: the numpy restatement of the two native ops (oracle/img_oracle.py) against the reference's own outputs
(tests/golden/ops_upfirdn_fused.npz, produced by oracle/gen_golden_img.py from op/upfirdn2d.py, op/fused_act.py)."""
import os

import numpy as np
import pytest

from oracle import img_oracle as IO

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ops_upfirdn_fused.npz")
CASES = ["up2_fir", "down2_fir", "convdown_pad", "odd_k3", "crop", "wide"]


@pytest.mark.parametrize("name", CASES)
def test_upfirdn2d_oracle_matches_reference(name):
    g = np.load(GOLD)
    x, k, y = g[f"{name}_x"], g[f"{name}_k"], g[f"{name}_y"]
    up, down, p0, p1 = (int(v) for v in g[f"{name}_args"])
    n, c, h, w = x.shape
    out = IO.upfirdn2d_np(x.reshape(n * c, h, w, 1), k, up, up, down, down, p0, p1, p0, p1)
    assert out.shape == (n * c, y.shape[2], y.shape[3], 1)
    np.testing.assert_allclose(out.reshape(y.shape), y, rtol=1e-5, atol=1e-6)


def test_fused_leaky_relu_oracle_matches_reference():
    g = np.load(GOLD)
    out = IO.fused_leaky_relu_cpu_np(g["flr_x"], g["flr_b"], 2 ** 0.5)
    np.testing.assert_allclose(out, g["flr_y"], rtol=1e-6, atol=1e-7)


def test_upfirdn2d_oracle_identity_and_minor():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, 5, 6, 3)).astype(np.float32)
    one = np.ones((1, 1), np.float32)
    np.testing.assert_array_equal(IO.upfirdn2d_np(x, one, 1, 1, 1, 1, 0, 0, 0, 0), x)
    k = rng.standard_normal((3, 2)).astype(np.float32)
    full = IO.upfirdn2d_np(x, k, 2, 1, 1, 2, 1, 2, 0, 1)
    for c in range(3):  # minor channels are independent
        np.testing.assert_array_equal(full[..., c:c + 1], IO.upfirdn2d_np(x[..., c:c + 1], k, 2, 1, 1, 2, 1, 2, 0, 1))
