"""GPU parity of idb_upfirdn2d / idb_fused_bias_act (through the C ABI via iddiff_b200.op) against the reference's
outputs (golden) and the numpy oracle, including the edge cases the op admits: negative pads, up and down together,
minor > 1, non-square FIR, large shapes (size-independent property: linearity + separability)."""
import os

import numpy as np
import pytest
import torch

from oracle import img_oracle as IO

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "ops_upfirdn_fused.npz")
CASES = ["up2_fir", "down2_fir", "convdown_pad", "odd_k3", "crop", "wide"]


@pytest.mark.parametrize("name", CASES)
def test_upfirdn2d_vs_reference_golden(name):
    from iddiff_b200 import op
    g = np.load(GOLD)
    x, k, y = g[f"{name}_x"], g[f"{name}_k"], g[f"{name}_y"]
    up, down, p0, p1 = (int(v) for v in g[f"{name}_args"])
    out = op.upfirdn2d(torch.from_numpy(x).cuda(), torch.from_numpy(k).cuda(), up=up, down=down, pad=(p0, p1))
    assert tuple(out.shape) == y.shape
    np.testing.assert_allclose(out.cpu().numpy(), y, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("shape,kshape,args", [
    ((3, 9, 13, 4), (4, 4), (2, 2, 1, 1, 2, 1, 2, 1)),
    ((2, 17, 5, 1), (3, 2), (1, 3, 2, 1, 0, 2, 1, 0)),
    ((1, 33, 70, 7), (4, 4), (1, 1, 2, 2, 1, 1, 1, 1)),
    ((2, 6, 6, 2), (2, 2), (2, 2, 3, 3, -1, 0, 0, -2)),
    ((1, 40, 40, 130), (4, 4), (2, 2, 1, 1, 2, 1, 2, 1)),
])
def test_upfirdn2d_op_vs_oracle(shape, kshape, args):
    from iddiff_b200 import op
    rng = np.random.default_rng(5)
    x = rng.standard_normal(shape).astype(np.float32)
    k = rng.standard_normal(kshape).astype(np.float32)
    want = IO.upfirdn2d_np(x.astype(np.float64), k.astype(np.float64), *args)
    out = op.upfirdn2d_op(torch.from_numpy(x).cuda(), torch.from_numpy(k).cuda(), *args)
    assert tuple(out.shape) == want.shape
    np.testing.assert_allclose(out.cpu().numpy(), want, rtol=2e-5, atol=2e-6)


def test_upfirdn2d_full_size_properties():
    """StyleGAN-size activation (128 x 128 x 64 x 64): up-by-2 FIR of a constant image is the constant away from the
    border (FIR gain 4 = up^2), and the op is linear."""
    from iddiff_b200 import op
    k1 = torch.tensor([1., 3., 3., 1.])
    k = (torch.outer(k1, k1) / 64 * 4).cuda()
    x = torch.full((16, 128, 64, 64), 0.75, device="cuda")
    y = op.upfirdn2d(x, k, up=2, pad=(2, 1))
    assert y.shape == (16, 128, 128, 128)
    assert torch.allclose(y[:, :, 2:-2, 2:-2], torch.full_like(y[:, :, 2:-2, 2:-2], 0.75), rtol=1e-6)
    a, b = torch.randn(4, 8, 64, 64, device="cuda"), torch.randn(4, 8, 64, 64, device="cuda")
    lhs = op.upfirdn2d(a + 2 * b, k, down=2, pad=(1, 1))
    rhs = op.upfirdn2d(a, k, down=2, pad=(1, 1)) + 2 * op.upfirdn2d(b, k, down=2, pad=(1, 1))
    assert torch.allclose(lhs, rhs, rtol=1e-4, atol=1e-5)


def test_fused_leaky_relu_vs_reference_golden():
    from iddiff_b200 import op
    g = np.load(GOLD)
    out = op.fused_leaky_relu(torch.from_numpy(g["flr_x"]).cuda(), torch.from_numpy(g["flr_b"]).cuda(), 0.2, 2 ** 0.5)
    np.testing.assert_allclose(out.cpu().numpy(), g["flr_y"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("shape", [(3, 6, 5, 7), (2, 8, 4, 4), (5, 3), (1, 4, 16, 16, 2)])
@pytest.mark.parametrize("act,grad", [(1, 0), (1, 1), (1, 2), (3, 0), (3, 1), (3, 2)])
def test_fused_bias_act_table(shape, act, grad):
    from iddiff_b200 import op
    rng = np.random.default_rng(3)
    x = rng.standard_normal(shape).astype(np.float32)
    b = rng.standard_normal(shape[1]).astype(np.float32)
    ref = rng.standard_normal(shape).astype(np.float32)
    want = IO.fused_bias_act_np(x, b, ref, act, grad, 0.3, 1.7)
    out = op.fused_bias_act(torch.from_numpy(x).cuda(), torch.from_numpy(b).cuda(), torch.from_numpy(ref).cuda(), act,
                            grad, 0.3, 1.7)
    np.testing.assert_array_equal(out.cpu().numpy(), want)
    nob = op.fused_bias_act(torch.from_numpy(x).cuda(), torch.empty(0, device="cuda"), torch.empty(0, device="cuda"),
                            act, 0, 0.3, 1.7)
    np.testing.assert_array_equal(nob.cpu().numpy(), IO.fused_bias_act_np(x, None, None, act, 0, 0.3, 1.7))


def test_ops_reject_cpu_tensors():
    from iddiff_b200 import op, _lib
    with pytest.raises(_lib.IdbError):
        op.upfirdn2d(torch.zeros(1, 1, 4, 4), torch.ones(2, 2))
    with pytest.raises(_lib.IdbError):
        op.fused_leaky_relu(torch.zeros(1, 2, 4), torch.zeros(2))
