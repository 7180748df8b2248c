"""GPU parity of the DSM training step on the sm_100a kernels (iddiff_b200/train.py, idb_train.cu)
against the reference's own objective evaluated by PyTorch autograd in fp64 / fp32:
  losses.py:164-188 (loss), models/utils.py:254-266 (VE: the network predicts the noise), losses.py:26-51 (warm-up, clip, Adam).
Tolerances: loss 1e-5 relative; gradients 3e-4 of the largest entry of each tensor (bf16-pair operands carry 2^-17 per
operand against fp32's 2^-24; measured ~2e-5); parameters after an optimiser step within 2 % of the step size."""
import copy

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from iddiff_b200 import models, sde_lib  # noqa: E402
from iddiff_b200.config import ConfigDict  # noqa: E402
from iddiff_b200.train import FcnTrainer, short_train  # noqa: E402


def _fcn(D, H, L, seed=0):
    torch.manual_seed(seed)
    cfg = ConfigDict(model=dict(name="fcn", state_size=D, hidden_layers=L, hidden_nodes=H, dropout=0.0, sigma_min=1e-2,
                                sigma_max=4.0, num_scales=1000, checkpoint_path=None),
                     training=dict(sde="vesde", batch_size=500), data=dict(use_data_mean=False))
    return models.create_model(cfg).cuda()


def _reference_loss(net, x0, t, z, sde, dtype):
    """losses.py:172-187 with models/utils.py:254-266 for the VE SDE, reduce_mean, likelihood_weighting False."""
    net = copy.deepcopy(net).to(dtype)
    x0, t, z = x0.to(dtype), t.to(dtype), z.to(dtype)
    mean, std = sde.marginal_prob(x0, t)
    xt = mean + std[:, None] * z
    labels = t * (sde.N - 1)
    out = net.mlp(torch.cat([xt, labels[:, None]], dim=1))          # models/fcn.py:35-40
    score = -out / std[:, None]
    losses = torch.square(score * std[:, None] + z).reshape(x0.shape[0], -1).mean(dim=-1)
    loss = losses.mean()
    loss.backward()
    lin = [m for m in net.mlp if isinstance(m, torch.nn.Linear)]
    return float(loss.detach()), [m.weight.grad for m in lin], [m.bias.grad for m in lin]


@pytest.mark.parametrize("D,H,L,B", [(24, 64, 2, 50), (200, 2048, 5, 500)])
def test_dsm_step_loss_and_gradients_vs_autograd(D, H, L, B):
    net = _fcn(D, H, L)
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    g = torch.Generator(device="cuda").manual_seed(3)
    x0 = torch.randn(B, D, generator=g, device="cuda")
    x0 = x0 / x0.norm(dim=1, keepdim=True)
    t = torch.rand(B, generator=g, device="cuda") * (1 - 1e-5) + 1e-5
    z = torch.randn(B, D, generator=g, device="cuda")
    tr = FcnTrainer(net, B, lr=2e-4, warmup=0, grad_clip=1.0)
    mean, std = sde.marginal_prob(torch.ones(B, 1, device="cuda"), t)
    tr.sample(x0, None, t, mean.reshape(-1).contiguous(), std.contiguous(), 999.0, z=z)
    tr.forward()
    loss = float(tr.loss_and_backward())
    torch.cuda.synchronize()
    ref_loss, gw, gb = _reference_loss(net, x0, t, z, sde, torch.float64)
    assert abs(loss - ref_loss) <= 1e-5 * abs(ref_loss), (loss, ref_loss)
    # the sampled input in the reference's fp32 arithmetic: x_t = mean + std * z, label = 999 t
    xt = x0 + std[:, None] * z
    assert torch.equal(tr.X[:B, :D], xt) and torch.equal(tr.X[:B, D], t * 999.0)
    assert float(tr.X[B:].abs().max() if tr.Mp > B else 0.0) == 0.0 and float(tr.X[:, D + 1:].abs().max()) == 0.0
    worst = 0.0
    for l in range(L + 2):
        n_out, n_in = tr.n_out[l], tr.n_in[l]
        dW = tr.dW[l][:n_out, :n_in].double()
        err = float((dW - gw[l]).abs().max() / gw[l].abs().max())
        errb = float((tr.db[l][:n_out].double() - gb[l]).abs().max() / gb[l].abs().max())
        worst = max(worst, err, errb)
        assert err < 3e-4 and errb < 3e-4, (l, err, errb)
        assert float(tr.dW[l][n_out:].abs().max() if tr.Pout[l] > n_out else 0.0) == 0.0     # padding carries no gradient
    print(f"D={D} H={H}: loss {loss:.6f} (fp64 {ref_loss:.6f}), worst relative gradient error {worst:.2e}")


def test_dsm_optimiser_steps_track_torch_adam():
    """20 steps with the reference's optimiser settings (Adam, warm-up 5, clip 1) on identical (batch, t, z) sequences:
    the loss trajectory follows PyTorch's to 2e-3 and the first update matches element-wise."""
    D, H, L, B = 24, 64, 2, 50
    net = _fcn(D, H, L, seed=1)
    ref = copy.deepcopy(net)
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    lr, warm = 1e-3, 5
    tr = FcnTrainer(net, B, lr=lr, warmup=warm, grad_clip=1.0)
    opt = torch.optim.Adam(ref.parameters(), lr=lr, betas=(0.9, 0.999), eps=1e-8)
    g = torch.Generator(device="cuda").manual_seed(5)
    data = torch.randn(400, D, generator=g, device="cuda")
    ours, theirs = [], []
    for step in range(1, 21):
        idx = torch.randint(0, 400, (B,), generator=g, device="cuda")
        t = torch.rand(B, generator=g, device="cuda") * (1 - 1e-5) + 1e-5
        z = torch.randn(B, D, generator=g, device="cuda")
        before = [p.detach().clone() for p in ref.parameters()]
        ours.append(float(tr.step(data, sde, idx=idx, t=t, z=z)))
        for gq in opt.param_groups:
            gq["lr"] = lr * min(step / warm, 1.0)
        x0 = data[idx]
        std = sde.marginal_prob(x0, t)[1]
        out = ref.mlp(torch.cat([x0 + std[:, None] * z, (t * 999.0)[:, None]], dim=1))
        loss = torch.square(out - z).mean()
        opt.zero_grad(set_to_none=True)
        loss.backward()
        grads = [p.grad.detach().clone() for p in ref.parameters()]
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        opt.step()
        theirs.append(float(loss))
        if step == 1:
            step_size = lr * (1 / warm)
            for p_o, p_r, b4, gr in zip(net.parameters(), ref.parameters(), before, grads):
                sel = gr.abs() > 1e-6                      # m / sqrt(v) = sign(g) at step 1: only defined away from g = 0
                assert float(((p_o - b4) - (p_r - b4))[sel].abs().max()) < 0.02 * step_size
    np.testing.assert_allclose(ours, theirs, rtol=2e-3)
    assert ours[-1] < ours[0]


def test_graph_replayed_step_equals_the_eager_step():
    """step_graph (one captured CUDA graph per step, counters on the device) against step() on the same draws: the torch
    generator is re-seeded so both runs see the same batch indices and times; Philox noise is keyed by (seed, step)."""
    D, H, L, B = 24, 64, 2, 50
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    data = torch.randn(400, D, generator=torch.Generator(device="cuda").manual_seed(9), device="cuda")
    outs = []
    for use_graph in (False, True):
        net = _fcn(D, H, L, seed=4)
        tr = FcnTrainer(net, B, lr=1e-3, warmup=3, grad_clip=1.0)
        torch.manual_seed(11)
        losses = []
        for _ in range(6):
            fn = tr.step_graph if use_graph else tr.step
            losses.append(float(fn(data, sde, seed=7)))
        outs.append((losses, [p.detach().clone() for p in net.parameters()]))
    assert outs[0][0] == outs[1][0]
    for a, b in zip(outs[0][1], outs[1][1]):
        assert torch.equal(a, b)


def test_cuda_short_training_recovers_the_sphere():
    """4000 DSM steps on the 50-sphere in R^200 entirely on the CUDA training step, then the estimation engine: every
    anchor reads ID 51 = k + 1 (what tools/train_fcn.py reaches with the PyTorch loop)."""
    from iddiff_b200 import data, engine as E
    from iddiff_b200.dim_reduction import estimate_anchors
    net = _fcn(200, 2048, 5)
    torch.manual_seed(2)
    pts = data.ksphere_points(50016, 1, 200, 50, 0.0, "random_isometry").float()
    import time
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run = short_train(net, pts[:50000].cuda(), 4000, batch=500, lr=2e-4, warmup=500)
    torch.cuda.synchronize()
    print(f"4000 steps in {time.perf_counter() - t0:.2f} s")
    assert run is not None and run < 0.5          # from ~1.0 at initialisation
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    r = estimate_anchors(E.FcnEngine.from_module(net, device="cuda"), pts[50000:].contiguous(), 1501, sde, 1e-5, seed=1)
    dims = r["dims"].tolist()
    print("loss", run, "dims", dims)
    assert sorted(dims)[len(dims) // 2] in (51, 52) and max(dims) <= 60
