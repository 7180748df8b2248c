#!/usr/bin/env python
"""Short-trained FCN score network for BASELINE configs 2-3 ("random-init and a short-trained variant"):
no checkpoint exists offline, so the weights are produced on the box with the reference's own objective --
denoising score matching for the VE SDE, losses.py:164-188 with models/utils.py:254-266: the network output is the
noise prediction, loss = mean_batch mean_D (model(x0 + sigma(t) z, 999 t) - z)^2, t ~ U(eps, 1), Adam
(configs/.../ksphere/50dim.py:109-114: lr 2e-5, warm-up, gradient clipping at 1) -- in plain PyTorch (training is
out of scope for the CUDA path; only the optimiser loop runs in torch), then handed to the sm_100a engine for the
intrinsic-dimension estimates.  Prints one JSON line per evaluation point and writes a Lightning-style checkpoint
(state_dict keys prefixed 'score_model.') that main.py --mode manifold_dimension --checkpoint_path loads.

  python tools/train_fcn.py --steps 20000 --eval-every 5000 --out gpurun_out/fcn_c2_short.ckpt
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def short_train(holder, train_points, steps, *, batch=500, lr=2e-4, warmup=500, sigma_min=1e-2, sigma_max=4.0,
                eps_t=1e-5, on_step=None):
    """Denoising score matching on `train_points` [n, D] (device) for `steps` Adam steps; updates holder in place.
    Returns the smoothed final loss."""
    net = holder.mlp                                     # nn.Sequential(Linear, Dropout(0), ELU, ...): fcn.py:16-30
    dev = train_points.device
    opt = torch.optim.Adam(net.parameters(), lr=lr, betas=(0.9, 0.999), eps=1e-8)
    tf32 = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True         # the optimiser loop only; the estimates never use torch
    log_ratio = float(torch.log(torch.tensor(sigma_max / sigma_min)))
    run = None
    try:
        for step in range(1, steps + 1):
            for g in opt.param_groups:
                g["lr"] = lr * min(1.0, step / max(warmup, 1))
            x0 = train_points[torch.randint(0, train_points.shape[0], (batch,), device=dev)]
            t = torch.rand(batch, device=dev) * (1.0 - eps_t) + eps_t
            std = sigma_min * torch.exp(t * log_ratio)
            z = torch.randn_like(x0)
            out = net(torch.cat([x0 + std[:, None] * z, (t * 999.0)[:, None]], dim=1))
            loss = torch.square(out - z).mean()
            opt.zero_grad(set_to_none=True)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(net.parameters(), 1.0)
            opt.step()
            if step % 50 == 0:
                v = float(loss.detach())
                run = v if run is None else 0.9 * run + 0.1 * v
            if on_step is not None:
                on_step(step, run)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = tf32
    return run


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--D", type=int, default=200)
    ap.add_argument("--manifold-dim", default="50", help="k, or k1,k2,... for a union of spheres (BASELINE config 3)")
    ap.add_argument("--angle-std", type=float, default=-1, help="> 0: non-uniform density on the sphere (KSphereDataset.py:92-94)")
    ap.add_argument("--hidden", type=int, default=2048)
    ap.add_argument("--layers", type=int, default=5)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--batch", type=int, default=500)
    ap.add_argument("--lr", type=float, default=2e-4)
    ap.add_argument("--warmup", type=int, default=500)
    ap.add_argument("--train-points", type=int, default=100000)
    ap.add_argument("--eval-every", type=int, default=5000)
    ap.add_argument("--eval-anchors", type=int, default=32)
    ap.add_argument("--K", type=int, default=1501)
    ap.add_argument("--sigma-min", type=float, default=1e-2)
    ap.add_argument("--sigma-max", type=float, default=4.0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default="")
    ap.add_argument("--torch", action="store_true", help="run the optimiser loop in PyTorch (the reference's own loop) "
                    "instead of the CUDA training step of iddiff_b200/train.py")
    args = ap.parse_args()

    from iddiff_b200 import data, engine as E, models, sde_lib
    from iddiff_b200.config import ConfigDict
    from iddiff_b200.dim_reduction import estimate_anchors
    dev = torch.device("cuda", 0)
    torch.manual_seed(args.seed)
    cfg = ConfigDict(model=dict(name="fcn", state_size=args.D, hidden_layers=args.layers, hidden_nodes=args.hidden,
                                dropout=0.0, sigma_min=args.sigma_min, sigma_max=args.sigma_max, num_scales=1000,
                                checkpoint_path=None),
                     training=dict(sde="vesde", batch_size=args.batch), data=dict(use_data_mean=False))
    holder = models.create_model(cfg).to(dev)           # owns the weights in the reference's state_dict layout
    mdims = [int(v) for v in str(args.manifold_dim).split(",")]
    S = len(mdims)
    per_train, per_eval = args.train_points // S, args.eval_anchors // S
    pts = data.ksphere_points(per_train + per_eval, S, args.D, mdims if S > 1 else mdims[0], 0.0, "random_isometry",
                              angle_std=args.angle_std).float().reshape(S, per_train + per_eval, args.D)
    train = pts[:, :per_train].reshape(-1, args.D).to(dev)
    anchors = pts[:, per_train:].reshape(-1, args.D).contiguous()      # component-major
    sde = sde_lib.VESDE(sigma_min=args.sigma_min, sigma_max=args.sigma_max, N=1000)
    eps_t = 1e-5

    def evaluate(step, loss, t_train):
        eng = E.FcnEngine.from_module(holder, device=dev)
        r = estimate_anchors(eng, anchors, args.K, sde, eps_t, seed=step)
        sv = r["singular_values"]
        comps = []
        for c, k in enumerate(mdims):
            dims = r["dims"][c * per_eval:(c + 1) * per_eval].tolist()
            a = c * per_eval
            comps.append({"true_manifold_dim": k, "dims_median": sorted(dims)[len(dims) // 2], "dims_min": min(dims),
                          "dims_max": max(dims), "dims_head": dims[:8],
                          "sv_first_anchor": {"0": float(sv[a, 0]), str(args.D - k - 2): float(sv[a, args.D - k - 2]),
                                              str(args.D - k - 1): float(sv[a, args.D - k - 1]),
                                              str(args.D - k): float(sv[a, args.D - k]), "last": float(sv[a, -1])}})
        print(json.dumps({"step": step, "loss": loss, "train_s": t_train, "D": args.D, "components": comps}), flush=True)

    evaluate(0, None, 0.0)
    t0 = time.perf_counter()

    def on_step(step, run):
        if step % args.eval_every == 0 or step == args.steps:
            torch.cuda.synchronize()
            evaluate(step, run, time.perf_counter() - t0)

    if args.torch:
        train_fn = short_train
    else:
        from iddiff_b200.train import short_train as train_fn
    train_fn(holder, train, args.steps, batch=args.batch, lr=args.lr, warmup=args.warmup, sigma_min=args.sigma_min,
             sigma_max=args.sigma_max, eps_t=eps_t, on_step=on_step)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        torch.save({"state_dict": {"score_model." + k: v.detach().cpu() for k, v in holder.state_dict().items()},
                    "config": {"D": args.D, "manifold_dim": mdims, "angle_std": args.angle_std, "steps": args.steps, "lr": args.lr,
                               "batch": args.batch, "seed": args.seed}}, args.out)


if __name__ == "__main__":
    main()
