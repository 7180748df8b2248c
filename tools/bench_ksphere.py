#!/usr/bin/env python
"""BASELINE configs[0] (the reference's own CPU-runnable case): 10-sphere embedded in R^100, analytic `ksphere_gt`
score (models/ksphere_gt.py:14-26), VE SDE sigma_min 1e-2 / sigma_max 4, t = 1e-5, K = 400 perturbations per anchor.
ID-estimates/s of the CUDA path on one B200 for N anchors (inputs resident), and -- with --cpu-baseline n -- the CPU
oracle of the same path (the reference's per-row scipy loop + torch.linalg.svd) on n anchors of the same workload.
Prints one JSON line.  Development tool; the headline bench is bench.py."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--anchors", type=int, default=4096)
    ap.add_argument("--K", type=int, default=400)
    ap.add_argument("--chunk", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--cpu-baseline", type=int, default=0)
    args = ap.parse_args()
    from iddiff_b200 import data, sde_lib
    from iddiff_b200.engine import KSphereEngine
    D, p = 100, 10
    torch.manual_seed(1)
    x0 = data.ksphere_points(args.anchors, 1, D, p, 0.0, "random_isometry").float().contiguous()
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    mc, std = sde.coefficients(1e-5)
    eng = KSphereEngine(p, "cuda:0")
    xd = x0.cuda()

    def step(seed):
        out = []
        for s in range(0, args.anchors, args.chunk):
            out.append(eng.estimate(xd[s:s + args.chunk], args.K, mc, std, seed=seed, anchor_id0=s))
        return out

    step(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        res = step(1 + i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    dims = torch.cat([r.dims for r in res]).cpu()
    sv = res[0].singular_values[0].cpu()
    out = {"workload": "BASELINE configs[0]: 10-sphere in R^100, ksphere_gt, K=400", "anchors": args.anchors,
           "K": args.K, "ms_per_step": ms, "id_estimates_per_s": args.anchors / ms * 1e3,
           "score_evals_per_s": args.anchors * args.K / ms * 1e3, "sv_anchor0": [float(sv[0]), float(sv[1]), float(sv[-1])],
           "dims_head": dims[:8].tolist()}
    if args.cpu_baseline:
        from oracle import id_oracle as O
        n = args.cpu_baseline
        torch.set_num_threads(os.cpu_count() or 1)
        params = dict(sigma_min=1e-2, sigma_max=4.0)
        score_fn = O.make_score_fn("ksphere_gt", "vesde", params, manifold_dim=p)
        t0 = time.perf_counter()
        cpu_dims = []
        for a in range(n):
            vec_t = torch.ones(args.K) * 1e-5
            mean, sd = O.marginal("vesde", x0[a].repeat(args.K, 1), vec_t, **params)
            xs = mean + sd[:, None] * torch.randn(args.K, D)
            cpu_dims.append(O.id_from_spectrum(O.centred_spectrum(score_fn(xs, vec_t))))
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": n / dt, "unit": "anchors/s", "cores": torch.get_num_threads(), "kind": "port",
                               "sample": f"{n} anchors of the same workload in {dt:.1f} s (oracle port: per-row scipy "
                                         "Bessel loop + torch.linalg.svd)", "dims_head": cpu_dims[:8]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
