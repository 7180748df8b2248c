#!/usr/bin/env python
"""DSM training step of the FCN (BASELINE configs[1] network: D = 200, H = 2048, L = 5, batch 500): ms per step of
  cuda-eager   iddiff_b200/train.py, one C-ABI call per kernel
  cuda-graph   the same step replayed as one CUDA graph
  torch        the reference's loop in PyTorch (tools/train_fcn.short_train; TF32 matmuls, eager autograd + Adam)
Prints one JSON line.  `--steps-only N`: just N eager CUDA steps (for an ncu launch list)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--steps-only", type=int, default=0)
    args = ap.parse_args()
    import bench
    from iddiff_b200 import sde_lib
    from iddiff_b200.train import FcnTrainer
    dev = torch.device("cuda", 0)
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    pts = bench.make_anchors(50000, seed=2).to(dev)

    def timed(fn, n):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e3

    tr = FcnTrainer(bench.make_model().to(dev), 500, lr=2e-4, warmup=500, grad_clip=1.0)
    if args.steps_only:
        for _ in range(args.steps_only):
            tr.step(pts, sde)
        torch.cuda.synchronize()
        return
    out = {"workload": "DSM training step, FCN D=200 H=2048 L=5, batch 500, Adam + warm-up + clip",
           "cuda_eager_ms": timed(lambda: tr.step(pts, sde), args.steps),
           "cuda_graph_ms": timed(lambda: tr.step_graph(pts, sde), args.steps)}
    import train_fcn
    m = bench.make_model().to(dev)
    train_fcn.short_train(m, pts, 50)          # cuBLAS / autograd warm-up outside the timed loop
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    train_fcn.short_train(m, pts, args.steps)
    torch.cuda.synchronize()
    out["torch_ms"] = (time.perf_counter() - t0) / args.steps * 1e3
    print(json.dumps(out))


if __name__ == "__main__":
    main()
