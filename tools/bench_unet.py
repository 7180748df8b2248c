#!/usr/bin/env python
"""Score-evals/s of the full-size DDPM U-Net (style_gan_base.py: nf=128, ch_mult (1,2,2,3,3), 4 res-blocks, attention at
16x16, 64x64x3, 116 M parameters, 79.7 GFLOP/image by SURVEY 8a a10) on one B200, with a per-stage breakdown.
Development tool (the headline bench is bench.py); prints one JSON line."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

FULL = dict(nf=128, ch_mult=(1, 2, 2, 3, 3), num_res_blocks=4, attn_resolutions=(16,), image_size=64,
            resamp_with_conv=True, centered=False, input_channels=3, output_channels=3, dropout=0.0, conditional=True,
            nonlinearity="swish")
GFLOP_PER_IMAGE = 79.7


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()
    from oracle import fixtures as FX
    from iddiff_b200 import unet_models, _lib
    from iddiff_b200.unet_engine import UNetEngine
    holder = unet_models.DDPM(FX.unet_config(FULL))
    with torch.no_grad():  # re-randomise the zero-initialised convolutions (appendix A.11)
        for k, p in holder.named_parameters():
            if p.dim() >= 2 and float(p.abs().max()) < 1e-3:
                unet_models.fan_avg_uniform_(p, 1.0)
    n_params = sum(p.numel() for p in holder.parameters())
    eng = UNetEngine.from_module(holder, "cuda:0")
    x = torch.rand(args.batch, 3, 64, 64, device="cuda")
    for _ in range(args.warmup):
        eng.forward(x, 0.00999)
    torch.cuda.synchronize()
    l0 = _lib.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        y = eng.forward(x, 0.00999)
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / args.steps
    print(json.dumps({"model": "ddpm full (style_gan_base)", "params_M": n_params / 1e6, "batch": args.batch,
                      "ms_per_forward": ms, "images_per_s": args.batch / ms * 1e3,
                      "tflops_algorithmic": args.batch * GFLOP_PER_IMAGE / ms, "launches_per_forward":
                      (_lib.launch_count - l0) / args.steps, "finite": bool(torch.isfinite(y).all()),
                      "out_absmax": float(y.abs().max())}))


if __name__ == "__main__":
    main()
