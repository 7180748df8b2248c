#!/usr/bin/env python
"""Score-evals/s and ID-estimates/s of the full-size image score networks on one B200 (BASELINE configs 4-5):
  ddpm     style_gan_base.py (model.name='ddpm'): nf=128, ch_mult (1,2,2,3,3), 4 res-blocks, attn@16, 116.1 M params, 79.7 GFLOP/img
  ncsnpp   same with model.name='ncsnpp' (BASELINE config 4): 126.3 M params, 88.6 GFLOP/img
  beatgans style_gan_64d_BeatGAN.py (BASELINE config 5): 87.5 M params, 37.4 GFLOP/img
(FLOPs per image: SURVEY 8a, FlopCounterMode of the reference.)  Random-init weights with the zero-initialised
convolutions re-randomised (appendix A.11); synthetic 64x64x3 images.  Development / profiling tool -- the headline
bench is bench.py.  Prints one JSON line.

  python tools/bench_unet.py --model ncsnpp --batch 128               # forward only
  python tools/bench_unet.py --model ncsnpp --anchors 1 --K 4096      # whole path: perturb -> net -> Gram -> spectrum -> ID
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from iddiff_b200 import synthetic as SY  # noqa: E402

MODELS = SY.FULL_MODELS


def build(model):
    return SY.build_full_engine(model, "cuda:0")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="ddpm", choices=list(MODELS))
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--anchors", type=int, default=0, help="> 0: time the whole manifold_dimension path")
    ap.add_argument("--K", type=int, default=4096)
    ap.add_argument("--cpu-baseline", type=int, default=0, help="time the oracle's CPU forward (reference arithmetic, "
                    "torch CPU fp32, all host cores) on this many images and report score-evals/s")
    args = ap.parse_args()
    from iddiff_b200 import _lib, engine as E
    eng, n_params, gflop = build(args.model)
    out = {"model": args.model, "params_M": n_params / 1e6, "gflop_per_image": gflop}
    C, S, _ = eng.image_shape
    D = C * S * S
    if gflop is None:   # measured from the oracle's forward with torch's FLOP counter (same arithmetic as the reference)
        from torch.utils.flop_counter import FlopCounterMode
        from oracle import img_oracle as IO
        state = {k: v.detach().cpu() for k, v in eng.h.state_dict().items()}
        with FlopCounterMode(display=False) as fc, torch.no_grad():
            IO.ddpm_forward(state, MODELS[args.model][1], torch.rand(1, C, S, S), torch.full((1,), 0.00999))
        gflop = out["gflop_per_image"] = fc.get_total_flops() / 1e9
    if args.anchors == 0:
        x = torch.rand(args.batch, C, S, S, device="cuda")
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
        out.update({"batch": args.batch, "ms_per_forward": ms, "score_evals_per_s": args.batch / ms * 1e3,
                    "tflops_algorithmic": args.batch * gflop / ms, "launches_per_forward":
                    (_lib.launch_count - l0) / args.steps, "finite": bool(torch.isfinite(y).all())})
    else:
        from iddiff_b200 import sde_lib
        sde = sde_lib.VESDE(sigma_min=0.009, sigma_max=50.0, N=1000)   # style_gan_base.py:100-101
        mc, std = sde.coefficients(1e-5)
        x0 = SY.smooth_images(args.anchors, C, S, 5).reshape(args.anchors, -1).cuda()
        eng.estimate(x0[:1], min(args.K, 2 * eng.IMAGE_CHUNK), mc, std, 0.00999, seed=0)   # warm-up (captures the forward graph)
        torch.cuda.synchronize()
        E.EVENT_LOG = {}
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        r = eng.estimate(x0, args.K, mc, std, 0.00999, seed=1)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        stages = {k: sum(a.elapsed_time(b) for a, b, _ in v) for k, v in E.EVENT_LOG.items()}
        E.EVENT_LOG = None
        out.update({"anchors": args.anchors, "K": args.K, "D": D, "n": min(args.K, D), "ms_total": ms,
                    "id_estimates_per_s": args.anchors / ms * 1e3, "score_evals_per_s": args.anchors * args.K / ms * 1e3,
                    "score_net_tflops_algorithmic": args.anchors * args.K * gflop / stages["score_net"],
                    "stages_ms": stages, "dims": r.dims.cpu().tolist(),
                    "sv_head": r.singular_values[0, :3].cpu().tolist(), "sv_tail": r.singular_values[0, -3:].cpu().tolist()})
    if args.cpu_baseline:
        import time
        from oracle import img_oracle as IO
        name, spec, _ = MODELS[args.model]
        fwd = {"squares": IO.ddpm_forward, "ddpm": IO.ddpm_forward, "ncsnpp": IO.ncsnpp_forward, "beatgans": IO.beatgans_forward}[args.model]
        state = {k: v.detach().cpu() for k, v in eng.h.state_dict().items()}
        torch.set_num_threads(os.cpu_count() or 1)
        xb = torch.rand(args.cpu_baseline, C, S, S)
        lab = torch.full((args.cpu_baseline,), 0.00999)
        with torch.no_grad():
            fwd(state, spec, xb[:1], lab[:1])
            t0 = time.perf_counter()
            yc = fwd(state, spec, xb, lab)
            dt = time.perf_counter() - t0
        yg = eng.forward(xb.cuda(), 0.00999).permute(0, 3, 1, 2).cpu()
        out["cpu_baseline"] = {"value": args.cpu_baseline / dt, "unit": "score-evals/s", "cores": torch.get_num_threads(),
                               "kind": "port", "sample": f"{args.cpu_baseline} images, one forward of the oracle restatement "
                               f"of the reference network (torch CPU fp32) in {dt:.1f} s",
                               "gpu_vs_cpu_max_abs_diff": float((yg - yc).abs().max()), "out_absmax": float(yc.abs().max())}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
