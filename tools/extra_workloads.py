"""The other BASELINE configs, measured inside bench.py's one driver-timed command (`extra_workloads` of its JSON line):

  C1  10-sphere in R^100, analytic ksphere_gt score, K = 400 (BASELINE configs[0], the reference's own CPU case)
  C4  64x64x3 images, NCSN++ U-Net at full size (126.3 M parameters, 88.6 GFLOP per image), K = 4096, one anchor
  C5  64x64x3 images, BeatGANs U-Net at full size (87.5 M, 37.4 GFLOP), K = 4096, one anchor; at N > 1 GPUs the same
      anchor with its K rows sharded over the ranks (partial Grams + one all-reduce)

Each entry: whole-path time with inputs resident in HBM (CUDA events), score-evals/s, stage times, a roofline object
for the score network (ALGORITHMIC FLOPs: K x GFLOP per image forward, SURVEY 6 / 8a, over the CUDA-event time of the
score-network stage, against MEASURED_PEAKS.json bf16_tflops_sustained) and a cpu_baseline: the reference's own module /
path on the host cores (kind "reference" when oracle/_ref or /root/reference is present, else the oracle port).
Random-init weights with the zero-initialised output convolutions re-randomised; synthetic anchors.
"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = fn()
    e1.record()
    torch.cuda.synchronize()
    return r, e0.elapsed_time(e1)


def ksphere_c1(dev, anchors=4096, K=400, chunk=1024, steps=3, cpu_anchors=24):
    from iddiff_b200 import data, sde_lib
    from iddiff_b200.engine import KSphereEngine
    D, p = 100, 10
    torch.manual_seed(1)
    x0 = data.ksphere_points(anchors, 1, D, p, 0.0, "random_isometry").float().contiguous()
    sde = sde_lib.VESDE(sigma_min=1e-2, sigma_max=4.0, N=1000)
    mc, std = sde.coefficients(1e-5)
    eng = KSphereEngine(p, dev)
    xd = x0.to(dev)

    def step(seed):
        return [eng.estimate(xd[s:s + chunk], K, mc, std, seed=seed, anchor_id0=s) for s in range(0, anchors, chunk)]

    step(0)
    res, ms = _timed(lambda: [step(1 + i) for i in range(steps)][-1])
    ms /= steps
    dims = torch.cat([r.dims for r in res]).cpu()
    out = {"workload": "BASELINE configs[0]: 10-sphere in R^100, analytic ksphere_gt score, VE sigma_min=0.01 t=1e-5, "
                       f"K={K} perturbations per anchor, {anchors} anchors per step",
           "ms_per_step": ms, "id_estimates_per_s": anchors / ms * 1e3, "score_evals_per_s": anchors * K / ms * 1e3,
           "dims_head": dims[:8].tolist(),
           "roofline": {"bound": "hbm", "kernel": "perturb + ksphere_score + centre/Gram (K x D fp32 rows written, read, read)",
                        "achieved": 3.0 * anchors * K * D * 4 / (ms / 1e3) / 1e9, "unit": "GB/s"}}
    # CPU: the reference's per-row scipy Bessel loop + torch.linalg.svd (oracle port of the same lines; the reference's
    # entry point needs its KSphere datamodule + Lightning module, timed for configs[1] in the headline arm instead)
    from oracle import id_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    params = dict(sigma_min=1e-2, sigma_max=4.0)
    score_fn = O.make_score_fn("ksphere_gt", "vesde", params, manifold_dim=p)
    t0 = time.perf_counter()
    for a in range(cpu_anchors):
        vec_t = torch.ones(K) * 1e-5
        mean, sd = O.marginal("vesde", x0[a].repeat(K, 1), vec_t, **params)
        xs = mean + sd[:, None] * torch.randn(K, D)
        O.id_from_spectrum(O.centred_spectrum(score_fn(xs, vec_t)))
    dt = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": cpu_anchors / dt, "unit": "anchors/s", "cores": torch.get_num_threads(), "kind": "port",
                           "sample": f"{cpu_anchors} anchors of the same workload in {dt:.1f} s (models/ksphere_gt.py:14-26 "
                                     "per-row scipy loop + torch.linalg.svd, oracle port)"}
    return out


def _cpu_forward_baseline(which, holder, n_images):
    state = {k: v.detach().cpu() for k, v in holder.state_dict().items()}
    from oracle import ref_runner as RR
    if RR.available():
        v, dt, cores, _, _ = RR.time_unet_forward(which, state, n_images)
        kind, what = "reference", "the reference's own nn.Module forward (autograd graph built, as dim_reduction.py:183 calls it)"
    else:
        from oracle import img_oracle as IO
        from iddiff_b200 import synthetic as SY
        _, spec, _ = SY.FULL_MODELS[which]
        fwd = {"ddpm": IO.ddpm_forward, "ncsnpp": IO.ncsnpp_forward, "beatgans": IO.beatgans_forward}[which]
        torch.set_num_threads(os.cpu_count() or 1)
        x = torch.rand(n_images, 3, 64, 64)
        lab = torch.full((n_images,), 0.00999)
        with torch.no_grad():
            fwd(state, spec, x[:1], lab[:1])
            t0 = time.perf_counter()
            fwd(state, spec, x, lab)
            dt = time.perf_counter() - t0
        v, cores, kind, what = n_images / dt, torch.get_num_threads(), "port", "oracle restatement of the network (torch CPU fp32)"
    return {"value": v, "unit": "score-evals/s", "cores": cores, "kind": kind,
            "sample": f"one forward of {n_images} images in {dt:.1f} s, {what}"}


_SDE = {"ncsnpp": (0.009, 50.0), "ddpm": (0.009, 50.0), "beatgans": (0.01, 50.0)}
_NAME = {"ncsnpp": "C4", "beatgans": "C5", "ddpm": "style_gan_base (DDPM)"}


def image_config(which, dev, peaks, K=4096, cpu_images=8, group=None, verify=False):
    """Whole path on ONE anchor (K x 12288 score matrix); with `group` (N > 1) the K rows are sharded over the ranks."""
    from iddiff_b200 import engine as E, sde_lib, synthetic as SY
    eng, n_params, gflop = SY.build_full_engine(which, dev)
    smin, smax = _SDE[which]
    sde = sde_lib.VESDE(sigma_min=smin, sigma_max=smax, N=1000)
    mc, std = sde.coefficients(1e-5)
    label = float((torch.ones(1) * 1e-5 * 999)[0])
    x0 = SY.smooth_images(1, 3, 64, 5).reshape(1, -1).to(dev)
    D = x0.shape[1]
    if group is None:
        run = lambda seed, k: eng.estimate(x0, k, mc, std, label, seed=seed)
    else:
        run = lambda seed, k: eng.estimate_row_sharded(x0, k, mc, std, label, seed=seed, group=group)
    run(0, min(K, 2 * eng.IMAGE_CHUNK if group is None else K))   # warm-up: two full image chunks, so the forward graph is captured here
    E.EVENT_LOG = {}
    r, ms = _timed(lambda: run(1, K))
    stages = {k: sum(a.elapsed_time(b) for a, b, _ in v) for k, v in E.EVENT_LOG.items()}
    E.EVENT_LOG = None
    world = 1
    if group is not None:
        import torch.distributed as dist
        world = dist.get_world_size(group)
        t = torch.tensor([ms] + [stages.get(k, 0.0) for k in sorted(stages)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        ms = float(t[0])
        stages = {k: float(t[1 + i]) for i, k in enumerate(sorted(stages))}
    peak = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
    ach = (K / world) * gflop / stages["score_net"]          # GFLOP / ms = TFLOP/s, per GPU
    sv = r.singular_values[0]
    out = {"workload": f"BASELINE {_NAME[which]}: 64x64x3 synthetic image anchor (D={D}), {which} U-Net at full size "
                       f"({n_params / 1e6:.1f} M parameters, {gflop} GFLOP per image), VE sigma_min={smin}, K={K} score "
                       f"evaluations per anchor, 1 anchor" + (f", K rows sharded over {world} GPUs" if world > 1 else ""),
           "ms_whole_path": ms, "id_estimates_per_s": 1e3 / ms, "score_evals_per_s": K / ms * 1e3,
           "stages_ms": stages, "n": min(K, D), "dims": r.dims.cpu().tolist(),
           "sv_head_tail": [float(sv[0]), float(sv[1]), float(sv[-2])],
           "finite": bool(torch.isfinite(sv).all()),
           "roofline": {"kernel": "conv_fp16x3_kernel (implicit-GEMM convolutions) + the passes between them = the "
                                  "score-network stage", "bound": "tensor", "achieved": ach, "peak": peak,
                        "unit": "TFLOP/s", "frac": ach / peak,
                        "note": "ALGORITHMIC FLOPs of the reference forward (FlopCounterMode, SURVEY 6) per GPU over the "
                                "CUDA-event time of the whole score-network stage; 3 tcgen05.mma per product"}}
    if group is not None and verify:
        # the sharded spectrum against the same anchor on ONE GPU (Philox rows are keyed globally, so only the summation
        # order of the Gram differs): every rank runs it, rank 0's difference is reported
        r1 = eng.estimate(x0, K, mc, std, label, seed=1)
        m = min(K, D) - (1 if K <= D else 0)
        a, b = r.singular_values[0, :m].double(), r1.singular_values[0, :m].double()
        out["max_rel_diff_vs_single_gpu"] = float(((a - b).abs() / b).max())
        out["dims_equal_single_gpu"] = r.dims.cpu().tolist() == r1.dims.cpu().tolist()
    if group is None and cpu_images:
        out["cpu_baseline"] = _cpu_forward_baseline(which, eng.h, cpu_images)
    del eng
    torch.cuda.empty_cache()
    return out
