#!/usr/bin/env python
"""Accuracy + speed attribution of the U-Net fusions (development tool): for each full-size model, the forward error
against the fp64 oracle output stored in tests/golden/unet_*_full.npz and the forward time at batch 512, with the fused
GroupNorm statistics / folded FIR switched on and off.  Prints one JSON line per (model, setting)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from iddiff_b200 import synthetic as SY, unet_models, unet_ops as U  # noqa: E402
from iddiff_b200.unet_engine import UNetEngine  # noqa: E402


def conv_accuracy():
    """One 3x3 convolution, Cin = 64 ... 1024: max error against fp64 of this kernel and of torch's fp32 convolution
    (TF32 off), in units of the output's largest magnitude -- how much of the network-level gap is the convolution."""
    import torch.nn.functional as F
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    for cin, hw in ((64, 32), (256, 16), (1024, 8)):
        g = torch.Generator().manual_seed(cin)
        x = torch.randn(8, cin, hw, hw, generator=g)
        x = F.silu(x)                                   # activations as the network feeds them (non-zero mean)
        w = (torch.rand(128, cin, 3, 3, generator=g) * 2 - 1) * (3.0 / (9 * cin)) ** 0.5
        ref64 = F.conv2d(x.double(), w.double(), padding=1)
        y32 = F.conv2d(x.cuda(), w.cuda(), padding=1).cpu()
        pc = U.pack_conv(w.cuda(), None)
        a = U.prep(x.permute(0, 2, 3, 1).contiguous().cuda())
        y = U.conv(a, pc).permute(0, 3, 1, 2).cpu()
        sc = float(ref64.abs().max())
        print(json.dumps({"conv_cin": cin, "terms": 9 * cin, "engine_err": float((y.double() - ref64).abs().max()) / sc,
                          "engine_rms": float((y.double() - ref64).pow(2).mean().sqrt()) / sc,
                          "torch_fp32_err": float((y32.double() - ref64).abs().max()) / sc,
                          "torch_fp32_rms": float((y32.double() - ref64).pow(2).mean().sqrt()) / sc,
                          "engine_mean_signed": float((y.double() - ref64).mean()) / sc}), flush=True)


def main():
    conv_accuracy()
    which_list = sys.argv[1].split(",") if len(sys.argv) > 1 else ["beatgans", "ddpm", "ncsnpp"]
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    for which in which_list:
        g = np.load(os.path.join(ROOT, "tests", "golden", f"unet_{which}_full.npz"))
        name, spec, gflop = SY.FULL_MODELS[which]
        holder = unet_models.get_model(name)(SY.unet_config(spec, name))
        holder.load_state_dict(SY.make_unet_state(holder, int(g["seed"])), strict=True)
        eng = UNetEngine.from_module(holder, "cuda:0")
        x = torch.from_numpy(g["x"]).cuda()
        xb = torch.rand(batch, 3, 64, 64, device="cuda")
        for stats, fir, silu in ((True, True, 1),):
            U.FUSED_GN_STATS, U.FUSED_FIR, U.SILU_ACT = stats, fir, silu
            got = eng.forward(x, float(g["label"])).permute(0, 3, 1, 2).cpu().numpy()
            for _ in range(2):
                eng.forward(xb, float(g["label"]))
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                eng.forward(xb, float(g["label"]))
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 3
            print(json.dumps({"model": which, "fused_gn_stats": stats, "fused_fir": fir, "silu_act": silu,
                              "err_vs_fp64": float(np.abs(got - g["y64"]).max()),
                              "err_vs_ref": float(np.abs(got - g["y"]).max()),
                              "ref_vs_fp64": float(np.abs(g["y"] - g["y64"]).max()), "ms_per_forward": ms,
                              "score_evals_per_s": batch / ms * 1e3,
                              "tflops_algorithmic": batch * gflop / ms}), flush=True)
        del eng
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
