#!/usr/bin/env python
"""Multi-GPU check + timing of the K-sharded path (one huge anchor over N GPUs):
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/check_kshard.py [--model beatgans] [--K 4096]
Rank 0 prints one JSON line: time of the sharded run, of the single-GPU run, and their maximum spectrum difference
(Philox noise is keyed by the global row, so the two must agree to rounding)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="beatgans")
    ap.add_argument("--K", default="4096", help="comma-separated list")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    world, rank = dist.get_world_size(), dist.get_rank()
    from iddiff_b200 import sde_lib, synthetic as FX
    eng, _, gflop = FX.build_full_engine(args.model, dev)
    sde = sde_lib.VESDE(sigma_min=0.009, sigma_max=50.0, N=1000)
    mc, std = sde.coefficients(1e-5)
    x0 = FX.smooth_images(1, 3, 64, 5).reshape(1, -1).to(dev)

    def timed(fn):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = fn(); e1.record()
        dist.barrier(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return r, float(t[0])

    eng.estimate_row_sharded(x0, 256, mc, std, 0.00999, seed=0)                      # warm-up (K < D path)
    from iddiff_b200 import engine as E
    from iddiff_b200.engine import spectrum_from_rows
    for K in [int(k) for k in str(args.K).split(",")]:
        E.EVENT_LOG = {}
        r_sh, ms_sh = timed(lambda: eng.estimate_row_sharded(x0, K, mc, std, 0.00999, seed=1))
        stages = {k: sum(a.elapsed_time(b) for a, b, _ in v) for k, v in E.EVENT_LOG.items()}
        E.EVENT_LOG = None
        r_ag, ms_ag = timed(lambda: eng.estimate_row_sharded(x0, K, mc, std, 0.00999, seed=1, gather_rows=True))
        out = {"model": args.model, "K": K, "n_gpus": world, "ms_partial_gram": ms_sh, "ms_gather_rows": ms_ag,
               "stages_ms_rank0": stages, "score_evals_per_s": K / ms_sh * 1e3}
        if rank == 0:
            S = eng.score_rows(x0, K, 0, mc, std, 0.00999, seed=1)
            r_1 = spectrum_from_rows(S, K, S.shape[2], -1.0 / std)
            m = min(K, S.shape[2]) - (1 if K <= S.shape[2] else 0)
            b1 = r_1.singular_values[0, :m].double()
            out["max_rel_diff_vs_single_gpu"] = float(((r_sh.singular_values[0, :m].double() - b1).abs() / b1).max())
            out["max_rel_diff_gather_rows_vs_single_gpu"] = float(((r_ag.singular_values[0, :m].double() - b1).abs() / b1).max())
            out["dims_equal"] = r_sh.dims.cpu().tolist() == r_1.dims.cpu().tolist()
            del S
            print(json.dumps(out), flush=True)
        dist.barrier()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
