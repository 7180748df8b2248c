#!/usr/bin/env python
"""Times the image-sized spectrum stage on one B200: K x D fp32 score matrix -> singular values (BASELINE configs 4-5)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iddiff_b200 import engine as E

for K, D in [(4096, 12288), (2048, 12288), (16768, 12288)][: int(sys.argv[1]) if len(sys.argv) > 1 else 2]:
    X = torch.randn(1, K, D, device="cuda") + 1.0
    E.EVENT_LOG = {}
    r = E.spectrum_from_rows(X, K, D, 1.0)
    torch.cuda.synchronize()
    out = {k: sum(a.elapsed_time(b) for a, b, _ in v) for k, v in E.EVENT_LOG.items()}
    E.EVENT_LOG = None
    n = min(K, D)
    ref = torch.linalg.svdvals((X[0] - X[0].mean(0, keepdim=True)).double()) if K * D <= 4096 * 12288 else None
    err = float(((r.singular_values[0, :n - 1].double() - ref[:n - 1]).abs() / ref[:n - 1]).max()) if ref is not None else None
    print(json.dumps({"K": K, "D": D, "n": n, "ms": out, "gram_tflops_fp64": (n * n * max(K, D)) / out["center_gram"] / 1e9,
                      "tridiag_GBps_algorithmic_16m2": 16.0 / 3.0 * n ** 3 / out["spectrum"] / 1e6, "max_rel_err_vs_fp64_svd": err}))
