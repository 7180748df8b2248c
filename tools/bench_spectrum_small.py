#!/usr/bin/env python
"""Times the per-anchor spectrum stage (idb_spectrum: one-CTA tridiagonalisation + bisection) on Gram matrices of the
FCN bench shape: A anchors, n = 200.  Development tool."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iddiff_b200 import engine as E

A, K, D = (int(v) for v in (sys.argv[1:4] + [128, 1501, 200][len(sys.argv) - 1:]))
X = torch.randn(A, K, D, device="cuda") * torch.logspace(0, -3, D, device="cuda")
for rep in range(3):
    E.EVENT_LOG = {}
    r = E.spectrum_from_rows(X, K, D, 1.0)
    torch.cuda.synchronize()
    out = {k: sum(a.elapsed_time(b) for a, b, _ in v) for k, v in E.EVENT_LOG.items()}
    E.EVENT_LOG = None
ref = torch.linalg.svdvals((X[0] - X[0].mean(0, keepdim=True)).double())
err = float(((r.singular_values[0].double() - ref).abs() / ref).max())
print(json.dumps({"anchors": A, "K": K, "D": D, "ms": out, "max_rel_err_vs_fp64_svd": err}))
