#!/usr/bin/env python
"""Times one tridiagonalisation variant of idb_spectrum_large on a random Gram matrix:  n variant [repeats]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iddiff_b200 import _lib
n, variant = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
g = torch.Generator(device="cuda").manual_seed(0)
B = torch.randn(n, n + 64, generator=g, device="cuda", dtype=torch.float64)
B *= torch.logspace(0, -3, n, dtype=torch.float64, device="cuda")[:, None]
G = B @ B.T
del B
lib = _lib.load()
ws = torch.empty(lib.idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
ms = []
for _ in range(reps):
    A = G.clone()
    sv = torch.empty(n, dtype=torch.float32, device="cuda")
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    _lib.call("idb_spectrum_large_variant", A.data_ptr(), n, n, sv.data_ptr(), ws.data_ptr(), variant, st)
    b.record()
    torch.cuda.synchronize()
    ms.append(a.elapsed_time(b))
print(json.dumps({"n": n, "variant": variant, "ms": ms, "sv0": float(sv[0]), "sv_last": float(sv[-1])}))
