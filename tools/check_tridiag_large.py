#!/usr/bin/env python
"""n = 12288 (K = 16768 rows of a 64x64x3 image): the two-class fused tridiagonalisation against the unfused two-pass
kernel on the same fp64 Gram matrix -- agreement of the singular values and the time of each."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iddiff_b200 import _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12288
variants = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [2, -1]
g = torch.Generator(device="cuda").manual_seed(0)
B = torch.randn(n, n + 64, generator=g, device="cuda", dtype=torch.float64)
B *= torch.logspace(0, -3, n, dtype=torch.float64, device="cuda")[:, None]
G = B @ B.T
del B
lib = _lib.load()
ws = torch.empty(lib.idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
out, ms = {}, {}
for v in variants:
    A = G.clone()
    sv = torch.empty(n, dtype=torch.float32, device="cuda")
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    _lib.call("idb_spectrum_large_variant", A.data_ptr(), n, n, sv.data_ptr(), ws.data_ptr(), v, st)
    b.record()
    torch.cuda.synchronize()
    out[v], ms[v] = sv.double(), a.elapsed_time(b)
ref = out[variants[-1]]
res = {"n": n, "ms": ms, "GBps_16m2": {v: 16.0 / 3.0 * n ** 3 / t / 1e6 for v, t in ms.items()}}
for v in variants[:-1]:
    res[f"max_rel_diff_{v}_vs_{variants[-1]}"] = float(((out[v] - ref).abs() / ref.clamp_min(1e-7 * float(ref[0]))).max())
print(json.dumps(res))
