"""Per-layer comparison of the CUDA FCN chain against an fp64 torch emulation (debug aid, GPU)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import fixtures as FX
from iddiff_b200 import sde_lib, layout
from iddiff_b200.engine import FcnEngine
from iddiff_b200._lib import call, ptr, stream_ptr

def elu(x): return torch.where(x > 0, x, torch.expm1(x))

name = sys.argv[1] if len(sys.argv) > 1 else "fcn_small_ve"
g = FX.golden(name); D, L, B, K = int(g["D"]), int(g["L"]), int(g["B"]), int(g["K"])
ws, bs = FX.fcn_state_to_lists(FX.golden_fcn_state(g), L)
dev = torch.device("cuda:0")
eng = FcnEngine([w.to(dev) for w in ws], [b.to(dev) for b in bs])
sde = sde_lib.VESDE(1e-2, 4.0, 1000); t = 1e-5
mc, std = sde.coefficients(t); label = float((torch.ones(1) * t * 999)[0])
x0 = torch.from_numpy(g["x0"]).to(dev); n = x0.shape[0]
if "z" in g.files:
    z = torch.from_numpy(g["z"]).reshape(n, -1, D)[:, :K].contiguous().to(dev)
else:
    from oracle import id_oracle as O
    nb, _, _ = O.k_schedule([D], B)
    z = FX.seeded_noise(int(g["seed"]) + 2, n, nb, B, [D]).reshape(n, nb * B, D)[:, :K].contiguous().to(dev)
A = n; Kp = eng.rows_per_anchor(K); M = A * Kp; st = stream_ptr()
e_hi = torch.empty(M * eng.Kc[0] * 2, dtype=torch.uint8, device=dev); e_lo = torch.empty_like(e_hi)
call("idb_perturb", ptr(x0), ptr(z), A, K, D, Kp, eng.Kc[0], mc, std, 0, 0, None, eng.fmt, eng.a_scale, ptr(e_hi), ptr(e_lo), st)
P = eng.anchor_chain(x0, mc, label)
# fp64 emulation
mean = (torch.tensor(mc, device=dev) * x0)
x = mean[:, None, :] + torch.tensor(std, device=dev) * z
d = torch.zeros(A, Kp, eng.Kc[0], dtype=torch.float64, device=dev); d[:, :K, :D] = (x - mean[:, None, :]).double()
DT = torch.float16 if eng.fmt == 1 else torch.bfloat16
def unslab(hi, lo, R, C): return (layout.from_slab(hi, R, C).view(DT).double() + layout.from_slab(lo, R, C).view(DT).double()) / eng.a_scale
got = unslab(e_hi, e_lo, M, eng.Kc[0]).reshape(A, Kp, -1)
print("input dev: max err", (got - d).abs().max().item(), "scale", d.abs().max().item())
a = torch.cat([mean.double(), torch.full((A, 1), label, dtype=torch.float64, device=dev)], 1)
for l in range(eng.n_layers):
    w64 = eng.w[l].double(); b64 = eng.b[l].double()
    a0 = a @ w64.T + b64
    perr = (P[l][:, :a0.shape[1]].double() - a0).abs().max().item()
    wd = w64[:, :D] if l == 0 else w64
    nd = d[:, :, :wd.shape[1]] @ wd.T
    if l < eng.n_layers - 1:
        o_hi = torch.empty(M * eng.Np[l] * 2, dtype=torch.uint8, device=dev); o_lo = torch.empty_like(o_hi)
        call("idb_fcn_hidden_layer", ptr(e_hi), ptr(e_lo), ptr(eng.w_hi[l]), ptr(eng.w_lo[l]), eng.fmt, eng.w_inv_scale[l] / eng.a_scale, eng.a_scale, M, eng.Np[l], eng.Kc[l],
             ptr(P[l]), Kp, ptr(o_hi), ptr(o_lo), st)
        torch.cuda.synchronize()
        got = unslab(o_hi, o_lo, M, eng.Np[l]).reshape(A, Kp, -1)
        ref = elu(a0[:, None, :] + nd) - elu(a0[:, None, :])
        err = (got[:, :, :ref.shape[2]] - ref)
        print(f"layer {l}: anchor err {perr:.3e}  dev err max {err.abs().max().item():.3e} rms {err.pow(2).mean().sqrt().item():.3e} "
              f"scale max {ref.abs().max().item():.3e} rms {ref.pow(2).mean().sqrt().item():.3e}  pad {got[:, :, ref.shape[2]:].abs().max().item() if got.shape[2] > ref.shape[2] else 0:.1e}")
        worst = err.abs().flatten().argmax().item()
        idx = np.unravel_index(worst, err.shape)
        print("   worst at", idx, "a0", a0[idx[0], idx[2]].item(), "nd", nd[idx].item(), "got", got[idx].item(), "ref", ref[idx].item())
        d = torch.zeros(A, Kp, eng.Np[l], dtype=torch.float64, device=dev); d[:, :, :ref.shape[2]] = ref
        a = elu(a0); e_hi, e_lo = o_hi, o_lo
    else:
        out = torch.empty(A, Kp, eng.Np[l], device=dev)
        call("idb_fcn_output_layer", ptr(e_hi), ptr(e_lo), ptr(eng.w_hi[l]), ptr(eng.w_lo[l]), eng.fmt, eng.w_inv_scale[l] / eng.a_scale, M, eng.Np[l], eng.Kc[l], ptr(out), st)
        torch.cuda.synchronize()
        err = out[:, :, :D].double() - nd
        print(f"out layer: anchor err {perr:.3e} dev err max {err.abs().max().item():.3e} rms {err.pow(2).mean().sqrt().item():.3e} scale rms {nd.pow(2).mean().sqrt().item():.3e}")
        c = (nd[:, :K] - nd[:, :K].mean(1, keepdim=True)) / std
        st64 = torch.linalg.svdvals(c)
        c2 = (out[:, :K, :D].double() - out[:, :K, :D].double().mean(1, keepdim=True)) / std
        s2 = torch.linalg.svdvals(c2)
        print("sv(gpu d_out, fp64 svd) vs truth max rel", ((s2 - st64).abs() / st64).max().item())
        res = eng.estimate(x0, K, mc, std, label, z=z)
        print("engine sv vs truth max rel", ((res.singular_values.double() - st64).abs() / st64).max().item(), "sweeps", res.sweeps.tolist())
