"""Device pipeline of the `manifold_dimension` hot path: thin Python over the C ABI.

    anchors x0 --perturb--> deviations (bf16 hi/lo slabs) --tcgen05 GEMM chain--> d_out (fp32)
            \\--fp32 anchor chain--> per-layer anchor pre-activations --^
    d_out --centre + Gram (fp64)--> G --Cholesky + one-sided Jacobi--> singular values --gap rule--> ID

Everything is batched over a chunk of anchors; nothing returns to the host until the spectra are
done (the reference moves every score batch to the CPU, dim_reduction.py:183).  torch is used for
device memory and streams only.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr


# Set to a dict to collect (start event, end event, algorithmic work) per stage -- bench.py reads it after a sync.
EVENT_LOG = None


class _stage:
    def __init__(self, name: str, work: float = 0.0):
        self.name, self.work = name, work

    def __enter__(self):
        self.start = None
        if EVENT_LOG is not None:
            self.start = torch.cuda.Event(enable_timing=True)
            self.start.record()

    def __exit__(self, *exc):
        if self.start is not None:
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            EVENT_LOG.setdefault(self.name, []).append((self.start, end, self.work))
        return False


def _ceil(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def _require_cuda(t: torch.Tensor, name: str):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
        raise _lib.IdbError(f"{name} must be a contiguous float32 CUDA tensor")


@dataclass
class SpectrumResult:
    singular_values: torch.Tensor  # [A, n] fp32, descending (device)
    dims: torch.Tensor             # [A] int32 (device); -1 where the reference raises ZeroDivisionError
    sweeps: torch.Tensor           # [A] int32 Jacobi sweeps
    scores: Optional[torch.Tensor] = None  # [A, K, D] fp32 raw scores when requested
    finite: Optional[torch.Tensor] = None  # [A] bool: the anchor's Gram diagonal (sums of squared scores) is finite


LARGE_N = 1024  # above this, one Gram matrix is solved by the whole GPU (idb_spectrum_large)
SPECTRUM_METHOD = _lib.SPECTRUM_TRIDIAG  # default solver; _lib.SPECTRUM_JACOBI selects the one-sided Jacobi


def _solve_large(G: torch.Tensor, n: int, sv_row: torch.Tensor, ws: torch.Tensor, st):
    with _stage("spectrum", n * n * 8.0 + n * 4.0):
        call("idb_spectrum_large", ptr(G), n, n, ptr(sv_row), ptr(ws), st)


def _read_out(sv: torch.Tensor, A: int, n: int, st) -> torch.Tensor:
    dims = torch.empty(A, dtype=torch.int32, device=sv.device)
    if n >= 3:
        call("idb_id_from_spectrum", ptr(sv), A, n, n, ptr(dims), st)
    else:
        dims.fill_(-2)
    return dims


def _gram_finite(G: torch.Tensor) -> torch.Tensor:
    """Guard, not arithmetic of the path: a NaN / inf score (an operand that left the fp16-pair range) makes the sum of
    squares on the Gram diagonal non-finite, while the bisection would still return finite garbage for it."""
    return torch.isfinite(torch.diagonal(G, dim1=-2, dim2=-1)).all(dim=-1)


def spectrum_from_gram(G: torch.Tensor, n: int, method: Optional[int] = None) -> SpectrumResult:
    """Singular values + IDs from centred Gram matrices G [A, n, n] fp64 (overwritten when n > LARGE_N)."""
    A = G.shape[0]
    dev, st = G.device, stream_ptr()
    finite = _gram_finite(G)
    sv = torch.empty(A, n, dtype=torch.float32, device=dev)
    sweeps = torch.zeros(A, dtype=torch.int32, device=dev)
    if n > LARGE_N:
        ws = torch.empty(_lib.load().idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device=dev)
        for a in range(A):
            _solve_large(G[a], n, sv[a], ws, st)
    else:
        ws = torch.empty(_lib.load().idb_spectrum_workspace_bytes(A, n), dtype=torch.uint8, device=dev)
        with _stage("spectrum", A * (n * n * 8.0 + n * 4.0)):
            call("idb_spectrum", ptr(G), A, n, n, SPECTRUM_METHOD if method is None else method, ptr(sv),
                 ptr(sweeps), ptr(ws), st)
    return SpectrumResult(sv, _read_out(sv, A, n, st), sweeps, finite=finite)


def spectrum_from_rows(X: torch.Tensor, K: int, D: int, scale: float, rows_per_anchor: Optional[int] = None,
                       method: Optional[int] = None) -> SpectrumResult:
    """dim_reduction.py:190-199 + plot_utils.py:224-229 on device.
    X: [A, rows_per_anchor, ld] fp32 (rows >= K and columns >= D are ignored)."""
    _require_cuda(X, "X")
    A, R, ld = X.shape
    if rows_per_anchor is not None and rows_per_anchor != R:
        raise _lib.IdbError("rows_per_anchor mismatch")
    dev = X.device
    st = stream_ptr()
    n = min(K, D)          # the smaller side of the centred score matrix carries all its singular values
    colmean = torch.empty(A, D, dtype=torch.float64, device=dev)
    large = n > LARGE_N
    G = torch.empty(1 if large else A, n, n, dtype=torch.float64, device=dev)
    if large:
        # image-sized matrices: one anchor at a time, the whole GPU on each (cooperative tridiagonalisation)
        sv = torch.empty(A, n, dtype=torch.float32, device=dev)
        finite = torch.empty(A, dtype=torch.bool, device=dev)
        ws = torch.empty(_lib.load().idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device=dev)
        for a in range(A):
            with _stage("center_gram", K * D * 4.0 + n * n * 8.0):
                if K < D:
                    call("idb_row_gram", ptr(X[a]), ld, K, D, float(scale), ptr(colmean[a]), ptr(G), st)
                else:
                    call("idb_center_gram", ptr(X[a]), R * ld, ld, 1, K, D, float(scale), ptr(colmean[a]), ptr(G), st)
            finite[a] = _gram_finite(G[0])
            _solve_large(G[0], n, sv[a], ws, st)
        return SpectrumResult(sv, _read_out(sv, A, n, st), torch.zeros(A, dtype=torch.int32, device=dev), finite=finite)
    with _stage("center_gram", A * (K * D * 4.0 + n * n * 8.0)):
        if K < D:
            for a in range(A):
                call("idb_row_gram", ptr(X[a]), ld, K, D, float(scale), ptr(colmean[a]), ptr(G[a]), st)
        else:
            call("idb_center_gram", ptr(X), R * ld, ld, A, K, D, float(scale), ptr(colmean), ptr(G), st)
    return spectrum_from_gram(G, n, method)


class FcnEngine:
    """models/fcn.py FCN (Linear(D+1,H), L x Linear(H,H), Linear(H,D), ELU between) evaluated for all K
    perturbations of a chunk of anchors.  Weights stay owned by the caller's module; this object keeps
    packed bf16 hi/lo slab copies (tensor-core operands) and fp32 references (anchor chain)."""

    # Deviations are sigma-sized; in the fp16-pair format they are multiplied by 2^6 so that the lo halves of
    # typical values stay normal fp16 numbers (absolute resolution 2^-24 / 2^6 ~ 9e-10, overflow at |e| > 1023).
    A_SCALE_FP16 = 64.0

    def __init__(self, weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor], fmt: int = _lib.FMT_FP16):
        if len(weights) < 2 or len(weights) != len(biases):
            raise _lib.IdbError("FCN needs >= 2 linear layers with biases")
        self.w = [w.detach().contiguous() for w in weights]
        self.b = [b.detach().contiguous() for b in biases]
        for t in self.w + self.b:
            _require_cuda(t, "FCN parameter")
        self.dev = self.w[0].device
        self._side = None
        self.D = self.w[-1].shape[0]
        if self.w[0].shape[1] != self.D + 1:
            raise _lib.IdbError("first layer must take state_size + 1 inputs (time column, models/fcn.py:35-38)")
        self.n_layers = len(self.w)
        # padded dims: layer l maps Kc[l] -> Np[l]
        self.in_real = [self.D] + [w.shape[0] for w in self.w[:-1]]
        self.out_real = [w.shape[0] for w in self.w]
        self.Np = [_ceil(n, 256) for n in self.out_real]
        self.Kc = [_ceil(self.D, 64)] + self.Np[:-1]
        st = stream_ptr()
        self.fmt = int(fmt)
        self.a_scale = self.A_SCALE_FP16 if self.fmt == _lib.FMT_FP16 else 1.0
        self.b_pad = []
        for l, b in enumerate(self.b):
            bp = torch.zeros(self.Np[l], dtype=torch.float32, device=self.dev)
            bp[:b.numel()] = b
            self.b_pad.append(bp)
        self.w_hi, self.w_lo, self.w_inv_scale = [], [], []
        for l, w in enumerate(self.w):
            scale = 1.0
            if self.fmt == _lib.FMT_FP16:
                # power of two that puts max|w| in [2^13, 2^14): hi and lo are normal fp16 numbers, products exact
                wmax = float(w.abs().max())
                scale = 2.0 ** (13 - math.frexp(wmax)[1] + 1) if wmax > 0 else 1.0
            self.w_inv_scale.append(1.0 / scale)
            nbytes = self.Np[l] * self.Kc[l] * 2
            hi = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
            lo = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
            call("idb_pack_weight", ptr(w), w.shape[0], self.in_real[l], w.shape[1], self.Np[l], self.Kc[l],
                 self.fmt, scale, ptr(hi), ptr(lo), st)
            self.w_hi.append(hi)
            self.w_lo.append(lo)

    @classmethod
    def from_module(cls, fcn_module: torch.nn.Module, device="cuda"):
        """Ingests a reference `models.fcn.FCN` (or any module whose `.mlp` is the same Sequential)."""
        lin = [m for m in fcn_module.mlp if isinstance(m, torch.nn.Linear)]
        return cls([m.weight.detach().to(device, torch.float32) for m in lin],
                   [m.bias.detach().to(device, torch.float32) for m in lin])

    GEMM_CHUNK = 256   # anchors per pass of the tensor-core chain (activation ping-pong buffers: 2 x 3.2 GB at K = 1501)

    def rows_per_anchor(self, K: int) -> int:
        """Rows of the operand slabs per anchor: the K perturbations packed DENSELY (a 128-row GEMM tile then spans at
        most two anchors and its epilogue looks the anchor up per row; no zero rows are multiplied: K = 1501 used to pad
        to 1536).  Below 128 rows an anchor is padded to one tile."""
        return max(int(K), 128)

    def bytes_per_anchor(self, K: int) -> int:
        """Per anchor of one estimate() call: fp32 deviations of the output layer, fp64 Gram + solver workspace; the
        activation buffers of the GEMM chain are bounded by GEMM_CHUNK, not by the call."""
        Kp = self.rows_per_anchor(K)
        return Kp * (self.Kc[0] * 4 + self.Np[-1] * 4) + self.D * self.D * 8 * 3

    def chain_bytes(self, K: int) -> int:
        return self.GEMM_CHUNK * self.rows_per_anchor(K) * max(self.Np[:-1]) * 2 * 2 * 2

    def anchor_chain(self, x0: torch.Tensor, mean_coeff: float, label: float) -> List[torch.Tensor]:
        """Pre-activations of every layer at the un-perturbed anchors, fp32: list of [>= A, Np[l]].
        Tensor-core version: the anchors are one more (small) batch through the same split-precision GEMM."""
        A = x0.shape[0]
        Mp = _ceil(A, 128)
        st = stream_ptr()
        dev = self.dev
        a_hi = torch.empty(Mp * self.Kc[0] * 2, dtype=torch.uint8, device=dev)
        a_lo = torch.empty_like(a_hi)
        # anchor activations are O(1): no scaling needed to keep fp16 hi/lo normal (scale 1)
        call("idb_pack_weight", ptr(x0), A, self.D, self.D, Mp, self.Kc[0], self.fmt, float(mean_coeff), ptr(a_hi),
             ptr(a_lo), st)
        P = []
        for l in range(self.n_layers):
            if l == 0:  # time column of cat([x, t]) folded into the bias (t is one constant, dim_reduction.py:173-174)
                bias = torch.zeros(self.Np[0], dtype=torch.float32, device=dev)
                bias[:self.out_real[0]] = self.b[0] + self.w[0][:, self.D] * label
            else:
                bias = self.b_pad[l]
            out = torch.empty(Mp, self.Np[l], dtype=torch.float32, device=dev)
            last = l == self.n_layers - 1
            o_hi = None if last else torch.empty(Mp * self.Np[l] * 2, dtype=torch.uint8, device=dev)
            o_lo = None if last else torch.empty_like(o_hi)
            call("idb_fcn_anchor_layer", ptr(a_hi), ptr(a_lo), ptr(self.w_hi[l]), ptr(self.w_lo[l]), self.fmt,
                 self.w_inv_scale[l], 1.0, Mp, self.Np[l], self.Kc[l], ptr(bias), ptr(out), ptr(o_hi), ptr(o_lo), st)
            P.append(out)
            a_hi, a_lo = o_hi, o_lo
        return P

    def anchor_chain_fp32(self, x0: torch.Tensor, mean_coeff: float, label: float) -> List[torch.Tensor]:
        """Plain-fp32 SIMT version of anchor_chain (cross-check)."""
        A = x0.shape[0]
        st = stream_ptr()
        P = []
        for l in range(self.n_layers):
            out = torch.zeros(A, self.Np[l], dtype=torch.float32, device=self.dev)
            if l == 0:
                call("idb_anchor_linear", ptr(x0), self.D, 0, float(mean_coeff), ptr(self.w[0]), self.D + 1,
                     ptr(self.b[0]), self.w[0].data_ptr() + 4 * self.D, self.D + 1, float(label), ptr(out),
                     self.Np[0], A, self.out_real[0], self.D, st)
            else:
                call("idb_anchor_linear", ptr(P[-1]), self.Np[l - 1], 1, 1.0, ptr(self.w[l]), self.w[l].shape[1],
                     ptr(self.b[l]), None, 0, 0.0, ptr(out), self.Np[l], A, self.out_real[l], self.in_real[l], st)
            P.append(out)
        return P

    def deviations(self, x0: torch.Tensor, K: int, mean_coeff: float, std: float, label: float,
                   z: Optional[torch.Tensor] = None, seed: int = 0, anchor_id0: int = 0, P=None, d_out=None):
        """Runs perturb + anchor chain + GEMM chain.  Returns (d_out [A, Kp, Np_last] fp32, a_out [A, Np_last]).
        P: anchor pre-activations of these anchors when the caller already ran the anchor chain (one chain for all
        the chunks of a call); d_out: where to write (a slice of the caller's buffer)."""
        _require_cuda(x0, "x0")
        A, D = x0.shape
        if D != self.D:
            raise _lib.IdbError(f"anchor dimension {D} != model state_size {self.D}")
        if z is not None:
            _require_cuda(z, "z")
            if tuple(z.shape) != (A, K, D):
                raise _lib.IdbError("z must be [A, K, D]")
        Kp = self.rows_per_anchor(K)
        M = _ceil(A * Kp, 128)          # slab rows: anchors back to back, zero rows up to the tile boundary
        st = stream_ptr()
        dev = self.dev
        e_hi = torch.empty(M * self.Kc[0] * 2, dtype=torch.uint8, device=dev)
        e_lo = torch.empty(M * self.Kc[0] * 2, dtype=torch.uint8, device=dev)
        with _stage("perturb", 2.0 * M * self.Kc[0] * 2):
            call("idb_perturb", ptr(x0), ptr(z), A, K, D, Kp, self.Kc[0], float(mean_coeff), float(std), int(seed),
                 int(anchor_id0), None, self.fmt, self.a_scale, ptr(e_hi), ptr(e_lo), st)
        if P is None:
            with _stage("anchor_chain", sum(2.0 * A * i * o for i, o in zip(self.in_real, self.out_real))):
                P = self.anchor_chain(x0, mean_coeff, label)
        for l in range(self.n_layers - 1):
            o_hi = torch.empty(M * self.Np[l] * 2, dtype=torch.uint8, device=dev)
            o_lo = torch.empty(M * self.Np[l] * 2, dtype=torch.uint8, device=dev)
            with _stage("hidden_first" if l == 0 else "hidden", 2.0 * A * K * self.in_real[l] * self.out_real[l]):
                call("idb_fcn_hidden_layer", ptr(e_hi), ptr(e_lo), ptr(self.w_hi[l]), ptr(self.w_lo[l]), self.fmt,
                     self.w_inv_scale[l] / self.a_scale, self.a_scale, M, self.Np[l], self.Kc[l], ptr(P[l]), Kp, A,
                     ptr(o_hi), ptr(o_lo), st)
            e_hi, e_lo = o_hi, o_lo
        l = self.n_layers - 1
        if d_out is None:
            d_out = torch.empty(M, self.Np[l], dtype=torch.float32, device=dev)
        elif tuple(d_out.shape) != (M, self.Np[l]) or not d_out.is_contiguous():
            raise _lib.IdbError("deviations: d_out must be a contiguous [ceil128(A * rows_per_anchor), Np] buffer")
        with _stage("output", 2.0 * A * K * self.in_real[l] * self.out_real[l]):
            call("idb_fcn_output_layer", ptr(e_hi), ptr(e_lo), ptr(self.w_hi[l]), ptr(self.w_lo[l]), self.fmt,
                 self.w_inv_scale[l] / self.a_scale, M, self.Np[l], self.Kc[l], ptr(d_out), st)
        return d_out[:A * Kp].view(A, Kp, self.Np[l]), P[-1]

    def join(self):
        """Makes the current stream wait for every spectrum stage deferred with estimate(..., defer=True)."""
        if self._side is not None:
            torch.cuda.current_stream(self.dev).wait_stream(self._side)

    def estimate(self, x0: torch.Tensor, K: int, mean_coeff: float, std: float, label: float,
                 z: Optional[torch.Tensor] = None, seed: int = 0, anchor_id0: int = 0,
                 want_scores: bool = False, defer: bool = False, chunk: Optional[int] = None) -> SpectrumResult:
        """All anchors of x0 in one call: ONE anchor chain, the tensor-core chain in passes of `chunk` anchors
        (default GEMM_CHUNK), then ONE centre/Gram + eigen-solve + read-out over all of them (the one-CTA-per-anchor
        stages fill the SMs in ceil(A / 148) waves instead of one wave per chunk).
        defer=True: the spectrum stage is enqueued on a second CUDA stream; the caller must call join() before
        reading the results on its own stream."""
        A = x0.shape[0]
        ck = min(A, chunk or self.GEMM_CHUNK)
        if A <= ck:
            d_out, a_out = self.deviations(x0, K, mean_coeff, std, label, z, seed, anchor_id0)
        else:
            _require_cuda(x0, "x0")
            with _stage("anchor_chain", sum(2.0 * A * i * o for i, o in zip(self.in_real, self.out_real))):
                P = self.anchor_chain(x0, mean_coeff, label)
            Kp = self.rows_per_anchor(K)
            # one flat buffer for all passes: a pass writes whole 128-row tiles, so its zero tail rows spill into the next
            # pass's region and are overwritten by it (same stream, in order); 128 rows of slack close the last one
            flat = torch.empty(_ceil(A * Kp, 128) + 128, self.Np[-1], dtype=torch.float32, device=self.dev)
            for s in range(0, A, ck):
                e = min(A, s + ck)
                self.deviations(x0[s:e], K, mean_coeff, std, label, None if z is None else z[s:e], seed,
                                anchor_id0 + s, P=[q[s:] for q in P], d_out=flat[s * Kp:s * Kp + _ceil((e - s) * Kp, 128)])
            d_out = flat[:A * Kp].view(A, Kp, self.Np[-1])
            a_out = P[-1]
        # centring removes the anchor part exactly: spectrum of (-(a + d)/std - mean) = spectrum of (d - mean(d))/std
        if defer and not want_scores:
            main = torch.cuda.current_stream(self.dev)
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.dev)
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side):
                res = spectrum_from_rows(d_out, K, self.D, 1.0 / std)
            d_out.record_stream(self._side)
            for t in (res.singular_values, res.dims, res.sweeps, res.finite):
                t.record_stream(main)
            return res
        res = spectrum_from_rows(d_out, K, self.D, 1.0 / std)
        if want_scores:
            A, Kp, ld = d_out.shape
            sc = torch.empty(A, K, self.D, dtype=torch.float32, device=self.dev)
            call("idb_fcn_scores", ptr(a_out), a_out.shape[1], ptr(d_out), ld, A, K, Kp, self.D, float(std), ptr(sc),
                 stream_ptr())
            res.scores = sc
        return res


class KSphereEngine:
    """models/ksphere_gt.py analytic score + models/utils.py:169-175 wrapper."""

    def __init__(self, manifold_dim: int, device="cuda"):
        self.p = int(manifold_dim)
        self.dev = torch.device(device)

    def bytes_per_anchor(self, K: int, D: int) -> int:
        return K * D * 8 + D * D * 24

    def estimate(self, x0: torch.Tensor, K: int, mean_coeff: float, std: float, label: float = 0.0,
                 z: Optional[torch.Tensor] = None, seed: int = 0, anchor_id0: int = 0,
                 want_scores: bool = False) -> SpectrumResult:
        _require_cuda(x0, "x0")
        A, D = x0.shape
        if z is not None:
            _require_cuda(z, "z")
        st = stream_ptr()
        x = torch.empty(A, K, D, dtype=torch.float32, device=x0.device)
        call("idb_perturb", ptr(x0), ptr(z), A, K, D, 0, 0, float(mean_coeff), float(std), int(seed), int(anchor_id0),
             ptr(x), 0, 1.0, None, None, st)
        sc = torch.empty_like(x)
        call("idb_ksphere_gt_score", ptr(x), None, float(std), A * K, D, self.p, ptr(sc), st)
        res = spectrum_from_rows(sc, K, D, 1.0)
        if want_scores:
            res.scores = sc
        return res
