"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's `manifold_dimension`
hot path (GBATZOLIS/ID-diff).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product package
(iddiff_b200/) never does and fails loudly when its CUDA library is missing.

Parity status: PINNED.  Every function below is checked against the reference's own
code, imported from /root/reference in the build container through oracle/ref_shims.py,
by oracle/gen_golden.py; the resulting vectors are committed under tests/golden/ and
re-checked by tests/test_oracle_golden.py (CPU, no reference needed).  The reference
itself ships no tests or golden vectors.

The restatement uses the same numerical back-ends the reference uses on CPU
(ATen/MKL `F.linear`, LAPACK `torch.linalg.svd`, `scipy.special.ive`) so that it
reproduces the reference to the last bit on the same machine; it re-expresses the
control flow and arithmetic, it does not call any reference code.

Each function cites the reference file:line (relative to the reference root) it follows.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# SDE marginals  (sde_lib.py)
# --------------------------------------------------------------------------------------


def ve_marginal(x: torch.Tensor, t: torch.Tensor, sigma_min: float, sigma_max: float):
    """VESDE.marginal_prob, sde_lib.py:342-347: mean = x, std = smin*(smax/smin)**t (dtype of t)."""
    smin = torch.tensor(sigma_min).type_as(t)
    smax = torch.tensor(sigma_max).type_as(t)
    std = smin * (smax / smin) ** t
    return x, std


def vp_marginal(x: torch.Tensor, t: torch.Tensor, beta_min: float, beta_max: float):
    """VPSDE.marginal_prob, sde_lib.py:250-254."""
    log_mean_coeff = -0.25 * t ** 2 * (beta_max - beta_min) - 0.5 * t * beta_min
    mean = torch.exp(log_mean_coeff[(...,) + (None,) * len(x.shape[1:])]) * x
    std = torch.sqrt(1.0 - torch.exp(2.0 * log_mean_coeff))
    return mean, std


def subvp_marginal(x: torch.Tensor, t: torch.Tensor, beta_min: float, beta_max: float):
    """subVPSDE.marginal_prob, sde_lib.py:300-304 (std is NOT square-rooted in the reference)."""
    log_mean_coeff = -0.25 * t ** 2 * (beta_max - beta_min) - 0.5 * t * beta_min
    mean = torch.exp(log_mean_coeff)[(...,) + (None,) * len(x.shape[1:])] * x
    std = 1 - torch.exp(2.0 * log_mean_coeff)
    return mean, std


def snr_marginal(x: torch.Tensor, t: torch.Tensor, a=2, b=3, c=6, m0=-10, m1=5):
    """SNRSDE.marginal_prob, sde_lib.py:153-186 (gamma(t)=a t + b t^c, normalised log-SNR)."""
    gamma = lambda u: a * u + b * u ** c
    norm = (m1 - m0) / (gamma(1) - gamma(0))
    log_snr = -(m0 + norm * (gamma(t) - gamma(0)))
    snr = torch.exp(log_snr)
    alpha = torch.sqrt(snr / (1 + snr))[(...,) + (None,) * len(x.shape[1:])]
    mean = alpha * x
    std = torch.sqrt(1 / (1 + snr))
    return mean, std


SAMPLING_EPS = {"vesde": 1e-5, "vpsde": 1e-3, "subvpsde": 1e-3, "snrsde": 1e-3}
"""BaseSdeGenerativeModel.configure_sde, lightning_modules/BaseSdeGenerativeModel.py:27-47."""


def marginal(sde: str, x, t, **p):
    sde = sde.lower()
    if sde == "vesde":
        return ve_marginal(x, t, p["sigma_min"], p["sigma_max"])
    if sde == "vpsde":
        return vp_marginal(x, t, p["beta_min"], p["beta_max"])
    if sde == "subvpsde":
        return subvp_marginal(x, t, p["beta_min"], p["beta_max"])
    if sde == "snrsde":
        return snr_marginal(x, t)
    raise NotImplementedError(sde)


# --------------------------------------------------------------------------------------
# K schedule  (dim_reduction.py:166-171,185-188)
# --------------------------------------------------------------------------------------


def k_schedule(anchor_shape: Sequence[int], batchsize: int) -> Tuple[int, int, int]:
    """Returns (num_batches, extra_in_last_batch, K) exactly as the reference's integer arithmetic:
    `x` is ONE sample, so `x.shape[1:]` drops the first data axis (dim_reduction.py:162,166)."""
    ambient_dim = math.prod(tuple(anchor_shape)[1:])
    num_batches = ambient_dim // batchsize + 1
    extra = ambient_dim - (ambient_dim // batchsize) * batchsize
    num_batches *= 4
    K = (num_batches - 1) * batchsize + extra
    return num_batches, extra, K


# --------------------------------------------------------------------------------------
# Score networks
# --------------------------------------------------------------------------------------


def fcn_forward(weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor],
                x: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    """FCN.forward, models/fcn.py:32-40 with layers :16-30 (dropout is identity in eval):
    cat([x, t]) -> Linear -> ELU -> L x (Linear -> ELU) -> Linear."""
    h = torch.cat([x, t.unsqueeze(-1)], dim=1)
    n = len(weights)
    for i, (w, b) in enumerate(zip(weights, biases)):
        h = F.linear(h, w, b)
        if i < n - 1:
            h = F.elu(h)
    return h


def ksphere_gt_forward(batch: torch.Tensor, sigmas: torch.Tensor, manifold_dim: int) -> torch.Tensor:
    """KSphereGT.forward, models/ksphere_gt.py:14-26: per-row numpy loop,
    (x/r) * (ive(p/2,k)/ive(p/2-1,k) - r) / sigma^2 with k = r/sigma^2 (all numpy scalars,
    dtype follows numpy promotion of the float32 inputs)."""
    from scipy.special import ive
    b = batch.detach().cpu().numpy()
    s = sigmas.detach().cpu().numpy()
    p = manifold_dim
    out = []
    for x, sigma in zip(b, s):
        r = np.linalg.norm(x)
        k = r / sigma ** 2
        vm = -ive(p / 2, k) / ive(p / 2 - 1, k)
        out.append((x / r) * (-vm - r) / sigma ** 2)
    return torch.from_numpy(np.stack(out))


def make_score_fn(model_kind: str, sde: str, sde_params: Dict, N: int = 1000,
                  fcn_params: Optional[Tuple[Sequence[torch.Tensor], Sequence[torch.Tensor]]] = None,
                  manifold_dim: Optional[int] = None) -> Callable:
    """get_score_fn (unconditional, continuous), models/utils.py:156-175,238-277.
    * KSphereGT: std is passed instead of labels, output returned unchanged (:169-175).
    * otherwise labels = t*(N-1), score = -model(x, labels)/std (:238-277)."""

    def std_of(x, t):
        return marginal(sde, torch.zeros_like(x), t, **sde_params)[1]

    if model_kind == "ksphere_gt":
        def score_fn(x, t):
            return ksphere_gt_forward(x, std_of(x, t), manifold_dim)
        return score_fn
    if model_kind == "fcn":
        ws, bs = fcn_params

        def score_fn(x, t):
            labels = t * (N - 1)
            out = fcn_forward(ws, bs, x, labels)
            std = std_of(x, t)
            return -out / std[(...,) + (None,) * len(x.shape[1:])]
        return score_fn
    raise NotImplementedError(model_kind)


# --------------------------------------------------------------------------------------
# Spectrum and read-out
# --------------------------------------------------------------------------------------


def centred_spectrum(scores: torch.Tensor) -> List[float]:
    """dim_reduction.py:190-199: flatten, subtract column means, full SVD, s.tolist()."""
    scores = torch.flatten(scores, start_dim=1)
    means = scores.mean(dim=0, keepdim=True)
    normalized = scores - means
    _, s, _ = torch.linalg.svd(normalized)
    return s.tolist()


def id_from_spectrum(s: Sequence[float]) -> int:
    """plot_utils.py:224-229 (duplicate :175-182): gap rule in Python floats.
    Raises ZeroDivisionError when s[1]==s[2], like the reference's list comprehension does."""
    s = list(s)
    norm_factor = s[1] - s[2]
    diff = [(s[i] - s[i + 1]) / norm_factor for i in range(1, len(s) - 1)]
    x = np.asarray(diff)
    e_x = np.exp(x - np.max(x))
    soft = e_x / e_x.sum(axis=0)
    return int(len(soft) - soft.argmax())


def extract_sing_vals(svd: Dict, mode: str = "first"):
    """plot_utils.py:197-205."""
    sv = svd["singular_values"]
    if mode == "first":
        return [sv[0]]
    if mode == "all":
        return sv
    if mode == "mean":
        return [np.mean(sv, axis=0)]
    raise ValueError(mode)


# --------------------------------------------------------------------------------------
# Driver  (dim_reduction.py:116-211)
# --------------------------------------------------------------------------------------


def manifold_dimension(anchor_batches: Sequence[torch.Tensor], num_datapoints: int,
                       score_fn: Callable, sde: str, sde_params: Dict,
                       randn_like: Callable = torch.randn_like,
                       capture: Optional[Dict] = None) -> Dict:
    """The two hot loops of get_manifold_dimension, dim_reduction.py:150-206, for anchors that
    arrive as loader batches (list of [B, ...] tensors).  Replicates the N-1 anchor quirk
    (:159-164), the K schedule (:166-171), the perturbation draw (:178-182), the last-batch
    truncation (:185-188), centring and SVD (:190-199).  `randn_like` is injectable so tests can
    replay recorded noise; `capture` (optional dict) receives per-anchor raw scores."""
    t_eps = SAMPLING_EPS[sde.lower()]
    singular_values = []
    idx = 0
    for orig_batch in anchor_batches:
        batchsize = orig_batch.size(0)
        if idx + 1 >= num_datapoints:
            break
        for x in orig_batch:
            if idx + 1 >= num_datapoints:
                break
            num_batches, extra, _ = k_schedule(x.shape, batchsize)
            xr = x.repeat([batchsize] + [1 for _ in range(len(x.shape))])
            vec_t = torch.ones(xr.size(0)) * t_eps
            scores = []
            for i in range(1, num_batches + 1):
                batch = xr.clone()
                mean, std = marginal(sde, batch, vec_t, **sde_params)
                z = randn_like(batch)
                batch = mean + std[(...,) + (None,) * len(batch.shape[1:])] * z
                score = score_fn(batch, vec_t).detach().cpu()
                scores.append(score if i < num_batches else score[:extra])
            scores = torch.cat(scores, dim=0)
            if capture is not None:
                capture.setdefault("scores", []).append(torch.flatten(scores, start_dim=1))
            singular_values.append(centred_spectrum(scores))
            idx += 1
    return {"singular_values": singular_values}


# --------------------------------------------------------------------------------------
# Synthetic anchors (KSphereDataset.py:21-95) -- input generator for the BASELINE configs
# --------------------------------------------------------------------------------------


def ksphere_data(n_samples: int, n_spheres: int, ambient_dim: int, manifold_dim, noise_std: float,
                 embedding_type: str, radii=(), angle_std: float = -1) -> torch.Tensor:
    """KSphereDataset.generate_data / sample_sphere, lightning_data_modules/KSphereDataset.py:21-95.
    Uses the torch global generator exactly where the reference does."""
    radii = list(radii) if len(radii) else [1] * n_spheres
    mdims = [manifold_dim] * n_spheres if isinstance(manifold_dim, int) else list(manifold_dim)
    data = []
    for i in range(n_spheres):
        md = mdims[i]
        if angle_std == -1:
            new = torch.randn((n_samples, md + 1))
            new = new / torch.linalg.norm(new, dim=1)[:, None]
        else:
            ang = angle_std * torch.randn((n_samples, md))
            rows = []
            for a in ang:
                xs, sp = [], 1
                for j in range(len(a)):
                    xs.append(sp * torch.cos(a[j]))
                    sp = sp * torch.sin(a[j])
                xs.append(sp)
                rows.append(torch.stack(xs)[None, ...])
            new = torch.cat(rows, dim=0)
        new = new * radii[i]
        if embedding_type == "random_isometry":
            g = torch.Generator().manual_seed(0)
            emb = torch.randn(size=(ambient_dim, md + 1), generator=g)
            q, _ = np.linalg.qr(emb.numpy())
            q = torch.from_numpy(q)
            new = (q @ new.T).T
        elif embedding_type == "first":
            new = torch.cat([new, torch.zeros([n_samples, ambient_dim - new.shape[1]])], dim=1)
        elif embedding_type == "separating":
            if n_spheres * (md + 1) > ambient_dim:
                raise RuntimeError("Cant fit that many spheres.")
            new = torch.cat([torch.zeros((n_samples, i * (md + 1))), new], dim=1)
            new = torch.cat([new, torch.zeros([n_samples, ambient_dim - new.shape[1]])], dim=1)
        elif embedding_type == "along_axis":
            if (n_spheres - 1) + (md + 1) > ambient_dim:
                raise RuntimeError("Cant fit that many spheres.")
            new = torch.cat([torch.zeros((n_samples, i)), new], dim=1)
            new = torch.cat([new, torch.zeros([n_samples, ambient_dim - new.shape[1]])], dim=1)
        else:
            raise RuntimeError("Unknown embedding type.")
        new = new + noise_std * torch.randn_like(new)
        data.append(new)
    return torch.cat(data, dim=0)


# --------------------------------------------------------------------------------------
# fp64 evaluation of the same restatement -- measures the reference's own fp32 noise floor
# --------------------------------------------------------------------------------------


def fcn_centred_spectrum_fp64(weights, biases, x0: torch.Tensor, z: torch.Tensor, sde: str, sde_params: Dict,
                              N: int = 1000) -> np.ndarray:
    """Singular values of the centred FCN score matrix with every operation AFTER the perturbation draw
    in float64.  The perturbed points x = mean + std*z are formed in float32 exactly as
    dim_reduction.py:178-182 does (they are the network INPUT, identical for every implementation);
    fcn.py:32-40, the -1/std scaling (models/utils.py:264-267), centring and the SVD
    (dim_reduction.py:190-199) then run in float64.  |reference - this| is the rounding noise of the
    reference's fp32 arithmetic for that input, the floor below which 'matching the reference' is
    not defined.  x0 [n, D], z [n, K, D] float32 -> [n, D] float64."""
    t_eps = SAMPLING_EPS[sde.lower()]
    n, K, D = z.shape
    out = []
    for a in range(n):
        vec_t = torch.ones(K) * t_eps
        batch = x0[a].repeat(K, 1)
        mean, std = marginal(sde, batch, vec_t, **sde_params)
        x = mean + std[:, None] * z[a]
        labels = vec_t * (N - 1)
        h = torch.cat([x.double(), labels.double().unsqueeze(-1)], dim=1)
        for i, (w, b) in enumerate(zip(weights, biases)):
            h = h @ w.double().T + b.double()
            if i < len(weights) - 1:
                h = torch.where(h > 0, h, torch.expm1(h))
        score = -h / std.double()[:, None]
        c = score - score.mean(dim=0, keepdim=True)
        out.append(torch.linalg.svdvals(c).numpy())
    return np.stack(out)
