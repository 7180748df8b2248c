"""Host-side SDE marginals for the `manifold_dimension` path.

Mirrors the part of the reference's `sde_lib` this path touches -- the class names, constructor
arguments, `.N` and `marginal_prob(x, t) -> (mean, std)` -- so configs and call sites carry over:
    VESDE     sde_lib.py:316-347      VPSDE   sde_lib.py:222-254
    subVPSDE  sde_lib.py:276-304      SNRSDE  sde_lib.py:153-186
Drift/diffusion/reverse/discretize/prior sampling belong to samplers and likelihoods and are out of
scope.

On this path `t` is one constant for every row of every batch (dim_reduction.py:173-174), so the
engine asks for two fp32 scalars, `coefficients(t) -> (mean_coeff, std)`, computed with the same
float32 torch ops as `marginal_prob` (bit-identical to the reference's per-row values), and the
per-element work happens in the CUDA perturbation kernel.
"""
from __future__ import annotations

import torch


class SDE:
    T = 1.0   # end time of every SDE of sde_lib.py (`.T` properties :44-46, :159-161, :228-230, :322-324)

    def __init__(self, N: int):
        self.N = N

    def marginal_prob(self, x: torch.Tensor, t: torch.Tensor):
        raise NotImplementedError

    def coefficients(self, t: float):
        """(mean_coeff, std) as Python floats holding exact float32 values, for x -> mean_coeff*x + std*z."""
        tt = torch.ones(1, dtype=torch.float32) * t
        mean, std = self.marginal_prob(torch.ones(1, 1, dtype=torch.float32), tt)
        return float(mean.reshape(-1)[0]), float(std.reshape(-1)[0])

    @staticmethod
    def _bcast(v, x):
        return v[(...,) + (None,) * len(x.shape[1:])]


class VESDE(SDE):
    def __init__(self, sigma_min=0.01, sigma_max=50, N=1000, data_mean=None):
        super().__init__(N)
        self.sigma_min, self.sigma_max, self.data_mean = sigma_min, sigma_max, data_mean

    def marginal_prob(self, x, t):
        # the two constants as tensors of t's type (same float32 arithmetic as before), cached per (device, dtype) so that
        # a captured training step (train.FcnTrainer.step_graph) performs no host-to-device copy
        key = (t.device, t.dtype)
        c = getattr(self, "_consts", {}).get(key)
        if c is None:
            c = (torch.tensor(self.sigma_min).type_as(t), torch.tensor(self.sigma_max).type_as(t))
            self.__dict__.setdefault("_consts", {})[key] = c
        smin, smax = c
        return x, smin * (smax / smin) ** t


class VPSDE(SDE):
    def __init__(self, beta_min=0.1, beta_max=20, N=1000):
        super().__init__(N)
        self.beta_0, self.beta_1 = beta_min, beta_max

    def _lmc(self, t):
        return -0.25 * t ** 2 * (self.beta_1 - self.beta_0) - 0.5 * t * self.beta_0

    def marginal_prob(self, x, t):
        lmc = self._lmc(t)
        return torch.exp(self._bcast(lmc, x)) * x, torch.sqrt(1.0 - torch.exp(2.0 * lmc))


class subVPSDE(VPSDE):
    def marginal_prob(self, x, t):
        lmc = self._lmc(t)
        # the reference does not take the square root here (sde_lib.py:300-304); replicated
        return self._bcast(torch.exp(lmc), x) * x, 1 - torch.exp(2.0 * lmc)


class SNRSDE(SDE):
    def __init__(self, N, gamma=None, a=2, b=3, c=6, minus_log_SNR_0=-10, minus_log_SNR_1=5):
        super().__init__(N)
        self.gamma = gamma if gamma is not None else (lambda u: a * u + b * u ** c)
        self.m0, self.m1 = minus_log_SNR_0, minus_log_SNR_1
        self.norm = (self.m1 - self.m0) / (self.gamma(1) - self.gamma(0))

    def marginal_prob(self, x, t):
        log_snr = -(self.m0 + self.norm * (self.gamma(t) - self.gamma(0)))
        snr = torch.exp(log_snr)
        return self._bcast(torch.sqrt(snr / (1 + snr)), x) * x, torch.sqrt(1 / (1 + snr))


def configure_sde(config):
    """BaseSdeGenerativeModel.configure_sde, lightning_modules/BaseSdeGenerativeModel.py:27-47
    -> (sde, sampling_eps)."""
    name = config.training.sde.lower()
    m = config.model
    if name == "vpsde":
        return VPSDE(beta_min=m.beta_min, beta_max=m.beta_max, N=m.num_scales), 1e-3
    if name == "subvpsde":
        return subVPSDE(beta_min=m.beta_min, beta_max=m.beta_max, N=m.num_scales), 1e-3
    if name == "vesde":
        # data.use_data_mean (BaseSdeGenerativeModel.py:36-41) loads a dataset mean that only shifts prior_sampling
        # (sde_lib.py:368-371); marginal_prob -- all this path uses -- does not read it, so the flag changes nothing here
        return VESDE(sigma_min=m.sigma_min, sigma_max=m.sigma_max, N=m.num_scales), 1e-5
    if name == "snrsde":
        return SNRSDE(N=m.num_scales), 1e-3
    raise NotImplementedError(f"SDE {config.training.sde} unknown.")
