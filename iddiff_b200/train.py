"""Denoising-score-matching training step of the FCN score network on the sm_100a kernels.

Reference: `losses.py:164-188` (unconditional `loss_fn`), `models/utils.py:254-266` (VE score wrapper: the network predicts
the noise), `losses.py:26-51` (Adam, linear warm-up, `clip_grad_norm_`), `BaseSdeGenerativeModel.training_step` (:60-63).
One step =

    sample   t ~ U(eps, T), x_t = mean(t) x0 + std(t) z                       idb_dsm_sample (Philox z, input layout [x_t, (N-1) t])
    forward  pre_l = H_{l-1} W_l^T + b_l, H_l = ELU(pre_l)                    idb_fcn_anchor_layer   (tcgen05, bf16 pairs)
    loss     mean (pre_L - z)^2, dZ_L = 2 (pre_L - z) / (B D)                 idb_dsm_loss
    backward dW_l = dZ_l^T H_{l-1}; dH_{l-1} = dZ_l W_l; dZ_{l-1} = dH_{l-1} ELU'(pre_{l-1})
                                                                              idb_train_pack (plain / transposed operands,
                                                                              activation derivative fused) + idb_fcn_output_layer
             db_l = column sums of dZ_l                                       idb_column_sums_f32
    update   clip by the global norm, Adam                                    idb_grad_sumsq + idb_adam_step (no host sync)

The weights stay owned by the caller's `nn.Module` (updated in place); this object keeps their operand slabs (plain and
transposed, re-packed after every update), the Adam state and the activation / gradient buffers.  torch supplies device
memory, the uniform draw of `t` and the batch indices only.  There is no CPU path.
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr

FMT = _lib.FMT_BF16   # bf16 pairs: 16 significand bits with the fp32 exponent range, so gradients need no loss scaling


def _ceil(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class FcnTrainer:
    def __init__(self, fcn_module: torch.nn.Module, batch: int, *, lr: float = 2e-4, beta1: float = 0.9,
                 beta2: float = 0.999, eps: float = 1e-8, warmup: int = 0, grad_clip: float = 1.0):
        self.lin: List[torch.nn.Linear] = [m for m in fcn_module.mlp if isinstance(m, torch.nn.Linear)]
        if len(self.lin) < 2:
            raise _lib.IdbError("FcnTrainer: models/fcn.py FCN expected (>= 2 Linear layers in .mlp)")
        w0 = self.lin[0].weight
        if not w0.is_cuda or w0.dtype != torch.float32:
            raise _lib.IdbError("FcnTrainer: the module must hold float32 CUDA parameters (no CPU path)")
        self.dev = w0.device
        self.B, self.Mp = int(batch), _ceil(int(batch), 128)
        self.D = self.lin[-1].weight.shape[0]
        if self.lin[0].weight.shape[1] != self.D + 1:
            raise _lib.IdbError("FcnTrainer: first layer must take state_size + 1 inputs (models/fcn.py:35-38)")
        self.lr, self.beta1, self.beta2, self.eps = float(lr), float(beta1), float(beta2), float(eps)
        self.warmup, self.grad_clip = int(warmup), float(grad_clip)
        self.n_in = [l.weight.shape[1] for l in self.lin]
        self.n_out = [l.weight.shape[0] for l in self.lin]
        self.Pin = [_ceil(n, 256) for n in self.n_in]      # every feature extent is padded to the GEMM's N tile
        self.Pout = [_ceil(n, 256) for n in self.n_out]
        dev, f32, u8 = self.dev, torch.float32, torch.uint8
        z = lambda *s, dtype=f32: torch.zeros(*s, dtype=dtype, device=dev)
        L = len(self.lin)
        # operand slabs of the weights: plain [Pout][Pin] (forward) and transposed [Pin][Pout] (dH = dZ W)
        self.w = [(z(self.Pout[l] * self.Pin[l] * 2, dtype=u8), z(self.Pout[l] * self.Pin[l] * 2, dtype=u8)) for l in range(L)]
        self.wt = [(z(self.Pout[l] * self.Pin[l] * 2, dtype=u8), z(self.Pout[l] * self.Pin[l] * 2, dtype=u8)) for l in range(L)]
        self.bias = [z(self.Pout[l]) for l in range(L)]
        # activations: X / pre_l fp32, A-form operand of H_{l-1} (forward), T-form operand of H_{l-1} (dW)
        self.X = z(self.Mp, self.Pin[0])
        self.zbuf = z(self.B, self.D)
        self.pre = [z(self.Mp, self.Pout[l]) for l in range(L)]
        self.actA = [(z(self.Mp * self.Pin[l] * 2, dtype=u8), z(self.Mp * self.Pin[l] * 2, dtype=u8)) for l in range(L)]
        self.actT = [(z(self.Pin[l] * self.Mp * 2, dtype=u8), z(self.Pin[l] * self.Mp * 2, dtype=u8)) for l in range(L)]
        # gradients
        self.dZ = [z(self.Mp, self.Pout[l]) for l in range(L)]
        self.dZA = [(z(self.Mp * self.Pout[l] * 2, dtype=u8), z(self.Mp * self.Pout[l] * 2, dtype=u8)) for l in range(L)]
        self.dZT = [(z(self.Pout[l] * self.Mp * 2, dtype=u8), z(self.Pout[l] * self.Mp * 2, dtype=u8)) for l in range(L)]
        self.dH = [z(self.Mp, self.Pin[l]) for l in range(L)]
        self.dW = [z(self.Pout[l], self.Pin[l]) for l in range(L)]
        self.db = [z(self.Pout[l]) for l in range(L)]
        self.m_w = [z(self.n_out[l], self.n_in[l]) for l in range(L)]
        self.v_w = [z(self.n_out[l], self.n_in[l]) for l in range(L)]
        self.m_b = [z(self.n_out[l]) for l in range(L)]
        self.v_b = [z(self.n_out[l]) for l in range(L)]
        self.normsq = z(1, dtype=torch.float64)
        self.loss = z(1)
        self.loss_scratch = z(64, dtype=torch.float64)
        self.step_count = 0                                   # host mirror of the device counter below
        self.step_dev = z(1, dtype=torch.int64)               # optimiser step, read by the sampling and Adam kernels
        self._graph = None
        self.repack()

    # ---- weights -> operand slabs (after construction and after every update)
    def repack(self):
        st = stream_ptr()
        for l, lin in enumerate(self.lin):
            w = lin.weight.detach()
            call("idb_train_pack", ptr(w), w.stride(0), self.n_out[l], self.n_in[l], None, 0, 0, 0, FMT, 1.0, self.Pout[l],
                 self.Pin[l], ptr(self.w[l][0]), ptr(self.w[l][1]), None, 0, st)
            call("idb_train_pack", ptr(w), w.stride(0), self.n_out[l], self.n_in[l], None, 0, 0, 1, FMT, 1.0, self.Pout[l],
                 self.Pin[l], ptr(self.wt[l][0]), ptr(self.wt[l][1]), None, 0, st)
            self.bias[l][:self.n_out[l]].copy_(lin.bias.detach())

    # ---- one step
    def sample(self, data: torch.Tensor, idx: Optional[torch.Tensor], t: torch.Tensor, mean_coeff: torch.Tensor,
               std: torch.Tensor, label_scale: float, z: Optional[torch.Tensor] = None, seed: int = 0):
        call("idb_dsm_sample", ptr(data), ptr(idx), ptr(t), ptr(mean_coeff), ptr(std), ptr(z), self.B, self.D, self.Mp,
             self.Pin[0], float(label_scale), int(seed), 0, ptr(self.step_dev), ptr(self.X), ptr(self.zbuf), stream_ptr())

    def forward(self):
        """X -> pre_L (the noise prediction, fp32 [Mp][Pout_L]); keeps every pre-activation for the backward pass."""
        st = stream_ptr()
        L = len(self.lin)
        call("idb_train_pack", ptr(self.X), self.Pin[0], self.B, self.n_in[0], None, 0, 0, 0, FMT, 1.0, self.Mp, self.Pin[0],
             ptr(self.actA[0][0]), ptr(self.actA[0][1]), None, 0, st)
        for l in range(L):
            last = l == L - 1
            nxt = (None, None) if last else self.actA[l + 1]
            call("idb_fcn_anchor_layer", ptr(self.actA[l][0]), ptr(self.actA[l][1]), ptr(self.w[l][0]), ptr(self.w[l][1]), FMT,
                 1.0, 1.0, self.Mp, self.Pout[l], self.Pin[l], ptr(self.bias[l]), ptr(self.pre[l]), ptr(nxt[0]), ptr(nxt[1]), st)
        return self.pre[-1]

    def loss_and_backward(self):
        st = stream_ptr()
        L = len(self.lin)
        call("idb_dsm_loss", ptr(self.pre[-1]), self.Pout[-1], ptr(self.zbuf), self.B, self.D, self.Mp, ptr(self.dZ[-1]),
             ptr(self.loss), ptr(self.loss_scratch), st)
        for l in range(L - 1, -1, -1):
            dz = self.dZ[l]
            # operands of this layer's two products: dZ_l as [batch][features] and as [features][batch]
            if l > 0:
                call("idb_train_pack", ptr(dz), self.Pout[l], self.B, self.n_out[l], None, 0, 0, 0, FMT, 1.0, self.Mp,
                     self.Pout[l], ptr(self.dZA[l][0]), ptr(self.dZA[l][1]), None, 0, st)
            call("idb_train_pack", ptr(dz), self.Pout[l], self.B, self.n_out[l], None, 0, 0, 1, FMT, 1.0, self.Mp, self.Pout[l],
                 ptr(self.dZT[l][0]), ptr(self.dZT[l][1]), None, 0, st)
            # H_{l-1} as [features][batch]: the network input for l = 0, ELU(pre_{l-1}) otherwise
            if l == 0:
                call("idb_train_pack", ptr(self.X), self.Pin[0], self.B, self.n_in[0], None, 0, 0, 1, FMT, 1.0, self.Mp,
                     self.Pin[0], ptr(self.actT[0][0]), ptr(self.actT[0][1]), None, 0, st)
            else:
                call("idb_train_pack", ptr(self.pre[l - 1]), self.Pout[l - 1], self.B, self.n_out[l - 1], None, 0, 1, 1, FMT,
                     1.0, self.Mp, self.Pin[l], ptr(self.actT[l][0]), ptr(self.actT[l][1]), None, 0, st)
            # dW_l [Pout][Pin] = dZ_l^T H_{l-1}: contraction over the (padded) batch
            call("idb_fcn_output_layer", ptr(self.dZT[l][0]), ptr(self.dZT[l][1]), ptr(self.actT[l][0]), ptr(self.actT[l][1]),
                 FMT, 1.0, self.Pout[l], self.Pin[l], self.Mp, ptr(self.dW[l]), st)
            call("idb_column_sums_f32", ptr(dz), self.Pout[l], self.B, self.n_out[l], ptr(self.db[l]), st)
            if l > 0:
                # dH_{l-1} [Mp][Pin] = dZ_l W_l, then dZ_{l-1} = dH_{l-1} * ELU'(pre_{l-1}) (fp32, for the next round)
                call("idb_fcn_output_layer", ptr(self.dZA[l][0]), ptr(self.dZA[l][1]), ptr(self.wt[l][0]), ptr(self.wt[l][1]),
                     FMT, 1.0, self.Mp, self.Pin[l], self.Pout[l], ptr(self.dH[l]), st)
                call("idb_train_pack", ptr(self.dH[l]), self.Pin[l], self.B, self.n_in[l], ptr(self.pre[l - 1]),
                     self.Pout[l - 1], 2, 0, FMT, 1.0, self.Mp, self.Pin[l], None, None, ptr(self.dZ[l - 1]), self.Pout[l - 1], st)
        return self.loss

    def update(self):
        """clip_grad_norm_ + Adam with linear warm-up (losses.py:40-50), then the operand slabs are refreshed.  The step
        counter, the warmed-up rate, the bias corrections and the clip coefficient all live on the device."""
        st = stream_ptr()
        self.normsq.zero_()
        clip = self.grad_clip >= 0
        if clip:
            for l in range(len(self.lin)):
                call("idb_grad_sumsq", ptr(self.dW[l]), self.Pin[l], self.n_out[l], self.n_in[l], ptr(self.normsq), st)
                call("idb_grad_sumsq", ptr(self.db[l]), self.n_out[l], 1, self.n_out[l], ptr(self.normsq), st)
        nq = ptr(self.normsq) if clip else None
        for l, lin in enumerate(self.lin):
            w, b = lin.weight.detach(), lin.bias.detach()
            call("idb_adam_step", ptr(w), w.stride(0), ptr(self.dW[l]), self.Pin[l], ptr(self.m_w[l]), ptr(self.v_w[l]),
                 self.n_out[l], self.n_in[l], self.lr, self.warmup, self.beta1, self.beta2, self.eps, 0, ptr(self.step_dev),
                 self.grad_clip, nq, st)
            call("idb_adam_step", ptr(b), self.n_out[l], ptr(self.db[l]), self.n_out[l], ptr(self.m_b[l]), ptr(self.v_b[l]), 1,
                 self.n_out[l], self.lr, self.warmup, self.beta1, self.beta2, self.eps, 0, ptr(self.step_dev), self.grad_clip,
                 nq, st)
        self.repack()

    def _one_step(self, data, sde, eps_t, idx, t, z, seed):
        self.step_dev.add_(1)
        if idx is None and data.shape[0] != self.B:
            idx = torch.randint(0, data.shape[0], (self.B,), device=self.dev)
        if t is None:
            t = torch.rand(self.B, device=self.dev) * (sde.T - eps_t) + eps_t
        mean, std = sde.marginal_prob(torch.ones(self.B, 1, device=self.dev), t)
        self.sample(data, idx, t.contiguous(), mean.reshape(-1).contiguous(), std.contiguous(), float(sde.N - 1), z, seed)
        self.forward()
        self.loss_and_backward()
        self.update()

    def step(self, data: torch.Tensor, sde, *, eps_t: float = 1e-5, idx: Optional[torch.Tensor] = None,
             t: Optional[torch.Tensor] = None, z: Optional[torch.Tensor] = None, seed: int = 0) -> torch.Tensor:
        """One training step on a batch of `data` rows (losses.py:172-187): returns the loss (a device scalar)."""
        self.step_count += 1
        self._one_step(data, sde, eps_t, idx, t, z, seed)
        return self.loss

    def step_graph(self, data: torch.Tensor, sde, *, eps_t: float = 1e-5, seed: int = 0) -> torch.Tensor:
        """The same step with internally drawn batch indices, times and Philox noise, captured ONCE as a CUDA graph
        (~120 launches: sampling, 7 forward GEMMs, 13 backward GEMMs, packing, norm, Adam, re-packing) and replayed: the
        step counter, learning-rate warm-up and bias corrections are read from device memory, torch's generator feeds the
        captured uniform / integer draws.  `data` must stay at the same address between calls."""
        self.step_count += 1
        if self._graph is None or self._graph[1] != (data.data_ptr(), tuple(data.shape), float(eps_t), int(seed)):
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(torch.cuda.current_stream(self.dev))
            with torch.cuda.stream(side):          # warm-up run outside the capture (lazy kernel attributes); counted
                self._one_step(data, sde, eps_t, None, None, None, seed)
            torch.cuda.current_stream(self.dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._one_step(data, sde, eps_t, None, None, None, seed)
            self._graph = (graph, (data.data_ptr(), tuple(data.shape), float(eps_t), int(seed)))
            return self.loss                        # the warm-up run WAS this step; the capture itself executes nothing
        self._graph[0].replay()
        return self.loss


def short_train(holder: torch.nn.Module, train_points: torch.Tensor, steps: int, *, batch: int = 500, lr: float = 2e-4,
                warmup: int = 500, sigma_min: float = 1e-2, sigma_max: float = 4.0, eps_t: float = 1e-5, seed: int = 0,
                on_step=None, graph: bool = True):
    """`steps` DSM steps on `train_points` [n, D] (device) through the CUDA training step; updates `holder` in place and
    returns the smoothed final loss (same contract as tools/train_fcn.short_train, which runs the reference's loop in
    PyTorch)."""
    from . import sde_lib
    sde = sde_lib.VESDE(sigma_min=sigma_min, sigma_max=sigma_max, N=1000)
    tr = FcnTrainer(holder, batch, lr=lr, warmup=warmup, grad_clip=1.0)
    run = None
    for step in range(1, steps + 1):
        loss = tr.step_graph(train_points, sde, eps_t=eps_t, seed=seed) if graph else \
            tr.step(train_points, sde, eps_t=eps_t, seed=seed)
        if step % 50 == 0:
            v = float(loss)
            run = v if run is None else 0.9 * run + 0.1 * v
        if on_step is not None:
            on_step(step, run)
    return run
