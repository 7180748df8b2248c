"""Drop-in for the reference's `dim_reduction.get_manifold_dimension` (dim_reduction.py:116-211).

Same signature, same config keys, same output (`{'singular_values': list[list[float]]}` returned or pickled
to {log_path}/{log_name}/svd/{name}.pkl), same quirks (N-1 anchors, K from the loader batch size, file
'None.pkl' when name is None).  The body is different: anchors are gathered from the loader, moved to the
device once, and every stage -- perturbation, score network, centring, spectrum, gap read-out -- runs on
the GPU for a whole chunk of anchors; only the spectra come back.
"""
from __future__ import annotations

import math
import os
import pickle
from pathlib import Path
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib, sde_lib
from .config import lookup
from .engine import FcnEngine, KSphereEngine, SpectrumResult


def k_schedule(anchor_shape: Sequence[int], batchsize: int):
    """(num_batches, extra_in_last_batch, K) with the reference's integer arithmetic, dim_reduction.py:166-171,
    185-188: `x` is ONE sample there, so `x.shape[1:]` drops the first DATA axis."""
    ambient_dim = math.prod(tuple(anchor_shape)[1:])
    num_batches = ambient_dim // batchsize + 1
    extra = ambient_dim - (ambient_dim // batchsize) * batchsize
    num_batches *= 4
    return num_batches, extra, (num_batches - 1) * batchsize + extra


def gather_anchors(loader, num_datapoints: int):
    """The anchors the reference's two loops visit (dim_reduction.py:154-164): idx+1 >= num_datapoints breaks,
    so num_datapoints-1 anchors are used.  Returns [(anchor tensor, loader batch size B)], B decides K."""
    out, idx = [], 0
    for orig_batch in loader:
        batchsize = orig_batch.size(0)
        if idx + 1 >= num_datapoints:
            break
        for x in orig_batch:
            if idx + 1 >= num_datapoints:
                break
            out.append((x, batchsize))
            idx += 1
    return out


def batch_size_runs(sizes: Sequence[int]):
    """Maximal runs of CONSECUTIVE anchors that share a loader batch size (hence K): [(B, [i, i+1, ...])].  One
    device batch per run keeps the Philox key of every anchor equal to its global index, whatever the grouping."""
    runs: List = []
    for i, B in enumerate(sizes):
        if runs and runs[-1][0] == B:
            runs[-1][1].append(i)
        else:
            runs.append((B, [i]))
    return runs


def build_engine(config, score_model: Optional[torch.nn.Module] = None, device="cuda"):
    """Recognises the model class by config.model.name and ingests its weights into a device engine."""
    name = config.model.name
    if name == "ksphere_gt":
        return KSphereEngine(config.data.manifold_dim, device=device)
    if name == "fcn":
        if score_model is None:
            from . import models
            score_model = models.create_model(config)
            path = lookup(config, "model.checkpoint_path")
            if path:
                models.load_checkpoint_into(score_model, path)
        return FcnEngine.from_module(score_model, device=device)
    if name in ("ddpm", "ncsnpp", "BeatGANsUNetModel"):
        from . import models, unet_models  # noqa: F401  (registers 'ddpm')
        from .unet_engine import UNetEngine
        if score_model is None:
            score_model = models.create_model(config)
            path = lookup(config, "model.checkpoint_path")
            if path:
                models.load_checkpoint_into(score_model, path)
        elif not hasattr(score_model, "plan"):
            # a live reference module (training callback, callbacks.py:403-432): same state_dict layout -> our holder
            holder = models.create_model(config)
            holder.load_state_dict(score_model.state_dict(), strict=True)
            score_model = holder
        eng = UNetEngine.from_module(score_model, device=device)
        eng.image_shape = tuple(config.data.shape)
        return eng
    raise NotImplementedError(f"model.name={name!r}: 'fcn', 'ksphere_gt', 'ddpm' and 'ncsnpp' are built so far (DESIGN.md scope)")


def sde_coefficients(sde, t: float):
    """(mean_coeff, std) as exact float32 values for x -> mean_coeff*x + std*z, from ANY object with the
    reference's `marginal_prob(x, t)` (sde_lib.py): evaluated on a ones tensor with the same float32 torch ops."""
    if hasattr(sde, "coefficients"):
        return sde.coefficients(t)
    tt = torch.ones(1, dtype=torch.float32) * t
    mean, std = sde.marginal_prob(torch.ones(1, 1, dtype=torch.float32), tt)
    return float(mean.reshape(-1)[0]), float(std.reshape(-1)[0])


def shard_range(n_anchors: int, rank: int, world: int):
    """Contiguous anchor shard [lo, hi) of this rank; the last ranks take the remainder-free tail."""
    per = (n_anchors + world - 1) // world
    lo = min(rank * per, n_anchors)
    return lo, min(lo + per, n_anchors)


def pack_results(sv: torch.Tensor, dims: torch.Tensor) -> torch.Tensor:
    """[A, n] fp32 spectra + [A] int32 IDs -> ONE [A, n + 1] fp32 buffer (IDs bit-cast into the last column), so the
    path's single collective carries both."""
    out = sv.new_empty(sv.shape[0], sv.shape[1] + 1)
    out[:, :-1].copy_(sv)
    out[:, -1].copy_(dims.to(torch.int32).view(torch.float32))
    return out


def unpack_results(packed: torch.Tensor):
    return packed[:, :-1], packed[:, -1].contiguous().view(torch.int32)


def all_gather_spectra(sv: torch.Tensor, dims: torch.Tensor, n_total: int, group=None):
    """The path's only collective, ONE all-gather: every rank contributes its packed [n_local, n + 1] results (shards
    padded to the common size), everyone receives [n_total, n] spectra / [n_total] IDs in global anchor order."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    per = (n_total + world - 1) // world
    packed = pack_results(sv, dims)
    pad = per - packed.shape[0]
    if pad:
        packed = torch.cat([packed, packed.new_zeros(pad, packed.shape[1])])
    out = packed.new_empty(world * per, packed.shape[1])
    dist.all_gather_into_tensor(out, packed.contiguous(), group=group)
    g_sv, g_dims = unpack_results(out[:n_total])
    return g_sv, g_dims


def all_gather_rows(S_loc: torch.Tensor, K: int, group=None) -> torch.Tensor:
    """K-sharding of ONE huge anchor (SURVEY 8e second row): every rank holds the score rows [lo, hi) =
    shard_range(K, rank, world) of S [A, K, D]; one all-gather (NCCL over NVLink on GPUs, gloo in the CPU tests) gives
    every rank the full [A, K, D] matrix in global row order.  Shards are padded to the common size for the collective."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    A, k_loc, D = S_loc.shape
    per = (K + world - 1) // world
    if k_loc < per:
        S_loc = torch.cat([S_loc, S_loc.new_zeros(A, per - k_loc, D)], dim=1)
    buf = S_loc.new_empty(world * A, per, D)   # concatenation along dim 0 (the form both NCCL and gloo accept)
    dist.all_gather_into_tensor(buf, S_loc.contiguous(), group=group)
    return buf.view(world, A, per, D).permute(1, 0, 2, 3).reshape(A, world * per, D)[:, :K].contiguous()


class _DeviceGramOps:
    """The three Gram pieces of the K-sharded mode on the device (C ABI); tests/test_dist_cpu.py swaps in a host
    stand-in to check the collective bookkeeping under gloo."""

    @staticmethod
    def column_sums(X: torch.Tensor) -> torch.Tensor:
        k, D = X.shape
        out = torch.empty(D, dtype=torch.float64, device=X.device)
        _lib.call("idb_column_sums", _lib.ptr(X), 0, X.stride(0), 1, k, D, _lib.ptr(out), _lib.stream_ptr())
        return out

    @staticmethod
    def gram_about(X: torch.Tensor, center: torch.Tensor, scale: float) -> torch.Tensor:
        k, D = X.shape
        G = torch.empty(D, D, dtype=torch.float64, device=X.device)
        _lib.call("idb_gram_about", _lib.ptr(X), 0, X.stride(0), 1, k, D, float(scale), _lib.ptr(center), _lib.ptr(G),
                  _lib.stream_ptr())
        return G

    @staticmethod
    def row_gram(X: torch.Tensor, scale: float) -> torch.Tensor:
        K, d = X.shape
        G = torch.empty(K, K, dtype=torch.float64, device=X.device)
        mean = torch.empty(d, dtype=torch.float64, device=X.device)
        _lib.call("idb_row_gram", _lib.ptr(X), X.stride(0), K, d, float(scale), _lib.ptr(mean), _lib.ptr(G),
                  _lib.stream_ptr())
        return G


def rows_to_column_panels(S_loc: torch.Tensor, K: int, group=None) -> torch.Tensor:
    """Re-shards ONE anchor's score matrix from row shards to column panels with one all-to-all: in, this rank's rows
    [k_loc, D] (rows shard_range(K, rank, world)); out, ALL K rows of its columns shard_range(D, rank, world) -> [K, d_loc]."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    k_loc, D = S_loc.shape
    rows = [shard_range(K, q, world) for q in range(world)]
    cols = [shard_range(D, q, world) for q in range(world)]
    assert k_loc == rows[rank][1] - rows[rank][0]
    d_loc = cols[rank][1] - cols[rank][0]
    send = torch.cat([S_loc[:, lo:hi].reshape(-1) for lo, hi in cols])
    recv = S_loc.new_empty(K * d_loc)
    dist.all_to_all_single(recv, send, [(hi - lo) * d_loc for lo, hi in rows], [k_loc * (hi - lo) for lo, hi in cols],
                           group=group)
    return recv.view(K, d_loc)      # source ranks arrive in rank order = global row order


def sharded_centred_gram(S_loc: torch.Tensor, K: int, scale: float, group=None, ops=_DeviceGramOps) -> torch.Tensor:
    """SURVEY 8e, "one huge anchor": the centred Gram matrix of dim_reduction.py:193-197 from row shards of S, formed on
    the smaller side, with PARTIAL Grams and one all-reduce instead of gathering the rows:
      K >= D  column sums all-reduced (D doubles) -> global mean; every rank adds up (s_k - mean)(s_k - mean)^T over ITS
              rows; one all-reduce of the D x D fp64 partials.
      K <  D  one all-to-all turns row shards into column panels; every rank forms the K x K row-side Gram of its panel
              (column means are local to a panel); one all-reduce of the K x K fp64 partials.
    S_loc [k_loc, D] fp32, this rank's rows shard_range(K, rank, world).  Returns G [n, n] fp64 on every rank."""
    import torch.distributed as dist
    D = S_loc.shape[1]
    if K >= D:
        mean = ops.column_sums(S_loc)
        dist.all_reduce(mean, group=group)
        mean /= K
        G = ops.gram_about(S_loc, mean, scale)
    else:
        G = ops.row_gram(rows_to_column_panels(S_loc, K, group), scale)
    dist.all_reduce(G, group=group)
    return G


def estimate_anchors(engine, anchors: torch.Tensor, K: int, sde, t: float, *, seed: int = 0,
                     z: Optional[torch.Tensor] = None, chunk: Optional[int] = None, anchor_id0: int = 0,
                     mem_budget_bytes: int = 24 << 30, gather_group=None,
                     n_total: Optional[int] = None, gemm_chunk: Optional[int] = None) -> Dict[str, torch.Tensor]:
    """anchors [N, D] (host or device) -> {'singular_values': [N, n] fp32, 'dims': [N] int32, 'sweeps'} on the
    HOST.  Host anchors are copied to the device chunk by chunk (pinned staging), results copied back once.
    With `gather_group` (anchors sharded over the ranks of a process group, `n_total` anchors in all) the spectra and
    IDs of ALL ranks are returned: one all-gather on the device before the copy back."""
    dev = engine.dev
    N, D = anchors.shape
    mc, std = sde_coefficients(sde, t)
    label = float((torch.ones(1, dtype=torch.float32) * t * (sde.N - 1))[0])
    if chunk is None:
        per = engine.bytes_per_anchor(K) if isinstance(engine, FcnEngine) else engine.bytes_per_anchor(K, D)
        if isinstance(engine, FcnEngine):
            mem_budget_bytes = max(mem_budget_bytes - engine.chain_bytes(K), per)
        if not isinstance(engine, (FcnEngine, KSphereEngine)):
            mem_budget_bytes = max(mem_budget_bytes, 96 << 30)  # image score matrices: size chunks for 180 GB HBM3e
        # the budgets above are caps: on a shared or partly occupied device the chunk shrinks to what is actually free
        # (60 % of it: the U-Net activation pool and the CUDA-graph private pools live beside the score matrices)
        free_bytes, _ = torch.cuda.mem_get_info(dev)
        mem_budget_bytes = min(mem_budget_bytes, int(free_bytes * 0.6))
        chunk = max(1, min(N, mem_budget_bytes // max(per, 1)))
    svs, dims, sweeps, finite = [], [], [], []
    for s in range(0, N, chunk):
        x0 = anchors[s:s + chunk]
        if not x0.is_cuda:
            x0 = x0.contiguous().float()
            x0 = (x0 if x0.is_pinned() else x0.pin_memory()).to(dev, non_blocking=True)
        zc = None if z is None else z[s:s + chunk].to(dev, non_blocking=True).contiguous()
        extra = {"chunk": gemm_chunk} if (gemm_chunk and isinstance(engine, FcnEngine)) else {}
        r: SpectrumResult = engine.estimate(x0.contiguous(), K, mc, std, label, z=zc, seed=seed,
                                            anchor_id0=anchor_id0 + s, **extra)
        svs.append(r.singular_values)
        dims.append(r.dims)
        sweeps.append(r.sweeps)
        if r.finite is not None:
            finite.append(r.finite)
    ok_dev = torch.cat(finite).all() if finite else None
    out = {"singular_values": torch.cat(svs), "dims": torch.cat(dims), "sweeps": torch.cat(sweeps)}
    if gather_group is not None:
        out["singular_values"], out["dims"] = all_gather_spectra(out["singular_values"], out["dims"],
                                                                 N if n_total is None else n_total, gather_group)
    host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True) for k, v in out.items()}
    for k in out:
        host[k].copy_(out[k], non_blocking=True)
    ok_host = None
    if ok_dev is not None:
        ok_host = torch.empty((), dtype=torch.bool, pin_memory=True)
        ok_host.copy_(ok_dev, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    if (ok_host is not None and not bool(ok_host)) or not bool(torch.isfinite(host["singular_values"]).all()):
        # the fp16 hi/lo operand split is unsaturated: |deviation| or |activation| * 64 beyond the fp16 range (> ~1023)
        # turns into inf - inf = NaN in the tensor-core chain; the bisection would return finite garbage for such a
        # Gram matrix and the read-out a full-rank ID, so the Gram diagonal is checked on the device
        raise _lib.IdbError("non-finite spectrum: the score matrix of at least one anchor holds NaN / inf -- an operand "
                            "left the fp16-pair range (|value| > ~1023 after scaling); normalise the data or run the "
                            "FCN engine with fmt=FMT_BF16")
    return host


def get_manifold_dimension(config, name=None, return_svd=False, *, score_model=None, loader=None, seed=None,
                           noise=None):
    """Reference signature `get_manifold_dimension(config, name=None, return_svd=False)`; the keyword-only extras
    let a caller pass live weights (the training callback, callbacks.py:403-432), an existing loader, a Philox
    seed, or recorded noise `noise[i]` = [K, D] per anchor (parity runs)."""
    save_path = os.path.join(config.logging.log_path, config.logging.log_name, "svd")
    Path(save_path).mkdir(parents=True, exist_ok=True)

    if loader is None:
        from .data import create_lightning_datamodule
        dm = create_lightning_datamodule(config)
        dm.setup()
        loader = dm.train_dataloader()
    sde, sampling_eps = sde_lib.configure_sde(config)
    device = torch.device(lookup(config, "device", "cuda"))
    if device.type != "cuda":
        raise _lib.IdbError("iddiff_b200 runs on CUDA devices only; the CPU path is the reference itself")
    engine = build_engine(config, score_model, device)

    num_datapoints = lookup(config, "dim_estimation.num_datapoints")
    if num_datapoints is None:
        num_datapoints = lookup(config, "logging.svd_points")
    if num_datapoints is None:  # the reference leaves the name unbound here (appendix A.8)
        raise NameError("name 'num_datapoints' is not defined")

    picked = gather_anchors(loader, num_datapoints)
    singular_values: List[List[float]] = [None] * len(picked)
    dims: List[int] = [None] * len(picked)
    # consecutive anchors that share a loader batch size share K: each run is one device batch
    for B, idxs in batch_size_runs([B for _, B in picked]):
        x = torch.stack([picked[i][0] for i in idxs]).float()
        if x.dim() > 2:  # image anchors [C, H, W]: rows of the score matrix are the NCHW flatten (dim_reduction.py:191)
            x = x.flatten(1)
        _, _, K = k_schedule(picked[idxs[0]][0].shape, B)
        z = None if noise is None else torch.stack([torch.as_tensor(noise[i]) for i in idxs]).float()
        res = estimate_anchors(engine, x, K, sde, sampling_eps, seed=0 if seed is None else seed, z=z,
                               anchor_id0=idxs[0])
        sv = res["singular_values"].double().tolist()  # fp32 values widened, like s.tolist() at dim_reduction.py:198
        for j, i in enumerate(idxs):
            singular_values[i] = sv[j]
            dims[i] = int(res["dims"][j])
    info = {"singular_values": singular_values}
    get_manifold_dimension.last_dims = dims  # device read-out (plot_utils.py:224-229), for callers that want it
    if return_svd:
        return info
    with open(os.path.join(save_path, f"{name}.pkl"), "wb") as f:
        pickle.dump(info, f)


def get_conditional_manifold_dimension(config, name=None, *, score_model=None, loader=None, seed=None, label=1,
                                       times=None, noise=None):
    """Drop-in for the reference's `get_conditional_manifold_dimension(config, name=None)` (dim_reduction.py:12-114):
    the same spectra for 12 diffusion times `linspace(sampling_eps, 0.3, 12)`, only anchors whose label equals 1
    (`:57-58`), from the VALIDATION loader with labels (`:19-22`), N-1 anchors (`:53-54,60-61`), one directory per
    time `svd/%.3f/` holding images.pkl, labels_svd.pkl, labels.pkl (`:101-114`).  All 12 time slices of an anchor
    group reuse one engine; per slice the whole group is one device batch.  The reference draws fresh noise per anchor
    and per time slice (`:79`): the Philox seed of slice i is `seed + i * 0x9E3779B1`; `noise[i][j]` = recorded [K, D]
    noise of anchor j in slice i (parity runs)."""
    import numpy as np
    log_path, log_name = config.logging.log_path, config.logging.log_name
    Path(os.path.join(log_path, log_name, "svd")).mkdir(parents=True, exist_ok=True)
    config.data.return_labels = True
    if loader is None:
        from .data import create_lightning_datamodule
        dm = create_lightning_datamodule(config)
        dm.setup()
        loader = dm.val_dataloader()
    loader = list(loader)
    sde, sampling_eps = sde_lib.configure_sde(config)
    device = torch.device(lookup(config, "device", "cuda"))
    if device.type != "cuda":
        raise _lib.IdbError("iddiff_b200 runs on CUDA devices only; the CPU path is the reference itself")
    engine = build_engine(config, score_model, device)
    num_datapoints = lookup(config, "dim_estimation.num_datapoints", 26)
    if times is None:
        times = torch.linspace(sampling_eps, 0.3, 12)

    # the anchors every time slice visits (the reference re-walks the loader per slice; the selection is the same)
    picked, idx = [], 0
    if loader and not (isinstance(loader[0], (tuple, list)) and len(loader[0]) == 2):
        raise _lib.IdbError("get_conditional_manifold_dimension needs a validation loader of (batch, labels) pairs "
                            "(config.data.return_labels, dim_reduction.py:19-22,48); this datamodule has no labels")
    for orig_batch, orig_labels in loader:
        batchsize = orig_batch.size(0)
        if idx + 1 >= num_datapoints:
            break
        for x, y in zip(orig_batch, orig_labels):
            if y.item() != label:
                continue
            if idx + 1 >= num_datapoints:
                break
            picked.append((x, batchsize, y.item()))
            idx += 1
    out = {}
    for ti, t_slice in enumerate(times):
        t = float(t_slice)
        t_save_path = os.path.join(log_path, log_name, "svd", "%.3f" % t)
        Path(t_save_path).mkdir(parents=True, exist_ok=True)
        singular_values = [None] * len(picked)
        for B, idxs in batch_size_runs([p[1] for p in picked]):
            x = torch.stack([picked[i][0] for i in idxs]).float()
            _, _, K = k_schedule(picked[idxs[0]][0].shape, B)
            z = None if noise is None else torch.stack([torch.as_tensor(noise[ti][i]) for i in idxs]).float()
            res = estimate_anchors(engine, x.flatten(1), K, sde, t, z=z,
                                   seed=((0 if seed is None else seed) + ti * 0x9E3779B1) & 0x7FFFFFFFFFFFFFFF,
                                   anchor_id0=idxs[0])
            sv = res["singular_values"].double().tolist()
            for j, i in enumerate(idxs):
                singular_values[i] = sv[j]
        imgs = [p[0].permute(1, 2, 0) if p[0].dim() == 3 else p[0] for p in picked]
        imgs = torch.stack(imgs).numpy() if imgs else np.zeros((0,))
        with open(os.path.join(t_save_path, "images.pkl"), "wb") as f:
            pickle.dump({"images": imgs}, f)
        with open(os.path.join(t_save_path, "labels_svd.pkl"), "wb") as f:
            pickle.dump({"singular_values": singular_values}, f)
        with open(os.path.join(t_save_path, "labels.pkl"), "wb") as f:
            pickle.dump({"labels": [p[2] for p in picked]}, f)
        out["%.3f" % t] = singular_values
    get_conditional_manifold_dimension.last = out
