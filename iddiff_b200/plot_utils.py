"""The part of the reference's plot_utils that turns spectra into intrinsic dimensions, without matplotlib
(plot_utils.py:158-238 in the reference).  Two implementations of the same rule:
  * `dims_on_device(svd)` -- the CUDA read-out kernel (idb_id_from_spectrum);
  * `dims_from_spectra(svd)` -- the reference's own few lines of Python-float arithmetic, for consumers of a
    stored svd pickle on a machine without a GPU (this is host bookkeeping on n floats, not a fallback of
    the hot path: the hot path's read-out is the kernel).
"""
from __future__ import annotations

import numpy as np


def softmax(x):
    """plot_utils.py:211-214."""
    e_x = np.exp(x - np.max(x))
    return e_x / e_x.sum(axis=0)


def extract_sing_vals(svd, mode="first"):
    """plot_utils.py:197-205: 'first' | 'all' | 'mean' aggregation of the per-anchor spectra."""
    sv = svd["singular_values"]
    if mode == "first":
        out = [sv[0]]
    elif mode == "all":
        out = sv
    elif mode == "mean":
        out = [np.mean(sv, axis=0)]
    else:
        raise ValueError(mode)
    return out


def dim_from_spectrum(s, tail=None):
    """plot_utils.py:173-182 / :226-229: normalised consecutive gaps (the first gap is skipped), softmax, then -- as
    plot_distribution(tail=) does at :178-179 -- only the LAST `tail` entries of the SOFTMAX are kept before the argmax
    (the spectrum itself is never truncated), dim = len(soft) - argmax."""
    norm_factor = s[1] - s[2]
    diff = [(s[i] - s[i + 1]) / norm_factor for i in range(1, len(s) - 1)]
    soft = softmax(np.asarray(diff))
    if tail:
        soft = soft[-tail:]
    return int(len(soft) - soft.argmax())


def dims_from_spectra(svd, mode="all", tail=None):
    """What plot_dims / plot_distribution return as `dims` (plot_utils.py:158-195,207-238)."""
    return [dim_from_spectrum(list(s), tail) for s in extract_sing_vals(svd, mode)]


def plot_distribution(svd, mode="first", return_tensor=False, tail=None):
    """The reference's signature (plot_utils.py:158); the figure is not drawn (no matplotlib on this path): returns dims."""
    return dims_from_spectra(svd, mode, tail)


def plot_dims(svd, title="Histogram of dimensions", tick_step=2, tick_start=1):
    """plot_utils.py:207-238 returns (figure, dims); the figure is None here."""
    return None, dims_from_spectra(svd, "all")


def dims_on_device(svd, device="cuda"):
    import torch
    from ._lib import call, ptr, stream_ptr
    sv = torch.tensor(svd["singular_values"], dtype=torch.float32, device=device).contiguous()
    A, n = sv.shape
    dims = torch.empty(A, dtype=torch.int32, device=device)
    call("idb_id_from_spectrum", ptr(sv), A, n, n, ptr(dims), stream_ptr())
    out = dims.cpu().tolist()
    if any(d == -1 for d in out):
        raise ZeroDivisionError("float division by zero")  # what the reference's list comprehension raises
    return out
