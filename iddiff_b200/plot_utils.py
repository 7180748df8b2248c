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


def extract_sing_vals(svd, mode="first", tail=None):
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
    if tail is not None:
        out = [s[-tail:] for s in out]
    return out


def dim_from_spectrum(s):
    """plot_utils.py:226-229: normalised consecutive gaps (the first gap is skipped), softmax, argmax."""
    norm_factor = s[1] - s[2]
    diff = [(s[i] - s[i + 1]) / norm_factor for i in range(1, len(s) - 1)]
    soft = softmax(np.asarray(diff))
    return int(len(soft) - soft.argmax())


def dims_from_spectra(svd, mode="all"):
    """What plot_dims / plot_distribution return as `dims` (plot_utils.py:158-195,207-238)."""
    return [dim_from_spectrum(list(s)) for s in extract_sing_vals(svd, mode)]


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
