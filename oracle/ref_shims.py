"""TEST INFRASTRUCTURE ONLY -- shims that let the *real* reference (/root/reference) be
imported in the build container, so that golden vectors can be generated from the
reference's own code (oracle/gen_golden.py) and the numpy/torch restatement in
oracle/id_oracle.py can be pinned against it.

Nothing in the product package imports this file.  /root/reference does not exist on
the GPU box; only the committed fixtures under tests/golden/ travel.

The three shims are exactly the ones listed in SURVEY.md section 8(c):
  * `pytorch_lightning` stub (LightningModule = nn.Module + .device/.dtype/.log/...)
  * `ml_collections` stub  (ConfigDict with attribute access and dotted-key lookup)
  * numpy-2 compat: `torch.from_numpy` passes Tensors through
    (KSphereDataset.py:42-43 calls np.linalg.qr on a Tensor, which returns Tensors
    under numpy 2.x / torch 2.11).
"""
import sys
import types

import torch
import torch.nn as nn

import os

# the reference itself in the build container; on the GPU box (no /root/reference) the copy made by oracle/build_ref.sh
REFERENCE_ROOT = "/root/reference" if os.path.isdir("/root/reference") else \
    os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "dim_reduction.py"))


class _ConfigDict(dict):
    """Minimal ml_collections.ConfigDict: attribute access + dotted-key __getitem__."""

    def __init__(self, initial=None, **kw):
        super().__init__()
        if initial:
            for k, v in dict(initial).items():
                self[k] = v
        for k, v in kw.items():
            self[k] = v

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name)

    def __setattr__(self, name, value):
        self[name] = value

    def __getitem__(self, key):
        if isinstance(key, str) and "." in key and not dict.__contains__(self, key):
            head, rest = key.split(".", 1)
            return dict.__getitem__(self, head)[rest]
        return dict.__getitem__(self, key)

    def get(self, key, default=None):
        try:
            return self[key]
        except (KeyError, TypeError):
            return default


class _LightningModule(nn.Module):
    def __init__(self, *a, **k):
        super().__init__()

    @property
    def device(self):
        try:
            return next(self.parameters()).device
        except StopIteration:
            return torch.device("cpu")

    @property
    def dtype(self):
        try:
            return next(self.parameters()).dtype
        except StopIteration:
            return torch.float32

    def save_hyperparameters(self, *a, **k):
        pass

    def log(self, *a, **k):
        pass

    def load_from_checkpoint(self, path=None, *a, **k):
        # dim_reduction.py:128 calls this on an *instance*; in-memory weights are kept when
        # path is None, otherwise a Lightning-style {'state_dict': ...} file is loaded.
        if path is not None:
            ckpt = torch.load(path, map_location="cpu", weights_only=False)
            self.load_state_dict(ckpt.get("state_dict", ckpt))
        return self


class _LightningDataModule:
    def __init__(self, *a, **k):
        pass


class _Callback:
    pass


def install():
    """Install the shims into sys.modules and put the reference on sys.path."""
    if "pytorch_lightning" not in sys.modules:
        pl = types.ModuleType("pytorch_lightning")
        pl.LightningModule = _LightningModule
        pl.LightningDataModule = _LightningDataModule
        pl.Callback = _Callback
        cb = types.ModuleType("pytorch_lightning.callbacks")
        cb.Callback = _Callback
        pl.callbacks = cb
        sys.modules["pytorch_lightning"] = pl
        sys.modules["pytorch_lightning.callbacks"] = cb
    if "ml_collections" not in sys.modules:
        mc = types.ModuleType("ml_collections")
        mc.ConfigDict = _ConfigDict
        sys.modules["ml_collections"] = mc
    if not getattr(torch.from_numpy, "_idb_patched", False):
        _orig = torch.from_numpy

        def _from_numpy(a):
            return a if isinstance(a, torch.Tensor) else _orig(a)

        _from_numpy._idb_patched = True
        torch.from_numpy = _from_numpy
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


ConfigDict = _ConfigDict
