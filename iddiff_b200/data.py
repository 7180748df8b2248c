"""Anchor sources (host side, plain PyTorch -- inputs, not kernels), registered under the reference's
datamodule names (lightning_data_modules/utils.py:5-30).  'KSphere' follows
lightning_data_modules/KSphereDataset.py:21-95,106-135; 'Tensor' serves caller-supplied anchors."""
from __future__ import annotations

import numpy as np
import torch
from torch.utils.data import DataLoader, random_split

_DATAMODULES = {}


def register_lightning_datamodule(name):
    def deco(cls):
        _DATAMODULES[name] = cls
        return cls
    return deco


def create_lightning_datamodule(config):
    return _DATAMODULES[config.data.datamodule](config)


def _embed(new, i, n_samples, ambient_dim, md, n_spheres, kind):
    if kind == "random_isometry":
        gen = torch.Generator().manual_seed(0)  # hard-coded seed, KSphereDataset.py:40
        q, _ = np.linalg.qr(torch.randn(size=(ambient_dim, md + 1), generator=gen).numpy())
        return (torch.from_numpy(q) @ new.T).T
    pad = lambda t: torch.cat([t, torch.zeros([n_samples, ambient_dim - t.shape[1]])], dim=1)
    if kind == "first":
        return pad(new)
    if kind == "separating":
        if n_spheres * (md + 1) > ambient_dim:
            raise RuntimeError("Cant fit that many spheres.")
        return pad(torch.cat([torch.zeros((n_samples, i * (md + 1))), new], dim=1))
    if kind == "along_axis":
        if (n_spheres - 1) + (md + 1) > ambient_dim:
            raise RuntimeError("Cant fit that many spheres.")
        return pad(torch.cat([torch.zeros((n_samples, i)), new], dim=1))
    raise RuntimeError("Unknown embedding type.")


def ksphere_points(n_samples, n_spheres, ambient_dim, manifold_dim, noise_std, embedding_type, radii=(),
                   angle_std=-1):
    """Union of k-spheres drawn from the torch global generator in the reference's order of draws."""
    radii = list(radii) if len(radii) else [1] * n_spheres
    mdims = [manifold_dim] * n_spheres if isinstance(manifold_dim, int) else list(manifold_dim)
    out = []
    for i in range(n_spheres):
        md = mdims[i]
        if angle_std == -1:
            pts = torch.randn((n_samples, md + 1))
            pts = pts / torch.linalg.norm(pts, dim=1)[:, None]
        else:  # polar angles ~ N(0, angle_std): non-uniform density (KSphereDataset.py:66-80,92-94)
            ang = angle_std * torch.randn((n_samples, md))
            sin_prod = torch.cumprod(torch.sin(ang), dim=1)
            lead = torch.cat([torch.ones(n_samples, 1), sin_prod[:, :-1]], dim=1) * torch.cos(ang)
            pts = torch.cat([lead, sin_prod[:, -1:]], dim=1)
        pts = _embed(pts * radii[i], i, n_samples, ambient_dim, md, n_spheres, embedding_type)
        out.append(pts + noise_std * torch.randn_like(pts))
    return torch.cat(out, dim=0)


@register_lightning_datamodule("KSphere")
class KSphereDataModule:
    def __init__(self, config):
        self.config = config

    def setup(self, stage=None):
        d = self.config.data
        g = lambda k, dflt=None: d.get(k, dflt) if hasattr(d, "get") else getattr(d, k, dflt)
        self.data = ksphere_points(g("data_samples"), g("n_spheres"), g("ambient_dim"), g("manifold_dim"),
                                   g("noise_std"), g("embedding_type"), g("radii", []), g("angle_std", -1))
        n = len(self.data)
        split = g("split", [1.0, 0.0, 0.0])
        lens = [int(split[0] * n), int(split[1] * n), int(split[2] * n)]
        self.train_data, self.valid_data, self.test_data = random_split(self.data, lens)

    def train_dataloader(self):
        t = self.config.training
        return DataLoader(self.train_data, batch_size=t.batch_size, num_workers=getattr(t, "workers", 0),
                          shuffle=True)


@register_lightning_datamodule("Tensor")
class TensorDataModule:
    """Anchors supplied by the caller as config.data.anchors ([N, *shape] tensor); batches in order."""

    def __init__(self, config):
        self.config = config

    def setup(self, stage=None):
        self.data = torch.as_tensor(self.config.data.anchors, dtype=torch.float32)

    def train_dataloader(self):
        B = self.config.training.batch_size
        return [self.data[i:i + B] for i in range(0, len(self.data), B)]

    def val_dataloader(self):
        """(batch, labels) pairs when config.data.return_labels is set (dim_reduction.py:19-22 of the reference reads
        the VALIDATION loader with labels for the conditional variant); labels come from config.data.labels."""
        B = getattr(getattr(self.config, "validation", None), "batch_size", None) or self.config.training.batch_size
        if not getattr(self.config.data, "return_labels", False):
            return [self.data[i:i + B] for i in range(0, len(self.data), B)]
        labels = torch.as_tensor(self.config.data.labels)
        return [(self.data[i:i + B], labels[i:i + B]) for i in range(0, len(self.data), B)]
