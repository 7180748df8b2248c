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

    def _eval_loader(self, data, section, shuffle):
        v = getattr(self.config, section, None)
        B = getattr(v, "batch_size", None) or self.config.training.batch_size
        return DataLoader(data, batch_size=B, num_workers=getattr(v, "workers", 0) or 0, shuffle=shuffle)

    def val_dataloader(self):   # KSphereDataset.py:131-132 (shuffle=True there too)
        return self._eval_loader(self.valid_data, "validation", True)

    def test_dataloader(self):  # KSphereDataset.py:134-135
        return self._eval_loader(self.test_data, "eval", False)


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


# ---------------------------------------------------------------------------------------------------------------
# The paper's synthetic image manifolds (configs/dimension_estimation/paper/image_data/{squares,gaussian_blobs}):
# `data.datamodule = 'Synthetic'`, `data.dataset_type = 'FixedSquaresManifold' | 'FixedGaussiansManifold'`
# (lightning_data_modules/SyntheticDataset.py:81-183,388-431).  Both are driven by Python's `random` seeded with
# config.seed; the draws below happen in the reference's order (geometry first, then per image one draw per shape), so
# the same seed gives the same images.  Generation is vectorised over the images instead of the reference's per-pixel
# Python loops; accumulation stays in fp32 in the shapes' order, which is what the element-wise `+=` there does.
def fixed_squares_images(seed, num_samples, num_squares, square_range, img_size):
    """FixedSquaresManifold: `num_squares` squares at fixed places, each image paints every square with its own uniform
    intensity (overlaps add) -> [num_samples, 1, S, S] fp32; the manifold dimension is num_squares."""
    import random
    random.seed(seed)
    info = []
    for _ in range(num_squares):
        side = random.choice(square_range)
        start, finish = (side + 1) // 2, img_size - (side + 1) // 2
        x = random.choice(np.arange(start, finish))
        y = random.choice(np.arange(start, finish))
        info.append((int(x), int(y), int(side)))
    c = torch.tensor([[random.random() for _ in range(num_squares)] for _ in range(num_samples)], dtype=torch.float64)
    c = c.to(torch.float32)                     # `img[...] += c` rounds the Python float to the image's fp32
    img = torch.zeros(num_samples, img_size, img_size)
    for i, (x, y, side) in enumerate(info):
        x0, y0 = x - ((side + 1) // 2 - 1), y - ((side + 1) // 2 - 1)
        img[:, x0:x0 + side, y0:y0 + side] += c[:, i, None, None]
    return img.unsqueeze(1)


def fixed_gaussians_images(seed, num_samples, num_gaussians, std_range, img_size):
    """FixedGaussiansManifold: `num_gaussians` bumps at fixed centres (sampled without replacement), each image draws
    every bump's width uniformly from std_range; the sum is rescaled to [0, 1] per image."""
    import random
    random.seed(seed)
    pairs = [[i, j] for i in range(img_size) for j in range(img_size)]
    centers = random.sample(pairs, k=num_gaussians)
    xx, yy = torch.meshgrid((torch.arange(img_size), torch.arange(img_size)), indexing="ij")
    data = []
    for _ in range(num_samples):
        img = torch.zeros(img_size, img_size)
        for cx, cy in centers:
            std = random.uniform(std_range[0], std_range[1])
            c = 1 / (np.sqrt(2 * np.pi) * std)
            d = -1 / (2 * std ** 2)
            arg = d * ((xx - cx) ** 2 + (yy - cy) ** 2)                  # int64 * Python float -> fp32 tensor
            bump = torch.from_numpy(np.exp(arg.numpy()))                 # NumPy's fp32 exp, as np.exp(tensor) there
            bump *= c
            img += bump
        lo, hi = torch.min(img), torch.max(img)
        img -= lo
        img /= hi - lo
        data.append(img.to(torch.float32).unsqueeze(0))
    return torch.stack(data)


@register_lightning_datamodule("Synthetic")
class SyntheticDataModule:
    def __init__(self, config):
        self.config = config

    def setup(self, stage=None):
        d = self.config.data
        kind = d.dataset_type
        seed = self.config.seed
        if kind == "FixedSquaresManifold":
            self.data = fixed_squares_images(seed, d.data_samples, d.num_squares, list(d.square_range), d.image_size)
        elif kind == "FixedGaussiansManifold":
            self.data = fixed_gaussians_images(seed, d.data_samples, d.num_gaussians, list(d.std_range), d.image_size)
        else:
            raise NotImplementedError(f"Synthetic dataset_type {kind!r}: the dimension-estimation configs use "
                                      "FixedSquaresManifold and FixedGaussiansManifold")
        n = len(self.data)
        split = list(d.split)
        lens = [int(split[0] * n), int(split[1] * n), int(split[2] * n)]
        self.train_data, self.valid_data, self.test_data = random_split(self.data, lens)

    def train_dataloader(self):
        t = self.config.training
        return DataLoader(self.train_data, batch_size=t.batch_size, num_workers=getattr(t, "workers", 0), shuffle=True)

    def val_dataloader(self):
        v = getattr(self.config, "validation", None)
        B = getattr(v, "batch_size", None) or self.config.training.batch_size
        return DataLoader(self.valid_data, batch_size=B, num_workers=getattr(v, "workers", 0) or 0, shuffle=False)
