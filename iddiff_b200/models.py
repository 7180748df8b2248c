"""Weight holders for the score networks of the path, registered under the reference's model names
(models/utils.py:27-47 registry; 'fcn' = models/fcn.py:6-30, 'ksphere_gt' = models/ksphere_gt.py:7-12).
They only OWN parameters in the reference's state_dict layout (so Lightning checkpoints load); the
forward pass lives in the CUDA engine -- calling them on tensors raises."""
from __future__ import annotations

import torch
import torch.nn as nn

_MODELS = {}


def register_model(name):
    def deco(cls):
        if name in _MODELS:
            raise ValueError(f"Already registered model with name: {name}")
        _MODELS[name] = cls
        return cls
    return deco


def get_model(name):
    return _MODELS[name]


def create_model(config):
    return get_model(config.model.name)(config)


@register_model("fcn")
class FCN(nn.Module):
    def __init__(self, config):
        super().__init__()
        m = config.model
        layers = [nn.Linear(m.state_size + 1, m.hidden_nodes), nn.Dropout(m.dropout), nn.ELU()]
        for _ in range(m.hidden_layers):
            layers += [nn.Linear(m.hidden_nodes, m.hidden_nodes), nn.Dropout(m.dropout), nn.ELU()]
        layers.append(nn.Linear(m.hidden_nodes, m.state_size))
        self.mlp = nn.Sequential(*layers)  # same indices as the reference: Linear at mlp.0, mlp.3, ...

    def forward(self, x, t):
        raise RuntimeError("iddiff_b200.models.FCN holds weights only; evaluate it through engine.FcnEngine "
                           "(there is no PyTorch/CPU path for this network here)")


@register_model("ksphere_gt")
class KSphereGT(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.manifold_dim = config.data.manifold_dim
        self.dummy = nn.Linear(1, 1)

    def forward(self, batch, sigmas):
        raise RuntimeError("iddiff_b200.models.KSphereGT is evaluated by engine.KSphereEngine only")


def load_checkpoint_into(model: nn.Module, path: str):
    """Lightning checkpoint layout (BaseSdeGenerativeModel.py:17,21): state_dict keys prefixed 'score_model.'."""
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    state = ckpt.get("state_dict", ckpt)
    own = {k[len("score_model."):]: v for k, v in state.items() if k.startswith("score_model.")}
    model.load_state_dict(own if own else state)
    return model
