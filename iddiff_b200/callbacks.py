"""Training-time caller of the path: the reference's `ScoreSpectrumVisualization` callback
(lightning_callbacks/callbacks.py:403-432) re-loads `checkpoints/best/last.ckpt` from disk every `svd_frequency`
epochs and runs get_manifold_dimension on it.  This mirror keeps the hook name, the config keys
(`logging.svd_frequency`, `logging.save_svd`, `logging.svd_points`), the pickle name `svd_{epoch}.pkl` and the logged
scalar 'dim' = mean of plot_distribution's dims -- but hands the LIVE `pl_module.score_model` weights to the CUDA
engine (keyword `score_model=`), so there is no checkpoint round trip.  The matplotlib images the reference also
logs are outside the hot path.  No pytorch_lightning import: any object with this hook works
as a Lightning callback (duck typing), and `register_callback` mirrors lightning_callbacks/utils.py."""
from __future__ import annotations

import logging
import os
import pickle

from . import dim_reduction, plot_utils

_CALLBACKS = {}


def register_callback(name):
    def deco(cls):
        if name in _CALLBACKS:
            raise ValueError(f"Already registered callback with name: {name}")
        _CALLBACKS[name] = cls
        return cls
    return deco


def get_callback_by_name(name):
    return _CALLBACKS[name]


@register_callback("ScoreSpectrumVisualization")
class ScoreSpectrumVisualization:
    def __init__(self, show_evolution=False):
        self.evolution = False
        self.last_dims = None

    def on_validation_epoch_end(self, trainer, pl_module):
        config = pl_module.config
        if (pl_module.current_epoch + 1) % config.logging.svd_frequency != 0:
            return
        name = f"svd_{pl_module.current_epoch}"
        try:
            if getattr(config.logging, "save_svd", False):
                dim_reduction.get_manifold_dimension(config=config, name=name, return_svd=False,
                                                     score_model=pl_module.score_model)
                path = os.path.join(config.logging.log_path, config.logging.log_name, "svd", f"{name}.pkl")
                with open(path, "rb") as f:
                    svd = pickle.load(f)
            else:
                svd = dim_reduction.get_manifold_dimension(config=config, name=name, return_svd=True,
                                                           score_model=pl_module.score_model)
            dims = plot_utils.dims_from_spectra(svd)           # plot_distribution(svd, mode='all') -> dims
            self.last_dims = dims
            pl_module.log("dim", sum(dims) / max(len(dims), 1), on_step=False, on_epoch=True, prog_bar=True,
                          logger=True)
        except Exception as e:  # the reference swallows failures the same way (callbacks.py:429-431)
            logging.warning("Could not create a score spectrum")
            logging.error(e)
