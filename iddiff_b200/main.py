"""CLI with the reference's flags for this path (main.py:17-26,32-43,70-71):
    python -m iddiff_b200.main --config <file.py|file.pkl> --mode manifold_dimension|conditional_manifold_dimension [--checkpoint_path P] [--log_path L]
"""
import argparse
import importlib.util
import os
import pickle
import sys
import types

from . import run_lib
from .config import ConfigDict


def _ensure_ml_collections():
    """The reference's config files start with `import ml_collections` (e.g. configs/dimension_estimation/paper/
    euclidean_data/ksphere/50dim.py:1,8).  Where the package is not installed, a module of that name exposing this
    package's ConfigDict stand-in lets those files load unmodified."""
    try:
        import ml_collections  # noqa: F401
        return
    except ImportError:
        pass
    mod = types.ModuleType("ml_collections")
    mod.ConfigDict = ConfigDict
    sub = types.ModuleType("ml_collections.config_dict")
    sub.ConfigDict = ConfigDict
    mod.config_dict = sub
    sys.modules["ml_collections"] = mod
    sys.modules["ml_collections.config_dict"] = sub


def read_config(path):
    """configs/utils.py:6-10: import the file and call get_config(); .pkl = pickled ConfigDict (main.py:32-34)."""
    if path.endswith(".pkl"):
        with open(path, "rb") as f:
            return pickle.load(f)
    _ensure_ml_collections()
    # configs that build on others import them as `configs.<...>` (style_gan_2d.py:7): make the tree that holds the
    # file's top-level `configs` package importable while the file is executed
    root, added = os.path.dirname(os.path.abspath(path)), None
    while root != os.path.dirname(root):
        if os.path.basename(root) == "configs":
            added = os.path.dirname(root)
            break
        root = os.path.dirname(root)
    cached = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "configs" or k.startswith("configs.")}
    if added is not None:
        sys.path.insert(0, added)
    try:
        spec = importlib.util.spec_from_file_location("idb_user_config", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod.get_config()
    finally:
        if added is not None and added in sys.path:
            sys.path.remove(added)
        for k in [k for k in sys.modules if k == "configs" or k.startswith("configs.")]:
            del sys.modules[k]
        sys.modules.update(cached)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True)
    ap.add_argument("--mode", required=True, choices=["manifold_dimension", "conditional_manifold_dimension"])
    ap.add_argument("--checkpoint_path", default=None)
    ap.add_argument("--log_path", default=None)
    a = ap.parse_args(argv)
    config = read_config(a.config)
    if a.checkpoint_path is not None:
        config.model.checkpoint_path = a.checkpoint_path
    if a.log_path is not None:
        config.logging.log_path = a.log_path
    if a.mode == "conditional_manifold_dimension":   # main.py:72-73
        run_lib.get_conditional_manifold_dimension(config)
    else:
        run_lib.get_manifold_dimension(config)


if __name__ == "__main__":
    main()
