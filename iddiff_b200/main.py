"""CLI with the reference's flags for this path (main.py:17-26,32-43,70-71):
    python -m iddiff_b200.main --config <file.py|file.pkl> --mode manifold_dimension|conditional_manifold_dimension [--checkpoint_path P] [--log_path L]
"""
import argparse
import importlib.util
import pickle

from . import run_lib


def read_config(path):
    """configs/utils.py:6-10: import the file and call get_config(); .pkl = pickled ConfigDict (main.py:32-34)."""
    if path.endswith(".pkl"):
        with open(path, "rb") as f:
            return pickle.load(f)
    spec = importlib.util.spec_from_file_location("idb_user_config", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.get_config()


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
