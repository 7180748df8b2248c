"""run_lib.get_manifold_dimension wrapper (run_lib.py:324-325 in the reference)."""
from . import dim_reduction


def get_manifold_dimension(config, name=None):
    dim_reduction.get_manifold_dimension(config, name)
