"""run_lib.get_manifold_dimension wrapper (run_lib.py:324-325 in the reference)."""
from . import dim_reduction


def get_manifold_dimension(config, name=None):
    dim_reduction.get_manifold_dimension(config, name)


def get_conditional_manifold_dimension(config, name=None):
    """run_lib.py:327-328 in the reference."""
    dim_reduction.get_conditional_manifold_dimension(config, name)
