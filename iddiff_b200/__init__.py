"""iddiff_b200 -- B200-native `manifold_dimension` hot path of GBATZOLIS/ID-diff.

The package holds only what that path needs: csrc/ (sm_100a CUDA kernels + the C ABI of
include/iddiff_b200.h), lib/ (the built libiddiff_b200.so) and the Python host code that mirrors the
reference's interface for the path (dim_reduction.get_manifold_dimension, run_lib, sde_lib, the gap
read-out of plot_utils).  There is no CPU fallback: importing is cheap, but every compute call
raises if the CUDA library is missing.
"""
__version__ = "0.1.0"
