"""trase_b200 — B200-native (sm_100a) differentiable Gaussian rasterizer for TRASE.

Drop-in for the reference's ``diff_gaussian_rasterization`` package
(reference: submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py):

    from trase_b200.diff_gaussian_rasterization import (
        GaussianRasterizationSettings, GaussianRasterizer)

The compute path is hand-written CUDA behind a C ABI (include/trase_b200.h,
trase_b200/libtrase_b200.so).  There is no CPU fallback: importing the native
library fails loudly if it has not been built (python -m trase_b200.build).
"""
__version__ = "0.1.0"
