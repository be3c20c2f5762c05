"""Drop-in replacement for the reference's ``diff_gaussian_rasterization`` package
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py).

Put ``ml-hugs_b200/`` on ``sys.path`` ahead of the reference install and
``hugs/renderer/gs_renderer.py:11-14`` imports this instead, unchanged.
"""
from hugs_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    rasterize_gaussians,
    _RasterizeGaussians,
    cpu_deep_copy_tuple,
    _C,
)
