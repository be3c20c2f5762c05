"""hugs_b200 -- B200-native splat hot path for HUGS (host side).

Importing this package loads libhugs_b200.so; it raises if the CUDA library has not been built.
"""
from . import _native
from .rasterizer import (
    GaussianRasterizationSettings,
    GaussianRasterizer,
    rasterize_gaussians,
    rasterize_gaussians_forward,
    rasterize_gaussians_backward,
    mark_visible,
    ShGradSink,
    attach_sh_grad_sink,
    detach_sh_grad_sink,
    GradSink,
    attach_grad_sink,
    detach_grad_sink,
)
from .lbs import lbs_extra, deform_gaussians, lbs_deform_frames, matrix_to_quaternion

__all__ = [
    "GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
    "rasterize_gaussians_forward", "rasterize_gaussians_backward", "mark_visible",
    "ShGradSink", "attach_sh_grad_sink", "detach_sh_grad_sink", "GradSink", "attach_grad_sink",
    "detach_grad_sink",
    "lbs_extra", "deform_gaussians", "lbs_deform_frames", "matrix_to_quaternion",
]
