"""Drop-in replacement for the reference's `diff_gaussian_rasterization` Python package
(DGR/diff_gaussian_rasterization/__init__.py:17-196), backed by libgsb200.so (hand-written sm_100a
kernels behind a C ABI) instead of the pybind module `_C`.

Same names, argument order, defaults, exceptions and return values:
    GaussianRasterizationSettings(image_height, image_width, tanfovx, tanfovy, bg, scale_modifier,
                                  viewmatrix, projmatrix, sh_degree, campos, prefiltered)
    GaussianRasterizer(raster_settings)(means3D, means2D, opacities, shs=None, colors_precomp=None,
                                        scales=None, rotations=None, cov3D_precomp=None)
        -> (color [3,H,W], radii [P] int32, depth [1,H,W]), differentiable w.r.t. means3D, means2D,
           colors_precomp, opacities, scales, rotations, cov3D_precomp (depth and radii carry no gradient,
           DGR/README.md:13).
The spherical-harmonics colour path is dead code in the reference repository (every caller passes
colors_precomp with sh_degree = 0: shared.py:32,121) and is not provided: passing `shs` raises.
"""
import os
import sys
from typing import NamedTuple

import torch
import torch.nn as nn

_pkg_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _pkg_root not in sys.path:
    sys.path.insert(0, _pkg_root)

from gsb200 import raster as _raster  # noqa: E402


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        if sh is not None and sh.numel() != 0:
            raise NotImplementedError("spherical-harmonics colours are not provided by the B200 path; "
                                      "pass colors_precomp (the only mode the reference repository uses)")
        rs = raster_settings
        color, depth, radii, state = _raster.forward(
            means3D, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs.bg, rs.scale_modifier,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, rs.prefiltered,
            reuse_geometry=True)
        # Like the reference (DGR/py:82-84) the inputs go through save_for_backward, so autograd's version check catches
        # an in-place change between forward and backward (the kernels read them again through raw pointers), and the
        # state lives as long as the graph does: backward may run more than once (retain_graph=True).
        ctx.save_for_backward(means3D, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
        ctx.state = state
        ctx.num_rendered = state.num_rendered
        ctx.has_cov = cov3Ds_precomp is not None and cov3Ds_precomp.numel() != 0
        ctx.has_sr = scales is not None and scales.numel() != 0
        ctx.mark_non_differentiable(radii)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, _grad_depth):
        _ = ctx.saved_tensors   # raises if an input was modified in place since the forward
        g = _raster.backward(ctx.state, grad_out_color)
        empty = None
        grads = (
            g["means3D"],
            g["means2D"],
            empty,                                   # sh
            g["colors"],
            g["opacities"],
            g["scales"] if ctx.has_sr else empty,
            g["rotations"] if ctx.has_sr else empty,
            g["cov3D"] if ctx.has_cov else empty,
            None,
        )
        return grads


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _raster.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, rs)
