"""Camera-batched training step on the hot path (the additional batched entry point of SURVEY.md
section 7/8b): all cameras of a timestep against one replicated Gaussian set --

    raster forward (n_cams) -> fused L1+SSIM loss + dL/dimage -> raster backward (gradients summed over
    cameras) -> densification statistics per camera (densify.py:154-162 + external.py:113-124)

with no per-camera Python work except the statistics launch.  Parity is per render call
(tests/test_raster_gpu.py::test_camera_batch_matches_single).
"""
import ctypes as C

import torch

from . import _lib as L
from . import loss as loss_mod
from . import raster


class DensifyStats:
    """DensificationVariables (shared.py:20-26): visibility_count, mean_2d_gradients_accumulated, max_2d_radii."""

    def __init__(self, P, device):
        z = lambda: torch.zeros(P, dtype=torch.float32, device=device)
        self.visibility_count, self.mean_2d_gradients_accumulated, self.max_2d_radii = z(), z(), z()
        self.gaussian_is_visible_mask = torch.zeros(P, dtype=torch.bool, device=device)

    def accumulate(self, radii, grad_means2D):
        P = radii.numel()
        with torch.cuda.device(radii.device):
            rc = L.lib().gsb_densify_accumulate(P, L.ptr(radii), L.ptr(grad_means2D), L.ptr(self.max_2d_radii),
                                                L.ptr(self.mean_2d_gradients_accumulated),
                                                L.ptr(self.visibility_count), L.ptr(self.gaussian_is_visible_mask),
                                                L.stream_ptr())
        L.check(rc, "gsb_densify_accumulate")

    def accumulate_batch(self, radii, grad_means2D):
        """radii [n,P], grad_means2D [n,P,3]: all cameras of the batch in one launch."""
        n, P = radii.shape
        with torch.cuda.device(radii.device):
            rc = L.lib().gsb_densify_accumulate_batch(P, n, L.ptr(radii), L.ptr(grad_means2D), L.ptr(self.max_2d_radii),
                                                      L.ptr(self.mean_2d_gradients_accumulated),
                                                      L.ptr(self.visibility_count),
                                                      L.ptr(self.gaussian_is_visible_mask), L.stream_ptr())
        L.check(rc, "gsb_densify_accumulate_batch")


def timestep_step(act, cams, targets, bg, stats=None, w_l1=0.8, w_ssim=0.2):
    """act: activated parameters (means3D, colors_precomp, opacities, scales, rotations) on the GPU;
    cams: gsb200.scene.stack_cameras(...); targets [n,3,H,W].
    Returns (loss_stats [n,3], grads dict summed over cameras, radii [n,P], RasterState)."""
    n = targets.size(0)
    color, depth, radii, st = raster.forward(
        act["means3D"], act["colors_precomp"], act["opacities"], act["scales"], act["rotations"], None, bg, 1.0,
        cams["viewmatrix"], cams["projmatrix"], cams["tanfovx"], cams["tanfovy"], cams["image_height"],
        cams["image_width"], n_cams=n)
    if n == 1:
        color, radii = color[None], radii[None]
    loss_stats, dL = loss_mod.l1_ssim_fused(color, targets, w_l1, w_ssim, grad=True)
    grads = raster.backward(st, dL, need_cov3D_grad=False)
    if stats is not None:
        g2 = grads["means2D"] if n > 1 else grads["means2D"][None]
        stats.accumulate_batch(radii.contiguous(), g2.contiguous())
    return loss_stats, grads, radii, st
