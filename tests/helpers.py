"""Shared helpers for the parity tests."""
import numpy as np
import torch

from gsb200 import scene


def scene_inputs(P, W, H, seed=0, n_cams=1, device="cpu", scale_override=None, extent=1.5):
    g = scene.make_gaussians(P, seed=seed, scale_override=scale_override, extent=extent)
    a = {k: v.to(device).contiguous() for k, v in scene.activate(g).items()}
    cams = scene.make_cameras(n_cams, W, H, seed=seed)
    return g, a, cams


def ref_forward(refdgr, a, cam, bg, device, cov3D_precomp=None, scale_modifier=1.0):
    """Call the unmodified reference extension the way DGR/diff_gaussian_rasterization/__init__.py:55-79 does."""
    e = torch.Tensor([])
    args = (bg, a["means3D"], a["colors_precomp"], a["opacities"],
            a["scales"] if cov3D_precomp is None else e, a["rotations"] if cov3D_precomp is None else e,
            scale_modifier, cov3D_precomp if cov3D_precomp is not None else e,
            cam["viewmatrix"].to(device), cam["projmatrix"].to(device), cam["tanfovx"], cam["tanfovy"],
            cam["image_height"], cam["image_width"], e, 0, cam["campos"].to(device), False)
    num_rendered, color, radii, geom, binning, img, depth = refdgr.rasterize_gaussians(*args)
    return dict(num_rendered=num_rendered, color=color, radii=radii, geom=geom, binning=binning, img=img, depth=depth)


def ref_backward(refdgr, a, cam, bg, device, fw, dL_dpix, cov3D_precomp=None, scale_modifier=1.0):
    e = torch.Tensor([])
    args = (bg, a["means3D"], fw["radii"], a["colors_precomp"],
            a["scales"] if cov3D_precomp is None else e, a["rotations"] if cov3D_precomp is None else e,
            scale_modifier, cov3D_precomp if cov3D_precomp is not None else e,
            cam["viewmatrix"].to(device), cam["projmatrix"].to(device), cam["tanfovx"], cam["tanfovy"], dL_dpix, e, 0,
            cam["campos"].to(device), fw["geom"], fw["num_rendered"], fw["binning"], fw["img"])
    dm2, dcol, dop, dm3, dcov, dsh, dsc, drot = refdgr.rasterize_gaussians_backward(*args)
    return {"means2D": dm2, "colors": dcol, "opacities": dop, "means3D": dm3, "cov3D": dcov, "scales": dsc,
            "rotations": drot}


def grad_close(a, b, rel=1e-4, floor=1e-6, spread=0.0):
    """north_star tolerance for parameter gradients: 1e-4 relative (atomic order differs), with an
    absolute floor for entries that are sums of cancelling terms: `floor` x the tensor's largest
    magnitude, or 16 x `spread`, the reference's OWN run-to-run difference on the same inputs (its
    float atomics are unordered), whichever is larger."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    tol = rel * np.maximum(np.abs(a), np.abs(b)) + max(floor * max(np.abs(b).max(), 1e-30), 16.0 * spread)
    bad = np.abs(a - b) > tol
    return int(bad.sum()), float(np.abs(a - b).max())


def to_np(t):
    return t.detach().cpu().numpy()


class RefRasterize(torch.autograd.Function):
    """The reference's autograd node (DGR/diff_gaussian_rasterization/__init__.py:40-132) around the
    unmodified extension module `refdgr` -- used to time / check the reference through its own call pattern."""

    @staticmethod
    def forward(ctx, refdgr, means3D, means2D, colors, opacities, scales, rotations, cam, bg):
        e = torch.Tensor([])
        dev = means3D.device
        args = (bg, means3D, colors, opacities, scales, rotations, 1.0, e, cam["viewmatrix"].to(dev),
                cam["projmatrix"].to(dev), cam["tanfovx"], cam["tanfovy"], cam["image_height"], cam["image_width"], e, 0,
                cam["campos"].to(dev), False)
        n, color, radii, geom, binning, img, depth = refdgr.rasterize_gaussians(*args)
        ctx.refdgr, ctx.cam, ctx.n, ctx.bg = refdgr, cam, n, bg
        ctx.save_for_backward(colors, means3D, scales, rotations, radii, geom, binning, img)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth

    @staticmethod
    def backward(ctx, g_color, _r, _d):
        colors, means3D, scales, rotations, radii, geom, binning, img = ctx.saved_tensors
        e = torch.Tensor([])
        cam, dev = ctx.cam, means3D.device
        args = (ctx.bg, means3D, radii, colors, scales, rotations, 1.0, e, cam["viewmatrix"].to(dev),
                cam["projmatrix"].to(dev), cam["tanfovx"], cam["tanfovy"], g_color.contiguous(), e, 0,
                cam["campos"].to(dev), geom, ctx.n, binning, img)
        dm2, dcol, dop, dm3, dcov, dsh, dsc, drot = ctx.refdgr.rasterize_gaussians_backward(*args)
        return None, dm3, dm2, dcol, dop, dsc, drot, None, None
