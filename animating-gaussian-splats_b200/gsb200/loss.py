"""Fused L1 + SSIM photometric loss (gsb_ssim_l1).

Drop-in names for the reference's loss pieces:
  calc_ssim(img1, img2, window_size=11, size_average=True)   external.py:68-76
  l1_ssim_loss(img, target, w_l1=0.8, w_ssim=0.2)            densify.py:127-129 / train.py:362-364 mix
Both are differentiable w.r.t. the first image (the render); the target carries no gradient, as in
the reference's use.  Inputs are [3,H,W] or [n,3,H,W] float32 CUDA tensors.
"""
import torch

from . import _lib as L


def _run(img, target, w_l1, w_ssim, want_grad, arena=None, tag="", rows=None):
    L.require_cuda(img, "img")
    squeeze = img.ndim == 3
    x = (img[None] if squeeze else img).contiguous().float()
    y = (target[None] if squeeze else target).contiguous().float()
    if x.shape != y.shape or x.ndim != 4 or x.size(1) != 3:
        raise RuntimeError("expected matching [n,3,H,W] (or [3,H,W]) images")
    n, _, H, W = x.shape
    lib = L.lib()
    if arena is not None:   # buffers that live with the arena: stable addresses, no allocation per step
        scratch = arena.output("ssim_scratch", (lib.gsb_ssim_scratch_bytes(n, H, W),), torch.uint8)
        out = arena.output(tag + "loss_stats", (n, 3), torch.float32)
        grad = arena.output(tag + "dL_dimage", tuple(x.shape), torch.float32) if want_grad else None
    else:
        scratch = torch.empty(lib.gsb_ssim_scratch_bytes(n, H, W), dtype=torch.uint8, device=x.device)
        out = torch.empty((n, 3), dtype=torch.float32, device=x.device)
        grad = torch.empty_like(x) if want_grad else None
    with L.device_guard(x.device):
        if rows is not None:
            if rows.dtype != torch.int32 or tuple(rows.shape) != (n, 4) or not rows.is_cuda or not rows.is_contiguous():
                raise RuntimeError("rows must be a contiguous int32 CUDA tensor [n, 4]")
            rc = lib.gsb_ssim_l1_rows(n, H, W, L.ptr(x), L.ptr(y), L.ptr(rows), float(w_l1), float(w_ssim), 1.0, L.ptr(out),
                                      L.ptr(grad), L.ptr(scratch), L.stream_ptr())
        else:
            rc = lib.gsb_ssim_l1(n, H, W, L.ptr(x), L.ptr(y), float(w_l1), float(w_ssim), 1.0, L.ptr(out), L.ptr(grad),
                                 L.ptr(scratch), L.stream_ptr())
    L.check(rc, "gsb_ssim_l1")
    return out, grad, squeeze


class _L1SSIM(torch.autograd.Function):
    """Returns per-image (l1, ssim, w_l1*l1 + w_ssim*(1-ssim)); the backward is computed eagerly in the
    same pass (one extra kernel) for each of the three outputs' linear combination."""

    @staticmethod
    def forward(ctx, img, target, w_l1, w_ssim, which):
        # gradient of the requested scalar: l1 -> (1,0), ssim -> (0,-1) in (w_l1, w_ssim) terms
        gw = {"loss": (w_l1, w_ssim), "ssim": (0.0, -1.0), "l1": (1.0, 0.0)}[which]
        need = img.requires_grad
        # ONE launch chain: the l1 and ssim columns do not depend on the weights, and for "loss" the gradient weights
        # are the mix weights, so the run that produces the gradient also produces the value that is returned
        if need:
            out, grad, squeeze = _run(img, target, gw[0], gw[1], True)
        else:
            out, grad, squeeze = _run(img, target, w_l1, w_ssim, False)
        ctx.save_for_backward(grad)
        ctx.squeeze = squeeze
        col = {"l1": 0, "ssim": 1, "loss": 2}[which]
        return out[:, col].clone()

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        if grad is None:
            return None, None, None, None, None
        gi = grad * g.view(-1, 1, 1, 1)
        return (gi[0] if ctx.squeeze else gi), None, None, None, None


def l1_ssim_loss(img, target, w_l1=0.8, w_ssim=0.2):
    """0.8 * l1_loss(img, target) + 0.2 * (1 - calc_ssim(img, target)), per image (scalar for [3,H,W])."""
    r = _L1SSIM.apply(img, target, w_l1, w_ssim, "loss")
    return r[0] if img.ndim == 3 else r


def calc_ssim(img1, img2, window_size=11, size_average=True):
    """external.calc_ssim (external.py:68-76).  Only the configuration the reference uses exists:
    window 11, sigma 1.5, mean over all elements."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("the reference only ever calls calc_ssim with window_size=11, size_average=True")
    r = _L1SSIM.apply(img1, img2, 0.0, 1.0, "ssim")
    return r[0] if img1.ndim == 3 else r.mean()


def l1_ssim_fused(img, target, w_l1=0.8, w_ssim=0.2, grad=True, arena=None, tag="", rows=None):
    """Non-autograd fast path used by the training step: one call, returns (stats [n,3], dL/dimg or None).
    arena (gsb200.raster.RasterArena): outputs and scratch live in the arena instead of fresh allocations (`tag`
    names the loss numbers and dL/dimg of a second loss in the same step; the scratch is shared, the calls run in order).
    rows (int32 [n, 4] = loss_y0, loss_y1, write_y0, write_y1): image bands of tile-band sharding (gsb_ssim_l1_rows)."""
    out, g, squeeze = _run(img, target, w_l1, w_ssim, grad, arena, tag, rows)
    return out, (g[0] if (squeeze and g is not None) else g)
