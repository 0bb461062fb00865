"""Geometry reuse between two renders of the same Gaussians through the same camera (gsb200.raster "Geometry
reuse", gsb_raster_forward_colors): the second render must be bit-identical to a fresh one, any change of a
geometry input (in place or a new tensor) must miss, and gradients / a short training run must not change."""
import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


def _setup(P=6000, W=128, H=96, seed=0):
    from gsb200 import scene
    dev = torch.device("cuda")
    raw = {k: torch.nn.Parameter(v.to(dev)) for k, v in scene.make_gaussians(P, seed=seed).items()}
    cams = scene.make_cameras(2, W, H, seed=seed)
    settings = [scene.settings_for(c, dev) for c in cams]
    return raw, settings


def _render(raw, settings, colors):
    from diff_gaussian_rasterization import GaussianRasterizer
    from gsb200 import optim
    a = optim.create_render_arguments(raw)
    a["colors_precomp"] = colors
    if a["means2D"].requires_grad:
        a["means2D"].retain_grad()
    im, radii, depth = GaussianRasterizer(settings)(**a)
    return im, radii, depth, a["means2D"]


def test_second_render_reuses_geometry_bit_exactly():
    from gsb200 import raster
    raw, settings = _setup()
    raster.invalidate_geometry_cache()
    h0 = raster.geom_cache_hits
    with torch.no_grad():
        im1, r1, d1, _ = _render(raw, settings[0], raw["colors"])
        im2, r2, d2, _ = _render(raw, settings[0], raw["segmentation_masks"])      # hit
        assert raster.geom_cache_hits == h0 + 1
        raster._GEOM_CACHE_ON = False
        try:
            want, rw, dw, _ = _render(raw, settings[0], raw["segmentation_masks"])  # fresh
        finally:
            raster._GEOM_CACHE_ON = True
        assert torch.equal(im2, want) and torch.equal(r2, rw) and torch.equal(d2, dw)
        assert torch.equal(r1, r2) and torch.equal(d1, d2) and not torch.equal(im1, im2)
        # a different camera, an in-place edit, or a new tensor with the same values must all miss
        h = raster.geom_cache_hits
        _render(raw, settings[1], raw["colors"])
        assert raster.geom_cache_hits == h
        _render(raw, settings[1], raw["segmentation_masks"])
        assert raster.geom_cache_hits == h + 1
        raw["means"].add_(0.05)
        im3, _, _, _ = _render(raw, settings[1], raw["segmentation_masks"])
        assert raster.geom_cache_hits == h + 1
        raster._GEOM_CACHE_ON = False
        try:
            want3, _, _, _ = _render(raw, settings[1], raw["segmentation_masks"])
        finally:
            raster._GEOM_CACHE_ON = True
        assert torch.equal(im3, want3)
        clone = {k: torch.nn.Parameter(v.detach().clone()) for k, v in raw.items()}
        h = raster.geom_cache_hits
        _render(clone, settings[1], clone["segmentation_masks"])
        assert raster.geom_cache_hits == h


def _loop_grads(raw, settings, cache_on):
    from gsb200 import raster, loss as gloss
    raster._GEOM_CACHE_ON = cache_on
    raster.invalidate_geometry_cache()
    try:
        for p in raw.values():
            p.grad = None
        tgt = torch.rand(3, 96, 128, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
        im, _, _, m2d = _render(raw, settings, raw["colors"])
        sg, _, _, _ = _render(raw, settings, raw["segmentation_masks"])
        total = gloss.l1_ssim_loss(im, tgt) + 3 * gloss.l1_ssim_loss(sg, (tgt > 0.5).float())
        total.backward()
        out = {k: p.grad.clone() for k, p in raw.items() if p.grad is not None}
        out["means2D"] = m2d.grad.clone()
        return float(total.detach()), out
    finally:
        raster._GEOM_CACHE_ON = True


def test_gradients_unchanged_by_reuse():
    raw, settings = _setup(seed=2)
    l0, g0 = _loop_grads(raw, settings[0], False)
    l1, g1 = _loop_grads(raw, settings[0], True)
    assert l0 == l1
    assert set(g0) == set(g1)
    for k in g0:   # same kernels on the same inputs; float atomics reorder between runs
        bad, worst = helpers.grad_close(helpers.to_np(g1[k]), helpers.to_np(g0[k]), rel=1e-4, floor=2e-6)
        assert bad == 0, (k, bad, worst)


def test_fused_adam_invalidates():
    """FusedAdam writes parameters through raw pointers; it must bump torch's version counters, otherwise the next
    iteration would reuse stale geometry."""
    from gsb200 import raster, optim
    raw, settings = _setup(seed=3)
    opt = optim.FusedAdam([{"params": [v], "name": k, "lr": 1e-2} for k, v in raw.items()], lr=0.0, eps=1e-15)
    im, _, _, _ = _render(raw, settings[0], raw["colors"])
    im.sum().backward()
    v0 = raw["means"]._version
    opt.step()
    assert raw["means"]._version > v0
    h = raster.geom_cache_hits
    with torch.no_grad():
        after, _, _, _ = _render(raw, settings[0], raw["colors"])
    assert raster.geom_cache_hits == h and not torch.equal(after, im)
