"""The drop-in Python surface (DGR/diff_gaussian_rasterization/__init__.py:134-196): same call form as the
reference's callers (shared.py:29-42, densify.py:114-133), autograd through raw parameters."""
import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


def test_render_like_the_callers_do(refdgr):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer as Renderer
    from gsb200 import scene
    dev = "cuda"
    P, W, H = 8000, 200, 136
    g = scene.make_gaussians(P, seed=3)
    params = {k: torch.nn.Parameter(v.to(dev).contiguous().requires_grad_(True)) for k, v in g.items()}
    cam = scene.make_cameras(1, W, H, seed=3)[0]
    settings = scene.settings_for(cam, dev)
    assert isinstance(settings, GaussianRasterizationSettings) and settings._fields[0] == "image_height"

    def render_args():  # create_render_arguments, shared.py:29-42
        return {
            "means3D": params["means"], "colors_precomp": params["colors"],
            "rotations": torch.nn.functional.normalize(params["rotation_quaternions"]),
            "opacities": torch.sigmoid(params["opacity_logits"]), "scales": torch.exp(params["log_scales"]),
            "means2D": torch.zeros_like(params["means"], requires_grad=True, device="cuda") + 0,
        }

    a = render_args()
    a["means2D"].retain_grad()   # densify.py:119
    image, radii, depth = Renderer(raster_settings=settings)(**a)
    assert image.shape == (3, H, W) and depth.shape == (1, H, W) and radii.shape == (P,) and radii.dtype == torch.int32
    target = torch.rand((3, H, W), generator=torch.Generator().manual_seed(0)).to(dev)
    loss = (image - target).abs().mean()
    loss.backward()
    assert a["means2D"].grad.shape == (P, 3) and float(a["means2D"].grad[:, 2].abs().max()) == 0.0

    # the same through the unmodified reference extension
    act = {k: v.detach() for k, v in a.items()}
    bg = settings.bg
    fw = helpers.ref_forward(refdgr, act, cam, bg, dev)
    assert torch.equal(radii, fw["radii"])
    assert (image.detach() - fw["color"]).abs().max().item() <= 1e-5
    dL = torch.sign(fw["color"] - target) / target.numel()
    gr = helpers.ref_backward(refdgr, act, cam, bg, dev, fw, dL)
    nbad, maxerr = helpers.grad_close(helpers.to_np(a["means2D"].grad), helpers.to_np(gr["means2D"]))
    assert nbad == 0, maxerr
    # chain rule of the activations, applied to the reference's gradients with torch
    ref_params = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    torch.autograd.backward(
        [ref_params["means"], ref_params["colors"], torch.nn.functional.normalize(ref_params["rotation_quaternions"]),
         torch.sigmoid(ref_params["opacity_logits"]), torch.exp(ref_params["log_scales"])],
        [gr["means3D"], gr["colors"], gr["rotations"], gr["opacities"], gr["scales"]])
    for k in ("means", "colors", "rotation_quaternions", "opacity_logits", "log_scales"):
        nbad, maxerr = helpers.grad_close(helpers.to_np(params[k].grad), helpers.to_np(ref_params[k].grad), floor=2e-6)
        assert nbad == 0, (k, nbad, maxerr)


def test_argument_validation_and_mark_visible():
    from diff_gaussian_rasterization import GaussianRasterizer
    from gsb200 import scene
    dev = "cuda"
    cam = scene.make_cameras(1, 64, 64)[0]
    r = GaussianRasterizer(scene.settings_for(cam, dev))
    m = torch.zeros(10, 3, device=dev)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=m[:, :1])
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(means3D=m, means2D=m, opacities=m[:, :1], colors_precomp=m)
    with pytest.raises(RuntimeError, match="num_points, 3"):
        r(means3D=torch.zeros(10, 4, device=dev), means2D=m, opacities=m[:, :1], colors_precomp=m, scales=m,
          rotations=torch.zeros(10, 4, device=dev))
    vis = r.markVisible(torch.tensor([[0.0, 0.0, 0.0], [100.0, 100.0, 100.0]], device=dev))
    assert vis.dtype == torch.bool and vis.tolist() == [True, False]


def _params(P, seed, dev):
    from gsb200 import scene
    g = scene.make_gaussians(P, seed=seed)
    return {k: torch.nn.Parameter(v.to(dev).contiguous().requires_grad_(True)) for k, v in g.items()}


def _args(params):
    return {"means3D": params["means"], "colors_precomp": params["colors"],
            "rotations": torch.nn.functional.normalize(params["rotation_quaternions"]),
            "opacities": torch.sigmoid(params["opacity_logits"]), "scales": torch.exp(params["log_scales"]),
            "means2D": torch.zeros_like(params["means"], requires_grad=True) + 0}


def test_backward_twice_with_retain_graph():
    """The reference keeps its state in save_for_backward (DGR/py:84), so its node survives retain_graph=True; so
    does this one, and both passes give the same gradients."""
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    from gsb200 import scene
    dev = "cuda"
    params = _params(3000, 4, dev)
    cam = scene.make_cameras(1, 96, 80, seed=4)[0]
    image, _, _ = Renderer(raster_settings=scene.settings_for(cam, dev))(**_args(params))
    loss = (image * torch.linspace(0, 1, image.numel(), device=dev).view_as(image)).sum()
    loss.backward(retain_graph=True)
    first = {k: v.grad.clone() for k, v in params.items() if v.grad is not None}
    for v in params.values():
        v.grad = None
    loss.backward()
    for k, g in first.items():
        nbad, worst = helpers.grad_close(helpers.to_np(params[k].grad), helpers.to_np(g), rel=1e-5, floor=1e-6)
        assert nbad == 0, (k, worst)


def test_inplace_change_between_forward_and_backward_is_caught():
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    from gsb200 import scene
    dev = "cuda"
    params = _params(500, 5, dev)
    cam = scene.make_cameras(1, 64, 64, seed=5)[0]
    a = _args(params)
    image, _, _ = Renderer(raster_settings=scene.settings_for(cam, dev))(**a)
    with torch.no_grad():
        a["scales"].mul_(1.5)   # e.g. an optimiser step issued before backward
    with pytest.raises(RuntimeError, match="modified by an inplace operation"):
        image.sum().backward()


def test_prefiltered_violation_raises():
    """prefiltered=True with a Gaussian behind the camera: the reference executes __trap() on the device
    (cr/auxiliary.h:154-161); here the call returns GSB_EPREFILTER -> RuntimeError and the context survives."""
    from gsb200 import raster, scene
    dev = "cuda"
    _, a, cams = helpers.scene_inputs(2000, 64, 64, seed=6, device=dev)
    cam = cams[0]
    bg = torch.zeros(3, device=dev)
    # in_frustum only objects to p_view.z <= 0.2 (cr/auxiliary.h:150-161): put a few Gaussians behind the camera
    a["means3D"][:5] = 1.5 * cam["campos"].to(dev) + 0.01 * a["means3D"][:5]
    call = lambda pre: raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None,
                                      bg, 1.0, cam["viewmatrix"].to(dev), cam["projmatrix"].to(dev), cam["tanfovx"],
                                      cam["tanfovy"], 64, 64, prefiltered=pre)
    _, _, radii, _ = call(False)
    assert int((radii == 0).sum()) > 0          # the scene does have culled Gaussians
    with pytest.raises(RuntimeError, match="filtered although prefiltered"):
        call(True)
    color, _, _, _ = call(False)                # the device is still usable
    assert torch.isfinite(color).all()
    vis = torch.ones_like(radii, dtype=torch.bool)   # and the promise holds once those Gaussians are filtered out
    vis[:5] = False
    sub = {k: v[vis].contiguous() for k, v in a.items()}
    c2, _, r2, _ = raster.forward(sub["means3D"], sub["colors_precomp"], sub["opacities"], sub["scales"], sub["rotations"],
                                  None, bg, 1.0, cam["viewmatrix"].to(dev), cam["projmatrix"].to(dev), cam["tanfovx"],
                                  cam["tanfovy"], 64, 64, prefiltered=True)
    assert torch.equal(r2, radii[vis]) and (c2 - color).abs().max().item() <= 1e-5


def test_colors_per_camera_batch_vs_reference(refdgr):
    """Batched call with one colour table per camera: every camera's image and colour gradient equal the reference
    extension's single-camera call with that table; geometry gradients are the sums."""
    from gsb200 import raster, scene
    dev = "cuda"
    P, W, H, n = 6000, 160, 112, 3
    _, a, cams = helpers.scene_inputs(P, W, H, seed=8, n_cams=n, device=dev)
    sc = scene.stack_cameras(cams, dev)
    bg = torch.tensor([0.3, 0.1, 0.2], device=dev)
    cols = torch.rand((n, P, 3), generator=torch.Generator().manual_seed(1)).to(dev)
    color, depth, radii, st = raster.forward(a["means3D"], cols, a["opacities"], a["scales"], a["rotations"], None, bg, 1.0,
                                             sc["viewmatrix"], sc["projmatrix"], sc["tanfovx"], sc["tanfovy"], H, W,
                                             n_cams=n, colors_per_camera=True)
    dL = torch.randn(color.shape, generator=torch.Generator().manual_seed(2)).to(dev)
    g = raster.backward(st, dL, need_cov3D_grad=False)
    sums = None
    for c in range(n):
        ac = dict(a, colors_precomp=cols[c].contiguous())
        fw = helpers.ref_forward(refdgr, ac, cams[c], bg, dev)
        assert torch.equal(radii[c], fw["radii"])
        assert (color[c] - fw["color"]).abs().max().item() <= 1e-5
        gr = helpers.ref_backward(refdgr, ac, cams[c], bg, dev, fw, dL[c].contiguous())
        nbad, worst = helpers.grad_close(helpers.to_np(g["colors"][c]), helpers.to_np(gr["colors"]))
        assert nbad == 0, (c, worst)
        nbad, worst = helpers.grad_close(helpers.to_np(g["means2D"][c]), helpers.to_np(gr["means2D"]))
        assert nbad == 0, (c, worst)
        cur = {k: gr[k].double() for k in ("means3D", "opacities", "scales", "rotations")}
        sums = cur if sums is None else {k: sums[k] + cur[k] for k in cur}
    for k, v in sums.items():
        nbad, worst = helpers.grad_close(helpers.to_np(g[k]), helpers.to_np(v), floor=2e-6)
        assert nbad == 0, (k, nbad, worst)
