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
