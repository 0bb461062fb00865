"""Fused activations + Adam.  CPU: the numpy oracle is pinned against torch itself
(autograd of normalize/sigmoid/exp, torch.optim.Adam with the reference's settings).  GPU: the kernels against
the oracle and against torch on the device; FusedAdam survives densify-style state surgery."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import adam_ref  # noqa: E402


def _raw(P, seed=0):
    g = torch.Generator().manual_seed(seed)
    return {"rotation_quaternions": torch.randn((P, 4), generator=g), "opacity_logits": 2 * torch.randn((P, 1), generator=g),
            "log_scales": torch.randn((P, 3), generator=g) - 3}


def _torch_activate(raw):
    return (torch.nn.functional.normalize(raw["rotation_quaternions"]), torch.sigmoid(raw["opacity_logits"]),
            torch.exp(raw["log_scales"]))


def test_oracle_activation_matches_torch():
    raw = {k: v.clone().requires_grad_(True) for k, v in _raw(5000).items()}
    rot, opa, sca = _torch_activate(raw)
    g = torch.Generator().manual_seed(1)
    dr, do, ds = torch.randn(rot.shape, generator=g), torch.randn(opa.shape, generator=g), torch.randn(sca.shape, generator=g)
    (rot * dr).sum().add((opa * do).sum()).add((sca * ds).sum()).backward()
    n = lambda t: t.detach().numpy()
    o_rot, o_opa, o_sca = adam_ref.activate(n(raw["rotation_quaternions"]), n(raw["opacity_logits"]), n(raw["log_scales"]))
    np.testing.assert_allclose(o_rot, n(rot), rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(o_opa, n(opa), rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(o_sca, n(sca), rtol=1e-6, atol=0)
    gq, gl, gs = adam_ref.activate_backward(n(raw["rotation_quaternions"]), n(dr), n(do), n(ds), o_opa, o_sca)
    np.testing.assert_allclose(gq, n(raw["rotation_quaternions"].grad), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gl, n(raw["opacity_logits"].grad), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(gs, n(raw["log_scales"].grad), rtol=1e-6, atol=0)


def test_oracle_adam_matches_torch_adam():
    g = torch.Generator().manual_seed(0)
    p = torch.randn((1000, 3), generator=g)
    param = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([{"params": [param], "name": "means", "lr": 0.00016 * 3.7}], lr=0.0, eps=1e-15)
    po, m, v = p.numpy().copy(), np.zeros((1000, 3), np.float32), np.zeros((1000, 3), np.float32)
    for step in range(1, 6):
        grad = torch.randn((1000, 3), generator=g) * (10.0 ** (step - 3))
        param.grad = grad.clone()
        opt.step()
        po, m, v = adam_ref.adam_step(po, grad.numpy(), m, v, step, 0.00016 * 3.7, eps=1e-15)
        np.testing.assert_allclose(po, param.detach().numpy(), rtol=2e-6, atol=1e-9)
        np.testing.assert_allclose(m, opt.state[param]["exp_avg"].numpy(), rtol=2e-6, atol=2e-7 * np.abs(m).max())
        np.testing.assert_allclose(v, opt.state[param]["exp_avg_sq"].numpy(), rtol=2e-6, atol=1e-20)


@pytest.mark.gpu
def test_fused_activation_vs_oracle_and_torch():
    from gsb200 import optim
    raw_cpu = _raw(30011, seed=4)
    raw = {k: v.cuda().requires_grad_(True) for k, v in raw_cpu.items()}
    args = optim.create_render_arguments({**raw, "means": torch.zeros(30011, 3, device="cuda"),
                                          "colors": torch.zeros(30011, 3, device="cuda")})
    assert set(args) == {"means3D", "colors_precomp", "rotations", "opacities", "scales", "means2D"}
    assert args["means2D"].requires_grad and float(args["means2D"].abs().sum()) == 0.0
    n = lambda t: t.detach().cpu().numpy()
    o = adam_ref.activate(*(n(raw_cpu[k]) for k in ("rotation_quaternions", "opacity_logits", "log_scales")))
    for got, want in zip((args["rotations"], args["opacities"], args["scales"]), o):
        np.testing.assert_allclose(n(got), want, rtol=2e-6, atol=1e-7)
    g = torch.Generator().manual_seed(2)
    ds = [torch.randn(t.shape, generator=g).cuda() for t in (args["rotations"], args["opacities"], args["scales"])]
    sum((t * d).sum() for t, d in zip((args["rotations"], args["opacities"], args["scales"]), ds)).backward()
    ref = {k: v.cuda().requires_grad_(True) for k, v in raw_cpu.items()}
    sum((t * d).sum() for t, d in zip(_torch_activate(ref), ds)).backward()
    for k in ref:
        np.testing.assert_allclose(n(raw[k].grad), n(ref[k].grad), rtol=2e-5, atol=1e-6)
    # a raw table that does not need a gradient gets none
    raw2 = {k: v.cuda() for k, v in raw_cpu.items()}
    raw2["opacity_logits"].requires_grad_(True)
    r2 = optim.activate(raw2)
    r2[1].sum().backward()
    assert raw2["opacity_logits"].grad is not None and raw2["log_scales"].grad is None


@pytest.mark.gpu
def test_fused_adam_matches_torch_adam_and_oracle():
    from gsb200 import optim
    gen = torch.Generator().manual_seed(0)
    shapes = {"means": (20001, 3), "colors": (20001, 3), "segmentation_masks": (20001, 3), "rotation_quaternions": (20001, 4),
              "opacity_logits": (20001, 1), "log_scales": (20001, 3), "camera_matrices": (27, 3, 3), "camera_center": (27, 3)}
    init = {k: torch.randn(s, generator=gen) for k, s in shapes.items()}
    mine = {k: torch.nn.Parameter(v.clone().cuda()) for k, v in init.items()}
    ref = {k: torch.nn.Parameter(v.clone().cuda()) for k, v in init.items()}
    o_mine = optim.create_optimizer(mine, 3.7)
    lrs = {g["name"]: g["lr"] for g in o_mine.param_groups}
    o_ref = torch.optim.Adam([{"params": [v], "name": k, "lr": lrs[k]} for k, v in ref.items()], lr=0.0, eps=1e-15)
    ora = {k: (v.numpy().copy(), np.zeros(shapes[k], np.float32), np.zeros(shapes[k], np.float32)) for k, v in init.items()}
    for step in range(1, 5):
        for k in shapes:
            if k == "camera_center" and step == 2:   # a parameter without a gradient is skipped, its step does not advance
                mine[k].grad = None; ref[k].grad = None
                continue
            gr = torch.randn(shapes[k], generator=gen) * 0.01
            mine[k].grad = gr.cuda(); ref[k].grad = gr.cuda()
            s_k = step - (1 if (k == "camera_center" and step > 2) else 0)
            ora[k] = adam_ref.adam_step(ora[k][0], gr.numpy(), ora[k][1], ora[k][2], s_k, lrs[k], eps=1e-15)
        o_mine.step(); o_ref.step()
        for k in shapes:
            np.testing.assert_allclose(mine[k].detach().cpu().numpy(), ref[k].detach().cpu().numpy(), rtol=2e-6, atol=2e-7, err_msg=k)
            np.testing.assert_allclose(mine[k].detach().cpu().numpy(), ora[k][0], rtol=2e-6, atol=2e-7, err_msg=k)
            sm, sr = o_mine.state[mine[k]], o_ref.state[ref[k]]
            np.testing.assert_allclose(sm["exp_avg"].cpu().numpy(), sr["exp_avg"].cpu().numpy(), rtol=2e-6,
                                       atol=2e-7 * float(sr["exp_avg"].abs().max()))
            np.testing.assert_allclose(sm["exp_avg_sq"].cpu().numpy(), sr["exp_avg_sq"].cpu().numpy(), rtol=2e-6, atol=1e-20)
            assert int(sm["step"]) == int(sr["step"])
    assert torch.equal(mine["segmentation_masks"].detach().cpu(), init["segmentation_masks"])   # lr 0: never moves


@pytest.mark.gpu
def test_fused_adam_through_densify_surgery():
    """densify_gaussians rewrites parameters and Adam moments in the optimiser; FusedAdam keeps stepping afterwards."""
    from gsb200 import densify as gd, optim, scene
    dev = "cuda"
    P = 3000
    params = {k: torch.nn.Parameter(v.to(dev).contiguous()) for k, v in scene.make_gaussians(P, seed=5).items()}
    opt = optim.create_optimizer(params, 3.0)
    for v in params.values():
        v.grad = torch.randn_like(v) * 1e-3
    opt.step()
    dv = gd.DensificationVariables(visibility_count=torch.ones(P, device=dev),
                                   mean_2d_gradients_accumulated=torch.rand(P, device=dev) * 1e-3,
                                   max_2d_radii=torch.zeros(P, device=dev),
                                   gaussian_is_visible_mask=torch.ones(P, dtype=torch.bool, device=dev))
    dv.means_2d = type("G", (), {"grad": torch.rand(P, 3, device=dev) * 1e-3})()
    gd.densify_gaussians(params, dv, 3.0, opt, 600)
    P2 = params["means"].size(0)
    assert P2 != P
    for k, v in params.items():
        assert opt.state[v]["exp_avg"].shape == v.shape, k
        v.grad = torch.randn_like(v) * 1e-3
    before = params["means"].detach().clone()
    opt.step()
    assert not torch.equal(before, params["means"].detach())
    assert torch.isfinite(params["means"]).all()
