"""Front end of the first-frame optimisation: gsb200.first_frame over a synthetic
sequence directory in the reference's layout.  The fused loop is compared, iteration by iteration, with the same
loop written in stock torch ops (torch activations, conv2d SSIM, masked-index statistics, torch.optim.Adam)
around the reference CUDA extension when it is built, else around the drop-in rasterizer."""
import math
from random import Random

import numpy as np
import pytest
import torch

import helpers
import oracle_lib

pytestmark = pytest.mark.gpu


def _make_sequence(root, n_cams=4, W=96, H=64, P=3000, seed=0):
    """Ground-truth Gaussians rendered with the drop-in rasterizer -> ims/seg files; noisy init_pt_cld.npz."""
    from gsb200 import io as gio, scene
    from diff_gaussian_rasterization import GaussianRasterizer
    dev = torch.device("cuda")
    gt = scene.make_gaussians(P, seed=seed, extent=1.0)
    gt["opacity_logits"] = gt["opacity_logits"].abs() + 1.0
    act = {k: v.to(dev) for k, v in scene.activate(gt).items()}
    rng = np.random.default_rng(seed)
    meta = {"w": W, "h": H, "k": [[]], "w2c": [[]], "fn": [[]]}
    images, masks = [[]], [[]]
    for c in range(n_cams):
        az = 2 * math.pi * c / n_cams
        w2c = scene.look_at_w2c((3.0 * math.cos(az), 3.0 * math.sin(az), 1.2))
        fx = 0.8 * W
        meta["w2c"][0].append(w2c.tolist())
        meta["k"][0].append([[fx, 0.0, W / 2], [0.0, fx, H / 2], [0.0, 0.0, 1.0]])
        meta["fn"][0].append(f"{c}/000000.jpg")
        s = gio.create_render_settings(W, H, meta["k"][0][c], meta["w2c"][0][c], device=dev)
        zeros = torch.zeros_like(act["means3D"])
        im, _, _ = GaussianRasterizer(s)(means3D=act["means3D"], means2D=zeros, colors_precomp=act["colors_precomp"],
                                         opacities=act["opacities"], scales=act["scales"], rotations=act["rotations"])
        sg, _, _ = GaussianRasterizer(s)(means3D=act["means3D"], means2D=zeros,
                                         colors_precomp=gt["segmentation_masks"].to(dev), opacities=act["opacities"],
                                         scales=act["scales"], rotations=act["rotations"])
        images[0].append((im.clamp(0, 1).permute(1, 2, 0) * 255).round().byte().cpu().numpy())
        masks[0].append((sg[0] > 0.5).byte().cpu().numpy())
    cloud = np.concatenate([gt["means"].numpy() + rng.normal(0, 0.02, size=(P, 3)), gt["colors"].numpy(),
                            gt["segmentation_masks"][:, :1].numpy()], axis=1)
    gio.write_sequence(root, meta, images, masks, cloud)
    return meta


def _stock_loop(root, iterations, seed, ref):
    """densify.py:201-248 in stock torch ops (no fused kernels except the rasterizer under test/oracle)."""
    import bench
    from gsb200 import io as gio
    from diff_gaussian_rasterization import GaussianRasterizer
    dev = torch.device("cuda")
    meta = gio.load_dataset_metadata(root)
    params = gio.initialize_parameters(root, device=dev)
    radius = gio.get_scene_radius(meta)
    lrs = {"means": 0.00016 * radius, "colors": 0.0025, "segmentation_masks": 0.0, "rotation_quaternions": 0.001,
           "opacity_logits": 0.05, "log_scales": 0.001, "camera_matrices": 1e-4, "camera_center": 1e-4}
    opt = torch.optim.Adam([{"params": [v], "name": k, "lr": lrs[k]} for k, v in params.items()], lr=0.0, eps=1e-15)
    views = gio.load_timestep_views(meta, 0, root, device=dev)
    P = params["means"].shape[0]
    acc, cnt, mx = torch.zeros(P, device=dev), torch.zeros(P, device=dev), torch.zeros(P, device=dev)
    rng, buf, losses = Random(seed), [], []

    def render(view, colors):
        a = {"means3D": params["means"], "colors_precomp": colors,
             "rotations": torch.nn.functional.normalize(params["rotation_quaternions"]),
             "opacities": torch.sigmoid(params["opacity_logits"]), "scales": torch.exp(params["log_scales"]),
             "means2D": torch.zeros_like(params["means"], requires_grad=True) + 0}
        a["means2D"].retain_grad()
        s = view.render_settings
        if ref is not None:
            cam = {"viewmatrix": s.viewmatrix, "projmatrix": s.projmatrix, "tanfovx": s.tanfovx, "tanfovy": s.tanfovy,
                   "image_width": s.image_width, "image_height": s.image_height, "campos": s.campos}
            im, radii, _ = helpers.RefRasterize.apply(ref, a["means3D"], a["means2D"], a["colors_precomp"],
                                                      a["opacities"], a["scales"], a["rotations"], cam, s.bg)
        else:
            im, radii, _ = GaussianRasterizer(s)(**a)
        return im, radii, a["means2D"]

    for i in range(iterations):
        # densify.py:108-111 as the reference behaves: its local rebinding leaves the caller's buffer empty, so every
        # iteration draws uniformly with replacement from all views
        draw = buf if buf else list(views)
        view = draw.pop(rng.randint(0, len(draw) - 1))
        im, radii, m2d = render(view, params["colors"])
        l_img = bench.torch_ssim_l1(im, view.image)
        sg, _, _ = render(view, params["segmentation_masks"])
        total = l_img + 3 * bench.torch_ssim_l1(sg, view.segmentation_mask)
        total.backward()
        losses.append(float(total.detach()))
        with torch.no_grad():
            pos = radii > 0
            mx[pos] = torch.max(radii[pos].float(), mx[pos])
            acc[pos] += torch.norm(m2d.grad[pos, :2], dim=-1)
            cnt[pos] += 1
            opt.step()
            opt.zero_grad(set_to_none=True)
    return params, losses, (acc, cnt, mx)


def test_densify_loop_matches_stock_torch_loop(tmp_path):
    from gsb200 import first_frame
    _make_sequence(tmp_path)
    ref = oracle_lib.load_reference_ext()
    n = 24   # before the first densify event (i = 500): pure render / loss / statistics / Adam trajectory
    losses = []
    params = first_frame.densify(tmp_path, iteration_count=n, seed=7, export=False,
                                 log=lambda d: losses.append(float(d["loss/total"])))
    want_params, want_losses, _ = _stock_loop(tmp_path, n, 7, ref)
    assert np.allclose(losses, want_losses, rtol=2e-4, atol=1e-6), (losses[-3:], want_losses[-3:])
    # Adam's first steps move every element by ~lr regardless of gradient size, so compare in units of lr * steps
    # (rotation_quaternions are left out: the cloud starts with isotropic scales, so their gradient is rounding
    # noise around an exact zero, which Adam normalises to +-lr per step in both loops)
    for k, lr in (("means", 0.00016 * 3.3), ("colors", 0.0025), ("opacity_logits", 0.05), ("log_scales", 0.001)):
        d = (params[k].data - want_params[k].data).abs()
        assert float(d.mean()) < 0.02 * lr * n, (k, float(d.mean()), float(d.max()))
    assert torch.equal(params["segmentation_masks"].data, want_params["segmentation_masks"].data)   # lr 0


def test_densify_through_events_exports_parameter_file(tmp_path):
    """600+ iterations cross the densify events at i = 500 and 600: the cloud is rebuilt, the loss keeps falling, and
    the exported file reads back through the reference's loader contract."""
    from gsb200 import first_frame, io as gio
    _make_sequence(tmp_path, n_cams=3, W=64, H=48, P=1500)
    log = []
    params = first_frame.densify(tmp_path, iteration_count=620, seed=1,
                                 log=lambda d: log.append((float(d["loss/total"]), d["gaussian_count"])))
    first, last = np.mean([l for l, _ in log[:30]]), np.mean([l for l, _ in log[-30:]])
    assert last < 0.8 * first, (first, last)
    counts = [c for _, c in log]
    assert counts[0] == 1500 and counts[-1] != 1500   # clone / split / prune changed the row count
    n = params["means"].shape[0]
    assert all(params[k].shape[0] == n for k in gio.PARAMETER_KEYS[:6])
    back = gio.load_densified_initial_parameters(tmp_path.parent, tmp_path.name, device="cuda")
    for k in gio.PARAMETER_KEYS:
        assert torch.equal(back[k].data, params[k].data) and not back[k].requires_grad


def test_batched_iteration_matches_sum_of_single_camera_iterations(tmp_path):
    """The camera-batched schedule's gradients = the sum over cameras of the reference-shaped iteration's."""
    from gsb200 import first_frame, io as gio, step as gstep
    _make_sequence(tmp_path, n_cams=3, W=64, H=48, P=1500)
    dev = torch.device("cuda")
    meta = gio.load_dataset_metadata(tmp_path)
    params = gio.initialize_parameters(tmp_path, device=dev)
    with torch.no_grad():   # leave the symmetric initial state so that every table has a gradient signal
        params["rotation_quaternions"].add_(0.3 * torch.randn_like(params["rotation_quaternions"]))
        params["log_scales"].add_(0.2 * torch.randn_like(params["log_scales"]))
    views = gio.load_timestep_views(meta, 0, tmp_path, device=dev)
    dv = first_frame.create_densification_variables(params)
    want_acc = torch.zeros(1500, device=dev)
    for v in views:
        total, _, _, dv = first_frame.calculate_image_and_segmentation_loss(params, v, dv)
        total.backward()
        vis = dv.gaussian_is_visible_mask
        want_acc[vis] += torch.norm(dv.means_2d.grad[vis, :2], dim=-1)
    want = {k: params[k].grad.clone() for k in gio.PARAMETER_KEYS[:6]}
    for p in params.values():
        p.grad = None
    cams, images, masks = gio.stack_views(views)
    stats = gstep.DensifyStats(1500, dev)
    first_frame.batched_iteration(params, cams, images, masks, torch.zeros(3, device=dev), stats)
    for k in gio.PARAMETER_KEYS[:6]:
        bad, worst = helpers.grad_close(helpers.to_np(params[k].grad), helpers.to_np(want[k]), rel=2e-4, floor=2e-6)
        assert bad == 0, (k, bad, worst)
    assert torch.allclose(stats.mean_2d_gradients_accumulated, want_acc, rtol=1e-4, atol=1e-9)


def test_engine_step_equals_batched_iteration_plus_fused_adam(tmp_path):
    """FirstFrameEngine (flat record -> statistics -> activation adjoint + Adam in one kernel) against the same
    iteration through the separate pieces: batched_iteration (two timestep_steps + autograd activations) + FusedAdam."""
    from gsb200 import first_frame, io as gio, optim as goptim, step as gstep
    _make_sequence(tmp_path, n_cams=3, W=64, H=48, P=1500)
    dev = torch.device("cuda")
    meta = gio.load_dataset_metadata(tmp_path)
    radius = gio.get_scene_radius(meta)
    views = gio.load_timestep_views(meta, 0, tmp_path, device=dev)
    cams, images, masks = gio.stack_views(views)

    def fresh():
        params = gio.initialize_parameters(tmp_path, device=dev)
        with torch.no_grad():
            g = torch.Generator(device="cuda").manual_seed(5)
            params["rotation_quaternions"].add_(0.3 * torch.randn(params["rotation_quaternions"].shape, device=dev, generator=g))
            params["log_scales"].add_(0.2 * torch.randn(params["log_scales"].shape, device=dev, generator=g))
        return params, goptim.create_optimizer(params, radius)

    pa, oa = fresh()
    stats = gstep.DensifyStats(1500, dev)
    for _ in range(3):
        first_frame.batched_iteration(pa, cams, images, masks, torch.zeros(3, device=dev), stats)
        oa.step()
        oa.zero_grad(set_to_none=True)
    for graph in (False, True):
        pb, ob = fresh()
        eng = first_frame.FirstFrameEngine(pb, ob, 3, 64, 48, dev, graph=False)
        eng.set_views(cams, images, masks)
        eng.rebuild()
        if graph:   # capture without the warm-up steps of StepGraph: build the graph by hand around one iteration
            eng._iteration(); eng._iteration()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                eng._iteration()
            g.replay()
        else:
            for _ in range(3):
                eng.step(check_every=0)
        torch.cuda.synchronize()
        assert eng.adam.sync_state() == 3
        for k, lr in (("means", 0.00016 * radius), ("colors", 0.0025), ("opacity_logits", 0.05), ("log_scales", 0.001),
                      ("rotation_quaternions", 0.001)):
            d = (pa[k].data - pb[k].data).abs()
            assert float(d.max()) < 0.05 * lr * 3 + 1e-7, (graph, k, float(d.max()), float(d.mean()))
            sa, sb = oa.state[pa[k]], ob.state[pb[k]]
            assert torch.allclose(sa["exp_avg"], sb["exp_avg"], rtol=2e-3, atol=1e-9 + 1e-4 * float(sa["exp_avg"].abs().max())), k
        assert torch.equal(pa["segmentation_masks"].data, pb["segmentation_masks"].data)
        assert torch.equal(eng.stats.visibility_count, stats.visibility_count)
        assert torch.equal(eng.stats.max_2d_radii, stats.max_2d_radii)
        assert torch.allclose(eng.stats.mean_2d_gradients_accumulated, stats.mean_2d_gradients_accumulated, rtol=1e-3, atol=1e-9)
        eng.close()


def test_run_first_frame_crosses_densify_events(tmp_path):
    """The captured loop through two densify events (compressed schedule): row counts change, the engine is rebuilt for
    the new count, the loss falls, Adam's step count survives the surgery."""
    from gsb200 import first_frame, io as gio, optim as goptim
    _make_sequence(tmp_path, n_cams=3, W=64, H=48, P=1500)
    dev = torch.device("cuda")
    meta = gio.load_dataset_metadata(tmp_path)
    radius = gio.get_scene_radius(meta)
    params = gio.initialize_parameters(tmp_path, device=dev)
    opt = goptim.create_optimizer(params, radius)
    cams, images, masks = gio.stack_views(gio.load_timestep_views(meta, 0, tmp_path, device=dev))
    idx = list(range(0, 60)) + [440 + j for j in range(61)] + [540 + j for j in range(61)]   # ..., 500, ..., 600
    log = []
    params, timeline = first_frame.run_first_frame(params, opt, radius, cams, images, masks, len(idx), dev,
                                                   log=lambda d: log.append((float(d["loss/image"]), d["gaussian_count"])),
                                                   iteration_index=lambda k: idx[k])
    assert [e["i"] for e in timeline] == [500, 600]
    assert timeline[0]["before"] == 1500 and timeline[0]["after"] != 1500
    assert params["means"].shape[0] == timeline[-1]["after"]
    n = params["means"].shape[0]
    assert all(params[k].shape[0] == n for k in gio.PARAMETER_KEYS[:6])
    first, last = np.mean([l for l, _ in log[:20]]), np.mean([l for l, _ in log[-20:]])
    assert last < 0.9 * first, (first, last)
    steps = {int(opt.state[params[k]]["step"]) for k in gio.PARAMETER_KEYS[:6]}
    assert len(steps) == 1 and steps.pop() >= len(idx)
