"""Front end of the first-frame ("densify") optimisation on the B200 path: the caller
of densify.py:201-248 with every device-side step routed through this package -- fused activations, the drop-in
rasterizer, fused L1+SSIM, the densification-statistics and row-surgery kernels, one-launch Adam -- and
the reference's files in (train_meta.json, ims/, seg/, init_pt_cld.npz) and out
(densified_initial_gaussian_cloud_parameters.pth).

Two schedules:
  * densify(...)          the reference's: ONE randomly drawn camera per iteration, colour + segmentation
                          render, total = image + 3 segmentation, densify hook, Adam (densify.py:218-247);
  * densify_batched(...)  all cameras of timestep 0 per optimiser step through the camera-batched entry point
                          (a different optimisation schedule, SURVEY.md 8(e) "semantics caveat"; used by the
                          multi-GPU path where rank r renders cameras r, r+N, ...).
No wandb, no tqdm: `log` is an optional callable receiving a dict per iteration.
"""
from pathlib import Path
from random import Random

import torch

from . import densify as gdens
from . import io as gio
from . import loss as gloss
from . import optim as goptim


def create_densification_variables(gaussian_cloud_parameters):   # densify.py:89-107
    n, dev = gaussian_cloud_parameters["means"].shape[0], gaussian_cloud_parameters["means"].device
    z = lambda: torch.zeros(n, dtype=torch.float32, device=dev)
    return gdens.DensificationVariables(visibility_count=z(), mean_2d_gradients_accumulated=z(), max_2d_radii=z())


def get_random_element(input_list, fallback_list, rng):   # densify.py:110-113
    if not input_list:
        input_list.extend(fallback_list)
    return input_list.pop(rng.randint(0, len(input_list) - 1))


def _render(parameters, view, colors=None, want_means2d=False):
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    arguments = goptim.create_render_arguments(parameters)
    if colors is not None:
        arguments["colors_precomp"] = colors
    if want_means2d:
        arguments["means2D"].retain_grad()
    image, radii, _ = Renderer(raster_settings=view.render_settings)(**arguments)
    return image, radii, arguments["means2D"]


def calculate_image_and_segmentation_loss(gaussian_cloud_parameters, target_view, densification_variables):
    """densify.py:116-187: (total, image_loss, segmentation_loss, densification_variables); the colour render's
    means2D carrier and radii feed the densification statistics."""
    image, radii, means_2d = _render(gaussian_cloud_parameters, target_view, want_means2d=True)
    image_loss = gloss.l1_ssim_loss(image, target_view.image)
    mask, _, _ = _render(gaussian_cloud_parameters, target_view,
                         colors=gaussian_cloud_parameters["segmentation_masks"])
    segmentation_loss = gloss.l1_ssim_loss(mask, target_view.segmentation_mask)
    densification_variables.means_2d = means_2d
    gdens.update_max_2d_radii_and_visibility_mask(radii=radii, densification_variables=densification_variables)
    return image_loss + 3 * segmentation_loss, image_loss, segmentation_loss, densification_variables


def densify(sequence_path, iteration_count=30_000, log=None, seed=None, device="cuda", export=True):
    """densify.py:201-248.  Returns the parameter dict (also written to the sequence directory)."""
    sequence_path = Path(sequence_path)
    dataset_metadata = gio.load_dataset_metadata(sequence_path)
    parameters = gio.initialize_parameters(sequence_path, device=device)
    scene_radius = gio.get_scene_radius(dataset_metadata)
    optimizer = goptim.create_optimizer(parameters=parameters, scene_radius=scene_radius)
    densification_variables = create_densification_variables(parameters)
    timestep_views = gio.load_timestep_views(dataset_metadata, 0, sequence_path, device=device)
    rng, buffer = Random(seed), []
    for i in range(iteration_count):
        view = get_random_element(buffer, timestep_views, rng)
        total, image_loss, segmentation_loss, densification_variables = calculate_image_and_segmentation_loss(
            parameters, view, densification_variables)
        total.backward()
        if log is not None:
            log({"i": i, "loss/image": image_loss.detach(), "loss/segmentation": segmentation_loss.detach(),
                 "loss/total": total.detach(), "gaussian_count": parameters["means"].shape[0]})
        with torch.no_grad():
            gdens.densify_gaussians(gaussian_cloud_parameters=parameters,
                                    densification_variables=densification_variables, scene_radius=scene_radius,
                                    optimizer=optimizer, i=i)
            optimizer.step()
            optimizer.zero_grad(set_to_none=True)
    if export:
        gio.export_parameters(sequence_path, parameters)
    return parameters


def batched_iteration(parameters, cams, images, masks, bg, stats=None):
    """One camera-batched iteration over the cameras in `cams`: colour render against `images`, segmentation
    render against `masks` (weight 3, densify.py:177), both through gsb200.step.timestep_step; the summed
    gradients are chained through the fused activation adjoint and left in `.grad` of the raw parameters.
    Only the colour render feeds the densification statistics (densify.py:147-150).
    Returns (loss numbers [n,3] colour, loss numbers [n,3] segmentation, radii [n,P])."""
    from . import step as gstep
    rot, opa, sca = goptim.activate(parameters)
    with torch.no_grad():
        act = {"means3D": parameters["means"].data, "colors_precomp": parameters["colors"].data,
               "rotations": rot.detach(), "opacities": opa.detach(), "scales": sca.detach()}
        loss_c, g_c, radii, _ = gstep.timestep_step(act, cams, images, bg, stats)
        act["colors_precomp"] = parameters["segmentation_masks"].data
        loss_s, g_s, _, _ = gstep.timestep_step(act, cams, masks, bg, None, w_l1=3 * 0.8, w_ssim=3 * 0.2)
        P = parameters["means"].shape[0]
        parameters["means"].grad = g_c["means3D"] + g_s["means3D"]
        parameters["colors"].grad = g_c["colors"].reshape(P, 3)
        parameters["segmentation_masks"].grad = g_s["colors"].reshape(P, 3)
    torch.autograd.backward([rot, opa, sca], [g_c["rotations"] + g_s["rotations"],
                                              (g_c["opacities"] + g_s["opacities"]).reshape(opa.shape),
                                              g_c["scales"] + g_s["scales"]])
    return loss_c, loss_s, radii


def _is_densify_event(i):   # external.py:218-221
    return 500 <= i <= 5000 and i % 100 == 0


def densify_batched(sequence_path, step_count, log=None, device="cuda", export=True):
    """First-frame optimisation with ALL cameras of timestep 0 per optimiser step.  Under torch.distributed
    rank r renders cameras r, r + world, ...; the raw-parameter gradients are summed with one allreduce per
    step and the densification statistics when a densify event reads them, so every rank applies the same
    update and the same clone / split / prune decisions (the split draw is seeded identically)."""
    import torch.distributed as dist
    from . import dist as gdist
    from . import step as gstep
    sequence_path = Path(sequence_path)
    rank, n_ranks = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)
    dataset_metadata = gio.load_dataset_metadata(sequence_path)
    parameters = gio.initialize_parameters(sequence_path, device=device)
    scene_radius = gio.get_scene_radius(dataset_metadata)
    optimizer = goptim.create_optimizer(parameters=parameters, scene_radius=scene_radius)
    views = gio.load_timestep_views(dataset_metadata, 0, sequence_path, device=device)
    mine = [views[c] for c in gdist.shard_cameras(len(views), rank, n_ranks)]
    cams, images, masks = gio.stack_views(mine)
    dev = images.device
    bg = torch.zeros(3, dtype=torch.float32, device=dev)
    stats = gstep.DensifyStats(parameters["means"].shape[0], dev)
    raw_keys = ("means", "colors", "segmentation_masks", "rotation_quaternions", "opacity_logits", "log_scales")
    for i in range(step_count):
        loss_c, loss_s, _ = batched_iteration(parameters, cams, images, masks, bg, stats)
        if n_ranks > 1:
            flat = torch.cat([parameters[k].grad.reshape(parameters[k].shape[0], -1) for k in raw_keys], dim=1)
            dist.all_reduce(flat)
            o = 0
            for k in raw_keys:
                w = parameters[k][0].numel()
                parameters[k].grad = flat[:, o:o + w].reshape(parameters[k].shape).contiguous()
                o += w
        if log is not None:
            log({"i": i, "loss/image": loss_c[:, 2].sum(), "loss/segmentation": loss_s[:, 2].sum() / 3,
                 "gaussian_count": parameters["means"].shape[0]})
        with torch.no_grad():
            acc, cnt, rad = stats.mean_2d_gradients_accumulated, stats.visibility_count, stats.max_2d_radii
            if n_ranks > 1 and _is_densify_event(i):   # global statistics, local accumulators untouched
                acc, cnt, rad = acc.clone(), cnt.clone(), rad.clone()
                dist.all_reduce(acc); dist.all_reduce(cnt); dist.all_reduce(rad, op=dist.ReduceOp.MAX)
                torch.manual_seed(1_000_003 + i)   # same split samples on every rank
            # the batched statistics kernel already accumulated this step: an all-false visibility mask makes
            # densify's own accumulate (external.py:113-124) a no-op
            dv = gdens.DensificationVariables(cnt, acc, rad, torch.zeros_like(stats.gaussian_is_visible_mask))
            dv.means_2d = type("G", (), {"grad": torch.zeros_like(parameters["means"])})()
            gdens.densify_gaussians(parameters, dv, scene_radius, optimizer, i)
            if dv.visibility_count is not cnt:   # an event rebuilt the tables: statistics restart from zero
                stats = gstep.DensifyStats(parameters["means"].shape[0], dev)
            optimizer.step()
            optimizer.zero_grad(set_to_none=True)
    if export and rank == 0:
        gio.export_parameters(sequence_path, parameters)
    return parameters
