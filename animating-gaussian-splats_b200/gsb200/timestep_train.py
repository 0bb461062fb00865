"""Per-timestep training step on the B200 path (the train.py side): the caller of
train.py:700-790 with every device-side piece from this package --

    DeformationNetwork (tcgen05 forward + backward)  ->  means / rotation_quaternions of the timestep
    camera-batched rasterizer forward + fused L1/SSIM + rasterizer backward over the sampled views (ONE launch chain)
    rigidity loss once per step (weighted by the number of views, train.py:373-377 evaluates it per view)
    Adam on the network's parameters

calculate_loss mirrors train.py:385-433: total = sum_v (0.8 l1_v + 0.2 (1 - ssim_v)) + 3 sum_v rigidity.
"""
import torch

from . import deform, io as gio, optim, rigidity
from . import step as gstep

_KEYS = ("means", "rotation_quaternions", "colors", "opacity_logits", "log_scales")


class _BatchedImageLoss(torch.autograd.Function):
    """sum over the views of 0.8 l1 + 0.2 (1 - ssim) as one autograd node: the forward runs the camera-batched
    render + loss + rasterizer backward chain (gsb200.step.timestep_step), so the gradients already exist when autograd
    asks for them."""

    @staticmethod
    def forward(ctx, means, quats, colors, logits, log_scales, cams, targets, bg, exchange=None):
        raw = {"rotation_quaternions": quats.detach(), "opacity_logits": logits.detach(), "log_scales": log_scales.detach()}
        with torch.enable_grad():
            leaves = {k: v.detach().requires_grad_(True) for k, v in raw.items()}
            rot, opa, sca = optim.activate(leaves)
        act = {"means3D": means.detach().contiguous(), "colors_precomp": colors.detach().contiguous(),
               "rotations": rot.detach(), "opacities": opa.detach(), "scales": sca.detach()}
        if exchange is not None:
            # view-sharded step (gsb200.dist.Exchange): this rank renders its share of the step's views (possibly
            # none) into the flat record, ONE exchange sums the records, every rank continues with the full gradient
            from . import dist as gdist
            import torch.distributed as dist
            p = means.shape[0]
            if targets is not None and targets.shape[0] > 0:
                loss_stats, flat, _, _ = gstep.timestep_step(act, cams, targets, bg, None, flat=exchange.flat)
                total = loss_stats[:, 2].sum()
            else:
                exchange.flat.zero_()
                loss_stats, total = None, torch.zeros((), dtype=torch.float32, device=means.device)
            exchange.allreduce()
            if dist.is_initialized() and exchange.world > 1:
                total = total.clone()
                dist.all_reduce(total, group=exchange.group)
            u = gdist.unpack(exchange.flat)
            g = {k: u[k].contiguous() for k in ("means3D", "colors", "rotations", "opacities", "scales")}
            torch.autograd.backward([rot, opa, sca], [g["rotations"], g["opacities"].reshape(opa.shape), g["scales"]])
            ctx.grads = (g["means3D"], leaves["rotation_quaternions"].grad, g["colors"].reshape(p, 3),
                         leaves["opacity_logits"].grad, leaves["log_scales"].grad)
            ctx.per_view = loss_stats
            return total
        loss_stats, g, _, _ = gstep.timestep_step(act, cams, targets, bg, None)
        # chain the activated-parameter gradients to the raw tables through the fused activation adjoint
        torch.autograd.backward([rot, opa, sca], [g["rotations"], g["opacities"].reshape(opa.shape), g["scales"]])
        p = means.shape[0]
        ctx.grads = (g["means3D"], leaves["rotation_quaternions"].grad, g["colors"].reshape(p, 3),
                     leaves["opacity_logits"].grad, leaves["log_scales"].grad)
        ctx.per_view = loss_stats
        return loss_stats[:, 2].sum()

    @staticmethod
    def backward(ctx, go):
        need = ctx.needs_input_grad
        out = tuple((g * go if need[i] else None) for i, g in enumerate(ctx.grads))
        ctx.grads = None
        return out + (None, None, None, None)


def batched_image_loss(gaussian_cloud_parameters, cams, targets, bg=None, exchange=None):
    p = gaussian_cloud_parameters
    if bg is None:
        bg = torch.zeros(3, dtype=torch.float32, device=p["means"].device)
    return _BatchedImageLoss.apply(p["means"], p["rotation_quaternions"], p["colors"], p["opacity_logits"],
                                   p["log_scales"], cams, targets, bg, exchange)


def calculate_loss(gaussian_cloud_parameters, target_views, initial_neighbor_info, previous_timestep_foreground_info,
                   exchange=None, view_count=None):
    """train.py:385-433.  target_views: a list of gsb200.io.View, or the (cams, images) pair gsb200.io.stack_views
    returns.  Returns (total, image_loss, rigidity_loss) with rigidity_loss the single-evaluation value.
    exchange / view_count: the view-sharded form -- target_views are THIS rank's share (images may be None or empty),
    view_count the step's total number of views over all ranks."""
    if isinstance(target_views, (list, tuple)) and len(target_views) and isinstance(target_views[0], gio.View):
        cams, images, _ = gio.stack_views(target_views)
    else:
        cams, images = target_views
    n_views = view_count if view_count is not None else images.shape[0]
    image_loss = batched_image_loss(gaussian_cloud_parameters, cams, images, exchange=exchange)
    rigid = rigidity.calculate_rigidity_loss(gaussian_cloud_parameters, initial_neighbor_info,
                                             previous_timestep_foreground_info)
    return image_loss + 3 * n_views * rigid, image_loss, rigid


def train_step(deformation_network, optimizer, initial_gaussian_cloud_parameters,
               encoded_normalized_initial_means_and_rotations, encoded_normalized_previous_means_and_rotations,
               previous_timestep_foreground_info, initial_neighbor_info, target_views, timestep, timestep_count,
               learning_rate_scheduler=None, exchange=None, view_count=None):
    """One iteration of train.py:745-790.  Returns (total loss, the timestep's encoded means/rotations, its foreground info)
    -- the latter two are the "previous timestep" inputs of the next call."""
    updated = deform.update_gaussian_cloud_parameters(
        deformation_network, initial_gaussian_cloud_parameters, encoded_normalized_initial_means_and_rotations,
        encoded_normalized_previous_means_and_rotations, timestep, timestep_count)
    total, _, _ = calculate_loss(updated, target_views, initial_neighbor_info, previous_timestep_foreground_info,
                                 exchange=exchange, view_count=view_count)
    with torch.no_grad():   # train.py:766-772: the next step's "previous" state, detached
        detached = {k: (v.detach() if torch.is_tensor(v) else v) for k, v in updated.items()}
        encoded = deform.normalize_and_encode_means_and_rotations(detached)
        foreground = rigidity.create_foreground_info(detached, initial_neighbor_info.indices)
    total.backward()
    optimizer.step()
    if learning_rate_scheduler is not None:
        learning_rate_scheduler.step()
    optimizer.zero_grad()
    return total.detach(), encoded, foreground
