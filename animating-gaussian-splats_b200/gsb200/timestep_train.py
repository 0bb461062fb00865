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


class ViewArena:
    """Host-free rendering for the per-timestep loop.  The blocking forward of the reference interface reads the instance
    count back every call (cr/rasterizer_impl.cu:281-282), which stops the host in the middle of every optimiser step --
    behind the network's forward -- so it cannot issue ahead; with the views sharded over several GPUs the step is then
    bound by the host.  This keeps one gsb200.raster.RasterArena per (Gaussians, views, image size): sized by ONE
    blocking render with `slack` x the instances it needed, then every step renders through the capacity-guarded
    asynchronous forward.  The call's status words travel to pinned host memory behind the step (no wait); poll()
    reads the copies that have arrived, grows the arena when a step used more than `grow_at` of it, and raises
    ArenaOverflow if a step did not fit (its images were background and its image gradients zero).

    The arena is sized from the FIRST view set of a given size, so `slack` has to cover how much the instance count
    differs between the view sets that follow: 1.5 is enough for 5 views sampled from a 27-camera rig (their sum varies
    by ~10 %), NOT for a single camera that changes every timestep (one camera of that rig needs 1.5x the instances of
    another) -- when the views are sharded so thinly, pass a larger slack or keep the blocking forward (view_arena=None)."""

    def __init__(self, slack=1.5, grow_at=0.8):
        self.slack, self.grow_at = float(slack), float(grow_at)
        self.arenas = {}          # (P, views, W, H, device) -> RasterArena: a rank's share of the views varies by one
        self.arena, self.key = None, None
        self.pending = []
        self.steps_monitored = 0

    def get(self, act, cams, n, bg):
        from . import raster
        P, dev = act["means3D"].shape[0], act["means3D"].device
        W, H = cams["image_width"], cams["image_height"]
        key = (P, n, W, H, dev)
        self.poll()
        if key not in self.arenas:
            _, _, _, st = raster.forward(act["means3D"], act["colors_precomp"], act["opacities"], act["scales"],
                                         act["rotations"], None, bg, 1.0, cams["viewmatrix"], cams["projmatrix"],
                                         cams["tanfovx"], cams["tanfovy"], H, W, n_cams=n, entries=cams.get("entries"))
            self.arenas[key] = raster.RasterArena(dev).reserve(P, n, W, H, int(self.slack * st.num_rendered) + 4096)
        self.arena, self.key = self.arenas[key], key
        if cams.get("tanfov_dev") is None:
            cams["tanfov_dev"] = raster.tanfov_table(cams["tanfovx"], cams["tanfovy"], n, dev)
        return self.arena

    def monitor(self, state):
        t = state.status_tensor()
        if t is None:
            return
        host = torch.empty(4, dtype=torch.int32, pin_memory=True)
        host.copy_(t, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self.pending.append((ev, host, int(self.arena.capacity), self.key))
        self.steps_monitored += 1

    def poll(self, wait=False):
        from . import raster
        while self.pending and (wait or self.pending[0][0].query()):
            ev, host, capacity, key = self.pending.pop(0)
            ev.synchronize()
            err, used, need = int(host[0]), int(host[1]), int(host[2])
            arena = self.arenas.get(key)
            if err == raster.L.GSB_ENOMEM:
                self.arenas.pop(key, None)          # the next step with these views sizes a new arena
                raise raster.ArenaOverflow(need, capacity)
            raster.L.check(err, "gsb_raster_forward_async")
            if arena is not None and used > self.grow_at * capacity and capacity == arena.capacity:
                P, n, W, H, _ = key
                torch.cuda.synchronize(arena.device)   # rare: nothing may still be using the buffers that are replaced
                arena.reserve(P, n, W, H, int(self.slack * used) + 4096)


class _BatchedImageLoss(torch.autograd.Function):
    """sum over the views of 0.8 l1 + 0.2 (1 - ssim) as one autograd node: the forward runs the camera-batched
    render + loss + rasterizer backward chain (gsb200.step.timestep_step), so the gradients already exist when autograd
    asks for them."""

    @staticmethod
    def forward(ctx, means, quats, colors, logits, log_scales, cams, targets, bg, exchange=None, view_arena=None):
        raw = {"rotation_quaternions": quats.detach(), "opacity_logits": logits.detach(), "log_scales": log_scales.detach()}
        with torch.enable_grad():
            leaves = {k: v.detach().requires_grad_(True) for k, v in raw.items()}
            rot, opa, sca = optim.activate(leaves)
        act = {"means3D": means.detach().contiguous(), "colors_precomp": colors.detach().contiguous(),
               "rotations": rot.detach(), "opacities": opa.detach(), "scales": sca.detach()}
        arena = None
        if view_arena is not None and targets is not None and targets.shape[0] > 0:
            arena = view_arena.get(act, cams, targets.shape[0], bg)
        if exchange is not None:
            # view-sharded step (gsb200.dist.Exchange): this rank renders its share of the step's views (possibly
            # none) into the flat record, ONE exchange sums the records, every rank continues with the full gradient
            from . import dist as gdist
            import torch.distributed as dist
            p = means.shape[0]
            if targets is not None and targets.shape[0] > 0:
                loss_stats, flat, _, st = gstep.timestep_step(act, cams, targets, bg, None, flat=exchange.flat, arena=arena)
                if arena is not None:
                    view_arena.monitor(st)
                    loss_stats = loss_stats.clone()     # the arena's output buffers belong to the next step
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
        loss_stats, g, _, st = gstep.timestep_step(act, cams, targets, bg, None, arena=arena)
        if arena is not None:
            view_arena.monitor(st)
            loss_stats = loss_stats.clone()
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
        return out + (None, None, None, None, None)


def batched_image_loss(gaussian_cloud_parameters, cams, targets, bg=None, exchange=None, view_arena=None):
    p = gaussian_cloud_parameters
    if bg is None:
        bg = torch.zeros(3, dtype=torch.float32, device=p["means"].device)
    return _BatchedImageLoss.apply(p["means"], p["rotation_quaternions"], p["colors"], p["opacity_logits"],
                                   p["log_scales"], cams, targets, bg, exchange, view_arena)


def calculate_loss(gaussian_cloud_parameters, target_views, initial_neighbor_info, previous_timestep_foreground_info,
                   exchange=None, view_count=None, view_arena=None):
    """train.py:385-433.  target_views: a list of gsb200.io.View, or the (cams, images) pair gsb200.io.stack_views
    returns.  Returns (total, image_loss, rigidity_loss) with rigidity_loss the single-evaluation value.
    exchange / view_count: the view-sharded form -- target_views are THIS rank's share (images may be None or empty),
    view_count the step's total number of views over all ranks.  view_arena (ViewArena): render without the blocking
    read-back of the instance count."""
    if isinstance(target_views, (list, tuple)) and len(target_views) and isinstance(target_views[0], gio.View):
        cams, images, _ = gio.stack_views(target_views)
    else:
        cams, images = target_views
    n_views = view_count if view_count is not None else images.shape[0]
    image_loss = batched_image_loss(gaussian_cloud_parameters, cams, images, exchange=exchange, view_arena=view_arena)
    rigid = rigidity.calculate_rigidity_loss(gaussian_cloud_parameters, initial_neighbor_info,
                                             previous_timestep_foreground_info)
    return image_loss + 3 * n_views * rigid, image_loss, rigid


def train_step(deformation_network, optimizer, initial_gaussian_cloud_parameters,
               encoded_normalized_initial_means_and_rotations, encoded_normalized_previous_means_and_rotations,
               previous_timestep_foreground_info, initial_neighbor_info, target_views, timestep, timestep_count,
               learning_rate_scheduler=None, exchange=None, view_count=None, view_arena=None):
    """One iteration of train.py:745-790.  Returns (total loss, the timestep's encoded means/rotations, its foreground info)
    -- the latter two are the "previous timestep" inputs of the next call."""
    updated = deform.update_gaussian_cloud_parameters(
        deformation_network, initial_gaussian_cloud_parameters, encoded_normalized_initial_means_and_rotations,
        encoded_normalized_previous_means_and_rotations, timestep, timestep_count)
    total, _, _ = calculate_loss(updated, target_views, initial_neighbor_info, previous_timestep_foreground_info,
                                 exchange=exchange, view_count=view_count, view_arena=view_arena)
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
