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
    # The reference REBINDS its local name to a copy of the fallback list, so the caller's buffer stays empty and
    # every iteration draws uniformly, with replacement, from all views; reproduced as is.
    if not input_list:
        input_list = fallback_list.copy()
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


class FirstFrameEngine:
    """One optimiser step of the camera-batched first-frame optimisation with the host out of the loop: fused
    activations -> asynchronous forward through a reused arena -> L1/SSIM -> backward into the flat step record ->
    segmentation re-blend (same geometry, weight 3, densify.py:136-150,177) -> its backward ADDED to the record ->
    ONE exchange over the ranks -> densification statistics from the record -> activation adjoint + Adam from the
    record; the ~45 launches are captured once per Gaussian count into a CUDA graph and replayed.  Under
    torch.distributed every rank renders its own cameras (set_views) against the replicated parameters; after the
    exchange all ranks hold the same record, so parameters, Adam moments and statistics stay bit-identical.
    A densify event changes P: call rebuild() afterwards (new exchange block, arena, graph)."""

    def __init__(self, parameters, optimizer, n_views, width, height, device, graph=True, two_table=True):
        self.parameters, self.optimizer = parameters, optimizer
        # two_table: colours and segmentation colours are blended by ONE forward walk and back-propagated by ONE
        # backward walk (gsb_raster_args.colors2); False: colour render, then a re-blend of the same geometry with the
        # segmentation colours and a second backward (the reference's two renders, densify.py:116-150)
        self.two_table = two_table
        self.n, self.W, self.H, self.dev = int(n_views), int(width), int(height), torch.device(device)
        self.use_graph = graph
        n, dev = self.n, self.dev
        f = dict(dtype=torch.float32, device=dev)
        self.cams = {"viewmatrix": torch.zeros((n, 4, 4), **f), "projmatrix": torch.zeros((n, 4, 4), **f),
                     "tanfov_dev": torch.ones((n, 2), **f), "tanfovx": [1.0] * n, "tanfovy": [1.0] * n,
                     "image_width": self.W, "image_height": self.H}
        self.images = torch.zeros((n, 3, self.H, self.W), **f)
        self.masks = torch.zeros((n, 3, self.H, self.W), **f)
        self.bg = torch.zeros(3, **f)
        self.ex = self.arena = self.graph = self.adam = self.stats = None
        self.steps_since_check = 0
        self._views_set = False

    # ---- inputs -------------------------------------------------------------------------------------------
    def set_views(self, cams, images, masks):
        """cams: gsb200.io.stack_views / gsb200.scene.stack_cameras dict of this rank's n_views cameras.  Device
        copies into the step's static slots (the captured graph reads them there)."""
        from . import raster
        self.cams["viewmatrix"].copy_(cams["viewmatrix"].reshape(self.n, 4, 4))
        self.cams["projmatrix"].copy_(cams["projmatrix"].reshape(self.n, 4, 4))
        self.cams["tanfov_dev"].copy_(raster.tanfov_table(cams["tanfovx"], cams["tanfovy"], self.n, self.dev))
        self.images.copy_(images)
        self.masks.copy_(masks)
        self._views_set = True

    # ---- (re)build for the current Gaussian count ---------------------------------------------------------
    def rebuild(self, capacity=None):
        from . import dist as gdist, raster, step as gstep
        if not self._views_set:
            raise RuntimeError("set_views() first: the arena is sized from a render of the views")
        p = self.parameters
        P = p["means"].shape[0]
        self.graph = None
        if self.adam is not None and self.adam.P == P:
            self.adam.sync_state()   # the device step count -> optimizer.state before a new FlatAdam reads it
        if self.ex is not None:
            self.ex.close()
        self.ex = gdist.Exchange(P, self.dev)
        self.stats = gstep.DensifyStats(P, self.dev)
        self.adam = goptim.FlatAdam(self.optimizer)
        f = dict(dtype=torch.float32, device=self.dev)
        self.act_out = (torch.empty((P, 4), **f), torch.empty((P, 1), **f), torch.empty((P, 3), **f))
        if capacity is None:   # one blocking render of the views sizes the instance arena
            goptim.activate_into(p, self.act_out)
            _, _, _, st = raster.forward(p["means"].data, p["colors"].data, self.act_out[1], self.act_out[2], self.act_out[0],
                                         None, self.bg, 1.0, self.cams["viewmatrix"], self.cams["projmatrix"],
                                         [float(x) for x in self.cams["tanfov_dev"][:, 0].cpu()],
                                         [float(x) for x in self.cams["tanfov_dev"][:, 1].cpu()], self.H, self.W, n_cams=self.n)
            capacity = int(1.3 * st.num_rendered) + 65536
            del st
        self.arena = raster.RasterArena(self.dev).reserve(P, self.n, self.W, self.H, capacity)
        self.state = None
        if self.use_graph:
            self.graph = gstep.StepGraph(self._iteration, device=self.dev)
            # the warm-up and capture passes are real optimiser steps on replicated state: every rank ran the same ones
        self.steps_since_check = 0
        return self

    def _iteration(self):
        from . import raster, step as gstep
        p = self.parameters
        rot, opa, sca = goptim.activate_into(p, self.act_out)
        act = {"means3D": p["means"].data, "colors_precomp": p["colors"].data, "rotations": rot, "opacities": opa, "scales": sca}
        flat = self.ex.flat
        if self.two_table:
            c = self.cams
            color, _, radii, st = raster.forward(act["means3D"], act["colors_precomp"], opa, sca, rot, None, self.bg, 1.0,
                                                 c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], self.H, self.W,
                                                 n_cams=self.n, arena=self.arena, tanfov_dev=c["tanfov_dev"],
                                                 colors2=p["segmentation_masks"].data)
            if self.n == 1:
                color = color[None]
            loss_c, dLc = gloss.l1_ssim_fused(color, self.images, 0.8, 0.2, grad=True, arena=self.arena)
            loss_s, dLs = gloss.l1_ssim_fused(st.color2, self.masks, 3 * 0.8, 3 * 0.2, grad=True, arena=self.arena, tag="seg_")
            raster.backward_flat(st, dLc, flat, scratch=self.arena.bwd_scratch, dL_dpix2=dLs)
        else:
            loss_c, _, radii, st = gstep.timestep_step(act, self.cams, self.images, self.bg, None, flat=flat, arena=self.arena)
            seg, _, st2 = raster.reblend(st, p["segmentation_masks"].data, arena=self.arena, name="seg")
            if self.n == 1:
                seg = seg[None]
            loss_s, dLs = gloss.l1_ssim_fused(seg, self.masks, 3 * 0.8, 3 * 0.2, grad=True, arena=self.arena, tag="seg_")
            raster.backward_flat(st2, dLs, flat, mode=1, scratch=self.arena.bwd_scratch)
        self.ex.allreduce()
        self.stats.apply_flat(flat)
        self.adam.step(flat)
        self.state = st
        return loss_c, loss_s

    def step(self, check_every=25):
        """One optimiser step.  Returns (colour loss numbers [n,3], segmentation loss numbers [n,3]) -- device tensors,
        valid until the next step.  Every `check_every` steps the forward's status words are read (one synchronisation):
        an arena overflow grows the arena, recaptures and raises nothing (the overflowing steps applied zero gradients)."""
        out = self.graph.replay() if self.graph is not None else self._iteration()
        self.steps_since_check += 1
        if check_every and self.steps_since_check >= check_every:
            self.check()
        return out

    def check(self):
        from . import raster
        self.steps_since_check = 0
        err, r, need = self.state.status()
        if err == L_ENOMEM:
            self.rebuild(capacity=int(1.5 * need) + 65536)
            return False
        if err != 0:
            raise RuntimeError(f"rasterizer status {err}")
        if r > 0.9 * self.arena.capacity:   # grow ahead of the overflow
            self.rebuild(capacity=int(1.5 * r) + 65536)
        return True

    def densification_variables(self):
        """The replicated statistics as the DensificationVariables densify_gaussians expects; this step's increments
        are already in them, so its own accumulate (external.py:113-124) is made a no-op with an all-false mask."""
        s = self.stats
        dv = gdens.DensificationVariables(s.visibility_count, s.mean_2d_gradients_accumulated, s.max_2d_radii,
                                          torch.zeros_like(s.gaussian_is_visible_mask))
        dv.means_2d = type("G", (), {"grad": torch.zeros((s.visibility_count.shape[0], 3), device=self.dev)})()
        return dv

    def close(self):
        self.graph = None
        if self.ex is not None:
            self.ex.close()
            self.ex = None


L_ENOMEM = -3


def densify_batched(sequence_path, step_count, log=None, device="cuda", export=True, graph=True, on_event=None):
    """First-frame optimisation with ALL cameras of timestep 0 per optimiser step (FirstFrameEngine).  Under
    torch.distributed rank r renders its share of the cameras; the step record (gradients of both renders, statistic
    increments, largest radius) is reduced by ONE exchange per step, so every rank applies the same update and -- the
    split draw being seeded identically -- the same clone / split / prune decisions (external.py:211-314).
    on_event(i, seconds, gaussian_count): called after each densify event."""
    import time
    import torch.distributed as dist
    from . import dist as gdist
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
    parameters, _ = run_first_frame(parameters, optimizer, scene_radius, cams, images, masks, step_count, dev, log=log,
                                    graph=graph, on_event=on_event)
    if export and rank == 0:
        gio.export_parameters(sequence_path, parameters)
    return parameters


def run_first_frame(parameters, optimizer, scene_radius, cams, images, masks, step_count, device, log=None, graph=True,
                    on_event=None, iteration_index=None, max_gaussians=None):
    """The loop of densify_batched on tensors already in memory (bench.py's C5 leg feeds it a synthetic sequence).
    iteration_index(k) -> the reference iteration number i of step k (default k): lets a bounded run visit every
    densify event of external.py:218-221 (i = 500, 600, ..., 5000) with fewer optimiser steps in between.
    max_gaussians: stop after the event that leaves more rows than this (the reference's schedule has no cap)."""
    import time
    n = images.shape[0]
    eng = FirstFrameEngine(parameters, optimizer, n, images.shape[3], images.shape[2], device, graph=graph)
    eng.set_views(cams, images, masks)
    eng.rebuild()
    timeline = []
    for k in range(step_count):
        i = k if iteration_index is None else iteration_index(k)
        loss_c, loss_s = eng.step()
        if log is not None:
            log({"i": i, "loss/image": loss_c[:, 2].sum(), "loss/segmentation": loss_s[:, 2].sum() / 3,
                 "gaussian_count": parameters["means"].shape[0]})
        if i <= 5000 and ((i >= 500 and i % 100 == 0) or (i > 0 and i % 3000 == 0)):
            with torch.no_grad():
                torch.cuda.synchronize(device)
                t0 = time.perf_counter()
                eng.check()
                eng.adam.sync_state()
                before = parameters["means"].shape[0]
                torch.manual_seed(1_000_003 + i)   # same split samples on every rank
                dv = eng.densification_variables()
                gdens.densify_gaussians(parameters, dv, scene_radius, optimizer, i)
                torch.cuda.synchronize(device)
                t1 = time.perf_counter()
                eng.rebuild()
                torch.cuda.synchronize(device)
                t2 = time.perf_counter()
            timeline.append({"i": i, "before": before, "after": parameters["means"].shape[0],
                             "event_ms": (t1 - t0) * 1e3, "rebuild_ms": (t2 - t1) * 1e3})
            if on_event is not None:
                on_event(i, t1 - t0, parameters["means"].shape[0])
            if max_gaussians is not None and parameters["means"].shape[0] > max_gaussians:
                break
    eng.check()
    eng.adam.sync_state()
    eng.close()
    return parameters, timeline
