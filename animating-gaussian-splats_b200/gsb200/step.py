"""Camera-batched training step on the hot path (the additional batched entry point of SURVEY.md
section 7/8b): all cameras of a timestep against one replicated Gaussian set --

    raster forward (n_cams) -> fused L1+SSIM loss + dL/dimage -> raster backward (gradients summed over
    cameras) -> densification statistics per camera (densify.py:154-162 + external.py:113-124)

with no per-camera Python work except the statistics launch.  Parity is per render call
(tests/test_raster_gpu.py::test_camera_batch_matches_single).
"""
import ctypes as C

import torch

from . import _lib as L
from . import loss as loss_mod
from . import raster


class DensifyStats:
    """DensificationVariables (shared.py:20-26): visibility_count, mean_2d_gradients_accumulated, max_2d_radii."""

    def __init__(self, P, device):
        z = lambda: torch.zeros(P, dtype=torch.float32, device=device)
        self.visibility_count, self.mean_2d_gradients_accumulated, self.max_2d_radii = z(), z(), z()
        self.gaussian_is_visible_mask = torch.zeros(P, dtype=torch.bool, device=device)

    def accumulate(self, radii, grad_means2D):
        P = radii.numel()
        with L.device_guard(radii.device):
            rc = L.lib().gsb_densify_accumulate(P, L.ptr(radii), L.ptr(grad_means2D), L.ptr(self.max_2d_radii),
                                                L.ptr(self.mean_2d_gradients_accumulated),
                                                L.ptr(self.visibility_count), L.ptr(self.gaussian_is_visible_mask),
                                                L.stream_ptr())
        L.check(rc, "gsb_densify_accumulate")

    def accumulate_batch(self, radii, grad_means2D):
        """radii [n,P], grad_means2D [n,P,3]: all cameras of the batch in one launch."""
        n, P = radii.shape
        with L.device_guard(radii.device):
            rc = L.lib().gsb_densify_accumulate_batch(P, n, L.ptr(radii), L.ptr(grad_means2D), L.ptr(self.max_2d_radii),
                                                      L.ptr(self.mean_2d_gradients_accumulated),
                                                      L.ptr(self.visibility_count),
                                                      L.ptr(self.gaussian_is_visible_mask), L.stream_ptr())
        L.check(rc, "gsb_densify_accumulate_batch")

    def apply_flat(self, flat):
        """Add the statistic columns of a flat step record (after the multi-GPU exchange: the sums / maximum over all
        ranks' cameras): accum += col 14, count += col 15, max_2d_radii = max(., col 16).  One launch."""
        P = flat.shape[0]
        with L.device_guard(flat.device):
            rc = L.lib().gsb_flat_apply_stats(P, L.ptr(flat), int(flat.shape[1]), L.ptr(self.mean_2d_gradients_accumulated),
                                              L.ptr(self.visibility_count), L.ptr(self.max_2d_radii), L.stream_ptr())
        L.check(rc, "gsb_flat_apply_stats")


def timestep_step(act, cams, targets, bg, stats=None, w_l1=0.8, w_ssim=0.2, flat=None, flat_mode=0, arena=None):
    """act: activated parameters (means3D, colors_precomp, opacities, scales, rotations) on the GPU;
    cams: gsb200.scene.stack_cameras(...); targets [n,3,H,W].
    Returns (loss_stats [n,3], grads dict summed over cameras, radii [n,P], RasterState).

    flat ([P, 20], e.g. gsb200.dist.Exchange.flat): the exchange form -- the backward writes the step's whole
    per-Gaussian record (gradients + statistic increments of THIS step, gsb200.dist.COLS) into it and `flat` is
    returned in place of the gradient dict; `stats` is left alone (apply the record with stats.apply_flat after the
    exchange).  arena (gsb200.raster.RasterArena): asynchronous forward -- no host synchronisation or allocation in the
    step, so the whole call can be captured into a CUDA graph (StepGraph)."""
    n = targets.size(0)
    color, depth, radii, st = raster.forward(
        act["means3D"], act["colors_precomp"], act["opacities"], act["scales"], act["rotations"], None, bg, 1.0,
        cams["viewmatrix"], cams["projmatrix"], cams["tanfovx"], cams["tanfovy"], cams["image_height"],
        cams["image_width"], n_cams=n, arena=arena, tanfov_dev=cams.get("tanfov_dev"), entries=cams.get("entries"))
    if n == 1:
        color, radii = color[None], radii[None]
    # cams["entries"] / cams["rows"]: tile-band sharding (gsb200.dist.shard_bands): the entries render bands of their
    # cameras, the loss covers the bands' own pixel rows
    loss_stats, dL = loss_mod.l1_ssim_fused(color, targets, w_l1, w_ssim, grad=True, arena=arena, rows=cams.get("rows"))
    if flat is not None:
        raster.backward_flat(st, dL, flat, mode=flat_mode, scratch=None if arena is None else arena.bwd_scratch)
        return loss_stats, flat, radii, st
    grads = raster.backward(st, dL, need_cov3D_grad=False)
    if stats is not None:
        g2 = grads["means2D"] if n > 1 else grads["means2D"][None]
        stats.accumulate_batch(radii.contiguous(), g2.contiguous())
    return loss_stats, grads, radii, st


class CameraGroups:
    """A rank's cameras as several batches that run CONCURRENTLY (one CUDA stream each): forward -> loss -> per-tile
    backward of every group overlap, so the few long tiles that end one group's blend kernels run beside another
    group's bulk instead of alone on the GPU -- what limits a step when a GPU holds only 3-4 cameras.  The short
    per-Gaussian passes that write the flat record are ordered on the calling stream (group 0 writes, the others
    add).  Every group owns an arena; capture the whole step() into a StepGraph: the fork / join becomes parallel
    branches of the graph."""

    def __init__(self, cams, targets, n_groups, P, device, capacities=None):
        n = targets.shape[0]
        self.dev = torch.device(device)
        n_groups = max(1, min(int(n_groups), n))
        bounds = [round(g * n / n_groups) for g in range(n_groups + 1)]
        self.groups = []
        W, H = cams["image_width"], cams["image_height"]
        for g in range(n_groups):
            ix = list(range(bounds[g], bounds[g + 1]))
            sub = {"viewmatrix": cams["viewmatrix"][ix].contiguous(), "projmatrix": cams["projmatrix"][ix].contiguous(),
                   "tanfovx": [cams["tanfovx"][i] for i in ix], "tanfovy": [cams["tanfovy"][i] for i in ix],
                   "image_width": W, "image_height": H}
            sub["tanfov_dev"] = raster.tanfov_table(sub["tanfovx"], sub["tanfovy"], len(ix), self.dev)
            self.groups.append({"ix": ix, "cams": sub, "targets": targets[ix].contiguous(), "arena": raster.RasterArena(self.dev),
                                "stream": torch.cuda.Stream(device=self.dev) if g > 0 else None})
        self.P = P
        self.capacities = capacities

    def reserve(self, act, bg, slack=1.25):
        """Size every group's arena from one blocking render of its cameras."""
        for g in self.groups:
            n = len(g["ix"])
            c = g["cams"]
            _, _, _, st = raster.forward(act["means3D"], act["colors_precomp"], act["opacities"], act["scales"], act["rotations"],
                                         None, bg, 1.0, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"],
                                         c["image_height"], c["image_width"], n_cams=n)
            g["arena"].reserve(self.P, n, c["image_width"], c["image_height"], int(slack * st.num_rendered) + 4096)
        return self

    def step(self, act, bg, flat, w_l1=0.8, w_ssim=0.2):
        """-> (list of per-group loss numbers [n_g, 3], flat, list of RasterStates)."""
        cur = torch.cuda.current_stream(self.dev)
        outs = []
        for g in self.groups:
            s = g["stream"]
            if s is not None:
                s.wait_stream(cur)
            with torch.cuda.stream(s if s is not None else cur):
                n = len(g["ix"])
                c = g["cams"]
                color, _, radii, st = raster.forward(act["means3D"], act["colors_precomp"], act["opacities"], act["scales"],
                                                     act["rotations"], None, bg, 1.0, c["viewmatrix"], c["projmatrix"],
                                                     c["tanfovx"], c["tanfovy"], c["image_height"], c["image_width"], n_cams=n,
                                                     arena=g["arena"], tanfov_dev=c["tanfov_dev"])
                if n == 1:
                    color = color[None]
                loss_stats, dL = loss_mod.l1_ssim_fused(color, g["targets"], w_l1, w_ssim, grad=True, arena=g["arena"])
                raster.backward_flat(st, dL, flat, mode=0, scratch=g["arena"].bwd_scratch, part="blend")
                outs.append((loss_stats, st))
        for k, g in enumerate(self.groups):   # join, then the short passes that touch flat, in order
            if g["stream"] is not None:
                cur.wait_stream(g["stream"])
        for k, g in enumerate(self.groups):
            raster.backward_flat(outs[k][1], None, flat, mode=0 if k == 0 else 2, scratch=g["arena"].bwd_scratch, part="finish")
        return [o[0] for o in outs], flat, [o[1] for o in outs]


class StepGraph:
    """A step captured once into a CUDA graph and replayed: `fn()` must issue only stream work on the current stream
    (asynchronous forward through a RasterArena, fused loss, flat backward, exchange kernel, statistics, optimiser),
    no host synchronisation and no host -> device copies of pageable memory.  The ~30 launches of a step then cost one
    graph launch on the host and back-to-back kernel starts on the device -- what is left of the per-step constant
    when a GPU holds 3-4 cameras.  Outputs are whatever tensors `fn` returns (static addresses: the arena's outputs,
    the flat buffer); inputs are updated in place (parameters by the optimiser, cameras / targets by copy_)."""

    def __init__(self, fn, warmup=2, device=None):
        self.fn = fn
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):   # first calls set kernel attributes, size pinned staging, fill caches
                fn()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn()

    def replay(self):
        self.graph.replay()
        return self.out


class HostFeed:
    """End-to-end front of timestep_step for callers whose step inputs live in HOST memory (raw parameter
    tables and target images in pinned buffers): every step copies its inputs host -> device, runs the
    activations (create_render_arguments, shared.py:29-42) and the step, and reads the per-camera loss
    numbers back.  The copies of step k+1 are issued on a side stream into the other half of a double
    buffer while step k computes, so the PCIe transfer (91 MB at 300 k Gaussians x 27 cameras) hides
    behind the rasterizer instead of preceding it."""

    RAW_KEYS = ("means", "colors", "rotation_quaternions", "opacity_logits", "log_scales")

    def __init__(self, host_raw, host_targets, device):
        self.dev = device
        self.host_raw = {k: host_raw[k] for k in self.RAW_KEYS}
        self.host_targets = host_targets
        for t in list(self.host_raw.values()) + [host_targets]:
            if not t.is_pinned():
                raise RuntimeError("HostFeed needs pinned host buffers")
        self.bufs = [({k: torch.empty_like(v, device=device) for k, v in self.host_raw.items()},
                      torch.empty_like(host_targets, device=device)) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=device)
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.loss_hosts = [torch.empty((host_targets.size(0), 3), dtype=torch.float32).pin_memory() for _ in range(2)]
        self.loss_host = self.loss_hosts[0]
        self.done = [torch.cuda.Event(), torch.cuda.Event()]   # step k's kernels and loss read-back finished
        self.graphs, self.act_out = [None, None], [None, None]
        self.k = 0
        self.h2d_bytes = sum(v.numel() * v.element_size() for v in self.host_raw.values()) + \
            host_targets.numel() * host_targets.element_size()
        self.d2h_bytes = self.loss_host.numel() * 4
        self._prefetch(0)

    def _prefetch(self, k):
        raw, tgt = self.bufs[k % 2]
        # the buffer's previous consumer is step k-2: the copy stream waits for its `done` event
        if k >= 2:
            self.copy_stream.wait_event(self.done[k % 2])
        with torch.cuda.stream(self.copy_stream):
            for key, h in self.host_raw.items():
                raw[key].copy_(h, non_blocking=True)
            tgt.copy_(self.host_targets, non_blocking=True)
            self.ready[k % 2].record(self.copy_stream)

    def step(self, cams, bg, stats=None, after=None, wait=True, flat=None, arena=None, graph=False):
        """One end-to-end step.  flat / arena: the exchange form and the host-free forward of timestep_step (`after`
        then receives the flat record in place of the gradient dict).  graph=True (needs flat and arena): activations
        + step + `after` are captured once per buffer half and replayed, so the host issues two copies, one graph
        launch and the loss read-back per step; cams / bg / after must then be the same objects every step.  `after(loss_stats, grads, radii, state)` runs on the compute stream before the
        loss read-back (e.g. the multi-GPU exchange).  Returns the pinned [n,3] loss buffer (l1, ssim, mix) of this
        step.  wait=True blocks until it is filled.  wait=False returns at once -- the way a training loop that only
        logs the loss calls it: the host goes on to issue the next step while this one's backward still runs; the
        buffer is valid after wait() and until the step after next is issued (two buffers alternate)."""
        from . import optim
        k = self.k
        raw, tgt = self.bufs[k % 2]
        cur = torch.cuda.current_stream(self.dev)
        if graph and (flat is None or arena is None):
            raise RuntimeError("HostFeed.step(graph=True) needs the flat record and an arena (host-free step)")
        if graph and self.graphs[k % 2] is None:
            if self.act_out[k % 2] is None:
                self.act_out[k % 2] = (torch.empty_like(raw["rotation_quaternions"]), torch.empty_like(raw["opacity_logits"]),
                                       torch.empty_like(raw["log_scales"]))
            out_a = self.act_out[k % 2]

            def body():
                rot, opa, sca = optim.activate_into(raw, out_a)
                act = {"means3D": raw["means"], "colors_precomp": raw["colors"], "rotations": rot, "opacities": opa, "scales": sca}
                r = timestep_step(act, cams, tgt, bg, stats, flat=flat, arena=arena)
                if after is not None:
                    after(*r)
                return r

            cur.wait_event(self.ready[k % 2])   # the half's first copy must have landed before the warm-up runs read it
            self.graphs[k % 2] = StepGraph(body, device=self.dev)
        cur.wait_event(self.ready[k % 2])
        self._prefetch(k + 1)
        if graph:
            loss_stats, grads, radii, st = self.graphs[k % 2].replay()
        else:
            rot, opa, sca = optim.activate(raw)   # one launch (gsb_activate)
            act = {"means3D": raw["means"], "colors_precomp": raw["colors"], "rotations": rot, "opacities": opa, "scales": sca}
            loss_stats, grads, radii, st = timestep_step(act, cams, tgt, bg, stats, flat=flat, arena=arena)
            if after is not None:
                after(loss_stats, grads, radii, st)
        out = self.loss_hosts[k % 2]
        out.copy_(loss_stats, non_blocking=True)
        self.done[k % 2].record(cur)
        self.loss_host = out
        self.k += 1
        if wait:
            self.done[k % 2].synchronize()
        return out

    def wait(self):
        """Block until the last issued step (kernels and loss read-back) has finished; returns its loss buffer."""
        if self.k > 0:
            self.done[(self.k - 1) % 2].synchronize()
        return self.loss_host
