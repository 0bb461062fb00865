"""Fused activations and optimiser step: the two steps either side of the
rasterizer in every iteration of the reference's loops.

  create_render_arguments(params)   shared.py:29-42 -- same dict (means3D, colors_precomp, rotations, opacities,
                                    scales, means2D carrier), but the three activations and their autograd are one
                                    kernel each (gsb_activate / gsb_activate_backward).
  FusedAdam                         torch.optim.Adam as create_optimizer configures it (densify.py:68-86): same
                                    constructor, param_groups and per-parameter state (step, exp_avg, exp_avg_sq),
                                    so densify_gaussians' optimiser surgery works unchanged; step() is ONE launch
                                    over all parameter tables (gsb_adam_step).
Both are CUDA-only like the rest of gsb200.
"""
import ctypes as C

import torch

from . import _lib as L


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, quats, logits, log_scales):
        L.require_cuda(quats, "rotation_quaternions")
        q, lg, ls = quats.contiguous().float(), logits.contiguous().float(), log_scales.contiguous().float()
        P = q.size(0)
        if q.shape != (P, 4) or lg.numel() != P or ls.shape != (P, 3):
            raise RuntimeError("expected rotation_quaternions [P,4], opacity_logits [P,1], log_scales [P,3]")
        rot, opa, sca = torch.empty_like(q), torch.empty_like(lg), torch.empty_like(ls)
        with L.device_guard(q.device):
            rc = L.lib().gsb_activate(P, L.ptr(q), L.ptr(lg), L.ptr(ls), L.ptr(rot), L.ptr(opa), L.ptr(sca), L.stream_ptr())
        L.check(rc, "gsb_activate")
        ctx.save_for_backward(q, opa, sca)
        return rot, opa, sca

    @staticmethod
    def backward(ctx, d_rot, d_opa, d_sca):
        q, opa, sca = ctx.saved_tensors
        P = q.size(0)
        need = ctx.needs_input_grad
        c = lambda t: None if t is None else t.contiguous().float()
        d_rot, d_opa, d_sca = c(d_rot), c(d_opa), c(d_sca)
        gq = torch.empty_like(q) if need[0] else None
        gl = torch.empty_like(opa) if need[1] else None
        gs = torch.empty_like(sca) if need[2] else None
        with L.device_guard(q.device):
            rc = L.lib().gsb_activate_backward(P, L.ptr(q), L.ptr(opa), L.ptr(sca), L.ptr(d_rot), L.ptr(d_opa), L.ptr(d_sca),
                                               L.ptr(gq), L.ptr(gl), L.ptr(gs), L.stream_ptr())
        L.check(rc, "gsb_activate_backward")
        return gq, gl, gs


def activate(params):
    """(rotations, opacities, scales) of a raw parameter dict, differentiable.  The outputs remember which raw
    tensors (object + version counter) they were computed from, so gsb200.raster can recognise two activations of
    the same parameters as the same geometry (see "Geometry reuse" there)."""
    q, lg, ls = params["rotation_quaternions"], params["opacity_logits"], params["log_scales"]
    out = _Activate.apply(q, lg, ls)
    rot, opa, sca = out
    rot._gsb_src = (("rot", 0), (q, q._version))
    opa._gsb_src = (("opa", 0), (lg, lg._version))
    sca._gsb_src = (("sca", 0), (ls, ls._version))
    return out


def create_render_arguments(gaussian_cloud_parameters):
    """shared.py:29-42.  The means2D carrier (zeros + 0, requires grad) is kept: callers read its .grad."""
    p = gaussian_cloud_parameters
    rot, opa, sca = activate(p)
    return {
        "means3D": p["means"],
        "colors_precomp": p["colors"],
        "rotations": rot,
        "opacities": opa,
        "scales": sca,
        "means2D": torch.zeros_like(p["means"], requires_grad=True) + 0,
    }


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam(params, lr, betas, eps) without weight decay / amsgrad / maximize, stepping every
    parameter table in one kernel launch.  State layout equals torch's (state[p] = {step, exp_avg, exp_avg_sq})."""

    MAX_GROUPS = 16

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1):
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        batches = {}   # (device, betas, eps) -> list of (param, grad, state, lr)
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                L.require_cuda(p, "parameter")
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam needs contiguous float32 parameters")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] = st["step"] + 1 if torch.is_tensor(st["step"]) else st["step"] + 1
                key = (p.device, tuple(group["betas"]), float(group["eps"]))
                batches.setdefault(key, []).append((p, p.grad.contiguous(), st, float(group["lr"])))
        lib = L.lib()
        for (dev, betas, eps), items in batches.items():
            for k in range(0, len(items), self.MAX_GROUPS):
                chunk = items[k:k + self.MAX_GROUPS]
                arr = (L.AdamGroup * len(chunk))()
                for a, (p, g, st, lr) in zip(arr, chunk):
                    a.param, a.grad = p.data_ptr(), g.data_ptr()
                    a.exp_avg, a.exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                    a.numel, a.step, a.lr = p.numel(), int(st["step"]), lr
                with L.device_guard(dev):
                    rc = lib.gsb_adam_step(len(chunk), arr, betas[0], betas[1], eps, L.stream_ptr())
                L.check(rc, "gsb_adam_step")
                for p, _, st, _ in chunk:   # the kernel wrote through raw pointers: tell torch the contents changed
                    torch.autograd.graph.increment_version(p)
                    torch.autograd.graph.increment_version(st["exp_avg"])
                    torch.autograd.graph.increment_version(st["exp_avg_sq"])
        return loss


def activate_into(params, out):
    """The activations without autograd, into caller-owned tensors out = (rotations, opacities, scales) -- the form
    a captured step uses (static addresses, one launch)."""
    q, lg, ls = params["rotation_quaternions"].data, params["opacity_logits"].data, params["log_scales"].data
    rot, opa, sca = out
    with L.device_guard(q.device):
        rc = L.lib().gsb_activate(q.size(0), L.ptr(q), L.ptr(lg), L.ptr(ls), L.ptr(rot), L.ptr(opa), L.ptr(sca), L.stream_ptr())
    L.check(rc, "gsb_activate")
    return out


class FlatAdam:
    """The optimiser step of the first-frame loop straight from the flat step record (gsb_adam_flat): activation
    adjoint + Adam over the six per-Gaussian tables in one launch, step count on the device, so the call can sit
    inside a captured step.  It works ON an optimizer built by create_optimizer (FusedAdam or torch.optim.Adam):
    parameters, learning rates, betas / eps and the exp_avg / exp_avg_sq tensors are that optimizer's, so the densify
    surgery of external.py:127-204 and a later optimizer.step() see a consistent state after sync_state()."""

    TABLES = ("means", "colors", "segmentation_masks", "rotation_quaternions", "opacity_logits", "log_scales")

    def __init__(self, optimizer):
        groups = {g["name"]: g for g in optimizer.param_groups}
        self.optimizer = optimizer
        self.args = L.FlatAdamArgs()
        self.keep = []
        steps = set()
        first = groups[self.TABLES[0]]
        self.betas, self.eps = tuple(first["betas"]), float(first["eps"])
        for k, name in enumerate(self.TABLES):
            g = groups[name]
            p = g["params"][0]
            L.require_cuda(p, name)
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("FlatAdam needs contiguous float32 parameters")
            if tuple(g["betas"]) != self.betas or float(g["eps"]) != self.eps:
                raise RuntimeError("FlatAdam needs one (betas, eps) for all tables")
            st = optimizer.state[p]
            if len(st) == 0:
                st["step"] = torch.tensor(0.0)
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            steps.add(int(st["step"]))
            self.args.param[k], self.args.exp_avg[k], self.args.exp_avg_sq[k] = p.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
            self.args.lr[k] = float(g["lr"])
            self.keep += [p, st["exp_avg"], st["exp_avg_sq"]]
        if len(steps) != 1:
            raise RuntimeError(f"FlatAdam needs one step count for all tables, got {sorted(steps)}")
        self.P = groups["means"]["params"][0].shape[0]
        self.args.P = self.P
        self.device = groups["means"]["params"][0].device
        self.step_dev = torch.full((1,), steps.pop(), dtype=torch.int64, device=self.device)

    def step(self, flat):
        with L.device_guard(self.device):
            rc = L.lib().gsb_adam_flat(C.byref(self.args), L.ptr(flat), int(flat.shape[1]), L.ptr(self.step_dev), self.betas[0],
                                       self.betas[1], self.eps, L.stream_ptr())
        L.check(rc, "gsb_adam_flat")

    def sync_state(self):
        """Publish the device step count to the optimizer's per-parameter state (blocks) and bump the tensors' torch
        version counters (the kernel wrote them through raw pointers)."""
        n = int(self.step_dev.item())
        for t in self.keep:
            torch.autograd.graph.increment_version(t)
        groups = {g["name"]: g for g in self.optimizer.param_groups}
        for name in self.TABLES:
            self.optimizer.state[groups[name]["params"][0]]["step"] = torch.tensor(float(n))
        return n


def create_optimizer(parameters, scene_radius):
    """densify.py:68-86 with the fused step."""
    learning_rates = {"means": 0.00016 * scene_radius, "colors": 0.0025, "segmentation_masks": 0.0,
                      "rotation_quaternions": 0.001, "opacity_logits": 0.05, "log_scales": 0.001,
                      "camera_matrices": 1e-4, "camera_center": 1e-4}
    return FusedAdam([{"params": [value], "name": name, "lr": learning_rates[name]} for name, value in parameters.items()],
                     lr=0.0, eps=1e-15)
