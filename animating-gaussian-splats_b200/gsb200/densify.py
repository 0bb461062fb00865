"""Densification on the GPU (replaces external.py:113-314 of the reference).

Drop-in names: DensificationVariables (shared.py:20-26), update_max_2d_radii_and_visibility_mask
(densify.py:154-162), accumulate_mean_2d_gradients (external.py:113-124), densify_gaussians
(external.py:211-314, same arguments: parameter dict, DensificationVariables, scene radius, torch Adam
optimizer, iteration).  The row surgery the reference performs with boolean-mask indexing and torch.cat
over every parameter and Adam-moment tensor -- four re-allocations of 18 tables per event -- is ONE pass here:
every final row is a function of one source row, so a plan kernel decides per source row which of its up to four
output rows survive, four scans rank them, and one apply kernel writes all 18 new tables (gsb_densify_plan /
gsb_densify_apply).  The one host synchronisation per event is the read-back of four row counts that size the new
tensors (the reference synchronises inside every mask indexing).
"""
import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib as L

KEYS = ["means", "colors", "segmentation_masks", "rotation_quaternions", "opacity_logits", "log_scales"]
_STATS = ("mean_2d_gradients_accumulated", "visibility_count", "max_2d_radii")


@dataclass
class DensificationVariables:  # shared.py:20-26
    visibility_count: torch.Tensor
    mean_2d_gradients_accumulated: torch.Tensor
    max_2d_radii: torch.Tensor
    gaussian_is_visible_mask: torch.Tensor = None
    means_2d: torch.Tensor = None


def _stream():
    return L.stream_ptr()


def update_max_2d_radii_and_visibility_mask(radii, densification_variables):
    """densify.py:154-162 as tensor expressions without a nonzero() sync."""
    dv = densification_variables
    pos = radii > 0
    dv.max_2d_radii = torch.where(pos, torch.maximum(radii.to(dv.max_2d_radii.dtype), dv.max_2d_radii),
                                  dv.max_2d_radii)
    dv.gaussian_is_visible_mask = pos


def accumulate_mean_2d_gradients(densification_variables):
    """external.py:113-124: accum[vis] += ||grad[vis, :2]||, count[vis] += 1 (one kernel)."""
    dv = densification_variables
    grad = dv.means_2d.grad.contiguous()
    vis = dv.gaussian_is_visible_mask.contiguous()
    P = vis.numel()
    with L.device_guard(vis.device):
        rc = L.lib().gsb_densify_accumulate_masked(P, L.ptr(vis), L.ptr(grad), L.ptr(dv.mean_2d_gradients_accumulated),
                                                   L.ptr(dv.visibility_count), _stream())
    L.check(rc, "gsb_densify_accumulate_masked")


def _fused_event(tables, stats, scene_radius, i, samples):
    """Clone / split / remove parents / prune as one plan + one apply launch chain over all 18 tables
    (gsb_densify_plan / gsb_densify_apply); one read-back of four row counts sizes the new tables."""
    lib = L.lib()
    p0 = tables["means"][0]
    dev, P = p0.device, p0.shape[0]
    src = []
    for j in range(3):
        for k in KEYS:
            t = tables[k][j]
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.float().contiguous()
                tables[k][j] = t
            src.append(t)
    scratch = torch.empty(lib.gsb_densify_plan_scratch_bytes(P), dtype=torch.uint8, device=dev)
    totals = torch.zeros(4, dtype=torch.int32, device=dev)
    thr = 0.25 if i == 5000 else 0.005                      # external.py:290-296
    with L.device_guard(dev):
        rc = lib.gsb_densify_plan(P, L.ptr(stats["mean_2d_gradients_accumulated"].contiguous()),
                                  L.ptr(stats["visibility_count"].contiguous()), L.ptr(src[4]), L.ptr(src[5]), 0.0002,
                                  0.01 * scene_radius, thr, 1 if i >= 3000 else 0, 0.1 * scene_radius, L.ptr(scratch),
                                  L.ptr(totals), _stream())
    L.check(rc, "gsb_densify_plan")
    K0, K1, K2, S = (int(x) for x in totals.cpu())          # the event's one host synchronisation
    N = K0 + K1 + 2 * K2
    if S > 0:
        if samples is None:   # the reference's draw (external.py:258-261): same shapes, same generator consumption
            flags = scratch[:P]
            stds = torch.exp(src[5])[(flags & 8) != 0].repeat(2, 1)
            samples = torch.normal(mean=torch.zeros_like(stds), std=stds)
        samples = samples.contiguous().float()
    dst = [torch.empty((N,) + tuple(t.shape[1:]), dtype=torch.float32, device=dev) for t in src]
    arr_t = C.c_void_p * 18
    hs = (C.c_int32 * 4)(K0, K1, K2, S)
    with L.device_guard(dev):
        rc = lib.gsb_densify_apply(P, L.ptr(scratch), hs, arr_t(*[t.data_ptr() if t.numel() else None for t in src]),
                                   arr_t(*[t.data_ptr() if t.numel() else None for t in dst]), L.ptr(samples) if S > 0 else None,
                                   _stream()) if (P > 0 and N > 0) else 0
    L.check(rc, "gsb_densify_apply")
    for j in range(3):
        for n_, k in enumerate(KEYS):
            tables[k][j] = dst[6 * j + n_]
    for k in _STATS:
        stats[k] = torch.zeros(N, dtype=torch.float32, device=dev)


def densify_step(tables, stats, visible, means2d_grad, scene_radius, i, samples=None):
    """The functional core on plain tensors: tables[name] = [param, exp_avg, exp_avg_sq],
    stats = dict(visibility_count, mean_2d_gradients_accumulated, max_2d_radii).  In place on the dicts.
    `samples` ([2S, 3]) replaces the torch.normal draw of external.py:261."""
    if i > 5000:
        return
    dv = DensificationVariables(stats["visibility_count"], stats["mean_2d_gradients_accumulated"],
                                stats["max_2d_radii"], visible)
    dv.means_2d = type("G", (), {"grad": means2d_grad})()
    accumulate_mean_2d_gradients(dv)
    if i >= 500 and i % 100 == 0:
        _fused_event(tables, stats, scene_radius, i, samples)
    if i > 0 and i % 3000 == 0:  # opacity reset (external.py:306-314): logit(0.01), moments zeroed
        p = tables["opacity_logits"][0]
        v = torch.log((torch.ones_like(p) * 0.01) / (1 - torch.ones_like(p) * 0.01))
        tables["opacity_logits"] = [v, torch.zeros_like(v), torch.zeros_like(v)]


def densify_gaussians(gaussian_cloud_parameters, densification_variables, scene_radius, optimizer, i, samples=None):
    """external.densify_gaussians (external.py:211-314) with the same argument list (+ optional `samples`).
    Rebuilds the nn.Parameters and the Adam state of the six per-Gaussian tables in place in the dict /
    optimizer; camera_matrices / camera_center are left alone (external.py:180,237,255)."""
    if i > 5000:
        return
    params, dv = gaussian_cloud_parameters, densification_variables
    groups = {g["name"]: g for g in optimizer.param_groups}
    tables, states = {}, {}
    for k in KEYS:
        p = groups[k]["params"][0]
        st = optimizer.state.get(p, None)
        states[k] = st
        if st is not None and "exp_avg" in st:
            tables[k] = [p.data, st["exp_avg"], st["exp_avg_sq"]]
        else:
            tables[k] = [p.data, torch.zeros_like(p.data), torch.zeros_like(p.data)]
    stats = {k: getattr(dv, k) for k in _STATS}
    before = {k: tables[k][0] for k in KEYS}
    densify_step(tables, stats, dv.gaussian_is_visible_mask, dv.means_2d.grad, scene_radius, i, samples)
    for k in _STATS:
        setattr(dv, k, stats[k])
    for k in KEYS:
        if tables[k][0] is before[k]:
            continue
        old = groups[k]["params"][0]
        st = states[k]
        new_p = torch.nn.Parameter(tables[k][0].requires_grad_(True))
        if st is not None:
            del optimizer.state[old]
            st["exp_avg"], st["exp_avg_sq"] = tables[k][1], tables[k][2]
            optimizer.state[new_p] = st
        groups[k]["params"][0] = new_p
        params[k] = new_p
