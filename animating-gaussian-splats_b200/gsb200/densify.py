"""Densification on the GPU (replaces external.py:113-314 of the reference).

Drop-in names: DensificationVariables (shared.py:20-26), update_max_2d_radii_and_visibility_mask
(densify.py:154-162), accumulate_mean_2d_gradients (external.py:113-124), densify_gaussians
(external.py:211-314, same arguments: parameter dict, DensificationVariables, scene radius, torch Adam
optimizer, iteration).  The row surgery the reference performs with boolean-mask indexing and torch.cat
over every parameter and Adam-moment tensor is done here by a mask scan (exclusive rank, one launch
chain per mask) and order-preserving row gathers through the C ABI; the split transform is one kernel.
The thresholds are cheap elementwise torch expressions; the only host syncs are the row counts that
size the new tensors (the reference syncs for the same reason inside its mask indexing).
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


class _Mask:
    """A row mask with its exclusive rank and count (gsb_mask_scan)."""

    def __init__(self, mask):
        self.mask = mask.contiguous()
        P = self.mask.numel()
        dev = mask.device
        self.rank = torch.empty(P, dtype=torch.int32, device=dev)
        cnt = torch.zeros(1, dtype=torch.int32, device=dev)
        lib = L.lib()
        scratch = torch.empty(lib.gsb_scan_scratch_bytes(max(P, 1)), dtype=torch.uint8, device=dev)
        with L.device_guard(dev):
            rc = lib.gsb_mask_scan(P, L.ptr(self.mask), L.ptr(self.rank), L.ptr(cnt), L.ptr(scratch), _stream())
        L.check(rc, "gsb_mask_scan")
        self.count = int(cnt.item())   # sizes the new tensors: the one host sync per mask

    def gather_into(self, src, dst):
        """dst[rank[i]] = src[i] for selected rows; dst is a [count, row] view."""
        if self.count == 0:
            return
        src = src.contiguous()
        row = src[0].numel()
        with L.device_guard(src.device):
            rc = L.lib().gsb_gather_rows(src.shape[0], L.ptr(self.mask), L.ptr(self.rank), L.ptr(src), L.ptr(dst),
                                         row, _stream())
        L.check(rc, "gsb_gather_rows")


def _append_selected(t, m, repeat=1):
    """torch.cat((t, t[mask].repeat(repeat, 1))) without the intermediate tensors."""
    N = t.shape[0]
    out = torch.empty((N + repeat * m.count,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    out[:N].copy_(t)
    for r in range(repeat):
        m.gather_into(t, out[N + r * m.count: N + (r + 1) * m.count])
    return out


def _append_zeros(t, n):
    out = torch.zeros((t.shape[0] + n,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    out[: t.shape[0]].copy_(t)
    return out


def _compact(t, keep):
    out = torch.empty((keep.count,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    keep.gather_into(t, out)
    return out


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
    dev = visible.device
    grad_thr = 0.0002
    if i >= 500 and i % 100 == 0:
        avg = stats["mean_2d_gradients_accumulated"] / stats["visibility_count"]
        avg = torch.where(avg.isnan(), torch.zeros_like(avg), avg)
        max_s = torch.exp(tables["log_scales"][0]).max(dim=1).values
        hot = avg >= grad_thr
        # clone (external.py:232-238): appended after the existing rows, Adam moments extended with zeros
        clone = _Mask(hot & (max_s <= 0.01 * scene_radius))
        N0 = avg.shape[0]
        # split candidates are evaluated on the N0 + C rows; clones carry padded gradient 0 (external.py:241-250)
        split_mask = torch.zeros(N0 + clone.count, dtype=torch.bool, device=dev)
        split_mask[:N0] = hot & (max_s > 0.01 * scene_radius)
        for k in KEYS:
            p, m, v = tables[k]
            tables[k] = [_append_selected(p, clone), _append_zeros(m, clone.count), _append_zeros(v, clone.count)]
        split = _Mask(split_mask)
        S = split.count
        N1 = N0 + clone.count
        if samples is None:
            stds = torch.exp(tables["log_scales"][0])[split_mask].repeat(2, 1)
            samples = torch.normal(mean=torch.zeros_like(stds), std=stds)
        samples = samples.contiguous().float()
        new = {}
        for k in KEYS:
            new[k] = _append_selected(tables[k][0], split, repeat=2)
        if S > 0:
            with L.device_guard(dev):
                rc = L.lib().gsb_split_rows(N1, S, L.ptr(split.mask), L.ptr(split.rank), L.ptr(tables["means"][0]),
                                            L.ptr(tables["rotation_quaternions"][0].contiguous()),
                                            L.ptr(tables["log_scales"][0]), L.ptr(samples),
                                            L.ptr(new["means"][N1:]), L.ptr(new["log_scales"][N1:]), _stream())
            L.check(rc, "gsb_split_rows")
        for k in KEYS:
            tables[k] = [new[k], _append_zeros(tables[k][1], 2 * S), _append_zeros(tables[k][2], 2 * S)]
        n_pts = N1 + 2 * S
        # remove the split parents (external.py:280-288); statistics restart from zero (external.py:273-279)
        keep = torch.ones(n_pts, dtype=torch.bool, device=dev)
        keep[:N1] = ~split_mask
        keep = _Mask(keep)
        for k in KEYS:
            tables[k] = [_compact(t, keep) for t in tables[k]]
        for k in _STATS:
            stats[k] = torch.zeros(keep.count, dtype=torch.float32, device=dev)
        # prune (external.py:290-302)
        thr = 0.25 if i == 5000 else 0.005
        to_remove = (torch.sigmoid(tables["opacity_logits"][0]) < thr).squeeze(-1)
        if i >= 3000:
            to_remove = to_remove | (torch.exp(tables["log_scales"][0]).max(dim=1).values > 0.1 * scene_radius)
        keep = _Mask(~to_remove)
        for k in KEYS:
            tables[k] = [_compact(t, keep) for t in tables[k]]
        for k in _STATS:
            stats[k] = torch.zeros(keep.count, dtype=torch.float32, device=dev)
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
