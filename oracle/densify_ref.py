"""densify_ref.py -- CPU ORACLE (test infrastructure, NOT product code) for the densification step.

A plain-PyTorch, device-agnostic restatement of the reference's densification
(/root/reference/external.py:113-314): gradient accumulation, clone, split (n = 2), removal of split
parents, prune, opacity reset, with the Adam-moment surgery of cat_params_to_optimizer /
remove_points / update_params_and_optimizer (external.py:127-204) expressed on plain tensors.
Pinned by tests/test_densify.py against tests/golden/densify_*.npz, which were produced by the
UNMODIFIED reference code (tests/golden/make_golden_densify.py).  Only tests/ may import this file.

State layout: tables[name] = [param, exp_avg, exp_avg_sq] for the six per-Gaussian tables
(means, colors, segmentation_masks, rotation_quaternions, opacity_logits, log_scales);
stats = dict(visibility_count, mean_2d_gradients_accumulated, max_2d_radii).
"""
import torch

KEYS = ["means", "colors", "segmentation_masks", "rotation_quaternions", "opacity_logits", "log_scales"]


def build_rotation(q):  # external.py:27-46
    norm = torch.sqrt(q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2] + q[:, 3] * q[:, 3])
    q = q / norm[:, None]
    rot = torch.zeros((q.size(0), 3, 3), dtype=q.dtype, device=q.device)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rot[:, 0, 0] = 1 - 2 * (y * y + z * z)
    rot[:, 0, 1] = 2 * (x * y - r * z)
    rot[:, 0, 2] = 2 * (x * z + r * y)
    rot[:, 1, 0] = 2 * (x * y + r * z)
    rot[:, 1, 1] = 1 - 2 * (x * x + z * z)
    rot[:, 1, 2] = 2 * (y * z - r * x)
    rot[:, 2, 0] = 2 * (x * z - r * y)
    rot[:, 2, 1] = 2 * (y * z + r * x)
    rot[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return rot


def _cat(tables, new):  # cat_params_to_optimizer, external.py:146-173: moments extended with zeros
    for k, v in new.items():
        p, m, s = tables[k]
        tables[k] = [torch.cat((p, v), 0), torch.cat((m, torch.zeros_like(v)), 0), torch.cat((s, torch.zeros_like(v)), 0)]


def _remove(tables, stats, to_remove):  # remove_points, external.py:176-204
    keep = ~to_remove
    for k in KEYS:
        tables[k] = [t[keep] for t in tables[k]]
    for k in ("mean_2d_gradients_accumulated", "visibility_count", "max_2d_radii"):
        stats[k] = stats[k][keep]


def densify_step(tables, stats, visible, means2d_grad, scene_radius, i, samples=None):
    """external.densify_gaussians (external.py:211-314) on plain tensors; `samples` replaces the
    torch.normal draw of external.py:261 (it is N(0, 1) * std when None)."""
    if i > 5000:
        return
    # accumulate_mean_2d_gradients, external.py:113-124
    stats["mean_2d_gradients_accumulated"][visible] += torch.norm(means2d_grad[visible, :2], dim=-1)
    stats["visibility_count"][visible] += 1
    grad_thr = 0.0002
    if i >= 500 and i % 100 == 0:
        avg = stats["mean_2d_gradients_accumulated"] / stats["visibility_count"]
        avg[avg.isnan()] = 0.0
        max_s = torch.exp(tables["log_scales"][0]).max(dim=1).values
        to_clone = (avg >= grad_thr) & (max_s <= 0.01 * scene_radius)
        _cat(tables, {k: tables[k][0][to_clone] for k in KEYS})
        n_pts = tables["means"][0].shape[0]
        padded = torch.zeros(n_pts, dtype=avg.dtype, device=avg.device)
        padded[: avg.shape[0]] = avg
        to_split = (padded >= grad_thr) & (torch.exp(tables["log_scales"][0]).max(dim=1).values > 0.01 * scene_radius)
        n = 2
        new = {k: tables[k][0][to_split].repeat(n, 1) for k in KEYS}
        stds = torch.exp(tables["log_scales"][0])[to_split].repeat(n, 1)
        if samples is None:
            samples = torch.normal(mean=torch.zeros_like(stds), std=stds)
        rots = build_rotation(tables["rotation_quaternions"][0][to_split]).repeat(n, 1, 1)
        new["means"] = new["means"] + torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1)
        new["log_scales"] = torch.log(torch.exp(new["log_scales"]) / (0.8 * n))
        _cat(tables, new)
        n_pts = tables["means"][0].shape[0]
        for k in ("mean_2d_gradients_accumulated", "visibility_count", "max_2d_radii"):
            stats[k] = torch.zeros(n_pts, dtype=avg.dtype, device=avg.device)
        to_remove = torch.cat((to_split, torch.zeros(n * int(to_split.sum()), dtype=torch.bool, device=avg.device)))
        _remove(tables, stats, to_remove)
        thr = 0.25 if i == 5000 else 0.005
        to_remove = (torch.sigmoid(tables["opacity_logits"][0]) < thr).squeeze(-1)
        if i >= 3000:
            to_remove = to_remove | (torch.exp(tables["log_scales"][0]).max(dim=1).values > 0.1 * scene_radius)
        _remove(tables, stats, to_remove)
    if i > 0 and i % 3000 == 0:  # update_params_and_optimizer, external.py:127-143
        p = tables["opacity_logits"][0]
        v = torch.log((torch.ones_like(p) * 0.01) / (1 - torch.ones_like(p) * 0.01))
        tables["opacity_logits"] = [v, torch.zeros_like(v), torch.zeros_like(v)]
