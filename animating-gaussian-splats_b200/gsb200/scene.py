"""Synthetic scenes and cameras with the reference's conventions (no dataset ships with either repo).

Cameras follow create_render_settings (shared.py:64-124): viewmatrix = w2c^T, OpenGL-style projection
with near 1 / far 100, projmatrix = w2c^T @ proj^T, tanfov = dim / (2 f), campos = inv(w2c)[:3, 3],
bg = 0, scale_modifier = 1, prefiltered = False.  Gaussians follow initialize_parameters
(densify.py:22-57): log_scales = log(sqrt(mean squared distance to the 3 nearest neighbours)), colours in
[0, 1], opacity logits around 0, raw (unnormalised) quaternions; activations as create_render_arguments
(shared.py:29-42).  Everything is seeded and built on the CPU so the same bytes reach every
implementation under test.
"""
import math

import numpy as np
import torch


def make_gaussians(P, seed=0, extent=1.5, scale_override=None):
    """Raw (pre-activation) parameter dict, CPU float32 tensors, the 8-entry layout of densify.py:29-57
    minus the camera matrices."""
    rng = np.random.default_rng(seed)
    means = rng.uniform(-extent, extent, size=(P, 3)).astype(np.float32)
    if scale_override is not None:
        log_scales = np.full((P, 1), math.log(scale_override), dtype=np.float32)
    else:
        try:
            from scipy.spatial import cKDTree
            d, _ = cKDTree(means).query(means, k=4)
            msd = np.clip((d[:, 1:] ** 2).mean(-1), 1e-7, None)
        except Exception:  # scipy missing: use the expected value for a uniform cloud
            msd = np.full((P,), (0.75 * (2 * extent) ** 3 / (math.pi * P)) ** (2.0 / 3.0))
        log_scales = np.log(np.sqrt(msd))[:, None].astype(np.float32)
    log_scales = np.tile(log_scales, (1, 3)) + rng.normal(0, 0.3, size=(P, 3)).astype(np.float32)
    quats = rng.normal(0, 1, size=(P, 4)).astype(np.float32)
    return {
        "means": torch.from_numpy(means),
        "colors": torch.from_numpy(rng.uniform(0, 1, size=(P, 3)).astype(np.float32)),
        "segmentation_masks": torch.from_numpy(
            np.eye(3, dtype=np.float32)[rng.integers(0, 2, size=P)]),  # one-hot fg/bg like densify.py:41-44
        "rotation_quaternions": torch.from_numpy(quats),
        "opacity_logits": torch.from_numpy(rng.normal(0, 1.5, size=(P, 1)).astype(np.float32)),
        "log_scales": torch.from_numpy(log_scales.astype(np.float32)),
    }


def activate(params):
    """create_render_arguments (shared.py:29-42) without the means2D carrier."""
    return {
        "means3D": params["means"],
        "colors_precomp": params["colors"],
        "rotations": torch.nn.functional.normalize(params["rotation_quaternions"]),
        "opacities": torch.sigmoid(params["opacity_logits"]),
        "scales": torch.exp(params["log_scales"]),
    }


def look_at_w2c(eye, target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0)):
    eye = np.asarray(eye, np.float64)
    f = np.asarray(target, np.float64) - eye
    f /= np.linalg.norm(f)
    r = np.cross(f, np.asarray(up, np.float64))
    r /= np.linalg.norm(r)
    d = np.cross(f, r)  # camera y points down (OpenCV convention, as the Panoptic w2c matrices)
    R = np.stack([r, d, f])
    w2c = np.eye(4)
    w2c[:3, :3] = R
    w2c[:3, 3] = -R @ eye
    return w2c


def make_camera(W, H, w2c, fx=None, fy=None, near=1.0, far=100.0):
    """One camera in the reference's packing; CPU float32 tensors."""
    fx = 0.7 * W if fx is None else fx
    fy = fx if fy is None else fy
    cx, cy = W / 2.0, H / 2.0
    w2c_t = torch.tensor(np.asarray(w2c), dtype=torch.float32)
    campos = torch.inverse(w2c_t)[:3, 3].contiguous()
    view = w2c_t.unsqueeze(0).transpose(1, 2).contiguous()
    proj = torch.tensor([[2 * fx / W, 0.0, -(W - 2 * cx) / W, 0.0],
                         [0.0, 2 * fy / H, -(H - 2 * cy) / H, 0.0],
                         [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                         [0.0, 0.0, 1.0, 0.0]], dtype=torch.float32).unsqueeze(0).transpose(1, 2)
    return {
        "image_width": W, "image_height": H,
        "tanfovx": W / (2 * fx), "tanfovy": H / (2 * fy),
        "viewmatrix": view, "projmatrix": view.bmm(proj).contiguous(), "campos": campos,
    }


def make_cameras(n, W, H, radius=3.0, seed=0):
    """n cameras on a dome of the given radius looking at the origin."""
    rng = np.random.default_rng(1000 + seed)
    cams = []
    for i in range(n):
        az = 2 * math.pi * (i + 0.25 * rng.uniform()) / n
        el = math.radians(15 + 40 * ((i * 7) % 5) / 4.0)
        eye = (radius * math.cos(el) * math.cos(az), radius * math.cos(el) * math.sin(az), radius * math.sin(el))
        cams.append(make_camera(W, H, look_at_w2c(eye)))
    return cams


def settings_for(cam, device, bg=None):
    """GaussianRasterizationSettings for a camera dict (shared.py:112-124)."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=cam["image_height"], image_width=cam["image_width"], tanfovx=cam["tanfovx"],
        tanfovy=cam["tanfovy"],
        bg=torch.zeros(3, dtype=torch.float32, device=device) if bg is None else bg.to(device),
        scale_modifier=1.0, viewmatrix=cam["viewmatrix"].to(device), projmatrix=cam["projmatrix"].to(device),
        sh_degree=0, campos=cam["campos"].to(device), prefiltered=False)


def stack_cameras(cams, device):
    """Batched camera tables for the n_cams entry points."""
    return {
        "viewmatrix": torch.cat([c["viewmatrix"] for c in cams]).to(device).contiguous(),
        "projmatrix": torch.cat([c["projmatrix"] for c in cams]).to(device).contiguous(),
        "tanfovx": [c["tanfovx"] for c in cams], "tanfovy": [c["tanfovy"] for c in cams],
        "image_width": cams[0]["image_width"], "image_height": cams[0]["image_height"],
    }
