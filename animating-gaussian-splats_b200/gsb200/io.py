"""On-disk formats either side of the hot path, so that real Panoptic-style
sequences flow through the B200 path and its artefacts stay interchangeable with the reference's:

  <sequence>/train_meta.json   {"w", "h", "k"[t][c][3][3], "w2c"[t][c][4][4], "fn"[t][c]}   shared.py:127-171
  <sequence>/ims/<fn>          colour image (uint8 HxWx3)                                     shared.py:154-161
  <sequence>/seg/<fn .png>     foreground mask (0/1)                                          shared.py:131-143
  <sequence>/init_pt_cld.npz   "data" [N,7] = x y z r g b seg                                 densify.py:22-57
  <sequence>/densified_initial_gaussian_cloud_parameters.pth   dict of 8 nn.Parameter         densify.py:190-198,
                                                                                              train.py:155-163

Function names and argument lists follow the reference (create_render_settings, load_timestep_views,
initialize_parameters, get_scene_radius, export_parameters, load_densified_initial_parameters); each takes
an extra `device` so the same code builds CPU tensors for the format tests.  The 3-nearest-neighbour
distances of initialize_parameters come from scipy's cKDTree instead of open3d's KDTreeFlann
(shared.py:45-61): both return exact squared Euclidean distances, computed here in float64 like open3d.
"""
import json
from dataclasses import dataclass
from pathlib import Path

import numpy as np
import torch

PARAMETER_FILE = "densified_initial_gaussian_cloud_parameters.pth"
PARAMETER_KEYS = ("means", "colors", "segmentation_masks", "rotation_quaternions", "opacity_logits", "log_scales",
                  "camera_matrices", "camera_center")
CAMERA_COUNT_LIMIT = 50   # densify.py:25


@dataclass
class View:   # shared.py:12-17
    camera_index: int
    render_settings: object
    image: torch.Tensor
    segmentation_mask: torch.Tensor


def load_dataset_metadata(sequence_path):
    with (Path(sequence_path) / "train_meta.json").open() as file:   # densify.py:203-205
        return json.load(file)


def create_render_settings(image_width, image_height, intrinsic_matrix, extrinsic_matrix,
                           near_clipping_plane_distance=1.0, far_clipping_plane_distance=100, device="cuda"):
    """shared.py:64-124: viewmatrix = w2c^T, projmatrix = w2c^T @ P^T with the OpenGL-style P, tanfov = dim / 2f,
    campos = inv(w2c)[:3, 3], bg 0, scale_modifier 1, sh_degree 0, prefiltered False.  All float32 arithmetic
    happens in torch in the reference's order (inverse and bmm on float32 tensors)."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    fx, fy = intrinsic_matrix[0][0], intrinsic_matrix[1][1]
    cx, cy = intrinsic_matrix[0][2], intrinsic_matrix[1][2]
    n, f = near_clipping_plane_distance, far_clipping_plane_distance
    w2c = torch.tensor(extrinsic_matrix).to(device).float()
    camera_center = torch.inverse(w2c)[:3, 3]
    w2c = w2c.unsqueeze(0).transpose(1, 2)
    projection = torch.tensor([[2 * fx / image_width, 0.0, -(image_width - 2 * cx) / image_width, 0.0],
                               [0.0, 2 * fy / image_height, -(image_height - 2 * cy) / image_height, 0.0],
                               [0.0, 0.0, f / (f - n), -(f * n) / (f - n)],
                               [0.0, 0.0, 1.0, 0.0]]).to(device).float().unsqueeze(0).transpose(1, 2)
    return GaussianRasterizationSettings(
        image_height=image_height, image_width=image_width,
        tanfovx=image_width / (2 * fx), tanfovy=image_height / (2 * fy),
        bg=torch.tensor([0, 0, 0], dtype=torch.float32, device=device), scale_modifier=1.0,
        viewmatrix=w2c, projmatrix=w2c.bmm(projection), sh_degree=0, campos=camera_center, prefiltered=False)


def _read_image(path):
    from PIL import Image
    with Image.open(path) as im:
        return np.array(im)


def load_timestep_views(dataset_metadata, timestep, sequence_path, device="cuda"):
    """shared.py:127-171: one View per camera of the timestep; image [3,H,W] in [0,1], segmentation mask
    stacked as (fg, 0, 1 - fg) [3,H,W]."""
    sequence_path = Path(sequence_path)
    views = []
    for camera_index, filename in enumerate(dataset_metadata["fn"][timestep]):
        seg = torch.tensor(_read_image(sequence_path / "seg" / filename.replace(".jpg", ".png"))
                           .astype(np.float32)).float().to(device)
        image = torch.tensor(_read_image(sequence_path / "ims" / filename)).float().to(device).permute(2, 0, 1) / 255
        views.append(View(
            camera_index=camera_index,
            render_settings=create_render_settings(
                image_width=dataset_metadata["w"], image_height=dataset_metadata["h"],
                intrinsic_matrix=dataset_metadata["k"][timestep][camera_index],
                extrinsic_matrix=dataset_metadata["w2c"][timestep][camera_index], device=device),
            image=image,
            segmentation_mask=torch.stack((seg, torch.zeros_like(seg), 1 - seg))))
    return views


def stack_views(views, device=None):
    """Camera-batched form of a list of Views for gsb200.step.timestep_step: camera tables as
    gsb200.scene.stack_cameras builds them plus the stacked targets [n,3,H,W] and masks [n,3,H,W]."""
    dev = device if device is not None else views[0].image.device
    s = [v.render_settings for v in views]
    cams = {
        "viewmatrix": torch.cat([x.viewmatrix for x in s]).to(dev).contiguous(),
        "projmatrix": torch.cat([x.projmatrix for x in s]).to(dev).contiguous(),
        "tanfovx": [float(x.tanfovx) for x in s], "tanfovy": [float(x.tanfovy) for x in s],
        "image_width": int(s[0].image_width), "image_height": int(s[0].image_height),
    }
    images = torch.stack([v.image for v in views]).to(dev).contiguous()
    masks = torch.stack([v.segmentation_mask for v in views]).to(dev).contiguous()
    return cams, images, masks


def knn_squared_distances(points, k):
    """Squared distances to the k nearest neighbours, self excluded (shared.py:45-61, float64)."""
    from scipy.spatial import cKDTree
    pts = np.ascontiguousarray(points, np.float64)
    d, idx = cKDTree(pts).query(pts, k=k + 1, workers=-1)   # same result on all host cores
    return idx[:, 1:], d[:, 1:] ** 2


def initialize_parameters(sequence_path, device="cuda"):
    """densify.py:22-57: the 8-entry parameter dict from init_pt_cld.npz."""
    cloud = np.load(Path(sequence_path) / "init_pt_cld.npz")["data"]
    seg = cloud[:, 6]
    n = seg.shape[0]
    _, sq = knn_squared_distances(cloud[:, :3], 3)
    tables = {
        "means": cloud[:, :3],
        "colors": cloud[:, 3:6],
        "segmentation_masks": np.stack((seg, np.zeros_like(seg), 1 - seg), -1),
        "rotation_quaternions": np.tile([1, 0, 0, 0], (n, 1)),
        "opacity_logits": np.zeros((n, 1)),
        "log_scales": np.tile(np.log(np.sqrt(sq.mean(-1).clip(min=0.0000001)))[..., None], (1, 3)),
        "camera_matrices": np.zeros((CAMERA_COUNT_LIMIT, 3)),
        "camera_center": np.zeros((CAMERA_COUNT_LIMIT, 3)),
    }
    return {k: torch.nn.Parameter(torch.tensor(v).to(device).float().contiguous().requires_grad_(True))
            for k, v in tables.items()}


def get_scene_radius(dataset_metadata):
    """densify.py:60-65: 1.1 x the largest distance of a timestep-0 camera centre from their mean."""
    centers = np.linalg.inv(dataset_metadata["w2c"][0])[:, :3, 3]
    return 1.1 * np.max(np.linalg.norm(centers - np.mean(centers, 0)[None], axis=-1))


def export_parameters(sequence_path, parameters):
    """densify.py:190-198 without the wandb upload: torch.save of the dict of nn.Parameter."""
    path = Path(sequence_path) / PARAMETER_FILE
    torch.save(parameters, path)
    return path


def load_densified_initial_parameters(data_directory_path, sequence_name, device="cuda"):
    """train.py:155-163: the dict comes back with requires_grad off.  The file holds pickled nn.Parameter
    objects, so weights_only must be off; map_location lets a CPU process read a file written on a GPU."""
    parameters = torch.load(Path(data_directory_path) / sequence_name / PARAMETER_FILE, map_location=device,
                            weights_only=False)
    missing = [k for k in PARAMETER_KEYS if k not in parameters]
    if missing:
        raise KeyError(f"{PARAMETER_FILE}: missing entries {missing}")
    for parameter in parameters.values():
        parameter.requires_grad = False
    return parameters


def write_sequence(sequence_path, metadata, images, masks, point_cloud):
    """Writes a sequence directory in the reference's layout (used by the tests and the synthetic-data
    front end): images[t][c] uint8 HxWx3, masks[t][c] uint8 HxW in {0,1}, point_cloud [N,7]."""
    from PIL import Image
    root = Path(sequence_path)
    for sub in ("ims", "seg"):
        (root / sub).mkdir(parents=True, exist_ok=True)
    for t, names in enumerate(metadata["fn"]):
        for c, fn in enumerate(names):
            for sub, arr, name in (("ims", images[t][c], fn), ("seg", masks[t][c], fn.replace(".jpg", ".png"))):
                (root / sub / name).parent.mkdir(parents=True, exist_ok=True)
                Image.fromarray(np.asarray(arr)).save(root / sub / name)
    with (root / "train_meta.json").open("w") as file:
        json.dump(metadata, file)
    np.savez(root / "init_pt_cld.npz", data=np.asarray(point_cloud))
