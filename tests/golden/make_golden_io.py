"""Generates tests/golden/io_sequence.npz by running the UNMODIFIED reference loaders
(/root/reference/shared.py: create_render_settings, load_timestep_views; /root/reference/densify.py:
initialize_parameters, get_scene_radius) in the build container on a tiny synthetic sequence.

The reference moves everything to the GPU (`.cuda()`, device="cuda"); this container has none, so while the
reference code runs `torch.Tensor.cuda` is the identity and the device keyword is stripped from torch
factory calls.  open3d is absent: the script installs a stand-in module whose KDTreeFlann answers
search_knn_vector_3d from scipy's cKDTree (exact squared Euclidean distances in float64, the quantity
open3d returns).  wandb / tqdm are stubbed because densify.py imports them.  The sequence files themselves
(JSON, JPEG/PNG bytes, npz) are stored in the fixture so that the test reads exactly the bytes the reference
read.  Run:  python tests/golden/make_golden_io.py
"""
import io
import json
import os
import sys
import tempfile
import types
from collections import namedtuple
from pathlib import Path

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "animating-gaussian-splats_b200"))


def synthetic_sequence(root, T=2, C=3, W=32, H=24, N=200, seed=0):
    from gsb200 import io as gio
    from gsb200 import scene
    rng = np.random.default_rng(seed)
    meta = {"w": W, "h": H, "k": [], "w2c": [], "fn": []}
    images, masks = [], []
    for t in range(T):
        ks, ws, fns, ims, sgs = [], [], [], [], []
        for c in range(C):
            az = 2 * np.pi * (c + 0.1 * t) / C
            eye = (3.0 * np.cos(az), 3.0 * np.sin(az), 1.0 + 0.2 * c)
            ws.append(scene.look_at_w2c(eye).tolist())
            fx = 0.7 * W * (1 + 0.05 * c)
            ks.append([[fx, 0.0, W / 2 + 0.5 * c], [0.0, fx * 1.01, H / 2 - 0.25 * c], [0.0, 0.0, 1.0]])
            fns.append(f"{c}/{t:06d}.jpg")
            ims.append(rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8))
            sgs.append((rng.uniform(size=(H, W)) > 0.5).astype(np.uint8))
        meta["k"].append(ks); meta["w2c"].append(ws); meta["fn"].append(fns)
        images.append(ims); masks.append(sgs)
    cloud = np.concatenate([rng.uniform(-1, 1, size=(N, 3)), rng.uniform(0, 1, size=(N, 3)),
                            (rng.uniform(size=(N, 1)) > 0.4).astype(np.float64)], axis=1)
    cloud[1, :3] = cloud[0, :3]   # coincident points: exercises the clip(min=1e-7) of densify.py:49
    cloud[2, :3] = cloud[0, :3]
    cloud[3, :3] = cloud[0, :3]
    gio.write_sequence(root, meta, images, masks, cloud)


def install_stubs():
    from scipy.spatial import cKDTree
    o3d = types.ModuleType("open3d")
    o3d.geometry = types.SimpleNamespace()
    o3d.utility = types.SimpleNamespace()

    class PointCloud:
        points = None
    o3d.geometry.PointCloud = PointCloud
    o3d.utility.Vector3dVector = lambda a: np.asarray(a, np.float64)

    class KDTreeFlann:
        def __init__(self, pc):
            self.tree = cKDTree(pc.points)

        def search_knn_vector_3d(self, p, k):
            d, i = self.tree.query(p, k=k)
            return k, list(i), list(d ** 2)
    o3d.geometry.KDTreeFlann = KDTreeFlann
    sys.modules["open3d"] = o3d
    dgr = types.ModuleType("diff_gaussian_rasterization")
    dgr.GaussianRasterizationSettings = namedtuple(
        "GaussianRasterizationSettings", "image_height image_width tanfovx tanfovy bg scale_modifier viewmatrix "
        "projmatrix sh_degree campos prefiltered")
    dgr.GaussianRasterizer = object
    sys.modules["diff_gaussian_rasterization"] = dgr
    sys.modules["wandb"] = types.ModuleType("wandb")
    tq = types.ModuleType("tqdm"); tq.tqdm = lambda x, **k: x
    sys.modules["tqdm"] = tq


def main():
    tmp = Path(tempfile.mkdtemp())
    synthetic_sequence(tmp)
    install_stubs()
    sys.path.insert(0, "/root/reference")
    orig = {n: getattr(torch, n) for n in ("zeros", "ones", "tensor")}
    orig_cuda = torch.Tensor.cuda

    def strip(fn):
        def w(*a, **k):
            k.pop("device", None)
            return fn(*a, **k)
        return w
    for n, f in orig.items():
        setattr(torch, n, strip(f))
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        import shared
        import densify
        meta = json.load(open(tmp / "train_meta.json"))
        out = {}
        for t in range(len(meta["fn"])):
            for v in shared.load_timestep_views(meta, t, tmp):
                s, p = v.render_settings, f"t{t}_c{v.camera_index}_"
                out[p + "viewmatrix"] = s.viewmatrix.numpy(); out[p + "projmatrix"] = s.projmatrix.numpy()
                out[p + "campos"] = s.campos.numpy(); out[p + "bg"] = s.bg.numpy()
                out[p + "scalars"] = np.array([s.image_height, s.image_width, s.tanfovx, s.tanfovy, s.scale_modifier,
                                               s.sh_degree, float(s.prefiltered)], np.float64)
                out[p + "image"] = v.image.numpy(); out[p + "segmentation_mask"] = v.segmentation_mask.numpy()
        params = densify.initialize_parameters(tmp)
        for k, v in params.items():
            assert isinstance(v, torch.nn.Parameter) and v.requires_grad
            out["param_" + k] = v.detach().numpy()
        out["scene_radius"] = np.float64(densify.get_scene_radius(meta))
    finally:
        for n, f in orig.items():
            setattr(torch, n, f)
        torch.Tensor.cuda = orig_cuda
    files = {}
    for path in sorted(tmp.rglob("*")):
        if path.is_file():
            files["file:" + str(path.relative_to(tmp))] = np.frombuffer(path.read_bytes(), np.uint8)
    np.savez_compressed(os.path.join(HERE, "io_sequence.npz"), **out, **files)
    print("wrote io_sequence.npz:", len(out), "arrays,", len(files), "files,",
          os.path.getsize(os.path.join(HERE, "io_sequence.npz")), "bytes")


if __name__ == "__main__":
    main()
