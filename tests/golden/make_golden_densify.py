"""Generates tests/golden/densify_*.npz by running the UNMODIFIED reference densification code
(/root/reference/external.py: densify_gaussians, external.py:211-314) in the build container.

The reference hard-codes device="cuda" (external.py:32,242,260,272-282); this container has no GPU, so
the script strips that keyword inside torch factory calls while the reference code runs (a shim around
torch.zeros / torch.ones / torch.tensor; the reference source is imported as is).  torch.normal is
wrapped only to RECORD the samples it returns, so that implementations under test can be fed the same
draws.  Run:  python tests/golden/make_golden_densify.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
KEYS = ["means", "colors", "segmentation_masks", "rotation_quaternions", "opacity_logits", "log_scales"]


def make_state(N, seed, scene_radius):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g)
    params = {
        "means": (r(N, 3) - 0.5) * 3.0,
        "colors": r(N, 3),
        "segmentation_masks": torch.nn.functional.one_hot((r(N) > 0.5).long() * 2, 3).float(),
        "rotation_quaternions": torch.randn(N, 4, generator=g),
        "opacity_logits": torch.randn(N, 1, generator=g) * 3.0 - 1.0,
        # scales straddling 0.01 * scene_radius (clone vs split) with a few above 0.1 * scene_radius
        "log_scales": torch.log(0.01 * scene_radius * torch.exp(torch.randn(N, 3, generator=g) * 1.2)),
        "camera_matrices": torch.zeros(50, 3), "camera_center": torch.zeros(50, 3),
    }
    params = {k: torch.nn.Parameter(v.float().contiguous().requires_grad_(True)) for k, v in params.items()}
    opt = torch.optim.Adam([{"params": [v], "name": k, "lr": 1e-3} for k, v in params.items()], lr=0.0, eps=1e-15)
    for v in params.values():   # one optimiser step so that Adam moments exist and are non-trivial
        v.grad = torch.randn(v.shape, generator=g)
    opt.step()
    opt.zero_grad()
    visible = r(N) > 0.3
    stats = dict(visibility_count=(r(N) * 40).floor() * visible, mean_2d_gradients_accumulated=r(N) * 0.02 * visible,
                 max_2d_radii=(r(N) * 30).floor(), visible=r(N) > 0.4, means2d_grad=torch.randn(N, 3, generator=g) * 1e-3)
    return params, opt, stats


def snapshot(params, opt, prefix):
    out = {}
    for k in KEYS:
        p = params[k]
        out[f"{prefix}_{k}"] = p.detach().numpy().copy()
        st = opt.state.get(p, None)
        if st is not None and "exp_avg" in st:
            out[f"{prefix}_m_{k}"] = st["exp_avg"].numpy().copy()
            out[f"{prefix}_v_{k}"] = st["exp_avg_sq"].numpy().copy()
    return out


def main():
    sys.modules["open3d"] = types.ModuleType("open3d")
    dgr = types.ModuleType("diff_gaussian_rasterization")
    dgr.GaussianRasterizationSettings = object
    sys.modules["diff_gaussian_rasterization"] = dgr
    sys.path.insert(0, "/root/reference")
    import external
    import shared

    recorded = []
    orig = {n: getattr(torch, n) for n in ("zeros", "ones", "tensor", "normal")}

    def strip(fn):
        def w(*a, **k):
            k.pop("device", None)
            return fn(*a, **k)
        return w

    def normal_rec(*a, **k):
        s = orig["normal"](*a, **k)
        recorded.append(s.clone())
        return s

    torch.cuda.empty_cache = lambda: None
    for n in ("zeros", "ones", "tensor"):
        setattr(torch, n, strip(orig[n]))
    torch.normal = normal_rec
    try:
        scene_radius = 3.3
        for name, N, seed, i in [("i600", 4000, 1, 600), ("i700_accum_only", 1500, 2, 750), ("i3000", 3000, 3, 3000),
                                 ("i5000", 2500, 4, 5000), ("i6000_inactive", 500, 5, 6000)]:
            torch.manual_seed(100 + seed)
            params, opt, st = make_state(N, seed, scene_radius)
            out = snapshot(params, opt, "in")
            dv = shared.DensificationVariables(visibility_count=st["visibility_count"].clone(),
                                               mean_2d_gradients_accumulated=st["mean_2d_gradients_accumulated"].clone(),
                                               max_2d_radii=st["max_2d_radii"].clone(),
                                               gaussian_is_visible_mask=st["visible"].clone(),
                                               means_2d=types.SimpleNamespace(grad=st["means2d_grad"].clone()))
            for k, v in st.items():
                out[f"in_stat_{k}"] = v.numpy().copy()
            recorded.clear()
            with torch.no_grad():
                external.densify_gaussians(params, dv, scene_radius, opt, i)
            out.update(snapshot(params, opt, "out"))
            out["out_stat_visibility_count"] = dv.visibility_count.numpy().copy()
            out["out_stat_mean_2d_gradients_accumulated"] = dv.mean_2d_gradients_accumulated.numpy().copy()
            out["out_stat_max_2d_radii"] = dv.max_2d_radii.numpy().copy()
            out["samples"] = recorded[0].numpy().copy() if recorded else np.zeros((0, 3), np.float32)
            out["i"] = np.int64(i); out["scene_radius"] = np.float64(scene_radius)
            np.savez_compressed(os.path.join(HERE, f"densify_{name}.npz"), **out)
            print(name, "N", N, "->", params["means"].shape[0], "samples", out["samples"].shape)
    finally:
        for n, f in orig.items():
            setattr(torch, n, f)


if __name__ == "__main__":
    main()
