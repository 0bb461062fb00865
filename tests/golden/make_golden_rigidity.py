"""Generates tests/golden/rigidity_*.npz by running the UNMODIFIED reference rigidity loss
(/root/reference/train.py: calculate_rigidity_loss :317-351, create_foreground_info :226-247) on CPU and
differentiating it with torch autograd.

The reference hard-codes device="cuda" (shared.py:39, external.py:32): the keyword is stripped from torch.zeros /
torch.zeros_like while the reference code runs.  imageio, wandb, tqdm, open3d and the rasterizer package are
absent or irrelevant here and are stubbed (train.py imports them at module level; none is on this path).
Run:  python tests/golden/make_golden_rigidity.py
"""
import os
import sys
import types
from collections import namedtuple

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def install_stubs():
    for name in ("imageio", "wandb", "open3d"):
        sys.modules[name] = types.ModuleType(name)
    tq = types.ModuleType("tqdm"); tq.tqdm = lambda x, **k: x
    sys.modules["tqdm"] = tq
    dgr = types.ModuleType("diff_gaussian_rasterization")
    dgr.GaussianRasterizationSettings = namedtuple("GaussianRasterizationSettings", "a")
    dgr.GaussianRasterizer = object
    sys.modules["diff_gaussian_rasterization"] = dgr


def case(P, K, seed, motion):
    g = torch.Generator().manual_seed(seed)
    means0 = (torch.rand(P, 3, generator=g) - 0.5) * 2.0
    quats0 = torch.randn(P, 4, generator=g)
    seg = torch.zeros(P, 3)
    fg = torch.rand(P, generator=g) > 0.35
    seg[:, 0] = fg.float(); seg[:, 2] = 1 - seg[:, 0]
    F = int(fg.sum())
    # neighbours among the foreground points (any index set is legal for the loss; real runs use 20-NN)
    m_fg = means0[fg]
    d2 = ((m_fg[:, None] - m_fg[None]) ** 2).sum(-1)
    d2.fill_diagonal_(float("inf"))
    sq, idx = torch.topk(d2, K, largest=False)
    weights = torch.exp(-2000 * sq.double()).float()           # train.py:176-181
    if seed % 2:                                               # also exercise weights that are not tiny
        weights = torch.rand(F, K, generator=g)
    means = (means0 + motion * torch.randn(P, 3, generator=g)).requires_grad_(True)
    quats = (quats0 + motion * 5 * torch.randn(P, 4, generator=g)).requires_grad_(True)
    return means0, quats0, seg, idx.long().contiguous(), weights.contiguous(), means, quats


def main():
    install_stubs()
    sys.path.insert(0, "/root/reference")
    orig = {n: getattr(torch, n) for n in ("zeros", "zeros_like")}

    def strip(fn):
        def w(*a, **k):
            k.pop("device", None)
            return fn(*a, **k)
        return w
    for n, f in orig.items():
        setattr(torch, n, strip(f))
    try:
        import train
        for name, (P, K, seed, motion) in {"small": (300, 20, 0, 0.01), "weights": (500, 20, 1, 0.02),
                                           "k5": (200, 5, 2, 0.05)}.items():
            means0, quats0, seg, idx, weights, means, quats = case(P, K, seed, motion)
            z3 = torch.zeros(P, 3)
            prev = train.create_foreground_info(
                {"means": means0, "rotation_quaternions": quats0, "segmentation_masks": seg}, idx)
            params = {"means": means, "rotation_quaternions": quats, "segmentation_masks": seg, "colors": z3,
                      "opacity_logits": torch.zeros(P, 1), "log_scales": z3}
            loss = train.calculate_rigidity_loss(params, train.NeighborInfo(weights=weights, indices=idx), prev)
            loss.backward()
            np.savez_compressed(
                os.path.join(HERE, f"rigidity_{name}.npz"), means=means.detach().numpy(), quats=quats.detach().numpy(),
                seg=seg.numpy(), nbr=idx.numpy(), weights=weights.numpy(),
                prev_inv_rot=prev.inverted_rotations.numpy(), prev_offsets=prev.offsets_to_neighbors.numpy(),
                means0=means0.numpy(), quats0=quats0.numpy(), loss=np.float64(loss.item()),
                d_means=means.grad.numpy(), d_quats=quats.grad.numpy())
            print(name, "F", int((seg[:, 0] > 0.5).sum()), "loss", loss.item())
    finally:
        for n, f in orig.items():
            setattr(torch, n, f)


if __name__ == "__main__":
    main()
