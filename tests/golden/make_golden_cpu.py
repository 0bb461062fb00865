"""Generates tests/golden/ssim_*.npz by importing the UNMODIFIED reference loss code
(/root/reference/external.py: calc_ssim, external.py:68-110) on the CPU of the build container.
`external` imports `shared`, which imports open3d and diff_gaussian_rasterization (shared.py:6,9); neither
is needed by calc_ssim, so both are stubbed in sys.modules.  Run:  python tests/golden/make_golden_cpu.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    sys.modules["open3d"] = types.ModuleType("open3d")
    dgr = types.ModuleType("diff_gaussian_rasterization")
    dgr.GaussianRasterizationSettings = object
    sys.modules["diff_gaussian_rasterization"] = dgr
    sys.path.insert(0, "/root/reference")
    import external  # the reference's own file, unmodified

    cases = [("a", 3, 40, 56, 0), ("b", 3, 23, 17, 1), ("c", 3, 64, 64, 2)]
    for name, C, H, W, seed in cases:
        g = torch.Generator().manual_seed(seed)
        img = torch.rand((C, H, W), generator=g)
        # smooth-ish second image so that SSIM is not degenerate
        tgt = (img + 0.25 * torch.rand((C, H, W), generator=g)).clamp(0, 1)
        if name == "c":
            tgt = img.clone()  # identical images: ssim = 1, l1 = 0, sign(0) = 0 gradient for L1
        x = img.clone().requires_grad_(True)
        ssim = external.calc_ssim(x, tgt)
        l1 = torch.nn.functional.l1_loss(x, tgt)
        loss = 0.8 * l1 + 0.2 * (1.0 - ssim)   # densify.py:127-129
        loss.backward()
        np.savez_compressed(os.path.join(HERE, f"ssim_{name}.npz"), img=img.numpy(), target=tgt.numpy(),
                            ssim=np.float64(ssim.item()), l1=np.float64(l1.item()), loss=np.float64(loss.item()),
                            grad=x.grad.numpy())
        print(name, "ssim", ssim.item(), "l1", l1.item(), "loss", loss.item())


if __name__ == "__main__":
    main()
