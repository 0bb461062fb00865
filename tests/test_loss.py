"""L1 + SSIM parity.  CPU: the C oracle against golden vectors produced by the reference's own
external.calc_ssim (tests/golden/make_golden_cpu.py).  GPU: the fused kernels (through the C ABI)
against the same golden vectors and against the oracle at Panoptic size."""
import glob
import os

import numpy as np
import pytest
import torch

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ssim_*.npz")))
IDS = [os.path.basename(p) for p in GOLDEN]


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_oracle_ssim_matches_reference_golden(oracle, path):
    z = np.load(path)
    out, g = oracle.ssim_l1(z["img"], z["target"], 0.8, 0.2, grad=True)
    assert abs(out[0] - float(z["l1"])) <= 1e-6
    assert abs(out[1] - float(z["ssim"])) <= 1e-5
    assert abs(out[2] - float(z["loss"])) <= 1e-6
    scale = np.abs(z["grad"]).max() + 1e-12
    assert np.abs(g - z["grad"]).max() <= 1e-4 * scale + 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_fused_ssim_matches_reference_golden(path):
    from gsb200 import loss
    z = np.load(path)
    x = torch.from_numpy(z["img"]).cuda().requires_grad_(True)
    y = torch.from_numpy(z["target"]).cuda()
    l = loss.l1_ssim_loss(x, y)
    l.backward()
    assert abs(l.item() - float(z["loss"])) <= 1e-6
    scale = np.abs(z["grad"]).max() + 1e-12
    assert np.abs(x.grad.cpu().numpy() - z["grad"]).max() <= 1e-4 * scale + 1e-9
    s = loss.calc_ssim(x.detach(), y)
    assert abs(s.item() - float(z["ssim"])) <= 1e-5
    # drop-in composition used by the callers: 0.8 * l1 + 0.2 * (1 - calc_ssim)  (densify.py:127-129)
    x2 = x.detach().clone().requires_grad_(True)
    l2 = 0.8 * torch.nn.functional.l1_loss(x2, y) + 0.2 * (1.0 - loss.calc_ssim(x2, y))
    l2.backward()
    assert abs(l2.item() - float(z["loss"])) <= 1e-6
    assert np.abs(x2.grad.cpu().numpy() - z["grad"]).max() <= 1e-4 * scale + 1e-9


@pytest.mark.gpu
def test_fused_ssim_batch_vs_oracle(oracle):
    from gsb200 import loss
    g = torch.Generator().manual_seed(5)
    n, H, W = 3, 360, 640   # Panoptic image size, ragged against the 16x16 tiles (360 = 22.5 tiles)
    img = torch.rand((n, 3, H, W), generator=g)
    tgt = (img + 0.2 * torch.rand((n, 3, H, W), generator=g)).clamp(0, 1)
    stats, grad = loss.l1_ssim_fused(img.cuda(), tgt.cuda(), 0.8, 0.2, grad=True)
    for i in range(n):
        out, go = oracle.ssim_l1(img[i].numpy(), tgt[i].numpy(), 0.8, 0.2, grad=True)
        assert np.abs(stats[i].cpu().numpy() - out).max() <= 2e-6
        assert np.abs(grad[i].cpu().numpy() - go).max() <= 1e-4 * np.abs(go).max()
