"""Tile-band sharding (gsb200.dist.shard_bands): the records of the bands of a timestep, rendered as separate batches
(one per virtual rank) with halo rows and band-restricted losses, add up to the record of the whole cameras."""
import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_cams,world,H", [(3, 4, 120), (1, 3, 100), (5, 8, 96)])
def test_band_records_add_up_to_whole_cameras(n_cams, world, H):
    from gsb200 import dist as gd, scene, step
    dev = torch.device("cuda", 0)
    P, W = 15000, 200
    g, a, cams = helpers.scene_inputs(P, W, H, seed=11, n_cams=n_cams, device=dev)
    sc = scene.stack_cameras(cams, dev)
    bg = torch.tensor([0.05, 0.1, 0.15], device=dev)
    targets = torch.rand((n_cams, 3, H, W), generator=torch.Generator().manual_seed(3)).to(dev)
    # the whole cameras in one batch: the answer
    full = torch.zeros((P, 20), device=dev)
    loss_full, _, radii_full, st = step.timestep_step(a, sc, targets, bg, None, flat=full)
    color_full = torch.empty(0)
    from gsb200 import raster
    color_full, _, _, _ = raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None, bg,
                                         1.0, sc["viewmatrix"], sc["projmatrix"], sc["tanfovx"], sc["tanfovy"], H, W, n_cams=n_cams)
    if n_cams == 1:
        color_full = color_full[None]
    bands, split = gd.shard_bands(gd.tile_row_costs(st), world)
    assert sorted((c, r) for b in bands for c, r0, r1 in b for r in range(r0, r1)) == \
        [(c, r) for c in range(n_cams) for r in range((H + 15) // 16)]
    assert len(split) >= 1
    width = gd.flat_width(len(split))
    total = torch.zeros((P, width), device=dev, dtype=torch.float64)
    maxr = torch.zeros(P, device=dev)
    loss_sum = torch.zeros(n_cams, device=dev, dtype=torch.float64)
    for mine in bands:
        if not mine:
            continue
        sub, tg = gd.band_batch(sc, targets, mine, split, dev)
        flat = torch.full((P, width), float("nan"), device=dev)
        loss_b, _, radii_b, st_b = step.timestep_step(a, sub, tg, bg, None, flat=flat)
        torch.cuda.synchronize()
        assert torch.isfinite(flat).all()
        total += flat.double()
        maxr = torch.maximum(maxr, flat[:, 16])
        n_b = len(mine)
        color_b, _, _, _ = raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None, bg,
                                          1.0, sub["viewmatrix"], sub["projmatrix"], sub["tanfovx"], sub["tanfovy"], H, W,
                                          n_cams=n_b, entries=sub["entries"])
        if n_b == 1:
            color_b = color_b[None]
        for k, (cam, r0, r1) in enumerate(mine):
            loss_sum[cam] += float(loss_b[k, 2])
            assert torch.equal(radii_b.reshape(n_b, P)[k], radii_full.reshape(n_cams, P)[cam])   # whole-camera state
            y0, y1 = max(16 * r0 - 5, 0), min(16 * r1 + 5, H)   # band + the SSIM halo: identical pixels
            assert torch.equal(color_b[k][:, y0:y1], color_full[cam][:, y0:y1]), (cam, r0, r1)
    got = total.float()
    bad, worst = helpers.grad_close(helpers.to_np(got[:, :14]), helpers.to_np(full[:, :14]), rel=1e-4, floor=1e-5)
    assert bad == 0, (bad, worst)
    assert torch.equal(got[:, 15], full[:, 15])          # visibility: every camera counted once
    assert torch.equal(maxr, full[:, 16])
    # the statistic: whole cameras' norms (col 14) + the norm of every split camera's summed gradient
    norm = got[:, 14].clone()
    for j in range(len(split)):
        norm += torch.sqrt(got[:, 20 + 2 * j] ** 2 + got[:, 21 + 2 * j] ** 2)
    assert torch.allclose(norm, full[:, 14], rtol=1e-4, atol=1e-9 + 1e-5 * float(full[:, 14].max()))
    assert torch.allclose(loss_sum.float(), loss_full[:, 2], rtol=0, atol=2e-6)
    # gsb_flat_apply_stats on the wide record == on the whole-camera record
    s_full, s_band = step.DensifyStats(P, dev), step.DensifyStats(P, dev)
    s_full.apply_flat(full)
    rec = got.contiguous(); rec[:, 16] = maxr
    s_band.apply_flat(rec)
    assert torch.equal(s_band.visibility_count, s_full.visibility_count) and torch.equal(s_band.max_2d_radii, s_full.max_2d_radii)
    assert torch.allclose(s_band.mean_2d_gradients_accumulated, s_full.mean_2d_gradients_accumulated, rtol=1e-4,
                          atol=1e-9 + 1e-5 * float(full[:, 14].max()))
