"""The camera-batched step and its host-fed front (gsb200.step): the end-to-end path with HOST buffers must
produce the loss numbers of the device-resident step on the same bytes, step after step (double buffer)."""
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


def test_host_feed_matches_device_step():
    from gsb200 import scene, step
    dev = torch.device("cuda", 0)
    P, W, H, n = 20000, 160, 96, 3
    g, a, cams = helpers.scene_inputs(P, W, H, seed=3, n_cams=n, device=dev)
    sc = scene.stack_cameras(cams, dev)
    bg = torch.zeros(3, device=dev)
    targets = torch.rand((n, 3, H, W), generator=torch.Generator().manual_seed(1))
    host_raw = {k: g[k].contiguous().pin_memory() for k in step.HostFeed.RAW_KEYS}
    host_targets = targets.pin_memory()
    want, grads, radii, st = step.timestep_step(a, sc, targets.to(dev), bg, None)
    torch.cuda.synchronize()
    feed = step.HostFeed(host_raw, host_targets, dev)
    seen = []
    for _ in range(3):   # both halves of the double buffer, and re-use of the first
        seen.append(feed.step(sc, bg, None).clone())
    for got in seen:
        assert torch.allclose(got, want.cpu(), rtol=0, atol=1e-6), (got, want)
    # non-blocking steps: the host runs ahead, results are read after wait()
    for i in range(5):
        out = feed.step(sc, bg, None, wait=False)
        if i % 2 == 1:
            assert feed.wait() is out
            assert torch.allclose(out, want.cpu(), rtol=0, atol=1e-6), (i, out, want)
    assert torch.allclose(feed.wait(), want.cpu(), rtol=0, atol=1e-6)
    assert feed.h2d_bytes == sum(v.numel() * 4 for v in host_raw.values()) + targets.numel() * 4
    assert feed.d2h_bytes == n * 3 * 4


def test_host_feed_rejects_pageable_memory():
    from gsb200 import step
    dev = torch.device("cuda", 0)
    raw = {k: torch.zeros(4, 3) for k in step.HostFeed.RAW_KEYS}
    with pytest.raises(RuntimeError):
        step.HostFeed(raw, torch.zeros(1, 3, 16, 16), dev)


def test_host_feed_graph_mode_matches_eager():
    """HostFeed with the step captured per buffer half (graph=True): same loss numbers and flat record as the eager
    host-fed step, over both halves of the double buffer and their re-use."""
    from gsb200 import raster, scene, step
    dev = torch.device("cuda", 0)
    P, W, H, n = 20000, 160, 96, 3
    g, a, cams = helpers.scene_inputs(P, W, H, seed=3, n_cams=n, device=dev)
    sc = scene.stack_cameras(cams, dev)
    sc["tanfov_dev"] = raster.tanfov_table(sc["tanfovx"], sc["tanfovy"], n, dev)
    bg = torch.zeros(3, device=dev)
    targets = torch.rand((n, 3, H, W), generator=torch.Generator().manual_seed(1))
    host_raw = {k: g[k].contiguous().pin_memory() for k in step.HostFeed.RAW_KEYS}
    host_targets = targets.pin_memory()
    flat_e = torch.zeros((P, 20), device=dev)
    want, _, _, st = step.timestep_step(a, sc, targets.to(dev), bg, None, flat=flat_e)
    arena = raster.RasterArena(dev).reserve(P, n, W, H, capacity=int(1.5 * st.num_rendered))
    flat = torch.zeros((P, 20), device=dev)
    stats = step.DensifyStats(P, dev)
    after = lambda *r: stats.apply_flat(flat)
    feed = step.HostFeed(host_raw, host_targets, dev)
    for i in range(5):
        out = feed.step(sc, bg, None, after, wait=True, flat=flat, arena=arena, graph=True)
        assert torch.allclose(out, want.cpu(), rtol=0, atol=1e-6), (i, out, want)
        bad, worst = helpers.grad_close(helpers.to_np(flat[:, :14]), helpers.to_np(flat_e[:, :14]), rel=1e-4, floor=1e-5)
        assert bad == 0, (i, bad, worst)
    # the hook ran inside the graph on every replay and on the warm-up passes of the two captures (2 x 2)
    assert torch.equal(stats.visibility_count, (5 + 4) * flat_e[:, 15])
