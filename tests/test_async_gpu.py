"""The host-free form of the step: asynchronous capacity-guarded forward through a caller-owned arena, the flat step
record written by the backward, statistics from the record, CUDA-graph replay, pair counts, and the one-kernel
exchange over peer-mapped memory.  Checker: the blocking path (itself pinned bit-exactly to the reference extension in
test_raster_gpu.py) and the C oracle."""
import os
import sys

import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _setup(P=20000, W=200, H=120, n=3, seed=5):
    from gsb200 import scene
    dev = torch.device("cuda", 0)
    g, a, cams = helpers.scene_inputs(P, W, H, seed=seed, n_cams=n, device=dev)
    sc = scene.stack_cameras(cams, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    targets = torch.rand((n, 3, H, W), generator=torch.Generator().manual_seed(seed)).to(dev)
    return dev, g, a, cams, sc, bg, targets


def _fwd(a, sc, bg, n, **kw):
    from gsb200 import raster
    return raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None, bg, 1.0,
                          sc["viewmatrix"], sc["projmatrix"], sc["tanfovx"], sc["tanfovy"], sc["image_height"],
                          sc["image_width"], n_cams=n, **kw)


@pytest.mark.parametrize("P,W,H,n", [(20000, 200, 120, 3), (3000, 1920, 1080, 1), (500, 64, 64, 1)])
def test_async_forward_equals_blocking(P, W, H, n):
    """Same bytes out of the capacity-guarded path: images, radii, instance lists, ranges, n_contrib."""
    from gsb200 import raster
    dev, g, a, cams, sc, bg, _ = _setup(P, W, H, n)
    color, depth, radii, st = _fwd(a, sc, bg, n)
    want = {k: v.clone() for k, v in st.tensors().items()}
    color, depth, radii = color.clone(), depth.clone(), radii.clone()
    R = st.num_rendered
    arena = raster.RasterArena(dev).reserve(P, n, W, H, capacity=int(1.3 * R) + 1000)
    c2, d2, r2, st2 = _fwd(a, sc, bg, n, arena=arena)
    assert st2.ctx.num_rendered == -1
    assert st2.status() == (0, R, 0)
    assert st2.num_rendered == R
    assert torch.equal(c2, color) and torch.equal(d2, depth) and torch.equal(r2, radii)
    got = st2.tensors()
    for k in ("point_list", "ranges", "n_contrib", "final_T", "seg_start", "gauss_order", "point_offsets"):
        assert torch.equal(got[k], want[k]), k
    # the arena is reused: a second call with other colours overwrites the same buffers and stays correct
    a2 = dict(a, colors_precomp=a["colors_precomp"].flip(0).contiguous())
    c_ref, _, _, _ = _fwd(a2, sc, bg, n)
    c3, _, _, _ = _fwd(a2, sc, bg, n, arena=arena)
    assert torch.equal(c3, c_ref)


def test_async_forward_overflow_is_reported_and_harmless():
    from gsb200 import _lib, raster
    dev, g, a, cams, sc, bg, _ = _setup()
    n = 3
    _, _, _, st = _fwd(a, sc, bg, n)
    R = st.num_rendered
    arena = raster.RasterArena(dev).reserve(a["means3D"].shape[0], n, 200, 120, capacity=R // 2)
    guard = torch.full((4096,), 7, dtype=torch.uint8, device=dev)   # something allocated right behind the arena
    color, depth, radii, st2 = _fwd(a, sc, bg, n, arena=arena)
    torch.cuda.synchronize()
    assert st2.status() == (_lib.GSB_ENOMEM, R, R)
    with pytest.raises(raster.ArenaOverflow) as e:
        st2.check()
    assert e.value.needed == R and e.value.capacity == R // 2
    # every tile list is empty: pure background, and the backward runs without touching anything out of bounds
    assert torch.equal(color, bg.view(1, 3, 1, 1).expand_as(color))
    flat = torch.empty((a["means3D"].shape[0], 20), device=dev)
    raster.backward_flat(st2, torch.ones_like(color), flat)
    torch.cuda.synchronize()
    assert float(flat[:, :14].abs().max()) == 0.0
    assert bool((guard == 7).all())
    # the caller's recovery: a larger arena
    arena.reserve(a["means3D"].shape[0], n, 200, 120, capacity=e.value.needed)
    _, _, _, st3 = _fwd(a, sc, bg, n, arena=arena)
    assert st3.status() == (0, R, 0)


def test_flat_backward_matches_separate_tables():
    """The flat record == the gradient tables of gsb_raster_backward + the statistics kernel on zeroed accumulators."""
    from gsb200 import raster, step
    dev, g, a, cams, sc, bg, targets = _setup()
    n, P = 3, a["means3D"].shape[0]
    color, depth, radii, st = _fwd(a, sc, bg, n)
    dL = torch.randn(color.shape, generator=torch.Generator().manual_seed(2)).to(dev)
    grads = raster.backward(st, dL, need_cov3D_grad=False)
    stats = step.DensifyStats(P, dev)
    stats.accumulate_batch(radii.contiguous(), grads["means2D"].contiguous())
    flat = torch.full((P, 20), float("nan"), device=dev)
    m2d = torch.full((n, P, 3), float("nan"), device=dev)
    raster.backward_flat(st, dL, flat, means2D_out=m2d)
    torch.cuda.synchronize()
    from gsb200 import dist as gd
    u = gd.unpack(flat)
    for k in ("means3D", "colors", "rotations", "opacities", "scales"):
        bad, worst = helpers.grad_close(helpers.to_np(u[k]), helpers.to_np(grads[k]).reshape(P, -1), rel=1e-5, floor=1e-6)
        assert bad == 0, (k, bad, worst)
    assert torch.equal(u["visibility_count"], stats.visibility_count)
    assert torch.equal(u["max_2d_radii"], stats.max_2d_radii)
    assert torch.allclose(u["mean_2d_gradients_accumulated"], stats.mean_2d_gradients_accumulated, rtol=1e-5, atol=1e-9)
    assert float(u["segmentation_colors"].abs().max()) == 0.0
    vis = (radii > 0)[..., None].expand(n, P, 3)
    bad, worst = helpers.grad_close(helpers.to_np(m2d[vis]), helpers.to_np(grads["means2D"][vis]), rel=1e-5, floor=1e-6)
    assert bad == 0, (bad, worst)
    assert float(m2d[~vis].abs().max()) == 0.0
    # statistics from the record (gsb_flat_apply_stats) == the per-camera kernel, twice in a row
    s2 = step.DensifyStats(P, dev)
    s2.apply_flat(flat); s2.apply_flat(flat)
    stats.accumulate_batch(radii.contiguous(), grads["means2D"].contiguous())
    assert torch.equal(s2.visibility_count, stats.visibility_count) and torch.equal(s2.max_2d_radii, stats.max_2d_radii)
    assert torch.allclose(s2.mean_2d_gradients_accumulated, stats.mean_2d_gradients_accumulated, rtol=1e-5, atol=1e-9)

    # mode 1: a second render of the same Gaussians with another colour table (densify.py:136-150)
    seg = torch.rand((P, 3), generator=torch.Generator().manual_seed(9)).to(dev)
    a2 = dict(a, colors_precomp=seg)
    color2, _, _, stb = _fwd(a2, sc, bg, n)
    dL2 = torch.randn(color.shape, generator=torch.Generator().manual_seed(3)).to(dev)
    g2 = raster.backward(stb, dL2, need_cov3D_grad=False)
    before = flat.clone()
    raster.backward_flat(stb, dL2, flat, mode=1)
    u2, ub = gd.unpack(flat), gd.unpack(before)
    for k in ("means3D", "rotations", "opacities", "scales"):
        want = helpers.to_np(ub[k]) + helpers.to_np(g2[k]).reshape(P, -1)
        bad, worst = helpers.grad_close(helpers.to_np(u2[k]), want, rel=1e-5, floor=1e-6)
        assert bad == 0, (k, bad, worst)
    bad, worst = helpers.grad_close(helpers.to_np(u2["segmentation_colors"]), helpers.to_np(g2["colors"]), rel=1e-5, floor=1e-6)
    assert bad == 0, (bad, worst)
    for k in ("colors", "mean_2d_gradients_accumulated", "visibility_count", "max_2d_radii"):
        assert torch.equal(u2[k], ub[k]), k


def test_step_graph_replays_the_eager_step():
    """timestep_step through arena + flat record, captured once and replayed, tracks parameter updates made in place."""
    from gsb200 import raster, scene, step
    dev, g, a, cams, sc, bg, targets = _setup()
    n, P = 3, a["means3D"].shape[0]
    sc["tanfov_dev"] = raster.tanfov_table(sc["tanfovx"], sc["tanfovy"], n, dev)
    _, _, _, st0 = _fwd(a, sc, bg, n)
    arena = raster.RasterArena(dev).reserve(P, n, 200, 120, capacity=int(1.5 * st0.num_rendered))
    flat = torch.zeros((P, 20), device=dev)
    graph = step.StepGraph(lambda: step.timestep_step(a, sc, targets, bg, None, flat=flat, arena=arena))
    for k in range(3):
        if k:   # an optimiser would do this in place between replays
            a["means3D"].add_(0.002 * torch.randn(a["means3D"].shape, generator=torch.Generator().manual_seed(k)).to(dev))
        loss_g, flat_g, radii_g, st_g = graph.replay()
        torch.cuda.synchronize()
        loss_g, flat_g, radii_g = loss_g.clone(), flat_g.clone(), radii_g.clone()
        assert st_g.status()[0] == 0
        flat_e = torch.zeros((P, 20), device=dev)
        loss_e, _, radii_e, _ = step.timestep_step(a, sc, targets, bg, None, flat=flat_e)
        assert torch.equal(radii_g, radii_e)
        assert torch.allclose(loss_g, loss_e, rtol=0, atol=1e-6)
        bad, worst = helpers.grad_close(helpers.to_np(flat_g[:, :14]), helpers.to_np(flat_e[:, :14]), rel=1e-4, floor=1e-5)
        assert bad == 0, (k, bad, worst)
        assert torch.equal(flat_g[:, 15:17], flat_e[:, 15:17])


def test_pair_counts(oracle):
    """B (walked) == sum of n_contrib; B_alpha (composited) == the oracle's count on the same scene."""
    from gsb200 import raster
    dev = torch.device("cuda", 0)
    P, W, H = 4000, 128, 96
    g, a, cams = helpers.scene_inputs(P, W, H, seed=0, device=dev)
    cam = cams[0]
    bg = torch.zeros(3, device=dev)
    _, _, _, st = raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None, bg,
                                 1.0, cam["viewmatrix"].to(dev), cam["projmatrix"].to(dev), cam["tanfovx"], cam["tanfovy"], H, W)
    counts = raster.count_pairs(st).cpu()
    assert int(counts[0, 0]) == int(st.tensors()["n_contrib"].long().sum())
    n = lambda k: a[k].cpu().numpy()
    ref = oracle.forward(n("means3D"), n("colors_precomp"), n("opacities"), n("scales"), n("rotations"), np.zeros(3, np.float32),
                         cam["viewmatrix"].numpy(), cam["projmatrix"].numpy(), cam["tanfovx"], cam["tanfovy"], H, W)
    assert int(counts[0, 0]) == int(ref["blend_count"][0])
    # the CPU cannot reproduce MUFU.EX2 bit for bit: pixels flagged fragile may differ by a pair or two
    assert abs(int(counts[0, 1]) - int(ref["blend_count"][1])) <= 2 * int(ref["fragile"].sum()) + 2
    assert 0 < int(counts[0, 1]) <= int(counts[0, 0])


def _exchange_worker(rank, world, port, P, out_dir):
    for p in (ROOT, os.path.join(ROOT, "animating-gaussian-splats_b200")):
        sys.path.insert(0, p)
    import torch.distributed as dist
    from gsb200 import dist as gd
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)   # host side channel only (handles, barriers)
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    ex = gd.Exchange(P, dev)
    assert ex.mode == "nvlink"
    outs = []
    for it in range(3):   # epochs advance inside the kernel: repeated launches, no host reset
        gen = torch.Generator().manual_seed(100 * it + rank)
        ex.flat.copy_(torch.randn((P, gd.WIDTH), generator=gen).to(dev))
        ex.allreduce()
        torch.cuda.synchronize()
        outs.append(ex.flat.cpu().numpy().copy())
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), *outs)
    ex.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_exchange_kernel_two_processes(tmp_path, world):
    """The one-kernel all-reduce over CUDA-IPC peer mappings: `world` processes (one per GPU when the box has that
    many, else sharing cuda:0 -- the kernels then meet through time slicing), SUM with MAX on the radius column,
    bit-identical replicas."""
    import torch.multiprocessing as mp
    from gsb200 import dist as gd
    P = 30011
    port = 29500 + (os.getpid() % 1500) + world
    mp.spawn(_exchange_worker, args=(world, port, P, str(tmp_path)), nprocs=world, join=True)
    got = [np.load(tmp_path / f"r{r}.npz") for r in range(world)]
    for it in range(3):
        parts = [torch.randn((P, gd.WIDTH), generator=torch.Generator().manual_seed(100 * it + r)).numpy() for r in range(world)]
        want = parts[0].copy()
        for p in parts[1:]:   # rank order, like the kernel
            want = want + p
        want[:, gd.MAX_COL] = np.max(np.stack([p[:, gd.MAX_COL] for p in parts]), axis=0)
        for r in range(world):
            assert np.array_equal(got[r][f"arr_{it}"], got[0][f"arr_{it}"]), (it, r)
        assert np.array_equal(got[0][f"arr_{it}"], want), it


@pytest.mark.parametrize("n_groups", [2, 3])
def test_camera_groups_equal_one_batch(n_groups):
    """The rank's cameras as concurrent groups (streams, one arena each, flat record written by group 0 and added to by
    the others) == the same cameras as one batch; also through a captured graph (fork / join as graph branches)."""
    from gsb200 import raster, step
    dev, g, a, cams, sc, bg, targets = _setup(n=3)
    P = a["means3D"].shape[0]
    flat_one = torch.zeros((P, 20), device=dev)
    loss_one, _, _, _ = step.timestep_step(a, sc, targets, bg, None, flat=flat_one)
    cg = step.CameraGroups(sc, targets, n_groups, P, dev).reserve(a, bg)
    flat = torch.full((P, 20), float("nan"), device=dev)
    for graph in (None, step.StepGraph(lambda: cg.step(a, bg, flat))):
        flat.fill_(float("nan"))
        losses, _, sts = cg.step(a, bg, flat) if graph is None else graph.replay()
        torch.cuda.synchronize()
        assert all(s.status()[0] == 0 for s in sts)
        assert torch.allclose(torch.cat(losses), loss_one, rtol=0, atol=1e-6)
        bad, worst = helpers.grad_close(helpers.to_np(flat[:, :15]), helpers.to_np(flat_one[:, :15]), rel=1e-4, floor=1e-5)
        assert bad == 0, (n_groups, bad, worst)
        assert torch.equal(flat[:, 15:17], flat_one[:, 15:17])
        assert float(flat[:, 17:].abs().max()) == 0.0


def test_two_table_walk_equals_two_renders():
    """colors2: both images from one forward walk are bit-identical to two separate renders, and the one-walk backward
    of both image gradients equals mode 0 on the first render + mode 1 on the second."""
    from gsb200 import raster
    dev, g, a, cams, sc, bg, targets = _setup()
    n, P = 3, a["means3D"].shape[0]
    seg = torch.rand((P, 3), generator=torch.Generator().manual_seed(9)).to(dev)
    c1, d1, r1, st1 = _fwd(a, sc, bg, n)
    c2, _, _, st2 = _fwd(dict(a, colors_precomp=seg), sc, bg, n)
    c1, c2 = c1.clone(), c2.clone()
    cf, df, rf, stf = _fwd(a, sc, bg, n, colors2=seg)
    assert torch.equal(cf, c1) and torch.equal(stf.color2, c2) and torch.equal(df, d1) and torch.equal(rf, r1)
    assert torch.equal(stf.tensors()["n_contrib"], st1.tensors()["n_contrib"])
    dL1 = torch.randn(c1.shape, generator=torch.Generator().manual_seed(2)).to(dev)
    dL2 = torch.randn(c1.shape, generator=torch.Generator().manual_seed(3)).to(dev)
    want = torch.zeros((P, 20), device=dev)
    raster.backward_flat(st1, dL1, want)
    raster.backward_flat(st2, dL2, want, mode=1)
    got = torch.full((P, 20), float("nan"), device=dev)
    raster.backward_flat(stf, dL1, got, dL_dpix2=dL2)
    torch.cuda.synchronize()
    for a_, b_ in ((0, 14), (17, 20)):
        bad, worst = helpers.grad_close(helpers.to_np(got[:, a_:b_]), helpers.to_np(want[:, a_:b_]), rel=1e-4, floor=1e-5)
        assert bad == 0, (a_, bad, worst)
    assert torch.equal(got[:, 15:17], want[:, 15:17])
    assert torch.allclose(got[:, 14], want[:, 14], rtol=1e-5, atol=1e-9)
    # the asynchronous path too
    R = st1.num_rendered
    arena = raster.RasterArena(dev).reserve(P, n, 200, 120, capacity=int(1.3 * R))
    ca, _, _, sta = _fwd(a, sc, bg, n, arena=arena, colors2=seg)
    assert torch.equal(ca, c1) and torch.equal(sta.color2, c2)
