"""N > 1 host logic on CPU: world_size-2 gloo ranks shard the cameras, pack per-rank gradients and
statistic increments into the flat buffer, run the one allreduce and must all end with the single-rank sum."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def fake_camera_grads(cam, P):
    g = torch.Generator().manual_seed(1000 + cam)
    r = lambda *s: torch.randn(*s, generator=g)
    vis = torch.rand(P, generator=g) > 0.3
    grads = {"means3D": r(P, 3), "colors": r(P, 3), "rotations": r(P, 4), "opacities": r(P, 1), "scales": r(P, 3)}
    stats = {"mean_2d_gradients_accumulated": r(P).abs() * vis, "visibility_count": vis.float()}
    radii = (torch.rand(P, generator=g) * 30).floor() * vis
    return grads, stats, radii


def worker(rank, world, port, n_cams, P, out_dir):
    sys.path.insert(0, os.path.join(ROOT, "animating-gaussian-splats_b200"))
    from gsb200 import dist as gd
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = gd.shard_cameras(n_cams, rank, world)
    acc = {k: torch.zeros(P, gd.COLS[k][1] - gd.COLS[k][0]) for k in ("means3D", "colors", "rotations", "opacities", "scales")}
    st = {"mean_2d_gradients_accumulated": torch.zeros(P), "visibility_count": torch.zeros(P)}
    maxr = torch.zeros(P)
    for c in mine:
        g, s, r = fake_camera_grads(c, P)
        for k in g:
            acc[k] += g[k]
        for k in s:
            st[k] += s[k]
        maxr = torch.maximum(maxr, r)
    # one step's record: gradients, THIS step's statistic increments, this step's largest radius
    flat = gd.pack(torch.zeros(P, gd.WIDTH), acc, st, maxr)
    ex = gd.Exchange(P, "cpu")          # CPU tensors: the torch.distributed form of the exchange
    ex.flat.copy_(flat)
    out = gd.unpack(ex.allreduce())
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), maxr=out["max_2d_radii"].numpy(), cams=np.array(mine),
             **{k: v.numpy() for k, v in out.items() if k not in ("max_2d_radii", "segmentation_colors")})
    dist.destroy_process_group()


def test_two_rank_camera_sharding_allreduce(tmp_path):
    world, n_cams, P = 2, 5, 257
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(worker, args=(world, port, n_cams, P, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    assert sorted(list(r0["cams"]) + list(r1["cams"])) == list(range(n_cams))
    # single-rank sum over all cameras
    want = None
    maxr = torch.zeros(P)
    for c in range(n_cams):
        g, s, r = fake_camera_grads(c, P)
        cur = {**{k: v.reshape(P, -1) for k, v in g.items()}, **{k: v.reshape(P, 1) for k, v in s.items()}}
        want = cur if want is None else {k: want[k] + cur[k] for k in cur}
        maxr = torch.maximum(maxr, r)
    for k, v in want.items():
        for rk in (r0, r1):
            assert np.allclose(rk[k].reshape(P, -1), v.numpy(), rtol=1e-5, atol=1e-6), k
    assert np.array_equal(r0["maxr"], maxr.numpy()) and np.array_equal(r1["maxr"], maxr.numpy())
    # replicas end bit-identical
    for k in want:
        assert np.array_equal(r0[k], r1[k])


def test_shard_cameras_cost_balanced():
    """LPT assignment: a partition of the cameras, identical on every rank, never worse than round-robin here."""
    from gsb200 import dist as gd
    costs = [1.0 + 0.6 * ((7 * c) % 5) / 4 + (1.5 if c % 8 == 0 else 0.0) for c in range(27)]
    for world in (1, 2, 4, 8):
        parts = [gd.shard_cameras(27, r, world, costs) for r in range(world)]
        assert sorted(c for p in parts for c in p) == list(range(27))
        assert parts == [gd.shard_cameras(27, r, world, list(costs)) for r in range(world)]
        lpt = max(sum(costs[c] for c in p) for p in parts)
        rr = max(sum(costs[c] for c in gd.shard_cameras(27, r, world)) for r in range(world))
        assert lpt <= rr + 1e-12
        assert lpt <= sum(costs) / world + max(costs)
    assert gd.shard_cameras(27, 3, 8) == list(range(3, 27, 8))


def row_shard_worker(rank, world, port, n, out_dir):
    sys.path.insert(0, os.path.join(ROOT, "animating-gaussian-splats_b200"))
    from gsb200 import deform
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = deform.RowShard()
    r0, r1 = sh.bounds(n)
    full = torch.arange(n * 7, dtype=torch.float32).view(n, 7)
    local = full[r0:r1].clone().requires_grad_(True)
    gathered = deform._GatherRows.apply(local, sh, n)          # [n, 7] on every rank
    w = torch.linspace(0.5, 1.5, n)[:, None]
    (gathered * w).sum().backward()                           # every rank evaluates the same loss on the gathered rows
    sums = torch.full((2, 4), float(rank + 1), dtype=torch.float64)
    sh.all_reduce(sums)
    np.savez(os.path.join(out_dir, f"shard{rank}.npz"), bounds=np.array([r0, r1]), gathered=gathered.detach().numpy(),
             grad=local.grad.numpy(), sums=sums.numpy())
    dist.destroy_process_group()


def test_row_shard_collectives_three_ranks(tmp_path):
    """gsb200.deform.RowShard on CPU tensors: contiguous row slices that cover [0, n) with a shorter last slice, the
    all-gather of the slices and its adjoint (a rank's slice of the gradient), the column-sum all-reduce."""
    world, n = 3, 100
    port = 29500 + (os.getpid() % 2000) + 7
    mp.spawn(row_shard_worker, args=(world, port, n, str(tmp_path)), nprocs=world, join=True)
    got = [np.load(tmp_path / f"shard{r}.npz") for r in range(world)]
    full = np.arange(n * 7, dtype=np.float32).reshape(n, 7)
    w = np.linspace(0.5, 1.5, n, dtype=np.float32)[:, None]
    edges = [tuple(g["bounds"]) for g in got]
    assert edges[0][0] == 0 and edges[-1][1] == n and all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
    assert edges[0][1] - edges[0][0] == 34 and edges[2][1] - edges[2][0] == 32      # ceil(100 / 3) = 34, 34, 32
    for r, g in enumerate(got):
        assert np.array_equal(g["gathered"], full)
        r0, r1 = edges[r]
        assert np.allclose(g["grad"], np.broadcast_to(w[r0:r1], (r1 - r0, 7)))
        assert np.array_equal(g["sums"], np.full((2, 4), 6.0))


def test_shard_bands_partition_properties():
    """Tile-band sharding host logic (gsb200.dist.shard_bands / band_tables / flat_width) on random cost tables: the
    bands tile every (camera, tile row) exactly once in order, a rank holds at most two partial cameras, exactly the
    cameras held by more than one rank get a split slot, the loads stay within one row's cost of the ideal share, and the
    entry / row tables carry one halo tile row per open edge with visibility counted by the band that holds row 0."""
    sys.path.insert(0, os.path.join(ROOT, "animating-gaussian-splats_b200"))
    import random
    from gsb200 import dist as gd
    rng = random.Random(5)
    for trial in range(60):
        n_cams, gy, world = rng.randint(1, 9), rng.randint(1, 23), rng.randint(1, 8)
        costs = [[rng.uniform(0.1, 3.0) for _ in range(gy)] for _ in range(n_cams)]
        bands, slots = gd.shard_bands(costs, world)
        assert len(bands) == world
        seen = []
        for mine in bands:
            partial = 0
            for cam, r0, r1 in mine:
                assert 0 <= r0 < r1 <= gy
                partial += (r1 - r0) < gy
                seen += [(cam, r) for r in range(r0, r1)]
            assert partial <= 2
        assert seen == [(c, r) for c in range(n_cams) for r in range(gy)]           # exactly once, in order
        owners = {}
        for rk, mine in enumerate(bands):
            for cam, _, _ in mine:
                owners.setdefault(cam, set()).add(rk)
        split = sorted(c for c, o in owners.items() if len(o) > 1)
        assert sorted(slots) == split and sorted(slots.values()) == list(range(len(split)))
        total = sum(map(sum, costs))
        biggest = max(map(max, costs))
        for mine in bands:
            load = sum(costs[cam][r] for cam, r0, r1 in mine for r in range(r0, r1))
            assert load <= total / world + biggest + 1e-9
        assert gd.flat_width(len(split)) % 4 == 0 and gd.flat_width(len(split)) >= gd.WIDTH + 2 * len(split)
        height = 16 * gy - rng.randint(0, 15)
        for mine in bands:
            if not mine:
                continue
            ent, rows = gd.band_tables(mine, slots, gy, height, "cpu")
            for (cam, r0, r1), e, rw in zip(mine, ent.tolist(), rows.tolist()):
                assert e[0] == max(r0 - 1, 0) and e[1] == min(r1 + 1, gy)
                assert e[2] == slots.get(cam, -1) and e[3] == (1 if r0 == 0 else 0)
                assert rw == [16 * r0, min(16 * r1, height), 16 * e[0], min(16 * e[1], height)]
