"""Per-timestep training step (gsb200.timestep_train, the train.py:700-790 caller): the batched image loss node
against the per-view drop-in autograd path, the total loss composition of train.py:385-433, and a short training run
through DeformationNetwork -> rasterizer -> losses -> Adam."""
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


def _scene(P=5000, n_cams=3, W=96, H=64, seed=0):
    from gsb200 import scene
    dev = torch.device("cuda")
    raw = {k: v.to(dev) for k, v in scene.make_gaussians(P, seed=seed, extent=1.0).items()}
    cams = scene.make_cameras(n_cams, W, H, seed=seed)
    return raw, cams, scene.stack_cameras(cams, dev), [scene.settings_for(c, dev) for c in cams]


def test_batched_image_loss_matches_per_view_autograd():
    from diff_gaussian_rasterization import GaussianRasterizer
    from gsb200 import loss as gloss, optim, timestep_train as tt
    raw, cams, stacked, settings = _scene()
    targets = torch.rand(3, 3, 64, 96, device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
    keys = ("means", "rotation_quaternions", "colors", "opacity_logits", "log_scales")
    a = {k: (raw[k].clone().requires_grad_(True) if k in keys else raw[k]) for k in raw}
    b = {k: (raw[k].clone().requires_grad_(True) if k in keys else raw[k]) for k in raw}
    total = tt.batched_image_loss(a, stacked, targets)
    (2.0 * total).backward()
    want = 0
    for i, s in enumerate(settings):
        args = optim.create_render_arguments(b)
        im, _, _ = GaussianRasterizer(s)(**args)
        want = want + gloss.l1_ssim_loss(im, targets[i])
    (2.0 * want).backward()
    assert abs(float(total.detach()) - float(want.detach())) < 1e-5 * abs(float(want.detach()))
    for k in keys:
        bad, worst = helpers.grad_close(helpers.to_np(a[k].grad), helpers.to_np(b[k].grad), rel=1e-4, floor=2e-6)
        assert bad == 0, (k, bad, worst)


def test_train_steps_reduce_the_loss():
    from gsb200 import deform, rigidity, timestep_train as tt, io as gio
    torch.manual_seed(0)
    raw, cams, stacked, settings = _scene(P=4000, n_cams=4, seed=1)
    dev = raw["means"].device
    # targets: the same cloud translated by 3 cm -- a motion the network can learn
    moved = dict(raw); moved["means"] = raw["means"] + torch.tensor([0.03, 0.0, 0.0], device=dev)
    from gsb200 import step as gstep, optim
    rot, opa, sca = optim.activate(moved)
    with torch.no_grad():
        from gsb200 import raster
        targets, _, _, _ = raster.forward(moved["means"], moved["colors"], opa, sca, rot, None, torch.zeros(3, device=dev), 1.0,
                                          stacked["viewmatrix"], stacked["projmatrix"], stacked["tanfovx"], stacked["tanfovy"],
                                          64, 96, n_cams=4)
    init = {k: v.clone() for k, v in raw.items()}
    ninfo = rigidity.create_neighbor_info(init, k=8)
    enc0 = deform.normalize_and_encode_means_and_rotations(init)
    fg0 = rigidity.create_foreground_info(init, ninfo.indices)
    net = deform.DeformationNetwork(64, 1, precise=True).to(dev)
    opt = torch.optim.Adam(net.parameters(), lr=2e-3)
    losses, enc, fg = [], enc0, fg0
    for it in range(25):
        loss, enc_next, fg_next = tt.train_step(net, opt, init, enc0, enc0, fg0, ninfo, (stacked, targets), 1, 1)
        losses.append(float(loss))
    assert all(torch.isfinite(torch.tensor(losses)))
    assert enc_next.shape == (4000, 92) and fg_next.offsets_to_neighbors.shape[1:] == (8, 3)
    assert sum(losses[-5:]) / 5 < 0.9 * sum(losses[:5]) / 5, losses


def test_view_arena_renders_without_read_back_and_matches_the_blocking_path():
    """ViewArena: the step through the capacity-guarded asynchronous forward equals the blocking path (same kernels),
    over several steps with a changing number of views; a far too small arena is reported by poll() as ArenaOverflow."""
    from gsb200 import raster, timestep_train as tt
    raw, cams, stacked, settings = _scene(P=6000, n_cams=3)
    keys = ("means", "rotation_quaternions", "colors", "opacity_logits", "log_scales")
    gen = torch.Generator(device="cuda").manual_seed(4)
    va = tt.ViewArena()
    for n in (3, 2, 3, 3):
        sub = {"viewmatrix": stacked["viewmatrix"][:n].contiguous(), "projmatrix": stacked["projmatrix"][:n].contiguous(),
               "tanfovx": stacked["tanfovx"][:n], "tanfovy": stacked["tanfovy"][:n], "image_width": 96, "image_height": 64}
        targets = torch.rand(n, 3, 64, 96, device="cuda", generator=gen)
        a = {k: (raw[k].clone().requires_grad_(True) if k in keys else raw[k]) for k in raw}
        b = {k: (raw[k].clone().requires_grad_(True) if k in keys else raw[k]) for k in raw}
        la = tt.batched_image_loss(a, dict(sub), targets, view_arena=va)
        la.backward()
        lb = tt.batched_image_loss(b, dict(sub), targets)
        lb.backward()
        assert float(la.detach()) == float(lb.detach())
        for k in keys:
            bad, worst = helpers.grad_close(helpers.to_np(a[k].grad), helpers.to_np(b[k].grad), rel=1e-4, floor=2e-6)
            assert bad == 0, (k, bad, worst)
    va.poll(wait=True)
    assert va.steps_monitored == 4 and len(va.arenas) == 2 and not va.pending
    # an arena that cannot hold the instances: the step renders background, poll() reports it
    key = next(k for k in va.arenas if k[1] == 3)
    small = raster.RasterArena(raw["means"].device).reserve(6000, 3, 96, 64, 64)
    va.arenas[key] = small
    sub = {"viewmatrix": stacked["viewmatrix"], "projmatrix": stacked["projmatrix"], "tanfovx": stacked["tanfovx"],
           "tanfovy": stacked["tanfovy"], "image_width": 96, "image_height": 64}
    tt.batched_image_loss({k: v for k, v in raw.items()}, sub, torch.rand(3, 3, 64, 96, device="cuda", generator=gen), view_arena=va)
    with pytest.raises(raster.ArenaOverflow):
        va.poll(wait=True)
    assert key not in va.arenas


def _sharded_case(dev):
    """The same small training problem in every process: a cloud, 4 views, targets = the cloud moved by 3 cm."""
    from gsb200 import deform, optim, raster, rigidity
    torch.manual_seed(0)
    raw, cams, stacked, settings = _scene(P=4000, n_cams=4, seed=1)
    moved = dict(raw); moved["means"] = raw["means"] + torch.tensor([0.03, 0.0, 0.0], device=dev)
    rot, opa, sca = optim.activate(moved)
    with torch.no_grad():
        targets, _, _, _ = raster.forward(moved["means"], moved["colors"], opa, sca, rot, None, torch.zeros(3, device=dev), 1.0,
                                          stacked["viewmatrix"], stacked["projmatrix"], stacked["tanfovx"], stacked["tanfovy"],
                                          64, 96, n_cams=4)
    init = {k: v.clone() for k, v in raw.items()}
    ninfo = rigidity.create_neighbor_info(init, k=8)
    enc0 = deform.normalize_and_encode_means_and_rotations(init)
    fg0 = rigidity.create_foreground_info(init, ninfo.indices)
    torch.manual_seed(3)
    net = deform.DeformationNetwork(64, 1, precise=True).to(dev)
    return stacked, targets.clone(), init, ninfo, enc0, fg0, net


def _sharded_train_worker(rank, world, port, steps, out_dir):
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "animating-gaussian-splats_b200"), os.path.join(root, "tests")):
        sys.path.insert(0, p)
    import torch.distributed as dist
    from gsb200 import dist as gd, timestep_train as tt
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    stacked, targets, init, ninfo, enc0, fg0, net = _sharded_case(dev)
    net.set_row_shard()
    opt = torch.optim.Adam(net.parameters(), lr=2e-3)
    ex = gd.Exchange(init["means"].shape[0], dev)
    va = tt.ViewArena()
    mine = list(range(rank, 4, world))
    sub = {"viewmatrix": stacked["viewmatrix"][mine].contiguous(), "projmatrix": stacked["projmatrix"][mine].contiguous(),
           "tanfovx": [stacked["tanfovx"][i] for i in mine], "tanfovy": [stacked["tanfovy"][i] for i in mine],
           "image_width": 96, "image_height": 64}
    losses = []
    for _ in range(steps):
        loss, _, _ = tt.train_step(net, opt, init, enc0, enc0, fg0, ninfo, (sub, targets[mine].contiguous()), 1, 1,
                                   exchange=ex, view_count=4, view_arena=va)
        losses.append(float(loss))
    va.poll(wait=True)
    torch.save({"losses": losses, "state": {k: v.cpu() for k, v in net.state_dict().items()}}, os.path.join(out_dir, f"t{rank}.pt"))
    ex.close()
    dist.destroy_process_group()


def test_two_process_training_equals_one_process(tmp_path):
    """train.py's step with the views sharded over two processes (one flat-record exchange per step), the network
    row-sharded (BatchNorm sums, [N, 7] output, dense-layer gradients exchanged) and host-free rendering, against the same
    steps in this process: the loss follows the same trajectory and both ranks hold bit-identical networks."""
    import os
    import torch.multiprocessing as mp
    from gsb200 import timestep_train as tt
    steps = 6
    port = 29500 + (os.getpid() % 1500) + 20
    mp.spawn(_sharded_train_worker, args=(2, port, steps, str(tmp_path)), nprocs=2, join=True)
    got = [torch.load(tmp_path / f"t{r}.pt") for r in range(2)]
    dev = torch.device("cuda")
    stacked, targets, init, ninfo, enc0, fg0, net = _sharded_case(dev)
    opt = torch.optim.Adam(net.parameters(), lr=2e-3)
    want = []
    for _ in range(steps):
        loss, _, _ = tt.train_step(net, opt, init, enc0, enc0, fg0, ninfo, (stacked, targets), 1, 1)
        want.append(float(loss))
    assert got[0]["losses"] == got[1]["losses"]
    for k, v in got[0]["state"].items():
        assert torch.equal(v, got[1]["state"][k]), k
    for a, b in zip(got[0]["losses"], want):
        assert abs(a - b) <= 2e-3 * abs(b), (got[0]["losses"], want)
