"""DeformationNetwork forward on the tcgen05 path against the same architecture in stock
torch ops (train.py:57-127 restated below with nn.Linear / nn.BatchNorm1d / gelu, FP32 matmuls, training-mode
BatchNorm like the reference).  Tolerances: 3xTF32 path 1e-4, TF32 path 2e-2, both relative to the output's RMS --
BatchNorm divides by the batch standard deviation in every block, which carries the matmul rounding forward."""
import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu


class _Block(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.fc1 = nn.Linear(d, d, bias=False); self.bn1 = nn.BatchNorm1d(d)
        self.fc2 = nn.Linear(d, d, bias=False); self.bn2 = nn.BatchNorm1d(d)

    def forward(self, x):
        y = nn.functional.gelu(self.bn1(self.fc1(x)))
        return nn.functional.gelu(self.bn2(self.fc2(y)) + x)


class _StockNetwork(nn.Module):
    def __init__(self, h, blocks):
        super().__init__()
        self.fc_in = nn.Linear(192, h)
        self.residual_blocks = nn.Sequential(*(_Block(h) for _ in range(blocks)))
        self.fc_out = nn.Linear(h, 7)

    def forward(self, init, x):
        return self.fc_out(self.residual_blocks(self.fc_in(x))) + init


def _stock_posenc(x, f):
    freq = ((torch.ones(f) * 2).pow(torch.arange(f)) * torch.pi).to(x.device)
    e = (freq * x[:, :, None]).repeat_interleave(2, 2)
    e[:, :, ::2] = torch.sin(e[:, :, ::2])
    e[:, :, 1::2] = torch.cos(e[:, :, ::2])
    return e.permute(0, 2, 1).flatten(start_dim=1)


def test_positional_encoding_and_normalisation():
    from gsb200 import deform
    g = torch.Generator().manual_seed(0)
    means = (torch.randn(5000, 3, generator=g) * 1.3).cuda()
    rots = torch.randn(5000, 4, generator=g).cuda()
    got = deform.normalize_and_encode_means_and_rotations({"means": means, "rotation_quaternions": rots})

    def norm(t):
        t = t - t.min(dim=0).values
        return 2.0 * t / t.max(dim=0).values - 1.0
    want = torch.cat((_stock_posenc(norm(means), 10), _stock_posenc(norm(rots), 4)), dim=1)
    assert got.shape == (5000, 92)
    # |f u| reaches 2^9 pi: a 1-ulp difference in the float32 argument moves sin by up to 2^9 pi 2^-24 ~ 1e-4
    assert (got - want).abs().max().item() < 3e-4
    assert (got[:, :6] - want[:, :6]).abs().max().item() < 2e-6   # lowest frequency: tight
    prog = torch.full((5000, 1), 7 / 150, device="cuda")
    assert (deform.positional_encoding(prog, 4) - _stock_posenc(prog, 4)).abs().max().item() < 2e-6


@pytest.mark.parametrize("rows,h,blocks", [(20_000, 128, 3), (4097, 64, 1), (300_000, 256, 2)])
@pytest.mark.parametrize("precise", [True, False])
def test_forward_matches_stock_torch(rows, h, blocks, precise):
    from gsb200 import deform
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(rows % 13)
    ref = _StockNetwork(h, blocks).cuda().train()
    with torch.no_grad():   # non-trivial affine parameters and running buffers
        for m in ref.modules():
            if isinstance(m, nn.BatchNorm1d):
                m.weight.uniform_(0.5, 1.5); m.bias.normal_(0, 0.3)
                m.running_mean.normal_(0, 0.1); m.running_var.uniform_(0.5, 2.0)
    ours = deform.DeformationNetwork(h, blocks, precise=precise).cuda()
    ours.load_state_dict(ref.state_dict())   # same parameter names as the reference module
    x = torch.randn(rows, 192, device="cuda")
    init = torch.randn(rows, 7, device="cuda")
    with torch.no_grad():
        want = ref(init, x)
    got = ours(init, network_input=x)
    rms = (want - init).pow(2).mean().sqrt().item()
    err = (got - want).abs().max().item() / rms
    assert err < (1e-4 if precise else 2e-2), err
    for a, b in zip(ours.modules(), ref.modules()):
        if isinstance(a, nn.BatchNorm1d):
            tol = 1e-5 if precise else 2e-3
            assert torch.allclose(a.running_mean, b.running_mean, rtol=tol, atol=tol)
            assert torch.allclose(a.running_var, b.running_var, rtol=tol, atol=tol)
            assert int(a.num_batches_tracked) == int(b.num_batches_tracked) == 1
    # state_dicts stay interchangeable after a forward
    ref.load_state_dict(ours.state_dict())


def test_update_gaussian_cloud_parameters_shapes():
    from gsb200 import deform
    torch.manual_seed(1)
    n = 10_000
    p = {"means": torch.randn(n, 3, device="cuda"), "rotation_quaternions": torch.randn(n, 4, device="cuda"),
         "colors": torch.rand(n, 3, device="cuda")}
    enc = deform.normalize_and_encode_means_and_rotations(p)
    net = deform.DeformationNetwork(64, 2).cuda()
    out = deform.update_gaussian_cloud_parameters(net, p, enc, enc, timestep=3, timestep_count=10)
    assert out["means"].shape == (n, 3) and out["rotation_quaternions"].shape == (n, 4)
    assert out["colors"] is p["colors"]
    assert torch.isfinite(out["means"]).all() and not torch.equal(out["means"], p["means"])


@pytest.mark.parametrize("precise", [True, False])
@pytest.mark.parametrize("rows,h,blocks", [(20_000, 128, 2), (5000, 64, 1), (100_000, 256, 3)])
def test_backward_matches_stock_torch(rows, h, blocks, precise):
    """Gradients of every parameter and of the [N,7] input against torch autograd on the same architecture in FP32.
    precise=True (the default: 3xTF32 forward and input gradients, FP32 weight gradients -- the reference's precision):
    2e-3 of each gradient's RMS; precise=False (TF32 everywhere, the opt-in fast mode): 3e-2 (BatchNorm's 1/std carries
    the TF32 rounding of the forward into every later layer, as in the forward test)."""
    from gsb200 import deform
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(rows % 11)
    ref = _StockNetwork(h, blocks).cuda().train()
    with torch.no_grad():
        for m in ref.modules():
            if isinstance(m, nn.BatchNorm1d):
                m.weight.uniform_(0.5, 1.5); m.bias.normal_(0, 0.3)
    ours = deform.DeformationNetwork(h, blocks, precise=precise).cuda()
    assert deform.DeformationNetwork(h, blocks).precise is True
    ours.load_state_dict(ref.state_dict())
    x = torch.randn(rows, 192, device="cuda")
    init_a = torch.randn(rows, 7, device="cuda", requires_grad=True)
    init_b = init_a.detach().clone().requires_grad_(True)
    w = torch.randn(rows, 7, device="cuda")
    (ref(init_a, x) * w).sum().backward()
    out = ours(init_b, network_input=x)
    assert out.requires_grad
    (out * w).sum().backward()
    assert torch.allclose(init_b.grad, init_a.grad)
    for (name, p), (_, q) in zip(ours.named_parameters(), ref.named_parameters()):
        assert p.grad is not None, name
        rms = q.grad.pow(2).mean().sqrt().item() + 1e-12
        err = (p.grad - q.grad).abs().max().item() / rms
        assert err < (2e-3 if precise else 3e-2), (name, precise, err)


def test_training_steps_follow_stock_torch():
    """Ten Adam steps on a regression target: the loss trajectory follows the stock network's."""
    from gsb200 import deform
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(5)
    ref = _StockNetwork(64, 2).cuda().train()
    ours = deform.DeformationNetwork(64, 2, precise=True).cuda()
    ours.load_state_dict(ref.state_dict())
    x = torch.randn(8192, 192, device="cuda"); init = torch.zeros(8192, 7, device="cuda")
    target = torch.randn(8192, 7, device="cuda") * 0.1
    oa, ob = torch.optim.Adam(ref.parameters(), lr=1e-3), torch.optim.Adam(ours.parameters(), lr=1e-3)
    la, lb = [], []
    for _ in range(10):
        for net, opt, log, call in ((ref, oa, la, lambda: ref(init, x)), (ours, ob, lb, lambda: ours(init, network_input=x))):
            opt.zero_grad()
            loss = (call() - target).pow(2).mean()
            loss.backward()
            opt.step()
            log.append(float(loss.detach()))
    assert lb[-1] < lb[0]
    assert max(abs(a - b) / a for a, b in zip(la, lb)) < 2e-2, (la, lb)


def _row_shard_worker(rank, world, port, n, h, blocks, out_dir):
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "animating-gaussian-splats_b200")):
        sys.path.insert(0, p)
    import torch.distributed as dist
    from gsb200 import deform
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    case = torch.load(os.path.join(out_dir, "case.pt"))
    net = deform.DeformationNetwork(h, blocks).to(dev)
    net.load_state_dict(case["state"])
    net.set_row_shard()
    assert net.shard is not None and net.shard.world == world
    p = {k: v.to(dev) for k, v in case["params"].items()}
    enc0, encp = case["enc0"].to(dev), case["encp"].to(dev)
    w = case["w"].to(dev)
    outs = {}
    for it in range(2):   # twice: the running statistics move, the cached batch row count is reused
        net.zero_grad(set_to_none=True)
        upd = deform.update_gaussian_cloud_parameters(net, p, enc0, encp, timestep=3, timestep_count=10)
        full = torch.cat((upd["means"], upd["rotation_quaternions"]), dim=1)
        assert full.shape == (n, 7)
        (full * w).sum().backward()
        outs[f"out{it}"] = full.detach().cpu()
        for name, q in net.named_parameters():
            outs[f"g{it}.{name}"] = q.grad.detach().cpu()
    for name, b in net.named_buffers():
        outs[f"buf.{name}"] = b.detach().cpu()
    torch.save(outs, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 20_001), (3, 4097)])
def test_row_sharded_network_equals_replicated(tmp_path, world, n):
    """The network on row slices (one process per slice; BatchNorm column sums, the [N, 7] output and the dense-layer
    gradients exchanged) against the same network on all rows in this process: outputs, every parameter gradient and
    the BatchNorm running statistics agree to summation order, and the ranks agree bit for bit."""
    import os
    import torch.multiprocessing as mp
    from gsb200 import deform
    h, blocks = 128, 2
    torch.manual_seed(n)
    ref = deform.DeformationNetwork(h, blocks).cuda()
    with torch.no_grad():
        for m in ref.modules():
            if isinstance(m, nn.BatchNorm1d):
                m.weight.uniform_(0.5, 1.5); m.bias.normal_(0, 0.3)
    params = {"means": torch.randn(n, 3), "rotation_quaternions": torch.randn(n, 4), "colors": torch.rand(n, 3)}
    enc0, encp, w = torch.randn(n, 92), torch.randn(n, 92), torch.randn(n, 7)
    torch.save({"state": {k: v.cpu() for k, v in ref.state_dict().items()}, "params": params, "enc0": enc0, "encp": encp, "w": w},
               tmp_path / "case.pt")
    port = 29500 + (os.getpid() % 1500) + 10 + world
    mp.spawn(_row_shard_worker, args=(world, port, n, h, blocks, str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    p = {k: v.cuda() for k, v in params.items()}
    want = {}
    for it in range(2):
        ref.zero_grad(set_to_none=True)
        upd = deform.update_gaussian_cloud_parameters(ref, p, enc0.cuda(), encp.cuda(), timestep=3, timestep_count=10)
        full = torch.cat((upd["means"], upd["rotation_quaternions"]), dim=1)
        (full * w.cuda()).sum().backward()
        want[f"out{it}"] = full.detach().cpu()
        for name, q in ref.named_parameters():
            want[f"g{it}.{name}"] = q.grad.detach().cpu()
    for name, b in ref.named_buffers():
        want[f"buf.{name}"] = b.detach().cpu()
    assert set(want) == set(got[0])
    for k, v in want.items():
        for r in range(1, world):
            assert torch.equal(got[r][k], got[0][k]), (k, r)            # replicas bit-identical
        a, b = got[0][k].double(), v.double()
        scale = b.pow(2).mean().sqrt().item() + 1e-12
        assert (a - b).abs().max().item() / scale < 2e-4, (k, (a - b).abs().max().item() / scale)
