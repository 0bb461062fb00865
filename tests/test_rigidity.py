"""Rigidity loss.  CPU: the numpy oracle against outputs of the reference's own
calculate_rigidity_loss + autograd (tests/golden/rigidity_*.npz, made by tests/golden/make_golden_rigidity.py).
GPU: gsb_rigidity_loss / gsb_foreground_info through gsb200.rigidity against the goldens and the oracle.
Tolerances: loss 1e-5 rel, gradients 1e-4 rel (north_star's bar for parameter gradients) with an absolute floor
of 2e-6 x the largest magnitude (float32 sums of cancelling terms)."""
import os

import numpy as np
import pytest
import torch

from oracle import rigidity_ref

GOLD = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["small", "weights", "k5"]


def _load(name):
    g = np.load(os.path.join(GOLD, f"rigidity_{name}.npz"))
    return g, np.nonzero(g["seg"][:, 0] > 0.5)[0]


def _close(a, b, rel=1e-4, floor=2e-6):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    tol = rel * np.maximum(np.abs(a), np.abs(b)) + floor * max(np.abs(b).max(), 1e-30)
    return int((np.abs(a - b) > tol).sum()), float(np.abs(a - b).max())


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
    g, fg = _load(name)
    loss, dm, dq = rigidity_ref.rigidity_loss(g["means"], g["quats"], fg, g["nbr"], g["weights"], g["prev_inv_rot"],
                                              g["prev_offsets"])
    assert abs(loss - float(g["loss"])) <= 1e-5 * float(g["loss"])
    assert _close(dm, g["d_means"])[0] == 0 and _close(dq, g["d_quats"])[0] == 0
    inv, off = rigidity_ref.foreground_info(g["means0"], g["quats0"], fg, g["nbr"])
    assert np.allclose(inv, g["prev_inv_rot"], rtol=0, atol=5e-7) and np.allclose(off, g["prev_offsets"], rtol=0, atol=5e-7)


def test_oracle_gradient_is_the_derivative():
    """Central differences in float64 on a few coordinates (the oracle's analytic chain stands on its own)."""
    g, fg = _load("weights")
    args = (fg, g["nbr"], g["weights"], g["prev_inv_rot"], g["prev_offsets"])
    _, dm, dq = rigidity_ref.rigidity_loss(g["means"], g["quats"], *args)
    rng = np.random.default_rng(0)
    for table, grad, k in ((0, dm, 3), (1, dq, 4)):
        for _ in range(6):
            r, c = int(rng.choice(fg)), int(rng.integers(k))
            t = [g["means"].astype(np.float64), g["quats"].astype(np.float64)]
            h = 1e-6
            t[table][r, c] += h
            up = rigidity_ref.rigidity_loss(t[0], t[1], *args)[0]
            t[table][r, c] -= 2 * h
            dn = rigidity_ref.rigidity_loss(t[0], t[1], *args)[0]
            assert abs((up - dn) / (2 * h) - grad[r, c]) <= 1e-5 * abs(grad[r, c]) + 1e-9


def _gpu_case(g):
    dev = torch.device("cuda")
    t = lambda k, dt=torch.float32: torch.tensor(g[k]).to(dev).to(dt).contiguous()
    params = {"means": t("means").requires_grad_(True), "rotation_quaternions": t("quats").requires_grad_(True),
              "segmentation_masks": t("seg")}
    return params, t("nbr", torch.int64), t("weights")


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_kernel_matches_reference_golden(name):
    from gsb200 import rigidity as R
    g, fg = _load(name)
    params, nbr, w = _gpu_case(g)
    prev = R.ForegroundInfo(inverted_rotations=torch.tensor(g["prev_inv_rot"]).cuda(),
                            offsets_to_neighbors=torch.tensor(g["prev_offsets"]).cuda())   # reference layout [F,K,3]
    loss = R.calculate_rigidity_loss(params, R.NeighborInfo(weights=w, indices=nbr), prev)
    assert loss.dtype == torch.float32 and loss.dim() == 0
    assert abs(float(loss.detach()) - float(g["loss"])) <= 1e-5 * float(g["loss"])
    (3.0 * loss).backward()   # the weight train.py:425 gives the term
    bad, worst = _close(params["means"].grad.cpu().numpy() / 3.0, g["d_means"])
    assert bad == 0, (bad, worst)
    bad, worst = _close(params["rotation_quaternions"].grad.cpu().numpy() / 3.0, g["d_quats"])
    assert bad == 0, (bad, worst)
    # background rows get no gradient
    bgrows = np.nonzero(g["seg"][:, 0] <= 0.5)[0]
    assert not params["means"].grad[bgrows].any() and not params["rotation_quaternions"].grad[bgrows].any()


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_foreground_info_kernel(name):
    from gsb200 import rigidity as R
    g, fg = _load(name)
    dev = torch.device("cuda")
    params = {"means": torch.tensor(g["means0"]).to(dev), "rotation_quaternions": torch.tensor(g["quats0"]).to(dev),
              "segmentation_masks": torch.tensor(g["seg"]).to(dev)}
    nbr = torch.tensor(g["nbr"]).to(dev)
    info = R.create_foreground_info(params, nbr)
    assert info.offsets_to_neighbors.shape == g["prev_offsets"].shape
    assert np.allclose(info.inverted_rotations.cpu().numpy(), g["prev_inv_rot"], rtol=0, atol=3e-7)
    assert np.array_equal(info.offsets_to_neighbors.cpu().numpy(), g["prev_offsets"])   # plain float32 subtraction
    # and the strided [K,3,F] storage feeds the loss kernel directly
    p2, _, w = _gpu_case(g)
    a = R.calculate_rigidity_loss(p2, R.NeighborInfo(weights=w, indices=nbr), info)
    assert abs(float(a.detach()) - float(g["loss"])) <= 1e-5 * float(g["loss"])


@pytest.mark.gpu
def test_full_size_properties_and_no_grad():
    """BASELINE-sized cloud (300 k Gaussians, ~60% foreground, K = 20): unmoved cloud -> loss at the sqrt(1e-20) floor;
    a rigid motion of the whole cloud leaves the loss at the floor as well (the loss measures non-rigidity);
    the oracle agrees on a random sub-problem; the no-grad path launches the loss-only kernel."""
    from gsb200 import rigidity as R
    dev = torch.device("cuda")
    gen = torch.Generator().manual_seed(5)
    P, K = 300_000, 20
    means = (torch.rand(P, 3, generator=gen) - 0.5) * 3
    quats = torch.randn(P, 4, generator=gen)
    seg = torch.zeros(P, 3); seg[:, 0] = (torch.rand(P, generator=gen) > 0.4).float(); seg[:, 2] = 1 - seg[:, 0]
    F = int(seg[:, 0].sum())
    # foreground points along a random walk (2 cm steps), neighbours = the 10 predecessors and 10 successors on the
    # walk: offsets of a few cm like a real 20-NN table, at BASELINE's size
    walk = torch.cumsum(0.02 * torch.randn(F, 3, generator=gen), 0)
    means[seg[:, 0] > 0.5] = walk - walk.mean(0)
    steps = torch.tensor([d for d in range(-10, 11) if d != 0])
    nbr = torch.arange(F)[:, None] + steps[None]
    nbr = torch.where((nbr < 0) | (nbr >= F), torch.arange(F)[:, None] - steps[None], nbr)   # reflect at the ends
    w = torch.rand(F, K, generator=gen)
    base = {"means": means.to(dev), "rotation_quaternions": quats.to(dev), "segmentation_masks": seg.to(dev)}
    ninfo = R.NeighborInfo(weights=w.to(dev), indices=nbr.to(dev))
    prev = R.create_foreground_info(base, ninfo.indices)
    with torch.no_grad():
        still = float(R.calculate_rigidity_loss(base, ninfo, prev))
    assert 1e-10 <= still < 1e-6, still   # sqrt(1e-20) floor + float32 rounding of q (x) conj(q) on ~1.5 m offsets
    # rigid motion: rotate every mean and every orientation by the same rotation Q, translate by t
    Q = torch.tensor([0.9, 0.1, -0.3, 0.2]); Q = Q / Q.norm()
    Rm = torch.tensor(rigidity_ref.build_rotation(Q[None].numpy().astype(np.float64))[0][0], dtype=torch.float32)
    qn = torch.nn.functional.normalize(quats)
    moved_q = torch.tensor(rigidity_ref.quat_mult(np.tile(Q.numpy(), (P, 1)), qn.numpy()))
    moved = {"means": (means @ Rm.T + torch.tensor([0.3, -0.2, 0.1])).to(dev), "rotation_quaternions": moved_q.float().to(dev),
             "segmentation_masks": base["segmentation_masks"]}
    with torch.no_grad():
        rigid = float(R.calculate_rigidity_loss(moved, ninfo, prev))
    assert rigid < 5e-6, rigid   # float32 rounding of metre-sized coordinates; a non-rigid 1 cm jitter gives ~1e-2
    jit = {"means": (means + 0.01 * torch.randn(P, 3, generator=gen)).to(dev).requires_grad_(True),
           "rotation_quaternions": quats.to(dev).requires_grad_(True), "segmentation_masks": base["segmentation_masks"]}
    loss = R.calculate_rigidity_loss(jit, ninfo, prev)
    loss.backward()
    assert float(loss.detach()) > 1e-3
    # oracle on the full problem (numpy, ~1 s)
    fg = np.nonzero(seg[:, 0].numpy() > 0.5)[0]
    want, dm, dq = rigidity_ref.rigidity_loss(jit["means"].detach().cpu().numpy(), quats.numpy(), fg, nbr.numpy(), w.numpy(),
                                              prev.inverted_rotations.cpu().numpy(),
                                              prev.offsets_to_neighbors.cpu().numpy())
    assert abs(float(loss.detach()) - want) <= 1e-5 * want
    assert _close(jit["means"].grad.cpu().numpy(), dm)[0] == 0
    # the quaternion gradient is a projection (dq - q (q . dq)) / |q|: components nearly parallel to q cancel, so the
    # absolute floor is taken per row (2e-5 x the row's largest component) instead of per table
    got = jit["rotation_quaternions"].grad.cpu().numpy().astype(np.float64)
    tol = 1e-4 * np.abs(dq) + 2e-5 * np.abs(dq).max(axis=1, keepdims=True) + 1e-12
    assert int((np.abs(got - dq) > tol).sum()) == 0
