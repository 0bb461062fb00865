"""GPU parity tests of the rasterizer path, through the C ABI (ctypes -> libgsb200.so).

Checkers: (A) the unmodified reference CUDA extension built into oracle/_ref (bit-exact targets:
radii, tiles_touched, depths, means2D, conic, cov3D, sorted keys + values, tile ranges, n_contrib;
image/depth <= 1e-5 abs; gradients <= 1e-4 rel) and (B) the C oracle (same, except pixels the oracle
flags `fragile`: a CPU cannot reproduce MUFU.EX2 bit for bit).
"""
import numpy as np
import pytest
import torch

import helpers
import oracle_lib
from gsb200 import raster

pytestmark = pytest.mark.gpu

CASES = [
    # name, P, W, H, seed, bg, scale_override
    ("c1_tiny", 16000, 256, 256, 0, (0.0, 0.0, 0.0), None),
    ("ragged_bg", 5000, 250, 131, 1, (0.2, 0.5, 0.9), None),
    ("big_splats", 800, 320, 200, 2, (0.0, 0.0, 0.0), 0.4),
    ("one", 1, 64, 64, 3, (1.0, 1.0, 1.0), 0.3),
    ("panoptic_shape", 60000, 640, 360, 4, (0.0, 0.0, 0.0), None),
]


def run_mine(a, cam, bg, cov=None):
    color, depth, radii, st = raster.forward(
        a["means3D"], a["colors_precomp"], a["opacities"], None if cov is not None else a["scales"],
        None if cov is not None else a["rotations"], cov, bg, 1.0, cam["viewmatrix"].cuda(), cam["projmatrix"].cuda(),
        cam["tanfovx"], cam["tanfovy"], cam["image_height"], cam["image_width"])
    return color, depth, radii, st


@pytest.mark.parametrize("name,P,W,H,seed,bgv,so", CASES)
def test_forward_vs_reference_ext(refdgr, name, P, W, H, seed, bgv, so):
    dev = "cuda"
    _, a, cams = helpers.scene_inputs(P, W, H, seed=seed, device=dev, scale_override=so)
    cam = cams[0]
    bg = torch.tensor(bgv, dtype=torch.float32, device=dev)
    fw = helpers.ref_forward(refdgr, a, cam, bg, dev)
    color, depth, radii, st = run_mine(a, cam, bg)
    torch.cuda.synchronize()
    R = fw["num_rendered"]
    assert st.num_rendered == R
    ref = oracle_lib.decode_reference_buffers(P, H, W, R, fw["geom"], fw["binning"], fw["img"])
    mine = st.tensors()
    assert torch.equal(radii, fw["radii"])
    vis = fw["radii"] > 0
    assert torch.equal(mine["tiles_touched"][0], ref["tiles_touched"])
    # our scan runs in depth order (gauss_order): a stable sort of the visible Gaussians by depth bits
    order = mine["gauss_order"][0].long()
    assert torch.equal(torch.sort(order).values, torch.arange(P, device=dev))
    assert torch.equal(mine["point_offsets"][0], torch.cumsum(ref["tiles_touched"][order], 0).int())
    dkey = torch.where(vis, ref["depths"].view(torch.int32).long() & 0xFFFFFFFF, torch.full_like(vis, 0xFFFFFFFF, dtype=torch.long))
    assert torch.equal(order, torch.sort(dkey, stable=True).indices)
    for k in ("depths", "means2D", "conic_opacity", "cov3D"):
        m, r = mine[k][0][vis], ref[k][vis]
        assert torch.equal(m.view(torch.int32), r.view(torch.int32)), k
    if R > 0:
        assert torch.equal(mine["point_list_keys"], ref["point_list_keys"])
        assert torch.equal(mine["point_list"], ref["point_list"])
    assert torch.equal(mine["ranges"], ref["ranges"])
    assert torch.equal(mine["n_contrib"][0], ref["n_contrib"])
    assert (mine["final_T"][0] - ref["final_T"]).abs().max().item() <= 1e-5
    assert (color - fw["color"]).abs().max().item() <= 1e-5   # north_star: image within 1e-5 abs
    assert (depth - fw["depth"]).abs().max().item() <= 1e-5


@pytest.mark.parametrize("name,P,W,H,seed,bgv,so", CASES)
def test_backward_vs_reference_ext(refdgr, name, P, W, H, seed, bgv, so):
    dev = "cuda"
    _, a, cams = helpers.scene_inputs(P, W, H, seed=seed, device=dev, scale_override=so)
    cam = cams[0]
    bg = torch.tensor(bgv, dtype=torch.float32, device=dev)
    gen = torch.Generator(device="cpu").manual_seed(seed)
    dL = torch.randn((3, H, W), generator=gen).to(dev)
    fw = helpers.ref_forward(refdgr, a, cam, bg, dev)
    gr = helpers.ref_backward(refdgr, a, cam, bg, dev, fw, dL)
    gr2 = helpers.ref_backward(refdgr, a, cam, bg, dev, fw, dL)   # the reference's own atomic-order spread
    color, depth, radii, st = run_mine(a, cam, bg)
    gm = raster.backward(st, dL)
    torch.cuda.synchronize()
    for k in ("means2D", "colors", "opacities", "means3D", "scales", "rotations"):
        spread = float((gr[k] - gr2[k]).abs().max())
        nbad, maxerr = helpers.grad_close(helpers.to_np(gm[k]), helpers.to_np(gr[k]), spread=spread)
        assert nbad == 0, (k, nbad, maxerr, spread)


def test_cov3d_precomp_path(refdgr):
    dev = "cuda"
    P, W, H = 4000, 200, 120
    _, a, cams = helpers.scene_inputs(P, W, H, seed=7, device=dev)
    cam = cams[0]
    bg = torch.zeros(3, device=dev)
    # take the covariances from a scale/rotation run
    _, _, _, st0 = run_mine(a, cam, bg)
    cov = st0.tensors()["cov3D"][0].clone()
    fw = helpers.ref_forward(refdgr, a, cam, bg, dev, cov3D_precomp=cov)
    color, depth, radii, st = run_mine(a, cam, bg, cov=cov)
    assert torch.equal(radii, fw["radii"])
    assert (color - fw["color"]).abs().max().item() <= 1e-5
    dL = torch.randn((3, H, W), generator=torch.Generator().manual_seed(1)).to(dev)
    gr = helpers.ref_backward(refdgr, a, cam, bg, dev, fw, dL, cov3D_precomp=cov)
    gr2 = helpers.ref_backward(refdgr, a, cam, bg, dev, fw, dL, cov3D_precomp=cov)
    gm = raster.backward(st, dL)
    for k in ("means2D", "colors", "opacities", "means3D", "cov3D"):
        spread = float((gr[k] - gr2[k]).abs().max())
        nbad, maxerr = helpers.grad_close(helpers.to_np(gm[k]), helpers.to_np(gr[k]), spread=spread)
        assert nbad == 0, (k, nbad, maxerr, spread)


@pytest.mark.parametrize("name,P,W,H,seed,bgv,so", CASES[:4])
def test_forward_backward_vs_c_oracle(oracle, name, P, W, H, seed, bgv, so):
    dev = "cuda"
    _, a, cams = helpers.scene_inputs(P, W, H, seed=seed, device=dev, scale_override=so)
    cam = cams[0]
    bg = torch.tensor(bgv, dtype=torch.float32, device=dev)
    color, depth, radii, st = run_mine(a, cam, bg)
    n = lambda k: helpers.to_np(a[k])
    o = oracle.forward(n("means3D"), n("colors_precomp"), n("opacities"), n("scales"), n("rotations"),
                       np.asarray(bgv, np.float32), cam["viewmatrix"].numpy(), cam["projmatrix"].numpy(),
                       cam["tanfovx"], cam["tanfovy"], H, W)
    mine = st.tensors()
    assert st.num_rendered == o["num_rendered"]
    assert np.array_equal(helpers.to_np(radii), o["radii"])
    vis = o["radii"] > 0
    assert np.array_equal(helpers.to_np(mine["tiles_touched"][0]).view(np.uint32), o["tiles_touched"])
    order = helpers.to_np(mine["gauss_order"][0])
    assert np.array_equal(helpers.to_np(mine["point_offsets"][0]).view(np.uint32), np.cumsum(o["tiles_touched"][order], dtype=np.uint32))
    for k in ("depths", "means2D", "conic_opacity", "cov3D"):
        assert np.array_equal(helpers.to_np(mine[k][0])[vis].view(np.uint32), o[k][vis].view(np.uint32)), k
    if o["num_rendered"]:
        assert np.array_equal(helpers.to_np(mine["point_list_keys"]).view(np.uint64), o["keys"])
        assert np.array_equal(helpers.to_np(mine["point_list"]).view(np.uint32), o["point_list"])
    assert np.array_equal(helpers.to_np(mine["ranges"]).view(np.uint32), o["ranges"])
    solid = o["fragile"] == 0
    assert np.array_equal(helpers.to_np(mine["n_contrib"][0]).view(np.uint32)[solid], o["n_contrib"][solid])
    assert np.abs(helpers.to_np(color) - o["color"])[:, solid].max() <= 1e-5
    assert np.abs(helpers.to_np(depth) - o["depth"])[:, solid].max() <= 1e-5
    # backward against the oracle on pixels whose forward decisions are solid
    dL = torch.randn((3, H, W), generator=torch.Generator().manual_seed(seed))
    dL = dL * torch.from_numpy(solid.astype(np.float32))
    go = oracle.backward(o, dL.numpy())
    gm = raster.backward(st, dL.to(dev))
    for k in ("means2D", "colors", "opacities", "means3D", "scales", "rotations"):
        nbad, maxerr = helpers.grad_close(helpers.to_np(gm[k]), go[k], rel=2e-4, floor=2e-6)
        assert nbad == 0, (k, nbad, maxerr)


def test_empty_and_culled():
    dev = "cuda"
    cam = helpers.scene.make_cameras(1, 64, 48)[0]
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    # P == 0 short-circuits to zeros (DGR/rasterize_points.cu:81)
    z = lambda *s: torch.zeros(*s, device=dev)
    color, depth, radii, st = raster.forward(z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 4), None, bg, 1.0,
                                             cam["viewmatrix"].cuda(), cam["projmatrix"].cuda(), cam["tanfovx"],
                                             cam["tanfovy"], 48, 64)
    assert color.shape == (3, 48, 64) and float(color.abs().max()) == 0.0 and radii.numel() == 0
    # everything behind the camera: background image, default depth 15, radii 0, zero gradients
    P = 100
    means = torch.zeros(P, 3, device=dev) + torch.tensor([50.0, 50.0, 50.0], device=dev)
    q = torch.zeros(P, 4, device=dev); q[:, 0] = 1
    color, depth, radii, st = raster.forward(means, torch.rand(P, 3, device=dev), torch.full((P, 1), 0.5, device=dev),
                                             torch.full((P, 3), 0.1, device=dev), q, None, bg, 1.0,
                                             cam["viewmatrix"].cuda(), cam["projmatrix"].cuda(), cam["tanfovx"],
                                             cam["tanfovy"], 48, 64)
    assert st.num_rendered == 0 and int(radii.abs().max()) == 0
    assert torch.allclose(color, bg.view(3, 1, 1).expand(3, 48, 64)) and float((depth - 15.0).abs().max()) == 0.0
    g = raster.backward(st, torch.ones(3, 48, 64, device=dev))
    assert all(float(v.abs().max()) == 0.0 for v in g.values() if v is not None)


def test_mark_visible(oracle):
    dev = "cuda"
    _, a, cams = helpers.scene_inputs(3000, 64, 64, seed=5, device=dev, extent=4.0)
    cam = cams[0]
    got = raster.mark_visible(a["means3D"], cam["viewmatrix"].cuda(), cam["projmatrix"].cuda())
    want = oracle.mark_visible(helpers.to_np(a["means3D"]), cam["viewmatrix"].numpy())
    assert np.array_equal(helpers.to_np(got), want)
    assert 0 < want.sum() < want.size


def test_camera_batch_matches_single(refdgr):
    """n_cams > 1 entry point: per-camera outputs identical to single-camera calls, gradients = sum."""
    dev = "cuda"
    P, W, H, n = 20000, 160, 96, 3
    _, a, cams = helpers.scene_inputs(P, W, H, seed=11, n_cams=n, device=dev)
    bg = torch.zeros(3, device=dev)
    sc = helpers.scene.stack_cameras(cams, dev)
    color, depth, radii, st = raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"],
                                             a["rotations"], None, bg, 1.0, sc["viewmatrix"], sc["projmatrix"],
                                             sc["tanfovx"], sc["tanfovy"], H, W, n_cams=n)
    dL = torch.randn((n, 3, H, W), generator=torch.Generator().manual_seed(3)).to(dev)
    gb = raster.backward(st, dL, need_cov3D_grad=False)
    acc = None
    for i, cam in enumerate(cams):
        c1, d1, r1, s1 = run_mine(a, cam, bg)
        assert torch.equal(radii[i], r1)
        assert torch.equal(color[i], c1) and torch.equal(depth[i], d1)
        assert torch.equal(st.tensors()["n_contrib"][i], s1.tensors()["n_contrib"][0])
        g1 = raster.backward(s1, dL[i])
        nbad, maxerr = helpers.grad_close(helpers.to_np(gb["means2D"][i]), helpers.to_np(g1["means2D"]))
        assert nbad == 0, ("means2D", i, nbad, maxerr)
        if acc is None:
            acc = {k: g1[k].double() for k in ("colors", "opacities", "means3D", "scales", "rotations")}
        else:
            for k in acc:
                acc[k] += g1[k].double()
    for k in acc:
        nbad, maxerr = helpers.grad_close(helpers.to_np(gb[k]), helpers.to_np(acc[k]), rel=2e-4)
        assert nbad == 0, (k, nbad, maxerr)


def test_c4_full_size_forward_only(refdgr):
    """BASELINE.json configs[3]: 1 M Gaussians at 1920x1080, forward only (sort / tile-range stress: 8160
    tiles, 45 sort bits in the reference).  One of the eight cameras, full size, against the reference."""
    dev = "cuda"
    P, W, H = 1_000_000, 1920, 1080
    g = helpers.scene.make_gaussians(P, seed=40, scale_override=0.012)
    a = {k: v.to(dev).contiguous() for k, v in helpers.scene.activate(g).items()}
    cam = helpers.scene.make_cameras(8, W, H, seed=40)[3]
    bg = torch.zeros(3, device=dev)
    fw = helpers.ref_forward(refdgr, a, cam, bg, dev)
    color, depth, radii, st = run_mine(a, cam, bg)
    R = fw["num_rendered"]
    assert st.num_rendered == R and R > 5_000_000
    assert st.ctx.sort_bits == 45
    ref = oracle_lib.decode_reference_buffers(P, H, W, R, fw["geom"], fw["binning"], fw["img"])
    mine = st.tensors()
    assert torch.equal(radii, fw["radii"])
    assert torch.equal(mine["point_list_keys"], ref["point_list_keys"])
    assert torch.equal(mine["point_list"], ref["point_list"])
    assert torch.equal(mine["ranges"], ref["ranges"])
    assert torch.equal(mine["n_contrib"][0], ref["n_contrib"])
    assert (color - fw["color"]).abs().max().item() <= 1e-5
    assert (depth - fw["depth"]).abs().max().item() <= 1e-5
    # size-independent properties at full size: ranges tile the list; keys ascending
    rg = mine["ranges"].long()
    assert int((rg[:, 1] - rg[:, 0]).sum()) == R
    assert bool((mine["point_list_keys"][1:] >= mine["point_list_keys"][:-1]).all())


def test_many_gaussians_unpacked_sort_path(refdgr):
    """More than 2^22 Gaussians at 257..1024 tiles (10-bit digit): the one-pass instance sort cannot stage digit and
    Gaussian id as one 32-bit word and takes the separate key / payload path.  Binning bit-exact against the reference."""
    dev = "cuda"
    P, W, H = (1 << 22) + 1234, 512, 384
    g = helpers.scene.make_gaussians(P, seed=41, scale_override=0.004)
    a = {k: v.to(dev).contiguous() for k, v in helpers.scene.activate(g).items()}
    cam = helpers.scene.make_cameras(4, W, H, seed=41)[1]
    bg = torch.zeros(3, device=dev)
    fw = helpers.ref_forward(refdgr, a, cam, bg, dev)
    color, depth, radii, st = run_mine(a, cam, bg)
    R = fw["num_rendered"]
    assert st.num_rendered == R and R > 100_000
    assert int(fw["radii"].argmax()) >= 0 and int((fw["radii"] > 0).nonzero().max()) >= (1 << 22)  # ids beyond 22 bits occur
    ref = oracle_lib.decode_reference_buffers(P, H, W, R, fw["geom"], fw["binning"], fw["img"])
    mine = st.tensors()
    assert torch.equal(radii, fw["radii"])
    assert torch.equal(mine["point_list"], ref["point_list"])
    assert torch.equal(mine["ranges"], ref["ranges"])
    assert torch.equal(mine["n_contrib"][0], ref["n_contrib"])
    assert (color - fw["color"]).abs().max().item() <= 1e-5
