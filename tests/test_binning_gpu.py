"""Scan / radix sort entry points of the C ABI against torch: exact integer results, stability, edge sizes."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from gsb200 import _lib as L
    return L, L.lib()


@pytest.mark.parametrize("n", [1, 7, 2048, 2049, 100_003, 3_000_000])
def test_inclusive_scan(n):
    L, lib = _lib()
    x = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda")
    out = torch.empty_like(x)
    scratch = torch.empty(lib.gsb_scan_scratch_bytes(n), dtype=torch.uint8, device="cuda")
    L.check(lib.gsb_inclusive_scan_u32(n, L.ptr(x), L.ptr(out), L.ptr(scratch), L.stream_ptr()), "scan")
    assert torch.equal(out, torch.cumsum(x, 0).int())


@pytest.mark.parametrize("n,end_bit", [(1, 42), (4095, 41), (4097, 45), (1_000_001, 42), (300_000, 37)])
def test_sort_pairs_u64_stable(n, end_bit):
    L, lib = _lib()
    g = torch.Generator(device="cuda").manual_seed(n)
    # few distinct keys in the low bits so that ties (stability) are exercised; garbage above end_bit is ignored
    keys = torch.randint(0, 1 << 20, (n,), dtype=torch.int64, device="cuda", generator=g) << (end_bit - 20)
    keys = keys | torch.randint(0, 4, (n,), dtype=torch.int64, device="cuda", generator=g)
    keys = keys | (torch.randint(0, 2, (n,), dtype=torch.int64, device="cuda", generator=g) << (end_bit + 3))
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    keys_copy = keys.clone()
    temp = torch.empty(lib.gsb_sort_temp_bytes(n), dtype=torch.uint8, device="cuda")
    L.check(lib.gsb_sort_pairs(n, L.ptr(keys), L.ptr(ko), L.ptr(vals), L.ptr(vo), end_bit, L.ptr(temp), L.stream_ptr()), "sort")
    masked = keys & ((1 << end_bit) - 1)
    order = torch.sort(masked, stable=True).indices
    assert torch.equal(vo.long(), order)
    assert torch.equal(ko, keys[order])
    assert torch.equal(keys, keys_copy)   # the input is never written


@pytest.mark.parametrize("n,end_bit", [(5, 10), (8191, 15), (2_000_000, 13)])
def test_sort_pairs_u32_stable(n, end_bit):
    L, lib = _lib()
    g = torch.Generator(device="cuda").manual_seed(n)
    keys = torch.randint(0, 1 << end_bit, (n,), dtype=torch.int32, device="cuda", generator=g)
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    temp = torch.empty(lib.gsb_sort_temp_bytes(n), dtype=torch.uint8, device="cuda")
    L.check(lib.gsb_sort_pairs_u32(n, L.ptr(keys), L.ptr(ko), L.ptr(vals), L.ptr(vo), end_bit, L.ptr(temp), L.stream_ptr()), "sort32")
    order = torch.sort(keys.long(), stable=True).indices
    assert torch.equal(vo.long(), order) and torch.equal(ko, keys[order])


def test_tile_range_coverage_property():
    """sum(range.y - range.x) == R and ranges tile the sorted list without gaps."""
    import helpers
    from gsb200 import raster
    _, a, cams = helpers.scene_inputs(30000, 320, 200, seed=9, device="cuda")
    cam = cams[0]
    _, _, _, st = raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None,
                                 torch.zeros(3, device="cuda"), 1.0, cam["viewmatrix"].cuda(), cam["projmatrix"].cuda(),
                                 cam["tanfovx"], cam["tanfovy"], 200, 320)
    t = st.tensors()
    rg = t["ranges"].long()
    assert int((rg[:, 1] - rg[:, 0]).sum()) == st.num_rendered
    nz = rg[rg[:, 1] > rg[:, 0]]
    assert torch.equal(nz[1:, 0], nz[:-1, 1]) and int(nz[0, 0]) == 0 and int(nz[-1, 1]) == st.num_rendered
    keys = t["point_list_keys"]
    assert bool((keys[1:] >= keys[:-1]).all())
