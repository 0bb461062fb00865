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


def _seg_sort_case(key_bits, n, seg_lens, end_bit, digit_bits, iota):
    """Segmented wide-digit onesweep (csrc/segsort.cu) against a stable torch sort per segment."""
    L, lib = _lib()
    g = torch.Generator(device="cuda").manual_seed(n + end_bit + digit_bits)
    kdt = torch.int16 if key_bits == 16 else torch.int32
    keys = torch.randint(0, 1 << end_bit, (max(n, 1),), dtype=torch.int64, device="cuda", generator=g)[:n]
    if end_bit > 12:   # few distinct low bits: ties exercise stability across passes
        keys = (keys >> 6 << 6) | torch.randint(0, 3, (n,), dtype=torch.int64, device="cuda", generator=g)
    kin = keys
    kin_t = keys.to(kdt)   # wraps: the bit pattern is what the kernels see
    vals = None if iota else torch.randint(0, 1 << 30, (n,), dtype=torch.int32, device="cuda", generator=g)
    seg = torch.tensor([0] + list(torch.tensor(seg_lens).cumsum(0)), dtype=torch.int32, device="cuda")
    assert int(seg[-1]) == n
    n_seg = len(seg_lens)
    passes = (end_bit + digit_bits - 1) // digit_bits
    ko = torch.empty_like(kin_t) if passes > 1 or not iota else None
    vo = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    temp = torch.empty(lib.gsb_seg_sort_temp_bytes(n, n_seg, key_bits // 8, end_bit, digit_bits), dtype=torch.uint8, device="cuda")
    fn = lib.gsb_seg_sort_u16 if key_bits == 16 else lib.gsb_seg_sort_u32
    L.check(fn(n, n_seg, L.ptr(seg), L.ptr(kin_t), L.ptr(ko), L.ptr(vals), L.ptr(vo), end_bit, digit_bits, L.ptr(temp),
               L.stream_ptr()), "seg_sort")
    torch.cuda.synchronize()
    segc = seg.cpu().tolist()
    for s in range(n_seg):
        a, b = segc[s], segc[s + 1]
        if a == b:
            continue
        order = torch.sort(kin[a:b], stable=True).indices
        want_v = order.int() if iota else vals[a:b][order]
        assert torch.equal(vo[a:b], want_v), (s, a, b)
        if ko is not None:
            assert torch.equal(ko[a:b], kin_t[a:b][order]), (s, a, b)


@pytest.mark.parametrize("n_seg,digit_bits,end_bit", [(1, 10, 10), (3, 10, 10), (27, 10, 10), (5, 8, 8), (2, 8, 13),
                                                       (4, 5, 10), (3, 5, 5), (1, 10, 7), (6, 8, 16)])
def test_seg_sort_u16(n_seg, digit_bits, end_bit):
    gen = torch.Generator().manual_seed(n_seg * 100 + end_bit)
    lens = torch.randint(0, 20000, (n_seg,), generator=gen).tolist()
    if n_seg >= 3:
        lens[1] = 0            # an empty segment
        lens[2] = 4096         # exactly one sort tile
    lens[-1] = max(lens[-1], 1)
    _seg_sort_case(16, sum(lens), lens, end_bit, digit_bits, iota=False)


@pytest.mark.parametrize("n_seg,P,digit_bits,iota", [(1, 1, 11, True), (1, 100_003, 11, True), (4, 4097, 11, True),
                                                      (27, 30_000, 11, True), (3, 12_345, 8, True), (2, 50_000, 8, False)])
def test_seg_sort_u32_depth_keys(n_seg, P, digit_bits, iota):
    _seg_sort_case(32, n_seg * P, [P] * n_seg, 32, digit_bits, iota)


def _seg_sort_keys(keys, seg_lens, end_bit, digit_bits, key_bits=16):
    """Given keys (no random draw): segmented sort against a stable torch sort per segment."""
    L, lib = _lib()
    n = keys.numel()
    kdt = torch.int16 if key_bits == 16 else torch.int32
    kin_t = keys.to(kdt)
    vals = torch.arange(n, dtype=torch.int32, device="cuda").flip(0).contiguous()
    seg = torch.tensor([0] + list(torch.tensor(seg_lens).cumsum(0)), dtype=torch.int32, device="cuda")
    n_seg = len(seg_lens)
    passes = (end_bit + digit_bits - 1) // digit_bits
    ko = torch.empty_like(kin_t) if passes > 1 else None
    vo = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    temp = torch.empty(lib.gsb_seg_sort_temp_bytes(n, n_seg, key_bits // 8, end_bit, digit_bits), dtype=torch.uint8, device="cuda")
    fn = lib.gsb_seg_sort_u16 if key_bits == 16 else lib.gsb_seg_sort_u32
    L.check(fn(n, n_seg, L.ptr(seg), L.ptr(kin_t), L.ptr(ko), L.ptr(vals), L.ptr(vo), end_bit, digit_bits, L.ptr(temp),
               L.stream_ptr()), "seg_sort")
    torch.cuda.synchronize()
    segc = seg.cpu().tolist()
    for s in range(n_seg):
        a, b = segc[s], segc[s + 1]
        if a < b:
            order = torch.sort(keys[a:b], stable=True).indices
            assert torch.equal(vo[a:b], vals[a:b][order]), (s, a, b)
            if ko is not None:
                assert torch.equal(ko[a:b], kin_t[a:b][order]), (s, a, b)


@pytest.mark.parametrize("digit_bits,end_bit", [(10, 10), (8, 13), (5, 10)])
def test_seg_sort_fully_skewed_keys(digit_bits, end_bit):
    """Every instance of a camera in ONE tile (one digit value holds the whole segment: 16-bit per-warp counters and
    per-tile u16 rows at their limits, every lane of every warp colliding on one counter), one different tile per
    camera, and a camera whose instances alternate between the two extreme tile ids."""
    top = (1 << end_bit) - 1
    lens = [70_001, 4096 * 16, 33_333, 5, 131_072]
    parts = [torch.full((lens[0],), 0, dtype=torch.int64), torch.full((lens[1],), top, dtype=torch.int64),
             torch.full((lens[2],), 517 & top, dtype=torch.int64), torch.full((lens[3],), 1, dtype=torch.int64),
             torch.arange(lens[4], dtype=torch.int64) % 2 * top]
    _seg_sort_keys(torch.cat(parts).cuda(), lens, end_bit, digit_bits)


def test_seg_sort_depth_keys_all_equal():
    """u32 depth keys with one value per camera (a wall of Gaussians at one depth): order must stay the input order."""
    P = 50_000
    keys = torch.cat([torch.full((P,), 0x3F800000, dtype=torch.int64), torch.full((P,), 0xFFFFFFFF, dtype=torch.int64)]).cuda()
    _seg_sort_keys(keys, [P, P], 32, 8, key_bits=32)


def test_seg_sort_large_one_pass():
    """C2-like instance sort shape: a few long segments, 10-bit digits, one pass, payload only."""
    lens = [1_500_000, 0, 2_300_001, 999_999]
    _seg_sort_case(16, sum(lens), lens, 10, 10, iota=False)


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
