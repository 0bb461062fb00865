"""gsb_inclusive_scan_u32_static: the allocation-free variant of the scan (scratch in a file-scope device array of the
library) against the existing entry point with a caller scratch buffer -- bit-exact (integer work), on two calls in a
row (the static array keeps the previous call's contents), at sizes up to the stated maximum, and the rejection beyond
it.  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ITEMS = 1 << 24


def _lib():
    from gsb200 import _lib as L
    return L, L.lib()


def _scan_ref(L, lib, x):
    out = torch.empty_like(x)
    scratch = torch.empty(lib.gsb_scan_scratch_bytes(x.numel()), dtype=torch.uint8, device="cuda")
    L.check(lib.gsb_inclusive_scan_u32(x.numel(), L.ptr(x), L.ptr(out), L.ptr(scratch), L.stream_ptr()), "scan")
    return out


def test_static_scan_matches_existing_entry_point_twice_in_a_row():
    L, lib = _lib()
    g = torch.Generator(device="cuda").manual_seed(7)
    # two calls in a row at different sizes: the second, shorter call finds the first one's status words in the array
    sizes = [3_000_000, 100_003, 2049, 1, MAX_ITEMS]
    xs = [torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda", generator=g) for n in sizes]
    refs = [_scan_ref(L, lib, x) for x in xs]
    outs = [torch.full_like(x, -1) for x in xs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    for x, out in zip(xs, outs):           # back to back on one stream, nothing in between
        L.check(lib.gsb_inclusive_scan_u32_static(x.numel(), L.ptr(x), L.ptr(out), L.stream_ptr()), "static scan")
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before     # torch's allocator was not asked either
    for x, out, ref in zip(xs, outs, refs):
        assert torch.equal(out, ref)
        assert torch.equal(out, torch.cumsum(x, 0).int())
    # the same input again, after everything above went through the array
    again = torch.full_like(xs[0], -1)
    L.check(lib.gsb_inclusive_scan_u32_static(sizes[0], L.ptr(xs[0]), L.ptr(again), L.stream_ptr()), "static scan")
    assert torch.equal(again, refs[0])


def test_static_scan_rejects_sizes_beyond_the_maximum():
    L, lib = _lib()
    x = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = lib.gsb_inclusive_scan_u32_static(MAX_ITEMS + 1, L.ptr(x), L.ptr(x), L.stream_ptr())
    assert rc == L.GSB_EINVAL
    assert b"GSB_SCAN_STATIC_MAX_ITEMS" in lib.gsb_last_error()
    assert lib.gsb_inclusive_scan_u32_static(0, None, None, L.stream_ptr()) == 0
    torch.cuda.synchronize()
