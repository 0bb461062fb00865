"""CPU: the C-ABI library loads and exports every symbol include/gsb200.h declares (no compute calls)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "gsb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsb_[a-z0-9_]+)\s*\(", src)) - {"gsb_alloc_fn"})


def test_library_exports_every_declared_symbol():
    from gsb200 import _lib
    lib = _lib.lib()
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(_lib.SYMBOLS), set(names) ^ set(_lib.SYMBOLS)
    assert lib.gsb_version().decode().startswith("gsb200")
    # pure host helpers are callable without a GPU
    assert [lib.gsb_higher_msb(n) for n in (256, 920, 8160)] == [9, 10, 13]
    assert lib.gsb_geom_bytes(1000, 2) > 2 * 1000 * 60
    assert lib.gsb_binning_bytes(10_000, 920, 1) >= 10_000 * 10
    assert lib.gsb_binning_bytes(10_000, 8160, 1) >= 10_000 * 18


def test_struct_layout_matches_header(tmp_path):
    """sizeof / offsetof of the two PODs as gcc sees the header == the ctypes mirrors."""
    import ctypes as C
    import subprocess
    from gsb200 import _lib
    fields = {"gsb_raster_args": [n for n, _ in _lib.RasterArgs._fields_],
              "gsb_raster_ctx": [n for n, _ in _lib.RasterCtx._fields_ if n != "_pad"]}
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "gsb200.h"', 'int main(void){']
    for st, fs in fields.items():
        src.append(f'printf("%zu\\n", sizeof({st}));')
        for f in fs:
            src.append(f'printf("%zu\\n", offsetof({st}, {f}));')
    src.append('return 0;}')
    c = tmp_path / "layout.c"
    c.write_text("\n".join(src))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = []
    for cls, st in ((_lib.RasterArgs, "gsb_raster_args"), (_lib.RasterCtx, "gsb_raster_ctx")):
        want.append(C.sizeof(cls))
        want += [getattr(cls, f).offset for f in fields[st]]
    assert got == want


def test_no_cpu_fallback():
    """The product path must fail loudly without CUDA tensors."""
    import torch
    from gsb200 import raster
    z = torch.zeros(4, 3)
    with pytest.raises(RuntimeError):
        raster.forward(z, z, torch.zeros(4, 1), z, torch.zeros(4, 4), None, torch.zeros(3), 1.0, torch.eye(4)[None],
                       torch.eye(4)[None], 1.0, 1.0, 16, 16)
