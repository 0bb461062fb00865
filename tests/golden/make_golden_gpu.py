"""Generates tests/golden/raster_*.npz by running the UNMODIFIED reference CUDA extension
(oracle/_ref/refdgr/_C.so, built by oracle/Makefile from /root/reference) on a B200.

    gpurun -- 'python tests/golden/make_golden_gpu.py gpurun_out/golden'   # then copy into tests/golden/

The reference ships no golden vectors of its own; these files are what pins the
C oracle (tests/test_oracle_golden.py) and, through it, the CUDA path.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "animating-gaussian-splats_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import helpers  # noqa: E402
import oracle_lib  # noqa: E402

CASES = [  # name, P, W, H, seed, bg, scale_override
    ("small", 3000, 96, 64, 21, (0.0, 0.0, 0.0), None),
    ("ragged_bg", 1500, 90, 50, 22, (0.3, 0.6, 0.1), None),
    ("big_splats", 300, 112, 80, 23, (0.0, 0.0, 0.0), 0.35),
]


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    ref = oracle_lib.load_reference_ext()
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    dev = "cuda"
    for name, P, W, H, seed, bgv, so in CASES:
        _, a, cams = helpers.scene_inputs(P, W, H, seed=seed, device=dev, scale_override=so)
        cam = cams[0]
        bg = torch.tensor(bgv, dtype=torch.float32, device=dev)
        fw = helpers.ref_forward(ref, a, cam, bg, dev)
        dL = torch.randn((3, H, W), generator=torch.Generator().manual_seed(seed)).to(dev)
        g1 = helpers.ref_backward(ref, a, cam, bg, dev, fw, dL)
        g2 = helpers.ref_backward(ref, a, cam, bg, dev, fw, dL)  # second run: atomic-order spread
        torch.cuda.synchronize()
        R = fw["num_rendered"]
        d = oracle_lib.decode_reference_buffers(P, H, W, R, fw["geom"], fw["binning"], fw["img"])
        out = {f"in_{k}": v.cpu().numpy() for k, v in a.items()}
        out.update(in_bg=np.asarray(bgv, np.float32), in_view=cam["viewmatrix"].numpy(), in_proj=cam["projmatrix"].numpy(),
                   in_tanfovx=np.float64(cam["tanfovx"]), in_tanfovy=np.float64(cam["tanfovy"]),
                   in_H=np.int64(H), in_W=np.int64(W), in_dL=dL.cpu().numpy(), num_rendered=np.int64(R),
                   color=fw["color"].cpu().numpy(), depth=fw["depth"].cpu().numpy(), radii=fw["radii"].cpu().numpy())
        for k, v in d.items():
            out[f"st_{k}"] = v.cpu().numpy()
        for k, v in g1.items():
            out[f"g_{k}"] = v.cpu().numpy()
            out[f"gspread_{k}"] = np.float64((v - g2[k]).abs().max().item())
        path = os.path.join(outdir, f"raster_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "R =", R, "visible =", int((fw["radii"] > 0).sum()), "->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
