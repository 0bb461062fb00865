"""Test-side access to the checkers under oracle/: the C restatement (libgs_oracle.so) and, when it
was built, the unmodified reference CUDA extension (oracle/_ref/refdgr/_C.so).  Only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline / reference legs import this module."""
import ctypes as C
import importlib.util
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "_build", "libgs_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "refdgr", "_C.so")

_f = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u8 = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u64 = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_d = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


def build_oracle():
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
            os.path.join(ROOT, "oracle", "gs_oracle.c")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], stdout=subprocess.DEVNULL)


class Oracle:
    def __init__(self, path):
        self.l = C.CDLL(path)
        l = self.l
        l.gso_preprocess.restype = C.c_int64
        l.gso_preprocess.argtypes = [C.c_int, _f, C.c_void_p, C.c_float, C.c_void_p, _f, C.c_void_p, _f, _f, C.c_int,
                                     C.c_int, C.c_float, C.c_float, _i, _f, _f, _f, _f, _u, _u]
        l.gso_mark_visible.argtypes = [C.c_int, _f, _f, _u8]
        l.gso_higher_msb.restype = C.c_uint32
        l.gso_higher_msb.argtypes = [C.c_uint32]
        l.gso_bin.argtypes = [C.c_int, C.c_int64, _i, _f, _f, _u, C.c_int, C.c_int, _u64, _u, _u64, _u, _u]
        l.gso_render_fwd.argtypes = [C.c_int, C.c_int, _u, _u, _f, _f, _f, _f, _f, _f, _f, _f, _u, _u8, _u64]
        l.gso_render_bwd.argtypes = [C.c_int, C.c_int, _u, _u, _f, _f, _f, _f, _f, _u, _f, C.c_int, _f, _f, _f, _f]
        l.gso_preprocess_bwd.argtypes = [C.c_int, _f, _i, C.c_void_p, C.c_float, C.c_void_p, _f, _f, _f, C.c_int,
                                         C.c_int, C.c_float, C.c_float, _f, _f, _f, _f, C.c_void_p, C.c_void_p]
        l.gso_ssim_l1.argtypes = [C.c_int, C.c_int, C.c_int, _f, _f, C.c_float, C.c_float, _d, C.c_void_p]
        l.gso_densify_accumulate.argtypes = [C.c_int, _i, _f, _f, _f, _f, _u8]
        l.gso_set_threads.restype = C.c_int
        l.gso_set_threads.argtypes = [C.c_int]

    def set_threads(self, n=0):
        """OpenMP team size of the oracle's loops (n <= 0: query).  Returns the size in effect."""
        return int(self.l.gso_set_threads(int(n)))

    # ---- forward ------------------------------------------------------------------------------
    def forward(self, means3D, colors, opacities, scales, rotations, bg, view, proj, tanfovx, tanfovy, H, W,
                scale_modifier=1.0, cov3D_precomp=None):
        f32 = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.float32))
        means3D, colors, opacities, bg = f32(means3D), f32(colors), f32(opacities).reshape(-1), f32(bg)
        view, proj = f32(view).reshape(16), f32(proj).reshape(16)
        P = means3D.shape[0]
        scales = f32(scales) if scales is not None else None
        rotations = f32(rotations) if rotations is not None else None
        cov_pre = f32(cov3D_precomp) if cov3D_precomp is not None else None
        vp = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        o = {
            "radii": np.zeros(P, np.int32), "means2D": np.zeros((P, 2), np.float32), "depths": np.zeros(P, np.float32),
            "cov3D": np.zeros((P, 6), np.float32), "conic_opacity": np.zeros((P, 4), np.float32),
            "tiles_touched": np.zeros(P, np.uint32), "point_offsets": np.zeros(P, np.uint32),
        }
        R = self.l.gso_preprocess(P, means3D, vp(scales), scale_modifier, vp(rotations), opacities, vp(cov_pre), view,
                                  proj, W, H, tanfovx, tanfovy, o["radii"], o["means2D"], o["depths"], o["cov3D"],
                                  o["conic_opacity"], o["tiles_touched"], o["point_offsets"])
        if cov_pre is not None:
            o["cov3D"] = cov_pre
        o["num_rendered"] = int(R)
        gx, gy = (W + 15) // 16, (H + 15) // 16
        n = max(R, 1)
        o["keys_unsorted"] = np.zeros(n, np.uint64); o["vals_unsorted"] = np.zeros(n, np.uint32)
        o["keys"] = np.zeros(n, np.uint64); o["point_list"] = np.zeros(n, np.uint32)
        o["ranges"] = np.zeros((gx * gy, 2), np.uint32)
        self.l.gso_bin(P, R, o["radii"], o["means2D"], o["depths"], o["point_offsets"], W, H, o["keys_unsorted"],
                       o["vals_unsorted"], o["keys"], o["point_list"], o["ranges"])
        for k in ("keys_unsorted", "vals_unsorted", "keys", "point_list"):
            o[k] = o[k][:R]
        o["color"] = np.zeros((3, H, W), np.float32); o["depth"] = np.zeros((1, H, W), np.float32)
        o["final_T"] = np.zeros((H, W), np.float32); o["n_contrib"] = np.zeros((H, W), np.uint32)
        o["fragile"] = np.zeros((H, W), np.uint8); o["blend_count"] = np.zeros(2, np.uint64)
        self.l.gso_render_fwd(W, H, o["ranges"], o["point_list"] if R else np.zeros(1, np.uint32), o["means2D"], colors,
                              o["conic_opacity"], o["depths"], bg, o["color"], o["depth"], o["final_T"], o["n_contrib"],
                              o["fragile"], o["blend_count"])
        o["_in"] = dict(means3D=means3D, colors=colors, opacities=opacities, scales=scales, rotations=rotations, bg=bg,
                        view=view, proj=proj, tanfovx=tanfovx, tanfovy=tanfovy, H=H, W=W, mod=scale_modifier,
                        cov_pre=cov_pre)
        return o

    def backward(self, o, dL_dpix):
        i = o["_in"]
        P, H, W = i["means3D"].shape[0], i["H"], i["W"]
        dL_dpix = np.ascontiguousarray(dL_dpix, dtype=np.float32)
        g = {"means2D": np.zeros((P, 3), np.float32), "conic": np.zeros((P, 4), np.float32),
             "opacities": np.zeros((P, 1), np.float32), "colors": np.zeros((P, 3), np.float32),
             "means3D": np.zeros((P, 3), np.float32), "cov3D": np.zeros((P, 6), np.float32),
             "scales": np.zeros((P, 3), np.float32), "rotations": np.zeros((P, 4), np.float32)}
        pl = o["point_list"] if o["num_rendered"] else np.zeros(1, np.uint32)
        self.l.gso_render_bwd(W, H, o["ranges"], pl, o["means2D"], i["colors"], o["conic_opacity"], i["bg"],
                              o["final_T"], o["n_contrib"], dL_dpix, P, g["means2D"], g["conic"], g["opacities"],
                              g["colors"])
        vp = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        has_sr = i["scales"] is not None
        self.l.gso_preprocess_bwd(P, i["means3D"], o["radii"], vp(i["scales"]), i["mod"], vp(i["rotations"]),
                                  np.ascontiguousarray(o["cov3D"]), i["view"], i["proj"], W, H, i["tanfovx"],
                                  i["tanfovy"], g["means2D"], g["conic"], g["means3D"], g["cov3D"],
                                  vp(g["scales"]) if has_sr else None, vp(g["rotations"]) if has_sr else None)
        return g

    def mark_visible(self, means3D, view):
        m = np.ascontiguousarray(means3D, dtype=np.float32)
        out = np.zeros(m.shape[0], np.uint8)
        self.l.gso_mark_visible(m.shape[0], m, np.ascontiguousarray(view, dtype=np.float32).reshape(16), out)
        return out.astype(bool)

    def ssim_l1(self, img1, img2, w_l1=0.8, w_ssim=0.2, grad=True):
        a = np.ascontiguousarray(img1, dtype=np.float32); b = np.ascontiguousarray(img2, dtype=np.float32)
        Cc, H, W = a.shape
        out = np.zeros(3, np.float64)
        g = np.zeros_like(a) if grad else None
        self.l.gso_ssim_l1(Cc, H, W, a, b, w_l1, w_ssim, out, g.ctypes.data_as(C.c_void_p) if grad else None)
        return out, g

    def densify_accumulate(self, radii, grad2d, max_r, accum, count):
        P = radii.shape[0]
        vis = np.zeros(P, np.uint8)
        self.l.gso_densify_accumulate(P, np.ascontiguousarray(radii, np.int32), np.ascontiguousarray(grad2d, np.float32),
                                      max_r, accum, count, vis)
        return vis.astype(bool)

    def higher_msb(self, n):
        return int(self.l.gso_higher_msb(n))


_oracle = None


def load():
    global _oracle
    if _oracle is None:
        build_oracle()
        _oracle = Oracle(ORACLE_SO)
    return _oracle


def load_reference_ext():
    """Import oracle/_ref/refdgr/_C.so (pybind module `_C` of the unmodified reference) or None."""
    if not os.path.exists(REF_SO):
        return None
    import torch  # noqa: F401  (the extension links against libtorch)
    spec = importlib.util.spec_from_file_location("_C", REF_SO)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def decode_reference_buffers(P, H, W, R, geom, binning, img):
    """Re-run the reference's fromChunk layout (cr/rasterizer_impl.cu:155-194; 128-B aligned bump
    allocation) over the byte tensors it returned.  Returns torch views (on the buffers' device)."""
    import torch

    def take(buf, off, dtype, count):
        off = (off + 127) // 128 * 128
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        return buf[off:off + nbytes].view(dtype), off + nbytes

    out = {}
    o = 0
    out["depths"], o = take(geom, o, torch.float32, P)
    _, o = take(geom, o, torch.uint8, 3 * P)
    out["radii"], o = take(geom, o, torch.int32, P)
    out["means2D"], o = take(geom, o, torch.float32, 2 * P)
    out["cov3D"], o = take(geom, o, torch.float32, 6 * P)
    out["conic_opacity"], o = take(geom, o, torch.float32, 4 * P)
    _, o = take(geom, o, torch.float32, 3 * P)
    out["tiles_touched"], o = take(geom, o, torch.int32, P)
    # scan temp (size known only to CUB) and point_offsets follow; offsets are re-derived by cumsum
    out["means2D"] = out["means2D"].view(P, 2); out["cov3D"] = out["cov3D"].view(P, 6)
    out["conic_opacity"] = out["conic_opacity"].view(P, 4)
    N = H * W
    o = 0
    out["final_T"], o = take(img, o, torch.float32, N)
    out["n_contrib"], o = take(img, o, torch.int32, N)
    rg, o = take(img, o, torch.int32, 2 * N)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["ranges"] = rg.view(N, 2)[:tiles]
    out["final_T"] = out["final_T"].view(H, W); out["n_contrib"] = out["n_contrib"].view(H, W)
    if R > 0:
        o = 0
        out["point_list"], o = take(binning, o, torch.int32, R)
        out["point_list_unsorted"], o = take(binning, o, torch.int32, R)
        out["point_list_keys"], o = take(binning, o, torch.int64, R)
        out["point_list_keys_unsorted"], o = take(binning, o, torch.int64, R)
    return out
