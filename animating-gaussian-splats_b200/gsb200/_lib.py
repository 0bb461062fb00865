"""ctypes binding of libgsb200.so (include/gsb200.h).

This is the whole "torch extension": tensors are turned into raw device pointers, the current CUDA
stream is passed through, and non-zero return codes raise.  There is NO fallback: if the library is
missing or no CUDA device is present, every compute entry point raises.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GSB200_LIB") or os.path.join(os.path.dirname(_HERE), "libgsb200.so")

GSB_OK, GSB_EINVAL, GSB_ECUDA, GSB_ENOMEM, GSB_EPREFILTER = 0, -1, -2, -3, -4
FLAT_WIDTH, FLAT_COL_GRADNORM, FLAT_COL_VISCOUNT, FLAT_COL_MAXRADIUS = 20, 14, 15, 16
_ERR = {GSB_EINVAL: "invalid argument", GSB_ECUDA: "CUDA error", GSB_ENOMEM: "buffer allocation failed",
        GSB_EPREFILTER: "Point is filtered although prefiltered is set. This shouldn't happen!"}

ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)


class RasterArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("n_cams", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
        ("prefiltered", C.c_int32), ("color_cam_stride", C.c_int32), ("scale_modifier", C.c_float),
        ("tanfovx_host", C.POINTER(C.c_float)), ("tanfovy_host", C.POINTER(C.c_float)),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("colors", C.c_void_p), ("opacities", C.c_void_p),
        ("scales", C.c_void_p), ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p),
        ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("entries", C.c_void_p),
        ("colors2", C.c_void_p), ("out_color2", C.c_void_p),
    ]


class RasterCtx(C.Structure):
    _fields_ = [
        ("num_rendered", C.c_int64),
        ("P", C.c_int32), ("n_cams", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
        ("tiles_x", C.c_int32), ("tiles_y", C.c_int32), ("sort_bits", C.c_int32), ("tile_bits", C.c_int32),
        ("depths", C.c_void_p), ("radii", C.c_void_p), ("means2D", C.c_void_p), ("cov3D", C.c_void_p),
        ("conic_opacity", C.c_void_p), ("tiles_touched", C.c_void_p),
        ("gauss_order", C.c_void_p), ("point_offsets", C.c_void_p), ("seg_start", C.c_void_p),
        ("final_T", C.c_void_p), ("n_contrib", C.c_void_p), ("ranges", C.c_void_p),
        ("point_list", C.c_void_p), ("point_list_unsorted", C.c_void_p), ("tile_keys_unsorted", C.c_void_p),
        ("tile_keys_sorted", C.c_void_p), ("sort_temp", C.c_void_p),
        ("capacity", C.c_int64), ("status", C.c_void_p),
    ]


class AdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_int64), ("step", C.c_int64), ("lr", C.c_double)]


class FlatAdamArgs(C.Structure):
    _fields_ = [("P", C.c_int32), ("param", C.c_void_p * 6), ("exp_avg", C.c_void_p * 6), ("exp_avg_sq", C.c_void_p * 6),
                ("lr", C.c_double * 6)]


# name -> (restype, argtypes): every symbol include/gsb200.h declares
SYMBOLS = {
    "gsb_version": (C.c_char_p, []),
    "gsb_last_error": (C.c_char_p, []),
    "gsb_geom_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gsb_image_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "gsb_binning_bytes": (C.c_size_t, [C.c_int64, C.c_int32, C.c_int32]),
    "gsb_higher_msb": (C.c_uint32, [C.c_uint32]),
    "gsb_raster_forward": (C.c_int, [C.POINTER(RasterArgs), ALLOC_FN, ALLOC_FN, ALLOC_FN, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.POINTER(RasterCtx), C.c_void_p]),
    "gsb_raster_backward": (C.c_int, [C.POINTER(RasterArgs), C.POINTER(RasterCtx)] + [C.c_void_p] * 10),
    "gsb_raster_forward_colors": (C.c_int, [C.POINTER(RasterArgs), C.POINTER(RasterCtx), C.c_void_p, C.c_void_p,
                                            C.c_void_p]),
    "gsb_raster_forward_async": (C.c_int, [C.POINTER(RasterArgs), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                           C.c_void_p, C.c_size_t, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.POINTER(RasterCtx), C.c_void_p]),
    "gsb_raster_count_pairs": (C.c_int, [C.POINTER(RasterArgs), C.POINTER(RasterCtx), C.c_void_p, C.c_void_p]),
    "gsb_backward_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gsb_raster_backward_flat": (C.c_int, [C.POINTER(RasterArgs), C.POINTER(RasterCtx), C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "gsb_raster_backward_flat2": (C.c_int, [C.POINTER(RasterArgs), C.POINTER(RasterCtx), C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "gsb_flat_apply_stats": (C.c_int, [C.c_int32, C.c_void_p, C.c_int32] + [C.c_void_p] * 4),
    "gsb_ssim_l1_rows": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                   C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsb_comm_bytes": (C.c_size_t, [C.c_size_t]),
    "gsb_comm_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]),
    "gsb_comm_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "gsb_comm_close": (C.c_int, [C.c_void_p]),
    "gsb_comm_free": (C.c_int, [C.c_void_p]),
    "gsb_allreduce_flat": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(C.c_void_p), C.c_size_t, C.c_int64, C.c_int32,
                                     C.c_int32, C.c_int32, C.c_void_p]),
    "gsb_mark_visible": (C.c_int, [C.c_int32] + [C.c_void_p] * 5),
    "gsb_ssim_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "gsb_ssim_l1": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                              C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsb_densify_accumulate": (C.c_int, [C.c_int32] + [C.c_void_p] * 7),
    "gsb_densify_accumulate_batch": (C.c_int, [C.c_int32, C.c_int32] + [C.c_void_p] * 7),
    "gsb_densify_plan_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "gsb_densify_plan": (C.c_int, [C.c_int32] + [C.c_void_p] * 4 + [C.c_float, C.c_float, C.c_float, C.c_int32, C.c_float,
                                   C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsb_densify_apply": (C.c_int, [C.c_int32, C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                    C.c_void_p, C.c_void_p]),
    "gsb_densify_accumulate_masked": (C.c_int, [C.c_int32] + [C.c_void_p] * 5),
    "gsb_split_rows": (C.c_int, [C.c_int32, C.c_int32] + [C.c_void_p] * 9),
    "gsb_scan_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "gsb_mask_scan": (C.c_int, [C.c_int32] + [C.c_void_p] * 5),
    "gsb_gather_rows": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "gsb_inclusive_scan_u32": (C.c_int, [C.c_int32] + [C.c_void_p] * 4),
    "gsb_inclusive_scan_u32_static": (C.c_int, [C.c_int32] + [C.c_void_p] * 3),
    "gsb_sort_temp_bytes": (C.c_size_t, [C.c_int64]),
    "gsb_sort_pairs": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                 C.c_void_p]),
    "gsb_seg_sort_temp_bytes": (C.c_size_t, [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "gsb_seg_sort_u16": (C.c_int, [C.c_int64, C.c_int32] + [C.c_void_p] * 5 + [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "gsb_seg_sort_u32": (C.c_int, [C.c_int64, C.c_int32] + [C.c_void_p] * 5 + [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "gsb_activate": (C.c_int, [C.c_int32] + [C.c_void_p] * 7),
    "gsb_activate_backward": (C.c_int, [C.c_int32] + [C.c_void_p] * 10),
    "gsb_adam_step": (C.c_int, [C.c_int32, C.POINTER(AdamGroup), C.c_double, C.c_double, C.c_double, C.c_void_p]),
    "gsb_adam_flat": (C.c_int, [C.POINTER(FlatAdamArgs), C.c_void_p, C.c_int32, C.c_void_p, C.c_double, C.c_double, C.c_double,
                                C.c_void_p]),
    "gsb_rigidity_loss": (C.c_int, [C.c_int32, C.c_int32] + [C.c_void_p] * 5 + [C.c_int64, C.c_int64, C.c_void_p,
                                    C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int64] +
                          [C.c_void_p] * 4),
    "gsb_foreground_info": (C.c_int, [C.c_int32, C.c_int32] + [C.c_void_p] * 4 + [C.c_int64, C.c_int64, C.c_void_p,
                                      C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "gsb_linear_tf32": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                  C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "gsb_linear_tf32x3_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gsb_linear_tf32x3": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                    C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsb_linear_residual": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    "gsb_linear_wgrad_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gsb_linear_wgrad_tf32": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                        C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "gsb_linear_wgrad_tf32x3": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                          C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "gsb_linear_thin_dx": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                     C.c_void_p, C.c_int64, C.c_void_p]),
    "gsb_linear_thin_wgrad_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gsb_linear_thin_wgrad": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsb_bn_gelu": (C.c_int, [C.c_int64, C.c_int64, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float,
                              C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_float,
                              C.c_int32, C.c_void_p]),
    "gsb_bn_gelu_backward": (C.c_int, [C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                       C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsb_positional_encoding": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "gsb_profile_enable": (C.c_int, [C.c_int]),
    "gsb_profile_read": (C.c_int, [C.POINTER(C.c_float), C.c_int32]),
}
PROFILE_STAGES = ["fwd_begin", "preprocess", "depth_sort", "scan", "d2h_sync", "duplicate_keys", "sort", "tile_ranges",
                  "render_fwd", "bwd_begin", "bwd_zero", "render_bwd", "preprocess_bwd"]

_lib = None


def lib():
    """Load libgsb200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU or PyTorch fallback)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc, what):
    if rc != GSB_OK:
        detail = lib().gsb_last_error().decode() if rc == GSB_ECUDA else ""
        raise RuntimeError(f"{what}: {_ERR.get(rc, rc)} {detail}".rstrip())


def ptr(t):
    """Device pointer of a tensor (None / empty -> NULL, the reference's nullptr convention,
    DGR/rasterize_points.cu:97-103)."""
    if t is None or t.numel() == 0:
        return None
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    """The current device's current CUDA stream as a void*.  Raw C accessors: torch.cuda.current_stream() builds
    a Stream object (~15 us) and this is called once per kernel launch."""
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))


class _NoGuard:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NO_GUARD = _NoGuard()


def device_guard(dev):
    """`with torch.cuda.device(dev)` only when dev is not already the current device (the common case)."""
    idx = dev.index if dev.index is not None else torch._C._cuda_getDevice()
    if idx == torch._C._cuda_getDevice():
        return _NO_GUARD
    return torch.cuda.device(idx)


def require_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: gsb200 has no CPU path")
