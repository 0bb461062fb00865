"""Tensor-level wrappers of gsb_raster_forward / gsb_raster_backward / gsb_mark_visible.

Mirrors the torch glue of the reference (RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA /
markVisible, DGR/rasterize_points.cu:35-215): allocate outputs, hand raw pointers to the C ABI, keep the
three state buffers alive for backward.  `n_cams > 1` is the additional batched entry point: all
cameras of a timestep in one call (one preprocess launch, ONE sort over (camera, tile, depth) keys,
one blend launch), gradients summed over the cameras.
"""
import ctypes as C

import torch

from . import _lib as L


class RasterState:
    """What the reference returns to Python as (num_rendered, geomBuffer, binningBuffer, imgBuffer)."""

    __slots__ = ("ctx", "args", "geom", "binning", "img", "keep", "tan_x", "tan_y", "_status", "color2")

    @property
    def num_rendered(self):
        """R.  After an asynchronous forward it is known on the device only: reading it here synchronises."""
        if int(self.ctx.num_rendered) < 0:
            err, r, _ = self.status()
            if err == 0:
                self.ctx.num_rendered = r
            return r
        return int(self.ctx.num_rendered)

    def status_tensor(self):
        """The call's device status words (int32 [4]: error code, num_rendered, instances needed, -) or None (P == 0)."""
        if not self.ctx.status:
            return None
        return self._view(self.ctx.status, self.geom, torch.int32, (4,))

    def status(self):
        """(error code, num_rendered, instances needed on overflow) -- blocks until the forward has run."""
        t = self.status_tensor()
        if t is None:
            return 0, 0, 0
        e, r, need, _ = (int(x) for x in t.cpu())
        return e, r, need

    def check(self):
        """Raise if the (asynchronous) forward recorded an error: ArenaOverflow when the instances did not fit."""
        e, r, need = self.status()
        if e == L.GSB_ENOMEM:
            raise ArenaOverflow(need, int(self.ctx.capacity))
        L.check(e, "gsb_raster_forward_async")

    # ---- debug accessors for the bit-exact tests: views into the state buffers ------------------
    def _view(self, addr, base_t, dtype, shape):
        off = addr - base_t.data_ptr()
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        return base_t[off:off + nbytes].view(dtype).view(*shape)

    def tensors(self):
        c = self.ctx
        n, P = c.n_cams, c.P
        R = self.num_rendered
        tiles = c.tiles_x * c.tiles_y * n
        hw = c.width * c.height
        out = {
            "depths": self._view(c.depths, self.geom, torch.float32, (n, P)),
            "radii": self._view(c.radii, self.geom, torch.int32, (n, P)),
            "means2D": self._view(c.means2D, self.geom, torch.float32, (n, P, 2)),
            "cov3D": self._view(c.cov3D, self.geom, torch.float32, (1, P, 6)).expand(n, P, 6),
            "conic_opacity": self._view(c.conic_opacity, self.geom, torch.float32, (n, P, 4)),
            "tiles_touched": self._view(c.tiles_touched, self.geom, torch.int32, (n, P)),
            "point_offsets": self._view(c.point_offsets, self.geom, torch.int32, (n, P)),
            "gauss_order": self._view(c.gauss_order, self.geom, torch.int32, (n, P)),
            "seg_start": self._view(c.seg_start, self.geom, torch.int32, (n + 1,)),
            "final_T": self._view(c.final_T, self.img, torch.float32, (n, c.height, c.width)),
            "n_contrib": self._view(c.n_contrib, self.img, torch.int32, (n, c.height, c.width)),
            "ranges": self._view(c.ranges, self.img, torch.int32, (tiles, 2)),
        }
        if R > 0:
            out["point_list"] = self._view(c.point_list, self.binning, torch.int32, (R,))
            out["point_list_unsorted"] = self._view(c.point_list_unsorted, self.binning, torch.int32, (R,))
            out["tile_keys_unsorted"] = self._view(c.tile_keys_unsorted, self.binning, torch.int16, (R,))
            # Sorted tile ids.  With <= 1024 tiles per camera they are never materialised (the sort is one pass
            # and the ranges come from its histogram), so they are reassembled from the ranges: instance i
            # belongs to the tile whose [start, end) holds i.
            rg = out["ranges"].long()
            out["point_list_tiles"] = torch.repeat_interleave(torch.arange(tiles, device=rg.device), rg[:, 1] - rg[:, 0])
            if c.tile_keys_sorted:
                tiles_per_cam = c.tiles_x * c.tiles_y
                local = self._view(c.tile_keys_sorted, self.binning, torch.int16, (R,)).long() & 0xFFFF
                seg = out["seg_start"].long()
                cam_of = torch.repeat_interleave(torch.arange(n, device=rg.device), seg[1:] - seg[:-1])
                out["point_list_tiles_stored"] = cam_of * tiles_per_cam + local
            # the reference's sorted 64-bit keys, reassembled: (tile << 32) | bits(depth of the Gaussian)
            pl = out["point_list"].long()
            tiles_per_cam = c.tiles_x * c.tiles_y
            cam = out["point_list_tiles"] // tiles_per_cam
            dbits = out["depths"].view(torch.int32).view(-1)[cam * P + pl].long() & 0xFFFFFFFF
            out["point_list_keys"] = (out["point_list_tiles"] << 32) | dbits
        return out


class ArenaOverflow(RuntimeError):
    """An asynchronous forward produced more (Gaussian, tile) instances than its arena holds."""

    def __init__(self, needed, capacity):
        super().__init__(f"rasterizer arena overflow: {needed} instances, capacity {capacity}")
        self.needed, self.capacity = needed, capacity


class RasterArena:
    """Caller-owned state buffers of the asynchronous forward (the reference allocates and grows three torch byte
    tensors per call, DGR/rasterize_points.cu:27-33,73-78): allocated once, grown on demand, reused by every step.
    `capacity` bounds the (Gaussian, tile) instances of a call; size it from an earlier render of the same scene
    (reserve(..., capacity=int(1.25 * R))).  One arena serves one forward state at a time: the state of the previous
    call is overwritten by the next."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.geom = self.img = self.binning = self.bwd_scratch = None
        self.capacity = 0
        self.outputs = {}

    def _grow(self, name, nbytes):
        t = getattr(self, name)
        if t is None or t.numel() < nbytes:
            t = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
            setattr(self, name, t)
        return t

    def reserve(self, P, n_cams, W, H, capacity):
        lib = L.lib()
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        self._grow("geom", lib.gsb_geom_bytes(P, n_cams))
        self._grow("img", lib.gsb_image_bytes(W, H, n_cams))
        self.capacity = max(int(capacity), self.capacity if self.binning is not None else 0)
        self._grow("binning", lib.gsb_binning_bytes(self.capacity, tiles, n_cams))
        self._grow("bwd_scratch", lib.gsb_backward_scratch_bytes(P, n_cams))
        return self

    def output(self, name, shape, dtype):
        """Output tensors that live as long as the arena (stable addresses for CUDA-graph replays)."""
        t = self.outputs.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            self.outputs[name] = t
        return t


def tanfov_table(tanfovx, tanfovy, n_cams, device):
    """DEVICE float [n_cams][2] for the asynchronous forward (build once per camera set, not per step)."""
    return torch.tensor(list(zip(_as_list(tanfovx, n_cams), _as_list(tanfovy, n_cams))), dtype=torch.float32,
                        device=device).contiguous()


# ------------------------------------------------------------------------------------------------
# Geometry reuse.  Everything a forward call computes before the blend depends only on (means3D, opacities,
# scales, rotations / cov3D, cameras, image size); densify.py:116-150 renders the SAME Gaussians through the SAME
# camera twice per iteration (colours, then segmentation colours), so the second call can take the first call's
# state and run the blend kernel alone: no preprocess, no sorts, no host synchronisation.  A hit requires every
# geometry input to be provably unchanged: the same tensor object at the same torch version counter, or an output
# of gsb200.optim.activate whose raw inputs are the same objects at the same versions (`_gsb_src`).  The cache
# keeps the previous call's tensors alive, so a recycled address can never alias.  GSB_GEOM_CACHE=0 disables it.
# ------------------------------------------------------------------------------------------------
import os as _os

_GEOM_CACHE_ON = _os.environ.get("GSB_GEOM_CACHE", "1") != "0"
_geom_cache = None   # (ident list, scalar key, RasterState, radii)
geom_cache_hits = 0


def _ident(t):
    """(object, version) pairs that pin down the contents of a geometry input."""
    if t is None or t.numel() == 0:   # the reference's "not provided" convention: None or an empty tensor
        return None
    src = getattr(t, "_gsb_src", None)
    if src is not None:
        return src
    return ((t, t._version),)


def _same_ident(a, b):
    if a is None or b is None:
        return a is None and b is None
    return len(a) == len(b) and all(x[0] is y[0] and x[1] == y[1] for x, y in zip(a, b))


def invalidate_geometry_cache():
    global _geom_cache
    _geom_cache = None


def _f32c(t, name):
    if t is None or t.numel() == 0:
        return None
    L.require_cuda(t, name)
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32")
    return t.contiguous()


def _as_list(x, n):
    if isinstance(x, (list, tuple)):
        assert len(x) == n
        return [float(v) for v in x]
    return [float(x)] * n


def forward(means3D, colors, opacities, scales, rotations, cov3D_precomp, bg, scale_modifier, viewmatrix,
            projmatrix, tanfovx, tanfovy, image_height, image_width, prefiltered=False, n_cams=1,
            colors_per_camera=False, reuse_geometry=False, arena=None, tanfov_dev=None, entries=None, colors2=None):
    """-> (color [n,3,H,W] or [3,H,W], depth, radii, RasterState).  Single camera squeezes the batch dim.
    reuse_geometry: allow this call to take (and to leave behind) pre-blend state, see "Geometry reuse" above; the
    drop-in autograd node passes True, the batched step does not (every step there has new parameters).
    arena (RasterArena, reserved by the caller): the asynchronous form -- no host synchronisation, no allocation,
    CUDA-graph capturable; outputs live in the arena; tanfov_dev = tanfov_table(...) (built here if None, which
    costs a host -> device copy and is therefore not capturable).
    entries (int32 CUDA tensor [n_cams, 4] = tile_row0, tile_row1, split_slot, count_visibility): tile-band sharding --
    every batch entry renders only those tile rows of its camera (include/gsb200.h: gsb_raster_args.entries).
    colors2 ([P, 3]): a second colour table blended in the same walk; its image is left in state.color2 ([n, 3, H, W])
    and backward_flat(..., dL_dpix2=) back-propagates both images at once."""
    global _geom_cache, geom_cache_hits
    if means3D.ndim != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # DGR/rasterize_points.cu:56-58
    L.require_cuda(means3D, "means3D")
    dev = means3D.device
    P = means3D.size(0)
    H, W = int(image_height), int(image_width)
    idents = [_ident(t) for t in (means3D, opacities, scales, rotations, cov3D_precomp, viewmatrix, projmatrix)]
    scalars = (P, n_cams, H, W, float(scale_modifier), bool(prefiltered), tuple(_as_list(tanfovx, n_cams)),
               tuple(_as_list(tanfovy, n_cams)), dev,
               torch._C._cuda_getCurrentRawStream(dev.index if dev.index is not None else torch._C._cuda_getDevice()))
    cached = _geom_cache
    use_cache = _GEOM_CACHE_ON and reuse_geometry
    hit = (use_cache and cached is not None and P > 0 and cached[1] == scalars and
           all(_same_ident(a, b) for a, b in zip(cached[0], idents)))
    st = RasterState()
    st.keep = [_f32c(means3D, "means3D"), _f32c(colors, "colors_precomp"), _f32c(opacities, "opacities"),
               _f32c(scales, "scales"), _f32c(rotations, "rotations"), _f32c(cov3D_precomp, "cov3D_precomp"),
               _f32c(bg, "bg"), _f32c(viewmatrix, "viewmatrix"), _f32c(projmatrix, "projmatrix")]
    m3, col, opa, sca, rot, cov, bgc, vm, pm = st.keep
    st.tan_x = (C.c_float * n_cams)(*_as_list(tanfovx, n_cams))
    st.tan_y = (C.c_float * n_cams)(*_as_list(tanfovy, n_cams))
    a = L.RasterArgs()
    a.P, a.n_cams, a.width, a.height = P, n_cams, W, H
    a.prefiltered = 1 if prefiltered else 0
    a.color_cam_stride = 3 * P if colors_per_camera else 0
    a.scale_modifier = float(scale_modifier)
    a.tanfovx_host = C.cast(st.tan_x, C.POINTER(C.c_float))
    a.tanfovy_host = C.cast(st.tan_y, C.POINTER(C.c_float))
    a.background, a.means3D, a.colors, a.opacities = L.ptr(bgc), L.ptr(m3), L.ptr(col), L.ptr(opa)
    a.scales, a.rotations, a.cov3D_precomp = L.ptr(sca), L.ptr(rot), L.ptr(cov)
    a.viewmatrix, a.projmatrix = L.ptr(vm), L.ptr(pm)
    st.color2 = None
    if colors2 is not None and P > 0:
        c2 = _f32c(colors2, "colors2")
        if colors_per_camera or tuple(c2.shape) != (P, 3):
            raise RuntimeError("colors2 must be one [P, 3] table shared by the cameras")
        st.color2 = (arena.output("color2", (n_cams, 3, H, W), torch.float32) if arena is not None
                     else torch.empty((n_cams, 3, H, W), dtype=torch.float32, device=dev))
        st.keep += [c2, st.color2]
        a.colors2, a.out_color2 = L.ptr(c2), L.ptr(st.color2)
    if entries is not None:
        if entries.dtype != torch.int32 or tuple(entries.shape) != (n_cams, 4) or not entries.is_cuda or not entries.is_contiguous():
            raise RuntimeError("entries must be a contiguous int32 CUDA tensor [n_cams, 4]")
        st.keep.append(entries)
        a.entries = L.ptr(entries)
    st.args = a
    st.ctx = L.RasterCtx()

    st.geom = st.binning = st.img = None
    if arena is not None and P > 0:
        if tanfov_dev is None:
            tanfov_dev = tanfov_table(tanfovx, tanfovy, n_cams, dev)
        st.keep.append(tanfov_dev)
        color = arena.output("color", (n_cams, 3, H, W), torch.float32)
        depth = arena.output("depth", (n_cams, 1, H, W), torch.float32)
        radii = arena.output("radii", (n_cams, P), torch.int32)
        st.geom, st.binning, st.img = arena.geom, arena.binning, arena.img
        with L.device_guard(dev):
            rc = L.lib().gsb_raster_forward_async(
                C.byref(a), L.ptr(tanfov_dev), L.ptr(arena.geom), arena.geom.numel(), L.ptr(arena.img),
                arena.img.numel(), L.ptr(arena.binning), arena.binning.numel(), int(arena.capacity), L.ptr(color),
                L.ptr(depth), L.ptr(radii), C.byref(st.ctx), L.stream_ptr())
        L.check(rc, "gsb_raster_forward_async")
        if n_cams == 1:
            return color[0], depth[0], radii[0], st
        return color, depth, radii, st
    # outputs (DGR/rasterize_points.cu:67-69); every pixel / slot is written by the kernels
    color = torch.empty((n_cams, 3, H, W), dtype=torch.float32, device=dev)
    depth = torch.empty((n_cams, 1, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((n_cams, P), dtype=torch.int32, device=dev)
    if hit:
        prev = cached[2]
        C.memmove(C.byref(st.ctx), C.byref(prev.ctx), C.sizeof(L.RasterCtx))
        st.geom, st.binning, st.img = prev.geom, prev.binning, prev.img
        radii = cached[3]
        with L.device_guard(dev):
            rc = L.lib().gsb_raster_forward_colors(C.byref(a), C.byref(st.ctx), L.ptr(color), L.ptr(depth), L.stream_ptr())
        L.check(rc, "gsb_raster_forward_colors")
        geom_cache_hits += 1
    elif P == 0:
        color.zero_(); depth.zero_()
        st.ctx.P, st.ctx.n_cams, st.ctx.width, st.ctx.height = 0, n_cams, W, H
    else:
        bufs = {}

        def make_cb(slot):
            def cb(_user, nbytes):
                t = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=dev)
                bufs[slot] = t
                return t.data_ptr()
            return L.ALLOC_FN(cb)

        cbs = [make_cb("geom"), make_cb("binning"), make_cb("img")]
        with L.device_guard(dev):
            rc = L.lib().gsb_raster_forward(C.byref(a), cbs[0], cbs[1], cbs[2], None, L.ptr(color), L.ptr(depth),
                                            L.ptr(radii), C.byref(st.ctx), L.stream_ptr())
        L.check(rc, "gsb_raster_forward")
        st.geom, st.binning, st.img = bufs.get("geom"), bufs.get("binning"), bufs.get("img")
        if use_cache:
            _geom_cache = (idents, scalars, st, radii)
    if n_cams == 1:
        return color[0], depth[0], radii[0], st
    return color, depth, radii, st


def backward(st, dL_dpix, need_cov3D_grad=True):
    """-> dict of gradients (summed over the cameras of the batch where the parameter is shared)."""
    a, c = st.args, st.ctx
    P, n = a.P, a.n_cams
    dev = dL_dpix.device
    f = dict(dtype=torch.float32, device=dev)
    g = {
        "means2D": torch.empty((n, P, 3), **f), "conic": torch.empty((n, P, 4), **f),
        "opacities": torch.empty((P, 1), **f),
        "colors": torch.empty((n, P, 3) if a.color_cam_stride else (P, 3), **f),
        "means3D": torch.empty((P, 3), **f),
        "cov3D": torch.empty((P, 6), **f) if need_cov3D_grad else None,
        "scales": torch.empty((P, 3), **f), "rotations": torch.empty((P, 4), **f),
    }
    if P > 0:
        dpix = dL_dpix.contiguous()
        if dpix.dtype != torch.float32:
            dpix = dpix.float()
        with L.device_guard(dev):
            rc = L.lib().gsb_raster_backward(C.byref(a), C.byref(c), L.ptr(dpix), L.ptr(g["means2D"]), L.ptr(g["conic"]),
                                             L.ptr(g["opacities"]), L.ptr(g["colors"]), L.ptr(g["means3D"]),
                                             L.ptr(g["cov3D"]), L.ptr(g["scales"]), L.ptr(g["rotations"]),
                                             L.stream_ptr())
        L.check(rc, "gsb_raster_backward")
    if n == 1:
        g["means2D"] = g["means2D"][0]
        g["conic"] = g["conic"][0]
    return g


def reblend(st, colors, arena=None, name="color2"):
    """Blend an existing forward state again with another colour table (gsb_raster_forward_colors): the second render
    of an iteration (densify.py:136-150) needs no preprocess, no sorts and no instance lists of its own.
    -> (color [n,3,H,W] or [3,H,W], depth, RasterState sharing st's buffers, for backward / backward_flat)."""
    P, n, H, W = st.args.P, st.args.n_cams, st.args.height, st.args.width
    col = _f32c(colors, "colors_precomp")
    st2 = RasterState()
    st2.ctx = st.ctx
    st2.geom, st2.binning, st2.img = st.geom, st.binning, st.img
    st2.tan_x, st2.tan_y = st.tan_x, st.tan_y
    st2.color2 = None
    st2.keep = list(st.keep) + [col]
    a = L.RasterArgs()
    C.memmove(C.byref(a), C.byref(st.args), C.sizeof(L.RasterArgs))
    a.colors = L.ptr(col)
    a.color_cam_stride = 0
    st2.args = a
    dev = col.device
    if arena is not None:
        color = arena.output(name, (n, 3, H, W), torch.float32)
        depth = arena.output(name + "_depth", (n, 1, H, W), torch.float32)
    else:
        color = torch.empty((n, 3, H, W), dtype=torch.float32, device=dev)
        depth = torch.empty((n, 1, H, W), dtype=torch.float32, device=dev)
    if P > 0:
        with L.device_guard(dev):
            rc = L.lib().gsb_raster_forward_colors(C.byref(a), C.byref(st2.ctx), L.ptr(color), L.ptr(depth), L.stream_ptr())
        L.check(rc, "gsb_raster_forward_colors")
    else:
        color.zero_(); depth.zero_()
    if n == 1:
        return color[0], depth[0], st2
    return color, depth, st2


_FLAT_PART = {"all": 0, "blend": 0x100, "finish": 0x200}


def backward_flat(st, dL_dpix, flat, mode=0, scratch=None, means2D_out=None, part="all", dL_dpix2=None):
    """The step's backward in its exchange form (gsb_raster_backward_flat): the per-Gaussian record -- gradients,
    statistic increments, largest radius -- goes into `flat` [P, 20] (layout: gsb200.dist.COLS).  mode 0 writes every
    column; mode 1 (a second render of the same Gaussians with another colour table) adds the geometry gradients and
    writes the colour gradient into cols 17-19; mode 2 (another camera group of the same step) adds to every column
    (col 16: max).  part: "blend" runs only the per-tile backward into `scratch`, "finish" only the per-Gaussian pass
    from `scratch` into flat (dL_dpix may be None) -- for callers that overlap several groups on streams.
    scratch: uint8 tensor of gsb_backward_scratch_bytes (allocated here if
    None).  means2D_out: optional [n, P, 3] for the per-camera screen-space gradients."""
    a, c = st.args, st.ctx
    P, n = a.P, a.n_cams
    if P == 0:
        return flat
    if (flat.dtype != torch.float32 or flat.ndim != 2 or flat.shape[0] != P or flat.shape[1] < L.FLAT_WIDTH or flat.shape[1] % 4
            or not flat.is_contiguous()):
        raise RuntimeError(f"flat must be a contiguous float32 [P, {L.FLAT_WIDTH} + 4 k] tensor")
    L.require_cuda(flat, "flat")
    lib = L.lib()
    if scratch is None:
        scratch = torch.empty(lib.gsb_backward_scratch_bytes(P, n), dtype=torch.uint8, device=flat.device)
    dpix = None
    if dL_dpix is not None:
        dpix = dL_dpix.contiguous()
        if dpix.dtype != torch.float32:
            dpix = dpix.float()
    if dL_dpix2 is not None:   # both images of a two-table forward (colors2) in one walk
        if st.color2 is None or mode != 0 or part != "all":
            raise RuntimeError("dL_dpix2 needs a forward with colors2, mode 0 and part 'all'")
        d2 = dL_dpix2.contiguous().float()
        with L.device_guard(flat.device):
            rc = lib.gsb_raster_backward_flat2(C.byref(a), C.byref(c), L.ptr(dpix), L.ptr(d2), L.ptr(scratch), L.ptr(flat),
                                               int(flat.shape[1]), L.ptr(means2D_out), L.stream_ptr())
        L.check(rc, "gsb_raster_backward_flat2")
        return flat
    with L.device_guard(flat.device):
        rc = lib.gsb_raster_backward_flat(C.byref(a), C.byref(c), L.ptr(dpix), L.ptr(scratch), L.ptr(flat), int(flat.shape[1]),
                                          int(mode) | _FLAT_PART[part], L.ptr(means2D_out), L.stream_ptr())
    L.check(rc, "gsb_raster_backward_flat")
    return flat


def count_pairs(st):
    """Work counters per camera of a forward state: int64 tensor [n_cams, 6] = walked pairs (sum of n_contrib, B of
    SURVEY.md 8(d)), composited pairs (alpha >= 1/255 and blended, B_alpha), pairs the forward's inner loop evaluated,
    lane slots it issued (evaluated / slots = lane utilisation), list entries staged after tile culling, list entries."""
    n = st.args.n_cams
    dev = st.geom.device
    counts = torch.zeros((n, 6), dtype=torch.int64, device=dev)
    with L.device_guard(dev):
        rc = L.lib().gsb_raster_count_pairs(C.byref(st.args), C.byref(st.ctx), L.ptr(counts), L.stream_ptr())
    L.check(rc, "gsb_raster_count_pairs")
    return counts


def mark_visible(means3D, viewmatrix, projmatrix):
    """bool[P]: p_view.z > 0.2 (DGR/rasterize_points.cu:196-215)."""
    L.require_cuda(means3D, "means3D")
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P > 0:
        m = means3D.contiguous().float()
        v = viewmatrix.contiguous().float()
        p = projmatrix.contiguous().float()
        with L.device_guard(means3D.device):
            rc = L.lib().gsb_mark_visible(P, L.ptr(m), L.ptr(v), L.ptr(p), L.ptr(present), L.stream_ptr())
        L.check(rc, "gsb_mark_visible")
    return present
