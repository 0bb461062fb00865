"""Camera-sharded data parallelism (the reference has no multi-GPU code).

One process per GPU.  The cameras of a timestep are dealt over the ranks (round-robin, or cost-balanced with
shard_cameras(costs=)); every rank rasterizes its cameras against the replicated Gaussian set; its backward writes
the step's whole per-Gaussian record -- parameter gradients, the densification-statistic increments of THIS step
and the step's largest radius -- into ONE flat [P, 20] buffer (gsb_raster_backward_flat), and ONE exchange reduces
it in place over the ranks: SUM for every column except col 16 (max_2d_radii), which takes MAX.  The densification
statistic is the per-camera norm of the means2D gradient, so norms are taken BEFORE the reduction
(external.py:113-124).  After the exchange every rank holds identical gradients and statistics, so the replicated
parameters stay bit-identical without a broadcast.

Exchange.allreduce() runs the hand-written NVLink kernel (csrc/comm.cu: every rank's flat buffer is mapped by all
peers through CUDA IPC; slice r is pulled from all peers, reduced and pushed back by rank r in one launch); on CPU
tensors (gloo tests) or with GSB_EXCHANGE=nccl it is torch.distributed: one SUM all_reduce + one MAX on the radius
column.
"""
import ctypes as C
import os

import torch
import torch.distributed as dist

from . import _lib as L

# column layout of the flat buffer (include/gsb200.h: gsb_raster_backward_flat)
COLS = {"means3D": (0, 3), "colors": (3, 6), "rotations": (6, 10), "opacities": (10, 11), "scales": (11, 14),
        "mean_2d_gradients_accumulated": (14, 15), "visibility_count": (15, 16), "max_2d_radii": (16, 17),
        "segmentation_colors": (17, 20)}
WIDTH = L.FLAT_WIDTH
MAX_COL = L.FLAT_COL_MAXRADIUS
_SCALARS = ("mean_2d_gradients_accumulated", "visibility_count", "max_2d_radii")


def shard_cameras(n_cams, rank, world, costs=None):
    """Indices of the cameras rank `rank` renders.  Without costs: round-robin (camera c -> rank c % world).
    With per-camera costs (camera_costs of the previous step or timestep; every rank must pass the same numbers):
    longest-processing-time-first -- cameras in decreasing cost, each to the least loaded rank.  A step ends with
    an allreduce, so it lasts as long as its slowest rank, and cameras of one rig differ in cost by tens of percent
    (distance to the subject); ranks may end up with different camera counts."""
    if costs is None:
        return list(range(rank, n_cams, world))
    costs = [float(c) for c in costs]
    assert len(costs) == n_cams
    load = [0.0] * world
    mine = [[] for _ in range(world)]
    for c in sorted(range(n_cams), key=lambda i: (-costs[i], i)):
        r = min(range(world), key=lambda j: (load[j], j))
        load[r] += costs[c]
        mine[r].append(c)
    return sorted(mine[rank])


# cost model of one camera of a step, in ms, fitted to the per-stage times of the C2 workload on a B200
# (bench.py stage_ms): blend forward + backward per walked (pixel, Gaussian) pair, key emission + instance sort per
# instance, and the per-Gaussian / per-pixel stages that every camera pays alike.
COST_PER_PAIR_MS = 3.3e-9
COST_PER_INSTANCE_MS = 8.7e-9
COST_PER_CAMERA_MS = 0.044


def camera_costs(state):
    """Estimated cost of each camera of a rendered batch (RasterState of raster.forward): list of floats."""
    c = state.ctx
    n = c.n_cams
    walked = state._view(c.n_contrib, state.img, torch.int32, (n, c.height * c.width)).long().sum(dim=1).double()
    seg = state._view(c.seg_start, state.geom, torch.int32, (n + 1,)).long()
    inst = (seg[1:] - seg[:-1]).double()
    cost = COST_PER_PAIR_MS * walked + COST_PER_INSTANCE_MS * inst + COST_PER_CAMERA_MS
    return [float(x) for x in cost.cpu()]


# ------------------------------------------------------------------------------------------------
# Tile-band sharding (SURVEY.md 8(e) "finer option"): the unit of work is a tile ROW of a camera, so 27 cameras (or
# train.py's 5) divide evenly over any number of ranks.  Legal because the loss is a plain sum over views and, inside
# a view, over pixels (train.py:402-418; L1 and the SSIM map are means over pixels).  A rank renders the tile rows of
# its band plus one halo row on each side (the 11x11 SSIM window reaches 5 pixels across), evaluates the loss on its
# own pixel rows only, and back-propagates through everything it rendered; the ranks' records add up to the whole
# camera's.  The one non-additive quantity is the densification statistic -- the NORM of a camera's screen-space
# gradient (external.py:113-124) -- so a camera whose rows are split travels as its two gradient components in an extra
# column pair of the flat record and the norm is taken after the exchange (gsb_flat_apply_stats).
# ------------------------------------------------------------------------------------------------
def tile_row_costs(state):
    """Cost of every tile row of every camera of a rendered batch: [n_cams][tiles_y] floats (same model as
    camera_costs, the per-camera constant spread over the rows)."""
    c = state.ctx
    n, H, W, gy = c.n_cams, c.height, c.width, c.tiles_y
    nc = state._view(c.n_contrib, state.img, torch.int32, (n, H, W)).long().sum(dim=2)          # per pixel row
    pad = gy * 16 - H
    if pad:
        nc = torch.cat([nc, torch.zeros((n, pad), dtype=nc.dtype, device=nc.device)], dim=1)
    walked = nc.view(n, gy, 16).sum(dim=2).double()
    seg = state._view(c.seg_start, state.geom, torch.int32, (n + 1,)).long()
    inst = (seg[1:] - seg[:-1]).double()
    share = walked / walked.sum(dim=1, keepdim=True).clamp_min(1.0)
    cost = COST_PER_PAIR_MS * walked + (COST_PER_INSTANCE_MS * inst)[:, None] * share + COST_PER_CAMERA_MS / gy
    return [[float(x) for x in row] for row in cost.cpu()]


def shard_bands(row_costs, world):
    """Contiguous cost-balanced partition of the (camera, tile row) sequence.  row_costs [n_cams][tiles_y] (every rank
    must pass the same numbers).  -> (bands, split_slots): bands[rank] = [(camera, row0, row1), ...] with at most two
    partial cameras per rank; split_slots = {camera: slot} for the cameras whose rows ended up on more than one rank."""
    n_cams, gy = len(row_costs), len(row_costs[0])
    flat_cost = [float(c) for row in row_costs for c in row]
    total = sum(flat_cost)
    cuts, run, k = [0], 0.0, 1
    for i, c in enumerate(flat_cost):   # unit i goes to the rank whose share its midpoint falls into
        mid = run + 0.5 * c
        while k < world and mid >= k * total / world:
            cuts.append(i)
            k += 1
        run += c
    while len(cuts) < world:
        cuts.append(len(flat_cost))
    cuts.append(len(flat_cost))
    bands = []
    for r in range(world):
        a, b = cuts[r], cuts[r + 1]
        mine = []
        cam = a // gy
        while cam * gy < b:
            r0, r1 = max(a - cam * gy, 0), min(b - cam * gy, gy)
            if r1 > r0:
                mine.append((cam, r0, r1))
            cam += 1
        bands.append(mine)
    owners = {}
    for r, mine in enumerate(bands):
        for cam, r0, r1 in mine:
            owners.setdefault(cam, []).append(r)
    split = [cam for cam in sorted(owners) if len(owners[cam]) > 1]
    return bands, {cam: j for j, cam in enumerate(split)}


def flat_width(n_split):
    """Floats per row of the flat record with n_split split-camera column pairs (a multiple of 4)."""
    return WIDTH + 4 * ((2 * n_split + 3) // 4)


def band_tables(mine, split_slots, tiles_y, height, device):
    """Device tables of one rank's bands: entries [n, 4] for raster.forward (RENDER rows = band +- one halo tile row,
    split slot, whether this entry counts the camera's visibility: the band that starts at row 0 does) and rows [n, 4]
    for the loss (loss pixel rows, rows of dL/dimage to write = the rendered rows)."""
    ent, rows = [], []
    for cam, r0, r1 in mine:
        h0, h1 = max(r0 - 1, 0), min(r1 + 1, tiles_y)
        ent.append([h0, h1, split_slots.get(cam, -1), 1 if r0 == 0 else 0])
        rows.append([16 * r0, min(16 * r1, height), 16 * h0, min(16 * h1, height)])
    return (torch.tensor(ent, dtype=torch.int32, device=device).contiguous(),
            torch.tensor(rows, dtype=torch.int32, device=device).contiguous())


def band_batch(cams, targets, mine, split_slots, device):
    """One rank's batch under tile-band sharding: `cams` = gsb200.scene.stack_cameras / gsb200.io.stack_views tables of ALL
    cameras, targets [n_cams, 3, H, W] (or None), mine = shard_bands(...)[0][rank].  -> (cams dict for
    gsb200.step.timestep_step incl. "entries" / "rows" / "tanfov_dev", targets of the batch's cameras)."""
    from . import raster
    ix = [cam for cam, _, _ in mine]
    H, W = cams["image_height"], cams["image_width"]
    sub = {"viewmatrix": cams["viewmatrix"][ix].contiguous(), "projmatrix": cams["projmatrix"][ix].contiguous(),
           "tanfovx": [cams["tanfovx"][i] for i in ix], "tanfovy": [cams["tanfovy"][i] for i in ix],
           "image_width": W, "image_height": H}
    sub["tanfov_dev"] = raster.tanfov_table(sub["tanfovx"], sub["tanfovy"], len(ix), device)
    sub["entries"], sub["rows"] = band_tables(mine, split_slots, (H + 15) // 16, H, device)
    return sub, (targets[ix].contiguous() if targets is not None else None)


def pack(flat, grads, stats_delta=None, max_radii=None):
    """Slice-copy form of the flat record for tensors that were NOT produced by gsb_raster_backward_flat (the gloo
    tests, callers holding separate gradient tables): gradients, THIS step's statistic increments (not running
    totals) and this step's largest radius per Gaussian into flat [P, 20]."""
    P = flat.shape[0]
    flat.zero_()
    for k in ("means3D", "colors", "rotations", "opacities", "scales"):
        a, b = COLS[k]
        flat[:, a:b] = grads[k].reshape(P, b - a)
    if "segmentation_colors" in grads:
        flat[:, 17:20] = grads["segmentation_colors"].reshape(P, 3)
    if stats_delta is not None:
        flat[:, 14] = stats_delta["mean_2d_gradients_accumulated"]
        flat[:, 15] = stats_delta["visibility_count"]
    if max_radii is not None:
        flat[:, 16] = max_radii
    return flat


def unpack(flat):
    """Views into the flat record: name -> [P, k] (the three statistics: [P])."""
    out = {}
    for k, (a, b) in COLS.items():
        v = flat[:, a:b]
        out[k] = v[:, 0] if k in _SCALARS else v
    return out


def allreduce(flat, group=None):
    """The exchange through torch.distributed (gloo on CPU, NCCL on GPUs): SUM of the record, MAX of the radius
    column.  In place; returns flat."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        rad = flat[:, MAX_COL].contiguous()
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(rad, op=dist.ReduceOp.MAX, group=group)
        flat[:, MAX_COL] = rad
    return flat


class _DevMem:
    """A raw device allocation exposed through the CUDA array interface (so torch can alias it without owning it)."""

    def __init__(self, ptr, nfloats):
        self.__cuda_array_interface__ = {"shape": (nfloats,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class Exchange:
    """The flat step record of this rank in peer-mapped memory + the one-kernel all-reduce over it.

        ex = Exchange(P, device)                # collective: every rank of the (default) group calls it
        ... gsb_raster_backward_flat writes ex.flat ...
        ex.allreduce()                          # in place; afterwards ex.flat is identical on every rank

    world == 1 (or no process group): a plain device buffer, allreduce() does nothing."""

    def __init__(self, P, device, group=None, width=WIDTH):
        self.P, self.device, self.group = int(P), torch.device(device), group
        self.width = int(width)   # > 20: split-camera column pairs of tile-band sharding (flat_width)
        if self.width < WIDTH or self.width % 4:
            raise ValueError("flat width must be 20 + 4 k")
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.mode = os.environ.get("GSB_EXCHANGE", "nvlink") if (self.world > 1 and self.device.type == "cuda") else "local"
        self.block = None
        self.peer_blocks = []
        self.payload = self.P * self.width * 4
        if self.mode == "nvlink":
            lib = L.lib()
            self.nbytes = lib.gsb_comm_bytes(self.payload)
            blk = C.c_void_p()
            handle = C.create_string_buffer(64)
            with L.device_guard(self.device):
                L.check(lib.gsb_comm_alloc(self.nbytes, C.byref(blk), handle), "gsb_comm_alloc")
            self.block = blk.value
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle.raw), group=group)
            self.ptrs = (C.c_void_p * self.world)()
            for r, h in enumerate(handles):
                if r == self.rank:
                    self.ptrs[r] = self.block
                    continue
                p = C.c_void_p()
                with L.device_guard(self.device):
                    L.check(lib.gsb_comm_open(h, C.byref(p)), "gsb_comm_open")
                self.peer_blocks.append(p.value)
                self.ptrs[r] = p.value
            self._mem = _DevMem(self.block, self.P * self.width)
            self.flat = torch.as_tensor(self._mem, device=self.device).view(self.P, self.width)
            dist.barrier(group=group)   # every rank has mapped every block before anyone launches
        else:
            self.flat = torch.zeros((self.P, self.width), dtype=torch.float32, device=self.device)

    def allreduce(self):
        if self.world == 1:
            return self.flat
        if self.mode != "nvlink":
            return allreduce(self.flat, self.group)
        with L.device_guard(self.device):
            rc = L.lib().gsb_allreduce_flat(self.rank, self.world, self.ptrs, self.payload, self.P * (self.width // 4),
                                            self.width // 4, MAX_COL // 4, MAX_COL % 4, L.stream_ptr())
        L.check(rc, "gsb_allreduce_flat")
        return self.flat

    def close(self):
        if self.block is None:
            return
        torch.cuda.synchronize(self.device)
        if dist.is_initialized():
            dist.barrier(group=self.group)   # no peer is still inside a kernel that touches this block
        lib = L.lib()
        self.flat = None
        for p in self.peer_blocks:
            lib.gsb_comm_close(p)
        lib.gsb_comm_free(self.block)
        self.block, self.peer_blocks = None, []
