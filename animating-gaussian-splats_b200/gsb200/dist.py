"""Camera-sharded data parallelism (the reference has no multi-GPU code).

One process per GPU.  The cameras of a timestep are dealt round-robin (camera c -> rank c % world); every
rank rasterizes its cameras against the replicated Gaussian set; parameter gradients and the summed
densification statistics are packed into ONE flat [P, 16] buffer and summed with ONE allreduce (NCCL over
NVLink on GPUs, gloo in the CPU tests); max_2d_radii takes a MAX allreduce.  The densification
statistic is the per-camera norm of the means2D gradient, so norms are taken BEFORE the reduction
(external.py:113-124).  After the reduction every rank holds identical gradients and statistics, so
the replicated parameters stay bit-identical without a broadcast.
"""
import torch
import torch.distributed as dist

# column layout of the flat buffer
COLS = {"means3D": (0, 3), "colors": (3, 6), "rotations": (6, 10), "opacities": (10, 11), "scales": (11, 14),
        "mean_2d_gradients_accumulated": (14, 15), "visibility_count": (15, 16)}
WIDTH = 16


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


def pack(flat, grads, stats_delta=None):
    """Write this rank's camera-summed gradients (and statistic increments) into flat [P, 16].
    CUDA tensors go through one gsb_pack_flat launch; CPU tensors (gloo tests) through slice copies."""
    if flat.is_cuda:
        from . import _lib as L
        c = lambda t: t.contiguous()
        acc = c(stats_delta["mean_2d_gradients_accumulated"]) if stats_delta is not None else None
        cnt = c(stats_delta["visibility_count"]) if stats_delta is not None else None
        keep = [c(grads[k]) for k in ("means3D", "colors", "rotations", "opacities", "scales")]
        with L.device_guard(flat.device):
            rc = L.lib().gsb_pack_flat(flat.shape[0], L.ptr(keep[0]), L.ptr(keep[1]), L.ptr(keep[2]), L.ptr(keep[3]),
                                       L.ptr(keep[4]), L.ptr(acc), L.ptr(cnt), L.ptr(flat), L.stream_ptr())
        L.check(rc, "gsb_pack_flat")
        return flat
    for k in ("means3D", "colors", "rotations", "opacities", "scales"):
        a, b = COLS[k]
        flat[:, a:b] = grads[k].reshape(flat.shape[0], b - a)
    if stats_delta is not None:
        flat[:, 14] = stats_delta["mean_2d_gradients_accumulated"]
        flat[:, 15] = stats_delta["visibility_count"]
    else:
        flat[:, 14:16] = 0
    return flat


def unpack(flat):
    out = {}
    for k, (a, b) in COLS.items():
        v = flat[:, a:b]
        out[k] = v[:, 0] if k in ("mean_2d_gradients_accumulated", "visibility_count") else v
    return out


def allreduce(flat, max_2d_radii=None, group=None):
    """The step's exchange: one SUM allreduce of the flat buffer (+ one MAX for max_2d_radii)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if max_2d_radii is not None:
            dist.all_reduce(max_2d_radii, op=dist.ReduceOp.MAX, group=group)
    return flat
