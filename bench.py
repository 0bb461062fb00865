#!/usr/bin/env python
"""bench.py -- headline benchmark of the rasterizer + L1/SSIM + densify-statistics hot path.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference ...                     (the CPU arm: C oracle on the host cores)
    python bench.py --impl reference_cuda ...                (extra: the unmodified reference CUDA extension)

Workload (BASELINE.json configs[1] / configs[2]): one Panoptic-shaped synthetic timestep, 300 000
Gaussians, 27 cameras 640x360.  One STEP = one pass of the hot path over all 27 cameras:
    rasterize forward (batched over the rank's cameras) -> fused L1+SSIM loss and dL/dimage ->
    rasterize backward (gradients summed over cameras) -> densification statistics per camera
and, for N > 1, ONE NCCL allreduce of the flat gradient + statistics buffer (cameras are sharded
c -> rank c % N; total work is fixed, so scaling is "strong").
metric = train iters/sec, one iteration = one camera's fwd + loss + bwd (the unit of densify.py:218-247),
so value = 27 / step time, whole job.  Gaussian-pixel blends/sec (fwd+bwd) is reported beside it.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "animating-gaussian-splats_b200")
for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402


# ------------------------------------------------------------------------------------------------
def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference_cuda"])
    ap.add_argument("--gaussians", type=int, default=300_000)
    ap.add_argument("--cams", type=int, default=27)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--height", type=int, default=360)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-reference-cuda", action="store_true")
    ap.add_argument("--no-c5", action="store_true", help="skip the sequence + densify leg (BASELINE.json configs[4])")
    ap.add_argument("--profile-only", action="store_true", help="one warm step + 2 steps, for ncu")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p.get("sm_max_mhz", 1965.0)), "measured"
    except Exception:
        return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: an NVML polling thread (nvidia_ml_py), started
    before the warm-up so that its set-up cost stays out of the timed steps; only samples taken between
    begin() and end() count.  (An nvidia-smi child process needs ~0.3 s to come up, longer than the timed region,
    and its start-up was seen to stall kernel launches for tens of ms.)"""
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index=0, period=0.002):
        self.index, self.period = index, period
        self.samples, self.reasons, self.active, self.stop_flag = [], 0, False, False
        self.h = self.nv = self.thread = None
        self.max_mhz = None

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.nv = nv
            self.h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
        except Exception:
            self.nv = None
            return
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _reasons(self):
        nv = self.nv
        for name in ("nvmlDeviceGetCurrentClocksEventReasons", "nvmlDeviceGetCurrentClocksThrottleReasons"):
            fn = getattr(nv, name, None)
            if fn is not None:
                try:
                    return int(fn(self.h))
                except Exception:
                    pass
        return 0

    def _run(self):
        nv = self.nv
        while not self.stop_flag:
            if self.active:
                try:
                    self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                    self.reasons |= self._reasons()
                except Exception:
                    pass
            time.sleep(self.period)

    def begin(self):
        self.active = True

    def end(self):
        self.active = False

    def stop(self):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["NVML unavailable"]}
        self.stop_flag = True
        self.thread.join(timeout=2)
        names = sorted(k for k, b in self.BITS.items() if self.reasons & b)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": names, "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
def build_workload(args):
    """Seeded synthetic timestep (CPU tensors): raw parameters, cameras, and the raw parameters of the
    perturbed copy whose renders are the target images."""
    from gsb200 import scene
    g = scene.make_gaussians(args.gaussians, seed=0)
    cams = scene.make_cameras(args.cams, args.width, args.height, seed=0)
    rng = np.random.default_rng(7)
    tgt = {k: v.clone() for k, v in g.items()}
    tgt["means"] = tgt["means"] + torch.from_numpy(rng.normal(0, 0.01, size=tuple(g["means"].shape)).astype(np.float32))
    tgt["colors"] = (tgt["colors"] + torch.from_numpy(rng.normal(0, 0.1, size=tuple(g["colors"].shape)).astype(np.float32))).clamp(0, 1)
    return g, tgt, cams


def render_targets(tgt_raw, cams_stacked, n, bg, dev):
    from gsb200 import raster, scene
    a = {k: v.to(dev).contiguous() for k, v in scene.activate(tgt_raw).items()}
    color, _, _, _ = raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None,
                                    bg, 1.0, cams_stacked["viewmatrix"], cams_stacked["projmatrix"],
                                    cams_stacked["tanfovx"], cams_stacked["tanfovy"], cams_stacked["image_height"],
                                    cams_stacked["image_width"], n_cams=n)
    return (color if n > 1 else color[None]).clone()


# ------------------------------------------------------------------------------------------------
def _host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def full_record(act, cams, targets, bg, dev):
    """The N = 1 result of the step over ALL cameras on this rank: the flat record the exchange must reproduce."""
    from gsb200 import step as step_mod, dist as gd
    flat = torch.zeros((act["means3D"].shape[0], gd.WIDTH), dtype=torch.float32, device=dev)
    step_mod.timestep_step(act, cams, targets, bg, None, flat=flat)
    return flat


def record_parity(got, want):
    """north_star tolerances on the exchanged record: gradients 1e-4 relative (absolute floor 1e-5 of the column's
    largest magnitude -- float atomics are unordered), visibility counts and radii exact."""
    from gsb200 import dist as gd
    g, w = got.double(), want.double()
    if got.shape[1] > gd.WIDTH:   # tile-band sharding: split cameras carry their summed 2-D gradient, the norm is taken here
        g = g[:, :gd.WIDTH].clone()
        for j in range((got.shape[1] - gd.WIDTH) // 2):
            g[:, 14] += torch.sqrt(got[:, gd.WIDTH + 2 * j].double() ** 2 + got[:, gd.WIDTH + 2 * j + 1].double() ** 2)
        got = got[:, :gd.WIDTH]
    out = {"ok": True}
    worst = 0.0
    for k, (a, b) in gd.COLS.items():
        if k in ("visibility_count", "max_2d_radii"):
            eq = bool(torch.equal(got[:, a:b], want[:, a:b]))
            out[k + "_exact"] = eq
            out["ok"] = out["ok"] and eq
            continue
        d = (g[:, a:b] - w[:, a:b]).abs()
        scale = torch.maximum(g[:, a:b].abs(), w[:, a:b].abs())
        floor = 1e-5 * float(w[:, a:b].abs().max()) + 1e-30
        rel = float((d / (scale + floor / 1e-4)).max())   # d <= 1e-4 * scale + floor  <=>  rel <= 1e-4
        worst = max(worst, rel)
        out["ok"] = out["ok"] and rel <= 1e-4
    out["max_rel"] = worst
    return out


def run_ours(args):
    import torch.distributed as dist
    from gsb200 import _lib, raster, scene, step as step_mod, dist as gd
    import ctypes as C

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl ours needs a CUDA device (no CPU fallback exists)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.lib()
    use_graph = os.environ.get("GSB_BENCH_GRAPH", "1") != "0"

    g_raw, tgt_raw, cams = build_workload(args)
    P = args.gaussians
    bg = torch.zeros(3, device=dev)
    # camera -> rank.  N > 1: cost-balanced (gsb200.dist.shard_cameras with per-camera costs): every rank runs the
    # whole 27-camera step once, untimed -- its forward state gives the per-camera costs (what a trainer takes from the
    # previous timestep's renders; rank 0's numbers are broadcast so all ranks agree) and its flat record is the N = 1
    # answer that the exchanged record is checked against below.
    raw_keys = ["means", "colors", "rotation_quaternions", "opacity_logits", "log_scales"]
    host_raw = {k: g_raw[k].contiguous().pin_memory() for k in raw_keys}
    dev_raw = {k: torch.empty_like(v, device=dev) for k, v in host_raw.items()}
    for k in raw_keys:
        dev_raw[k].copy_(host_raw[k])
    act = {k: v.contiguous() for k, v in scene.activate(dev_raw).items()}   # the activated tables every leg below uses
    cam_costs = None
    want_flat = None
    bands = split = None
    full_counts = None
    # tile-band sharding is built and parity-checked (GSB_BENCH_BANDS=1) but whole cameras are the default here: at 27
    # cameras x 300 k Gaussians every extra batch entry costs ~0.06 ms of per-Gaussian work and the split-camera columns
    # widen the exchanged record (24 -> 43 MB at N = 8), which together outweigh the better balance (1.276 vs 1.265 ms
    # at N = 8, 2.879 vs 2.889 ms at N = 2); bands are what lets FEWER cameras than ranks use every GPU
    use_bands = world > 1 and os.environ.get("GSB_BENCH_BANDS", "1" if args.cams < world else "0") != "0"
    if world > 1:
        all_sc = scene.stack_cameras(cams, dev)
        all_targets = render_targets(tgt_raw, all_sc, args.cams, bg, dev)
        want_flat = torch.zeros((P, gd.WIDTH), dtype=torch.float32, device=dev)
        _, _, _, st0 = step_mod.timestep_step(act, all_sc, all_targets, bg, None, flat=want_flat)
        full_counts = raster.count_pairs(st0).sum(dim=0).tolist()[:2] + [st0.num_rendered]
        if use_bands:
            # tile-band sharding: the unit is a tile row of a camera, cut where the cumulative cost crosses k / N
            rc_ = torch.tensor(gd.tile_row_costs(st0), dtype=torch.float64, device=dev)
            dist.broadcast(rc_, 0)
            bands, split = gd.shard_bands([[float(x) for x in row] for row in rc_.cpu()], world)
        elif os.environ.get("GSB_BENCH_ROUND_ROBIN", "0") != "1":
            ct = torch.tensor(gd.camera_costs(st0), dtype=torch.float64, device=dev)
            dist.broadcast(ct, 0)
            cam_costs = [float(x) for x in ct.cpu()]
        del st0
    if use_bands:
        mine = bands[rank]
        my_cam_ids = [c for c, _, _ in mine]
        sc, targets = gd.band_batch(all_sc, all_targets, mine, split, dev)
        targets = targets.clone()
        n_my = len(mine)
        flat_w = gd.flat_width(len(split))
    else:
        my_cam_ids = gd.shard_cameras(args.cams, rank, world, cam_costs)
        mine = None
        n_my = len(my_cam_ids)
        sc = scene.stack_cameras([cams[i] for i in my_cam_ids], dev)
        sc["tanfov_dev"] = raster.tanfov_table(sc["tanfovx"], sc["tanfovy"], n_my, dev)
        targets = render_targets(tgt_raw, sc, n_my, bg, dev)
        flat_w = gd.WIDTH
    my_cams = [cams[i] for i in my_cam_ids]
    if world > 1:
        del all_sc, all_targets

    # host (pinned) copies for the end-to-end leg
    host_targets = targets.cpu().pin_memory()
    dev_targets = torch.empty_like(targets)
    dev_targets.copy_(host_targets)

    stats = step_mod.DensifyStats(P, dev)
    # the step record ([P, 20]: gradients + this step's statistic increments + largest radius) lives in peer-mapped
    # memory; the backward writes it, ONE kernel reduces it over the ranks in place
    ex = gd.Exchange(P, dev, width=flat_w)
    # arena of the host-free forward: capacity from an untimed render of this rank's cameras
    _, _, _, st_probe = step_mod.timestep_step(act, sc, dev_targets, bg, None)
    R_probe = st_probe.num_rendered
    del st_probe
    arena = raster.RasterArena(dev).reserve(P, n_my, args.width, args.height, capacity=int(1.25 * R_probe) + 4096)

    # GSB_BENCH_GROUPS=k: render the rank's cameras as k concurrent groups (gsb200.step.CameraGroups)
    # measured at 3 and 4 cameras per GPU: SLOWER than one batch (1.14-1.35 vs 0.95-1.02 ms): each group pays the
    # launch-bound per-batch kernels (depth sort, scan, set-up) again, which costs more than the overlapped tails save;
    # one batch is the default
    n_groups = int(os.environ.get("GSB_BENCH_GROUPS", "0")) or 1
    cg = step_mod.CameraGroups(sc, dev_targets, n_groups, P, dev).reserve(act, bg) if n_groups > 1 else None

    def one_step():
        """The product step: forward -> fused loss -> backward into the flat record -> exchange -> statistics."""
        if cg is not None:
            loss_stats, flat, sts = cg.step(act, bg, ex.flat)
            st = sts[0]
        else:
            loss_stats, flat, radii, st = step_mod.timestep_step(act, sc, dev_targets, bg, None, flat=ex.flat, arena=arena)
        ex.allreduce()
        stats.apply_flat(ex.flat)
        return loss_stats, flat, st

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- N > 1: the exchanged record must equal the N = 1 record --------------------------------------
    parity = None
    loss_stats, flat, st = one_step()
    torch.cuda.synchronize()
    err, R_dev, _ = st.status()
    if err != 0:
        raise RuntimeError(f"asynchronous forward failed: status {err}")
    if world > 1:
        parity = record_parity(ex.flat, want_flat)
        parity["transport"] = "one-kernel NVLink all-reduce over CUDA-IPC peer memory" if ex.mode == "nvlink" else ex.mode
        ref0 = ex.flat.clone()
        dist.broadcast(ref0, 0)
        parity["replicas_bit_identical"] = bool(torch.equal(ref0, ex.flat))
        ok = torch.tensor([1.0 if (parity["ok"] and parity["replicas_bit_identical"]) else 0.0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        parity["ok"] = bool(ok.item() == 1.0)
        del want_flat, ref0
        print(f"nccl_parity: {'ok' if parity['ok'] else 'FAILED'}, max_rel={parity['max_rel']:.3e}", file=sys.stderr)
        if not parity["ok"]:
            raise RuntimeError(f"exchanged record differs from the single-GPU record: {parity}")

    graph = step_mod.StepGraph(one_step) if use_graph else None
    run_step = graph.replay if graph is not None else one_step

    if args.profile_only:
        for _ in range(3):
            one_step()
        for _ in range(3):
            run_step()
        torch.cuda.synchronize()
        return

    # ---- device-resident timing -------------------------------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        run_step()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.begin()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    e0.record()
    marks[0].record()
    for k in range(args.steps):
        run_step()
        marks[k + 1].record()
    e1.record()
    sync_all()
    sampler.end()
    per_step = [marks[k].elapsed_time(marks[k + 1]) for k in range(args.steps)]
    print("per-step ms:", " ".join(f"{x:.2f}" for x in per_step), file=sys.stderr)
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    if st.status()[0] != 0:
        raise RuntimeError("arena overflow inside the timed region")

    def timed_loop(fn, iters=10):
        fn()
        sync_all()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        sync_all()
        tt_ = torch.tensor([a.elapsed_time(b) / iters], dtype=torch.float64, device=dev)
        own = float(tt_.item())
        if world > 1:
            dist.all_reduce(tt_, op=dist.ReduceOp.MAX)
        timed_loop.own = own
        return float(tt_.item())

    # the same step issued launch by launch from Python (no graph): what the graph removes
    ms_eager = timed_loop(one_step) if use_graph else ms_step

    # ---- N > 1: what the exchange costs (each piece timed alone, max over ranks) -----------------
    exchange = None
    if world > 1:
        def local_step():
            if cg is not None:
                cg.step(act, bg, ex.flat)
            else:
                step_mod.timestep_step(act, sc, dev_targets, bg, None, flat=ex.flat, arena=arena)
        g_local = step_mod.StepGraph(local_step) if use_graph else None
        exchange = {
            "allreduce_kernel": timed_loop(ex.allreduce),
            "apply_stats": timed_loop(lambda: stats.apply_flat(ex.flat)),
            "step_without_exchange": timed_loop(g_local.replay if g_local is not None else local_step),
            "bytes": ex.flat.numel() * 4, "transport": parity["transport"],
        }
        own_local = timed_loop.own
        flat_n = torch.zeros_like(ex.flat)
        exchange["nccl_allreduce_sum_same_bytes"] = timed_loop(lambda: dist.all_reduce(flat_n))
        del flat_n
        per_rank = [None] * world
        dist.all_gather_object(per_rank, {"rank": rank, "cameras": my_cam_ids, "bands": mine, "step_without_exchange": own_local,
                                          "est_cost": None if cam_costs is None else sum(cam_costs[c] for c in my_cam_ids)})
        exchange["per_rank"] = per_rank

    # ---- end to end: host buffers in, loss out, every step ------------------------------------
    # gsb200.step.HostFeed: each step's inputs (raw parameter tables + target images, pinned host memory) are
    # copied host -> device (double-buffered, the copy of step k+1 overlaps step k), activations + step run on
    # the device, the per-camera loss numbers come back to pinned host memory; gradients stay on the device (the
    # optimiser consumes them there), only the 3 loss numbers per camera return to the host.
    feed = step_mod.HostFeed(host_raw, host_targets, dev)
    E2E_BLOCKING = os.environ.get("GSB_BENCH_E2E_BLOCKING", "0") == "1"

    def after(loss_stats_, flat_, radii_, st_):
        ex.allreduce()
        stats.apply_flat(ex.flat)

    def e2e_step():
        # non-blocking: the loss numbers are copied to pinned host memory every step, the host does not wait for
        # them before issuing the next step (a training loop only logs them); GSB_BENCH_E2E_BLOCKING=1 waits
        feed.step(sc, bg, None, after, wait=E2E_BLOCKING, flat=ex.flat, arena=arena, graph=use_graph)

    for _ in range(3):   # both buffer halves captured (every rank captures at the same steps: the warm-ups run the exchange)
        e2e_step()
    sync_all()
    sampler.begin()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    sync_all()
    sampler.end()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_e2e = float(t.item()) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    h2d = feed.h2d_bytes
    d2h = feed.d2h_bytes

    # ---- per-stage timing of one step (CUDA events inside the library, on the launching stream) --
    lib.gsb_profile_enable(1)
    stage_acc = np.zeros(len(_lib.PROFILE_STAGES))
    reps = 3
    buf = (C.c_float * len(_lib.PROFILE_STAGES))()
    for _ in range(reps):   # the rank's cameras as ONE batch: the stage events of concurrent groups would interleave
        loss_stats, flat, radii_, st = step_mod.timestep_step(act, sc, dev_targets, bg, None, flat=ex.flat, arena=arena)
        torch.cuda.synchronize()
        lib.gsb_profile_read(buf, len(_lib.PROFILE_STAGES))
        stage_acc += np.array(list(buf))
    lib.gsb_profile_enable(0)
    stages = {k: float(v / reps) for k, v in zip(_lib.PROFILE_STAGES, stage_acc) if not k.endswith("_begin")}

    # ---- workload statistics ---------------------------------------------------------------------
    pairs = raster.count_pairs(st).sum(dim=0)                     # (walked = sum of n_contrib, composited)
    walked, blended = int(pairs[0].item()), int(pairs[1].item())
    fwd_work = {"evaluated_pairs": int(pairs[2]), "lane_slots": int(pairs[3]), "lane_utilisation": float(pairs[2]) / max(float(pairs[3]), 1.0),
                "entries_staged": int(pairs[4]), "entries_listed": int(pairs[5])}
    R = st.num_rendered
    tile_bits = int(st.ctx.tile_bits)
    cnt = torch.tensor([walked, R, blended], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(cnt)
    walked_all, R_all, blended_all = int(cnt[0].item()), int(cnt[1].item()), int(cnt[2].item())
    rendered_walked = walked_all   # what the ranks actually walked (bands: the halo rows are rendered twice)
    if full_counts is not None:    # the metric's pair counts are those of the N = 1 step
        walked_all, blended_all, R_all = int(full_counts[0]), int(full_counts[1]), int(full_counts[2])
    peak_mem = torch.cuda.max_memory_allocated(dev)
    extras = os.environ.get("GSB_BENCH_EXTRAS", "1") != "0"

    # ---- C5: the sequence + densify config, on all ranks (after the headline measurements) ------------------
    c5 = None
    if extras and not args.no_c5:
        del graph
        graph = None
        torch.cuda.empty_cache()
        try:
            c5 = c5_sequence(args, world, rank, dev)
            torch.cuda.empty_cache()
            # The reference's schedule grows this synthetic cloud by ~3% per event (150 k -> ~420 k before the final
            # prune), not to the ~2 M of BASELINE.json configs[4]; the 2 M operating point is measured from a cloud of
            # that size: optimiser steps, two densify events, three timesteps.
            big = c5_sequence(args, world, rank, dev, P0=2_000_000, P_gt=2_400_000, gt_scale=0.006, steps_per_event=4,
                              events=(500, 600), timesteps=3, steps_per_timestep=2)
            c5["at_2M_gaussians"] = {k: big[k] for k in ("gaussians_start", "gaussians_end", "gaussian_count_over_events",
                                                          "first_frame", "timesteps", "peak_memory_bytes_per_gpu",
                                                          "setup_seconds")}
        except Exception as ex_:
            import traceback
            traceback.print_exc(file=sys.stderr)
            c5 = dict(c5 or {}, error=repr(ex_)[:300])
        torch.cuda.empty_cache()

    if rank != 0:
        ex.close()
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, sm_max_mhz, peak_src = peaks()
    traffic = {}
    try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
        if (args.gaussians, args.cams, args.width, args.height, world) == (300_000, 27, 640, 360, 1):
            for name in ("r2_traffic.json", "r1_traffic.json"):
                fp = os.path.join(ROOT, "profiles", name)
                if os.path.exists(fp):
                    with open(fp) as f:
                        traffic = json.load(f)["kernels"]
                    break
    except Exception:
        traffic = {}
    tr = lambda name: (traffic.get(name, {}).get("dram_bytes_per_launch") if traffic else None)
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    fp32_peak_tflops = n_sm * 128 * 2 * sm_max_mhz * 1e6 / 1e12
    peak_note = f"{n_sm} SMs x 128 lanes x 2 x {sm_max_mhz:.0f} MHz ({peak_src} sm_max_mhz)"
    # dominant kernel: backward blend.  SURVEY 8(d): 70 FP32 flop per CONTRIBUTING pair (alpha >= 1/255 and blended,
    # counted on the GPU by gsb_raster_count_pairs); the walked-pair figure (sum of n_contrib) is kept beside it.
    bwd_ms = stages["render_bwd"]
    ach = lambda pairs_, flop, ms_: (pairs_ * flop) / (ms_ * 1e-3) / 1e12 if ms_ > 0 else 0.0
    achieved = ach(blended, 70.0, bwd_ms)
    roofline = {"kernel": "render_bwd2_kernel", "bound": "fp32", "achieved": achieved, "peak": fp32_peak_tflops,
                "unit": "TFLOP/s", "frac": achieved / fp32_peak_tflops, "traffic": tr("render_bwd2_kernel"),
                "peak_source": peak_note, "ms": bwd_ms, "pairs": blended,
                "pairs_definition": "contributing (pixel, Gaussian) pairs B_alpha, 70 flop each (SURVEY 8d)",
                "frac_walked": ach(walked, 70.0, bwd_ms) / fp32_peak_tflops, "walked_pairs": walked}
    fwd_ms = stages["render_fwd"]
    roofline_fwd = {"kernel": "render_fwd_kernel", "bound": "fp32", "achieved": ach(walked, 24.0, fwd_ms),
                    "peak": fp32_peak_tflops, "unit": "TFLOP/s", "frac": ach(walked, 24.0, fwd_ms) / fp32_peak_tflops,
                    "traffic": tr("render_fwd_kernel"), "peak_source": peak_note, "ms": fwd_ms, "pairs": walked,
                    "pairs_definition": "walked pairs B = sum of n_contrib, 24 flop each (SURVEY 8d)"}
    # the HBM-bound kernel with the most bytes: the instance sort (segmented onesweep on 16-bit camera-local
    # tile ids).  Algorithmic bytes per pass: the histogram reads the key (2); the scatter reads key (2) + payload (4)
    # and writes the payload (4) (+ the key (2) when another pass follows).  DESIGN.md section 5.
    tile_digit = int(os.environ.get("GSB_TILE_DIGIT", "0")) or (5 if tile_bits <= 5 else 8 if tile_bits <= 8 else 10 if tile_bits <= 10 else 8)
    passes = (max(tile_bits, 1) + tile_digit - 1) // tile_digit
    depth_digit = 11 if os.environ.get("GSB_DEPTH_DIGIT") == "11" else 8
    gpasses = (32 + depth_digit - 1) // depth_digit
    sort_bytes = R * (12 * passes + (2 * passes if passes > 1 else 0))
    sort_ms = stages["sort"]
    sort_name = (f"(seg_tile_hist + seg_chunk_scan + seg_scatter)<u16,{tile_digit}> x{passes}: instance sort on camera-local tile ids")
    roofline_hbm = {"kernel": sort_name, "bound": "hbm",
                    "achieved": sort_bytes / (sort_ms * 1e-3) / 1e9 if sort_ms > 0 else 0.0, "peak": hbm_peak,
                    "unit": "GB/s", "ms": sort_ms, "instances": R,
                    "traffic": tr("seg_sort_instances"), "algorithmic_bytes": sort_bytes}
    roofline_hbm["frac"] = roofline_hbm["achieved"] / hbm_peak
    # cam params, preprocess, depth sort (setup + 3 kernels per pass), scan, setup, key emission, instance sort (3 per
    # pass), ranges, (heaviest-first tile order for short launches), blend fwd, ssim fwd/finalize/bwd, blend bwd,
    # preprocess bwd (writes the flat record), exchange kernel (N > 1), statistics from the record
    tiles_total = ((args.width + 15) // 16) * ((args.height + 15) // 16) * n_my
    launches = (1 + 1 + (1 + 3 * gpasses) + 1 + 1 + 1 + 3 * passes + 1 + (1 if tiles_total < 12 * 148 * 6 else 0) + 1 + 3 + 1 + 1 +
                (1 if world > 1 else 0) + 1)

    out = {
        "metric": "train iters/sec (one iter = one camera fwd + L1/SSIM + bwd; 27 cameras per step)",
        "value": args.cams / (ms_step * 1e-3), "unit": "iters/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"Panoptic-shaped timestep: {P} gaussians, {args.cams} cameras {args.width}x{args.height}, "
                               "fwd + L1/SSIM + bwd + densify stats (BASELINE.json configs[1])",
                   "cameras_per_rank": n_my, "parallelism": f"camera-sharded dp{world}",
                   "camera_assignment": ("tile-band sharding: contiguous cost-balanced partition of the (camera, tile row) sequence, "
                                         "one halo tile row per band edge, loss on the band's own pixel rows "
                                         f"({len(split)} split cameras, flat record {flat_w} floats wide)" if use_bands else
                                         "round-robin" if cam_costs is None else
                                         "cost-balanced (LPT on per-camera walked pairs + instances of an untimed calibration render)"),
                   "walked_pairs_rendered_incl_halo": rendered_walked,
                   "camera_groups": n_groups,
                   "step_issue": ("one CUDA graph per step (gsb200.step.StepGraph): host-free forward through a reused arena, "
                                  "flat step record, exchange kernel, statistics") if use_graph else "launch by launch from Python",
                   "l2": "per-step working set (state + instance lists) is several hundred MB > 126 MB L2; no explicit flush"},
        "blends_per_sec": 2.0 * walked_all / (ms_step * 1e-3), "walked_pairs_per_step": walked_all,
        "contributing_pairs_per_step": blended_all, "num_rendered_per_step": R_all,
        "e2e": {"value": args.cams / (ms_e2e * 1e-3), "unit": "iters/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "returns": "the per-camera loss numbers only (3 floats per camera); gradients and statistics stay on the "
                           "device, where the optimiser and the densify step consume them"},
        "gpu_launches": launches * args.steps, "ms_per_step_eager": ms_eager, "peak_memory_bytes": int(peak_mem),
        "forward_work": fwd_work, "stage_ms": stages, "roofline": roofline, "roofline_fwd": roofline_fwd, "roofline_hbm": roofline_hbm, "clocks": clocks,
    }
    if exchange is not None:
        out["exchange_ms"] = exchange
    if parity is not None:
        out["nccl_parity"] = parity
    if c5 is not None:
        out["c5_sequence"] = c5
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args, g_raw, cams, host_targets)
    if world == 1 and not args.no_reference_cuda:
        out["reference_cuda"] = reference_cuda(args, act, my_cams, dev_targets, bg, dev)
    # free the headline step's buffers before the extra legs
    del graph, feed, arena
    ex.close()
    torch.cuda.empty_cache()
    if world == 1 and not args.no_reference_cuda and extras:
        for key, fn in (("c4_forward_throughput", lambda: c4_forward(args)),
                        ("densify_loop_shape", lambda: densify_loop_leg(args)),
                        ("densify_event", lambda: densify_event_leg(args)),
                        ("rigidity_loss", lambda: rigidity_shaped(args)),
                        ("deformation_forward", lambda: deformation_shaped(args)),
                        ("train_step_shape", lambda: train_step_shaped(args))):
            try:
                out[key] = fn()
            except Exception as ex_:  # never lose the headline line to an extra measurement
                out[key] = {"error": repr(ex_)[:300]}
            torch.cuda.empty_cache()
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def c5_sequence(args, world, rank, dev, P0=150_000, P_gt=600_000, steps_per_event=6, timesteps=20, steps_per_timestep=3,
                views_per_step=5, cap=2_600_000, init_noise=0.01, steps_after_reset=None, gt_scale=None, color_noise=0.0,
                events=None):
    """BASELINE.json configs[4]: a synthetic 150-timestep sequence; first-frame optimisation with the reference's
    densify schedule (external.py:211-314: events at i = 500, 600, ..., 5000, opacity reset at 3000) growing the
    cloud from 150 k Gaussians, camera-sharded over the ranks with ONE exchange per step (gsb200.first_frame:
    FirstFrameEngine / run_first_frame), then train.py-shaped timesteps (train.py:728-776: DeformationNetwork ->
    5 sampled views -> rigidity -> Adam) on the densified cloud with the views sharded over the ranks.
    Bounded: every densify event of the schedule is visited, with `steps_per_event` optimiser steps in front of each
    (the reference runs 100), and `timesteps` of the 150 timesteps are trained with `steps_per_timestep` steps each.
    Collective: every rank calls it."""
    import math
    import random
    import time
    import torch.distributed as dist
    from gsb200 import scene, raster, dist as gd, first_frame, optim as goptim, deform, rigidity as R, timestep_train as tt
    T_total = 150
    W, H, n_cams = args.width, args.height, args.cams
    torch.cuda.reset_peak_memory_stats(dev)
    t_begin = time.perf_counter()
    # ---- the sequence: ground-truth Gaussians that move rigidly (+ a little noise); images are their renders ------
    gt = scene.make_gaussians(P_gt, seed=11, scale_override=gt_scale)
    gt["opacity_logits"] = gt["opacity_logits"].abs() + 0.5
    steps_after_reset = steps_after_reset or 2 * steps_per_event
    cams = scene.make_cameras(n_cams, W, H, seed=11)
    mine = gd.shard_cameras(n_cams, rank, world)
    sc = scene.stack_cameras([cams[i] for i in mine], dev)
    bg = torch.zeros(3, device=dev)
    gt_act = {k: v.to(dev).contiguous() for k, v in scene.activate(gt).items()}
    gt_seg = gt["segmentation_masks"].to(dev).contiguous()
    rng = np.random.default_rng(5)
    noise_dir = torch.from_numpy(rng.normal(0, 1, size=(P_gt, 3)).astype(np.float32)).to(dev)

    def gt_means(t):
        ang = 0.15 * math.sin(2 * math.pi * t / T_total)
        c, s_ = math.cos(ang), math.sin(ang)
        rot = torch.tensor([[c, -s_, 0.0], [s_, c, 0.0], [0.0, 0.0, 1.0]], device=dev)
        return (gt_act["means3D"] @ rot.T + 0.002 * math.sin(2 * math.pi * t / 37.0) * noise_dir).contiguous()

    def render_gt(t, sub_sc, n, colors):
        a = dict(gt_act, means3D=gt_means(t))
        col, _, _, _ = raster.forward(a["means3D"], colors, a["opacities"], a["scales"], a["rotations"], None, bg, 1.0,
                                      sub_sc["viewmatrix"], sub_sc["projmatrix"], sub_sc["tanfovx"], sub_sc["tanfovy"], H, W,
                                      n_cams=n)
        return (col if n > 1 else col[None]).clone()

    images = render_gt(0, sc, len(mine), gt_act["colors_precomp"]).clamp(0, 1)
    masks = (render_gt(0, sc, len(mine), gt_seg) > 0.5).float()
    # ---- initial cloud: a noisy subsample of the ground truth, initialised like densify.py:22-57 --------------------
    sel = torch.from_numpy(rng.choice(P_gt, size=P0, replace=False))
    means0 = gt["means"][sel] + torch.from_numpy(rng.normal(0, init_noise, size=(P0, 3)).astype(np.float32))
    from scipy.spatial import cKDTree
    d, _ = cKDTree(means0.numpy()).query(means0.numpy(), k=4, workers=-1)
    ls = np.log(np.sqrt(np.clip((d[:, 1:] ** 2).mean(-1), 1e-7, None)))[:, None].astype(np.float32)
    quat = torch.zeros(P0, 4); quat[:, 0] = 1.0
    col0 = gt["colors"][sel].clone()
    if color_noise > 0:
        col0 = (col0 + torch.from_numpy(rng.normal(0, color_noise, size=(P0, 3)).astype(np.float32))).clamp(0, 1)
    raw = {"means": means0, "colors": col0, "segmentation_masks": gt["segmentation_masks"][sel].clone(),
           "rotation_quaternions": quat, "opacity_logits": torch.zeros(P0, 1), "log_scales": torch.from_numpy(np.tile(ls, (1, 3)))}
    params = {k: torch.nn.Parameter(v.to(dev).contiguous().requires_grad_(True)) for k, v in raw.items()}
    cam_centers = np.stack([np.linalg.inv(c["viewmatrix"][0].numpy().T)[:3, 3] for c in cams])
    scene_radius = 1.1 * float(np.max(np.linalg.norm(cam_centers - cam_centers.mean(0), axis=1)))   # densify.py:60-65
    opt = goptim.create_optimizer(params, scene_radius)
    setup_s = time.perf_counter() - t_begin

    # ---- phase 1: first-frame optimisation through every densify event of the schedule -------------------------
    idx = []
    for e in (events if events is not None else range(500, 5001, 100)):   # after the opacity reset at 3000 the opacities need steps to come back before the
        n_e = steps_per_event if e <= 3000 else steps_after_reset   # final prune at 5000 (threshold 0.25): twice as many there
        idx += [e - n_e + 1 + j for j in range(n_e)]
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    # the schedule has no cap of its own: the run stops densifying once the cloud has outgrown `cap`
    _, timeline = first_frame.run_first_frame(params, opt, scene_radius, sc, images, masks, len(idx), dev,
                                              iteration_index=lambda k: idx[k], max_gaussians=cap)
    torch.cuda.synchronize(dev)
    counts = [(0, P0)] + [(e["i"], e["after"]) for e in timeline]
    stop = {"at": timeline[-1]["i"] if (timeline and timeline[-1]["after"] > cap) else None}
    phase1_s = time.perf_counter() - t0
    P1 = params["means"].shape[0]
    events = len(counts) - 1
    steps1 = sum(1 for i in idx if stop["at"] is None or i <= stop["at"])
    # a clean timing of the optimiser step at the final size (graph replays, no events in between)
    eng = first_frame.FirstFrameEngine(params, opt, len(mine), W, H, dev)
    eng.set_views(sc, images, masks)
    eng.rebuild()
    for _ in range(3):
        eng.step(check_every=0)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        eng.step(check_every=0)
    e1.record()
    torch.cuda.synchronize(dev)
    tms = torch.tensor([e0.elapsed_time(e1) / 10], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    final_step_ms = float(tms.item())
    eng.check()
    eng.adam.sync_state()
    # replicas must be bit-identical after all the events
    identical = True
    if world > 1:
        for k in ("means", "opacity_logits", "log_scales"):
            ref = params[k].data.clone()
            dist.broadcast(ref, 0)
            identical = identical and bool(torch.equal(ref, params[k].data))
        flag = torch.tensor([1.0 if identical else 0.0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        identical = bool(flag.item() == 1.0)
    eng.close()
    del eng
    torch.cuda.empty_cache()

    # ---- phase 2: train.py-shaped timesteps on the densified cloud, views sharded over the ranks ---------------------
    init_p = {k: v.data.detach().contiguous() for k, v in params.items()}
    t1 = time.perf_counter()
    ninfo = R.create_neighbor_info(init_p, k=20)
    knn_s = time.perf_counter() - t1
    enc0 = deform.normalize_and_encode_means_and_rotations(init_p)
    fg_prev = R.create_foreground_info(init_p, ninfo.indices)
    enc_prev = enc0
    torch.manual_seed(0)
    net = deform.DeformationNetwork(256, 4, precise=True).to(dev)
    if world > 1 and os.environ.get("GSB_BENCH_NET_SHARD", "1") != "0":
        net.set_row_shard()     # every rank runs the network on its slice of the Gaussians (gsb200.deform row sharding)
    nopt = goptim.FusedAdam(net.parameters(), lr=1e-4)   # torch.optim.Adam's update, every table in two launches
    ex = gd.Exchange(P1, dev)
    # host-free rendering (no read-back of the instance count inside the step) where a rank renders several of the
    # step's views; with the views spread over several ranks a rank holds ONE camera that changes from timestep to
    # timestep, and cameras differ by more than the arena's slack in instances (an 8-GPU run overflowed an arena sized
    # from the first camera: 898 k instances against a capacity of 587 k), so those runs keep the blocking forward
    varena = tt.ViewArena() if world == 1 else None
    all_sc = scene.stack_cameras(cams, dev)
    prng = random.Random(0)
    ts_list = [1 + (k * (T_total - 2)) // max(timesteps - 1, 1) for k in range(timesteps)]
    train_ms = []
    losses = []
    for t in ts_list:
        picks = sorted(prng.sample(range(n_cams), views_per_step))
        my_picks = picks[rank::world]
        if my_picks:
            sub = {"viewmatrix": all_sc["viewmatrix"][my_picks].contiguous(), "projmatrix": all_sc["projmatrix"][my_picks].contiguous(),
                   "tanfovx": [all_sc["tanfovx"][i] for i in my_picks], "tanfovy": [all_sc["tanfovy"][i] for i in my_picks],
                   "image_width": W, "image_height": H}
            tgt = render_gt(t, sub, len(my_picks), gt_act["colors_precomp"]).clamp(0, 1)
        else:
            sub, tgt = None, None
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        e0.record()
        for _ in range(steps_per_timestep):
            loss, enc_t, fg_t = tt.train_step(net, nopt, init_p, enc0, enc_prev, fg_prev, ninfo, (sub, tgt), t, T_total,
                                              exchange=ex, view_count=views_per_step, view_arena=varena)
        e1.record()
        torch.cuda.synchronize(dev)
        enc_prev, fg_prev = enc_t, fg_t
        tm = torch.tensor([e0.elapsed_time(e1) / steps_per_timestep], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        train_ms.append(float(tm.item()))
        losses.append(float(loss))
    if varena is not None:
        varena.poll(wait=True)   # every step's status words: raises if a step's instances had not fit
    ex.close()
    peak = torch.tensor([float(torch.cuda.max_memory_allocated(dev))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(peak, op=dist.ReduceOp.MAX)
    ev_ms = [e["event_ms"] for e in timeline] if timeline else []
    rb_ms = [e["rebuild_ms"] for e in timeline] if timeline else []
    return {
        "what": f"synthetic {T_total}-timestep sequence, {n_cams} cameras {W}x{H}: first-frame optimisation from {P0} Gaussians "
                f"through the densify schedule of external.py:211-314 ({steps_per_event} optimiser steps before each of the "
                f"events at i = 500..3000, {steps_after_reset} after the opacity reset; the reference runs 100), then {timesteps} train.py-shaped timesteps x "
                f"{steps_per_timestep} steps ({views_per_step} sampled views, DeformationNetwork 256 x 4, rigidity K = 20) "
                f"(BASELINE.json configs[4])",
        "n_gpus": world, "gaussians_start": P0, "gaussians_end": P1, "densify_events": events,
        "stopped_at_event": stop["at"], "gaussian_count_over_events": counts,
        "first_frame": {"optimiser_steps": steps1, "seconds_wall": phase1_s,
                        "iters_per_s_wall": steps1 * n_cams / phase1_s,
                        "ms_per_step_at_final_size": final_step_ms, "iters_per_s_at_final_size": n_cams / (final_step_ms * 1e-3),
                        "ms_per_densify_event_mean": float(np.mean(ev_ms)) if ev_ms else None,
                        "ms_per_densify_event_max": float(np.max(ev_ms)) if ev_ms else None,
                        "ms_per_rebuild_mean": float(np.mean(rb_ms)) if rb_ms else None,
                        "replicas_bit_identical": identical},
        "timesteps": {"count": timesteps, "ms_per_step_mean": float(np.mean(train_ms)), "ms_per_step_first": train_ms[0],
                      "steps_per_s": 1e3 / float(np.mean(train_ms)), "loss_first": losses[0], "loss_last": losses[-1],
                      "knn_seconds": knn_s, "network": ("row-sharded over the ranks (BatchNorm column sums, the [N, 7] output and the dense-layer gradients exchanged)"
                                  if net.shard is not None else "one GPU") + "; precise mode (3xTF32 products); views sharded"},
        "peak_memory_bytes_per_gpu": int(peak.item()), "setup_seconds": setup_s,
    }


def densify_loop_leg(args):
    return {"what": "densify.py:201-248 shape: 1 camera/iteration, colour + segmentation render, 2x L1/SSIM, "
                    "backward, densify statistics, Adam",
            "unit": "iters/s", "ours": densify_shaped(args, "ours"), "ours_graph": densify_shaped(args, "ours_graph"),
            "reference_cuda": densify_shaped(args, "reference_cuda")}


def deformation_shaped(args, hidden=256, blocks=4, iters=10):
    """DeformationNetwork forward and forward + backward (train.py:80-110; 192 -> H, `blocks` residual blocks, H -> 7,
    training-mode BatchNorm) over all Gaussians.  "ours_tf32" / "ours_3xtf32": gsb200.deform on the hand-written
    tcgen05 GEMM; "torch_fp32" (the reference's setting: torch's default keeps FP32 matmuls) and "torch_tf32"
    (allow_tf32): the same architecture in stock torch ops.  Also the 192 -> H layer alone with its HBM roofline."""
    from gsb200 import deform
    from gsb200.linear import linear
    from torch import nn
    dev = torch.device("cuda", torch.cuda.current_device())
    n = args.gaussians

    class Block(nn.Module):
        def __init__(self, d):
            super().__init__()
            self.fc1 = nn.Linear(d, d, bias=False); self.bn1 = nn.BatchNorm1d(d)
            self.fc2 = nn.Linear(d, d, bias=False); self.bn2 = nn.BatchNorm1d(d)

        def forward(self, x):
            y = nn.functional.gelu(self.bn1(self.fc1(x)))
            return nn.functional.gelu(self.bn2(self.fc2(y)) + x)

    class Stock(nn.Module):
        def __init__(self):
            super().__init__()
            self.fc_in = nn.Linear(192, hidden)
            self.residual_blocks = nn.Sequential(*(Block(hidden) for _ in range(blocks)))
            self.fc_out = nn.Linear(hidden, 7)

        def forward(self, init, x):
            return self.fc_out(self.residual_blocks(self.fc_in(x))) + init

    torch.manual_seed(0)
    stock = Stock().to(dev).train()
    x = torch.randn(n, 192, device=dev); init = torch.randn(n, 7, device=dev)

    def timed(fn):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(iters):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    res = {"what": f"DeformationNetwork forward, {n} Gaussians, hidden {hidden}, {blocks} residual blocks", "unit": "ms"}
    for precise in (False, True):
        net = deform.DeformationNetwork(hidden, blocks, precise=precise).to(dev)
        net.load_state_dict(stock.state_dict())
        res["ours_3xtf32" if precise else "ours_tf32"] = timed(lambda: net(init, network_input=x))
    with torch.no_grad():
        for name, flag in (("torch_fp32", False), ("torch_tf32", True)):
            torch.backends.cuda.matmul.allow_tf32 = flag
            res[name] = timed(lambda: stock(init, x))
        torch.backends.cuda.matmul.allow_tf32 = False
    # forward + backward (one training step of the network without the optimiser), gradient of sum(out * w)
    wgt = torch.randn(n, 7, device=dev)

    def train(net_call, net):
        net.zero_grad(set_to_none=True)
        (net_call() * wgt).sum().backward()

    for precise in (False, True):
        net = deform.DeformationNetwork(hidden, blocks, precise=precise).to(dev)
        net.load_state_dict(stock.state_dict())
        res["fwd_bwd_ours_precise" if precise else "fwd_bwd_ours_tf32"] = timed(
            lambda: train(lambda: net(init, network_input=x), net))
    for name, flag in (("fwd_bwd_torch_fp32", False), ("fwd_bwd_torch_tf32", True)):
        torch.backends.cuda.matmul.allow_tf32 = flag
        res[name] = timed(lambda: train(lambda: stock(init, x), stock))
    torch.backends.cuda.matmul.allow_tf32 = False
    w, b = stock.fc_in.weight.detach(), stock.fc_in.bias.detach()
    y = torch.empty(n, hidden, device=dev)
    ms = timed(lambda: linear(x, w, b, out=y))
    peak = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mp = json.load(f)
        peak = float(mp["hbm_gbs"])
    except Exception:
        pass
    byts = n * (192 + hidden) * 4
    res["fc_in_layer"] = {"kernel": "linear_tf32_kernel (TMA + tcgen05.mma kind::tf32, TMEM accumulators)", "ms": ms,
                          "bound": "hbm", "algorithmic_bytes": byts, "achieved": byts / (ms * 1e-3) / 1e9,
                          "peak": peak, "unit": "GB/s", "frac": (byts / (ms * 1e-3) / 1e9 / peak) if peak else None,
                          "tflops": 2.0 * n * 192 * hidden / (ms * 1e-3) / 1e12}
    # weight gradient of a hidden layer, dW[H, H] = G^T A over the n rows: TF32 and 3xTF32 on the tcgen05 kernel against
    # torch (cuBLAS) at the same precisions; and the 7-wide output layer's backward against its two torch products
    from gsb200.linear import linear_wgrad, thin_backward
    gmat, amat = torch.randn(n, hidden, device=dev), torch.randn(n, hidden, device=dev)
    wg = {"what": f"dW[{hidden},{hidden}] = G^T A over {n} rows", "unit": "ms",
          "ours_tf32": timed(lambda: linear_wgrad(gmat, amat)),
          "ours_3xtf32": timed(lambda: linear_wgrad(gmat, amat, precise=True)),
          "torch_fp32": timed(lambda: gmat.t() @ amat)}
    torch.backends.cuda.matmul.allow_tf32 = True
    wg["torch_tf32"] = timed(lambda: gmat.t() @ amat)
    torch.backends.cuda.matmul.allow_tf32 = False
    wg["algorithmic_bytes"] = 2 * n * hidden * 4
    wg["ours_3xtf32_gbs"] = wg["algorithmic_bytes"] / (wg["ours_3xtf32"] * 1e-3) / 1e9
    res["hidden_layer_wgrad"] = wg
    g7, w7 = torch.randn(n, 7, device=dev), torch.randn(7, hidden, device=dev)
    res["output_layer_backward"] = {"what": f"dX = G W, dW = G^T A, db = sum G for the 7-wide output layer, {n} rows", "unit": "ms",
                                    "ours": timed(lambda: thin_backward(g7, w7, amat)),
                                    "ours_dx_only": timed(lambda: thin_backward(g7, w7, amat, parts="dx")),
                                    "ours_wgrad_only": timed(lambda: thin_backward(g7, w7, amat, parts="dw")),
                                    "torch_fp32": timed(lambda: (g7 @ w7, g7.t() @ amat, g7.sum(0)))}
    return res


def train_step_shaped(args, hidden=256, blocks=4, views=5, neighbours=20, iters=8):
    """SURVEY.md 8(d) metric (ii), train.py's shape (train.py:745-790): one optimiser step of the DeformationNetwork =
    network forward over all Gaussians -> 5 sampled cameras rendered, 0.8 L1 + 0.2 (1 - SSIM) each -> rigidity term ->
    backward -> Adam.  "ours": gsb200.timestep_train.train_step (tcgen05 network, camera-batched rasterizer, fused loss,
    fused rigidity once per step).  "stock": the same step composed the reference's way -- unmodified reference
    extension camera by camera, conv2d SSIM, the rigidity formula in torch ops once per view, nn.Linear/BatchNorm1d
    network with FP32 matmuls, torch Adam."""
    import random
    import helpers
    import oracle_lib
    from torch import nn
    from gsb200 import scene, deform, optim as goptim, rigidity as R, timestep_train as tt
    dev = torch.device("cuda", torch.cuda.current_device())
    g_raw, tgt_raw, cams = build_workload(args)
    sc = scene.stack_cameras(cams, dev)
    bg = torch.zeros(3, device=dev)
    targets = render_targets(tgt_raw, sc, args.cams, bg, dev)
    init = {k: v.to(dev).contiguous() for k, v in g_raw.items()}
    ninfo = R.create_neighbor_info(init, k=neighbours)
    enc0 = deform.normalize_and_encode_means_and_rotations(init)
    fg0 = R.create_foreground_info(init, ninfo.indices)
    rng = random.Random(0)
    picks = [sorted(rng.sample(range(args.cams), views)) for _ in range(iters + 3)]

    def sub(ix):
        return ({"viewmatrix": sc["viewmatrix"][ix].contiguous(), "projmatrix": sc["projmatrix"][ix].contiguous(),
                 "tanfovx": [sc["tanfovx"][i] for i in ix], "tanfovy": [sc["tanfovy"][i] for i in ix],
                 "image_width": sc["image_width"], "image_height": sc["image_height"]}, targets[ix].contiguous())

    def timed(fn):
        for k in range(3):
            fn(k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for k in range(iters):
            fn(3 + k)
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    torch.manual_seed(0)
    net = deform.DeformationNetwork(hidden, blocks, precise=True).to(dev)
    opt = goptim.FusedAdam(net.parameters(), lr=1e-4)   # torch.optim.Adam's update, every table in two launches
    last = {}
    varena = tt.ViewArena()   # host-free rendering (no read-back of the instance count inside the step)

    def ours(k):
        loss, _, _ = tt.train_step(net, opt, init, enc0, enc0, fg0, ninfo, sub(picks[k]), 7, 150, view_arena=varena)
        last["ours"] = loss

    res = {"what": f"one train.py step: DeformationNetwork ({hidden} x {blocks} blocks) over {args.gaussians} Gaussians, "
                   f"{views} of {args.cams} cameras {args.width}x{args.height}, rigidity K={neighbours}, Adam",
           "unit": "ms per step", "ours": timed(ours)}
    res["ours_steps_per_s"] = 1e3 / res["ours"]
    res["network_matmuls"] = ("precise (default): 3xTF32 forward, input and weight gradients (FP32-level accuracy) -- compare with "
                              "`stock` (torch FP32); ours_tf32: TF32 everywhere (opt-in fast mode)")
    torch.manual_seed(0)
    net = deform.DeformationNetwork(hidden, blocks, precise=False).to(dev)
    opt = goptim.FusedAdam(net.parameters(), lr=1e-4)
    res["ours_tf32"] = timed(ours)
    varena.poll(wait=True)
    ref = oracle_lib.load_reference_ext()
    if ref is None:
        return res

    class Block(nn.Module):
        def __init__(self, d):
            super().__init__()
            self.fc1 = nn.Linear(d, d, bias=False); self.bn1 = nn.BatchNorm1d(d)
            self.fc2 = nn.Linear(d, d, bias=False); self.bn2 = nn.BatchNorm1d(d)

        def forward(self, x):
            y = nn.functional.gelu(self.bn1(self.fc1(x)))
            return nn.functional.gelu(self.bn2(self.fc2(y)) + x)

    class Stock(nn.Module):
        def __init__(self):
            super().__init__()
            self.fc_in = nn.Linear(192, hidden)
            self.residual_blocks = nn.Sequential(*(Block(hidden) for _ in range(blocks)))
            self.fc_out = nn.Linear(hidden, 7)

        def forward(self, i7, x):
            return self.fc_out(self.residual_blocks(self.fc_in(x))) + i7

    torch.backends.cuda.matmul.allow_tf32 = False
    snet = Stock().to(dev).train()
    sopt = torch.optim.Adam(snet.parameters(), lr=1e-4)
    fgm = init["segmentation_masks"][:, 0] > 0.5
    prev_off = fg0.offsets_to_neighbors.contiguous()
    i7 = torch.cat((init["means"], init["rotation_quaternions"]), dim=1)
    xin = torch.cat((enc0, enc0, deform.positional_encoding(torch.full((args.gaussians, 1), 7 / 150, device=dev), 4)), dim=1)

    def stock_rigidity(means, quats):
        m = means[fgm]
        q = torch.nn.functional.normalize(quats)[fgm]
        a, b = q.unbind(-1), fg0.inverted_rotations.unbind(-1)
        rel = torch.stack([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
                           a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                           a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
                           a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]], -1)
        r, x, y, z = (rel / rel.norm(dim=-1, keepdim=True)).unbind(-1)
        rot = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                           2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                           2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
        off = m[ninfo.indices] - m[:, None]
        loc = (rot.transpose(2, 1)[:, None] @ off[..., None]).squeeze(-1)
        return torch.sqrt(((loc - prev_off) ** 2).sum(-1) * ninfo.weights + 1e-20).mean()

    def stock(k):
        delta = snet(i7, xin)
        means = init["means"] + delta[:, :3] * 0.01
        quats = init["rotation_quaternions"] + delta[:, 3:] * 0.01
        total = 0
        for c in picks[k]:
            a = {"means3D": means, "colors_precomp": init["colors"], "rotations": torch.nn.functional.normalize(quats),
                 "opacities": torch.sigmoid(init["opacity_logits"]), "scales": torch.exp(init["log_scales"]),
                 "means2D": torch.zeros_like(means, requires_grad=True) + 0}
            im, _, _ = helpers.RefRasterize.apply(ref, a["means3D"], a["means2D"], a["colors_precomp"], a["opacities"],
                                                  a["scales"], a["rotations"], cams[c], bg)
            total = total + torch_ssim_l1(im, targets[c]) + 3 * stock_rigidity(means, quats)
        total.backward()
        sopt.step()
        sopt.zero_grad()
        last["stock"] = total.detach()

    res["stock"] = timed(stock)
    res["stock_steps_per_s"] = 1e3 / res["stock"]
    return res


def rigidity_shaped(args, views=5, iters=20):
    """the rigidity term of one train.py step (train.py:317-351) -- loss + gradient w.r.t.
    means and rotation_quaternions for ~60% foreground Gaussians x 20 neighbours.  "ours": gsb200.rigidity, one
    fused launch per step.  "stock_torch": the same formula in plain torch ops with autograd, evaluated once per
    sampled view like the reference's loop (train.py:373-377 inside :402-412; 5 views, train.py:753)."""
    from gsb200 import rigidity as R
    dev = torch.device("cuda", torch.cuda.current_device())
    gen = torch.Generator().manual_seed(11)
    P, K = args.gaussians, 20
    seg = torch.zeros(P, 3); seg[:, 0] = (torch.rand(P, generator=gen) > 0.4).float(); seg[:, 2] = 1 - seg[:, 0]
    F = int(seg[:, 0].sum())
    means = (torch.rand(P, 3, generator=gen) - 0.5) * 3
    walk = torch.cumsum(0.02 * torch.randn(F, 3, generator=gen), 0)
    means[seg[:, 0] > 0.5] = walk - walk.mean(0)
    steps_ = torch.tensor([d for d in range(-10, 11) if d != 0])
    nbr = torch.arange(F)[:, None] + steps_[None]
    nbr = torch.where((nbr < 0) | (nbr >= F), torch.arange(F)[:, None] - steps_[None], nbr)
    base = {"means": means.to(dev), "rotation_quaternions": torch.randn(P, 4, generator=gen).to(dev),
            "segmentation_masks": seg.to(dev)}
    ninfo = R.NeighborInfo(weights=torch.rand(F, K, generator=gen).to(dev), indices=nbr.to(dev))
    prev = R.create_foreground_info(base, ninfo.indices)
    cur = {"means": (base["means"] + 0.01 * torch.randn(P, 3, device=dev)).requires_grad_(True),
           "rotation_quaternions": base["rotation_quaternions"].clone().requires_grad_(True),
           "segmentation_masks": base["segmentation_masks"]}
    prev_off = prev.offsets_to_neighbors.contiguous()
    fg = base["segmentation_masks"][:, 0] > 0.5

    def ours():
        (views * R.calculate_rigidity_loss(cur, ninfo, prev)).backward()

    def stock_one():
        m = cur["means"][fg]
        q = torch.nn.functional.normalize(cur["rotation_quaternions"])[fg]
        a, b = q.unbind(-1), prev.inverted_rotations.unbind(-1)
        rel = torch.stack([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
                           a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                           a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
                           a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]], -1)
        r, x, y, z = (rel / rel.norm(dim=-1, keepdim=True)).unbind(-1)
        rot = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                           2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                           2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
        off = m[ninfo.indices] - m[:, None]
        loc = (rot.transpose(2, 1)[:, None] @ off[..., None]).squeeze(-1)
        return torch.sqrt(((loc - prev_off) ** 2).sum(-1) * ninfo.weights + 1e-20).mean()

    def stock():
        torch.stack([stock_one() for _ in range(views)]).sum().backward()

    res = {"what": f"rigidity term of one train.py step: F={F} foreground Gaussians x K={K}, loss + gradients, "
                   f"{views} views per step", "unit": "ms per step"}
    for name, fn in (("ours", ours), ("stock_torch", stock)):
        for _ in range(3):
            fn()
        cur["means"].grad = None; cur["rotation_quaternions"].grad = None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(iters):
            fn()
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / iters
        res[name + "_loss"] = float((R.calculate_rigidity_loss(cur, ninfo, prev) if name == "ours" else stock_one()).detach())
    return res


# ------------------------------------------------------------------------------------------------
def cpu_step(o, act_np, cam, target_np, H, W):
    """One camera of the workload on the CPU oracle: forward, L1+SSIM (+ gradient), backward."""
    r = o.forward(act_np["means3D"], act_np["colors_precomp"], act_np["opacities"], act_np["scales"],
                  act_np["rotations"], np.zeros(3, np.float32), cam["viewmatrix"].numpy(), cam["projmatrix"].numpy(),
                  cam["tanfovx"], cam["tanfovy"], H, W)
    _, dL = o.ssim_l1(r["color"], target_np, 0.8, 0.2, grad=True)
    o.backward(r, dL)
    return r


def cpu_baseline(args, g_raw, cams, host_targets, budget_s=12.0):
    """The C oracle (OpenMP, all host threads) on a bounded sample of the workload: whole cameras of the timestep,
    one after the other (each with its own target image), until about `budget_s` seconds of CPU work are done."""
    import oracle_lib
    from gsb200 import scene
    o = oracle_lib.load()
    threads = o.set_threads(_host_threads())   # explicit: launchers export OMP_NUM_THREADS=1
    act_np = {k: v.numpy() for k, v in scene.activate(g_raw).items()}
    n_avail = min(len(cams), host_targets.size(0))
    done = 0
    t0 = time.perf_counter()
    while done < n_avail and (done == 0 or time.perf_counter() - t0 < budget_s):
        cpu_step(o, act_np, cams[done], host_targets[done].numpy(), args.height, args.width)
        done += 1
    dt = time.perf_counter() - t0
    return {"value": done / dt, "unit": "iters/s", "cores": threads, "kind": "port",
            "sample": f"cameras 0..{done - 1} of the workload ({args.gaussians} gaussians, {args.width}x{args.height}): "
                      f"fwd + L1/SSIM + bwd on the C oracle (OpenMP, {threads} threads), {done} iterations in {dt:.1f} s"}


def run_reference(args):
    """--impl reference: the CPU arm.  The reference ships no CPU implementation of this path, so the
    C restatement under oracle/ (OpenMP, all host threads) is what is timed; each step is a bounded
    sample of the workload: ONE camera (fwd + L1/SSIM + bwd)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle_lib
    from gsb200 import scene
    o = oracle_lib.load()
    # torch.distributed.run exports OMP_NUM_THREADS=1 and libgomp read it when torch loaded it: set the team size
    # explicitly to the cores this process may use, and report the number actually in effect
    threads = o.set_threads(_host_threads())
    g_raw, tgt_raw, cams = build_workload(args)
    act_np = {k: v.numpy() for k, v in scene.activate(g_raw).items()}
    tgt_np = {k: v.numpy() for k, v in scene.activate(tgt_raw).items()}
    H, W = args.height, args.width
    tr = o.forward(tgt_np["means3D"], tgt_np["colors_precomp"], tgt_np["opacities"], tgt_np["scales"], tgt_np["rotations"],
                   np.zeros(3, np.float32), cams[0]["viewmatrix"].numpy(), cams[0]["projmatrix"].numpy(),
                   cams[0]["tanfovx"], cams[0]["tanfovy"], H, W)
    target = tr["color"]
    for _ in range(min(args.warmup, 1)):
        cpu_step(o, act_np, cams[0], target, H, W)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_step(o, act_np, cams[0], target, H, W)
    dt = (time.perf_counter() - t0) / args.steps
    v = 1.0 / dt
    sample = (f"camera 0 of the workload ({args.gaussians} gaussians, {W}x{H}) per step: fwd + L1/SSIM + bwd on the "
              f"C oracle (OpenMP, {threads} threads); the reference has no CPU path for the rasterizer")
    print(json.dumps({
        "impl": "reference", "metric": "train iters/sec (one iter = one camera fwd + L1/SSIM + bwd; 27 cameras per step)",
        "value": v, "unit": "iters/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 1),
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"Panoptic-shaped timestep: {args.gaussians} gaussians, {args.cams} cameras "
                               f"{W}x{H}, fwd + L1/SSIM + bwd + densify stats (BASELINE.json configs[1])",
                   "sample": sample},
        "cpu_baseline": {"value": v, "unit": "iters/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------
def torch_ssim_l1(img, target):
    """Restatement of external.calc_ssim + the 0.8/0.2 mix (external.py:49-110, densify.py:127-129) in
    stock PyTorch ops -- what the reference runs next to its CUDA rasterizer."""
    import torch.nn.functional as F
    g = torch.tensor([np.exp(-((x - 5) ** 2) / (2 * 1.5 ** 2)) for x in range(11)], dtype=torch.float32)
    g = (g / g.sum()).unsqueeze(1)
    w = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0).expand(3, 1, 11, 11).contiguous().to(img.device)
    mu1 = F.conv2d(img, w, padding=5, groups=3); mu2 = F.conv2d(target, w, padding=5, groups=3)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(img * img, w, padding=5, groups=3) - mu1_sq
    s2 = F.conv2d(target * target, w, padding=5, groups=3) - mu2_sq
    s12 = F.conv2d(img * target, w, padding=5, groups=3) - mu12
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu12 + c1) * (2 * s12 + c2)) / ((mu1_sq + mu2_sq + c1) * (s1 + s2 + c2))).mean()
    return 0.8 * (img - target).abs().mean() + 0.2 * (1.0 - ssim)


def reference_cuda(args, act, cams, targets, bg, dev, passes=10):
    """The unmodified reference CUDA extension (oracle/_ref) on the same GPU, same inputs, its own
    per-camera call pattern: forward -> torch L1/SSIM -> autograd -> backward, camera by camera.  Timed like the
    headline: warm-up, >= 10 passes over the cameras, CUDA events, its own clock / throttle sample."""
    import oracle_lib
    import helpers
    ref = oracle_lib.load_reference_ext()
    if ref is None:
        return {"unavailable": "oracle/_ref/refdgr/_C.so not built"}

    def one_pass():
        for i, cam in enumerate(cams):
            fw = helpers.ref_forward(ref, act, cam, bg, dev)
            img = fw["color"].detach().requires_grad_(True)
            loss = torch_ssim_l1(img, targets[i])
            loss.backward()
            helpers.ref_backward(ref, act, cam, bg, dev, fw, img.grad)

    if os.environ.get("GSB_REF_PROFILE") == "1":   # tools/prof_r2.sh: one warm pass + one pass under ncu
        one_pass(); torch.cuda.synchronize()
        one_pass(); torch.cuda.synchronize()
        return {"profile_only": True}
    sampler = ClockSampler(dev.index or 0)
    sampler.start()
    for _ in range(2):
        one_pass()
    torch.cuda.synchronize()
    passes = max(passes, 10)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.begin()
    e0.record()
    for _ in range(passes):
        one_pass()
    e1.record()
    torch.cuda.synchronize()
    sampler.end()
    ms = e0.elapsed_time(e1) / passes
    return {"value": len(cams) / (ms * 1e-3), "unit": "iters/s", "ms_per_step": ms, "passes": passes, "warmup_passes": 2,
            "clocks": sampler.stop(),
            "what": "unmodified reference extension (sm_100a build, CUB sort) + stock torch L1/SSIM, camera by camera; "
                    "per-kernel breakdown: profiles/r2_reference_stages.md"}


def c4_forward(args, gaussians=1_000_000, n_cams=8, W=1920, H=1080, iters=10):
    """BASELINE.json configs[3]: 1 M Gaussians, 1920x1080, 8 cameras, forward-only render throughput.  8160 tiles per
    camera > 1024, so the instance sort takes the path only this config exercises: two 8-bit passes on the 16-bit
    camera-local tile ids + tile_ranges_seg_kernel.  "ours": all 8 cameras in one host-free batched call replayed
    from a CUDA graph; "reference_cuda": the unmodified extension camera by camera on the same B200."""
    import ctypes as C
    import helpers
    import oracle_lib
    from gsb200 import _lib, raster, scene, step as step_mod
    dev = torch.device("cuda", torch.cuda.current_device())
    g = scene.make_gaussians(gaussians, seed=1)
    cams = scene.make_cameras(n_cams, W, H, seed=1)
    a = {k: v.to(dev).contiguous() for k, v in scene.activate(g).items()}
    sc = scene.stack_cameras(cams, dev)
    td = raster.tanfov_table(sc["tanfovx"], sc["tanfovy"], n_cams, dev)
    bg = torch.zeros(3, device=dev)

    def fwd(**kw):
        return raster.forward(a["means3D"], a["colors_precomp"], a["opacities"], a["scales"], a["rotations"], None, bg, 1.0,
                              sc["viewmatrix"], sc["projmatrix"], sc["tanfovx"], sc["tanfovy"], H, W, n_cams=n_cams, **kw)

    color_sync, _, _, st = fwd()
    R = st.num_rendered
    tile_bits = int(st.ctx.tile_bits)
    cam0 = color_sync[0].clone()
    del st, color_sync
    arena = raster.RasterArena(dev).reserve(gaussians, n_cams, W, H, capacity=int(1.1 * R) + 4096)
    graph = step_mod.StepGraph(lambda: fwd(arena=arena, tanfov_dev=td))

    def timed(fn, n):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    ms = timed(graph.replay, iters)
    color, _, _, st = graph.out
    if st.status()[0] != 0:
        raise RuntimeError("arena overflow in the C4 leg")
    lib = _lib.lib()
    lib.gsb_profile_enable(1)
    buf = (C.c_float * len(_lib.PROFILE_STAGES))()
    acc = np.zeros(len(_lib.PROFILE_STAGES))
    for _ in range(3):
        fwd(arena=arena, tanfov_dev=td)
        torch.cuda.synchronize()
        lib.gsb_profile_read(buf, len(_lib.PROFILE_STAGES))
        acc += np.array(list(buf))
    lib.gsb_profile_enable(0)
    stages = {k: float(v / 3) for k, v in zip(_lib.PROFILE_STAGES, acc) if not k.endswith("_begin") and v > 0}
    pairs = raster.count_pairs(st).sum(dim=0)
    hbm_peak, sm_max_mhz, _ = peaks()
    digit = 8
    passes = (tile_bits + digit - 1) // digit
    # per pass: histogram reads the key (2 B); scatter reads key + payload (6 B), writes payload (4 B) and the key (2 B;
    # the last pass too: tile_ranges_seg_kernel needs the sorted ids); then the ranges kernel reads the sorted ids (2 B)
    sort_bytes = R * 14 * passes
    out = {"what": f"{gaussians} Gaussians, {n_cams} cameras {W}x{H}, forward only (BASELINE.json configs[3])",
           "ms_per_pass": ms, "renders_per_s": n_cams / (ms * 1e-3), "num_rendered": R, "tile_bits": tile_bits,
           "sort_bits_reference": 32 + int(oracle_lib.load().higher_msb(((W + 15) // 16) * ((H + 15) // 16))),
           "walked_pairs": int(pairs[0]), "contributing_pairs": int(pairs[1]),
           "blends_per_sec_fwd": float(pairs[0]) / (ms * 1e-3), "stage_ms": stages,
           "roofline_hbm": {"kernel": f"(seg_tile_hist + seg_chunk_scan + seg_scatter)<u16,{digit}> x{passes} (two-pass instance sort)",
                            "bound": "hbm", "algorithmic_bytes": sort_bytes, "ms": stages.get("sort"),
                            "achieved": sort_bytes / (stages["sort"] * 1e-3) / 1e9 if stages.get("sort") else None,
                            "peak": hbm_peak, "unit": "GB/s",
                            "frac": sort_bytes / (stages["sort"] * 1e-3) / 1e9 / hbm_peak if stages.get("sort") else None},
           "roofline_fwd": {"kernel": "render_fwd_kernel", "bound": "fp32", "ms": stages.get("render_fwd"),
                            "frac": (float(pairs[0]) * 24.0 / (stages["render_fwd"] * 1e-3) / 1e12) /
                                    (torch.cuda.get_device_properties(dev).multi_processor_count * 256 * sm_max_mhz * 1e-6)
                            if stages.get("render_fwd") else None}}
    ref = oracle_lib.load_reference_ext()
    if ref is not None:
        def ref_pass():
            for cam in cams:
                helpers.ref_forward(ref, a, cam, bg, dev)
        fw0 = helpers.ref_forward(ref, a, cams[0], bg, dev)
        out["max_abs_diff_camera0"] = float((fw0["color"] - cam0).abs().max())
        del fw0
        ms_ref = timed(ref_pass, 5)
        out["reference_cuda"] = {"ms_per_pass": ms_ref, "renders_per_s": n_cams / (ms_ref * 1e-3)}
        out["speedup_vs_reference_cuda"] = ms_ref / ms
    return out


def densify_event_leg(args):
    """One densify event (external.py:211-314: clone / split / remove parents / prune over the six tables and both
    Adam moments) at 300 k, 1 M and 2 M Gaussians: ms per event and the achieved bandwidth against SURVEY 8(d)'s
    2 x 204 B per surviving Gaussian."""
    import time
    from gsb200 import densify as gdens, optim as goptim, scene
    dev = torch.device("cuda", torch.cuda.current_device())
    hbm_peak, _, _ = peaks()
    res = {"what": "densify_gaussians at i = 1000: ~12% cloned, ~8% split, 3% pruned; wall clock around the call with "
                   "device synchronisation (the call reads row counts back)", "unit": "ms per event"}
    for P in (300_000, 1_000_000, 2_000_000):
        gen = torch.Generator().manual_seed(P)
        raw = {"means": torch.randn(P, 3, generator=gen), "colors": torch.rand(P, 3, generator=gen),
               "segmentation_masks": torch.rand(P, 3, generator=gen), "rotation_quaternions": torch.randn(P, 4, generator=gen),
               "opacity_logits": torch.randn(P, 1, generator=gen) * 1.5 + 2.0,
               "log_scales": torch.log(torch.where(torch.rand(P, 1, generator=gen) < 0.4, 0.02, 0.002) * torch.ones(P, 3))}
        times = []
        for rep in range(3):
            params = {k: torch.nn.Parameter(v.to(dev).contiguous().requires_grad_(True)) for k, v in raw.items()}
            opt = goptim.FusedAdam([{"params": [v], "name": k, "lr": 1e-4} for k, v in params.items()], lr=0.0, eps=1e-15)
            for v in params.values():
                v.grad = torch.zeros_like(v)
            opt.step()
            hot = (torch.rand(P, generator=gen) < 0.2).float().to(dev)
            dv = gdens.DensificationVariables(visibility_count=torch.ones(P, device=dev),
                                              mean_2d_gradients_accumulated=hot * 0.001, max_2d_radii=torch.zeros(P, device=dev),
                                              gaussian_is_visible_mask=torch.zeros(P, dtype=torch.bool, device=dev))
            dv.means_2d = type("G", (), {"grad": torch.zeros(P, 3, device=dev)})()
            torch.manual_seed(0)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            with torch.no_grad():
                gdens.densify_gaussians(params, dv, 1.0, opt, 1000)
            torch.cuda.synchronize()
            times.append((time.perf_counter() - t0) * 1e3)
            after = params["means"].shape[0]
            del params, opt, dv
        ms = min(times)
        res[str(P)] = {"ms": ms, "after": after, "algorithmic_bytes": 2 * 204 * after,
                       "achieved_gbs": 2 * 204 * after / (ms * 1e-3) / 1e9, "frac_of_hbm": 2 * 204 * after / (ms * 1e-3) / 1e9 / hbm_peak}
    return res


def densify_shaped(args, impl, iters=30, warm=5):
    """The reference's first-frame loop shape (densify.py:201-248): ONE camera per iteration, two renders
    (colour + segmentation colours), 0.8 L1 + 0.2 (1-SSIM) each, total = image + 3 segmentation, backward,
    densification statistics, Adam.  impl = "ours" (drop-in GaussianRasterizer + fused loss + statistics
    kernels) or "reference_cuda" (unmodified extension through its autograd node + stock torch loss and
    masked-index statistics).  Returns iterations / second (CUDA events)."""
    from gsb200 import scene, loss as gloss, densify as gdens
    import helpers
    dev = torch.device("cuda", torch.cuda.current_device())
    g_raw, tgt_raw, cams = build_workload(args)
    sc = scene.stack_cameras(cams, dev)
    bg = torch.zeros(3, device=dev)
    targets = render_targets(tgt_raw, sc, args.cams, bg, dev)
    seg_targets = (targets > 0.5).float()
    params = {k: torch.nn.Parameter(v.to(dev).contiguous().requires_grad_(True)) for k, v in g_raw.items()}
    groups = [{"params": [v], "name": k, "lr": 1e-4} for k, v in params.items()]
    if impl == "ours_graph":
        # the same iteration with the host out of the loop: gsb200.first_frame.FirstFrameEngine with one view slot --
        # fused activations, host-free forward, segmentation re-blend on the same geometry, flat record, statistics and
        # activation adjoint + Adam from the record, all replayed from one CUDA graph; the camera of the iteration is
        # copied into the slot (device to device) before each replay
        from gsb200 import first_frame, optim as goptim
        opt = goptim.FusedAdam(groups, lr=0.0, eps=1e-15)
        eng = first_frame.FirstFrameEngine(params, opt, 1, args.width, args.height, dev)
        one_cam = [{"viewmatrix": sc["viewmatrix"][i:i + 1], "projmatrix": sc["projmatrix"][i:i + 1],
                    "tanfovx": [sc["tanfovx"][i]], "tanfovy": [sc["tanfovy"][i]]} for i in range(args.cams)]
        td = [torch.tensor([[sc["tanfovx"][i], sc["tanfovy"][i]]], dtype=torch.float32, device=dev) for i in range(args.cams)]
        eng.set_views(one_cam[0], targets[0:1], seg_targets[0:1])
        eng.rebuild()
        # arena sized for the most expensive camera, not just camera 0
        eng.rebuild(capacity=2 * eng.arena.capacity)

        def one_g(it):
            i = it % args.cams
            eng.cams["viewmatrix"].copy_(one_cam[i]["viewmatrix"].reshape(1, 4, 4))
            eng.cams["projmatrix"].copy_(one_cam[i]["projmatrix"].reshape(1, 4, 4))
            eng.cams["tanfov_dev"].copy_(td[i])
            eng.images.copy_(targets[i:i + 1])
            eng.masks.copy_(seg_targets[i:i + 1])
            eng.step(check_every=0)

        for it in range(warm):
            one_g(it)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for it in range(iters):
            one_g(warm + it)
        e1.record()
        torch.cuda.synchronize()
        ok = eng.state.status()[0] == 0
        eng.close()
        if not ok:
            raise RuntimeError("arena overflow in the densify-shaped graph loop")
        return iters / (e0.elapsed_time(e1) * 1e-3)
    if impl == "ours":   # the fused step of gsb200.optim (one launch over all tables); the reference arm keeps torch's
        from gsb200 import optim as goptim
        opt = goptim.FusedAdam(groups, lr=0.0, eps=1e-15)
    else:
        opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    P = args.gaussians
    dv = gdens.DensificationVariables(visibility_count=torch.zeros(P, device=dev),
                                      mean_2d_gradients_accumulated=torch.zeros(P, device=dev),
                                      max_2d_radii=torch.zeros(P, device=dev))
    if impl == "ours":
        from diff_gaussian_rasterization import GaussianRasterizer as Renderer
        settings = [scene.settings_for(c, dev) for c in cams]
    else:
        import oracle_lib
        ref = oracle_lib.load_reference_ext()
        if ref is None:
            return None

    def render(i, colors):
        if impl == "ours":   # fused activations (gsb200.optim.create_render_arguments, shared.py:29-42)
            a = goptim.create_render_arguments(params)
            a["colors_precomp"] = colors
            a["means2D"].retain_grad()
            im, radii, _ = Renderer(raster_settings=settings[i])(**a)
            return im, radii, a["means2D"]
        a = {"means3D": params["means"], "colors_precomp": colors,
             "rotations": torch.nn.functional.normalize(params["rotation_quaternions"]),
             "opacities": torch.sigmoid(params["opacity_logits"]), "scales": torch.exp(params["log_scales"]),
             "means2D": torch.zeros_like(params["means"], requires_grad=True) + 0}
        a["means2D"].retain_grad()
        if impl == "ours":
            im, radii, _ = Renderer(raster_settings=settings[i])(**a)
        else:
            im, radii, _ = helpers.RefRasterize.apply(ref, a["means3D"], a["means2D"], a["colors_precomp"], a["opacities"],
                                                      a["scales"], a["rotations"], cams[i], bg)
        return im, radii, a["means2D"]

    def loss_fn(im, tgt):
        return gloss.l1_ssim_loss(im, tgt) if impl == "ours" else torch_ssim_l1(im, tgt)

    def one(it):
        i = it % args.cams
        im, radii, m2d = render(i, params["colors"])
        l_img = loss_fn(im, targets[i])
        seg, _, _ = render(i, params["segmentation_masks"])
        total = l_img + 3 * loss_fn(seg, seg_targets[i])
        total.backward()
        with torch.no_grad():
            dv.means_2d = m2d
            if impl == "ours":
                gdens.update_max_2d_radii_and_visibility_mask(radii, dv)
                gdens.accumulate_mean_2d_gradients(dv)
            else:  # densify.py:154-162 + external.py:113-124 as the reference writes them
                pos = radii > 0
                dv.max_2d_radii[pos] = torch.max(radii[pos], dv.max_2d_radii[pos])
                dv.gaussian_is_visible_mask = pos
                dv.mean_2d_gradients_accumulated[pos] += torch.norm(m2d.grad[pos, :2], dim=-1)
                dv.visibility_count[pos] += 1
        opt.step()
        opt.zero_grad(set_to_none=True)

    for it in range(warm):
        one(it)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(iters):
        one(warm + it)
    e1.record()
    torch.cuda.synchronize()
    return iters / (e0.elapsed_time(e1) * 1e-3)


def run_reference_cuda(args):
    from gsb200 import scene
    dev = torch.device("cuda", 0)
    g_raw, tgt_raw, cams = build_workload(args)
    sc = scene.stack_cameras(cams, dev)
    bg = torch.zeros(3, device=dev)
    targets = render_targets(tgt_raw, sc, args.cams, bg, dev)
    act = {k: v.to(dev).contiguous() for k, v in scene.activate(g_raw).items()}
    r = reference_cuda(args, act, cams, targets, bg, dev, passes=max(args.steps, 1))
    r["impl"] = "reference_cuda"
    print(json.dumps(r))


if __name__ == "__main__":
    # Libraries (NCCL's version banner, torchrun notices) write to stdout; the contract is ONE JSON line
    # there.  Send everything else to stderr and print the result line on the real stdout.
    _real_stdout = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)
    _print = print

    def print(*a, **k):  # noqa: A001
        k.setdefault("file", _real_stdout)
        _print(*a, **k)
        _real_stdout.flush()

    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.impl == "reference_cuda":
        run_reference_cuda(a)
    else:
        run_ours(a)
