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

    g_raw, tgt_raw, cams = build_workload(args)
    # camera -> rank.  N > 1: cost-balanced (gsb200.dist.shard_cameras with per-camera costs): every rank renders
    # all cameras once, untimed, and estimates each camera's cost from its walked pairs and instance count (what a
    # trainer takes from the previous timestep's renders); rank 0's numbers are broadcast so all ranks agree.
    cam_costs = None
    if world > 1 and os.environ.get("GSB_BENCH_ROUND_ROBIN", "0") != "1":
        all_sc = scene.stack_cameras(cams, dev)
        a0 = {k: v.to(dev).contiguous() for k, v in scene.activate(g_raw).items()}
        _, _, _, st0 = raster.forward(a0["means3D"], a0["colors_precomp"], a0["opacities"], a0["scales"], a0["rotations"],
                                      None, torch.zeros(3, device=dev), 1.0, all_sc["viewmatrix"], all_sc["projmatrix"],
                                      all_sc["tanfovx"], all_sc["tanfovy"], all_sc["image_height"], all_sc["image_width"],
                                      n_cams=args.cams)
        ct = torch.tensor(gd.camera_costs(st0), dtype=torch.float64, device=dev)
        dist.broadcast(ct, 0)
        cam_costs = [float(x) for x in ct.cpu()]
        del st0, a0, all_sc
    my_cam_ids = gd.shard_cameras(args.cams, rank, world, cam_costs)
    my_cams = [cams[i] for i in my_cam_ids]
    n_my = len(my_cams)
    sc = scene.stack_cameras(my_cams, dev)
    bg = torch.zeros(3, device=dev)
    targets = render_targets(tgt_raw, sc, n_my, bg, dev)
    P = args.gaussians

    # host (pinned) copies for the end-to-end leg
    raw_keys = ["means", "colors", "rotation_quaternions", "opacity_logits", "log_scales"]
    host_raw = {k: g_raw[k].contiguous().pin_memory() for k in raw_keys}
    host_targets = targets.cpu().pin_memory()
    dev_raw = {k: torch.empty_like(v, device=dev) for k, v in host_raw.items()}
    dev_targets = torch.empty_like(targets)
    loss_host = torch.empty((n_my, 3), dtype=torch.float32).pin_memory()

    stats = step_mod.DensifyStats(P, dev)
    # flat buffer for the one allreduce: 17 gradient floats + 2 summed statistics per Gaussian
    flat = torch.zeros((P, gd.WIDTH), dtype=torch.float32, device=dev) if world > 1 else None

    def activate():
        return {k: v.contiguous() for k, v in scene.activate(dev_raw).items()}

    def one_step(act):
        loss_stats, grads, radii, st = step_mod.timestep_step(act, sc, dev_targets, bg, stats)
        if world > 1:
            gd.pack(flat, grads, {"mean_2d_gradients_accumulated": stats.mean_2d_gradients_accumulated,
                                  "visibility_count": stats.visibility_count})
            gd.allreduce(flat, stats.max_2d_radii)
        return loss_stats, grads, st

    for k in raw_keys:
        dev_raw[k].copy_(host_raw[k])
    dev_targets.copy_(host_targets)
    act = activate()

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.profile_only:
        for _ in range(3):
            one_step(act)
        torch.cuda.synchronize()
        return

    # ---- device-resident timing -------------------------------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        loss_stats, grads, st = one_step(act)
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.begin()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    e0.record()
    marks[0].record()
    for k in range(args.steps):
        loss_stats, grads, st = one_step(act)
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

    # ---- N > 1: what the exchange costs (each piece timed alone, max over ranks) -----------------
    exchange = None
    if world > 1:
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
            dist.all_reduce(tt_, op=dist.ReduceOp.MAX)
            timed_loop.own = own
            return float(tt_.item())
        sd = {"mean_2d_gradients_accumulated": stats.mean_2d_gradients_accumulated,
              "visibility_count": stats.visibility_count}
        exchange = {
            "pack": timed_loop(lambda: gd.pack(flat, grads, sd)),
            "allreduce_sum": timed_loop(lambda: dist.all_reduce(flat)),
            "allreduce_max": timed_loop(lambda: dist.all_reduce(stats.max_2d_radii, op=dist.ReduceOp.MAX)),
            "step_without_exchange": timed_loop(lambda: step_mod.timestep_step(act, sc, dev_targets, bg, stats)),
            "bytes_sum": flat.numel() * 4, "bytes_max": stats.max_2d_radii.numel() * 4,
        }
        per_rank = [None] * world
        dist.all_gather_object(per_rank, {"rank": rank, "cameras": my_cam_ids, "step_without_exchange": timed_loop.own,
                                          "est_cost": None if cam_costs is None else sum(cam_costs[c] for c in my_cam_ids)})
        exchange["per_rank"] = per_rank

    # ---- end to end: host buffers in, loss out, every step ------------------------------------
    # gsb200.step.HostFeed: each step's inputs (raw parameter tables + target images, pinned host memory) are
    # copied host -> device (double-buffered, the copy of step k+1 overlaps step k), activations + step run on
    # the device, the per-camera loss numbers come back to pinned host memory; one stream sync per step.
    feed = step_mod.HostFeed(host_raw, host_targets, dev)
    E2E_BLOCKING = os.environ.get("GSB_BENCH_E2E_BLOCKING", "0") == "1"

    def after(loss_stats, grads, radii, st_):
        if world > 1:
            gd.pack(flat, grads, {"mean_2d_gradients_accumulated": stats.mean_2d_gradients_accumulated,
                                  "visibility_count": stats.visibility_count})
            gd.allreduce(flat, stats.max_2d_radii)

    def e2e_step():
        # non-blocking: the loss numbers are copied to pinned host memory every step, the host does not wait for
        # them before issuing the next step (a training loop only logs them); GSB_BENCH_E2E_BLOCKING=1 waits
        feed.step(sc, bg, stats, after, wait=E2E_BLOCKING)

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
    for _ in range(reps):
        loss_stats, grads, st = one_step(act)
        torch.cuda.synchronize()
        lib.gsb_profile_read(buf, len(_lib.PROFILE_STAGES))
        stage_acc += np.array(list(buf))
    lib.gsb_profile_enable(0)
    stages = {k: float(v / reps) for k, v in zip(_lib.PROFILE_STAGES, stage_acc) if not k.endswith("_begin")}

    # ---- workload statistics ---------------------------------------------------------------------
    tt = st.tensors()
    walked = int(tt["n_contrib"].long().sum().item())             # sum over pixels of n_contrib
    R = st.num_rendered
    cnt = torch.tensor([walked, R], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(cnt)
    walked_all, R_all = int(cnt[0].item()), int(cnt[1].item())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, sm_max_mhz, peak_src = peaks()
    traffic = {}
    try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
        if (args.gaussians, args.cams, args.width, args.height, world) == (300_000, 27, 640, 360, 1):
            with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as f:
                traffic = json.load(f)["kernels"]
    except Exception:
        traffic = {}
    tr = lambda name: (traffic.get(name, {}).get("dram_bytes_per_launch") if traffic else None)
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    fp32_peak_tflops = n_sm * 128 * 2 * sm_max_mhz * 1e6 / 1e12
    # dominant kernel: backward blend.  Algorithmic work: 70 FP32 flop per walked pair.
    bwd_ms = stages["render_bwd"]
    achieved = (walked * 70.0) / (bwd_ms * 1e-3) / 1e12 if bwd_ms > 0 else 0.0
    roofline = {"kernel": "render_bwd2_kernel", "bound": "fp32", "achieved": achieved, "peak": fp32_peak_tflops,
                "unit": "TFLOP/s", "frac": achieved / fp32_peak_tflops, "traffic": tr("render_bwd2_kernel"),
                "peak_source": f"{n_sm} SMs x 128 lanes x 2 x {sm_max_mhz:.0f} MHz ({peak_src} sm_max_mhz)",
                "ms": bwd_ms, "pairs": walked}
    # the HBM-bound kernel with the most bytes: the instance sort (segmented onesweep on 16-bit camera-local
    # tile ids).  Algorithmic bytes per pass: the histogram reads the key (2); the scatter reads key (2) + payload (4)
    # and writes the payload (4) (+ the key (2) when another pass follows).  DESIGN.md section 5.
    tile_digit = int(os.environ.get("GSB_TILE_DIGIT", "0")) or (5 if st.ctx.tile_bits <= 5 else 8 if st.ctx.tile_bits <= 8 else 10 if st.ctx.tile_bits <= 10 else 8)
    passes = (max(st.ctx.tile_bits, 1) + tile_digit - 1) // tile_digit
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
    # preprocess, depth sort (setup + 3 kernels per pass), scan, setup, key emission, instance sort (3 per pass),
    # ranges, blend fwd, ssim fwd/finalize/bwd, blend bwd, preprocess bwd, densify statistics
    launches = 1 + (1 + 3 * gpasses) + 1 + 1 + 1 + 3 * passes + 1 + 1 + 3 + 1 + 1 + 1

    out = {
        "metric": "train iters/sec (one iter = one camera fwd + L1/SSIM + bwd; 27 cameras per step)",
        "value": args.cams / (ms_step * 1e-3), "unit": "iters/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"Panoptic-shaped timestep: {P} gaussians, {args.cams} cameras {args.width}x{args.height}, "
                               "fwd + L1/SSIM + bwd + densify stats (BASELINE.json configs[1])",
                   "cameras_per_rank": n_my, "parallelism": f"camera-sharded dp{world}",
                   "camera_assignment": ("round-robin" if cam_costs is None else
                                         "cost-balanced (LPT on per-camera walked pairs + instances of an untimed calibration render)"),
                   "l2": "per-step working set (state + instance lists) is several hundred MB > 126 MB L2; no explicit flush"},
        "blends_per_sec": 2.0 * walked_all / (ms_step * 1e-3), "walked_pairs_per_step": walked_all,
        "num_rendered_per_step": R_all,
        "e2e": {"value": args.cams / (ms_e2e * 1e-3), "unit": "iters/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches * args.steps,
        "stage_ms": stages, "roofline": roofline, "roofline_hbm": roofline_hbm, "clocks": clocks,
    }
    if exchange is not None:
        out["exchange_ms"] = exchange
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args, g_raw, cams, host_targets)
    if world == 1 and not args.no_reference_cuda:
        out["reference_cuda"] = reference_cuda(args, act, my_cams, dev_targets, bg, dev)
        try:
            out["densify_loop_shape"] = {
                "what": "densify.py:201-248 shape: 1 camera/iteration, colour + segmentation render, 2x L1/SSIM, "
                        "backward, densify statistics, torch Adam; through the drop-in Python surface",
                "unit": "iters/s", "ours": densify_shaped(args, "ours"),
                "reference_cuda": densify_shaped(args, "reference_cuda")}
        except Exception as ex:  # never lose the headline line to the extra measurement
            out["densify_loop_shape"] = {"error": repr(ex)[:300]}
        try:
            out["rigidity_loss"] = rigidity_shaped(args)
        except Exception as ex:
            out["rigidity_loss"] = {"error": repr(ex)[:300]}
        try:
            out["deformation_forward"] = deformation_shaped(args)
        except Exception as ex:
            out["deformation_forward"] = {"error": repr(ex)[:300]}
        try:
            out["train_step_shape"] = train_step_shaped(args)
        except Exception as ex:
            out["train_step_shape"] = {"error": repr(ex)[:300]}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


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
        res["fwd_bwd_ours_3xtf32_fwd" if precise else "fwd_bwd_ours_tf32"] = timed(
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
    from gsb200 import scene, deform, rigidity as R, timestep_train as tt
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
    opt = torch.optim.Adam(net.parameters(), lr=1e-4)
    last = {}

    def ours(k):
        loss, _, _ = tt.train_step(net, opt, init, enc0, enc0, fg0, ninfo, sub(picks[k]), 7, 150)
        last["ours"] = loss

    res = {"what": f"one train.py step: DeformationNetwork ({hidden} x {blocks} blocks) over {args.gaussians} Gaussians, "
                   f"{views} of {args.cams} cameras {args.width}x{args.height}, rigidity K={neighbours}, Adam",
           "unit": "ms per step", "ours": timed(ours)}
    res["ours_steps_per_s"] = 1e3 / res["ours"]
    res["network_matmuls"] = "3xTF32 forward (FP32-level), TF32 backward"
    torch.manual_seed(0)
    net = deform.DeformationNetwork(hidden, blocks, precise=False).to(dev)
    opt = torch.optim.Adam(net.parameters(), lr=1e-4)
    res["ours_tf32_forward"] = timed(ours)
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
    act_np = {k: v.numpy() for k, v in scene.activate(g_raw).items()}
    n_avail = min(len(cams), host_targets.size(0))
    done = 0
    t0 = time.perf_counter()
    while done < n_avail and (done == 0 or time.perf_counter() - t0 < budget_s):
        cpu_step(o, act_np, cams[done], host_targets[done].numpy(), args.height, args.width)
        done += 1
    dt = time.perf_counter() - t0
    return {"value": done / dt, "unit": "iters/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"cameras 0..{done - 1} of the workload ({args.gaussians} gaussians, {args.width}x{args.height}): "
                      f"fwd + L1/SSIM + bwd on the C oracle (OpenMP), {done} iterations in {dt:.1f} s"}


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
              f"C oracle (OpenMP); the reference has no CPU path for the rasterizer")
    print(json.dumps({
        "impl": "reference", "metric": "train iters/sec (one iter = one camera fwd + L1/SSIM + bwd; 27 cameras per step)",
        "value": v, "unit": "iters/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 1),
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"Panoptic-shaped timestep: {args.gaussians} gaussians, {args.cams} cameras "
                               f"{W}x{H}, fwd + L1/SSIM + bwd + densify stats (BASELINE.json configs[1])",
                   "sample": sample},
        "cpu_baseline": {"value": v, "unit": "iters/s", "cores": os.cpu_count(), "kind": "port", "sample": sample},
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


def reference_cuda(args, act, cams, targets, bg, dev, passes=2):
    """The unmodified reference CUDA extension (oracle/_ref) on the same GPU, same inputs, its own
    per-camera call pattern: forward -> torch L1/SSIM -> autograd -> backward, camera by camera."""
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

    one_pass()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(passes):
        one_pass()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / passes
    return {"value": len(cams) / (ms * 1e-3), "unit": "iters/s", "ms_per_step": ms,
            "what": "unmodified reference extension (sm_100a build, CUB sort) + stock torch L1/SSIM, camera by camera"}


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
