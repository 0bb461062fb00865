#!/usr/bin/env python
"""Scratch driver: the first-frame phase of bench.c5_sequence with other scene parameters (growth of the cloud).
usage: c5_probe.py P_gt init_noise steps_per_event [gt_scale|0] [color_noise]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
sys.argv = [sys.argv[0]] + sys.argv[1:]
P_gt, noise, spe = int(sys.argv[1]), float(sys.argv[2]), int(sys.argv[3])
gs = float(sys.argv[4]) if len(sys.argv) > 4 and float(sys.argv[4]) > 0 else None
cn = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
class A: pass
a = A(); a.width, a.height, a.cams = 640, 360, 27
torch.cuda.set_device(0)
r = bench.c5_sequence(a, 1, 0, torch.device("cuda", 0), P_gt=P_gt, init_noise=noise, steps_per_event=spe, timesteps=2, gt_scale=gs, color_noise=cn)
print(json.dumps({"end": r["gaussians_end"], "counts": r["gaussian_count_over_events"][::5], "ff": r["first_frame"]}))
