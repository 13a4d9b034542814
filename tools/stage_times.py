#!/usr/bin/env python3
"""Per-stage device times of FullPathTracer (trace / surface / sort / shade) on one workload, plus a checksum of the accumulated
frame so that kernel tuning can be checked for bit-identity run to run. Usage: python tools/stage_times.py [workload] [W H] [frames]"""
import hashlib
import importlib
import os
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
pt = importlib.import_module("hiprt-path-tracer_b200")
wl = importlib.import_module("hiprt-path-tracer_b200.workloads")

name = sys.argv[1] if len(sys.argv) > 1 else "cornell"
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 8
cfg = getattr(wl, name)(W, H)
ctx = pt.Context(0)
g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
for _ in range(3):
    g.render_frame()
g.synchronize()
lanes = g.path_lanes
if lanes == 1:
    ctx.path_profile_enable(True)       # per-stage events need a single lane: the stages of concurrent lanes overlap
fpt = cam = 0.0
for _ in range(frames):
    g.render_frame()
    g.synchronize()
    fpt += g.last_kernel_ms("FullPathTracer")
    cam += g.last_kernel_ms("CameraRays")
stage = {k: round(v / frames, 3) for k, v in ctx.path_profile_read(True).items()} if lanes == 1 else f"(lanes={lanes})"
digest = hashlib.sha1(g.download("pixels").tobytes()).hexdigest()[:16]
print(f"{name} {W}x{H}: FullPathTracer {fpt / frames:.3f} ms  CameraRays {cam / frames:.3f} ms  stages {stage}  pixels sha1 {digest}")
