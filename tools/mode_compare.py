#!/usr/bin/env python3
"""A/B of the three FullPathTracer shapes (megakernel / fused wavefront / split wavefront) on one workload:
bit-equality of the accumulated frame and ms per sample pass. Usage: python tools/mode_compare.py [workload] [W H] [spp]"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
pt = importlib.import_module("hiprt-path-tracer_b200")
wl = importlib.import_module("hiprt-path-tracer_b200.workloads")

name = sys.argv[1] if len(sys.argv) > 1 else "cornell"
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 8
cfg = getattr(wl, name)(W, H)
ctx = pt.Context(0)
g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
frames = {}
MODES = os.environ.get("MODES", "megakernel,fused,split,phased").split(",")
for mode in MODES:
    ctx.set_path_mode(mode)
    g.reset()
    g.resize(W, H)
    wl.apply_settings(g.render_settings, cfg["settings"])
    ms = []
    for s in range(spp):
        g.render_frame()
        g.synchronize()
        ms.append(g.last_kernel_ms("FullPathTracer"))
    frames[mode] = g.download("pixels").copy()
    print(f"{mode:10s} FullPathTracer ms/pass: first {ms[0]:.3f}  median {sorted(ms)[len(ms)//2]:.3f}  mean/px {frames[mode].mean()/spp:.6f}")
for mode in MODES[1:]:
    same = np.array_equal(frames[MODES[0]], frames[mode])
    print(f"{mode} == {MODES[0]} bit for bit: {same}")
    if not same:
        d = np.abs(frames[MODES[0]] - frames[mode])
        print("  max abs diff", d.max(), "pixels differing", int((d.max(axis=2) > 0).sum()))
