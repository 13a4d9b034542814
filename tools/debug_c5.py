#!/usr/bin/env python3
"""Which feature combination makes the GPU and the oracle differ on the nested-dielectrics-complex scene (debug aid)."""
import importlib, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import numpy as np
pt = importlib.import_module("hiprt-path-tracer_b200"); wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
from oracle import ref
W, H = 128, 96
ctx = pt.Context(0)
for thin, disp, adaptive, mode in ((0, 1, 0, "phased"), (1, 0, 0, "phased"), (1, 1, 0, "phased"), (1, 1, 0, "megakernel"), (0, 1, 1, "phased")):
    cfg = wl.golden_scene("nested-dielectrics-complex", W, H, adaptive=bool(adaptive), thin_film=bool(thin))
    if disp:
        for m in cfg["scene"].materials:
            if m.base.specular_transmission > 0.0 and m.base.ior > 1.05:
                m.base.dispersion_scale = 1.0
    ctx.set_path_mode(mode)
    r = wl.setup_oracle(ref, pt.structs, cfg); g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    r.render(1); g.render_frame(); g.synchronize()
    a, b = r.pixels(), g.download("pixels")
    bad = (np.abs(a - b) / np.maximum(a, 1e-3) > 1e-2).any(axis=2)
    va, vb = r.buffer("ray_volume_states"), g.download("ray_volume_states")
    print(f"thin {thin} disp {disp} adaptive {adaptive} {mode}: means {a.mean():.5f} {b.mean():.5f} pixels >1% off {int(bad.sum())}; G-buffer volume states differ in {int((va[..., :28] != vb[..., :28]).any(axis=-1).sum())} px", file=sys.stderr)
    g.close()
ctx.set_path_mode("phased")
