#!/usr/bin/env python3
"""Development probe (run under gpurun): oracle vs CUDA on the Cornell box -- hit ids on logged ray batches, G-buffer,
per-sample radiance -- plus a first throughput number. The pytest suite under tests/ is the durable version of this."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
pt = importlib.import_module("hiprt-path-tracer_b200")
sc = importlib.import_module("hiprt-path-tracer_b200.scene")
from oracle import ref  # noqa: E402

S = pt.structs


def configure(rs, bounces=8):
    rs.nb_bounces = bounces
    rs.do_alpha_testing = 0
    rs.enable_adaptive_sampling = 0
    rs.samples_per_frame = 1


def main():
    W = int(os.environ.get("W", 256))
    H = int(os.environ.get("H", 256))
    SPP = int(os.environ.get("SPP", 4))
    scene = sc.load_scene(os.path.join(ROOT, "tests", "golden", "cornell_pbr.npz"))
    luts = pt.load_brdf_luts()
    cam = pt.scene_camera(scene, W, H)

    r = ref.RefRenderer("mis")
    r.set_luts(luts)
    r.set_scene(scene)
    r.set_camera(cam)
    configure(r.struct("render_settings", S.HIPRTRenderSettings))
    r.resize(W, H)

    ctx = pt.Context(0)
    g = pt.GPURenderer(ctx, options=dict(lss="mis", ess="none"), seed_schedule="cpu")
    g.set_brdf_luts(luts)
    g.set_scene(scene)
    print("bvh", ctx.bvh_info(g.bvh))
    g.set_camera(cam)
    configure(g.render_settings)
    g.resize(W, H)

    # ---- ray-batch parity on rays logged from the oracle's own render ----
    r.ray_log_begin(4_000_000, 1)
    t0 = time.time()
    r.render(1)
    t_ref = time.time() - t0
    log = r.ray_log_end(4_000_000)
    n = len(log["rays"])
    prim, t, uv = ctx.trace_closest(g.bvh, log["rays"], log["last_prim"])
    # The log holds closest-hit queries (trace_ray, shadow rays and shadow-light rays all call BVH::intersect)
    mism = prim != log["hit_prim"]
    print(f"ray batch: {n} rays, primID mismatches {mism.sum()} ({mism.mean():.2e}); t bit-equal on matches: "
          f"{(t[~mism].view(np.uint32) == log['t'][~mism].view(np.uint32)).mean():.6f}")
    if mism.any():
        idx = np.nonzero(mism)[0][:10]
        for i in idx:
            print("  ray", i, log["rays"][i], "last", log["last_prim"][i], "ref", log["hit_prim"][i], log["t"][i], "gpu", prim[i], t[i])

    # ---- frame parity, sample by sample ----
    g.render_frame()
    g.synchronize()
    for name in ("first_hit_prim_index", "camera_ray_hit"):
        a, b = r.buffer(name), g.download(name)
        print(name, "mismatch", int((a != b).sum()), "of", a.size)
    for name in ("shading_normals", "first_hits", "view_directions"):
        a, b = r.buffer(name), g.download(name)
        hit = r.buffer("camera_ray_hit") == 1
        m = hit if name != "view_directions" else np.ones_like(hit)
        print(name, "max abs diff", float(np.abs(a[m] - b[m]).max()), "bit-equal frac", float((a[m].view(np.uint32) == b[m].view(np.uint32)).mean()))
    pa, pb = r.pixels(), g.download("pixels")
    d = np.abs(pa - pb)
    print("sample 0: max abs diff", float(d.max()), "mean abs diff", float(d.mean()), "bit-equal px", float((pa.view(np.uint32) == pb.view(np.uint32)).all(axis=2).mean()),
          "ref mean", pa.mean(), "gpu mean", pb.mean())
    bad = np.argwhere(d.max(axis=2) > 1e-3)
    print("pixels with diff > 1e-3:", len(bad), bad[:5].tolist())

    for s in range(1, SPP):
        r.render(1)
        g.render_frame()
    g.synchronize()
    pa, pb = r.pixels() / SPP, g.download("pixels") / SPP
    rel = np.abs(pa - pb) / np.maximum(pa, 1e-3)
    print(f"{SPP} spp: mean relative error {float(rel.mean()):.3e}, mean ref {pa.mean():.5f} gpu {pb.mean():.5f}, px identical {float((pa == pb).all(axis=2).mean()):.4f}")

    # ---- throughput: 1080p, GPU seed schedule, counting build for rays then timing build ----
    W2, H2 = 1920, 1080
    g.set_seed_schedule("gpu")
    g.set_camera(pt.scene_camera(scene, W2, H2))
    g.resize(W2, H2)
    ctx.counters_enable(True)
    ctx.counters_read(True)
    g.render_frame()
    g.synchronize()
    c = ctx.counters_read(True)
    ctx.counters_enable(False)
    print("1080p counters / sample:", c, "nodes/ray", c["nodes"] / max(1, c["rays"]), "tris/ray", c["triangles"] / max(1, c["rays"]))
    g.render_settings.samples_per_frame = 8
    for _ in range(2):
        g.render_frame()
    g.synchronize()
    ms = g.last_frame_ms()
    print(f"1080p: {ms / 8:.3f} ms/sample -> {c['rays'] / (ms / 8 * 1e-3) / 1e6:.1f} Mrays/s, {8 / (ms * 1e-3):.1f} samples/px/s; oracle {W}x{H} 1 spp in {t_ref:.2f}s")


if __name__ == "__main__":
    main()
