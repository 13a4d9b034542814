#!/usr/bin/env python3
"""A small tour of every kernel family for compute-sanitizer (tools/r02_sanitizer.sh): fused sample batches and the per-pass launches (MIS),
RIS, ReSTIR DI (fused and separate passes), adaptive sampling with nested dielectrics, the device BVH builder, the ray-batch entry points
(per-thread, ray queue in every loop flavour, static scratch), ray binning, tone mapping and the per-function debug kernels. Tiny frames: the
tools slow kernels down 10-100x."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
pt = importlib.import_module("hiprt-path-tracer_b200")
wl = importlib.import_module("hiprt-path-tracer_b200.workloads")


def render(ctx, cfg, frames=2, spf=1, in_flight=8, lanes=1, mode="phased"):
    ctx.set_path_mode(mode)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="gpu")
    g.set_samples_in_flight(in_flight)
    g.set_path_lanes(lanes)
    g.render_settings.samples_per_frame = spf
    for _ in range(frames):
        g.render_frame()
    g.synchronize()
    mean = float(g.download("pixels").mean())
    g.close()
    ctx.set_path_mode("phased")
    print(f"  {cfg['name'][:70]}: mode {mode}, {frames} frames x {spf} passes, {in_flight} in flight -> mean {mean:.4f}", flush=True)
    assert np.isfinite(mean)


def main():
    ctx = pt.Context(0)
    W, H = 64, 48
    render(ctx, wl.cornell(W, H), frames=2, spf=4, in_flight=4, lanes=1)
    ctx.set_ray_bin_bits(4)
    render(ctx, wl.synthetic_small(W, H), frames=1, spf=2, in_flight=2)
    ctx.set_ray_bin_bits(-1)
    render(ctx, wl.cornell(W, H), frames=1, spf=2, in_flight=0)
    for mode in ("megakernel", "fused", "split"):
        render(ctx, wl.cornell(32, 24), frames=1, mode=mode)
    ris = wl.cornell(W, H)
    ris["options"] = dict(lss="ris", ess="none", ggx=1, later=3)
    render(ctx, ris, frames=1)
    render(ctx, wl.many_lights(W, H, boxes=60, lights=800), frames=3)
    render(ctx, wl.many_lights(W, H, boxes=60, lights=800, fused=False, spatial_passes=2), frames=3)
    ctx.set_restir_phases(False)         # the single-kernel ReSTIR DI passes (per-thread traversal); phased above
    render(ctx, wl.many_lights(W, H, boxes=60, lights=800), frames=2)
    ctx.set_restir_phases(True)
    render(ctx, wl.adaptive_full(W, H, envmap_size=(128, 64)), frames=3)
    # ray batches + builders
    scene = wl.synthetic_small(W, H)["scene"]
    rng = np.random.default_rng(0)
    n = 20000
    lo, hi = scene.vertices.min(axis=0), scene.vertices.max(axis=0)
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.ascontiguousarray(np.concatenate([o, d], axis=1), np.float32)
    t_max = rng.uniform(0.1, 30.0, n).astype(np.float32)
    for device in (False, True):
        bvh = ctx.build_bvh(scene.vertices, scene.indices, device=device)
        prim, t, uv = ctx.trace_closest(bvh, rays)
        for flavour in (0, 5, 6, 7):
            ctx.set_trace_flavour(flavour)
            qprim, qt, quv = ctx.trace_queue(bvh, rays)
            assert np.array_equal(prim, qprim)
            ctx.trace_queue(bvh, rays, t_max)
        ctx.set_trace_flavour(-1)
        ctx.trace_shadow(bvh, rays, t_max)
        d_rays = ctx.upload(rays)
        d_prim, d_t, d_uv = ctx.malloc(4 * n), ctx.malloc(4 * n), ctx.malloc(8 * n)
        ctx.trace_queue_static(bvh, d_rays, None, None, n, d_prim, d_t, d_uv, None)
        ctx.sync()
        assert np.array_equal(ctx.download(d_prim, (n,), np.int32), prim)
        for p in (d_rays, d_prim, d_t, d_uv):
            ctx.free(p)
        ctx.destroy_bvh(bvh)
        print(f"  ray batches on the {'device' if device else 'host'}-built tree: {int((prim >= 0).sum())} hits of {n}", flush=True)
    ctx.close()
    print("sanitizer workload done")


if __name__ == "__main__":
    main()
