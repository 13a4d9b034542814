#!/usr/bin/env python3
"""Bakes the directional-albedo tables on the GPU (hiprt-path-tracer_b200/baker.py -> libb200pt.so) and reports time and samples/s per table.
   python tools/bake_tables.py [--out DIR] [--tables conductor,glossy,glass,thin_glass,fresnel] [--scale K] [--chunks N]
--scale K divides every 3D table's ior axis by K (profiling under ncu); --out writes the reference's .hdr files."""
import argparse
import importlib
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--tables", default="conductor,glossy,glass,thin_glass")
    ap.add_argument("--scale", type=int, default=1)
    ap.add_argument("--chunks", type=int, default=0)
    ap.add_argument("--ggx", type=int, default=0)
    args = ap.parse_args()
    pt = importlib.import_module("hiprt-path-tracer_b200")
    bk = importlib.import_module("hiprt-path-tracer_b200.baker")
    ctx = pt.Context(0)
    if args.out:
        os.makedirs(args.out, exist_ok=True)
    baker = bk.GPUBaker(ctx, ggx_sample_function=args.ggx, chunks=args.chunks, output_directory=args.out)
    k = args.scale
    jobs = {
        "conductor": (0, "ggx_conductor", lambda: baker.bake_ggx_conductor_directional_albedo(output_filename=bk.get_GGX_conductor_Ess_filename()), (128, 128, 1)),
        "fresnel": (1, "ggx_fresnel", lambda: baker.bake_ggx_fresnel_directional_albedo(bk.GGXFresnelDirectionalAlbedoSettings(256, 256, 256 // k), bk.get_GGX_fresnel_Ess_filename()), (256, 256, 256 // k)),
        "glossy": (2, "glossy_dielectric", lambda: baker.bake_glossy_dielectric_directional_albedo(bk.GlossyDielectricDirectionalAlbedoSettings(128, 64, 128 // k), bk.get_glossy_dielectric_Ess_filename()), (128, 64, 128 // k)),
        "glass": (3, "ggx_glass", lambda: baker.bake_ggx_glass_directional_albedo(bk.GGXGlassDirectionalAlbedoSettings(256, 16, 128 // k), bk.get_GGX_glass_Ess_filename()), (256, 16, 128 // k)),
        "thin_glass": (5, "ggx_thin_glass", lambda: baker.bake_ggx_thin_glass_directional_albedo(bk.GGXThinGlassDirectionalAlbedoSettings(32, 32, 96 // k), bk.get_GGX_thin_glass_Ess_filename()), (32, 32, 96 // k)),
    }
    report = {}
    for name in args.tables.split(","):
        kind, key, run, (sx, sy, sz) = jobs[name]
        run()
        plan = bk.bake_plan(kind, sx, sy, sz, 65536, baker.options)
        samples = plan.texels * plan.nb_kernel_launch * plan.iterations_per_kernel * (2 if name == "glass" else 1)
        ms = baker.last_kernel_ms[key]
        report[name] = {"size": [sx, sy, sz], "samples_per_texel": plan.nb_kernel_launch * plan.iterations_per_kernel, "threads_per_texel": plan.chunks,
                        "ms": round(ms, 2), "gsamples_per_s": round(samples / ms / 1e6, 2)}
    print(json.dumps(report))
    ctx.close()


if __name__ == "__main__":
    main()
