#!/usr/bin/env python3
"""Generates the golden vectors under tests/golden/ from the reference's own CPU kernels (oracle/_ref).

The reference ships no tests, golden images or known-answer vectors, so the pins are produced
here, in the build container, by running the reference code itself; this script is the committed generator."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
pt = importlib.import_module("hiprt-path-tracer_b200")
wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
from oracle import ref  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def cornell_golden(variant, options_lss, spp=4, W=64, H=48):
    cfg = wl.cornell(W, H)
    cfg["oracle_variant"] = variant
    r = wl.setup_oracle(ref, pt.structs, cfg)
    r.ray_log_begin(400_000, 1)
    r.render(1)
    log = r.ray_log_end(400_000)
    first = r.pixels().copy()
    prim = r.buffer("first_hit_prim_index").copy()
    normals = r.buffer("shading_normals").copy()
    r.render(spp - 1)
    keep = slice(0, 60_000)
    np.savez_compressed(os.path.join(OUT, f"cornell_{W}x{H}_{variant}.npz"), width=W, height=H, spp=spp, lss=options_lss,
                        pixels_1spp=first, pixels=r.pixels(), first_hit_prim_index=prim, shading_normals=normals,
                        rays=log["rays"][keep], last_prim=log["last_prim"][keep], hit_prim=log["hit_prim"][keep], hit_t=log["t"][keep], hit_uv=log["uv"][keep],
                        ray_count=len(log["rays"]))
    print(variant, "rays logged", len(log["rays"]), "mean radiance", r.pixels().mean() / spp)


def kats():
    lib = ref.load("mis")
    import ctypes as C
    n = 64
    u32, f32, idx = np.empty(n, np.uint32), np.empty(n, np.float32), np.empty(n, np.int32)
    lib.ref_xorshift(C.c_uint(42), C.c_int(n), u32.ctypes.data_as(C.c_void_p), f32.ctypes.data_as(C.c_void_p), idx.ctypes.data_as(C.c_void_p), C.c_int(1000))
    hashes = np.array([lib.ref_wang_hash(C.c_uint(i)) & 0xFFFFFFFF for i in range(64)], dtype=np.uint32)
    # glass correction exponent on a grid (tuning table of MicrofacetEnergyCompensation.h:88-635)
    r = ref.RefRenderer("mis")
    rough = np.linspace(0.0, 1.0, 41, dtype=np.float32)
    etas = np.array([1.011, 1.015, 1.02, 1.025, 1.05, 1.1, 1.15, 1.3, 1.45, 1.5, 1.7, 2.0, 2.2, 2.4, 2.7, 3.0], dtype=np.float32)
    corr = np.array([[lib.ref_glass_correction_exponent(r.h, C.c_float(float(a)), C.c_float(float(e))) for a in rough] for e in etas], dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, "kats.npz"), xorshift_u32=u32, xorshift_float=f32, xorshift_index_1000=idx, wang_hash_0_63=hashes,
                        glass_rough=rough, glass_etas=etas, glass_correction=corr)
    print("kats: wang_hash(1) =", hashes[1], "xorshift(42) ->", u32[:2])


if __name__ == "__main__":
    cornell_golden("mis", 3)
    cornell_golden("ris", 4)
    kats()
