#!/usr/bin/env python3
"""Golden vectors of the LUT bakers (tests/golden/bake_tables.npz), made by the reference's own bake kernels compiled for the CPU
(oracle/ref_bake.cpp -> oracle/_ref/libhippt_ref_bake_*.so). Run in the build container only (needs /root/reference for the oracle build).

  small/<kind>_<sampler>   every kernel on a small grid (the reference integrates 1e8 / texels samples per texel whatever the requested count)
  slice/<kind>_z<k>        two ior slices of each full-size 3D table the reference ships (data/BRDFsData), at the shipped resolution and sample count
  full/conductor           the whole 128 x 128 conductor table
  full/thin_glass          the whole 32 x 32 x 96 thin-glass table (2.7 % of its texels are NaN in the reference's own arithmetic, see tests/test_bake.py)
The GPU tests compare b200pt_bake_* with these (same seeds, same samples: differences are libm ulps), and the full-size GPU tables with the
shipped LUTs of data/brdf_luts.npz (RGBE-quantised, so to 1/256 relative)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from oracle import ref  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "bake_tables.npz")
KINDS = {"conductor": 0, "fresnel": 1, "glossy": 2, "glass": 3, "glass_inverse": 4, "thin_glass": 5}
SMALL = {"conductor": (16, 16, 1), "fresnel": (12, 8, 8), "glossy": (12, 8, 8), "glass": (12, 8, 8), "glass_inverse": (12, 8, 8), "thin_glass": (12, 8, 8)}
FULL = {"glossy": (128, 64, 128), "glass": (256, 16, 128), "glass_inverse": (256, 16, 128)}
SLICES = {"glossy": (5, 100), "glass": (10, 90), "glass_inverse": (10, 90)}


def main():
    out = {}
    for name, kind in KINDS.items():
        for sampler in ("vndf", "caps") if name in ("conductor", "glass") else ("vndf",):
            sx, sy, sz = SMALL[name]
            t = time.time()
            table, ipk, launches = ref.bake(kind, sx, sy, sz, 4096, sampler=sampler)
            out[f"small/{name}_{sampler}"] = table
            out[f"small/{name}_{sampler}_plan"] = np.array([sx, sy, sz, 4096, ipk, launches], np.int32)
            print(f"small {name} {sampler}: {sx}x{sy}x{sz}, {ipk} samples x {launches} launches, {time.time() - t:.1f} s, range {table.min():.4f} .. {table.max():.4f}", flush=True)
    for name, (sx, sy, sz) in FULL.items():
        for z in SLICES[name]:
            t = time.time()
            table, ipk, launches = ref.bake(KINDS[name], sx, sy, sz, 65536, z_range=(z, z + 1))
            out[f"slice/{name}_z{z}"] = table[z]
            print(f"slice {name} z={z}: {ipk} samples x {launches} launches, {time.time() - t:.1f} s", flush=True)
    t = time.time()
    table, ipk, launches = ref.bake(0, 128, 128, 1, 65536)
    out["full/conductor"] = table[0]
    print(f"full conductor: {ipk} x {launches}, {time.time() - t:.1f} s", flush=True)
    t = time.time()
    table, ipk, launches = ref.bake(5, 32, 32, 96, 65536)
    out["full/thin_glass"] = table
    print(f"full thin glass: {ipk} x {launches}, {time.time() - t:.1f} s, NaN texels {int((~np.isfinite(table)).sum())} of {table.size}", flush=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
