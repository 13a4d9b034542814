#!/usr/bin/env python3
"""Packs the reference's pre-baked directional-albedo LUTs (data/BRDFsData/**/*.hdr, 481 Radiance RGBE slices) and the
sheen LTC table (src/Device/includes/BSDFs/SheenLTCFittedParameters.h) into data/brdf_luts.npz.

These tables are INPUT DATA of the hot path ("LUTs ship pre-baked ... consumed read-only");
both the oracle and the CUDA path receive them as float arrays. Decoding goes through the reference's own loader
(Image32Bit::read_image_hdr with flipY = true, first channel, as CPURenderer::setup_brdfs_data does,
src/Renderer/CPURenderer.cpp:93-132) via oracle/_ref. Run in the build container only (needs /root/reference)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from oracle import ref  # noqa: E402

REF_DATA = "/root/reference/data/BRDFsData"
OUT = os.path.join(os.path.dirname(__file__), "..", "data", "brdf_luts.npz")


def load_hdr(lib, path, w, h):
    out = np.empty((h, w), np.float32)
    ww, hh = C.c_int(), C.c_int()
    rc = lib.ref_load_hdr(path.encode(), 1, 1, out.ctypes.data_as(C.c_void_p), out.size, C.byref(ww), C.byref(hh))
    assert rc == 0 and (ww.value, hh.value) == (w, h), (path, rc, ww.value, hh.value)
    return out


def main():
    lib = ref.load("mis")
    luts = {}
    sheen = np.empty((32, 32, 3), np.float32)
    lib.ref_sheen_ltc_table(sheen.ctypes.data_as(C.c_void_p))
    luts["sheen_ltc"] = sheen
    luts["conductor"] = load_hdr(lib, f"{REF_DATA}/GGX/GGX_Conductor_128x128.hdr", 128, 128)
    luts["glossy"] = np.stack([load_hdr(lib, f"{REF_DATA}/GlossyDielectrics/{i}Glossy_Ess_128x64x128.hdr", 128, 64) for i in range(128)])
    luts["glass"] = np.stack([load_hdr(lib, f"{REF_DATA}/GGX/Glass/{i}GGX_Glass_Ess_256x16x128.hdr", 256, 16) for i in range(128)])
    luts["glass_inverse"] = np.stack([load_hdr(lib, f"{REF_DATA}/GGX/Glass/{i}inv_GGX_Glass_Ess_256x16x128.hdr", 256, 16) for i in range(128)])
    luts["thin_glass"] = np.stack([load_hdr(lib, f"{REF_DATA}/GGX/Glass/{i}GGX_Thin_Glass_Ess_32x32x96.hdr", 32, 32) for i in range(96)])
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **luts)
    for k, v in luts.items():
        print(k, v.shape, float(v.min()), float(v.max()))
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
