"""Pre-baked BRDF look-up tables consumed (read-only) by the principled BSDF: data/brdf_luts.npz, packed from the
reference's data/BRDFsData by tools/pack_luts.py (dimensions: src/Renderer/Baker/GPUBakerConstants.h:14-31)."""
import os

import numpy as np

LUT_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "data", "brdf_luts.npz")
_cache = None


def load_brdf_luts():
    global _cache
    if _cache is None:
        z = np.load(LUT_PATH)
        _cache = {k: np.ascontiguousarray(z[k], dtype=np.float32) for k in z.files}
    return _cache
