#!/usr/bin/env python3
"""Which G-buffer material fields differ between the oracle and the CUDA path on the textured Cornell scene (debug aid)."""
import importlib, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
pt = importlib.import_module("hiprt-path-tracer_b200"); wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
from oracle import ref
import test_gpu_parity as T
W, H = 160, 96
cfg = T._textured_cornell(wl, W, H)
ctx = pt.Context(0)
r = wl.setup_oracle(ref, pt.structs, cfg); g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
r.render(1); g.render_frame(); g.synchronize()
hit = r.buffer("camera_ray_hit") == 1
print("prim equal", np.array_equal(r.buffer("first_hit_prim_index"), g.download("first_hit_prim_index")))
ma, mb = r.buffer("materials")[hit], g.download("materials")[hit]
fa, fb = ma.view(np.float32), mb.view(np.float32)
diff = (ma.view(np.uint32) != mb.view(np.uint32))
print("pixels differing", diff.any(axis=1).mean())
fields = [(n, getattr(pt.structs.SimplifiedRendererMaterial, n).offset) for n, _ in pt.structs.SimplifiedRendererMaterial._fields_]
for word in np.nonzero(diff.any(axis=0))[0]:
    name = [n for n, o in fields if o <= word * 4][-1]
    rows = np.nonzero(diff[:, word])[0]
    print(f"word {word} ({name}): {len(rows)} pixels; e.g. ref {fa[rows[0], word]!r} gpu {fb[rows[0], word]!r}; max abs diff {np.nanmax(np.abs(fa[rows, word] - fb[rows, word]))}")
prim = r.buffer("first_hit_prim_index")[hit]
mat = cfg["scene"].material_indices[prim]
rows = np.nonzero(diff.any(axis=1))[0]
print("materials of differing pixels:", np.unique(mat[rows], return_counts=True))
sn = (r.buffer("shading_normals")[hit] != g.download("shading_normals")[hit]).any(axis=1)
print("shading normals differing:", sn.mean(), np.unique(mat[sn], return_counts=True))
g.close()
