#!/usr/bin/env python3
"""Converts the glTF scenes the reference ships (data/GLTFs) into npz fixtures under tests/golden/ with the loader
shim of hiprt-path-tracer_b200/scene.py, so that GPU-box runs (no /root/reference there) use the very same arrays."""
import importlib
import os
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
pt = importlib.import_module("hiprt-path-tracer_b200")
sc = importlib.import_module("hiprt-path-tracer_b200.scene")

for name in ("cornell_pbr", "nested-dielectrics", "nested-dielectrics-complex", "multi-dispersion"):
    s = pt.load_gltf(f"/root/reference/data/GLTFs/{name}.gltf")
    out = os.path.join(ROOT, "tests", "golden", f"{name}.npz")
    sc.save_scene(s, out)
    t = sc.load_scene(out)
    assert t.triangle_count == s.triangle_count and t.material_count == s.material_count
    print(name, s.triangle_count, "tris", s.material_count, "materials", len(s.emissive), "emissive", os.path.getsize(out), "bytes")
