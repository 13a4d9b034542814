#!/usr/bin/env python3
"""Writes a workload's triangles for tools/bvh_host_sah.cpp: python tools/dump_scene.py synthetic /tmp/c3.bin"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
scene = getattr(wl, sys.argv[1])(1920, 1080)["scene"]
v = np.ascontiguousarray(scene.vertices, np.float32)
i = np.ascontiguousarray(scene.indices, np.int32).reshape(-1)
with open(sys.argv[2], "wb") as f:
    np.array([len(v), len(i) // 3], np.int32).tofile(f)
    v.tofile(f)
    i.tofile(f)
print(sys.argv[2], len(v), "vertices", len(i) // 3, "triangles")
