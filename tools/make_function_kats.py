#!/usr/bin/env python3
"""Per-function known-answer vectors generated from the reference's own CPU kernels (oracle/_ref): bsdf_dispatcher_eval /
_sample with both GGX samplers on the 64 random principled materials of the C3 generator (+ thin-film and dispersion variants), envmap_sample
through the alias table of a procedural HDR, thin_film_fresnel. Writes tests/golden/function_kats.npz (inputs AND the reference's outputs):
tests/test_oracle_golden.py re-derives the outputs from the oracle, tests/test_gpu_function_kats.py checks the CUDA device functions
(b200pt_debug_*) against them."""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
pt = importlib.import_module("hiprt-path-tracer_b200")
wl = importlib.import_module("hiprt-path-tracer_b200.workloads")

OUT = os.path.join(ROOT, "tests", "golden", "function_kats.npz")


def kat_config():
    """The scene whose materials / envmap / LUTs the vectors are taken on (small cut of C3: 64 random principled materials + 256x128 envmap)."""
    cfg = wl.synthetic_small(64, 48)
    scene = cfg["scene"]
    import copy
    scene = copy.copy(scene)
    scene.materials = [copy.deepcopy(m) for m in scene.materials]
    for i, m in enumerate(scene.materials):
        if i % 4 == 1:
            m.base.thin_film, m.base.thin_film_thickness, m.base.thin_film_ior = 1.0, 300.0 + 40.0 * i, 1.3 + 0.01 * (i % 7)
        if m.base.specular_transmission > 0.0 and m.base.ior > 1.05 and i % 2 == 0:
            m.base.dispersion_scale, m.base.dispersion_abbe_number = 1.0, 20.0
    cfg["scene"] = scene
    return cfg


def kat_inputs(ref, cfg, seed=20261105):
    """Shading inputs as the path tracer itself produces them: the G-buffer of one CameraRays pass of the reference over the KAT scene
    (per hit pixel: 256-byte material, 36-byte volume state after trace_ray's interior-stack push, shading / geometric normal, view direction),
    plus seeded light directions, generator seeds and thin-film arguments. Arbitrary volume states are NOT valid inputs: with the default
    state (incident material -1) the reference indexes its material buffer out of bounds."""
    r = wl.setup_oracle(ref, pt.structs, dict(cfg, oracle_variant="mis"))
    r.render(1)
    hit = r.buffer("camera_ray_hit").astype(bool).reshape(-1)
    pick = np.nonzero(hit)[0]
    rng = np.random.default_rng(seed)
    n = len(pick)
    normal = r.buffer("shading_normals").reshape(-1, 3)[pick]
    geometric = r.buffer("geometric_normals").reshape(-1, 3)[pick]
    view = r.buffer("view_directions").reshape(-1, 3)[pick]
    light = rng.normal(size=(n, 3))
    light /= np.linalg.norm(light, axis=1, keepdims=True)
    light = np.where((rng.random(n) < 0.7)[:, None] & (np.sum(light * normal, axis=1) < 0)[:, None], -light, light)      # mostly above the surface
    volume_states = np.ascontiguousarray(r.buffer("ray_volume_states").reshape(-1, 36)[pick])
    pad = pt.structs.RayVolumeState._pad0
    volume_states[:, pad.offset:pad.offset + pad.size] = 0          # struct padding: uninitialised in the reference, read by nobody
    return {
        "materials": np.ascontiguousarray(r.buffer("materials").reshape(-1, 256)[pick]),
        "volume_states": volume_states,
        "eval_vecs": np.concatenate([view, normal, light], axis=1).astype(np.float32),
        "sample_vecs": np.concatenate([view, normal, geometric], axis=1).astype(np.float32),
        "seeds": rng.integers(1, 2**32 - 1, n, dtype=np.uint64).astype(np.uint32),
        "film_args": np.stack([rng.choice([1.0, 1.33, 1.5], n), rng.uniform(0.01, 1.0, n)], axis=1).astype(np.float32),
    }


def crash_safe_shading(variant, n):
    """bsdf eval / sample of items 0..n-1 by the reference, in fresh worker processes (this script with --worker): the reference's shading
    code has undefined behaviour on some inputs (uninitialised locals, unguarded table lookups: DESIGN.md section 7) and can take the
    process down. The worker appends every finished item to a progress file; when it dies, the item it was on is marked invalid and a
    new worker resumes after it. (A forked child is not an option: the reference's results differ after fork().)"""
    import subprocess
    import tempfile
    tmp = tempfile.mkdtemp(prefix="kats_")
    progress = os.path.join(tmp, "progress.bin")
    valid = np.ones(n, bool)
    start = 0
    rows = {}
    while start < n:
        open(progress, "wb").close()
        subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", variant, str(start), str(n), progress], capture_output=True)
        raw = np.fromfile(progress, dtype=np.float32)
        per = 1 + 4 + 7 + 1
        done = len(raw) // per
        for k in range(done):
            row = raw[k * per:(k + 1) * per]
            rows[int(row[0])] = row[1:]
        if start + done < n:
            valid[start + done] = False           # the item the worker died on
            start = start + done + 1
        else:
            start = n
    ev, sa, sd = np.zeros((n, 4), np.float32), np.zeros((n, 7), np.float32), np.zeros(n, np.uint32)
    for i, row in rows.items():
        ev[i], sa[i], sd[i] = row[0:4], row[4:11], row[11:12].view(np.uint32)[0]
    return {"eval": ev, "sample": sa, "sample_seeds": sd}, valid


def worker(variant, start, n, progress):
    from oracle import ref
    cfg = kat_config()
    inputs = kat_inputs(ref, cfg)
    r = oracle_state(ref, cfg, variant)
    per_item, vs = inputs["materials"], inputs["volume_states"]
    with open(progress, "ab", buffering=0) as f:
        for i in range(start, n):
            ev, _ = r.bsdf_eval(per_item[i:i + 1], vs[i:i + 1], inputs["eval_vecs"][i:i + 1])
            sa, _, sd = r.bsdf_sample(per_item[i:i + 1], vs[i:i + 1], inputs["sample_vecs"][i:i + 1], inputs["seeds"][i:i + 1])
            row = np.concatenate([np.array([i], np.float32), ev[0], sa[0], sd.view(np.float32)]).astype(np.float32)
            f.write(row.tobytes())


def oracle_state(ref, cfg, variant):
    r = wl.setup_oracle(ref, pt.structs, dict(cfg, oracle_variant=variant))
    r.render(1)          # configures the render data exactly as a frame does
    return r


def oracle_outputs(ref, cfg, inputs, variant):
    """The reference's answers: dict of arrays. variant 'mis' (VNDF) or 'mis_caps' (spherical caps) decides the sampler of bsdf_sample."""
    r = oracle_state(ref, cfg, variant)
    lib = r.lib
    n = len(inputs["seeds"])
    per_item = inputs["materials"]
    out = {}
    got, valid = crash_safe_shading(variant, n)
    out.update(got)
    out["valid"] = valid
    seeds = inputs["seeds"].copy()
    env = np.empty((n, 8), np.float32)
    lib.ref_envmap_sample(r.h, seeds.ctypes.data_as(C.c_void_p), C.c_longlong(n), env.ctypes.data_as(C.c_void_p))
    out["envmap"], out["envmap_seeds"] = env, seeds
    film = np.empty((n, 3), np.float32)
    for i in range(n):
        lib.ref_thin_film_fresnel(per_item[i].ctypes.data_as(C.c_void_p), C.c_float(float(inputs["film_args"][i, 0])), C.c_float(float(inputs["film_args"][i, 1])), film[i].ctypes.data_as(C.c_void_p))
    out["film"] = film
    return out


def main():
    from oracle import ref
    cfg = kat_config()
    inputs = kat_inputs(ref, cfg)
    vndf = oracle_outputs(ref, cfg, inputs, "mis")
    caps = oracle_outputs(ref, cfg, inputs, "mis_caps")
    valid = vndf["valid"] & caps["valid"]
    assert np.array_equal(vndf["eval"][valid], caps["eval"][valid], equal_nan=True), "bsdf eval does not depend on the GGX sampler"
    print("items:", len(valid), "; items on which the reference itself crashed (excluded):", int((~valid).sum()))
    np.savez_compressed(OUT, **inputs, valid=valid, eval=vndf["eval"], sample_vndf=vndf["sample"], sample_vndf_seeds=vndf["sample_seeds"],
                        sample_caps=caps["sample"], sample_caps_seeds=caps["sample_seeds"], envmap=vndf["envmap"], envmap_seeds=vndf["envmap_seeds"], film=vndf["film"])
    f = vndf["eval"]
    print("function_kats:", OUT, os.path.getsize(OUT) // 1024, "KiB; eval finite on", int(np.isfinite(f).all(axis=1).sum()), "pdf > 0 on", int((f[:, 3] > 0).sum()), "of", len(f),
          "; sampled pdf > 0 on", int((vndf["sample"][:, 3] > 0).sum()), "; samplers differ on", int((vndf["sample"] != caps["sample"]).any(axis=1).sum()))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--worker":
        worker(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5])
    else:
        main()
