"""The benchmark / parity configurations of BASELINE.json, as inputs: scene arrays, camera,
kernel options and run-time settings. The same description is applied to the CUDA renderer and to the oracle."""
import os

from . import procedural
from . import scene as sc
from .luts import load_brdf_luts

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def cornell(width, height):
    """C1 / C2: data/GLTFs/cornell_pbr.gltf (8096 triangles, 9 materials), uniform sky 0.5, 8 bounces, MIS light+BSDF,
    alpha testing off (RNG-stream parity), adaptive sampling off."""
    scene = sc.load_scene(os.path.join(ROOT, "tests", "golden", "cornell_pbr.npz"))
    return {
        "name": f"cornell_pbr {width}x{height}, 8 bounces, LSS_MIS_LIGHT_BSDF, uniform sky",
        "scene": scene, "camera": sc.scene_camera(scene, width, height), "width": width, "height": height,
        "options": dict(lss="mis", ess="none", ggx=0, later=3), "oracle_variant": "mis", "envmap": None,
        "settings": dict(nb_bounces=8, do_alpha_testing=0, enable_adaptive_sampling=0, samples_per_frame=1),
    }


def golden_scene(name, width, height, lss="mis", adaptive=False, thin_film=False, nb_bounces=8, alpha_opacity=None):
    """C5-shaped parity inputs: the reference's own test scenes (tests/golden/<name>.npz: nested-dielectrics,
    nested-dielectrics-complex, multi-dispersion, cornell_pbr), optionally with per-pixel adaptive sampling and with a
    thin-film coating put on every non-emissive material. No emissive geometry -> uniform sky 0.5 lights the scene."""
    scene = sc.load_scene(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    if thin_film:
        for i, m in enumerate(scene.materials):
            if i not in set(int(scene.material_indices[t]) for t in scene.emissive):
                m.base.thin_film = 1.0
                m.base.thin_film_ior = 1.3 + 0.05 * (i % 4)
                m.base.thin_film_thickness = 300.0 + 90.0 * i
    if alpha_opacity is not None:
        # stochastic alpha testing (filter_function): every second non-emissive material lets part of the rays through
        emissive_materials = set(int(scene.material_indices[t]) for t in scene.emissive)
        for i, m in enumerate(scene.materials):
            if i not in emissive_materials and i % 2 == 0:
                m.base.alpha_opacity = alpha_opacity
    settings = dict(nb_bounces=nb_bounces, do_alpha_testing=int(alpha_opacity is not None), enable_adaptive_sampling=int(adaptive), samples_per_frame=1)
    if adaptive:
        settings.update(adaptive_sampling_min_samples=4, adaptive_sampling_noise_threshold=1.0)
    return {
        "name": f"{name} {width}x{height}, {nb_bounces} bounces, {lss}" + (", adaptive sampling" if adaptive else "") + (", thin film" if thin_film else ""),
        "scene": scene, "camera": sc.scene_camera(scene, width, height), "width": width, "height": height,
        "options": dict(lss=lss, ess="none", ggx=0, later=3), "oracle_variant": lss, "envmap": None, "settings": settings,
    }


def adaptive(width, height):
    """C5: per-pixel adaptive sampling on the reference's nested-dielectrics-complex scene with a thin-film coating on every
    non-emissive material and spectral dispersion switched on for the transmissive ones (4K in BASELINE.json: bench.py
    --workload adaptive --width 3840 --height 2160; N GPUs shard sample passes and all-reduce the accumulation, squared-luminance
    and sample-count buffers)."""
    cfg = golden_scene("nested-dielectrics-complex", width, height, adaptive=True, thin_film=True)
    for m in cfg["scene"].materials:
        if m.base.specular_transmission > 0.0 and m.base.ior > 1.05:     # dispersion around an index of 1 makes the reference produce NaN samples
            m.base.dispersion_scale = 1.0
            m.base.dispersion_abbe_number = 20.0
    cfg["name"] = f"nested-dielectrics-complex {width}x{height}, 8 bounces, LSS_MIS_LIGHT_BSDF, adaptive sampling (min 4 samples, threshold 1.0), thin film + dispersion"
    return cfg


def merge_scenes(a, b, offset_b=(0.0, 0.0, 0.0)):
    """a UNION b (b translated by offset_b): vertex, material and emissive indices of b shifted; a's camera kept."""
    import numpy as np
    out = sc.Scene()
    nv, nm = len(a.vertices), len(a.materials)
    out.vertices = np.concatenate([a.vertices, (b.vertices + np.asarray(offset_b, np.float32)[None, :]).astype(np.float32)])
    out.indices = np.concatenate([a.indices, (b.indices + nv).astype(np.int32)])
    out.normals = np.concatenate([a.normals, b.normals])
    out.has_normals = np.concatenate([a.has_normals, b.has_normals])
    out.texcoords = np.concatenate([a.texcoords, b.texcoords])
    out.material_indices = np.concatenate([a.material_indices, (b.material_indices + nm).astype(np.int32)])
    out.materials = list(a.materials) + list(b.materials)
    out.textures = list(a.textures) + list(b.textures)
    out.camera = a.camera
    out.name = f"{a.name}+{b.name}"
    out.compute_emissive_triangles()
    return out


# C5's fixed table of the material fields a glTF cannot carry: per transmissive material, in scene order, cycling
C5_PRIORITIES = (1, 2, 3)
C5_ABSORPTION = ((0.9, 0.3, 0.3), (0.3, 0.9, 0.4), (0.35, 0.45, 0.95), (0.95, 0.85, 0.3))


def adaptive_full(width, height, envmap_size=(2048, 1024), thin_film_on_glass=True):
    """C5 as SURVEY.md 8(d) specifies it: nested-dielectrics-complex (3088 triangles) UNION multi-dispersion (9002 triangles, set beside it),
    the procedural HDR envmap of C3 sampled through its alias table with BSDF MIS, and on the transmissive materials the fields the glTF
    cannot carry, from a fixed table: dielectric_priority 1..3 (ISS_WITH_PRIORITIES, stack size 3), absorption colour at distance 1,
    thin film (thickness 500 nm, ior 1.3), dispersion (scale 1, Abbe number 20); per-pixel adaptive sampling at the reference's defaults
    (min 64 samples, noise threshold 0.3, RenderSettings.h:126-133). 4K in BASELINE.json: bench.py --workload adaptive_full --width 3840
    --height 2160. thin_film_on_glass=False leaves the coating off the transmissive materials: thin film on glass under total internal
    reflection is undefined behaviour in the reference, so parity is pinned without it and bounded with it."""
    key = ("adaptive_full", envmap_size)
    if key not in _CACHE:
        a = sc.load_scene(os.path.join(ROOT, "tests", "golden", "nested-dielectrics-complex.npz"))
        b = sc.load_scene(os.path.join(ROOT, "tests", "golden", "multi-dispersion.npz"))
        extent = a.vertices.max(axis=0) - a.vertices.min(axis=0)
        shift = (float(a.vertices.max(axis=0)[0] - b.vertices.min(axis=0)[0] + 0.05 * extent[0]), float(a.vertices.min(axis=0)[1] - b.vertices.min(axis=0)[1]), 0.0)
        _CACHE[key] = (merge_scenes(a, b, shift), procedural.procedural_envmap(*envmap_size))
    scene, env = _CACHE[key]
    import copy
    scene = copy.copy(scene)
    scene.materials = [copy.deepcopy(m) for m in scene.materials]
    k = 0
    for m in scene.materials:
        if m.base.specular_transmission > 0.0:
            m.base.dielectric_priority = C5_PRIORITIES[k % len(C5_PRIORITIES)]
            c = C5_ABSORPTION[k % len(C5_ABSORPTION)]
            m.base.absorption_color.r, m.base.absorption_color.g, m.base.absorption_color.b = c
            m.base.absorption_at_distance = 1.0
            if m.base.ior > 1.05:            # dispersion around an index of 1 makes the reference produce NaN samples
                m.base.dispersion_scale = 1.0
                m.base.dispersion_abbe_number = 20.0
            if thin_film_on_glass:
                m.base.thin_film, m.base.thin_film_thickness, m.base.thin_film_ior = 1.0, 500.0, 1.3
            k += 1
        elif not sc.material_is_emissive(m):
            m.base.thin_film, m.base.thin_film_thickness, m.base.thin_film_ior = 1.0, 500.0, 1.3
    return {
        "name": f"nested-dielectrics-complex + multi-dispersion ({scene.triangle_count} triangles) {width}x{height}, 8 bounces, LSS_MIS_LIGHT_BSDF, envmap {envmap_size[0]}x{envmap_size[1]} alias table + BSDF MIS, "
                f"priorities 1..3 + absorption + dispersion + thin film{'' if thin_film_on_glass else ' (not on glass)'}, adaptive sampling (min 64 samples, threshold 0.3)",
        "scene": scene, "camera": sc.scene_camera(scene, width, height), "width": width, "height": height,
        "options": dict(lss="mis", ess="alias_table", ggx=0, later=3), "oracle_variant": "mis",
        "envmap": {"texels": env, "alias_probas": None, "alias_alias": None, "total_sum": 0.0},
        "settings": dict(nb_bounces=8, do_alpha_testing=0, enable_adaptive_sampling=1, adaptive_sampling_min_samples=64, adaptive_sampling_noise_threshold=0.3, samples_per_frame=1),
    }


_CACHE = {}


def synthetic(width, height, quads=1581, spheres=3906, subdivisions=3, ggx=0, envmap_size=(2048, 1024)):
    """C3: procedural ~10 M-triangle terrain + spheres (9 998 802 triangles at the defaults), 64 random principled materials,
    procedural HDR envmap sampled through its alias table with BSDF MIS, no emissive geometry."""
    key = ("synthetic", quads, spheres, subdivisions, envmap_size)
    if key not in _CACHE:
        _CACHE[key] = (procedural.terrain_and_spheres(quads, spheres, subdivisions), procedural.procedural_envmap(*envmap_size))
    scene, env = _CACHE[key]
    return {
        "name": f"synthetic terrain+spheres {scene.triangle_count} triangles {width}x{height}, 8 bounces, LSS_MIS_LIGHT_BSDF, envmap {envmap_size[0]}x{envmap_size[1]} alias table + BSDF MIS",
        "scene": scene, "camera": sc.scene_camera(scene, width, height), "width": width, "height": height,
        "options": dict(lss="mis", ess="alias_table", ggx=ggx, later=3), "oracle_variant": "mis" if ggx == 0 else "mis_caps",
        "envmap": {"texels": env, "alias_probas": None, "alias_alias": None, "total_sum": 0.0},
        "settings": dict(nb_bounces=8, do_alpha_testing=0, enable_adaptive_sampling=0, samples_per_frame=1),
    }


def synthetic_small(width, height, ggx=0):
    """A parity-sized cut of C3 (~37 k triangles, 256x128 envmap) that the CPU oracle renders in seconds."""
    cfg = synthetic(width, height, quads=96, spheres=60, subdivisions=2, ggx=ggx, envmap_size=(256, 128))
    return cfg


def many_lights(width, height, boxes=2000, lights=100_000, fused=True, temporal=True, spatial=True, spatial_passes=1):
    """C4: closed room + diffuse boxes + 100 000 small emissive triangles; ReSTIR DI with light presampling and the fused
    spatiotemporal pass, RIS on later bounces. fused=False runs the stand-alone TemporalReuse / SpatialReuse passes
    (ReSTIRDIRenderPass.cpp:332-474); spatial_passes > 1 adds SpatialReuse passes (after the fused pass: GPU seed schedule only)."""
    key = ("many_lights", boxes, lights)
    if key not in _CACHE:
        _CACHE[key] = procedural.many_lights_room(boxes, lights)
    scene = _CACHE[key]
    return {
        "name": f"many-light room {scene.triangle_count} triangles ({len(scene.emissive)} emissive) {width}x{height}, 8 bounces, LSS_RESTIR_DI (presampled lights, "
                + ("fused spatiotemporal" if fused else "+".join(n for n, on in (("temporal", temporal), ("spatial", spatial)) if on) + " reuse passes")
                + (f", {spatial_passes} spatial passes" if spatial_passes > 1 else "") + "), later bounces RIS",
        "scene": scene, "camera": sc.scene_camera(scene, width, height), "width": width, "height": height,
        "options": dict(lss="restir_di", ess="none", ggx=0, later=3), "oracle_variant": "restir", "envmap": None,
        "settings": {"nb_bounces": 8, "do_alpha_testing": 0, "enable_adaptive_sampling": 0, "samples_per_frame": 1,
                     "restir_di_settings.do_fused_spatiotemporal": int(fused), "restir_di_settings.temporal_pass.do_temporal_reuse_pass": int(temporal),
                     "restir_di_settings.spatial_pass.do_spatial_reuse_pass": int(spatial), "restir_di_settings.spatial_pass.number_of_passes": spatial_passes},
    }


def apply_settings(render_settings, settings):
    """Keys may be dotted to reach nested settings, e.g. "restir_di_settings.spatial_pass.reuse_neighbor_count"."""
    for k, v in settings.items():
        target = render_settings
        *parents, leaf = k.split(".")
        for name in parents:
            target = getattr(target, name)
        setattr(target, leaf, v)


def setup_gpu(pt, ctx, cfg, seed_schedule="gpu"):
    g = pt.GPURenderer(ctx, options=cfg["options"], seed_schedule=seed_schedule)
    g.set_brdf_luts(load_brdf_luts())
    g.set_scene(cfg["scene"])
    if cfg["envmap"] is not None:
        e = cfg["envmap"]
        g.set_envmap(e["texels"], e["alias_probas"], e["alias_alias"], e["total_sum"])
    g.set_camera(cfg["camera"])
    apply_settings(g.render_settings, cfg["settings"])
    g.resize(cfg["width"], cfg["height"])
    return g


def setup_oracle(ref, structs, cfg):
    """ORACLE: only tests/, smoke() and bench.py's cpu_baseline / --impl reference legs call this."""
    r = ref.RefRenderer(cfg["oracle_variant"])
    r.set_luts(load_brdf_luts())
    r.set_scene(cfg["scene"])
    if cfg["envmap"] is not None:
        r.set_envmap(cfg["envmap"]["texels"])
    r.set_camera(cfg["camera"])
    apply_settings(r.struct("render_settings", structs.HIPRTRenderSettings), cfg["settings"])
    r.resize(cfg["width"], cfg["height"])
    return r
