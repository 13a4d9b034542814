"""The benchmark / parity configurations of BASELINE.json, as inputs: scene arrays, camera,
kernel options and run-time settings. The same description is applied to the CUDA renderer and to the oracle."""
import os

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


def apply_settings(render_settings, settings):
    for k, v in settings.items():
        setattr(render_settings, k, v)


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
