#!/usr/bin/env python3
"""ORACLE / TEST INFRASTRUCTURE ONLY.

Build recipe for oracle/_ref/libhippt_ref_<variant>.so : the reference's own CPU kernels
(/root/reference/src/Device/**, src/Renderer/{BVH,Triangle}.cpp, src/Image/Image.cpp) compiled
*where they lie* with g++ -O2 -fopenmp -ffp-contract=off, driven by oracle/ref_harness.cpp.
Nothing of the reference is copied into the repository; all outputs go to oracle/_ref/
(git-ignored, but shipped to the GPU box as a prebuilt binary).

Kernel options (src/HostDeviceCommon/KernelOptions.h:85-368) are fixed #defines without #ifndef
guards on the CPU, so each variant gets a generated overlay header placed earlier on the -I path:
the overlay is produced at build time by re-writing the '#define <Option> <value>' lines of the
reference header and lives only under oracle/_ref/.
"""
import os
import re
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("HIPPT_REFERENCE", "/root/reference")

# name -> {kernel option macro: value}
VARIANTS = {
    # configs C1/C2: MIS light+BSDF, no envmap importance sampling needed (uniform sky) but alias table kept compiled in
    "mis": {"DirectLightSamplingStrategy": "LSS_MIS_LIGHT_BSDF", "EnvmapSamplingStrategy": "ESS_ALIAS_TABLE"},
    # config C3 second GGX sampler
    "mis_caps": {"DirectLightSamplingStrategy": "LSS_MIS_LIGHT_BSDF", "EnvmapSamplingStrategy": "ESS_ALIAS_TABLE",
                 "PrincipledBSDFAnisotropicGGXSampleFunction": "GGX_VNDF_SPHERICAL_CAPS"},
    # reference CPU default
    "ris": {"DirectLightSamplingStrategy": "LSS_RIS_BSDF_AND_LIGHT", "EnvmapSamplingStrategy": "ESS_ALIAS_TABLE"},
    # config C4
    "restir": {"DirectLightSamplingStrategy": "LSS_RESTIR_DI", "EnvmapSamplingStrategy": "ESS_NO_SAMPLING"},
    # the rest of the permutation matrix of SURVEY.md 8(b): spherical-cap GGX sampling under RIS / ReSTIR DI, ReSTIR DI with the alias table
    "ris_caps": {"DirectLightSamplingStrategy": "LSS_RIS_BSDF_AND_LIGHT", "EnvmapSamplingStrategy": "ESS_ALIAS_TABLE",
                 "PrincipledBSDFAnisotropicGGXSampleFunction": "GGX_VNDF_SPHERICAL_CAPS"},
    "restir_caps": {"DirectLightSamplingStrategy": "LSS_RESTIR_DI", "EnvmapSamplingStrategy": "ESS_NO_SAMPLING",
                    "PrincipledBSDFAnisotropicGGXSampleFunction": "GGX_VNDF_SPHERICAL_CAPS"},
    "restir_alias": {"DirectLightSamplingStrategy": "LSS_RESTIR_DI", "EnvmapSamplingStrategy": "ESS_ALIAS_TABLE"},
    "restir_alias_caps": {"DirectLightSamplingStrategy": "LSS_RESTIR_DI", "EnvmapSamplingStrategy": "ESS_ALIAS_TABLE",
                          "PrincipledBSDFAnisotropicGGXSampleFunction": "GGX_VNDF_SPHERICAL_CAPS"},
}

# bake oracle name -> the variant whose KernelOptions overlay it is compiled with (only PrincipledBSDFAnisotropicGGXSampleFunction matters there)
BAKE_VARIANTS = {"vndf": "mis", "caps": "mis_caps"}

CXXFLAGS = ["-std=c++20", "-O2", "-fopenmp", "-ffp-contract=off", "-fPIC", "-w"]


def includes(overlay=None):
    inc = []
    if overlay:
        inc.append(overlay)
    inc += [
        os.path.join(REF, "src"),
        os.path.join(REF, "thirdparties/HIPRT-Fork"),
        os.path.join(REF, "thirdparties/HIPRT-Fork/contrib/Orochi"),
        os.path.join(REF, "thirdparties/glm"),
        os.path.join(REF, "thirdparties/stbi"),
        os.path.join(REF, "thirdparties/imgui"),
        os.path.join(REF, "thirdparties/tinyexr"),
        HERE,
    ]
    return [f"-I{p}" for p in inc]


def run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise SystemExit(f"oracle build failed: {cmd[-1]}")


def make_overlay(variant, options):
    src = open(os.path.join(REF, "src/HostDeviceCommon/KernelOptions.h")).read()
    for macro, value in options.items():
        pat = re.compile(r"^(#define\s+" + re.escape(macro) + r")\s+\S+\s*$", re.M)
        if not pat.search(src):
            raise SystemExit(f"kernel option {macro} not found in the reference header")
        src = pat.sub(lambda m: f"{m.group(1)} {value}", src)
    d = os.path.join(OUT, f"overlay_{variant}", "HostDeviceCommon")
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, "KernelOptions.h")
    if not os.path.exists(path) or open(path).read() != src:
        open(path, "w").write(src)
    return os.path.dirname(d)


def newer(target, deps):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps if os.path.exists(d))


def build(variants=None, force=False):
    if not os.path.isdir(REF):
        print(f"[oracle] {REF} absent: using prebuilt files in {OUT} (if any)")
        return False
    os.makedirs(OUT, exist_ok=True)
    variants = variants or list(VARIANTS)

    # Objects shared by every variant (they do not depend on the options that the variants change)
    common = {
        "BVH.o": os.path.join(REF, "src/Renderer/BVH.cpp"),
        "Triangle.o": os.path.join(REF, "src/Renderer/Triangle.cpp"),
        "Image.o": os.path.join(REF, "src/Image/Image.cpp"),
        "ImGuiLogger.o": os.path.join(REF, "src/UI/ImGui/ImGuiLogger.cpp"),
        "imgui.o": os.path.join(REF, "thirdparties/imgui/imgui.cpp"),
        "imgui_draw.o": os.path.join(REF, "thirdparties/imgui/imgui_draw.cpp"),
        "imgui_tables.o": os.path.join(REF, "thirdparties/imgui/imgui_tables.cpp"),
        "imgui_widgets.o": os.path.join(REF, "thirdparties/imgui/imgui_widgets.cpp"),
        "stubs.o": os.path.join(HERE, "ref_stubs.cpp"),
    }
    jobs = []
    for obj, src in common.items():
        o = os.path.join(OUT, obj)
        if force or not newer(o, [src]):
            jobs.append(["g++", *CXXFLAGS, *includes(), "-c", src, "-o", o])

    harness_src = [os.path.join(HERE, "ref_harness.cpp"), os.path.join(HERE, "ref_harness_passes.inl")]
    harness_objs = {}
    for v in variants:
        overlay = make_overlay(v, VARIANTS[v])
        o = os.path.join(OUT, f"harness_{v}.o")
        harness_objs[v] = o
        if force or not newer(o, harness_src + [os.path.join(overlay, "HostDeviceCommon/KernelOptions.h")]):
            jobs.append(["g++", *CXXFLAGS, *includes(overlay), "-c", harness_src[0], "-o", o])

    # the reference's LUT bake kernels (src/Device/kernels/Baking/*.h) behind oracle/ref_bake.cpp, once per GGX sampling routine
    bake_src = os.path.join(HERE, "ref_bake.cpp")
    bake_objs = {}
    for name, v in BAKE_VARIANTS.items():
        overlay = make_overlay(v, VARIANTS[v])
        o = os.path.join(OUT, f"bake_{name}.o")
        bake_objs[name] = o
        if force or not newer(o, [bake_src, os.path.join(overlay, "HostDeviceCommon/KernelOptions.h")]):
            jobs.append(["g++", *CXXFLAGS, *includes(overlay), "-c", bake_src, "-o", o])

    # the reference's 8-bit texture loader behind oracle/ref_image.cpp
    image_src = os.path.join(HERE, "ref_image.cpp")
    image_obj = os.path.join(OUT, "ref_image.o")
    if force or not newer(image_obj, [image_src]):
        jobs.append(["g++", *CXXFLAGS, *includes(), "-c", image_src, "-o", image_obj])

    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        list(ex.map(run, jobs))

    wrap = "-Wl,--wrap=_ZNK3BVH9intersectERK8hiprtRayR7HitInfoPv"
    for v in variants:
        so = os.path.join(OUT, f"libhippt_ref_{v}.so")
        deps = [harness_objs[v]] + [os.path.join(OUT, o) for o in common]
        if force or not newer(so, deps):
            run(["g++", "-shared", "-fopenmp", wrap, "-o", so, *deps])
        print(f"[oracle] {so}")
    so = os.path.join(OUT, "libhippt_ref_image.so")
    deps = [image_obj] + [os.path.join(OUT, c) for c in common]
    if force or not newer(so, deps):
        run(["g++", "-shared", "-fopenmp", "-o", so, *deps])
    print(f"[oracle] {so}")
    for name, o in bake_objs.items():
        so = os.path.join(OUT, f"libhippt_ref_bake_{name}.so")
        deps = [o] + [os.path.join(OUT, c) for c in common]
        if force or not newer(so, deps):
            run(["g++", "-shared", "-fopenmp", "-o", so, *deps])
        print(f"[oracle] {so}")
    return True


if __name__ == "__main__":
    build([a for a in sys.argv[1:] if not a.startswith("--")] or None, force="--force" in sys.argv)
