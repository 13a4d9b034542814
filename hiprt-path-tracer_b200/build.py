#!/usr/bin/env python3
"""Builds hiprt-path-tracer_b200/libb200pt.so for sm_100a with nvcc (cross-compiles without a GPU).

One translation unit per FullPathTracer permutation (csrc/variants.h) and per build flavour (timing / counting),
compiled in parallel; objects are cached under hiprt-path-tracer_b200/build/ and rebuilt when a source is newer.
"""
import os
import re
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libb200pt.so")
HOST_LIB = os.path.join(HERE, "libb200pt_host.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

# -fmad=false: shading arithmetic is evaluated exactly like the CPU oracle (g++ -ffp-contract=off); the traversal's
# box tests use explicit fused multiply-adds, its triangle test explicit unfused ones.
NVCC_FLAGS = ["-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false", "-Xfatbin", "-compress-all",
              "-Xcompiler", "-fPIC,-fopenmp,-fvisibility=hidden", "-I", os.path.join(HERE, "..", "include")]


def variants():
    text = open(os.path.join(CSRC, "variants.h")).read()
    return [tuple(int(x) for x in m) for m in re.findall(r"M\((\d+),\s*(\d+),\s*(\d+),\s*(\d+)\)", text.split("PT_FOR_EACH_VARIANT(M)")[1])]


def sources_mtime():
    newest = 0.0
    for d in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(d):
            newest = max(newest, os.path.getmtime(os.path.join(d, f)))
    return newest


def stale(obj):
    """An object is rebuilt when it, or its dependency file (nvcc -MD), is missing, or when a file it includes is newer."""
    dep = obj[:-2] + ".d"
    if not (os.path.exists(obj) and os.path.exists(dep)):
        return True
    built = os.path.getmtime(obj)
    text = open(dep).read().replace("\\\n", " ")
    deps = text.split(":", 1)[1].split() if ":" in text else []
    for d in deps:
        if d.startswith(("/usr/", "/opt/")):
            continue
        if not os.path.exists(d) or os.path.getmtime(d) > built:
            return True
    return False


def run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError("b200pt build failed: " + cmd[-1])
    return r.stdout + r.stderr


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    newest = sources_mtime()
    jobs = []   # (object, command)

    def add(obj, src, extra=()):
        o = os.path.join(OBJ, obj)
        if force or stale(o):
            jobs.append((o, [NVCC, *NVCC_FLAGS, *extra, "-MD", "-MF", o[:-2] + ".d", "-c", os.path.join(CSRC, src), "-o", o]))
        return o

    objects = [add("bvh8_build.o", "bvh8_build.cpp", ["-x", "cu"]), add("bvh8_build_device.o", "bvh8_build_device.cu"), add("capi.o", "b200pt_capi.cu"),
               add("bake.o", "b200pt_bake.cu")]
    for suffix, defs in (("t", []), ("c", ["-DPT_COUNTERS"])):
        objects.append(add(f"kernels_{suffix}.o", "b200pt_kernels.cu", defs))
        for (lss, ess, later, ggx) in variants():
            objects.append(add(f"fpt_{lss}_{ess}_{later}_{ggx}_{suffix}.o", "b200pt_fpt_variant.cu",
                               defs + [f"-DPT_LSS={lss}", f"-DPT_ESS={ess}", f"-DPT_LATER={later}", f"-DPT_GGX={ggx}"]))
            if lss in (3, 4, 5):
                # the phase kernels of the phased MIS pipeline, with the BSDF entry points inlined (see b200pt_phase_variant.cu)
                objects.append(add(f"phase_{lss}_{ess}_{later}_{ggx}_{suffix}.o", "b200pt_phase_variant.cu",
                                   defs + [f"-DPT_LSS={lss}", f"-DPT_ESS={ess}", f"-DPT_LATER={later}", f"-DPT_GGX={ggx}", "-DPT_BSDF_INLINE"]))

    # slowest translation units first (the phase kernels take ~2 minutes each)
    jobs.sort(key=lambda j: (0 if "/phase_" in j[0] else 1 if "/fpt_" in j[0] else 2))
    if jobs:
        with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
            for out in ex.map(lambda j: run(j[1]), jobs):
                if verbose and out.strip():
                    print(out)
    if jobs or not os.path.exists(LIB):
        run([NVCC, "-shared", "-o", LIB, *objects, "-Xcompiler", "-fopenmp", "-lgomp"])

    # Host layer (C++ mirror of the reference's GPURenderer) above the C ABI: plain g++, links against libb200pt.so
    host_dir = os.path.join(HERE, "host")
    host_src = [os.path.join(host_dir, f) for f in ("GPURenderer.cpp", "GPUBaker.cpp", "host_capi.cpp")]
    host_newest = max(os.path.getmtime(os.path.join(host_dir, f)) for f in os.listdir(host_dir))
    host_newest = max(host_newest, os.path.getmtime(LIB), newest)
    if force or not os.path.exists(HOST_LIB) or os.path.getmtime(HOST_LIB) < host_newest:
        run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-I", os.path.join(HERE, "..", "include"), "-I", host_dir,
             *host_src, "-o", HOST_LIB, "-L", HERE, "-lb200pt", "-pthread", "-Wl,-rpath,$ORIGIN"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
