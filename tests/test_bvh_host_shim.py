"""BVH8 builder + traversal logic against brute force, compiled for the host with an intrinsics shim (no GPU needed)."""
import os
import subprocess

import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
CSRC = os.path.join(ROOT, "hiprt-path-tracer_b200", "csrc")


@pytest.fixture(scope="module")
def shim_binary(tmp_path_factory):
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("CUDA headers not installed")
    out = str(tmp_path_factory.mktemp("shim") / "traverse_host_test")
    cmd = ["g++", "-std=c++17", "-O1", "-fopenmp", "-I", cuda_inc, "-I", CSRC, "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "host_shim", "traverse_host_test.cpp"), os.path.join(CSRC, "bvh8_build.cpp"), "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return out


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_traversal_equals_brute_force(shim_binary, mode):
    r = subprocess.run([shim_binary, str(mode)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.fixture(scope="module")
def queue_shim_binary(tmp_path_factory):
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("CUDA headers not installed")
    out = str(tmp_path_factory.mktemp("shim") / "trace_queue_host_test")
    cmd = ["g++", "-std=c++17", "-O1", "-fopenmp", "-I", cuda_inc, "-I", CSRC, "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "host_shim", "trace_queue_host_test.cpp"), os.path.join(CSRC, "bvh8_build.cpp"), "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return out


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_ray_queue_loop_flavours_equal_brute_force(queue_shim_binary, mode):
    """Both loops of the ray-queue traversal kernel (classic and software-pipelined, csrc/trace_queue.cuh) as a warp of one lane on
    the host: closest hits (id, t bit-equal, triangle slot) and occlusion flags against a brute-force loop over all triangles."""
    r = subprocess.run([queue_shim_binary, str(mode)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.fixture(scope="module")
def lbvh_shim_binary(tmp_path_factory):
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("CUDA headers not installed")
    out = str(tmp_path_factory.mktemp("shim") / "lbvh_host_test")
    cmd = ["g++", "-std=c++17", "-O1", "-I", cuda_inc, "-I", CSRC, "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "host_shim", "lbvh_host_test.cpp"), "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return out


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4, 10, 11, 12, 13, 14])
def test_device_builder_steps_equal_brute_force(lbvh_shim_binary, mode):
    """The per-element functions of the DEVICE BVH8 builder (Morton codes, Karras radix tree or PLOC clustering, bottom-up boxes + collapse
    costs, emission; csrc/bvh8_build_device.cuh) in serial loops on the host: every triangle emitted once, traversal equals brute force, depth
    fits the stack. Modes: soup, height field, one triangle, duplicated axis-aligned triangles (equal Morton codes, flat boxes), two triangles;
    10 + mode: the same scenes with the binary topology from PLOC (every iteration merges, the last merge is node 0, the depth-first
    re-ordering of the triangles is a permutation)."""
    r = subprocess.run([lbvh_shim_binary, str(mode)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr



# ---- the same three shims under AddressSanitizer + UndefinedBehaviorSanitizer -------------------------------------------------------
# compute-sanitizer is closed on the GPU pool of this project (it answers "closed on this pool and stays closed"), so out-of-bounds
# and use-of-uninitialised checks of the traversal stacks, the ray-queue loops and the device builder's per-element functions run
# here, on the host compilation of the very same headers (csrc/traverse.cuh, trace_queue.cuh, bvh8_build_device.cuh).
SANITIZE = ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer", "-g"]


@pytest.fixture(scope="module")
def sanitized_shims(tmp_path_factory):
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("CUDA headers not installed")
    probe = tmp_path_factory.mktemp("asan") / "probe.cpp"
    probe.write_text("int main() { return 0; }\n")
    if subprocess.run(["g++", *SANITIZE, str(probe), "-o", str(probe) + ".out"], capture_output=True).returncode != 0:
        pytest.skip("g++ has no AddressSanitizer runtime here")
    out = {}
    for name, extra in (("traverse_host_test", [os.path.join(CSRC, "bvh8_build.cpp")]), ("trace_queue_host_test", [os.path.join(CSRC, "bvh8_build.cpp")]), ("lbvh_host_test", [])):
        binary = str(tmp_path_factory.mktemp("asan") / name)
        cmd = ["g++", "-std=c++17", "-O1", "-fopenmp", *SANITIZE, "-I", cuda_inc, "-I", CSRC, "-I", os.path.join(ROOT, "include"),
               os.path.join(ROOT, "tests", "host_shim", name + ".cpp"), *extra, "-o", binary]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-3000:]
        out[name] = binary
    return out


@pytest.mark.parametrize("name,mode", [("traverse_host_test", 0), ("traverse_host_test", 3), ("trace_queue_host_test", 0), ("trace_queue_host_test", 1), ("lbvh_host_test", 0),
                                       ("lbvh_host_test", 3), ("lbvh_host_test", 4), ("lbvh_host_test", 10), ("lbvh_host_test", 13)])
def test_host_shims_are_clean_under_asan_and_ubsan(sanitized_shims, name, mode):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1", OMP_NUM_THREADS="4")
    r = subprocess.run([sanitized_shims[name], str(mode)], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0 and "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr, (r.stdout[-1500:] + r.stderr[-3000:])
