"""ORACLE / TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/_ref/libhippt_ref_<variant>.so (the reference's own
CPU kernels behind the C API of oracle/ref_harness.cpp). Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product path never does."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def lib_path(variant):
    return os.path.join(REF_DIR, f"libhippt_ref_{variant}.so")


def available(variant="mis"):
    return os.path.exists(lib_path(variant))


_libs = {}


def load(variant="mis"):
    if variant not in _libs:
        path = lib_path(variant)
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `python oracle/build_ref.py {variant}` where /root/reference exists")
        l = C.CDLL(path)
        l.ref_variant.restype = C.c_char_p
        l.ref_create.restype = C.c_void_p
        for name in ("ref_render_data", "ref_render_settings", "ref_world_settings", "ref_bsdfs_data"):
            getattr(l, name).restype = C.c_void_p
            getattr(l, name).argtypes = [C.c_void_p]
        l.ref_set_scene.restype = C.c_double
        l.ref_render.restype = C.c_double
        l.ref_ray_count.restype = C.c_ulonglong
        l.ref_get_buffer.restype = C.c_longlong
        l.ref_ray_log_end.restype = C.c_longlong
        l.ref_glass_correction_exponent.restype = C.c_float
        l.ref_glass_correction_exponent.argtypes = [C.c_void_p, C.c_float, C.c_float]
        _libs[variant] = l
    return _libs[variant]


def image_lib_path():
    return os.path.join(REF_DIR, "libhippt_ref_image.so")


def read_image_8bit(path, output_channels):
    """Image8Bit::read_image(path, output_channels, flipY=false) of the reference (stb_image), oracle/ref_image.cpp.
    Returns (texels[h, w, c] uint8, is_constant_color(5), sample_rgba32f((0, 0)))."""
    if "image" not in _libs:
        if not os.path.exists(image_lib_path()):
            raise FileNotFoundError(f"{image_lib_path()} missing: run `python oracle/build_ref.py` where /root/reference exists")
        _libs["image"] = C.CDLL(image_lib_path())
    l = _libs["image"]
    w, h, const = C.c_int(), C.c_int(), C.c_int()
    rgba = (C.c_float * 4)()
    if l.ref_read_image_8bit(path.encode(), output_channels, None, C.c_longlong(0), C.byref(w), C.byref(h), C.byref(const), rgba) != 0:
        raise IOError(f"the reference's loader cannot read {path}")
    out = np.empty((h.value, w.value, output_channels), np.uint8)
    rc = l.ref_read_image_8bit(path.encode(), output_channels, out.ctypes.data_as(C.c_void_p), C.c_longlong(out.size), C.byref(w), C.byref(h), None, None)
    assert rc == 0
    return out, bool(const.value), np.array(list(rgba), np.float32)


def bake_lib_path(sampler="vndf"):
    return os.path.join(REF_DIR, f"libhippt_ref_bake_{sampler}.so")


def bake(kind, size_cos_theta, size_roughness, size_ior, sample_count, sampler="vndf", z_range=None, plan_only=False):
    """The reference's bake kernel `kind` (0 conductor, 1 Fresnel, 2 glossy dielectric, 3 glass entering, 4 glass exiting, 5 thin glass)
    driven like GPUBakerKernel::bake_internal (oracle/ref_bake.cpp). Returns (table[z, y, x], iterations_per_kernel, nb_kernel_launch);
    z_range = (begin, end) integrates only those ior slices (the rest of the table stays zero), plan_only returns table = None."""
    key = "bake_" + sampler
    if key not in _libs:
        path = bake_lib_path(sampler)
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `python oracle/build_ref.py` where /root/reference exists")
        _libs[key] = C.CDLL(path)
    sz = 1 if kind == 0 else size_ior
    out = None if plan_only else np.zeros((sz, size_roughness, size_cos_theta), np.float32)
    ipk, launches = C.c_int(), C.c_int()
    z0, z1 = z_range if z_range is not None else (0, sz)
    rc = _libs[key].ref_bake(kind, size_cos_theta, size_roughness, sz, sample_count, z0, z1, out.ctypes.data_as(C.c_void_p) if out is not None else None,
                             C.byref(ipk), C.byref(launches))
    if rc != 0:
        raise ValueError("ref_bake rejected its arguments")
    return out, ipk.value, launches.value


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


BUFFERS = {
    "first_hit_prim_index": (0, np.int32, 1), "shading_normals": (1, np.float32, 3), "geometric_normals": (2, np.float32, 3),
    "view_directions": (3, np.float32, 3), "first_hits": (4, np.float32, 3), "camera_ray_hit": (5, np.uint8, 1),
    "materials": (6, np.uint8, 256), "ray_volume_states": (7, np.uint8, 36), "pixel_active": (8, np.uint8, 1),
    "pixel_sample_count": (9, np.int32, 1), "pixel_squared_luminance": (10, np.float32, 1), "pixel_converged_sample_count": (11, np.int32, 1),
    "denoiser_albedo": (12, np.float32, 3), "denoiser_normals": (13, np.float32, 3), "restir_output_reservoirs": (14, np.uint8, 36),
    "restir_initial_candidates": (15, np.uint8, 36), "presampled_lights": (16, np.uint8, 48),
}


class RefRenderer:
    """The reference's CPURenderer (src/Renderer/CPURenderer.cpp) as restated by oracle/ref_harness.cpp."""

    def __init__(self, variant="mis"):
        self.lib = load(variant)
        self.variant = variant
        self.h = C.c_void_p(self.lib.ref_create())
        self._keep = []
        self.width = self.height = 0

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def struct(self, name, ctype):
        """Live view on the harness' HIPRTRenderSettings / WorldSettings / BRDFsData / HIPRTRenderData."""
        addr = getattr(self.lib, f"ref_{name}")(self.h)
        return ctype.from_address(addr)

    def set_luts(self, luts):
        arrs = [np.ascontiguousarray(luts[k], dtype=np.float32) for k in ("sheen_ltc", "conductor", "glossy", "glass", "glass_inverse", "thin_glass")]
        self.lib.ref_set_luts(self.h, *[_p(a) for a in arrs])

    def set_scene(self, scene):
        for tex in scene.textures:
            t = np.ascontiguousarray(tex, dtype=np.uint8)
            self.lib.ref_add_texture(self.h, _p(t), C.c_int(t.shape[1]), C.c_int(t.shape[0]), C.c_int(t.shape[2]))
        a = scene.arrays()
        return self.lib.ref_set_scene(self.h, _p(a["vertices"]), C.c_int(len(a["vertices"])), _p(a["indices"]), C.c_int(len(a["indices"]) // 3),
                                      _p(a["normals"]), _p(a["has_normals"]), _p(a["texcoords"]), _p(a["material_indices"]), _p(a["materials"]),
                                      C.c_int(scene.material_count), _p(a["emissive"]), C.c_int(len(a["emissive"])))

    def set_envmap(self, texels):
        texels = np.ascontiguousarray(texels, dtype=np.float32)
        h, w, ch = texels.shape
        self.lib.ref_set_envmap(self.h, _p(texels), C.c_int(w), C.c_int(h), C.c_int(ch))
        probas, alias, total = np.empty(w * h, np.float32), np.empty(w * h, np.int32), C.c_float()
        self.lib.ref_get_alias_table(self.h, _p(probas), _p(alias), C.byref(total))
        return probas, alias, total.value

    def set_camera(self, camera):
        self.lib.ref_set_camera(self.h, C.byref(camera))

    def resize(self, width, height):
        self.lib.ref_resize(self.h, C.c_int(width), C.c_int(height))
        self.width, self.height = width, height

    def render(self, nsamples, passes=7):
        return self.lib.ref_render(self.h, C.c_int(nsamples), C.c_int(passes))

    def pixels(self):
        out = np.empty((self.height, self.width, 3), np.float32)
        self.lib.ref_get_pixels(self.h, _p(out))
        return out

    def tonemap(self, gamma=2.2, exposure=1.0):
        """CPURenderer::tonemap + the 8-bit quantisation of write_image_png (ref_harness_passes.inl: ref_tonemap)."""
        f = np.empty((self.height, self.width, 3), np.float32)
        u = np.empty((self.height, self.width, 3), np.uint8)
        self.lib.ref_tonemap(self.h, C.c_float(gamma), C.c_float(exposure), _p(f), _p(u))
        return f, u

    def buffer(self, name):
        which, dtype, comps = BUFFERS[name]
        n = -self.lib.ref_get_buffer(self.h, C.c_int(which), None, C.c_longlong(0))
        out = np.empty(n // np.dtype(dtype).itemsize, dtype=dtype)
        self.lib.ref_get_buffer(self.h, C.c_int(which), _p(out), C.c_longlong(out.nbytes))
        if name == "presampled_lights":
            return out.reshape(-1, comps)
        return out.reshape(self.height, self.width, comps) if comps > 1 else out.reshape(self.height, self.width)

    def ray_count(self, reset=False):
        n = self.lib.ref_ray_count()
        if reset:
            self.lib.ref_ray_count_reset()
        return n

    def ray_log_begin(self, capacity, stride=1):
        self.lib.ref_ray_log_begin(C.c_longlong(capacity), C.c_int(stride))

    def ray_log_end(self, capacity):
        raw = np.empty((capacity, 11), np.float32)
        n = self.lib.ref_ray_log_end(_p(raw), C.c_longlong(capacity))
        raw = raw[:n]
        ints = raw.view(np.int32)
        return {"rays": raw[:, 0:6].copy(), "last_prim": ints[:, 6].copy(), "hit_prim": ints[:, 7].copy(), "t": raw[:, 8].copy(), "uv": raw[:, 9:11].copy()}

    def trace_closest(self, rays, last_prim=None):
        rays = np.ascontiguousarray(rays, np.float32)
        n = len(rays)
        prim, t, uv = np.empty(n, np.int32), np.empty(n, np.float32), np.empty((n, 2), np.float32)
        lp = np.ascontiguousarray(last_prim, np.int32) if last_prim is not None else None
        self.lib.ref_trace_closest(self.h, _p(rays), _p(lp), C.c_longlong(n), _p(prim), _p(t), _p(uv))
        return prim, t, uv

    def trace_shadow(self, rays, t_max, last_prim=None):
        rays = np.ascontiguousarray(rays, np.float32)
        n = len(rays)
        out = np.empty(n, np.uint8)
        lp = np.ascontiguousarray(last_prim, np.int32) if last_prim is not None else None
        self.lib.ref_trace_shadow(self.h, _p(rays), _p(np.ascontiguousarray(t_max, np.float32)), _p(lp), C.c_longlong(n), _p(out))
        return out

    def camera_rays(self, xy, width, height):
        xy = np.ascontiguousarray(xy, np.float32)
        out = np.empty((len(xy), 6), np.float32)
        self.lib.ref_camera_rays(self.h, _p(xy), C.c_longlong(len(xy)), C.c_int(width), C.c_int(height), _p(out))
        return out

    def bsdf_eval(self, materials, volume_states, vecs):
        n = len(vecs)
        out = np.empty((n, 4), np.float32)
        vs = np.ascontiguousarray(volume_states, np.uint8).copy()
        self.lib.ref_bsdf_eval(self.h, _p(np.ascontiguousarray(materials, np.uint8)), _p(vs), _p(np.ascontiguousarray(vecs, np.float32)), C.c_longlong(n), _p(out))
        return out, vs

    def bsdf_sample(self, materials, volume_states, vecs, seeds):
        n = len(vecs)
        out = np.empty((n, 7), np.float32)
        vs = np.ascontiguousarray(volume_states, np.uint8).copy()
        sd = np.ascontiguousarray(seeds, np.uint32).copy()
        self.lib.ref_bsdf_sample(self.h, _p(np.ascontiguousarray(materials, np.uint8)), _p(vs), _p(np.ascontiguousarray(vecs, np.float32)), _p(sd), C.c_longlong(n), _p(out))
        return out, vs, sd
