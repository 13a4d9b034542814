"""GPUBaker: host-side mirror of the reference's LUT baker (src/Renderer/Baker/GPUBaker.{h,cpp}, GPUBakerKernel.cpp) above the C ABI of
include/b200pt_bake.h. Same method names, same settings structs, same output files (one Radiance .hdr slice per ior index, named as
GPUBakerConstants.h:39-82 names them, written unflipped like GPUBakerKernel.cpp:107-126) -- what data/BRDFsData ships. There is no CPU path:
every bake is a CUDA launch through libb200pt.so.

The reference bakes on a worker thread and polls is_*_bake_complete(); here a bake is one asynchronous launch on the context's main stream
and the methods return the table after synchronising, so the is_*_complete() queries are always true afterwards (kept for drop-in callers)."""
import ctypes as C
import os

import numpy as np

from . import _capi
from .image_io import write_hdr

KIND_CONDUCTOR, KIND_FRESNEL, KIND_GLOSSY, KIND_GLASS_ENTERING, KIND_GLASS_EXITING, KIND_THIN_GLASS = range(6)


class GGXConductorDirectionalAlbedoSettings(C.Structure):
    """src/Renderer/Baker/GGXConductorDirectionalAlbedoSettings.h:11-17"""
    _fields_ = [("texture_size_cos_theta", C.c_int), ("texture_size_roughness", C.c_int), ("integration_sample_count", C.c_int)]

    def __init__(self, texture_size_cos_theta=128, texture_size_roughness=128, integration_sample_count=65536):
        super().__init__(texture_size_cos_theta, texture_size_roughness, integration_sample_count)


class GGXFresnelDirectionalAlbedoSettings(C.Structure):
    """src/Renderer/Baker/GGXFresnelDirectionalAlbedoSettings.h:11-18"""
    _fields_ = [("texture_size_cos_theta", C.c_int), ("texture_size_roughness", C.c_int), ("texture_size_ior", C.c_int), ("integration_sample_count", C.c_int)]

    def __init__(self, texture_size_cos_theta=256, texture_size_roughness=256, texture_size_ior=256, integration_sample_count=65536):
        super().__init__(texture_size_cos_theta, texture_size_roughness, texture_size_ior, integration_sample_count)


class GlossyDielectricDirectionalAlbedoSettings(C.Structure):
    """src/Renderer/Baker/GlossyDielectricDirectionalAlbedoSettings.h:11-18"""
    _fields_ = [("texture_size_cos_theta_o", C.c_int), ("texture_size_roughness", C.c_int), ("texture_size_ior", C.c_int), ("integration_sample_count", C.c_int)]

    def __init__(self, texture_size_cos_theta_o=128, texture_size_roughness=64, texture_size_ior=128, integration_sample_count=65536):
        super().__init__(texture_size_cos_theta_o, texture_size_roughness, texture_size_ior, integration_sample_count)


class GGXGlassDirectionalAlbedoSettings(C.Structure):
    """src/Renderer/Baker/GGXGlassDirectionalAlbedoSettings.h:11-18"""
    _fields_ = [("texture_size_cos_theta_o", C.c_int), ("texture_size_roughness", C.c_int), ("texture_size_ior", C.c_int), ("integration_sample_count", C.c_int)]

    def __init__(self, texture_size_cos_theta_o=256, texture_size_roughness=16, texture_size_ior=128, integration_sample_count=65536):
        super().__init__(texture_size_cos_theta_o, texture_size_roughness, texture_size_ior, integration_sample_count)


class GGXThinGlassDirectionalAlbedoSettings(C.Structure):
    """src/Renderer/Baker/GGXThinGlassDirectionalAlbedoSettings.h:11-18"""
    _fields_ = [("texture_size_cos_theta_o", C.c_int), ("texture_size_roughness", C.c_int), ("texture_size_ior", C.c_int), ("integration_sample_count", C.c_int)]

    def __init__(self, texture_size_cos_theta_o=32, texture_size_roughness=32, texture_size_ior=96, integration_sample_count=65536):
        super().__init__(texture_size_cos_theta_o, texture_size_roughness, texture_size_ior, integration_sample_count)


class BakeOptions(C.Structure):
    """b200pt_bake_options (include/b200pt_bake.h)"""
    _fields_ = [("ggx_sample_function", C.c_int), ("GGX_masking_shadowing", C.c_int), ("chunks", C.c_int)]


class BakePlan(C.Structure):
    """b200pt_bake_plan (include/b200pt_bake.h)"""
    _fields_ = [("iterations_per_kernel", C.c_int), ("nb_kernel_launch", C.c_int), ("nb_samples", C.c_int), ("chunks", C.c_int), ("texels", C.c_longlong)]


# GPUBakerConstants.h:39-82
def get_GGX_conductor_Ess_filename(size_cos_theta=128, size_roughness=128):
    return f"GGX_Conductor_{size_cos_theta}x{size_roughness}.hdr"


def get_GGX_fresnel_Ess_filename(size_cos_theta=256, size_roughness=256, size_ior=256):
    return f"GGX_Fresnel_{size_cos_theta}x{size_roughness}x{size_ior}.hdr"


def get_glossy_dielectric_Ess_filename(size_cos_theta=128, size_roughness=64, size_ior=128):
    return f"Glossy_Ess_{size_cos_theta}x{size_roughness}x{size_ior}.hdr"


def get_GGX_glass_Ess_filename(size_cos_theta=256, size_roughness=16, size_ior=128):
    return f"GGX_Glass_Ess_{size_cos_theta}x{size_roughness}x{size_ior}.hdr"


def get_GGX_thin_glass_Ess_filename(size_cos_theta=32, size_roughness=32, size_ior=96):
    return f"GGX_Thin_Glass_Ess_{size_cos_theta}x{size_roughness}x{size_ior}.hdr"


def get_GGX_glass_inv_Ess_filename(size_cos_theta=256, size_roughness=16, size_ior=128):
    return "inv_" + get_GGX_glass_Ess_filename(size_cos_theta, size_roughness, size_ior)


def bake_plan(kind, size_cos_theta, size_roughness, size_ior, integration_sample_count, options=None):
    """b200pt_bake_plan_for: the reference's launch arithmetic for a table (host only, no device needed)."""
    plan = BakePlan()
    _capi.check(_capi.lib().b200pt_bake_plan_for(C.c_int(kind), C.c_int(size_cos_theta), C.c_int(size_roughness), C.c_int(size_ior), C.c_int(integration_sample_count),
                                                 C.byref(options) if options is not None else None, C.byref(plan)), "b200pt_bake_plan_for")
    return plan


def write_table(table, output_filename, directory="."):
    """GPUBakerKernel.cpp:107-126: a 3D table is one single-channel .hdr per ior slice, '<slice index><name>'; 2D tables are one file."""
    paths = []
    if table.ndim == 2:
        table = table[None]
        names = [output_filename]
    else:
        names = [f"{i}{output_filename}" for i in range(table.shape[0])]
    for sl, name in zip(table, names):
        path = os.path.join(directory, name)
        # NaN texels (the thin-glass table has some by the reference's own arithmetic) have no RGBE encoding: written as 0
        write_hdr(path, np.repeat(np.nan_to_num(sl, nan=0.0, posinf=0.0, neginf=0.0)[:, :, None], 3, axis=2))
        paths.append(path)
    return paths


class GPUBaker:
    """src/Renderer/Baker/GPUBaker.h:18-57. `ctx` is a renderer.Context (the reference takes its GPURenderer for the device context)."""

    def __init__(self, ctx, ggx_sample_function=0, GGX_masking_shadowing=0, chunks=0, output_directory=None):
        self.ctx = ctx
        self.lib = _capi.lib()
        self.options = BakeOptions(ggx_sample_function, GGX_masking_shadowing, chunks)
        self.output_directory = output_directory
        self.last_kernel_ms = {}
        self._complete = set()

    def _run(self, name, entry, settings, shape, outputs=1):
        n = int(np.prod(shape))
        bufs = [self.ctx.malloc(4 * n) for _ in range(outputs)]
        try:
            ev = [C.c_void_p(), C.c_void_p()]
            for e in ev:
                _capi.check(self.lib.b200pt_event_create(self.ctx.handle, C.byref(e)), "b200pt_event_create")
            _capi.check(self.lib.b200pt_event_record(self.ctx.handle, ev[0], None), "b200pt_event_record")
            _capi.check(entry(self.ctx.handle, C.byref(settings), C.byref(self.options), *bufs, None), name)
            _capi.check(self.lib.b200pt_event_record(self.ctx.handle, ev[1], None), "b200pt_event_record")
            self.ctx.sync()
            ms = C.c_float()
            _capi.check(self.lib.b200pt_event_elapsed_ms(self.ctx.handle, ev[0], ev[1], C.byref(ms)), "b200pt_event_elapsed_ms")
            self.last_kernel_ms[name] = ms.value
            for e in ev:
                self.lib.b200pt_event_destroy(self.ctx.handle, e)
            tables = [self.ctx.download(b, shape, np.float32) for b in bufs]
        finally:
            for b in bufs:
                self.ctx.free(b)
        self._complete.add(name)
        return tables

    def _maybe_write(self, table, output_filename):
        if output_filename and self.output_directory is not None:
            write_table(table, output_filename, self.output_directory)

    def bake_ggx_conductor_directional_albedo(self, bake_settings=None, output_filename=None):
        s = bake_settings or GGXConductorDirectionalAlbedoSettings()
        (t,) = self._run("ggx_conductor", self.lib.b200pt_bake_ggx_conductor_directional_albedo, s, (s.texture_size_roughness, s.texture_size_cos_theta))
        self._maybe_write(t, output_filename)
        return t

    def bake_ggx_fresnel_directional_albedo(self, bake_settings=None, output_filename=None):
        s = bake_settings or GGXFresnelDirectionalAlbedoSettings()
        (t,) = self._run("ggx_fresnel", self.lib.b200pt_bake_ggx_fresnel_directional_albedo, s, (s.texture_size_ior, s.texture_size_roughness, s.texture_size_cos_theta))
        self._maybe_write(t, output_filename)
        return t

    def bake_glossy_dielectric_directional_albedo(self, bake_settings=None, output_filename=None):
        s = bake_settings or GlossyDielectricDirectionalAlbedoSettings()
        (t,) = self._run("glossy_dielectric", self.lib.b200pt_bake_glossy_dielectric_directional_albedo, s,
                         (s.texture_size_ior, s.texture_size_roughness, s.texture_size_cos_theta_o))
        self._maybe_write(t, output_filename)
        return t

    def bake_ggx_glass_directional_albedo(self, bake_settings=None, output_filename=None):
        """Returns (entering, exiting); the exiting table is written as 'inv_' + output_filename (GPUBaker.cpp:74-83)."""
        s = bake_settings or GGXGlassDirectionalAlbedoSettings()
        entering, exiting = self._run("ggx_glass", self.lib.b200pt_bake_ggx_glass_directional_albedo, s,
                                      (s.texture_size_ior, s.texture_size_roughness, s.texture_size_cos_theta_o), outputs=2)
        self._maybe_write(entering, output_filename)
        self._maybe_write(exiting, "inv_" + output_filename if output_filename else None)
        return entering, exiting

    def bake_ggx_thin_glass_directional_albedo(self, bake_settings=None, output_filename=None):
        s = bake_settings or GGXThinGlassDirectionalAlbedoSettings()
        (t,) = self._run("ggx_thin_glass", self.lib.b200pt_bake_ggx_thin_glass_directional_albedo, s,
                         (s.texture_size_ior, s.texture_size_roughness, s.texture_size_cos_theta_o))
        self._maybe_write(t, output_filename)
        return t

    def is_ggx_conductor_directional_albedo_bake_complete(self):
        return "ggx_conductor" in self._complete

    def is_ggx_fresnel_directional_albedo_bake_complete(self):
        return "ggx_fresnel" in self._complete

    def is_glossy_dielectric_directional_albedo_bake_complete(self):
        return "glossy_dielectric" in self._complete

    def is_ggx_glass_directional_albedo_bake_complete(self):
        return "ggx_glass" in self._complete

    def is_ggx_thin_glass_directional_albedo_bake_complete(self):
        return "ggx_thin_glass" in self._complete
