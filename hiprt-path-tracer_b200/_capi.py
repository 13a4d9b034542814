"""ctypes binding of libb200pt.so (include/b200pt.h) and libb200pt_host.so (include/b200pt_host.h).

There is no fallback of any kind: if the libraries are missing, or no CUDA device is present when a context is
created, an exception is raised."""
import ctypes as C
import os

from . import structs as S

HERE = os.path.dirname(os.path.abspath(__file__))
# B200PT_LIB_DIR: a directory holding another build of BOTH libraries (tools/trace_sweep.sh puts tuning variants there)
LIB_DIR = os.environ.get("B200PT_LIB_DIR") or HERE
LIB_PATH = os.path.join(LIB_DIR, "libb200pt.so")
HOST_LIB_PATH = os.path.join(LIB_DIR, "libb200pt_host.so")


class B200ptError(RuntimeError):
    pass


_lib = None
_host = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200ptError(f"{LIB_PATH} is missing: run `python __graft_entry__.py build` (no CPU fallback exists)")
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        _lib.b200pt_last_error.restype = C.c_char_p
        _lib.b200pt_version.restype = C.c_char_p
        _lib.b200pt_bvh_device_handle.restype = C.c_void_p
        _lib.b200pt_bvh_device_handle.argtypes = [C.c_void_p]
    return _lib


def host():
    global _host
    if _host is None:
        lib()
        if not os.path.exists(HOST_LIB_PATH):
            raise B200ptError(f"{HOST_LIB_PATH} is missing: run `python __graft_entry__.py build`")
        _host = C.CDLL(HOST_LIB_PATH)
        _host.b200pt_host_last_error.restype = C.c_char_p
        for name, res in (("b200pt_renderer_render_settings", S.HIPRTRenderSettings), ("b200pt_renderer_world_settings", S.WorldSettings),
                          ("b200pt_renderer_bsdfs_data", S.BRDFsData), ("b200pt_renderer_render_data", S.HIPRTRenderData)):
            getattr(_host, name).restype = C.POINTER(res)
            getattr(_host, name).argtypes = [C.c_void_p]
        _host.b200pt_renderer_bvh.restype = C.c_void_p
        _host.b200pt_renderer_bvh.argtypes = [C.c_void_p]
        _host.b200pt_renderer_device_pointer.restype = C.c_void_p
        _host.b200pt_renderer_device_pointer.argtypes = [C.c_void_p, C.c_int]
        _host.b200pt_baker_last_table.restype = C.c_longlong
        _host.b200pt_baker_last_table.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p]
        _host.b200pt_renderer_download.restype = C.c_longlong
        _host.b200pt_renderer_download.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_longlong]
    return _host


def check(code, what=""):
    if code != 0:
        msg = lib().b200pt_last_error().decode(errors="replace")
        raise B200ptError(f"{what} failed ({code}): {msg}")


def check_host(code, what=""):
    if code != 0:
        msg = host().b200pt_host_last_error().decode(errors="replace")
        raise B200ptError(f"{what} failed ({code}): {msg}")


# Every symbol include/b200pt.h and include/b200pt_host.h declare (checked by tests/test_capi_symbols.py)
API_SYMBOLS = [
    "b200pt_device_count", "b200pt_ctx_create", "b200pt_ctx_destroy", "b200pt_ctx_make_current", "b200pt_last_error", "b200pt_version",
    "b200pt_malloc", "b200pt_free", "b200pt_memset", "b200pt_memcpy_h2d", "b200pt_memcpy_d2h", "b200pt_memcpy_d2d",
    "b200pt_host_alloc_pinned", "b200pt_host_free_pinned", "b200pt_stream_create", "b200pt_stream_destroy", "b200pt_stream_query",
    "b200pt_stream_sync", "b200pt_event_create", "b200pt_event_destroy", "b200pt_event_record", "b200pt_event_elapsed_ms",
    "b200pt_tex2d_create", "b200pt_tex3d_create", "b200pt_tex_destroy", "b200pt_tex2d_array_create", "b200pt_bvh_build",
    "b200pt_bvh_destroy", "b200pt_bvh_attach_shading_data", "b200pt_bvh_device_handle", "b200pt_bvh_info", "b200pt_bvh_get_stats", "b200pt_kernel_variant_select", "b200pt_set_path_mode", "b200pt_get_path_mode", "b200pt_path_profile_enable", "b200pt_path_profile_read", "b200pt_launch_CameraRays",
    "b200pt_launch_FullPathTracer", "b200pt_launch_FullPathTracer_lane", "b200pt_launch_PathTraceSamples", "b200pt_stream_wait_event", "b200pt_launch_ReSTIR_DI_LightsPresampling", "b200pt_launch_ReSTIR_DI_InitialCandidates",
    "b200pt_launch_ReSTIR_DI_SpatiotemporalReuse", "b200pt_launch_ReSTIR_DI_TemporalReuse", "b200pt_launch_ReSTIR_DI_SpatialReuse", "b200pt_tonemap", "b200pt_trace_closest", "b200pt_trace_shadow", "b200pt_trace_queue", "b200pt_trace_queue_static", "b200pt_set_trace_flavour", "b200pt_set_ray_bin_bits", "b200pt_set_restir_phases", "b200pt_debug_bsdf_eval", "b200pt_debug_bsdf_sample", "b200pt_debug_envmap_sample", "b200pt_debug_thin_film_fresnel", "b200pt_counters_enable",
    "b200pt_counters_read",
]
# include/b200pt_comm.h (multi-GPU plumbing; part of libb200pt.so)
COMM_API_SYMBOLS = [
    "b200pt_comm_tile", "b200pt_comm_restir_halo_rows", "b200pt_comm_get_unique_id", "b200pt_comm_create", "b200pt_comm_destroy", "b200pt_comm_rank", "b200pt_comm_world",
    "b200pt_comm_gather_tiles", "b200pt_comm_exchange_halo", "b200pt_comm_allreduce_sum", "b200pt_comm_allreduce_host_f64", "b200pt_comm_barrier",
    "b200pt_comm_balanced_bounds", "b200pt_comm_set_column_bounds",
]
# include/b200pt_bake.h (directional-albedo LUT bakers; part of libb200pt.so)
BAKE_API_SYMBOLS = [
    "b200pt_bake_plan_for", "b200pt_bake_ggx_conductor_directional_albedo", "b200pt_bake_ggx_fresnel_directional_albedo",
    "b200pt_bake_glossy_dielectric_directional_albedo", "b200pt_bake_ggx_glass_directional_albedo", "b200pt_bake_ggx_thin_glass_directional_albedo",
]
HOST_API_SYMBOLS = [
    "b200pt_host_last_error", "b200pt_renderer_create", "b200pt_renderer_destroy", "b200pt_renderer_set_seed_schedule", "b200pt_renderer_set_rng_seed",
    "b200pt_renderer_set_kernel_options", "b200pt_renderer_set_brdf_luts", "b200pt_renderer_add_texture", "b200pt_renderer_set_scene",
    "b200pt_renderer_set_envmap", "b200pt_host_compute_alias_table", "b200pt_renderer_set_camera", "b200pt_renderer_resize", "b200pt_renderer_set_tile", "b200pt_renderer_set_tile_halo", "b200pt_renderer_set_path_lanes", "b200pt_renderer_path_lanes", "b200pt_renderer_set_samples_in_flight", "b200pt_renderer_renders_sample_batches",
    "b200pt_renderer_reset", "b200pt_renderer_render_settings", "b200pt_renderer_world_settings", "b200pt_renderer_bsdfs_data",
    "b200pt_renderer_render_data", "b200pt_renderer_bvh", "b200pt_renderer_render_frame", "b200pt_renderer_synchronize",
    "b200pt_renderer_last_frame_ms", "b200pt_renderer_last_kernel_ms", "b200pt_renderer_download", "b200pt_renderer_device_pointer",
    "b200pt_baker_create", "b200pt_baker_destroy", "b200pt_baker_set_options", "b200pt_baker_bake", "b200pt_baker_is_complete", "b200pt_baker_wait", "b200pt_baker_last_table",
    "b200pt_host_write_hdr_1channel",
]
