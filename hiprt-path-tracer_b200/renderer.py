"""Thin Python handles over the C ABI: Context (b200pt.h) and GPURenderer (b200pt_host.h, the C++ re-hosting of the
reference's GPURenderer). All compute happens in the CUDA library; numpy is only used to carry host arrays."""
import ctypes as C

import weakref

import numpy as np

from . import _capi
from . import structs as S

DOWNLOAD = {
    "pixels": (0, np.float32, 3), "first_hit_prim_index": (1, np.int32, 1), "shading_normals": (2, np.float32, 3),
    "geometric_normals": (3, np.float32, 3), "view_directions": (4, np.float32, 3), "first_hits": (5, np.float32, 3),
    "camera_ray_hit": (6, np.uint8, 1), "materials": (7, np.uint8, 256), "ray_volume_states": (8, np.uint8, 36),
    "pixel_active": (9, np.uint8, 1), "pixel_sample_count": (10, np.int32, 1), "pixel_squared_luminance": (11, np.float32, 1),
    "pixel_converged_sample_count": (12, np.int32, 1), "denoiser_albedo": (13, np.float32, 3), "denoiser_normals": (14, np.float32, 3),
    "restir_output_reservoirs": (15, np.uint8, 36), "restir_initial_candidates": (16, np.uint8, 36), "presampled_lights": (17, np.uint8, 48),
}


def _ptr(a, ctype=C.c_void_p):
    return a.ctypes.data_as(ctype) if a is not None else None


class Context:
    """b200pt_ctx: one per GPU (replaces HIPRTOrochiCtx, src/HIPRT-Orochi/HIPRTOrochiCtx.h:20-66)."""

    def __init__(self, device=0):
        self.lib = _capi.lib()
        self.handle = C.c_void_p()
        _capi.check(self.lib.b200pt_ctx_create(C.c_int(device), C.byref(self.handle)), "b200pt_ctx_create")
        self.device = device

    def close(self):
        if self.handle:
            # renderers hold device buffers of this context: destroy the ones still alive first (a renderer leaked by a failed
            # test would otherwise be collected after the context and free through a dangling b200pt_ctx)
            for r in list(getattr(self, "_renderers", ())):
                r.close()
            self.lib.b200pt_ctx_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- memory -----------------------------------------------------------------------------
    def malloc(self, nbytes):
        p = C.c_void_p()
        _capi.check(self.lib.b200pt_malloc(self.handle, C.c_size_t(nbytes), C.byref(p)), "b200pt_malloc")
        return p

    def free(self, dptr):
        _capi.check(self.lib.b200pt_free(self.handle, dptr), "b200pt_free")

    def upload(self, array):
        array = np.ascontiguousarray(array)
        p = self.malloc(max(array.nbytes, 1))
        _capi.check(self.lib.b200pt_memcpy_h2d(self.handle, p, _ptr(array), C.c_size_t(array.nbytes), None), "b200pt_memcpy_h2d")
        self.sync()
        return p

    def download(self, dptr, shape, dtype):
        out = np.empty(shape, dtype=dtype)
        _capi.check(self.lib.b200pt_memcpy_d2h(self.handle, _ptr(out), dptr, C.c_size_t(out.nbytes), None), "b200pt_memcpy_d2h")
        self.sync()
        return out

    def sync(self):
        _capi.check(self.lib.b200pt_stream_sync(self.handle, None), "b200pt_stream_sync")

    # ---- acceleration structure + ray batches ------------------------------------------------
    def build_bvh(self, vertices, indices, device=False):
        """device: False = the host binned-SAH builder, True / "ploc" = the device builder (PLOC clustering), "lbvh" = the device builder with the radix tree."""
        vertices = np.ascontiguousarray(vertices, dtype=np.float32)
        indices = np.ascontiguousarray(indices, dtype=np.int32)
        bvh = C.c_void_p()
        _capi.check(self.lib.b200pt_bvh_build(self.handle, _ptr(vertices), C.c_int(len(vertices)), _ptr(indices), C.c_int(len(indices) // 3 if indices.ndim == 1 else len(indices)),
                                              C.c_uint((3 if device == "lbvh" else 1) if device else 0), C.byref(bvh)), "b200pt_bvh_build")
        return bvh

    def bvh_info(self, bvh):
        nodes, tris, sah, secs = C.c_uint(), C.c_uint(), C.c_float(), C.c_double()
        _capi.check(self.lib.b200pt_bvh_info(bvh, C.byref(nodes), C.byref(tris), C.byref(sah), C.byref(secs)), "b200pt_bvh_info")
        st = S.BvhStats()
        _capi.check(self.lib.b200pt_bvh_get_stats(bvh, C.byref(st)), "b200pt_bvh_get_stats")
        return {"nodes": nodes.value, "triangles": tris.value, "sah_cost": sah.value, "build_seconds": secs.value, "max_depth": st.max_depth,
                "node_fill": st.node_fill, "built_on_device": bool(st.built_on_device), "device_build_seconds": st.device_build_seconds}

    def destroy_bvh(self, bvh):
        self.lib.b200pt_bvh_destroy(self.handle, bvh)

    def trace_closest(self, bvh, rays, last_prim=None):
        """rays: (n, 6) float32 host array. Returns (prim, t, uv) host arrays; the query itself runs on the GPU."""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = len(rays)
        d_rays = self.upload(rays)
        d_last = self.upload(np.ascontiguousarray(last_prim, dtype=np.int32)) if last_prim is not None else None
        d_prim, d_t, d_uv = self.malloc(4 * n), self.malloc(4 * n), self.malloc(8 * n)
        _capi.check(self.lib.b200pt_trace_closest(self.handle, bvh, d_rays, d_last, C.c_longlong(n), d_prim, d_t, d_uv, None), "b200pt_trace_closest")
        out = self.download(d_prim, (n,), np.int32), self.download(d_t, (n,), np.float32), self.download(d_uv, (n, 2), np.float32)
        for p in (d_rays, d_last, d_prim, d_t, d_uv):
            if p is not None:
                self.free(p)
        return out

    def trace_queue(self, bvh, rays, t_max=None, last_prim=None):
        """The same queries through the ray-queue traversal kernel FullPathTracer uses (b200pt_trace_queue). t_max None: closest hit,
        returns (prim, t, uv); else any-hit, returns the occlusion flags."""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = len(rays)
        d_rays = self.upload(rays)
        d_tmax = self.upload(np.ascontiguousarray(t_max, dtype=np.float32)) if t_max is not None else None
        d_last = self.upload(np.ascontiguousarray(last_prim, dtype=np.int32)) if last_prim is not None else None
        if t_max is None:
            d_prim, d_t, d_uv, d_occ = self.malloc(4 * n), self.malloc(4 * n), self.malloc(8 * n), None
        else:
            d_prim, d_t, d_uv, d_occ = None, None, None, self.malloc(n)
        _capi.check(self.lib.b200pt_trace_queue(self.handle, bvh, d_rays, d_tmax, d_last, C.c_longlong(n), d_prim, d_t, d_uv, d_occ, None), "b200pt_trace_queue")
        if t_max is None:
            out = self.download(d_prim, (n,), np.int32), self.download(d_t, (n,), np.float32), self.download(d_uv, (n, 2), np.float32)
        else:
            out = self.download(d_occ, (n,), np.uint8)
        for p in (d_rays, d_tmax, d_last, d_prim, d_t, d_uv, d_occ):
            if p is not None:
                self.free(p)
        return out

    STATIC_MAX_RAYS = 262144     # B200PT_TRACE_QUEUE_STATIC_MAX_RAYS (include/b200pt.h)

    def trace_queue_static(self, bvh, d_rays, d_t_max, d_last_prim, n, d_prim, d_t, d_uv, d_occluded):
        """b200pt_trace_queue_static on DEVICE buffers the caller owns (the entry point itself allocates nothing). Asynchronous."""
        _capi.check(self.lib.b200pt_trace_queue_static(self.handle, bvh, d_rays, d_t_max, d_last_prim, C.c_longlong(n), d_prim, d_t, d_uv, d_occluded, None), "b200pt_trace_queue_static")

    def trace_shadow(self, bvh, rays, t_max, last_prim=None):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = len(rays)
        d_rays, d_tmax = self.upload(rays), self.upload(np.ascontiguousarray(t_max, dtype=np.float32))
        d_last = self.upload(np.ascontiguousarray(last_prim, dtype=np.int32)) if last_prim is not None else None
        d_out = self.malloc(n)
        _capi.check(self.lib.b200pt_trace_shadow(self.handle, bvh, d_rays, d_tmax, d_last, C.c_longlong(n), d_out, None), "b200pt_trace_shadow")
        out = self.download(d_out, (n,), np.uint8)
        for p in (d_rays, d_tmax, d_last, d_out):
            if p is not None:
                self.free(p)
        return out

    # ---- counters ---------------------------------------------------------------------------
    def set_path_mode(self, mode):
        """'megakernel' | 'fused' | 'split' (b200pt_set_path_mode): the shape FullPathTracer runs in."""
        m = {"megakernel": 0, "fused": 1, "split": 2, "phased": 3}[mode] if isinstance(mode, str) else int(mode)
        _capi.check(self.lib.b200pt_set_path_mode(self.handle, C.c_int(m)), "b200pt_set_path_mode")

    def path_profile_enable(self, enable=True):
        _capi.check(self.lib.b200pt_path_profile_enable(self.handle, C.c_int(1 if enable else 0)), "b200pt_path_profile_enable")

    def path_profile_read(self, reset=True):
        """ms per stage of the wavefront FullPathTracer since the last reset."""
        out = (C.c_double * 4)()
        _capi.check(self.lib.b200pt_path_profile_read(self.handle, out, C.c_int(1 if reset else 0)), "b200pt_path_profile_read")
        return {"trace": out[0], "surface": out[1], "sort": out[2], "shade": out[3]}

    def set_trace_flavour(self, flavour):
        """0 classic, 5 / 6 / 7 pipelined at that many blocks per SM, 1 pipelined default, -1 environment / built-in default."""
        _capi.check(self.lib.b200pt_set_trace_flavour(self.handle, C.c_int(flavour)), "b200pt_set_trace_flavour")

    def set_ray_bin_bits(self, bits):
        """Ray binning before traversal: 2^bits cells per axis, 0 off, -1 environment / built-in default (b200pt_set_ray_bin_bits)."""
        _capi.check(self.lib.b200pt_set_ray_bin_bits(self.handle, C.c_int(bits)), "b200pt_set_ray_bin_bits")

    def set_restir_phases(self, enable=True):
        """ReSTIR DI InitialCandidates / SpatiotemporalReuse as collect -> ray-queue traversal -> resolve phases (default) or as single kernels
        with per-thread traversal (b200pt_set_restir_phases); the reservoirs are the same bits."""
        _capi.check(self.lib.b200pt_set_restir_phases(self.handle, C.c_int(1 if enable else 0)), "b200pt_set_restir_phases")

    def counters_enable(self, enable=True):
        _capi.check(self.lib.b200pt_counters_enable(self.handle, C.c_int(1 if enable else 0)), "b200pt_counters_enable")

    def counters_read(self, reset=True):
        c = S.Counters()
        _capi.check(self.lib.b200pt_counters_read(self.handle, C.byref(c), C.c_int(1 if reset else 0)), "b200pt_counters_read")
        return {"rays": c.rays, "nodes": c.nodes, "triangles": c.triangles, "stack_overflows": c.stack_overflows,
                "queue_rays": c.queue_rays, "queue_nodes": c.queue_nodes, "queue_triangles": c.queue_triangles}


class GPURenderer:
    """Handle on b200pt_host::GPURenderer (hiprt-path-tracer_b200/host/GPURenderer.h)."""

    def __init__(self, ctx, options=None, seed_schedule="gpu"):
        self.ctx = ctx
        self.host = _capi.host()
        self.handle = C.c_void_p()
        _capi.check_host(self.host.b200pt_renderer_create(ctx.handle, C.byref(self.handle)), "b200pt_renderer_create")
        if not hasattr(ctx, "_renderers"):
            ctx._renderers = weakref.WeakSet()
        ctx._renderers.add(self)
        self.set_seed_schedule(seed_schedule)
        if options is not None:
            self.set_kernel_options(**options)

    def close(self):
        if self.handle:
            self.host.b200pt_renderer_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_rng_seed(self, seed):
        """Host generator seed (42 = the reference's); resets the accumulation. One per rank under sample-pass sharding."""
        _capi.check_host(self.host.b200pt_renderer_set_rng_seed(self.handle, C.c_uint(seed)), "b200pt_renderer_set_rng_seed")

    def set_seed_schedule(self, schedule):
        self.host.b200pt_renderer_set_seed_schedule(self.handle, C.c_int(1 if schedule == "cpu" else 0))

    def set_kernel_options(self, lss="mis", ess="none", ggx=0, later=3):
        o = S.Options(S.LSS[lss] if isinstance(lss, str) else lss, S.ESS[ess] if isinstance(ess, str) else ess, ggx, later)
        _capi.check_host(self.host.b200pt_renderer_set_kernel_options(self.handle, C.byref(o)), "b200pt_renderer_set_kernel_options")

    def set_brdf_luts(self, luts):
        arrs = [np.ascontiguousarray(luts[k], dtype=np.float32) for k in ("sheen_ltc", "conductor", "glossy", "glass", "glass_inverse", "thin_glass")]
        _capi.check_host(self.host.b200pt_renderer_set_brdf_luts(self.handle, *[_ptr(a) for a in arrs]), "b200pt_renderer_set_brdf_luts")

    def set_scene(self, scene):
        for tex in scene.textures:
            t = np.ascontiguousarray(tex, dtype=np.uint8)
            self.host.b200pt_renderer_add_texture(self.handle, _ptr(t), C.c_int(t.shape[1]), C.c_int(t.shape[0]), C.c_int(t.shape[2]))
        a = scene.arrays()
        _capi.check_host(self.host.b200pt_renderer_set_scene(
            self.handle, _ptr(a["vertices"]), C.c_int(len(a["vertices"])), _ptr(a["indices"]), C.c_int(len(a["indices"]) // 3),
            _ptr(a["normals"]), _ptr(a["has_normals"]), _ptr(a["texcoords"]), _ptr(a["material_indices"]), _ptr(a["materials"]),
            C.c_int(scene.material_count), _ptr(a["emissive"]), C.c_int(len(a["emissive"]))), "b200pt_renderer_set_scene")

    def set_envmap(self, texels, alias_probas=None, alias_alias=None, total_sum=0.0):
        """Without an alias table the host layer builds it (GPURenderer.cpp: compute_envmap_alias_table)."""
        texels = np.ascontiguousarray(texels, dtype=np.float32)
        h, w, ch = texels.shape
        probas = None if alias_probas is None else np.ascontiguousarray(alias_probas, dtype=np.float32)
        alias = None if alias_alias is None else np.ascontiguousarray(alias_alias, dtype=np.int32)
        _capi.check_host(self.host.b200pt_renderer_set_envmap(self.handle, _ptr(texels), C.c_int(w), C.c_int(h), C.c_int(ch),
                                                              _ptr(probas), _ptr(alias), C.c_float(total_sum)), "b200pt_renderer_set_envmap")

    def set_camera(self, camera):
        self.host.b200pt_renderer_set_camera(self.handle, C.byref(camera))

    def resize(self, width, height):
        _capi.check_host(self.host.b200pt_renderer_resize(self.handle, C.c_int(width), C.c_int(height)), "b200pt_renderer_resize")
        self.width, self.height = width, height

    def set_tile(self, tile, halo=0):
        """tile = (x0, y0, x1, y1) or None; halo = rows either side CameraRays also renders (ReSTIR DI tile sharding)."""
        if tile is None:
            self.host.b200pt_renderer_set_tile(self.handle, None)
        else:
            t = S.Tile(*tile)
            self.host.b200pt_renderer_set_tile(self.handle, C.byref(t))
        self.host.b200pt_renderer_set_tile_halo(self.handle, C.c_int(halo))

    def set_samples_in_flight(self, samples, keep_g_buffer=True):
        """Fused sample batches (b200pt_renderer_set_samples_in_flight): 0 = the two launch points per pass."""
        _capi.check(self.host.b200pt_renderer_set_samples_in_flight(self.handle, C.c_int(samples), C.c_int(1 if keep_g_buffer else 0)), "b200pt_renderer_set_samples_in_flight")

    @property
    def renders_sample_batches(self):
        return bool(self.host.b200pt_renderer_renders_sample_batches(self.handle))

    def set_path_lanes(self, lanes):
        """FullPathTracer as 1..4 concurrent band launches (trace of one band under the shade phases of another)."""
        _capi.check(self.host.b200pt_renderer_set_path_lanes(self.handle, C.c_int(lanes)), "b200pt_renderer_set_path_lanes")

    @property
    def path_lanes(self):
        return int(self.host.b200pt_renderer_path_lanes(self.handle))

    def reset(self):
        self.host.b200pt_renderer_reset(self.handle)

    @property
    def render_settings(self):
        return self.host.b200pt_renderer_render_settings(self.handle).contents

    @property
    def world_settings(self):
        return self.host.b200pt_renderer_world_settings(self.handle).contents

    @property
    def bsdfs_data(self):
        return self.host.b200pt_renderer_bsdfs_data(self.handle).contents

    @property
    def render_data(self):
        return self.host.b200pt_renderer_render_data(self.handle).contents

    @property
    def bvh(self):
        return C.c_void_p(self.host.b200pt_renderer_bvh(self.handle))

    def render_frame(self):
        _capi.check_host(self.host.b200pt_renderer_render_frame(self.handle), "b200pt_renderer_render_frame")

    def synchronize(self):
        _capi.check_host(self.host.b200pt_renderer_synchronize(self.handle), "b200pt_renderer_synchronize")

    def last_frame_ms(self):
        ms = C.c_float()
        _capi.check_host(self.host.b200pt_renderer_last_frame_ms(self.handle, C.byref(ms)), "b200pt_renderer_last_frame_ms")
        return ms.value

    def last_kernel_ms(self, which):
        ms = C.c_float()
        idx = {"CameraRays": 0, "FullPathTracer": 1, "ReSTIR_DI": 2}[which]
        _capi.check_host(self.host.b200pt_renderer_last_kernel_ms(self.handle, C.c_int(idx), C.byref(ms)), "b200pt_renderer_last_kernel_ms")
        return ms.value

    def device_pointer(self, name):
        return self.host.b200pt_renderer_device_pointer(self.handle, C.c_int(DOWNLOAD[name][0]))

    def tonemap(self, gamma=2.2, exposure=1.0, flip_y=False):
        """Headless output (the reference's main.cpp:91-97: cpu_renderer.tonemap(2.2f, 1.0f), then write_image_png): the accumulated
        frame tone-mapped on the device (b200pt_tonemap). Returns ([H, W, 3] float32, [H, W, 3] uint8)."""
        self.synchronize()
        ctx, n = self.ctx, self.width * self.height
        d_float, d_u8 = ctx.malloc(n * 12), ctx.malloc(n * 3)
        try:
            samples = int(self.render_settings.sample_number) if self.render_settings.accumulate else 0
            _capi.check(ctx.lib.b200pt_tonemap(ctx.handle, C.c_void_p(self.device_pointer("pixels")), C.c_int(self.width), C.c_int(self.height), C.c_int(samples),
                                               C.c_float(gamma), C.c_float(exposure), C.c_int(int(flip_y)), d_float, d_u8, None), "b200pt_tonemap")
            return ctx.download(d_float, (self.height, self.width, 3), np.float32), ctx.download(d_u8, (self.height, self.width, 3), np.uint8)
        finally:
            ctx.free(d_float)
            ctx.free(d_u8)

    def download(self, name):
        which, dtype, comps = DOWNLOAD[name]
        n = self.width * self.height
        need = -self.host.b200pt_renderer_download(self.handle, which, None, 0)
        if need <= 0:
            return None
        out = np.empty(need // np.dtype(dtype).itemsize, dtype=dtype)
        got = self.host.b200pt_renderer_download(self.handle, which, _ptr(out), out.nbytes)
        if got < 0:
            _capi.check_host(int(got), "b200pt_renderer_download")
        if name == "presampled_lights":
            return out.reshape(-1, comps)
        return out.reshape(self.height, self.width, comps) if comps > 1 else out.reshape(self.height, self.width)
