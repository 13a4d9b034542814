// The render kernels of the hot path: CameraRays and FullPathTracer (megakernel form, one thread per pixel).
// reference: src/Device/kernels/CameraRays.h:19-179, src/Device/kernels/FullPathTracer.h:74-327,
//            src/Device/includes/AdaptiveSampling.h:11-104, RussianRoulette.h:14-49, HIPRTCamera.h:27-48.
#pragma once

#include "envmap.cuh"
#include "ris.cuh"

namespace pt
{
struct Tile { int x0, y0, x1, y1; };

// ---- adaptive sampling (AdaptiveSampling.h) ----------------------------------------------------
PT_DEV bool has_access_to_adaptive_sampling_buffers(const HIPRTRenderSettings& s)
{
    bool has_access = false;
    has_access |= s.stop_pixel_noise_threshold > 0.0f;
    has_access |= (bool)s.enable_adaptive_sampling;
    has_access &= (bool)s.accumulate;
    return has_access;
}

PT_DEV bool do_render_low_resolution(const HIPRTRenderSettings& s) { return s.wants_render_low_resolution && s.allow_render_low_resolution && s.accumulate; }

PT_DEV float get_pixel_confidence_interval(const HIPRTRenderData& rd, int pixel_index, int pixel_sample_count, float& average_luminance)
{
    float luminance = RGB(ld3_rw(rd.buffers.pixels, pixel_index)).luminance();
    average_luminance = luminance / (pixel_sample_count + 1);
    float squared_luminance = rd.aux_buffers.pixel_squared_luminance[pixel_index];
    float pixel_variance = (squared_luminance - luminance * average_luminance) / (pixel_sample_count + 1);
    return 1.96f * sqrtf(pixel_variance) / sqrtf(pixel_sample_count + 1);
}

PT_DEV bool adaptive_sampling(const HIPRTRenderData& rd, int pixel_index, bool& pixel_converged)
{
    const HIPRTRenderSettings& s = rd.render_settings;
    const AuxiliaryBuffers& aux = rd.aux_buffers;
    if (!has_access_to_adaptive_sampling_buffers(s))
        return true;

    if (s.enable_adaptive_sampling)
    {
        if (aux.pixel_converged_sample_count[pixel_index] != -1)
            return false;
        int pixel_sample_count = aux.pixel_sample_count[pixel_index];
        if (pixel_sample_count > s.adaptive_sampling_min_samples)
        {
            float average_luminance;
            float confidence_interval = get_pixel_confidence_interval(rd, pixel_index, pixel_sample_count, average_luminance);
            bool pixel_needs_sampling = confidence_interval > s.adaptive_sampling_noise_threshold * average_luminance;
            if (!pixel_needs_sampling)
            {
                if (aux.pixel_converged_sample_count[pixel_index] == -1)
                    aux.pixel_converged_sample_count[pixel_index] = pixel_sample_count;
                return false;
            }
        }
        return true;
    }
    else if (s.stop_pixel_noise_threshold > 0.0f && s.enable_pixel_stop_noise_threshold)
    {
        int pixel_sample_count = aux.pixel_sample_count[pixel_index];
        float average_luminance;
        float confidence_interval = get_pixel_confidence_interval(rd, pixel_index, pixel_sample_count, average_luminance);
        pixel_converged = (confidence_interval <= s.stop_pixel_noise_threshold * average_luminance) && (s.sample_number > 1);
        int current_converged_count = aux.pixel_converged_sample_count[pixel_index];
        if (pixel_converged && current_converged_count == -1)
            aux.pixel_converged_sample_count[pixel_index] = pixel_sample_count;
        else if (!pixel_converged)
            aux.pixel_converged_sample_count[pixel_index] = -1;
    }
    return true;
}

// ---- G-buffer element copies (256-byte material, 36-byte volume state) -------------------------
PT_DEV void store_material(SimplifiedRendererMaterial* dst, const Material& m)
{
    const uint4* s = reinterpret_cast<const uint4*>(&m);
    uint4* d = reinterpret_cast<uint4*>(dst);   // 256-byte elements of a cudaMalloc'd array: 16-byte aligned
#pragma unroll
    for (int i = 0; i < 16; i++)
        d[i] = s[i];
}
PT_DEV void load_material(Material& m, const SimplifiedRendererMaterial* src)
{
    const uint4* s = reinterpret_cast<const uint4*>(src);
    uint4* d = reinterpret_cast<uint4*>(&m);
#pragma unroll
    for (int i = 0; i < 16; i++)
        d[i] = s[i];
}
PT_DEV void copy_volume_state(RayVolumeState* dst, const RayVolumeState* src)
{
    const uint32_t* s = reinterpret_cast<const uint32_t*>(src);
    uint32_t* d = reinterpret_cast<uint32_t*>(dst);
#pragma unroll
    for (int i = 0; i < 9; i++)
        d[i] = s[i];
}

PT_DEV void reset_reservoir(ReSTIRDIReservoir* r)
{
    r->M = 0; r->weight_sum = 0.0f; r->UCW = 0.0f;
    r->sample.emissive_triangle_index = -1;
    r->sample.point_on_light_source.x = 0.0f; r->sample.point_on_light_source.y = 0.0f; r->sample.point_on_light_source.z = 0.0f;
    r->sample.target_function = 0.0f;
    r->sample.flags = 0;
}

// reset_render (CameraRays.h:19-43)
PT_DEV void reset_render(const HIPRTRenderData& rd, uint32_t pixel_index)
{
    if (rd.render_settings.accumulate && rd.aux_buffers.restir_reservoir_buffer_1 != nullptr)
    {
        reset_reservoir(rd.aux_buffers.restir_reservoir_buffer_1 + pixel_index);
        reset_reservoir(rd.aux_buffers.restir_reservoir_buffer_2 + pixel_index);
        reset_reservoir(rd.aux_buffers.restir_reservoir_buffer_3 + pixel_index);
    }
    if (has_access_to_adaptive_sampling_buffers(rd.render_settings))
    {
        rd.aux_buffers.pixel_sample_count[pixel_index] = 0;
        rd.aux_buffers.pixel_squared_luminance[pixel_index] = 0;
        rd.aux_buffers.pixel_converged_sample_count[pixel_index] = -1;
    }
}

// HIPRTCamera::get_camera_ray (HIPRTCamera.h:27-48)
PT_DEV void get_camera_ray(const HIPRTCamera& cam, float x, float y, int2 res, float3& origin, float3& direction)
{
    float x_ndc = x / res.x * 2 - 1;
    float y_ndc = y / res.y * 2 - 1;
    const Mat4& inv_view = as_mat4(cam.inverse_view);
    float3 ray_origin = matrix_X_point(inv_view, f3(0.0f, 0.0f, 0.0f));
    float3 p_vs = matrix_X_point(as_mat4(cam.inverse_projection), f3(x_ndc, y_ndc, -1.0f));
    float3 p_ws = matrix_X_point(inv_view, p_vs);
    origin = ray_origin;
    direction = normalize(p_ws - ray_origin);
}

PT_DEV unsigned int pixel_seed(const HIPRTRenderData& rd, uint32_t pixel_index)
{
    if (rd.render_settings.freeze_random)
        return wang_hash(pixel_index + 1);
    return wang_hash((pixel_index + 1) * (rd.render_settings.sample_number + 1) * rd.random_seed);
}

template <bool USE_PREV_GBUFFER>
PT_DEV void camera_rays_pixel(const HIPRTRenderData& rd, int2 res, uint32_t x, uint32_t y)
{
    uint32_t pixel_index = x + y * res.x;

    if (do_render_low_resolution(rd.render_settings))
    {
        int res_scaling = rd.render_settings.render_low_resolution_scaling;
        if (x % res_scaling != 0 || y % res_scaling != 0)
        {
            rd.aux_buffers.pixel_active[pixel_index] = false;
            return;
        }
        pixel_index /= res_scaling;
    }

    if (USE_PREV_GBUFFER && rd.render_settings.restir_di_settings.temporal_pass.do_temporal_reuse_pass)
    {
        // use_prev_frame_g_buffer(): LSS_RESTIR_DI && do_temporal_reuse_pass (RenderSettings.h:232-238)
        st3(rd.g_buffer_prev_frame.geometric_normals, pixel_index, ld3_rw(rd.g_buffer.geometric_normals, pixel_index));
        st3(rd.g_buffer_prev_frame.shading_normals, pixel_index, ld3_rw(rd.g_buffer.shading_normals, pixel_index));
        {
            Material m;
            load_material(m, rd.g_buffer.materials + pixel_index);
            store_material(rd.g_buffer_prev_frame.materials + pixel_index, m);
        }
        st3(rd.g_buffer_prev_frame.first_hits, pixel_index, ld3_rw(rd.g_buffer.first_hits, pixel_index));
        rd.g_buffer_prev_frame.first_hit_prim_index[pixel_index] = rd.g_buffer.first_hit_prim_index[pixel_index];
        copy_volume_state(rd.g_buffer_prev_frame.ray_volume_states + pixel_index, rd.g_buffer.ray_volume_states + pixel_index);
        st3(rd.g_buffer_prev_frame.view_directions, pixel_index, ld3_rw(rd.g_buffer.view_directions, pixel_index));
        rd.g_buffer_prev_frame.camera_ray_hit[pixel_index] = rd.g_buffer.camera_ray_hit[pixel_index];
    }

    if (rd.render_settings.sample_number == 0 || rd.render_settings.need_to_reset)
        reset_render(rd, pixel_index);

    bool pixel_converged = false;
    bool sampling_needed = adaptive_sampling(rd, pixel_index, pixel_converged);

    if (pixel_converged || !sampling_needed)
    {
        if (rd.render_settings.do_update_status_buffers)
            atomicAdd(rd.aux_buffers.stop_noise_threshold_converged_count, 1u);
    }

    if (has_access_to_adaptive_sampling_buffers(rd.render_settings))
    {
        if (!sampling_needed)
        {
            RGB px(ld3_rw(rd.buffers.pixels, pixel_index));
            px = px / rd.render_settings.sample_number * (rd.render_settings.sample_number + 1);
            st3(rd.buffers.pixels, pixel_index, to_f3(px));
            rd.aux_buffers.pixel_active[pixel_index] = false;
            return;
        }
        else
            rd.aux_buffers.pixel_sample_count[pixel_index]++;
    }

    Rng rng(pixel_seed(rd, pixel_index));

    float x_dir = (x + 0.5f);
    float y_dir = (y + 0.5f);
    if (rd.current_camera.do_jittering)
    {
        x_dir += rng() - 0.5f;
        y_dir += rng() - 0.5f;
    }

    float3 origin, direction;
    get_camera_ray(rd.current_camera, x_dir, y_dir, res, origin, direction);

    RayPayload payload;
    payload.throughput = RGB(1.0f);
    payload.ray_color = RGB(0.0f);
    payload.next_ray_state = 0;
    volume_state_init(payload.volume_state);

    HitInfo hit;
    bool intersection_found = trace_ray(rd, origin, direction, payload, hit, -1, rng);

    if (intersection_found)
    {
        if (material_is_emissive(payload.material) && dot(-direction, hit.geometric_normal) < 0)
        {
            hit.geometric_normal = -hit.geometric_normal;
            hit.shading_normal = -hit.shading_normal;
        }
        st3(rd.g_buffer.geometric_normals, pixel_index, hit.geometric_normal);
        st3(rd.g_buffer.shading_normals, pixel_index, hit.shading_normal);
        store_material(rd.g_buffer.materials + pixel_index, payload.material);
        st3(rd.g_buffer.first_hits, pixel_index, hit.inter_point);
        copy_volume_state(rd.g_buffer.ray_volume_states + pixel_index, &payload.volume_state);
    }

    rd.g_buffer.first_hit_prim_index[pixel_index] = intersection_found ? hit.primitive_index : -1;
    st3(rd.g_buffer.view_directions, pixel_index, -direction);
    rd.g_buffer.camera_ray_hit[pixel_index] = intersection_found;
    rd.aux_buffers.pixel_active[pixel_index] = true;

    if (rd.render_settings.do_update_status_buffers)
        rd.aux_buffers.still_one_ray_active[0] = 1;
}
} // namespace pt
