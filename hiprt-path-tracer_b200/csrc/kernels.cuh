// The render kernels of the hot path: CameraRays and FullPathTracer (megakernel form, one thread per pixel).
// reference: src/Device/kernels/CameraRays.h:19-179, src/Device/kernels/FullPathTracer.h:74-327,
//            src/Device/includes/AdaptiveSampling.h:11-104, RussianRoulette.h:14-49, HIPRTCamera.h:27-48.
#pragma once

#include "envmap.cuh"

namespace pt
{
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

} // namespace pt
#include "restir_di.cuh"
namespace pt {

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

// Seed of a pixel's generator in the sample pass (sample_number, random_seed): CameraRays.h:127-132, FullPathTracer.h:124-129
PT_DEV unsigned int pixel_seed_of_pass(const HIPRTRenderData& rd, uint32_t pixel_index, int sample_number, unsigned int random_seed)
{
    if (rd.render_settings.freeze_random)
        return wang_hash(pixel_index + 1);
    return wang_hash((pixel_index + 1) * (sample_number + 1) * random_seed);
}

PT_DEV unsigned int pixel_seed(const HIPRTRenderData& rd, uint32_t pixel_index)
{
    return pixel_seed_of_pass(rd, pixel_index, rd.render_settings.sample_number, rd.random_seed);
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
            alignas(16) Material m;
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

// do_russian_roulette (RussianRoulette.h:14-49). Returns false if the path is killed.
PT_DEV bool do_russian_roulette(const HIPRTRenderSettings& s, int bounce, RGB& throughput, RGB current_weight, Rng& rng)
{
    if (bounce >= s.russian_roulette_min_depth && s.use_russian_roulette)
    {
        float survive_probability = 0.0f;
        if (s.path_russian_roulette_method == MAX_THROUGHPUT)
            survive_probability = throughput.max_component();
        else if (s.path_russian_roulette_method == ARNOLD_2014)
        {
            survive_probability = (throughput * current_weight).max_component() / throughput.max_component();
            survive_probability = sqrtf(survive_probability);
        }
        survive_probability = minf(survive_probability, 1.0f);
        if (rng() > survive_probability)
            return false;
        float throughput_increase = 1.0f / survive_probability;
        if (s.russian_roulette_throughput_clamp > 0.0f)
            throughput_increase = minf(throughput_increase, s.russian_roulette_throughput_clamp);
        throughput *= throughput_increase;
    }
    return true;
}

// sample_one_light (Lights.h:277-321) with sample_many_lights (:222-242) and sample_one_light_ReSTIR_DI (:244-275)
template <int LSS, int LATER, int GGX_SAMPLER>
PT_DEV RGB sample_one_light(const HIPRTRenderData& rd, RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng, int2 pixel_coords, int2 res, int bounce)
{
    if (rd.buffers.emissive_triangles_count == 0 && !(rd.world_settings.ambient_light_type == AMBIENT_ENVMAP && LSS == 5))
        return RGB(0.0f);
    if (rd.bsdfs_data.white_furnace_mode && rd.bsdfs_data.white_furnace_mode_turn_off_emissives)
        return RGB(0.0f);
    if (material_is_emissive(payload.material))
    {
        if (payload.material.emissive_texture_used && bounce > 0)
            return material_emission(payload.material);
        return RGB(0.0f);
    }
    if (LSS == 0)
        return RGB(0.0f);

    RGB direct;
    if (LSS == 5 && bounce == 0)
        return sample_light_ReSTIR_DI<GGX_SAMPLER>(rd, payload, hit, view_direction, rng, pixel_coords, res);

    const int strategy = LSS == 5 ? (LATER == 0 ? 1 : LATER == 1 ? 2 : LATER == 2 ? 3 : 4) : LSS;
    for (int i = 0; i < rd.render_settings.number_of_light_samples; i++)
    {
        if (strategy == 1)
            direct += sample_one_light_no_MIS(rd, payload, hit, view_direction, rng);
        else if (strategy == 2)
            direct += sample_one_light_bsdf<GGX_SAMPLER>(rd, payload, hit, view_direction, rng);
        else if (strategy == 3)
            direct += sample_one_light_MIS<GGX_SAMPLER>(rd, payload, hit, view_direction, rng);
        else
            direct += sample_lights_RIS<GGX_SAMPLER>(rd, payload, hit, view_direction, rng);
    }
    return direct / rd.render_settings.number_of_light_samples;
}

PT_DEV bool sanity_check(const HIPRTRenderData& rd, RGB& ray_color, float sampled_wavelength, uint32_t pixel_index, int sample_number)
{
    bool invalid = false;
    if (sampled_wavelength == 0.0f)
        invalid |= ray_color.r < 0 || ray_color.g < 0 || ray_color.b < 0;
    invalid |= ray_color.has_NaN();
    if (invalid)
    {
        if (rd.render_settings.display_NaNs)
        {
            // debug_set_final_color (FullPathTracer.h:27-33)
            RGB c(1.0e30f, 0.0f, 1.0e30f);
            if (sample_number != 0)
                c = c * sample_number;
            st3(rd.buffers.pixels, pixel_index, to_f3(c));
        }
        else
            ray_color = RGB(0.0f);
    }
    return !invalid;
}

// One path vertex with an intersection (FullPathTracer.h:165-241): NEE + envmap sampling, emission, BSDF sample,
// russian roulette, throughput update. Returns false when the path ends here. Shared by the megakernel and the
// wavefront kernels so that both evaluate exactly the same arithmetic in the same RNG order.
// First half of a vertex: direct lighting (NEE + envmap sampling with their MIS rays), emission, the three clamps.
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void shade_vertex_direct(const HIPRTRenderData& rd, RayPayload& payload, HitInfo& hit, float3 direction, Rng& rng, int2 pixel_coords, int2 res, int bounce)
{
    if (material_is_emissive(payload.material) && dot(-direction, hit.geometric_normal) < 0)
    {
        hit.geometric_normal = -hit.geometric_normal;
        hit.shading_normal = -hit.shading_normal;
    }

    RGB light_direct = sample_one_light<LSS, LATER, GGX_SAMPLER>(rd, payload, hit, -direction, rng, pixel_coords, res, bounce);
    RGB envmap_direct = sample_environment_map<LSS, ESS, GGX_SAMPLER>(rd, payload, hit, -direction, bounce, rng);

    light_direct = clamp_light_contribution(light_direct, rd.render_settings.direct_contribution_clamp, bounce == 0);
    envmap_direct = clamp_light_contribution(envmap_direct, rd.render_settings.envmap_contribution_clamp, bounce == 0);

    if (LSS == 0)
    {
        RGB hit_emission = material_emission(payload.material);
        hit_emission = clamp_light_contribution(hit_emission, rd.render_settings.indirect_contribution_clamp, bounce > 0);
        payload.ray_color += hit_emission * payload.throughput;
    }
    else
    {
        if (bounce == 0)
            payload.ray_color += material_emission(payload.material);
        RGB indirect = (light_direct + envmap_direct) * payload.throughput;
        payload.ray_color += clamp_light_contribution(indirect, rd.render_settings.indirect_contribution_clamp, bounce > 0);
    }
}

// Second half: BSDF sample, russian roulette, throughput update. `hit` carries the normals as shade_vertex_direct left
// them (flipped towards the viewer on emissive back faces). Returns false when the path ends here.
template <int GGX_SAMPLER>
PT_DEV bool shade_vertex_bounce(const HIPRTRenderData& rd, RayPayload& payload, const HitInfo& hit, float3& origin, float3& direction, Rng& rng, int bounce)
{
    float bsdf_pdf;
    float3 bounce_direction;
    RGB bsdf_color = bsdf_sample<GGX_SAMPLER>(rd, payload.material, payload.volume_state, -direction, hit.shading_normal, hit.geometric_normal, bounce_direction, bsdf_pdf, rng);
    RGB throughput_attenuation = bsdf_color * fabsf(dot(bounce_direction, hit.shading_normal)) / bsdf_pdf;

    if (bsdf_pdf <= 0.0f)
        return false;
    if (!do_russian_roulette(rd.render_settings, bounce, payload.throughput, throughput_attenuation, rng))
        return false;

    payload.throughput *= get_dispersion_ray_color(payload.volume_state.sampled_wavelength, payload.material.dispersion_scale);
    payload.throughput *= throughput_attenuation;
    payload.next_ray_state = 0;

    origin = hit.inter_point;
    direction = bounce_direction;
    return true;
}

template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV bool shade_vertex(const HIPRTRenderData& rd, RayPayload& payload, HitInfo& hit, float3& origin, float3& direction, Rng& rng, int2 pixel_coords, int2 res, int bounce)
{
    shade_vertex_direct<LSS, ESS, LATER, GGX_SAMPLER>(rd, payload, hit, direction, rng, pixel_coords, res, bounce);
    return shade_vertex_bounce<GGX_SAMPLER>(rd, payload, hit, origin, direction, rng, bounce);
}

// The ray left the scene (FullPathTracer.h:243-286): sky / envmap radiance
template <int ESS>
PT_DEV void miss_vertex(const HIPRTRenderData& rd, RayPayload& payload, float3 direction, int bounce)
{
    RGB skysphere_color;
    if (rd.world_settings.ambient_light_type == AMBIENT_UNIFORM || rd.bsdfs_data.white_furnace_mode)
        skysphere_color = rgb(rd.world_settings.uniform_light_color);
    else if (rd.world_settings.ambient_light_type == AMBIENT_ENVMAP)
    {
        if (ESS == 0 || bounce == 0)
        {
            skysphere_color = eval_envmap_no_pdf(rd.world_settings, direction);
            bool unscale = ESS == 0 ? (!rd.world_settings.envmap_scale_background_intensity && bounce == 0) : !rd.world_settings.envmap_scale_background_intensity;
            if (unscale)
                skysphere_color /= rd.world_settings.envmap_intensity;
        }
    }

    skysphere_color = clamp_light_contribution(skysphere_color, rd.render_settings.envmap_contribution_clamp, true);
    RGB indirect = skysphere_color * payload.throughput;
    payload.ray_color += clamp_light_contribution(indirect, rd.render_settings.indirect_contribution_clamp, bounce > 0);
    payload.next_ray_state = 1;
}

// End of a path (FullPathTracer.h:288-326): sanity check, then accumulation of the sample into the frame buffers.
// The denoiser AOVs only depend on the camera hit, so they are re-read from the G-buffer here.
// `sample_number` / `aov_counter` are the render settings' values of the sample pass the colour belongs to (a sample batch
// accumulates several passes in one kernel, in pass order: wf_accumulate_pixel).
PT_DEV void accumulate_sample_at(const HIPRTRenderData& rd, uint32_t pixel_index, RGB ray_color, float sampled_wavelength, float3 denoiser_normal, RGB denoiser_albedo,
                                 const int sample_number, const int aov_counter)
{
    if (!sanity_check(rd, ray_color, sampled_wavelength, pixel_index, sample_number))
        return;

    float squared_luminance_of_samples = 0.0f;
    squared_luminance_of_samples += ray_color.luminance() * ray_color.luminance();
    rd.aux_buffers.still_one_ray_active[0] = 1;

    if (has_access_to_adaptive_sampling_buffers(rd.render_settings))
        rd.aux_buffers.pixel_squared_luminance[pixel_index] += squared_luminance_of_samples;

    if (sample_number == 0)
        st3(rd.buffers.pixels, pixel_index, to_f3(ray_color));
    else
        st3(rd.buffers.pixels, pixel_index, ld3_rw(rd.buffers.pixels, pixel_index) + to_f3(ray_color));

    const float counter = (float)aov_counter;
    if (rd.aux_buffers.denoiser_albedo != nullptr)
    {
        if (sample_number == 0)
            st3(rd.aux_buffers.denoiser_albedo, pixel_index, to_f3(denoiser_albedo));
        else
        {
            RGB prev(ld3_rw(rd.aux_buffers.denoiser_albedo, pixel_index));
            st3(rd.aux_buffers.denoiser_albedo, pixel_index, to_f3((prev * counter + denoiser_albedo) / (counter + 1.0f)));
        }
    }
    if (rd.aux_buffers.denoiser_normals != nullptr)
    {
        if (sample_number == 0)
            st3(rd.aux_buffers.denoiser_normals, pixel_index, denoiser_normal);
        else
        {
            float3 accumulated = (ld3_rw(rd.aux_buffers.denoiser_normals, pixel_index) * counter + denoiser_normal) / (counter + 1.0f);
            float normal_length = length(accumulated);
            if (!is_zero(normal_length))
                st3(rd.aux_buffers.denoiser_normals, pixel_index, accumulated / normal_length);
        }
    }
}

PT_DEV void accumulate_sample(const HIPRTRenderData& rd, uint32_t pixel_index, RGB ray_color, float sampled_wavelength, float3 denoiser_normal, RGB denoiser_albedo)
{
    accumulate_sample_at(rd, pixel_index, ray_color, sampled_wavelength, denoiser_normal, denoiser_albedo, rd.render_settings.sample_number,
                         rd.render_settings.denoiser_AOV_accumulation_counter);
}

// Denoiser AOVs of a pixel from its G-buffer entry (FullPathTracer.h:167-171: first hit's shading normal + base colour)
PT_DEV void denoiser_aovs_from_g_buffer(const HIPRTRenderData& rd, uint32_t pixel_index, float3& denoiser_normal, RGB& denoiser_albedo)
{
    denoiser_normal = f3(0.0f, 0.0f, 0.0f);
    denoiser_albedo = RGB(0.0f);
    if (rd.g_buffer.camera_ray_hit[pixel_index] == 1)
    {
        denoiser_normal += normalize(ld3_rw(rd.g_buffer.shading_normals, pixel_index));
        const ColorRGB32F& bc = rd.g_buffer.materials[pixel_index].base_color;
        denoiser_albedo += RGB(bc.r, bc.g, bc.b);
    }
}

// Megakernel form: one thread walks the whole path of its pixel.
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void full_path_tracer_pixel(const HIPRTRenderData& rd, int2 res, uint32_t x, uint32_t y)
{
    uint32_t pixel_index = x + y * res.x;
    if (!rd.aux_buffers.pixel_active[pixel_index])
        return;

    int nb_bounces = rd.render_settings.nb_bounces;
    if (do_render_low_resolution(rd.render_settings))
        nb_bounces = mini(3, nb_bounces);

    Rng rng(pixel_seed(rd, pixel_index));

    RGB denoiser_albedo(0.0f);
    float3 denoiser_normal = f3(0.0f, 0.0f, 0.0f);

    HitInfo hit;
    hit.inter_point = ld3_rw(rd.g_buffer.first_hits, pixel_index);
    hit.geometric_normal = normalize(ld3_rw(rd.g_buffer.geometric_normals, pixel_index));
    hit.shading_normal = normalize(ld3_rw(rd.g_buffer.shading_normals, pixel_index));
    hit.primitive_index = rd.g_buffer.first_hit_prim_index[pixel_index];

    float3 origin = f3(0.0f, 0.0f, 0.0f);
    float3 direction = normalize(-ld3_rw(rd.g_buffer.view_directions, pixel_index));
    bool intersection_found = rd.g_buffer.camera_ray_hit[pixel_index] == 1;

    RayPayload payload;
    payload.throughput = RGB(1.0f);
    payload.ray_color = RGB(0.0f);
    payload.next_ray_state = 0;
    load_material(payload.material, rd.g_buffer.materials + pixel_index);
    copy_volume_state(&payload.volume_state, rd.g_buffer.ray_volume_states + pixel_index);

    for (int bounce = 0; bounce < nb_bounces + 1; bounce++)
    {
        if (bounce > 0)
            intersection_found = trace_ray(rd, origin, direction, payload, hit, hit.primitive_index, rng);

        if (intersection_found)
        {
            if (bounce == 0)
            {
                denoiser_normal += hit.shading_normal;
                denoiser_albedo += rgb(payload.material.base_color);
            }
            if (!shade_vertex<LSS, ESS, LATER, GGX_SAMPLER>(rd, payload, hit, origin, direction, rng, make_int2(x, y), res, bounce))
                break;
        }
        else
        {
            miss_vertex<ESS>(rd, payload, direction, bounce);
            break;
        }
    }

    accumulate_sample(rd, pixel_index, payload.ray_color, payload.volume_state.sampled_wavelength, denoiser_normal, denoiser_albedo);
}
} // namespace pt
