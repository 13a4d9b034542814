// ORACLE / TEST INFRASTRUCTURE ONLY. Included at the end of ref_harness.cpp.
// Render loop, ReSTIR DI pass driver, ray-batch and BSDF known-answer entry points.

namespace
{
// src/Renderer/CPURenderer.cpp:436-485
void restir_presampling_pass(RefCtx* c)
{
    if (ReSTIR_DI_DoLightsPresampling != KERNEL_OPTION_TRUE)
        return;

    HIPRTRenderData& rd = c->rd;
    LightPresamplingParameters p;
    p.number_of_subsets = rd.render_settings.restir_di_settings.light_presampling.number_of_subsets;
    p.subset_size = rd.render_settings.restir_di_settings.light_presampling.subset_size;
    p.out_light_samples = c->presampled_lights.data();
    p.emissive_triangles_count = rd.buffers.emissive_triangles_count;
    p.emissive_triangles_indices = rd.buffers.emissive_triangles_indices;
    p.triangles_indices = rd.buffers.triangles_indices;
    p.vertices_positions = rd.buffers.vertices_positions;
    p.material_indices = rd.buffers.material_indices;
    p.materials = rd.buffers.materials_buffer;
    p.world_settings = rd.world_settings;
    p.freeze_random = rd.render_settings.freeze_random;
    p.sample_number = rd.render_settings.sample_number;
    p.random_seed = c->rng.xorshift32();
    p.envmap_sampling_probability = rd.render_settings.restir_di_settings.initial_candidates.envmap_candidate_probability;

    // Serial in the reference (:481-484); every index is independent (seeded per index) so a parallel loop gives the same buffer
    int total = p.number_of_subsets * p.subset_size;
#pragma omp parallel for schedule(static)
    for (int index = 0; index < total; index++)
        ReSTIR_DI_LightsPresampling(p, index);
}

// src/Renderer/CPURenderer.cpp:487-501
void restir_initial_candidates_pass(RefCtx* c)
{
    HIPRTRenderData& rd = c->rd;
    rd.random_seed = c->rng.xorshift32();
    rd.render_settings.restir_di_settings.light_presampling.light_samples = c->presampled_lights.data();
    rd.render_settings.restir_di_settings.initial_candidates.output_reservoirs = c->initial_candidates_reservoirs.data();
    for_each_pixel(c->res, [&](int x, int y) { ReSTIR_DI_InitialCandidates(rd, c->res, x, y); });
}

// src/Renderer/CPURenderer.cpp:401-434 with the configure_* helpers of :503-633
void restir_di(RefCtx* c)
{
    HIPRTRenderData& rd = c->rd;
    ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;

    restir_presampling_pass(c);
    restir_initial_candidates_pass(c);

    if (rs.do_fused_spatiotemporal)
    {
        // :535-552
        rd.random_seed = c->rng.xorshift32();
        rs.temporal_pass.permutation_sampling_random_bits = c->rng.xorshift32();
        rs.temporal_pass.input_reservoirs = c->odd_frame ? c->spatial_output_1.data() : c->spatial_output_2.data();
        rs.temporal_pass.output_reservoirs = nullptr;
        // :590-608
        rs.spatial_pass.input_reservoirs = rs.temporal_pass.input_reservoirs;
        rs.spatial_pass.output_reservoirs = c->odd_frame ? c->spatial_output_2.data() : c->spatial_output_1.data();

        for_each_pixel(c->res, [&](int x, int y) { ReSTIR_DI_SpatiotemporalReuse(rd, c->res, x, y); });
    }
    else
    {
        if (rs.temporal_pass.do_temporal_reuse_pass)
        {
            // :503-533
            rd.random_seed = c->rng.xorshift32();
            rs.temporal_pass.permutation_sampling_random_bits = c->rng.xorshift32();
            rs.temporal_pass.input_reservoirs = rs.restir_output_reservoirs;
            if (rs.spatial_pass.do_spatial_reuse_pass)
                rs.temporal_pass.output_reservoirs = c->initial_candidates_reservoirs.data();
            else
                rs.temporal_pass.output_reservoirs = c->odd_frame ? c->spatial_output_1.data() : c->spatial_output_2.data();

            for_each_pixel(c->res, [&](int x, int y) { ReSTIR_DI_TemporalReuse(rd, c->res, x, y); });
        }

        if (rs.spatial_pass.do_spatial_reuse_pass)
        {
            for (int pass = 0; pass < rs.spatial_pass.number_of_passes; pass++)
            {
                // :554-588
                rd.random_seed = c->rng.xorshift32();
                if (pass == 0)
                {
                    rs.spatial_pass.input_reservoirs = rs.temporal_pass.do_temporal_reuse_pass ? rs.temporal_pass.output_reservoirs : rs.initial_candidates.output_reservoirs;
                    rs.spatial_pass.output_reservoirs = c->spatial_output_1.data();
                }
                else if ((pass & 1) == 0)
                {
                    rs.spatial_pass.input_reservoirs = c->spatial_output_2.data();
                    rs.spatial_pass.output_reservoirs = c->spatial_output_1.data();
                }
                else
                {
                    rs.spatial_pass.input_reservoirs = c->spatial_output_1.data();
                    rs.spatial_pass.output_reservoirs = c->spatial_output_2.data();
                }

                for_each_pixel(c->res, [&](int x, int y) { ReSTIR_DI_SpatialReuse(rd, c->res, x, y); });
            }
        }
    }

    // :619-633
    if (rs.spatial_pass.do_spatial_reuse_pass)
        rs.restir_output_reservoirs = rs.spatial_pass.output_reservoirs;
    else if (rs.temporal_pass.do_temporal_reuse_pass)
        rs.restir_output_reservoirs = rs.temporal_pass.output_reservoirs;
    else
        rs.restir_output_reservoirs = rs.initial_candidates.output_reservoirs;

    c->odd_frame = !c->odd_frame;
}
} // namespace

extern "C" {

// CPURenderer::render, src/Renderer/CPURenderer.cpp:264-296. 'nsamples' plays samples_per_frame.
// passes_mask: bit0 CameraRays, bit1 ReSTIR DI, bit2 FullPathTracer (7 = the reference loop).
double ref_render(void* h, int nsamples, int passes_mask)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    HIPRTRenderData& rd = c->rd;
    HIPRTCamera camera = rd.current_camera;

    auto t0 = std::chrono::steady_clock::now();
    for (int frame = 1; frame <= nsamples; frame++)
    {
        rd.render_settings.do_update_status_buffers = true;

        // update(), :298-309
        c->still_one_ray_active = 0;
        c->stop_noise_threshold_count.store(0);
        // update_render_data(), :311-315 (static camera)
        rd.prev_camera = rd.current_camera;
        rd.current_camera = camera;

        if (passes_mask & 1)
            for_each_pixel(c->res, [&](int x, int y) { CameraRays(rd, c->res, x, y); });
#if DirectLightSamplingStrategy == LSS_RESTIR_DI
        if (passes_mask & 2)
            restir_di(c);
#endif
        if (passes_mask & 4)
            for_each_pixel(c->res, [&](int x, int y) { FullPathTracer(rd, c->res, x, y); });

        if (rd.render_settings.accumulate)
            rd.render_settings.sample_number++;
        rd.random_seed = c->rng.xorshift32();
        rd.render_settings.need_to_reset = false;
    }
    c->last_render_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return c->last_render_seconds;
}

unsigned int ref_current_seed(void* h) { return reinterpret_cast<RefCtx*>(h)->rd.random_seed; }
int ref_sample_number(void* h) { return reinterpret_cast<RefCtx*>(h)->rd.render_settings.sample_number; }

void ref_get_pixels(void* h, float* out)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    std::memcpy(out, c->pixels.data(), c->pixels.size() * sizeof(ColorRGB32F));
}

// which: 0 first_hit_prim_index(int) 1 shading_normals 2 geometric_normals 3 view_directions 4 first_hits 5 camera_ray_hit(u8)
//        6 materials(256 B) 7 ray_volume_states(36 B) 8 pixel_active(u8) 9 pixel_sample_count(int) 10 pixel_squared_luminance(float)
//        11 pixel_converged_sample_count(int) 12 denoiser_albedo 13 denoiser_normals 14 restir output reservoirs 15 initial candidates
//        16 presampled lights
long long ref_get_buffer(void* h, int which, void* out, long long capacity_bytes)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    const void* src = nullptr;
    size_t bytes = 0;
    switch (which)
    {
    case 0: src = c->g_buffer.first_hit_prim_index.data(); bytes = c->g_buffer.first_hit_prim_index.size() * sizeof(int); break;
    case 1: src = c->g_buffer.shading_normals.data(); bytes = c->g_buffer.shading_normals.size() * sizeof(float3); break;
    case 2: src = c->g_buffer.geometric_normals.data(); bytes = c->g_buffer.geometric_normals.size() * sizeof(float3); break;
    case 3: src = c->g_buffer.view_directions.data(); bytes = c->g_buffer.view_directions.size() * sizeof(float3); break;
    case 4: src = c->g_buffer.first_hits.data(); bytes = c->g_buffer.first_hits.size() * sizeof(float3); break;
    case 5: src = c->g_buffer.cameray_ray_hit.data(); bytes = c->g_buffer.cameray_ray_hit.size(); break;
    case 6: src = c->g_buffer.materials.data(); bytes = c->g_buffer.materials.size() * sizeof(SimplifiedRendererMaterial); break;
    case 7: src = c->g_buffer.ray_volume_states.data(); bytes = c->g_buffer.ray_volume_states.size() * sizeof(RayVolumeState); break;
    case 8: src = c->pixel_active.data(); bytes = c->pixel_active.size(); break;
    case 9: src = c->pixel_sample_count.data(); bytes = c->pixel_sample_count.size() * sizeof(int); break;
    case 10: src = c->pixel_squared_luminance.data(); bytes = c->pixel_squared_luminance.size() * sizeof(float); break;
    case 11: src = c->pixel_converged_sample_count.data(); bytes = c->pixel_converged_sample_count.size() * sizeof(int); break;
    case 12: src = c->denoiser_albedo.data(); bytes = c->denoiser_albedo.size() * sizeof(ColorRGB32F); break;
    case 13: src = c->denoiser_normals.data(); bytes = c->denoiser_normals.size() * sizeof(float3); break;
    case 14: src = c->rd.render_settings.restir_di_settings.restir_output_reservoirs; bytes = c->pixels.size() * sizeof(ReSTIRDIReservoir); break;
    case 15: src = c->initial_candidates_reservoirs.data(); bytes = c->initial_candidates_reservoirs.size() * sizeof(ReSTIRDIReservoir); break;
    case 16: src = c->presampled_lights.data(); bytes = c->presampled_lights.size() * sizeof(ReSTIRDIPresampledLight); break;
    default: return -1;
    }
    if ((long long)bytes > capacity_bytes) return -(long long)bytes;
    std::memcpy(out, src, bytes);
    return (long long)bytes;
}

// CPURenderer::tonemap (src/Renderer/CPURenderer.cpp:663-686) on a copy of the accumulation buffer, with the reference's own
// ColorRGB32F exp / pow (HostDeviceCommon/Color.h), then the 8-bit quantisation of Image32Bit::write_image_png
// (src/Image/Image.cpp:427-437: clamp(0, 255, v * 255) truncated to unsigned char). out_float / out_u8: W*H*3 each, may be null.
void ref_tonemap(void* h, float gamma, float exposure, float* out_float, unsigned char* out_u8)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    const size_t n = c->pixels.size();
    for (size_t index = 0; index < n; index++)
    {
        ColorRGB32F hdr_color = c->pixels[index];
        if (c->rd.render_settings.accumulate)
            hdr_color = hdr_color / float(c->rd.render_settings.sample_number);
        ColorRGB32F tone_mapped = ColorRGB32F(1.0f) - exp(-hdr_color * exposure);
        tone_mapped = pow(tone_mapped, 1.0f / gamma);
        const float v[3] = {tone_mapped.r, tone_mapped.g, tone_mapped.b};
        for (int k = 0; k < 3; k++)
        {
            if (out_float) out_float[index * 3 + k] = v[k];
            if (out_u8) out_u8[index * 3 + k] = (unsigned char)hippt::clamp(0.0f, 255.0f, v[k] * 255.0f);
        }
    }
}

unsigned long long ref_ray_count() { return g_ray_count.load(); }
void ref_ray_count_reset() { g_ray_count.store(0); }

void ref_ray_log_begin(long long capacity, int stride)
{
    delete[] g_log;
    g_log = new RayLogEntry[capacity];
    g_log_capacity = capacity;
    g_log_stride = stride < 1 ? 1 : stride;
    g_log_cursor.store(0);
}

// Returns the number of entries copied (11 x 4 bytes each: o.xyz d.xyz last_prim hit_prim t u v)
long long ref_ray_log_end(void* out, long long capacity_entries)
{
    long long n = std::min(g_log_cursor.load(), g_log_capacity);
    n = std::min(n, capacity_entries);
    if (out != nullptr && n > 0)
        std::memcpy(out, g_log, n * sizeof(RayLogEntry));
    delete[] g_log;
    g_log = nullptr;
    g_log_capacity = 0;
    return n;
}

// Closest hit of BVH::intersect (src/Renderer/BVH.h:132-227) through intersect_scene_cpu (src/Device/includes/Intersect.h:87-108)
// rays: n x 6 floats (origin, direction). Alpha testing is whatever render_settings.do_alpha_testing says.
void ref_trace_closest(void* h, const float* rays, const int* last_prim, long long n, int* out_prim, float* out_t, float* out_uv)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
#pragma omp parallel for schedule(dynamic, 1024)
    for (long long i = 0; i < n; i++)
    {
        hiprtRay ray;
        ray.origin = make_float3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
        ray.direction = make_float3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
        Xorshift32Generator rng(42);
        hiprtHit hit = intersect_scene_cpu(c->rd, ray, last_prim ? last_prim[i] : -1, rng);
        out_prim[i] = hit.hasHit() ? (int)hit.primID : -1;
        out_t[i] = hit.hasHit() ? hit.t : -1.0f;
        if (out_uv) { out_uv[i * 2 + 0] = hit.uv.x; out_uv[i * 2 + 1] = hit.uv.y; }
    }
}

// evaluate_shadow_ray, src/Device/includes/Intersect.h:224-286 (CPU branch)
void ref_trace_shadow(void* h, const float* rays, const float* t_max, const int* last_prim, long long n, unsigned char* out_occluded)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
#pragma omp parallel for schedule(dynamic, 1024)
    for (long long i = 0; i < n; i++)
    {
        hiprtRay ray;
        ray.origin = make_float3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
        ray.direction = make_float3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
        Xorshift32Generator rng(42);
        out_occluded[i] = evaluate_shadow_ray(c->rd, ray, t_max[i], last_prim ? last_prim[i] : -1, rng) ? 1 : 0;
    }
}

// HIPRTCamera::get_camera_ray, src/HostDeviceCommon/HIPRTCamera.h:27-48
void ref_camera_rays(void* h, const float* xy, long long n, int w, int hgt, float* out_rays)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    for (long long i = 0; i < n; i++)
    {
        hiprtRay r = c->rd.current_camera.get_camera_ray(xy[i * 2], xy[i * 2 + 1], make_int2(w, hgt));
        out_rays[i * 6 + 0] = r.origin.x; out_rays[i * 6 + 1] = r.origin.y; out_rays[i * 6 + 2] = r.origin.z;
        out_rays[i * 6 + 3] = r.direction.x; out_rays[i * 6 + 4] = r.direction.y; out_rays[i * 6 + 5] = r.direction.z;
    }
}

// ---- known-answer entry points --------------------------------------------------------------
unsigned int ref_wang_hash(unsigned int x) { return wang_hash(x); }

void ref_xorshift(unsigned int seed, int n, unsigned int* out_u32, float* out_float, int* out_index, int index_range)
{
    Xorshift32Generator a(seed), b(seed), d(seed);
    for (int i = 0; i < n; i++)
    {
        out_u32[i] = a.xorshift32();
        out_float[i] = b();
        out_index[i] = d.random_index(index_range);
    }
}

// bsdf_dispatcher_eval, src/Device/includes/Dispatcher.h:18-44. vecs: n x 9 floats (view, shading normal, to_light), all world space.
// materials: n x 256 B; volume states: n x 36 B (in/out). out: n x 4 floats (rgb, pdf)
void ref_bsdf_eval(void* h, const void* materials, void* volume_states, const float* vecs, long long n, float* out)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    const SimplifiedRendererMaterial* mats = reinterpret_cast<const SimplifiedRendererMaterial*>(materials);
    RayVolumeState* vs = reinterpret_cast<RayVolumeState*>(volume_states);
    for (long long i = 0; i < n; i++)
    {
        const float* v = vecs + i * 9;
        float pdf;
        Xorshift32Generator rng(42);
        ColorRGB32F f = bsdf_dispatcher_eval(c->rd, mats[i], vs[i], make_float3(v[0], v[1], v[2]), make_float3(v[3], v[4], v[5]), make_float3(v[3], v[4], v[5]),
                                             make_float3(v[6], v[7], v[8]), pdf, rng);
        out[i * 4 + 0] = f.r; out[i * 4 + 1] = f.g; out[i * 4 + 2] = f.b; out[i * 4 + 3] = pdf;
    }
}

// bsdf_dispatcher_sample, src/Device/includes/Dispatcher.h:46-68. vecs: n x 9 floats (view, shading normal, geometric normal).
// seeds: n (in: generator state, out: state after sampling). out: n x 7 floats (rgb, pdf, direction)
void ref_bsdf_sample(void* h, const void* materials, void* volume_states, const float* vecs, unsigned int* seeds, long long n, float* out)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    const SimplifiedRendererMaterial* mats = reinterpret_cast<const SimplifiedRendererMaterial*>(materials);
    RayVolumeState* vs = reinterpret_cast<RayVolumeState*>(volume_states);
    for (long long i = 0; i < n; i++)
    {
        const float* v = vecs + i * 9;
        float pdf;
        float3 dir = make_float3(0, 0, 0);
        Xorshift32Generator rng(seeds[i]);
        ColorRGB32F f = bsdf_dispatcher_sample(c->rd, mats[i], vs[i], make_float3(v[0], v[1], v[2]), make_float3(v[3], v[4], v[5]), make_float3(v[6], v[7], v[8]),
                                               dir, pdf, rng);
        seeds[i] = rng.m_state.seed;
        out[i * 7 + 0] = f.r; out[i * 7 + 1] = f.g; out[i * 7 + 2] = f.b; out[i * 7 + 3] = pdf;
        out[i * 7 + 4] = dir.x; out[i * 7 + 5] = dir.y; out[i * 7 + 6] = dir.z;
    }
}

// get_intersection_material, src/Device/includes/Material.h:47-96
void ref_intersection_material(void* h, int material_index, float u, float v, void* out_256)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    SimplifiedRendererMaterial m = get_intersection_material(c->rd, material_index, make_float2(u, v));
    std::memcpy(out_256, &m, sizeof(m));
}

// envmap_sample + envmap_eval, src/Device/includes/Envmap.h:77-149. out: n x 8 floats (radiance rgb, pdf, direction xyz, eval pdf of that direction)
void ref_envmap_sample(void* h, unsigned int* seeds, long long n, float* out)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    for (long long i = 0; i < n; i++)
    {
        Xorshift32Generator rng(seeds[i]);
        float3 dir; float pdf, pdf2;
        ColorRGB32F r = envmap_sample(c->rd.world_settings, dir, pdf, rng);
        envmap_eval(c->rd, dir, pdf2);
        seeds[i] = rng.m_state.seed;
        out[i * 8 + 0] = r.r; out[i * 8 + 1] = r.g; out[i * 8 + 2] = r.b; out[i * 8 + 3] = pdf;
        out[i * 8 + 4] = dir.x; out[i * 8 + 5] = dir.y; out[i * 8 + 6] = dir.z; out[i * 8 + 7] = pdf2;
    }
}

float ref_glass_correction_exponent(void* h, float roughness, float relative_eta)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    return GGX_glass_energy_conservation_get_correction_exponent(c->rd, roughness, relative_eta);
}

// thin_film_fresnel, src/Device/includes/BSDFs/ThinFilm.h:112-219
void ref_thin_film_fresnel(const void* material_256, float ambient_ior, float HoL, float* out_rgb)
{
    ColorRGB32F f = thin_film_fresnel(*reinterpret_cast<const SimplifiedRendererMaterial*>(material_256), ambient_ior, HoL);
    out_rgb[0] = f.r; out_rgb[1] = f.g; out_rgb[2] = f.b;
}

} // extern "C"
