// The three ReSTIR DI passes as per-pixel / per-sample device bodies.
// reference: src/Device/kernels/ReSTIR/DI/LightsPresampling.h:22-128, InitialCandidates.h:24-506,
//            FusedSpatiotemporalReuse.h:39-584 (default kernel options, see restir_di.cuh).
#pragma once

#include "kernels.cuh"

namespace pt
{
// ---- K3: ReSTIR_DI_LightsPresampling -----------------------------------------------------------------------------------------
template <int ESS>
PT_DEV void restir_presample_light(const LightPresamplingParameters& p, uint32_t x)
{
    if (p.emissive_triangles_count == 0 && p.world_settings.ambient_light_type != AMBIENT_ENVMAP)
        return;
    if (x >= (uint32_t)(p.subset_size * p.number_of_subsets))
        return;

    unsigned int seed = p.freeze_random ? wang_hash(x + 1) : wang_hash((x + 1) * (p.sample_number + 1) * p.random_seed);
    Rng rng(seed);

    float envmap_candidate_probability = 0.0f;
    if (p.world_settings.ambient_light_type == AMBIENT_ENVMAP)
        envmap_candidate_probability = p.emissive_triangles_count == 0 ? 1.0f : p.envmap_sampling_probability;

    ReSTIRDIPresampledLight light;
    light.emissive_triangle_index = -1;
    light.point_on_light_source.x = light.point_on_light_source.y = light.point_on_light_source.z = 0.0f;
    light.light_source_normal.x = light.light_source_normal.y = light.light_source_normal.z = 0.0f;
    light.radiance.r = light.radiance.g = light.radiance.b = 0.0f;
    light.pdf = 0.0f;
    light.flags = RESTIR_DI_FLAGS_NONE;
    light._pad0[0] = light._pad0[1] = light._pad0[2] = 0;

    if (rng() < envmap_candidate_probability)
    {
        // presample_envmap (:22-35): the direction is stored in envmap space
        light.flags |= RESTIR_DI_FLAGS_ENVMAP_SAMPLE;
        float3 direction;
        RGB radiance = envmap_sample<ESS == 0 ? 2 : ESS>(p.world_settings, direction, light.pdf, rng);
        set_f3(light.point_on_light_source, matrix_X_vec(as_mat4(p.world_settings.world_to_envmap_matrix), direction));
        set_rgb(light.radiance, radiance);
        light.pdf *= envmap_candidate_probability;
    }
    else
    {
        // presample_emissive_triangle (:37-77)
        const float light_sampling_probability = 1.0f - envmap_candidate_probability;
        int random_index = rng.random_index(p.emissive_triangles_count);
        int triangle_index = __ldg(p.emissive_triangles_indices + random_index);
        const int* idx = p.triangles_indices + (size_t)triangle_index * 3;
        float3 A = ld3(p.vertices_positions, __ldg(idx)), B = ld3(p.vertices_positions, __ldg(idx + 1)), C = ld3(p.vertices_positions, __ldg(idx + 2));
        float rand_1 = rng();
        float rand_2 = rng();
        float sqrt_r1 = sqrtf(rand_1);
        float u = 1.0f - sqrt_r1;
        float v = (1.0f - rand_2) * sqrt_r1;
        float3 AB = B - A, AC = C - A;
        float3 random_point_on_triangle = A + AB * u + AC * v;
        float3 normal = cross(AB, AC);
        float length_normal = length(normal);
        if (length_normal > 1.0e-6f)
        {
            float triangle_area = length_normal * 0.5f;
            set_f3(light.point_on_light_source, random_point_on_triangle);
            set_f3(light.light_source_normal, normal / length_normal);
            light.emissive_triangle_index = triangle_index;
            light.pdf = 1.0f / triangle_area;
            light.pdf /= p.emissive_triangles_count;
            light.pdf *= light_sampling_probability;
            const RendererMaterial& m = p.materials[__ldg(p.material_indices + triangle_index)];
            set_rgb(light.radiance, rgb(m.base.emission) * m.base.emission_strength);
        }
    }
    uint32_t* dst = reinterpret_cast<uint32_t*>(p.out_light_samples + x);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&light);
#pragma unroll
    for (int i = 0; i < 12; i++)
        dst[i] = src[i];
}

// ---- K4: ReSTIR_DI_InitialCandidates -----------------------------------------------------------------------------------------
PT_DEV int cantor_pairing_function(int x, int y) { return (x + y + 1) * (x + y) / 2 + y; }

// use_presampled_light_candidate (InitialCandidates.h:29-91)
PT_DEV ReSTIRDISample restir_use_presampled_light_candidate(const HIPRTRenderData& rd, int2 pixel_coords, float3 evaluated_point, float3 shading_normal, RGB& out_sample_radiance,
                                                           float& out_sample_cosine_term, float& out_sample_pdf, float& out_distance_to_light, float3& out_to_light_direction, Rng& rng)
{
    const LightPresamplingSettings& lp = rd.render_settings.restir_di_settings.light_presampling;
    int tile_index_seed = cantor_pairing_function(pixel_coords.x / lp.tile_size, pixel_coords.y / lp.tile_size);
    Rng subset_rng(rd.random_seed * (tile_index_seed + 1));
    int random_subset_index = subset_rng.random_index(lp.number_of_subsets);
    int random_light_index_in_subset = rng.random_index(lp.subset_size);
    int light_sample_index = random_subset_index * lp.subset_size + random_light_index_in_subset;

    ReSTIRDIPresampledLight presampled;
    {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(lp.light_samples + light_sample_index);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&presampled);
#pragma unroll
        for (int i = 0; i < 12; i++)
            dst[i] = __ldg(src + i);
    }

    ReSTIRDISample light_sample;
    light_sample.emissive_triangle_index = presampled.emissive_triangle_index;
    light_sample.point_on_light_source = presampled.point_on_light_source;
    light_sample.target_function = 0.0f;
    light_sample.flags = presampled.flags;
    light_sample._pad0[0] = light_sample._pad0[1] = light_sample._pad0[2] = 0;
    out_sample_radiance = rgb(presampled.radiance);
    out_sample_pdf = presampled.pdf;

    if (light_sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE)
    {
        out_to_light_direction = matrix_X_vec(as_mat4(rd.world_settings.envmap_to_world_matrix), to_f3(light_sample.point_on_light_source));
        out_distance_to_light = 1.0e35f;
    }
    else
    {
        out_to_light_direction = to_f3(light_sample.point_on_light_source) - evaluated_point;
        out_to_light_direction = out_to_light_direction / (out_distance_to_light = length(out_to_light_direction));
    }
    out_sample_cosine_term = dot(shading_normal, out_to_light_direction);

    if (!(light_sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE))
    {
        float cosine_at_light_source = fabsf(dot(to_f3(presampled.light_source_normal), -out_to_light_direction));
        out_sample_pdf *= out_distance_to_light * out_distance_to_light;
        out_sample_pdf /= cosine_at_light_source;
        if (!check_minimum_light_contribution(rd.render_settings.minimum_light_contribution, out_sample_radiance * out_sample_cosine_term / out_sample_pdf))
            light_sample.target_function = -42.0f;
    }
    return light_sample;
}

// The state of a pixel between the phases of the phased pass (RestirScratch::ic_state, 4 x float4): the reservoir after the light
// candidates, the generator, the BSDF candidate's value and pdf and whether it refracts (its direction is in the queued ray)
PT_DEV void restir_ic_state_store(const RestirSites& sites, const ReSTIRDIReservoir& r, const Rng& rng, RGB bsdf_color, float bsdf_sample_pdf, bool refraction_sampled)
{
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&r);
    float4* st = sites.scratch->ic_state + sites.pixel;
    const uint32_t n = sites.scratch->capacity;
    st[0] = make_float4(__uint_as_float(w[0]), __uint_as_float(w[1]), __uint_as_float(w[2]), __uint_as_float(w[3]));
    st[n] = make_float4(__uint_as_float(w[4]), __uint_as_float(w[5]), __uint_as_float(w[6]), __uint_as_float(w[7]));
    st[2 * n] = make_float4(__uint_as_float(w[8]), __uint_as_float(rng.state), bsdf_sample_pdf, __uint_as_float(refraction_sampled ? 1u : 0u));
    st[3 * n] = make_float4(bsdf_color.r, bsdf_color.g, bsdf_color.b, 0.0f);
}
PT_DEV void restir_ic_state_load(const RestirSites& sites, ReSTIRDIReservoir& r, Rng& rng, RGB& bsdf_color, float& bsdf_sample_pdf, bool& refraction_sampled)
{
    uint32_t* w = reinterpret_cast<uint32_t*>(&r);
    const float4* st = sites.scratch->ic_state + sites.pixel;
    const uint32_t n = sites.scratch->capacity;
    const float4 a = st[0], b = st[n], c = st[2 * n], d = st[3 * n];
    w[0] = __float_as_uint(a.x); w[1] = __float_as_uint(a.y); w[2] = __float_as_uint(a.z); w[3] = __float_as_uint(a.w);
    w[4] = __float_as_uint(b.x); w[5] = __float_as_uint(b.y); w[6] = __float_as_uint(b.z); w[7] = __float_as_uint(b.w);
    w[8] = __float_as_uint(c.x);
    rng.state = __float_as_uint(c.y);
    bsdf_sample_pdf = c.z;
    refraction_sampled = __float_as_uint(c.w) != 0u;
    bsdf_color = RGB(d.x, d.y, d.z);
}

// MODE (restir_di.cuh): INLINE = the whole pass, rays traced per thread. Phased, with one BSDF candidate per pixel: COLLECT = light
// candidates + BSDF sample, the BSDF candidate's ray queued (site 0, closest hit), state saved; RESOLVE = the BSDF candidate finished
// from the traced hit, reservoir closed and stored, its visibility ray queued (site 1; status 3) -- restir_initial_candidates_visibility
// applies the answer to the stored reservoir.
template <int GGX_SAMPLER, int MODE = RESTIR_INLINE>
PT_DEV void restir_initial_candidates_pixel(const HIPRTRenderData& rd, int2 res, uint32_t x, uint32_t y, const RestirSites sites = RestirSites())
{
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 0u;
    if (rd.buffers.emissive_triangles_count == 0 && rd.world_settings.ambient_light_type != AMBIENT_ENVMAP)
        return;
    const uint32_t pixel_index = x + y * res.x;
    const uint32_t status = MODE == RESTIR_RESOLVE ? sites.scratch->status[sites.pixel] : 0u;
    if (MODE == RESTIR_RESOLVE)
    {
        if (status == 0u)
            return;
        sites.scratch->status[sites.pixel] = 0u;
    }

    RayPayload payload;
    if (MODE != RESTIR_RESOLVE)
    {
        load_material(payload.material, rd.g_buffer.materials + pixel_index);
        if (material_is_emissive(payload.material))
            return;
    }
    Rng rng(pixel_seed(rd, pixel_index));
    if (!rd.aux_buffers.pixel_active[pixel_index] || !rd.g_buffer.camera_ray_hit[pixel_index])
        return;

    HitInfo hit;
    hit.geometric_normal = ld3_rw(rd.g_buffer.geometric_normals, pixel_index);
    hit.shading_normal = ld3_rw(rd.g_buffer.shading_normals, pixel_index);
    hit.inter_point = ld3_rw(rd.g_buffer.first_hits, pixel_index);
    hit.primitive_index = rd.g_buffer.first_hit_prim_index[pixel_index];
    float3 view_direction = f3(0.0f);
    if (MODE != RESTIR_RESOLVE)
    {
        copy_volume_state(&payload.volume_state, rd.g_buffer.ray_volume_states + pixel_index);
        view_direction = ld3_rw(rd.g_buffer.view_directions, pixel_index);
    }

    // sample_initial_candidates (:413-446)
    const InitialCandidatesSettings& ic = rd.render_settings.restir_di_settings.initial_candidates;
    const bool low_res = do_render_low_resolution(rd.render_settings);
    const int nb_light_candidates = low_res ? mini(1, ic.number_of_initial_light_candidates) : ic.number_of_initial_light_candidates;
    const int nb_bsdf_candidates = low_res ? mini(1, ic.number_of_initial_bsdf_candidates) : ic.number_of_initial_bsdf_candidates;
    float envmap_candidate_probability = 0.0f;
    if (rd.world_settings.ambient_light_type == AMBIENT_ENVMAP)
        envmap_candidate_probability = rd.buffers.emissive_triangles_count == 0 ? 1.0f : ic.envmap_candidate_probability;

    ReSTIRDIReservoir reservoir = empty_reservoir();
    const float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f * 1.0f;

    // sample_light_candidates (:185-279), presampled lights
    for (int i = 0; i < nb_light_candidates && MODE != RESTIR_RESOLVE; i++)
    {
        RGB sample_radiance;
        float sample_cosine_term = 0.0f, sample_pdf = 0.0f, distance_to_light = 0.0f;
        float3 to_light_direction = f3(0.0f);
        ReSTIRDISample light_sample = restir_use_presampled_light_candidate(rd, make_int2(x, y), evaluated_point, hit.shading_normal * 1.0f, sample_radiance, sample_cosine_term, sample_pdf,
                                                                            distance_to_light, to_light_direction, rng);
        if (light_sample.target_function == -42.0f)
        {
            reservoir.M++;
            continue;
        }
        float candidate_weight = 0.0f;
        if (sample_cosine_term > 0.0f && sample_pdf > 0.0f)
        {
            float bsdf_pdf;
            RayVolumeState volume_state = payload.volume_state;
            RGB bsdf_contribution = bsdf_eval(rd, payload.material, volume_state, view_direction, hit.shading_normal, to_light_direction, bsdf_pdf);
            RGB light_contribution = bsdf_contribution * sample_radiance * sample_cosine_term;
            float target_function = light_contribution.luminance();
            if (check_minimum_light_contribution(rd.render_settings.minimum_light_contribution, light_contribution / sample_pdf / bsdf_pdf))
            {
                float mis_weight = power_heuristic(sample_pdf, nb_light_candidates, bsdf_pdf, nb_bsdf_candidates);
                candidate_weight = mis_weight * target_function / sample_pdf;
                light_sample.target_function = target_function;
            }
        }
        reservoir_add_one_candidate(reservoir, light_sample, candidate_weight, rng);
    }

    // sample_bsdf_candidates (:281-411)
    for (int i = 0; i < nb_bsdf_candidates; i++)
    {
        float bsdf_sample_pdf = 0.0f;
        float3 sampled_direction;
        RGB bsdf_color;
        bool refraction_sampled;
        if (MODE != RESTIR_RESOLVE)
        {
            RayVolumeState trash = payload.volume_state;
            bsdf_color = bsdf_sample<GGX_SAMPLER>(rd, payload.material, trash, view_direction, hit.shading_normal, hit.geometric_normal, sampled_direction, bsdf_sample_pdf, rng);
            refraction_sampled = dot(sampled_direction, view_direction) < 0.0f;
            if (!(bsdf_sample_pdf > 0.0f))
                continue;
        }
        if (MODE == RESTIR_COLLECT)
        {
            sites.emit(0, hit.inter_point, sampled_direction, 1.0e35f, hit.primitive_index, 0);
            restir_ic_state_store(sites, reservoir, rng, bsdf_color, bsdf_sample_pdf, refraction_sampled);
            sites.scratch->status[sites.pixel] = 2u;
            return;
        }

        ShadowLightRayHitInfo light_hit;
        bool hit_found;
        if (MODE == RESTIR_RESOLVE)
        {
            if (status != 2u)
                break;       // the BSDF sample had no pdf: no candidate
            restir_ic_state_load(sites, reservoir, rng, bsdf_color, bsdf_sample_pdf, refraction_sampled);
            const float4 d = sites.scratch->ray_d[sites.ray_id(0)];
            sampled_direction = f3(d.x, d.y, d.z);
            const float4 h = sites.scratch->hit_tuvp[sites.pixel];
            Hit closest;
            closest.t = h.x; closest.u = h.y; closest.v = h.z; closest.prim = __float_as_int(h.w);
            closest.tri_slot = (int)sites.scratch->hit_tri_slot[sites.pixel];
            hit_found = shadow_light_ray_resolve(rd, 1.0e35f, light_hit, closest);
        }
        else
            hit_found = evaluate_shadow_light_ray(rd, hit.inter_point, sampled_direction, 1.0e35f, light_hit, hit.primitive_index, rng);
        if (hit_found && !light_hit.hit_emission.is_black())
        {
            float cosine_at_evaluated_point = fabsf(dot(hit.shading_normal, sampled_direction));
            RGB light_contribution = bsdf_color * light_hit.hit_emission * cosine_at_evaluated_point;
            float target_function = light_contribution.luminance();
            float light_pdf = 0.0f;
            if (!refraction_sampled)
                light_pdf = pdf_of_emissive_triangle_hit(rd, light_hit, sampled_direction);
            if (!check_minimum_light_contribution(rd.render_settings.minimum_light_contribution, light_contribution / light_pdf / bsdf_sample_pdf))
            {
                reservoir.M++;
                continue;
            }
            light_pdf *= (1.0f - envmap_candidate_probability);
            float mis_weight = power_heuristic(bsdf_sample_pdf, nb_bsdf_candidates, light_pdf, nb_light_candidates);
            float candidate_weight = mis_weight * target_function / bsdf_sample_pdf;

            ReSTIRDISample s;
            s.emissive_triangle_index = light_hit.hit_prim_index;
            set_f3(s.point_on_light_source, hit.inter_point + sampled_direction * light_hit.hit_distance);
            s.target_function = target_function;
            s.flags = RESTIR_DI_FLAGS_UNOCCLUDED | (refraction_sampled ? RESTIR_DI_FLAGS_BSDF_REFRACTION : 0);
            s._pad0[0] = s._pad0[1] = s._pad0[2] = 0;
            reservoir_add_one_candidate(reservoir, s, candidate_weight, rng);
        }
        else if (!hit_found && rd.world_settings.ambient_light_type == AMBIENT_ENVMAP)
        {
            float cosine_at_evaluated_point = maxf(0.0f, dot(hit.shading_normal, sampled_direction));
            if (cosine_at_evaluated_point > 0.0f)
            {
                float envmap_pdf;
                RGB envmap_radiance = envmap_eval(rd, sampled_direction, envmap_pdf);
                RGB envmap_contribution = bsdf_color * envmap_radiance * cosine_at_evaluated_point;
                if (!check_minimum_light_contribution(rd.render_settings.minimum_light_contribution, envmap_contribution / envmap_pdf / bsdf_sample_pdf))
                {
                    reservoir.M++;
                    continue;
                }
                float target_function = envmap_contribution.luminance();
                envmap_pdf *= envmap_candidate_probability;
                float mis_weight = power_heuristic(bsdf_sample_pdf, nb_bsdf_candidates, envmap_pdf, nb_light_candidates);
                float candidate_weight = mis_weight * target_function / bsdf_sample_pdf;

                ReSTIRDISample s;
                s.emissive_triangle_index = -1;
                set_f3(s.point_on_light_source, matrix_X_vec(as_mat4(rd.world_settings.world_to_envmap_matrix), sampled_direction));
                s.target_function = target_function;
                s.flags = RESTIR_DI_FLAGS_ENVMAP_SAMPLE | RESTIR_DI_FLAGS_UNOCCLUDED;
                s._pad0[0] = s._pad0[1] = s._pad0[2] = 0;
                reservoir_add_one_candidate(reservoir, s, candidate_weight, rng);
            }
        }
    }

    if (MODE == RESTIR_COLLECT)
    {
        // no BSDF candidate to trace: the reservoir of the light candidates goes to the next phase as it is
        restir_ic_state_store(sites, reservoir, rng, RGB(0.0f), 0.0f, false);
        sites.scratch->status[sites.pixel] = 1u;
        return;
    }
    if (MODE == RESTIR_RESOLVE && status == 1u)
    {
        RGB unused_color;
        float unused_pdf;
        bool unused_flag;
        restir_ic_state_load(sites, reservoir, rng, unused_color, unused_pdf, unused_flag);
    }

    reservoir_end(reservoir);
    reservoir.M = 1;
    // ReSTIR_DI_DoVisibilityReuse
    if (MODE == RESTIR_RESOLVE)
    {
        if (reservoir.UCW > 0.0f && !(reservoir.sample.flags & RESTIR_DI_FLAGS_UNOCCLUDED))
            sites.scratch->status[sites.pixel] = 3u;
        restir_visibility_reuse<RESTIR_COLLECT>(rd, reservoir, hit.inter_point + hit.shading_normal * 1.0e-4f, hit.primitive_index, rng, sites, 1);
    }
    else
        restir_visibility_reuse(rd, reservoir, hit.inter_point + hit.shading_normal * 1.0e-4f, hit.primitive_index, rng);
    store_reservoir(rd.render_settings.restir_di_settings.initial_candidates.output_reservoirs + pixel_index, reservoir);
}

// Last phase of the phased pass: the visibility of the candidate a pixel kept (ReSTIR_DI_visibility_reuse's two outcomes)
PT_DEV void restir_initial_candidates_visibility(const HIPRTRenderData& rd, int2 res, uint32_t x, uint32_t y, const RestirSites sites)
{
    if (sites.scratch->status[sites.pixel] != 3u)
        return;
    ReSTIRDIReservoir* out = rd.render_settings.restir_di_settings.initial_candidates.output_reservoirs + (x + y * res.x);
    if (sites.occluded(1))
        out->UCW = -1.0f;
    else
        out->sample.flags |= RESTIR_DI_FLAGS_UNOCCLUDED;
}

// ---- K5: ReSTIR_DI_SpatiotemporalReuse (fused temporal + spatial pass) ------------------------------------------------------------
// do_include_spatial_visibility_term_or_not (FusedSpatiotemporalReuse.h:39-63); reuse_neighbor_count is the kernel's local
// value (the reference overwrites it in its by-value copy of the render data on disocclusion, :186-187)
PT_DEV bool restir_include_spatial_visibility(const HIPRTRenderData& rd, int current_neighbor_index, int reuse_neighbor_count)
{
    const SpatialPassSettings& sp = rd.render_settings.restir_di_settings.spatial_pass;
    bool is_last_pass = sp.spatial_pass_index == sp.number_of_passes - 1;
    bool include = sp.do_visibility_only_last_pass && is_last_pass;
    include |= !sp.do_visibility_only_last_pass;
    include &= current_neighbor_index < sp.neighbor_visibility_count;
    include &= current_neighbor_index != reuse_neighbor_count;
    return include;     // ReSTIR_DI_SpatialTargetFunctionVisibility == TRUE
}

// MODE (restir_di.cuh): INLINE = the whole pass with per-thread shadow rays; COLLECT = neighbour search + every target function before
// visibility (stored per site) + the shadow rays queued; RESOLVE = neighbour search again (same generator, same draws), then the
// resampling with the stored target functions and the traced visibilities. Sites: launch.h (RestirScratch).
template <int MODE = RESTIR_INLINE>
PT_DEV void restir_spatiotemporal_pixel(const HIPRTRenderData& rd, int2 res, uint32_t x, uint32_t y, const RestirSites sites = RestirSites())
{
    const uint32_t center_pixel_index = x + y * res.x;
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 0u;
    if (!rd.aux_buffers.pixel_active[center_pixel_index] || !rd.g_buffer.camera_ray_hit[center_pixel_index])
        return;
    if (MODE == RESTIR_RESOLVE && sites.scratch->status[sites.pixel] == 0u)
        return;

    const ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;
    Rng rng(pixel_seed(rd, center_pixel_index));
    const bool prev_g = use_prev_frame_g_buffer(rd);

    ReSTIRSurface center;
    get_pixel_surface<MODE != RESTIR_RESOLVE>(rd, center_pixel_index, false, center);
    if (MODE != RESTIR_RESOLVE && material_is_emissive(center.material))
        return;
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 1u;

    if (rs.temporal_pass.temporal_buffer_clear_requested)
        store_reservoir(rs.temporal_pass.input_reservoirs + center_pixel_index, empty_reservoir());

    // load_temporal_neighbor_data (:69-97)
    ReSTIRDIReservoir temporal_neighbor_reservoir = empty_reservoir();
    ReSTIRSurface temporal_neighbor_surface;
    temporal_neighbor_surface.shading_point = f3(0.0f);
    temporal_neighbor_surface.shading_normal = f3(0.0f);
    int3 temporal = restir_find_temporal_neighbor_index(rd, ld3_rw(rd.g_buffer.first_hits, center_pixel_index), center.shading_normal, res, center_pixel_index, rng);
    if (temporal.x != -1 && !rd.render_settings.freeze_random)
    {
        temporal_neighbor_reservoir = load_reservoir(rs.temporal_pass.input_reservoirs + temporal.x);
        if (temporal_neighbor_reservoir.M != 0)
            get_pixel_surface<MODE != RESTIR_RESOLVE>(rd, temporal.x, prev_g, temporal_neighbor_surface);
    }
    const int2 temporal_position = make_int2(temporal.y, temporal.z);

    int reused_neighbors_count = rs.spatial_pass.reuse_neighbor_count;
    if ((temporal.x == -1 || temporal_neighbor_reservoir.M <= 1) && rs.spatial_pass.do_disocclusion_reuse_boost)
        reused_neighbors_count = rs.spatial_pass.disocclusion_reuse_count;

    float rotation_theta = rs.spatial_pass.do_neighbor_rotation ? kTwoPi * rng() : 0.0f;
    float2 cos_sin_theta_rotation = make_float2(cosf(rotation_theta), sinf(rotation_theta));

    // count_valid_spatiotemporal_neighbors (:113-132)
    int neighbor_heuristics_cache = 0, valid_neighbors_count = 0, valid_neighbors_M_sum = 0;
    for (int neighbor_index = 0; neighbor_index < reused_neighbors_count; neighbor_index++)
    {
        int neighbor_pixel_index = restir_spatial_neighbor_pixel_index(rd, neighbor_index, reused_neighbors_count, rs.spatial_pass.reuse_radius, temporal_position, res, cos_sin_theta_rotation);
        if (neighbor_pixel_index == -1)
            continue;
        if (!restir_check_neighbor_similarity_heuristics(rd, neighbor_pixel_index, center_pixel_index, center.shading_point, center.shading_normal, prev_g))
            continue;
        valid_neighbors_M_sum += rs.spatial_pass.input_reservoirs[neighbor_pixel_index].M;
        valid_neighbors_count++;
        neighbor_heuristics_cache |= (1 << neighbor_index);
    }
    if (temporal.x != -1 && temporal_neighbor_reservoir.M > 0)
    {
        valid_neighbors_count++;
        valid_neighbors_M_sum += temporal_neighbor_reservoir.M;
    }

    ReSTIRDIReservoir out = empty_reservoir();
    const ReSTIRDIReservoir initial_candidates_reservoir = load_reservoir(rs.initial_candidates.output_reservoirs + center_pixel_index);
    PairwiseDefensiveMISWeight mis_weight_function;
    const bool update_mc = initial_candidates_reservoir.M > 0 && initial_candidates_reservoir.UCW > 0.0f;

    // temporal neighbour (:206-290)
    if (temporal.x != -1 && temporal_neighbor_reservoir.M > 0)
    {
        float target_function_at_center = 0.0f;
        if (temporal_neighbor_reservoir.UCW > 0.0f)
            target_function_at_center = restir_evaluate_target_function<true, MODE>(rd, temporal_neighbor_reservoir.sample, center, rng, sites, 0);

        float jacobian_determinant = 1.0f;
        if (target_function_at_center > 0.0f && temporal_neighbor_reservoir.UCW > 0.0f && !(temporal_neighbor_reservoir.sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE))
        {
            jacobian_determinant = restir_jacobian_reconnection_shift(rd, temporal_neighbor_reservoir, center.shading_point, temporal_neighbor_surface.shading_point - temporal_neighbor_surface.shading_normal * 1.0e-4f);
            if (jacobian_determinant == -1.0f)
                jacobian_determinant = 0.0f;
        }
        float w = mis_weight_function.template get<MODE>(rd, temporal_neighbor_reservoir, initial_candidates_reservoir, target_function_at_center, temporal.x, valid_neighbors_count, valid_neighbors_M_sum,
                                                         update_mc, false, rng, sites, 1);
        if (MODE != RESTIR_COLLECT && reservoir_combine_with(out, temporal_neighbor_reservoir, w, target_function_at_center, jacobian_determinant, rng))
            out.sample.flags |= RESTIR_DI_FLAGS_UNOCCLUDED;       // ReSTIR_DI_BiasCorrectionUseVisibility == TRUE
    }

    // spatial neighbours around the temporal position, the canonical (initial candidates) sample last (:292-470)
    const ReSTIRDIReservoir* spatial_input = rs.spatial_pass.input_reservoirs;
    int final_visibility_site = -1;       // phased: the site whose ray is the end-of-pass visibility ray of the sample `out` holds
    int start_index = valid_neighbors_M_sum == 0 ? reused_neighbors_count : 0;
    for (int spatial_neighbor_index = start_index; spatial_neighbor_index < reused_neighbors_count + 1; spatial_neighbor_index++)
    {
        const bool canonical = spatial_neighbor_index == reused_neighbors_count;
        if (!canonical && reused_neighbors_count <= 32)
            if ((neighbor_heuristics_cache & (1 << spatial_neighbor_index)) == 0)
                continue;

        int neighbor_pixel_index = canonical ? (int)center_pixel_index
                                             : restir_spatial_neighbor_pixel_index(rd, spatial_neighbor_index, reused_neighbors_count, rs.spatial_pass.reuse_radius, temporal_position, res, cos_sin_theta_rotation);
        if (neighbor_pixel_index == -1)
            continue;
        if (!canonical && reused_neighbors_count > 32)
            if (!restir_check_neighbor_similarity_heuristics(rd, neighbor_pixel_index, center_pixel_index, center.shading_point, center.shading_normal, prev_g))
                continue;

        ReSTIRDIReservoir neighbor_reservoir = canonical ? initial_candidates_reservoir : load_reservoir(spatial_input + neighbor_pixel_index);

        float target_function_at_center = 0.0f;
        const bool do_neighbor_target_function_visibility = restir_include_spatial_visibility(rd, spatial_neighbor_index, reused_neighbors_count);
        if (neighbor_reservoir.UCW > 0.0f)
        {
            if (canonical)
                target_function_at_center = neighbor_reservoir.sample.target_function;
            else if (do_neighbor_target_function_visibility)
                target_function_at_center = restir_evaluate_target_function<true, MODE>(rd, neighbor_reservoir.sample, center, rng, sites, 2 + 2 * spatial_neighbor_index);
            else
            {
                target_function_at_center = restir_evaluate_target_function<false, MODE>(rd, neighbor_reservoir.sample, center, rng, sites, 2 + 2 * spatial_neighbor_index);
                if (MODE == RESTIR_COLLECT)
                {
                    // if this sample is the one the pass keeps, the end-of-pass visibility reuse traces it from the pixel: that ray takes this site
                    float distance_to_light;
                    float3 sample_direction;
                    restir_visibility_reuse_ray(rd, neighbor_reservoir.sample, center.shading_point, sample_direction, distance_to_light);
                    sites.emit_any_hit(2 + 2 * spatial_neighbor_index, center.shading_point, sample_direction, distance_to_light, center.last_hit_primitive_index);
                }
            }
        }

        float jacobian_determinant = 1.0f;
        if (target_function_at_center > 0.0f && neighbor_reservoir.UCW > 0.0f && !canonical && !(neighbor_reservoir.sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE))
        {
            const GBuffer& g = prev_g ? rd.g_buffer_prev_frame : rd.g_buffer;
            float3 neighbor_shading_point = ld3_rw(g.first_hits, neighbor_pixel_index) + ld3_rw(g.shading_normals, neighbor_pixel_index) * 1.0e-4f;
            jacobian_determinant = restir_jacobian_reconnection_shift(rd, neighbor_reservoir, center.shading_point, neighbor_shading_point);
            if (jacobian_determinant == -1.0f && MODE != RESTIR_COLLECT)     // COLLECT saw the target function before visibility: the neighbour may still be resampled
            {
                out.M += neighbor_reservoir.M;
                continue;
            }
        }

        float mis_weight;
        if (neighbor_reservoir.UCW == 0.0f && !update_mc)
            mis_weight = 1.0f;
        else
            mis_weight = mis_weight_function.template get<MODE>(rd, neighbor_reservoir, initial_candidates_reservoir, target_function_at_center, neighbor_pixel_index, valid_neighbors_count,
                                                                valid_neighbors_M_sum, update_mc, canonical, rng, sites, 3 + 2 * spatial_neighbor_index);

        if (MODE != RESTIR_COLLECT && reservoir_combine_with(out, neighbor_reservoir, mis_weight, target_function_at_center, jacobian_determinant, rng))
        {
            final_visibility_site = -1;
            if (do_neighbor_target_function_visibility)
                out.sample.flags |= RESTIR_DI_FLAGS_UNOCCLUDED;
            else if (canonical)
                out.sample.flags |= neighbor_reservoir.sample.flags & RESTIR_DI_FLAGS_UNOCCLUDED;
            else
            {
                out.sample.flags &= ~RESTIR_DI_FLAGS_UNOCCLUDED;
                final_visibility_site = 2 + 2 * spatial_neighbor_index;
            }
        }
    }
    if (MODE == RESTIR_COLLECT)
        return;

    // pairwise MIS: nothing left to normalise (SpatiotemporalNormalizationWeight.h:270-278)
    reservoir_end_with_normalization(out, 1.0f, 1.0f);

    // end-of-pass visibility reuse (:560-575)
    if (rs.temporal_pass.do_temporal_reuse_pass || rs.spatial_pass.number_of_passes - 1 != rs.spatial_pass.spatial_pass_index)
    {
        if (MODE == RESTIR_RESOLVE && final_visibility_site >= 0)
            restir_visibility_reuse<RESTIR_RESOLVE>(rd, out, center.shading_point, center.last_hit_primitive_index, rng, sites, final_visibility_site);
        else
            restir_visibility_reuse(rd, out, center.shading_point, center.last_hit_primitive_index, rng);     // RESOLVE: a sample whose ray was not queued (none by construction)
    }

    if (rs.m_cap > 0)
        out.M = mini(out.M, rs.m_cap);
    store_reservoir(rs.spatial_pass.output_reservoirs + center_pixel_index, out);
}

// ---- ReSTIR_DI_TemporalReuse (TemporalReuse.h:48-306): the non-fused configuration's temporal pass ---------------------------------
// Combines last frame's reservoir at the back-projected pixel with this frame's initial candidates (pairwise-defensive MIS,
// bias correction with visibility). Pixels without a usable temporal neighbour copy their initial candidates through.
// Phased (MODE, restir_di.cuh): COLLECT stores nothing but the site values, the queued rays and the pixel's status (0 nothing to do,
// 1 resample, 2 copy the initial candidates through); RESOLVE repeats the neighbour search (same draws) and writes the output.
// Sites: 0 = the temporal neighbour's sample seen from the pixel, 1 = the pixel's sample seen from the temporal neighbour's surface.
template <int MODE = RESTIR_INLINE>
PT_DEV void restir_temporal_pixel(const HIPRTRenderData& rd, int2 res, uint32_t x, uint32_t y, const RestirSites sites = RestirSites())
{
    const uint32_t center_pixel_index = x + y * res.x;
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 0u;
    if (!rd.aux_buffers.pixel_active[center_pixel_index] || !rd.g_buffer.camera_ray_hit[center_pixel_index])
        return;
    const uint32_t status = MODE == RESTIR_RESOLVE ? sites.scratch->status[sites.pixel] : 0u;
    if (MODE == RESTIR_RESOLVE && status == 0u)
        return;

    const ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;
    Rng rng(pixel_seed(rd, center_pixel_index));

    if (rs.temporal_pass.temporal_buffer_clear_requested)
        store_reservoir(rs.temporal_pass.input_reservoirs + center_pixel_index, empty_reservoir());

    ReSTIRSurface center;
    get_pixel_surface<MODE != RESTIR_RESOLVE>(rd, center_pixel_index, false, center);
    if (MODE != RESTIR_RESOLVE && material_is_emissive(center.material))
        return;
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 2u;

    const ReSTIRDIReservoir initial_candidates_reservoir = load_reservoir(rs.initial_candidates.output_reservoirs + center_pixel_index);
    ReSTIRDIReservoir* output = rs.temporal_pass.output_reservoirs + center_pixel_index;

    int temporal_neighbor_pixel_index = restir_find_temporal_neighbor_index(rd, ld3_rw(rd.g_buffer.first_hits, center_pixel_index), center.shading_normal, res, center_pixel_index, rng).x;
    if (temporal_neighbor_pixel_index == -1 || rd.render_settings.freeze_random || (MODE == RESTIR_RESOLVE && status == 2u))
    {
        if (MODE != RESTIR_COLLECT)
            store_reservoir(output, initial_candidates_reservoir);       // :88-97 (disocclusion, or frozen random: same samples every frame); RESOLVE: every copy-through case
        return;
    }
    const ReSTIRDIReservoir temporal_neighbor_reservoir = load_reservoir(rs.temporal_pass.input_reservoirs + temporal_neighbor_pixel_index);
    if (temporal_neighbor_reservoir.M == 0)
    {
        if (MODE != RESTIR_COLLECT)
            store_reservoir(output, initial_candidates_reservoir);       // :100-108
        return;
    }
    ReSTIRSurface temporal_neighbor_surface;
    get_pixel_surface<MODE != RESTIR_RESOLVE>(rd, temporal_neighbor_pixel_index, use_prev_frame_g_buffer(rd), temporal_neighbor_surface);
    if (MODE != RESTIR_RESOLVE && material_is_emissive(temporal_neighbor_surface.material))
    {
        if (MODE != RESTIR_COLLECT)
            store_reservoir(output, initial_candidates_reservoir);       // :113-120
        return;
    }
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 1u;

    ReSTIRDIReservoir out = empty_reservoir();
    TemporalPairwiseDefensiveMISWeight mis_weight_function;

    // temporal neighbour (:136-207)
    {
        float target_function_at_center = 0.0f;
        if (temporal_neighbor_reservoir.UCW > 0.0f)
            target_function_at_center = restir_evaluate_target_function<true, MODE>(rd, temporal_neighbor_reservoir.sample, center, rng, sites, 0);

        float jacobian_determinant = 1.0f;
        if (target_function_at_center > 0.0f && temporal_neighbor_reservoir.UCW > 0.0f && !(temporal_neighbor_reservoir.sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE))
        {
            jacobian_determinant = restir_jacobian_reconnection_shift(rd, temporal_neighbor_reservoir, center.shading_point, temporal_neighbor_surface.shading_point - temporal_neighbor_surface.shading_normal * 1.0e-4f);
            if (jacobian_determinant == -1.0f)
                jacobian_determinant = 0.0f;
        }
        float w = mis_weight_function.template get<MODE>(rd, temporal_neighbor_reservoir, initial_candidates_reservoir, temporal_neighbor_surface, target_function_at_center, false, rng, sites, 1);
        if (MODE == RESTIR_COLLECT)
            return;
        if (reservoir_combine_with(out, temporal_neighbor_reservoir, w, target_function_at_center, jacobian_determinant, rng))
            out.sample.flags |= RESTIR_DI_FLAGS_UNOCCLUDED;           // ReSTIR_DI_BiasCorrectionUseVisibility == TRUE
    }

    // initial candidates (:210-247)
    {
        float w = mis_weight_function.get(rd, temporal_neighbor_reservoir, initial_candidates_reservoir, temporal_neighbor_surface, 0.0f, true, rng);
        if (reservoir_combine_with(out, initial_candidates_reservoir, w, initial_candidates_reservoir.sample.target_function, 1.0f, rng))
            out.sample.flags |= RESTIR_DI_FLAGS_UNOCCLUDED;
    }

    reservoir_end_with_normalization(out, 1.0f, 1.0f);
    if (rs.m_cap > 0)
        out.M = mini(out.M, rs.m_cap);
    store_reservoir(output, out);
}

// ---- ReSTIR_DI_SpatialReuse (SpatialReuse.h:64-348): a spatial pass on its own -- the non-fused configuration, and the passes
// 1..number_of_passes-1 that follow the fused pass on the GPU (ReSTIRDIRenderPass.cpp:492-501). Neighbours on the rotated
// Hammersley disk around the pixel ITSELF, read from this frame's G-buffer; the centre reservoir is the pass input at the pixel.
// Phased (MODE): as restir_spatiotemporal_pixel; sites 2 + 2 i / 3 + 2 i for spatial neighbour i, site 0 = the end-of-pass visibility ray
// of the pixel's own input sample when it arrives without the UNOCCLUDED flag.
template <int MODE = RESTIR_INLINE>
PT_DEV void restir_spatial_pixel(const HIPRTRenderData& rd, int2 res, uint32_t x, uint32_t y, const RestirSites sites = RestirSites())
{
    const uint32_t center_pixel_index = x + y * res.x;
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 0u;
    if (!rd.aux_buffers.pixel_active[center_pixel_index] || !rd.g_buffer.camera_ray_hit[center_pixel_index])
        return;
    if (MODE == RESTIR_RESOLVE && sites.scratch->status[sites.pixel] == 0u)
        return;

    const ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;
    Rng rng(pixel_seed(rd, center_pixel_index));
    const ReSTIRDIReservoir* spatial_input = rs.spatial_pass.input_reservoirs;
    const int2 center_pixel_coords = make_int2(x, y);

    ReSTIRSurface center;
    get_pixel_surface<MODE != RESTIR_RESOLVE>(rd, center_pixel_index, false, center);
    if (MODE != RESTIR_RESOLVE && material_is_emissive(center.material))
        return;
    if (MODE == RESTIR_COLLECT)
        sites.scratch->status[sites.pixel] = 1u;

    float rotation_theta = rs.spatial_pass.do_neighbor_rotation ? kTwoPi * rng() : 0.0f;
    float2 cos_sin_theta_rotation = make_float2(cosf(rotation_theta), sinf(rotation_theta));

    const ReSTIRDIReservoir center_pixel_reservoir = load_reservoir(spatial_input + center_pixel_index);
    int reused_neighbors_count = rs.spatial_pass.reuse_neighbor_count;
    if (center_pixel_reservoir.M <= 1 && rs.spatial_pass.do_disocclusion_reuse_boost)
        reused_neighbors_count = rs.spatial_pass.disocclusion_reuse_count;

    // count_valid_spatial_neighbors (Utils.h:382-402)
    int neighbor_heuristics_cache = 0, valid_neighbors_count = 0, valid_neighbors_M_sum = 0;
    for (int neighbor_index = 0; neighbor_index < reused_neighbors_count; neighbor_index++)
    {
        int neighbor_pixel_index = restir_spatial_neighbor_pixel_index(rd, neighbor_index, reused_neighbors_count, rs.spatial_pass.reuse_radius, center_pixel_coords, res, cos_sin_theta_rotation);
        if (neighbor_pixel_index == -1)
            continue;
        if (!restir_check_neighbor_similarity_heuristics(rd, neighbor_pixel_index, center_pixel_index, center.shading_point, center.shading_normal, false))
            continue;
        valid_neighbors_M_sum += spatial_input[neighbor_pixel_index].M;
        valid_neighbors_count++;
        neighbor_heuristics_cache |= (1 << neighbor_index);
    }

    ReSTIRDIReservoir out = empty_reservoir();
    PairwiseDefensiveMISWeight mis_weight_function;
    mis_weight_function.neighbor_surface_of_current_frame = true;
    const bool update_mc = center_pixel_reservoir.M > 0 && center_pixel_reservoir.UCW > 0.0f;

    int final_visibility_site = -1;       // phased: the site whose ray is the end-of-pass visibility ray of the sample `out` holds
    int start_index = valid_neighbors_M_sum == 0 ? reused_neighbors_count : 0;
    for (int neighbor_index = start_index; neighbor_index < reused_neighbors_count + 1; neighbor_index++)
    {
        const bool canonical = neighbor_index == reused_neighbors_count;
        if (!canonical && reused_neighbors_count <= 32)
            if ((neighbor_heuristics_cache & (1 << neighbor_index)) == 0)
                continue;

        int neighbor_pixel_index = restir_spatial_neighbor_pixel_index(rd, neighbor_index, reused_neighbors_count, rs.spatial_pass.reuse_radius, center_pixel_coords, res, cos_sin_theta_rotation);
        if (neighbor_pixel_index == -1)
            continue;
        if (!canonical && reused_neighbors_count > 32)
            if (!restir_check_neighbor_similarity_heuristics(rd, neighbor_pixel_index, center_pixel_index, center.shading_point, center.shading_normal, false))
                continue;

        ReSTIRDIReservoir neighbor_reservoir = canonical ? center_pixel_reservoir : load_reservoir(spatial_input + neighbor_pixel_index);

        float target_function_at_center = 0.0f;
        const bool do_neighbor_target_function_visibility = restir_include_spatial_visibility(rd, neighbor_index, reused_neighbors_count);
        if (neighbor_reservoir.UCW > 0.0f)
        {
            if (canonical)
                target_function_at_center = neighbor_reservoir.sample.target_function;
            else if (do_neighbor_target_function_visibility)
                target_function_at_center = restir_evaluate_target_function<true, MODE>(rd, neighbor_reservoir.sample, center, rng, sites, 2 + 2 * neighbor_index);
            else
                target_function_at_center = restir_evaluate_target_function<false, MODE>(rd, neighbor_reservoir.sample, center, rng, sites, 2 + 2 * neighbor_index);
            if (MODE == RESTIR_COLLECT && (canonical ? !(neighbor_reservoir.sample.flags & RESTIR_DI_FLAGS_UNOCCLUDED) : !do_neighbor_target_function_visibility))
            {
                // if this sample is the one the pass keeps, the end-of-pass visibility reuse traces it from the pixel: that ray takes this site
                float distance_to_light;
                float3 sample_direction;
                restir_visibility_reuse_ray(rd, neighbor_reservoir.sample, center.shading_point, sample_direction, distance_to_light);
                sites.emit_any_hit(canonical ? 0 : 2 + 2 * neighbor_index, center.shading_point, sample_direction, distance_to_light, center.last_hit_primitive_index);
            }
        }

        float jacobian_determinant = 1.0f;
        if (target_function_at_center > 0.0f && neighbor_reservoir.UCW > 0.0f && !canonical && !(neighbor_reservoir.sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE))
        {
            // Utils.h:199-202: the neighbour's first hit WITHOUT the normal offset
            jacobian_determinant = restir_jacobian_reconnection_shift(rd, neighbor_reservoir, center.shading_point, ld3_rw(rd.g_buffer.first_hits, neighbor_pixel_index));
            if (jacobian_determinant == -1.0f && MODE != RESTIR_COLLECT)     // COLLECT saw the target function before visibility: the neighbour may still be resampled
            {
                out.M += neighbor_reservoir.M;
                continue;
            }
        }

        float mis_weight = mis_weight_function.template get<MODE>(rd, neighbor_reservoir, center_pixel_reservoir, target_function_at_center, neighbor_pixel_index, valid_neighbors_count, valid_neighbors_M_sum,
                                                                  update_mc, canonical, rng, sites, 3 + 2 * neighbor_index);

        if (MODE != RESTIR_COLLECT && reservoir_combine_with(out, neighbor_reservoir, mis_weight, target_function_at_center, jacobian_determinant, rng))
        {
            final_visibility_site = -1;
            if (do_neighbor_target_function_visibility)
                out.sample.flags |= RESTIR_DI_FLAGS_UNOCCLUDED;
            else if (canonical)
            {
                out.sample.flags |= neighbor_reservoir.sample.flags & RESTIR_DI_FLAGS_UNOCCLUDED;
                final_visibility_site = 0;
            }
            else
            {
                out.sample.flags &= ~RESTIR_DI_FLAGS_UNOCCLUDED;
                final_visibility_site = 2 + 2 * neighbor_index;
            }
        }
    }
    if (MODE == RESTIR_COLLECT)
        return;

    reservoir_end_with_normalization(out, 1.0f, 1.0f);

    // end-of-pass visibility reuse (:262-336)
    if (rs.temporal_pass.do_temporal_reuse_pass || rs.spatial_pass.number_of_passes - 1 != rs.spatial_pass.spatial_pass_index)
    {
        if (MODE == RESTIR_RESOLVE && final_visibility_site >= 0)
            restir_visibility_reuse<RESTIR_RESOLVE>(rd, out, center.shading_point, center.last_hit_primitive_index, rng, sites, final_visibility_site);
        else
            restir_visibility_reuse(rd, out, center.shading_point, center.last_hit_primitive_index, rng);     // RESOLVE: a sample whose ray was not queued (none by construction)
    }

    if (rs.m_cap > 0)
        out.M = mini(out.M, rs.m_cap);
    store_reservoir(rs.spatial_pass.output_reservoirs + center_pixel_index, out);
}
} // namespace pt
