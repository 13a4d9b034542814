// ReSTIR DI device code: reservoirs, target function, reconnection-shift jacobian, neighbour selection and similarity
// heuristics, pairwise-defensive MIS weights, final shading of the output reservoir.
// reference: src/Device/includes/ReSTIR/DI/{Reservoir.h:20-172, Surface.h:13-71, Utils.h:17-478, SpatiotemporalMISWeight.h:193-290,
//            SpatiotemporalNormalizationWeight.h:270-278, FinalShading.h:16-115}, src/Device/includes/Sampling.h:25-55.
// Built for the reference's default ReSTIR DI kernel options (KernelOptions.h:270-366), which are what config C4 uses:
// presampled lights, visibility reuse, bias correction with visibility, spatial target-function visibility, pairwise-
// defensive MIS weights, no visibility in the initial target function.
#pragma once

#include "ris.cuh"
#include "launch.h"

namespace pt
{
enum { RESTIR_DI_FLAGS_NONE = 0, RESTIR_DI_FLAGS_ENVMAP_SAMPLE = 1, RESTIR_DI_FLAGS_BSDF_REFRACTION = 2, RESTIR_DI_FLAGS_UNOCCLUDED = 4 };

PT_DEV float3 to_f3(const b200pt_float3& v) { return f3(v.x, v.y, v.z); }
PT_DEV void set_f3(b200pt_float3& d, float3 v) { d.x = v.x; d.y = v.y; d.z = v.z; }

// 36-byte reservoirs move as nine 32-bit words (the arrays are only 4-byte aligned)
PT_DEV ReSTIRDIReservoir load_reservoir(const ReSTIRDIReservoir* p)
{
    ReSTIRDIReservoir r;
    const uint32_t* s = reinterpret_cast<const uint32_t*>(p);
    uint32_t* d = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
    for (int i = 0; i < 9; i++)
        d[i] = s[i];
    return r;
}
PT_DEV void store_reservoir(ReSTIRDIReservoir* p, const ReSTIRDIReservoir& r)
{
    uint32_t* d = reinterpret_cast<uint32_t*>(p);
    const uint32_t* s = reinterpret_cast<const uint32_t*>(&r);
#pragma unroll
    for (int i = 0; i < 9; i++)
        d[i] = s[i];
}
PT_DEV ReSTIRDIReservoir empty_reservoir()
{
    ReSTIRDIReservoir r;
    r.M = 0; r.weight_sum = 0.0f; r.UCW = 0.0f;
    r.sample.emissive_triangle_index = -1;
    r.sample.point_on_light_source.x = r.sample.point_on_light_source.y = r.sample.point_on_light_source.z = 0.0f;
    r.sample.target_function = 0.0f;
    r.sample.flags = 0;
    r.sample._pad0[0] = r.sample._pad0[1] = r.sample._pad0[2] = 0;
    return r;
}

// ReSTIRDIReservoir::add_one_candidate / combine_with / end / end_with_normalization (Reservoir.h:34-97)
PT_DEV void reservoir_add_one_candidate(ReSTIRDIReservoir& r, const ReSTIRDISample& new_sample, float weight, Rng& rng)
{
    r.M++;
    r.weight_sum += weight;
    if (rng() < weight / r.weight_sum)
        r.sample = new_sample;
}
PT_DEV bool reservoir_combine_with(ReSTIRDIReservoir& r, const ReSTIRDIReservoir& other, float mis_weight, float target_function, float jacobian_determinant, Rng& rng)
{
    if (other.UCW <= 0.0f)
    {
        r.M += other.M;
        return false;
    }
    float reservoir_sample_weight = mis_weight * target_function * other.UCW * jacobian_determinant;
    r.M += other.M;
    r.weight_sum += reservoir_sample_weight;
    if (rng() < reservoir_sample_weight / r.weight_sum)
    {
        r.sample = other.sample;
        r.sample.target_function = target_function;
        return true;
    }
    return false;
}
PT_DEV void reservoir_end(ReSTIRDIReservoir& r)
{
    if (r.weight_sum == 0.0f)
        r.UCW = 0.0f;
    else
        r.UCW = 1.0f / r.sample.target_function * r.weight_sum;
}
PT_DEV void reservoir_end_with_normalization(ReSTIRDIReservoir& r, float normalization_numerator, float normalization_denominator)
{
    if (r.weight_sum == 0.0f || r.weight_sum < 1.0e-10f || r.weight_sum > 1.0e10f || normalization_denominator == 0.0f || normalization_numerator == 0.0f)
        r.UCW = 0.0f;
    else
        r.UCW = 1.0f / r.sample.target_function * r.weight_sum * normalization_numerator / normalization_denominator;
    r.M = mini(r.M, 1000000);
}

// ---- Sampling.h:25-55 ------------------------------------------------------------------------------------------------
PT_DEV float radical_inverse_base_2(unsigned int index)
{
    index = (index << 16u) | (index >> 16u);
    index = ((index & 0x55555555u) << 1u) | ((index & 0xAAAAAAAAu) >> 1u);
    index = ((index & 0x33333333u) << 2u) | ((index & 0xCCCCCCCCu) >> 2u);
    index = ((index & 0x0F0F0F0Fu) << 4u) | ((index & 0xF0F0F0F0u) >> 4u);
    index = ((index & 0x00FF00FFu) << 8u) | ((index & 0xFF00FF00u) >> 8u);
    return (float)index * 2.3283064365386963e-10f;
}
PT_DEV float2 sample_hammersley_2D(unsigned int number_of_points, unsigned int point_index)
{
    return make_float2((float)point_index / (float)number_of_points, radical_inverse_base_2(point_index));
}
PT_DEV float2 sample_in_disk_uv(float radius, float2 uv)
{
    float r_sqrt_v = radius * sqrtf(uv.y);
    return make_float2(r_sqrt_v * cosf(kTwoPi * uv.x), r_sqrt_v * sinf(kTwoPi * uv.x));
}

// ---- Surface.h ---------------------------------------------------------------------------------------------------------
struct ReSTIRSurface
{
    alignas(16) Material material;
    RayVolumeState ray_volume_state;
    int last_hit_primitive_index;
    float3 view_direction, shading_normal, geometric_normal, shading_point;
};

PT_DEV bool use_prev_frame_g_buffer(const HIPRTRenderData& rd) { return rd.render_settings.restir_di_settings.temporal_pass.do_temporal_reuse_pass; }

// FULL = false: only what the resolve phase of a phased pass reads (points, normals, primitive), not the 256-byte material
template <bool FULL = true>
PT_DEV void get_pixel_surface(const HIPRTRenderData& rd, int pixel_index, bool previous_frame, ReSTIRSurface& s)
{
    const GBuffer& g = previous_frame ? rd.g_buffer_prev_frame : rd.g_buffer;
    if (FULL)
    {
        load_material(s.material, g.materials + pixel_index);
        copy_volume_state(&s.ray_volume_state, g.ray_volume_states + pixel_index);
        s.view_direction = ld3_rw(g.view_directions, pixel_index);
    }
    s.last_hit_primitive_index = g.first_hit_prim_index[pixel_index];
    s.shading_normal = ld3_rw(g.shading_normals, pixel_index);
    s.geometric_normal = ld3_rw(g.geometric_normals, pixel_index);
    s.shading_point = ld3_rw(g.first_hits, pixel_index) + s.shading_normal * 1.0e-4f;
}

PT_DEV RGB restir_sample_emission(const HIPRTRenderData& rd, const ReSTIRDISample& sample, float3 sample_direction)
{
    if (sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE)
    {
        float envmap_pdf;
        return envmap_eval(rd, sample_direction, envmap_pdf);
    }
    return material_buffer_emission(rd, __ldg(rd.buffers.material_indices + sample.emissive_triangle_index));
}

// ---- phased passes (launch.h: RestirScratch): a pass runs as COLLECT (everything up to its visibility rays, which are queued) ->
// ray-queue traversal -> RESOLVE (the resampling itself, with the traced answers). INLINE is the unphased pass: one kernel that
// traces per thread. Which rays a pass needs does not depend on the answers of earlier ones (the samples and surfaces are inputs,
// no random number is drawn between the neighbour search and the resampling), so COLLECT can queue them all.
enum { RESTIR_INLINE = 0, RESTIR_COLLECT = 1, RESTIR_RESOLVE = 2 };
struct RestirSites
{
    const RestirScratch* scratch = nullptr;
    uint32_t pixel = 0;        // pixel inside the launch tile
    PT_DEV uint32_t ray_id(int site) const { return (uint32_t)site * scratch->capacity + pixel; }
    PT_DEV void emit(int site, float3 origin, float3 direction, float t_max, int last_prim, int counts_index) const
    {
        const uint32_t id = ray_id(site);
        scratch->ray_o[id] = make_float4(origin.x, origin.y, origin.z, t_max);
        scratch->ray_d[id] = make_float4(direction.x, direction.y, direction.z, __int_as_float(last_prim));
        // warp-aggregated append (the lanes of the warp that are here together take consecutive entries)
        const unsigned mask = __activemask();
        uint32_t lane;
        asm("mov.u32 %0, %%laneid;" : "=r"(lane));
        const int leader = __ffs((int)mask) - 1;
        uint32_t base = 0;
        if ((int)lane == leader)
            base = atomicAdd(scratch->counts + counts_index, (uint32_t)__popc(mask));
        base = __shfl_sync(mask, base, leader);
        scratch->queue[base + (uint32_t)__popc(mask & ((1u << lane) - 1u))] = id;
    }
    PT_DEV void emit_any_hit(int site, float3 origin, float3 direction, float t_max, int last_prim) const { emit(site, origin, direction, t_max, last_prim, 2); }
    PT_DEV bool occluded(int site) const { return scratch->ray_o[ray_id(site)].w == -1.0f; }
};

// ReSTIR_DI_evaluate_target_function<withVisibility> (Utils.h:17-132)
template <bool WITH_VISIBILITY, int MODE = RESTIR_INLINE>
PT_DEV_NOINLINE float restir_evaluate_target_function(const HIPRTRenderData& rd, const ReSTIRDISample& sample, const ReSTIRSurface& surface, Rng& rng, const RestirSites sites = RestirSites(),
                                                      int site = 0)
{
    if (MODE == RESTIR_RESOLVE)
    {
        // the value COLLECT stored for this site, times the traced visibility
        float target_function = sites.scratch->target[sites.ray_id(site)];
        if (WITH_VISIBILITY && target_function != 0.0f)
        {
            bool visible = !sites.occluded(site);
            target_function *= visible;
        }
        return target_function;
    }
    if (MODE == RESTIR_COLLECT)
        sites.scratch->target[sites.ray_id(site)] = 0.0f;
    if (sample.emissive_triangle_index == -1 && !(sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE))
        return 0.0f;

    float bsdf_pdf;
    float distance_to_light = 1.0e35f;
    float3 sample_direction;
    if (sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE)
        sample_direction = matrix_X_vec(as_mat4(rd.world_settings.envmap_to_world_matrix), to_f3(sample.point_on_light_source));
    else if (WITH_VISIBILITY)
    {
        sample_direction = to_f3(sample.point_on_light_source) - surface.shading_point;
        sample_direction = sample_direction / (distance_to_light = length(sample_direction));
    }
    else
        sample_direction = normalize(to_f3(sample.point_on_light_source) - surface.shading_point);

    float cosine_term = maxf(0.0f, dot(surface.shading_normal, sample_direction));
    if (cosine_term == 0.0f)
        return 0.0f;

    RayVolumeState trash_volume_state = surface.ray_volume_state;
    RGB bsdf_color = bsdf_eval(rd, surface.material, trash_volume_state, surface.view_direction, surface.shading_normal, sample_direction, bsdf_pdf);
    RGB sample_emission = restir_sample_emission(rd, sample, sample_direction);

    float target_function = (bsdf_color * sample_emission * cosine_term).luminance();
    if (target_function == 0.0f)
        return 0.0f;
    if (MODE == RESTIR_COLLECT)
    {
        sites.scratch->target[sites.ray_id(site)] = target_function;
        if (WITH_VISIBILITY)
            sites.emit_any_hit(site, surface.shading_point, sample_direction, distance_to_light, surface.last_hit_primitive_index);
        return target_function;
    }
    if (WITH_VISIBILITY)
    {
        bool visible = !evaluate_shadow_ray(rd, surface.shading_point, sample_direction, distance_to_light, surface.last_hit_primitive_index, rng);
        target_function *= visible;
    }
    return target_function;
}

// ReSTIR_DI_visibility_reuse (Utils.h:138-170)
// The shadow ray ReSTIR_DI_visibility_reuse traces for a sample seen from shading_point
PT_DEV void restir_visibility_reuse_ray(const HIPRTRenderData& rd, const ReSTIRDISample& sample, float3 shading_point, float3& sample_direction, float& distance_to_light)
{
    if (sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE)
    {
        sample_direction = matrix_X_vec(as_mat4(rd.world_settings.envmap_to_world_matrix), to_f3(sample.point_on_light_source));
        distance_to_light = 1.0e35f;
    }
    else
    {
        sample_direction = to_f3(sample.point_on_light_source) - shading_point;
        sample_direction /= (distance_to_light = length(sample_direction));
    }
}

// COLLECT queues the ray at `site` and leaves the reservoir alone (the answer is applied by a later phase); RESOLVE applies the answer
template <int MODE = RESTIR_INLINE>
PT_DEV void restir_visibility_reuse(const HIPRTRenderData& rd, ReSTIRDIReservoir& reservoir, float3 shading_point, int last_primitive_hit_index, Rng& rng, const RestirSites sites = RestirSites(),
                                    int site = 0)
{
    if (reservoir.UCW <= 0.0f)
        return;
    else if (reservoir.sample.flags & RESTIR_DI_FLAGS_UNOCCLUDED)
        return;

    bool visible;
    if (MODE == RESTIR_RESOLVE)
        visible = !sites.occluded(site);
    else
    {
        float distance_to_light;
        float3 sample_direction;
        restir_visibility_reuse_ray(rd, reservoir.sample, shading_point, sample_direction, distance_to_light);
        if (MODE == RESTIR_COLLECT)
        {
            sites.emit_any_hit(site, shading_point, sample_direction, distance_to_light, last_primitive_hit_index);
            return;
        }
        visible = !evaluate_shadow_ray(rd, shading_point, sample_direction, distance_to_light, last_primitive_hit_index, rng);
    }
    if (!visible)
        reservoir.UCW = -1.0f;
    else
        reservoir.sample.flags |= RESTIR_DI_FLAGS_UNOCCLUDED;
}

// get_triangle_normal_non_normalized (LightUtils.h:55-65)
PT_DEV float3 triangle_normal_non_normalized(const HIPRTRenderData& rd, int triangle_index)
{
    int3 vi = triangle_vertex_indices(rd, triangle_index);
    float3 A = ld3(rd.buffers.vertices_positions, vi.x);
    float3 AB = ld3(rd.buffers.vertices_positions, vi.y) - A;
    float3 AC = ld3(rd.buffers.vertices_positions, vi.z) - A;
    return cross(AB, AC);
}

// get_jacobian_determinant_reconnection_shift (Utils.h:172-197)
PT_DEV float restir_jacobian_reconnection_shift(const HIPRTRenderData& rd, const ReSTIRDIReservoir& neighbor_reservoir, float3 center_pixel_shading_point, float3 neighbor_shading_point)
{
    float distance_to_light_at_center, distance_to_light_at_neighbor;
    float3 to_light_direction_at_center = to_f3(neighbor_reservoir.sample.point_on_light_source) - center_pixel_shading_point;
    float3 to_light_direction_at_neighbor = to_f3(neighbor_reservoir.sample.point_on_light_source) - neighbor_shading_point;
    to_light_direction_at_center /= (distance_to_light_at_center = length(to_light_direction_at_center));
    to_light_direction_at_neighbor /= (distance_to_light_at_neighbor = length(to_light_direction_at_neighbor));

    float3 light_source_normal = normalize(triangle_normal_non_normalized(rd, neighbor_reservoir.sample.emissive_triangle_index));
    float cosine_light_source_at_center = fabsf(dot(-to_light_direction_at_center, light_source_normal));
    float cosine_light_source_at_neighbor = fabsf(dot(-to_light_direction_at_neighbor, light_source_normal));
    float cosine_ratio = cosine_light_source_at_center / cosine_light_source_at_neighbor;
    float distance_squared_ratio = (distance_to_light_at_neighbor * distance_to_light_at_neighbor) / (distance_to_light_at_center * distance_to_light_at_center);
    float jacobian = cosine_ratio * distance_squared_ratio;

    const float jacobian_clamp = 20.0f;
    if (jacobian > jacobian_clamp || jacobian < 1.0f / jacobian_clamp || jacobian != jacobian)
        return -1.0f;
    return jacobian;
}

// ---- neighbour similarity heuristics (Utils.h:199-275). normal_similarity_heuristic returns true when the heuristic is
// ENABLED: the reference's inverted flag (Utils.h:218-224) is part of the behaviour to reproduce. -----------------------------
PT_DEV bool restir_check_neighbor_similarity_heuristics(const HIPRTRenderData& rd, int neighbor_pixel_index, int center_pixel_index, float3 current_shading_point, float3 current_normal, bool previous_frame)
{
    const ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;
    float3 neighbor_world_space_point = f3(0.0f);
    float neighbor_roughness = 0.0f;
    float current_material_roughness = 0.0f;
    if (previous_frame)
    {
        if (rs.use_plane_distance_heuristic)
            neighbor_world_space_point = ld3_rw(rd.g_buffer_prev_frame.first_hits, neighbor_pixel_index);
        if (rs.use_roughness_similarity_heuristic)
            neighbor_roughness = rd.g_buffer_prev_frame.materials[neighbor_pixel_index].roughness;
    }
    else
    {
        neighbor_world_space_point = ld3_rw(rd.g_buffer.first_hits, neighbor_pixel_index);
        neighbor_roughness = rd.g_buffer.materials[neighbor_pixel_index].roughness;
    }
    if (rs.use_roughness_similarity_heuristic)
        current_material_roughness = rd.g_buffer.materials[center_pixel_index].roughness;

    bool plane_distance_passed = true;
    if (rs.use_plane_distance_heuristic)
    {
        float3 direction_between_points = neighbor_world_space_point - current_shading_point;
        plane_distance_passed = fabsf(dot(direction_between_points, current_normal)) < rs.plane_distance_threshold;
    }
    bool normal_similarity_passed = true;
    if (!rs.use_normal_similarity_heuristic)
        normal_similarity_passed = dot(current_normal, ld3_rw(rd.g_buffer.shading_normals, neighbor_pixel_index)) > rs.normal_similarity_angle_precomp;
    bool roughness_similarity_passed = true;
    if (rs.use_roughness_similarity_heuristic)
        roughness_similarity_passed = fabsf(neighbor_roughness - current_material_roughness) < rs.roughness_similarity_threshold;

    const SimplifiedRendererMaterial* nm = (previous_frame ? rd.g_buffer_prev_frame.materials : rd.g_buffer.materials) + neighbor_pixel_index;
    Material e;
    e.emissive_texture_used = nm->emissive_texture_used;
    e.emission_strength = nm->emission_strength;
    e.emission = nm->emission;
    bool neighbor_is_emissive = material_is_emissive(e);
    return plane_distance_passed && normal_similarity_passed && roughness_similarity_passed && !neighbor_is_emissive;
}

// get_spatial_neighbor_pixel_index (Utils.h:301-366); rng_converged_neighbor_reuse is a by-value generator seeded with random_seed
PT_DEV int restir_spatial_neighbor_pixel_index(const HIPRTRenderData& rd, int neighbor_number, int neighbor_reuse_count, int neighbor_reuse_radius, int2 center_pixel_coords, int2 res, float2 cos_sin_theta_rotation)
{
    if (neighbor_number == neighbor_reuse_count)
        return center_pixel_coords.x + center_pixel_coords.y * res.x;

    const SpatialPassSettings& sp = rd.render_settings.restir_di_settings.spatial_pass;
    float2 uv = sample_hammersley_2D(neighbor_reuse_count + 1, neighbor_number + 1);
    float2 neighbor_offset_in_disk = sample_in_disk_uv((float)neighbor_reuse_radius, uv);
    float cos_theta = cos_sin_theta_rotation.x, sin_theta = cos_sin_theta_rotation.y;
    float2 rotated = make_float2(neighbor_offset_in_disk.x * cos_theta - neighbor_offset_in_disk.y * sin_theta, neighbor_offset_in_disk.x * sin_theta + neighbor_offset_in_disk.y * cos_theta);
    int2 offset = make_int2((int)rotated.x, (int)rotated.y);
    int2 p = sp.debug_neighbor_location ? make_int2(center_pixel_coords.x + 15, center_pixel_coords.y) : make_int2(center_pixel_coords.x + offset.x, center_pixel_coords.y + offset.y);
    if (p.x < 0 || p.x >= res.x || p.y < 0 || p.y >= res.y)
        return -1;
    int neighbor_pixel_index = p.x + p.y * res.x;

    if (rd.render_settings.enable_adaptive_sampling && rd.render_settings.sample_number >= rd.render_settings.adaptive_sampling_min_samples)
    {
        if (sp.allow_converged_neighbors_reuse)
        {
            Rng rng_converged_neighbor_reuse(rd.random_seed);
            if (rng_converged_neighbor_reuse() > sp.converged_neighbor_reuse_probability)
                if (rd.aux_buffers.pixel_converged_sample_count[neighbor_pixel_index] != -1)
                    return -1;
        }
        else if (rd.aux_buffers.pixel_converged_sample_count[neighbor_pixel_index] != -1)
            return -1;
    }
    return neighbor_pixel_index;
}

PT_DEV int2 restir_apply_permutation_sampling(int2 pixel_position, int random_bits)
{
    int2 offset = make_int2(random_bits & 3, (random_bits >> 2) & 3);
    pixel_position.x += offset.x; pixel_position.y += offset.y;
    pixel_position.x ^= 3; pixel_position.y ^= 3;
    pixel_position.x -= offset.x; pixel_position.y -= offset.y;
    return pixel_position;
}

// find_temporal_neighbor_index (Utils.h:426-478): x = pixel index or -1, (y, z) = back-projected pixel position
PT_DEV int3 restir_find_temporal_neighbor_index(const HIPRTRenderData& rd, float3 current_shading_point, float3 current_normal, int2 resolution, int center_pixel_index, Rng& rng)
{
    const TemporalPassSettings& tp = rd.render_settings.restir_di_settings.temporal_pass;
    float3 prev_xyz = matrix_X_point(as_mat4(rd.prev_camera.view_projection), current_shading_point);
    float2 prev = make_float2(prev_xyz.x, prev_xyz.y);
    prev.x += 1.0f; prev.y += 1.0f;
    prev.x *= 0.5f; prev.y *= 0.5f;
    float2 prev_pixel_float = make_float2(prev.x * resolution.x, prev.y * resolution.y);
    prev_pixel_float.x -= 0.5f; prev_pixel_float.y -= 0.5f;

    int temporal_neighbor_index = -1;
    for (int i = 0; i < tp.max_neighbor_search_count + 1; i++)
    {
        float2 offset = make_float2(0.0f, 0.0f);
        if (i > 0)
        {
            // make_float2(rng() - 0.5f, rng() - 0.5f): the order of the two draws is unspecified in C++; the oracle build (g++)
            // evaluates arguments right to left, so y takes the first number (either order samples the same square)
            float oy = rng() - 0.5f;
            float ox = rng() - 0.5f;
            offset = make_float2(ox * tp.neighbor_search_radius, oy * tp.neighbor_search_radius);
        }
        int2 pos = make_int2((int)roundf(prev_pixel_float.x + offset.x), (int)roundf(prev_pixel_float.y + offset.y));
        if (tp.use_permutation_sampling && i == 0)
            pos = restir_apply_permutation_sampling(pos, tp.permutation_sampling_random_bits);
        if (pos.x < 0 || pos.x >= resolution.x || pos.y < 0 || pos.y >= resolution.y)
            continue;
        temporal_neighbor_index = pos.x + pos.y * resolution.x;
        if (restir_check_neighbor_similarity_heuristics(rd, temporal_neighbor_index, center_pixel_index, current_shading_point, current_normal, use_prev_frame_g_buffer(rd)))
            break;
        temporal_neighbor_index = -1;
    }
    return make_int3(temporal_neighbor_index, (int)roundf(prev_pixel_float.x), (int)roundf(prev_pixel_float.y));
}

// ReSTIRDISpatiotemporalResamplingMISWeight<PAIRWISE_MIS_DEFENSIVE> (SpatiotemporalMISWeight.h:193-290)
struct PairwiseDefensiveMISWeight
{
    float mc = 0.0f;
    // the stand-alone spatial pass (SpatialMISWeight.h:216) reads the neighbour's surface from THIS frame's G-buffer, the fused
    // pass (SpatiotemporalMISWeight.h:241) from last frame's when the temporal pass is enabled; the arithmetic is the same
    bool neighbor_surface_of_current_frame = false;

    // `site`: the ray site of the canonical sample seen from this neighbour's surface (phased passes)
    template <int MODE = RESTIR_INLINE>
    PT_DEV float get(const HIPRTRenderData& rd, const ReSTIRDIReservoir& reservoir_being_resampled, const ReSTIRDIReservoir& center_pixel_reservoir, float target_function_at_center,
                     int neighbor_pixel_index, int valid_neighbors_count, int valid_neighbors_M_sum, bool update_mc, bool resample_canonical, Rng& rng,
                     const RestirSites sites = RestirSites(), int site = 0)
    {
        const bool confidence = rd.render_settings.restir_di_settings.use_confidence_weights;
        if (!resample_canonical)
        {
            float target_function_at_neighbor = reservoir_being_resampled.sample.target_function;
            float reservoir_resampled_M = confidence ? reservoir_being_resampled.M : 1;
            float center_reservoir_M = confidence ? center_pixel_reservoir.M : 1;
            float neighbors_confidence_sum = confidence ? valid_neighbors_M_sum : 1;
            float valid_neighbor_division_term = confidence ? 1 : valid_neighbors_count;

            float nume = target_function_at_neighbor * reservoir_resampled_M;
            float denom = target_function_at_neighbor * neighbors_confidence_sum + target_function_at_center / valid_neighbor_division_term * center_reservoir_M;
            float mi = 0.0f;
            if (denom != 0.0f)
                mi = nume / denom;
            if (confidence)
                mi *= neighbors_confidence_sum / (neighbors_confidence_sum + center_reservoir_M);

            if (update_mc)
            {
                ReSTIRSurface neighbor_pixel_surface;
                if (MODE != RESTIR_RESOLVE)
                    get_pixel_surface(rd, neighbor_pixel_index, neighbor_surface_of_current_frame ? false : use_prev_frame_g_buffer(rd), neighbor_pixel_surface);
                float at_neighbor = restir_evaluate_target_function<true, MODE>(rd, center_pixel_reservoir.sample, neighbor_pixel_surface, rng, sites, site);
                float at_center = center_pixel_reservoir.sample.target_function;
                float nume_mc = at_center / valid_neighbor_division_term * center_reservoir_M;
                float denom_mc = at_neighbor * neighbors_confidence_sum + at_center / valid_neighbor_division_term * center_reservoir_M;
                float confidence_multiplier = 1.0f;
                if (confidence)
                    confidence_multiplier = reservoir_resampled_M / (center_reservoir_M + neighbors_confidence_sum);
                if (denom_mc != 0.0f)
                    mc += nume_mc / denom_mc * confidence_multiplier;
            }
            return confidence ? mi : mi / (valid_neighbors_count + 1);
        }
        if (mc == 0.0f)
            return 1.0f;
        if (confidence)
            return mc + (float)center_pixel_reservoir.M / (center_pixel_reservoir.M + valid_neighbors_M_sum);
        return (1 + mc) / (valid_neighbors_count + 1);
    }
};

// ReSTIRDITemporalResamplingMISWeight<PAIRWISE_MIS_DEFENSIVE> (TemporalMISWeight.h:204-279): the two-sample case of the
// pairwise weights, temporal neighbour first (accumulates mc), initial candidates second
struct TemporalPairwiseDefensiveMISWeight
{
    float mc = 0.0f;

    template <int MODE = RESTIR_INLINE>
    PT_DEV float get(const HIPRTRenderData& rd, const ReSTIRDIReservoir& temporal_neighbor_reservoir, const ReSTIRDIReservoir& initial_candidates_reservoir,
                     const ReSTIRSurface& temporal_neighbor_surface, float neighbor_sample_target_function_at_center, bool resample_initial_candidates, Rng& rng,
                     const RestirSites sites = RestirSites(), int site = 0)
    {
        const bool confidence = rd.render_settings.restir_di_settings.use_confidence_weights;
        if (!resample_initial_candidates)
        {
            float target_function_at_neighbor = temporal_neighbor_reservoir.sample.target_function;
            float target_function_at_center = neighbor_sample_target_function_at_center;
            float temporal_neighbor_M = confidence ? temporal_neighbor_reservoir.M : 1;
            float center_reservoir_M = confidence ? initial_candidates_reservoir.M : 1;
            float neighbors_confidence_sum = confidence ? temporal_neighbor_reservoir.M : 1;

            float nume = target_function_at_neighbor * temporal_neighbor_M;
            float denom = target_function_at_neighbor * neighbors_confidence_sum + target_function_at_center * center_reservoir_M;
            float mi = denom == 0.0f ? 0.0f : (nume / denom);
            if (confidence)
                mi *= neighbors_confidence_sum / (neighbors_confidence_sum + center_reservoir_M);

            float at_neighbor = restir_evaluate_target_function<true, MODE>(rd, initial_candidates_reservoir.sample, temporal_neighbor_surface, rng, sites, site);
            float at_center = initial_candidates_reservoir.sample.target_function;
            float nume_mc = at_center * center_reservoir_M;
            float denom_mc = at_neighbor * neighbors_confidence_sum + at_center * center_reservoir_M;
            float confidence_multiplier = 1.0f;
            if (confidence)
                confidence_multiplier = neighbors_confidence_sum / (neighbors_confidence_sum + center_reservoir_M);
            if (denom_mc != 0.0f)
                mc += nume_mc / denom_mc * confidence_multiplier;
            return confidence ? mi : mi * 0.5f;
        }
        if (mc == 0.0f)
            return 1.0f;
        if (confidence)
            return mc + (float)initial_candidates_reservoir.M / (initial_candidates_reservoir.M + temporal_neighbor_reservoir.M);
        return (1.0f + mc) * 0.5f;
    }
};

// ---- final shading of the output reservoir (FinalShading.h:16-115) ---------------------------------------------------------
PT_DEV RGB evaluate_ReSTIR_DI_reservoir(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, const ReSTIRDIReservoir& reservoir, Rng& rng)
{
    RGB final_color;
    if (reservoir.UCW <= 0.0f)
        return RGB(0.0f);

    const ReSTIRDISample& sample = reservoir.sample;
    float distance_to_light;
    float3 shadow_ray_direction;
    if (sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE)
    {
        shadow_ray_direction = matrix_X_vec(as_mat4(rd.world_settings.envmap_to_world_matrix), to_f3(sample.point_on_light_source));
        distance_to_light = 1.0e35f;
    }
    else
    {
        shadow_ray_direction = to_f3(sample.point_on_light_source) - hit.inter_point;
        shadow_ray_direction = shadow_ray_direction / (distance_to_light = length(shadow_ray_direction));
    }

    bool in_shadow = false;
    if (sample.flags & RESTIR_DI_FLAGS_UNOCCLUDED)
        in_shadow = false;
    else if (rd.render_settings.restir_di_settings.do_final_shading_visibility)
        in_shadow = evaluate_shadow_ray(rd, hit.inter_point, shadow_ray_direction, distance_to_light, hit.primitive_index, rng);

    if (!in_shadow)
    {
        float bsdf_pdf;
        RayVolumeState trash_volume_state = payload.volume_state;
        RGB bsdf_color = bsdf_eval(rd, payload.material, trash_volume_state, view_direction, hit.shading_normal, shadow_ray_direction, bsdf_pdf);
        float cosine_at_evaluated_point = dot(hit.shading_normal, shadow_ray_direction);
        if (sample.flags & RESTIR_DI_FLAGS_BSDF_REFRACTION)
            cosine_at_evaluated_point = fabsf(cosine_at_evaluated_point);
        if (cosine_at_evaluated_point > 0.0f)
        {
            RGB sample_emission = restir_sample_emission(rd, sample, shadow_ray_direction);
            final_color = bsdf_color * reservoir.UCW * sample_emission * cosine_at_evaluated_point;
        }
    }
    return final_color;
}

// sample_light_ReSTIR_DI (FinalShading.h:101-115): validate_reservoir writes UCW = 0 back for envmap samples without an envmap
template <int GGX_SAMPLER>
PT_DEV RGB sample_light_ReSTIR_DI(const HIPRTRenderData& rd, RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng, int2 pixel_coords, int2 res)
{
    int pixel_index = pixel_coords.x + pixel_coords.y * res.x;
    ReSTIRDIReservoir* out = rd.render_settings.restir_di_settings.restir_output_reservoirs + pixel_index;
    ReSTIRDIReservoir reservoir = load_reservoir(out);
    if ((reservoir.sample.flags & RESTIR_DI_FLAGS_ENVMAP_SAMPLE) && rd.world_settings.ambient_light_type != AMBIENT_ENVMAP)
    {
        reservoir.UCW = 0.0f;
        out->UCW = 0.0f;
    }
    return evaluate_ReSTIR_DI_reservoir(rd, payload, hit, view_direction, reservoir, rng);
}
} // namespace pt
