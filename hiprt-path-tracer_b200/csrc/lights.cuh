// Next-event estimation on emissive triangles (uniform pick + MIS with a BSDF sample) and on the environment map
// (alias-table texel pick + MIS).
// reference: src/Device/includes/LightUtils.h:13-142, Lights.h:22-321, Envmap.h:29-246.
#pragma once

#include "intersect.cuh"

namespace pt
{
struct LightSourceInformation
{
    int emissive_triangle_index;
    float3 light_source_normal;
    float light_area;
    RGB emission;
};

// uniform_sample_one_emissive_triangle (LightUtils.h:13-53)
PT_DEV float3 uniform_sample_one_emissive_triangle(const HIPRTRenderData& rd, Rng& rng, float& pdf, LightSourceInformation& light_info)
{
    int random_index = rng.random_index(rd.buffers.emissive_triangles_count);
    int triangle_index = __ldg(rd.buffers.emissive_triangles_indices + random_index);
    int3 vi = triangle_vertex_indices(rd, triangle_index);
    float3 A = ld3(rd.buffers.vertices_positions, vi.x);
    float3 B = ld3(rd.buffers.vertices_positions, vi.y);
    float3 C = ld3(rd.buffers.vertices_positions, vi.z);

    float rand_1 = rng();
    float rand_2 = rng();
    float sqrt_r1 = sqrtf(rand_1);
    float u = 1.0f - sqrt_r1;
    float v = (1.0f - rand_2) * sqrt_r1;

    float3 AB = B - A;
    float3 AC = C - A;
    float3 random_point_on_triangle = A + AB * u + AC * v;

    float3 normal = cross(AB, AC);
    float length_normal = length(normal);
    if (length_normal <= 1.0e-6f)
    {
        pdf = 0.0f;
        return f3(0.0f, 0.0f, 0.0f);
    }

    light_info.emissive_triangle_index = triangle_index;
    light_info.light_source_normal = normal / length_normal;
    light_info.light_area = length_normal * 0.5f;
    light_info.emission = material_buffer_emission(rd, __ldg(rd.buffers.material_indices + triangle_index));

    pdf = 1.0f / light_info.light_area;
    pdf /= rd.buffers.emissive_triangles_count;
    return random_point_on_triangle;
}

PT_DEV float triangle_area(const HIPRTRenderData& rd, int triangle_index)
{
    int3 vi = triangle_vertex_indices(rd, triangle_index);
    float3 A = ld3(rd.buffers.vertices_positions, vi.x);
    float3 AB = ld3(rd.buffers.vertices_positions, vi.y) - A;
    float3 AC = ld3(rd.buffers.vertices_positions, vi.z) - A;
    return length(cross(AB, AC)) * 0.5f;
}

// clamp_light_contribution (LightUtils.h:79-91)
PT_DEV RGB clamp_light_contribution(RGB light_contribution, float clamp_max_value, bool clamp_condition)
{
    if (!light_contribution.has_NaN() && clamp_max_value > 0.0f && clamp_condition)
        light_contribution.clamp(-clamp_max_value, clamp_max_value);
    return light_contribution;
}

// pdf_of_emissive_triangle_hit (LightUtils.h:101-120)
PT_DEV float pdf_of_emissive_triangle_hit(const HIPRTRenderData& rd, const ShadowLightRayHitInfo& light_hit_info, float3 ray_direction)
{
    float light_area = triangle_area(rd, light_hit_info.hit_prim_index);
    float pdf = 1.0f / light_area;
    pdf /= rd.buffers.emissive_triangles_count;
    float cosine_light_source = fabsf(dot(light_hit_info.hit_shading_normal, -ray_direction));
    pdf *= light_hit_info.hit_distance * light_hit_info.hit_distance;
    pdf /= cosine_light_source;
    return pdf;
}

PT_DEV bool check_minimum_light_contribution(float minimum_contribution, RGB contribution)
{
    if (minimum_contribution > 0.0f)
        return !(contribution.r < minimum_contribution && contribution.g < minimum_contribution && contribution.b < minimum_contribution);
    return true;
}

// sample_one_light_no_MIS (Lights.h:22-65)
PT_DEV RGB sample_one_light_no_MIS(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng)
{
    float light_sample_pdf;
    LightSourceInformation light_info;
    RGB radiance;
    float3 random_light_point = uniform_sample_one_emissive_triangle(rd, rng, light_sample_pdf, light_info);
    if (!(light_sample_pdf > 0.0f))
        return RGB(0.0f);

    float3 shadow_ray_origin = hit.inter_point + hit.shading_normal * 1.0e-4f;
    float3 shadow_ray_direction = random_light_point - shadow_ray_origin;
    float distance_to_light = length(shadow_ray_direction);
    float3 dir = shadow_ray_direction / distance_to_light;

    float dot_light_source = fabsf(dot(light_info.light_source_normal, -dir));
    if (dot_light_source > 0.0f)
    {
        bool in_shadow = evaluate_shadow_ray(rd, shadow_ray_origin, dir, distance_to_light, hit.primitive_index, rng);
        if (!in_shadow)
        {
            float bsdf_pdf;
            RayVolumeState trash = payload.volume_state;
            RGB bsdf_color = bsdf_eval(rd, payload.material, trash, view_direction, hit.shading_normal, dir, bsdf_pdf);
            if (bsdf_pdf != 0.0f)
            {
                light_sample_pdf *= distance_to_light * distance_to_light;
                light_sample_pdf /= dot_light_source;
                float cosine_term = maxf(dot(hit.shading_normal, dir), 0.0f);
                radiance = light_info.emission * cosine_term * bsdf_color / light_sample_pdf;
            }
        }
    }
    return radiance;
}

// sample_one_light_bsdf (Lights.h:67-113)
template <int GGX_SAMPLER>
PT_DEV RGB sample_one_light_bsdf(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng)
{
    bool inside_surface = dot(view_direction, hit.geometric_normal) < 0;
    float inside_surface_multiplier = inside_surface ? -1.0f : 1.0f;
    RGB bsdf_radiance = RGB(0.0f);

    float direction_pdf;
    float3 sampled_direction;
    RayVolumeState trash = payload.volume_state;
    RGB bsdf_color = bsdf_sample<GGX_SAMPLER>(rd, payload.material, trash, view_direction, hit.shading_normal, hit.geometric_normal, sampled_direction, direction_pdf, rng);

    bool refraction_sampled = dot(sampled_direction, hit.shading_normal * inside_surface_multiplier) < 0;
    if (direction_pdf > 0.0f)
    {
        float3 origin = hit.inter_point + hit.shading_normal * 1.0e-4f;
        if (refraction_sampled)
            origin = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier * -1.0f;

        ShadowLightRayHitInfo light_hit;
        bool inter_found = evaluate_shadow_light_ray(rd, origin, sampled_direction, 1.0e35f, light_hit, hit.primitive_index, rng);
        if (inter_found && !light_hit.hit_emission.is_black())
        {
            float cosine_term = maxf(0.0f, dot(hit.shading_normal, sampled_direction));
            bsdf_radiance = bsdf_color * cosine_term * light_hit.hit_emission / direction_pdf;
        }
    }
    return bsdf_radiance;
}

// sample_one_light_MIS (Lights.h:115-220)
template <int GGX_SAMPLER>
PT_DEV RGB sample_one_light_MIS(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng)
{
    bool inside_surface = dot(view_direction, hit.geometric_normal) < 0;
    float inside_surface_multiplier = inside_surface ? -1.0f : 1.0f;
    float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier;

    float light_sample_pdf;
    RGB light_source_radiance_mis;
    LightSourceInformation light_info;
    float3 random_light_point = uniform_sample_one_emissive_triangle(rd, rng, light_sample_pdf, light_info);
    if (light_sample_pdf <= 0.0f)
        return RGB(0.0f);

    float3 shadow_ray_direction = random_light_point - evaluated_point;
    float distance_to_light = length(shadow_ray_direction);
    float3 dir = shadow_ray_direction / distance_to_light;

    float dot_light_source = fabsf(dot(light_info.light_source_normal, -dir));
    if (dot_light_source > 0.0f)
    {
        bool in_shadow = evaluate_shadow_ray(rd, evaluated_point, dir, distance_to_light, hit.primitive_index, rng);
        if (!in_shadow)
        {
            float bsdf_pdf;
            RayVolumeState trash = payload.volume_state;
            RGB bsdf_color = bsdf_eval(rd, payload.material, trash, view_direction, hit.shading_normal, dir, bsdf_pdf);
            if (bsdf_pdf != 0.0f)
            {
                light_sample_pdf *= distance_to_light * distance_to_light;
                light_sample_pdf /= dot_light_source;
                float mis_weight = balance_heuristic(light_sample_pdf, bsdf_pdf);
                float cosine_term = maxf(dot(hit.shading_normal, dir), 0.0f);
                light_source_radiance_mis = bsdf_color * cosine_term * light_info.emission * mis_weight / light_sample_pdf;
            }
        }
    }

    RGB bsdf_radiance_mis;
    float direction_pdf;
    float3 sampled_bsdf_direction;
    float3 bsdf_shadow_ray_origin = evaluated_point;
    RayVolumeState trash = payload.volume_state;
    RGB bsdf_color = bsdf_sample<GGX_SAMPLER>(rd, payload.material, trash, view_direction, hit.shading_normal, hit.geometric_normal, sampled_bsdf_direction, direction_pdf, rng);
    bool refraction_sampled = dot(sampled_bsdf_direction, hit.shading_normal * inside_surface_multiplier) < 0;
    if (refraction_sampled)
        bsdf_shadow_ray_origin = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier * -1.0f;

    if (direction_pdf > 0)
    {
        ShadowLightRayHitInfo light_hit;
        bool inter_found = evaluate_shadow_light_ray(rd, bsdf_shadow_ray_origin, sampled_bsdf_direction, 1.0e35f, light_hit, hit.primitive_index, rng);
        if (inter_found && !light_hit.hit_emission.is_black())
        {
            float light_pdf = pdf_of_emissive_triangle_hit(rd, light_hit, sampled_bsdf_direction);
            float mis_weight = balance_heuristic(direction_pdf, light_pdf);
            float cosine_term = fabsf(dot(hit.shading_normal, sampled_bsdf_direction));
            bsdf_radiance_mis = bsdf_color * cosine_term * light_hit.hit_emission * mis_weight / direction_pdf;
        }
    }
    return light_source_radiance_mis + bsdf_radiance_mis;
}
} // namespace pt
