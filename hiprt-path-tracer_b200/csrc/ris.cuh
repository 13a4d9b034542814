// Resampled importance sampling of direct lighting: N light candidates + M BSDF candidates, one shadow ray.
// This is the reference's default DirectLightSamplingStrategy (LSS_RIS_BSDF_AND_LIGHT, KernelOptions.h:232) and the
// default strategy of bounces > 0 under ReSTIR DI (KernelOptions.h:355).
// reference: src/Device/includes/RIS/RIS.h:18-302, RIS/RIS_Reservoir.h:20-114.
#pragma once

#include "envmap.cuh"

namespace pt
{
struct RISSample
{
    int emissive_triangle_index = -1;
    float3 point_on_light_source = {0.0f, 0.0f, 0.0f};
    float target_function = 0.0f;
    bool is_bsdf_sample = false;
    RGB bsdf_sample_contribution;
    float bsdf_sample_cosine_term = 0.0f;
};

struct RISReservoir
{
    unsigned int M = 0;
    float weight_sum = 0.0f;
    float UCW = 0.0f;
    RISSample sample;

    PT_DEV void add_one_candidate(const RISSample& new_sample, float weight, Rng& rng)
    {
        M++;
        weight_sum += weight;
        if (rng() < weight / weight_sum)
            sample = new_sample;
    }

    PT_DEV void end()
    {
        if (weight_sum == 0.0f)
            UCW = 0.0f;
        else
            UCW = 1.0f / sample.target_function * weight_sum;
    }
};

// evaluate_reservoir_sample (RIS.h:18-80)
PT_DEV RGB evaluate_reservoir_sample(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, const RISReservoir& reservoir, Rng& rng)
{
    RGB final_color;
    if (reservoir.UCW <= 0.0f)
        return RGB(0.0f);

    const RISSample& sample = reservoir.sample;
    bool in_shadow;
    float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f;
    float3 shadow_ray_direction = sample.point_on_light_source - evaluated_point;
    float distance_to_light = length(shadow_ray_direction);
    float3 dir = shadow_ray_direction / distance_to_light;
    if (sample.is_bsdf_sample)
        in_shadow = false;
    else
        in_shadow = evaluate_shadow_ray(rd, evaluated_point, dir, distance_to_light, hit.primitive_index, rng);

    if (!in_shadow)
    {
        float bsdf_pdf;
        float cosine_at_evaluated_point;
        RGB bsdf_color;
        RayVolumeState trash = payload.volume_state;
        if (sample.is_bsdf_sample)
        {
            bsdf_color = sample.bsdf_sample_contribution;
            cosine_at_evaluated_point = sample.bsdf_sample_cosine_term;
        }
        else
        {
            bsdf_color = bsdf_eval(rd, payload.material, trash, view_direction, hit.shading_normal, dir, bsdf_pdf);
            cosine_at_evaluated_point = maxf(0.0f, dot(hit.shading_normal, dir));
        }

        if (cosine_at_evaluated_point > 0.0f)
        {
            int material_index = __ldg(rd.buffers.material_indices + sample.emissive_triangle_index);
            RGB sample_emission = material_buffer_emission(rd, material_index);
            final_color = bsdf_color * reservoir.UCW * sample_emission * cosine_at_evaluated_point;
        }
    }
    return final_color;
}

// sample_bsdf_and_lights_RIS_reservoir (RIS.h:82-290), RISUseVisiblityTargetFunction = FALSE, cut at the points where it traces a ray, so
// that the per-thread form below and the phased wavefront (wavefront_mis.cuh: R1 / R2) run the SAME pieces in the same order:
//   ris_light_candidates        the light-candidate loop (RIS.h:101-190)
//   ris_bsdf_candidate_sample   BSDF sample of one BSDF candidate and its ray (RIS.h:192-215)       -> closest-hit ray
//   ris_bsdf_candidate_finish   what the ray hit -> candidate sample + weight (RIS.h:217-280)
struct RISBsdfCandidate
{
    float pdf;
    float3 direction, origin;
    RGB color;
    bool refraction;
};

PT_DEV void ris_light_candidates(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, float3 evaluated_point, float inside_surface_multiplier,
                                 int nb_light_candidates, int nb_bsdf_candidates, Rng& rng, RISReservoir& reservoir)
{
    const HIPRTRenderSettings& s = rd.render_settings;
    for (int i = 0; i < nb_light_candidates; i++)
    {
        float light_sample_pdf;
        LightSourceInformation light_info;
        light_info.emissive_triangle_index = -1;

        float target_function = 0.0f;
        float candidate_weight = 0.0f;
        float3 random_light_point = uniform_sample_one_emissive_triangle(rd, rng, light_sample_pdf, light_info);
        if (light_sample_pdf > 0.0f)
        {
            float3 to_light = random_light_point - evaluated_point;
            float distance_to_light = length(to_light);
            to_light = to_light / distance_to_light;
            float cosine_at_light_source = fabsf(dot(light_info.light_source_normal, -to_light));
            float cosine_at_evaluated_point = maxf(0.0f, dot(hit.shading_normal * inside_surface_multiplier, to_light));
            if (cosine_at_evaluated_point > 0.0f && cosine_at_light_source > 1.0e-6f)
            {
                light_sample_pdf *= distance_to_light * distance_to_light;
                light_sample_pdf /= cosine_at_light_source;

                float bsdf_pdf = 0.0f;
                bool contributes_enough = check_minimum_light_contribution(s.minimum_light_contribution, light_info.emission / light_sample_pdf);
                if (!contributes_enough)
                    target_function = 0.0f;
                else
                {
                    RayVolumeState trash = payload.volume_state;
                    RGB bsdf_color = bsdf_eval(rd, payload.material, trash, view_direction, hit.shading_normal, to_light, bsdf_pdf);
                    RGB light_contribution = bsdf_color * light_info.emission * cosine_at_evaluated_point;
                    contributes_enough = check_minimum_light_contribution(s.minimum_light_contribution, light_contribution / bsdf_pdf / light_sample_pdf);
                    if (!contributes_enough)
                        target_function = 0.0f;
                    else
                        target_function = light_contribution.luminance();
                }

                float mis_weight = balance_heuristic(light_sample_pdf, (float)nb_light_candidates, bsdf_pdf, (float)nb_bsdf_candidates);
                candidate_weight = mis_weight * target_function / light_sample_pdf;
            }
        }

        RISSample light_sample;
        light_sample.is_bsdf_sample = false;
        light_sample.point_on_light_source = random_light_point;
        light_sample.target_function = target_function;
        light_sample.emissive_triangle_index = light_info.emissive_triangle_index;
        reservoir.add_one_candidate(light_sample, candidate_weight, rng);
    }
}

template <int GGX_SAMPLER>
PT_DEV RISBsdfCandidate ris_bsdf_candidate_sample(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, float3 evaluated_point,
                                                  float inside_surface_multiplier, Rng& rng)
{
    RISBsdfCandidate c;
    c.pdf = 0.0f;
    c.origin = evaluated_point;
    RayVolumeState trash = payload.volume_state;
    c.color = bsdf_sample<GGX_SAMPLER>(rd, payload.material, trash, view_direction, hit.shading_normal, hit.geometric_normal, c.direction, c.pdf, rng);
    c.refraction = dot(c.direction, hit.shading_normal * inside_surface_multiplier) < 0;
    if (c.refraction)
        c.origin = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier * -1.0f;
    return c;
}

PT_DEV void ris_bsdf_candidate_finish(const HIPRTRenderData& rd, const HitInfo& hit, const RISBsdfCandidate& c, bool hit_found, const ShadowLightRayHitInfo& light_hit, int nb_light_candidates,
                                      int nb_bsdf_candidates, RISSample& bsdf_RIS_sample, float& candidate_weight)
{
    const HIPRTRenderSettings& s = rd.render_settings;
    if (hit_found && !light_hit.hit_emission.is_black())
    {
        float cosine_at_evaluated_point = fabsf(dot(hit.shading_normal, c.direction));
        RGB light_contribution = c.color * light_hit.hit_emission * cosine_at_evaluated_point;
        float target_function = light_contribution.luminance();

        float light_pdf = pdf_of_emissive_triangle_hit(rd, light_hit, c.direction);
        light_pdf *= c.refraction ? 0.0f : 1.0f;

        bool contributes_enough = check_minimum_light_contribution(s.minimum_light_contribution, light_contribution / light_pdf / c.pdf);
        if (!contributes_enough)
            target_function = 0.0f;

        float mis_weight = balance_heuristic(c.pdf, (float)nb_bsdf_candidates, light_pdf, (float)nb_light_candidates);
        candidate_weight = mis_weight * target_function / c.pdf;

        bsdf_RIS_sample.emissive_triangle_index = light_hit.hit_prim_index;
        bsdf_RIS_sample.point_on_light_source = c.origin + c.direction * light_hit.hit_distance;
        bsdf_RIS_sample.is_bsdf_sample = true;
        bsdf_RIS_sample.bsdf_sample_contribution = c.color;
        bsdf_RIS_sample.bsdf_sample_cosine_term = cosine_at_evaluated_point;
        bsdf_RIS_sample.target_function = target_function;
    }
}

PT_DEV void ris_candidate_counts(const HIPRTRenderData& rd, int& nb_light_candidates, int& nb_bsdf_candidates)
{
    const HIPRTRenderSettings& s = rd.render_settings;
    bool low_res = s.wants_render_low_resolution && s.allow_render_low_resolution && s.accumulate;
    nb_light_candidates = low_res ? 1 : s.ris_settings.number_of_light_candidates;
    nb_bsdf_candidates = low_res ? 1 : s.ris_settings.number_of_bsdf_candidates;
}

template <int GGX_SAMPLER>
PT_DEV RISReservoir sample_bsdf_and_lights_RIS_reservoir(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng)
{
    bool inside_surface = dot(view_direction, hit.geometric_normal) < 0;
    float inside_surface_multiplier = inside_surface ? -1.0f : 1.0f;
    float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier;

    int nb_light_candidates, nb_bsdf_candidates;
    ris_candidate_counts(rd, nb_light_candidates, nb_bsdf_candidates);

    RISReservoir reservoir;
    ris_light_candidates(rd, payload, hit, view_direction, evaluated_point, inside_surface_multiplier, nb_light_candidates, nb_bsdf_candidates, rng, reservoir);

    for (int i = 0; i < nb_bsdf_candidates; i++)
    {
        float candidate_weight = 0.0f;
        RISBsdfCandidate c = ris_bsdf_candidate_sample<GGX_SAMPLER>(rd, payload, hit, view_direction, evaluated_point, inside_surface_multiplier, rng);
        RISSample bsdf_RIS_sample;
        if (c.pdf > 0.0f)
        {
            ShadowLightRayHitInfo light_hit;
            bool hit_found = evaluate_shadow_light_ray(rd, c.origin, c.direction, 1.0e35f, light_hit, hit.primitive_index, rng);
            ris_bsdf_candidate_finish(rd, hit, c, hit_found, light_hit, nb_light_candidates, nb_bsdf_candidates, bsdf_RIS_sample, candidate_weight);
        }
        reservoir.add_one_candidate(bsdf_RIS_sample, candidate_weight, rng);
    }

    reservoir.end();
    return reservoir;
}

// sample_lights_RIS (RIS.h:292-302)
template <int GGX_SAMPLER>
PT_DEV RGB sample_lights_RIS(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng)
{
    if (rd.buffers.emissive_triangles_count == 0)
        return RGB(0.0f);
    RISReservoir reservoir = sample_bsdf_and_lights_RIS_reservoir<GGX_SAMPLER>(rd, payload, hit, view_direction, rng);
    return evaluate_reservoir_sample(rd, payload, hit, view_direction, reservoir, rng);
}
} // namespace pt
