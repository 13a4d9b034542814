// Environment-map lighting: alias-table / CDF texel sampling, evaluation, MIS with a BSDF sample.
// reference: src/Device/includes/Envmap.h:29-246, Texture.h:196-206.
#pragma once

#include "lights.cuh"

namespace pt
{
PT_DEV const Mat4& as_mat4(const b200pt_float4x4& m) { return reinterpret_cast<const Mat4&>(m); }

PT_DEV RGB sample_environment_map_texture(const WorldSettings& ws, float2 uv)
{
    const b200pt_tex2d* tex = reinterpret_cast<const b200pt_tex2d*>(ws.envmap);
    b200pt_tex2d t;
    t.texels = tex->texels; t.width = tex->width; t.height = tex->height; t.channels = tex->channels; t.bytes_per_channel = tex->bytes_per_channel;
    RGBA c = sample_tex2d_rgba(t, uv, false);
    return RGB(c.r, c.g, c.b) * ws.envmap_intensity;
}

// eval_envmap_no_pdf (Envmap.h:29-38); 'direction' in world space
PT_DEV RGB eval_envmap_no_pdf(const WorldSettings& ws, float3 direction)
{
    float3 rotated = matrix_X_vec(as_mat4(ws.world_to_envmap_matrix), direction);
    float u = 0.5f + atan2f(rotated.z, rotated.x) * kInv2Pi;
    float v = 0.5f + asinf(rotated.y) * kInvPi;
    return sample_environment_map_texture(ws, make_float2(u, 1.0f - v));
}

// envmap_cdf_search (Envmap.h:40-75)
PT_DEV void envmap_cdf_search(const WorldSettings& ws, float value, int& x, int& y)
{
    unsigned int lower = 0;
    int upper = ws.envmap_height - 1;
    int x_index = ws.envmap_width - 1;
    while (lower < (unsigned int)upper)
    {
        int y_index = (lower + upper) / 2;
        int idx = y_index * ws.envmap_width + x_index;
        if (value < __ldg(ws.envmap_cdf + idx))
            upper = y_index;
        else
            lower = y_index + 1;
    }
    y = (int)max(min(lower, ws.envmap_height), 0u);

    lower = 0;
    upper = ws.envmap_width - 1;
    int y_index = y;
    while (lower < (unsigned int)upper)
    {
        int x_idx = (lower + upper) / 2;
        int idx = y_index * ws.envmap_width + x_idx;
        if (value < __ldg(ws.envmap_cdf + idx))
            upper = x_idx;
        else
            lower = x_idx + 1;
    }
    x = (int)max(min(lower, ws.envmap_width), 0u);
}

// envmap_sample (Envmap.h:77-124). ESS: 1 binary search, 2 alias table
template <int ESS>
PT_DEV RGB envmap_sample(const WorldSettings& ws, float3& sampled_direction, float& envmap_pdf, Rng& rng)
{
    int x, y;
    float total_sum = ws.envmap_total_sum;
    if (ESS == 1)
        envmap_cdf_search(ws, rng() * total_sum, x, y);
    else
    {
        int random_index = rng.random_index(ws.envmap_height * ws.envmap_width);
        float probability = __ldg(ws.alias_table_probas + random_index);
        if (rng() > probability)
            random_index = __ldg(ws.alias_table_alias + random_index);
        y = (int)(random_index / ws.envmap_width);
        x = (int)(random_index - y * ws.envmap_width);
    }

    float u = (float)x / ws.envmap_width;
    float v = (float)y / ws.envmap_height;
    float phi = u * kTwoPi;
    float theta = maxf(1.0e-5f, v * kPi);

    float cos_theta = cosf(theta);
    float sin_theta = sinf(theta);
    sampled_direction = f3(-sin_theta * cosf(phi), -cos_theta, -sin_theta * sinf(phi));
    sampled_direction = matrix_X_vec(as_mat4(ws.envmap_to_world_matrix), sampled_direction);

    RGB radiance = sample_environment_map_texture(ws, make_float2(u, 1.0f - v));
    envmap_pdf = radiance.luminance() / (total_sum * ws.envmap_intensity);
    envmap_pdf *= ws.envmap_width * ws.envmap_height;
    envmap_pdf /= (kTwoPiPi * sin_theta);
    return radiance;
}

// envmap_eval (Envmap.h:130-149)
PT_DEV RGB envmap_eval(const HIPRTRenderData& rd, float3 direction, float& pdf)
{
    const WorldSettings& ws = rd.world_settings;
    RGB radiance = eval_envmap_no_pdf(ws, direction);
    float theta = acosf(-direction.y);
    float sin_theta = sinf(theta);
    pdf = radiance.luminance() / (ws.envmap_total_sum * ws.envmap_intensity);
    pdf *= ws.envmap_width * ws.envmap_height;
    pdf /= (kTwoPiPi * sin_theta);
    return radiance;
}

// sample_environment_map_with_mis (Envmap.h:151-218), EnvmapSamplingDoBSDFMIS = TRUE
template <int ESS, int GGX_SAMPLER>
PT_DEV RGB sample_environment_map_with_mis(const HIPRTRenderData& rd, const Material& material, const RayVolumeState& volume_state, const HitInfo& hit, float3 view_direction, Rng& rng)
{
    float envmap_pdf;
    float3 sampled_direction;
    RGB envmap_color = envmap_sample<ESS>(rd.world_settings, sampled_direction, envmap_pdf, rng);
    RGB envmap_mis_contribution;

    float cosine_term = dot(hit.shading_normal, sampled_direction);
    if (envmap_pdf > 0.0f && cosine_term > 0.0f)
    {
        bool in_shadow = evaluate_shadow_ray(rd, hit.inter_point, sampled_direction, 1.0e35f, hit.primitive_index, rng);
        if (!in_shadow)
        {
            float bsdf_pdf;
            RayVolumeState trash = volume_state;
            RGB bsdf_color = bsdf_eval(rd, material, trash, view_direction, hit.shading_normal, sampled_direction, bsdf_pdf);
            float mis_weight = balance_heuristic(envmap_pdf, bsdf_pdf);
            envmap_mis_contribution = bsdf_color * cosine_term * mis_weight * envmap_color / envmap_pdf;
        }
    }

    float bsdf_sample_pdf;
    float3 bsdf_sampled_dir;
    RayVolumeState trash = volume_state;
    RGB bsdf_color = bsdf_sample<GGX_SAMPLER>(rd, material, trash, view_direction, hit.shading_normal, hit.geometric_normal, bsdf_sampled_dir, bsdf_sample_pdf, rng);
    RGB bsdf_mis_contribution;

    cosine_term = fabsf(dot(hit.shading_normal, bsdf_sampled_dir));
    if (bsdf_sample_pdf > 0.0f)
    {
        bool in_shadow = evaluate_shadow_ray(rd, hit.inter_point, bsdf_sampled_dir, 1.0e35f, hit.primitive_index, rng);
        if (!in_shadow)
        {
            float envmap_eval_pdf;
            RGB envmap_radiance = envmap_eval(rd, bsdf_sampled_dir, envmap_eval_pdf);
            if (envmap_eval_pdf > 0.0f)
            {
                float mis_weight = balance_heuristic(bsdf_sample_pdf, envmap_eval_pdf);
                bsdf_mis_contribution = envmap_radiance * mis_weight * cosine_term * bsdf_color / bsdf_sample_pdf;
            }
        }
    }
    return bsdf_mis_contribution + envmap_mis_contribution;
}

// sample_environment_map (Envmap.h:220-246)
template <int LSS, int ESS, int GGX_SAMPLER>
PT_DEV RGB sample_environment_map(const HIPRTRenderData& rd, const RayPayload& payload, const HitInfo& hit, float3 view_direction, int bounce, Rng& rng)
{
    const WorldSettings& ws = rd.world_settings;
    if (ws.ambient_light_type != AMBIENT_ENVMAP || rd.bsdfs_data.white_furnace_mode)
        return RGB(0.0f);
    if (material_is_emissive(payload.material))
        return RGB(0.0f);
    if (ws.envmap_intensity <= 0.0f)
        return RGB(0.0f);
    if (bounce == 0 && LSS == 5)
        return RGB(0.0f);
    if (ESS == 0)
        return RGB(0.0f);
    return sample_environment_map_with_mis<ESS == 0 ? 2 : ESS, GGX_SAMPLER>(rd, payload.material, payload.volume_state, hit, view_direction, rng);
}
} // namespace pt
