// Layered principled BSDF: coat / sheen / metal x2 / glass / specular / diffuse lobes with layer throughput
// attenuation, lobe-CDF sampling followed by a full evaluation.
// reference: src/Device/includes/BSDFs/Principled.h:47-1193, BSDFs/Lambertian.h:14-30, Dispatcher.h:18-68.
// None of the lobes includes the cosine term (Principled.h:43-44).
#pragma once

#include "bsdf_energy.cuh"

namespace pt
{
PT_DEV float incident_medium_ior_of(const HIPRTRenderData& rd, int mat_index)
{
    return mat_index == B200PT_MAX_MATERIAL_INDEX ? 1.0f : material_ior(rd, mat_index);
}

// principled_specular_fresnel (Principled.h:109-136)
PT_DEV RGB principled_specular_fresnel(const Material& material, float relative_specular_ior, float cos_theta_i)
{
    float layer_above_IOR = material.ior / relative_specular_ior;
    RGB F_specular;
    if (material.thin_film < 1.0f)
        F_specular = RGB(full_fresnel_dielectric(cos_theta_i, relative_specular_ior));
    RGB F_thin_film;
    if (material.thin_film > 0.0f)
        F_thin_film = thin_film_fresnel(material, layer_above_IOR, cos_theta_i);
    return lerp(F_specular, F_thin_film, material.thin_film);
}

// principled_glass_eval (Principled.h:207-388)
PT_BSDF_FN RGB principled_glass_eval(const HIPRTRenderData& rd, const Material& material, RayVolumeState& vs, float3 wo, float3 wi, float& pdf)
{
    pdf = 0.0f;
    float NoV = wo.z;
    float NoL = wi.z;
    if (fabsf(NoL) < 1.0e-8f)
        return RGB(0.0f);

    bool reflecting = NoL * NoV > 0;

    float eta_i = incident_medium_ior_of(rd, vs.incident_mat_index);
    float eta_t = incident_medium_ior_of(rd, vs.outgoing_mat_index);
    eta_i = compute_dispersion_ior(material.dispersion_abbe_number, material.dispersion_scale, eta_i, fabsf(vs.sampled_wavelength));
    eta_t = compute_dispersion_ior(material.dispersion_abbe_number, material.dispersion_scale, eta_t, fabsf(vs.sampled_wavelength));

    float relative_eta = eta_t / eta_i;
    if (fabsf(relative_eta - 1.0f) < 1.0e-5f)
        relative_eta = 1.0f + 1.0e-5f;

    float3 h;
    if (reflecting)
        h = wi + wo;
    else
    {
        if (material.thin_walled)
            h = wi * f3(1.0f, 1.0f, -1.0f) + wo;
        else
            h = wi * relative_eta + wo;
    }
    h = normalize(h);
    if (h.z < 0.0f)
        h = -h;

    float HoL = dot(wi, h);
    float HoV = dot(wo, h);
    if (HoL * NoL < 0.0f || HoV * NoV < 0.0f)
        return RGB(0.0f);

    float compensation_term = ggx_energy_compensation_dielectrics(rd, material, vs, eta_t, eta_i, relative_eta, wo.z);

    RGB F_thin_film;
    if (material.thin_film > 0.0f)
        F_thin_film = thin_film_fresnel(material, eta_i, HoV);
    RGB F_no_thin_film;
    if (material.thin_film < 1.0f)
        F_no_thin_film = RGB(full_fresnel_dielectric(HoV, relative_eta));
    RGB F = lerp(F_no_thin_film, F_thin_film, material.thin_film);

    float f_reflect_proba = F.luminance();
    float roughness = get_thin_walled_roughness(material.thin_walled, material.roughness, relative_eta);
    if (f_reflect_proba < 1.0f && material.thin_film == 0.0f && material.thin_walled && roughness < 0.1f)
        f_reflect_proba += square(1.0f - f_reflect_proba) * f_reflect_proba / (1.0f - square(f_reflect_proba));

    RGB color;
    if (reflecting)
    {
        color = torrance_sparrow_GGX_eval_single(rd, roughness, material.anisotropy, F, wo, wi, h, pdf);
        color /= compensation_term;
        pdf *= f_reflect_proba;
    }
    else
    {
        float dot_prod = HoL + HoV / relative_eta;
        float dot_prod2 = dot_prod * dot_prod;
        float denom = dot_prod2 * NoL * NoV;

        float alpha_x, alpha_y;
        get_alphas(roughness, material.anisotropy, alpha_x, alpha_y);

        float D = GGX_anisotropic(alpha_x, alpha_y, h);
        float G1_V = G1_Smith(alpha_x, alpha_y, wo);
        float G1_L = G1_Smith(alpha_x, alpha_y, wi);
        float G2 = G1_V * G1_L;

        float dwm_dwi = fabsf(HoL) / dot_prod2;
        float D_pdf = G1_V / fabsf(NoV) * D * fabsf(HoV);
        pdf = dwm_dwi * D_pdf;
        pdf *= 1.0f - f_reflect_proba;

        color = rgb(material.base_color) * D * (RGB(1.0f) - F) * G2 * fabsf(HoL * HoV / denom);
        if (material.thin_walled)
            color *= rgb(material.base_color);
        color /= compensation_term;

        if (material.thin_walled)
            interior_stack_pop(vs, vs.inside_material);
        else if (vs.incident_mat_index != B200PT_MAX_MATERIAL_INDEX)
        {
            // Beer-Lambert absorption of the volume the ray just left (the incident material)
            size_t off = offsetof(SimplifiedRendererMaterial, absorption_color);
            RGB absorption_color(material_field_f(rd, vs.incident_mat_index, off), material_field_f(rd, vs.incident_mat_index, off + 4), material_field_f(rd, vs.incident_mat_index, off + 8));
            if (!absorption_color.is_white())
            {
                float absorption_at_distance = material_field_f(rd, vs.incident_mat_index, offsetof(SimplifiedRendererMaterial, absorption_at_distance));
                RGB absorption_coefficient = log(absorption_color) / absorption_at_distance;
                color = color * exp(absorption_coefficient * vs.distance_in_volume);
            }
            vs.distance_in_volume = 0.0f;
            if (vs.inside_material)
                interior_stack_pop(vs, vs.inside_material);
        }
    }
    return color;
}

// principled_glass_sample (Principled.h:393-484). Result is in the local (rotated) shading frame.
template <int GGX_SAMPLER>
PT_DEV float3 principled_glass_sample(const HIPRTRenderData& rd, const Material& material, RayVolumeState& vs, float3 wo, Rng& rng)
{
    float eta_i = incident_medium_ior_of(rd, vs.incident_mat_index);
    float eta_t = incident_medium_ior_of(rd, vs.outgoing_mat_index);
    eta_i = compute_dispersion_ior(material.dispersion_abbe_number, material.dispersion_scale, eta_i, fabsf(vs.sampled_wavelength));
    eta_t = compute_dispersion_ior(material.dispersion_abbe_number, material.dispersion_scale, eta_t, fabsf(vs.sampled_wavelength));

    float relative_eta = eta_t / eta_i;
    if (fabsf(relative_eta - 1.0f) < 1.0e-5f)
        relative_eta = 1.0f + 1.0e-5f;

    float roughness = get_thin_walled_roughness(material.thin_walled, material.roughness, relative_eta);
    float alpha_x, alpha_y;
    get_alphas(roughness, material.anisotropy, alpha_x, alpha_y);

    float3 m = GGX_sample_microfacet<GGX_SAMPLER>(wo, alpha_x, alpha_y, rng);
    float HoV = dot(wo, m);

    RGB F_thin_film;
    if (material.thin_film > 0.0f)
        F_thin_film = thin_film_fresnel(material, eta_i, HoV);
    RGB F_no_thin_film;
    if (material.thin_film < 1.0f)
        F_no_thin_film = RGB(full_fresnel_dielectric(HoV, relative_eta));
    RGB F = lerp(F_no_thin_film, F_thin_film, material.thin_film);
    float f_reflect_proba = F.luminance();
    if (f_reflect_proba < 1.0f && material.thin_film == 0.0f && material.thin_walled && roughness < 0.1f)
        f_reflect_proba += square(1.0f - f_reflect_proba) * f_reflect_proba / (1.0f - square(f_reflect_proba));

    float rand_1 = rng();
    // Uninitialised in the reference when refraction fails (total internal reflection): returned as is there.
    float3 sampled = f3(0.0f, 0.0f, 0.0f);
    if (rand_1 < f_reflect_proba)
    {
        sampled = reflect_ray(wo, m);
        interior_stack_pop(vs, false);
    }
    else
    {
        if (dot(m, wo) < 0.0f)
            m = -m;
        if (material.thin_walled)
        {
            float3 reflected = reflect_ray(wo, m);
            reflected.z *= -1.0f;
            interior_stack_pop(vs, false);
            return reflected;
        }
        refract_ray(wo, m, sampled, relative_eta);
    }
    return sampled;
}

// principled_coat_compute_darkening (Principled.h:493-525)
PT_DEV RGB principled_coat_compute_darkening(const Material& material, float relative_eta, float view_dir_fresnel)
{
    if (material.coat_darkening == 0.0f)
        return RGB(1.0f);
    float Kr = 1.0f - (1.0f - fresnel_hemispherical_albedo(relative_eta)) / (relative_eta * relative_eta);
    float Ks = view_dir_fresnel;
    float K = lerpf(Ks, Kr, material.roughness);
    RGB base_albedo = (rgb(material.base_color) + rgb(material.sheen_color) * material.sheen) / (1.0f + material.sheen);
    RGB darkening = (1.0f - K) / (RGB(1.0f) - base_albedo * K);
    darkening = lerp(RGB(1.0f), darkening, material.coat * material.coat_darkening);
    return darkening;
}

// principled_specular_compute_darkening (Principled.h:717-739)
PT_DEV RGB principled_specular_compute_darkening(const Material& material, float relative_eta)
{
    if (material.specular_darkening == 0.0f)
        return RGB(1.0f);
    float K = 1.0f - (1.0f - fresnel_hemispherical_albedo(relative_eta)) / (relative_eta * relative_eta);
    RGB base_albedo = rgb(material.base_color);
    RGB darkening = (1.0f - K) / (RGB(1.0f) - base_albedo * K);
    darkening = lerp(RGB(1.0f), darkening, material.specular * material.specular_darkening);
    return darkening;
}

// internal_eval_coat_layer (Principled.h:530-630)
PT_DEV RGB eval_coat_layer(const HIPRTRenderData& rd, const Material& material, float3 wo, float3 wi, float3 h, float incident_ior, float coat_weight, bool refracting, float coat_proba, RGB& layers_throughput, float& cumulative_pdf)
{
    if (coat_weight > 0.0f || refracting)
    {
        float coat_pdf = 0.0f;
        RGB contribution;
        if (!refracting)
        {
            contribution = torrance_sparrow_GGX_eval_dielectric(rd, material.coat_roughness, material.coat_anisotropy, material.coat_ior, incident_ior, wo, wi, h, coat_pdf);
            contribution *= coat_weight;
            contribution *= layers_throughput;
        }
        cumulative_pdf += coat_pdf * coat_proba;

        RGB layer_below_attenuation = RGB(1.0f);
        layer_below_attenuation *= 1.0f - full_fresnel_dielectric(fabsf(wi.z), incident_ior, material.coat_ior);
        float view_dir_fresnel = full_fresnel_dielectric(fabsf(wo.z), incident_ior, material.coat_ior);
        layer_below_attenuation *= 1.0f - view_dir_fresnel;

        RGB absorption = rgb(material.coat_medium_absorption);
        if (!absorption.is_white())
        {
            float incident_refracted_angle = maxf(1.0e-6f, sqrtf(1.0f - (1.0f - wi.z * wi.z) / (material.coat_ior * material.coat_ior)));
            float outgoing_refracted_angle = maxf(1.0e-6f, sqrtf(1.0f - (1.0f - wo.z * wo.z) / (material.coat_ior * material.coat_ior)));
            float traveled_distance_angle = 1.0f / incident_refracted_angle + 1.0f / outgoing_refracted_angle;
            RGB coat_absorption = exp(-(RGB(1.0f) - pow(sqrt(absorption), traveled_distance_angle)) * material.coat_medium_thickness);
            layer_below_attenuation *= coat_absorption;
        }

        layer_below_attenuation *= principled_coat_compute_darkening(material, material.coat_ior / incident_ior, view_dir_fresnel);
        layer_below_attenuation = lerp(RGB(1.0f), layer_below_attenuation, material.coat);
        layers_throughput *= layer_below_attenuation;
        return contribution;
    }
    return RGB(0.0f);
}

// internal_eval_sheen_layer (Principled.h:632-654)
PT_DEV RGB eval_sheen_layer(const HIPRTRenderData& rd, const Material& material, float3 wo, float3 wi, float sheen_weight, float sheen_proba, RGB& layers_throughput, float& cumulative_pdf)
{
    if (sheen_weight > 0.0f)
    {
        float sheen_reflectance, sheen_pdf;
        RGB contribution = sheen_ltc_eval(rd, material, wi, wo, sheen_pdf, sheen_reflectance);
        contribution *= sheen_weight;
        contribution *= layers_throughput;
        cumulative_pdf += sheen_pdf * sheen_proba;
        layers_throughput *= 1.0f - material.sheen * sheen_reflectance;
        return contribution;
    }
    return RGB(0.0f);
}

// internal_eval_metal_layer + principled_metallic_eval (Principled.h:71-80, 656-679).
// The thin-film Fresnel is evaluated even for thin_film == 0 in the reference (lerp weight 0); it is skipped here
// when the weight is exactly 0 and the metal Fresnel is finite, which yields the same bits ((1-0)*F + 0*x == F).
PT_DEV RGB eval_metal_layer(const HIPRTRenderData& rd, const Material& material, float3 wo, float3 wi, float3 h, float roughness, float anisotropy, float incident_ior, float metal_weight, float metal_proba, const RGB& layers_throughput, float& cumulative_pdf)
{
    if (metal_weight > 0.0f)
    {
        float HoL = clampf(1.0e-8f, 1.0f, dot(h, wi));
        RGB F_metal = adobe_f82_tint_fresnel(rgb(material.base_color), rgb(material.metallic_F82), rgb(material.metallic_F90), material.metallic_F90_falloff_exponent, HoL);
        RGB F = F_metal;
        if (material.thin_film != 0.0f)
            F = lerp(F_metal, thin_film_fresnel(material, incident_ior, HoL), material.thin_film);
        else
            F = lerp(F_metal, RGB(0.0f), 0.0f);

        float metal_pdf;
        RGB contribution = torrance_sparrow_GGX_eval_ms(rd, roughness, anisotropy, F, wo, wi, h, metal_pdf);
        contribution *= metal_weight;
        contribution *= layers_throughput;
        cumulative_pdf += metal_pdf * metal_proba;
        return contribution;
    }
    return RGB(0.0f);
}

// internal_eval_glass_layer (Principled.h:681-699)
PT_DEV RGB eval_glass_layer(const HIPRTRenderData& rd, const Material& material, RayVolumeState& vs, float3 wo, float3 wi, float glass_weight, float glass_proba, const RGB& layers_throughput, float& cumulative_pdf)
{
    if (glass_weight > 0.0f)
    {
        float glass_pdf;
        RGB contribution = principled_glass_eval(rd, material, vs, wo, wi, glass_pdf);
        contribution *= glass_weight;
        contribution *= layers_throughput;
        cumulative_pdf += glass_pdf * glass_proba;
        return contribution;
    }
    return RGB(0.0f);
}

// internal_eval_specular_layer (Principled.h:741-817)
PT_DEV RGB eval_specular_layer(const HIPRTRenderData& rd, const Material& material, float3 wo, float3 wi, float3 h, float incident_medium_ior, float specular_weight, float specular_proba, RGB& layers_throughput, float& cumulative_pdf)
{
    if (specular_weight > 0.0f)
    {
        float relative_ior = principled_specular_relative_ior(material, incident_medium_ior);
        float specular_pdf;
        RGB F = principled_specular_fresnel(material, relative_ior, dot(wi, h));
        RGB contribution = torrance_sparrow_GGX_eval_single(rd, material.roughness, material.anisotropy, F, wo, wi, h, specular_pdf);

        if (fabsf(relative_ior - 1.0f) > 1.0e-3f)
        {
            contribution *= lerp(RGB(1.0f), material.specular_tint * rgb(material.specular_color), material.specular);
            contribution *= specular_weight;
            contribution *= layers_throughput;

            RGB layer_below_attenuation(1.0f);
            layer_below_attenuation *= RGB(1.0f) - principled_specular_fresnel(material, relative_ior, wi.z);
            RGB view_dir_fresnel = principled_specular_fresnel(material, relative_ior, wo.z);
            layer_below_attenuation *= RGB(1.0f) - view_dir_fresnel;
            layer_below_attenuation *= principled_specular_compute_darkening(material, relative_ior);
            layer_below_attenuation = lerp(RGB(1.0f), layer_below_attenuation, material.specular);
            layers_throughput *= layer_below_attenuation;
        }
        cumulative_pdf += specular_pdf * specular_proba;
        return contribution;
    }
    return RGB(0.0f);
}

// internal_eval_diffuse_layer (Principled.h:819-837) with the Lambertian lobe (Lambertian.h:14-23)
PT_DEV RGB eval_diffuse_layer(const Material& material, float3 wi, float diffuse_weight, float diffuse_proba, const RGB& layers_throughput, float& cumulative_pdf)
{
    if (diffuse_weight > 0.0f)
    {
        float diffuse_pdf = 0.0f;
        RGB contribution(0.0f);
        if (!(wi.z <= 0.0f))
        {
            diffuse_pdf = wi.z * kInvPi;
            contribution = rgb(material.base_color) * kInvPi;
        }
        contribution *= diffuse_weight;
        contribution *= layers_throughput;
        cumulative_pdf += diffuse_pdf * diffuse_proba;
        return contribution;
    }
    return RGB(0.0f);
}

struct LobeWeights
{
    float coat, sheen, metal_1, metal_2, specular, diffuse, glass;
};

// principled_bsdf_get_lobes_weights (Principled.h:863-895)
PT_DEV LobeWeights principled_lobe_weights(const Material& material, bool outside_object)
{
    LobeWeights w;
    float outside = outside_object ? 1.0f : 0.0f;
    w.coat = material.coat * outside;
    w.sheen = material.sheen * outside;
    w.metal_1 = material.metallic * outside;
    w.metal_2 = material.metallic * outside;
    w.metal_1 = lerpf(w.metal_1, 0.0f, material.second_roughness_weight);
    w.metal_2 = lerpf(0.0f, w.metal_2, material.second_roughness_weight);
    w.specular = (1.0f - material.metallic) * (1.0f - material.specular_transmission) * material.specular * outside;
    w.diffuse = (1.0f - material.metallic) * (1.0f - material.specular_transmission) * outside;
    w.glass = !outside_object ? 1.0f : (1.0f - material.metallic) * material.specular_transmission;
    return w;
}

// principled_bsdf_get_lobes_sampling_proba (Principled.h:951-972)
PT_DEV LobeWeights principled_lobe_probabilities(const LobeWeights& w)
{
    float normalize_factor = 1.0f / (w.coat + w.sheen + w.metal_1 + w.metal_2 + w.specular + w.diffuse + w.glass);
    LobeWeights p;
    p.coat = w.coat * normalize_factor;
    p.sheen = w.sheen * normalize_factor;
    p.metal_1 = w.metal_1 * normalize_factor;
    p.metal_2 = w.metal_2 * normalize_factor;
    p.specular = w.specular * normalize_factor;
    p.diffuse = w.diffuse * normalize_factor;
    p.glass = w.glass * normalize_factor;
    return p;
}

// principled_bsdf_eval (Principled.h:974-1063)
PT_BSDF_FN RGB principled_bsdf_eval(const HIPRTRenderData& rd, const Material& material, RayVolumeState& vs, float3 view_direction, float3 shading_normal, float3 to_light_direction, float& pdf)
{
    pdf = 0.0f;

    bool outside_object = dot(view_direction, shading_normal) > 0 || material.thin_walled;
    bool refracting = dot(shading_normal, to_light_direction) < 0.0f && outside_object;
    if (dot(view_direction, shading_normal) < 0.0f)
        shading_normal = -shading_normal;

    float3 T, B;
    build_ONB(shading_normal, T, B);
    float3 wo = world_to_local_frame(T, B, shading_normal, view_direction);
    float3 wi = world_to_local_frame(T, B, shading_normal, to_light_direction);
    float3 h = normalize(wo + wi);

    float3 TR, BR;
    build_rotated_ONB(shading_normal, TR, BR, material.anisotropy_rotation * kPi);
    float3 wo_r = world_to_local_frame(TR, BR, shading_normal, view_direction);
    float3 wi_r = world_to_local_frame(TR, BR, shading_normal, to_light_direction);
    float3 h_r = normalize(wo_r + wi_r);

    LobeWeights w = principled_lobe_weights(material, outside_object);
    float incident_medium_ior = incident_medium_ior_of(rd, vs.incident_mat_index);
    LobeWeights p = principled_lobe_probabilities(w);

    RGB layers_throughput = RGB(1.0f);
    RGB final_color = RGB(0.0f);
    float not_refracting = refracting ? 0.0f : 1.0f;

    final_color += eval_coat_layer(rd, material, wo, wi, h, incident_medium_ior, w.coat, refracting, p.coat, layers_throughput, pdf);
    final_color += eval_sheen_layer(rd, material, wo, wi, w.sheen, p.sheen, layers_throughput, pdf);
    final_color += eval_metal_layer(rd, material, wo_r, wi_r, h_r, material.roughness, material.anisotropy, incident_medium_ior, w.metal_1 * not_refracting, p.metal_1, layers_throughput, pdf);
    final_color += eval_metal_layer(rd, material, wo_r, wi_r, h_r, material.second_roughness, material.anisotropy, incident_medium_ior, w.metal_2 * not_refracting, p.metal_2, layers_throughput, pdf);
    // glass before the glossy base: the specular layer must not attenuate the glass lobe (Principled.h:1038-1046)
    final_color += eval_glass_layer(rd, material, vs, wo_r, wi_r, w.glass, p.glass, layers_throughput, pdf);

    // internal_eval_glossy_base (Principled.h:843-858)
    RGB glossy_base = RGB(0.0f);
    glossy_base += eval_specular_layer(rd, material, wo_r, wi_r, h_r, incident_medium_ior, w.specular * not_refracting, p.specular, layers_throughput, pdf);
    glossy_base += eval_diffuse_layer(material, wi, w.diffuse * not_refracting, p.diffuse, layers_throughput, pdf);
    float glossy_base_compensation = energy_compensation_glossy_base(rd, material, incident_medium_ior, wo.z);
    final_color += glossy_base / glossy_base_compensation;

    if (rd.bsdfs_data.clearcoat_compensation_approximation)
        final_color /= energy_compensation_clearcoat_lobe(rd, material, incident_medium_ior, wo.z);

    return final_color;
}

// principled_bsdf_sample (Principled.h:1065-1193)
template <int GGX_SAMPLER>
PT_BSDF_FN RGB principled_bsdf_sample(const HIPRTRenderData& rd, const Material& material, RayVolumeState& vs, float3 view_direction, float3 shading_normal, float3 geometric_normal, float3& output_direction, float& pdf, Rng& rng)
{
    pdf = 0.0f;
    float3 normal = shading_normal;

    // principled_bsdf_get_lobes_weights_fringe_fix (Principled.h:903-949)
    bool outside_object = dot(view_direction, normal) > 0 || material.thin_walled;
    float glass_weight_probe = (1.0f - material.metallic) * material.specular_transmission;
    if (is_zero(glass_weight_probe) && !outside_object)
    {
        normal = reflect_ray(shading_normal, geometric_normal);
        outside_object = true;
    }
    LobeWeights w = principled_lobe_weights(material, outside_object);
    if (!outside_object)
        w.glass = 1.0f;
    LobeWeights p = principled_lobe_probabilities(w);

    float cdf[6];
    cdf[0] = p.coat;
    cdf[1] = cdf[0] + p.sheen;
    cdf[2] = cdf[1] + p.metal_1;
    cdf[3] = cdf[2] + p.metal_2;
    cdf[4] = cdf[3] + p.specular;
    cdf[5] = cdf[4] + p.diffuse;

    float rand_1 = rng();
    bool sampling_glass_lobe = rand_1 > cdf[5];
    if (sampling_glass_lobe)
    {
        float dot_shading = dot(view_direction, shading_normal);
        float dot_geometric = dot(view_direction, geometric_normal);
        if (dot_shading * dot_geometric < 0)
            normal = reflect_ray(shading_normal, geometric_normal);
    }

    if (!sampling_glass_lobe)
        interior_stack_pop(vs, false);

    if (dot(view_direction, normal) < 0)
        normal = -normal;

    float3 TR, BR;
    build_rotated_ONB(normal, TR, BR, material.anisotropy_rotation * kPi);
    float3 wo_r = world_to_local_frame(TR, BR, normal, view_direction);

    // Left untouched when no branch below writes it (cannot happen: the branches are exhaustive)
    output_direction = f3(0.0f, 0.0f, 0.0f);
    if (rand_1 < cdf[0])
    {
        float3 TC, BC;
        build_rotated_ONB(normal, TC, BC, material.coat_anisotropy_rotation * kPi);
        float3 wo_c = world_to_local_frame(TC, BC, normal, view_direction);
        output_direction = local_to_world_frame(TC, BC, normal, microfacet_GGX_sample_reflection<GGX_SAMPLER>(material.coat_roughness, material.coat_anisotropy, wo_c, rng));
    }
    else if (rand_1 < cdf[1])
    {
        float3 T, B;
        build_ONB(normal, T, B);
        float3 wo = world_to_local_frame(T, B, normal, view_direction);
        output_direction = local_to_world_frame(T, B, normal, sheen_ltc_sample(rd, material, wo, rng));
    }
    else if (rand_1 < cdf[2])
        output_direction = local_to_world_frame(TR, BR, normal, microfacet_GGX_sample_reflection<GGX_SAMPLER>(material.roughness, material.anisotropy, wo_r, rng));
    else if (rand_1 < cdf[3])
        output_direction = local_to_world_frame(TR, BR, normal, microfacet_GGX_sample_reflection<GGX_SAMPLER>(material.second_roughness, material.anisotropy, wo_r, rng));
    else if (rand_1 < cdf[4])
        output_direction = local_to_world_frame(TR, BR, normal, microfacet_GGX_sample_reflection<GGX_SAMPLER>(material.roughness, material.anisotropy, wo_r, rng));
    else if (rand_1 < cdf[5])
        output_direction = cosine_weighted_sample_around_normal(normal, rng);
    else
        output_direction = local_to_world_frame(TR, BR, normal, principled_glass_sample<GGX_SAMPLER>(rd, material, vs, wo_r, rng));

    if (dot(output_direction, shading_normal) < 0 && !sampling_glass_lobe)
        return RGB(0.0f);

    // Evaluated with the ORIGINAL shading normal so that eval() can tell inside from outside (Principled.h:1187-1192)
    return principled_bsdf_eval(rd, material, vs, view_direction, shading_normal, output_direction, pdf);
}

// bsdf_dispatcher_eval / _sample with BSDFOverride == BSDF_NONE and no strong energy conservation (Dispatcher.h:18-68)
PT_DEV RGB bsdf_eval(const HIPRTRenderData& rd, const Material& material, RayVolumeState& vs, float3 view_direction, float3 shading_normal, float3 to_light_direction, float& pdf)
{
    return principled_bsdf_eval(rd, material, vs, view_direction, shading_normal, to_light_direction, pdf);
}

template <int GGX_SAMPLER>
PT_DEV RGB bsdf_sample(const HIPRTRenderData& rd, const Material& material, RayVolumeState& vs, float3 view_direction, float3 shading_normal, float3 geometric_normal, float3& sampled_direction, float& pdf, Rng& rng)
{
    return principled_bsdf_sample<GGX_SAMPLER>(rd, material, vs, view_direction, shading_normal, geometric_normal, sampled_direction, pdf, rng);
}
} // namespace pt
