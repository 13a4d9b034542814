// Multiple-scattering energy compensation terms that read the pre-baked directional-albedo LUTs.
// reference: src/Device/includes/BSDFs/MicrofacetEnergyCompensation.h:88-703 (glass),
//            src/Device/includes/BSDFs/PrincipledEnergyCompensation.h:13-112 (glossy base, clearcoat).
// LUT dimensions: src/Renderer/Baker/GPUBakerConstants.h:14-31.
#pragma once

#include "bsdf_microfacet.cuh"
#include "ggx_glass_correction_table.inc"

namespace pt
{
PT_DEV float glass_correction_row(const float (*row)[2], float roughness)
{
    // Same comparison chain as the reference: <= 0, <= 0.1f, ..., <= 1.0f; above 1 nothing is assigned there
    // (undefined in the reference; the last segment's end value is used here).
    const float starts[11] = {0.0f, 0.0f, 0.1f, 0.2f, 0.3f, 0.4f, 0.5f, 0.6f, 0.7f, 0.8f, 0.9f};
    const float ends[11] = {0.0f, 0.1f, 0.2f, 0.3f, 0.4f, 0.5f, 0.6f, 0.7f, 0.8f, 0.9f, 1.0f};
    for (int s = 0; s < 11; s++)
    {
        if (roughness <= ends[s])
        {
            float a = row[s][0], b = row[s][1];
            if (a == b)
                return a;
            return lerpf(a, b, (roughness - starts[s]) / 0.1f);
        }
    }
    return row[10][1];
}

// GGX_glass_energy_conservation_get_correction_exponent (MicrofacetEnergyCompensation.h:88-635)
PT_DEV float glass_correction_exponent(float roughness, float relative_eta)
{
    if (is_zero(roughness) || fabsf(1.0f - relative_eta) < 1.0e-3f)
        return 2.5f;

    // lower bound: interval (k_i, k_{i+1}] selects row i of the lower table. Outside (1.01, 3.0] the reference leaves
    // both values uninitialised (undefined behaviour); row 0 / row 8 are used here.
    float lower_bound = kGlassCorrectionLowerKnots[0];
    int lower_row = 0;
#pragma unroll
    for (int i = 0; i < 9; i++)
    {
        if (relative_eta > kGlassCorrectionLowerKnots[i] && relative_eta <= kGlassCorrectionHigherKnots[i + 1])
        {
            lower_bound = kGlassCorrectionLowerKnots[i];
            lower_row = i;
        }
    }
    if (relative_eta > 3.0f)
    {
        lower_bound = kGlassCorrectionLowerKnots[8];
        lower_row = 8;
    }
    float lower_correction = glass_correction_row(kGlassCorrectionLower[lower_row], roughness);

    int higher_row = 9;
#pragma unroll
    for (int i = 8; i >= 0; i--)
        if (relative_eta <= kGlassCorrectionHigherKnots[i])
            higher_row = i;
    float higher_bound = kGlassCorrectionHigherKnots[higher_row];
    float higher_correction = glass_correction_row(kGlassCorrectionHigher[higher_row], roughness);
    if (higher_row == 0)
        return higher_correction;   // relative_eta <= 1.01: no lower knot exists (uninitialised read in the reference)

    return lerpf(lower_correction, higher_correction, (relative_eta - lower_bound) / (higher_bound - lower_bound));
}

// get_GGX_energy_compensation_dielectrics (MicrofacetEnergyCompensation.h:637-703), multiple scattering ON
PT_DEV float ggx_energy_compensation_dielectrics(const HIPRTRenderData& rd, const Material& material, const RayVolumeState& vs, float eta_t, float eta_i, float relative_eta, float NoV)
{
    float compensation_term = 1.0f;
    if (material.thin_film < 1.0f)
    {
        bool inside_object = vs.inside_material;
        float relative_eta_for_correction = inside_object ? 1.0f / relative_eta : relative_eta;
        float exponent_correction = 2.5f;
        if (!material.thin_walled)
            exponent_correction = glass_correction_exponent(material.roughness, relative_eta_for_correction);

        float view_direction_tex_fetch = powf(maxf(1.0e-3f, NoV), 1.0f / exponent_correction);
        float F0 = F0_from_eta(eta_t, eta_i);
        float F0_remapped = sqrtf(sqrtf(F0));
        float3 uvw = f3(view_direction_tex_fetch, material.roughness, F0_remapped);
        if (!material.thin_walled)
            compensation_term = sample_lut3d_r(inside_object ? rd.bsdfs_data.GGX_Ess_glass_inverse : rd.bsdfs_data.GGX_Ess_glass, uvw);
        else
            compensation_term = sample_lut3d_r(rd.bsdfs_data.GGX_Ess_thin_glass, uvw);

        compensation_term = lerpf(compensation_term, 1.0f, material.thin_film * material.roughness);
    }
    return compensation_term;
}

// principled_specular_relative_ior (Principled.h:143-177)
PT_DEV float principled_specular_relative_ior(const Material& material, float incident_medium_ior)
{
    float incident_layer_ior = lerpf(incident_medium_ior, material.coat_ior, material.coat);
    float relative_ior = material.ior / incident_layer_ior;
    if (relative_ior < 1.0f)
        relative_ior = 1.0f / relative_ior;
    return relative_ior;
}

// get_principled_energy_compensation_glossy_base (PrincipledEnergyCompensation.h:13-38)
PT_DEV float energy_compensation_glossy_base(const HIPRTRenderData& rd, const Material& material, float incident_medium_ior, float NoV)
{
    float relative_ior = principled_specular_relative_ior(material, incident_medium_ior);
    if (fabsf(relative_ior - 1.0f) < 1.0e-3f)
        relative_ior += 1.0e-3f;

    float view_dir_remapped = powf(NoV, 1.0f / 2.5f);
    float F0_remapped = sqrtf(sqrtf(F0_from_eta_t_and_relative_ior(material.ior, relative_ior)));
    float3 uvw = f3(view_dir_remapped, material.roughness, F0_remapped);
    float ms = sample_lut3d_r(rd.bsdfs_data.glossy_dielectric_Ess, uvw);

    float ms_compensation = lerpf(1.0f, ms, material.specular);
    ms_compensation = lerpf(ms_compensation, 1.0f, material.thin_film);
    return ms_compensation;
}

// get_principled_energy_compensation_clearcoat_lobe (PrincipledEnergyCompensation.h:54-81)
PT_DEV float energy_compensation_clearcoat_lobe(const HIPRTRenderData& rd, const Material& material, float incident_medium_ior, float NoV)
{
    if (material.coat == 0.0f)
        return 1.0f;

    if (fabsf(material.coat_ior / incident_medium_ior - 1.0f) < 1.0e-3f)
        incident_medium_ior += 1.0e-3f;

    float view_dir_remapped = powf(NoV, 1.0f / 2.5f);
    float F0_remapped = sqrtf(sqrtf(F0_from_eta(material.coat_ior, incident_medium_ior)));
    float3 uvw = f3(view_dir_remapped, material.coat_roughness, F0_remapped);
    float ms = sample_lut3d_r(rd.bsdfs_data.glossy_dielectric_Ess, uvw);

    float ms_compensation = lerpf(1.0f, ms, material.coat * (1.0f - material.specular_transmission));
    ms_compensation = lerpf(ms_compensation, 1.0f, material.thin_film);
    return ms_compensation;
}
} // namespace pt
