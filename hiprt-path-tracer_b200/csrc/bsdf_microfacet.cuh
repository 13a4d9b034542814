// GGX microfacet lobe: NDF, Smith masking, visible-normal sampling (Heitz 2018 and spherical caps), conductor
// multiple-scattering compensation, thin-film interference Fresnel, sheen LTC.
// reference: src/Device/includes/BSDFs/Microfacet.h:25-254, MicrofacetEnergyCompensation.h:25-51,
//            BSDFs/ThinFilm.h:13-219, BSDFs/SheenLTC.h:24-150.
#pragma once

#include "bsdf_common.cuh"

namespace pt
{
constexpr float kGGXDotProductsClamp = 1.0e-3f;

PT_DEV float GGX_anisotropic(float alpha_x, float alpha_y, float3 h)
{
    float denom = (h.x * h.x) / (alpha_x * alpha_x) + (h.y * h.y) / (alpha_y * alpha_y) + (h.z * h.z);
    return 1.0f / (kPi * alpha_x * alpha_y * denom * denom);
}

PT_DEV float G1_Smith_lambda(float alpha_x, float alpha_y, float3 w)
{
    float ax = w.x * alpha_x;
    float ay = w.y * alpha_y;
    return (-1.0f + sqrtf(1.0f + (ax * ax + ay * ay) / (w.z * w.z))) * 0.5f;
}

PT_DEV float G1_Smith(float alpha_x, float alpha_y, float3 w) { return 1.0f / (1.0f + G1_Smith_lambda(alpha_x, alpha_y, w)); }

// get_GGX_energy_compensation_conductors (MicrofacetEnergyCompensation.h:25-51), Fresnel-weighted (DoFresnel = TRUE)
PT_DEV RGB ggx_energy_compensation_conductors(const HIPRTRenderData& rd, RGB F0, float roughness, float3 local_view)
{
    float Ess = sample_lut2d_r(rd.bsdfs_data.GGX_conductor_Ess, maxf(0.0f, local_view.z), roughness);
    float kms = (1.0f - Ess) / Ess;
    return RGB(1.0f) + kms * F0;
}

// torrance_sparrow_GGX_eval<0> (Microfacet.h:86-128)
PT_DEV RGB torrance_sparrow_GGX_eval_single(const HIPRTRenderData& rd, float roughness, float anisotropy, RGB F, float3 wo, float3 wi, float3 h, float& out_pdf)
{
    out_pdf = 0.0f;
    float alpha_x, alpha_y;
    get_alphas(roughness, anisotropy, alpha_x, alpha_y);

    float D = GGX_anisotropic(alpha_x, alpha_y, h);
    float lambda_V = G1_Smith_lambda(alpha_x, alpha_y, wo);
    float G1V = 1.0f / (1.0f + lambda_V);
    float HoL = maxf(kGGXDotProductsClamp, dot(wo, h));
    float Dvisible = G1V * D * HoL / wo.z;

    float NoV = maxf(kGGXDotProductsClamp, fabsf(wo.z));
    float NoL = maxf(kGGXDotProductsClamp, fabsf(wi.z));

    out_pdf = Dvisible / (4.0f * dot(wo, h));
    if (out_pdf == 0.0f)
        return RGB(0.0f);

    float lambda_L = G1_Smith_lambda(alpha_x, alpha_y, wi);
    if (rd.bsdfs_data.GGX_masking_shadowing == HeightUncorrelated)
    {
        float G1L = 1.0f / (1.0f + lambda_L);
        float G2 = G1V * G1L;
        return F * D * G2 / (4.0f * NoL * NoV);
    }
    float G2 = 1.0f / (1.0f + lambda_V + lambda_L);
    return F * D * G2 / (4.0f * NoL * NoV);
}

// torrance_sparrow_GGX_eval<1>: the compensation term is evaluated FIRST in the reference (Microfacet.h:130-137)
PT_DEV RGB torrance_sparrow_GGX_eval_ms(const HIPRTRenderData& rd, float roughness, float anisotropy, RGB F, float3 wo, float3 wi, float3 h, float& out_pdf)
{
    RGB ms = ggx_energy_compensation_conductors(rd, F, roughness, wo);
    RGB single = torrance_sparrow_GGX_eval_single(rd, roughness, anisotropy, F, wo, wi, h, out_pdf);
    return single * ms;
}

// Dielectric-Fresnel overload used by the coat lobe (Microfacet.h:143-150)
PT_DEV RGB torrance_sparrow_GGX_eval_dielectric(const HIPRTRenderData& rd, float roughness, float anisotropy, float material_ior, float incident_ior, float3 wo, float3 wi, float3 h, float& out_pdf)
{
    float HoL = clampf(1.0e-8f, 1.0f, dot(h, wi));
    RGB F = RGB(full_fresnel_dielectric(HoL, incident_ior, material_ior));
    return torrance_sparrow_GGX_eval_ms(rd, roughness, anisotropy, F, wo, wi, h, out_pdf);
}

template <int GGX_SAMPLER>
PT_DEV float3 GGX_sample_microfacet(float3 wo, float alpha_x, float alpha_y, Rng& rng)
{
    float r1 = rng();
    float r2 = rng();
    float3 Vh = normalize(f3(alpha_x * wo.x, alpha_y * wo.y, wo.z));
    if (GGX_SAMPLER == 0)
    {
        // GGX_VNDF_sample (Microfacet.h:162-188)
        float lensq = Vh.x * Vh.x + Vh.y * Vh.y;
        float3 T1 = lensq > 0.0f ? f3(-Vh.y, Vh.x, 0.0f) / sqrtf(lensq) : f3(1.0f, 0.0f, 0.0f);
        float3 T2 = cross(Vh, T1);

        float r = sqrtf(r1);
        float phi = kTwoPi * r2;
        float t1 = r * cosf(phi);
        float t2 = r * sinf(phi);
        float s = 0.5f * (1.0f + Vh.z);
        t2 = (1.0f - s) * sqrtf(1.0f - t1 * t1) + s * t2;

        float3 Nh = t1 * T1 + t2 * T2 + sqrtf(maxf(0.0f, 1.0f - t1 * t1 - t2 * t2)) * Vh;
        return normalize(f3(alpha_x * Nh.x, alpha_y * Nh.y, maxf(0.0f, Nh.z)));
    }
    else
    {
        // GGX_VNDF_spherical_caps_sample (Microfacet.h:197-218)
        float phi = kTwoPi * r1;
        float z = (1.0f - r2) * (1.0f + Vh.z) - Vh.z;
        float sinTheta = sqrtf(clampf(0.0f, 1.0f, 1.0f - z * z));
        float x = sinTheta * cosf(phi);
        float y = sinTheta * sinf(phi);
        float3 Nh = f3(x, y, z) + Vh;
        return normalize(f3(alpha_x * Nh.x, alpha_y * Nh.y, Nh.z));
    }
}

// microfacet_GGX_sample_reflection (Microfacet.h:241-254)
template <int GGX_SAMPLER>
PT_DEV float3 microfacet_GGX_sample_reflection(float roughness, float anisotropy, float3 wo, Rng& rng)
{
    int below_normal = (wo.z < 0) ? -1 : 1;
    float alpha_x, alpha_y;
    get_alphas(roughness, anisotropy, alpha_x, alpha_y);
    float3 m = GGX_sample_microfacet<GGX_SAMPLER>(wo * (float)below_normal, alpha_x, alpha_y, rng);
    float3 sampled = reflect_ray(wo, m * (float)below_normal);
    return normalize(sampled);
}

// ---- thin-film interference (ThinFilm.h) ------------------------------------------------------
PT_DEV RGB eval_sensitivity(float opd, float shift)
{
    float phase = 2.0f * kPi * opd * 1.0e-6f;
    float3 val = f3(5.4856e-13f, 4.4201e-13f, 5.2481e-13f);
    float3 pos = f3(1.6810e+06f, 1.7953e+06f, 2.2084e+06f);
    float3 var = f3(4.3278e+09f, 9.3046e+09f, 6.6121e+09f);
    // val * sqrt(2 pi var) * cos(pos * phase + shift) * exp(-1 * var * phase * phase), component-wise, left to right
    float3 a = val * f3(sqrtf(2.0f * kPi * var.x), sqrtf(2.0f * kPi * var.y), sqrtf(2.0f * kPi * var.z));
    float3 cs = f3(cosf(pos.x * phase + shift), cosf(pos.y * phase + shift), cosf(pos.z * phase + shift));
    float3 ex = f3(expf(-1.0f * var.x * phase * phase), expf(-1.0f * var.y * phase * phase), expf(-1.0f * var.z * phase * phase));
    float3 xyz = a * cs * ex;
    xyz.x += 9.7470e-14f * sqrtf(2.0f * kPi * 4.5282e+09f) * cosf(2.2399e+06f * phase + shift) * expf(-4.5282e+09f * phase * phase);
    return RGB(xyz / 1.0685e-7f);
}

PT_DEV void fresnel_phase(float cos_theta_i, float eta1, float eta2, float kappa2, float& phi_par, float& phi_perp)
{
    float sinThetaSqr = 1.0f - square(cos_theta_i);
    float A = square(eta2) * (1.0f - square(kappa2)) - square(eta1) * sinThetaSqr;
    float B = sqrtf(square(A) + square(2 * square(eta2) * kappa2));
    // The reference writes sqrt((A + B) / 2.0): a double division by two (exact) and a double sqrt rounded to float.
    // For square roots 53 >= 2*24+2 bits makes that double rounding innocuous, so the fp32 sqrt gives the same bits.
    float U = sqrtf((A + B) / 2.0f);
    float V = sqrtf((B - A) / 2.0f);

    float phi_perp_y = 2.0f * eta1 * V * cos_theta_i;
    float phi_perp_x = square(U) + square(V) - square(eta1 * cos_theta_i);
    phi_perp = atan2f(phi_perp_y, phi_perp_x);

    float phi_par_y = 2.0f * eta1 * square(eta2) * cos_theta_i * (2 * kappa2 * U - (1.0f - square(kappa2)) * V);
    float phi_par_x = square(square(eta2) * (1.0f + square(kappa2)) * cos_theta_i) - square(eta1) * (square(U) + square(V));
    phi_par = atan2f(phi_par_y, phi_par_x);
}

PT_DEV void fresnel_conductor(float cos_theta_i, float eta, float k, float& Rp2, float& Rs2)
{
    float cos2 = cos_theta_i * cos_theta_i;
    float sin2 = 1.0f - cos2;
    float temp1 = eta * eta - k * k - sin2;
    float a2pb2 = sqrtf(temp1 * temp1 + 4.0f * k * k * eta * eta);
    float a = sqrtf(0.5f * (a2pb2 + temp1));
    float term1 = a2pb2 + cos2;
    float term2 = 2.0f * a * cos_theta_i;
    Rs2 = (term1 - term2) / (term1 + term2);
    Rs2 = clampf(0.0f, 1.0f, Rs2);
    float term3 = a2pb2 * cos2 + sin2 * sin2;
    float term4 = term2 * sin2;
    Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
    Rp2 = clampf(0.0f, 1.0f, Rp2);
}

PT_DEV RGB RGB_hue_shift(RGB color, float hue_shift_degrees)
{
    if (hue_shift_degrees == 0.0f)
        return color;
    // double-precision matrix entries rounded to float, as in the reference (ThinFilm.h:87-110)
    float cosA = cosf(hue_shift_degrees / 180.0f * kPi);
    float sinA = sinf(hue_shift_degrees / 180.0f * kPi);
    double c = cosA, s = sinA, k = ::sqrt(1. / 3.);
    float m00 = (float)(c + (1.0 - c) / 3.0);
    float m01 = (float)(1. / 3. * (1.0 - c) - k * s);
    float m02 = (float)(1. / 3. * (1.0 - c) + k * s);
    float m11 = (float)(c + 1. / 3. * (1.0 - c));
    RGB out;
    out.r = color.r * m00 + color.g * m01 + color.b * m02;
    out.g = color.r * m02 + color.g * m11 + color.b * m01;
    out.b = color.r * m01 + color.g * m02 + color.b * m11;
    out.clamp(0.0f, 1.0f);
    return out;
}

// thin_film_fresnel (ThinFilm.h:112-219)
PT_BSDF_FN RGB thin_film_fresnel(const Material& material, float ambient_IOR, float HoL)
{
    float eta1 = ambient_IOR;
    float eta2 = material.thin_film_ior;
    float eta3 = material.thin_film_do_ior_override ? material.thin_film_base_ior_override : material.ior;
    float kappa3 = material.thin_film_do_ior_override ? material.thin_film_kappa_3 : 0.0f;

    float R12p = 0.0f, R12s = 0.0f, T121p = 0.0f, T121s = 0.0f, R23p = 0.0f, R23s = 0.0f, cos_theta_2 = 0.0f;
    float cos_theta_transmission_2 = 1.0f - (1.0f - square(HoL)) * square(eta1 / eta2);
    if (cos_theta_transmission_2 <= 0.0f)
    {
        R12s = 1.0f;
        R12p = 1.0f;
        T121p = 0.0f;
        T121s = 0.0f;
    }
    else
    {
        cos_theta_2 = sqrtf(cos_theta_transmission_2);
        fresnel_conductor(HoL, eta2 / eta1, 0.0f, R12p, R12s);
        fresnel_conductor(cos_theta_2, eta3 / eta2, kappa3, R23p, R23s);
        T121p = 1.0f - R12p;
        T121s = 1.0f - R12s;
    }

    float D = material.thin_film_thickness / 1000.0f * cos_theta_2;

    float phi21p, phi21s, phi23p, phi23s;
    fresnel_phase(HoL, eta1, eta2, 0.0f, phi21p, phi21s);
    fresnel_phase(cos_theta_2, eta2, eta3, kappa3, phi23p, phi23s);
    phi21p = kPi - phi21p;
    phi21s = kPi - phi21s;

    float r123p = sqrtf(R12p * R23p);
    float r123s = sqrtf(R12s * R23s);

    float Rs = (square(T121p) * R23p) / (1.0f - R12p * R23p);
    float C0 = R12p + Rs;
    RGB I = RGB(C0);
    RGB Sm;

    float Cm = Rs - T121p;
    for (int m = 1; m <= 2; ++m)
    {
        Cm *= r123p;
        Sm = 2.0f * eval_sensitivity(m * D, m * (phi23p + phi21p));
        I += Cm * Sm;
    }

    Rs = (square(T121s) * R23s) / (1.0f - R12s * R23s);
    C0 = R12s + Rs;
    I += RGB(C0);

    Cm = Rs - T121s;
    for (int m = 1; m <= 2; ++m)
    {
        Cm *= r123s;
        Sm = 2.0f * eval_sensitivity(m * D, m * (phi23s + phi21s));
        I += Cm * Sm;
    }

    I *= 0.5f;

    float r = 2.3646381f * I.r - 0.8965361f * I.g - 0.4680737f * I.b;
    float g = -0.5151664f * I.r + 1.4264000f * I.g + 0.0887608f * I.b;
    float b = 0.0052037f * I.r - 0.0144081f * I.g + 1.0092106f * I.b;
    I = RGB(r, g, b);
    I.clamp(0.0f, 1.0f);
    return RGB_hue_shift(I, material.thin_film_hue_shift_degrees);
}

// ---- sheen LTC (SheenLTC.h) --------------------------------------------------------------------
PT_DEV RGB read_LTC_parameters(const HIPRTRenderData& rd, float roughness, float cos_theta)
{
    return sample_lut2d_rgb(rd.bsdfs_data.sheen_ltc_parameters_texture, cos_theta, 1.0f - roughness);
}

PT_DEV float get_phi(float3 direction)
{
    float p = atan2f(direction.y, direction.x);
    if (p < 0.0f)
        p += kTwoPi;
    return p;
}

PT_DEV float3 rotate_vector(float3 vec, float3 axis, float angle)
{
    float sin_angle = sinf(angle);
    float cos_angle = cosf(angle);
    return vec * cos_angle + axis * dot(vec, axis) * (1.0f - cos_angle) + sin_angle * cross(axis, vec);
}

PT_DEV float eval_ltc(float3 wi_standard, RGB AiBiRi)
{
    float3 w = f3(wi_standard.x * AiBiRi.r + wi_standard.z * AiBiRi.g, wi_standard.y * AiBiRi.r, wi_standard.z);
    float len = length(w);
    w /= len;
    float M_inv_determinant = AiBiRi.r * AiBiRi.r;
    float jacobian = M_inv_determinant / (len * len * len);
    return w.z * kInvPi * jacobian;
}

PT_DEV RGB sheen_ltc_eval(const HIPRTRenderData& rd, const Material& material, float3 wi, float3 wo, float& out_pdf, float& out_sheen_reflectance)
{
    if (wo.z <= 0.0f || wi.z <= 0.0f)
    {
        out_pdf = 0.0f;
        if (wo.z > 0.0f)
            out_sheen_reflectance = read_LTC_parameters(rd, material.sheen_roughness, wo.z).b;
        else
            out_sheen_reflectance = 0.0f;
        return RGB(0.0f);
    }
    float phi = get_phi(wo);
    float3 wi_standard = rotate_vector(wi, f3(0.0f, 0.0f, 1.0f), -phi);
    RGB AiBiRi = read_LTC_parameters(rd, material.sheen_roughness, wo.z);
    float Do = eval_ltc(wi_standard, AiBiRi);
    out_pdf = Do;
    out_sheen_reflectance = AiBiRi.b;
    return rgb(material.sheen_color) * AiBiRi.b * Do / wi.z;
}

PT_DEV float3 sheen_ltc_sample(const HIPRTRenderData& rd, const Material& material, float3 wo, Rng& rng)
{
    float3 cosine_sample = cosine_weighted_sample_z_up_frame(rng);
    RGB AiBiRi = read_LTC_parameters(rd, material.sheen_roughness, wo.z);
    float Ai_inv = 1.0f / AiBiRi.r;
    float Bi = AiBiRi.g;
    float3 w = normalize(f3(cosine_sample.x * Ai_inv - cosine_sample.z * Bi * Ai_inv, cosine_sample.y * Ai_inv, cosine_sample.z));
    return rotate_vector(w, f3(0.0f, 0.0f, 1.0f), get_phi(wo));
}
} // namespace pt
