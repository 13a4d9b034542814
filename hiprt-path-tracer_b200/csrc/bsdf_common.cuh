// Building blocks of the layered principled BSDF: frames, sampling helpers, Fresnel terms, dispersion.
// reference: src/Device/includes/ONB.h:18-76, Sampling.h:75-217, Fresnel.h:11-163, Dispersion.h:433-529.
#pragma once

#include "material.cuh"
#include "rng.cuh"

namespace pt
{
// ---- orthonormal bases (ONB.h) ---------------------------------------------------------------
PT_DEV void build_ONB(float3 N, float3& T, float3& B)
{
    if (N.z < -0.99998796f)
    {
        T = f3(0.0f, -1.0f, 0.0f);
        B = f3(-1.0f, 0.0f, 0.0f);
        return;
    }
    float nxa = -N.x / (1.0f + N.z);
    T = f3(1.0f + N.x * nxa, nxa * N.y, -N.x);
    B = f3(T.y, 1.0f - N.y * N.y / (1.0f + N.z), -N.y);
}

PT_DEV void build_rotated_ONB(float3 N, float3& T, float3& B, float basis_rotation)
{
    float3 up = fabsf(N.z) < 0.9999999f ? f3(0.0f, 0.0f, 1.0f) : f3(1.0f, 0.0f, 0.0f);
    T = normalize(cross(up, N));
    float c = cosf(basis_rotation), s = sinf(basis_rotation);
    T = T * c + cross(N, T) * s + N * dot(N, T) * (1.0f - c);
    B = cross(N, T);
}

PT_DEV float3 local_to_world_frame(float3 T, float3 B, float3 N, float3 V) { return normalize(V.x * T + V.y * B + V.z * N); }
PT_DEV float3 world_to_local_frame(float3 T, float3 B, float3 N, float3 V) { return normalize(f3(dot(V, T), dot(V, B), dot(V, N))); }
PT_DEV float3 local_to_world_frame(float3 N, float3 V)
{
    float3 T, B;
    build_ONB(N, T, B);
    return local_to_world_frame(T, B, N, V);
}

// ---- MIS heuristics (Sampling.h:75-131). The three-strategy balance heuristic takes an INT pdf_c, as the reference does.
PT_DEV float power_heuristic(float pdf_a, int nb_a, float pdf_b, int nb_b)
{
    float p_a_sqr = (nb_a * pdf_a) * (nb_a * pdf_a);
    float p_b_sqr = (nb_b * pdf_b) * (nb_b * pdf_b);
    return nb_a * pdf_a * pdf_a / (p_a_sqr + p_b_sqr);
}
PT_DEV float power_heuristic(float pdf_a, float pdf_b) { return power_heuristic(pdf_a, 1, pdf_b, 1); }
PT_DEV float balance_heuristic(float pdf_a, float nb_a, float pdf_b, float nb_b) { return pdf_a / (nb_a * pdf_a + nb_b * pdf_b); }
PT_DEV float balance_heuristic(float pdf_a, float pdf_b) { return balance_heuristic(pdf_a, 1.0f, pdf_b, 1.0f); }
PT_DEV float balance_heuristic3(float pdf_a, int nb_a, float pdf_b, int nb_b, int pdf_c, int nb_c) { return pdf_a / (nb_a * pdf_a + nb_b * pdf_b + nb_c * pdf_c); }

PT_DEV float3 reflect_ray(float3 ray_direction, float3 surface_normal)
{
    return -ray_direction + 2.0f * dot(ray_direction, surface_normal) * surface_normal;
}

PT_DEV bool refract_ray(float3 ray_direction, float3 surface_normal, float3& refract_direction, float relative_eta)
{
    float NoI = dot(ray_direction, surface_normal);
    float sin_theta_i_2 = 1.0f - NoI * NoI;
    float root_term = 1.0f - sin_theta_i_2 / (relative_eta * relative_eta);
    if (root_term < 0.0f)
        return false;
    float cos_theta_t = sqrtf(root_term);
    refract_direction = -ray_direction / relative_eta + (NoI / relative_eta - cos_theta_t) * surface_normal;
    return true;
}

PT_DEV float3 cosine_weighted_sample_around_normal(float3 normal, Rng& rng)
{
    float rand_1 = rng();
    float rand_2 = 2.0f * rng() - 1.0f;
    if (rand_1 < 1.0e-8f && rand_2 < -0.999999f && normal.z > 0.999999f)
    {
        rand_1 += 1.0e-7f;
        rand_2 += 1.0e-7f;
    }
    float theta = kTwoPi * rand_1;
    float k = sqrtf(1.0f - rand_2 * rand_2);
    float3 sphere_point = f3(k * cosf(theta), k * sinf(theta), rand_2);
    return normalize(normal + sphere_point);
}

PT_DEV float3 cosine_weighted_sample_z_up_frame(Rng& rng)
{
    float r1 = rng();
    float r2 = rng();
    float phi = kTwoPi * r1;
    float cos_theta = sqrtf(r2);
    float sin_theta = sqrtf(1 - cos_theta * cos_theta);
    return normalize(f3(cosf(phi) * sin_theta, sinf(phi) * sin_theta, cos_theta));
}

// ---- Fresnel (Fresnel.h) ---------------------------------------------------------------------
PT_DEV float F0_from_eta(float eta_t, float eta_i)
{
    float nume = (eta_t - eta_i);
    float denom = (eta_t + eta_i);
    return (nume * nume) / (denom * denom);
}
PT_DEV float F0_from_eta_t_and_relative_ior(float eta_t, float relative_eta) { return F0_from_eta(eta_t, eta_t / relative_eta); }

PT_DEV float full_fresnel_dielectric(float cos_theta_i, float relative_eta)
{
    float sin_theta_i2 = 1.0f - cos_theta_i * cos_theta_i;
    float sin_theta_t2 = sin_theta_i2 / (relative_eta * relative_eta);
    if (sin_theta_t2 >= 1.0f)
        return 1.0f;
    float cos_theta_t = sqrtf(1.0f - sin_theta_t2);
    float r_parallel = (relative_eta * cos_theta_i - cos_theta_t) / (relative_eta * cos_theta_i + cos_theta_t);
    float r_perpendicular = (cos_theta_i - relative_eta * cos_theta_t) / (cos_theta_i + relative_eta * cos_theta_t);
    return (r_parallel * r_parallel + r_perpendicular * r_perpendicular) / 2;
}
PT_DEV float full_fresnel_dielectric(float cos_theta_i, float eta_i, float eta_t) { return full_fresnel_dielectric(cos_theta_i, eta_t / eta_i); }

PT_DEV RGB adobe_f82_tint_fresnel(RGB F0, RGB F82, RGB F90, float F90_falloff_exponent, float cos_theta)
{
    RGB base_term = F0 + (F90 - F0) * powf(1.0f - cos_theta, F90_falloff_exponent);
    float lazanyi_correction = cos_theta * pow_6(1.0f - cos_theta);

    const float cos_theta_max = 1.0f / 7.0f;
    const float denom_a = cos_theta_max * pow_6(1.0f - cos_theta_max);

    RGB nume_a = (F0 + (F90 - F0) * powf(1.0f - cos_theta_max, F90_falloff_exponent)) * (RGB(1.0f) - F82);
    RGB a = nume_a / denom_a;

    RGB F = base_term - a * lazanyi_correction;
    F.clamp(0.0f, 1.0f);
    return F;
}

PT_DEV float fresnel_hemispherical_albedo(float relative_eta)
{
    return logf((10893.0f * relative_eta - 1438.2f) / (-774.4f * square(relative_eta) + 10212.0f * relative_eta + 1.0f));
}

// ---- spectral dispersion (Dispersion.h:433-529) ------------------------------------------------
PT_DEV RGB wavelength_to_RGB_fit(float wavelength)
{
    RGB c;
    if (wavelength < 463.0f)
    {
        c.r = -1.2776028240727566e-01f / (1.0f + expf((wavelength - 4.2680623367293401e+02f) / 8.2197460736637176e+00f)) + -1.3925673552505122e-11f * expf((wavelength - 45.0f) / 1.8175459086411596e+01f);
        c.r += 1.2898689750552100e-01f;
    }
    else if (wavelength > 553.0f)
    {
        c.r = 1.7963649137825513e+01f * (1.0f / 2.6577826611702449e+01f) * expf(-0.5f * square((wavelength - 6.0625724092824566e+02f) * (1.0f / 2.6577826611702449e+01f)));
        c.r += 2.5574660155104657e-03f;
    }
    else
        c.r = 0.0f;

    c.g = 3.4962267376163049e+02f * expf(-0.5f * square((wavelength - 5.4209217455705152e+02f) / -2.9598170255834638e+01f));
    c.g /= wavelength;
    c.b = expf(3.2987659944421112e+03f + (-2.0975839709372405e+05f / wavelength) - 4.6368268395094020e+02f * logf(wavelength));
    return c * 10.0f;
}

PT_DEV float sample_wavelength_uniformly(Rng& rng)
{
    float r = rng();
    return r * (830 - 360) + 360;
}

PT_DEV float compute_dispersion_ior(float dispersion_abbe_number, float dispersion_scale, float base_IOR, float wavelength)
{
    if (dispersion_scale == 0.0f)
        return base_IOR;
    const float SQUARE_587_6 = 334777.96f;
    const float POW_MIN2_LAMBDA_F_MINUS_LAMBDA_C = 0.00000191038851931481f;
    float abbe_number = dispersion_abbe_number / dispersion_scale;
    float B = (base_IOR - 1.0f) / (abbe_number * POW_MIN2_LAMBDA_F_MINUS_LAMBDA_C);
    float A = base_IOR - B / (SQUARE_587_6);
    return A + B / (wavelength * wavelength);
}

PT_DEV RGB get_dispersion_ray_color(float& wavelength, float dispersion_scale)
{
    if (dispersion_scale == 0.0f)
        return RGB(1.0f);
    if (wavelength >= 0.0f)
        return RGB(1.0f);
    wavelength *= -1.0f;
    return wavelength_to_RGB_fit(wavelength);
}
} // namespace pt
