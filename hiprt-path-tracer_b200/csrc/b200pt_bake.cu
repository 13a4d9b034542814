// Directional-albedo LUT bakers: Monte-Carlo integration of the GGX conductor / dielectric-Fresnel / glossy
// dielectric / rough glass (entering, exiting) / thin glass lobes over the (cos theta_o, roughness[, F0]) grid the principled BSDF's energy
// compensation looks up (bsdf_energy.cuh).
// reference: src/Device/kernels/Baking/{GGXConductor,GGXFresnel,GlossyDielectric,GGXGlass,GGXThinGlass}DirectionalAlbedo.h,
//            launched by src/Renderer/Baker/GPUBakerKernel.cpp:62-95.
//
// Shape on B200: the reference cuts a bake into nb_kernel_launch launches of iterations_per_kernel samples (a display GPU must not run one
// kernel for seconds) and accumulates into the table in global memory with a read-modify-write per sample. Here a bake is one launch: a
// thread owns (texel, chunk of that texel's launches), re-seeds at every launch boundary exactly like the reference (wang_hash(texel + 1) *
// launch number), keeps the sum in a register, and the <= 16 chunks of a texel -- threads of one block -- are added in launch order through
// shared memory. The table is written once (texels * 4 bytes): the kernel is bound by fp32 / SFU issue (about 150-400 instructions per
// sample, no memory traffic), so the only sizing rule is "enough threads for 148 SMs", which `chunks` provides for the small tables.
// Arithmetic follows the reference statement by statement (-fmad=false, like every shading file here).
#include "bake.h"
#include "bsdf_microfacet.cuh"

#include "../../include/b200pt_bake.h"

namespace pt
{
namespace
{
constexpr int kBakeTexelsPerBlock = 32;         // with chunks > 1: one warp of texels per chunk row
constexpr int kBakeTexelsPerBlockUnsplit = 128; // chunks == 1: plain 128-thread blocks (32-thread blocks would cap an SM at 32 resident warps)
constexpr int kBakeMaxChunks = 16;   // b200pt_capi_bake.inl clamps b200pt_bake_options::chunks to this

struct BakeTexel
{
    float roughness;
    float relative_ior;
    float3 local_view_direction;
};

// The grid parametrisation shared by the five kernels (e.g. GGXFresnelDirectionalAlbedo.h:43-61): roughness linear in y, cos(theta_o)
// linear in x (to the power 2.5 for every table but the conductor's and the thin-glass one), F0 = (z / (size - 1))^4 -> relative ior.
template <int KIND>
PT_DEV BakeTexel bake_texel(const BakeLaunch& p, unsigned int x, unsigned int y, unsigned int z)
{
    BakeTexel t;
    float roughness = 1.0f / (p.size_y - 1) * y;
    t.roughness = maxf(roughness, 1.0e-4f);

    float cos_theta_o = 1.0f / (p.size_x - 1) * x;
    cos_theta_o = maxf(kGGXDotProductsClamp, cos_theta_o);
    if (KIND != B200PT_BAKE_GGX_CONDUCTOR && KIND != B200PT_BAKE_GGX_THIN_GLASS)
        cos_theta_o = powf(cos_theta_o, 2.5f);
    float sin_theta_o = sinf(acosf(cos_theta_o));

    t.relative_ior = 1.0f;
    if (KIND != B200PT_BAKE_GGX_CONDUCTOR)
    {
        float F0 = 1.0f / (p.size_z - 1) * z;
        F0 *= F0;
        F0 *= F0;
        float sqrt_F0 = sqrtf(clampf(0.0f, 0.99f, F0));
        t.relative_ior = (1.0f + sqrt_F0) / (1.0f - sqrt_F0);
    }
    // make_float3(cos(0) * sin_theta_o, sin(0) * sin_theta_o, cos_theta_o)
    t.local_view_direction = normalize(f3(1.0f * sin_theta_o, 0.0f * sin_theta_o, cos_theta_o));
    if (KIND == B200PT_BAKE_GGX_GLASS_EXITING)
        t.relative_ior = 1.0f / t.relative_ior;
    return t;
}

// torrance_sparrow_GGX_eval<0>(HIPRTRenderData(), ...).r with a single-channel Fresnel term
PT_DEV float ggx_eval_r(int masking, float roughness, float F, float3 wo, float3 wi, float3 h, float& pdf)
{
    HIPRTRenderData rd;
    rd.bsdfs_data.GGX_masking_shadowing = masking;
    return torrance_sparrow_GGX_eval_single(rd, roughness, 0.0f, RGB(F), wo, wi, h, pdf).r;
}

// GGXConductorDirectionalAlbedo.h:62-76 / GGXFresnelDirectionalAlbedo.h:67-82: one sample; false = the reference's `continue`
template <int KIND, int SAMPLER>
PT_DEV bool bake_sample_reflection(const BakeTexel& t, int masking, Rng& rng, float& albedo)
{
    float3 wo = t.local_view_direction;
    float3 wi = microfacet_GGX_sample_reflection<SAMPLER>(t.roughness, 0.0f, wo, rng);
    if (wi.z < 0)
        return false;
    float F = KIND == B200PT_BAKE_GGX_FRESNEL ? full_fresnel_dielectric(wi.z, t.relative_ior) : 1.0f;
    float eval_pdf;
    albedo = ggx_eval_r(masking, t.roughness, F, wo, wi, normalize(wo + wi), eval_pdf);
    albedo /= eval_pdf;
    albedo *= wi.z;
    return true;
}

// GlossyDielectricDirectionalAlbedo.h:72-120
template <int SAMPLER>
PT_DEV bool bake_sample_glossy(const BakeTexel& t, int masking, Rng& rng, float& albedo)
{
    float3 wo = t.local_view_direction;
    float rand_lobe = rng();
    float3 wi;
    if (rand_lobe < 0.5f)
    {
        wi = microfacet_GGX_sample_reflection<SAMPLER>(t.roughness, 0.0f, wo, rng);
        if (wi.z < 0)
            return false;
    }
    else
        wi = cosine_weighted_sample_z_up_frame(rng);

    float3 h = normalize(wo + wi);
    float total_pdf = 0.0f;

    float F = full_fresnel_dielectric(dot(h, wi), t.relative_ior);
    float eval_pdf_specular;
    float specular = ggx_eval_r(masking, t.roughness, F, wo, wi, h, eval_pdf_specular);
    total_pdf += eval_pdf_specular * 0.5f;
    float specular_layer_throughput = 1.0f;
    specular_layer_throughput *= 1.0f - full_fresnel_dielectric(wi.z, t.relative_ior);
    specular_layer_throughput *= 1.0f - full_fresnel_dielectric(wo.z, t.relative_ior);

    // lambertian_brdf_eval with a white base colour (Lambertian.h:14-23)
    float eval_pdf_diffuse = 0.0f;
    float diffuse = 0.0f;
    if (!(wi.z <= 0.0f))
    {
        eval_pdf_diffuse = wi.z * kInvPi;
        diffuse = 1.0f * kInvPi;
    }
    total_pdf += eval_pdf_diffuse * 0.5f;
    diffuse *= specular_layer_throughput;

    float final_albedo = specular + diffuse;
    final_albedo *= wi.z;
    final_albedo /= total_pdf;
    albedo = final_albedo;
    return true;
}

// GGX_glass_E_sample (GGXGlassDirectionalAlbedo.h:135-168) and thin_glass_sample (GGXThinGlassDirectionalAlbedo.h:32-86)
template <bool THIN, int SAMPLER>
PT_DEV float3 bake_glass_sample(float relative_ior, float roughness, float3 wo, Rng& rng)
{
    if (fabsf(relative_ior - 1.0f) < 1.0e-5f)
        relative_ior = 1.0f + 1.0e-5f;
    float alpha_x, alpha_y;
    get_alphas(roughness, 0.0f, alpha_x, alpha_y);
    float3 m = GGX_sample_microfacet<SAMPLER>(wo, alpha_x, alpha_y, rng);

    float F = full_fresnel_dielectric(dot(wo, m), relative_ior);
    if (THIN && roughness < 0.1f)
        F += square(1.0f - F) * F / (1.0f - square(F));
    float rand_1 = rng();

    if (rand_1 < F)
        return reflect_ray(wo, m);
    if (dot(m, wo) < 0.0f)
        m = -m;
    if (THIN)
    {
        // both interfaces in one layer: the view direction reflected, then mirrored through the surface plane
        float3 reflected = reflect_ray(wo, m);
        reflected.z *= -1.0f;
        return reflected;
    }
    // refract_ray fails only under total internal reflection, where F = 1 and this branch is not reached
    float3 refracted = f3(0.0f, 0.0f, 0.0f);
    refract_ray(wo, m, refracted, relative_ior);
    return refracted;
}

// GGX_glass_E_eval (GGXGlassDirectionalAlbedo.h:37-130) and thin_glass_eval (GGXThinGlassDirectionalAlbedo.h:88-199)
template <bool THIN>
PT_DEV float bake_glass_eval(int masking, float relative_ior, float roughness, float3 wo, float3 wi, float& pdf)
{
    pdf = 0.0f;
    float NoV = wo.z;
    float NoL = wi.z;
    if (fabsf(NoL) < 1.0e-8f)
        return 0.0f;
    bool reflecting = NoL * NoV > 0;
    if (fabsf(relative_ior - 1.0f) < 1.0e-5f)
        relative_ior = 1.0f + 1.0e-5f;

    float3 h;
    if (reflecting)
        h = wi + wo;
    else if (THIN)
        h = wi * f3(1.0f, 1.0f, -1.0f) + wo;
    else
        h = wi * relative_ior + wo;
    h = normalize(h);
    if (h.z < 0.0f)
        h = -h;

    float HoL = dot(wi, h);
    float HoV = dot(wo, h);
    if (HoL * NoL < 0.0f || HoV * NoV < 0.0f)
        return 0.0f;

    float F = full_fresnel_dielectric(HoV, relative_ior);
    if (THIN && roughness < 0.1f)
        F += square(1.0f - F) * F / (1.0f - square(F));
    if (reflecting)
    {
        float albedo = ggx_eval_r(masking, roughness, F, wo, wi, h, pdf);
        pdf *= F;
        return albedo;
    }
    float dot_prod = HoL + HoV / relative_ior;
    float dot_prod2 = dot_prod * dot_prod;
    float denom = dot_prod2 * NoL * NoV;

    float alpha_x, alpha_y;
    get_alphas(roughness, 0.0f, alpha_x, alpha_y);
    float D = GGX_anisotropic(alpha_x, alpha_y, h);
    float G1_V = G1_Smith(alpha_x, alpha_y, wo);
    float G1_L = G1_Smith(alpha_x, alpha_y, wi);
    float G2 = G1_V * G1_L;

    float dwm_dwi = fabsf(HoL) / dot_prod2;
    float D_pdf = G1_V / fabsf(NoV) * D * fabsf(HoV);
    pdf = dwm_dwi * D_pdf;
    pdf *= (1.0f - F);
    return D * (1.0f - F) * G2 * fabsf(HoL * HoV / denom);
}

// glass_directional_albedo_integration (GGXGlassDirectionalAlbedo.h:198-211) / GGXThinGlassDirectionalAlbedoBake (...ThinGlass...h:241-255)
template <bool THIN, int SAMPLER>
PT_DEV bool bake_sample_glass(const BakeTexel& t, int masking, Rng& rng, float& albedo)
{
    float roughness = THIN ? get_thin_walled_roughness(true, t.roughness, t.relative_ior) : t.roughness;
    float3 wi = bake_glass_sample<THIN, SAMPLER>(t.relative_ior, roughness, t.local_view_direction, rng);
    float eval_pdf = 0.0f;
    albedo = bake_glass_eval<THIN>(masking, t.relative_ior, roughness, t.local_view_direction, wi, eval_pdf);
    if (eval_pdf == 0.0f)
        return false;
    albedo /= eval_pdf;
    albedo *= fabsf(wi.z);
    return true;
}

template <int KIND, int SAMPLER>
__global__ void __launch_bounds__(kBakeTexelsPerBlock * kBakeMaxChunks) k_bake(BakeLaunch p, float* __restrict__ out)
{
    __shared__ float partial[kBakeMaxChunks][kBakeTexelsPerBlock];

    const long long texels = (long long)p.size_x * p.size_y * p.size_z;
    const long long texel = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int chunk = threadIdx.y;
    float sum = 0.0f;
    if (texel < texels)
    {
        const unsigned int pixel_index = (unsigned int)texel;
        const unsigned int x = pixel_index % (unsigned int)p.size_x;
        const unsigned int y = (pixel_index / (unsigned int)p.size_x) % (unsigned int)p.size_y;
        const unsigned int z = pixel_index / ((unsigned int)p.size_x * (unsigned int)p.size_y);
        const BakeTexel t = bake_texel<KIND>(p, x, y, z);
        const float nb_samples = (float)p.nb_samples;

        const int per_chunk = (p.nb_kernel_launch + p.chunks - 1) / p.chunks;
        const int first = chunk * per_chunk;
        const int last = min(first + per_chunk, p.nb_kernel_launch);
        const unsigned int hash = wang_hash(pixel_index + 1);
        for (int launch = first; launch < last; launch++)
        {
            // current_iteration = launch + 1 (GPUBakerKernel.cpp:88)
            Rng rng(hash * (unsigned int)(launch + 1));
            for (int sample = 0; sample < p.iterations_per_kernel; sample++)
            {
                float albedo;
                bool ok;
                if (KIND == B200PT_BAKE_GGX_CONDUCTOR || KIND == B200PT_BAKE_GGX_FRESNEL)
                    ok = bake_sample_reflection<KIND, SAMPLER>(t, p.masking, rng, albedo);
                else if (KIND == B200PT_BAKE_GLOSSY_DIELECTRIC)
                    ok = bake_sample_glossy<SAMPLER>(t, p.masking, rng, albedo);
                else
                    ok = bake_sample_glass<KIND == B200PT_BAKE_GGX_THIN_GLASS, SAMPLER>(t, p.masking, rng, albedo);
                if (ok)
                    sum += albedo / nb_samples;
            }
        }
    }
    if (p.chunks == 1)
    {
        if (texel < texels)
            out[texel] = sum;
        return;
    }
    partial[chunk][threadIdx.x] = sum;
    __syncthreads();
    if (chunk == 0 && texel < texels)
    {
        float total = partial[0][threadIdx.x];
        for (int c = 1; c < p.chunks; c++)
            total += partial[c][threadIdx.x];
        out[texel] = total;
    }
}

template <int KIND>
cudaError_t launch_bake_kind(const BakeLaunch& p, float* d_out, cudaStream_t stream)
{
    const long long texels = (long long)p.size_x * p.size_y * p.size_z;
    if (p.chunks < 1 || p.chunks > kBakeMaxChunks || texels <= 0 || texels > 0x7fffffffLL)
        return cudaErrorInvalidValue;
    const int texels_per_block = p.chunks == 1 ? kBakeTexelsPerBlockUnsplit : kBakeTexelsPerBlock;
    dim3 block(texels_per_block, p.chunks, 1);
    dim3 grid((unsigned int)((texels + texels_per_block - 1) / texels_per_block), 1, 1);
    if (p.ggx_sampler == 0)
        k_bake<KIND, 0><<<grid, block, 0, stream>>>(p, d_out);
    else
        k_bake<KIND, 1><<<grid, block, 0, stream>>>(p, d_out);
    return cudaGetLastError();
}
} // namespace

cudaError_t launch_bake(const BakeLaunch& p, float* d_out, cudaStream_t stream)
{
    switch (p.kind)
    {
    case B200PT_BAKE_GGX_CONDUCTOR: return launch_bake_kind<B200PT_BAKE_GGX_CONDUCTOR>(p, d_out, stream);
    case B200PT_BAKE_GGX_FRESNEL: return launch_bake_kind<B200PT_BAKE_GGX_FRESNEL>(p, d_out, stream);
    case B200PT_BAKE_GLOSSY_DIELECTRIC: return launch_bake_kind<B200PT_BAKE_GLOSSY_DIELECTRIC>(p, d_out, stream);
    case B200PT_BAKE_GGX_GLASS_ENTERING: return launch_bake_kind<B200PT_BAKE_GGX_GLASS_ENTERING>(p, d_out, stream);
    case B200PT_BAKE_GGX_GLASS_EXITING: return launch_bake_kind<B200PT_BAKE_GGX_GLASS_EXITING>(p, d_out, stream);
    case B200PT_BAKE_GGX_THIN_GLASS: return launch_bake_kind<B200PT_BAKE_GGX_THIN_GLASS>(p, d_out, stream);
    }
    return cudaErrorInvalidValue;
}
} // namespace pt
