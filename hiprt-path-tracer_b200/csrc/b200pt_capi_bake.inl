// C ABI of the directional-albedo LUT bakers (include/b200pt_bake.h); included by b200pt_capi.cu. Kernels: b200pt_bake.cu.
#include "bake.h"

#include "../../include/b200pt_bake.h"

#include <algorithm>
#include <cmath>

namespace
{
constexpr int kBakeMaxChunksHost = 16;   // = kBakeMaxChunks of b200pt_bake.cu

// GPUBakerKernel::bake_internal (src/Renderer/Baker/GPUBakerKernel.cpp:82-83) for the host side of the plan, and the nb_samples each kernel
// recomputes for itself: conductor (GGXConductorDirectionalAlbedo.h:58-60), Fresnel (GGXFresnelDirectionalAlbedo.h:63-65: cos x roughness
// only, the ior axis is missing there), glossy (GlossyDielectricDirectionalAlbedo.h:68-70), glass (GGXGlassDirectionalAlbedo.h:194-196),
// thin glass (GGXThinGlassDirectionalAlbedo.h:237-238: from the launch's own iteration count). float arithmetic as written there.
int bake_plan(int kind, int sx, int sy, int sz, int sample_count, const b200pt_bake_options* options, b200pt_bake_plan* out)
{
    if (kind < B200PT_BAKE_GGX_CONDUCTOR || kind > B200PT_BAKE_GGX_THIN_GLASS)
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "b200pt_bake: unknown table kind");
    if (kind == B200PT_BAKE_GGX_CONDUCTOR)
        sz = 1;
    if (sx < 2 || sy < 2 || (kind != B200PT_BAKE_GGX_CONDUCTOR && sz < 2) || sample_count < 1)
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "b200pt_bake: every table axis needs at least 2 texels and the sample count must be positive");
    long long texels = (long long)sx * sy * sz;
    if (texels > 0x7fffffffLL)
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "b200pt_bake: more than 2^31 - 1 texels");
    const float elements = (float)B200PT_BAKE_COMPUTE_ELEMENT_PER_BAKE_KERNEL_LAUNCH;

    int iterations_per_kernel = (int)floorf(fmaxf(1.0f, elements / (sx * sy * sz)));
    int nb_kernel_launch = (int)ceilf(sample_count / (float)iterations_per_kernel);

    int nb_samples;
    if (kind == B200PT_BAKE_GGX_THIN_GLASS)
        nb_samples = (int)ceilf(sample_count / (float)iterations_per_kernel) * iterations_per_kernel;
    else
    {
        int device_elements = (kind == B200PT_BAKE_GGX_CONDUCTOR || kind == B200PT_BAKE_GGX_FRESNEL) ? sx * sy : sx * sy * sz;
        int device_iterations = (int)floorf(fmaxf(1.0f, elements / device_elements));
        int device_launches = (int)ceilf(sample_count / (float)device_iterations);
        nb_samples = device_launches * device_iterations;
    }

    int chunks = options ? options->chunks : 0;
    if (chunks < 0 || chunks > kBakeMaxChunksHost)
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "b200pt_bake_options::chunks must be 0 (automatic) or 1..16");
    if (chunks == 0)
    {
        // enough threads for 148 SMs x 1024: small tables split a texel's launches, large ones keep the reference's summation order
        const long long fill = 148LL * 1024;
        chunks = 1;
        while (chunks < kBakeMaxChunksHost && texels * chunks < fill && chunks * 2 <= nb_kernel_launch)
            chunks *= 2;
        if (chunks > 1)
        {
            // ... and among the splits up to two fills of the GPU, the one whose longest chunk is shortest (11 launches over 8 threads leave
            // two of them idle and the others with 2 launches each; over 11 threads every thread runs one)
            const int most = (int)std::min<long long>(kBakeMaxChunksHost, std::max<long long>(chunks, 2 * fill / texels));
            int best = chunks, best_longest = (nb_kernel_launch + chunks - 1) / chunks;
            for (int c = chunks + 1; c <= most && c <= nb_kernel_launch; c++)
            {
                const int longest = (nb_kernel_launch + c - 1) / c;
                if (longest < best_longest)
                {
                    best = c;
                    best_longest = longest;
                }
            }
            chunks = best;
        }
    }
    if (chunks > nb_kernel_launch)
        chunks = nb_kernel_launch;
    out->iterations_per_kernel = iterations_per_kernel;
    out->nb_kernel_launch = nb_kernel_launch;
    out->nb_samples = nb_samples;
    out->chunks = chunks;
    out->texels = texels;
    return B200PT_SUCCESS;
}

int bake_launch(b200pt_ctx* ctx, int kind, int sx, int sy, int sz, int sample_count, const b200pt_bake_options* options, float* d_out, b200pt_stream stream, const char* what)
{
    REQUIRE(ctx != nullptr && d_out != nullptr);
    b200pt_bake_plan plan;
    int rc = bake_plan(kind, sx, sy, sz, sample_count, options, &plan);
    if (rc != B200PT_SUCCESS)
        return rc;
    int sampler = options ? options->ggx_sample_function : 0;
    int masking = options ? options->GGX_masking_shadowing : 0;
    if (sampler < 0 || sampler > 1 || masking < 0 || masking > 1)
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "b200pt_bake_options: ggx_sample_function and GGX_masking_shadowing are 0 or 1");
    CUDA_TRY(cudaSetDevice(ctx->device));
    pt::BakeLaunch launch{kind, sx, sy, kind == B200PT_BAKE_GGX_CONDUCTOR ? 1 : sz, plan.iterations_per_kernel, plan.nb_kernel_launch, plan.nb_samples, plan.chunks, sampler, masking};
    cudaError_t e = pt::launch_bake(launch, d_out, stream ? reinterpret_cast<cudaStream_t>(stream) : ctx->main_stream);
    if (e != cudaSuccess)
        return cuda_fail(e, what);
    return B200PT_SUCCESS;
}
} // namespace

extern "C" int b200pt_bake_plan_for(int kind, int size_cos_theta, int size_roughness, int size_ior, int integration_sample_count, const b200pt_bake_options* options,
                                    b200pt_bake_plan* out_plan)
{
    REQUIRE(out_plan != nullptr);
    return bake_plan(kind, size_cos_theta, size_roughness, size_ior, integration_sample_count, options, out_plan);
}

extern "C" int b200pt_bake_ggx_conductor_directional_albedo(b200pt_ctx* ctx, const GGXConductorDirectionalAlbedoSettings* s, const b200pt_bake_options* options, float* d_out,
                                                            b200pt_stream stream)
{
    REQUIRE(s != nullptr);
    return bake_launch(ctx, B200PT_BAKE_GGX_CONDUCTOR, s->texture_size_cos_theta, s->texture_size_roughness, 1, s->integration_sample_count, options, d_out, stream,
                       "GGXConductorDirectionalAlbedoBake");
}

extern "C" int b200pt_bake_ggx_fresnel_directional_albedo(b200pt_ctx* ctx, const GGXFresnelDirectionalAlbedoSettings* s, const b200pt_bake_options* options, float* d_out,
                                                          b200pt_stream stream)
{
    REQUIRE(s != nullptr);
    return bake_launch(ctx, B200PT_BAKE_GGX_FRESNEL, s->texture_size_cos_theta, s->texture_size_roughness, s->texture_size_ior, s->integration_sample_count, options, d_out, stream,
                       "GGXFresnelDirectionalAlbedoBake");
}

extern "C" int b200pt_bake_glossy_dielectric_directional_albedo(b200pt_ctx* ctx, const GlossyDielectricDirectionalAlbedoSettings* s, const b200pt_bake_options* options,
                                                                float* d_out, b200pt_stream stream)
{
    REQUIRE(s != nullptr);
    return bake_launch(ctx, B200PT_BAKE_GLOSSY_DIELECTRIC, s->texture_size_cos_theta_o, s->texture_size_roughness, s->texture_size_ior, s->integration_sample_count, options, d_out,
                       stream, "GlossyDielectricDirectionalAlbedoBake");
}

extern "C" int b200pt_bake_ggx_glass_directional_albedo(b200pt_ctx* ctx, const GGXGlassDirectionalAlbedoSettings* s, const b200pt_bake_options* options, float* d_out_entering,
                                                        float* d_out_exiting, b200pt_stream stream)
{
    REQUIRE(s != nullptr && (d_out_entering != nullptr || d_out_exiting != nullptr));
    if (d_out_entering != nullptr)
    {
        int rc = bake_launch(ctx, B200PT_BAKE_GGX_GLASS_ENTERING, s->texture_size_cos_theta_o, s->texture_size_roughness, s->texture_size_ior, s->integration_sample_count, options,
                             d_out_entering, stream, "GGXGlassDirectionalAlbedoBakeEntering");
        if (rc != B200PT_SUCCESS)
            return rc;
    }
    if (d_out_exiting != nullptr)
        return bake_launch(ctx, B200PT_BAKE_GGX_GLASS_EXITING, s->texture_size_cos_theta_o, s->texture_size_roughness, s->texture_size_ior, s->integration_sample_count, options,
                           d_out_exiting, stream, "GGXGlassDirectionalAlbedoBakeExiting");
    return B200PT_SUCCESS;
}

extern "C" int b200pt_bake_ggx_thin_glass_directional_albedo(b200pt_ctx* ctx, const GGXThinGlassDirectionalAlbedoSettings* s, const b200pt_bake_options* options, float* d_out,
                                                             b200pt_stream stream)
{
    REQUIRE(s != nullptr);
    return bake_launch(ctx, B200PT_BAKE_GGX_THIN_GLASS, s->texture_size_cos_theta_o, s->texture_size_roughness, s->texture_size_ior, s->integration_sample_count, options, d_out,
                       stream, "GGXThinGlassDirectionalAlbedoBake");
}
