/*
 * b200pt_bake.h -- directional-albedo look-up-table bakers, part of libb200pt.so.
 *
 * The principled BSDF's energy compensation reads pre-integrated directional albedos (BRDFsData, src/HostDeviceCommon/BSDFsData.h) that the
 * reference regenerates with its GPUBaker (src/Renderer/Baker/GPUBaker.cpp:38-96): one Monte-Carlo integration kernel per table
 * (src/Device/kernels/Baking/*.h), launched nb_kernel_launch times from GPUBakerKernel::bake_internal (src/Renderer/Baker/GPUBakerKernel.cpp:62-95)
 * because a desktop GPU must not run one kernel for seconds. Here a table is ONE launch: a thread owns a texel and walks the reference's
 * launches itself -- same seeds (wang_hash(texel + 1) * launch number), same samples per launch, same order of the float sum -- so with
 * `chunks` = 1 a texel is the reference's arithmetic statement for statement; with `chunks` > 1 the launches of a texel are split over
 * that many threads of one block (small tables do not fill 148 SMs otherwise) and their partial sums are added in launch order, which
 * regroups the float sum and nothing else. Nothing is allocated; the output is overwritten (the reference's ZeroingKernel is not needed).
 *
 * The settings structs are the reference's (same members, same defaults); the entry points are named after GPUBaker's methods. The output
 * layout is the reference's bake buffer: texel (x, y, z) at x + y * size_cos_theta + z * size_cos_theta * size_roughness, x = cos(theta_o),
 * y = roughness, z = F0 (ior) axis; GPUBakerKernel.cpp:107-126 writes one .hdr slice per z, hiprt-path-tracer_b200/baker.py does the same.
 */
#ifndef B200PT_BAKE_H
#define B200PT_BAKE_H

#include "b200pt.h"

#ifdef __cplusplus
extern "C" {
#endif

/* src/Renderer/Baker/GPUBakerConstants.h:14-36 */
#define B200PT_BAKE_COMPUTE_ELEMENT_PER_BAKE_KERNEL_LAUNCH 100000000

/* src/Renderer/Baker/GGXConductorDirectionalAlbedoSettings.h:11-17 */
typedef struct GGXConductorDirectionalAlbedoSettings
{
    int texture_size_cos_theta B200PT_DEFAULT(128);
    int texture_size_roughness B200PT_DEFAULT(128);
    int integration_sample_count B200PT_DEFAULT(65536);
} GGXConductorDirectionalAlbedoSettings;

/* src/Renderer/Baker/GGXFresnelDirectionalAlbedoSettings.h:11-18 */
typedef struct GGXFresnelDirectionalAlbedoSettings
{
    int texture_size_cos_theta B200PT_DEFAULT(256);
    int texture_size_roughness B200PT_DEFAULT(256);
    int texture_size_ior B200PT_DEFAULT(256);
    int integration_sample_count B200PT_DEFAULT(65536);
} GGXFresnelDirectionalAlbedoSettings;

/* src/Renderer/Baker/GlossyDielectricDirectionalAlbedoSettings.h:11-18 */
typedef struct GlossyDielectricDirectionalAlbedoSettings
{
    int texture_size_cos_theta_o B200PT_DEFAULT(128);
    int texture_size_roughness B200PT_DEFAULT(64);
    int texture_size_ior B200PT_DEFAULT(128);
    int integration_sample_count B200PT_DEFAULT(65536);
} GlossyDielectricDirectionalAlbedoSettings;

/* src/Renderer/Baker/GGXGlassDirectionalAlbedoSettings.h:11-18 */
typedef struct GGXGlassDirectionalAlbedoSettings
{
    int texture_size_cos_theta_o B200PT_DEFAULT(256);
    int texture_size_roughness B200PT_DEFAULT(16);
    int texture_size_ior B200PT_DEFAULT(128);
    int integration_sample_count B200PT_DEFAULT(65536);
} GGXGlassDirectionalAlbedoSettings;

/* src/Renderer/Baker/GGXThinGlassDirectionalAlbedoSettings.h:11-18 */
typedef struct GGXThinGlassDirectionalAlbedoSettings
{
    int texture_size_cos_theta_o B200PT_DEFAULT(32);
    int texture_size_roughness B200PT_DEFAULT(32);
    int texture_size_ior B200PT_DEFAULT(96);
    int integration_sample_count B200PT_DEFAULT(65536);
} GGXThinGlassDirectionalAlbedoSettings;

/* What the reference fixes at kernel-compile time or leaves at its default, and how one texel's launches are spread over threads. */
typedef struct b200pt_bake_options
{
    int ggx_sample_function B200PT_DEFAULT(0);   /* PrincipledBSDFAnisotropicGGXSampleFunction (KernelOptions.h:155): 0 = GGX_VNDF_SAMPLING, 1 = GGX_VNDF_SPHERICAL_CAPS */
    int GGX_masking_shadowing B200PT_DEFAULT(0); /* BRDFsData::GGX_masking_shadowing of the default-constructed HIPRTRenderData the kernels evaluate with (BSDFsData.h:62) */
    int chunks B200PT_DEFAULT(0);                /* threads per texel: 1 = the reference's summation order exactly, 2..16 = a texel's launches split in order over that many threads, 0 = chosen from the table size */
} b200pt_bake_options;

/* The launch plan GPUBakerKernel::bake_internal derives from a table's resolution (GPUBakerKernel.cpp:82-83) and the sample count the KERNEL
 * divides by (each kernel recomputes it; GGXFresnelDirectionalAlbedo.h:63-65 leaves the ior axis out, so its tables sum to slightly less
 * than the mean -- restated as it is). samples_run = nb_kernel_launch * iterations_per_kernel is what a texel really integrates. */
typedef struct b200pt_bake_plan
{
    int iterations_per_kernel;
    int nb_kernel_launch;
    int nb_samples;          /* the divisor */
    int chunks;              /* as resolved from b200pt_bake_options::chunks */
    long long texels;
} b200pt_bake_plan;

typedef enum b200pt_bake_kind
{
    B200PT_BAKE_GGX_CONDUCTOR = 0,
    B200PT_BAKE_GGX_FRESNEL = 1,
    B200PT_BAKE_GLOSSY_DIELECTRIC = 2,
    B200PT_BAKE_GGX_GLASS_ENTERING = 3,
    B200PT_BAKE_GGX_GLASS_EXITING = 4,
    B200PT_BAKE_GGX_THIN_GLASS = 5
} b200pt_bake_kind;

/* Host only, no device needed. options may be NULL (defaults). */
B200PT_API int b200pt_bake_plan_for(int kind, int size_cos_theta, int size_roughness, int size_ior, int integration_sample_count, const b200pt_bake_options* options,
                                    b200pt_bake_plan* out_plan);

/* GPUBaker::bake_ggx_conductor_directional_albedo (GPUBaker.cpp:38-43) + kernel GGXConductorDirectionalAlbedoBake (Baking/GGXConductorDirectionalAlbedo.h:32-77).
 * d_out: texture_size_cos_theta * texture_size_roughness floats. Asynchronous on `stream`. */
B200PT_API int b200pt_bake_ggx_conductor_directional_albedo(b200pt_ctx* ctx, const GGXConductorDirectionalAlbedoSettings* bake_settings, const b200pt_bake_options* options,
                                                            float* d_out, b200pt_stream stream);
/* GPUBaker::bake_ggx_fresnel_directional_albedo (GPUBaker.cpp:50-55) + GGXFresnelDirectionalAlbedoBake (Baking/GGXFresnelDirectionalAlbedo.h:25-84). */
B200PT_API int b200pt_bake_ggx_fresnel_directional_albedo(b200pt_ctx* ctx, const GGXFresnelDirectionalAlbedoSettings* bake_settings, const b200pt_bake_options* options, float* d_out,
                                                          b200pt_stream stream);
/* GPUBaker::bake_glossy_dielectric_directional_albedo (GPUBaker.cpp:62-67) + GlossyDielectricDirectionalAlbedoBake (Baking/GlossyDielectricDirectionalAlbedo.h:29-129). */
B200PT_API int b200pt_bake_glossy_dielectric_directional_albedo(b200pt_ctx* ctx, const GlossyDielectricDirectionalAlbedoSettings* bake_settings, const b200pt_bake_options* options,
                                                                float* d_out, b200pt_stream stream);
/* GPUBaker::bake_ggx_glass_directional_albedo (GPUBaker.cpp:74-83): two tables, entering the medium and leaving it (relative ior inverted),
 * kernels GGXGlassDirectionalAlbedoBakeEntering / Exiting (Baking/GGXGlassDirectionalAlbedo.h:36-244). Either output may be NULL. */
B200PT_API int b200pt_bake_ggx_glass_directional_albedo(b200pt_ctx* ctx, const GGXGlassDirectionalAlbedoSettings* bake_settings, const b200pt_bake_options* options, float* d_out_entering,
                                                        float* d_out_exiting, b200pt_stream stream);
/* GPUBaker::bake_ggx_thin_glass_directional_albedo (GPUBaker.cpp:90-95) + GGXThinGlassDirectionalAlbedoBake (Baking/GGXThinGlassDirectionalAlbedo.h:32-256). */
B200PT_API int b200pt_bake_ggx_thin_glass_directional_albedo(b200pt_ctx* ctx, const GGXThinGlassDirectionalAlbedoSettings* bake_settings, const b200pt_bake_options* options,
                                                             float* d_out, b200pt_stream stream);

#ifdef __cplusplus
}
#endif
#endif
