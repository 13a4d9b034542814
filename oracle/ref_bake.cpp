// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
//
// The reference's directional-albedo bake kernels (src/Device/kernels/Baking/*.h), compiled *unmodified* for the CPU from where they lie
// under /root/reference (their `#ifndef __KERNELCC__` signatures take the texel coordinates as arguments), driven the way
// GPUBakerKernel::bake_internal drives them on the GPU (src/Renderer/Baker/GPUBakerKernel.cpp:62-95): buffer zeroed, iterations_per_kernel
// and nb_kernel_launch from the table's resolution, launch i run with current_iteration = i + 1. Texels are independent, so the launch loop
// is inside the texel loop here (one OpenMP iteration per texel); each texel's sequence of additions is the reference's.
// Built by oracle/build_ref.py into oracle/_ref/libhippt_ref_bake_<variant>.so (variant = which GGX sampling routine KernelOptions.h selects).

#include "Device/kernels/Baking/GGXConductorDirectionalAlbedo.h"
#include "Device/kernels/Baking/GGXFresnelDirectionalAlbedo.h"
#include "Device/kernels/Baking/GlossyDielectricDirectionalAlbedo.h"
#include "Device/kernels/Baking/GGXGlassDirectionalAlbedo.h"
#include "Device/kernels/Baking/GGXThinGlassDirectionalAlbedo.h"

#include <cmath>
#include <cstring>
#include <omp.h>

// kinds: 0 conductor, 1 fresnel, 2 glossy dielectric, 3 glass entering, 4 glass exiting, 5 thin glass (= b200pt_bake_kind)
// z_begin / z_end: only the ior slices [z_begin, z_end) are integrated (the others stay zero) -- a full-size table takes the CPU an hour, two
// of its slices a minute; the launch plan is the full table's either way.
extern "C" int ref_bake(int kind, int sx, int sy, int sz, int sample_count, int z_begin, int z_end, float* out, int* out_iterations_per_kernel, int* out_nb_kernel_launch)
{
    if (kind < 0 || kind > 5 || sx < 2 || sy < 2 || (kind != 0 && sz < 2) || sample_count < 1)
        return -1;
    if (kind == 0)
        sz = 1;
    int3 bake_resolution = make_int3(sx, sy, sz);
    if (out != nullptr)
        std::memset(out, 0, sizeof(float) * (size_t)sx * sy * sz);   // ZeroingKernel

    // GPUBakerKernel.cpp:82-83
    int iterations_per_kernel = floor(hippt::max(1.0f, (float)GPUBakerConstants::COMPUTE_ELEMENT_PER_BAKE_KERNEL_LAUNCH / (bake_resolution.x * bake_resolution.y * bake_resolution.z)));
    int nb_kernel_launch = ceil(sample_count / (float)iterations_per_kernel);
    if (out_iterations_per_kernel)
        *out_iterations_per_kernel = iterations_per_kernel;
    if (out_nb_kernel_launch)
        *out_nb_kernel_launch = nb_kernel_launch;

    GGXConductorDirectionalAlbedoSettings conductor;
    conductor.texture_size_cos_theta = sx; conductor.texture_size_roughness = sy; conductor.integration_sample_count = sample_count;
    GGXFresnelDirectionalAlbedoSettings fresnel;
    fresnel.texture_size_cos_theta = sx; fresnel.texture_size_roughness = sy; fresnel.texture_size_ior = sz; fresnel.integration_sample_count = sample_count;
    GlossyDielectricDirectionalAlbedoSettings glossy;
    glossy.texture_size_cos_theta_o = sx; glossy.texture_size_roughness = sy; glossy.texture_size_ior = sz; glossy.integration_sample_count = sample_count;
    GGXGlassDirectionalAlbedoSettings glass;
    glass.texture_size_cos_theta_o = sx; glass.texture_size_roughness = sy; glass.texture_size_ior = sz; glass.integration_sample_count = sample_count;
    GGXThinGlassDirectionalAlbedoSettings thin;
    thin.texture_size_cos_theta_o = sx; thin.texture_size_roughness = sy; thin.texture_size_ior = sz; thin.integration_sample_count = sample_count;

    if (out == nullptr)
        return 0;   // plan only
    z_begin = z_begin < 0 ? 0 : z_begin;
    z_end = z_end > sz ? sz : z_end;
    const long long first = (long long)sx * sy * z_begin, texels = (long long)sx * sy * z_end;
#pragma omp parallel for schedule(dynamic, 4)
    for (long long texel = first; texel < texels; texel++)
    {
        int x = (int)(texel % sx), y = (int)((texel / sx) % sy), z = (int)(texel / ((long long)sx * sy));
        for (int i = 0; i < nb_kernel_launch; i++)
        {
            int current_iteration = i + 1;
            switch (kind)
            {
            case 0: GGXConductorDirectionalAlbedoBake(iterations_per_kernel, current_iteration, conductor, out, x, y); break;
            case 1: GGXFresnelDirectionalAlbedoBake(iterations_per_kernel, current_iteration, fresnel, out, x, y, z); break;
            case 2: GlossyDielectricDirectionalAlbedoBake(iterations_per_kernel, current_iteration, glossy, out, x, y, z); break;
            case 3: GGXGlassDirectionalAlbedoBakeEntering(iterations_per_kernel, current_iteration, glass, out, x, y, z); break;
            case 4: GGXGlassDirectionalAlbedoBakeExiting(iterations_per_kernel, current_iteration, glass, out, x, y, z); break;
            case 5: GGXThinGlassDirectionalAlbedoBake(iterations_per_kernel, current_iteration, thin, out, x, y, z); break;
            }
        }
    }
    return 0;
}
