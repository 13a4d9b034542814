// Launch interface of the directional-albedo bakers (b200pt_bake.cu), used by the C ABI in b200pt_capi_bake.inl.
#pragma once

#include <cuda_runtime.h>

namespace pt
{
struct BakeLaunch
{
    int kind;                    // b200pt_bake_kind
    int size_x, size_y, size_z;  // cos(theta_o), roughness, ior (1 for the conductor table)
    int iterations_per_kernel;   // samples per reference launch
    int nb_kernel_launch;        // reference launches per texel
    int nb_samples;              // the divisor the reference kernel applies
    int chunks;                  // threads per texel (1..32)
    int ggx_sampler;             // 0 = VNDF, 1 = spherical caps
    int masking;                 // GGXMaskingShadowingFlavor
};

cudaError_t launch_bake(const BakeLaunch& plan, float* d_out, cudaStream_t stream);
} // namespace pt
