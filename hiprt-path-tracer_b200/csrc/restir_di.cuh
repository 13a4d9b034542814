// ReSTIR DI device code (final shading of the output reservoir + the three passes). Filled in by restir_di_*.cuh.
#pragma once

#include "ris.cuh"

namespace pt
{
// sample_light_ReSTIR_DI (src/Device/includes/ReSTIR/DI/FinalShading.h:16-115) -- defined in restir_di_shading.cuh
template <int GGX_SAMPLER>
PT_DEV RGB sample_light_ReSTIR_DI(const HIPRTRenderData& rd, RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng, int2 pixel_coords, int2 res);
} // namespace pt

#include "restir_di_shading.cuh"
