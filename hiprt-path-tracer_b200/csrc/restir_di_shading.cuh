// PLACEHOLDER until the ReSTIR DI passes land: final shading of restir_output_reservoirs.
#pragma once

namespace pt
{
template <int GGX_SAMPLER>
PT_DEV RGB sample_light_ReSTIR_DI(const HIPRTRenderData& rd, RayPayload& payload, const HitInfo& hit, float3 view_direction, Rng& rng, int2 pixel_coords, int2 res)
{
    return RGB(0.0f);
}
} // namespace pt
