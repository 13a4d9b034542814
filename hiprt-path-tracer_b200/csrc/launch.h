// Host-callable launchers of the two kernel builds (timing build: suffix _t, counting build: suffix _c).
#pragma once

#include "../../include/b200pt.h"
#include "bvh8.h"

#include <cuda_runtime.h>

namespace pt
{
struct VariantOptions { int lss, ess, later, ggx; };
struct Tile { int x0, y0, x1, y1; };
struct TraversalCounters { unsigned long long rays, nodes, triangles; };

#ifdef PT_COUNTERS
#define PT_SUFFIXED(name) name##_c
#define PT_VARIANT_NAME_OF(name, L, E, A, G) name##_##L##_##E##_##A##_##G##_c
#else
#define PT_SUFFIXED(name) name##_t
#define PT_VARIANT_NAME_OF(name, L, E, A, G) name##_##L##_##E##_##A##_##G##_t
#endif
#define PT_VARIANT_NAME_EXPAND(name, L, E, A, G) PT_VARIANT_NAME_OF(name, L, E, A, G)
#define PT_VARIANT_NAME(name) PT_VARIANT_NAME_EXPAND(name, PT_LSS, PT_ESS, PT_LATER, PT_GGX)

#define PT_DECLARE_LAUNCHERS(S)                                                                                                        \
    bool variant_available##S(const VariantOptions& opt);                                                                              \
    cudaError_t launch_CameraRays##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream);   \
    cudaError_t launch_FullPathTracer##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream); \
    cudaError_t launch_trace_closest##S(const Bvh8Device& bvh, const float* rays, const int* last_prim, long long n, int* out_prim, float* out_t, float* out_uv, TraversalCounters* counters, cudaStream_t stream); \
    cudaError_t launch_trace_shadow##S(const Bvh8Device& bvh, const float* rays, const float* t_max, const int* last_prim, long long n, unsigned char* out_occluded, TraversalCounters* counters, cudaStream_t stream);

PT_DECLARE_LAUNCHERS(_t)
PT_DECLARE_LAUNCHERS(_c)
} // namespace pt
