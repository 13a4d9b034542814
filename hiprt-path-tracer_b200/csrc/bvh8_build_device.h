// Device-side BVH8 construction (see bvh8_build_device.cuh / .cu): the opt-in GPU path of b200pt_bvh_build.
#pragma once

#include "bvh8.h"

#include <cuda_runtime.h>

namespace pt
{
struct Bvh8DeviceBuild
{
    Bvh8Node* nodes = nullptr;             // device arrays, owned by the caller (cudaFree)
    Bvh8Triangle* triangles = nullptr;
    uint32_t node_count = 0, triangle_count = 0;
    int max_depth = 0;
    unsigned long long child_slots_used = 0;
    float sah_cost = 0.0f;
    double device_seconds = 0.0;           // CUDA-event time of the build with the geometry already on the device
    int ploc_iterations = 0;               // clustering iterations (0: radix tree)
    float scene_min[3] = {0.0f, 0.0f, 0.0f}, scene_max[3] = {0.0f, 0.0f, 0.0f};
};

// vertices: 3 floats per vertex, indices: 3 ints per triangle, HOST arrays (as b200pt_bvh_build receives them); triangle_count >= 1
// algorithm: how the binary topology is found. PLOC (default of b200pt_bvh_build): mutual-nearest-neighbour clustering within `ploc_radius`
// positions along the Morton curve; LBVH: the Karras radix tree (faster to build, ~10 % higher SAH cost)
enum { BVH8_DEVICE_LBVH = 0, BVH8_DEVICE_PLOC = 1 };
cudaError_t build_bvh8_device(const float* vertices, int vertex_count, const int* indices, int triangle_count, cudaStream_t stream, Bvh8DeviceBuild* out, int algorithm = BVH8_DEVICE_PLOC,
                              int ploc_radius = 16);
} // namespace pt
