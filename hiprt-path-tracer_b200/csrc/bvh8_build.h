// Host-side BVH8 construction (see bvh8_build.cpp).
#pragma once

#include "bvh8.h"

#include <vector>

namespace pt
{
struct Bvh8Host
{
    std::vector<Bvh8Node> nodes;
    std::vector<Bvh8Triangle> triangles;
    float scene_min[3];
    float scene_max[3];
    float sah_cost = 0.0f;
    int max_depth = 0;                 // levels of wide nodes (a root-only tree has depth 1): bounds the traversal stack
    unsigned long long child_slots_used = 0;   // occupied child slots over all nodes (fill = child_slots_used / (8 * nodes))
};

// vertices: 3 floats per vertex; indices: 3 ints per triangle (the reference's Scene::vertices_positions /
// Scene::triangle_indices, src/Scene/SceneParser.h:80-131)
Bvh8Host build_bvh8(const float* vertices, int vertex_count, const int* indices, int triangle_count);
} // namespace pt
