// Compressed 8-wide BVH layout shared by the host builder and the device traversal.
//
// Replaces the HIPRT BVH4 of the reference (128-byte BoxNode, 64-byte TriangleNode pairs:
// thirdparties/HIPRT-Fork/hiprt/impl/BvhNode.h:105-159) that src/HIPRT-Orochi/HIPRTScene.h:59-86 builds.
// Layout after Ylitie, Karras, Laine, "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs"
// (HPG 2017): 80-byte nodes = five 16-byte vector loads, child boxes quantised to 8 bits on a per-node
// power-of-two grid (boxes only ever GROW, so traversal stays conservative and the closest hit is decided by the
// triangle routine alone), 48-byte triangles = three 16-byte loads holding the UN-transformed vertices so that
// the reference's Moeller-Trumbore arithmetic (src/Renderer/Triangle.h:20-63) can be reproduced bit for bit.
#pragma once

#include <stdint.h>

namespace pt
{
struct alignas(16) Bvh8Node
{
    float px, py, pz;        // origin of the quantisation grid (= node box min)
    uint8_t ex, ey, ez;      // biased float exponents of the grid cell size per axis
    uint8_t imask;           // bit s set: slot s holds an internal child
    uint32_t child_base;     // index of the first internal child (children are contiguous, in slot order)
    uint32_t tri_base;       // index of the first triangle of this node's leaves (<= 24 per node)
    uint8_t meta[8];         // empty: 0; internal: 0b001_11000 + slot; leaf: (unary count << 5) | first triangle offset
    uint8_t qlo_x[8], qlo_y[8], qlo_z[8];
    uint8_t qhi_x[8], qhi_y[8], qhi_z[8];
};
static_assert(sizeof(Bvh8Node) == 80, "BVH8 node must be 80 bytes");

struct alignas(16) Bvh8Triangle
{
    float v0x, v0y, v0z; int prim_id;     // index of the triangle in the scene's index buffer
    float v1x, v1y, v1z; int pad1;        // builders write 0; b200pt_bvh_attach_shading_data: the triangle's material index
    float v2x, v2y, v2z; int pad2;        //                                                    has_vertex_normals of vertex A (0 / 1)
};
static_assert(sizeof(Bvh8Triangle) == 48, "BVH8 triangle must be 48 bytes");

// Device-visible descriptor; its device address is what HIPRTRenderData::geom carries.
struct Bvh8Device
{
    const Bvh8Node* nodes;
    const Bvh8Triangle* triangles;
    uint32_t node_count;
    uint32_t triangle_count;
    float scene_min[3];
    float scene_max[3];
    // Per-triangle-SLOT shading records (b200pt_bvh_attach_shading_data; null until attached): 64 bytes = four 16-byte loads holding the three
    // vertex normals (9 floats), the three texcoords (6 floats) and one spare word -- what SURFACE otherwise gathers through the index buffer
    // from three scattered vertices (indices, 3 normals, 3 texcoords, a flag byte, the material index: ~10 sectors per hit against 2).
    const void* shading;
};

constexpr int kBvh8MaxLeafTriangles = 3;
constexpr int kTraversalStackSize = 48;   // uint2 entries per ray (spills to local memory past the shared part)
} // namespace pt
