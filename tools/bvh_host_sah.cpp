// Host BVH8 builder on a scene dumped by tools/dump_scene.py (int32 vertex count, int32 triangle count, float32 xyz per vertex, int32 indices):
// prints SAH cost, node count, fill and depth for the builder's current environment settings (B200PT_BVH_SWEEP_BELOW, B200PT_BVH_PRIMITIVE_COST).
//   g++ -std=c++17 -O2 -fopenmp -I /usr/local/cuda/include -I hiprt-path-tracer_b200/csrc -I include tools/bvh_host_sah.cpp hiprt-path-tracer_b200/csrc/bvh8_build.cpp -o /tmp/bvh_host_sah
#include "bvh8_build.h"
#include <chrono>
#include <cstdio>
#include <vector>
int main(int argc, char** argv)
{
    if (argc < 2) return 2;
    FILE* f = fopen(argv[1], "rb");
    if (!f) return 2;
    int nv = 0, nt = 0;
    if (fread(&nv, 4, 1, f) != 1 || fread(&nt, 4, 1, f) != 1) return 2;
    std::vector<float> v(3 * (size_t)nv);
    std::vector<int> idx(3 * (size_t)nt);
    if (fread(v.data(), 4, v.size(), f) != v.size() || fread(idx.data(), 4, idx.size(), f) != idx.size()) return 2;
    fclose(f);
    auto t0 = std::chrono::steady_clock::now();
    pt::Bvh8Host b = pt::build_bvh8(v.data(), nv, idx.data(), nt);
    double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("triangles %d nodes %zu sah %.3f fill %.3f depth %d build %.2f s\n", nt, b.nodes.size(), b.sah_cost, b.child_slots_used / (8.0 * b.nodes.size()), b.max_depth, s);
    return 0;
}
