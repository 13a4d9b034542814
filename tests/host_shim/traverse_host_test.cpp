// TEST INFRASTRUCTURE: compiles csrc/traverse.cuh + the BVH8 builder for the HOST with a shim of the few CUDA intrinsics
// they use, and checks the traversal against a brute-force loop over all triangles with the same triangle routine.
// This validates the builder (conservative quantisation, slot/bit encoding) where no GPU is available; it is not a
// product path (the product only ever runs the CUDA build).
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#undef __device__
#undef __forceinline__
#undef __noinline__
#define __device__
#define __forceinline__ inline
#define __noinline__
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline float __uint_as_float(unsigned u) { float f; memcpy(&f, &u, 4); return f; }
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
#include "traverse.cuh"
#include "bvh8_build.h"
using namespace pt;

static float frand() { return rand() / (float)RAND_MAX; }

int main(int argc, char** argv)
{
    int mode = argc > 1 ? atoi(argv[1]) : 0;   // 0 random soup, 1 axis-aligned grid of quads (flat boxes on the grid), 2 empty, 3 one triangle
    int N = mode == 2 ? 0 : mode == 3 ? 1 : 3000;
    std::vector<float> v;
    std::vector<int> idx;
    srand(1234 + mode);
    for (int i = 0; i < N; i++)
    {
        if (mode == 1)
        {
            // axis-aligned right triangles lying exactly in planes x|y|z = integer
            int axis = i % 3;
            float c[3] = {(float)(rand() % 8), (float)(rand() % 8), (float)(rand() % 8)};
            float p[3][3];
            for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) p[k][a] = c[a];
            p[1][(axis + 1) % 3] += 1.0f;
            p[2][(axis + 2) % 3] += 1.0f;
            for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) v.push_back(p[k][a]);
        }
        else
        {
            float cx = frand() * 10, cy = frand() * 10, cz = frand() * 10, s = mode == 3 ? 5.0f : 0.5f;
            for (int k = 0; k < 3; k++) { v.push_back(cx + frand() * s); v.push_back(cy + frand() * s); v.push_back(cz + frand() * s); }
        }
        idx.push_back(3 * i); idx.push_back(3 * i + 1); idx.push_back(3 * i + 2);
    }
    Bvh8Host h = build_bvh8(v.data(), (int)v.size() / 3, idx.data(), N);
    Bvh8Device d;
    d.nodes = h.nodes.data();
    d.triangles = h.triangles.data();
    if ((int)h.triangles.size() != N) { printf("FAIL triangle count %zu\n", h.triangles.size()); return 1; }
    int bad = 0, hits = 0, shadow_bad = 0;
    for (int r = 0; r < 20000; r++)
    {
        float3 o = make_float3(frand() * 10 - 1, frand() * 10 - 1, frand() * 10 - 1);
        float3 dir = make_float3(frand() - 0.5f, frand() - 0.5f, frand() - 0.5f);
        if (r % 7 == 0) dir.x = 0.0f;          // axis-parallel components
        if (r % 11 == 0) dir.y = 0.0f;
        if (r % 13 == 0) dir.z = -0.0f;        // negative zero: its reciprocal is negative although it does not compare below zero
        if (r % 17 == 0) dir.x = -0.0f;
        float l = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
        if (l == 0.0f) continue;
        dir.x /= l; dir.y /= l; dir.z /= l;
        int last = N > 0 ? rand() % N : -1;
        Hit hit;
        bvh8_trace<false>(d, o, dir, INFINITY, last, hit, nullptr);
        int best = -1;
        float bt = INFINITY;
        for (int t = 0; t < N; t++)
        {
            if (t == last) continue;
            float tt, u, vv;
            float3 a = make_float3(v[9 * t], v[9 * t + 1], v[9 * t + 2]), b = make_float3(v[9 * t + 3], v[9 * t + 4], v[9 * t + 5]), c = make_float3(v[9 * t + 6], v[9 * t + 7], v[9 * t + 8]);
            if (intersect_triangle_reference(o, dir, a, b, c, tt, u, vv) && tt < bt) { bt = tt; best = t; }
        }
        if (best >= 0) hits++;
        if (best != hit.prim && !(best >= 0 && hit.prim >= 0 && hit.t == bt)) bad++;   // exact distance ties excluded
        float tmax = frand() * 12.0f;
        Hit sh;
        bool occluded = bvh8_trace<true>(d, o, dir, tmax - 1.0e-4f, last, sh, nullptr);
        if (occluded != (best >= 0 && bt < tmax - 1.0e-4f)) shadow_bad++;
    }
    printf("mode %d nodes %zu tris %zu hits %d closest_mismatches %d shadow_mismatches %d\n", mode, h.nodes.size(), h.triangles.size(), hits, bad, shadow_bad);
    return (bad || shadow_bad) ? 1 : 0;
}
