// TEST INFRASTRUCTURE: compiles csrc/trace_queue.cuh (both loop flavours of the ray-queue traversal kernel) + the BVH8 builder for
// the HOST with a shim of the CUDA intrinsics they use -- a "warp" of ONE lane -- and checks every ray against a brute-force loop
// over all triangles with the same triangle routine. This pins the queue / stack / pipelining logic where no GPU is available
// (the warp-level scheduling itself only runs on the GPU: tests/test_gpu_parity.py); it is not a product path.
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
static struct { unsigned x, y, z; } threadIdx = {0, 0, 0}, blockDim = {1, 1, 1};
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline float __uint_as_float(unsigned u) { float f; memcpy(&f, &u, 4); return f; }
static inline unsigned __float_as_uint(float f) { unsigned u; memcpy(&u, &f, 4); return u; }
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
static inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline unsigned __byte_perm(unsigned x, unsigned y, unsigned s)
{
    const unsigned long long pool = ((unsigned long long)y << 32) | x;
    unsigned r = 0;
    for (int i = 0; i < 4; i++)
        r |= (unsigned)((pool >> (8 * ((s >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
    return r;
}
// a warp of one lane
static inline unsigned __ballot_sync(unsigned, bool p) { return p ? 1u : 0u; }
template <class T> static inline T __shfl_sync(unsigned, T v, int) { return v; }
static inline uint32_t atomicAdd(uint32_t* p, uint32_t v) { uint32_t old = *p; *p += v; return old; }
#include "trace_queue.cuh"
#include "bvh8_build.h"
using namespace pt;

static float frand() { return rand() / (float)RAND_MAX; }

int main(int argc, char** argv)
{
    const int mode = argc > 1 ? atoi(argv[1]) : 0;   // 0 random soup, 1 height field (long thin deep tree), 2 empty scene, 3 axis-aligned quads
    const int N = mode == 2 ? 0 : mode == 1 ? 2 * 90 * 90 : 4000;
    std::vector<float> v;
    std::vector<int> idx;
    srand(4321 + mode);
    if (mode == 1)
    {
        const int q = 90;
        std::vector<float> hgt((q + 1) * (q + 1));
        for (float& h : hgt) h = frand() * 0.6f;
        for (int y = 0; y < q; y++)
            for (int x = 0; x < q; x++)
            {
                auto P = [&](int i, int j, float* p) { p[0] = i * 0.1f; p[1] = hgt[j * (q + 1) + i]; p[2] = j * 0.1f; };
                float a[3], b[3], c[3], d[3];
                P(x, y, a); P(x + 1, y, b); P(x + 1, y + 1, c); P(x, y + 1, d);
                const float* tris[2][3] = {{a, b, c}, {a, c, d}};
                for (auto& t : tris)
                    for (int k = 0; k < 3; k++)
                        for (int e = 0; e < 3; e++)
                            v.push_back(t[k][e]);
            }
    }
    else
        for (int i = 0; i < N; i++)
        {
            if (mode == 3)
            {
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
                float cx = frand() * 10, cy = frand() * 10, cz = frand() * 10, s = 0.5f;
                for (int k = 0; k < 3; k++) { v.push_back(cx + frand() * s); v.push_back(cy + frand() * s); v.push_back(cz + frand() * s); }
            }
        }
    for (int i = 0; i < N; i++) { idx.push_back(3 * i); idx.push_back(3 * i + 1); idx.push_back(3 * i + 2); }
    Bvh8Host h = build_bvh8(v.data(), (int)v.size() / 3, idx.data(), N);
    Bvh8Device d;
    d.nodes = h.nodes.data();
    d.triangles = h.triangles.data();
    if (2 * h.max_depth + 1 > kTraversalStackSize) { printf("FAIL depth %d\n", h.max_depth); return 1; }

    const int R = 12000;
    std::vector<float4> ray_o(R), ray_d(R), hit(R);
    std::vector<uint32_t> ids(R), slot(R);
    std::vector<int> best(R, -1);
    std::vector<float> best_t(R, INFINITY), tmax(R);
    const float span = mode == 1 ? 9.0f : 10.0f;
    for (int r = 0; r < R; r++)
    {
        float3 o = make_float3(frand() * span, mode == 1 ? frand() * 1.5f : frand() * span, frand() * span);
        float3 dir = make_float3(frand() - 0.5f, (frand() - 0.5f) * (mode == 1 ? 0.3f : 1.0f), frand() - 0.5f);
        if (r % 7 == 0) dir.x = 0.0f;
        if (r % 11 == 0) dir.z = 0.0f;
        if (r % 13 == 0) dir.z = -0.0f;     // negative zero: its reciprocal is negative although it does not compare below zero
        if (r % 17 == 0) dir.x = -0.0f;
        float l = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
        if (l == 0.0f) { dir = make_float3(0, -1, 0); l = 1.0f; }
        dir.x /= l; dir.y /= l; dir.z /= l;
        const int last = N > 0 ? rand() % N : -1;
        tmax[r] = frand() * 12.0f;
        ray_o[r] = make_float4(o.x, o.y, o.z, tmax[r]);
        ray_d[r] = make_float4(dir.x, dir.y, dir.z, __int_as_float(last));
        ids[r] = (uint32_t)(R - 1 - r);     // the queue order is not the ray order
        for (int t = 0; t < N; t++)
        {
            if (t == last) continue;
            float tt, u, vv;
            float3 a = make_float3(v[9 * t], v[9 * t + 1], v[9 * t + 2]), b = make_float3(v[9 * t + 3], v[9 * t + 4], v[9 * t + 5]), c = make_float3(v[9 * t + 6], v[9 * t + 7], v[9 * t + 8]);
            if (intersect_triangle_reference(o, dir, a, b, c, tt, u, vv) && tt < best_t[r]) { best_t[r] = tt; best[r] = t; }
        }
    }

    int failures = 0;
    uint2 shared_stack[kSharedStackEntries];
    for (int flavour = 0; flavour < 2; flavour++)
    {
        // closest hit
        uint32_t count = R, cursor = 0;
        for (auto& x : hit) x = make_float4(-7.0f, 0, 0, __int_as_float(-7));
        RayQueueView q = {ids.data(), &count, &cursor, ray_o.data(), ray_d.data(), hit.data(), slot.data(), nullptr};
        if (flavour) trace_queue_body_pipelined<false>(d, q, shared_stack, nullptr, 1u); else trace_queue_body<false>(d, q, shared_stack, nullptr, 1u);
        int bad = 0, hits = 0;
        for (int r = 0; r < R; r++)
        {
            const int prim = __float_as_int(hit[r].w);
            if (best[r] >= 0) hits++;
            if (prim == -7) { bad++; continue; }                                             // never written
            if (prim != best[r] && !(best[r] >= 0 && prim >= 0 && hit[r].x == best_t[r])) bad++;   // exact distance ties excluded
            if (prim >= 0 && hit[r].x != best_t[r]) bad++;
            if (prim >= 0 && h.triangles[slot[r]].prim_id != prim) bad++;
        }
        // any hit
        std::vector<float4> o2 = ray_o;
        cursor = 0;
        RayQueueView qa = {ids.data(), &count, &cursor, o2.data(), ray_d.data(), nullptr, nullptr, o2.data()};
        if (flavour) trace_queue_body_pipelined<true>(d, qa, shared_stack, nullptr, 1u); else trace_queue_body<true>(d, qa, shared_stack, nullptr, 1u);
        int shadow_bad = 0;
        for (int r = 0; r < R; r++)
        {
            const bool occluded = o2[r].w == -1.0f;
            if (occluded != (best[r] >= 0 && best_t[r] < tmax[r] - 1.0e-4f)) shadow_bad++;
        }
        printf("mode %d flavour %s nodes %zu depth %d fill %.2f tris %zu hits %d closest_mismatches %d shadow_mismatches %d\n", mode, flavour ? "pipelined" : "classic", h.nodes.size(), h.max_depth,
               h.child_slots_used / (8.0 * h.nodes.size()), h.triangles.size(), hits, bad, shadow_bad);
        failures += bad + shadow_bad;
    }
    return failures ? 1 : 0;
}
