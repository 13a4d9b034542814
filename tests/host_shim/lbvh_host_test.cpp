// TEST INFRASTRUCTURE: the per-element functions of the DEVICE BVH8 builder (csrc/bvh8_build_device.cuh: Morton codes, Karras radix
// tree, bottom-up boxes, collapse to 8-wide nodes) run here in serial loops on the host, and the tree they produce is checked by
// traversal (csrc/traverse.cuh through the intrinsics shim) against a brute-force loop over all triangles. The CUDA kernels of
// bvh8_build_device.cu call the same functions with a thread index; this pins their logic where no GPU is available.
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
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
#include "bvh8_build_device.cuh"
using namespace pt;

static float frand() { return rand() / (float)RAND_MAX; }

int main(int argc, char** argv)
{
    const int mode_arg = argc > 1 ? atoi(argv[1]) : 0;
    const bool ploc = mode_arg >= 10;                // 10 + mode: the binary topology by PLOC clustering instead of the radix tree
    const int ploc_radius = argc > 2 ? atoi(argv[2]) : 8;
    const int mode = mode_arg % 10;                  // 0 random soup, 1 height field, 2 one triangle, 3 axis-aligned quads with duplicates, 4 two triangles
    const int N = mode == 2 ? 1 : mode == 4 ? 2 : mode == 1 ? 2 * 70 * 70 : 3000;
    std::vector<float> v;
    std::vector<int> idx;
    srand(99 + mode);
    if (mode == 1)
    {
        const int q = 70;
        std::vector<float> hgt((q + 1) * (q + 1));
        for (float& h : hgt) h = frand() * 0.6f;
        for (int y = 0; y < q; y++)
            for (int x = 0; x < q; x++)
            {
                auto P = [&](int i, int j, float* p) { p[0] = i * 0.1f; p[1] = hgt[j * (q + 1) + i]; p[2] = j * 0.1f; };
                float a[3], b[3], c[3], d[3];
                P(x, y, a); P(x + 1, y, b); P(x + 1, y + 1, c); P(x, y + 1, d);
                const float* tris[2][3] = {{a, b, c}, {a, c, d}};
                for (auto& t : tris) for (int k = 0; k < 3; k++) for (int e = 0; e < 3; e++) v.push_back(t[k][e]);
            }
    }
    else
        for (int i = 0; i < N; i++)
        {
            if (mode == 3)
            {
                // exact duplicates (equal Morton codes) and flat boxes
                int axis = i % 3;
                float c[3] = {(float)(rand() % 5), (float)(rand() % 5), (float)(rand() % 5)};
                float p[3][3];
                for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) p[k][a] = c[a];
                p[1][(axis + 1) % 3] += 1.0f;
                p[2][(axis + 2) % 3] += 1.0f;
                for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) v.push_back(p[k][a]);
            }
            else
            {
                float cx = frand() * 10, cy = frand() * 10, cz = frand() * 10, s = mode == 2 ? 5.0f : 0.5f;
                for (int k = 0; k < 3; k++) { v.push_back(cx + frand() * s); v.push_back(cy + frand() * s); v.push_back(cz + frand() * s); }
            }
        }
    for (int i = 0; i < N; i++) { idx.push_back(3 * i); idx.push_back(3 * i + 1); idx.push_back(3 * i + 2); }

    // ---- the builder's steps, serially ----
    const uint32_t n = (uint32_t)N;
    float smin[3] = {INFINITY, INFINITY, INFINITY}, smax[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (size_t i = 0; i < v.size(); i++) { smin[i % 3] = fminf(smin[i % 3], v[i]); smax[i % 3] = fmaxf(smax[i % 3], v[i]); }
    std::vector<float> plo(3 * n), phi(3 * n);
    std::vector<uint64_t> codes(n);
    for (uint32_t t = 0; t < n; t++)
        codes[t] = lbvh::triangle_box_and_code(t, v.data(), idx.data(), smin, smax, &plo[3 * t], &phi[3 * t]);
    std::vector<uint32_t> order(n);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return codes[a] < codes[b]; });
    std::vector<uint64_t> keys(n);
    for (uint32_t i = 0; i < n; i++) keys[i] = codes[order[i]];
    const size_t nodes2 = 2 * (size_t)n - 1;
    std::vector<int> left(n), right(n), parent(nodes2, -1);
    std::vector<uint32_t> first(nodes2), last(nodes2), visits(n, 0u);
    std::vector<float> lo(3 * nodes2), hi(3 * nodes2);
    std::vector<float> cost(7 * nodes2);
    std::vector<lbvh::Choice> choice(nodes2);
    lbvh::Tree t = {n, keys.data(), order.data(), left.data(), right.data(), parent.data(), first.data(), last.data(), lo.data(), hi.data(), visits.data(), cost.data(), choice.data()};
    std::vector<uint32_t> reordered(n);
    int ploc_iterations = 0;
    if (ploc && n > 1)
    {
        // the kernels of bvh8_build_device.cu, serially: init, then (nearest, flags, exclusive scan, merge) until one cluster is left
        std::vector<int> cluster[2] = {std::vector<int>(n), std::vector<int>(n)};
        std::vector<float> box[2] = {std::vector<float>(6 * (size_t)n), std::vector<float>(6 * (size_t)n)};
        std::vector<uint32_t> nearest(n), subtree(nodes2, 0u);
        std::vector<uint64_t> flags(n), scanned(n);
        for (uint32_t i = 0; i < n; i++)
        {
            cluster[0][i] = (int)(n - 1 + i);
            subtree[n - 1 + i] = 1;
            for (int a = 0; a < 3; a++) { box[0][6 * i + a] = plo[3 * order[i] + a]; box[0][6 * i + 3 + a] = phi[3 * order[i] + a]; }
        }
        uint32_t clusters = n;
        int next_free = (int)n - 1;
        while (clusters > 1)
        {
            const int in = ploc_iterations & 1, out = in ^ 1;
            lbvh::Ploc p = {clusters, ploc_radius, cluster[in].data(), box[in].data(), nearest.data(), flags.data(), scanned.data(), cluster[out].data(), box[out].data(), subtree.data(), next_free};
            for (uint32_t i = 0; i < clusters; i++) lbvh::ploc_nearest(p, i);
            for (uint32_t i = 0; i < clusters; i++) lbvh::ploc_flags(p, i);
            uint64_t sum = 0;
            for (uint32_t i = 0; i < clusters; i++) { scanned[i] = sum; sum += flags[i]; }
            for (uint32_t i = 0; i < clusters; i++) lbvh::ploc_merge(p, t, i);
            const uint32_t merges = (uint32_t)(sum >> 32);
            if (merges == 0) { printf("FAIL: PLOC iteration %d merged nothing\n", ploc_iterations); return 1; }
            clusters = (uint32_t)(sum & 0xFFFFFFFFull);
            next_free -= (int)merges;
            ploc_iterations++;
        }
        if (next_free != 0 || parent[0] != -1) { printf("FAIL: PLOC left %d unused internal nodes, root parent %d\n", next_free, parent[0]); return 1; }
        // depth-first leaf layout, triangles moved, leaves renumbered (k_ploc_first_leaf / _move_leaves / _renumber / _leaf_parents)
        for (size_t v2 = 0; v2 < nodes2; v2++) first[v2] = lbvh::ploc_first_leaf(t, subtree.data(), (int)v2);
        std::vector<int> leaf_parent(n);
        std::vector<uint32_t> new_position(n);
        for (uint32_t i = 0; i < n; i++) { const uint32_t q = first[n - 1 + i]; reordered[q] = order[i]; leaf_parent[q] = parent[n - 1 + i]; new_position[i] = q; }
        for (uint32_t v2 = 0; v2 + 1 < n; v2++)
        {
            const int leaf0 = (int)n - 1;
            if (left[v2] >= leaf0) left[v2] = leaf0 + (int)new_position[left[v2] - leaf0];
            if (right[v2] >= leaf0) right[v2] = leaf0 + (int)new_position[right[v2] - leaf0];
            last[v2] = first[v2] + subtree[v2] - 1;
        }
        for (uint32_t q = 0; q < n; q++) parent[n - 1 + q] = leaf_parent[q];
        std::vector<int> once(n, 0);
        for (uint32_t q = 0; q < n; q++) once[reordered[q]]++;
        for (uint32_t q = 0; q < n; q++) if (once[q] != 1) { printf("FAIL: the depth-first order is not a permutation\n"); return 1; }
        t.prim = reordered.data();
    }
    else
        for (int i = 0; i + 1 < (int)n; i++)
            lbvh::build_internal_node(t, i);
    for (uint32_t i = 0; i < n; i++)
        lbvh::refit_from_leaf(t, i, plo.data(), phi.data());
    std::vector<Bvh8Node> nodes(std::max<size_t>(n, 1));
    std::vector<Bvh8Triangle> tris(n);
    uint32_t node_count = 1, tri_count = 0, slots = 0;
    float sah[2] = {0.0f, 0.0f};
    lbvh::Wide w = {nodes.data(), tris.data(), &node_count, &tri_count, &slots, sah, v.data(), idx.data()};
    std::vector<lbvh::WorkItem> queue(1), next(std::max<size_t>(n, 1));
    queue[0] = {n == 1 ? 0 : 0, 0u};          // the root: internal node 0, or leaf 0 (= node n - 1 = 0) of a one-triangle scene
    const float root_area = std::max(1e-30f, lbvh::half_area(&lo[0], &hi[0]));
    int depth = 0;
    while (!queue.empty())
    {
        depth++;
        uint32_t next_count = 0;
        for (const lbvh::WorkItem& item : queue)
            lbvh::collapse_wide_node(t, w, item, next.data(), &next_count, root_area);
        queue.assign(next.begin(), next.begin() + next_count);
    }
    if (tri_count != n) { printf("FAIL: %u of %u triangles emitted\n", tri_count, n); return 1; }
    std::vector<int> seen(n, 0);
    for (uint32_t i = 0; i < n; i++) seen[tris[i].prim_id]++;
    for (uint32_t i = 0; i < n; i++) if (seen[i] != 1) { printf("FAIL: triangle %u emitted %d times\n", i, seen[i]); return 1; }

    Bvh8Device d;
    d.nodes = nodes.data();
    d.triangles = tris.data();
    int bad = 0, hits = 0, shadow_bad = 0;
    for (int r = 0; r < 20000; r++)
    {
        float3 o = make_float3(frand() * 10 - 1, frand() * (mode == 1 ? 2.0f : 10.0f) - (mode == 1 ? 0.5f : 1.0f), frand() * 10 - 1);
        float3 dir = make_float3(frand() - 0.5f, frand() - 0.5f, frand() - 0.5f);
        if (r % 7 == 0) dir.x = 0.0f;
        if (r % 13 == 0) dir.z = -0.0f;
        float l = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
        if (l == 0.0f) continue;
        dir.x /= l; dir.y /= l; dir.z /= l;
        int lastp = rand() % N;
        Hit hit;
        bvh8_trace<false>(d, o, dir, INFINITY, lastp, hit, nullptr);
        int best = -1;
        float bt = INFINITY;
        for (int k = 0; k < N; k++)
        {
            if (k == lastp) continue;
            float tt, u, vv;
            float3 a = make_float3(v[9 * k], v[9 * k + 1], v[9 * k + 2]), b = make_float3(v[9 * k + 3], v[9 * k + 4], v[9 * k + 5]), c = make_float3(v[9 * k + 6], v[9 * k + 7], v[9 * k + 8]);
            if (intersect_triangle_reference(o, dir, a, b, c, tt, u, vv) && tt < bt) { bt = tt; best = k; }
        }
        if (best >= 0) hits++;
        if (best != hit.prim && !(best >= 0 && hit.prim >= 0 && hit.t == bt)) bad++;
        float tmax = frand() * 12.0f;
        Hit sh;
        bool occluded = bvh8_trace<true>(d, o, dir, tmax - 1.0e-4f, lastp, sh, nullptr);
        if (occluded != (best >= 0 && bt < tmax - 1.0e-4f)) shadow_bad++;
    }
    printf("mode %d (%s, %d clustering iterations): %u wide nodes, depth %d, fill %.2f, sah %.2f, hits %d, closest_mismatches %d, shadow_mismatches %d\n", mode, ploc ? "PLOC" : "radix tree",
           ploc_iterations, node_count, depth, slots / (8.0 * node_count), sah[0] + sah[1], hits, bad, shadow_bad);
    return (bad || shadow_bad || 2 * depth + 1 > kTraversalStackSize) ? 1 : 0;
}
