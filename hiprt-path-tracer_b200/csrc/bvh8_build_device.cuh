// Device-side construction of the compressed BVH8 (layout: bvh8.h): Morton-code LBVH -> bottom-up boxes -> area-guided collapse to
// 8-wide nodes with <= 3 triangles per leaf slot -> octant-aware slot assignment -> 8-bit quantisation.
//
// Replaces, as the opt-in GPU path of b200pt_bvh_build (flag B200PT_BVH_BUILD_DEVICE), what the reference gets from
// hiprtBuildGeometry on the device (src/HIPRT-Orochi/HIPRTScene.h:59-86; builders thirdparties/HIPRT-Fork/hiprt/impl/LbvhBuilder.h,
// PlocBuilder.h): the binary tree is the radix tree of Karras, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d
// Trees" (HPG 2012); the collapse follows Ylitie, Karras, Laine (HPG 2017), section 3.2, with the greedy surface-area rule instead of the
// host builder's SAH-optimal dynamic programme. The host builder (bvh8_build.cpp, binned SAH) stays the quality baseline and the default.
//
// Every step is a per-element function (`one thread = one triangle / one binary node / one wide node`) that compiles for the device AND
// for the host: the CUDA kernels in bvh8_build_device.cu call them with the thread index, tests/host_shim/lbvh_host_test.cpp calls them
// in a serial loop and checks the resulting tree by traversal against brute force (no GPU in the build container).
#pragma once

#include "bvh8.h"

#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define LBVH_HD __host__ __device__ __forceinline__
#else
#define LBVH_HD inline
#endif

namespace pt
{
namespace lbvh
{
struct Choice;
// Binary radix tree over n sorted primitives: internal nodes 0 .. n-2, leaf of sorted position i is node (n - 1) + i. Root = 0.
struct Tree
{
    uint32_t n;                  // primitives
    const uint64_t* keys;        // sorted Morton codes
    const uint32_t* prim;        // sorted position -> scene triangle index
    int* left;                   // [n-1] children of the internal nodes
    int* right;
    int* parent;                 // [2n-1]
    uint32_t* first;             // [2n-1] range of sorted positions below the node
    uint32_t* last;
    float* lo;                   // [2n-1][3] boxes
    float* hi;
    uint32_t* visits;            // [n-1] arrival counters of the bottom-up pass (zeroed)
    float* cost;                 // [2n-1][7] collapse costs: cheapest forest of <= i + 1 slots for the node's subtree (see refit_from_leaf)
    Choice* choice;              // [2n-1] decisions of the collapse dynamic programme
};

// Decisions of the collapse DP for one binary node (the host builder's CollapseChoice, bvh8_build.cpp): used[i-1] = slots actually taken
// when i are granted (1 = the node itself, as a leaf slot or as a wide node: `leaf`); left_slots[j-2] = slots handed to the left child
// when the node's subtree is split over j slots
struct Choice
{
    uint8_t used[7];
    uint8_t left_slots[7];
    uint8_t leaf;
    uint8_t pad;
};

LBVH_HD uint64_t expand_bits_21(uint64_t v)      // 21 bits -> every third bit of 63
{
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x1F00000000FFFFull;
    v = (v | (v << 16)) & 0x1F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}

// Box of triangle t and the 63-bit Morton code of its box centre inside the scene box
LBVH_HD uint64_t triangle_box_and_code(uint32_t t, const float* vertices, const int* indices, const float* scene_min, const float* scene_max, float* box_lo, float* box_hi)
{
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++)
    {
        const float v0 = vertices[3 * (size_t)indices[3 * (size_t)t + 0] + a], v1 = vertices[3 * (size_t)indices[3 * (size_t)t + 1] + a], v2 = vertices[3 * (size_t)indices[3 * (size_t)t + 2] + a];
        lo[a] = fminf(v0, fminf(v1, v2));
        hi[a] = fmaxf(v0, fmaxf(v1, v2));
        box_lo[a] = lo[a];
        box_hi[a] = hi[a];
    }
    // ONE cell size for the three axes (the largest extent): a Morton split then halves a cube, and a flat scene (a terrain: y extent
    // of a few percent of x and z) is split along x and z until its cells are as thin as it is, instead of being sliced into layers
    // (per-axis normalisation: SAH 58.5 on the 10 M-triangle scene against 43.5 of the host's binned-SAH tree)
    const float extent = fmaxf(scene_max[0] - scene_min[0], fmaxf(scene_max[1] - scene_min[1], scene_max[2] - scene_min[2]));
    uint64_t code = 0;
    for (int a = 0; a < 3; a++)
    {
        float u = extent > 0.0f ? (0.5f * (lo[a] + hi[a]) - scene_min[a]) / extent : 0.0f;
        u = fminf(fmaxf(u, 0.0f), 1.0f);
        const uint64_t q = (uint64_t)fminf(u * 2097152.0f, 2097151.0f);
        code |= expand_bits_21(q) << a;
    }
    return code;
}

LBVH_HD int clz64(uint64_t x)
{
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return x == 0 ? 64 : __builtin_clzll(x);
#endif
}

// Length of the common prefix of the keys at sorted positions i and j (equal keys are told apart by their position); -1 outside
LBVH_HD int common_prefix(const Tree& t, int i, int j)
{
    if (j < 0 || j >= (int)t.n)
        return -1;
    const uint64_t a = t.keys[i], b = t.keys[j];
    if (a != b)
        return clz64(a ^ b);
    return 64 + clz64((uint64_t)((uint32_t)i ^ (uint32_t)j) << 32);
}

// Karras 2012, algorithm 3: internal node i -> its range, split and children
LBVH_HD void build_internal_node(const Tree& t, int i)
{
    const int d = common_prefix(t, i, i + 1) - common_prefix(t, i, i - 1) >= 0 ? 1 : -1;
    const int delta_min = common_prefix(t, i, i - d);
    int l_max = 2;
    while (common_prefix(t, i, i + l_max * d) > delta_min)
        l_max *= 2;
    int l = 0;
    for (int step = l_max / 2; step >= 1; step /= 2)
        if (common_prefix(t, i, i + (l + step) * d) > delta_min)
            l += step;
    const int j = i + l * d;
    const int delta_node = common_prefix(t, i, j);
    int s = 0;
    for (int div = 2, step = (l + 1) / 2;; div *= 2, step = (l + div - 1) / div)
    {
        if (common_prefix(t, i, i + (s + step) * d) > delta_node)
            s += step;
        if (step <= 1)
            break;
    }
    const int gamma = i + s * d + (d < 0 ? -1 : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    const int leaf0 = (int)t.n - 1;
    const int left = lo == gamma ? leaf0 + gamma : gamma;
    const int right = hi == gamma + 1 ? leaf0 + gamma + 1 : gamma + 1;
    t.left[i] = left;
    t.right[i] = right;
    t.parent[left] = i;
    t.parent[right] = i;
    t.first[i] = (uint32_t)lo;
    t.last[i] = (uint32_t)hi;
    if (i == 0)
        t.parent[0] = -1;
}

#ifdef __CUDA_ARCH__
#define LBVH_ATOMIC_INC(p) atomicAdd((p), 1u)
#define LBVH_ATOMIC_ADD(p, v) atomicAdd((p), (v))
#define LBVH_FENCE() __threadfence()
// boxes written by OTHER threads of the same launch are read through L2: a line of the box arrays holds several nodes, and a copy of it
// that this SM cached in L1 before a neighbour's box was written would be stale
#define LBVH_LOAD_SHARED(p) __ldcg(p)
#else
// host emulation: the loops of tests/host_shim/lbvh_host_test.cpp are serial
LBVH_HD uint32_t lbvh_host_add(uint32_t* p, uint32_t v) { const uint32_t old = *p; *p += v; return old; }
#define LBVH_ATOMIC_INC(p) ::pt::lbvh::lbvh_host_add((p), 1u)
#define LBVH_ATOMIC_ADD(p, v) ::pt::lbvh::lbvh_host_add((p), (v))
#define LBVH_FENCE()
#define LBVH_LOAD_SHARED(p) (*(p))
#endif

constexpr float kCollapsePrimitiveCost = 0.4f;     // c_prim of the host builder (c_node = 1)

LBVH_HD float half_area(const float* lo, const float* hi)
{
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// Bottom-up boxes and collapse costs: leaf i sets its box and range, then walks up; the SECOND arrival at an internal node merges the children's boxes
LBVH_HD void refit_from_leaf(const Tree& t, uint32_t i, const float* prim_lo, const float* prim_hi)
{
    const int leaf = (int)t.n - 1 + (int)i;
    const uint32_t p = t.prim[i];
    for (int a = 0; a < 3; a++)
    {
        t.lo[3 * (size_t)leaf + a] = prim_lo[3 * (size_t)p + a];
        t.hi[3 * (size_t)leaf + a] = prim_hi[3 * (size_t)p + a];
    }
    t.first[leaf] = i;
    t.last[leaf] = i;
    {
        // a single triangle: always a leaf slot
        const float leaf_cost = half_area(t.lo + 3 * (size_t)leaf, t.hi + 3 * (size_t)leaf) * 1.0f * kCollapsePrimitiveCost;
        Choice ch;
        for (int k = 0; k < 7; k++) { t.cost[7 * (size_t)leaf + k] = leaf_cost; ch.used[k] = 1; ch.left_slots[k] = 1; }
        ch.leaf = 1; ch.pad = 0;
        t.choice[leaf] = ch;
    }
    int node = t.parent[leaf];
    while (node >= 0)
    {
        LBVH_FENCE();
        if (LBVH_ATOMIC_INC(t.visits + node) == 0u)
            return;                                   // the sibling subtree is not done yet: its last thread continues
        LBVH_FENCE();
        const int l = t.left[node], r = t.right[node];
        for (int a = 0; a < 3; a++)
        {
            t.lo[3 * (size_t)node + a] = fminf(LBVH_LOAD_SHARED(t.lo + 3 * (size_t)l + a), LBVH_LOAD_SHARED(t.lo + 3 * (size_t)r + a));
            t.hi[3 * (size_t)node + a] = fmaxf(LBVH_LOAD_SHARED(t.hi + 3 * (size_t)l + a), LBVH_LOAD_SHARED(t.hi + 3 * (size_t)r + a));
        }
        // SAH-optimal collapse (Ylitie, Karras, Laine 2017, section 3.2; the same programme as the host builder):
        //   leaf(n) = A(n) count(n) c_prim (count <= 3), internal(n) = A(n) c_node + distribute(n, 8), C(n, 1) = min(leaf, internal),
        //   C(n, i) = min(distribute(n, i), C(n, i-1)), distribute(n, j) = min over 0 < k < j of C(left, k) + C(right, j - k)
        {
            const float area = half_area(t.lo + 3 * (size_t)node, t.hi + 3 * (size_t)node);
            const int count = (int)(t.last[node] - t.first[node] + 1u);
            const float leaf_cost = count <= kBvh8MaxLeafTriangles ? area * (float)count * kCollapsePrimitiveCost : INFINITY;
            float cl[7], cr[7];
            for (int k = 0; k < 7; k++)
            {
                cl[k] = LBVH_LOAD_SHARED(t.cost + 7 * (size_t)l + k);
                cr[k] = LBVH_LOAD_SHARED(t.cost + 7 * (size_t)r + k);
            }
            Choice ch;
            float distribute[9];
            for (int j = 2; j <= 8; j++)
            {
                float best = INFINITY;
                int best_k = 1;
                for (int k = 1; k < j; k++)
                {
                    if (k > 7 || j - k > 7)
                        continue;
                    const float v = cl[k - 1] + cr[j - k - 1];
                    if (v < best) { best = v; best_k = k; }
                }
                distribute[j] = best;
                ch.left_slots[j - 2] = (uint8_t)best_k;
            }
            const float internal_cost = area * 1.0f + distribute[8];
            ch.leaf = leaf_cost <= internal_cost ? 1 : 0;
            ch.pad = 0;
            float c[7];
            c[0] = ch.leaf ? leaf_cost : internal_cost;
            ch.used[0] = 1;
            for (int i = 2; i <= 7; i++)
            {
                if (distribute[i] < c[i - 2]) { c[i - 1] = distribute[i]; ch.used[i - 1] = (uint8_t)i; }
                else { c[i - 1] = c[i - 2]; ch.used[i - 1] = ch.used[i - 2]; }
            }
            for (int k = 0; k < 7; k++)
                t.cost[7 * (size_t)node + k] = c[k];
            t.choice[node] = ch;
        }
        node = t.parent[node];
    }
}

// ---- PLOC (Meister & Bittner 2018, "Parallel Locally-Ordered Clustering for BVH Construction"): the binary topology by agglomerative
// clustering along the Morton curve instead of the radix tree. Every iteration each cluster looks for the neighbour within `radius`
// positions whose union with it has the smallest surface area; mutual nearest neighbours merge into a new internal node; the clusters
// are compacted (order kept) and the loop repeats until one cluster is left. The globally cheapest pair is always mutual (ties go to the
// lower position on both sides), so every iteration merges at least one pair. Costs ~10 ms more than the radix tree on 10 M triangles
// and gives a tree whose SAH is a few percent from the host builder's binned-SAH sweep (the radix tree: ~10 %).
struct Ploc
{
    uint32_t count;              // clusters of this iteration
    int radius;
    const int* cluster;          // [count] binary node of cluster i
    const float* box;            // [count][6] lo xyz, hi xyz of cluster i
    uint32_t* nearest;           // [count]
    uint64_t* flags;             // [count] low word 1: the position survives (unmerged cluster, or the lower position of a merging pair); high word 1: it merges
    const uint64_t* scanned;     // exclusive prefix sums of flags: low word = position in the next iteration, high word = rank among this iteration's merges
    int* next_cluster;           // output of the iteration
    float* next_box;
    uint32_t* subtree_count;     // [2n-1] leaves below a binary node
    int next_free_node;          // internal nodes are numbered downwards from n - 2, so that the last merge creates node 0, the root
};

LBVH_HD void ploc_nearest(const Ploc& p, uint32_t i)
{
    const float* b = p.box + 6 * (size_t)i;
    const int begin = (int)i - p.radius < 0 ? 0 : (int)i - p.radius;
    const int end = (int)i + p.radius >= (int)p.count ? (int)p.count - 1 : (int)i + p.radius;
    float best = INFINITY;
    uint32_t best_j = i;
    for (int j = begin; j <= end; j++)
    {
        if (j == (int)i)
            continue;
        const float* o = p.box + 6 * (size_t)j;
        const float lo[3] = {fminf(b[0], o[0]), fminf(b[1], o[1]), fminf(b[2], o[2])};
        const float hi[3] = {fmaxf(b[3], o[3]), fmaxf(b[4], o[4]), fmaxf(b[5], o[5])};
        const float d = half_area(lo, hi);
        if (d < best)          // ascending j: ties keep the lower position
        {
            best = d;
            best_j = (uint32_t)j;
        }
    }
    p.nearest[i] = best_j;
}

LBVH_HD void ploc_flags(const Ploc& p, uint32_t i)
{
    const uint32_t j = p.nearest[i];
    const bool mutual = j != i && p.nearest[j] == i;
    p.flags[i] = (mutual && i > j ? 0ull : 1ull) | (mutual && i < j ? 1ull << 32 : 0ull);
}

LBVH_HD void ploc_merge(const Ploc& p, const Tree& t, uint32_t i)
{
    const uint64_t f = p.flags[i];
    if ((f & 1ull) == 0ull)
        return;                  // the upper position of a merging pair: its partner writes the new cluster
    const uint32_t out = (uint32_t)(p.scanned[i] & 0xFFFFFFFFull);
    const float* b = p.box + 6 * (size_t)i;
    float* nb = p.next_box + 6 * (size_t)out;
    if ((f >> 32) == 0ull)
    {
        p.next_cluster[out] = p.cluster[i];
        for (int a = 0; a < 6; a++)
            nb[a] = b[a];
        return;
    }
    const uint32_t j = p.nearest[i];
    const float* o = p.box + 6 * (size_t)j;
    const int node = p.next_free_node - 1 - (int)(p.scanned[i] >> 32);
    const int left = p.cluster[i], right = p.cluster[j];
    t.left[node] = left;
    t.right[node] = right;
    t.parent[left] = node;
    t.parent[right] = node;
    if (node == 0)
        t.parent[0] = -1;
    p.subtree_count[node] = p.subtree_count[left] + p.subtree_count[right];
    for (int a = 0; a < 3; a++)
    {
        nb[a] = fminf(b[a], o[a]);
        nb[3 + a] = fmaxf(b[3 + a], o[3 + a]);
    }
    p.next_cluster[out] = node;
}

// Position of the first leaf below binary node v when the leaves are laid out in depth-first order (left before right): the sum, over the
// ancestors entered through their right child, of the leaves below the left sibling. The radix tree has this property by construction
// (a node covers a contiguous range of the sorted triangles); after PLOC the triangles are re-ordered to restore it.
LBVH_HD uint32_t ploc_first_leaf(const Tree& t, const uint32_t* subtree_count, int v)
{
    uint32_t first = 0;
    for (int parent = t.parent[v]; parent >= 0; v = parent, parent = t.parent[parent])
        if (t.right[parent] == v)
            first += subtree_count[t.left[parent]];
    return first;
}

// Output of the collapse
struct Wide
{
    Bvh8Node* nodes;             // capacity: n (a wide node holds >= 2 children unless it is the root of a one-triangle scene)
    Bvh8Triangle* triangles;     // n
    uint32_t* node_count;        // allocation cursors
    uint32_t* triangle_count;
    uint32_t* slots_used;        // statistics: occupied child slots
    float* sah;                  // [2]: sum of internal / leaf area ratios (atomicAdd on float: statistics only)
    const float* vertices;
    const int* indices;
};

struct WorkItem { int binary_node; uint32_t wide_node; };

// One wide node: choose its <= 8 children in the binary subtree, assign slots, quantise, emit triangles, queue the internal children.
// Returns the number of items appended to next_queue (their wide nodes are allocated contiguously, in slot order).
LBVH_HD void collapse_wide_node(const Tree& t, const Wide& w, const WorkItem item, WorkItem* next_queue, uint32_t* next_count, float root_area)
{
    const int leaf0 = (int)t.n - 1;
    auto count_of = [&](int node) { return (int)(t.last[node] - t.first[node] + 1u); };

    // ---- children: the forest the SAH-optimal collapse chose for the binary subtree (refit_from_leaf), as in the host builder's emission
    int cand[8];
    bool cand_leaf[8];
    int ncand = 0;
    if (item.binary_node >= leaf0 || count_of(item.binary_node) <= kBvh8MaxLeafTriangles)
    {
        cand_leaf[ncand] = true;                           // only the root of a tiny scene comes here
        cand[ncand++] = item.binary_node;
    }
    else
    {
        int work_node[16], work_slots[16];
        int nwork = 0;
        work_node[nwork] = item.binary_node; work_slots[nwork++] = 8;
        bool at_root = true;
        while (nwork > 0)
        {
            nwork--;
            const int node = work_node[nwork], slots = work_slots[nwork];
            const Choice ch = t.choice[node];
            const int used = at_root ? 8 : (node >= leaf0 ? 1 : ch.used[slots - 1]);
            if (used <= 1)
            {
                cand_leaf[ncand] = node >= leaf0 || ch.leaf != 0;
                cand[ncand++] = node;
            }
            else
            {
                const int k = ch.left_slots[used - 2];
                work_node[nwork] = t.right[node]; work_slots[nwork++] = used - k;
                work_node[nwork] = t.left[node]; work_slots[nwork++] = k;
            }
            at_root = false;
        }
    }

    // ---- node box and quantisation grid
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < ncand; i++)
        for (int a = 0; a < 3; a++)
        {
            lo[a] = fminf(lo[a], t.lo[3 * (size_t)cand[i] + a]);
            hi[a] = fmaxf(hi[a], t.hi[3 * (size_t)cand[i] + a]);
        }
    Bvh8Node node;
    {
        uint32_t* words = reinterpret_cast<uint32_t*>(&node);
        for (int k = 0; k < 20; k++)
            words[k] = 0u;
    }
    node.px = lo[0]; node.py = lo[1]; node.pz = lo[2];
    uint8_t exps[3];
    double scale[3];
    for (int a = 0; a < 3; a++)
    {
        const float extent = hi[a] - lo[a];
        int e = 1;                                          // biased exponent; 2^(1-127) for flat nodes
        if (extent > 0.0f)
        {
            int ex;
            frexp((double)(extent / 255.0f), &ex);          // extent/255 = m 2^ex, m in [0.5, 1) => 2^ex >= extent/255
            e = ex + 127 < 1 ? 1 : (ex + 127 > 254 ? 254 : ex + 127);
            while (e < 254 && (double)extent / ldexp(1.0, e - 127) > 255.0)
                e++;
        }
        exps[a] = (uint8_t)e;
        scale[a] = ldexp(1.0, e - 127);
    }
    node.ex = exps[0]; node.ey = exps[1]; node.ez = exps[2];

    // ---- octant-aware slot assignment: greedy maximum of dot(child centre - node centre, slot direction)
    int slot_of[8];
    bool slot_used[8] = {false, false, false, false, false, false, false, false};
    bool done[8] = {false, false, false, false, false, false, false, false};
    float centre[3];
    for (int a = 0; a < 3; a++)
        centre[a] = 0.5f * (lo[a] + hi[a]);
    for (int round = 0; round < ncand; round++)
    {
        float best = -INFINITY;
        int bc = -1, bs = -1;
        for (int i = 0; i < ncand; i++)
        {
            if (done[i])
                continue;
            float d[3];
            for (int a = 0; a < 3; a++)
                d[a] = 0.5f * (t.lo[3 * (size_t)cand[i] + a] + t.hi[3 * (size_t)cand[i] + a]) - centre[a];
            for (int s = 0; s < 8; s++)
            {
                if (slot_used[s])
                    continue;
                const float score = ((s & 1) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 4) ? d[2] : -d[2]);
                if (score > best) { best = score; bc = i; bs = s; }
            }
        }
        done[bc] = true;
        slot_used[bs] = true;
        slot_of[bc] = bs;
    }
    int cand_in_slot[8] = {-1, -1, -1, -1, -1, -1, -1, -1};
    bool leaf_in_slot[8] = {false, false, false, false, false, false, false, false};
    for (int i = 0; i < ncand; i++)
    {
        cand_in_slot[slot_of[i]] = cand[i];
        leaf_in_slot[slot_of[i]] = cand_leaf[i];
    }

    // ---- allocate the internal children (contiguous, slot order) and the triangles of the leaf slots (contiguous)
    uint32_t internal = 0, tris = 0;
    for (int s = 0; s < 8; s++)
    {
        const int c = cand_in_slot[s];
        if (c < 0)
            continue;
        if (leaf_in_slot[s])
            tris += (uint32_t)count_of(c);
        else
            internal++;
    }
    const uint32_t child_base = internal ? LBVH_ATOMIC_ADD(w.node_count, internal) : 0u;
    const uint32_t tri_base = tris ? LBVH_ATOMIC_ADD(w.triangle_count, tris) : 0u;
    const uint32_t queue_base = internal ? LBVH_ATOMIC_ADD(next_count, internal) : 0u;
    node.child_base = child_base;
    node.tri_base = tri_base;
    LBVH_ATOMIC_ADD(w.slots_used, (uint32_t)ncand);

    uint32_t next_internal = 0, tri_offset = 0;
    float sah_internal = 0.0f, sah_leaf = 0.0f;
    uint8_t* qlo[3] = {node.qlo_x, node.qlo_y, node.qlo_z};
    uint8_t* qhi[3] = {node.qhi_x, node.qhi_y, node.qhi_z};
    for (int s = 0; s < 8; s++)
    {
        const int c = cand_in_slot[s];
        if (c < 0)
            continue;                                       // meta 0 = empty slot
        const float ratio = half_area(t.lo + 3 * (size_t)c, t.hi + 3 * (size_t)c) / root_area;
        if (leaf_in_slot[s])
        {
            const int count = count_of(c);
            node.meta[s] = (uint8_t)((((1u << count) - 1u) << 5) | tri_offset);
            for (uint32_t i = t.first[c]; i <= t.last[c]; i++)
            {
                const int prim = (int)t.prim[i];
                const float* a = w.vertices + 3 * (size_t)w.indices[3 * (size_t)prim + 0];
                const float* b = w.vertices + 3 * (size_t)w.indices[3 * (size_t)prim + 1];
                const float* cc = w.vertices + 3 * (size_t)w.indices[3 * (size_t)prim + 2];
                Bvh8Triangle tri;
                tri.v0x = a[0]; tri.v0y = a[1]; tri.v0z = a[2]; tri.prim_id = prim;
                tri.v1x = b[0]; tri.v1y = b[1]; tri.v1z = b[2]; tri.pad1 = 0;
                tri.v2x = cc[0]; tri.v2y = cc[1]; tri.v2z = cc[2]; tri.pad2 = 0;
                w.triangles[tri_base + tri_offset + (i - t.first[c])] = tri;
            }
            tri_offset += (uint32_t)count;
            sah_leaf += ratio * (float)count;
        }
        else
        {
            node.imask |= (uint8_t)(1u << s);
            node.meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
            next_queue[queue_base + next_internal].binary_node = c;
            next_queue[queue_base + next_internal].wide_node = child_base + next_internal;
            next_internal++;
            sah_internal += ratio;
        }
        // conservative quantisation in double: floor for the lower bound, ceil for the upper bound
        for (int a = 0; a < 3; a++)
        {
            const double l = ((double)t.lo[3 * (size_t)c + a] - (double)lo[a]) / scale[a];
            const double h = ((double)t.hi[3 * (size_t)c + a] - (double)lo[a]) / scale[a];
            qlo[a][s] = (uint8_t)fmin(255.0, fmax(0.0, floor(l)));
            qhi[a][s] = (uint8_t)fmin(255.0, fmax(0.0, ceil(h)));
        }
    }
    w.nodes[item.wide_node] = node;
#ifdef __CUDA_ARCH__
    atomicAdd(w.sah + 0, sah_internal);
    atomicAdd(w.sah + 1, sah_leaf);
#else
    w.sah[0] += sah_internal;
    w.sah[1] += sah_leaf;
#endif
}
} // namespace lbvh
} // namespace pt
