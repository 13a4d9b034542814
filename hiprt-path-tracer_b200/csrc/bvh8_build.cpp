// Host builder for the compressed BVH8: binned-SAH binary build (OpenMP tasks) -> SAH-optimal (dynamic programming) collapse to
// 8-wide nodes -> octant-aware slot assignment -> 8-bit quantisation (layout: bvh8.h).
//
// Replaces hiprtBuildGeometry(hiprtBuildFlagBitPreferHighQualityBuild) behind HIPRTGeometry::build_bvh
// (reference: src/HIPRT-Orochi/HIPRTScene.h:59-86). Triangles keep their ORIGINAL index (`prim_id`), because
// hit primitive ids are part of the parity contract with the reference's CPURenderer BVH
// (src/Renderer/BVH.h:132-227 reports the index into Scene::triangle_indices / 3).
#include "bvh8_build.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <omp.h>

namespace pt
{
namespace
{
struct Box
{
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    float hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    void grow(const float* p)
    {
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); }
    }
    void grow(const Box& b)
    {
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); }
    }
    float half_area() const
    {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (dx < 0.0f) return 0.0f;
        return dx * dy + dy * dz + dz * dx;
    }
};

struct Node2
{
    Box box;
    int left = -1, right = -1;   // children (internal)
    int first = 0, count = 0;    // range in the permuted primitive array (count = all primitives below)
};

struct Builder2
{
    const std::vector<Box>& prim_boxes;
    const std::vector<float>& centroids;   // 3 per primitive
    std::vector<int>& order;
    std::vector<Node2> nodes;
    std::atomic<int> node_cursor{0};

    Builder2(const std::vector<Box>& pb, const std::vector<float>& c, std::vector<int>& o) : prim_boxes(pb), centroids(c), order(o)
    {
        nodes.resize(std::max<size_t>(1, 2 * o.size()));
    }

    int alloc() { return node_cursor.fetch_add(1); }

    void build(int node_index, int first, int count)
    {
        Node2& node = nodes[node_index];
        node.first = first;
        node.count = count;

        Box box, cbox;
        for (int i = first; i < first + count; i++)
        {
            box.grow(prim_boxes[order[i]]);
            cbox.grow(&centroids[3 * (size_t)order[i]]);
        }
        node.box = box;
        if (count <= 1)
            return;

        constexpr int kBins = 16;
        int best_axis = -1, best_split = -1;
        float best_cost = FLT_MAX;
        for (int axis = 0; axis < 3; axis++)
        {
            float extent = cbox.hi[axis] - cbox.lo[axis];
            if (!(extent > 0.0f))
                continue;
            Box bin_box[kBins];
            int bin_count[kBins] = {0};
            float scale = kBins / extent;
            for (int i = first; i < first + count; i++)
            {
                int p = order[i];
                int b = std::min(kBins - 1, std::max(0, (int)((centroids[3 * (size_t)p + axis] - cbox.lo[axis]) * scale)));
                bin_box[b].grow(prim_boxes[p]);
                bin_count[b]++;
            }
            float right_area[kBins];
            Box acc;
            int n = 0;
            for (int b = kBins - 1; b > 0; b--)
            {
                acc.grow(bin_box[b]);
                n += bin_count[b];
                right_area[b] = acc.half_area() * n;
            }
            acc = Box();
            n = 0;
            for (int b = 0; b < kBins - 1; b++)
            {
                acc.grow(bin_box[b]);
                n += bin_count[b];
                if (n == 0 || n == count)
                    continue;
                float cost = acc.half_area() * n + right_area[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = b; }
            }
        }

        int mid;
        if (best_axis >= 0)
        {
            float extent = cbox.hi[best_axis] - cbox.lo[best_axis];
            float scale = kBins / extent;
            float lo = cbox.lo[best_axis];
            int* begin = order.data() + first;
            int* end = begin + count;
            int* split = std::partition(begin, end, [&](int p) {
                int b = std::min(kBins - 1, std::max(0, (int)((centroids[3 * (size_t)p + best_axis] - lo) * scale)));
                return b <= best_split;
            });
            mid = (int)(split - order.data());
        }
        else
            mid = first + count / 2;   // all centroids coincide: median split keeps the tree balanced
        if (mid == first || mid == first + count)
            mid = first + count / 2;

        int l = alloc(), r = alloc();
        nodes[node_index].left = l;
        nodes[node_index].right = r;
        int lc = mid - first, rc = count - lc;
        if (count > 32768)
        {
#pragma omp task default(shared) firstprivate(l, first, lc)
            build(l, first, lc);
#pragma omp task default(shared) firstprivate(r, mid, rc)
            build(r, mid, rc);
#pragma omp taskwait
        }
        else
        {
            build(l, first, lc);
            build(r, mid, rc);
        }
    }
};

uint8_t unary_bits(int count) { return (uint8_t)((1 << count) - 1); }

// Decisions of the collapse DP for one binary node: used[i-1] = slots actually taken when i are granted (1 = the node itself,
// as a leaf slot or as a wide node: `leaf`); left_slots[j-2] = slots handed to the left child when the node's subtree is split over j
struct CollapseChoice
{
    uint8_t used[7];
    uint8_t left_slots[7];
    uint8_t leaf = 0;
};
} // namespace

Bvh8Host build_bvh8(const float* vertices, int vertex_count, const int* indices, int triangle_count)
{
    (void)vertex_count;
    Bvh8Host out;
    const size_t n = (size_t)triangle_count;

    std::vector<Box> prim_boxes(n);
    std::vector<float> centroids(3 * n);
    std::vector<int> order(n);
#pragma omp parallel for schedule(static)
    for (long long t = 0; t < (long long)n; t++)
    {
        Box b;
        for (int k = 0; k < 3; k++)
            b.grow(vertices + 3 * (size_t)indices[3 * t + k]);
        prim_boxes[t] = b;
        // same centroid rule as the reference octree's insertion (Triangle::bbox_centroid, src/Renderer/Triangle.cpp:8-11)
        for (int a = 0; a < 3; a++)
            centroids[3 * t + a] = 0.5f * (b.lo[a] + b.hi[a]);
        order[t] = (int)t;
    }

    if (n == 0)
    {
        // One empty node so that traversal always has a valid root
        Bvh8Node root;
        std::memset(&root, 0, sizeof(root));
        out.nodes.push_back(root);
        for (int a = 0; a < 3; a++) { out.scene_min[a] = 0.0f; out.scene_max[a] = 0.0f; }
        return out;
    }

    Builder2 b2(prim_boxes, centroids, order);
    int root2 = b2.alloc();
#pragma omp parallel
#pragma omp single
    b2.build(root2, 0, (int)n);

    const std::vector<Node2>& n2 = b2.nodes;
    for (int a = 0; a < 3; a++) { out.scene_min[a] = n2[root2].box.lo[a]; out.scene_max[a] = n2[root2].box.hi[a]; }

    // ---- SAH-optimal collapse of the binary tree into 8-wide nodes (Ylitie, Karras, Laine 2017, section 3.2) --------------
    // cost[n][i-1] = cheapest way to represent the subtree of n as a forest of at most i children of some wide node:
    //   leaf(n)       = A(n) * count(n) * c_prim                      (count <= 3: one slot addresses at most 3 triangles)
    //   internal(n)   = A(n) * c_node + distribute(n, 8)              (n becomes a wide node of its own)
    //   C(n, 1)       = min(leaf, internal);   C(n, i) = min(distribute(n, i), C(n, i-1))
    //   distribute(n, j) = min over 0 < k < j of C(left, k) + C(right, j - k)
    // Children are allocated after their parents, so decreasing index order is bottom-up.
    const int node_total = b2.node_cursor.load();
    std::vector<CollapseChoice> choice((size_t)node_total);
    {
        static const float c_prim = [] { const char* v = std::getenv("B200PT_BVH_PRIMITIVE_COST"); float f = v ? (float)std::atof(v) : 0.0f; return f > 0.0f ? f : 0.4f; }();
        const float c_node = 1.0f;
        std::vector<float> cost((size_t)node_total * 7);
        for (int n_index = node_total - 1; n_index >= 0; n_index--)
        {
            const Node2& nd = n2[n_index];
            float* c = &cost[(size_t)n_index * 7];
            CollapseChoice& ch = choice[n_index];
            const float area = nd.box.half_area();
            const float leaf = nd.count <= kBvh8MaxLeafTriangles ? area * nd.count * c_prim : FLT_MAX;
            if (nd.left < 0)
            {
                for (int i = 0; i < 7; i++) { c[i] = leaf; ch.used[i] = 1; }
                ch.leaf = 1;
                continue;
            }
            const float* cl = &cost[(size_t)nd.left * 7];
            const float* cr = &cost[(size_t)nd.right * 7];
            float distribute[9];
            for (int j = 2; j <= 8; j++)
            {
                float best = FLT_MAX;
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
            const float internal = area * c_node + distribute[8];
            ch.leaf = leaf <= internal;
            c[0] = ch.leaf ? leaf : internal;
            ch.used[0] = 1;
            for (int i = 2; i <= 7; i++)
            {
                if (distribute[i] < c[i - 2]) { c[i - 1] = distribute[i]; ch.used[i - 1] = (uint8_t)i; }
                else { c[i - 1] = c[i - 2]; ch.used[i - 1] = ch.used[i - 2]; }
            }
        }
    }

    // ---- emit breadth-first -------------------------------------------------------------------
    out.nodes.reserve(n / 4 + 16);
    out.triangles.reserve(n);
    out.nodes.emplace_back();
    std::queue<std::pair<int, uint32_t>> todo;   // (binary node, wide node index)
    std::queue<int> todo_depth;
    todo.push({root2, 0u});
    todo_depth.push(1);

    double sah_internal = 0.0, sah_leaf = 0.0;
    const double root_area = std::max(1e-30f, n2[root2].box.half_area());

    while (!todo.empty())
    {
        auto [bin_index, wide_index] = todo.front();
        todo.pop();
        const int depth = todo_depth.front();
        todo_depth.pop();
        out.max_depth = std::max(out.max_depth, depth);

        // Children of this wide node: the forest the SAH-optimal collapse chose for the binary subtree (see collapse_costs)
        int cand[8];
        bool cand_is_leaf[8];
        int ncand = 0;
        if (n2[bin_index].left < 0 || (bin_index == root2 && n2[bin_index].count <= kBvh8MaxLeafTriangles))
        {
            cand_is_leaf[ncand] = true;
            cand[ncand++] = bin_index;
        }
        else
        {
            // explicit stack of (binary node, slots granted); a grant of 1 ends in the node's own leaf / internal decision
            std::pair<int, int> work[16];
            int nwork = 0;
            work[nwork++] = {bin_index, 8};
            bool at_root = true;
            while (nwork > 0)
            {
                auto [node, slots] = work[--nwork];
                const CollapseChoice& ch = choice[node];
                const int used = at_root ? 8 : (n2[node].left < 0 ? 1 : ch.used[slots - 1]);
                if (used <= 1)
                {
                    cand_is_leaf[ncand] = n2[node].left < 0 || ch.leaf;
                    cand[ncand++] = node;
                }
                else
                {
                    const int k = ch.left_slots[used - 2];
                    work[nwork++] = {n2[node].right, used - k};
                    work[nwork++] = {n2[node].left, k};
                }
                at_root = false;
            }
        }

        // Node box and quantisation grid
        Box box;
        for (int i = 0; i < ncand; i++)
            box.grow(n2[cand[i]].box);
        Bvh8Node node;
        std::memset(&node, 0, sizeof(node));
        node.px = box.lo[0]; node.py = box.lo[1]; node.pz = box.lo[2];
        uint8_t exps[3];
        float scale[3];
        for (int a = 0; a < 3; a++)
        {
            float extent = box.hi[a] - box.lo[a];
            int e = 1;   // biased exponent; 2^(1-127) for flat nodes
            if (extent > 0.0f)
            {
                int ex;
                std::frexp(extent / 255.0f, &ex);   // extent/255 = m * 2^ex, m in [0.5, 1)  =>  2^ex >= extent/255
                e = std::min(254, std::max(1, ex + 127));
                // guard against rounding in the division above
                while (e < 254 && (double)extent / std::ldexp(1.0, e - 127) > 255.0)
                    e++;
            }
            exps[a] = (uint8_t)e;
            scale[a] = std::ldexp(1.0f, e - 127);
        }
        node.ex = exps[0]; node.ey = exps[1]; node.ez = exps[2];

        // Octant-aware slot assignment: slot bit a set <=> child sits on the + side of the node centre along axis a.
        // Greedy maximum of dot(child centre - node centre, slot direction) (Ylitie et al. 2017, section 4.2).
        int slot_of[8];
        bool slot_used[8] = {false};
        bool cand_done[8] = {false};
        float centre[3];
        for (int a = 0; a < 3; a++)
            centre[a] = 0.5f * (box.lo[a] + box.hi[a]);
        for (int round = 0; round < ncand; round++)
        {
            float best = -FLT_MAX;
            int bc = -1, bs = -1;
            for (int i = 0; i < ncand; i++)
            {
                if (cand_done[i]) continue;
                const Box& cb = n2[cand[i]].box;
                float d[3];
                for (int a = 0; a < 3; a++)
                    d[a] = 0.5f * (cb.lo[a] + cb.hi[a]) - centre[a];
                for (int s = 0; s < 8; s++)
                {
                    if (slot_used[s]) continue;
                    float cost = ((s & 1) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 4) ? d[2] : -d[2]);
                    if (cost > best) { best = cost; bc = i; bs = s; }
                }
            }
            cand_done[bc] = true;
            slot_used[bs] = true;
            slot_of[bc] = bs;
        }
        int cand_in_slot[8];
        bool leaf_in_slot[8] = {false};
        for (int s = 0; s < 8; s++) cand_in_slot[s] = -1;
        for (int i = 0; i < ncand; i++) { cand_in_slot[slot_of[i]] = cand[i]; leaf_in_slot[slot_of[i]] = cand_is_leaf[i]; }

        node.child_base = (uint32_t)out.nodes.size();
        node.tri_base = (uint32_t)out.triangles.size();
        int tri_offset = 0;
        for (int s = 0; s < 8; s++)
        {
            int c = cand_in_slot[s];
            if (c < 0)
                continue;   // meta 0 = empty slot
            const Node2& child = n2[c];
            out.child_slots_used++;
            double area_ratio = child.box.half_area() / root_area;
            if (!leaf_in_slot[s])
            {
                node.imask |= (uint8_t)(1u << s);
                node.meta[s] = (uint8_t)((1u << 5) | (24u + s));
                todo.push({c, (uint32_t)out.nodes.size()});
                todo_depth.push(depth + 1);
                out.nodes.emplace_back();
                sah_internal += area_ratio;
            }
            else
            {
                node.meta[s] = (uint8_t)((unary_bits(child.count) << 5) | tri_offset);
                for (int i = child.first; i < child.first + child.count; i++)
                {
                    int prim = order[i];
                    Bvh8Triangle tri;
                    const float* a = vertices + 3 * (size_t)indices[3 * (size_t)prim + 0];
                    const float* b = vertices + 3 * (size_t)indices[3 * (size_t)prim + 1];
                    const float* cc = vertices + 3 * (size_t)indices[3 * (size_t)prim + 2];
                    tri.v0x = a[0]; tri.v0y = a[1]; tri.v0z = a[2]; tri.prim_id = prim;
                    tri.v1x = b[0]; tri.v1y = b[1]; tri.v1z = b[2]; tri.pad1 = 0;
                    tri.v2x = cc[0]; tri.v2y = cc[1]; tri.v2z = cc[2]; tri.pad2 = 0;
                    out.triangles.push_back(tri);
                }
                tri_offset += child.count;
                sah_leaf += area_ratio * child.count;
            }

            // conservative quantisation in double: floor for the lower bound, ceil for the upper bound
            uint8_t* qlo[3] = {node.qlo_x, node.qlo_y, node.qlo_z};
            uint8_t* qhi[3] = {node.qhi_x, node.qhi_y, node.qhi_z};
            for (int a = 0; a < 3; a++)
            {
                double lo = ((double)child.box.lo[a] - (double)box.lo[a]) / (double)scale[a];
                double hi = ((double)child.box.hi[a] - (double)box.lo[a]) / (double)scale[a];
                qlo[a][s] = (uint8_t)std::min(255.0, std::max(0.0, std::floor(lo)));
                qhi[a][s] = (uint8_t)std::min(255.0, std::max(0.0, std::ceil(hi)));
            }
        }
        out.nodes[wide_index] = node;
    }

    out.sah_cost = (float)(sah_internal * 1.0 + sah_leaf * 1.0);
    return out;
}
} // namespace pt
