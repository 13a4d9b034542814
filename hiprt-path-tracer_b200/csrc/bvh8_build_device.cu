// CUDA kernels and host driver of the device BVH8 builder (per-element functions: bvh8_build_device.cuh).
// Pipeline: scene box (reduction) -> triangle boxes + 63-bit Morton codes -> radix sort of (code, triangle) pairs -> binary topology (PLOC:
// mutual-nearest-neighbour clustering along the Morton curve, the default; or the Karras radix tree) -> bottom-up boxes + SAH-optimal
// collapse costs -> top-down emission of the 8-wide nodes, one launch per tree level.
// The radix sort and the prefix sums are cub's (CUDA toolkit header library): plumbing here, every tree-building step is ours.
#include "bvh8_build_device.cuh"
#include "bvh8_build_device.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <cstring>
#include <vector>

namespace pt
{
namespace
{
__device__ __forceinline__ void atomic_min_float(float* addr, float v)
{
    // ordered-int trick: valid for any mix of signs
    if (v >= 0.0f)
        atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
    else
        atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_float(float* addr, float v)
{
    if (v >= 0.0f)
        atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
    else
        atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}

// scene box over the vertices the triangles use (bounds[0..2] = min, [3..5] = max; initialised to +-FLT_MAX)
__global__ void __launch_bounds__(256) k_lbvh_scene_box(const float* __restrict__ vertices, const int* __restrict__ indices, uint32_t n, float* bounds)
{
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x)
        for (int k = 0; k < 3; k++)
        {
            const float* p = vertices + 3 * (size_t)indices[3 * (size_t)t + k];
            for (int a = 0; a < 3; a++)
            {
                lo[a] = fminf(lo[a], p[a]);
                hi[a] = fmaxf(hi[a], p[a]);
            }
        }
    for (int a = 0; a < 3; a++)
    {
        for (int offset = 16; offset > 0; offset >>= 1)
        {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xFFFFFFFFu, lo[a], offset));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xFFFFFFFFu, hi[a], offset));
        }
        if ((threadIdx.x & 31u) == 0u)
        {
            atomic_min_float(bounds + a, lo[a]);
            atomic_max_float(bounds + 3 + a, hi[a]);
        }
    }
}

__global__ void __launch_bounds__(256) k_lbvh_codes(const float* __restrict__ vertices, const int* __restrict__ indices, uint32_t n, const float* __restrict__ bounds, float* prim_lo,
                                                    float* prim_hi, uint64_t* codes, uint32_t* ids)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n)
        return;
    codes[t] = lbvh::triangle_box_and_code(t, vertices, indices, bounds, bounds + 3, prim_lo + 3 * (size_t)t, prim_hi + 3 * (size_t)t);
    ids[t] = t;
}

__global__ void __launch_bounds__(256) k_lbvh_internal(lbvh::Tree tree)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1u < tree.n)
        lbvh::build_internal_node(tree, (int)i);
}

__global__ void __launch_bounds__(256) k_lbvh_refit(lbvh::Tree tree, const float* __restrict__ prim_lo, const float* __restrict__ prim_hi)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < tree.n)
        lbvh::refit_from_leaf(tree, i, prim_lo, prim_hi);
}

__global__ void __launch_bounds__(128) k_lbvh_collapse(lbvh::Tree tree, lbvh::Wide wide, const lbvh::WorkItem* __restrict__ queue, uint32_t count, lbvh::WorkItem* next_queue,
                                                       uint32_t* next_count, const float* __restrict__ root_box /* scene box: min[3], max[3] */)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count)
        return;
    const float root_area = fmaxf(1.0e-30f, lbvh::half_area(root_box, root_box + 3));
    lbvh::collapse_wide_node(tree, wide, queue[i], next_queue, next_count, root_area);
}

// ---- PLOC (bvh8_build_device.cuh) ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_ploc_init(lbvh::Tree tree, const float* __restrict__ prim_lo, const float* __restrict__ prim_hi, int* cluster, float* box, uint32_t* subtree_count)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= tree.n)
        return;
    const uint32_t p = tree.prim[i];
    cluster[i] = (int)(tree.n - 1u + i);
    subtree_count[tree.n - 1u + i] = 1u;
    for (int a = 0; a < 3; a++)
    {
        box[6 * (size_t)i + a] = prim_lo[3 * (size_t)p + a];
        box[6 * (size_t)i + 3 + a] = prim_hi[3 * (size_t)p + a];
    }
}
__global__ void __launch_bounds__(256) k_ploc_nearest(lbvh::Ploc p)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.count)
        lbvh::ploc_nearest(p, i);
}
__global__ void __launch_bounds__(256) k_ploc_flags(lbvh::Ploc p)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.count)
        lbvh::ploc_flags(p, i);
}
__global__ void __launch_bounds__(256) k_ploc_merge(lbvh::Ploc p, lbvh::Tree tree)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.count)
        lbvh::ploc_merge(p, tree, i);
}
// Depth-first leaf layout of the clustered tree: first leaf of every node, then the triangles move to that order and the leaves are renumbered
__global__ void __launch_bounds__(256) k_ploc_first_leaf(lbvh::Tree tree, const uint32_t* __restrict__ subtree_count)
{
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < 2u * tree.n - 1u)
        tree.first[v] = lbvh::ploc_first_leaf(tree, subtree_count, (int)v);
}
__global__ void __launch_bounds__(256) k_ploc_move_leaves(lbvh::Tree tree, uint32_t* new_prim, int* new_leaf_parent, uint32_t* new_position)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= tree.n)
        return;
    const uint32_t leaf = tree.n - 1u + i, q = tree.first[leaf];
    new_prim[q] = tree.prim[i];
    new_leaf_parent[q] = tree.parent[leaf];
    new_position[i] = q;
}
__global__ void __launch_bounds__(256) k_ploc_renumber(lbvh::Tree tree, const uint32_t* __restrict__ subtree_count, const uint32_t* __restrict__ new_position)
{
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v + 1u >= tree.n)
        return;                  // internal nodes: 0 .. n - 2
    const int leaf0 = (int)tree.n - 1;
    const int l = tree.left[v], r = tree.right[v];
    if (l >= leaf0)
        tree.left[v] = leaf0 + (int)new_position[l - leaf0];
    if (r >= leaf0)
        tree.right[v] = leaf0 + (int)new_position[r - leaf0];
    tree.last[v] = tree.first[v] + subtree_count[v] - 1u;
}
__global__ void __launch_bounds__(256) k_ploc_leaf_parents(lbvh::Tree tree, const int* __restrict__ new_leaf_parent)
{
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < tree.n)
        tree.parent[tree.n - 1u + q] = new_leaf_parent[q];
}

struct DeviceArena
{
    std::vector<void*> blocks;
    cudaError_t error = cudaSuccess;
    template <class T> T* alloc(size_t count)
    {
        void* p = nullptr;
        if (error == cudaSuccess)
            error = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
        if (error == cudaSuccess)
            blocks.push_back(p);
        return static_cast<T*>(p);
    }
    ~DeviceArena()
    {
        for (void* p : blocks)
            cudaFree(p);
    }
};
} // namespace

// vertices / indices: HOST arrays, as b200pt_bvh_build receives them. On success *out_nodes / *out_triangles are device arrays the caller
// owns (cudaFree), trimmed to the node / triangle counts.
cudaError_t build_bvh8_device(const float* vertices, int vertex_count, const int* indices, int triangle_count, cudaStream_t stream, Bvh8DeviceBuild* out, int algorithm, int ploc_radius)
{
    *out = Bvh8DeviceBuild();
    const uint32_t n = (uint32_t)triangle_count;
    if (n == 0)
        return cudaErrorInvalidValue;          // the caller builds the one-empty-node tree of an empty scene on the host
    DeviceArena arena;
    float* d_vertices = arena.alloc<float>(3 * (size_t)vertex_count);
    int* d_indices = arena.alloc<int>(3 * (size_t)n);
    float* d_bounds = arena.alloc<float>(6);
    float* d_prim_lo = arena.alloc<float>(3 * (size_t)n);
    float* d_prim_hi = arena.alloc<float>(3 * (size_t)n);
    uint64_t* d_codes = arena.alloc<uint64_t>(n);
    uint64_t* d_keys = arena.alloc<uint64_t>(n);
    uint32_t* d_ids = arena.alloc<uint32_t>(n);
    uint32_t* d_prim = arena.alloc<uint32_t>(n);
    const size_t nodes2 = 2 * (size_t)n - 1;
    int* d_left = arena.alloc<int>(n);
    int* d_right = arena.alloc<int>(n);
    int* d_parent = arena.alloc<int>(nodes2);
    uint32_t* d_first = arena.alloc<uint32_t>(nodes2);
    uint32_t* d_last = arena.alloc<uint32_t>(nodes2);
    float* d_lo = arena.alloc<float>(3 * nodes2);
    float* d_hi = arena.alloc<float>(3 * nodes2);
    uint32_t* d_visits = arena.alloc<uint32_t>(n);
    float* d_cost = arena.alloc<float>(7 * nodes2);
    lbvh::Choice* d_choice = arena.alloc<lbvh::Choice>(nodes2);
    lbvh::WorkItem* d_queue[2] = {arena.alloc<lbvh::WorkItem>(n), arena.alloc<lbvh::WorkItem>(n)};
    uint32_t* d_counters = arena.alloc<uint32_t>(8);      // [0] node count, [1] triangle count, [2] slots used, [3] next-queue count, [4..5] sah (float)
    if (arena.error != cudaSuccess)
        return arena.error;
    Bvh8Node* d_nodes = nullptr;
    Bvh8Triangle* d_triangles = nullptr;
    cudaError_t e = cudaMalloc(&d_nodes, sizeof(Bvh8Node) * (size_t)n);
    if (e == cudaSuccess)
        e = cudaMalloc(&d_triangles, sizeof(Bvh8Triangle) * (size_t)n);
    auto fail = [&](cudaError_t err) {
        cudaFree(d_nodes);
        cudaFree(d_triangles);
        return err;
    };
    if (e != cudaSuccess)
        return fail(e);

    cudaEvent_t ev0, ev1;
    cudaEventCreate(&ev0);
    cudaEventCreate(&ev1);
    e = cudaMemcpyAsync(d_vertices, vertices, sizeof(float) * 3 * (size_t)vertex_count, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d_indices, indices, sizeof(int) * 3 * (size_t)n, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess)
        return fail(e);
    cudaEventRecord(ev0, stream);          // the build proper: geometry already on the device

    const float init_bounds[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    cudaMemcpyAsync(d_bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, stream);
    const unsigned grid_n = (n + 255u) / 256u;
    k_lbvh_scene_box<<<std::min(grid_n, 148u * 16u), 256, 0, stream>>>(d_vertices, d_indices, n, d_bounds);
    k_lbvh_codes<<<grid_n, 256, 0, stream>>>(d_vertices, d_indices, n, d_bounds, d_prim_lo, d_prim_hi, d_codes, d_ids);
    size_t temp_bytes = 0;
    e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, d_codes, d_keys, d_ids, d_prim, (int)n, 0, 63, stream);
    if (e != cudaSuccess)
        return fail(e);
    void* d_temp = arena.alloc<unsigned char>(temp_bytes);
    if (arena.error != cudaSuccess)
        return fail(arena.error);
    e = cub::DeviceRadixSort::SortPairs(d_temp, temp_bytes, d_codes, d_keys, d_ids, d_prim, (int)n, 0, 63, stream);
    if (e != cudaSuccess)
        return fail(e);

    lbvh::Tree tree = {n, d_keys, d_prim, d_left, d_right, d_parent, d_first, d_last, d_lo, d_hi, d_visits, d_cost, d_choice};
    cudaMemsetAsync(d_visits, 0, sizeof(uint32_t) * (size_t)n, stream);
    cudaMemsetAsync(d_parent, 0xFF, sizeof(int) * nodes2, stream);                  // -1: the root (and the only node of a one-triangle scene) has no parent
    int ploc_iterations = 0;
    if (n > 1 && algorithm == BVH8_DEVICE_PLOC)
    {
        int* d_cluster[2] = {arena.alloc<int>(n), arena.alloc<int>(n)};
        float* d_box[2] = {arena.alloc<float>(6 * (size_t)n), arena.alloc<float>(6 * (size_t)n)};
        uint32_t* d_nearest = arena.alloc<uint32_t>(n);
        uint64_t* d_flags = arena.alloc<uint64_t>(n);
        uint64_t* d_scanned = arena.alloc<uint64_t>(n);
        uint32_t* d_subtree = arena.alloc<uint32_t>(nodes2);
        int* d_leaf_parent = arena.alloc<int>(n);
        uint32_t* d_new_position = arena.alloc<uint32_t>(n);
        size_t scan_bytes = 0;
        e = cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_flags, d_scanned, (int)n, stream);
        if (e != cudaSuccess)
            return fail(e);
        void* d_scan_temp = arena.alloc<unsigned char>(scan_bytes);
        if (arena.error != cudaSuccess)
            return fail(arena.error);
        k_ploc_init<<<grid_n, 256, 0, stream>>>(tree, d_prim_lo, d_prim_hi, d_cluster[0], d_box[0], d_subtree);
        uint32_t clusters = n;
        int next_free = (int)n - 1;          // internal nodes n - 2 .. 0
        while (clusters > 1)
        {
            const int in = ploc_iterations & 1, outb = in ^ 1;
            lbvh::Ploc p = {clusters, ploc_radius, d_cluster[in], d_box[in], d_nearest, d_flags, d_scanned, d_cluster[outb], d_box[outb], d_subtree, next_free};
            const unsigned grid_c = (clusters + 255u) / 256u;
            k_ploc_nearest<<<grid_c, 256, 0, stream>>>(p);
            k_ploc_flags<<<grid_c, 256, 0, stream>>>(p);
            e = cub::DeviceScan::ExclusiveSum(d_scan_temp, scan_bytes, d_flags, d_scanned, (int)clusters, stream);
            if (e != cudaSuccess)
                return fail(e);
            k_ploc_merge<<<grid_c, 256, 0, stream>>>(p, tree);
            uint64_t last[2];
            e = cudaMemcpyAsync(&last[0], d_scanned + (clusters - 1u), sizeof(uint64_t), cudaMemcpyDeviceToHost, stream);
            if (e == cudaSuccess)
                e = cudaMemcpyAsync(&last[1], d_flags + (clusters - 1u), sizeof(uint64_t), cudaMemcpyDeviceToHost, stream);
            if (e == cudaSuccess)
                e = cudaStreamSynchronize(stream);
            if (e != cudaSuccess)
                return fail(e);
            const uint64_t total = last[0] + last[1];
            const uint32_t merges = (uint32_t)(total >> 32);
            if (merges == 0u || ++ploc_iterations > 4096)
                return fail(cudaErrorUnknown);          // cannot happen: the cheapest pair is always mutual
            clusters = (uint32_t)(total & 0xFFFFFFFFull);
            next_free -= (int)merges;
        }
        if (next_free != 0)
            return fail(cudaErrorUnknown);
        // leaves in depth-first order: every node covers a contiguous range of the triangle order again
        k_ploc_first_leaf<<<(unsigned)((nodes2 + 255) / 256), 256, 0, stream>>>(tree, d_subtree);
        k_ploc_move_leaves<<<grid_n, 256, 0, stream>>>(tree, d_ids, d_leaf_parent, d_new_position);
        k_ploc_renumber<<<grid_n, 256, 0, stream>>>(tree, d_subtree, d_new_position);
        k_ploc_leaf_parents<<<grid_n, 256, 0, stream>>>(tree, d_leaf_parent);
        tree.prim = d_ids;
    }
    else if (n > 1)
        k_lbvh_internal<<<grid_n, 256, 0, stream>>>(tree);
    k_lbvh_refit<<<grid_n, 256, 0, stream>>>(tree, d_prim_lo, d_prim_hi);

    const uint32_t init_counters[8] = {1u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    cudaMemcpyAsync(d_counters, init_counters, sizeof(init_counters), cudaMemcpyHostToDevice, stream);
    lbvh::Wide wide = {d_nodes, d_triangles, d_counters + 0, d_counters + 1, d_counters + 2, reinterpret_cast<float*>(d_counters + 4), d_vertices, d_indices};
    const lbvh::WorkItem root = {0, 0u};      // binary node 0 = the root (for n == 1: the single leaf, node n - 1 = 0)
    cudaMemcpyAsync(d_queue[0], &root, sizeof(root), cudaMemcpyHostToDevice, stream);
    uint32_t count = 1;
    int depth = 0;
    while (count > 0)
    {
        depth++;
        cudaMemsetAsync(d_counters + 3, 0, sizeof(uint32_t), stream);
        k_lbvh_collapse<<<(count + 127u) / 128u, 128, 0, stream>>>(tree, wide, d_queue[(depth - 1) & 1], count, d_queue[depth & 1], d_counters + 3, d_bounds);
        e = cudaMemcpyAsync(&count, d_counters + 3, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess)
            return fail(e);
        if (depth > 64)
            return fail(cudaErrorUnknown);
    }
    cudaEventRecord(ev1, stream);
    uint32_t counters[8];
    float bounds[6];
    e = cudaMemcpyAsync(counters, d_counters, sizeof(counters), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(bounds, d_bounds, sizeof(bounds), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess)
        return fail(e);
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    // trim: the emission arrays were sized for the worst case (n nodes)
    {
        Bvh8Node* exact_nodes = nullptr;
        if (cudaMalloc(&exact_nodes, sizeof(Bvh8Node) * (size_t)counters[0]) == cudaSuccess)
        {
            e = cudaMemcpyAsync(exact_nodes, d_nodes, sizeof(Bvh8Node) * (size_t)counters[0], cudaMemcpyDeviceToDevice, stream);
            if (e == cudaSuccess)
                e = cudaStreamSynchronize(stream);
            if (e != cudaSuccess)
            {
                cudaFree(exact_nodes);
                return fail(e);
            }
            cudaFree(d_nodes);
            d_nodes = exact_nodes;
        }
        else
            cudaGetLastError();      // keep the large array
    }
    out->nodes = d_nodes;
    out->triangles = d_triangles;
    out->node_count = counters[0];
    out->triangle_count = counters[1];
    out->max_depth = depth;
    out->child_slots_used = counters[2];
    float sah[2];
    memcpy(sah, counters + 4, sizeof(sah));
    out->sah_cost = sah[0] + sah[1];
    out->device_seconds = ms * 1.0e-3;
    out->ploc_iterations = ploc_iterations;
    for (int a = 0; a < 3; a++)
    {
        out->scene_min[a] = bounds[a];
        out->scene_max[a] = bounds[3 + a];
    }
    return out->triangle_count == n ? cudaSuccess : fail(cudaErrorUnknown);
}
} // namespace pt
