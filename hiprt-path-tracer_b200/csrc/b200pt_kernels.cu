// __global__ entry points and launchers shared by every permutation: CameraRays, ray-batch queries, dispatch of the
// per-permutation FullPathTracer translation units. Compiled twice: timing build (suffix _t) and, with -DPT_COUNTERS,
// counting build (suffix _c) that accumulates rays / nodes / triangles for Mrays/s and the node-fetch roofline.
#include "kernels.cuh"
#include "trace_queue.cuh"
#include "launch.h"
#include "variants.h"

namespace pt
{
template <int LSS>
__global__ void __launch_bounds__(64) PT_SUFFIXED(k_CameraRays)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    camera_rays_pixel<LSS == 5>(rd, res, x, y);
}

__global__ void PT_SUFFIXED(k_trace_closest)(Bvh8Device bvh, const float* __restrict__ rays, const int* __restrict__ last_prim, long long n, int* out_prim, float* out_t, float* out_uv, TraversalCounters* counters)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    float3 o = f3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
    float3 d = f3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
    Hit hit;
    bvh8_trace<false>(bvh, o, d, INFINITY, last_prim ? last_prim[i] : -1, hit, counters);
    out_prim[i] = hit.prim;
    out_t[i] = hit.t;
    if (out_uv)
    {
        out_uv[i * 2 + 0] = hit.u;
        out_uv[i * 2 + 1] = hit.v;
    }
}

__global__ void PT_SUFFIXED(k_trace_shadow)(Bvh8Device bvh, const float* __restrict__ rays, const float* __restrict__ t_max, const int* __restrict__ last_prim, long long n, unsigned char* out_occluded, TraversalCounters* counters)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    float3 o = f3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
    float3 d = f3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
    Hit hit;
    out_occluded[i] = bvh8_trace<true>(bvh, o, d, t_max[i] - 1.0e-4f, last_prim ? last_prim[i] : -1, hit, counters) ? 1 : 0;
}

// FLAVOUR 0: the classic loop (load, wait, test) at kTraceBlocksPerSM resident blocks; FLAVOUR 5 / 6 / 7: the pipelined loop
// (trace_queue_body_pipelined) compiled for that many resident blocks per SM (95 / 80 / 72 registers: the loads in flight stay
// live across the box tests, so fewer warps fit, each with more memory-level parallelism); FLAVOUR 2: the classic loop WITH the
// distance cull at pop time (trace_queue.cuh: PT_TRACE_POP_CULL; measured slower, kept for A/B). B200PT_TRACE_PIPELINE selects at run time.
template <bool ANY_HIT, int FLAVOUR>
__global__ void __launch_bounds__(kTraceBlock, FLAVOUR >= 5 ? FLAVOUR : kTraceBlocksPerSM) PT_SUFFIXED(k_trace_queue)(const Bvh8Device* __restrict__ bvh, const __grid_constant__ RayQueueView q,
                                                                                                              TraversalCounters* counters, unsigned refill_idle_lanes)
{
    __shared__ uint2 shared_stack[kSharedStackEntries * kTraceBlock];
    const Bvh8Device b = *bvh;
    if (FLAVOUR >= 5)
        trace_queue_body_pipelined<ANY_HIT>(b, q, shared_stack, counters, refill_idle_lanes);
    else if (FLAVOUR == 2)
        trace_queue_body<ANY_HIT, true>(b, q, shared_stack, counters, refill_idle_lanes);
    else
        trace_queue_body<ANY_HIT>(b, q, shared_stack, counters, refill_idle_lanes);
}

// Two independent queues in one launch: every warp drains the any-hit queue, then the closest-hit queue. A launch of this kernel
// cannot end before its longest ray has walked its dependent node fetches (~40 us on B200 whatever the queue length,
// profiles/r01_launches_c2_bench_profile.txt): the NEE shadow rays and the NEE light rays of a bounce do not depend on each
// other, so the second queue's rays fill the lanes the first queue's stragglers leave idle instead of paying that tail twice.
template <int FLAVOUR>
__global__ void __launch_bounds__(kTraceBlock, FLAVOUR >= 5 ? FLAVOUR : kTraceBlocksPerSM) PT_SUFFIXED(k_trace_queue_pair)(const Bvh8Device* __restrict__ bvh, const __grid_constant__ RayQueueView q_any,
                                                                                                                   const __grid_constant__ RayQueueView q_closest, TraversalCounters* counters,
                                                                                                                   unsigned refill_idle_lanes)
{
    __shared__ uint2 shared_stack[kSharedStackEntries * kTraceBlock];
    const Bvh8Device b = *bvh;
    if (FLAVOUR >= 5)
    {
        trace_queue_body_pipelined<true>(b, q_any, shared_stack, counters, refill_idle_lanes);
        trace_queue_body_pipelined<false>(b, q_closest, shared_stack, counters, refill_idle_lanes);
    }
    else if (FLAVOUR == 2)
    {
        trace_queue_body<true, true>(b, q_any, shared_stack, counters, refill_idle_lanes);
        trace_queue_body<false, true>(b, q_closest, shared_stack, counters, refill_idle_lanes);
    }
    else
    {
        trace_queue_body<true>(b, q_any, shared_stack, counters, refill_idle_lanes);
        trace_queue_body<false>(b, q_closest, shared_stack, counters, refill_idle_lanes);
    }
}

// scheduling knobs of the kernel (trace_queue.cuh), overridable for sweeps (tools/trace_sweep.sh); results do not depend on them
static unsigned trace_knob_refill() { static const unsigned v = [] { const char* e = getenv("B200PT_TRACE_REFILL_IDLE"); int n = e ? atoi(e) : 0; return n > 0 && n <= 32 ? (unsigned)n : kRefillIdleLanes; }(); return v; }
static int trace_knob_flavour()
{
    static const int from_env = [] { const char* e = getenv("B200PT_TRACE_PIPELINE"); return e ? atoi(e) : PT_TRACE_PIPELINED_DEFAULT; }();
    const int n = g_trace_flavour >= 0 ? g_trace_flavour : from_env;
    return n == 5 || n == 6 || n == 7 || n == 2 ? n : (n == 1 ? 6 : 0);
}
static unsigned trace_knob_grid(unsigned blocks)
{
    static const unsigned pct = [] { const char* v = getenv("B200PT_TRACE_GRID_PCT"); int n = v ? atoi(v) : 0; return n > 0 ? (unsigned)n : 100u; }();
    // `blocks` is sized for kTraceBlocksPerSM resident blocks per SM: a flavour with fewer resident blocks starts fewer
    const int flavour = trace_knob_flavour();
    const unsigned resident = flavour >= 5 ? blocks * (unsigned)flavour / (unsigned)kTraceBlocksPerSM : blocks;
    const unsigned scaled = resident * pct / 100u;
    return scaled > 0u ? scaled : 1u;
}

cudaError_t PT_SUFFIXED(launch_trace_queue_pair)(const void* bvh_device_handle, const RayQueueView& q_any, const RayQueueView& q_closest, unsigned blocks, TraversalCounters* counters, cudaStream_t stream)
{
    const Bvh8Device* bvh = reinterpret_cast<const Bvh8Device*>(bvh_device_handle);
    blocks = trace_knob_grid(blocks);
    const unsigned refill = trace_knob_refill();
    switch (trace_knob_flavour())
    {
    case 5: PT_SUFFIXED(k_trace_queue_pair)<5><<<blocks, kTraceBlock, 0, stream>>>(bvh, q_any, q_closest, counters, refill); break;
    case 6: PT_SUFFIXED(k_trace_queue_pair)<6><<<blocks, kTraceBlock, 0, stream>>>(bvh, q_any, q_closest, counters, refill); break;
    case 7: PT_SUFFIXED(k_trace_queue_pair)<7><<<blocks, kTraceBlock, 0, stream>>>(bvh, q_any, q_closest, counters, refill); break;
    case 2: PT_SUFFIXED(k_trace_queue_pair)<2><<<blocks, kTraceBlock, 0, stream>>>(bvh, q_any, q_closest, counters, refill); break;
    default: PT_SUFFIXED(k_trace_queue_pair)<0><<<blocks, kTraceBlock, 0, stream>>>(bvh, q_any, q_closest, counters, refill); break;
    }
    return cudaGetLastError();
}

template <bool ANY_HIT>
static void launch_trace_queue_flavour(const Bvh8Device* bvh, const RayQueueView& q, unsigned blocks, TraversalCounters* counters, unsigned refill, cudaStream_t stream)
{
    switch (trace_knob_flavour())
    {
    case 5: PT_SUFFIXED(k_trace_queue)<ANY_HIT, 5><<<blocks, kTraceBlock, 0, stream>>>(bvh, q, counters, refill); break;
    case 6: PT_SUFFIXED(k_trace_queue)<ANY_HIT, 6><<<blocks, kTraceBlock, 0, stream>>>(bvh, q, counters, refill); break;
    case 7: PT_SUFFIXED(k_trace_queue)<ANY_HIT, 7><<<blocks, kTraceBlock, 0, stream>>>(bvh, q, counters, refill); break;
    case 2: PT_SUFFIXED(k_trace_queue)<ANY_HIT, 2><<<blocks, kTraceBlock, 0, stream>>>(bvh, q, counters, refill); break;
    default: PT_SUFFIXED(k_trace_queue)<ANY_HIT, 0><<<blocks, kTraceBlock, 0, stream>>>(bvh, q, counters, refill); break;
    }
}

cudaError_t PT_SUFFIXED(launch_trace_queue)(const void* bvh_device_handle, const RayQueueView& q, bool any_hit, unsigned blocks, TraversalCounters* counters, cudaStream_t stream)
{
    const Bvh8Device* bvh = reinterpret_cast<const Bvh8Device*>(bvh_device_handle);
    blocks = trace_knob_grid(blocks);
    if (any_hit)
        launch_trace_queue_flavour<true>(bvh, q, blocks, counters, trace_knob_refill(), stream);
    else
        launch_trace_queue_flavour<false>(bvh, q, blocks, counters, trace_knob_refill(), stream);
    return cudaGetLastError();
}

// ---- ray binning: counting sort of a ray queue by (origin cell in the scene's bounding box, direction octant) -----------------------
// Why: rays of bounce >= 1 reach the traversal kernel in path-slot order grouped by material signature, i.e. spatially random; ncu on
// the 10 M-triangle scene (profiles/r01_trace_queue_closest_c3_v2.txt) shows the kernel waiting on dependent node / triangle fetches
// (long_scoreboard 6.5 warps per issue, L1 hit 52 %, 19.5 of 32 lanes per instruction). Rays that start in the same cell and head into
// the same octant walk the same upper nodes and neighbouring leaves: a warp's 32 rays then share cache lines and branches. Only the
// ORDER in which rays are traced changes; each ray's hit record is the same (tests: bit-identical frames with binning on and off).
// Three small kernels per queue: histogram, exclusive scan (one block), scatter with per-bin atomic cursors.
__device__ __forceinline__ uint32_t spread_bits_3(uint32_t v)     // abcdef -> a00b00c00d00e00f
{
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

__device__ __forceinline__ uint32_t ray_bin_key(const Bvh8Device& b, const float4 o, const float4 d, const int bits)
{
    const float cells = (float)(1 << bits);
    const int top = (1 << bits) - 1;
    const float ex = b.scene_max[0] - b.scene_min[0], ey = b.scene_max[1] - b.scene_min[1], ez = b.scene_max[2] - b.scene_min[2];
    const int cx = min(max((int)((o.x - b.scene_min[0]) * (cells / fmaxf(ex, 1.0e-20f))), 0), top);
    const int cy = min(max((int)((o.y - b.scene_min[1]) * (cells / fmaxf(ey, 1.0e-20f))), 0), top);
    const int cz = min(max((int)((o.z - b.scene_min[2]) * (cells / fmaxf(ez, 1.0e-20f))), 0), top);
    const uint32_t cell = spread_bits_3((uint32_t)cx) | (spread_bits_3((uint32_t)cy) << 1) | (spread_bits_3((uint32_t)cz) << 2);
    const uint32_t octant = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    return (cell << 3) | octant;
}

__global__ void __launch_bounds__(256) PT_SUFFIXED(k_bin_count)(const Bvh8Device* __restrict__ bvh, const __grid_constant__ RayQueueView q, uint32_t* hist, uint32_t* keys, int bits)
{
    const Bvh8Device b = *bvh;
    const uint32_t n = *q.count;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const uint32_t id = q.ids[i];
        const size_t at = (size_t)id * (q.ray_stride_minus_1 + 1u);
        const uint32_t key = ray_bin_key(b, q.ray_o[at], q.ray_d[at], bits);
        keys[i] = key;
        // one atomic per distinct key of the warp: neighbouring queue entries often share a cell, and a popular cell's counter would
        // otherwise serialise the whole grid (first version: 2 M rays of the Cornell box in a few hundred cells, +4 ms)
        const unsigned peers = __match_any_sync(__activemask(), key);
        if ((threadIdx.x & 31u) == (uint32_t)(__ffs((int)peers) - 1))
            atomicAdd(hist + key, (uint32_t)__popc(peers));
    }
}

// Exclusive scan of `bins` counters in place by ONE block: per-thread chunk sums, block scan of the sums, chunk rewrite
__global__ void __launch_bounds__(1024) PT_SUFFIXED(k_bin_scan)(uint32_t* hist, uint32_t bins)
{
    __shared__ uint32_t partial[1024];
    const uint32_t t = threadIdx.x;
    const uint32_t chunk = (bins + 1023u) / 1024u;
    const uint32_t begin = min(t * chunk, bins), end = min(begin + chunk, bins);
    uint32_t sum = 0;
    for (uint32_t i = begin; i < end; i++)
        sum += hist[i];
    partial[t] = sum;
    __syncthreads();
    for (uint32_t offset = 1; offset < 1024u; offset <<= 1)
    {
        const uint32_t add = t >= offset ? partial[t - offset] : 0u;
        __syncthreads();
        partial[t] += add;
        __syncthreads();
    }
    uint32_t running = partial[t] - sum;
    for (uint32_t i = begin; i < end; i++)
    {
        const uint32_t c = hist[i];
        hist[i] = running;
        running += c;
    }
}

__global__ void __launch_bounds__(256) PT_SUFFIXED(k_bin_scatter)(const __grid_constant__ RayQueueView q, uint32_t* hist, const uint32_t* __restrict__ keys, uint32_t* out_ids)
{
    const uint32_t n = *q.count;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const uint32_t id = q.ids[i];
        const uint32_t key = keys[i];
        const unsigned peers = __match_any_sync(__activemask(), key);
        const uint32_t lane = threadIdx.x & 31u;
        const int leader = __ffs((int)peers) - 1;
        uint32_t base = 0;
        if ((int)lane == leader)
            base = atomicAdd(hist + key, (uint32_t)__popc(peers));
        base = __shfl_sync(peers, base, leader);
        out_ids[base + (uint32_t)__popc(peers & ((1u << lane) - 1u))] = id;
    }
}

// Regroups the queue: on return q.ids is out_ids (the same ids in bin order); count and cursor are untouched.
cudaError_t PT_SUFFIXED(launch_bin_rays)(const void* bvh_device_handle, RayQueueView& q, uint32_t* hist, uint32_t* keys, uint32_t* out_ids, int bits, unsigned sm_count, cudaStream_t stream)
{
    if (bits <= 0 || hist == nullptr || out_ids == nullptr || keys == nullptr)
        return cudaSuccess;
    bits = bits > kRayBinBitsMax ? kRayBinBitsMax : bits;
    const uint32_t bins = 1u << (3 * bits + 3);
    const Bvh8Device* bvh = reinterpret_cast<const Bvh8Device*>(bvh_device_handle);
    cudaError_t e = cudaMemsetAsync(hist, 0, sizeof(uint32_t) * bins, stream);
    if (e != cudaSuccess)
        return e;
    const unsigned grid = sm_count * 8u;
    PT_SUFFIXED(k_bin_count)<<<grid, 256, 0, stream>>>(bvh, q, hist, keys, bits);
    PT_SUFFIXED(k_bin_scan)<<<1, 1024, 0, stream>>>(hist, bins);
    PT_SUFFIXED(k_bin_scatter)<<<grid, 256, 0, stream>>>(q, hist, keys, out_ids);
    q.ids = out_ids;
    return cudaGetLastError();
}

// ---- a ray batch through the ray-queue kernel (b200pt_trace_queue): pack -> k_trace_queue -> unpack -------------------------------
__global__ void __launch_bounds__(256) PT_SUFFIXED(k_batch_pack)(const float* __restrict__ rays, const float* __restrict__ t_max, const int* __restrict__ last_prim, uint32_t n,
                                                                float4* ray_o, float4* ray_d, uint32_t* ids, float4* hit_tuvp, uint32_t* count_and_cursor)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0)
    {
        count_and_cursor[0] = n;
        count_and_cursor[1] = 0u;
    }
    if (i >= n)
        return;
    ray_o[i] = make_float4(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2], t_max ? t_max[i] : 0.0f);
    ray_d[i] = make_float4(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5], __int_as_float(last_prim ? last_prim[i] : -1));
    ids[i] = i;
    if (hit_tuvp)
        hit_tuvp[i] = make_float4(-1.0f, 0.0f, 0.0f, __int_as_float(-1));
}

__global__ void __launch_bounds__(256) PT_SUFFIXED(k_batch_unpack)(const float4* __restrict__ ray_o, const float4* __restrict__ hit_tuvp, uint32_t n, int* out_prim, float* out_t, float* out_uv,
                                                                  unsigned char* out_occluded)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    if (out_occluded)
        out_occluded[i] = ray_o[i].w == -1.0f ? 1 : 0;      // the any-hit kernel's flag (trace_queue_store_hit)
    else
    {
        const float4 h = hit_tuvp[i];
        out_prim[i] = __float_as_int(h.w);
        out_t[i] = h.x;
        if (out_uv)
        {
            out_uv[i * 2 + 0] = h.y;
            out_uv[i * 2 + 1] = h.z;
        }
    }
}

cudaError_t PT_SUFFIXED(launch_trace_batch_queue)(const void* bvh_device_handle, const float* rays, const float* t_max, const int* last_prim, long long n, int* out_prim, float* out_t, float* out_uv,
                                                  unsigned char* out_occluded, unsigned sm_count, TraversalCounters* counters, cudaStream_t stream)
{
    const bool any_hit = t_max != nullptr;
    const uint32_t count = (uint32_t)n;
    float4 *ray_o = nullptr, *ray_d = nullptr, *hit_tuvp = nullptr;
    uint32_t *ids = nullptr, *tri_slot = nullptr, *count_and_cursor = nullptr;
    cudaError_t e = cudaMalloc(&ray_o, sizeof(float4) * n);
    if (e == cudaSuccess) e = cudaMalloc(&ray_d, sizeof(float4) * n);
    if (e == cudaSuccess) e = cudaMalloc(&ids, sizeof(uint32_t) * n);
    if (e == cudaSuccess) e = cudaMalloc(&count_and_cursor, sizeof(uint32_t) * 2);
    if (e == cudaSuccess && !any_hit) e = cudaMalloc(&hit_tuvp, sizeof(float4) * n);
    if (e == cudaSuccess && !any_hit) e = cudaMalloc(&tri_slot, sizeof(uint32_t) * n);
    if (e == cudaSuccess)
    {
        const unsigned grid = (count + 255u) / 256u;
        PT_SUFFIXED(k_batch_pack)<<<grid, 256, 0, stream>>>(rays, t_max, last_prim, count, ray_o, ray_d, ids, hit_tuvp, count_and_cursor);
        RayQueueView q = {ids, count_and_cursor, count_and_cursor + 1, ray_o, ray_d, hit_tuvp, tri_slot, any_hit ? ray_o : nullptr};
        e = PT_SUFFIXED(launch_trace_queue)(bvh_device_handle, q, any_hit, sm_count * (unsigned)kTraceGridBlocksPerSM, counters, stream);
        if (e == cudaSuccess)
        {
            PT_SUFFIXED(k_batch_unpack)<<<grid, 256, 0, stream>>>(ray_o, hit_tuvp, count, out_prim, out_t, out_uv, any_hit ? out_occluded : nullptr);
            e = cudaGetLastError();
        }
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
    }
    cudaFree(ray_o); cudaFree(ray_d); cudaFree(ids); cudaFree(count_and_cursor); cudaFree(hit_tuvp); cudaFree(tri_slot);
    return e;
}

static inline dim3 pixel_grid(const Tile& t) { return dim3((t.x1 - t.x0 + 7) / 8, (t.y1 - t.y0 + 7) / 8, 1); }

#define PT_DECLARE_VARIANT(LSS, ESS, LATER, GGX) cudaError_t PT_VARIANT_NAME_OF(launch_FullPathTracer, LSS, ESS, LATER, GGX)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const WavefrontState* wf, int mode, WavefrontProfile* profile, const SampleBatch* batch);
PT_FOR_EACH_VARIANT(PT_DECLARE_VARIANT)

#define PT_DECLARE_RESTIR(LSS, ESS, LATER, GGX)                                                                                                                    \
    cudaError_t PT_VARIANT_NAME_OF(launch_ReSTIR_DI_LightsPresampling, LSS, ESS, LATER, GGX)(const LightPresamplingParameters* p, cudaStream_t stream);              \
    cudaError_t PT_VARIANT_NAME_OF(launch_ReSTIR_DI_InitialCandidates, LSS, ESS, LATER, GGX)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch);    \
    cudaError_t PT_VARIANT_NAME_OF(launch_ReSTIR_DI_SpatiotemporalReuse, LSS, ESS, LATER, GGX)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch); \
    cudaError_t PT_VARIANT_NAME_OF(launch_ReSTIR_DI_TemporalReuse, LSS, ESS, LATER, GGX)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch);        \
    cudaError_t PT_VARIANT_NAME_OF(launch_ReSTIR_DI_SpatialReuse, LSS, ESS, LATER, GGX)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch);
#define PT_DECLARE_RESTIR_IF_5(LSS, ESS, LATER, GGX) PT_DECLARE_RESTIR_##LSS(LSS, ESS, LATER, GGX)
#define PT_DECLARE_RESTIR_3(LSS, ESS, LATER, GGX)
#define PT_DECLARE_RESTIR_4(LSS, ESS, LATER, GGX)
#define PT_DECLARE_RESTIR_5(LSS, ESS, LATER, GGX) PT_DECLARE_RESTIR(LSS, ESS, LATER, GGX)
PT_FOR_EACH_VARIANT(PT_DECLARE_RESTIR_IF_5)

// ReSTIR DI passes exist for the LSS_RESTIR_DI permutations only; other options get cudaErrorInvalidDeviceFunction
#define PT_CASE_RESTIR_3(CALL, LSS, ESS, LATER, GGX)
#define PT_CASE_RESTIR_4(CALL, LSS, ESS, LATER, GGX)
#define PT_CASE_RESTIR_5(CALL, LSS, ESS, LATER, GGX) if (opt.lss == LSS && opt.ess == ESS && opt.later == LATER && opt.ggx == GGX) return CALL;

cudaError_t PT_SUFFIXED(launch_ReSTIR_DI_LightsPresampling)(const VariantOptions& opt, const LightPresamplingParameters* p, cudaStream_t stream)
{
#define PT_CASE_PRESAMPLE(LSS, ESS, LATER, GGX) PT_CASE_RESTIR_##LSS(PT_VARIANT_NAME_OF(launch_ReSTIR_DI_LightsPresampling, LSS, ESS, LATER, GGX)(p, stream), LSS, ESS, LATER, GGX)
    PT_FOR_EACH_VARIANT(PT_CASE_PRESAMPLE)
    return cudaErrorInvalidDeviceFunction;
}

cudaError_t PT_SUFFIXED(launch_ReSTIR_DI_InitialCandidates)(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
#define PT_CASE_INITIAL(LSS, ESS, LATER, GGX) PT_CASE_RESTIR_##LSS(PT_VARIANT_NAME_OF(launch_ReSTIR_DI_InitialCandidates, LSS, ESS, LATER, GGX)(rd, res, tile, stream, scratch), LSS, ESS, LATER, GGX)
    PT_FOR_EACH_VARIANT(PT_CASE_INITIAL)
    return cudaErrorInvalidDeviceFunction;
}

cudaError_t PT_SUFFIXED(launch_ReSTIR_DI_SpatiotemporalReuse)(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
#define PT_CASE_SPATIOTEMPORAL(LSS, ESS, LATER, GGX) PT_CASE_RESTIR_##LSS(PT_VARIANT_NAME_OF(launch_ReSTIR_DI_SpatiotemporalReuse, LSS, ESS, LATER, GGX)(rd, res, tile, stream, scratch), LSS, ESS, LATER, GGX)
    PT_FOR_EACH_VARIANT(PT_CASE_SPATIOTEMPORAL)
    return cudaErrorInvalidDeviceFunction;
}

cudaError_t PT_SUFFIXED(launch_ReSTIR_DI_TemporalReuse)(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
#define PT_CASE_TEMPORAL(LSS, ESS, LATER, GGX) PT_CASE_RESTIR_##LSS(PT_VARIANT_NAME_OF(launch_ReSTIR_DI_TemporalReuse, LSS, ESS, LATER, GGX)(rd, res, tile, stream, scratch), LSS, ESS, LATER, GGX)
    PT_FOR_EACH_VARIANT(PT_CASE_TEMPORAL)
    return cudaErrorInvalidDeviceFunction;
}

cudaError_t PT_SUFFIXED(launch_ReSTIR_DI_SpatialReuse)(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
#define PT_CASE_SPATIAL(LSS, ESS, LATER, GGX) PT_CASE_RESTIR_##LSS(PT_VARIANT_NAME_OF(launch_ReSTIR_DI_SpatialReuse, LSS, ESS, LATER, GGX)(rd, res, tile, stream, scratch), LSS, ESS, LATER, GGX)
    PT_FOR_EACH_VARIANT(PT_CASE_SPATIAL)
    return cudaErrorInvalidDeviceFunction;
}

bool PT_SUFFIXED(variant_available)(const VariantOptions& opt)
{
#define PT_CASE_AVAIL(LSS, ESS, LATER, GGX) if (opt.lss == LSS && opt.ess == ESS && (opt.lss != 5 || opt.later == LATER) && opt.ggx == GGX) return true;
    PT_FOR_EACH_VARIANT(PT_CASE_AVAIL)
    return false;
}

cudaError_t PT_SUFFIXED(launch_CameraRays)(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream)
{
    if (opt.lss == 5)
        PT_SUFFIXED(k_CameraRays)<5><<<pixel_grid(tile), dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
    else
        PT_SUFFIXED(k_CameraRays)<0><<<pixel_grid(tile), dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
    return cudaGetLastError();
}

cudaError_t PT_SUFFIXED(launch_FullPathTracer)(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const WavefrontState* wf, int mode, WavefrontProfile* profile)
{
#define PT_CASE_FPT(LSS, ESS, LATER, GGX) if (opt.lss == LSS && opt.ess == ESS && (opt.lss != 5 || opt.later == LATER) && opt.ggx == GGX) return PT_VARIANT_NAME_OF(launch_FullPathTracer, LSS, ESS, LATER, GGX)(rd, res, tile, stream, wf, mode, profile, nullptr);
    PT_FOR_EACH_VARIANT(PT_CASE_FPT)
    return cudaErrorInvalidDeviceFunction;
}

cudaError_t PT_SUFFIXED(launch_PathTraceSamples)(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const WavefrontState* wf, const SampleBatch* batch, WavefrontProfile* profile)
{
#define PT_CASE_BATCH(LSS, ESS, LATER, GGX) if (opt.lss == LSS && opt.ess == ESS && (opt.lss != 5 || opt.later == LATER) && opt.ggx == GGX) return PT_VARIANT_NAME_OF(launch_FullPathTracer, LSS, ESS, LATER, GGX)(rd, res, tile, stream, wf, WF_PHASED, profile, batch);
    PT_FOR_EACH_VARIANT(PT_CASE_BATCH)
    return cudaErrorInvalidDeviceFunction;
}

cudaError_t PT_SUFFIXED(launch_trace_closest)(const Bvh8Device& bvh, const float* rays, const int* last_prim, long long n, int* out_prim, float* out_t, float* out_uv, TraversalCounters* counters, cudaStream_t stream)
{
    if (n <= 0)
        return cudaSuccess;
    const int block = 128;
    PT_SUFFIXED(k_trace_closest)<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(bvh, rays, last_prim, n, out_prim, out_t, out_uv, counters);
    return cudaGetLastError();
}

cudaError_t PT_SUFFIXED(launch_trace_shadow)(const Bvh8Device& bvh, const float* rays, const float* t_max, const int* last_prim, long long n, unsigned char* out_occluded, TraversalCounters* counters, cudaStream_t stream)
{
    if (n <= 0)
        return cudaSuccess;
    const int block = 128;
    PT_SUFFIXED(k_trace_shadow)<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(bvh, rays, t_max, last_prim, n, out_occluded, counters);
    return cudaGetLastError();
}

#ifndef PT_COUNTERS
// ---- per-function known-answer entry points (b200pt_debug_*): the device functions of the shading path on
// caller-provided inputs, one thread per item, so that a failing image test can be bisected to a function. Layouts follow the oracle's
// ref_bsdf_eval / ref_bsdf_sample / ref_envmap_sample / ref_thin_film_fresnel (oracle/ref_harness_passes.inl).
__global__ void __launch_bounds__(64) k_debug_bsdf_eval(const __grid_constant__ HIPRTRenderData rd, const SimplifiedRendererMaterial* materials, RayVolumeState* volume_states, const float* vecs,
                                                        long long n, float* out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    alignas(16) Material m;
    load_material(m, materials + i);
    RayVolumeState vs;
    copy_volume_state(&vs, volume_states + i);
    const float* v = vecs + i * 9;
    float pdf;
    const RGB f = bsdf_eval(rd, m, vs, f3(v[0], v[1], v[2]), f3(v[3], v[4], v[5]), f3(v[6], v[7], v[8]), pdf);
    copy_volume_state(volume_states + i, &vs);
    out[i * 4 + 0] = f.r; out[i * 4 + 1] = f.g; out[i * 4 + 2] = f.b; out[i * 4 + 3] = pdf;
}

template <int GGX_SAMPLER>
__global__ void __launch_bounds__(64) k_debug_bsdf_sample(const __grid_constant__ HIPRTRenderData rd, const SimplifiedRendererMaterial* materials, RayVolumeState* volume_states, const float* vecs,
                                                          unsigned int* seeds, long long n, float* out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    alignas(16) Material m;
    load_material(m, materials + i);
    RayVolumeState vs;
    copy_volume_state(&vs, volume_states + i);
    const float* v = vecs + i * 9;
    float pdf;
    float3 dir = f3(0.0f, 0.0f, 0.0f);
    Rng rng(seeds[i]);
    const RGB f = bsdf_sample<GGX_SAMPLER>(rd, m, vs, f3(v[0], v[1], v[2]), f3(v[3], v[4], v[5]), f3(v[6], v[7], v[8]), dir, pdf, rng);
    seeds[i] = rng.state;
    copy_volume_state(volume_states + i, &vs);
    out[i * 7 + 0] = f.r; out[i * 7 + 1] = f.g; out[i * 7 + 2] = f.b; out[i * 7 + 3] = pdf;
    out[i * 7 + 4] = dir.x; out[i * 7 + 5] = dir.y; out[i * 7 + 6] = dir.z;
}

__global__ void __launch_bounds__(64) k_debug_envmap_sample(const __grid_constant__ HIPRTRenderData rd, unsigned int* seeds, long long n, float* out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    Rng rng(seeds[i]);
    float3 dir;
    float pdf, pdf2;
    const RGB r = envmap_sample<2>(rd.world_settings, dir, pdf, rng);      // ESS_ALIAS_TABLE
    envmap_eval(rd, dir, pdf2);
    seeds[i] = rng.state;
    out[i * 8 + 0] = r.r; out[i * 8 + 1] = r.g; out[i * 8 + 2] = r.b; out[i * 8 + 3] = pdf;
    out[i * 8 + 4] = dir.x; out[i * 8 + 5] = dir.y; out[i * 8 + 6] = dir.z; out[i * 8 + 7] = pdf2;
}

__global__ void __launch_bounds__(64) k_debug_thin_film_fresnel(const SimplifiedRendererMaterial* materials, const float* ior_and_cosine, long long n, float* out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    alignas(16) Material m;
    load_material(m, materials + i);
    const RGB f = thin_film_fresnel(m, ior_and_cosine[i * 2 + 0], ior_and_cosine[i * 2 + 1]);
    out[i * 3 + 0] = f.r; out[i * 3 + 1] = f.g; out[i * 3 + 2] = f.b;
}

cudaError_t launch_debug_function(int which, const HIPRTRenderData* rd, const void* materials, void* volume_states, const float* vecs, unsigned int* seeds, long long n, int ggx_sampler, float* out,
                                  cudaStream_t stream)
{
    if (n <= 0)
        return cudaSuccess;
    const unsigned grid = (unsigned)((n + 63) / 64);
    const SimplifiedRendererMaterial* mats = reinterpret_cast<const SimplifiedRendererMaterial*>(materials);
    RayVolumeState* vs = reinterpret_cast<RayVolumeState*>(volume_states);
    switch (which)
    {
    case 0: k_debug_bsdf_eval<<<grid, 64, 0, stream>>>(*rd, mats, vs, vecs, n, out); break;
    case 1:
        if (ggx_sampler == 1)
            k_debug_bsdf_sample<1><<<grid, 64, 0, stream>>>(*rd, mats, vs, vecs, seeds, n, out);
        else
            k_debug_bsdf_sample<0><<<grid, 64, 0, stream>>>(*rd, mats, vs, vecs, seeds, n, out);
        break;
    case 2: k_debug_envmap_sample<<<grid, 64, 0, stream>>>(*rd, seeds, n, out); break;
    case 3: k_debug_thin_film_fresnel<<<grid, 64, 0, stream>>>(mats, vecs, n, out); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
#endif
} // namespace pt
