// b200pt_trace_queue_static: the ray-batch query of b200pt_trace_queue with NO allocation of any kind.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it). b200pt_trace_queue owns temporary queue buffers (ray records,
// id list, hit records, work cursor) that it cudaMalloc's and frees per call; here the same scratch lives in ONE file-scope
// `static __device__` array of a fixed maximum size, B200PT_TRACE_QUEUE_STATIC_MAX_RAYS rays x 56 bytes + 256 (about 14 MiB), which
// every GPU that loads libb200pt.so holds until the process exits.
//
//   * In host code the array's name is a host-side stand-in, not a device address: the address is taken once per device, with that
//     device current, by cudaGetSymbolAddress (checked), and handed to the kernels as an ordinary pointer.
//   * Every call shares the array and finds the previous call's contents in it: the pack kernel rewrites everything a call reads (ray
//     records, ids, initial hit records) and zeroes the work cursor, so nothing has to be cleared by the caller.
//   * Calls that overlap on the GPU (different streams, or two host threads on one device) RACE on the array: serialise them.
//   * The launch is asynchronous on the given stream (nothing is freed afterwards, so nothing has to be waited for).
// The traversal itself is the ray-queue kernel's loop (trace_queue.cuh: trace_queue_body, classic flavour), so the outputs equal
// b200pt_trace_queue's bit for bit (tests/test_zz_no_alloc.py).
#include "trace_queue.cuh"

#include <mutex>

namespace pt
{
constexpr unsigned long long kStaticRays = B200PT_TRACE_QUEUE_STATIC_MAX_RAYS;
// layout: [0, 256) count + cursor | ray_o float4[N] | ray_d float4[N] | hit_tuvp float4[N] | ids u32[N] | tri_slot u32[N]
constexpr unsigned long long kStaticScratchBytes = 256ull + kStaticRays * (16ull + 16ull + 16ull + 4ull + 4ull);
static __device__ __align__(256) unsigned char g_trace_static_scratch[kStaticScratchBytes];

struct StaticScratchView
{
    uint32_t* count_and_cursor;
    float4 *ray_o, *ray_d, *hit_tuvp;
    uint32_t *ids, *tri_slot;
};

static StaticScratchView static_scratch_view(unsigned char* base)
{
    StaticScratchView v;
    v.count_and_cursor = reinterpret_cast<uint32_t*>(base);
    v.ray_o = reinterpret_cast<float4*>(base + 256);
    v.ray_d = v.ray_o + kStaticRays;
    v.hit_tuvp = v.ray_d + kStaticRays;
    v.ids = reinterpret_cast<uint32_t*>(v.hit_tuvp + kStaticRays);
    v.tri_slot = v.ids + kStaticRays;
    return v;
}

__global__ void __launch_bounds__(256) k_static_pack(const float* __restrict__ rays, const float* __restrict__ t_max, const int* __restrict__ last_prim, uint32_t n, StaticScratchView s)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0)
    {
        s.count_and_cursor[0] = n;      // queue length
        s.count_and_cursor[1] = 0u;     // work cursor: must start at zero, whatever the previous call left
    }
    if (i >= n)
        return;
    s.ray_o[i] = make_float4(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2], t_max ? t_max[i] : 0.0f);
    s.ray_d[i] = make_float4(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5], __int_as_float(last_prim ? last_prim[i] : -1));
    s.ids[i] = i;
    s.hit_tuvp[i] = make_float4(-1.0f, 0.0f, 0.0f, __int_as_float(-1));
    s.tri_slot[i] = 0xFFFFFFFFu;
}

template <bool ANY_HIT>
__global__ void __launch_bounds__(kTraceBlock, kTraceBlocksPerSM) k_static_trace(const Bvh8Device* __restrict__ bvh, const __grid_constant__ RayQueueView q)
{
    __shared__ uint2 shared_stack[kSharedStackEntries * kTraceBlock];
    const Bvh8Device b = *bvh;
    trace_queue_body<ANY_HIT>(b, q, shared_stack, nullptr, kRefillIdleLanes);
}

__global__ void __launch_bounds__(256) k_static_unpack(StaticScratchView s, uint32_t n, int* out_prim, float* out_t, float* out_uv, unsigned char* out_occluded)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    if (out_occluded)
        out_occluded[i] = s.ray_o[i].w == -1.0f ? 1 : 0;      // the any-hit loop's flag (trace_queue_store_hit)
    else
    {
        const float4 h = s.hit_tuvp[i];
        out_prim[i] = __float_as_int(h.w);
        out_t[i] = h.x;
        if (out_uv)
        {
            out_uv[i * 2 + 0] = h.y;
            out_uv[i * 2 + 1] = h.z;
        }
    }
}

// Device address of the scratch array, per device: cudaGetSymbolAddress with that device current, once.
constexpr int kMaxDevices = 64;
static std::mutex g_static_mutex;
static unsigned char* g_static_address[kMaxDevices] = {};
static int g_static_sms[kMaxDevices] = {};

static cudaError_t static_scratch_address(int device, unsigned char** out, int* out_sms)
{
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_address[device] == nullptr)
    {
        cudaError_t e = cudaSetDevice(device);
        if (e != cudaSuccess)
            return e;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_trace_static_scratch);
        if (e != cudaSuccess)
            return e;
        if (p == nullptr)
            return cudaErrorInvalidSymbol;
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess)
            return e;
        g_static_address[device] = static_cast<unsigned char*>(p);
        g_static_sms[device] = sms;
    }
    *out = g_static_address[device];
    *out_sms = g_static_sms[device];
    return cudaSuccess;
}
} // namespace pt

extern "C" int b200pt_trace_queue_static(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const float* d_t_max, const int* d_last_prim, long long n, int* d_out_prim, float* d_out_t,
                                         float* d_out_uv, unsigned char* d_out_occluded, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && bvh != nullptr && n >= 0);
    if ((unsigned long long)n > pt::kStaticRays)
    {
        char msg[224];
        snprintf(msg, sizeof(msg), "b200pt_trace_queue_static: %lld rays exceed the static scratch of B200PT_TRACE_QUEUE_STATIC_MAX_RAYS = %llu rays; split the batch or call b200pt_trace_queue",
                 n, pt::kStaticRays);
        return fail(B200PT_ERROR_INVALID_ARGUMENT, msg);
    }
    const bool any_hit = d_t_max != nullptr;
    REQUIRE(n == 0 || (d_rays != nullptr && (any_hit ? d_out_occluded != nullptr : (d_out_prim != nullptr && d_out_t != nullptr))));
    REQUIRE(ctx->device >= 0 && ctx->device < pt::kMaxDevices);
    if (n == 0)
        return B200PT_SUCCESS;
    unsigned char* base = nullptr;
    int sms = 0;
    cudaError_t e = pt::static_scratch_address(ctx->device, &base, &sms);
    if (e != cudaSuccess)
        return cuda_fail(e, "cudaGetSymbolAddress(trace-queue static scratch)");
    const pt::StaticScratchView s = pt::static_scratch_view(base);
    const uint32_t count = (uint32_t)n;
    const unsigned grid = (count + 255u) / 256u;
    cudaStream_t st = stream_of(ctx, stream);
    pt::k_static_pack<<<grid, 256, 0, st>>>(d_rays, d_t_max, d_last_prim, count, s);
    pt::RayQueueView q = {s.ids, s.count_and_cursor, s.count_and_cursor + 1, s.ray_o, s.ray_d, s.hit_tuvp, s.tri_slot, any_hit ? s.ray_o : nullptr};
    const unsigned blocks = (unsigned)sms * (unsigned)pt::kTraceBlocksPerSM;
    if (any_hit)
        pt::k_static_trace<true><<<blocks, pt::kTraceBlock, 0, st>>>(bvh->d_view, q);
    else
        pt::k_static_trace<false><<<blocks, pt::kTraceBlock, 0, st>>>(bvh->d_view, q);
    pt::k_static_unpack<<<grid, 256, 0, st>>>(s, count, d_out_prim, d_out_t, d_out_uv, any_hit ? d_out_occluded : nullptr);
    e = cudaGetLastError();
    if (e != cudaSuccess)
        return cuda_fail(e, "b200pt_trace_queue_static");
    return B200PT_SUCCESS;
}
