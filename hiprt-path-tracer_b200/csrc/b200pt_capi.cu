// Implementation of the C ABI declared in include/b200pt.h (see that header for the reference call sites each
// entry point replaces). No CPU fallback: every compute entry point needs a CUDA device.
#include "bvh8_build.h"
#include "bvh8_build_device.h"
#include "launch.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct b200pt_ctx
{
    int device = 0;
    cudaStream_t main_stream = nullptr;
    pt::VariantOptions options = {3, 0, 3, 0};
    bool counting = false;
    pt::TraversalCounters* d_counters = nullptr;
    int path_mode = pt::WF_PHASED;
    // wavefront scratch, one set per path lane (b200pt_launch_FullPathTracer_lane): lane launches run on concurrent streams
    static constexpr int kMaxPathLanes = 4;
    pt::WavefrontState wf[kMaxPathLanes] = {};
    std::vector<void*> wf_allocations[kMaxPathLanes];
    // scratch of the phased ReSTIR DI passes (launch.h: RestirScratch), sized for the largest tile launched so far
    pt::RestirScratch restir_scratch{};
    bool restir_phases = true;          // b200pt_set_restir_phases
    std::vector<void*> restir_allocations;
    // per-stage timing of FullPathTracer (b200pt_path_profile_enable / _read)
    bool profiling = false;
    std::vector<cudaEvent_t> profile_events;
    std::vector<int> profile_stages;
    pt::WavefrontProfile profile = {};
    double stage_ms[4] = {0.0, 0.0, 0.0, 0.0};
};

struct b200pt_bvh
{
    pt::Bvh8Device view;
    pt::Bvh8Device* d_view = nullptr;
    void* d_nodes = nullptr;
    void* d_triangles = nullptr;
    void* d_shading = nullptr;      // b200pt_bvh_attach_shading_data: 64 bytes per triangle slot
    float sah_cost = 0.0f;
    double build_seconds = 0.0;
    int max_depth = 0;
    float node_fill = 0.0f;
    bool built_on_device = false;
    double device_build_seconds = 0.0;
};

namespace pt { int g_trace_flavour = -1; int g_ray_bin_bits = -1; }

namespace
{
thread_local std::string g_last_error;

int fail(int code, const std::string& message)
{
    g_last_error = message;
    return code;
}

int cuda_fail(cudaError_t e, const char* what)
{
    int code = e == cudaErrorMemoryAllocation ? B200PT_ERROR_OUT_OF_MEMORY : (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? B200PT_ERROR_NO_DEVICE : B200PT_ERROR_CUDA;
    return fail(code, std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}

#define CUDA_TRY(expr)                         \
    do                                         \
    {                                          \
        cudaError_t e_ = (expr);               \
        if (e_ != cudaSuccess)                 \
            return cuda_fail(e_, #expr);       \
    } while (0)

#define REQUIRE(cond)                                                            \
    do                                                                           \
    {                                                                            \
        if (!(cond))                                                             \
            return fail(B200PT_ERROR_INVALID_ARGUMENT, "invalid argument: " #cond); \
    } while (0)

cudaStream_t stream_of(b200pt_ctx* ctx, b200pt_stream s) { return s ? reinterpret_cast<cudaStream_t>(s) : ctx->main_stream; }

pt::Tile tile_of(const b200pt_tile* tile, b200pt_int2 res)
{
    pt::Tile t = {0, 0, res.x, res.y};
    if (tile)
    {
        t.x0 = tile->x0 < 0 ? 0 : tile->x0;
        t.y0 = tile->y0 < 0 ? 0 : tile->y0;
        t.x1 = tile->x1 > res.x ? res.x : tile->x1;
        t.y1 = tile->y1 > res.y ? res.y : tile->y1;
    }
    return t;
}
} // namespace

extern "C" {

int b200pt_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char* b200pt_last_error(void) { return g_last_error.c_str(); }
const char* b200pt_version(void) { return "b200pt 0.2 (sm_100a; BVH8 ray-queue traversal; phased wavefront path tracer; ReSTIR DI)"; }

int b200pt_ctx_create(int device_index, b200pt_ctx** out_ctx)
{
    REQUIRE(out_ctx != nullptr);
    *out_ctx = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
    {
        cudaGetLastError();
        return fail(B200PT_ERROR_NO_DEVICE, "no CUDA device available: the b200pt hot path has no CPU fallback");
    }
    REQUIRE(device_index >= 0 && device_index < n);
    CUDA_TRY(cudaSetDevice(device_index));
    b200pt_ctx* ctx = new b200pt_ctx();
    ctx->device = device_index;
    e = cudaStreamCreateWithFlags(&ctx->main_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess)
    {
        delete ctx;
        return cuda_fail(e, "cudaStreamCreate");
    }
    e = cudaMalloc(&ctx->d_counters, sizeof(pt::TraversalCounters));
    if (e == cudaSuccess)
        e = cudaMemset(ctx->d_counters, 0, sizeof(pt::TraversalCounters));
    if (e != cudaSuccess)
    {
        cudaStreamDestroy(ctx->main_stream);
        delete ctx;
        return cuda_fail(e, "cudaMalloc(counters)");
    }
    // The megakernel keeps per-thread traversal stacks and a 256-byte material in local memory
    cudaDeviceSetLimit(cudaLimitStackSize, 8192);
    // B200PT_PATH_MODE=megakernel|fused|split overrides the default shape of FullPathTracer (A/B measurements)
    if (const char* m = getenv("B200PT_PATH_MODE"))
        ctx->path_mode = !strcmp(m, "megakernel") ? pt::WF_MEGAKERNEL : !strcmp(m, "fused") ? pt::WF_FUSED : !strcmp(m, "split") ? pt::WF_SPLIT : pt::WF_PHASED;
    *out_ctx = ctx;
    return B200PT_SUCCESS;
}

int b200pt_ctx_destroy(b200pt_ctx* ctx)
{
    if (!ctx)
        return B200PT_SUCCESS;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->main_stream);
    cudaFree(ctx->d_counters);
    for (auto& allocations : ctx->wf_allocations)
        for (void* p : allocations)
            cudaFree(p);
    for (void* p : ctx->restir_allocations)
        cudaFree(p);
    for (cudaEvent_t e : ctx->profile_events)
        cudaEventDestroy(e);
    cudaStreamDestroy(ctx->main_stream);
    delete ctx;
    return B200PT_SUCCESS;
}

int b200pt_ctx_make_current(b200pt_ctx* ctx)
{
    REQUIRE(ctx != nullptr);
    CUDA_TRY(cudaSetDevice(ctx->device));
    return B200PT_SUCCESS;
}

int b200pt_malloc(b200pt_ctx* ctx, size_t bytes, void** out_dptr)
{
    REQUIRE(ctx != nullptr && out_dptr != nullptr);
    *out_dptr = nullptr;
    if (bytes == 0)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaMalloc(out_dptr, bytes));
    return B200PT_SUCCESS;
}

int b200pt_free(b200pt_ctx* ctx, void* dptr)
{
    REQUIRE(ctx != nullptr);
    if (!dptr)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaFree(dptr));
    return B200PT_SUCCESS;
}

int b200pt_memset(b200pt_ctx* ctx, void* dptr, int value, size_t bytes, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && (dptr != nullptr || bytes == 0));
    if (bytes == 0)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaMemsetAsync(dptr, value, bytes, stream_of(ctx, stream)));
    return B200PT_SUCCESS;
}

// Headless output: CPURenderer::tonemap (src/Renderer/CPURenderer.cpp:663-686) = Utils::tonemap_hdr_image
// (src/Utils/Utils.cpp:35-52): 1 - exp(-hdr / samples * exposure), then pow(., 1 / gamma); the 8-bit output is the quantisation
// of Image32Bit::write_image_png (src/Image/Image.cpp:427-437). One thread per pixel, 12 B in, 12 B and/or 3 B out; flip_y
// writes row H-1-y (stbi_flip_vertically_on_write). Store-bound: nothing to tune beyond coalescing.
__global__ void __launch_bounds__(256) k_tonemap(const float* __restrict__ hdr, int width, int height, float samples, float exposure, float inv_gamma, int flip_y,
                                                 float* __restrict__ out_float, unsigned char* __restrict__ out_u8)
{
    const size_t n = (size_t)width * (size_t)height;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    {
        const size_t y = i / (size_t)width, x = i - y * (size_t)width;
        const size_t o = flip_y ? ((size_t)height - 1 - y) * (size_t)width + x : i;
#pragma unroll
        for (int k = 0; k < 3; k++)
        {
            const float c = hdr[i * 3 + k] / samples;      // a division, as in the reference: x * (1 / n) differs in the last bit
            float v = 1.0f - expf(-c * exposure);
            v = powf(v, inv_gamma);
            if (out_float)
                out_float[o * 3 + k] = v;
            if (out_u8)
                out_u8[o * 3 + k] = (unsigned char)fminf(fmaxf(v * 255.0f, 0.0f), 255.0f);
        }
    }
}

int b200pt_tonemap(b200pt_ctx* ctx, const float* d_hdr_rgb, int width, int height, int sample_number, float gamma, float exposure, int flip_y, float* d_out_rgb_float,
                   unsigned char* d_out_rgb8, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && d_hdr_rgb != nullptr && width > 0 && height > 0 && gamma > 0.0f && (d_out_rgb_float != nullptr || d_out_rgb8 != nullptr));
    REQUIRE(d_out_rgb_float != d_hdr_rgb || !flip_y);      // in place is allowed (CPURenderer::tonemap is), but not while flipping
    const float samples = sample_number > 0 ? (float)sample_number : 1.0f;       // sample_number <= 0: not accumulating, no scaling
    const size_t n = (size_t)width * (size_t)height;
    const unsigned blocks = (unsigned)((n + 255) / 256 < 148u * 16u ? (n + 255) / 256 : 148u * 16u);
    k_tonemap<<<blocks, 256, 0, stream_of(ctx, stream)>>>(d_hdr_rgb, width, height, samples, exposure, 1.0f / gamma, flip_y, d_out_rgb_float, d_out_rgb8);
    CUDA_TRY(cudaGetLastError());
    return B200PT_SUCCESS;
}

int b200pt_memcpy_h2d(b200pt_ctx* ctx, void* dst_dptr, const void* src_host, size_t bytes, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && (bytes == 0 || (dst_dptr != nullptr && src_host != nullptr)));
    if (bytes == 0)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaMemcpyAsync(dst_dptr, src_host, bytes, cudaMemcpyHostToDevice, stream_of(ctx, stream)));
    return B200PT_SUCCESS;
}

int b200pt_memcpy_d2h(b200pt_ctx* ctx, void* dst_host, const void* src_dptr, size_t bytes, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && (bytes == 0 || (dst_host != nullptr && src_dptr != nullptr)));
    if (bytes == 0)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaMemcpyAsync(dst_host, src_dptr, bytes, cudaMemcpyDeviceToHost, stream_of(ctx, stream)));
    return B200PT_SUCCESS;
}

int b200pt_memcpy_d2d(b200pt_ctx* ctx, void* dst_dptr, const void* src_dptr, size_t bytes, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && (bytes == 0 || (dst_dptr != nullptr && src_dptr != nullptr)));
    if (bytes == 0)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaMemcpyAsync(dst_dptr, src_dptr, bytes, cudaMemcpyDeviceToDevice, stream_of(ctx, stream)));
    return B200PT_SUCCESS;
}

int b200pt_host_alloc_pinned(b200pt_ctx* ctx, size_t bytes, void** out_host)
{
    REQUIRE(ctx != nullptr && out_host != nullptr);
    CUDA_TRY(cudaMallocHost(out_host, bytes));
    return B200PT_SUCCESS;
}

int b200pt_host_free_pinned(b200pt_ctx* ctx, void* host)
{
    REQUIRE(ctx != nullptr);
    if (host)
        CUDA_TRY(cudaFreeHost(host));
    return B200PT_SUCCESS;
}

int b200pt_stream_create(b200pt_ctx* ctx, b200pt_stream* out_stream)
{
    REQUIRE(ctx != nullptr && out_stream != nullptr);
    cudaStream_t s;
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *out_stream = reinterpret_cast<b200pt_stream>(s);
    return B200PT_SUCCESS;
}

int b200pt_stream_destroy(b200pt_ctx* ctx, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr);
    if (stream)
        CUDA_TRY(cudaStreamDestroy(reinterpret_cast<cudaStream_t>(stream)));
    return B200PT_SUCCESS;
}

int b200pt_stream_query(b200pt_ctx* ctx, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr);
    cudaError_t e = cudaStreamQuery(stream_of(ctx, stream));
    if (e == cudaSuccess)
        return B200PT_SUCCESS;
    if (e == cudaErrorNotReady)
        return B200PT_ERROR_NOT_READY;
    return cuda_fail(e, "cudaStreamQuery");
}

int b200pt_stream_sync(b200pt_ctx* ctx, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr);
    CUDA_TRY(cudaStreamSynchronize(stream_of(ctx, stream)));
    return B200PT_SUCCESS;
}

int b200pt_event_create(b200pt_ctx* ctx, b200pt_event* out_event)
{
    REQUIRE(ctx != nullptr && out_event != nullptr);
    cudaEvent_t ev;
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaEventCreate(&ev));
    *out_event = reinterpret_cast<b200pt_event>(ev);
    return B200PT_SUCCESS;
}

int b200pt_event_destroy(b200pt_ctx* ctx, b200pt_event event)
{
    REQUIRE(ctx != nullptr);
    if (event)
        CUDA_TRY(cudaEventDestroy(reinterpret_cast<cudaEvent_t>(event)));
    return B200PT_SUCCESS;
}

int b200pt_event_record(b200pt_ctx* ctx, b200pt_event event, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && event != nullptr);
    CUDA_TRY(cudaEventRecord(reinterpret_cast<cudaEvent_t>(event), stream_of(ctx, stream)));
    return B200PT_SUCCESS;
}

int b200pt_stream_wait_event(b200pt_ctx* ctx, b200pt_stream stream, b200pt_event event)
{
    REQUIRE(ctx != nullptr && event != nullptr);
    CUDA_TRY(cudaStreamWaitEvent(stream_of(ctx, stream), reinterpret_cast<cudaEvent_t>(event), 0));
    return B200PT_SUCCESS;
}

int b200pt_event_elapsed_ms(b200pt_ctx* ctx, b200pt_event start, b200pt_event stop, float* out_ms)
{
    REQUIRE(ctx != nullptr && start != nullptr && stop != nullptr && out_ms != nullptr);
    CUDA_TRY(cudaEventSynchronize(reinterpret_cast<cudaEvent_t>(stop)));
    CUDA_TRY(cudaEventElapsedTime(out_ms, reinterpret_cast<cudaEvent_t>(start), reinterpret_cast<cudaEvent_t>(stop)));
    return B200PT_SUCCESS;
}

} // extern "C"

#include "b200pt_capi_scene.inl"
#include "b200pt_capi_noalloc.inl"
#include "b200pt_capi_comm.inl"
#include "b200pt_capi_bake.inl"
