// Host side of the hot path above the C ABI: the launch loop and buffer ownership of the reference's GPURenderer,
// re-hosted on libb200pt (reference: src/Renderer/GPURenderer.{h,cpp}, src/HIPRT-Orochi/OrochiBuffer.h,
// src/Renderer/RenderPasses/ReSTIRDIRenderPass.cpp). Method names and semantics follow the reference so that the
// parity tests read like its own code; rendering is headless (plain device buffers instead of GL interop buffers).
#pragma once

#include "b200pt.h"

#include <stdexcept>
#include <string>
#include <vector>

namespace b200pt_host
{
// Xorshift32Generator (src/HostDeviceCommon/Xorshift.h:17-65), host copy for the per-launch seed draws
struct Xorshift32Generator
{
    unsigned int seed = 42;
    unsigned int xorshift32()
    {
        unsigned int x = seed;
        x ^= x << 13;
        x ^= x >> 17;
        x ^= x << 5;
        return seed = x;
    }
};

inline void check(int code, const char* what)
{
    if (code != 0)
        throw std::runtime_error(std::string(what) + ": " + b200pt_last_error());
}

// OrochiBuffer<T> (src/HIPRT-Orochi/OrochiBuffer.h:17-173): resize frees and reallocates, contents are lost
template <typename T>
class DeviceBuffer
{
public:
    DeviceBuffer() = default;
    explicit DeviceBuffer(b200pt_ctx* ctx) : m_ctx(ctx) {}
    DeviceBuffer(const DeviceBuffer&) = delete;
    DeviceBuffer& operator=(const DeviceBuffer&) = delete;
    ~DeviceBuffer() { free(); }

    void set_context(b200pt_ctx* ctx) { m_ctx = ctx; }
    void resize(size_t new_element_count)
    {
        free();
        if (new_element_count > 0)
            check(b200pt_malloc(m_ctx, sizeof(T) * new_element_count, reinterpret_cast<void**>(&m_data)), "b200pt_malloc");
        m_count = new_element_count;
    }
    void upload_data(const T* data) { check(b200pt_memcpy_h2d(m_ctx, m_data, data, sizeof(T) * m_count, nullptr), "b200pt_memcpy_h2d"); }
    void upload_data(const std::vector<T>& data)
    {
        if (data.size() != m_count)
            resize(data.size());
        if (m_count > 0)
            upload_data(data.data());
    }
    void memset_zero() { check(b200pt_memset(m_ctx, m_data, 0, sizeof(T) * m_count, nullptr), "b200pt_memset"); }
    std::vector<T> download_data() const
    {
        std::vector<T> out(m_count);
        if (m_count > 0)
        {
            check(b200pt_memcpy_d2h(m_ctx, out.data(), m_data, sizeof(T) * m_count, nullptr), "b200pt_memcpy_d2h");
            check(b200pt_stream_sync(m_ctx, nullptr), "b200pt_stream_sync");
        }
        return out;
    }
    T* get_device_pointer() const { return m_data; }
    size_t get_element_count() const { return m_count; }
    void free()
    {
        if (m_data)
            b200pt_free(m_ctx, m_data);
        m_data = nullptr;
        m_count = 0;
    }

private:
    b200pt_ctx* m_ctx = nullptr;
    T* m_data = nullptr;
    size_t m_count = 0;
};

// struct Scene (src/Scene/SceneParser.h:80-131), geometry + materials only
struct Scene
{
    std::vector<RendererMaterial> materials;
    std::vector<int> triangle_indices;
    std::vector<b200pt_float3> vertices_positions;
    std::vector<unsigned char> has_vertex_normals;
    std::vector<b200pt_float3> vertex_normals;
    std::vector<b200pt_float2> texcoords;
    std::vector<int> emissive_triangle_indices;
    std::vector<int> material_indices;

    struct Texture8 { std::vector<unsigned char> texels; int width = 0, height = 0, channels = 0; };
    std::vector<Texture8> textures;
};

struct GBufferStorage
{
    DeviceBuffer<SimplifiedRendererMaterial> materials;
    DeviceBuffer<b200pt_float3> shading_normals, geometric_normals, view_directions, first_hits;
    DeviceBuffer<int> first_hit_prim_index;
    DeviceBuffer<unsigned char> camera_ray_hit;
    DeviceBuffer<RayVolumeState> ray_volume_states;

    void set_context(b200pt_ctx* ctx);
    void resize(size_t n);
    GBuffer view() const;
};

enum class SeedSchedule
{
    GPU,   // a fresh seed before every kernel launch (src/Renderer/GPURenderer.cpp:466-486)
    CPU    // CPURenderer's schedule: seed 42 on the first sample, one draw after each sample, shared by the passes of a sample
           // (src/Renderer/CPURenderer.cpp:264-296) -- the mode parity runs use
};

void compute_envmap_alias_table(const float* texels, int width, int height, int channels, std::vector<float>& probas, std::vector<int>& alias, float& luminance_total_sum);

class GPURenderer
{
public:
    explicit GPURenderer(b200pt_ctx* ctx);
    ~GPURenderer();

    // BRDFsData LUTs (CPURenderer::setup_brdfs_data, src/Renderer/CPURenderer.cpp:93-132): float arrays, first channel only
    void set_brdf_luts(const float* sheen_ltc_32x32x3, const float* conductor_128x128, const float* glossy_128x64x128,
                       const float* glass_256x16x128, const float* glass_inverse_256x16x128, const float* thin_glass_32x32x96);
    void set_scene(const Scene& scene);                                               // GPURenderer.cpp:1041-1134
    void set_envmap(const float* texels, int width, int height, int channels,
                    const float* alias_probas, const int* alias_alias, float total_sum);   // GPURenderer.cpp:1136-1174
    void set_camera(const HIPRTCamera& camera);
    void set_kernel_options(const b200pt_options& options);                          // GPUKernelCompilerOptions -> prebuilt permutation
    void resize(int width, int height);                                               // GPURenderer.cpp:resize
    void set_tile(const b200pt_tile* tile);                                           // multi-GPU image shard (new)
    void set_tile_halo(int rows);
    // FullPathTracer as `lanes` (1..4) concurrent band launches on separate streams (b200pt_launch_FullPathTracer_lane): the
    // traversal kernels of one band run under the DRAM-bound shade phases of another. Bands are row tiles, so pixels are unchanged.
    void set_path_lanes(int lanes);
    // Sample batches (b200pt_launch_PathTraceSamples): render() runs the samples_per_frame passes of a frame as fused wavefronts of up
    // to `samples_in_flight` passes (CameraRays fused in) whenever the options allow it; 0 keeps the two launch points per pass.
    int samples_in_flight = 16;
    bool renders_sample_batches() const { return can_render_sample_batches(); }
    bool keep_g_buffer = true;      // the last pass of a batch leaves its camera hits in the G-buffer, as CameraRays would
    int path_lanes() const { return m_path_lanes; }                                                     // rows either side of the tile CameraRays also covers (ReSTIR DI)
    void reset();

    void update();                                                                    // GPURenderer.cpp:236-267
    void render();                                                                    // GPURenderer.cpp:408-464
    void synchronize_kernel();

    HIPRTRenderSettings& get_render_settings() { return m_render_data.render_settings; }
    WorldSettings& get_world_settings() { return m_render_data.world_settings; }
    BRDFsData& get_bsdfs_data() { return m_render_data.bsdfs_data; }
    HIPRTRenderData& get_render_data() { return m_render_data; }
    SeedSchedule seed_schedule = SeedSchedule::GPU;
    // Seed of the host generator the per-launch random_seed values are drawn from (the reference's m_rng starts at 42,
    // src/Renderer/GPURenderer.h). Multi-GPU sample-pass sharding gives every rank its own stream: the ranks then render
    // independent samples with the SAME sample_number, which adaptive sampling's thresholds are expressed in.
    unsigned int rng_seed = 42;

    std::vector<ColorRGB32F> download_framebuffer() const { return m_pixels.download_data(); }
    const DeviceBuffer<ColorRGB32F>& framebuffer() const { return m_pixels; }
    const GBufferStorage& g_buffer() const { return m_g_buffer; }
    b200pt_bvh* bvh() const { return m_bvh; }
    float last_frame_ms()
    {
        check(b200pt_event_elapsed_ms(m_ctx, m_frame_start, m_frame_stop, &m_last_frame_ms), "b200pt_event_elapsed_ms");
        return m_last_frame_ms;
    }
    // Per-kernel device times of the LAST sample of the last frame (the reference keeps a start/stop event pair per
    // kernel, src/Compiler/GPUKernel.cpp:202-211,233-245). which: 0 CameraRays, 1 FullPathTracer, 2 ReSTIR DI passes
    float last_kernel_ms(int which)
    {
        float ms = 0.0f;
        check(b200pt_event_elapsed_ms(m_ctx, m_kernel_events[2 * which], m_kernel_events[2 * which + 1], &ms), "b200pt_event_elapsed_ms");
        return ms;
    }
    int width() const { return m_res.x; }
    int height() const { return m_res.y; }

    // buffers the tests read back
    DeviceBuffer<unsigned char> m_pixel_active;
    DeviceBuffer<int> m_pixel_sample_count, m_pixel_converged_sample_count;
    DeviceBuffer<float> m_pixel_squared_luminance;
    DeviceBuffer<ColorRGB32F> m_denoiser_albedo;
    DeviceBuffer<b200pt_float3> m_denoiser_normals;
    DeviceBuffer<ReSTIRDIReservoir> m_restir_initial, m_restir_spatial_1, m_restir_spatial_2;
    DeviceBuffer<ReSTIRDIPresampledLight> m_restir_presampled_lights;

private:
    void update_render_data();
    void launch_camera_rays();
    void launch_ReSTIR_DI();
    void launch_path_tracing(const b200pt_sample_batch* batch = nullptr);
    bool can_render_sample_batches() const;
    void render_sample_batches();

    b200pt_ctx* m_ctx;
    b200pt_bvh* m_bvh = nullptr;
    b200pt_options m_options = {3, 0, 0, 3};
    HIPRTRenderData m_render_data;
    b200pt_int2 m_res = {0, 0};
    bool m_has_tile = false;
    b200pt_tile m_tile = {0, 0, 0, 0};
    int m_tile_halo = 0;
    int m_path_lanes = 1;
    b200pt_stream m_lane_streams[3] = {nullptr, nullptr, nullptr};
    b200pt_event m_lane_fork = nullptr;
    b200pt_event m_lane_join[3] = {nullptr, nullptr, nullptr};
    Xorshift32Generator m_rng;
    HIPRTCamera m_camera;
    bool m_restir_odd_frame = false;
    b200pt_event m_frame_start = nullptr, m_frame_stop = nullptr;
    b200pt_event m_kernel_events[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool m_time_kernels = false;
    float m_last_frame_ms = 0.0f;

    DeviceBuffer<ColorRGB32F> m_pixels;
    DeviceBuffer<unsigned char> m_still_one_ray_active;
    DeviceBuffer<unsigned int> m_stop_noise_threshold_count;
    GBufferStorage m_g_buffer, m_g_buffer_prev;

    DeviceBuffer<int> m_triangle_indices, m_material_indices, m_emissive_triangles;
    DeviceBuffer<b200pt_float3> m_vertices, m_normals;
    DeviceBuffer<b200pt_float2> m_texcoords;
    DeviceBuffer<unsigned char> m_has_vertex_normals;
    DeviceBuffer<RendererMaterial> m_materials;
    DeviceBuffer<b200pt_int2> m_textures_dims;
    std::vector<void*> m_texture_descriptors;
    void* m_texture_array = nullptr;
    void* m_luts[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    void* m_envmap = nullptr;
    DeviceBuffer<float> m_alias_probas;
    DeviceBuffer<int> m_alias_alias;
};
} // namespace b200pt_host
