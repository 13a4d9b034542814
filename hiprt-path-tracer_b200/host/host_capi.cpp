// C handle API over b200pt_host::GPURenderer (include/b200pt_host.h).
#include "GPUBaker.h"
#include "GPURenderer.h"
#include "b200pt_host.h"

#include <cstring>
#include <string>

using namespace b200pt_host;

struct b200pt_baker
{
    explicit b200pt_baker(b200pt_ctx* c) : baker(c) {}
    GPUBaker baker;
};

struct b200pt_renderer
{
    explicit b200pt_renderer(b200pt_ctx* c) : ctx(c), renderer(c) {}
    b200pt_ctx* ctx;
    GPURenderer renderer;
    Scene pending_scene;   // textures are added before set_scene
};

namespace
{
thread_local std::string g_host_error;

template <typename F>
int guarded(F&& f)
{
    try
    {
        f();
        return B200PT_SUCCESS;
    }
    catch (const std::exception& e)
    {
        g_host_error = e.what();
        return B200PT_ERROR_CUDA;
    }
}
} // namespace

extern "C" {

const char* b200pt_host_last_error(void) { return g_host_error.c_str(); }

int b200pt_renderer_create(b200pt_ctx* ctx, b200pt_renderer** out)
{
    if (!ctx || !out)
    {
        g_host_error = "invalid argument";
        return B200PT_ERROR_INVALID_ARGUMENT;
    }
    *out = nullptr;
    return guarded([&] { *out = new b200pt_renderer(ctx); });
}

int b200pt_renderer_destroy(b200pt_renderer* r)
{
    delete r;
    return B200PT_SUCCESS;
}

int b200pt_renderer_set_seed_schedule(b200pt_renderer* r, int seed_schedule)
{
    r->renderer.seed_schedule = seed_schedule == 1 ? SeedSchedule::CPU : SeedSchedule::GPU;
    return B200PT_SUCCESS;
}

int b200pt_renderer_set_rng_seed(b200pt_renderer* r, unsigned int seed)
{
    r->renderer.rng_seed = seed;
    r->renderer.reset();
    return B200PT_SUCCESS;
}

int b200pt_renderer_set_kernel_options(b200pt_renderer* r, const b200pt_options* options)
{
    return guarded([&] { r->renderer.set_kernel_options(*options); });
}

int b200pt_renderer_set_brdf_luts(b200pt_renderer* r, const float* sheen_ltc, const float* conductor, const float* glossy, const float* glass, const float* glass_inverse, const float* thin_glass)
{
    return guarded([&] { r->renderer.set_brdf_luts(sheen_ltc, conductor, glossy, glass, glass_inverse, thin_glass); });
}

int b200pt_renderer_add_texture(b200pt_renderer* r, const unsigned char* texels, int width, int height, int channels)
{
    Scene::Texture8 t;
    t.texels.assign(texels, texels + (size_t)width * height * channels);
    t.width = width; t.height = height; t.channels = channels;
    r->pending_scene.textures.push_back(std::move(t));
    return B200PT_SUCCESS;
}

int b200pt_renderer_set_scene(b200pt_renderer* r, const float* vertices, int vertex_count, const int* triangle_indices, int triangle_count,
                              const float* vertex_normals, const unsigned char* has_vertex_normals, const float* texcoords,
                              const int* material_indices, const RendererMaterial* materials, int material_count,
                              const int* emissive_triangle_indices, int emissive_triangle_count)
{
    return guarded([&] {
        Scene& s = r->pending_scene;
        const b200pt_float3* v = reinterpret_cast<const b200pt_float3*>(vertices);
        const b200pt_float3* n = reinterpret_cast<const b200pt_float3*>(vertex_normals);
        const b200pt_float2* t = reinterpret_cast<const b200pt_float2*>(texcoords);
        s.vertices_positions.assign(v, v + vertex_count);
        s.vertex_normals.assign(n, n + vertex_count);
        s.texcoords.assign(t, t + vertex_count);
        s.has_vertex_normals.assign(has_vertex_normals, has_vertex_normals + vertex_count);
        s.triangle_indices.assign(triangle_indices, triangle_indices + 3 * (size_t)triangle_count);
        s.material_indices.assign(material_indices, material_indices + triangle_count);
        s.materials.assign(materials, materials + material_count);
        s.emissive_triangle_indices.assign(emissive_triangle_indices, emissive_triangle_indices + emissive_triangle_count);
        r->renderer.set_scene(s);
        s = Scene();
    });
}

int b200pt_renderer_set_envmap(b200pt_renderer* r, const float* texels, int width, int height, int channels, const float* alias_probas, const int* alias_alias, float total_sum)
{
    return guarded([&] { r->renderer.set_envmap(texels, width, height, channels, alias_probas, alias_alias, total_sum); });
}

int b200pt_host_compute_alias_table(const float* texels, int width, int height, int channels, float* out_probas, int* out_alias, float* out_total_sum)
{
    return guarded([&] {
        if (!texels || width <= 0 || height <= 0 || channels < 1 || !out_probas || !out_alias || !out_total_sum)
            throw std::runtime_error("b200pt_host_compute_alias_table: invalid argument");
        std::vector<float> probas;
        std::vector<int> alias;
        b200pt_host::compute_envmap_alias_table(texels, width, height, channels, probas, alias, *out_total_sum);
        std::copy(probas.begin(), probas.end(), out_probas);
        std::copy(alias.begin(), alias.end(), out_alias);
    });
}

int b200pt_renderer_set_camera(b200pt_renderer* r, const HIPRTCamera* camera)
{
    r->renderer.set_camera(*camera);
    return B200PT_SUCCESS;
}

int b200pt_renderer_resize(b200pt_renderer* r, int width, int height) { return guarded([&] { r->renderer.resize(width, height); }); }
int b200pt_renderer_set_tile(b200pt_renderer* r, const b200pt_tile* tile) { r->renderer.set_tile(tile); return B200PT_SUCCESS; }
int b200pt_renderer_set_tile_halo(b200pt_renderer* r, int rows) { r->renderer.set_tile_halo(rows); return B200PT_SUCCESS; }
int b200pt_renderer_set_path_lanes(b200pt_renderer* r, int lanes) { return guarded([&] { r->renderer.set_path_lanes(lanes); }); }
int b200pt_renderer_path_lanes(b200pt_renderer* r) { return r->renderer.path_lanes(); }
int b200pt_renderer_set_samples_in_flight(b200pt_renderer* r, int samples, int keep_g_buffer)
{
    r->renderer.samples_in_flight = samples < 0 ? 0 : (samples > 16 ? 16 : samples);
    r->renderer.keep_g_buffer = keep_g_buffer != 0;
    return B200PT_SUCCESS;
}
int b200pt_renderer_renders_sample_batches(b200pt_renderer* r) { return r->renderer.renders_sample_batches() ? 1 : 0; }
int b200pt_renderer_reset(b200pt_renderer* r) { r->renderer.reset(); return B200PT_SUCCESS; }
HIPRTRenderSettings* b200pt_renderer_render_settings(b200pt_renderer* r) { return &r->renderer.get_render_settings(); }
WorldSettings* b200pt_renderer_world_settings(b200pt_renderer* r) { return &r->renderer.get_world_settings(); }
BRDFsData* b200pt_renderer_bsdfs_data(b200pt_renderer* r) { return &r->renderer.get_bsdfs_data(); }
HIPRTRenderData* b200pt_renderer_render_data(b200pt_renderer* r) { return &r->renderer.get_render_data(); }
b200pt_bvh* b200pt_renderer_bvh(b200pt_renderer* r) { return r->renderer.bvh(); }

int b200pt_renderer_render_frame(b200pt_renderer* r)
{
    return guarded([&] {
        r->renderer.update();
        r->renderer.render();
    });
}

int b200pt_renderer_synchronize(b200pt_renderer* r) { return guarded([&] { r->renderer.synchronize_kernel(); }); }

int b200pt_renderer_last_frame_ms(b200pt_renderer* r, float* out_ms) { return guarded([&] { *out_ms = r->renderer.last_frame_ms(); }); }

int b200pt_renderer_last_kernel_ms(b200pt_renderer* r, int which, float* out_ms)
{
    if (which < 0 || which > 2) return B200PT_ERROR_INVALID_ARGUMENT;
    return guarded([&] { *out_ms = r->renderer.last_kernel_ms(which); });
}

void* b200pt_renderer_device_pointer(b200pt_renderer* r, int which)
{
    GPURenderer& g = r->renderer;
    const HIPRTRenderData& rd = g.get_render_data();
    switch (which)
    {
    case 0: return g.framebuffer().get_device_pointer();
    case 1: return g.g_buffer().first_hit_prim_index.get_device_pointer();
    case 2: return g.g_buffer().shading_normals.get_device_pointer();
    case 3: return g.g_buffer().geometric_normals.get_device_pointer();
    case 4: return g.g_buffer().view_directions.get_device_pointer();
    case 5: return g.g_buffer().first_hits.get_device_pointer();
    case 6: return g.g_buffer().camera_ray_hit.get_device_pointer();
    case 7: return g.g_buffer().materials.get_device_pointer();
    case 8: return g.g_buffer().ray_volume_states.get_device_pointer();
    case 9: return g.m_pixel_active.get_device_pointer();
    case 10: return g.m_pixel_sample_count.get_device_pointer();
    case 11: return g.m_pixel_squared_luminance.get_device_pointer();
    case 12: return g.m_pixel_converged_sample_count.get_device_pointer();
    case 13: return g.m_denoiser_albedo.get_device_pointer();
    case 14: return g.m_denoiser_normals.get_device_pointer();
    case 15: return rd.render_settings.restir_di_settings.restir_output_reservoirs;
    case 16: return g.m_restir_initial.get_device_pointer();
    case 17: return g.m_restir_presampled_lights.get_device_pointer();
    default: return nullptr;
    }
}

long long b200pt_renderer_download(b200pt_renderer* r, int which, void* out_host, long long capacity_bytes)
{
    GPURenderer& g = r->renderer;
    size_t n = (size_t)g.width() * g.height();
    static const size_t elem[] = {12, 4, 12, 12, 12, 12, 1, 256, 36, 1, 4, 4, 4, 12, 12, 36, 36, 48};
    if (which < 0 || which > 17)
        return B200PT_ERROR_INVALID_ARGUMENT;
    size_t count = which == 17 ? g.m_restir_presampled_lights.get_element_count() : n;
    long long bytes = (long long)(count * elem[which]);
    void* src = b200pt_renderer_device_pointer(r, which);
    if (src == nullptr)
        return 0;
    if (bytes > capacity_bytes)
        return -bytes;
    int e = b200pt_memcpy_d2h(r->ctx, out_host, src, (size_t)bytes, nullptr);
    if (e == 0)
        e = b200pt_stream_sync(r->ctx, nullptr);
    if (e != 0)
    {
        g_host_error = b200pt_last_error();
        return e;
    }
    return bytes;
}

// ---- GPUBaker -----------------------------------------------------------------------------------------------------------------------
int b200pt_baker_create(b200pt_ctx* ctx, b200pt_baker** out)
{
    if (!ctx || !out)
    {
        g_host_error = "invalid argument";
        return B200PT_ERROR_INVALID_ARGUMENT;
    }
    *out = nullptr;
    return guarded([&] { *out = new b200pt_baker(ctx); });
}

int b200pt_baker_destroy(b200pt_baker* b)
{
    delete b;
    return B200PT_SUCCESS;
}

int b200pt_baker_set_options(b200pt_baker* b, int ggx_sample_function, int ggx_masking_shadowing, int chunks, const char* output_directory)
{
    if (!b)
    {
        g_host_error = "invalid argument";
        return B200PT_ERROR_INVALID_ARGUMENT;
    }
    b200pt_bake_options o;
    o.ggx_sample_function = ggx_sample_function;
    o.GGX_masking_shadowing = ggx_masking_shadowing;
    o.chunks = chunks;
    b->baker.set_options(o);
    b->baker.set_output_directory(output_directory ? output_directory : "");
    return B200PT_SUCCESS;
}

int b200pt_baker_bake(b200pt_baker* b, int table, int size_cos_theta, int size_roughness, int size_ior, int integration_sample_count, const char* output_filename)
{
    if (!b || table < B200PT_BAKE_GGX_CONDUCTOR || table > B200PT_BAKE_GGX_THIN_GLASS || table == B200PT_BAKE_GGX_GLASS_EXITING)
    {
        g_host_error = "invalid argument (table: a b200pt_bake_kind; the two glass tables are baked together through B200PT_BAKE_GGX_GLASS_ENTERING)";
        return B200PT_ERROR_INVALID_ARGUMENT;
    }
    // the plan validates the sizes before a worker is started
    b200pt_bake_plan plan;
    int rc = b200pt_bake_plan_for(table, size_cos_theta, size_roughness, size_ior, integration_sample_count, nullptr, &plan);
    if (rc != B200PT_SUCCESS)
    {
        g_host_error = b200pt_last_error();
        return rc;
    }
    const std::string name = output_filename ? output_filename : "";
    return guarded([&] {
        switch (table)
        {
        case B200PT_BAKE_GGX_CONDUCTOR: b->baker.bake_ggx_conductor_directional_albedo({size_cos_theta, size_roughness, integration_sample_count}, name); break;
        case B200PT_BAKE_GGX_FRESNEL: b->baker.bake_ggx_fresnel_directional_albedo({size_cos_theta, size_roughness, size_ior, integration_sample_count}, name); break;
        case B200PT_BAKE_GLOSSY_DIELECTRIC: b->baker.bake_glossy_dielectric_directional_albedo({size_cos_theta, size_roughness, size_ior, integration_sample_count}, name); break;
        case B200PT_BAKE_GGX_GLASS_ENTERING: b->baker.bake_ggx_glass_directional_albedo({size_cos_theta, size_roughness, size_ior, integration_sample_count}, name); break;
        default: b->baker.bake_ggx_thin_glass_directional_albedo({size_cos_theta, size_roughness, size_ior, integration_sample_count}, name); break;
        }
    });
}

int b200pt_baker_is_complete(b200pt_baker* b, int table)
{
    if (!b)
        return B200PT_ERROR_INVALID_ARGUMENT;
    switch (table)
    {
    case B200PT_BAKE_GGX_CONDUCTOR: return b->baker.is_ggx_conductor_directional_albedo_bake_complete() ? 1 : 0;
    case B200PT_BAKE_GGX_FRESNEL: return b->baker.is_ggx_fresnel_directional_albedo_bake_complete() ? 1 : 0;
    case B200PT_BAKE_GLOSSY_DIELECTRIC: return b->baker.is_glossy_dielectric_directional_albedo_bake_complete() ? 1 : 0;
    case B200PT_BAKE_GGX_GLASS_ENTERING:
    case B200PT_BAKE_GGX_GLASS_EXITING: return b->baker.is_ggx_glass_directional_albedo_bake_complete() ? 1 : 0;
    case B200PT_BAKE_GGX_THIN_GLASS: return b->baker.is_ggx_thin_glass_directional_albedo_bake_complete() ? 1 : 0;
    }
    return B200PT_ERROR_INVALID_ARGUMENT;
}

int b200pt_baker_wait(b200pt_baker* b)
{
    if (!b)
        return B200PT_ERROR_INVALID_ARGUMENT;
    b->baker.wait();
    for (int kind = 0; kind < 6; kind++)
        if (!b->baker.last_error(kind).empty())
        {
            g_host_error = b->baker.last_error(kind);
            return B200PT_ERROR_CUDA;
        }
    return B200PT_SUCCESS;
}

long long b200pt_baker_last_table(b200pt_baker* b, int kind, float* out, long long capacity_floats, float* out_kernel_ms)
{
    if (!b || kind < 0 || kind > 5 || !b200pt_baker_is_complete(b, kind))
        return 0;
    const std::vector<float>& t = b->baker.last_table(kind);
    if ((long long)t.size() > capacity_floats || out == nullptr)
        return -(long long)t.size();
    std::memcpy(out, t.data(), t.size() * sizeof(float));
    if (out_kernel_ms)
        *out_kernel_ms = b->baker.last_bake_ms(kind);
    return (long long)t.size();
}

int b200pt_host_write_hdr_1channel(const char* filename, const float* texels, int width, int height)
{
    if (!filename || !texels || width <= 0 || height <= 0)
    {
        g_host_error = "invalid argument";
        return B200PT_ERROR_INVALID_ARGUMENT;
    }
    if (!write_image_hdr_1channel(filename, texels, width, height))
    {
        g_host_error = std::string("cannot write ") + filename;
        return B200PT_ERROR_INVALID_ARGUMENT;
    }
    return B200PT_SUCCESS;
}

} // extern "C"
