// reference: src/Renderer/Baker/GPUBaker.cpp:12-96, src/Renderer/Baker/GPUBakerKernel.cpp:22-139 (see GPUBaker.h)
#include "GPUBaker.h"

#include "GPURenderer.h"

#include <cmath>
#include <cstdio>

namespace b200pt_host
{
bool write_image_hdr_1channel(const std::string& filename, const float* texels, int width, int height)
{
    FILE* f = std::fopen(filename.c_str(), "wb");
    if (!f)
        return false;
    std::fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", height, width);
    std::vector<unsigned char> row((size_t)width * 4);
    bool ok = true;
    for (int y = 0; y < height && ok; y++)
    {
        for (int x = 0; x < width; x++)
        {
            float v = texels[(size_t)y * width + x];
            unsigned char* p = &row[(size_t)x * 4];
            if (!std::isfinite(v) || v < 1.0e-32f)
            {
                p[0] = p[1] = p[2] = p[3] = 0;
                continue;
            }
            int e;
            float scale = std::frexp(v, &e) * 256.0f / v;      // shared exponent of the (only) channel
            unsigned char m = (unsigned char)(v * scale);
            p[0] = p[1] = p[2] = m;
            p[3] = (unsigned char)(e + 128);
        }
        ok = std::fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    return std::fclose(f) == 0 && ok;
}

GPUBaker::GPUBaker(b200pt_ctx* ctx) : m_ctx(ctx)
{
    check(b200pt_stream_create(m_ctx, &m_bake_stream), "b200pt_stream_create");
}

GPUBaker::~GPUBaker()
{
    wait();
    if (m_bake_stream)
        b200pt_stream_destroy(m_ctx, m_bake_stream);
}

void GPUBaker::wait()
{
    for (Job& job : m_jobs)
        if (job.worker.joinable())
            job.worker.join();
}

void GPUBaker::start(int kind, int size_x, int size_y, int size_z, int sample_count, const std::string& output_filename)
{
    Job& job = m_jobs[kind];
    if (job.worker.joinable())
        job.worker.join();                      // a second bake of the same table waits for the first
    job.complete = false;
    job.error.clear();
    const b200pt_bake_options options = m_options;
    const std::string path_prefix = m_directory.empty() ? std::string() : m_directory + "/";
    job.worker = std::thread([this, &job, kind, size_x, size_y, size_z, sample_count, output_filename, options, path_prefix] {
        const size_t texels = (size_t)size_x * size_y * size_z;
        float* device_buffer = nullptr;
        b200pt_event start = nullptr, stop = nullptr;
        try
        {
            check(b200pt_ctx_make_current(m_ctx), "b200pt_ctx_make_current");
            check(b200pt_malloc(m_ctx, texels * sizeof(float), reinterpret_cast<void**>(&device_buffer)), "b200pt_malloc");
            check(b200pt_event_create(m_ctx, &start), "b200pt_event_create");
            check(b200pt_event_create(m_ctx, &stop), "b200pt_event_create");
            check(b200pt_event_record(m_ctx, start, m_bake_stream), "b200pt_event_record");
            int rc = 0;
            switch (kind)
            {
            case B200PT_BAKE_GGX_CONDUCTOR:
            {
                GGXConductorDirectionalAlbedoSettings s{size_x, size_y, sample_count};
                rc = b200pt_bake_ggx_conductor_directional_albedo(m_ctx, &s, &options, device_buffer, m_bake_stream);
                break;
            }
            case B200PT_BAKE_GGX_FRESNEL:
            {
                GGXFresnelDirectionalAlbedoSettings s{size_x, size_y, size_z, sample_count};
                rc = b200pt_bake_ggx_fresnel_directional_albedo(m_ctx, &s, &options, device_buffer, m_bake_stream);
                break;
            }
            case B200PT_BAKE_GLOSSY_DIELECTRIC:
            {
                GlossyDielectricDirectionalAlbedoSettings s{size_x, size_y, size_z, sample_count};
                rc = b200pt_bake_glossy_dielectric_directional_albedo(m_ctx, &s, &options, device_buffer, m_bake_stream);
                break;
            }
            case B200PT_BAKE_GGX_GLASS_ENTERING:
            case B200PT_BAKE_GGX_GLASS_EXITING:
            {
                GGXGlassDirectionalAlbedoSettings s{size_x, size_y, size_z, sample_count};
                rc = b200pt_bake_ggx_glass_directional_albedo(m_ctx, &s, &options, kind == B200PT_BAKE_GGX_GLASS_ENTERING ? device_buffer : nullptr,
                                                              kind == B200PT_BAKE_GGX_GLASS_EXITING ? device_buffer : nullptr, m_bake_stream);
                break;
            }
            default:
            {
                GGXThinGlassDirectionalAlbedoSettings s{size_x, size_y, size_z, sample_count};
                rc = b200pt_bake_ggx_thin_glass_directional_albedo(m_ctx, &s, &options, device_buffer, m_bake_stream);
                break;
            }
            }
            check(rc, "b200pt_bake");
            check(b200pt_event_record(m_ctx, stop, m_bake_stream), "b200pt_event_record");
            job.table.resize(texels);
            check(b200pt_memcpy_d2h(m_ctx, job.table.data(), device_buffer, texels * sizeof(float), m_bake_stream), "b200pt_memcpy_d2h");
            check(b200pt_stream_sync(m_ctx, m_bake_stream), "b200pt_stream_sync");
            check(b200pt_event_elapsed_ms(m_ctx, start, stop, &job.ms), "b200pt_event_elapsed_ms");
            // GPUBakerKernel.cpp:107-126
            if (!output_filename.empty())
            {
                const size_t slice = (size_t)size_x * size_y;
                for (int z = 0; z < size_z; z++)
                {
                    const std::string name = path_prefix + (size_z > 1 ? std::to_string(z) + output_filename : output_filename);
                    if (!write_image_hdr_1channel(name, job.table.data() + (size_t)z * slice, size_x, size_y))
                        throw std::runtime_error("cannot write " + name);
                }
            }
        }
        catch (const std::exception& e)
        {
            job.error = e.what();
        }
        if (start)
            b200pt_event_destroy(m_ctx, start);
        if (stop)
            b200pt_event_destroy(m_ctx, stop);
        if (device_buffer)
            b200pt_free(m_ctx, device_buffer);
        job.complete = true;
    });
}

void GPUBaker::bake_ggx_conductor_directional_albedo(const GGXConductorDirectionalAlbedoSettings& s, const std::string& output_filename)
{
    start(B200PT_BAKE_GGX_CONDUCTOR, s.texture_size_cos_theta, s.texture_size_roughness, 1, s.integration_sample_count, output_filename);
}

void GPUBaker::bake_ggx_fresnel_directional_albedo(const GGXFresnelDirectionalAlbedoSettings& s, const std::string& output_filename)
{
    start(B200PT_BAKE_GGX_FRESNEL, s.texture_size_cos_theta, s.texture_size_roughness, s.texture_size_ior, s.integration_sample_count, output_filename);
}

void GPUBaker::bake_glossy_dielectric_directional_albedo(const GlossyDielectricDirectionalAlbedoSettings& s, const std::string& output_filename)
{
    start(B200PT_BAKE_GLOSSY_DIELECTRIC, s.texture_size_cos_theta_o, s.texture_size_roughness, s.texture_size_ior, s.integration_sample_count, output_filename);
}

void GPUBaker::bake_ggx_glass_directional_albedo(const GGXGlassDirectionalAlbedoSettings& s, const std::string& output_filename)
{
    start(B200PT_BAKE_GGX_GLASS_ENTERING, s.texture_size_cos_theta_o, s.texture_size_roughness, s.texture_size_ior, s.integration_sample_count, output_filename);
    start(B200PT_BAKE_GGX_GLASS_EXITING, s.texture_size_cos_theta_o, s.texture_size_roughness, s.texture_size_ior, s.integration_sample_count,
          output_filename.empty() ? output_filename : "inv_" + output_filename);
}

void GPUBaker::bake_ggx_thin_glass_directional_albedo(const GGXThinGlassDirectionalAlbedoSettings& s, const std::string& output_filename)
{
    start(B200PT_BAKE_GGX_THIN_GLASS, s.texture_size_cos_theta_o, s.texture_size_roughness, s.texture_size_ior, s.integration_sample_count, output_filename);
}
} // namespace b200pt_host
