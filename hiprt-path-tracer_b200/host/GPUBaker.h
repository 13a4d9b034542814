// Host side of the LUT bakers above the C ABI (include/b200pt_bake.h): the reference's GPUBaker (src/Renderer/Baker/GPUBaker.{h,cpp}) with its
// method names and its asynchronous behaviour -- bake_*() returns at once, a worker waits for the table, writes the .hdr files the way
// GPUBakerKernel::bake_internal does (src/Renderer/Baker/GPUBakerKernel.cpp:107-126: one single-channel Radiance file per ior slice,
// "<slice index><output_filename>", rows unflipped) and raises the flag is_*_bake_complete() polls. What is gone: the run-time kernel compilation,
// the zeroing kernel and the loop of nb_kernel_launch launches (one launch per table, csrc/b200pt_bake.cu).
#pragma once

#include "b200pt_bake.h"

#include <atomic>
#include <memory>
#include <string>
#include <thread>
#include <vector>

namespace b200pt_host
{
// Image32Bit::write_image_hdr of a 1-channel image (src/Image/Image.cpp:440-452 -> stbi_write_hdr): grey replicated into Greg Ward's RGBE,
// flat (no run-length encoding), rows top to bottom. Non-finite texels (the thin-glass table has NaN texels by the reference's own arithmetic,
// DESIGN 7.1) are written as 0. Returns false when the file cannot be written.
bool write_image_hdr_1channel(const std::string& filename, const float* texels, int width, int height);

class GPUBaker
{
public:
    // The reference takes its GPURenderer for the device context (GPUBaker.cpp:12-16); here the context itself. The baker owns one stream.
    explicit GPUBaker(b200pt_ctx* ctx);
    ~GPUBaker();                                          // waits for the bakes in flight
    GPUBaker(const GPUBaker&) = delete;
    GPUBaker& operator=(const GPUBaker&) = delete;

    void set_options(const b200pt_bake_options& options) { m_options = options; }
    void set_output_directory(const std::string& directory) { m_directory = directory; }   // the reference writes relative to the working directory

    void bake_ggx_conductor_directional_albedo(const GGXConductorDirectionalAlbedoSettings& bake_settings, const std::string& output_filename);
    bool is_ggx_conductor_directional_albedo_bake_complete() const { return m_jobs[0].complete; }

    void bake_ggx_fresnel_directional_albedo(const GGXFresnelDirectionalAlbedoSettings& bake_settings, const std::string& output_filename);
    bool is_ggx_fresnel_directional_albedo_bake_complete() const { return m_jobs[1].complete; }

    void bake_glossy_dielectric_directional_albedo(const GlossyDielectricDirectionalAlbedoSettings& bake_settings, const std::string& output_filename);
    bool is_glossy_dielectric_directional_albedo_bake_complete() const { return m_jobs[2].complete; }

    // two tables: entering the medium -> output_filename, leaving it -> "inv_" + output_filename (GPUBaker.cpp:74-83)
    void bake_ggx_glass_directional_albedo(const GGXGlassDirectionalAlbedoSettings& bake_settings, const std::string& output_filename);
    bool is_ggx_glass_directional_albedo_bake_complete() const { return m_jobs[3].complete && m_jobs[4].complete; }

    void bake_ggx_thin_glass_directional_albedo(const GGXThinGlassDirectionalAlbedoSettings& bake_settings, const std::string& output_filename);
    bool is_ggx_thin_glass_directional_albedo_bake_complete() const { return m_jobs[5].complete; }

    // additions for headless callers: block until every bake has finished; the kernel time of the last bake of a table (b200pt_bake_kind), the
    // error of a failed one (empty = none), the table itself (kept until the next bake of that kind)
    void wait();
    float last_bake_ms(int kind) const { return m_jobs[kind].ms; }
    const std::string& last_error(int kind) const { return m_jobs[kind].error; }
    const std::vector<float>& last_table(int kind) const { return m_jobs[kind].table; }

private:
    struct Job
    {
        std::thread worker;
        std::atomic<bool> complete{true};      // GPUBakerKernel::m_bake_complete starts true (GPUBakerKernel.h:31)
        float ms = 0.0f;
        std::string error;
        std::vector<float> table;
    };
    void start(int kind, int size_x, int size_y, int size_z, int sample_count, const std::string& output_filename);

    b200pt_ctx* m_ctx = nullptr;
    b200pt_stream m_bake_stream = nullptr;
    b200pt_bake_options m_options{};
    std::string m_directory;
    Job m_jobs[6];
};
} // namespace b200pt_host
