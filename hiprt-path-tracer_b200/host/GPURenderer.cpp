// See GPURenderer.h. reference: src/Renderer/GPURenderer.cpp:236-267 (update), :408-486 (render + launch points),
// :980-1039 (update_render_data), :1041-1174 (set_scene / set_envmap); CPU seed schedule: src/Renderer/CPURenderer.cpp:264-296.
#include "GPURenderer.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b200pt_host
{
void GBufferStorage::set_context(b200pt_ctx* ctx)
{
    materials.set_context(ctx); shading_normals.set_context(ctx); geometric_normals.set_context(ctx); view_directions.set_context(ctx);
    first_hits.set_context(ctx); first_hit_prim_index.set_context(ctx); camera_ray_hit.set_context(ctx); ray_volume_states.set_context(ctx);
}

void GBufferStorage::resize(size_t n)
{
    materials.resize(n); shading_normals.resize(n); geometric_normals.resize(n); view_directions.resize(n);
    first_hits.resize(n); first_hit_prim_index.resize(n); camera_ray_hit.resize(n); ray_volume_states.resize(n);
    if (n > 0)
    {
        materials.memset_zero(); shading_normals.memset_zero(); geometric_normals.memset_zero(); view_directions.memset_zero();
        first_hits.memset_zero(); first_hit_prim_index.memset_zero(); camera_ray_hit.memset_zero(); ray_volume_states.memset_zero();
    }
}

GBuffer GBufferStorage::view() const
{
    GBuffer g;
    g.materials = materials.get_device_pointer();
    g.first_hit_prim_index = first_hit_prim_index.get_device_pointer();
    g.shading_normals = shading_normals.get_device_pointer();
    g.geometric_normals = geometric_normals.get_device_pointer();
    g.view_directions = view_directions.get_device_pointer();
    g.first_hits = first_hits.get_device_pointer();
    g.camera_ray_hit = camera_ray_hit.get_device_pointer();
    g.ray_volume_states = ray_volume_states.get_device_pointer();
    return g;
}

GPURenderer::GPURenderer(b200pt_ctx* ctx) : m_ctx(ctx)
{
    std::memset(&m_camera, 0, sizeof(m_camera));
    m_camera.do_jittering = 1;
    DeviceBuffer<unsigned char>* u8[] = {&m_pixel_active, &m_still_one_ray_active, &m_has_vertex_normals};
    for (auto* b : u8) b->set_context(ctx);
    DeviceBuffer<int>* i32[] = {&m_pixel_sample_count, &m_pixel_converged_sample_count, &m_triangle_indices, &m_material_indices, &m_emissive_triangles, &m_alias_alias};
    for (auto* b : i32) b->set_context(ctx);
    m_pixel_squared_luminance.set_context(ctx); m_alias_probas.set_context(ctx);
    m_denoiser_albedo.set_context(ctx); m_pixels.set_context(ctx);
    m_denoiser_normals.set_context(ctx); m_vertices.set_context(ctx); m_normals.set_context(ctx);
    m_texcoords.set_context(ctx); m_materials.set_context(ctx); m_textures_dims.set_context(ctx);
    m_stop_noise_threshold_count.set_context(ctx);
    m_restir_initial.set_context(ctx); m_restir_spatial_1.set_context(ctx); m_restir_spatial_2.set_context(ctx); m_restir_presampled_lights.set_context(ctx);
    m_g_buffer.set_context(ctx); m_g_buffer_prev.set_context(ctx);

    m_still_one_ray_active.resize(1);
    m_stop_noise_threshold_count.resize(1);
    check(b200pt_event_create(ctx, &m_frame_start), "b200pt_event_create");
    check(b200pt_event_create(ctx, &m_frame_stop), "b200pt_event_create");
    for (b200pt_event& e : m_kernel_events)
        check(b200pt_event_create(ctx, &e), "b200pt_event_create");
    check(b200pt_kernel_variant_select(ctx, &m_options), "b200pt_kernel_variant_select");
    // two concurrent band launches by default: measured -13 % (Cornell) / -4.5 % (10 M triangles) of FullPathTracer time at 1080p
    const char* lanes = std::getenv("B200PT_PATH_LANES");
    set_path_lanes(lanes ? std::atoi(lanes) : 2);
    if (const char* v = std::getenv("B200PT_SAMPLES_IN_FLIGHT"))
        samples_in_flight = std::min(16, std::max(0, std::atoi(v)));
}

GPURenderer::~GPURenderer()
{
    b200pt_stream_sync(m_ctx, nullptr);
    if (m_bvh) b200pt_bvh_destroy(m_ctx, m_bvh);
    for (void* d : m_texture_descriptors) b200pt_tex_destroy(m_ctx, d);
    if (m_texture_array) b200pt_free(m_ctx, m_texture_array);
    for (void* l : m_luts) if (l) b200pt_tex_destroy(m_ctx, l);
    if (m_envmap) b200pt_tex_destroy(m_ctx, m_envmap);
    for (b200pt_stream st : m_lane_streams) if (st) b200pt_stream_destroy(m_ctx, st);
    for (b200pt_event e : m_lane_join) if (e) b200pt_event_destroy(m_ctx, e);
    if (m_lane_fork) b200pt_event_destroy(m_ctx, m_lane_fork);
    b200pt_event_destroy(m_ctx, m_frame_start);
    b200pt_event_destroy(m_ctx, m_frame_stop);
    for (b200pt_event e : m_kernel_events) b200pt_event_destroy(m_ctx, e);
}

void GPURenderer::set_kernel_options(const b200pt_options& options)
{
    check(b200pt_kernel_variant_select(m_ctx, &options), "b200pt_kernel_variant_select");
    m_options = options;
}

void GPURenderer::set_brdf_luts(const float* sheen, const float* conductor, const float* glossy, const float* glass, const float* glass_inverse, const float* thin_glass)
{
    for (void*& l : m_luts)
    {
        if (l) b200pt_tex_destroy(m_ctx, l);
        l = nullptr;
    }
    // dimensions: src/Renderer/Baker/GPUBakerConstants.h:14-31; sheen table: 32x32 RGB (SheenLTCFittedParameters.h)
    check(b200pt_tex2d_create(m_ctx, sheen, 32, 32, 3, 4, &m_luts[0]), "sheen LTC LUT");
    check(b200pt_tex2d_create(m_ctx, conductor, 128, 128, 1, 4, &m_luts[1]), "GGX conductor LUT");
    check(b200pt_tex3d_create(m_ctx, glossy, 128, 64, 128, 1, &m_luts[2]), "glossy dielectric LUT");
    check(b200pt_tex3d_create(m_ctx, glass, 256, 16, 128, 1, &m_luts[3]), "GGX glass LUT");
    check(b200pt_tex3d_create(m_ctx, glass_inverse, 256, 16, 128, 1, &m_luts[4]), "GGX glass inverse LUT");
    check(b200pt_tex3d_create(m_ctx, thin_glass, 32, 32, 96, 1, &m_luts[5]), "GGX thin glass LUT");
    BRDFsData& b = m_render_data.bsdfs_data;
    b.sheen_ltc_parameters_texture = m_luts[0];
    b.GGX_conductor_Ess = m_luts[1];
    b.glossy_dielectric_Ess = m_luts[2];
    b.GGX_Ess_glass = m_luts[3];
    b.GGX_Ess_glass_inverse = m_luts[4];
    b.GGX_Ess_thin_glass = m_luts[5];
}

void GPURenderer::set_scene(const Scene& scene)
{
    if (m_bvh)
    {
        b200pt_bvh_destroy(m_ctx, m_bvh);
        m_bvh = nullptr;
    }
    int ntris = (int)(scene.triangle_indices.size() / 3);
    check(b200pt_bvh_build(m_ctx, reinterpret_cast<const float*>(scene.vertices_positions.data()), (int)scene.vertices_positions.size(), scene.triangle_indices.data(), ntris, 0, &m_bvh), "b200pt_bvh_build");

    m_triangle_indices.upload_data(scene.triangle_indices);
    m_vertices.upload_data(scene.vertices_positions);
    m_normals.upload_data(scene.vertex_normals);
    m_has_vertex_normals.upload_data(scene.has_vertex_normals);
    m_texcoords.upload_data(scene.texcoords);
    m_material_indices.upload_data(scene.material_indices);
    m_materials.upload_data(scene.materials);
    m_emissive_triangles.upload_data(scene.emissive_triangle_indices);

    for (void* d : m_texture_descriptors) b200pt_tex_destroy(m_ctx, d);
    m_texture_descriptors.clear();
    if (m_texture_array) b200pt_free(m_ctx, m_texture_array);
    m_texture_array = nullptr;
    std::vector<b200pt_int2> dims;
    for (const Scene::Texture8& t : scene.textures)
    {
        void* d = nullptr;
        check(b200pt_tex2d_create(m_ctx, t.texels.data(), t.width, t.height, t.channels, 1, &d), "b200pt_tex2d_create");
        m_texture_descriptors.push_back(d);
        dims.push_back({t.width, t.height});
    }
    check(b200pt_tex2d_array_create(m_ctx, m_texture_descriptors.data(), (int)m_texture_descriptors.size(), &m_texture_array), "b200pt_tex2d_array_create");
    m_textures_dims.upload_data(dims);

    // Per-triangle-slot shading records (b200pt.h): what a hit needs of its three vertices, gathered once instead of per hit.
    // B200PT_SHADING_RECORDS=0 keeps the reference's buffers as the only source (A/B and the bit-identity test).
    const char* records = std::getenv("B200PT_SHADING_RECORDS");
    if (ntris > 0 && !(records && records[0] == '0'))
        check(b200pt_bvh_attach_shading_data(m_ctx, m_bvh, m_triangle_indices.get_device_pointer(), reinterpret_cast<const float*>(m_normals.get_device_pointer()),
                                             reinterpret_cast<const unsigned char*>(m_has_vertex_normals.get_device_pointer()),
                                             reinterpret_cast<const float*>(m_texcoords.get_device_pointer()), m_material_indices.get_device_pointer(), nullptr),
              "b200pt_bvh_attach_shading_data");

    RenderBuffers& b = m_render_data.buffers;
    m_render_data.geom = b200pt_bvh_device_handle(m_bvh);
    b.triangles_indices = m_triangle_indices.get_device_pointer();
    b.vertices_positions = m_vertices.get_device_pointer();
    b.has_vertex_normals = m_has_vertex_normals.get_device_pointer();
    b.vertex_normals = m_normals.get_device_pointer();
    b.texcoords = m_texcoords.get_device_pointer();
    b.material_indices = m_material_indices.get_device_pointer();
    b.materials_buffer = m_materials.get_device_pointer();
    b.emissive_triangles_count = (int)scene.emissive_triangle_indices.size();
    b.emissive_triangles_indices = m_emissive_triangles.get_device_pointer();
    b.material_textures = m_texture_array;
    b.textures_dims = m_textures_dims.get_device_pointer();
}

// Alias table of the envmap's luminance distribution (Vose's method, double precision, FIFO work lists), the table
// envmap_sample draws texels from. Same construction rule as the reference's Image32Bit::compute_alias_table
// (src/Image/Image.cpp:579-659) -- per-texel luminance in float over at most 3 channels (:454-474), normalised to
// mean 1, "< 1" goes to the small list, a large entry is re-queued as large only while it stays > 1 -- so the
// tables are identical entry for entry.
void compute_envmap_alias_table(const float* texels, int width, int height, int channels, std::vector<float>& probas, std::vector<int>& alias, float& luminance_total_sum)
{
    const size_t n = (size_t)width * height;
    std::vector<double> scaled(n);
    const float weights[3] = {0.3086f, 0.6094f, 0.0820f};
    double sum = 0.0;
    for (size_t i = 0; i < n; i++)
    {
        float luminance = 0.0f;
        for (int c = 0; c < (channels < 3 ? channels : 3); c++)
            luminance += texels[i * channels + c] * weights[c];
        scaled[i] = (double)luminance;
        sum += (double)luminance;
    }
    luminance_total_sum = (float)sum;
    for (double& v : scaled)
    {
        v /= sum;
        v *= (double)(width * height);
    }

    probas.assign(n, 0.0f);
    alias.assign(n, 0);
    // FIFO lists as index rings: an entry is pushed at most once per pop, so n + 1 slots suffice for each
    std::vector<int> small_list(n + 1), large_list(n + 1);
    size_t small_head = 0, small_tail = 0, large_head = 0, large_tail = 0;
    auto push = [&](std::vector<int>& list, size_t& tail, int v) { list[tail] = v; tail = tail == n ? 0 : tail + 1; };
    auto pop = [&](std::vector<int>& list, size_t& head) { int v = list[head]; head = head == n ? 0 : head + 1; return v; };
    for (size_t i = 0; i < n; i++)
    {
        if (scaled[i] < 1.0)
            push(small_list, small_tail, (int)i);
        else
            push(large_list, large_tail, (int)i);
    }
    while (small_head != small_tail && large_head != large_tail)
    {
        const int s = pop(small_list, small_head);
        const int l = pop(large_list, large_head);
        probas[s] = (float)scaled[s];
        alias[s] = l;
        scaled[l] = (scaled[l] + scaled[s]) - 1.0;
        if (scaled[l] > 1.0)
            push(large_list, large_tail, l);
        else
            push(small_list, small_tail, l);
    }
    while (large_head != large_tail)
        probas[pop(large_list, large_head)] = 1.0f;
    while (small_head != small_tail)
        probas[pop(small_list, small_head)] = 1.0f;
}

void GPURenderer::set_envmap(const float* texels, int width, int height, int channels, const float* alias_probas, const int* alias_alias, float total_sum)
{
    std::vector<float> own_probas;
    std::vector<int> own_alias;
    if (texels != nullptr && width > 0 && height > 0 && (alias_probas == nullptr || alias_alias == nullptr))
    {
        compute_envmap_alias_table(texels, width, height, channels, own_probas, own_alias, total_sum);
        alias_probas = own_probas.data();
        alias_alias = own_alias.data();
    }
    WorldSettings& w = m_render_data.world_settings;
    if (m_envmap) b200pt_tex_destroy(m_ctx, m_envmap);
    m_envmap = nullptr;
    if (width == 0 || height == 0 || texels == nullptr)
    {
        // GPURenderer::set_envmap with an empty image: uniform ambient light (GPURenderer.cpp:1142-1150)
        w.ambient_light_type = AMBIENT_UNIFORM;
        w.uniform_light_color = {0.5f, 0.5f, 0.5f};
        return;
    }
    check(b200pt_tex2d_create(m_ctx, texels, width, height, channels, 4, &m_envmap), "envmap upload");
    size_t n = (size_t)width * height;
    m_alias_probas.resize(n);
    m_alias_alias.resize(n);
    m_alias_probas.upload_data(alias_probas);
    m_alias_alias.upload_data(alias_alias);
    w.envmap = m_envmap;
    w.envmap_width = (unsigned)width;
    w.envmap_height = (unsigned)height;
    w.envmap_total_sum = total_sum;
    w.alias_table_probas = m_alias_probas.get_device_pointer();
    w.alias_table_alias = m_alias_alias.get_device_pointer();
    w.ambient_light_type = AMBIENT_ENVMAP;
}

void GPURenderer::set_camera(const HIPRTCamera& camera)
{
    m_camera = camera;
    m_render_data.current_camera = camera;
    m_render_data.prev_camera = camera;
}

void GPURenderer::set_tile(const b200pt_tile* tile)
{
    m_has_tile = tile != nullptr;
    if (tile) m_tile = *tile;
}

void GPURenderer::set_tile_halo(int rows) { m_tile_halo = rows > 0 ? rows : 0; }

void GPURenderer::set_path_lanes(int lanes)
{
    m_path_lanes = std::min(4, std::max(1, lanes));
    for (int i = 0; i + 1 < m_path_lanes; i++)
    {
        if (!m_lane_streams[i]) check(b200pt_stream_create(m_ctx, &m_lane_streams[i]), "b200pt_stream_create");
        if (!m_lane_join[i]) check(b200pt_event_create(m_ctx, &m_lane_join[i]), "b200pt_event_create");
    }
    if (m_path_lanes > 1 && !m_lane_fork) check(b200pt_event_create(m_ctx, &m_lane_fork), "b200pt_event_create");
}

void GPURenderer::resize(int width, int height)
{
    synchronize_kernel();
    m_res = {width, height};
    size_t n = (size_t)width * height;
    m_pixels.resize(n); m_pixels.memset_zero();
    m_pixel_active.resize(n); m_pixel_active.memset_zero();
    m_denoiser_albedo.resize(n); m_denoiser_albedo.memset_zero();
    m_denoiser_normals.resize(n); m_denoiser_normals.memset_zero();
    m_g_buffer.resize(n);
    reset();
}

void GPURenderer::reset()
{
    m_render_data.render_settings.sample_number = 0;
    m_render_data.render_settings.denoiser_AOV_accumulation_counter = 0;
    m_render_data.render_settings.need_to_reset = 1;
    m_render_data.random_seed = 42;
    m_rng = Xorshift32Generator();
    m_rng.seed = rng_seed != 0u ? rng_seed : 42u;
    m_restir_odd_frame = false;
}

void GPURenderer::synchronize_kernel() { check(b200pt_stream_sync(m_ctx, nullptr), "b200pt_stream_sync"); }

void GPURenderer::update()
{
    HIPRTRenderSettings& s = m_render_data.render_settings;
    size_t n = (size_t)m_res.x * m_res.y;

    // ReSTIRDIRenderPass::update (ReSTIRDIRenderPass.cpp:266-304): reservoirs exist only while ReSTIR DI is the strategy
    bool restir = m_options.direct_light_sampling_strategy == 5;
    if (restir && m_restir_initial.get_element_count() != n)
    {
        std::vector<ReSTIRDIReservoir> empty(n);
        m_restir_initial.upload_data(empty);
        m_restir_spatial_1.upload_data(empty);
        m_restir_spatial_2.upload_data(empty);
        size_t presampled = (size_t)s.restir_di_settings.light_presampling.number_of_subsets * s.restir_di_settings.light_presampling.subset_size;
        m_restir_presampled_lights.resize(presampled);
        m_restir_presampled_lights.memset_zero();
        // The buffers were reallocated (resolution change, or ReSTIR DI switched back on): the output pointer of the previous
        // frame refers to freed memory. Re-point it to the (empty) first spatial buffer, as the reference does whenever it
        // invalidates the reservoirs (ReSTIRDIRenderPass.cpp:195-207), and restart the ping-pong.
        s.restir_di_settings.restir_output_reservoirs = m_restir_spatial_1.get_device_pointer();
        m_restir_odd_frame = false;
    }
    else if (!restir && m_restir_initial.get_element_count() != 0)
    {
        m_restir_initial.free(); m_restir_spatial_1.free(); m_restir_spatial_2.free(); m_restir_presampled_lights.free();
        s.restir_di_settings.restir_output_reservoirs = nullptr;
        m_restir_odd_frame = false;
    }

    // internal_update_clear_device_status_buffers (GPURenderer.cpp:275-283)
    m_still_one_ray_active.memset_zero();
    m_stop_noise_threshold_count.memset_zero();

    // internal_update_prev_frame_g_buffer (:291-317)
    bool need_prev = restir && s.restir_di_settings.temporal_pass.do_temporal_reuse_pass;
    if (need_prev && m_g_buffer_prev.camera_ray_hit.get_element_count() != n)
        m_g_buffer_prev.resize(n);
    else if (!need_prev && m_g_buffer_prev.camera_ray_hit.get_element_count() != 0)
        m_g_buffer_prev.resize(0);

    // internal_update_adaptive_sampling_buffers (:319-354)
    bool adaptive = (s.stop_pixel_noise_threshold > 0.0f || s.enable_adaptive_sampling) && s.accumulate;
    if (adaptive && m_pixel_sample_count.get_element_count() != n)
    {
        m_pixel_sample_count.resize(n); m_pixel_sample_count.memset_zero();
        m_pixel_converged_sample_count.resize(n); m_pixel_converged_sample_count.memset_zero();
        m_pixel_squared_luminance.resize(n); m_pixel_squared_luminance.memset_zero();
    }
    else if (!adaptive && m_pixel_sample_count.get_element_count() != 0)
    {
        m_pixel_sample_count.free(); m_pixel_converged_sample_count.free(); m_pixel_squared_luminance.free();
    }

    update_render_data();

    s.do_update_status_buffers = 0;
    if (!s.accumulate)
        s.sample_number = 0;
}

void GPURenderer::update_render_data()
{
    // GPURenderer::update_render_data (GPURenderer.cpp:980-1039): re-point every device pointer
    HIPRTRenderData& rd = m_render_data;
    rd.buffers.pixels = m_pixels.get_device_pointer();
    rd.aux_buffers.pixel_active = m_pixel_active.get_device_pointer();
    rd.aux_buffers.denoiser_albedo = m_denoiser_albedo.get_device_pointer();
    rd.aux_buffers.denoiser_normals = m_denoiser_normals.get_device_pointer();
    rd.aux_buffers.pixel_sample_count = m_pixel_sample_count.get_device_pointer();
    rd.aux_buffers.pixel_converged_sample_count = m_pixel_converged_sample_count.get_device_pointer();
    rd.aux_buffers.pixel_squared_luminance = m_pixel_squared_luminance.get_device_pointer();
    rd.aux_buffers.still_one_ray_active = m_still_one_ray_active.get_device_pointer();
    rd.aux_buffers.stop_noise_threshold_converged_count = m_stop_noise_threshold_count.get_device_pointer();
    rd.aux_buffers.restir_reservoir_buffer_1 = m_restir_initial.get_device_pointer();
    rd.aux_buffers.restir_reservoir_buffer_2 = m_restir_spatial_1.get_device_pointer();
    rd.aux_buffers.restir_reservoir_buffer_3 = m_restir_spatial_2.get_device_pointer();
    rd.g_buffer = m_g_buffer.view();
    rd.g_buffer_prev_frame = m_g_buffer_prev.view();
    ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;
    rs.light_presampling.light_samples = m_restir_presampled_lights.get_device_pointer();
    rs.initial_candidates.output_reservoirs = m_restir_initial.get_device_pointer();
    if (rs.restir_output_reservoirs == nullptr)
        rs.restir_output_reservoirs = m_restir_spatial_1.get_device_pointer();
}

// Fused sample batches need options under which a pass does not read what the previous pass accumulated (b200pt.h)
bool GPURenderer::can_render_sample_batches() const
{
    const HIPRTRenderSettings& s = m_render_data.render_settings;
    const int lss = m_options.direct_light_sampling_strategy;
    if (samples_in_flight <= 0 || !(lss == 3 || (lss == 4 && s.ris_settings.number_of_bsdf_candidates == 1)))
        return false;
    int mode = -1;
    if (b200pt_get_path_mode(m_ctx, &mode) != 0 || mode != 3)
        return false;
    if (s.number_of_light_samples != 1 || s.do_alpha_testing || !s.accumulate)
        return false;
    if ((s.stop_pixel_noise_threshold > 0.0f || s.enable_adaptive_sampling) && s.accumulate)
        return false;
    if (s.wants_render_low_resolution && s.allow_render_low_resolution && s.accumulate)
        return false;
    return true;
}

// The sample loop of render() below with CameraRays fused into FullPathTracer and up to samples_in_flight passes per wavefront.
// Seeds, sample numbers and counters of every pass are drawn exactly as the per-pass loop draws them.
void GPURenderer::render_sample_batches()
{
    HIPRTRenderSettings& s = m_render_data.render_settings;
    int done = 0;
    while (done < s.samples_per_frame)
    {
        const int n = std::min(std::min(samples_in_flight, 16), s.samples_per_frame - done);
        b200pt_sample_batch batch;
        std::memset(&batch, 0, sizeof(batch));
        batch.count = n;
        batch.write_g_buffer = keep_g_buffer ? 1 : 0;
        m_render_data.prev_camera = m_render_data.current_camera;
        m_render_data.current_camera = m_camera;
        const bool last_batch = done + n == s.samples_per_frame;
        s.do_update_status_buffers = (seed_schedule == SeedSchedule::CPU || last_batch) ? 1 : s.do_update_status_buffers;
        // the render data snapshot of the launch carries the FIRST pass' sample number; every pass' own values travel in the batch
        const int first_sample_number = s.sample_number;
        const int first_aov_counter = s.denoiser_AOV_accumulation_counter;
        const unsigned int first_seed = m_render_data.random_seed;
        for (int i = 0; i < n; i++)
        {
            batch.sample_numbers[i] = s.sample_number;
            batch.aov_counters[i] = s.denoiser_AOV_accumulation_counter;
            if (seed_schedule == SeedSchedule::GPU)
            {
                batch.camera_seeds[i] = m_rng.xorshift32();      // launch_camera_rays
                batch.path_seeds[i] = m_rng.xorshift32();        // launch_path_tracing
                s.sample_number++;
                s.denoiser_AOV_accumulation_counter++;
            }
            else
            {
                batch.camera_seeds[i] = m_render_data.random_seed;
                batch.path_seeds[i] = m_render_data.random_seed;
                if (s.accumulate)
                    s.sample_number++;
                m_render_data.random_seed = m_rng.xorshift32();  // CPURenderer.cpp:284-286
            }
        }
        const int next_sample_number = s.sample_number;
        const int next_aov_counter = s.denoiser_AOV_accumulation_counter;
        const unsigned int next_seed = m_render_data.random_seed;
        s.sample_number = first_sample_number;
        s.denoiser_AOV_accumulation_counter = first_aov_counter;
        m_render_data.random_seed = seed_schedule == SeedSchedule::GPU ? batch.path_seeds[0] : first_seed;

        m_time_kernels = last_batch;
        if (m_time_kernels)
        {
            // CameraRays has no launch of its own here: its interval is empty, the fused wavefront is timed as FullPathTracer
            check(b200pt_event_record(m_ctx, m_kernel_events[0], nullptr), "b200pt_event_record");
            check(b200pt_event_record(m_ctx, m_kernel_events[1], nullptr), "b200pt_event_record");
            check(b200pt_event_record(m_ctx, m_kernel_events[4], nullptr), "b200pt_event_record");
            check(b200pt_event_record(m_ctx, m_kernel_events[5], nullptr), "b200pt_event_record");
            check(b200pt_event_record(m_ctx, m_kernel_events[2], nullptr), "b200pt_event_record");
        }
        launch_path_tracing(&batch);
        if (m_time_kernels)
            check(b200pt_event_record(m_ctx, m_kernel_events[3], nullptr), "b200pt_event_record");

        s.sample_number = next_sample_number;
        s.denoiser_AOV_accumulation_counter = next_aov_counter;
        m_render_data.random_seed = next_seed;
        s.need_to_reset = 0;
        done += n;
    }
}

void GPURenderer::render()
{
    HIPRTRenderSettings& s = m_render_data.render_settings;
    check(b200pt_event_record(m_ctx, m_frame_start, nullptr), "b200pt_event_record");
    if (can_render_sample_batches())
    {
        render_sample_batches();
        check(b200pt_event_record(m_ctx, m_frame_stop, nullptr), "b200pt_event_record");
        return;
    }
    for (int i = 1; i <= s.samples_per_frame; i++)
    {
        m_render_data.prev_camera = m_render_data.current_camera;
        m_render_data.current_camera = m_camera;

        if (seed_schedule == SeedSchedule::CPU)
            s.do_update_status_buffers = 1;                         // CPURenderer.cpp:273
        else if (i == s.samples_per_frame)
            s.do_update_status_buffers = 1;                         // GPURenderer.cpp:430-434

        m_time_kernels = i == s.samples_per_frame;
        if (m_time_kernels) check(b200pt_event_record(m_ctx, m_kernel_events[0], nullptr), "b200pt_event_record");
        launch_camera_rays();
        if (m_time_kernels) check(b200pt_event_record(m_ctx, m_kernel_events[1], nullptr), "b200pt_event_record");
        if (m_time_kernels) check(b200pt_event_record(m_ctx, m_kernel_events[4], nullptr), "b200pt_event_record");
        launch_ReSTIR_DI();
        if (m_time_kernels) check(b200pt_event_record(m_ctx, m_kernel_events[5], nullptr), "b200pt_event_record");
        if (m_time_kernels) check(b200pt_event_record(m_ctx, m_kernel_events[2], nullptr), "b200pt_event_record");
        launch_path_tracing();
        if (m_time_kernels) check(b200pt_event_record(m_ctx, m_kernel_events[3], nullptr), "b200pt_event_record");

        if (seed_schedule == SeedSchedule::CPU)
        {
            if (s.accumulate)
                s.sample_number++;
            m_render_data.random_seed = m_rng.xorshift32();          // CPURenderer.cpp:284-286
        }
        else
        {
            s.sample_number++;
            s.denoiser_AOV_accumulation_counter++;
        }
        s.need_to_reset = 0;
        s.restir_di_settings.temporal_pass.temporal_buffer_clear_requested = 0;
    }
    check(b200pt_event_record(m_ctx, m_frame_stop, nullptr), "b200pt_event_record");
}

void GPURenderer::launch_camera_rays()
{
    if (seed_schedule == SeedSchedule::GPU)
        m_render_data.random_seed = m_rng.xorshift32();
    // A ReSTIR DI tile also needs the G-buffer (and last frame's) of the halo rows its spatiotemporal pass reads: CameraRays is
    // deterministic per pixel, so the rank recomputes those rows instead of receiving 344 B/px from the neighbouring rank
    b200pt_tile camera_tile = m_tile;
    if (m_has_tile && m_tile_halo > 0 && m_tile.y1 > m_tile.y0)
    {
        camera_tile.y0 = std::max(0, m_tile.y0 - m_tile_halo);
        camera_tile.y1 = std::min(m_res.y, m_tile.y1 + m_tile_halo);
    }
    check(b200pt_launch_CameraRays(m_ctx, &m_render_data, m_res, m_has_tile ? &camera_tile : nullptr, nullptr), "b200pt_launch_CameraRays");
}

void GPURenderer::launch_path_tracing(const b200pt_sample_batch* batch)
{
    if (seed_schedule == SeedSchedule::GPU && batch == nullptr)
        m_render_data.random_seed = m_rng.xorshift32();
    b200pt_tile whole = m_has_tile ? m_tile : b200pt_tile{0, 0, m_res.x, m_res.y};
    const int columns = whole.x1 - whole.x0;
    int lanes = m_path_lanes;
    while (lanes > 1 && columns < 256 * lanes)
        lanes--;
    if (lanes <= 1)
    {
        if (batch)
            check(b200pt_launch_PathTraceSamples(m_ctx, &m_render_data, m_res, m_has_tile ? &m_tile : nullptr, nullptr, batch, 0, 1), "b200pt_launch_PathTraceSamples");
        else
            check(b200pt_launch_FullPathTracer(m_ctx, &m_render_data, m_res, m_has_tile ? &m_tile : nullptr, nullptr), "b200pt_launch_FullPathTracer");
        return;
    }
    // Bands are COLUMN ranges of whole 8-pixel blocks (a landscape frame is sky above and geometry below: row bands would
    // give one lane all the work); lane 0 runs on the main stream, the others fork from it and join back
    check(b200pt_event_record(m_ctx, m_lane_fork, nullptr), "b200pt_event_record");
    const int blocks = (columns + 7) / 8;
    for (int lane = lanes - 1; lane >= 0; lane--)
    {
        b200pt_tile band = whole;
        band.x0 = whole.x0 + 8 * ((blocks * lane) / lanes);
        band.x1 = lane + 1 == lanes ? whole.x1 : whole.x0 + 8 * ((blocks * (lane + 1)) / lanes);
        b200pt_stream st = lane == 0 ? nullptr : m_lane_streams[lane - 1];
        if (lane > 0)
            check(b200pt_stream_wait_event(m_ctx, st, m_lane_fork), "b200pt_stream_wait_event");
        if (batch)
            check(b200pt_launch_PathTraceSamples(m_ctx, &m_render_data, m_res, &band, st, batch, lane, lanes), "b200pt_launch_PathTraceSamples");
        else
            check(b200pt_launch_FullPathTracer_lane(m_ctx, &m_render_data, m_res, &band, st, lane, lanes), "b200pt_launch_FullPathTracer_lane");
        if (lane > 0)
            check(b200pt_event_record(m_ctx, m_lane_join[lane - 1], st), "b200pt_event_record");
    }
    for (int lane = 1; lane < lanes; lane++)
        check(b200pt_stream_wait_event(m_ctx, nullptr, m_lane_join[lane - 1]), "b200pt_stream_wait_event");
}

void GPURenderer::launch_ReSTIR_DI()
{
    if (m_options.direct_light_sampling_strategy != 5)
        return;

    // ReSTIRDIRenderPass::launch (ReSTIRDIRenderPass.cpp:306-352) / CPURenderer::ReSTIR_DI (CPURenderer.cpp:401-434):
    // presampling -> initial candidates -> fused spatiotemporal reuse, each with its own seed draw
    HIPRTRenderData& rd = m_render_data;
    ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;
    if (rs.do_fused_spatiotemporal && !(rs.temporal_pass.do_temporal_reuse_pass && rs.spatial_pass.do_spatial_reuse_pass))
        throw std::runtime_error("ReSTIR DI: do_fused_spatiotemporal needs both the temporal and the spatial reuse pass enabled (ReSTIRDIRenderPass.cpp:268-274)");
    const b200pt_tile* tile = m_has_tile ? &m_tile : nullptr;

    // ReSTIRDIRenderPass::launch_presampling_lights_pass / configure (ReSTIRDIRenderPass.cpp:354-386)
    LightPresamplingParameters p;
    std::memset(&p, 0, sizeof(p));
    p.number_of_subsets = rs.light_presampling.number_of_subsets;
    p.subset_size = rs.light_presampling.subset_size;
    p.out_light_samples = m_restir_presampled_lights.get_device_pointer();
    p.emissive_triangles_count = rd.buffers.emissive_triangles_count;
    p.emissive_triangles_indices = rd.buffers.emissive_triangles_indices;
    p.triangles_indices = rd.buffers.triangles_indices;
    p.vertices_positions = rd.buffers.vertices_positions;
    p.material_indices = rd.buffers.material_indices;
    p.materials = rd.buffers.materials_buffer;
    p.world_settings = rd.world_settings;
    p.freeze_random = rd.render_settings.freeze_random;
    p.sample_number = rd.render_settings.sample_number;
    p.random_seed = m_rng.xorshift32();
    p.envmap_sampling_probability = rs.initial_candidates.envmap_candidate_probability;
    check(b200pt_launch_ReSTIR_DI_LightsPresampling(m_ctx, &p, nullptr), "b200pt_launch_ReSTIR_DI_LightsPresampling");

    // initial candidates (ReSTIRDIRenderPass.cpp:388-403)
    rd.random_seed = m_rng.xorshift32();
    rs.light_presampling.light_samples = m_restir_presampled_lights.get_device_pointer();
    rs.initial_candidates.output_reservoirs = m_restir_initial.get_device_pointer();
    check(b200pt_launch_ReSTIR_DI_InitialCandidates(m_ctx, &rd, m_res, tile, nullptr), "b200pt_launch_ReSTIR_DI_InitialCandidates");

    ReSTIRDIReservoir* const spatial_1 = m_restir_spatial_1.get_device_pointer();
    ReSTIRDIReservoir* const spatial_2 = m_restir_spatial_2.get_device_pointer();
    // The CPU seed schedule mirrors CPURenderer::ReSTIR_DI (CPURenderer.cpp:401-434, configure_* :503-633), which never updates
    // spatial_pass_index and runs no extra spatial pass after the fused one; the GPU schedule mirrors ReSTIRDIRenderPass
    const bool cpu_schedule = seed_schedule == SeedSchedule::CPU;
    if (rs.do_fused_spatiotemporal)
    {
        // fused spatiotemporal reuse (ReSTIRDIRenderPass.cpp:378-390,433-455,476-503): last frame's output is this frame's temporal input
        rd.random_seed = m_rng.xorshift32();
        rs.temporal_pass.permutation_sampling_random_bits = m_rng.xorshift32();
        ReSTIRDIReservoir* previous = m_restir_odd_frame ? spatial_1 : spatial_2;
        rs.temporal_pass.input_reservoirs = previous;
        rs.temporal_pass.output_reservoirs = nullptr;
        rs.spatial_pass.input_reservoirs = previous;
        rs.spatial_pass.output_reservoirs = previous == spatial_1 ? spatial_2 : spatial_1;
        if (!cpu_schedule)
            rs.spatial_pass.spatial_pass_index = 0;
        check(b200pt_launch_ReSTIR_DI_SpatiotemporalReuse(m_ctx, &rd, m_res, tile, nullptr), "b200pt_launch_ReSTIR_DI_SpatiotemporalReuse");

        // the remaining spatial passes of number_of_passes (:492-501), ping-ponging between the two spatial buffers
        for (int pass = 1; !cpu_schedule && pass < rs.spatial_pass.number_of_passes; pass++)
        {
            if (m_has_tile && m_tile_halo > 0)
                throw std::runtime_error("ReSTIR DI tile sharding exchanges one halo per frame: number_of_passes > 1 is not supported on a tile with a halo");
            rs.spatial_pass.spatial_pass_index = pass;
            rd.random_seed = m_rng.xorshift32();
            rs.spatial_pass.input_reservoirs = rs.spatial_pass.output_reservoirs;
            rs.spatial_pass.output_reservoirs = rs.spatial_pass.input_reservoirs == spatial_1 ? spatial_2 : spatial_1;
            check(b200pt_launch_ReSTIR_DI_SpatialReuse(m_ctx, &rd, m_res, tile, nullptr), "b200pt_launch_ReSTIR_DI_SpatialReuse");
        }
        // With an even number of passes the output lands in last frame's buffer: the next frame's temporal input is
        // restir_output_reservoirs (ReSTIRDIRenderPass.cpp:384), so m_restir_odd_frame follows the buffer, not the frame count
        m_restir_odd_frame = rs.spatial_pass.output_reservoirs == spatial_1;
    }
    else
    {
        if (m_has_tile && m_tile_halo > 0)
            throw std::runtime_error("ReSTIR DI tile sharding with a halo is built for the fused spatiotemporal pass only");
        if (rs.temporal_pass.do_temporal_reuse_pass)
        {
            // configure_temporal_pass (ReSTIRDIRenderPass.cpp:332-366)
            rd.random_seed = m_rng.xorshift32();
            rs.temporal_pass.permutation_sampling_random_bits = m_rng.xorshift32();
            rs.temporal_pass.input_reservoirs = rs.restir_output_reservoirs;
            if (rs.spatial_pass.do_spatial_reuse_pass)
                rs.temporal_pass.output_reservoirs = m_restir_initial.get_device_pointer();     // in place: a pixel touches only its own entry
            else
                rs.temporal_pass.output_reservoirs = m_restir_odd_frame ? spatial_1 : spatial_2;
            check(b200pt_launch_ReSTIR_DI_TemporalReuse(m_ctx, &rd, m_res, tile, nullptr), "b200pt_launch_ReSTIR_DI_TemporalReuse");
        }
        if (rs.spatial_pass.do_spatial_reuse_pass)
        {
            for (int pass = 0; pass < rs.spatial_pass.number_of_passes; pass++)
            {
                // configure_spatial_pass (ReSTIRDIRenderPass.cpp:392-431)
                rd.random_seed = m_rng.xorshift32();
                if (!cpu_schedule)
                    rs.spatial_pass.spatial_pass_index = pass;
                if (pass == 0)
                {
                    rs.spatial_pass.input_reservoirs = rs.temporal_pass.do_temporal_reuse_pass ? rs.temporal_pass.output_reservoirs : rs.initial_candidates.output_reservoirs;
                    rs.spatial_pass.output_reservoirs = spatial_1;
                }
                else
                {
                    rs.spatial_pass.input_reservoirs = (pass & 1) == 0 ? spatial_2 : spatial_1;
                    rs.spatial_pass.output_reservoirs = (pass & 1) == 0 ? spatial_1 : spatial_2;
                }
                check(b200pt_launch_ReSTIR_DI_SpatialReuse(m_ctx, &rd, m_res, tile, nullptr), "b200pt_launch_ReSTIR_DI_SpatialReuse");
            }
        }
        m_restir_odd_frame = !m_restir_odd_frame;
    }

    // configure_output_buffer (ReSTIRDIRenderPass.cpp:550-570)
    if (rs.spatial_pass.do_spatial_reuse_pass || rs.do_fused_spatiotemporal)
        rs.restir_output_reservoirs = rs.spatial_pass.output_reservoirs;
    else if (rs.temporal_pass.do_temporal_reuse_pass)
        rs.restir_output_reservoirs = rs.temporal_pass.output_reservoirs;
    else
        rs.restir_output_reservoirs = rs.initial_candidates.output_reservoirs;
}
} // namespace b200pt_host
