// Second half of the C ABI: textures, acceleration structure, kernel launches, ray batches, counters.

extern "C" {

int b200pt_tex2d_create(b200pt_ctx* ctx, const void* host_texels, int width, int height, int channels, int bytes_per_channel, void** out_descriptor_dptr)
{
    REQUIRE(ctx != nullptr && host_texels != nullptr && out_descriptor_dptr != nullptr);
    REQUIRE(width > 0 && height > 0 && channels >= 1 && channels <= 4 && (bytes_per_channel == 1 || bytes_per_channel == 4));
    CUDA_TRY(cudaSetDevice(ctx->device));
    size_t bytes = (size_t)width * height * channels * bytes_per_channel;
    // descriptor and texels share one allocation: [b200pt_tex2d padded to 64 B][texels]
    unsigned char* block = nullptr;
    CUDA_TRY(cudaMalloc(&block, 64 + bytes));
    b200pt_tex2d desc;
    desc.texels = block + 64;
    desc.width = width;
    desc.height = height;
    desc.channels = channels;
    desc.bytes_per_channel = bytes_per_channel;
    cudaError_t e = cudaMemcpy(block, &desc, sizeof(desc), cudaMemcpyHostToDevice);
    if (e == cudaSuccess)
        e = cudaMemcpy(block + 64, host_texels, bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess)
    {
        cudaFree(block);
        return cuda_fail(e, "texture upload");
    }
    *out_descriptor_dptr = block;
    return B200PT_SUCCESS;
}

int b200pt_tex3d_create(b200pt_ctx* ctx, const float* host_texels, int width, int height, int depth, int channels, void** out_descriptor_dptr)
{
    REQUIRE(ctx != nullptr && host_texels != nullptr && out_descriptor_dptr != nullptr);
    REQUIRE(width > 0 && height > 0 && depth > 0 && channels >= 1 && channels <= 4);
    CUDA_TRY(cudaSetDevice(ctx->device));
    size_t bytes = (size_t)width * height * depth * channels * sizeof(float);
    unsigned char* block = nullptr;
    CUDA_TRY(cudaMalloc(&block, 64 + bytes));
    b200pt_tex3d desc;
    desc.texels = reinterpret_cast<const float*>(block + 64);
    desc.width = width;
    desc.height = height;
    desc.depth = depth;
    desc.channels = channels;
    cudaError_t e = cudaMemcpy(block, &desc, sizeof(desc), cudaMemcpyHostToDevice);
    if (e == cudaSuccess)
        e = cudaMemcpy(block + 64, host_texels, bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess)
    {
        cudaFree(block);
        return cuda_fail(e, "texture upload");
    }
    *out_descriptor_dptr = block;
    return B200PT_SUCCESS;
}

int b200pt_tex_destroy(b200pt_ctx* ctx, void* descriptor_dptr) { return b200pt_free(ctx, descriptor_dptr); }

int b200pt_tex2d_array_create(b200pt_ctx* ctx, void* const* descriptor_dptrs, int count, void** out_array_dptr)
{
    REQUIRE(ctx != nullptr && out_array_dptr != nullptr && count >= 0 && (count == 0 || descriptor_dptrs != nullptr));
    *out_array_dptr = nullptr;
    if (count == 0)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaSetDevice(ctx->device));
    b200pt_tex2d* array = nullptr;
    CUDA_TRY(cudaMalloc(&array, sizeof(b200pt_tex2d) * (size_t)count));
    for (int i = 0; i < count; i++)
    {
        cudaError_t e = cudaMemcpy(array + i, descriptor_dptrs[i], sizeof(b200pt_tex2d), cudaMemcpyDeviceToDevice);
        if (e != cudaSuccess)
        {
            cudaFree(array);
            return cuda_fail(e, "texture array upload");
        }
    }
    *out_array_dptr = array;
    return B200PT_SUCCESS;
}

int b200pt_bvh_build(b200pt_ctx* ctx, const float* vertices, int vertex_count, const int* triangle_indices, int triangle_count, unsigned flags, b200pt_bvh** out_bvh)
{
    REQUIRE(ctx != nullptr && out_bvh != nullptr && triangle_count >= 0 && vertex_count >= 0);
    REQUIRE(triangle_count == 0 || (vertices != nullptr && triangle_indices != nullptr));
    *out_bvh = nullptr;
    CUDA_TRY(cudaSetDevice(ctx->device));

    // B200PT_BVH_BUILD_DEVICE (flag, or the environment variable of the same name set to 1): Morton order + PLOC (or radix tree) + SAH-optimal collapse on the GPU
    // (bvh8_build_device.cu) instead of the host's binned-SAH builder, which stays the quality baseline and the default
    static const bool device_by_env = [] { const char* v = getenv("B200PT_BVH_BUILD_DEVICE"); return v && atoi(v) != 0; }();
    if (((flags & B200PT_BVH_BUILD_DEVICE) || device_by_env) && triangle_count > 0)
    {
        auto t0 = std::chrono::steady_clock::now();
        pt::Bvh8DeviceBuild built;
        static const int algorithm_by_env = [] { const char* v = getenv("B200PT_BVH_DEVICE_ALGORITHM"); return v == nullptr ? -1 : (strcmp(v, "lbvh") == 0 ? (int)pt::BVH8_DEVICE_LBVH : (int)pt::BVH8_DEVICE_PLOC); }();
        static const int ploc_radius = [] { const char* v = getenv("B200PT_PLOC_RADIUS"); int r = v ? atoi(v) : 0; return r > 0 ? (r > 256 ? 256 : r) : 16; }();
        const int algorithm = (flags & B200PT_BVH_BUILD_DEVICE_LBVH) ? (int)pt::BVH8_DEVICE_LBVH : (algorithm_by_env >= 0 ? algorithm_by_env : (int)pt::BVH8_DEVICE_PLOC);
        cudaError_t e = pt::build_bvh8_device(vertices, vertex_count, triangle_indices, triangle_count, ctx->main_stream, &built, algorithm, ploc_radius);
        if (e != cudaSuccess)
            return cuda_fail(e, "device BVH8 build");
        if (2 * built.max_depth + 1 > pt::kTraversalStackSize)
        {
            cudaFree(built.nodes);
            cudaFree(built.triangles);
            char msg[160];
            snprintf(msg, sizeof(msg), "BVH8 depth %d needs %d traversal-stack entries, kTraversalStackSize is %d", built.max_depth, 2 * built.max_depth + 1, pt::kTraversalStackSize);
            return fail(B200PT_ERROR_INVALID_ARGUMENT, msg);
        }
        b200pt_bvh* bvh = new b200pt_bvh();
        bvh->d_nodes = built.nodes;
        bvh->d_triangles = built.triangles;
        bvh->sah_cost = built.sah_cost;
        bvh->max_depth = built.max_depth;
        bvh->node_fill = built.node_count ? (float)((double)built.child_slots_used / (8.0 * (double)built.node_count)) : 0.0f;
        bvh->built_on_device = true;
        bvh->device_build_seconds = built.device_seconds;
        bvh->view.nodes = built.nodes;
        bvh->view.triangles = built.triangles;
        bvh->view.node_count = built.node_count;
        bvh->view.triangle_count = built.triangle_count;
        for (int a = 0; a < 3; a++)
        {
            bvh->view.scene_min[a] = built.scene_min[a];
            bvh->view.scene_max[a] = built.scene_max[a];
        }
        e = cudaMalloc(&bvh->d_view, sizeof(pt::Bvh8Device));
        if (e == cudaSuccess)
            e = cudaMemcpy(bvh->d_view, &bvh->view, sizeof(pt::Bvh8Device), cudaMemcpyHostToDevice);
        if (e != cudaSuccess)
        {
            cudaFree(bvh->d_nodes);
            cudaFree(bvh->d_triangles);
            cudaFree(bvh->d_view);
            delete bvh;
            return cuda_fail(e, "BVH descriptor upload");
        }
        bvh->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        *out_bvh = bvh;
        return B200PT_SUCCESS;
    }

    auto t0 = std::chrono::steady_clock::now();
    pt::Bvh8Host host = pt::build_bvh8(vertices, vertex_count, triangle_indices, triangle_count);
    double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    // The traversal stacks hold kTraversalStackSize entries: a ray of the ray-queue kernel defers at most one node group and one
    // triangle group per level. A deeper tree would drop subtrees silently, so it is refused here instead.
    if (2 * host.max_depth + 1 > pt::kTraversalStackSize)
    {
        char msg[160];
        snprintf(msg, sizeof(msg), "BVH8 depth %d needs %d traversal-stack entries, kTraversalStackSize is %d", host.max_depth, 2 * host.max_depth + 1, pt::kTraversalStackSize);
        return fail(B200PT_ERROR_INVALID_ARGUMENT, msg);
    }
    b200pt_bvh* bvh = new b200pt_bvh();
    bvh->sah_cost = host.sah_cost;
    bvh->build_seconds = seconds;
    bvh->max_depth = host.max_depth;
    bvh->node_fill = host.nodes.empty() ? 0.0f : (float)((double)host.child_slots_used / (8.0 * (double)host.nodes.size()));
    size_t node_bytes = host.nodes.size() * sizeof(pt::Bvh8Node);
    size_t tri_bytes = host.triangles.size() * sizeof(pt::Bvh8Triangle);
    cudaError_t e = cudaMalloc(&bvh->d_nodes, node_bytes);
    if (e == cudaSuccess)
        e = cudaMalloc(&bvh->d_triangles, tri_bytes > 0 ? tri_bytes : 16);
    if (e == cudaSuccess)
        e = cudaMalloc(&bvh->d_view, sizeof(pt::Bvh8Device));
    if (e == cudaSuccess)
        e = cudaMemcpy(bvh->d_nodes, host.nodes.data(), node_bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && tri_bytes > 0)
        e = cudaMemcpy(bvh->d_triangles, host.triangles.data(), tri_bytes, cudaMemcpyHostToDevice);
    bvh->view.nodes = reinterpret_cast<const pt::Bvh8Node*>(bvh->d_nodes);
    bvh->view.triangles = reinterpret_cast<const pt::Bvh8Triangle*>(bvh->d_triangles);
    bvh->view.node_count = (uint32_t)host.nodes.size();
    bvh->view.triangle_count = (uint32_t)host.triangles.size();
    for (int a = 0; a < 3; a++)
    {
        bvh->view.scene_min[a] = host.scene_min[a];
        bvh->view.scene_max[a] = host.scene_max[a];
    }
    if (e == cudaSuccess)
        e = cudaMemcpy(bvh->d_view, &bvh->view, sizeof(pt::Bvh8Device), cudaMemcpyHostToDevice);
    if (e != cudaSuccess)
    {
        cudaFree(bvh->d_nodes);
        cudaFree(bvh->d_triangles);
        cudaFree(bvh->d_view);
        delete bvh;
        return cuda_fail(e, "BVH upload");
    }
    *out_bvh = bvh;
    return B200PT_SUCCESS;
}

int b200pt_bvh_destroy(b200pt_ctx* ctx, b200pt_bvh* bvh)
{
    REQUIRE(ctx != nullptr);
    if (!bvh)
        return B200PT_SUCCESS;
    cudaSetDevice(ctx->device);
    cudaFree(bvh->d_nodes);
    cudaFree(bvh->d_triangles);
    cudaFree(bvh->d_shading);
    cudaFree(bvh->d_view);
    delete bvh;
    return B200PT_SUCCESS;
}

void* b200pt_bvh_device_handle(b200pt_bvh* bvh) { return bvh ? bvh->d_view : nullptr; }

// One thread per triangle slot of the BVH: gather what a hit on that triangle needs for shading (see Bvh8Device::shading, bvh8.h)
static __global__ void k_attach_shading(pt::Bvh8Triangle* triangles, float4* shading, uint32_t triangle_count, const int* __restrict__ triangle_indices,
                                        const float* __restrict__ vertex_normals, const unsigned char* __restrict__ has_vertex_normals, const float* __restrict__ texcoords,
                                        const int* __restrict__ material_indices)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= triangle_count)
        return;
    const int prim = triangles[slot].prim_id;
    const int a = triangle_indices[3 * (size_t)prim], b = triangle_indices[3 * (size_t)prim + 1], c = triangle_indices[3 * (size_t)prim + 2];
    const float* na = vertex_normals + 3 * (size_t)a;
    const float* nb = vertex_normals + 3 * (size_t)b;
    const float* nc = vertex_normals + 3 * (size_t)c;
    const float* ta = texcoords + 2 * (size_t)a;
    const float* tb = texcoords + 2 * (size_t)b;
    const float* tc = texcoords + 2 * (size_t)c;
    float4* out = shading + 4 * (size_t)slot;
    out[0] = make_float4(na[0], na[1], na[2], nb[0]);
    out[1] = make_float4(nb[1], nb[2], nc[0], nc[1]);
    out[2] = make_float4(nc[2], ta[0], ta[1], tb[0]);
    out[3] = make_float4(tb[1], tc[0], tc[1], 0.0f);
    triangles[slot].pad1 = material_indices[prim];
    triangles[slot].pad2 = has_vertex_normals[a] ? 1 : 0;
}

int b200pt_bvh_attach_shading_data(b200pt_ctx* ctx, b200pt_bvh* bvh, const int* d_triangle_indices, const float* d_vertex_normals, const unsigned char* d_has_vertex_normals,
                                   const float* d_texcoords, const int* d_material_indices, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && bvh != nullptr);
    CUDA_TRY(cudaSetDevice(ctx->device));
    if (d_triangle_indices == nullptr)
    {
        // detach: hits go back to the reference's buffers
        bvh->view.shading = nullptr;
        CUDA_TRY(cudaMemcpyAsync(bvh->d_view, &bvh->view, sizeof(pt::Bvh8Device), cudaMemcpyHostToDevice, stream_of(ctx, stream)));
        CUDA_TRY(cudaStreamSynchronize(stream_of(ctx, stream)));
        return B200PT_SUCCESS;
    }
    REQUIRE(d_vertex_normals != nullptr && d_has_vertex_normals != nullptr && d_texcoords != nullptr && d_material_indices != nullptr);
    const uint32_t n = bvh->view.triangle_count;
    if (n == 0)
        return B200PT_SUCCESS;
    if (bvh->d_shading == nullptr)
        CUDA_TRY(cudaMalloc(&bvh->d_shading, (size_t)n * 64));
    k_attach_shading<<<(n + 255) / 256, 256, 0, stream_of(ctx, stream)>>>(reinterpret_cast<pt::Bvh8Triangle*>(bvh->d_triangles), reinterpret_cast<float4*>(bvh->d_shading), n,
                                                                           d_triangle_indices, d_vertex_normals, d_has_vertex_normals, d_texcoords, d_material_indices);
    CUDA_TRY(cudaGetLastError());
    bvh->view.shading = bvh->d_shading;
    CUDA_TRY(cudaMemcpyAsync(bvh->d_view, &bvh->view, sizeof(pt::Bvh8Device), cudaMemcpyHostToDevice, stream_of(ctx, stream)));
    CUDA_TRY(cudaStreamSynchronize(stream_of(ctx, stream)));     // bvh->view is host memory of this call's lifetime only through the copy above
    return B200PT_SUCCESS;
}

int b200pt_bvh_info(b200pt_bvh* bvh, unsigned* out_node_count, unsigned* out_triangle_count, float* out_sah_cost, double* out_build_seconds)
{
    REQUIRE(bvh != nullptr);
    if (out_node_count) *out_node_count = bvh->view.node_count;
    if (out_triangle_count) *out_triangle_count = bvh->view.triangle_count;
    if (out_sah_cost) *out_sah_cost = bvh->sah_cost;
    if (out_build_seconds) *out_build_seconds = bvh->build_seconds;
    return B200PT_SUCCESS;
}

int b200pt_bvh_get_stats(b200pt_bvh* bvh, b200pt_bvh_stats* out_stats)
{
    REQUIRE(bvh != nullptr && out_stats != nullptr);
    out_stats->node_count = bvh->view.node_count;
    out_stats->triangle_count = bvh->view.triangle_count;
    out_stats->max_depth = bvh->max_depth;
    out_stats->built_on_device = bvh->built_on_device ? 1 : 0;
    out_stats->sah_cost = bvh->sah_cost;
    out_stats->node_fill = bvh->node_fill;
    out_stats->build_seconds = bvh->build_seconds;
    out_stats->device_build_seconds = bvh->device_build_seconds;
    return B200PT_SUCCESS;
}

int b200pt_kernel_variant_select(b200pt_ctx* ctx, const b200pt_options* options)
{
    REQUIRE(ctx != nullptr && options != nullptr);
    pt::VariantOptions v = {options->direct_light_sampling_strategy, options->envmap_sampling_strategy, options->restir_di_later_bounces_strategy, options->ggx_sample_function};
    if (!pt::variant_available_t(v))
    {
        char msg[256];
        snprintf(msg, sizeof(msg), "kernel permutation (LSS=%d, ESS=%d, later=%d, GGX=%d) is not compiled into this build (see csrc/variants.h)", v.lss, v.ess, v.later, v.ggx);
        return fail(B200PT_ERROR_VARIANT_NOT_COMPILED, msg);
    }
    ctx->options = v;
    return B200PT_SUCCESS;
}

// Grows the context's wavefront scratch (wavefront.cuh) to `paths` paths and `bounces` bounces. Reallocation waits for
// the stream: launches in flight may still use the old arrays.
extern "C" int b200pt_path_profile_read(b200pt_ctx* ctx, double* out_stage_ms, int reset);

static int ensure_wavefront(b200pt_ctx* ctx, int lane, uint32_t paths, uint32_t bounces)
{
    pt::WavefrontState& wf = ctx->wf[lane];
    std::vector<void*>& allocations = ctx->wf_allocations[lane];
    // The seven per-path float4 records in three groups (launch.h: WavefrontState::stride_*): one array of 32- / 64- / 32-byte records per group by
    // default (B200PT_PATH_RECORDS unset or 1); 0: seven separate arrays (round 1's layout); 2: the NEE records (64 B) and the ray-queue
    // kernel's hit records (32 B) grouped as well -- measured neutral on C3 and -2 % on C5 (gpurun calls 39 / 41), so not the default. Results do
    // not depend on the layout. Read at every call so that a test can switch within one process: a change of layout reallocates.
    const char* records_env = getenv("B200PT_PATH_RECORDS");
    const bool grouped = !(records_env && records_env[0] == '0');
    const bool grouped_rays = records_env && records_env[0] == '2';
    if (paths <= wf.capacity && bounces <= wf.max_bounces && wf.stride_ray == (grouped ? 2u : 1u) && wf.stride_nee == (grouped_rays ? 4u : 1u))
        return B200PT_SUCCESS;
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaDeviceSynchronize());
    for (void* p : allocations)
        cudaFree(p);
    allocations.clear();
    uint32_t capacity = paths > wf.capacity ? paths : wf.capacity;
    uint32_t max_bounces = bounces > wf.max_bounces ? bounces : wf.max_bounces;
    wf = pt::WavefrontState{};
    auto alloc = [&](size_t bytes) -> void* {
        void* p = nullptr;
        if (cudaMalloc(&p, bytes) != cudaSuccess)
            return nullptr;
        allocations.push_back(p);
        return p;
    };
    const size_t n = capacity;
    bool ok = true;
    struct Group { float4** fields[3]; int count; uint32_t stride; };
    const Group groups[] = {{{&wf.origin_rng, &wf.dir_prim, nullptr}, 2, 2u}, {{&wf.hit_point, &wf.hit_ns, &wf.hit_ng}, 3, 4u}, {{&wf.throughput, &wf.color, nullptr}, 2, 2u}};
    for (const Group& g : groups)
    {
        if (grouped)
        {
            float4* base = static_cast<float4*>(alloc(n * g.stride * sizeof(float4)));
            ok &= base != nullptr;
            for (int k = 0; k < g.count; k++)
                *g.fields[k] = base ? base + k : nullptr;
        }
        else
            for (int k = 0; k < g.count; k++)
                ok &= (*g.fields[k] = static_cast<float4*>(alloc(n * sizeof(float4)))) != nullptr;
    }
    float4** f4[] = {&wf.final_color, &wf.aov_normal, &wf.aov_albedo, &wf.ris_a, &wf.ris_b};
    for (float4** a : f4)
        ok &= (*a = static_cast<float4*>(alloc(n * sizeof(float4)))) != nullptr;
    if (grouped_rays)
    {
        // 64-byte NEE records: origin | t_max, pending contribution, direction | primitive, spare
        float4* base = static_cast<float4*>(alloc(4 * n * 4 * sizeof(float4)));
        ok &= base != nullptr;
        wf.nee_ray_o = base;
        wf.nee_pending = base ? base + 1 : nullptr;
        wf.nee_ray_d = base ? base + 2 : nullptr;
    }
    else
    {
        float4** nee[] = {&wf.nee_ray_o, &wf.nee_ray_d, &wf.nee_pending};
        for (float4** a : nee)
            ok &= (*a = static_cast<float4*>(alloc(4 * n * sizeof(float4)))) != nullptr;
    }
    ok &= (wf.volume = static_cast<uint32_t*>(alloc(n * 36))) != nullptr;
    ok &= (wf.materials = alloc(n * 256)) != nullptr;
    ok &= (wf.ray_queue[0] = static_cast<uint32_t*>(alloc(n * 4))) != nullptr;
    ok &= (wf.ray_queue[1] = static_cast<uint32_t*>(alloc(n * 4))) != nullptr;
    ok &= (wf.shade_queue = static_cast<uint32_t*>(alloc(n * 4))) != nullptr;
    ok &= (wf.sorted_queue = static_cast<uint32_t*>(alloc(n * 4))) != nullptr;
    ok &= (wf.sort_key = static_cast<uint8_t*>(alloc(n))) != nullptr;
    ok &= (wf.bucket_count = static_cast<uint32_t*>(alloc(sizeof(uint32_t) * pt::kSortBuckets))) != nullptr;
    if (ok)
        ok &= cudaMemset(wf.bucket_count, 0, sizeof(uint32_t) * pt::kSortBuckets) == cudaSuccess;
    if (grouped_rays)
    {
        // 32-byte hit records: t, u, v, primitive | triangle slot, 3 spare words
        float4* base = static_cast<float4*>(alloc(2 * n * 2 * sizeof(float4)));
        ok &= base != nullptr;
        wf.hit_tuvp = base;
        wf.hit_tri_slot = base ? reinterpret_cast<uint32_t*>(base + 1) : nullptr;
    }
    else
    {
        ok &= (wf.hit_tuvp = static_cast<float4*>(alloc(2 * n * sizeof(float4)))) != nullptr;
        ok &= (wf.hit_tri_slot = static_cast<uint32_t*>(alloc(2 * n * 4))) != nullptr;
    }
    ok &= (wf.any_queue = static_cast<uint32_t*>(alloc(3 * n * 4))) != nullptr;
    ok &= (wf.closest_queue = static_cast<uint32_t*>(alloc(n * 4))) != nullptr;
    ok &= (wf.bin_hist = static_cast<uint32_t*>(alloc(sizeof(uint32_t) * (pt::kRayBinsMax + 1u)))) != nullptr;
    ok &= (wf.binned_any = static_cast<uint32_t*>(alloc(3 * n * 4))) != nullptr;
    ok &= (wf.binned_ext = static_cast<uint32_t*>(alloc(n * 4))) != nullptr;
    ok &= (wf.bin_keys = static_cast<uint32_t*>(alloc(3 * n * 4))) != nullptr;
    ok &= (wf.counts = static_cast<uint32_t*>(alloc(sizeof(uint32_t) * (pt::kCountsPerBounce + pt::kCursorsPerBounce) * (max_bounces + 2)))) != nullptr;
    if (!ok)
    {
        cudaError_t e = cudaGetLastError();
        for (void* p : allocations)
            cudaFree(p);
        allocations.clear();
        wf = pt::WavefrontState{};
        return cuda_fail(e == cudaSuccess ? cudaErrorMemoryAllocation : e, "wavefront scratch allocation");
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    wf.sm_count = (uint32_t)sms;
    wf.cursors = wf.counts + pt::kCountsPerBounce * (max_bounces + 2);
    wf.capacity = capacity;
    wf.max_bounces = max_bounces;
    wf.stride_ray = grouped ? 2u : 1u;
    wf.stride_hit = grouped ? 4u : 1u;
    wf.stride_radiance = grouped ? 2u : 1u;
    wf.stride_nee = grouped_rays ? 4u : 1u;
    wf.stride_hitrec = grouped_rays ? 2u : 1u;
    wf.stride_hitrec_words = grouped_rays ? 8u : 1u;
    return B200PT_SUCCESS;
}

#define LAUNCH_PIXEL_KERNEL(NAME, ...)                                                                                             \
    REQUIRE(ctx != nullptr && render_data != nullptr && res.x > 0 && res.y > 0);                                                   \
    REQUIRE(render_data->geom != nullptr);                                                                                         \
    pt::Tile t = tile_of(tile, res);                                                                                               \
    if (t.x1 <= t.x0 || t.y1 <= t.y0)                                                                                              \
        return B200PT_SUCCESS;                                                                                                     \
    cudaError_t e;                                                                                                                 \
    if (ctx->counting)                                                                                                             \
    {                                                                                                                              \
        HIPRTRenderData copy = *render_data;                                                                                       \
        copy.hiprt_function_table = ctx->d_counters;                                                                               \
        e = pt::NAME##_c(ctx->options, &copy, make_int2(res.x, res.y), t, stream_of(ctx, stream) __VA_ARGS__);                     \
    }                                                                                                                              \
    else                                                                                                                           \
        e = pt::NAME##_t(ctx->options, render_data, make_int2(res.x, res.y), t, stream_of(ctx, stream) __VA_ARGS__);               \
    if (e != cudaSuccess)                                                                                                          \
        return cuda_fail(e, #NAME);                                                                                                \
    return B200PT_SUCCESS;

int b200pt_launch_CameraRays(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream)
{
    LAUNCH_PIXEL_KERNEL(launch_CameraRays)
}

int b200pt_set_path_mode(b200pt_ctx* ctx, int mode)
{
    REQUIRE(ctx != nullptr && mode >= pt::WF_MEGAKERNEL && mode <= pt::WF_PHASED);
    ctx->path_mode = mode;
    return B200PT_SUCCESS;
}

int b200pt_get_path_mode(b200pt_ctx* ctx, int* out_mode)
{
    REQUIRE(ctx != nullptr && out_mode != nullptr);
    *out_mode = ctx->path_mode;
    return B200PT_SUCCESS;
}

int b200pt_launch_FullPathTracer(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream)
{
    return b200pt_launch_FullPathTracer_lane(ctx, render_data, res, tile, stream, 0, 1);
}

int b200pt_launch_FullPathTracer_lane(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream, int lane, int lane_count)
{
    REQUIRE(ctx != nullptr && render_data != nullptr && res.x > 0 && res.y > 0);
    REQUIRE(lane_count >= 1 && lane_count <= b200pt_ctx::kMaxPathLanes && lane >= 0 && lane < lane_count);
    if (ctx->path_mode != pt::WF_MEGAKERNEL)
    {
        pt::Tile wt = tile_of(tile, res);
        if (wt.x1 > wt.x0 && wt.y1 > wt.y0)
        {
            REQUIRE(render_data->render_settings.nb_bounces >= 0);
            int rc = ensure_wavefront(ctx, lane, (uint32_t)(wt.x1 - wt.x0) * (uint32_t)(wt.y1 - wt.y0), (uint32_t)render_data->render_settings.nb_bounces);
            if (rc != B200PT_SUCCESS)
                return rc;
            ctx->wf[lane].grid_share = (uint32_t)lane_count;
            static const bool slot_materials = [] { const char* v = getenv("B200PT_SLOT_MATERIALS"); return v && atoi(v) != 0; }();
            ctx->wf[lane].force_slot_materials = slot_materials ? 1u : 0u;
            ctx->wf[lane].fused_camera = 0u;
            ctx->wf[lane].samples = 1u;
            ctx->wf[lane].tile_pixels = (uint32_t)(wt.x1 - wt.x0) * (uint32_t)(wt.y1 - wt.y0);
        }
    }
    if (ctx->profiling && lane_count > 1)
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "per-stage profiling needs a single path lane (the stages of concurrent lanes overlap)");
    if (ctx->profiling)
    {
        // fold the previous launch's intervals in, then hand the event pool to this launch
        int rc = b200pt_path_profile_read(ctx, nullptr, 0);
        if (rc != B200PT_SUCCESS)
            return rc;
        ctx->profile.count = 0;
    }
    pt::WavefrontProfile* profile = ctx->profiling ? &ctx->profile : nullptr;
    LAUNCH_PIXEL_KERNEL(launch_FullPathTracer, , &ctx->wf[lane], ctx->path_mode, profile)
}

int b200pt_launch_PathTraceSamples(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream, const b200pt_sample_batch* batch, int lane,
                                   int lane_count)
{
    REQUIRE(ctx != nullptr && render_data != nullptr && batch != nullptr && res.x > 0 && res.y > 0);
    REQUIRE(render_data->geom != nullptr);
    REQUIRE(lane_count >= 1 && lane_count <= b200pt_ctx::kMaxPathLanes && lane >= 0 && lane < lane_count);
    REQUIRE(batch->count >= 1 && batch->count <= pt::kMaxSamplesInFlight);
    const HIPRTRenderSettings& rs = render_data->render_settings;
    REQUIRE(rs.nb_bounces >= 0);
    // A pass of the batch must not depend on the previous pass' result, and the phased MIS pipeline must be the one that runs
    const bool adaptive_buffers = (rs.stop_pixel_noise_threshold > 0.0f || rs.enable_adaptive_sampling) && rs.accumulate;
    const bool low_resolution = rs.wants_render_low_resolution && rs.allow_render_low_resolution && rs.accumulate;
    const bool strategy_ok = ctx->options.lss == 3 || (ctx->options.lss == 4 && rs.ris_settings.number_of_bsdf_candidates == 1);
    if (!strategy_ok || ctx->path_mode != pt::WF_PHASED || rs.number_of_light_samples != 1 || rs.do_alpha_testing || adaptive_buffers || low_resolution)
        return fail(B200PT_ERROR_VARIANT_NOT_COMPILED, "b200pt_launch_PathTraceSamples: fused sample batches need LSS_MIS_LIGHT_BSDF or LSS_RIS_BSDF_AND_LIGHT (one BSDF candidate) in the phased wavefront shape with one light sample, "
                                                        "no alpha testing, no adaptive sampling / pixel noise threshold and no low-resolution rendering; launch CameraRays and FullPathTracer separately");
    pt::Tile t = tile_of(tile, res);
    if (t.x1 <= t.x0 || t.y1 <= t.y0)
        return B200PT_SUCCESS;
    const uint32_t tile_pixels = (uint32_t)(t.x1 - t.x0) * (uint32_t)(t.y1 - t.y0);
    REQUIRE((unsigned long long)tile_pixels * (unsigned)batch->count < (1ull << 30));     // NEE ray ids are 3 * capacity + slot in 32 bits
    int rc = ensure_wavefront(ctx, lane, tile_pixels * (uint32_t)batch->count, (uint32_t)rs.nb_bounces);
    if (rc != B200PT_SUCCESS)
        return rc;
    pt::WavefrontState& wf = ctx->wf[lane];
    wf.grid_share = (uint32_t)lane_count;
    wf.force_slot_materials = 0u;
    wf.fused_camera = 1u;
    wf.samples = (uint32_t)batch->count;
    wf.tile_pixels = tile_pixels;
    pt::SampleBatch b;
    memset(&b, 0, sizeof(b));
    b.count = (uint32_t)batch->count;
    for (int i = 0; i < batch->count; i++)
    {
        b.camera_seed[i] = batch->camera_seeds[i];
        b.path_seed[i] = batch->path_seeds[i];
        b.sample_number[i] = batch->sample_numbers[i];
        b.aov_counter[i] = batch->aov_counters[i];
    }
    b.write_g_buffer = batch->write_g_buffer ? 1u : 0u;
    if (ctx->profiling && lane_count > 1)
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "per-stage profiling needs a single path lane (the stages of concurrent lanes overlap)");
    if (ctx->profiling)
    {
        rc = b200pt_path_profile_read(ctx, nullptr, 0);
        if (rc != B200PT_SUCCESS)
            return rc;
        ctx->profile.count = 0;
    }
    pt::WavefrontProfile* profile = ctx->profiling ? &ctx->profile : nullptr;
    cudaError_t e;
    if (ctx->counting)
    {
        HIPRTRenderData copy = *render_data;
        copy.hiprt_function_table = ctx->d_counters;
        e = pt::launch_PathTraceSamples_c(ctx->options, &copy, make_int2(res.x, res.y), t, stream_of(ctx, stream), &wf, &b, profile);
    }
    else
        e = pt::launch_PathTraceSamples_t(ctx->options, render_data, make_int2(res.x, res.y), t, stream_of(ctx, stream), &wf, &b, profile);
    if (e != cudaSuccess)
        return cuda_fail(e, "launch_PathTraceSamples");
    return B200PT_SUCCESS;
}

int b200pt_path_profile_enable(b200pt_ctx* ctx, int enable)
{
    REQUIRE(ctx != nullptr);
    if (enable && ctx->profile_events.empty())
    {
        CUDA_TRY(cudaSetDevice(ctx->device));
        ctx->profile_events.resize(4096);
        ctx->profile_stages.resize(4096);
        for (cudaEvent_t& e : ctx->profile_events)
            CUDA_TRY(cudaEventCreate(&e));
        ctx->profile.events = ctx->profile_events.data();
        ctx->profile.stages = ctx->profile_stages.data();
        ctx->profile.capacity = (int)ctx->profile_events.size();
    }
    ctx->profiling = enable != 0;
    ctx->profile.count = 0;
    for (double& v : ctx->stage_ms)
        v = 0.0;
    return B200PT_SUCCESS;
}

int b200pt_path_profile_read(b200pt_ctx* ctx, double* out_stage_ms, int reset)
{
    REQUIRE(ctx != nullptr);
    if (ctx->profile.count > 1)
    {
        CUDA_TRY(cudaEventSynchronize(ctx->profile.events[ctx->profile.count - 1]));
        for (int i = 0; i + 1 < ctx->profile.count; i++)
        {
            float ms = 0.0f;
            CUDA_TRY(cudaEventElapsedTime(&ms, ctx->profile.events[i], ctx->profile.events[i + 1]));
            int s = ctx->profile.stages[i];
            if (s >= 0 && s < 4)
                ctx->stage_ms[s] += ms;
        }
    }
    ctx->profile.count = 0;
    if (out_stage_ms)
        for (int s = 0; s < 4; s++)
            out_stage_ms[s] = ctx->stage_ms[s];
    if (reset)
        for (double& v : ctx->stage_ms)
            v = 0.0;
    return B200PT_SUCCESS;
}

int b200pt_launch_ReSTIR_DI_LightsPresampling(b200pt_ctx* ctx, const LightPresamplingParameters* parameters, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && parameters != nullptr && parameters->out_light_samples != nullptr);
    cudaError_t e = pt::launch_ReSTIR_DI_LightsPresampling_t(ctx->options, parameters, stream_of(ctx, stream));
    if (e == cudaErrorInvalidDeviceFunction)
        return fail(B200PT_ERROR_VARIANT_NOT_COMPILED, "ReSTIR DI passes need a LSS_RESTIR_DI kernel permutation (b200pt_kernel_variant_select)");
    if (e != cudaSuccess)
        return cuda_fail(e, "ReSTIR_DI_LightsPresampling");
    return B200PT_SUCCESS;
}

// Grows the scratch of the phased ReSTIR DI passes to `pixels` pixels (same discipline as ensure_wavefront). A failed allocation is not an
// error: the passes then run as single kernels with per-thread traversal (scratch capacity 0).
static const pt::RestirScratch* ensure_restir_scratch(b200pt_ctx* ctx, uint32_t pixels)
{
    pt::RestirScratch& rs = ctx->restir_scratch;
    if (!ctx->restir_phases)
        return nullptr;
    if (pixels <= rs.capacity)
        return &rs;
    if (cudaSetDevice(ctx->device) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess)
        return nullptr;
    for (void* p : ctx->restir_allocations)
        cudaFree(p);
    ctx->restir_allocations.clear();
    rs = pt::RestirScratch{};
    bool ok = true;
    auto alloc = [&](size_t bytes) -> void* {
        void* p = nullptr;
        if (!ok || cudaMalloc(&p, bytes) != cudaSuccess)
        {
            ok = false;
            return nullptr;
        }
        ctx->restir_allocations.push_back(p);
        return p;
    };
    const size_t n = pixels, rays = n * (size_t)pt::kRestirSites;
    rs.ray_o = static_cast<float4*>(alloc(rays * sizeof(float4)));
    rs.ray_d = static_cast<float4*>(alloc(rays * sizeof(float4)));
    rs.target = static_cast<float*>(alloc(rays * sizeof(float)));
    rs.hit_tuvp = static_cast<float4*>(alloc(n * sizeof(float4)));
    rs.hit_tri_slot = static_cast<uint32_t*>(alloc(n * sizeof(uint32_t)));
    rs.ic_state = static_cast<float4*>(alloc(4 * n * sizeof(float4)));
    rs.status = static_cast<uint32_t*>(alloc(n * sizeof(uint32_t)));
    rs.queue = static_cast<uint32_t*>(alloc(rays * sizeof(uint32_t)));
    rs.counts = static_cast<uint32_t*>(alloc(4 * sizeof(uint32_t)));
    if (!ok)
    {
        cudaGetLastError();
        for (void* p : ctx->restir_allocations)
            cudaFree(p);
        ctx->restir_allocations.clear();
        rs = pt::RestirScratch{};
        return nullptr;
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    rs.sm_count = (uint32_t)sms;
    rs.capacity = pixels;
    return &rs;
}

static uint32_t restir_tile_pixels(const b200pt_tile* tile, b200pt_int2 res)
{
    const pt::Tile t = tile_of(tile, res);
    return t.x1 > t.x0 && t.y1 > t.y0 ? (uint32_t)(t.x1 - t.x0) * (uint32_t)(t.y1 - t.y0) : 0u;
}

#define LAUNCH_RESTIR_PIXEL_KERNEL(NAME, ...)                                                                                      \
    REQUIRE(ctx != nullptr && render_data != nullptr && res.x > 0 && res.y > 0);                                                   \
    REQUIRE(render_data->geom != nullptr);                                                                                         \
    pt::Tile t = tile_of(tile, res);                                                                                               \
    if (t.x1 <= t.x0 || t.y1 <= t.y0)                                                                                              \
        return B200PT_SUCCESS;                                                                                                     \
    cudaError_t e;                                                                                                                 \
    if (ctx->counting)                                                                                                             \
    {                                                                                                                              \
        HIPRTRenderData copy = *render_data;                                                                                       \
        copy.hiprt_function_table = ctx->d_counters;                                                                               \
        e = pt::NAME##_c(ctx->options, &copy, make_int2(res.x, res.y), t, stream_of(ctx, stream) __VA_ARGS__);                     \
    }                                                                                                                              \
    else                                                                                                                           \
        e = pt::NAME##_t(ctx->options, render_data, make_int2(res.x, res.y), t, stream_of(ctx, stream) __VA_ARGS__);               \
    if (e == cudaErrorInvalidDeviceFunction)                                                                                       \
        return fail(B200PT_ERROR_VARIANT_NOT_COMPILED, "ReSTIR DI passes need a LSS_RESTIR_DI kernel permutation (b200pt_kernel_variant_select)"); \
    if (e != cudaSuccess)                                                                                                          \
        return cuda_fail(e, #NAME);                                                                                                \
    return B200PT_SUCCESS;

int b200pt_launch_ReSTIR_DI_InitialCandidates(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream)
{
    const pt::RestirScratch* scratch = ctx != nullptr && res.x > 0 && res.y > 0 ? ensure_restir_scratch(ctx, restir_tile_pixels(tile, res)) : nullptr;
    LAUNCH_RESTIR_PIXEL_KERNEL(launch_ReSTIR_DI_InitialCandidates, , scratch)
}

int b200pt_launch_ReSTIR_DI_SpatiotemporalReuse(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream)
{
    const pt::RestirScratch* scratch = ctx != nullptr && res.x > 0 && res.y > 0 ? ensure_restir_scratch(ctx, restir_tile_pixels(tile, res)) : nullptr;
    LAUNCH_RESTIR_PIXEL_KERNEL(launch_ReSTIR_DI_SpatiotemporalReuse, , scratch)
}

int b200pt_launch_ReSTIR_DI_TemporalReuse(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream)
{
    const pt::RestirScratch* scratch = ctx != nullptr && res.x > 0 && res.y > 0 ? ensure_restir_scratch(ctx, restir_tile_pixels(tile, res)) : nullptr;
    LAUNCH_RESTIR_PIXEL_KERNEL(launch_ReSTIR_DI_TemporalReuse, , scratch)
}

int b200pt_launch_ReSTIR_DI_SpatialReuse(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream)
{
    const pt::RestirScratch* scratch = ctx != nullptr && res.x > 0 && res.y > 0 ? ensure_restir_scratch(ctx, restir_tile_pixels(tile, res)) : nullptr;
    LAUNCH_RESTIR_PIXEL_KERNEL(launch_ReSTIR_DI_SpatialReuse, , scratch)
}

int b200pt_trace_closest(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const int* d_last_prim, long long n, int* d_out_prim, float* d_out_t, float* d_out_uv, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && bvh != nullptr && n >= 0 && (n == 0 || (d_rays != nullptr && d_out_prim != nullptr && d_out_t != nullptr)));
    cudaError_t e = ctx->counting ? pt::launch_trace_closest_c(bvh->view, d_rays, d_last_prim, n, d_out_prim, d_out_t, d_out_uv, ctx->d_counters, stream_of(ctx, stream))
                                  : pt::launch_trace_closest_t(bvh->view, d_rays, d_last_prim, n, d_out_prim, d_out_t, d_out_uv, nullptr, stream_of(ctx, stream));
    if (e != cudaSuccess)
        return cuda_fail(e, "trace_closest");
    return B200PT_SUCCESS;
}

int b200pt_trace_shadow(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const float* d_t_max, const int* d_last_prim, long long n, unsigned char* d_out_occluded, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && bvh != nullptr && n >= 0 && (n == 0 || (d_rays != nullptr && d_t_max != nullptr && d_out_occluded != nullptr)));
    cudaError_t e = ctx->counting ? pt::launch_trace_shadow_c(bvh->view, d_rays, d_t_max, d_last_prim, n, d_out_occluded, ctx->d_counters, stream_of(ctx, stream))
                                  : pt::launch_trace_shadow_t(bvh->view, d_rays, d_t_max, d_last_prim, n, d_out_occluded, nullptr, stream_of(ctx, stream));
    if (e != cudaSuccess)
        return cuda_fail(e, "trace_shadow");
    return B200PT_SUCCESS;
}

int b200pt_trace_queue(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const float* d_t_max, const int* d_last_prim, long long n, int* d_out_prim, float* d_out_t, float* d_out_uv,
                       unsigned char* d_out_occluded, b200pt_stream stream)
{
    REQUIRE(ctx != nullptr && bvh != nullptr && n >= 0 && n < (1ll << 31));
    const bool any_hit = d_t_max != nullptr;
    REQUIRE(n == 0 || (d_rays != nullptr && (any_hit ? d_out_occluded != nullptr : (d_out_prim != nullptr && d_out_t != nullptr))));
    if (n == 0)
        return B200PT_SUCCESS;
    CUDA_TRY(cudaSetDevice(ctx->device));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    cudaError_t e = ctx->counting ? pt::launch_trace_batch_queue_c(bvh->d_view, d_rays, d_t_max, d_last_prim, n, d_out_prim, d_out_t, d_out_uv, d_out_occluded, (unsigned)sms, ctx->d_counters, stream_of(ctx, stream))
                                  : pt::launch_trace_batch_queue_t(bvh->d_view, d_rays, d_t_max, d_last_prim, n, d_out_prim, d_out_t, d_out_uv, d_out_occluded, (unsigned)sms, nullptr, stream_of(ctx, stream));
    if (e != cudaSuccess)
        return cuda_fail(e, "trace_queue");
    return B200PT_SUCCESS;
}

int b200pt_set_trace_flavour(b200pt_ctx* ctx, int flavour)
{
    REQUIRE(ctx != nullptr && (flavour == -1 || flavour == 0 || flavour == 1 || flavour == 2 || (flavour >= 5 && flavour <= 7)));
    pt::g_trace_flavour = flavour;
    return B200PT_SUCCESS;
}

static int debug_function(b200pt_ctx* ctx, int which, const HIPRTRenderData* rd, const void* materials, void* volume_states, const float* vecs, unsigned int* seeds, long long n, int ggx, float* out,
                          const char* what)
{
    REQUIRE(ctx != nullptr && n >= 0 && (n == 0 || out != nullptr));
    HIPRTRenderData none;
    memset(&none, 0, sizeof(none));
    cudaError_t e = pt::launch_debug_function(which, rd ? rd : &none, materials, volume_states, vecs, seeds, n, ggx, out, ctx->main_stream);
    if (e != cudaSuccess)
        return cuda_fail(e, what);
    return B200PT_SUCCESS;
}

int b200pt_debug_bsdf_eval(b200pt_ctx* ctx, const HIPRTRenderData* render_data, const void* d_materials, void* d_volume_states, const float* d_vecs, long long n, float* d_out)
{
    REQUIRE(render_data != nullptr && (n == 0 || (d_materials != nullptr && d_volume_states != nullptr && d_vecs != nullptr)));
    return debug_function(ctx, 0, render_data, d_materials, d_volume_states, d_vecs, nullptr, n, 0, d_out, "debug_bsdf_eval");
}

int b200pt_debug_bsdf_sample(b200pt_ctx* ctx, const HIPRTRenderData* render_data, const void* d_materials, void* d_volume_states, const float* d_vecs, unsigned int* d_seeds, long long n, int ggx_sampler,
                             float* d_out)
{
    REQUIRE(render_data != nullptr && (ggx_sampler == 0 || ggx_sampler == 1) && (n == 0 || (d_materials != nullptr && d_volume_states != nullptr && d_vecs != nullptr && d_seeds != nullptr)));
    return debug_function(ctx, 1, render_data, d_materials, d_volume_states, d_vecs, d_seeds, n, ggx_sampler, d_out, "debug_bsdf_sample");
}

int b200pt_debug_envmap_sample(b200pt_ctx* ctx, const HIPRTRenderData* render_data, unsigned int* d_seeds, long long n, float* d_out)
{
    REQUIRE(render_data != nullptr && render_data->world_settings.alias_table_probas != nullptr && (n == 0 || d_seeds != nullptr));
    return debug_function(ctx, 2, render_data, nullptr, nullptr, nullptr, d_seeds, n, 0, d_out, "debug_envmap_sample");
}

int b200pt_debug_thin_film_fresnel(b200pt_ctx* ctx, const void* d_materials, const float* d_ambient_ior_and_cosine, long long n, float* d_out)
{
    REQUIRE(n == 0 || (d_materials != nullptr && d_ambient_ior_and_cosine != nullptr));
    return debug_function(ctx, 3, nullptr, d_materials, nullptr, d_ambient_ior_and_cosine, nullptr, n, 0, d_out, "debug_thin_film_fresnel");
}

int b200pt_set_restir_phases(b200pt_ctx* ctx, int enable)
{
    REQUIRE(ctx != nullptr);
    ctx->restir_phases = enable != 0;
    return B200PT_SUCCESS;
}

int b200pt_set_ray_bin_bits(b200pt_ctx* ctx, int bits)
{
    REQUIRE(ctx != nullptr && bits >= -1 && bits <= pt::kRayBinBitsMax);
    pt::g_ray_bin_bits = bits;
    return B200PT_SUCCESS;
}

int b200pt_counters_enable(b200pt_ctx* ctx, int enable)
{
    REQUIRE(ctx != nullptr);
    ctx->counting = enable != 0;
    return B200PT_SUCCESS;
}

int b200pt_counters_read(b200pt_ctx* ctx, b200pt_counters* out, int reset)
{
    REQUIRE(ctx != nullptr && out != nullptr);
    CUDA_TRY(cudaStreamSynchronize(ctx->main_stream));
    pt::TraversalCounters c;
    CUDA_TRY(cudaMemcpy(&c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
    out->rays = c.rays;
    out->nodes = c.nodes;
    out->triangles = c.triangles;
    out->stack_overflows = c.overflows;
    out->queue_rays = c.queue_rays;
    out->queue_nodes = c.queue_nodes;
    out->queue_triangles = c.queue_triangles;
    if (reset)
        CUDA_TRY(cudaMemset(ctx->d_counters, 0, sizeof(c)));
    return B200PT_SUCCESS;
}

} // extern "C"
