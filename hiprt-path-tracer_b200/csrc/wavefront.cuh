// Wavefront form of FullPathTracer: launches per bounce over compacted queues of the paths that are still alive.
//
// Why: ncu on the megakernel (profiles/r01_fullpathtracer_megakernel_cornell1080p.txt) shows 8.5 of 32 threads active per
// instruction -- a warp waits for its longest path -- and the dominant stall is `no_instruction`: ~330 KB of SASS against a
// 32 KB L1.5 instruction cache, with every warp of an SM at a different place in it. Here every bounce is its own set of
// launches; surviving paths are appended to the next queue with one warp-aggregated atomic, so warps stay full, and the
// shade queue is grouped by a material signature so that the lanes of a warp (and the warps of an SM) walk the same
// BSDF branches. The per-path arithmetic and RNG order are the shared device functions of kernels.cuh (shade_vertex /
// miss_vertex / accumulate_sample): results are bit-identical to the megakernel and therefore to the reference
// (src/Device/kernels/FullPathTracer.h:99-327).
//
// Path state lives in HBM, indexed by the path's slot = its pixel's index inside the launch tile; queues only carry slot
// ids (4 B). Shapes, selected at run time (launch.h: WavefrontMode):
//   FUSED: bounce 0 shade kernel, then per bounce ONE kernel = trace_ray + shade_vertex (state: ray, throughput, colour, volume)
//   SPLIT: GEN (bounce 0: ray state from the G-buffer) or EXTEND (bounce > 0: trace_ray only, low register count) produce the
//          hit + material and its signature; SCAN + SCATTER group the queue by signature; SHADE consumes it.
#pragma once

#include "kernels.cuh"
#include "launch.h"

namespace pt
{
PT_DEV uint32_t lane_id()
{
    uint32_t l;
    asm("mov.u32 %0, %%laneid;" : "=r"(l));
    return l;
}

// Appends `value` to queue[*count ...] for every calling thread with pred set: one atomic per warp.
PT_DEV void queue_push(uint32_t* queue, uint32_t* count, bool pred, uint32_t value)
{
    const unsigned active = __activemask();
    const unsigned mask = __ballot_sync(active, pred);
    if (!pred)
        return;
    const uint32_t lane = lane_id();
    const int leader = __ffs((int)mask) - 1;
    uint32_t base = 0;
    if ((int)lane == leader)
        base = atomicAdd(count, (uint32_t)__popc(mask));
    base = __shfl_sync(mask, base, leader);
    queue[base + (uint32_t)__popc(mask & ((1u << lane) - 1u))] = value;
}

// Next queue index of this lane: the warp advances the launch's cursor by 32 (lane 0) and shares the base.
PT_DEV uint32_t wf_next_entry(uint32_t* cursor)
{
    __syncwarp();
    uint32_t base = 0;
    if (lane_id() == 0)
        base = atomicAdd(cursor, 32u);
    return __shfl_sync(0xFFFFFFFFu, base, 0) + lane_id();
}

// Which branches of the principled BSDF a material takes (lobe weights that are zero skip their layer:
// reference Principled.h:863-972). Paths are grouped by this value before shading.
PT_DEV uint32_t material_signature(const Material& m)
{
    uint32_t s = 0;
    s |= material_is_emissive(m) ? 1u : 0u;
    s |= m.coat > 0.0f ? 2u : 0u;
    s |= m.sheen > 0.0f ? 4u : 0u;
    s |= m.metallic > 0.0f ? 8u : 0u;
    s |= m.metallic < 1.0f ? 16u : 0u;
    s |= m.specular_transmission > 0.0f ? 32u : 0u;
    s |= m.thin_film > 0.0f ? 64u : 0u;
    return s;   // < kSortBuckets
}

// Records the slot's signature and counts it in the histogram of the current shade queue (one atomic per distinct key per warp)
PT_DEV void wf_classify(const WavefrontState& wf, uint32_t slot, const Material& m)
{
    const uint32_t key = material_signature(m);
    wf.sort_key[slot] = (uint8_t)key;
    const unsigned peers = __match_any_sync(__activemask(), key);
    if ((int)lane_id() == __ffs((int)peers) - 1)
        atomicAdd(wf.bucket_count + key, (uint32_t)__popc(peers));
}

struct PathRegs
{
    float3 origin, direction;
    Rng rng;
    HitInfo hit;
    RayPayload payload;
    PT_DEV PathRegs() : rng(0u) {}
};

// index of a path's entry in the per-path float4 records of each group (WavefrontState::stride_*)
PT_DEV size_t wf_ray_at(const WavefrontState& wf, uint32_t slot) { return (size_t)slot * wf.stride_ray; }
PT_DEV size_t wf_hit_at(const WavefrontState& wf, uint32_t slot) { return (size_t)slot * wf.stride_hit; }
PT_DEV size_t wf_radiance_at(const WavefrontState& wf, uint32_t slot) { return (size_t)slot * wf.stride_radiance; }
// ... of an id k (slot, or which * capacity + slot) in the NEE records and in the ray-queue kernel's hit records
PT_DEV size_t nee_at(const WavefrontState& wf, size_t k) { return k * wf.stride_nee; }
PT_DEV size_t hitrec_at(const WavefrontState& wf, size_t k) { return k * wf.stride_hitrec; }
PT_DEV size_t hitrec_word_at(const WavefrontState& wf, size_t k) { return k * wf.stride_hitrec_words; }

PT_DEV void wf_store_ray(const WavefrontState& wf, uint32_t slot, const PathRegs& p)
{
    wf.origin_rng[wf_ray_at(wf, slot)] = make_float4(p.origin.x, p.origin.y, p.origin.z, __uint_as_float(p.rng.state));
    wf.dir_prim[wf_ray_at(wf, slot)] = make_float4(p.direction.x, p.direction.y, p.direction.z, __int_as_float(p.hit.primitive_index));
    wf.throughput[wf_radiance_at(wf, slot)] = make_float4(p.payload.throughput.r, p.payload.throughput.g, p.payload.throughput.b, 0.0f);
    wf.color[wf_radiance_at(wf, slot)] = make_float4(p.payload.ray_color.r, p.payload.ray_color.g, p.payload.ray_color.b, 0.0f);
    copy_volume_state(reinterpret_cast<RayVolumeState*>(wf.volume + (size_t)slot * 9), &p.payload.volume_state);
}

PT_DEV void wf_load_ray(const WavefrontState& wf, uint32_t slot, PathRegs& p)
{
    const float4 o = wf.origin_rng[wf_ray_at(wf, slot)], d = wf.dir_prim[wf_ray_at(wf, slot)], t = wf.throughput[wf_radiance_at(wf, slot)], c = wf.color[wf_radiance_at(wf, slot)];
    p.origin = f3(o.x, o.y, o.z);
    p.rng.state = __float_as_uint(o.w);
    p.direction = f3(d.x, d.y, d.z);
    p.hit.primitive_index = __float_as_int(d.w);
    p.payload.throughput = RGB(t.x, t.y, t.z);
    p.payload.ray_color = RGB(c.x, c.y, c.z);
    p.payload.next_ray_state = 0;
    copy_volume_state(&p.payload.volume_state, reinterpret_cast<const RayVolumeState*>(wf.volume + (size_t)slot * 9));
}

// SPLIT mode: what the extend kernel hands to the shade kernel (the ray fields other than rng/prim/volume are unchanged)
PT_DEV void wf_store_hit(const WavefrontState& wf, uint32_t slot, const PathRegs& p)
{
    wf.origin_rng[wf_ray_at(wf, slot)].w = __uint_as_float(p.rng.state);          // trace_ray may have drawn a wavelength
    wf.hit_point[wf_hit_at(wf, slot)] = make_float4(p.hit.inter_point.x, p.hit.inter_point.y, p.hit.inter_point.z, __int_as_float(p.hit.primitive_index));
    // A material without texture overrides is a function of its index alone: the slot keeps the index (in the spare word of
    // the shading normal) and the shading kernels rebuild it from the material buffer (L1/L2 resident) where they load the
    // vertex; only textured hits write their 256-byte material to the slot. B200PT_SLOT_MATERIALS=1 (a debug switch read by
    // the launcher into wf.force_slot_materials) restores the copy for every hit.
    const int material_ref = wf.force_slot_materials ? 0 : p.hit.material_ref;
    wf.hit_ns[wf_hit_at(wf, slot)] = make_float4(p.hit.shading_normal.x, p.hit.shading_normal.y, p.hit.shading_normal.z, __int_as_float(material_ref));
    wf.hit_ng[wf_hit_at(wf, slot)] = make_float4(p.hit.geometric_normal.x, p.hit.geometric_normal.y, p.hit.geometric_normal.z, 0.0f);
    copy_volume_state(reinterpret_cast<RayVolumeState*>(wf.volume + (size_t)slot * 9), &p.payload.volume_state);
    if (material_ref == 0)
        store_material(reinterpret_cast<SimplifiedRendererMaterial*>(wf.materials) + slot, p.payload.material);
}

// The surface a path is shaded at: bounce 0 reads the camera hit straight from the G-buffer (FullPathTracer.h:136-151,
// the normals are re-normalised there), later bounces read what EXTEND stored.
PT_DEV void wf_load_hit(const HIPRTRenderData& rd, const WavefrontState& wf, uint32_t slot, uint32_t pixel_index, int bounce, PathRegs& p)
{
    if (bounce == 0 && !wf.fused_camera)      // fused: the camera hit went through SURFACE like any later hit (wf_camera_surface_path)
    {
        p.hit.inter_point = ld3_rw(rd.g_buffer.first_hits, pixel_index);
        p.hit.geometric_normal = normalize(ld3_rw(rd.g_buffer.geometric_normals, pixel_index));
        p.hit.shading_normal = normalize(ld3_rw(rd.g_buffer.shading_normals, pixel_index));
        p.hit.primitive_index = rd.g_buffer.first_hit_prim_index[pixel_index];
        load_material(p.payload.material, rd.g_buffer.materials + pixel_index);
        return;
    }
    const float4 hp = wf.hit_point[wf_hit_at(wf, slot)], ns = wf.hit_ns[wf_hit_at(wf, slot)], ng = wf.hit_ng[wf_hit_at(wf, slot)];
    p.hit.inter_point = f3(hp.x, hp.y, hp.z);
    p.hit.primitive_index = __float_as_int(hp.w);
    p.hit.shading_normal = f3(ns.x, ns.y, ns.z);
    p.hit.geometric_normal = f3(ng.x, ng.y, ng.z);
    const int material_ref = __float_as_int(ns.w);
    if (material_ref > 0)
        plain_intersection_material(rd, material_ref - 1, p.payload.material);
    else
        load_material(p.payload.material, reinterpret_cast<const SimplifiedRendererMaterial*>(wf.materials) + slot);
}

PT_DEV int wf_nb_bounces(const HIPRTRenderData& rd)
{
    int nb_bounces = rd.render_settings.nb_bounces;
    if (do_render_low_resolution(rd.render_settings))
        nb_bounces = mini(3, nb_bounces);
    return nb_bounces;
}

PT_DEV void wf_finish_path(const HIPRTRenderData& rd, const WavefrontState& wf, uint32_t slot, uint32_t pixel_index, const PathRegs& p)
{
    if (wf.fused_camera)
    {
        // sample batch: the samples of a pixel are accumulated in sample order by k_wf_accumulate once every path has ended
        wf.final_color[slot] = make_float4(p.payload.ray_color.r, p.payload.ray_color.g, p.payload.ray_color.b, p.payload.volume_state.sampled_wavelength);
        return;
    }
    float3 denoiser_normal;
    RGB denoiser_albedo;
    denoiser_aovs_from_g_buffer(rd, pixel_index, denoiser_normal, denoiser_albedo);
    accumulate_sample(rd, pixel_index, p.payload.ray_color, p.payload.volume_state.sampled_wavelength, denoiser_normal, denoiser_albedo);
}

PT_DEV void wf_slot_to_pixel(const WavefrontState& wf, const Tile& tile, int2 res, uint32_t slot, uint32_t& x, uint32_t& y, uint32_t& pixel_index)
{
    const uint32_t tw = (uint32_t)(tile.x1 - tile.x0);
    const uint32_t local = wf.samples > 1u ? slot % wf.tile_pixels : slot;     // slot = sample * tile_pixels + pixel inside the tile
    x = (uint32_t)tile.x0 + local % tw;
    y = (uint32_t)tile.y0 + local / tw;
    pixel_index = x + y * (uint32_t)res.x;
}

// Ray state of a fresh path at its camera hit (FullPathTracer.h:124-151)
PT_DEV void wf_init_path(const HIPRTRenderData& rd, uint32_t pixel_index, PathRegs& p)
{
    p.rng = Rng(pixel_seed(rd, pixel_index));
    p.hit.primitive_index = rd.g_buffer.first_hit_prim_index[pixel_index];
    p.origin = f3(0.0f, 0.0f, 0.0f);
    p.direction = normalize(-ld3_rw(rd.g_buffer.view_directions, pixel_index));
    p.payload.throughput = RGB(1.0f);
    p.payload.ray_color = RGB(0.0f);
    p.payload.next_ray_state = 0;
    copy_volume_state(&p.payload.volume_state, rd.g_buffer.ray_volume_states + pixel_index);
}

// ---- FUSED, bounce 0: the camera hit comes from the G-buffer ------------------------------------------------------
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void wf_bounce0_pixel(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, uint32_t x, uint32_t y)
{
    const uint32_t pixel_index = x + y * res.x;
    const uint32_t slot = (x - tile.x0) + (y - tile.y0) * (uint32_t)(tile.x1 - tile.x0);
    bool alive = false;
    if (rd.aux_buffers.pixel_active[pixel_index])
    {
        PathRegs p;
        wf_init_path(rd, pixel_index, p);
        wf_load_hit(rd, wf, slot, pixel_index, 0, p);
        if (rd.g_buffer.camera_ray_hit[pixel_index] == 1)
            alive = shade_vertex<LSS, ESS, LATER, GGX_SAMPLER>(rd, p.payload, p.hit, p.origin, p.direction, p.rng, make_int2(x, y), res, 0);
        else
            miss_vertex<ESS>(rd, p.payload, p.direction, 0);
        alive = alive && wf_nb_bounces(rd) > 0;
        if (alive)
            wf_store_ray(wf, slot, p);
        else
            wf_finish_path(rd, wf, slot, pixel_index, p);
    }
    queue_push(wf.ray_queue[0], wf.counts + kCountsPerBounce * 1, alive, slot);
}

// ---- FUSED: bounce >= 1, trace + shade in one kernel ------------------------------------------------------------
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void wf_bounce_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.ray_queue[(bounce - 1) & 1][i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    wf_load_ray(wf, slot, p);
    bool alive = false;
    if (trace_ray(rd, p.origin, p.direction, p.payload, p.hit, p.hit.primitive_index, p.rng))
        alive = shade_vertex<LSS, ESS, LATER, GGX_SAMPLER>(rd, p.payload, p.hit, p.origin, p.direction, p.rng, make_int2(x, y), res, bounce);
    else
        miss_vertex<ESS>(rd, p.payload, p.direction, bounce);
    alive = alive && bounce < wf_nb_bounces(rd);
    if (alive)
        wf_store_ray(wf, slot, p);
    else
        wf_finish_path(rd, wf, slot, pixel_index, p);
    queue_push(wf.ray_queue[bounce & 1], wf.counts + kCountsPerBounce * (bounce + 1), alive, slot);
}

// ---- SPLIT, bounce 0: GEN writes the ray state of every active pixel and queues the camera hits for shading -------------
template <int ESS>
PT_DEV void wf_gen_pixel(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, uint32_t x, uint32_t y)
{
    const uint32_t pixel_index = x + y * res.x;
    const uint32_t slot = (x - tile.x0) + (y - tile.y0) * (uint32_t)(tile.x1 - tile.x0);
    bool found = false;
    if (rd.aux_buffers.pixel_active[pixel_index])
    {
        PathRegs p;
        wf_init_path(rd, pixel_index, p);
        found = rd.g_buffer.camera_ray_hit[pixel_index] == 1;
        if (found)
        {
            wf_store_ray(wf, slot, p);
            load_material(p.payload.material, rd.g_buffer.materials + pixel_index);
            wf_classify(wf, slot, p.payload.material);
        }
        else
        {
            miss_vertex<ESS>(rd, p.payload, p.direction, 0);
            wf_finish_path(rd, wf, slot, pixel_index, p);
        }
    }
    queue_push(wf.shade_queue, wf.counts + 1, found, slot);
}

// ---- sample batches (b200pt_launch_PathTraceSamples): CameraRays fused into the wavefront, several sample passes in flight ---------
// What the reference runs as  for s in passes: CameraRays(seed_c[s]); FullPathTracer(seed_p[s])  (GPURenderer::render,
// src/Renderer/GPURenderer.cpp:408-464) runs here as ONE wavefront over slots = (pass, pixel): CAMERA generates every pass's primary
// rays (CameraRays.h:127-150), the ray-queue kernel traces them, CAMERA_SURFACE finishes CameraRays (hit -> surface, emissive flip,
// G-buffer of the last pass: CameraRays.h:152-170) and starts FullPathTracer on the same registers (FullPathTracer.h:124-151: fresh
// generator from the path seed, normals and direction re-normalised as when they are read back from the G-buffer), so the 345 B/px
// G-buffer round trip is not paid per pass; paths end in final_color[slot] and ACCUMULATE adds a pixel's passes in pass order
// (FullPathTracer.h:296-326). Same arithmetic, same generators, same order of additions: same bits as the separate launches.
// Only for options under which a pass does not depend on the previous one's result: no adaptive sampling / pixel noise threshold
// (CameraRays reads the accumulated statistics), no ReSTIR DI (temporal reuse), no low-resolution rendering.
PT_DEV void wf_camera_slot(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, const SampleBatch& batch, uint32_t slot)
{
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    const uint32_t pass = slot / wf.tile_pixels;

    Rng rng(pixel_seed_of_pass(rd, pixel_index, batch.sample_number[pass], batch.camera_seed[pass]));
    float x_dir = (x + 0.5f);
    float y_dir = (y + 0.5f);
    if (rd.current_camera.do_jittering)
    {
        x_dir += rng() - 0.5f;
        y_dir += rng() - 0.5f;
    }
    PathRegs p;
    get_camera_ray(rd.current_camera, x_dir, y_dir, res, p.origin, p.direction);
    p.rng = rng;
    p.hit.primitive_index = -1;             // the camera ray leaves from no primitive
    p.payload.throughput = RGB(1.0f);
    p.payload.ray_color = RGB(0.0f);
    p.payload.next_ray_state = 0;
    volume_state_init(p.payload.volume_state);
    wf_store_ray(wf, slot, p);
    wf.ray_queue[1][slot] = slot;           // every slot has a primary ray: the queue is the identity, its length is written by the launcher
    if (pass + 1 == batch.count)
    {
        rd.aux_buffers.pixel_active[pixel_index] = true;
        if (rd.render_settings.do_update_status_buffers)
            rd.aux_buffers.still_one_ray_active[0] = 1;
    }
}

template <int ESS>
PT_DEV void wf_camera_surface_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, const SampleBatch& batch, uint32_t i)
{
    const uint32_t slot = wf.ray_queue[1][i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    const uint32_t pass = slot / wf.tile_pixels;
    PathRegs p;
    wf_load_ray(wf, slot, p);
    const float4 h = wf.hit_tuvp[hitrec_at(wf, slot)];
    Hit first;
    first.t = h.x; first.u = h.y; first.v = h.z;
    first.prim = __float_as_int(h.w);
    first.tri_slot = (int)wf.hit_tri_slot[hitrec_word_at(wf, slot)];
    // ---- the rest of CameraRays (CameraRays.h:150-170)
    const bool found = trace_ray_from(rd, p.origin, p.direction, p.payload, p.hit, -1, p.rng, first, true);
    if (found && material_is_emissive(p.payload.material) && dot(-p.direction, p.hit.geometric_normal) < 0)
    {
        p.hit.geometric_normal = -p.hit.geometric_normal;
        p.hit.shading_normal = -p.hit.shading_normal;
    }
    if (batch.write_g_buffer && pass + 1 == batch.count)
    {
        if (found)
        {
            st3(rd.g_buffer.geometric_normals, pixel_index, p.hit.geometric_normal);
            st3(rd.g_buffer.shading_normals, pixel_index, p.hit.shading_normal);
            store_material(rd.g_buffer.materials + pixel_index, p.payload.material);
            st3(rd.g_buffer.first_hits, pixel_index, p.hit.inter_point);
            copy_volume_state(rd.g_buffer.ray_volume_states + pixel_index, &p.payload.volume_state);
        }
        rd.g_buffer.first_hit_prim_index[pixel_index] = found ? p.hit.primitive_index : -1;
        st3(rd.g_buffer.view_directions, pixel_index, -p.direction);
        rd.g_buffer.camera_ray_hit[pixel_index] = found;
    }
    // ---- the head of FullPathTracer (FullPathTracer.h:124-151): its own generator; what it reads back from the G-buffer it normalises
    p.rng = Rng(pixel_seed_of_pass(rd, pixel_index, batch.sample_number[pass], batch.path_seed[pass]));
    p.origin = f3(0.0f, 0.0f, 0.0f);
    p.direction = normalize(-(-p.direction));
    if (found)
    {
        p.hit.geometric_normal = normalize(p.hit.geometric_normal);
        p.hit.shading_normal = normalize(p.hit.shading_normal);
        wf.aov_normal[slot] = make_float4(p.hit.shading_normal.x, p.hit.shading_normal.y, p.hit.shading_normal.z, 1.0f);
        wf.aov_albedo[slot] = make_float4(p.payload.material.base_color.r, p.payload.material.base_color.g, p.payload.material.base_color.b, 0.0f);
        wf_store_ray(wf, slot, p);
        wf_store_hit(wf, slot, p);
        wf_classify(wf, slot, p.payload.material);
    }
    else
    {
        wf.aov_normal[slot] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        wf.aov_albedo[slot] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        miss_vertex<ESS>(rd, p.payload, p.direction, 0);
        wf_finish_path(rd, wf, slot, pixel_index, p);
    }
    queue_push(wf.shade_queue, wf.counts + 1, found, slot);
}

// The end of FullPathTracer for every pass of the batch, in pass order (FullPathTracer.h:288-326); one thread per pixel of the tile
PT_DEV void wf_accumulate_pixel(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, const SampleBatch& batch, uint32_t local)
{
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, local, x, y, pixel_index);
    for (uint32_t pass = 0; pass < batch.count; pass++)
    {
        const uint32_t slot = pass * wf.tile_pixels + local;
        const float4 c = wf.final_color[slot], n = wf.aov_normal[slot], a = wf.aov_albedo[slot];
        // aov_normal holds normalize(G-buffer shading normal), the value denoiser_aovs_from_g_buffer adds to zero
        const float3 denoiser_normal = n.w != 0.0f ? f3(0.0f, 0.0f, 0.0f) + f3(n.x, n.y, n.z) : f3(0.0f, 0.0f, 0.0f);
        const RGB denoiser_albedo = n.w != 0.0f ? RGB(0.0f) + RGB(a.x, a.y, a.z) : RGB(0.0f);
        accumulate_sample_at(rd, pixel_index, RGB(c.x, c.y, c.z), c.w, denoiser_normal, denoiser_albedo, batch.sample_number[pass], batch.aov_counter[pass]);
    }
}

// ---- SPLIT, bounce >= 1: EXTEND = closest hit + material ----------------------------------------------------------------
template <int ESS>
PT_DEV void wf_extend_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.ray_queue[(bounce - 1) & 1][i];
    PathRegs p;
    wf_load_ray(wf, slot, p);
    const bool found = trace_ray(rd, p.origin, p.direction, p.payload, p.hit, p.hit.primitive_index, p.rng);
    if (found)
    {
        wf_store_hit(wf, slot, p);
        wf_classify(wf, slot, p.payload.material);
    }
    else
    {
        uint32_t x, y, pixel_index;
        wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
        miss_vertex<ESS>(rd, p.payload, p.direction, bounce);
        wf_finish_path(rd, wf, slot, pixel_index, p);
    }
    queue_push(wf.shade_queue, wf.counts + kCountsPerBounce * bounce + 1, found, slot);
}

// ---- SPLIT: group the shade queue by material signature (counting sort: histogram came from wf_classify) ----------------
// One block of kSortBuckets threads: exclusive scan of the histogram in place -> scatter cursors.
PT_DEV void wf_scan_buckets(const WavefrontState& wf, uint32_t* smem)
{
    const uint32_t t = threadIdx.x;
    const uint32_t v = wf.bucket_count[t];
    smem[t] = v;
    __syncthreads();
    for (uint32_t offset = 1; offset < (uint32_t)kSortBuckets; offset <<= 1)
    {
        const uint32_t add = t >= offset ? smem[t - offset] : 0u;
        __syncthreads();
        smem[t] += add;
        __syncthreads();
    }
    wf.bucket_count[t] = smem[t] - v;
}

PT_DEV void wf_scatter_path(const WavefrontState& wf, uint32_t i)
{
    const uint32_t slot = wf.shade_queue[i];
    const uint32_t key = wf.sort_key[slot];
    const unsigned peers = __match_any_sync(__activemask(), key);
    const uint32_t lane = lane_id();
    const int leader = __ffs((int)peers) - 1;
    uint32_t base = 0;
    if ((int)lane == leader)
        base = atomicAdd(wf.bucket_count + key, (uint32_t)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    wf.sorted_queue[base + (uint32_t)__popc(peers & ((1u << lane) - 1u))] = slot;
}

// ---- SPLIT: SHADE one vertex of every queued path -------------------------------------------------------------------------
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void wf_shade_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    wf_load_ray(wf, slot, p);
    wf_load_hit(rd, wf, slot, pixel_index, bounce, p);
    bool alive = shade_vertex<LSS, ESS, LATER, GGX_SAMPLER>(rd, p.payload, p.hit, p.origin, p.direction, p.rng, make_int2(x, y), res, bounce);
    alive = alive && bounce < wf_nb_bounces(rd);
    if (alive)
        wf_store_ray(wf, slot, p);
    else
        wf_finish_path(rd, wf, slot, pixel_index, p);
    queue_push(wf.ray_queue[bounce & 1], wf.counts + kCountsPerBounce * (bounce + 1), alive, slot);
}

// ---- PHASED: SHADE split into its two halves, so that all warps of an SM run the same part of the BSDF code ---------------
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void wf_direct_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    wf_load_ray(wf, slot, p);
    wf_load_hit(rd, wf, slot, pixel_index, bounce, p);
    shade_vertex_direct<LSS, ESS, LATER, GGX_SAMPLER>(rd, p.payload, p.hit, p.direction, p.rng, make_int2(x, y), res, bounce);
    wf.origin_rng[wf_ray_at(wf, slot)].w = __uint_as_float(p.rng.state);
    wf.color[wf_radiance_at(wf, slot)] = make_float4(p.payload.ray_color.r, p.payload.ray_color.g, p.payload.ray_color.b, 0.0f);
}

template <int GGX_SAMPLER>
PT_DEV void wf_bsdf_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    wf_load_ray(wf, slot, p);
    wf_load_hit(rd, wf, slot, pixel_index, bounce, p);
    // the flip shade_vertex_direct applied to its copy of the normals (FullPathTracer.h:173-177)
    if (material_is_emissive(p.payload.material) && dot(-p.direction, p.hit.geometric_normal) < 0)
    {
        p.hit.geometric_normal = -p.hit.geometric_normal;
        p.hit.shading_normal = -p.hit.shading_normal;
    }
    bool alive = shade_vertex_bounce<GGX_SAMPLER>(rd, p.payload, p.hit, p.origin, p.direction, p.rng, bounce);
    alive = alive && bounce < wf_nb_bounces(rd);
    if (alive)
        wf_store_ray(wf, slot, p);
    else
        wf_finish_path(rd, wf, slot, pixel_index, p);
    queue_push(wf.ray_queue[bounce & 1], wf.counts + kCountsPerBounce * (bounce + 1), alive, slot);
}

// ---- PHASED, bounce >= 1: SURFACE = what trace_ray does with the closest hit the ray-queue kernel found --------------------
template <int ESS>
PT_DEV void wf_surface_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.ray_queue[(bounce - 1) & 1][i];
    PathRegs p;
    wf_load_ray(wf, slot, p);
    const float4 h = wf.hit_tuvp[hitrec_at(wf, slot)];
    Hit first;
    first.t = h.x; first.u = h.y; first.v = h.z;
    first.prim = __float_as_int(h.w);
    first.tri_slot = (int)wf.hit_tri_slot[hitrec_word_at(wf, slot)];
    const bool found = trace_ray_from(rd, p.origin, p.direction, p.payload, p.hit, p.hit.primitive_index, p.rng, first, true);
    if (found)
    {
        wf_store_hit(wf, slot, p);
        wf_classify(wf, slot, p.payload.material);
    }
    else
    {
        uint32_t x, y, pixel_index;
        wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
        miss_vertex<ESS>(rd, p.payload, p.direction, bounce);
        wf_finish_path(rd, wf, slot, pixel_index, p);
    }
    queue_push(wf.shade_queue, wf.counts + kCountsPerBounce * bounce + 1, found, slot);
}
} // namespace pt
