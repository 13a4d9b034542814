// Wavefront form of FullPathTracer: one launch per bounce over a compacted queue of the paths that are still alive.
//
// Why: ncu on the megakernel (profiles/r01_fullpathtracer_megakernel_cornell1080p.txt) shows 8.5 of 32 threads active per
// instruction -- a warp waits for its longest path. Here every bounce is its own launch; surviving paths are appended to
// the next bounce's queue with one warp-aggregated atomic, so warps stay full. The per-path arithmetic and RNG order are
// the shared device functions of kernels.cuh (shade_vertex / miss_vertex / accumulate_sample): results are bit-identical
// to the megakernel and therefore to the reference (src/Device/kernels/FullPathTracer.h:99-327).
//
// Path state lives in HBM, indexed by the path's slot = its pixel's index inside the launch tile; queues only carry slot
// ids (4 B). Two shapes, selected at run time (launch.h: WavefrontMode):
//   FUSED: bounce 0 shade kernel, then per bounce ONE kernel = trace_ray + shade_vertex (state: ray, throughput, colour, volume)
//   SPLIT: per bounce an EXTEND kernel (trace_ray only: low register count, traversal-bound) that writes hit + material,
//          and a SHADE kernel that consumes them.
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

struct PathRegs
{
    float3 origin, direction;
    Rng rng;
    HitInfo hit;
    RayPayload payload;
    PT_DEV PathRegs() : rng(0u) {}
};

PT_DEV void wf_store_ray(const WavefrontState& wf, uint32_t slot, const PathRegs& p)
{
    wf.origin_rng[slot] = make_float4(p.origin.x, p.origin.y, p.origin.z, __uint_as_float(p.rng.state));
    wf.dir_prim[slot] = make_float4(p.direction.x, p.direction.y, p.direction.z, __int_as_float(p.hit.primitive_index));
    wf.throughput[slot] = make_float4(p.payload.throughput.r, p.payload.throughput.g, p.payload.throughput.b, 0.0f);
    wf.color[slot] = make_float4(p.payload.ray_color.r, p.payload.ray_color.g, p.payload.ray_color.b, 0.0f);
    copy_volume_state(reinterpret_cast<RayVolumeState*>(wf.volume + (size_t)slot * 9), &p.payload.volume_state);
}

PT_DEV void wf_load_ray(const WavefrontState& wf, uint32_t slot, PathRegs& p)
{
    const float4 o = wf.origin_rng[slot], d = wf.dir_prim[slot], t = wf.throughput[slot], c = wf.color[slot];
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
    wf.origin_rng[slot].w = __uint_as_float(p.rng.state);          // trace_ray may have drawn a wavelength
    wf.hit_point[slot] = make_float4(p.hit.inter_point.x, p.hit.inter_point.y, p.hit.inter_point.z, __int_as_float(p.hit.primitive_index));
    wf.hit_ns[slot] = make_float4(p.hit.shading_normal.x, p.hit.shading_normal.y, p.hit.shading_normal.z, 0.0f);
    wf.hit_ng[slot] = make_float4(p.hit.geometric_normal.x, p.hit.geometric_normal.y, p.hit.geometric_normal.z, 0.0f);
    copy_volume_state(reinterpret_cast<RayVolumeState*>(wf.volume + (size_t)slot * 9), &p.payload.volume_state);
    store_material(reinterpret_cast<SimplifiedRendererMaterial*>(wf.materials) + slot, p.payload.material);
}

PT_DEV void wf_load_hit(const WavefrontState& wf, uint32_t slot, PathRegs& p)
{
    const float4 hp = wf.hit_point[slot], ns = wf.hit_ns[slot], ng = wf.hit_ng[slot];
    p.hit.inter_point = f3(hp.x, hp.y, hp.z);
    p.hit.primitive_index = __float_as_int(hp.w);
    p.hit.shading_normal = f3(ns.x, ns.y, ns.z);
    p.hit.geometric_normal = f3(ng.x, ng.y, ng.z);
    load_material(p.payload.material, reinterpret_cast<const SimplifiedRendererMaterial*>(wf.materials) + slot);
}

PT_DEV int wf_nb_bounces(const HIPRTRenderData& rd)
{
    int nb_bounces = rd.render_settings.nb_bounces;
    if (do_render_low_resolution(rd.render_settings))
        nb_bounces = mini(3, nb_bounces);
    return nb_bounces;
}

PT_DEV void wf_finish_path(const HIPRTRenderData& rd, uint32_t pixel_index, const PathRegs& p)
{
    float3 denoiser_normal;
    RGB denoiser_albedo;
    denoiser_aovs_from_g_buffer(rd, pixel_index, denoiser_normal, denoiser_albedo);
    accumulate_sample(rd, pixel_index, p.payload.ray_color, p.payload.volume_state.sampled_wavelength, denoiser_normal, denoiser_albedo);
}

PT_DEV void wf_slot_to_pixel(const Tile& tile, int2 res, uint32_t slot, uint32_t& x, uint32_t& y, uint32_t& pixel_index)
{
    const uint32_t tw = (uint32_t)(tile.x1 - tile.x0);
    x = (uint32_t)tile.x0 + slot % tw;
    y = (uint32_t)tile.y0 + slot / tw;
    pixel_index = x + y * (uint32_t)res.x;
}

// ---- bounce 0: the camera hit comes from the G-buffer (FullPathTracer.h:136-151) ------------------------------
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void wf_bounce0_pixel(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, uint32_t x, uint32_t y)
{
    const uint32_t pixel_index = x + y * res.x;
    const uint32_t slot = (x - tile.x0) + (y - tile.y0) * (uint32_t)(tile.x1 - tile.x0);
    bool alive = false;
    PathRegs p;
    const bool active = rd.aux_buffers.pixel_active[pixel_index];
    if (active)
    {
        p.rng = Rng(pixel_seed(rd, pixel_index));
        p.hit.inter_point = ld3_rw(rd.g_buffer.first_hits, pixel_index);
        p.hit.geometric_normal = normalize(ld3_rw(rd.g_buffer.geometric_normals, pixel_index));
        p.hit.shading_normal = normalize(ld3_rw(rd.g_buffer.shading_normals, pixel_index));
        p.hit.primitive_index = rd.g_buffer.first_hit_prim_index[pixel_index];
        p.origin = f3(0.0f, 0.0f, 0.0f);
        p.direction = normalize(-ld3_rw(rd.g_buffer.view_directions, pixel_index));
        p.payload.throughput = RGB(1.0f);
        p.payload.ray_color = RGB(0.0f);
        p.payload.next_ray_state = 0;
        load_material(p.payload.material, rd.g_buffer.materials + pixel_index);
        copy_volume_state(&p.payload.volume_state, rd.g_buffer.ray_volume_states + pixel_index);

        if (rd.g_buffer.camera_ray_hit[pixel_index] == 1)
            alive = shade_vertex<LSS, ESS, LATER, GGX_SAMPLER>(rd, p.payload, p.hit, p.origin, p.direction, p.rng, make_int2(x, y), res, 0);
        else
            miss_vertex<ESS>(rd, p.payload, p.direction, 0);
        alive = alive && wf_nb_bounces(rd) > 0;
        if (alive)
            wf_store_ray(wf, slot, p);
        else
            wf_finish_path(rd, pixel_index, p);
    }
    queue_push(wf.ray_queue[0], wf.counts + 2 * 1, alive, slot);
}

// ---- FUSED: bounce >= 1, trace + shade in one kernel ------------------------------------------------------------
template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void wf_bounce_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.ray_queue[(bounce - 1) & 1][i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(tile, res, slot, x, y, pixel_index);
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
        wf_finish_path(rd, pixel_index, p);
    queue_push(wf.ray_queue[bounce & 1], wf.counts + 2 * (bounce + 1), alive, slot);
}

// ---- SPLIT: extend (closest hit + material) and shade as two kernels --------------------------------------------
template <int ESS>
PT_DEV void wf_extend_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.ray_queue[(bounce - 1) & 1][i];
    PathRegs p;
    wf_load_ray(wf, slot, p);
    const bool found = trace_ray(rd, p.origin, p.direction, p.payload, p.hit, p.hit.primitive_index, p.rng);
    if (found)
        wf_store_hit(wf, slot, p);
    else
    {
        uint32_t x, y, pixel_index;
        wf_slot_to_pixel(tile, res, slot, x, y, pixel_index);
        miss_vertex<ESS>(rd, p.payload, p.direction, bounce);
        wf_finish_path(rd, pixel_index, p);
    }
    queue_push(wf.shade_queue, wf.counts + 2 * bounce + 1, found, slot);
}

template <int LSS, int ESS, int LATER, int GGX_SAMPLER>
PT_DEV void wf_shade_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.shade_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(tile, res, slot, x, y, pixel_index);
    PathRegs p;
    wf_load_ray(wf, slot, p);
    wf_load_hit(wf, slot, p);
    bool alive = shade_vertex<LSS, ESS, LATER, GGX_SAMPLER>(rd, p.payload, p.hit, p.origin, p.direction, p.rng, make_int2(x, y), res, bounce);
    alive = alive && bounce < wf_nb_bounces(rd);
    if (alive)
        wf_store_ray(wf, slot, p);
    else
        wf_finish_path(rd, pixel_index, p);
    queue_push(wf.ray_queue[bounce & 1], wf.counts + 2 * (bounce + 1), alive, slot);
}
} // namespace pt
