// PHASED wavefront, LSS_MIS_LIGHT_BSDF: the direct-lighting half of a vertex (shade_vertex_direct) cut into phases, one kernel
// each, so that every warp of an SM is inside the same few KB of code (ncu: `no_instruction` was the top stall of the
// monolithic shade kernel -- its hot path does not fit the 32 KB L1.5 instruction cache) and rays are traced by a small
// low-register kernel instead of from inside the BSDF code:
//
//   P1 nee_light_eval   pick an emissive triangle + point (3 draws), BSDF eval          -> shadow ray 0 + its contribution
//   P2 nee_bsdf_sample  BSDF sample of the light MIS pair                               -> closest-hit ray 1 + BSDF value
//   P3 env_light_eval   envmap texel by alias table / CDF (2 draws), BSDF eval          -> shadow ray 2 + its contribution
//   P4 env_bsdf_sample  BSDF sample of the envmap MIS pair, envmap_eval                 -> shadow ray 3 + its contribution
//   T  ray-queue kernel (trace_queue.cuh) over the compact id lists P1-P4 appended to: any-hit for rays 0,2,3, closest for ray 1
//   P5 bsdf             resolves ray 1 (emission + light pdf of what was hit), the three clamps, ray_color; then the BSDF
//                       sample, russian roulette and next ray of shade_vertex_bounce
//
// The phases run in the reference's program order per path and hand the Xorshift32 state on, so the random numbers each
// sampler sees are the ones it sees in the reference (src/Device/includes/Lights.h:115-220,277-321, Envmap.h:151-246,
// FullPathTracer.h:173-214); rays draw nothing (alpha testing is off on this path). The reference evaluates the BSDF of
// a light sample only when its shadow ray is unoccluded; here it is evaluated first and multiplied by the visibility:
// the same value whenever it is used.
#pragma once

#include "wavefront.cuh"

namespace pt
{
enum NeeSlot { NEE_LIGHT = 0, NEE_LIGHT_BSDF = 1, NEE_ENV = 2, NEE_ENV_BSDF = 3 };

PT_DEV size_t nee_index(const WavefrontState& wf, int which, uint32_t slot) { return (size_t)which * wf.capacity + slot; }

PT_DEV void nee_store_none(const WavefrontState& wf, int which, uint32_t slot)
{
    wf.nee_ray_o[nee_at(wf, nee_index(wf, which, slot))] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
}

// Records NEE ray `which` of the path and appends it to the bounce's any-hit (rays 0, 2, 3) or closest-hit (ray 1) id list
PT_DEV void nee_store(const WavefrontState& wf, int bounce, int which, uint32_t slot, float3 origin, float t_max, float3 direction, int last_prim, RGB pending, float pending_w)
{
    const size_t k = nee_index(wf, which, slot);
    wf.nee_ray_o[nee_at(wf, k)] = make_float4(origin.x, origin.y, origin.z, t_max);
    wf.nee_ray_d[nee_at(wf, k)] = make_float4(direction.x, direction.y, direction.z, __int_as_float(last_prim));
    wf.nee_pending[nee_at(wf, k)] = make_float4(pending.r, pending.g, pending.b, pending_w);
    if (which == NEE_LIGHT_BSDF)
        queue_push(wf.closest_queue, wf.counts + kCountsPerBounce * bounce + 3, true, (uint32_t)k);
    else
        queue_push(wf.any_queue, wf.counts + kCountsPerBounce * bounce + 2, true, (uint32_t)k);
}

// What sample_one_light does before it samples (Lights.h:277-300)
enum LightsPrelude { LIGHTS_NONE = 0, LIGHTS_EMISSION = 1, LIGHTS_SAMPLE = 2 };
PT_DEV int lights_prelude(const HIPRTRenderData& rd, const Material& material, int bounce)
{
    if (rd.buffers.emissive_triangles_count == 0)
        return LIGHTS_NONE;
    if (rd.bsdfs_data.white_furnace_mode && rd.bsdfs_data.white_furnace_mode_turn_off_emissives)
        return LIGHTS_NONE;
    if (material_is_emissive(material))
        return (material.emissive_texture_used && bounce > 0) ? LIGHTS_EMISSION : LIGHTS_NONE;
    return LIGHTS_SAMPLE;
}

// The early-outs of sample_environment_map (Envmap.h:220-246) for LSS_MIS_LIGHT_BSDF
template <int ESS>
PT_DEV bool envmap_prelude(const HIPRTRenderData& rd, const Material& material)
{
    const WorldSettings& ws = rd.world_settings;
    if (ESS == 0)
        return false;
    if (ws.ambient_light_type != AMBIENT_ENVMAP || rd.bsdfs_data.white_furnace_mode)
        return false;
    if (material_is_emissive(material))
        return false;
    return ws.envmap_intensity > 0.0f;
}

// Loads the vertex the way shade_vertex_direct sees it: ray state, hit, material, normals flipped on emissive back faces
PT_DEV void nee_load_vertex(const HIPRTRenderData& rd, const WavefrontState& wf, uint32_t slot, uint32_t pixel_index, int bounce, PathRegs& p)
{
    wf_load_ray(wf, slot, p);
    wf_load_hit(rd, wf, slot, pixel_index, bounce, p);
    if (material_is_emissive(p.payload.material) && dot(-p.direction, p.hit.geometric_normal) < 0)
    {
        p.hit.geometric_normal = -p.hit.geometric_normal;
        p.hit.shading_normal = -p.hit.shading_normal;
    }
}

PT_DEV void nee_store_rng(const WavefrontState& wf, uint32_t slot, const Rng& rng) { wf.origin_rng[wf_ray_at(wf, slot)].w = __uint_as_float(rng.state); }

// ---- P1: light sample + BSDF eval (sample_one_light_MIS, first half: Lights.h:115-170) ------------------------------------
PT_DEV void nee_light_eval_body(const HIPRTRenderData& rd, const WavefrontState& wf, int bounce, uint32_t slot, PathRegs& p)
{
    nee_store_none(wf, NEE_LIGHT, slot);
    // ray 1 doubles as the "P2 has work" flag: t_max = -2 tells P2 to skip its BSDF sample (light pdf <= 0: the
    // reference returns before sampling the BSDF, Lights.h:127-128)
    wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_LIGHT_BSDF, slot))] = make_float4(0.0f, 0.0f, 0.0f, -2.0f);
    if (lights_prelude(rd, p.payload.material, bounce) != LIGHTS_SAMPLE)
        return;

    const float3 view_direction = -p.direction;
    const HitInfo& hit = p.hit;
    bool inside_surface = dot(view_direction, hit.geometric_normal) < 0;
    float inside_surface_multiplier = inside_surface ? -1.0f : 1.0f;
    float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier;

    float light_sample_pdf;
    LightSourceInformation light_info;
    float3 random_light_point = uniform_sample_one_emissive_triangle(rd, p.rng, light_sample_pdf, light_info);
    nee_store_rng(wf, slot, p.rng);
    if (light_sample_pdf <= 0.0f)
        return;
    wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_LIGHT_BSDF, slot))].w = -1.0f;   // P2 runs

    float3 shadow_ray_direction = random_light_point - evaluated_point;
    float distance_to_light = length(shadow_ray_direction);
    float3 dir = shadow_ray_direction / distance_to_light;

    float dot_light_source = fabsf(dot(light_info.light_source_normal, -dir));
    if (dot_light_source > 0.0f)
    {
        float bsdf_pdf;
        RayVolumeState trash = p.payload.volume_state;
        RGB bsdf_color = bsdf_eval(rd, p.payload.material, trash, view_direction, hit.shading_normal, dir, bsdf_pdf);
        if (bsdf_pdf != 0.0f)
        {
            light_sample_pdf *= distance_to_light * distance_to_light;
            light_sample_pdf /= dot_light_source;
            float mis_weight = balance_heuristic(light_sample_pdf, bsdf_pdf);
            float cosine_term = maxf(dot(hit.shading_normal, dir), 0.0f);
            RGB light_source_radiance_mis = bsdf_color * cosine_term * light_info.emission * mis_weight / light_sample_pdf;
            nee_store(wf, bounce, NEE_LIGHT, slot, evaluated_point, distance_to_light, dir, hit.primitive_index, light_source_radiance_mis, 0.0f);
        }
    }
}

// ---- P2: BSDF sample of the light MIS pair (Lights.h:172-190) ----------------------------------------------------------------
template <int GGX_SAMPLER>
PT_DEV void nee_bsdf_sample_body(const HIPRTRenderData& rd, const WavefrontState& wf, int bounce, uint32_t slot, PathRegs& p)
{

    const float3 view_direction = -p.direction;
    const HitInfo& hit = p.hit;
    bool inside_surface = dot(view_direction, hit.geometric_normal) < 0;
    float inside_surface_multiplier = inside_surface ? -1.0f : 1.0f;
    float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier;

    float direction_pdf;
    float3 sampled_bsdf_direction;
    float3 bsdf_shadow_ray_origin = evaluated_point;
    RayVolumeState trash = p.payload.volume_state;
    RGB bsdf_color = bsdf_sample<GGX_SAMPLER>(rd, p.payload.material, trash, view_direction, hit.shading_normal, hit.geometric_normal, sampled_bsdf_direction, direction_pdf, p.rng);
    nee_store_rng(wf, slot, p.rng);
    bool refraction_sampled = dot(sampled_bsdf_direction, hit.shading_normal * inside_surface_multiplier) < 0;
    if (refraction_sampled)
        bsdf_shadow_ray_origin = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier * -1.0f;

    if (direction_pdf > 0)
    {
        float cosine_term = fabsf(dot(hit.shading_normal, sampled_bsdf_direction));
        nee_store(wf, bounce, NEE_LIGHT_BSDF, slot, bsdf_shadow_ray_origin, 1.0e35f, sampled_bsdf_direction, hit.primitive_index, bsdf_color * cosine_term, direction_pdf);
    }
}

// ---- P3: envmap sample + BSDF eval (sample_environment_map_with_mis, first half: Envmap.h:151-185) ---------------------------
template <int ESS>
PT_DEV void env_light_eval_body(const HIPRTRenderData& rd, const WavefrontState& wf, int bounce, uint32_t slot, PathRegs& p)
{
    nee_store_none(wf, NEE_ENV, slot);
    wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_ENV_BSDF, slot))] = make_float4(0.0f, 0.0f, 0.0f, -2.0f);
    if (!envmap_prelude<ESS>(rd, p.payload.material))
        return;
    wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_ENV_BSDF, slot))].w = -1.0f;   // P4 runs

    const float3 view_direction = -p.direction;
    const HitInfo& hit = p.hit;
    float envmap_pdf;
    float3 sampled_direction;
    RGB envmap_color = envmap_sample<ESS == 0 ? 2 : ESS>(rd.world_settings, sampled_direction, envmap_pdf, p.rng);
    nee_store_rng(wf, slot, p.rng);

    float cosine_term = dot(hit.shading_normal, sampled_direction);
    if (envmap_pdf > 0.0f && cosine_term > 0.0f)
    {
        float bsdf_pdf;
        RayVolumeState trash = p.payload.volume_state;
        RGB bsdf_color = bsdf_eval(rd, p.payload.material, trash, view_direction, hit.shading_normal, sampled_direction, bsdf_pdf);
        float mis_weight = balance_heuristic(envmap_pdf, bsdf_pdf);
        RGB envmap_mis_contribution = bsdf_color * cosine_term * mis_weight * envmap_color / envmap_pdf;
        nee_store(wf, bounce, NEE_ENV, slot, hit.inter_point, 1.0e35f, sampled_direction, hit.primitive_index, envmap_mis_contribution, 0.0f);
    }
}

// ---- P4: BSDF sample of the envmap MIS pair (Envmap.h:187-215) ----------------------------------------------------------------
template <int GGX_SAMPLER>
PT_DEV void env_bsdf_sample_body(const HIPRTRenderData& rd, const WavefrontState& wf, int bounce, uint32_t slot, PathRegs& p)
{

    const float3 view_direction = -p.direction;
    const HitInfo& hit = p.hit;
    float bsdf_sample_pdf;
    float3 bsdf_sampled_dir;
    RayVolumeState trash = p.payload.volume_state;
    RGB bsdf_color = bsdf_sample<GGX_SAMPLER>(rd, p.payload.material, trash, view_direction, hit.shading_normal, hit.geometric_normal, bsdf_sampled_dir, bsdf_sample_pdf, p.rng);
    nee_store_rng(wf, slot, p.rng);

    float cosine_term = fabsf(dot(hit.shading_normal, bsdf_sampled_dir));
    if (bsdf_sample_pdf > 0.0f)
    {
        float envmap_eval_pdf;
        RGB envmap_radiance = envmap_eval(rd, bsdf_sampled_dir, envmap_eval_pdf);
        if (envmap_eval_pdf > 0.0f)
        {
            float mis_weight = balance_heuristic(bsdf_sample_pdf, envmap_eval_pdf);
            RGB bsdf_mis_contribution = envmap_radiance * mis_weight * cosine_term * bsdf_color / bsdf_sample_pdf;
            nee_store(wf, bounce, NEE_ENV_BSDF, slot, hit.inter_point, 1.0e35f, bsdf_sampled_dir, hit.primitive_index, bsdf_mis_contribution, 0.0f);
        }
    }
}

// ---- kernel bodies: one phase per kernel, or an eval + sample pair on one loaded vertex ---------------------------------------------
// A phase that has work loads the vertex (ray, hit, 256-byte material: ~400 bytes); the paired forms load it once for both halves.
// The sample half runs only where the eval half left its flag (t_max = -1 in the pair's second ray record); it continues from the
// Xorshift32 state the eval half stored, which is the state already in p.
PT_DEV void nee_light_eval_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    nee_light_eval_body(rd, wf, bounce, slot, p);
}

template <int GGX_SAMPLER>
PT_DEV void nee_bsdf_sample_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    if (wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_LIGHT_BSDF, slot))].w != -1.0f)
        return;
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    nee_bsdf_sample_body<GGX_SAMPLER>(rd, wf, bounce, slot, p);
}

template <int ESS>
PT_DEV void env_light_eval_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    env_light_eval_body<ESS>(rd, wf, bounce, slot, p);
}

template <int GGX_SAMPLER>
PT_DEV void env_bsdf_sample_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    if (wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_ENV_BSDF, slot))].w != -1.0f)
        return;
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    env_bsdf_sample_body<GGX_SAMPLER>(rd, wf, bounce, slot, p);
}

template <int GGX_SAMPLER>
PT_DEV void nee_pair_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    nee_light_eval_body(rd, wf, bounce, slot, p);
    if (wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_LIGHT_BSDF, slot))].w == -1.0f)
        nee_bsdf_sample_body<GGX_SAMPLER>(rd, wf, bounce, slot, p);
}

template <int ESS, int GGX_SAMPLER>
PT_DEV void env_pair_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    env_light_eval_body<ESS>(rd, wf, bounce, slot, p);
    if (wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_ENV_BSDF, slot))].w == -1.0f)
        env_bsdf_sample_body<GGX_SAMPLER>(rd, wf, bounce, slot, p);
}

// ---- PHASED RIS (LSS_RIS_BSDF_AND_LIGHT, and the later bounces of LSS_RESTIR_DI with RESTIR_DI_LATER_BOUNCES_RIS) ------------------------
// sample_lights_RIS (RIS.h:292-302) traces twice in the middle of its random-number sequence: the BSDF candidate's ray (closest hit, then
// one more draw for the reservoir update) and the winner's shadow ray. Cut there:
//   R1 ris_candidates  light candidates (draws + BSDF evals), BSDF candidate sample                    -> closest-hit ray (NEE_LIGHT_BSDF record)
//   T  ray-queue kernel, closest
//   R2 ris_select      what the candidate hit -> reservoir update (1 draw), UCW, the winner's shadow ray -> any-hit ray (NEE_LIGHT record)
//   [P3 / P4 envmap phases], T any-hit, P5 resolves (ris_resolve_light)
// The pieces are the ones of the per-thread form (ris.cuh), in its order, on the same generator state handed from phase to phase: same bits.
// Only number_of_bsdf_candidates == 1 (the reference's default) runs phased; other counts keep the split shape.
// Reservoir between the phases: ris_a[slot] = weight_sum | target_function | M | emissive triangle, ris_b[slot] = point on the light | 0;
// after R2: ris_a = UCW | cosine of a BSDF sample | is-BSDF-sample flag | emissive triangle, ris_b = BSDF sample contribution.
template <int GGX_SAMPLER>
PT_DEV void ris_candidates_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    nee_store_none(wf, NEE_LIGHT, slot);
    const size_t kb = nee_index(wf, NEE_LIGHT_BSDF, slot);
    wf.nee_ray_o[nee_at(wf, kb)] = make_float4(0.0f, 0.0f, 0.0f, -2.0f);          // -2: RIS does not run at this vertex (R2 and the resolve skip it)
    if (lights_prelude(rd, p.payload.material, bounce) != LIGHTS_SAMPLE)
        return;

    const float3 view_direction = -p.direction;
    const HitInfo& hit = p.hit;
    bool inside_surface = dot(view_direction, hit.geometric_normal) < 0;
    float inside_surface_multiplier = inside_surface ? -1.0f : 1.0f;
    float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier;
    int nb_light_candidates, nb_bsdf_candidates;
    ris_candidate_counts(rd, nb_light_candidates, nb_bsdf_candidates);

    RISReservoir reservoir;
    ris_light_candidates(rd, p.payload, hit, view_direction, evaluated_point, inside_surface_multiplier, nb_light_candidates, nb_bsdf_candidates, p.rng, reservoir);
    RISBsdfCandidate c = ris_bsdf_candidate_sample<GGX_SAMPLER>(rd, p.payload, hit, view_direction, evaluated_point, inside_surface_multiplier, p.rng);
    nee_store_rng(wf, slot, p.rng);
    wf.ris_a[slot] = make_float4(reservoir.weight_sum, reservoir.sample.target_function, __uint_as_float(reservoir.M), __int_as_float(reservoir.sample.emissive_triangle_index));
    wf.ris_b[slot] = make_float4(reservoir.sample.point_on_light_source.x, reservoir.sample.point_on_light_source.y, reservoir.sample.point_on_light_source.z, 0.0f);
    wf.nee_ray_o[nee_at(wf, kb)].w = -1.0f;                                       // R2 runs; no candidate ray unless the sample has a pdf
    // the candidate's direction and colour travel in its ray record even when no ray is traced (pdf <= 0): R2 rebuilds the candidate from it
    wf.nee_ray_d[nee_at(wf, kb)] = make_float4(c.direction.x, c.direction.y, c.direction.z, __int_as_float(hit.primitive_index));
    wf.nee_pending[nee_at(wf, kb)] = make_float4(c.color.r, c.color.g, c.color.b, c.pdf);
    if (c.pdf > 0.0f)
        nee_store(wf, bounce, NEE_LIGHT_BSDF, slot, c.origin, 1.0e35f, c.direction, hit.primitive_index, c.color, c.pdf);
}

PT_DEV void ris_select_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i)
{
    const uint32_t slot = wf.sorted_queue[i];
    const size_t kb = nee_index(wf, NEE_LIGHT_BSDF, slot);
    const float4 o = wf.nee_ray_o[nee_at(wf, kb)];
    if (o.w == -2.0f)
        return;
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    const HitInfo& hit = p.hit;
    const float3 view_direction = -p.direction;
    bool inside_surface = dot(view_direction, hit.geometric_normal) < 0;
    float inside_surface_multiplier = inside_surface ? -1.0f : 1.0f;
    int nb_light_candidates, nb_bsdf_candidates;
    ris_candidate_counts(rd, nb_light_candidates, nb_bsdf_candidates);

    const float4 a = wf.ris_a[slot], b = wf.ris_b[slot];
    RISReservoir reservoir;
    reservoir.weight_sum = a.x;
    reservoir.sample.target_function = a.y;
    reservoir.M = __float_as_uint(a.z);
    reservoir.sample.emissive_triangle_index = __float_as_int(a.w);
    reservoir.sample.point_on_light_source = f3(b.x, b.y, b.z);
    reservoir.sample.is_bsdf_sample = false;

    const float4 d = wf.nee_ray_d[nee_at(wf, kb)], pc = wf.nee_pending[nee_at(wf, kb)];
    RISBsdfCandidate c;
    c.direction = f3(d.x, d.y, d.z);
    c.color = RGB(pc.x, pc.y, pc.z);
    c.pdf = pc.w;
    c.refraction = dot(c.direction, hit.shading_normal * inside_surface_multiplier) < 0;
    c.origin = c.refraction ? hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier * -1.0f
                            : hit.inter_point + hit.shading_normal * 1.0e-4f * inside_surface_multiplier;
    float candidate_weight = 0.0f;
    RISSample bsdf_RIS_sample;
    if (c.pdf > 0.0f)
    {
        const float4 h = wf.hit_tuvp[hitrec_at(wf, kb)];
        Hit light_ray_hit;
        light_ray_hit.t = h.x; light_ray_hit.u = h.y; light_ray_hit.v = h.z;
        light_ray_hit.prim = __float_as_int(h.w);
        light_ray_hit.tri_slot = (int)wf.hit_tri_slot[hitrec_word_at(wf, kb)];
        ShadowLightRayHitInfo light_hit;
        bool hit_found = shadow_light_ray_resolve(rd, 1.0e35f, light_hit, light_ray_hit);
        ris_bsdf_candidate_finish(rd, hit, c, hit_found, light_hit, nb_light_candidates, nb_bsdf_candidates, bsdf_RIS_sample, candidate_weight);
    }
    reservoir.add_one_candidate(bsdf_RIS_sample, candidate_weight, p.rng);
    reservoir.end();
    nee_store_rng(wf, slot, p.rng);

    // first half of evaluate_reservoir_sample (RIS.h:18-48): the winner's shadow ray
    wf.ris_a[slot] = make_float4(reservoir.UCW, reservoir.sample.bsdf_sample_cosine_term, reservoir.sample.is_bsdf_sample ? 1.0f : 0.0f, __int_as_float(reservoir.sample.emissive_triangle_index));
    wf.ris_b[slot] = make_float4(reservoir.sample.bsdf_sample_contribution.r, reservoir.sample.bsdf_sample_contribution.g, reservoir.sample.bsdf_sample_contribution.b, 0.0f);
    if (reservoir.UCW <= 0.0f || reservoir.sample.is_bsdf_sample)
        return;
    float3 evaluated_point = hit.inter_point + hit.shading_normal * 1.0e-4f;
    float3 shadow_ray_direction = reservoir.sample.point_on_light_source - evaluated_point;
    float distance_to_light = length(shadow_ray_direction);
    float3 dir = shadow_ray_direction / distance_to_light;
    nee_store(wf, bounce, NEE_LIGHT, slot, evaluated_point, distance_to_light, dir, hit.primitive_index, RGB(0.0f), 0.0f);
}

// second half of evaluate_reservoir_sample (RIS.h:50-80), from the traced shadow ray
PT_DEV RGB ris_resolve_light(const HIPRTRenderData& rd, const WavefrontState& wf, uint32_t slot, PathRegs& p)
{
    if (wf.nee_ray_o[nee_at(wf, nee_index(wf, NEE_LIGHT_BSDF, slot))].w == -2.0f)
        return RGB(0.0f);
    const float4 a = wf.ris_a[slot];
    const float UCW = a.x;
    RGB final_color;
    if (UCW <= 0.0f)
        return RGB(0.0f);
    const bool is_bsdf_sample = a.z != 0.0f;
    bool in_shadow = false;
    const size_t k = nee_index(wf, NEE_LIGHT, slot);
    if (!is_bsdf_sample)
        in_shadow = !(wf.nee_ray_o[nee_at(wf, k)].w > 0.0f);          // the any-hit kernel set t_max = -1 on occluded rays
    if (!in_shadow)
    {
        float bsdf_pdf;
        float cosine_at_evaluated_point;
        RGB bsdf_color;
        RayVolumeState trash = p.payload.volume_state;
        if (is_bsdf_sample)
        {
            const float4 b = wf.ris_b[slot];
            bsdf_color = RGB(b.x, b.y, b.z);
            cosine_at_evaluated_point = a.y;
        }
        else
        {
            const float4 d = wf.nee_ray_d[nee_at(wf, k)];
            const float3 dir = f3(d.x, d.y, d.z);
            bsdf_color = bsdf_eval(rd, p.payload.material, trash, -p.direction, p.hit.shading_normal, dir, bsdf_pdf);
            cosine_at_evaluated_point = maxf(0.0f, dot(p.hit.shading_normal, dir));
        }
        if (cosine_at_evaluated_point > 0.0f)
        {
            int material_index = __ldg(rd.buffers.material_indices + __float_as_int(a.w));
            RGB sample_emission = material_buffer_emission(rd, material_index);
            final_color = bsdf_color * UCW * sample_emission * cosine_at_evaluated_point;
        }
    }
    return final_color;
}

// ---- P5: finish shade_vertex_direct from the traced NEE rays (FullPathTracer.h:195-214), then shade_vertex_bounce --------------
// The any-hit kernel has set t_max = -1 on the occluded rays; ray 1's closest hit is in the hit records at [capacity + slot].
PT_DEV RGB nee_unoccluded_contribution(const WavefrontState& wf, int which, uint32_t slot)
{
    const size_t k = nee_index(wf, which, slot);
    if (!(wf.nee_ray_o[nee_at(wf, k)].w > 0.0f))
        return RGB(0.0f);
    const float4 c = wf.nee_pending[nee_at(wf, k)];
    return RGB(c.x, c.y, c.z);
}

template <int ESS, bool RIS>
PT_DEV void nee_resolve(const HIPRTRenderData& rd, const WavefrontState& wf, uint32_t slot, int bounce, PathRegs& p, bool lights_sampled, bool envmap_sampled)
{
    const Material& m = p.payload.material;
    RGB light_direct(0.0f);
    const int prelude = lights_prelude(rd, m, bounce);
    if (prelude == LIGHTS_EMISSION)
        light_direct = material_emission(m);
    else if (RIS && prelude == LIGHTS_SAMPLE && lights_sampled)
    {
        RGB direct;
        direct += ris_resolve_light(rd, wf, slot, p);
        light_direct = direct / rd.render_settings.number_of_light_samples;
    }
    else if (prelude == LIGHTS_SAMPLE && lights_sampled)
    {
        RGB light_source_radiance_mis = nee_unoccluded_contribution(wf, NEE_LIGHT, slot);
        RGB bsdf_radiance_mis;
        const size_t k = nee_index(wf, NEE_LIGHT_BSDF, slot);
        const float4 o = wf.nee_ray_o[nee_at(wf, k)];
        bool sampled = o.w != -2.0f;
        if (o.w > 0.0f)
        {
            const float4 d = wf.nee_ray_d[nee_at(wf, k)];
            const float3 sampled_bsdf_direction = f3(d.x, d.y, d.z);
            const float4 h = wf.hit_tuvp[hitrec_at(wf, k)];
            Hit light_ray_hit;
            light_ray_hit.t = h.x; light_ray_hit.u = h.y; light_ray_hit.v = h.z;
            light_ray_hit.prim = __float_as_int(h.w);
            light_ray_hit.tri_slot = (int)wf.hit_tri_slot[hitrec_word_at(wf, k)];
            ShadowLightRayHitInfo light_hit;
            bool inter_found = shadow_light_ray_resolve(rd, 1.0e35f, light_hit, light_ray_hit);
            if (inter_found && !light_hit.hit_emission.is_black())
            {
                const float4 c = wf.nee_pending[nee_at(wf, k)];
                const RGB bsdf_color_cosine(c.x, c.y, c.z);      // bsdf_color * cosine_term, the expression's first product
                const float direction_pdf = c.w;
                float light_pdf = pdf_of_emissive_triangle_hit(rd, light_hit, sampled_bsdf_direction);
                float mis_weight = balance_heuristic(direction_pdf, light_pdf);
                bsdf_radiance_mis = bsdf_color_cosine * light_hit.hit_emission * mis_weight / direction_pdf;
            }
        }
        RGB direct;
        if (sampled)
            direct += light_source_radiance_mis + bsdf_radiance_mis;
        else
            direct += RGB(0.0f);
        light_direct = direct / rd.render_settings.number_of_light_samples;
    }

    RGB envmap_direct(0.0f);
    if (envmap_sampled && envmap_prelude<ESS>(rd, m))
    {
        RGB envmap_mis_contribution = nee_unoccluded_contribution(wf, NEE_ENV, slot);
        RGB bsdf_mis_contribution = nee_unoccluded_contribution(wf, NEE_ENV_BSDF, slot);
        envmap_direct = bsdf_mis_contribution + envmap_mis_contribution;
    }

    light_direct = clamp_light_contribution(light_direct, rd.render_settings.direct_contribution_clamp, bounce == 0);
    envmap_direct = clamp_light_contribution(envmap_direct, rd.render_settings.envmap_contribution_clamp, bounce == 0);

    if (bounce == 0)
        p.payload.ray_color += material_emission(m);
    RGB indirect = (light_direct + envmap_direct) * p.payload.throughput;
    p.payload.ray_color += clamp_light_contribution(indirect, rd.render_settings.indirect_contribution_clamp, bounce > 0);
}

template <int ESS, int GGX_SAMPLER, bool RIS>
PT_DEV void wf_bsdf_resolved_path(const HIPRTRenderData& rd, int2 res, const Tile& tile, const WavefrontState& wf, int bounce, uint32_t i, bool lights_sampled, bool envmap_sampled)
{
    const uint32_t slot = wf.sorted_queue[i];
    uint32_t x, y, pixel_index;
    wf_slot_to_pixel(wf, tile, res, slot, x, y, pixel_index);
    PathRegs p;
    nee_load_vertex(rd, wf, slot, pixel_index, bounce, p);
    nee_resolve<ESS, RIS>(rd, wf, slot, bounce, p, lights_sampled, envmap_sampled);
    bool alive = shade_vertex_bounce<GGX_SAMPLER>(rd, p.payload, p.hit, p.origin, p.direction, p.rng, bounce);
    alive = alive && bounce < wf_nb_bounces(rd);
    if (alive)
        wf_store_ray(wf, slot, p);
    else
        wf_finish_path(rd, wf, slot, pixel_index, p);
    queue_push(wf.ray_queue[bounce & 1], wf.counts + kCountsPerBounce * (bounce + 1), alive, slot);
}
} // namespace pt
