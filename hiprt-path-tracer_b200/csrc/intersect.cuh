// Scene queries used by the kernels: closest hit with shading data, shadow rays, shadow-light rays.
// reference: src/Device/includes/Intersect.h:30-410, Texture.h:208-222 (uv_interpolate).
#pragma once

#include "bsdf_principled.cuh"
#include "traverse.cuh"

namespace pt
{
struct HitInfo
{
    float3 inter_point;
    float3 shading_normal;
    float3 geometric_normal;
    float2 texcoords;
    float2 uv;
    float t;
    int primitive_index;
    int material_ref;       // trace_ray: material index + 1 when the material has no texture override (plain_intersection_material), else 0
};

struct RayPayload
{
    RGB throughput;
    RGB ray_color;
    int next_ray_state;   // 0 BOUNCE, 1 MISSED
    alignas(16) Material material;   // moved with 16-byte vector loads/stores to and from the G-buffer
    RayVolumeState volume_state;
};

PT_DEV const Bvh8Device& scene_bvh(const HIPRTRenderData& rd) { return *reinterpret_cast<const Bvh8Device*>(rd.geom); }

// Counting build: the launcher stores the device address of the context's counter block in the (otherwise unused)
// HIPRTRenderData::hiprt_function_table field of its private copy of the struct.
#ifdef PT_COUNTERS
#define PT_COUNTERS_PTR reinterpret_cast<TraversalCounters*>(rd.hiprt_function_table)
#else
#define PT_COUNTERS_PTR nullptr
#endif

PT_DEV int3 triangle_vertex_indices(const HIPRTRenderData& rd, int primitive_index)
{
    const int* idx = rd.buffers.triangles_indices + (size_t)primitive_index * 3;
    return make_int3(__ldg(idx), __ldg(idx + 1), __ldg(idx + 2));
}

PT_DEV float2 ld2(const void* base, size_t index)
{
    const float* p = reinterpret_cast<const float*>(base) + index * 2;
    return make_float2(__ldg(p), __ldg(p + 1));
}

// What a hit needs of its triangle besides the geometry: the three vertices' normals and texcoords, whether vertex A has a normal, the
// material index (and the vertex indices, for normal mapping only). Two sources with the same values: the reference's buffers through the
// index buffer (RenderBuffers, src/HostDeviceCommon/RenderBuffers.h), or -- when b200pt_bvh_attach_shading_data has run -- the 64-byte record
// of the triangle's BVH slot plus the two spare words of its 48-byte triangle record.
struct TriangleShading
{
    float3 na, nb, nc;
    float2 ta, tb, tc;
    int material_index;
    int has_normals;          // has_vertex_normals[vertex A], the only one get_shading_normal looks at (Intersect.h:69)
    int prim;
    bool from_record;
};

PT_DEV TriangleShading load_triangle_shading(const HIPRTRenderData& rd, int prim, int tri_slot)
{
    TriangleShading ts;
    ts.prim = prim;
    const Bvh8Device& bvh = scene_bvh(rd);
    if (bvh.shading != nullptr && tri_slot >= 0)
    {
        const float4* sp = reinterpret_cast<const float4*>(bvh.shading) + (size_t)tri_slot * 4;
        const float4 s0 = __ldg(sp + 0), s1 = __ldg(sp + 1), s2 = __ldg(sp + 2), s3 = __ldg(sp + 3);
        const int* tw = reinterpret_cast<const int*>(bvh.triangles + tri_slot);
        ts.na = f3(s0.x, s0.y, s0.z); ts.nb = f3(s0.w, s1.x, s1.y); ts.nc = f3(s1.z, s1.w, s2.x);
        ts.ta = make_float2(s2.y, s2.z); ts.tb = make_float2(s2.w, s3.x); ts.tc = make_float2(s3.y, s3.z);
        ts.material_index = __ldg(tw + 7);
        ts.has_normals = __ldg(tw + 11);
        ts.from_record = true;
        return ts;
    }
    const int3 vi = triangle_vertex_indices(rd, prim);
    ts.na = ld3(rd.buffers.vertex_normals, vi.x); ts.nb = ld3(rd.buffers.vertex_normals, vi.y); ts.nc = ld3(rd.buffers.vertex_normals, vi.z);
    ts.ta = ld2(rd.buffers.texcoords, vi.x); ts.tb = ld2(rd.buffers.texcoords, vi.y); ts.tc = ld2(rd.buffers.texcoords, vi.z);
    ts.material_index = __ldg(rd.buffers.material_indices + prim);
    ts.has_normals = __ldg(rd.buffers.has_vertex_normals + vi.x);
    ts.from_record = false;
    return ts;
}

// uv_interpolate: data[B] * u + data[C] * v + data[A] * (1 - u - v)
PT_DEV float2 interpolate_texcoords(const HIPRTRenderData& rd, int3 vi, float2 uv)
{
    float2 a = ld2(rd.buffers.texcoords, vi.x), b = ld2(rd.buffers.texcoords, vi.y), c = ld2(rd.buffers.texcoords, vi.z);
    float w = 1.0f - uv.x - uv.y;
    return make_float2(b.x * uv.x + c.x * uv.y + a.x * w, b.y * uv.x + c.y * uv.y + a.y * w);
}

PT_DEV float2 interpolate_texcoords(const TriangleShading& ts, float2 uv)
{
    float w = 1.0f - uv.x - uv.y;
    return make_float2(ts.tb.x * uv.x + ts.tc.x * uv.y + ts.ta.x * w, ts.tb.y * uv.x + ts.tc.y * uv.y + ts.ta.y * w);
}

PT_DEV float3 interpolate_normals(const HIPRTRenderData& rd, int3 vi, float2 uv)
{
    float3 a = ld3(rd.buffers.vertex_normals, vi.x), b = ld3(rd.buffers.vertex_normals, vi.y), c = ld3(rd.buffers.vertex_normals, vi.z);
    return b * uv.x + c * uv.y + a * (1.0f - uv.x - uv.y);
}

PT_DEV float3 interpolate_normals(const TriangleShading& ts, float2 uv)
{
    return ts.nb * uv.x + ts.nc * uv.y + ts.na * (1.0f - uv.x - uv.y);
}

// normal_mapping (Intersect.h:30-62)
PT_DEV float3 normal_mapping(const HIPRTRenderData& rd, int normal_map_texture_index, int3 vi, float2 texcoords, float3 surface_normal)
{
    float2 P0_uv = ld2(rd.buffers.texcoords, vi.x), P1_uv = ld2(rd.buffers.texcoords, vi.y), P2_uv = ld2(rd.buffers.texcoords, vi.z);
    float2 d10 = make_float2(P1_uv.x - P0_uv.x, P1_uv.y - P0_uv.y);
    float2 d20 = make_float2(P2_uv.x - P0_uv.x, P2_uv.y - P0_uv.y);

    float3 P0 = ld3(rd.buffers.vertices_positions, vi.x), P1 = ld3(rd.buffers.vertices_positions, vi.y), P2 = ld3(rd.buffers.vertices_positions, vi.z);
    float3 e01 = P1 - P0;
    float3 e02 = P2 - P0;

    float det_inverse = 1.0f / (d10.x * d20.y - d10.y * d20.x);
    float3 T = (e01 * d20.y - e02 * d10.y) * det_inverse;
    float3 B = (e02 * d10.x - e01 * d20.x) * det_inverse;

    RGBA n = sample_tex2d_rgba(material_texture(rd, normal_map_texture_index), texcoords, false);
    float3 normal_tangent_space = normalize(f3(n.r - 0.5f, n.g - 0.5f, n.b - 0.5f));
    return local_to_world_frame(normalize(T), normalize(B), surface_normal, normal_tangent_space);
}

// get_shading_normal (Intersect.h:64-83)
PT_DEV float3 get_shading_normal(const HIPRTRenderData& rd, float3 geometric_normal, const TriangleShading& ts, float2 uv, float2 texcoords)
{
    float3 surface_normal;
    if (ts.has_normals)
        surface_normal = normalize(interpolate_normals(ts, uv));
    else
        surface_normal = geometric_normal;

    int normal_map = material_field_i(rd, ts.material_index, offsetof(RendererMaterial, normal_map_texture_index));
    if (normal_map != B200PT_NO_TEXTURE)
        surface_normal = normal_mapping(rd, normal_map, triangle_vertex_indices(rd, ts.prim), texcoords, surface_normal);   // (positions: by index, rare)
    return surface_normal;
}

// Geometric normal as the oracle produces it: Triangle::intersect normalises cross(e1, e2) (Triangle.h:53) and
// trace_ray normalises the result once more (Intersect.h:152).
PT_DEV float3 hit_geometric_normal(const HIPRTRenderData& rd, const Hit& hit)
{
    return normalize(normalize(bvh8_triangle_normal(scene_bvh(rd), hit.tri_slot)));
}

// filter_function with do_alpha_testing (FilterFunction.h:33-47): a candidate hit is kept iff rng() < alpha_opacity * alpha of
// the base colour texture at the hit. One number of the path's generator is drawn per candidate, in traversal order (so the
// stream, unlike the image statistics, depends on the acceleration structure -- also true between the reference's own CPU
// and GPU paths).
struct AlphaHitFilter
{
    const HIPRTRenderData& rd;
    Rng& rng;
    PT_DEV bool reject(int prim, float u, float v) const
    {
        const int material_index = __ldg(rd.buffers.material_indices + prim);
        const float base_alpha = base_color_alpha(rd, material_index, interpolate_texcoords(rd, triangle_vertex_indices(rd, prim), make_float2(u, v)));
        const float composited_alpha = material_field_f(rd, material_index, offsetof(SimplifiedRendererMaterial, alpha_opacity)) * base_alpha;
        return !(rng() < composited_alpha);
    }
};

template <bool ANY_HIT>
PT_DEV bool trace_filtered(const HIPRTRenderData& rd, float3 origin, float3 direction, float t_limit, int last_hit_primitive_index, Hit& hit, Rng& rng)
{
    if (rd.render_settings.do_alpha_testing)
        return bvh8_trace<ANY_HIT, AlphaHitFilter>(scene_bvh(rd), origin, direction, t_limit, last_hit_primitive_index, hit, PT_COUNTERS_PTR, AlphaHitFilter{rd, rng});
    return bvh8_trace<ANY_HIT>(scene_bvh(rd), origin, direction, t_limit, last_hit_primitive_index, hit, PT_COUNTERS_PTR);
}

// trace_ray (Intersect.h:114-219). Returns true if a hit was found. `first`: when have_first is set, the closest hit of
// (origin, direction) was already found by the ray-queue kernel (trace_queue.cuh) and is passed in instead of traced here;
// volume boundaries that are skipped (nested dielectrics) are re-traced in place either way.
PT_DEV_NOINLINE bool trace_ray_from(const HIPRTRenderData& rd, float3 origin, float3 direction, RayPayload& payload, HitInfo& out, int last_hit_primitive_index, Rng& rng, Hit hit, bool have_first)
{
    bool skipping_volume_boundary = false;
    do
    {
        if (have_first)
        {
            if (hit.prim < 0)
                return false;
        }
        else if (!trace_filtered<false>(rd, origin, direction, INFINITY, last_hit_primitive_index, hit, rng))
            return false;
        have_first = false;

        const TriangleShading ts = load_triangle_shading(rd, hit.prim, hit.tri_slot);
        float2 uv = make_float2(hit.u, hit.v);
        out.inter_point = origin + hit.t * direction;
        out.primitive_index = hit.prim;
        out.texcoords = interpolate_texcoords(ts, uv);
        out.geometric_normal = hit_geometric_normal(rd, hit);
        out.shading_normal = get_shading_normal(rd, out.geometric_normal, ts, uv, out.texcoords);
        out.t = hit.t;
        out.uv = uv;

        if (is_inside_volume(payload.volume_state))
            payload.volume_state.distance_in_volume += hit.t;

        int material_index = ts.material_index;
        out.material_ref = get_intersection_material(rd, material_index, out.texcoords, payload.material) ? material_index + 1 : 0;

        if ((!is_inside_volume(payload.volume_state) || payload.material.specular_transmission == 0.0f) && !payload.material.thin_walled)
        {
            out.geometric_normal *= dot(out.geometric_normal, -direction) < 0.0f ? -1.0f : 1.0f;
            out.shading_normal *= dot(out.shading_normal, out.geometric_normal) < 0.0f ? -1.0f : 1.0f;
            float NdotV = dot(out.shading_normal, -direction);
            out.shading_normal += (2.0f * clampf(0.0f, 1.0f, -NdotV)) * -direction;
        }

        skipping_volume_boundary = interior_stack_push(payload.volume_state, material_index, payload.material.dielectric_priority);
        if (skipping_volume_boundary)
        {
            origin = out.inter_point;
            // counted twice on skipped boundaries, as in the reference (Intersect.h:161-162, 203)
            payload.volume_state.distance_in_volume += hit.t;
        }
    } while (skipping_volume_boundary);

    if (payload.material.dispersion_scale > 0.0f && payload.material.specular_transmission > 0.0f && payload.volume_state.sampled_wavelength == 0.0f)
        payload.volume_state.sampled_wavelength = -sample_wavelength_uniformly(rng);

    return true;
}

PT_DEV bool trace_ray(const HIPRTRenderData& rd, float3 origin, float3 direction, RayPayload& payload, HitInfo& out, int last_hit_primitive_index, Rng& rng)
{
    Hit none;
    none.prim = -1; none.tri_slot = -1; none.t = -1.0f; none.u = 0.0f; none.v = 0.0f;
    return trace_ray_from(rd, origin, direction, payload, out, last_hit_primitive_index, rng, none, false);
}

// evaluate_shadow_ray (Intersect.h:224-254, the device branch): true if occluded; one any-hit query through the filter
PT_DEV bool evaluate_shadow_ray(const HIPRTRenderData& rd, float3 origin, float3 direction, float t_max, int last_hit_primitive_index, Rng& rng)
{
    Hit hit;
    return trace_filtered<true>(rd, origin, direction, t_max - 1.0e-4f, last_hit_primitive_index, hit, rng);
}

struct ShadowLightRayHitInfo
{
    int hit_prim_index;
    float hit_distance;
    float3 hit_shading_normal;
    RGB hit_emission;
};

// evaluate_shadow_light_ray (Intersect.h:293-410): closest hit, then the emission + shading normal of what was hit
// Second half of evaluate_shadow_light_ray: what the closest hit `hit` of the ray emits (Intersect.h:377-409)
PT_DEV_NOINLINE bool shadow_light_ray_resolve(const HIPRTRenderData& rd, float t_max, ShadowLightRayHitInfo& out, const Hit& hit)
{
    if (hit.prim < 0)
        return false;
    if (!(hit.t < t_max - 1.0e-4f))
        return false;

    const TriangleShading ts = load_triangle_shading(rd, hit.prim, hit.tri_slot);
    int material_index = ts.material_index;
    int emission_texture_index = material_field_i(rd, material_index, offsetof(RendererMaterial, emission_texture_index));
    float2 uv = make_float2(hit.u, hit.v);
    float2 texcoords = interpolate_texcoords(ts, uv);
    if (emission_texture_index != B200PT_NO_TEXTURE)
    {
        // get_material_property leaves the (uninitialised) output untouched for CONSTANT_EMISSIVE_TEXTURE; black here
        out.hit_emission = RGB(0.0f);
        if (has_texture(emission_texture_index))
        {
            RGBA c = sample_tex2d_rgba(material_texture(rd, emission_texture_index), texcoords, false);
            out.hit_emission = RGB(c.r, c.g, c.b);
        }
    }
    else
        out.hit_emission = material_buffer_emission(rd, material_index);

    float3 ng = normalize(normalize(bvh8_triangle_normal(scene_bvh(rd), hit.tri_slot)));
    out.hit_shading_normal = get_shading_normal(rd, ng, ts, uv, texcoords);
    out.hit_distance = hit.t;
    out.hit_prim_index = hit.prim;
    return true;
}

PT_DEV bool evaluate_shadow_light_ray(const HIPRTRenderData& rd, float3 origin, float3 direction, float t_max, ShadowLightRayHitInfo& out, int last_hit_primitive_index, Rng& rng)
{
    Hit hit;
    trace_filtered<false>(rd, origin, direction, INFINITY, last_hit_primitive_index, hit, rng);
    return shadow_light_ray_resolve(rd, t_max, out, hit);
}
} // namespace pt
