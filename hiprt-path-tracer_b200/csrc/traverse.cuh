// Software ray traversal of the compressed BVH8 for sm_100a (B200 has no RT cores).
//
// Replaces hiprtGeomTraversalClosestCustomStack / hiprtGeomTraversalAnyHitCustomStack + filter_function
// (reference: src/Device/includes/Intersect.h:120-145,227-255; src/Device/functions/FilterFunction.h:19-48).
// Semantics follow the reference's CPU oracle, which is the parity spec:
//   * triangle routine = Triangle::intersect (src/Renderer/Triangle.h:20-63): Moeller-Trumbore, reject |det| < 1e-7,
//     u in [0,1], v >= 0, u+v <= 1, t > 1e-7; evaluated with explicit round-to-nearest mul/add/div (never fused), in
//     the same order, so t,u,v are bit-identical to the g++ -ffp-contract=off build;
//   * closest hit = strict '<' on t over every accepted triangle (src/Renderer/BVH.h:175);
//   * a hit on `last_prim` is always filtered out (FilterFunction.h:22-30);
//   * shadow rays: occluded iff some accepted triangle has t < t_max - 1e-4 (Intersect.h:229,284).
// Box tests are conservative (quantised boxes only grow, slab results get a relative slack), so the set of tested
// triangles is a superset of the triangles that can win and the decision is the triangle routine's alone.
#pragma once

#include "bvh8.h"
#include "launch.h"
#include "vec.cuh"

namespace pt
{
struct Hit
{
    int prim;        // scene triangle index, -1 = miss
    int tri_slot;    // index in the BVH triangle array (to re-fetch the vertices)
    float t, u, v;
};

#ifdef PT_COUNTERS
#define PT_COUNT(x) x
#else
#define PT_COUNT(x)
#endif

// Replicates the sign bit of each byte over the byte. PRMT with selector nibbles 8..b (bit 3 = "replicate sign");
// __byte_perm() masks its selector to 3 bits, hence the inline PTX.
PT_DEV uint32_t sign_extend_s8x4(uint32_t x)
{
#ifdef __CUDA_ARCH__
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(0u), "r"(0x0000ba98u));
    return r;
#else
    return ((x & 0x80u) ? 0xFFu : 0u) | ((x & 0x8000u) ? 0xFF00u : 0u) | ((x & 0x800000u) ? 0xFF0000u : 0u) | ((x & 0x80000000u) ? 0xFF000000u : 0u);
#endif
}
PT_DEV float byte_to_float(uint32_t word, int byte) { return (float)((word >> (8 * byte)) & 0xFFu); }

// Triangle::intersect of the reference, bit for bit. Returns true when the reference would report a hit.
PT_DEV bool intersect_triangle_reference(float3 o, float3 d, float3 a, float3 b, float3 c, float& t, float& u, float& v)
{
    const float EPSILON = 0.0000001f;
    float e1x = __fsub_rn(b.x, a.x), e1y = __fsub_rn(b.y, a.y), e1z = __fsub_rn(b.z, a.z);
    float e2x = __fsub_rn(c.x, a.x), e2y = __fsub_rn(c.y, a.y), e2z = __fsub_rn(c.z, a.z);
    // h = cross(d, e2)
    float hx = __fsub_rn(__fmul_rn(d.y, e2z), __fmul_rn(d.z, e2y));
    float hy = __fsub_rn(__fmul_rn(d.z, e2x), __fmul_rn(d.x, e2z));
    float hz = __fsub_rn(__fmul_rn(d.x, e2y), __fmul_rn(d.y, e2x));
    float det = __fadd_rn(__fadd_rn(__fmul_rn(e1x, hx), __fmul_rn(e1y, hy)), __fmul_rn(e1z, hz));
    if (det > -EPSILON && det < EPSILON)
        return false;
    float f = __fdiv_rn(1.0f, det);
    float sx = __fsub_rn(o.x, a.x), sy = __fsub_rn(o.y, a.y), sz = __fsub_rn(o.z, a.z);
    u = __fmul_rn(f, __fadd_rn(__fadd_rn(__fmul_rn(sx, hx), __fmul_rn(sy, hy)), __fmul_rn(sz, hz)));
    if (u < 0.0f || u > 1.0f)
        return false;
    // q = cross(s, e1)
    float qx = __fsub_rn(__fmul_rn(sy, e1z), __fmul_rn(sz, e1y));
    float qy = __fsub_rn(__fmul_rn(sz, e1x), __fmul_rn(sx, e1z));
    float qz = __fsub_rn(__fmul_rn(sx, e1y), __fmul_rn(sy, e1x));
    v = __fmul_rn(f, __fadd_rn(__fadd_rn(__fmul_rn(d.x, qx), __fmul_rn(d.y, qy)), __fmul_rn(d.z, qz)));
    if (v < 0.0f || __fadd_rn(u, v) > 1.0f)
        return false;
    t = __fmul_rn(f, __fadd_rn(__fadd_rn(__fmul_rn(e2x, qx), __fmul_rn(e2y, qy)), __fmul_rn(e2z, qz)));
    return t > EPSILON;
}

// The any-hit callback of the reference (filter_function, src/Device/functions/FilterFunction.h:19-48) beyond the
// last-primitive rejection, which every traversal does: a FILTER sees each candidate hit that would otherwise be accepted and
// may reject it (alpha testing, AlphaHitFilter in intersect.cuh). NoHitFilter compiles away.
struct NoHitFilter
{
    PT_DEV bool reject(int, float, float) const { return false; }
};

// ANY_HIT = false: closest hit over t in (1e-7, +inf).  ANY_HIT = true: first accepted triangle with t < t_limit.
template <bool ANY_HIT, class FILTER = NoHitFilter>
PT_DEV_NOINLINE bool bvh8_trace(const Bvh8Device& bvh, float3 o, float3 d, float t_limit, int last_prim, Hit& hit, TraversalCounters* counters = nullptr, FILTER filter = FILTER())
{
    hit.prim = -1;
    hit.tri_slot = -1;
    hit.t = -1.0f;
    hit.u = 0.0f;
    hit.v = 0.0f;
    float t_best = t_limit;   // closest: +inf; any-hit: the occlusion distance

    // Reciprocal direction, clamped away from +-inf so that 0 * idir never yields NaN
    const float kTiny = 1.0e-20f;
    float idx = 1.0f / (fabsf(d.x) > kTiny ? d.x : copysignf(kTiny, d.x));
    float idy = 1.0f / (fabsf(d.y) > kTiny ? d.y : copysignf(kTiny, d.y));
    float idz = 1.0f / (fabsf(d.z) > kTiny ? d.z : copysignf(kTiny, d.z));
    // Slot s (bit a set = child on the + side of axis a) is visited first when the ray comes from that side
    // The signs are those of the RECIPROCALS: a component of -0.0 has a negative reciprocal (copysignf above) although it does not
    // compare below zero; taking near / far planes by `d < 0` then inverted the slab interval of that axis and culled every box
    // (found on the 10 M-triangle scene: 2 of 2.1e6 oracle rays, both with d.z = -0.0; tests/test_gpu_parity.py::test_c3_hit_ids_...).
    const uint32_t oct_inv = (idx > 0.0f ? 1u : 0u) | (idy > 0.0f ? 2u : 0u) | (idz > 0.0f ? 4u : 0u);
    const uint32_t oct_inv4 = oct_inv * 0x01010101u;
    const bool neg_x = idx < 0.0f, neg_y = idy < 0.0f, neg_z = idz < 0.0f;

    uint2 stack[kTraversalStackSize];
    int stack_size = 0;
    uint2 ngroup = make_uint2(0u, 0x80000000u);   // the root plays "internal child in the top-priority slot"

    PT_COUNT(unsigned int n_nodes = 0; unsigned int n_tris = 0; unsigned int n_overflows = 0;)

    while (true)
    {
        uint2 tgroup;
        if (ngroup.y > 0x00FFFFFFu)
        {
            const uint32_t hits = ngroup.y;
            const uint32_t imask = hits & 0xFFu;
            const uint32_t bit = 31u - (uint32_t)__clz((int)hits);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y > 0x00FFFFFFu)
            {
                // A full stack cannot happen on a tree whose wide depth d satisfies d <= kTraversalStackSize (one deferred group per
                // level; the builder reports d and b200pt_bvh_build refuses deeper trees). The counting build still counts it.
                if (stack_size < kTraversalStackSize)
                    stack[stack_size++] = ngroup;
                else { PT_COUNT(n_overflows++;) }
            }
            const uint32_t slot = (bit - 24u) ^ oct_inv;
            const uint32_t rank = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
            const uint32_t node_index = ngroup.x + rank;

            const uint4* np = reinterpret_cast<const uint4*>(bvh.nodes + node_index);
            const uint4 n0 = __ldg(np + 0), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            PT_COUNT(n_nodes++;)

            const float px = __uint_as_float(n0.x), py = __uint_as_float(n0.y), pz = __uint_as_float(n0.z);
            const float adx = __uint_as_float((n0.w & 0xFFu) << 23) * idx;
            const float ady = __uint_as_float(((n0.w >> 8) & 0xFFu) << 23) * idy;
            const float adz = __uint_as_float(((n0.w >> 16) & 0xFFu) << 23) * idz;
            const float ox = (px - o.x) * idx, oy = (py - o.y) * idy, oz = (pz - o.z) * idz;
            // Additive pad per axis, the rule of the ray-queue kernel (trace_queue.cuh): the rounding error of a plane distance
            // fma(q, a, o) scales with |o| and with the cell size |a|, not with the distance itself, so a purely relative slack
            // is not conservative for short rays that start far from the node origin. 2.55e-4 |a| + 1e-6 |o| = 1e-6 of the largest
            // plane distance of the axis (|t| <= |o| + 255 |a|).
            const float ex = __fmaf_rn(fabsf(adx), 2.55e-4f, fabsf(ox) * 1.0e-6f), ey = __fmaf_rn(fabsf(ady), 2.55e-4f, fabsf(oy) * 1.0e-6f),
                        ez = __fmaf_rn(fabsf(adz), 2.55e-4f, fabsf(oz) * 1.0e-6f);
            const float oxn = ox - ex, oyn = oy - ey, ozn = oz - ez, oxf = ox + ex, oyf = oy + ey, ozf = oz + ez;

            uint32_t hitmask = 0;
#pragma unroll
            for (int half = 0; half < 2; half++)
            {
                const uint32_t meta4 = half ? n1.w : n1.z;
                const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
                const uint32_t bit_index4 = (meta4 ^ (oct_inv4 & inner_mask4)) & 0x1F1F1F1Fu;
                const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;

                const uint32_t lox = half ? n2.y : n2.x, loy = half ? n2.w : n2.z, loz = half ? n3.y : n3.x;
                const uint32_t hix = half ? n3.w : n3.z, hiy = half ? n4.y : n4.x, hiz = half ? n4.w : n4.z;
                const uint32_t nearx = neg_x ? hix : lox, farx = neg_x ? lox : hix;
                const uint32_t neary = neg_y ? hiy : loy, fary = neg_y ? loy : hiy;
                const uint32_t nearz = neg_z ? hiz : loz, farz = neg_z ? loz : hiz;
#pragma unroll
                for (int j = 0; j < 4; j++)
                {
                    const float tnx = __fmaf_rn(byte_to_float(nearx, j), adx, oxn);
                    const float tny = __fmaf_rn(byte_to_float(neary, j), ady, oyn);
                    const float tnz = __fmaf_rn(byte_to_float(nearz, j), adz, ozn);
                    const float tfx = __fmaf_rn(byte_to_float(farx, j), adx, oxf);
                    const float tfy = __fmaf_rn(byte_to_float(fary, j), ady, oyf);
                    const float tfz = __fmaf_rn(byte_to_float(farz, j), adz, ozf);
                    const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
                    const float cmax = fminf(fminf(fminf(tfx, tfy), tfz), t_best);
                    if (cmin <= cmax)
                    {
                        const uint32_t child_bits = (child_bits4 >> (8 * j)) & 0xFFu;
                        const uint32_t bit_index = (bit_index4 >> (8 * j)) & 0xFFu;
                        hitmask |= child_bits << bit_index;
                    }
                }
            }
            ngroup.x = n1.x;
            ngroup.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
            tgroup.x = n1.y;
            tgroup.y = hitmask & 0x00FFFFFFu;
        }
        else
        {
            tgroup = make_uint2(0u, 0u);
        }

        while (tgroup.y != 0u)
        {
            const uint32_t tbit = (uint32_t)__ffs((int)tgroup.y) - 1u;
            tgroup.y &= tgroup.y - 1u;
            const uint32_t tri_index = tgroup.x + tbit;
            const float4* tp = reinterpret_cast<const float4*>(bvh.triangles + tri_index);
            const float4 v0 = __ldg(tp + 0), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
            PT_COUNT(n_tris++;)
            const int prim = __float_as_int(v0.w);
            if (prim == last_prim)
                continue;
            float t, u, v;
            if (intersect_triangle_reference(o, d, f3(v0.x, v0.y, v0.z), f3(v1.x, v1.y, v1.z), f3(v2.x, v2.y, v2.z), t, u, v))
            {
                // strict '<' as in the reference (BVH.h:132-227); among hits at exactly the same distance (a ray through a shared
                // edge) the lowest primitive id wins, so the result does not depend on the order the triangles were visited in
                if ((t < t_best || (!ANY_HIT && t == t_best && prim < hit.prim)) && !filter.reject(prim, u, v))
                {
                    t_best = t;
                    hit.prim = prim;
                    hit.tri_slot = (int)tri_index;
                    hit.t = t;
                    hit.u = u;
                    hit.v = v;
                    if (ANY_HIT)
                    {
                        PT_COUNT(if (counters) { atomicAdd(&counters->rays, 1ull); atomicAdd(&counters->nodes, (unsigned long long)n_nodes); atomicAdd(&counters->triangles, (unsigned long long)n_tris); if (n_overflows) atomicAdd(&counters->overflows, (unsigned long long)n_overflows); })
                        return true;
                    }
                }
            }
        }

        if (ngroup.y <= 0x00FFFFFFu)
        {
            if (stack_size == 0)
                break;
            ngroup = stack[--stack_size];
        }
    }

    PT_COUNT(if (counters) { atomicAdd(&counters->rays, 1ull); atomicAdd(&counters->nodes, (unsigned long long)n_nodes); atomicAdd(&counters->triangles, (unsigned long long)n_tris); if (n_overflows) atomicAdd(&counters->overflows, (unsigned long long)n_overflows); })
    return hit.prim >= 0;
}

// Unnormalised geometric normal cross(B-A, C-A) of a BVH triangle (Triangle.h:53 normalises it afterwards)
PT_DEV float3 bvh8_triangle_normal(const Bvh8Device& bvh, int tri_slot)
{
    const float4* tp = reinterpret_cast<const float4*>(bvh.triangles + tri_slot);
    const float4 v0 = __ldg(tp + 0), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
    float3 e1 = f3(v1.x - v0.x, v1.y - v0.y, v1.z - v0.z);
    float3 e2 = f3(v2.x - v0.x, v2.y - v0.y, v2.z - v0.z);
    return cross(e1, e2);
}
} // namespace pt
