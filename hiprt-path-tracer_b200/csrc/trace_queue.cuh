// Ray-queue traversal kernel of the wavefront path tracer: persistent warps, per-lane ray replacement.
//
// ncu on the 10 M-triangle scene (profiles/r01_launches_c3_phased_before_tracequeue.txt): the kernels that traced rays as
// one-ray-per-thread batches issued instructions for 6.5 of 32 lanes on average -- incoherent rays have very different
// lengths (a sky ray leaves after 3 nodes, a grazing ray visits 100) and a warp waited for its longest ray -- and every
// node visit paid 48 I2F.U8 conversions on the quarter-rate XU pipe. Here
//   * a lane whose ray ends takes the next ray of the queue at once (one warp-aggregated atomic per refill round), so the
//     lanes of a warp stay busy until the queue is empty (after Aila & Laine, HPG 2009, "persistent threads");
//   * quantised bounds become floats with one PRMT each (byte into the mantissa of 1.0f, bias folded into the per-node offsets);
//   * a lane tests at most one triangle per iteration, next to its node step (its node work goes on meanwhile); postponing the
//     triangle until several lanes of the warp have one pending (after Ylitie et al., HPG 2017) is built in but measured slower;
//   * the first kSharedStackEntries stack entries of each lane live in shared memory (lane-interleaved), the rest spills to local memory;
//   * the kernel only traverses: 16-byte ray records in, 16+4-byte hit records (or an occlusion flag) out; shading data is
//     fetched by the kernels that consume the hits.
// Per-ray arithmetic (slab test, visiting order, triangle routine, strict '<' closest rule, last-primitive filter) is the same
// as bvh8_trace (traverse.cuh), so the hits are identical and parity with the reference (src/Renderer/BVH.h:132-227,
// Triangle.h:20-63, FilterFunction.h:19-48) is unchanged.
#pragma once

#include "launch.h"
#include "traverse.cuh"

namespace pt
{
// Tunables (-D overrides exist for the sweeps of tools/trace_sweep.sh; the defaults are the measured best)
#ifndef PT_TRACE_SHARED_STACK
#define PT_TRACE_SHARED_STACK 8
#endif
#ifndef PT_TRACE_BLOCKS_PER_SM
#define PT_TRACE_BLOCKS_PER_SM 7
#endif
#ifndef PT_TRACE_REFILL_IDLE
#define PT_TRACE_REFILL_IDLE 8
#endif
#ifndef PT_TRACE_TRI_POSTPONE
#define PT_TRACE_TRI_POSTPONE 1
#endif
constexpr int kSharedStackEntries = PT_TRACE_SHARED_STACK;
constexpr int kTraceBlock = 128;
constexpr int kTraceBlocksPerSM = PT_TRACE_BLOCKS_PER_SM;         // 72 registers per thread; 64 (8 blocks) measured no faster
// Sweep on B200 (tools/trace_sweep.sh, profiles/r01_trace_sweep.txt), trace stage of a C3 pass: refill at 1 / 2 / 4 / 8 / 16 idle
// lanes 7.36 / 7.02 / 6.77 / 6.73 / 7.00 ms; triangle postponement below 1 (= none) / 4 / 8 / 16 / 24 lanes 6.47 / 6.59 / 6.76 /
// 7.17 / 8.22 ms. With one triangle per iteration and the node step of the other lanes going on in the same iteration, waiting
// for fuller triangle warps only delays the rays: postponement is off by default (POSTPONE = false compiles the vote out).
constexpr unsigned kRefillIdleLanes = PT_TRACE_REFILL_IDLE;       // refill as soon as this many lanes of a warp have no ray
constexpr int kTrianglePostponeBelow = PT_TRACE_TRI_POSTPONE;     // > 1: triangles are tested when at least this many lanes of the warp have one pending

// 1 + byte * 2^-15, exactly: the byte lands in mantissa bits 8..15 of 1.0f with one PRMT (result bytes, low to high:
// 0x00, byte j of word, 0x80, 0x3F). The slab test then evaluates plane = (1 + q 2^-15) * (2^15 a) + (o - 2^15 a), the same
// value as q * a + o up to the rounding of the folded offset (<= 2^-9 of a grid cell), which the per-node offsets below
// absorb on the conservative side -- no I2F (quarter-rate XU pipe) and no separate bias subtraction per plane.
// 'one' is 0x3F800000 held in a register (opaque_one_bits): PRMT has a single immediate slot, and with a literal constant
// ptxas spends it on the constant and re-materialises the SELECTOR in a register before every PRMT (one extra MOV per plane,
// 48 per node in the first version of this kernel, see profiles/r01_trace_queue_sass_notes.txt).
PT_DEV float quantised_byte_to_unit_float(uint32_t word, int byte, uint32_t one)
{
    return __uint_as_float(__byte_perm(word, one, 0x7604u + ((uint32_t)byte << 4)));
}
PT_DEV uint32_t opaque_one_bits(uint32_t queue_count)
{
    return 0x3F800000u | (queue_count >> 31);    // a queue holds < 2^31 rays: always 0x3F800000, but not a compile-time constant to ptxas
}

// Non-blocking hint: the line is on its way into L2/L1 while the lane does other work (ncu, profiles/r01_trace_queue_closest_c3_v2.txt:
// a third of the stall samples of this kernel were long-scoreboard waits on the triangle fetch, which misses L2 -- the 480 MB
// triangle array of the 10 M-triangle scene does not fit -- and another third on the node fetch).
PT_DEV void prefetch_global(const void* p)
{
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}

template <bool ANY_HIT, bool POSTPONE>
__device__ __forceinline__ void trace_queue_body(const Bvh8Device& bvh, const RayQueueView& q, uint2* shared_stack, TraversalCounters* counters,
                                                 const unsigned refill_idle_lanes = kRefillIdleLanes, const int triangle_postpone_below = kTrianglePostponeBelow)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = *q.count;
    uint2* const my_shared = shared_stack + threadIdx.x;          // entry k at my_shared[k * blockDim.x]
    uint2 local_stack[kTraversalStackSize - kSharedStackEntries];
    const uint32_t one = opaque_one_bits(n);

    bool active = false, exhausted = false;
    uint32_t ray_id = 0;
    float3 o = f3(0.0f), d = f3(0.0f);
    float idx = 0.0f, idy = 0.0f, idz = 0.0f, t_best = 0.0f;
    int last_prim = -1;
    uint32_t oct_inv = 0;
    Hit hit;
    hit.prim = -1; hit.tri_slot = -1; hit.t = -1.0f; hit.u = 0.0f; hit.v = 0.0f;
    uint2 ngroup = make_uint2(0u, 0u);   // (first child node, hit internal children << 24 | imask): node work of this lane
    uint2 tgroup = make_uint2(0u, 0u);   // (first triangle, hit-leaf triangle bits): triangle work of this lane
    int stack_size = 0;
    PT_COUNT(unsigned int n_nodes = 0; unsigned int n_tris = 0; unsigned int n_rays = 0;)

    while (true)
    {
        // ---- refill: lanes without a ray take the next queue entries -------------------------------------------------
        const unsigned idle_mask = __ballot_sync(0xFFFFFFFFu, !active);
        if (!exhausted && (__popc(idle_mask) >= refill_idle_lanes || idle_mask == 0xFFFFFFFFu))
        {
            uint32_t base = 0;
            if (lane == (uint32_t)(__ffs((int)idle_mask) - 1))
                base = atomicAdd(q.cursor, (uint32_t)__popc(idle_mask));
            base = __shfl_sync(0xFFFFFFFFu, base, __ffs((int)idle_mask) - 1);
            exhausted = base + (uint32_t)__popc(idle_mask) >= n;
            if (!active)
            {
                const uint32_t i = base + (uint32_t)__popc(idle_mask & ((1u << lane) - 1u));
                if (i < n)
                {
                    ray_id = q.ids[i];
                    const float4 ro = q.ray_o[ray_id], rdir = q.ray_d[ray_id];
                    o = f3(ro.x, ro.y, ro.z);
                    d = f3(rdir.x, rdir.y, rdir.z);
                    last_prim = __float_as_int(rdir.w);
                    t_best = ANY_HIT ? ro.w - 1.0e-4f : INFINITY;    // evaluate_shadow_ray: occluded iff t < t_max - 1e-4 (Intersect.h:229)
                    const float kTiny = 1.0e-20f;
                    idx = 1.0f / (fabsf(d.x) > kTiny ? d.x : copysignf(kTiny, d.x));
                    idy = 1.0f / (fabsf(d.y) > kTiny ? d.y : copysignf(kTiny, d.y));
                    idz = 1.0f / (fabsf(d.z) > kTiny ? d.z : copysignf(kTiny, d.z));
                    oct_inv = (d.x > 0.0f ? 1u : 0u) | (d.y > 0.0f ? 2u : 0u) | (d.z > 0.0f ? 4u : 0u);
                    hit.prim = -1; hit.tri_slot = -1; hit.t = -1.0f; hit.u = 0.0f; hit.v = 0.0f;
                    ngroup = make_uint2(0u, 0x80000000u);
                    tgroup = make_uint2(0u, 0u);
                    stack_size = 0;
                    active = true;
                    PT_COUNT(n_rays++;)
                }
            }
        }
        const unsigned active_mask = __ballot_sync(0xFFFFFFFFu, active);
        if (active_mask == 0u)
            break;
        if (!active)
            continue;

        // ---- one iteration: at most one node and at most one triangle of this lane's ray ------------------------------------
        bool finished = false;
        if (ngroup.y > 0x00FFFFFFu)
        {
            const uint32_t hits = ngroup.y;
            const uint32_t imask = hits & 0xFFu;
            const uint32_t bit = 31u - (uint32_t)__clz((int)hits);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y > 0x00FFFFFFu)
            {
                if (stack_size < kSharedStackEntries)
                    my_shared[stack_size * blockDim.x] = ngroup;
                else if (stack_size < kTraversalStackSize)
                    local_stack[stack_size - kSharedStackEntries] = ngroup;
                if (stack_size < kTraversalStackSize)
                    stack_size++;
            }
            const uint32_t slot = (bit - 24u) ^ oct_inv;
            const uint32_t rank = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
            const uint4* np = reinterpret_cast<const uint4*>(bvh.nodes + (ngroup.x + rank));
            const uint4 n0 = __ldg(np + 0), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            PT_COUNT(n_nodes++;)

            // cell size * 2^15 / direction, per axis; near / far offsets pushed 2^-7 of a cell outwards (see above)
            const float adx = __uint_as_float((n0.w & 0xFFu) << 23) * idx * 32768.0f;
            const float ady = __uint_as_float(((n0.w >> 8) & 0xFFu) << 23) * idy * 32768.0f;
            const float adz = __uint_as_float(((n0.w >> 16) & 0xFFu) << 23) * idz * 32768.0f;
            const float ox = (__uint_as_float(n0.x) - o.x) * idx - adx, oy = (__uint_as_float(n0.y) - o.y) * idy - ady, oz = (__uint_as_float(n0.z) - o.z) * idz - adz;
            // pad per axis: 2^-7 of a cell (2^-22 * 2^15 a) plus 1e-6 relative to the largest plane distance of the axis
            // (|t| <= |o| + 2 |2^15 a|), which covers the rounding of the offset, of the product and of the triangle routine's t
            const float ex = __fmaf_rn(fabsf(adx), 2.2384186e-6f, fabsf(ox) * 1.0e-6f), ey = __fmaf_rn(fabsf(ady), 2.2384186e-6f, fabsf(oy) * 1.0e-6f),
                        ez = __fmaf_rn(fabsf(adz), 2.2384186e-6f, fabsf(oz) * 1.0e-6f);
            const float oxn = ox - ex, oyn = oy - ey, ozn = oz - ez, oxf = ox + ex, oyf = oy + ey, ozf = oz + ez;
            const bool neg_x = d.x < 0.0f, neg_y = d.y < 0.0f, neg_z = d.z < 0.0f;
            const uint32_t oct_inv4 = oct_inv * 0x01010101u;

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
                    const float tnx = __fmaf_rn(quantised_byte_to_unit_float(nearx, j, one), adx, oxn);
                    const float tny = __fmaf_rn(quantised_byte_to_unit_float(neary, j, one), ady, oyn);
                    const float tnz = __fmaf_rn(quantised_byte_to_unit_float(nearz, j, one), adz, ozn);
                    const float tfx = __fmaf_rn(quantised_byte_to_unit_float(farx, j, one), adx, oxf);
                    const float tfy = __fmaf_rn(quantised_byte_to_unit_float(fary, j, one), ady, oyf);
                    const float tfz = __fmaf_rn(quantised_byte_to_unit_float(farz, j, one), adz, ozf);
                    const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
                    const float cmax = fminf(fminf(fminf(tfx, tfy), tfz), t_best);
                    if (cmin <= cmax)
                        hitmask |= ((child_bits4 >> (8 * j)) & 0xFFu) << ((bit_index4 >> (8 * j)) & 0xFFu);
                }
            }
            ngroup.x = n1.x;
            ngroup.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
            if ((hitmask & 0x00FFFFFFu) != 0u)
            {
                // the node handed out triangles: they become the lane's triangle work, older triangle work waits on the stack
                if (tgroup.y != 0u)
                {
                    if (stack_size < kSharedStackEntries)
                        my_shared[stack_size * blockDim.x] = tgroup;
                    else if (stack_size < kTraversalStackSize)
                        local_stack[stack_size - kSharedStackEntries] = tgroup;
                    if (stack_size < kTraversalStackSize)
                        stack_size++;
                }
                tgroup = make_uint2(n1.y, hitmask & 0x00FFFFFFu);
                // the triangles are tested in later iterations (one per iteration, possibly postponed): start fetching them now
                const char* first_triangle = reinterpret_cast<const char*>(bvh.triangles + (n1.y + (uint32_t)__ffs((int)tgroup.y) - 1u));
                prefetch_global(first_triangle);
                prefetch_global(first_triangle + 128);
            }
            if (ngroup.y > 0x00FFFFFFu)
            {
                // ... and the node the next iteration visits (the nearest hit child in ray-octant order)
                const uint32_t next_bit = 31u - (uint32_t)__clz((int)ngroup.y);
                const uint32_t next_slot = (next_bit - 24u) ^ oct_inv;
                const uint32_t next_rank = (uint32_t)__popc(ngroup.y & 0xFFu & ~(0xFFFFFFFFu << next_slot));
                const char* next_node = reinterpret_cast<const char*>(bvh.nodes + (ngroup.x + next_rank));
                prefetch_global(next_node);
                prefetch_global(next_node + 64);
            }
        }

        // One triangle per lane, and only when enough lanes of the warp have one: the triangle code runs with fuller warps
        // while the other lanes' triangle work waits (their node work goes on).
        bool test_triangle = tgroup.y != 0u;
        if (POSTPONE)
        {
            const unsigned triangle_lanes = __ballot_sync(active_mask, tgroup.y != 0u);
            test_triangle = test_triangle && __popc(triangle_lanes) >= min(triangle_postpone_below, __popc(active_mask));
        }
        if (test_triangle)
        {
            const uint32_t tbit = (uint32_t)__ffs((int)tgroup.y) - 1u;
            tgroup.y &= tgroup.y - 1u;
            const uint32_t tri_index = tgroup.x + tbit;
            const float4* tp = reinterpret_cast<const float4*>(bvh.triangles + tri_index);
            const float4 v0 = __ldg(tp + 0), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
            PT_COUNT(n_tris++;)
            const int prim = __float_as_int(v0.w);
            float t, u, v;
            if (prim != last_prim && intersect_triangle_reference(o, d, f3(v0.x, v0.y, v0.z), f3(v1.x, v1.y, v1.z), f3(v2.x, v2.y, v2.z), t, u, v) &&
                (t < t_best || (!ANY_HIT && t == t_best && prim < hit.prim)))     // equal distance: lowest primitive id, whatever the visiting order
            {
                t_best = t;
                hit.prim = prim; hit.tri_slot = (int)tri_index; hit.t = t; hit.u = u; hit.v = v;
                if (ANY_HIT)
                    finished = true;
            }
        }

        // Out of node work: take the next stack entry. A node group is taken even while triangles are still pending (they are
        // tested one per iteration, possibly postponed, and the lane would otherwise idle through the node step meanwhile);
        // a triangle group that was set aside is taken only once the current one is done.
        if (!finished && ngroup.y <= 0x00FFFFFFu)
        {
            if (stack_size == 0)
                finished = tgroup.y == 0u;
            else
            {
                const int top = stack_size - 1;
                const uint2 e = top < kSharedStackEntries ? my_shared[top * blockDim.x] : local_stack[top - kSharedStackEntries];
                if (e.y > 0x00FFFFFFu)
                {
                    ngroup = e;
                    stack_size = top;
                }
                else if (tgroup.y == 0u)
                {
                    tgroup = e;
                    stack_size = top;
                }
            }
        }

        if (finished)
        {
            if (ANY_HIT)
            {
                if (hit.prim >= 0)
                    q.occluded_flag[ray_id].w = -1.0f;
            }
            else
            {
                q.hit_tuvp[ray_id] = make_float4(hit.t, hit.u, hit.v, __int_as_float(hit.prim));
                q.hit_tri_slot[ray_id] = (uint32_t)hit.tri_slot;
            }
            active = false;
        }
    }
    PT_COUNT(if (counters && n_rays) { atomicAdd(&counters->rays, (unsigned long long)n_rays); atomicAdd(&counters->nodes, (unsigned long long)n_nodes); atomicAdd(&counters->triangles, (unsigned long long)n_tris); })
}
} // namespace pt
