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
//     triangle until several lanes of the warp have one pending (after Ylitie et al., HPG 2017) was measured slower at every
//     threshold in round 1 (profiles/r01_trace_sweep.txt) and is no longer built;
//   * PIPELINED flavour (trace_queue_body_pipelined): the loads of an iteration's node and triangle are issued one iteration
//     ahead, so the node fetch flies under the triangle test and the triangle fetch under the next box tests;
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
#ifndef PT_TRACE_PIPELINED_DEFAULT
#define PT_TRACE_PIPELINED_DEFAULT 0
#endif
#ifndef PT_TRACE_POP_CULL
#define PT_TRACE_POP_CULL 0
#endif
constexpr int kSharedStackEntries = PT_TRACE_SHARED_STACK;
constexpr int kTraceBlock = 128;
constexpr int kTraceBlocksPerSM = PT_TRACE_BLOCKS_PER_SM;         // 72 registers per thread; 64 (8 blocks) measured no faster
static_assert(kTraceBlocksPerSM == kTraceGridBlocksPerSM || PT_TRACE_BLOCKS_PER_SM != 7, "launch.h: kTraceGridBlocksPerSM sizes the persistent grid");
// Sweep on B200 (tools/trace_sweep.sh, profiles/r01_trace_sweep.txt), trace stage of a C3 pass: refill at 1 / 2 / 4 / 8 / 16 idle
// lanes 7.36 / 7.02 / 6.77 / 6.73 / 7.00 ms.
constexpr unsigned kRefillIdleLanes = PT_TRACE_REFILL_IDLE;       // refill as soon as this many lanes of a warp have no ray

// 1 + byte * 2^-15, exactly: the byte lands in mantissa bits 8..15 of 1.0f with one PRMT (result bytes, low to high:
// 0x00, byte j of word, 0x80, 0x3F). The slab test then evaluates plane = (1 + q 2^-15) * (2^15 a) + (o - 2^15 a), the same
// value as q * a + o up to the rounding of the folded offset (<= 2^-9 of a grid cell), which the per-node offsets below
// absorb on the conservative side -- no I2F (quarter-rate XU pipe) and no separate bias subtraction per plane.
// 'one' is 0x3F800000 held in a register (opaque_one_bits): PRMT has a single immediate slot, and with a literal constant
// ptxas spends it on the constant and re-materialises the SELECTOR in a register before every PRMT (one extra MOV per plane,
// 48 per node in the first version of this kernel; the census of the current SASS is profiles/r02_trace_queue_sass_notes.txt, tools/sass_notes.py).
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
#ifdef __CUDA_ARCH__
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
    (void)p;    // host build of the traversal logic (tests/host_shim)
#endif
}

// Triangle fetch policy (PT_TRACE_TRI_LOAD): the 48-byte triangle records of a big scene (480 MB on C3) are touched about once per ray batch,
// the 80-byte nodes of the upper levels by every ray. 0: plain read-only loads that allocate in L1 like the nodes (LDG.CONSTANT); 1: streaming
// loads (ld.global.cs: evict-first in L1 and L2) and the triangle prefetch into L2 only; 2: ld.global.nc.L1::no_allocate, prefetch into L2 only.
#ifndef PT_TRACE_TRI_LOAD
#define PT_TRACE_TRI_LOAD 0
#endif
PT_DEV float4 load_triangle_f4(const float4* p)
{
#if defined(__CUDA_ARCH__) && PT_TRACE_TRI_LOAD == 1
    return __ldcs(p);
#elif defined(__CUDA_ARCH__) && PT_TRACE_TRI_LOAD == 2
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}
PT_DEV void prefetch_triangle(const void* p)
{
#if defined(__CUDA_ARCH__) && PT_TRACE_TRI_LOAD != 0
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    prefetch_global(p);
#endif
}

// Per-ray constants of the slab test
struct RaySlabs
{
    float3 o, d;
    float idx, idy, idz;
    uint32_t oct_inv;
};

// The 8 box tests of one 80-byte node: bit (24 + s ^ oct_inv) of the result for a hit internal child in slot s (ray-octant order:
// the highest bit is the child to visit first), bits 0..23 = triangles of the hit leaves. Same arithmetic as round 1's kernel.
// WITH_TMIN: group_tmin = the smallest (padded, i.e. conservative) entry distance among the hit INTERNAL children: a lower bound of the
// distance of every triangle below any child of the group (+inf when no internal child is hit).
template <bool WITH_TMIN>
PT_DEV uint32_t node_hitmask_t(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4, const RaySlabs& r, const float t_best, const uint32_t one,
                               float& group_tmin)
{
    // cell size * 2^15 / direction, per axis; near / far offsets pushed 2^-7 of a cell outwards (see above)
    const float adx = __uint_as_float((n0.w & 0xFFu) << 23) * r.idx * 32768.0f;
    const float ady = __uint_as_float(((n0.w >> 8) & 0xFFu) << 23) * r.idy * 32768.0f;
    const float adz = __uint_as_float(((n0.w >> 16) & 0xFFu) << 23) * r.idz * 32768.0f;
    const float ox = (__uint_as_float(n0.x) - r.o.x) * r.idx - adx, oy = (__uint_as_float(n0.y) - r.o.y) * r.idy - ady, oz = (__uint_as_float(n0.z) - r.o.z) * r.idz - adz;
    // pad per axis: 2^-7 of a cell (2^-22 * 2^15 a) plus 1e-6 relative to the largest plane distance of the axis
    // (|t| <= |o| + 2 |2^15 a|), which covers the rounding of the offset, of the product and of the triangle routine's t
    const float ex = __fmaf_rn(fabsf(adx), 2.2384186e-6f, fabsf(ox) * 1.0e-6f), ey = __fmaf_rn(fabsf(ady), 2.2384186e-6f, fabsf(oy) * 1.0e-6f),
                ez = __fmaf_rn(fabsf(adz), 2.2384186e-6f, fabsf(oz) * 1.0e-6f);
    const float oxn = ox - ex, oyn = oy - ey, ozn = oz - ez, oxf = ox + ex, oyf = oy + ey, ozf = oz + ez;
    const bool neg_x = r.idx < 0.0f, neg_y = r.idy < 0.0f, neg_z = r.idz < 0.0f;     // signs of the reciprocals: -0.0 counts as negative (traverse.cuh)
    const uint32_t oct_inv4 = r.oct_inv * 0x01010101u;

    uint32_t hitmask = 0;
    float tmin = INFINITY;
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
            if (WITH_TMIN && cmin <= cmax && ((is_inner4 >> (8 * j + 4)) & 1u) != 0u)
                tmin = fminf(tmin, cmin);
        }
    }
    group_tmin = tmin;
    return hitmask;
}

PT_DEV uint32_t node_hitmask(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4, const RaySlabs& r, const float t_best, const uint32_t one)
{
    float unused;
    return node_hitmask_t<false>(n0, n1, n2, n3, n4, r, t_best, one, unused);
}

// Distance cull at pop time (PT_TRACE_POP_CULL): a node group waits on the stack while nearer subtrees are searched and t_best shrinks; without
// a distance it is fetched and box-tested again whatever the ray has found meanwhile. Bits 8..23 of a node group's second word are free (hit
// bits in 24..31, imask in 0..7): they hold the group's entry distance as the upper 16 bits of the float, i.e. truncated towards zero = rounded
// DOWN (distances are >= 0), so the bound stays conservative. A group whose bound exceeds t_best cannot hold a hit that is closer or equal:
// dropping it changes neither the hit nor the tie rule (equal distances are still visited), only the number of nodes fetched.
// MEASURED AND LEFT OFF (gpurun call 26, profiles/r02_pop_cull_and_surface_occupancy.txt): on C3 the cull removes 0.36 % of the node visits
// (17.815 -> 17.751 nodes per ray; C2 2.035 -> 2.031) -- a ray's deferred groups almost always start in front of its final hit, the per-child
// test against t_best when a node IS fetched already does the culling that matters -- while the 8 extra FMNMX per node and the pop loop cost the
// trace stage 4.9 % (83.3 -> 87.4 ms). Kept as FLAVOUR 2 of k_trace_queue (B200PT_TRACE_PIPELINE=2), hit records identical (parity tests run it).
PT_DEV uint32_t pack_group_tmin(const float tmin) { return (__float_as_uint(tmin) >> 8) & 0x00FFFF00u; }
PT_DEV float group_tmin_of(const uint2 group) { return __uint_as_float((group.y & 0x00FFFF00u) << 8); }

// Index of the node behind the highest internal-child bit of a node group (first child, hit bits << 24 | imask)
PT_DEV uint32_t group_next_node(const uint2 group, const uint32_t oct_inv, uint32_t& bit)
{
    bit = 31u - (uint32_t)__clz((int)group.y);
    const uint32_t slot = (bit - 24u) ^ oct_inv;
    return group.x + (uint32_t)__popc(group.y & 0xFFu & ~(0xFFFFFFFFu << slot));
}

// Per-lane traversal stack: the first kSharedStackEntries entries in shared memory (lane-interleaved), the rest in local memory.
// A wide tree of depth d needs at most 2 d entries (per level one deferred node group and one set-aside triangle group); the builder
// reports d and b200pt_bvh_build refuses trees with 2 d + 1 > kTraversalStackSize, so `full` cannot happen; the counting build counts it.
// On the device the shared part is addressed in the shared state space (32-bit window address, st.shared / ld.shared): through the
// generic pointer of the first version ptxas could not prove the address space and emitted generic ST.E.64 / LD.E.64 with 64-bit
// address arithmetic for every push and pop (profiles/r02_trace_queue_sass_notes.txt).
struct LaneStack
{
#ifdef __CUDA_ARCH__
    uint32_t shared_addr;      // shared-window byte address of this lane's entry 0; entry k is kTraceBlock * 8 bytes further per k
    PT_DEV void bind(uint2* lane_entry0) { shared_addr = (uint32_t)__cvta_generic_to_shared(lane_entry0); }
    // volatile, no memory clobber: the stack's own stores and loads keep their order, global loads may still move across them
    PT_DEV void shared_store(const int k, const uint2 e) { asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(shared_addr + (uint32_t)k * (kTraceBlock * 8u)), "r"(e.x), "r"(e.y)); }
    PT_DEV uint2 shared_load(const int k) const
    {
        uint2 e;
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e.x), "=r"(e.y) : "r"(shared_addr + (uint32_t)k * (kTraceBlock * 8u)));
        return e;
    }
#else
    uint2* shared_base;        // host emulation (tests/host_shim): entry k at shared_base[k * blockDim.x]
    PT_DEV void bind(uint2* lane_entry0) { shared_base = lane_entry0; }
    PT_DEV void shared_store(const int k, const uint2 e) { shared_base[k * blockDim.x] = e; }
    PT_DEV uint2 shared_load(const int k) const { return shared_base[k * blockDim.x]; }
#endif
    uint2 local[kTraversalStackSize - kSharedStackEntries];
    int size;
    PT_DEV bool push(const uint2 e)
    {
        if (size < kSharedStackEntries)
            shared_store(size, e);
        else if (size < kTraversalStackSize)
            local[size - kSharedStackEntries] = e;
        else
            return false;
        size++;
        return true;
    }
    PT_DEV uint2 top() const { return size - 1 < kSharedStackEntries ? shared_load(size - 1) : local[size - 1 - kSharedStackEntries]; }
};

PT_DEV void trace_queue_load_ray(const RayQueueView& q, const uint32_t ray_id, const bool any_hit, RaySlabs& r, int& last_prim, float& t_best)
{
    const size_t at = (size_t)ray_id * (q.ray_stride_minus_1 + 1u);
    const float4 ro = q.ray_o[at], rdir = q.ray_d[at];
    r.o = f3(ro.x, ro.y, ro.z);
    r.d = f3(rdir.x, rdir.y, rdir.z);
    last_prim = __float_as_int(rdir.w);
    t_best = any_hit ? ro.w - 1.0e-4f : INFINITY;    // evaluate_shadow_ray: occluded iff t < t_max - 1e-4 (Intersect.h:229)
    const float kTiny = 1.0e-20f;
    r.idx = 1.0f / (fabsf(r.d.x) > kTiny ? r.d.x : copysignf(kTiny, r.d.x));
    r.idy = 1.0f / (fabsf(r.d.y) > kTiny ? r.d.y : copysignf(kTiny, r.d.y));
    r.idz = 1.0f / (fabsf(r.d.z) > kTiny ? r.d.z : copysignf(kTiny, r.d.z));
    r.oct_inv = (r.idx > 0.0f ? 1u : 0u) | (r.idy > 0.0f ? 2u : 0u) | (r.idz > 0.0f ? 4u : 0u);
}

template <bool ANY_HIT>
PT_DEV void trace_queue_store_hit(const RayQueueView& q, const uint32_t ray_id, const Hit& hit)
{
    if (ANY_HIT)
    {
        if (hit.prim >= 0)
            q.occluded_flag[(size_t)ray_id * (q.ray_stride_minus_1 + 1u)].w = -1.0f;
    }
    else
    {
        q.hit_tuvp[(size_t)ray_id * (q.hit_stride_minus_1 + 1u)] = make_float4(hit.t, hit.u, hit.v, __int_as_float(hit.prim));
        q.hit_tri_slot[(size_t)ray_id * (q.hit_word_stride_minus_1 + 1u)] = (uint32_t)hit.tri_slot;
    }
}

#define PT_TRACE_FLUSH_COUNTERS()                                                                                                   \
    PT_COUNT(if (counters && n_rays) {                                                                                              \
        atomicAdd(&counters->rays, (unsigned long long)n_rays); atomicAdd(&counters->nodes, (unsigned long long)n_nodes);           \
        atomicAdd(&counters->triangles, (unsigned long long)n_tris);                                                                \
        atomicAdd(&counters->queue_rays, (unsigned long long)n_rays); atomicAdd(&counters->queue_nodes, (unsigned long long)n_nodes); \
        atomicAdd(&counters->queue_triangles, (unsigned long long)n_tris);                                                          \
        if (n_overflows) atomicAdd(&counters->overflows, (unsigned long long)n_overflows); })

template <bool ANY_HIT, bool POP_CULL = (PT_TRACE_POP_CULL != 0)>
__device__ __forceinline__ void trace_queue_body(const Bvh8Device& bvh, const RayQueueView& q, uint2* shared_stack, TraversalCounters* counters,
                                                 const unsigned refill_idle_lanes = kRefillIdleLanes)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = *q.count;
    LaneStack stack;
    stack.bind(shared_stack + threadIdx.x);
    stack.size = 0;
    const uint32_t one = opaque_one_bits(n);

    bool active = false, exhausted = false;
    uint32_t ray_id = 0;
    RaySlabs r;
    r.o = f3(0.0f); r.d = f3(0.0f); r.idx = r.idy = r.idz = 0.0f; r.oct_inv = 0;
    float t_best = 0.0f;
    int last_prim = -1;
    Hit hit;
    hit.prim = -1; hit.tri_slot = -1; hit.t = -1.0f; hit.u = 0.0f; hit.v = 0.0f;
    uint2 ngroup = make_uint2(0u, 0u);   // (first child node, hit internal children << 24 | imask): node work of this lane
    uint2 tgroup = make_uint2(0u, 0u);   // (first triangle, hit-leaf triangle bits): triangle work of this lane
    PT_COUNT(unsigned int n_nodes = 0; unsigned int n_tris = 0; unsigned int n_rays = 0; unsigned int n_overflows = 0; unsigned int n_culled = 0;)

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
                    trace_queue_load_ray(q, ray_id, ANY_HIT, r, last_prim, t_best);
                    hit.prim = -1; hit.tri_slot = -1; hit.t = -1.0f; hit.u = 0.0f; hit.v = 0.0f;
                    ngroup = make_uint2(0u, 0x80000000u);
                    tgroup = make_uint2(0u, 0u);
                    stack.size = 0;
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
            uint32_t bit;
            const uint32_t node_index = group_next_node(ngroup, r.oct_inv, bit);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y > 0x00FFFFFFu && !stack.push(ngroup)) { PT_COUNT(n_overflows++;) }
            const uint4* np = reinterpret_cast<const uint4*>(bvh.nodes + node_index);
            const uint4 n0 = __ldg(np + 0), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            PT_COUNT(n_nodes++;)
            float group_tmin;
            const uint32_t hitmask = node_hitmask_t<POP_CULL && !ANY_HIT>(n0, n1, n2, n3, n4, r, t_best, one, group_tmin);
            ngroup.x = n1.x;
            ngroup.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
            if (POP_CULL && !ANY_HIT)
                ngroup.y |= pack_group_tmin(group_tmin);
            if ((hitmask & 0x00FFFFFFu) != 0u)
            {
                // the node handed out triangles: they become the lane's triangle work, older triangle work waits on the stack
                if (tgroup.y != 0u && !stack.push(tgroup)) { PT_COUNT(n_overflows++;) }
                tgroup = make_uint2(n1.y, hitmask & 0x00FFFFFFu);
                // the triangles are tested in later iterations (one per iteration): start fetching them now
                const char* first_triangle = reinterpret_cast<const char*>(bvh.triangles + (n1.y + (uint32_t)__ffs((int)tgroup.y) - 1u));
                prefetch_triangle(first_triangle);
                prefetch_triangle(first_triangle + 128);
            }
            if (ngroup.y > 0x00FFFFFFu)
            {
                // ... and the node the next iteration visits (the nearest hit child in ray-octant order)
                uint32_t next_bit;
                const char* next_node = reinterpret_cast<const char*>(bvh.nodes + group_next_node(ngroup, r.oct_inv, next_bit));
                prefetch_global(next_node);
                prefetch_global(next_node + 64);
            }
        }

        if (tgroup.y != 0u)
        {
            const uint32_t tbit = (uint32_t)__ffs((int)tgroup.y) - 1u;
            tgroup.y &= tgroup.y - 1u;
            const uint32_t tri_index = tgroup.x + tbit;
            const float4* tp = reinterpret_cast<const float4*>(bvh.triangles + tri_index);
            const float4 v0 = load_triangle_f4(tp + 0), v1 = load_triangle_f4(tp + 1), v2 = load_triangle_f4(tp + 2);
            PT_COUNT(n_tris++;)
            const int prim = __float_as_int(v0.w);
            float t, u, v;
            if (prim != last_prim && intersect_triangle_reference(r.o, r.d, f3(v0.x, v0.y, v0.z), f3(v1.x, v1.y, v1.z), f3(v2.x, v2.y, v2.z), t, u, v) &&
                (t < t_best || (!ANY_HIT && t == t_best && prim < hit.prim)))     // equal distance: lowest primitive id, whatever the visiting order
            {
                t_best = t;
                hit.prim = prim; hit.tri_slot = (int)tri_index; hit.t = t; hit.u = u; hit.v = v;
                if (ANY_HIT)
                    finished = true;
            }
        }

        // Out of node work: take the next stack entry. A node group is taken even while triangles are still pending (they are
        // tested one per iteration and the lane would otherwise idle through the node step meanwhile); a triangle group that was
        // set aside is taken only once the current one is done.
        if (!finished && ngroup.y <= 0x00FFFFFFu)
        {
            // (any-hit rays never shrink t_best: nothing to cull)
            while (POP_CULL && !ANY_HIT && stack.size > 0)
            {
                const uint2 e = stack.top();
                if (e.y <= 0x00FFFFFFu || !(group_tmin_of(e) > t_best))
                    break;
                stack.size--;                   // every child of this group starts beyond the closest hit found since it was pushed
                PT_COUNT(n_culled++;)
            }
            if (stack.size == 0)
                finished = tgroup.y == 0u;
            else
            {
                const uint2 e = stack.top();
                if (e.y > 0x00FFFFFFu)
                {
                    ngroup = e;
                    stack.size--;
                }
                else if (tgroup.y == 0u)
                {
                    tgroup = e;
                    stack.size--;
                }
            }
        }

        if (finished)
        {
            trace_queue_store_hit<ANY_HIT>(q, ray_id, hit);
            active = false;
        }
    }
    PT_TRACE_FLUSH_COUNTERS()
}

// PIPELINED flavour: same refill, same per-ray arithmetic, same stack discipline, but every load is issued one iteration before its
// use. ncu on round 1's kernel (profiles/r01_trace_queue_closest_c3_v2.txt): 6.5 warps stalled on `long_scoreboard` per issue, a third
// of the samples on the triangle fetch (the 480 MB triangle array misses L2) and a third on the node fetch, at 49 % issue-active:
// a warp waited for its node, tested it, then waited for its triangle. Here an iteration
//   P1 tests the node whose five LDG.128 were issued in the previous iteration,
//   S1 picks the next node (nearest hit child, else the stack) and issues its loads,
//   P2 tests the triangle loaded in the previous iteration (the node loads of S1 fly meanwhile),
//   S2 picks the next triangle and issues its loads (they fly under the next iteration's P1).
// The visiting order of a ray differs slightly from the non-pipelined flavour (the stack is popped before the triangle of the
// iteration is tested); the closest hit does not depend on the order (strict '<' plus the lowest-id rule on ties), so the hit
// records are identical (tests/test_gpu_parity.py::test_trace_queue_flavours_agree).
template <bool ANY_HIT>
__device__ __forceinline__ void trace_queue_body_pipelined(const Bvh8Device& bvh, const RayQueueView& q, uint2* shared_stack, TraversalCounters* counters,
                                                           const unsigned refill_idle_lanes = kRefillIdleLanes)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = *q.count;
    LaneStack stack;
    stack.bind(shared_stack + threadIdx.x);
    stack.size = 0;
    const uint32_t one = opaque_one_bits(n);

    bool active = false, exhausted = false;
    uint32_t ray_id = 0;
    RaySlabs r;
    r.o = f3(0.0f); r.d = f3(0.0f); r.idx = r.idy = r.idz = 0.0f; r.oct_inv = 0;
    float t_best = 0.0f;
    int last_prim = -1;
    Hit hit;
    hit.prim = -1; hit.tri_slot = -1; hit.t = -1.0f; hit.u = 0.0f; hit.v = 0.0f;
    uint2 ngroup = make_uint2(0u, 0u);
    uint2 tgroup = make_uint2(0u, 0u);
    // loads in flight
    uint4 n0 = make_uint4(0u, 0u, 0u, 0u), n1 = n0, n2 = n0, n3 = n0, n4 = n0;
    float4 v0 = make_float4(0.0f, 0.0f, 0.0f, 0.0f), v1 = v0, v2 = v0;
    bool node_loaded = false, tri_loaded = false;
    uint32_t tri_loaded_index = 0;
    PT_COUNT(unsigned int n_nodes = 0; unsigned int n_tris = 0; unsigned int n_rays = 0; unsigned int n_overflows = 0;)

    while (true)
    {
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
                    trace_queue_load_ray(q, ray_id, ANY_HIT, r, last_prim, t_best);
                    hit.prim = -1; hit.tri_slot = -1; hit.t = -1.0f; hit.u = 0.0f; hit.v = 0.0f;
                    ngroup = make_uint2(0u, 0x80000000u);
                    tgroup = make_uint2(0u, 0u);
                    stack.size = 0;
                    node_loaded = false;
                    tri_loaded = false;
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

        bool finished = false;
        // ---- P1: box tests of the node loaded during the previous iteration ---------------------------------------------
        if (node_loaded)
        {
            node_loaded = false;
            PT_COUNT(n_nodes++;)
            const uint32_t hitmask = node_hitmask(n0, n1, n2, n3, n4, r, t_best, one);
            ngroup.x = n1.x;
            ngroup.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
            if ((hitmask & 0x00FFFFFFu) != 0u)
            {
                if (tgroup.y != 0u && !stack.push(tgroup)) { PT_COUNT(n_overflows++;) }
                tgroup = make_uint2(n1.y, hitmask & 0x00FFFFFFu);
            }
        }
        // ---- S1: next node = nearest hit child of the current group, else what the stack holds --------------------------
        if (ngroup.y <= 0x00FFFFFFu && stack.size > 0)
        {
            uint2 e = stack.top();
            if (e.y <= 0x00FFFFFFu && tgroup.y == 0u)
            {
                // a triangle group that was set aside: taken once the current one is done; the node group below it follows
                tgroup = e;
                stack.size--;
                if (stack.size > 0)
                    e = stack.top();
            }
            if (stack.size > 0 && e.y > 0x00FFFFFFu)
            {
                ngroup = e;
                stack.size--;
            }
        }
        if (ngroup.y > 0x00FFFFFFu)
        {
            uint32_t bit;
            const uint32_t node_index = group_next_node(ngroup, r.oct_inv, bit);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y > 0x00FFFFFFu && !stack.push(ngroup)) { PT_COUNT(n_overflows++;) }
            ngroup.y = 0u;                      // the rest of the group waits on the stack; the loaded node's children replace it in P1
            const uint4* np = reinterpret_cast<const uint4*>(bvh.nodes + node_index);
            n0 = __ldg(np + 0); n1 = __ldg(np + 1); n2 = __ldg(np + 2); n3 = __ldg(np + 3); n4 = __ldg(np + 4);
            node_loaded = true;
        }
        // ---- P2: the triangle loaded during the previous iteration ------------------------------------------------------
        if (tri_loaded)
        {
            tri_loaded = false;
            PT_COUNT(n_tris++;)
            const int prim = __float_as_int(v0.w);
            float t, u, v;
            if (prim != last_prim && intersect_triangle_reference(r.o, r.d, f3(v0.x, v0.y, v0.z), f3(v1.x, v1.y, v1.z), f3(v2.x, v2.y, v2.z), t, u, v) &&
                (t < t_best || (!ANY_HIT && t == t_best && prim < hit.prim)))
            {
                t_best = t;
                hit.prim = prim; hit.tri_slot = (int)tri_loaded_index; hit.t = t; hit.u = u; hit.v = v;
                if (ANY_HIT)
                    finished = true;
            }
        }
        // ---- S2: next triangle ------------------------------------------------------------------------------------------
        if (!finished && tgroup.y != 0u)
        {
            const uint32_t tbit = (uint32_t)__ffs((int)tgroup.y) - 1u;
            tgroup.y &= tgroup.y - 1u;
            tri_loaded_index = tgroup.x + tbit;
            const float4* tp = reinterpret_cast<const float4*>(bvh.triangles + tri_loaded_index);
            v0 = load_triangle_f4(tp + 0); v1 = load_triangle_f4(tp + 1); v2 = load_triangle_f4(tp + 2);
            tri_loaded = true;
        }
        if (!finished)
            finished = !node_loaded && !tri_loaded && stack.size == 0;   // (ngroup / tgroup are empty then: S1 / S2 would have loaded from them)

        if (finished)
        {
            trace_queue_store_hit<ANY_HIT>(q, ray_id, hit);
            active = false;
        }
    }
    PT_TRACE_FLUSH_COUNTERS()
}
} // namespace pt
