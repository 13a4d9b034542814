// One FullPathTracer permutation per translation unit: compile with -DPT_LSS= -DPT_ESS= -DPT_LATER= -DPT_GGX=
// Holds the megakernel and the wavefront kernels (wavefront.cuh) of the permutation plus their launcher.
#include <cstdlib>
#include "wavefront_mis.cuh"
#if PT_LSS == 5
#include "restir_di_kernels.cuh"
#endif

namespace pt
{
// Queue kernels: the queue length is only known on the device. Every warp of a grid sized to the SMs pulls the next 32
// queue entries with one atomic on a per-launch cursor (dynamic balance, and no mostly-empty 32k-block grids:
// those cost ~1.5 ns per empty block, 50 us per launch at 1080p). KID = cursor of this kernel within its bounce.
#define PT_FOR_EACH_QUEUED(I, COUNT, KID)                                                                           \
    const uint32_t n_queued_ = (COUNT);                                                                             \
    uint32_t* const cursor_ = wf.cursors + (uint32_t)bounce * kCursorsPerBounce + (KID);                            \
    for (uint32_t I = wf_next_entry(cursor_); __any_sync(0xFFFFFFFFu, I < n_queued_); I = wf_next_entry(cursor_))   \
        if (I < n_queued_)

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_FullPathTracer)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    full_path_tracer_pixel<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, x, y);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_bounce0)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    wf_bounce0_pixel<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, tile, wf, x, y);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_bounce)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce], 0)
        wf_bounce_path<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, tile, wf, bounce, i);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_gen)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    wf_gen_pixel<PT_ESS>(rd, res, tile, wf, x, y);
}

__global__ void __launch_bounds__(128) PT_VARIANT_NAME(k_wf_extend)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce], 0)
        wf_extend_path<PT_ESS>(rd, res, tile, wf, bounce, i);
}

__global__ void __launch_bounds__(kSortBuckets) PT_VARIANT_NAME(k_wf_scan)(const __grid_constant__ WavefrontState wf)
{
    __shared__ uint32_t smem[kSortBuckets];
    wf_scan_buckets(wf, smem);
}

__global__ void __launch_bounds__(256) PT_VARIANT_NAME(k_wf_scatter)(const __grid_constant__ WavefrontState wf, int bounce)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce + 1], 1)
        wf_scatter_path(wf, i);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_shade)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce + 1], 2)
        wf_shade_path<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, tile, wf, bounce, i);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_direct)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce + 1], 2)
        wf_direct_path<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, tile, wf, bounce, i);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_bsdf)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce + 1], 7)
        wf_bsdf_path<PT_GGX>(rd, res, tile, wf, bounce, i);
}

#ifndef PT_PAIR_PHASES_DEFAULT
#define PT_PAIR_PHASES_DEFAULT false
#endif
// The phase kernels of the phased pipelines (MIS: P1..P5; RIS: R1, R2, P3..P5) live in their own translation unit
// (b200pt_phase_variant.cu), compiled with the BSDF entry points inlined; this one launches them through launch_phase.
enum PhaseKernel { PHASE_NEE_LIGHT_EVAL, PHASE_NEE_BSDF_SAMPLE, PHASE_ENV_LIGHT_EVAL, PHASE_ENV_BSDF_SAMPLE, PHASE_BSDF_RESOLVED, PHASE_NEE_PAIR, PHASE_ENV_PAIR, PHASE_RIS_CANDIDATES, PHASE_RIS_SELECT };
cudaError_t PT_VARIANT_NAME(launch_phase)(int which, unsigned blocks, cudaStream_t stream, const HIPRTRenderData* rd, int2 res, Tile tile, const WavefrontState* wf, int bounce,
                                          bool lights_sampled, bool envmap_sampled);
// SURFACE waits on two dependent scattered fetches per path (triangle indices -> vertex normals / texcoords, material index -> material): ncu on C3
// shows 22 % issue-active with 12 warps per issue on a long scoreboard at 96 registers = 20 warps per SM (profiles/r02_shade_c3_final.txt).
// More resident warps by capping the registers did not pay: at 7 blocks per SM (72 registers) the stage went 24.0 -> 27.1 ms (gpurun call 19).
#ifndef PT_SURFACE_BLOCKS_PER_SM
#define PT_SURFACE_BLOCKS_PER_SM 1
#endif
__global__ void __launch_bounds__(128, PT_SURFACE_BLOCKS_PER_SM) PT_VARIANT_NAME(k_wf_surface)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce], 8)
        wf_surface_path<PT_ESS>(rd, res, tile, wf, bounce, i);
}

#if PT_LSS == 3 || PT_LSS == 4
// ---- sample batches (wavefront.cuh): camera rays of every pass in flight, camera hits, accumulation in pass order --------------------
__global__ void __launch_bounds__(128) PT_VARIANT_NAME(k_wf_camera)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf,
                                                                   const __grid_constant__ SampleBatch batch)
{
    const uint32_t total = wf.tile_pixels * batch.count;
    if (blockIdx.x == 0 && threadIdx.x == 0)
        wf.counts[0] = total;             // length of the primary-ray queue (identity), read by the ray-queue kernel and CAMERA_SURFACE
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < total; slot += gridDim.x * blockDim.x)
        wf_camera_slot(rd, res, tile, wf, batch, slot);
}

__global__ void __launch_bounds__(128, PT_SURFACE_BLOCKS_PER_SM) PT_VARIANT_NAME(k_wf_camera_surface)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf,
                                                                           const __grid_constant__ SampleBatch batch)
{
    const int bounce = 0;
    PT_FOR_EACH_QUEUED(i, wf.counts[0], 8)
        wf_camera_surface_path<PT_ESS>(rd, res, tile, wf, batch, i);
}

__global__ void __launch_bounds__(128) PT_VARIANT_NAME(k_wf_accumulate)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf,
                                                                       const __grid_constant__ SampleBatch batch)
{
    for (uint32_t local = blockIdx.x * blockDim.x + threadIdx.x; local < wf.tile_pixels; local += gridDim.x * blockDim.x)
        wf_accumulate_pixel(rd, res, tile, wf, batch, local);
}

#endif

// batch == nullptr: FullPathTracer alone, on the G-buffer a CameraRays launch left (the reference's two launch points).
// batch != nullptr: b200pt_launch_PathTraceSamples, `batch->count` passes of CameraRays + FullPathTracer as one wavefront (phased MIS only).
cudaError_t PT_VARIANT_NAME(launch_FullPathTracer)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const WavefrontState* wf, int mode, WavefrontProfile* profile,
                                                   const SampleBatch* batch)
{
    auto stage = [&](int s) {
        if (profile && profile->count < profile->capacity)
        {
            cudaEventRecord(profile->events[profile->count], stream);
            profile->stages[profile->count++] = s;
        }
    };
    dim3 grid((tile.x1 - tile.x0 + 7) / 8, (tile.y1 - tile.y0 + 7) / 8, 1);
    if (mode == WF_MEGAKERNEL || wf == nullptr)
    {
        PT_VARIANT_NAME(k_FullPathTracer)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
        return cudaGetLastError();
    }

    int nb_bounces = rd->render_settings.nb_bounces;
    if (rd->render_settings.wants_render_low_resolution && rd->render_settings.allow_render_low_resolution && rd->render_settings.accumulate)
        nb_bounces = nb_bounces < 3 ? nb_bounces : 3;
    const uint32_t paths = (uint32_t)(tile.x1 - tile.x0) * (uint32_t)(tile.y1 - tile.y0) * (batch ? batch->count : 1u);
    if (paths > wf->capacity || (uint32_t)nb_bounces > wf->max_bounces)
        return cudaErrorInvalidValue;
    if (batch && (mode != WF_PHASED || (PT_LSS != 3 && PT_LSS != 4) || wf->fused_camera == 0u || wf->samples != batch->count))
        return cudaErrorInvalidValue;

    // Queue lengths are only known on the device: queue kernels run grid-stride loops on a grid capped at a few blocks per
    // SM (an uncapped grid of mostly empty blocks costs ~1.5 ns per block: 50 us per launch at 1080p)
    // B200PT_QUEUE_BLOCKS_PER_SM: blocks of 64 threads per SM for all lanes together; default 12 per lane = what the phases' register cap admits
    // (16 for two lanes before: C3 1034 -> 1065 Mrays/s at 24, 1056 -> 1073 with three lanes at 36; profiles/r02_knob_sweep_lanes_queue_grid.txt)
    static const unsigned blocks_per_sm = [] { const char* v = getenv("B200PT_QUEUE_BLOCKS_PER_SM"); int n = v ? atoi(v) : 0; return n > 0 ? (unsigned)n : 0u; }();
    // Path lanes: when the host runs several band launches on concurrent streams (trace of one band under the shade phases of
    // another), each launch sizes its grids for its share of an SM so that blocks of both are resident together
    static const unsigned share_override = [] { const char* v = getenv("B200PT_GRID_SHARE"); int n = v ? atoi(v) : 0; return n > 0 ? (unsigned)n : 0u; }();
    const unsigned share = share_override ? share_override : (wf->grid_share > 0u ? wf->grid_share : 1u);
    const unsigned cap = wf->sm_count * (blocks_per_sm ? (blocks_per_sm + share - 1u) / share : 12u);
    auto blocks = [&](unsigned block) { unsigned g = (paths + block - 1) / block, c = cap * 64u / block; return g < c ? g : c; };
    cudaError_t e = cudaMemsetAsync(wf->counts, 0, sizeof(uint32_t) * (kCountsPerBounce + kCursorsPerBounce) * (wf->max_bounces + 2), stream);
    if (e != cudaSuccess)
        return e;
    if (mode == WF_FUSED)
    {
        PT_VARIANT_NAME(k_wf_bounce0)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *wf);
        for (int bounce = 1; bounce <= nb_bounces; bounce++)
            PT_VARIANT_NAME(k_wf_bounce)<<<blocks(64), 64, 0, stream>>>(*rd, res, tile, *wf, bounce);
        return cudaGetLastError();
    }
    // Which strategy the phases run: MIS light + BSDF (LSS 3), or RIS (LSS 4; LSS 5 with RIS on the later bounces, from bounce 1 on: bounce 0
    // shades from the ReSTIR DI reservoirs in the unphased kernels). Alpha testing would make shadow rays draw random numbers
    // (FilterFunction.h:42): then the phases cannot be reordered. RIS runs phased with one BSDF candidate (the default) only.
    constexpr bool kPhasesMis = PT_LSS == 3;
    constexpr bool kPhasesRis = PT_LSS == 4 || (PT_LSS == 5 && PT_LATER == 3);
    const HIPRTRenderSettings& rs = rd->render_settings;
    const bool low_resolution = rs.wants_render_low_resolution && rs.allow_render_low_resolution && rs.accumulate;
    static const bool ris_phases_enabled = [] { const char* v = getenv("B200PT_RIS_PHASES"); return v ? atoi(v) != 0 : true; }();
    const bool phased = mode == WF_PHASED && rs.number_of_light_samples == 1 && !rs.do_alpha_testing &&
                        (kPhasesMis || (kPhasesRis && ris_phases_enabled && !low_resolution && rs.ris_settings.number_of_bsdf_candidates == 1));
    if (batch && !(phased && (kPhasesMis || PT_LSS == 4)))
        return cudaErrorInvalidValue;
#ifdef PT_COUNTERS
    TraversalCounters* const counters = reinterpret_cast<TraversalCounters*>(rd->hiprt_function_table);
#else
    TraversalCounters* const counters = nullptr;
#endif
    const unsigned trace_blocks = wf->sm_count * (((unsigned)kTraceGridBlocksPerSM + share - 1u) / share);   // persistent: the resident blocks of 128 threads per SM (trace_queue.cuh), shared between lanes
    (void)counters; (void)trace_blocks;
    // Ray binning (launch_bin_rays): cells per axis = 2^bits, 0 = off. B200PT_RAY_BIN_BITS for both kinds of queue, or
    // B200PT_RAY_BIN_BITS_EXT / _NEE separately (sweeps: profiles/r02_ray_binning.txt)
#ifndef PT_RAY_BIN_BITS_DEFAULT
#define PT_RAY_BIN_BITS_DEFAULT 0
#endif
    static const int bin_bits_both = [] { const char* v = getenv("B200PT_RAY_BIN_BITS"); return v ? atoi(v) : PT_RAY_BIN_BITS_DEFAULT; }();
    static const int bin_bits_ext_env = [] { const char* v = getenv("B200PT_RAY_BIN_BITS_EXT"); return v ? atoi(v) : bin_bits_both; }();
    static const int bin_bits_nee_env = [] { const char* v = getenv("B200PT_RAY_BIN_BITS_NEE"); return v ? atoi(v) : bin_bits_both; }();
    const int bin_bits_ext = g_ray_bin_bits >= 0 ? g_ray_bin_bits : bin_bits_ext_env;
    const int bin_bits_nee = g_ray_bin_bits >= 0 ? g_ray_bin_bits : bin_bits_nee_env;
    (void)bin_bits_ext; (void)bin_bits_nee;
    for (int bounce = 0; bounce <= nb_bounces; bounce++)
    {
        uint32_t* const counts = wf->counts + kCountsPerBounce * bounce;
        uint32_t* const cursors = wf->cursors + kCursorsPerBounce * bounce;
        (void)counts; (void)cursors;
        stage(STAGE_SURFACE);
#if PT_LSS == 3 || PT_LSS == 4
        if (bounce == 0 && batch)
        {
            PT_VARIANT_NAME(k_wf_camera)<<<blocks(128), 128, 0, stream>>>(*rd, res, tile, *wf, *batch);
            stage(STAGE_TRACE);
            RayQueueView q = {wf->ray_queue[1], counts + 0, cursors + 0, wf->origin_rng, wf->dir_prim, wf->hit_tuvp, wf->hit_tri_slot, nullptr, wf->stride_ray - 1u, wf->stride_hitrec - 1u, wf->stride_hitrec_words - 1u};
            e = PT_SUFFIXED(launch_trace_queue)(rd->geom, q, false, trace_blocks, counters, stream);
            if (e != cudaSuccess)
                return e;
            stage(STAGE_SURFACE);
            PT_VARIANT_NAME(k_wf_camera_surface)<<<blocks(128), 128, 0, stream>>>(*rd, res, tile, *wf, *batch);
        }
        else
#endif
        if (bounce == 0)
            PT_VARIANT_NAME(k_wf_gen)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *wf);
        else if (phased)
        {
            stage(bin_bits_ext > 0 ? STAGE_SORT : STAGE_TRACE);
            RayQueueView q = {wf->ray_queue[(bounce - 1) & 1], counts + 0, cursors + 0, wf->origin_rng, wf->dir_prim, wf->hit_tuvp, wf->hit_tri_slot, nullptr, wf->stride_ray - 1u, wf->stride_hitrec - 1u, wf->stride_hitrec_words - 1u};
            // rays regrouped by origin cell and direction octant (launch_bin_rays); SURFACE then walks the same order: the rays of a
            // cell hit neighbouring triangles, so its vertex / normal gathers share cache lines too
            e = PT_SUFFIXED(launch_bin_rays)(rd->geom, q, wf->bin_hist, wf->bin_keys, wf->binned_ext, bin_bits_ext, wf->sm_count, stream);
            if (e != cudaSuccess)
                return e;
            stage(STAGE_TRACE);
            e = PT_SUFFIXED(launch_trace_queue)(rd->geom, q, false, trace_blocks, counters, stream);
            if (e != cudaSuccess)
                return e;
            stage(STAGE_SURFACE);
            WavefrontState wf_binned = *wf;
            wf_binned.ray_queue[(bounce - 1) & 1] = const_cast<uint32_t*>(q.ids);
            PT_VARIANT_NAME(k_wf_surface)<<<blocks(128), 128, 0, stream>>>(*rd, res, tile, wf_binned, bounce);
        }
        else
            PT_VARIANT_NAME(k_wf_extend)<<<blocks(128), 128, 0, stream>>>(*rd, res, tile, *wf, bounce);
        stage(STAGE_SORT);
        PT_VARIANT_NAME(k_wf_scan)<<<1, kSortBuckets, 0, stream>>>(*wf);
        PT_VARIANT_NAME(k_wf_scatter)<<<blocks(256), 256, 0, stream>>>(*wf, bounce);
        // the scatter cursors become the next histogram: zero them again
        e = cudaMemsetAsync(wf->bucket_count, 0, sizeof(uint32_t) * kSortBuckets, stream);
        if (e != cudaSuccess)
            return e;
        stage(STAGE_SHADE);
        if (phased && !(PT_LSS == 5 && bounce == 0))
        {
            const unsigned g64 = blocks(64);
            const bool lights = rd->buffers.emissive_triangles_count > 0;
            const bool envmap = PT_ESS != 0 && rd->world_settings.ambient_light_type == AMBIENT_ENVMAP;
            RayQueueView qa = {wf->any_queue, counts + 2, cursors + 6, wf->nee_ray_o, wf->nee_ray_d, nullptr, nullptr, wf->nee_ray_o, wf->stride_nee - 1u, 0u, 0u};
            RayQueueView qc = {wf->closest_queue, counts + 3, cursors + 9, wf->nee_ray_o, wf->nee_ray_d, wf->hit_tuvp, wf->hit_tri_slot, nullptr, wf->stride_nee - 1u,
                               wf->stride_hitrec - 1u, wf->stride_hitrec_words - 1u};
            if (kPhasesRis)
            {
                // R1 -> closest trace of the BSDF candidates' rays -> R2 -> envmap phases -> any-hit trace of the winners' shadow rays (+ the
                // envmap rays) -> resolve
                if (lights)
                {
                    PT_VARIANT_NAME(launch_phase)(PHASE_RIS_CANDIDATES, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
                    stage(STAGE_TRACE);
                    e = PT_SUFFIXED(launch_trace_queue)(rd->geom, qc, false, trace_blocks, counters, stream);
                    if (e != cudaSuccess)
                        return e;
                    stage(STAGE_SHADE);
                    PT_VARIANT_NAME(launch_phase)(PHASE_RIS_SELECT, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
                }
                if (envmap)
                {
                    PT_VARIANT_NAME(launch_phase)(PHASE_ENV_LIGHT_EVAL, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
                    PT_VARIANT_NAME(launch_phase)(PHASE_ENV_BSDF_SAMPLE, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
                }
                if (lights || envmap)
                {
                    stage(STAGE_TRACE);
                    e = PT_SUFFIXED(launch_trace_queue)(rd->geom, qa, true, trace_blocks, counters, stream);
                    if (e != cudaSuccess)
                        return e;
                    stage(STAGE_SHADE);
                }
                PT_VARIANT_NAME(launch_phase)(PHASE_BSDF_RESOLVED, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
                continue;
            }
            // B200PT_PAIR_PHASES=1: eval + sample of a MIS pair in one kernel on one loaded vertex (one ~400-byte vertex load
            // less per pair, twice the code per kernel); same arithmetic and RNG order, same bits
            static const bool pair_phases = [] { const char* v = getenv("B200PT_PAIR_PHASES"); return v ? atoi(v) != 0 : PT_PAIR_PHASES_DEFAULT; }();
            if (lights && pair_phases)
                PT_VARIANT_NAME(launch_phase)(PHASE_NEE_PAIR, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
            else if (lights)
            {
                PT_VARIANT_NAME(launch_phase)(PHASE_NEE_LIGHT_EVAL, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
                PT_VARIANT_NAME(launch_phase)(PHASE_NEE_BSDF_SAMPLE, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
            }
            if (envmap && pair_phases)
                PT_VARIANT_NAME(launch_phase)(PHASE_ENV_PAIR, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
            else if (envmap)
            {
                PT_VARIANT_NAME(launch_phase)(PHASE_ENV_LIGHT_EVAL, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
                PT_VARIANT_NAME(launch_phase)(PHASE_ENV_BSDF_SAMPLE, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
            }
            stage(bin_bits_nee > 0 ? STAGE_SORT : STAGE_TRACE);
            static const bool pair_trace = [] { const char* v = getenv("B200PT_PAIR_TRACE"); return v ? atoi(v) != 0 : true; }();
            if (lights || envmap)
            {
                e = PT_SUFFIXED(launch_bin_rays)(rd->geom, qa, wf->bin_hist, wf->bin_keys, wf->binned_any, bin_bits_nee, wf->sm_count, stream);
                if (e != cudaSuccess)
                    return e;
            }
            if (lights)
            {
                e = PT_SUFFIXED(launch_bin_rays)(rd->geom, qc, wf->bin_hist, wf->bin_keys, wf->binned_ext, bin_bits_nee, wf->sm_count, stream);
                if (e != cudaSuccess)
                    return e;
            }
            stage(STAGE_TRACE);
            if (lights && pair_trace)
            {
                // shadow rays (any hit) and light rays (closest hit) of the bounce in ONE launch: see k_trace_queue_pair
                e = PT_SUFFIXED(launch_trace_queue_pair)(rd->geom, qa, qc, trace_blocks, counters, stream);
                if (e != cudaSuccess)
                    return e;
            }
            else
            {
                if (lights || envmap)
                {
                    e = PT_SUFFIXED(launch_trace_queue)(rd->geom, qa, true, trace_blocks, counters, stream);
                    if (e != cudaSuccess)
                        return e;
                }
                if (lights)
                {
                    e = PT_SUFFIXED(launch_trace_queue)(rd->geom, qc, false, trace_blocks, counters, stream);
                    if (e != cudaSuccess)
                        return e;
                }
            }
            stage(STAGE_SHADE);
            PT_VARIANT_NAME(launch_phase)(PHASE_BSDF_RESOLVED, g64, stream, rd, res, tile, wf, bounce, lights, envmap);
        }
        else if (mode == WF_PHASED)
        {
            PT_VARIANT_NAME(k_wf_direct)<<<blocks(64), 64, 0, stream>>>(*rd, res, tile, *wf, bounce);
            PT_VARIANT_NAME(k_wf_bsdf)<<<blocks(64), 64, 0, stream>>>(*rd, res, tile, *wf, bounce);
        }
        else
            PT_VARIANT_NAME(k_wf_shade)<<<blocks(64), 64, 0, stream>>>(*rd, res, tile, *wf, bounce);
    }
#if PT_LSS == 3 || PT_LSS == 4
    if (batch)
    {
        stage(STAGE_SHADE);
        PT_VARIANT_NAME(k_wf_accumulate)<<<blocks(128), 128, 0, stream>>>(*rd, res, tile, *wf, *batch);
    }
#endif
    stage(STAGE_END);
    return cudaGetLastError();
}


#if PT_LSS == 5
// ---- ReSTIR DI passes (restir_di_kernels.cuh); compiled with the LSS_RESTIR_DI permutations only --------------------------------
__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_ReSTIR_DI_LightsPresampling)(const __grid_constant__ LightPresamplingParameters p)
{
    restir_presample_light<PT_ESS>(p, blockIdx.x * blockDim.x + threadIdx.x);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_ReSTIR_DI_InitialCandidates)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    restir_initial_candidates_pixel<PT_GGX>(rd, res, x, y);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_ReSTIR_DI_SpatiotemporalReuse)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    restir_spatiotemporal_pixel(rd, res, x, y);
}

// Phased passes (restir_di_kernels.cuh): COLLECT / RESOLVE kernels around launches of the ray-queue traversal kernel
// Registers of the COLLECT kernels are left uncapped (128-152): capped at 96 (10 blocks per SM, as the shade phases of b200pt_phase_variant.cu
// are) the fused pass's COLLECT spilled 160 B more and C4 went 645 -> 639 Mrays/s (B200, gpurun call 14)
#ifndef PT_RESTIR_COLLECT_BLOCKS_PER_SM
#define PT_RESTIR_COLLECT_BLOCKS_PER_SM 1
#endif
#define PT_RESTIR_PHASE_KERNEL(NAME, CALL, MIN_BLOCKS)                                                                                           \
    __global__ void __launch_bounds__(64, MIN_BLOCKS) PT_VARIANT_NAME(NAME)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ RestirScratch scratch) \
    {                                                                                                                                            \
        const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;                                                                      \
        const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;                                                                      \
        if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)                                                                                    \
            return;                                                                                                                              \
        RestirSites sites;                                                                                                                       \
        sites.scratch = &scratch;                                                                                                                \
        sites.pixel = (x - (uint32_t)tile.x0) + (y - (uint32_t)tile.y0) * (uint32_t)(tile.x1 - tile.x0);                                         \
        CALL(rd, res, x, y, sites);                                                                                                              \
    }
#define PT_IC_COLLECT restir_initial_candidates_pixel<PT_GGX, RESTIR_COLLECT>
#define PT_IC_RESOLVE restir_initial_candidates_pixel<PT_GGX, RESTIR_RESOLVE>
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_InitialCandidates_collect, PT_IC_COLLECT, PT_RESTIR_COLLECT_BLOCKS_PER_SM)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_InitialCandidates_resolve, PT_IC_RESOLVE, 1)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_InitialCandidates_visibility, restir_initial_candidates_visibility, 1)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_SpatiotemporalReuse_collect, restir_spatiotemporal_pixel<RESTIR_COLLECT>, PT_RESTIR_COLLECT_BLOCKS_PER_SM)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_SpatiotemporalReuse_resolve, restir_spatiotemporal_pixel<RESTIR_RESOLVE>, 1)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_TemporalReuse_collect, restir_temporal_pixel<RESTIR_COLLECT>, PT_RESTIR_COLLECT_BLOCKS_PER_SM)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_TemporalReuse_resolve, restir_temporal_pixel<RESTIR_RESOLVE>, 1)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_SpatialReuse_collect, restir_spatial_pixel<RESTIR_COLLECT>, PT_RESTIR_COLLECT_BLOCKS_PER_SM)
PT_RESTIR_PHASE_KERNEL(k_ReSTIR_DI_SpatialReuse_resolve, restir_spatial_pixel<RESTIR_RESOLVE>, 1)

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_ReSTIR_DI_TemporalReuse)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    restir_temporal_pixel(rd, res, x, y);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_ReSTIR_DI_SpatialReuse)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    restir_spatial_pixel(rd, res, x, y);
}

cudaError_t PT_VARIANT_NAME(launch_ReSTIR_DI_LightsPresampling)(const LightPresamplingParameters* p, cudaStream_t stream)
{
    const long long n = (long long)p->subset_size * p->number_of_subsets;
    if (n <= 0)
        return cudaSuccess;
    PT_VARIANT_NAME(k_ReSTIR_DI_LightsPresampling)<<<(unsigned)((n + 63) / 64), 64, 0, stream>>>(*p);
    return cudaGetLastError();
}

// Whether a ReSTIR DI pass runs phased (its rays through the ray-queue kernel): scratch large enough for the tile, no alpha testing
// (a shadow ray would draw random numbers: FilterFunction.h:42), full-resolution rendering; B200PT_RESTIR_PHASES=0 keeps the one-kernel passes
static bool restir_phased(const HIPRTRenderData* rd, Tile tile, const RestirScratch* scratch)
{
    static const bool enabled = [] { const char* v = getenv("B200PT_RESTIR_PHASES"); return v ? atoi(v) != 0 : true; }();
    const HIPRTRenderSettings& rs = rd->render_settings;
    const bool low_resolution = rs.wants_render_low_resolution && rs.allow_render_low_resolution && rs.accumulate;
    return enabled && scratch != nullptr && scratch->capacity >= (uint32_t)(tile.x1 - tile.x0) * (uint32_t)(tile.y1 - tile.y0) && !rs.do_alpha_testing && !low_resolution;
}

static cudaError_t restir_trace(const HIPRTRenderData* rd, const RestirScratch* scratch, bool any_hit, cudaStream_t stream)
{
#ifdef PT_COUNTERS
    TraversalCounters* const counters = reinterpret_cast<TraversalCounters*>(rd->hiprt_function_table);
#else
    TraversalCounters* const counters = nullptr;
#endif
    uint32_t* const counts = scratch->counts + (any_hit ? 2 : 0);
    RayQueueView q{scratch->queue, counts, counts + 1, scratch->ray_o, scratch->ray_d, scratch->hit_tuvp, scratch->hit_tri_slot, scratch->ray_o};
    return PT_SUFFIXED(launch_trace_queue)(rd->geom, q, any_hit, scratch->sm_count * (unsigned)kTraceGridBlocksPerSM, counters, stream);
}

cudaError_t PT_VARIANT_NAME(launch_ReSTIR_DI_InitialCandidates)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
    dim3 grid((tile.x1 - tile.x0 + 7) / 8, (tile.y1 - tile.y0 + 7) / 8, 1);
    if (!restir_phased(rd, tile, scratch) || rd->render_settings.restir_di_settings.initial_candidates.number_of_initial_bsdf_candidates > 1)
    {
        PT_VARIANT_NAME(k_ReSTIR_DI_InitialCandidates)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
        return cudaGetLastError();
    }
    cudaError_t e = cudaMemsetAsync(scratch->counts, 0, 4 * sizeof(uint32_t), stream);
    if (e != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_InitialCandidates_collect)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    if ((e = restir_trace(rd, scratch, false, stream)) != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_InitialCandidates_resolve)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    if ((e = restir_trace(rd, scratch, true, stream)) != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_InitialCandidates_visibility)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    return cudaGetLastError();
}

cudaError_t PT_VARIANT_NAME(launch_ReSTIR_DI_SpatiotemporalReuse)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
    dim3 grid((tile.x1 - tile.x0 + 7) / 8, (tile.y1 - tile.y0 + 7) / 8, 1);
    const SpatialPassSettings& sp = rd->render_settings.restir_di_settings.spatial_pass;
    const int neighbors = sp.do_disocclusion_reuse_boost && sp.disocclusion_reuse_count > sp.reuse_neighbor_count ? sp.disocclusion_reuse_count : sp.reuse_neighbor_count;
    if (!restir_phased(rd, tile, scratch) || neighbors > kRestirMaxNeighbors)
    {
        PT_VARIANT_NAME(k_ReSTIR_DI_SpatiotemporalReuse)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
        return cudaGetLastError();
    }
    cudaError_t e = cudaMemsetAsync(scratch->counts, 0, 4 * sizeof(uint32_t), stream);
    if (e != cudaSuccess)
        return e;
    // (COLLECT cut once more -- neighbour search per pixel, then one thread per (pixel, neighbour) for the target functions, 80 registers, BSDF
    // inlined -- was built and measured: search 0.24 ms + sites 3.32 ms against 3.09 ms for this kernel, bit-identical, C4 unchanged (gpurun call
    // 20): the kernel is bound by the instructions of its BSDF evaluations at ~18 lanes per instruction, not by their latency. Removed.)
    PT_VARIANT_NAME(k_ReSTIR_DI_SpatiotemporalReuse_collect)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    if ((e = restir_trace(rd, scratch, true, stream)) != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_SpatiotemporalReuse_resolve)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    return cudaGetLastError();
}

cudaError_t PT_VARIANT_NAME(launch_ReSTIR_DI_TemporalReuse)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
    dim3 grid((tile.x1 - tile.x0 + 7) / 8, (tile.y1 - tile.y0 + 7) / 8, 1);
    if (!restir_phased(rd, tile, scratch))
    {
        PT_VARIANT_NAME(k_ReSTIR_DI_TemporalReuse)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
        return cudaGetLastError();
    }
    cudaError_t e = cudaMemsetAsync(scratch->counts, 0, 4 * sizeof(uint32_t), stream);
    if (e != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_TemporalReuse_collect)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    if ((e = restir_trace(rd, scratch, true, stream)) != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_TemporalReuse_resolve)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    return cudaGetLastError();
}

cudaError_t PT_VARIANT_NAME(launch_ReSTIR_DI_SpatialReuse)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch)
{
    dim3 grid((tile.x1 - tile.x0 + 7) / 8, (tile.y1 - tile.y0 + 7) / 8, 1);
    const SpatialPassSettings& sp = rd->render_settings.restir_di_settings.spatial_pass;
    const int neighbors = sp.do_disocclusion_reuse_boost && sp.disocclusion_reuse_count > sp.reuse_neighbor_count ? sp.disocclusion_reuse_count : sp.reuse_neighbor_count;
    if (!restir_phased(rd, tile, scratch) || neighbors > kRestirMaxNeighbors)
    {
        PT_VARIANT_NAME(k_ReSTIR_DI_SpatialReuse)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
        return cudaGetLastError();
    }
    cudaError_t e = cudaMemsetAsync(scratch->counts, 0, 4 * sizeof(uint32_t), stream);
    if (e != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_SpatialReuse_collect)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    if ((e = restir_trace(rd, scratch, true, stream)) != cudaSuccess)
        return e;
    PT_VARIANT_NAME(k_ReSTIR_DI_SpatialReuse_resolve)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *scratch);
    return cudaGetLastError();
}
#endif
} // namespace pt
