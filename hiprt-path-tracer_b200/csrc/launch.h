// Host-callable launchers of the two kernel builds (timing build: suffix _t, counting build: suffix _c).
#pragma once

#include "../../include/b200pt.h"
#include "bvh8.h"

#include <cuda_runtime.h>

namespace pt
{
struct VariantOptions { int lss, ess, later, ggx; };
struct Tile { int x0, y0, x1, y1; };
// rays / nodes / triangles: every BVH query of the launch (per-thread traversals and the ray-queue kernel); queue_*: the ray-queue kernel's
// share (the kernel the node-fetch roofline is about); overflows: traversal-stack pushes that found the stack full (must stay 0).
struct TraversalCounters { unsigned long long rays, nodes, triangles, overflows, queue_rays, queue_nodes, queue_triangles; };

// How b200pt_launch_FullPathTracer runs the path (wavefront.cuh). Selected per context (b200pt_set_path_mode).
constexpr int kSortBuckets = 128;
// Loop flavour of the ray-queue traversal kernel (b200pt_set_trace_flavour): -1 = B200PT_TRACE_PIPELINE / the built-in default,
// 0 = classic, 5 / 6 / 7 = pipelined at that many resident blocks per SM. Process-wide; the hits do not depend on it.
extern int g_trace_flavour;
// Ray binning before traversal (b200pt_set_ray_bin_bits): -1 = B200PT_RAY_BIN_BITS / the built-in default, 0 = off, 1..6 = cells per axis 2^bits
extern int g_ray_bin_bits;
constexpr int kTraceGridBlocksPerSM = 7;   // resident blocks per SM of the ray-queue traversal kernel (= kTraceBlocksPerSM, trace_queue.cuh): its persistent grid
constexpr int kCursorsPerBounce = 12;
constexpr int kCountsPerBounce = 4;    // rays to extend, hits to shade, NEE any-hit rays, NEE closest-hit rays
enum WavefrontMode { WF_MEGAKERNEL = 0, WF_FUSED = 1, WF_SPLIT = 2, WF_PHASED = 3 };

// Scratch of the wavefront path tracer, owned by the context, sized for the pixels of the largest tile launched so far.
struct WavefrontState
{
    float4* origin_rng;      // ray origin xyz | Xorshift32 state
    float4* dir_prim;        // ray direction xyz | primitive the ray leaves from
    float4* throughput;
    float4* color;
    float4* hit_point;       // SPLIT only: intersection point | hit primitive
    float4* hit_ns;          // SPLIT only: shading normal
    float4* hit_ng;          // SPLIT only: geometric normal
    uint32_t* volume;        // RayVolumeState, 9 words per path
    void* materials;         // SPLIT only: 256-byte SimplifiedRendererMaterial per path
    uint32_t* ray_queue[2];  // slot ids of the paths to extend at the next bounce (ping-pong)
    uint32_t* shade_queue;   // SPLIT only: slot ids of the paths that hit something at this bounce (arrival order)
    uint32_t* sorted_queue;  // SPLIT only: the same ids, grouped by material signature (coherent BSDF branches per warp)
    uint8_t* sort_key;       // SPLIT only: material signature of the slot's current hit (kSortBuckets values)
    uint32_t* bucket_count;  // SPLIT only: kSortBuckets histogram of the current shade queue, then scatter cursors
    float4* nee_ray_o;       // PHASED only: 4 NEE rays per path (wavefront_mis.cuh), [which * capacity + slot]: origin | t_max (<0: none)
    float4* nee_ray_d;       //   direction | primitive the ray leaves from
    float4* nee_pending;     //   contribution if unoccluded (rays 0,2,3) or BSDF value * cosine | pdf (ray 1)
    float4* hit_tuvp;        // PHASED only: hit records of the ray-queue kernel, [slot] extension rays, [capacity + slot] NEE light rays
    uint32_t* hit_tri_slot;  //   BVH triangle slot of the hit
    uint32_t* any_queue;     // PHASED only: ids (which * capacity + slot) of the NEE shadow rays of this bounce
    uint32_t* closest_queue; // PHASED only: ids (capacity + slot) of the NEE light rays (closest hit + emission) of this bounce
    uint32_t* counts;        // [4*b + k]: k = 0 rays queued for bounce b, 1 hits queued for shading, 2 NEE any-hit rays, 3 NEE closest-hit rays
    uint32_t* cursors;       // kCursorsPerBounce work cursors per bounce (dynamic fetch of the queue kernels); follows `counts`
    uint32_t sm_count;       // SMs of the device (grid sizing of the queue kernels)
    uint32_t grid_share;     // how many FullPathTracer launches share the device at the same time (path lanes, >= 1): grids are sized 1/share
    uint32_t capacity;       // paths
    uint32_t max_bounces;    // counts holds 2 * (max_bounces + 2) entries
    uint32_t force_slot_materials;   // debug (B200PT_SLOT_MATERIALS=1): every hit writes its 256-byte material to the slot (wf_store_hit)
    // ---- sample batches (b200pt_launch_PathTraceSamples): CameraRays fused in, `samples` sample passes in flight together ----
    uint32_t fused_camera;   // 1: bounce 0 comes from the camera kernel + ray-queue traversal of this launch, not from the G-buffer
    uint32_t samples;        // sample passes in flight (>= 1); slot = sample * tile_pixels + index of the pixel inside the tile
    uint32_t tile_pixels;    // pixels of the launch tile
    float4* final_color;     // fused: [slot] final ray colour | sampled wavelength, accumulated in sample order by k_wf_accumulate
    float4* aov_normal;      // fused: [slot] denoiser normal of the camera hit | 1 if the camera ray hit something
    float4* aov_albedo;      // fused: [slot] denoiser albedo of the camera hit
    // ---- ray binning (launch_bin_rays): a queue's ids regrouped by origin cell and direction octant before it is traced ----
    uint32_t* bin_hist;      // kRayBinsMax + 1 counters / scatter cursors
    uint32_t* binned_any;    // 3 * capacity ids: the NEE any-hit queue in bin order
    uint32_t* binned_ext;    // capacity ids: the extension-ray queue (or the NEE closest-hit queue) in bin order
    uint32_t* bin_keys;      // 3 * capacity: bin of queue entry i (histogram pass -> scatter pass)
    float4* ris_a;           // phased RIS (wavefront_mis.cuh): the reservoir between R1, R2 and the resolve, 2 x float4 per slot
    float4* ris_b;
    // Layout of the seven per-path float4 records above, in three groups that are read and written together: RAY (origin_rng, dir_prim), HIT
    // (hit_point, hit_ns, hit_ng) and RADIANCE (throughput, color). Entry of `slot` = array[slot * stride of its group]. Strides 1 / 1 / 1: seven
    // separate arrays (round 1). Strides 2 / 4 / 2 (default): each group is one array of 32- / 64- / 32-byte records and the pointers of a group are
    // base + 0, 1, .. of it, so that a phase working on a material-sorted (scattered) queue moves whole 32-byte sectors it uses entirely instead
    // of one half-used sector per field, and does not pull the groups it does not read. (One 128-byte record per path was measured too: it is
    // what wins on scattered queues and loses on queues in slot order, where every load instruction of a warp then touches 32 lines.)
    uint32_t stride_ray, stride_hit, stride_radiance;
    // Likewise the NEE records (nee_ray_o | nee_pending, then nee_ray_d: 64-byte records, so that the resolve phase reads a ray's t_max flag
    // and its pending contribution out of ONE sector and the generating phases write two whole sectors) and the hit records of the ray-queue
    // kernel (hit_tuvp and hit_tri_slot in one 32-byte record: one sector written per ray and read per hit instead of two). Entry of id k =
    // nee_*[k * stride_nee], hit_tuvp[k * stride_hitrec], hit_tri_slot[k * stride_hitrec_words]; 1 / 1 / 1 = separate arrays.
    uint32_t stride_nee, stride_hitrec, stride_hitrec_words;
};
constexpr int kRayBinBitsMax = 6;                                   // cells per axis = 2^bits
constexpr uint32_t kRayBinsMax = 1u << (3 * kRayBinBitsMax + 3);    // cells x 8 direction octants

// The sample passes one b200pt_launch_PathTraceSamples call renders: pass s is CameraRays with random_seed = camera_seed[s] followed by
// FullPathTracer with random_seed = path_seed[s], both at sample_number[s] (GPURenderer::render: src/Renderer/GPURenderer.cpp:408-464).
constexpr int kMaxSamplesInFlight = 16;
struct SampleBatch
{
    uint32_t count;
    uint32_t camera_seed[kMaxSamplesInFlight];
    uint32_t path_seed[kMaxSamplesInFlight];
    int sample_number[kMaxSamplesInFlight];
    int aov_counter[kMaxSamplesInFlight];      // denoiser_AOV_accumulation_counter of the pass
    uint32_t write_g_buffer;                   // the last pass of the batch leaves its camera hits in the G-buffer, as CameraRays would
};

#ifdef PT_COUNTERS
#define PT_SUFFIXED(name) name##_c
#define PT_VARIANT_NAME_OF(name, L, E, A, G) name##_##L##_##E##_##A##_##G##_c
#else
#define PT_SUFFIXED(name) name##_t
#define PT_VARIANT_NAME_OF(name, L, E, A, G) name##_##L##_##E##_##A##_##G##_t
#endif
#define PT_VARIANT_NAME_EXPAND(name, L, E, A, G) PT_VARIANT_NAME_OF(name, L, E, A, G)
#define PT_VARIANT_NAME(name) PT_VARIANT_NAME_EXPAND(name, PT_LSS, PT_ESS, PT_LATER, PT_GGX)

// Optional per-stage timing of one FullPathTracer launch (b200pt_path_profile_*): the launcher records an event at every
// change of stage; the context sums the intervals per stage after the stream has drained.
enum WavefrontStage { STAGE_TRACE = 0, STAGE_SURFACE = 1, STAGE_SORT = 2, STAGE_SHADE = 3, STAGE_END = 4 };
struct WavefrontProfile
{
    cudaEvent_t* events;
    int* stages;      // stage that starts at events[i]
    int capacity;
    int count;
};

// Scratch of the PHASED ReSTIR DI passes (restir_di_kernels.cuh): a pass is cut at its rays, which go through the ray-queue traversal
// kernel instead of a per-thread traversal inside the pass. A ray of a pixel belongs to a SITE of the pass (which evaluation of the
// pass it serves); ray id = site * capacity + pixel inside the tile.
//   InitialCandidates:     site 0 = the BSDF candidate's ray (closest hit), site 1 = visibility of the chosen candidate
//   SpatiotemporalReuse:   site 0 = temporal neighbour's sample seen from the pixel, site 1 = the pixel's sample seen from the temporal
//                          neighbour's surface, sites 2 + 2 i / 3 + 2 i = the same pair for spatial neighbour i
//   TemporalReuse:         sites 0 and 1 as above;   SpatialReuse: sites 2 + 2 i / 3 + 2 i as above, site 0 = the pixel's own input sample
constexpr int kRestirMaxNeighbors = 5;                       // spatial neighbours per pixel the phased pass has sites for (more: unphased kernel)
constexpr int kRestirSites = 2 + 2 * kRestirMaxNeighbors;
struct RestirScratch
{
    float4* ray_o;           // [ray id] origin | t_max; the any-hit trace sets .w = -1 when the ray is occluded
    float4* ray_d;           // [ray id] direction | primitive the ray leaves from (int bits)
    float* target;           // [ray id] target function of the site before visibility (0: nothing to trace)
    float4* hit_tuvp;        // [pixel] closest hit of the initial-candidates BSDF ray
    uint32_t* hit_tri_slot;
    float4* ic_state;        // [k * capacity + pixel], k < 4: initial candidates between their phases (reservoir, generator, BSDF sample)
    uint32_t* status;        // [pixel] 0: the pass has nothing to do for the pixel (COLLECT's early returns), else what the next phase continues with
    uint32_t* queue;         // ray ids of the trace launch in flight (kRestirSites * capacity)
    uint32_t* counts;        // 2 x (queue length, work cursor)
    uint32_t capacity;       // pixels
    uint32_t sm_count;
};

// Input/output arrays of one launch of the ray-queue traversal kernel (trace_queue.cuh)
struct RayQueueView
{
    const uint32_t* ids;        // ray ids to trace
    const uint32_t* count;      // number of ids (device)
    uint32_t* cursor;           // work cursor of this launch (zeroed beforehand)
    const float4* ray_o;        // [id]: origin xyz | t_max (any-hit only)
    const float4* ray_d;        // [id]: direction xyz | primitive the ray leaves from (int bits)
    float4* hit_tuvp;           // closest: [id] t, u, v, primitive (int bits; -1 = miss)
    uint32_t* hit_tri_slot;     // closest: [id] index of the triangle in the BVH triangle array
    float4* occluded_flag;      // any-hit: the ray_o array; .w is set to -1 when the ray is occluded
    uint32_t ray_stride_minus_1; // ray_o / ray_d / occluded_flag entry of `id` = array[id * (1 + this)]: 0 for contiguous ray arrays (what aggregate initialisers
                                 // that stop at occluded_flag give), stride_ray - 1 / stride_nee - 1 for the grouped records (WavefrontState)
    uint32_t hit_stride_minus_1;      // likewise hit_tuvp[id * (1 + this)]
    uint32_t hit_word_stride_minus_1; // and hit_tri_slot[id * (1 + this)]
};

#define PT_DECLARE_LAUNCHERS(S)                                                                                                        \
    bool variant_available##S(const VariantOptions& opt);                                                                              \
    cudaError_t launch_CameraRays##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream);   \
    cudaError_t launch_FullPathTracer##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const WavefrontState* wf, int mode, WavefrontProfile* profile); \
    cudaError_t launch_PathTraceSamples##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const WavefrontState* wf, const SampleBatch* batch, WavefrontProfile* profile); \
    cudaError_t launch_ReSTIR_DI_LightsPresampling##S(const VariantOptions& opt, const LightPresamplingParameters* p, cudaStream_t stream);            \
    cudaError_t launch_ReSTIR_DI_InitialCandidates##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch); \
    cudaError_t launch_ReSTIR_DI_SpatiotemporalReuse##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch); \
    cudaError_t launch_ReSTIR_DI_TemporalReuse##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch); \
    cudaError_t launch_ReSTIR_DI_SpatialReuse##S(const VariantOptions& opt, const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const RestirScratch* scratch); \
    cudaError_t launch_trace_queue##S(const void* bvh_device_handle, const RayQueueView& q, bool any_hit, unsigned blocks, TraversalCounters* counters, cudaStream_t stream); \
    cudaError_t launch_bin_rays##S(const void* bvh_device_handle, RayQueueView& q, uint32_t* hist, uint32_t* keys, uint32_t* out_ids, int bits, unsigned sm_count, cudaStream_t stream); \
    cudaError_t launch_trace_queue_pair##S(const void* bvh_device_handle, const RayQueueView& q_any, const RayQueueView& q_closest, unsigned blocks, TraversalCounters* counters, cudaStream_t stream); \
    cudaError_t launch_trace_batch_queue##S(const void* bvh_device_handle, const float* rays, const float* t_max, const int* last_prim, long long n, int* out_prim, float* out_t, float* out_uv, unsigned char* out_occluded, unsigned sm_count, TraversalCounters* counters, cudaStream_t stream); \
    cudaError_t launch_trace_closest##S(const Bvh8Device& bvh, const float* rays, const int* last_prim, long long n, int* out_prim, float* out_t, float* out_uv, TraversalCounters* counters, cudaStream_t stream); \
    cudaError_t launch_trace_shadow##S(const Bvh8Device& bvh, const float* rays, const float* t_max, const int* last_prim, long long n, unsigned char* out_occluded, TraversalCounters* counters, cudaStream_t stream);

// b200pt_debug_*: which = 0 bsdf_eval, 1 bsdf_sample, 2 envmap_sample (alias table), 3 thin_film_fresnel (timing build only)
cudaError_t launch_debug_function(int which, const HIPRTRenderData* rd, const void* materials, void* volume_states, const float* vecs, unsigned int* seeds, long long n, int ggx_sampler, float* out,
                                  cudaStream_t stream);
PT_DECLARE_LAUNCHERS(_t)
PT_DECLARE_LAUNCHERS(_c)
} // namespace pt
