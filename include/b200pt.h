/*
 * b200pt.h -- thin C-ABI CUDA layer for the HIPRT-Path-Tracer render hot path on NVIDIA B200 (sm_100a).
 *
 * This library replaces, for the CameraRays -> [ReSTIR DI passes] -> FullPathTracer path only, what the reference
 * renderer obtains from Orochi + HIPRT + its run-time kernel compiler:
 *
 *   reference call site                                                  replaced by
 *   -------------------------------------------------------------------  ------------------------------------------
 *   HIPRTOrochiCtx::init            src/HIPRT-Orochi/HIPRTOrochiCtx.h:29-57   b200pt_ctx_create / _make_current
 *   OrochiBuffer<T>::resize/upload  src/HIPRT-Orochi/OrochiBuffer.h:91-147    b200pt_malloc / _free / _memcpy_*
 *   OrochiTexture / OrochiTexture3D src/HIPRT-Orochi/OrochiTexture.cpp:47-142 b200pt_tex2d_create / _tex3d_create
 *   OrochiEnvmap alias-table upload src/HIPRT-Orochi/OrochiEnvmap.cpp:53-76   b200pt_malloc + b200pt_memcpy_h2d
 *   HIPRTGeometry::build_bvh        src/HIPRT-Orochi/HIPRTScene.h:59-86       b200pt_bvh_build (compressed BVH8)
 *   GPUKernel::compile (NVRTC)      src/Compiler/GPUKernel.cpp:72-96          none: kernels are built offline for
 *                                                                             sm_100a; b200pt_kernel_variant_select
 *   GPUKernel::launch_asynchronous  src/Compiler/GPUKernel.cpp:169-245        b200pt_launch_<KernelName>
 *   oroStream* / oroEvent*          src/Renderer/GPURenderer.cpp:75-85,494-499 b200pt_stream_* / b200pt_event_*
 *
 * Conventions: every entry point returns 0 on success or a negative b200pt_error; it never exits the process and
 * never throws across the ABI (the reference's OROCHI_CHECK_ERROR logs and exit()s, HIPRTOrochiUtils.cpp:15-48).
 * b200pt_last_error() gives the message of the last failure on the calling thread. There is NO CPU fallback: without
 * a CUDA device every compute entry point fails with B200PT_ERROR_NO_DEVICE.
 * Kernel launches are asynchronous on the given stream; struct arguments are snapshotted at call time, exactly like
 * the by-value kernel arguments of the reference (src/Renderer/GPURenderer.cpp:468,482).
 */
#ifndef B200PT_H
#define B200PT_H

#include "b200pt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define B200PT_API __attribute__((visibility("default")))

typedef enum b200pt_error
{
    B200PT_SUCCESS = 0,
    B200PT_ERROR_INVALID_ARGUMENT = -1,
    B200PT_ERROR_NO_DEVICE = -2,
    B200PT_ERROR_CUDA = -3,
    B200PT_ERROR_OUT_OF_MEMORY = -4,
    B200PT_ERROR_VARIANT_NOT_COMPILED = -5,
    B200PT_ERROR_NOT_READY = -6
} b200pt_error;

typedef struct b200pt_ctx b200pt_ctx;       /* one per GPU (the reference has a single shared context, device 0) */
typedef struct b200pt_bvh b200pt_bvh;
typedef struct b200pt_stream_t* b200pt_stream;   /* NULL = the context's main stream (GPURenderer::m_main_stream) */
typedef struct b200pt_event_t* b200pt_event;

/* Device-visible texture descriptors: what RenderBuffers::material_textures[i], BRDFsData::* and
 * WorldSettings::envmap point to. Sampling follows the reference's CPU rule: nearest texel, repeat addressing,
 * v flipped, x = int(u * (w - 1)) (src/Image/Image.cpp:161-195, 492-526, 737-779). */
typedef struct b200pt_tex2d
{
    const void* texels;            /* device pointer, row-major, `channels` interleaved */
    int width, height;
    int channels;                  /* 1..4 */
    int bytes_per_channel;         /* 1 = unorm8 (Image8Bit), 4 = float32 (Image32Bit) */
} b200pt_tex2d;

typedef struct b200pt_tex3d
{
    const float* texels;           /* device pointer, depth slices of row-major width x height x channels */
    int width, height, depth;
    int channels;
} b200pt_tex3d;

/* Compile-time kernel options of the reference (src/HostDeviceCommon/KernelOptions.h:91-366) that select a prebuilt
 * kernel variant. Values are the reference's macro values. */
typedef struct b200pt_options
{
    int direct_light_sampling_strategy;   /* LSS_*: 0 none, 1 uniform one light, 2 BSDF, 3 MIS light+BSDF, 4 RIS, 5 ReSTIR DI */
    int envmap_sampling_strategy;         /* ESS_*: 0 no sampling, 1 binary search, 2 alias table */
    int ggx_sample_function;              /* 0 GGX_VNDF_SAMPLING, 1 GGX_VNDF_SPHERICAL_CAPS */
    int restir_di_later_bounces_strategy; /* RESTIR_DI_LATER_BOUNCES_*: 0 uniform, 1 BSDF, 2 MIS, 3 RIS */
} b200pt_options;

/* Image-space shard of a launch: pixels with x0 <= x < x1, y0 <= y < y1 are processed; pixel_index stays x + y * res.x
 * so that a sharded render is bit-identical to the unsharded one. NULL = the whole image. */
typedef struct b200pt_tile { int x0, y0, x1, y1; } b200pt_tile;

/* Traversal statistics of the counting build (for Mrays/s and the node-fetch roofline). */
typedef struct b200pt_counters
{
    unsigned long long rays;        /* BVH queries: trace_ray iterations + shadow rays + shadow-light rays */
    unsigned long long nodes;       /* 80-byte BVH8 nodes fetched */
    unsigned long long triangles;   /* 48-byte triangles fetched */
    unsigned long long stack_overflows;   /* traversal-stack pushes that found the stack full (a lost subtree): must be 0 */
    unsigned long long queue_rays;        /* the share of the ray-queue traversal kernel (k_trace_queue), the kernel the */
    unsigned long long queue_nodes;       /*   node-fetch roofline is reported for */
    unsigned long long queue_triangles;
} b200pt_counters;

/* Shape of a built acceleration structure. max_depth bounds the traversal stacks: the ray-queue kernel needs at most
 * 2 * max_depth entries of its 48; b200pt_bvh_build fails with B200PT_ERROR_INVALID_ARGUMENT for a deeper tree. */
typedef struct b200pt_bvh_stats
{
    unsigned node_count, triangle_count;
    int max_depth;
    int built_on_device;            /* 1: the device builder (B200PT_BVH_BUILD_DEVICE), 0: the host SAH builder */
    float sah_cost;
    float node_fill;                /* occupied child slots / (8 * node_count) */
    double build_seconds;           /* wall clock of b200pt_bvh_build (host build + upload, or upload + device build) */
    double device_build_seconds;    /* device builder only: CUDA-event time of the build with the geometry already on the device */
} b200pt_bvh_stats;

/* ---- context, memory, streams, events ------------------------------------------------------------------------ */
B200PT_API int b200pt_device_count(void);
B200PT_API int b200pt_ctx_create(int device_index, b200pt_ctx** out_ctx);
B200PT_API int b200pt_ctx_destroy(b200pt_ctx* ctx);
B200PT_API int b200pt_ctx_make_current(b200pt_ctx* ctx);                 /* oroCtxSetCurrent, GPURenderer.cpp:74 */
B200PT_API const char* b200pt_last_error(void);
B200PT_API const char* b200pt_version(void);

B200PT_API int b200pt_malloc(b200pt_ctx* ctx, size_t bytes, void** out_dptr);
B200PT_API int b200pt_free(b200pt_ctx* ctx, void* dptr);
B200PT_API int b200pt_memset(b200pt_ctx* ctx, void* dptr, int value, size_t bytes, b200pt_stream stream);
B200PT_API int b200pt_memcpy_h2d(b200pt_ctx* ctx, void* dst_dptr, const void* src_host, size_t bytes, b200pt_stream stream);
B200PT_API int b200pt_memcpy_d2h(b200pt_ctx* ctx, void* dst_host, const void* src_dptr, size_t bytes, b200pt_stream stream);
B200PT_API int b200pt_memcpy_d2d(b200pt_ctx* ctx, void* dst_dptr, const void* src_dptr, size_t bytes, b200pt_stream stream);
B200PT_API int b200pt_host_alloc_pinned(b200pt_ctx* ctx, size_t bytes, void** out_host);
B200PT_API int b200pt_host_free_pinned(b200pt_ctx* ctx, void* host);

B200PT_API int b200pt_stream_create(b200pt_ctx* ctx, b200pt_stream* out_stream);
B200PT_API int b200pt_stream_destroy(b200pt_ctx* ctx, b200pt_stream stream);
B200PT_API int b200pt_stream_query(b200pt_ctx* ctx, b200pt_stream stream);     /* 0 = idle, B200PT_ERROR_NOT_READY = busy */
B200PT_API int b200pt_stream_sync(b200pt_ctx* ctx, b200pt_stream stream);
B200PT_API int b200pt_event_create(b200pt_ctx* ctx, b200pt_event* out_event);
B200PT_API int b200pt_event_destroy(b200pt_ctx* ctx, b200pt_event event);
B200PT_API int b200pt_event_record(b200pt_ctx* ctx, b200pt_event event, b200pt_stream stream);
B200PT_API int b200pt_event_elapsed_ms(b200pt_ctx* ctx, b200pt_event start, b200pt_event stop, float* out_ms);

/* ---- textures ------------------------------------------------------------------------------------------------ */
/* Uploads host texels and returns the DEVICE address of a b200pt_tex2d / b200pt_tex3d descriptor.
 * 3-channel 8-bit images are accepted (the reference's GPU path rejects them, OrochiTexture.cpp:50-55; its CPU path does not). */
B200PT_API int b200pt_tex2d_create(b200pt_ctx* ctx, const void* host_texels, int width, int height, int channels, int bytes_per_channel, void** out_descriptor_dptr);
B200PT_API int b200pt_tex3d_create(b200pt_ctx* ctx, const float* host_texels, int width, int height, int depth, int channels, void** out_descriptor_dptr);
B200PT_API int b200pt_tex_destroy(b200pt_ctx* ctx, void* descriptor_dptr);
/* Builds the contiguous device array of b200pt_tex2d that RenderBuffers::material_textures points to. */
B200PT_API int b200pt_tex2d_array_create(b200pt_ctx* ctx, void* const* descriptor_dptrs, int count, void** out_array_dptr);

/* ---- acceleration structure ---------------------------------------------------------------------------------- */
/* vertices: 3 floats per vertex; triangle_indices: 3 ints per triangle (host pointers). The primitive index reported
 * by the kernels is the index into triangle_indices / 3, as in the reference. flags: 0, B200PT_BVH_BUILD_DEVICE, or that | B200PT_BVH_BUILD_DEVICE_LBVH.
 * Environment: B200PT_BVH_BUILD_DEVICE=1 forces the device builder, B200PT_BVH_DEVICE_ALGORITHM=lbvh|ploc and B200PT_PLOC_RADIUS (default 16) tune it. */
#define B200PT_BVH_BUILD_DEVICE 1u   /* flags: build on the GPU (Morton order -> PLOC clustering -> SAH-optimal collapse) instead of the host's binned-SAH builder */
#define B200PT_BVH_BUILD_DEVICE_LBVH 2u   /* with B200PT_BVH_BUILD_DEVICE: the Karras radix tree instead of PLOC (faster to build, ~10 % higher SAH cost) */
B200PT_API int b200pt_bvh_build(b200pt_ctx* ctx, const float* vertices, int vertex_count, const int* triangle_indices, int triangle_count, unsigned flags, b200pt_bvh** out_bvh);
B200PT_API int b200pt_bvh_destroy(b200pt_ctx* ctx, b200pt_bvh* bvh);
/* Optional, after the scene's vertex data is on the device (GPURenderer::set_scene, src/Renderer/GPURenderer.cpp:1041-1134 uploads it next to
 * the geometry build): gathers, per triangle SLOT of the BVH, the three vertex normals and texcoords (one 64-byte record) plus the material index
 * and vertex A's has_vertex_normals flag (the two spare words of the 48-byte triangle record). A hit then reads 2 + 2 sectors of its own slot
 * instead of following the index buffer to three scattered vertices (~10 sectors): same values, so results are bit-identical with and without.
 * The pointers are the device arrays RenderBuffers carries (triangles_indices, vertex_normals, has_vertex_normals, texcoords,
 * material_indices); call it again after changing any of them, or with d_triangle_indices = NULL to detach. No counterpart in the reference
 * (HIPRT hands out primitive ids only). Synchronous with respect to `stream`. */
B200PT_API int b200pt_bvh_attach_shading_data(b200pt_ctx* ctx, b200pt_bvh* bvh, const int* d_triangle_indices, const float* d_vertex_normals,
                                              const unsigned char* d_has_vertex_normals, const float* d_texcoords, const int* d_material_indices, b200pt_stream stream);
B200PT_API void* b200pt_bvh_device_handle(b200pt_bvh* bvh);             /* value for HIPRTRenderData::geom */
B200PT_API int b200pt_bvh_info(b200pt_bvh* bvh, unsigned* out_node_count, unsigned* out_triangle_count, float* out_sah_cost, double* out_build_seconds);
B200PT_API int b200pt_bvh_get_stats(b200pt_bvh* bvh, b200pt_bvh_stats* out_stats);

/* ---- kernels (names of the reference's kernels kept; src/Device/kernels) --------------------------------------- */
B200PT_API int b200pt_kernel_variant_select(b200pt_ctx* ctx, const b200pt_options* options);
/* Shape of FullPathTracer: 0 = megakernel (one launch, one thread per pixel, as the reference), 1 = wavefront with one
 * fused trace+shade launch per bounce over a compacted queue of live paths, 2 = wavefront with separate extend and
 * shade launches per bounce and the shade queue grouped by material signature, 3 = (default) as 2 with the shade step
 * cut into phases (light eval / BSDF samples / NEE rays / bounce), one small kernel each.
 * Same per-path arithmetic and RNG order in all four: same bits. */
B200PT_API int b200pt_set_path_mode(b200pt_ctx* ctx, int mode);
B200PT_API int b200pt_get_path_mode(b200pt_ctx* ctx, int* out_mode);
/* Device time of the stages of the wavefront FullPathTracer, summed over the launches since the last reset:
 * out_stage_ms[4] = { ray-queue traversal kernels, gen/surface (hit -> shading data), queue sort, shading phases }.
 * Measured with CUDA events on the launching stream between the stages' kernels. */
B200PT_API int b200pt_path_profile_enable(b200pt_ctx* ctx, int enable);
B200PT_API int b200pt_path_profile_read(b200pt_ctx* ctx, double* out_stage_ms, int reset);
B200PT_API int b200pt_launch_CameraRays(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream);
B200PT_API int b200pt_launch_FullPathTracer(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream);
/* New (no reference counterpart): the same launch as one of `lane_count` (<= 4) concurrent band launches. Each lane has its own
 * wavefront scratch and sizes its grids for 1/lane_count of an SM, so that launches issued on different streams for different
 * tiles overlap (the ray-queue traversal of one band is latency-bound, the shade phases of another are DRAM-bound).
 * b200pt_launch_FullPathTracer is lane 0 of 1. */
B200PT_API int b200pt_launch_FullPathTracer_lane(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream, int lane, int lane_count);
/* New (no reference counterpart; replaces the LOOP around the two launch points, src/Renderer/GPURenderer.cpp:408-464, when samples_per_frame
 * sample passes are rendered per frame): `count` consecutive sample passes -- pass i = CameraRays with random_seed = camera_seeds[i], then
 * FullPathTracer with random_seed = path_seeds[i], both at render_settings.sample_number = sample_numbers[i] and
 * denoiser_AOV_accumulation_counter = aov_counters[i] -- as ONE wavefront: CameraRays is fused in (the primary rays go through the ray-queue
 * traversal kernel; the 345 B/px G-buffer round trip per pass disappears) and all passes' paths are in flight together, so the per-launch
 * costs of the bounce chain are paid once per batch and a small tile (multi-GPU tile sharding) still fills the GPU. The frame buffers
 * receive the passes in pass order: same bits as `count` iterations of the two separate launches. With write_g_buffer the last pass
 * leaves its camera hits in the G-buffer as CameraRays would. Restrictions (B200PT_ERROR_VARIANT_NOT_COMPILED otherwise; use the separate
 * launches then): LSS_MIS_LIGHT_BSDF or LSS_RIS_BSDF_AND_LIGHT (with one BSDF candidate, the default), one light sample, no alpha testing, no adaptive sampling / pixel noise threshold, no ReSTIR DI, no
 * low-resolution rendering -- the options under which a pass does not read what the previous pass accumulated. */
typedef struct b200pt_sample_batch
{
    int count;                       /* 1..16 */
    unsigned camera_seeds[16];
    unsigned path_seeds[16];
    int sample_numbers[16];
    int aov_counters[16];
    int write_g_buffer;
} b200pt_sample_batch;
B200PT_API int b200pt_launch_PathTraceSamples(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream,
                                              const b200pt_sample_batch* batch, int lane, int lane_count);
/* cudaStreamWaitEvent: work queued on `stream` after this call starts after `event` (replaces oroStreamWaitEvent) */
B200PT_API int b200pt_stream_wait_event(b200pt_ctx* ctx, b200pt_stream stream, b200pt_event event);
B200PT_API int b200pt_launch_ReSTIR_DI_LightsPresampling(b200pt_ctx* ctx, const LightPresamplingParameters* parameters, b200pt_stream stream);
B200PT_API int b200pt_launch_ReSTIR_DI_InitialCandidates(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream);
B200PT_API int b200pt_launch_ReSTIR_DI_SpatiotemporalReuse(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream);
/* The stand-alone reuse passes: replace the launches of ReSTIR_DI_TemporalReuse (src/Device/kernels/ReSTIR/DI/TemporalReuse.h:48,
 * launched at src/Renderer/RenderPasses/ReSTIRDIRenderPass.cpp:368-376) and ReSTIR_DI_SpatialReuse (SpatialReuse.h:64, launched at
 * ReSTIRDIRenderPass.cpp:457-474 and, after the fused pass when number_of_passes > 1, :492-501). Buffers are the pointers inside
 * render_data->render_settings.restir_di_settings.{temporal_pass,spatial_pass}; the temporal pass may write in place over the
 * initial-candidates buffer (a pixel reads and writes only its own entry of it), a spatial pass must not alias input and output. */
B200PT_API int b200pt_launch_ReSTIR_DI_TemporalReuse(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream);
B200PT_API int b200pt_launch_ReSTIR_DI_SpatialReuse(b200pt_ctx* ctx, const HIPRTRenderData* render_data, b200pt_int2 res, const b200pt_tile* tile, b200pt_stream stream);

/* ---- headless output (replaces the GL display path) ------------------------------------------------ */
/* Replaces CPURenderer::tonemap (src/Renderer/CPURenderer.cpp:663-686) / Utils::tonemap_hdr_image (src/Utils/Utils.cpp:35-52):
 * out = pow(1 - exp(-(hdr / sample_number) * exposure), 1 / gamma) per channel; sample_number <= 0 = no scaling (not accumulating).
 * d_out_rgb_float (W*H*3 floats; may alias d_hdr_rgb when flip_y is 0) and/or d_out_rgb8 (W*H*3 bytes, quantised like
 * Image32Bit::write_image_png, src/Image/Image.cpp:427-437: clamp(0, 255, v * 255) truncated); flip_y writes row H-1-y. */
B200PT_API int b200pt_tonemap(b200pt_ctx* ctx, const float* d_hdr_rgb, int width, int height, int sample_number, float gamma, float exposure, int flip_y,
                              float* d_out_rgb_float, unsigned char* d_out_rgb8, b200pt_stream stream);

/* ---- ray batches: the traversal core on its own (parity pins on hit ids; roofline measurement) ------------------ */
/* rays: n x 6 floats (origin, direction) on the device. last_prim: n ints or NULL (-1). Outputs on the device. */
B200PT_API int b200pt_trace_closest(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const int* d_last_prim, long long n, int* d_out_prim, float* d_out_t, float* d_out_uv, b200pt_stream stream);
B200PT_API int b200pt_trace_shadow(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const float* d_t_max, const int* d_last_prim, long long n, unsigned char* d_out_occluded, b200pt_stream stream);
/* The same queries through the ray-queue traversal kernel FullPathTracer uses (persistent warps, per-lane ray replacement), so that
 * hit-id parity is pinned on the kernel the render runs. d_t_max == NULL: closest hit (d_out_prim / d_out_t / d_out_uv); else any-hit
 * (d_out_occluded). Synchronises the stream before it returns (it owns temporary queue buffers). */
B200PT_API int b200pt_trace_queue(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const float* d_t_max, const int* d_last_prim, long long n, int* d_out_prim, float* d_out_t,
                                  float* d_out_uv, unsigned char* d_out_occluded, b200pt_stream stream);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA graph
 * capture, tight serving loops, harnesses that forbid it). The same query as b200pt_trace_queue, same outputs bit for bit, but it allocates
 * NOTHING (no cudaMalloc / cudaFree / allocating library call): its queue scratch is one file-scope static __device__ array sized for
 * B200PT_TRACE_QUEUE_STATIC_MAX_RAYS rays (56 B per ray, about 14 MiB, held by every GPU that loads the library until exit). n beyond that
 * maximum is rejected with B200PT_ERROR_INVALID_ARGUMENT and a message naming the limit. Asynchronous on `stream`. Every call shares the array
 * (each call rewrites all it reads and zeroes its work cursor): calls that overlap on the GPU race -- serialise them. */
#define B200PT_TRACE_QUEUE_STATIC_MAX_RAYS 262144
B200PT_API int b200pt_trace_queue_static(b200pt_ctx* ctx, b200pt_bvh* bvh, const float* d_rays, const float* d_t_max, const int* d_last_prim, long long n, int* d_out_prim, float* d_out_t,
                                         float* d_out_uv, unsigned char* d_out_occluded, b200pt_stream stream);

/* Loop flavour of the ray-queue traversal kernel: 0 = classic (load, wait, test), 5 / 6 / 7 = software-pipelined (every load issued one
 * iteration ahead) compiled for that many resident blocks per SM, 1 = the pipelined default, -1 = B200PT_TRACE_PIPELINE / built-in default.
 * A scheduling choice only: hit records are identical in every flavour. Process-wide. */
B200PT_API int b200pt_set_trace_flavour(b200pt_ctx* ctx, int flavour);

/* Ray binning (new): before a queue of bounce / NEE rays is traced, its ids are regrouped by (origin cell in the scene's bounding box,
 * direction octant), 2^bits cells per axis (0 = off, max 6, -1 = B200PT_RAY_BIN_BITS / built-in default), so that the rays of a warp walk
 * the same part of the BVH. A scheduling choice only: every ray's hit record, hence the frame, is the same bits. Process-wide. */
B200PT_API int b200pt_set_ray_bin_bits(b200pt_ctx* ctx, int bits);

/* Phased ReSTIR DI passes (new; default on): b200pt_launch_ReSTIR_DI_InitialCandidates and _SpatiotemporalReuse run as
 * COLLECT kernel(s) -> ray-queue traversal of the pass's rays -> RESOLVE kernel instead of one kernel that traces per thread (the reference's
 * shape: ReSTIR_DI_InitialCandidates.h, FusedSpatiotemporalReuse.h). Needs no alpha testing, full-resolution rendering, <= 1 BSDF candidate
 * and <= 5 reused spatial neighbours; otherwise, or with enable = 0, or B200PT_RESTIR_PHASES=0, the single kernels run. The reservoirs a
 * pass writes are the same bits either way. The phased passes of one context share scratch: do not overlap them on different streams. */
B200PT_API int b200pt_set_restir_phases(b200pt_ctx* ctx, int enable);

/* ---- per-function known-answer entry points (the parity pins of SURVEY.md 8c item 3; not used by the render path) -------------------
 * The device functions of the shading path on caller-provided inputs, one item per thread, so that an image difference can be bisected to
 * a function. render_data supplies the LUT / envmap pointers (take it from a renderer that has rendered a frame). Device pointers; layouts
 * as the oracle's ref_bsdf_eval / ref_bsdf_sample / ref_envmap_sample / ref_thin_film_fresnel:
 *   bsdf_eval     bsdf_dispatcher_eval   (src/Device/includes/Dispatcher.h:18-44): materials n x 256 B (SimplifiedRendererMaterial), volume states
 *                 n x 36 B (in/out), vecs n x 9 floats (view, shading normal, to-light direction) -> out n x 4 floats (rgb, pdf)
 *   bsdf_sample   bsdf_dispatcher_sample (Dispatcher.h:46-68): vecs n x 9 (view, shading normal, geometric normal), seeds n (Xorshift32 state,
 *                 in/out), ggx_sampler 0 VNDF / 1 spherical caps (BSDFs/Microfacet.h:162-218) -> out n x 7 (rgb, pdf, sampled direction)
 *   envmap_sample envmap_sample + envmap_eval (Envmap.h:77-149), alias table: seeds n (in/out) -> out n x 8 (radiance rgb, pdf, direction, eval pdf)
 *   thin_film_fresnel (BSDFs/ThinFilm.h:112-219): materials n x 256 B, n x 2 floats (ambient ior, cosine) -> out n x 3 (rgb) */
B200PT_API int b200pt_debug_bsdf_eval(b200pt_ctx* ctx, const HIPRTRenderData* render_data, const void* d_materials, void* d_volume_states, const float* d_vecs, long long n, float* d_out);
B200PT_API int b200pt_debug_bsdf_sample(b200pt_ctx* ctx, const HIPRTRenderData* render_data, const void* d_materials, void* d_volume_states, const float* d_vecs, unsigned int* d_seeds, long long n,
                                        int ggx_sampler, float* d_out);
B200PT_API int b200pt_debug_envmap_sample(b200pt_ctx* ctx, const HIPRTRenderData* render_data, unsigned int* d_seeds, long long n, float* d_out);
B200PT_API int b200pt_debug_thin_film_fresnel(b200pt_ctx* ctx, const void* d_materials, const float* d_ambient_ior_and_cosine, long long n, float* d_out);

/* ---- counters (counting build of the same kernels; the timing build has them compiled out) --------------------- */
B200PT_API int b200pt_counters_enable(b200pt_ctx* ctx, int enable);
B200PT_API int b200pt_counters_read(b200pt_ctx* ctx, b200pt_counters* out, int reset);

#ifdef __cplusplus
}
#endif

#endif /* B200PT_H */
