/*
 * b200pt_host.h -- C handle API over b200pt_host::GPURenderer (hiprt-path-tracer_b200/host/GPURenderer.h), the C++
 * re-hosting of the reference's GPURenderer launch loop (src/Renderer/GPURenderer.cpp:236-486) on top of b200pt.h.
 * It exists so that non-C++ callers (the Python parity tests, bench.py) drive exactly the same host code a C++
 * integrator would; it adds no functionality of its own. Every function returns 0 or a negative b200pt_error and
 * sets b200pt_host_last_error().
 */
#ifndef B200PT_HOST_H
#define B200PT_HOST_H

#include "b200pt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200pt_renderer b200pt_renderer;

B200PT_API const char* b200pt_host_last_error(void);
B200PT_API int b200pt_renderer_create(b200pt_ctx* ctx, b200pt_renderer** out);
B200PT_API int b200pt_renderer_destroy(b200pt_renderer* r);
/* seed_schedule: 0 = GPU (fresh seed per launch, GPURenderer.cpp:466-486), 1 = CPU (CPURenderer.cpp:264-296) */
B200PT_API int b200pt_renderer_set_seed_schedule(b200pt_renderer* r, int seed_schedule);
/* Seed of the host xorshift32 generator the per-launch random_seed values come from (42 in the reference, nonzero); resets the
 * accumulation. One value per rank under sample-pass sharding: independent samples at equal sample_number on every rank. */
B200PT_API int b200pt_renderer_set_rng_seed(b200pt_renderer* r, unsigned int seed);
B200PT_API int b200pt_renderer_set_kernel_options(b200pt_renderer* r, const b200pt_options* options);
B200PT_API int b200pt_renderer_set_brdf_luts(b200pt_renderer* r, const float* sheen_ltc, const float* conductor, const float* glossy,
                                             const float* glass, const float* glass_inverse, const float* thin_glass);
B200PT_API int b200pt_renderer_add_texture(b200pt_renderer* r, const unsigned char* texels, int width, int height, int channels);
B200PT_API int b200pt_renderer_set_scene(b200pt_renderer* r, const float* vertices, int vertex_count, const int* triangle_indices, int triangle_count,
                                         const float* vertex_normals, const unsigned char* has_vertex_normals, const float* texcoords,
                                         const int* material_indices, const RendererMaterial* materials, int material_count,
                                         const int* emissive_triangle_indices, int emissive_triangle_count);
B200PT_API int b200pt_renderer_set_envmap(b200pt_renderer* r, const float* texels, int width, int height, int channels,
                                          const float* alias_probas, const int* alias_alias, float total_sum);
/* The alias table set_envmap builds when none is passed (replaces Image32Bit::compute_alias_table, src/Image/Image.cpp:579-659,
 * called from GPURenderer::set_envmap, src/Renderer/GPURenderer.cpp:1136-1174). Host-only: needs no device. */
B200PT_API int b200pt_host_compute_alias_table(const float* texels, int width, int height, int channels, float* out_probas, int* out_alias, float* out_total_sum);
B200PT_API int b200pt_renderer_set_camera(b200pt_renderer* r, const HIPRTCamera* camera);
B200PT_API int b200pt_renderer_resize(b200pt_renderer* r, int width, int height);
B200PT_API int b200pt_renderer_set_tile(b200pt_renderer* r, const b200pt_tile* tile);
/* Tile sharding with ReSTIR DI (new): CameraRays also covers `rows` rows above and below the tile, so that the
 * G-buffers the fused spatiotemporal pass reads around the tile are computed locally; only reservoirs are exchanged. */
B200PT_API int b200pt_renderer_set_tile_halo(b200pt_renderer* r, int rows);
/* FullPathTracer as 1..4 concurrent band launches on separate streams (new; see b200pt_launch_FullPathTracer_lane). Also read
 * from the environment variable B200PT_PATH_LANES at construction. Pixels do not change (bands are row tiles). */
B200PT_API int b200pt_renderer_set_path_lanes(b200pt_renderer* r, int lanes);
B200PT_API int b200pt_renderer_path_lanes(b200pt_renderer* r);
/* Fused sample batches (new; see b200pt_launch_PathTraceSamples): a frame's samples_per_frame passes run as wavefronts of up to `samples`
 * (0..16; 0 = off, default 16, environment B200PT_SAMPLES_IN_FLIGHT) passes with CameraRays fused in, whenever the options allow it
 * (b200pt_renderer_renders_sample_batches tells); same bits as the per-pass launches. keep_g_buffer: the last pass of a batch leaves its
 * camera hits in the G-buffer as CameraRays would (345 B/px of stores; switch off when nothing reads the G-buffer). */
B200PT_API int b200pt_renderer_set_samples_in_flight(b200pt_renderer* r, int samples, int keep_g_buffer);
B200PT_API int b200pt_renderer_renders_sample_batches(b200pt_renderer* r);
B200PT_API int b200pt_renderer_reset(b200pt_renderer* r);
/* Host-side structs the caller may edit between frames (copied by value into each launch) */
B200PT_API HIPRTRenderSettings* b200pt_renderer_render_settings(b200pt_renderer* r);
B200PT_API WorldSettings* b200pt_renderer_world_settings(b200pt_renderer* r);
B200PT_API BRDFsData* b200pt_renderer_bsdfs_data(b200pt_renderer* r);
B200PT_API HIPRTRenderData* b200pt_renderer_render_data(b200pt_renderer* r);
B200PT_API b200pt_bvh* b200pt_renderer_bvh(b200pt_renderer* r);
/* update() + render() of one frame = render_settings.samples_per_frame samples; asynchronous */
B200PT_API int b200pt_renderer_render_frame(b200pt_renderer* r);
B200PT_API int b200pt_renderer_synchronize(b200pt_renderer* r);
B200PT_API int b200pt_renderer_last_frame_ms(b200pt_renderer* r, float* out_ms);
/* device time of one kernel in the last sample of the last frame; which: 0 CameraRays, 1 FullPathTracer, 2 ReSTIR DI passes */
B200PT_API int b200pt_renderer_last_kernel_ms(b200pt_renderer* r, int which, float* out_ms);
/* which: 0 pixels(3 f32) 1 first_hit_prim_index(i32) 2 shading_normals 3 geometric_normals 4 view_directions 5 first_hits
 *        6 camera_ray_hit(u8) 7 materials(256 B) 8 ray_volume_states(36 B) 9 pixel_active(u8) 10 pixel_sample_count(i32)
 *        11 pixel_squared_luminance(f32) 12 pixel_converged_sample_count(i32) 13 denoiser_albedo 14 denoiser_normals
 *        15 restir output reservoirs(36 B) 16 restir initial candidates 17 presampled lights(48 B)
 * Returns the number of bytes copied, or the required size negated if capacity is too small. */
B200PT_API long long b200pt_renderer_download(b200pt_renderer* r, int which, void* out_host, long long capacity_bytes);
B200PT_API void* b200pt_renderer_device_pointer(b200pt_renderer* r, int which);

/* ---- b200pt_host::GPUBaker (hiprt-path-tracer_b200/host/GPUBaker.h): the reference's GPUBaker (src/Renderer/Baker/GPUBaker.cpp:12-96) ----------
 * table: a b200pt_bake_kind (b200pt_bake.h); B200PT_BAKE_GGX_GLASS_ENTERING bakes BOTH glass tables like GPUBaker::bake_ggx_glass_directional_albedo
 * (the exiting one goes to "inv_" + output_filename). b200pt_baker_bake returns at once; a worker writes the .hdr slices into output_directory
 * ("" / NULL: the working directory, as the reference does; NULL output_filename: no files) and raises the flag b200pt_baker_is_complete polls. */
typedef struct b200pt_baker b200pt_baker;
B200PT_API int b200pt_baker_create(b200pt_ctx* ctx, b200pt_baker** out);
B200PT_API int b200pt_baker_destroy(b200pt_baker* baker);                      /* waits for the bakes in flight */
B200PT_API int b200pt_baker_set_options(b200pt_baker* baker, int ggx_sample_function, int ggx_masking_shadowing, int chunks, const char* output_directory);
B200PT_API int b200pt_baker_bake(b200pt_baker* baker, int table, int size_cos_theta, int size_roughness, int size_ior, int integration_sample_count, const char* output_filename);
B200PT_API int b200pt_baker_is_complete(b200pt_baker* baker, int table);       /* 1 complete (also: never started), 0 in flight */
B200PT_API int b200pt_baker_wait(b200pt_baker* baker);                         /* every bake finished; B200PT_ERROR_CUDA + message if one of them failed */
/* The last finished table of that kind: floats copied (or the required count negated); out_kernel_ms may be NULL. */
B200PT_API long long b200pt_baker_last_table(b200pt_baker* baker, int kind, float* out, long long capacity_floats, float* out_kernel_ms);
/* The file writer the baker uses for a table slice (Image32Bit::write_image_hdr of a 1-channel image, src/Image/Image.cpp:440-452): Radiance RGBE,
 * grey replicated, rows top to bottom, non-finite texels written as 0. Host only. */
B200PT_API int b200pt_host_write_hdr_1channel(const char* filename, const float* texels, int width, int height);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_HOST_H */
