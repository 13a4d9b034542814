// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
//
// C-API driver around the *unmodified* reference CPU kernels of wuyakuma/HIPRT-Path-Tracer.
// This file is compiled by oracle/build_ref.py against the sources where they lie under
// /root/reference (nothing is copied), producing oracle/_ref/libhippt_ref_<variant>.so.
//
// What is restated here (host logic only; the device headers are compiled as they are):
//   * CPURenderer::CPURenderer / setup_brdfs_data / set_scene / set_envmap
//       (reference src/Renderer/CPURenderer.cpp:68-241) -- buffers are owned by RefCtx,
//       LUTs come in as float arrays instead of ../data/BRDFsData/*.hdr paths.
//   * CPURenderer::render / update / update_render_data / camera_rays_pass / tracing_pass
//       (src/Renderer/CPURenderer.cpp:264-316, 378-399, 656-661) with DEBUG_PIXEL off (:27).
//   * CPURenderer::ReSTIR_DI and its pass launchers (src/Renderer/CPURenderer.cpp:401-654).
//   * Scene::get_triangles (src/Scene/SceneParser.h:118-130).
// Ray counting / ray-batch logging is done by link-time --wrap of BVH::intersect
// (src/Renderer/BVH.cpp:66-69) so that the reference sources stay untouched.

#include "Device/kernels/CameraRays.h"
#include "Device/kernels/FullPathTracer.h"
#include "Device/kernels/ReSTIR/DI/LightsPresampling.h"
#include "Device/kernels/ReSTIR/DI/InitialCandidates.h"
#include "Device/kernels/ReSTIR/DI/FusedSpatiotemporalReuse.h"
#include "Device/kernels/ReSTIR/DI/TemporalReuse.h"
#include "Device/kernels/ReSTIR/DI/SpatialReuse.h"
#include "Device/includes/BSDFs/SheenLTCFittedParameters.h"
#include "Renderer/BVH.h"
#include "Renderer/CPURendererGBuffer.h"
#include "Image/Image.h"

#include <atomic>
#include <chrono>
#include <cstddef>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <omp.h>

// ---------------------------------------------------------------------------------------------
// Ray counter + optional ray log (filled by __wrap_ of BVH::intersect)
// ---------------------------------------------------------------------------------------------
struct RayLogEntry
{
    float ox, oy, oz, dx, dy, dz;
    int last_prim;
    int hit_prim;
    float t, u, v;
};

static std::atomic<unsigned long long> g_ray_count{0};
static std::atomic<long long> g_log_cursor{0};
static RayLogEntry* g_log = nullptr;
static long long g_log_capacity = 0;
static int g_log_stride = 1;   // keep 1 ray out of g_log_stride

extern "C" bool __real__ZNK3BVH9intersectERK8hiprtRayR7HitInfoPv(const BVH* self, const hiprtRay& ray, HitInfo& hit_info, void* payload);
extern "C" bool __wrap__ZNK3BVH9intersectERK8hiprtRayR7HitInfoPv(const BVH* self, const hiprtRay& ray, HitInfo& hit_info, void* payload)
{
    unsigned long long n = g_ray_count.fetch_add(1, std::memory_order_relaxed);
    bool hit = __real__ZNK3BVH9intersectERK8hiprtRayR7HitInfoPv(self, ray, hit_info, payload);
    if (g_log != nullptr && (n % (unsigned long long)g_log_stride) == 0)
    {
        long long slot = g_log_cursor.fetch_add(1, std::memory_order_relaxed);
        if (slot < g_log_capacity)
        {
            RayLogEntry& e = g_log[slot];
            e.ox = ray.origin.x; e.oy = ray.origin.y; e.oz = ray.origin.z;
            e.dx = ray.direction.x; e.dy = ray.direction.y; e.dz = ray.direction.z;
            e.last_prim = payload ? reinterpret_cast<FilterFunctionPayload*>(payload)->last_hit_primitive_index : -1;
            e.hit_prim = hit ? hit_info.primitive_index : -1;
            e.t = hit ? hit_info.t : -1.0f;
            e.u = hit ? hit_info.uv.x : 0.0f;
            e.v = hit ? hit_info.uv.y : 0.0f;
        }
    }
    return hit;
}

// ---------------------------------------------------------------------------------------------
// Context: owns every buffer HIPRTRenderData points to (CPURenderer members, CPURenderer.h)
// ---------------------------------------------------------------------------------------------
struct RefCtx
{
    HIPRTRenderData rd;
    int2 res = make_int2(0, 0);
    Xorshift32Generator rng = Xorshift32Generator(42);

    // scene
    std::vector<float3> vertices, normals;
    std::vector<float2> texcoords;
    std::vector<unsigned char> has_normals;
    std::vector<int> indices, material_indices, emissive;
    std::vector<RendererMaterial> materials;
    std::vector<Triangle> triangles;
    std::shared_ptr<BVH> bvh;
    std::vector<Image8Bit> textures;
    std::vector<int2> textures_dims;

    // LUTs
    Image32Bit sheen_ltc, conductor_ess;
    Image32Bit3D glossy_ess, glass_ess, glass_inv_ess, thin_glass_ess;

    // envmap
    Image32Bit envmap;
    std::vector<float> alias_probas, envmap_cdf;
    std::vector<int> alias_alias;

    // frame buffers
    std::vector<ColorRGB32F> pixels;
    std::vector<unsigned char> pixel_active;
    std::vector<ColorRGB32F> denoiser_albedo;
    std::vector<float3> denoiser_normals;
    std::vector<int> pixel_sample_count, pixel_converged_sample_count;
    std::vector<float> pixel_squared_luminance;
    unsigned char still_one_ray_active = 1;
    AtomicType<unsigned int> stop_noise_threshold_count;
    CPURendererGBuffer g_buffer, g_buffer_prev;

    // ReSTIR DI state (CPURenderer.h ReSTIRDIState)
    std::vector<ReSTIRDIReservoir> initial_candidates_reservoirs, spatial_output_1, spatial_output_2;
    std::vector<ReSTIRDIPresampledLight> presampled_lights;
    ReSTIRDIReservoir* output_reservoirs = nullptr;
    bool odd_frame = false;

    double last_render_seconds = 0.0;
};

static Image32Bit make_image(const float* data, int w, int h, int channels)
{
    std::vector<float> v(data, data + (size_t)w * h * channels);
    return Image32Bit(v, w, h, channels);
}

static Image32Bit3D make_image3d(const float* data, int w, int h, int d)
{
    std::vector<Image32Bit> slices(d);
    for (int z = 0; z < d; z++)
        slices[z] = make_image(data + (size_t)z * w * h, w, h, 1);
    return Image32Bit3D(slices);
}

template <typename F>
static void for_each_pixel(int2 res, F&& f)
{
    // src/Renderer/CPURenderer.cpp:378-389 (DEBUG_PIXEL == 0 branch)
#pragma omp parallel for schedule(dynamic)
    for (int y = 0; y < res.y; y++)
        for (int x = 0; x < res.x; x++)
            f(x, y);
}

extern "C" {

int ref_sizeof(const char* name)
{
    std::string n(name);
    if (n == "HIPRTRenderData") return sizeof(HIPRTRenderData);
    if (n == "RenderBuffers") return sizeof(RenderBuffers);
    if (n == "AuxiliaryBuffers") return sizeof(AuxiliaryBuffers);
    if (n == "BRDFsData") return sizeof(BRDFsData);
    if (n == "GBuffer") return sizeof(GBuffer);
    if (n == "HIPRTRenderSettings") return sizeof(HIPRTRenderSettings);
    if (n == "ReSTIRDISettings") return sizeof(ReSTIRDISettings);
    if (n == "WorldSettings") return sizeof(WorldSettings);
    if (n == "HIPRTCamera") return sizeof(HIPRTCamera);
    if (n == "RendererMaterial") return sizeof(RendererMaterial);
    if (n == "SimplifiedRendererMaterial") return sizeof(SimplifiedRendererMaterial);
    if (n == "RayVolumeState") return sizeof(RayVolumeState);
    if (n == "ReSTIRDIReservoir") return sizeof(ReSTIRDIReservoir);
    if (n == "ReSTIRDISample") return sizeof(ReSTIRDISample);
    if (n == "ReSTIRDIPresampledLight") return sizeof(ReSTIRDIPresampledLight);
    if (n == "LightPresamplingParameters") return sizeof(LightPresamplingParameters);
    if (n == "HitInfo") return sizeof(HitInfo);
    if (n == "hiprtRay") return sizeof(hiprtRay);
    if (n == "hiprtHit") return sizeof(hiprtHit);
    return -1;
}

#define REF_OFF(S, f) if (n == #S "." #f) return (int)offsetof(S, f)
int ref_offsetof(const char* name)
{
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Winvalid-offsetof"
    std::string n(name);
    REF_OFF(HIPRTRenderData, random_seed); REF_OFF(HIPRTRenderData, geom); REF_OFF(HIPRTRenderData, hiprt_function_table);
    REF_OFF(HIPRTRenderData, global_traversal_stack_buffer_size); REF_OFF(HIPRTRenderData, global_traversal_stack_buffer);
    REF_OFF(HIPRTRenderData, buffers); REF_OFF(HIPRTRenderData, bsdfs_data); REF_OFF(HIPRTRenderData, aux_buffers);
    REF_OFF(HIPRTRenderData, g_buffer); REF_OFF(HIPRTRenderData, g_buffer_prev_frame); REF_OFF(HIPRTRenderData, render_settings);
    REF_OFF(HIPRTRenderData, world_settings); REF_OFF(HIPRTRenderData, current_camera); REF_OFF(HIPRTRenderData, prev_camera);
    REF_OFF(HIPRTRenderData, cpu_only);
    REF_OFF(RenderBuffers, pixels); REF_OFF(RenderBuffers, triangles_indices); REF_OFF(RenderBuffers, vertices_positions);
    REF_OFF(RenderBuffers, has_vertex_normals); REF_OFF(RenderBuffers, vertex_normals); REF_OFF(RenderBuffers, texcoords);
    REF_OFF(RenderBuffers, material_indices); REF_OFF(RenderBuffers, materials_buffer); REF_OFF(RenderBuffers, emissive_triangles_count);
    REF_OFF(RenderBuffers, emissive_triangles_indices); REF_OFF(RenderBuffers, material_textures); REF_OFF(RenderBuffers, textures_dims);
    REF_OFF(BRDFsData, white_furnace_mode); REF_OFF(BRDFsData, white_furnace_mode_turn_off_emissives);
    REF_OFF(BRDFsData, sheen_ltc_parameters_texture); REF_OFF(BRDFsData, GGX_conductor_Ess); REF_OFF(BRDFsData, glossy_dielectric_Ess);
    REF_OFF(BRDFsData, GGX_Ess_glass); REF_OFF(BRDFsData, GGX_Ess_glass_inverse); REF_OFF(BRDFsData, GGX_Ess_thin_glass);
    REF_OFF(BRDFsData, use_hardware_tex_interpolation); REF_OFF(BRDFsData, clearcoat_compensation_approximation);
    REF_OFF(BRDFsData, GGX_masking_shadowing);
    REF_OFF(AuxiliaryBuffers, pixel_active); REF_OFF(AuxiliaryBuffers, denoiser_normals); REF_OFF(AuxiliaryBuffers, denoiser_albedo);
    REF_OFF(AuxiliaryBuffers, pixel_sample_count); REF_OFF(AuxiliaryBuffers, pixel_squared_luminance);
    REF_OFF(AuxiliaryBuffers, pixel_converged_sample_count); REF_OFF(AuxiliaryBuffers, still_one_ray_active);
    REF_OFF(AuxiliaryBuffers, stop_noise_threshold_converged_count); REF_OFF(AuxiliaryBuffers, restir_reservoir_buffer_1);
    REF_OFF(AuxiliaryBuffers, restir_reservoir_buffer_2); REF_OFF(AuxiliaryBuffers, restir_reservoir_buffer_3);
    REF_OFF(GBuffer, materials); REF_OFF(GBuffer, first_hit_prim_index); REF_OFF(GBuffer, shading_normals);
    REF_OFF(GBuffer, geometric_normals); REF_OFF(GBuffer, view_directions); REF_OFF(GBuffer, first_hits);
    REF_OFF(GBuffer, camera_ray_hit); REF_OFF(GBuffer, ray_volume_states);
    REF_OFF(HIPRTRenderSettings, need_to_reset); REF_OFF(HIPRTRenderSettings, do_update_status_buffers);
    REF_OFF(HIPRTRenderSettings, accumulate); REF_OFF(HIPRTRenderSettings, denoiser_AOV_accumulation_counter);
    REF_OFF(HIPRTRenderSettings, sample_number); REF_OFF(HIPRTRenderSettings, samples_per_frame); REF_OFF(HIPRTRenderSettings, nb_bounces);
    REF_OFF(HIPRTRenderSettings, use_russian_roulette); REF_OFF(HIPRTRenderSettings, russian_roulette_min_depth);
    REF_OFF(HIPRTRenderSettings, russian_roulette_throughput_clamp); REF_OFF(HIPRTRenderSettings, path_russian_roulette_method);
    REF_OFF(HIPRTRenderSettings, freeze_random); REF_OFF(HIPRTRenderSettings, display_NaNs);
    REF_OFF(HIPRTRenderSettings, allow_render_low_resolution); REF_OFF(HIPRTRenderSettings, wants_render_low_resolution);
    REF_OFF(HIPRTRenderSettings, render_low_resolution_scaling); REF_OFF(HIPRTRenderSettings, enable_adaptive_sampling);
    REF_OFF(HIPRTRenderSettings, adaptive_sampling_min_samples); REF_OFF(HIPRTRenderSettings, adaptive_sampling_noise_threshold);
    REF_OFF(HIPRTRenderSettings, enable_pixel_stop_noise_threshold); REF_OFF(HIPRTRenderSettings, stop_pixel_percentage_converged);
    REF_OFF(HIPRTRenderSettings, stop_pixel_noise_threshold); REF_OFF(HIPRTRenderSettings, direct_contribution_clamp);
    REF_OFF(HIPRTRenderSettings, envmap_contribution_clamp); REF_OFF(HIPRTRenderSettings, indirect_contribution_clamp);
    REF_OFF(HIPRTRenderSettings, minimum_light_contribution); REF_OFF(HIPRTRenderSettings, number_of_light_samples);
    REF_OFF(HIPRTRenderSettings, do_alpha_testing); REF_OFF(HIPRTRenderSettings, ris_settings); REF_OFF(HIPRTRenderSettings, restir_di_settings);
    REF_OFF(ReSTIRDISettings, initial_candidates); REF_OFF(ReSTIRDISettings, temporal_pass); REF_OFF(ReSTIRDISettings, spatial_pass);
    REF_OFF(ReSTIRDISettings, light_presampling); REF_OFF(ReSTIRDISettings, do_fused_spatiotemporal); REF_OFF(ReSTIRDISettings, m_cap);
    REF_OFF(ReSTIRDISettings, use_confidence_weights); REF_OFF(ReSTIRDISettings, use_normal_similarity_heuristic);
    REF_OFF(ReSTIRDISettings, normal_similarity_angle_degrees); REF_OFF(ReSTIRDISettings, normal_similarity_angle_precomp);
    REF_OFF(ReSTIRDISettings, use_plane_distance_heuristic); REF_OFF(ReSTIRDISettings, plane_distance_threshold);
    REF_OFF(ReSTIRDISettings, use_roughness_similarity_heuristic); REF_OFF(ReSTIRDISettings, roughness_similarity_threshold);
    REF_OFF(ReSTIRDISettings, do_final_shading_visibility); REF_OFF(ReSTIRDISettings, restir_output_reservoirs);
    REF_OFF(InitialCandidatesSettings, number_of_initial_light_candidates); REF_OFF(InitialCandidatesSettings, number_of_initial_bsdf_candidates);
    REF_OFF(InitialCandidatesSettings, envmap_candidate_probability); REF_OFF(InitialCandidatesSettings, output_reservoirs);
    REF_OFF(TemporalPassSettings, do_temporal_reuse_pass); REF_OFF(TemporalPassSettings, use_permutation_sampling);
    REF_OFF(TemporalPassSettings, permutation_sampling_random_bits); REF_OFF(TemporalPassSettings, max_neighbor_search_count);
    REF_OFF(TemporalPassSettings, neighbor_search_radius); REF_OFF(TemporalPassSettings, temporal_buffer_clear_requested);
    REF_OFF(TemporalPassSettings, input_reservoirs); REF_OFF(TemporalPassSettings, output_reservoirs);
    REF_OFF(SpatialPassSettings, do_spatial_reuse_pass); REF_OFF(SpatialPassSettings, spatial_pass_index);
    REF_OFF(SpatialPassSettings, number_of_passes); REF_OFF(SpatialPassSettings, reuse_radius); REF_OFF(SpatialPassSettings, reuse_neighbor_count);
    REF_OFF(SpatialPassSettings, do_disocclusion_reuse_boost); REF_OFF(SpatialPassSettings, disocclusion_reuse_count);
    REF_OFF(SpatialPassSettings, debug_neighbor_location); REF_OFF(SpatialPassSettings, do_neighbor_rotation);
    REF_OFF(SpatialPassSettings, allow_converged_neighbors_reuse); REF_OFF(SpatialPassSettings, converged_neighbor_reuse_probability);
    REF_OFF(SpatialPassSettings, do_visibility_only_last_pass); REF_OFF(SpatialPassSettings, neighbor_visibility_count);
    REF_OFF(SpatialPassSettings, input_reservoirs); REF_OFF(SpatialPassSettings, output_reservoirs);
    REF_OFF(LightPresamplingSettings, number_of_subsets); REF_OFF(LightPresamplingSettings, subset_size);
    REF_OFF(LightPresamplingSettings, tile_size); REF_OFF(LightPresamplingSettings, light_samples);
    REF_OFF(WorldSettings, ambient_light_type); REF_OFF(WorldSettings, uniform_light_color); REF_OFF(WorldSettings, envmap_width);
    REF_OFF(WorldSettings, envmap_height); REF_OFF(WorldSettings, envmap_intensity); REF_OFF(WorldSettings, envmap_scale_background_intensity);
    REF_OFF(WorldSettings, envmap); REF_OFF(WorldSettings, envmap_total_sum); REF_OFF(WorldSettings, envmap_cdf);
    REF_OFF(WorldSettings, alias_table_alias); REF_OFF(WorldSettings, alias_table_probas);
    REF_OFF(WorldSettings, envmap_to_world_matrix); REF_OFF(WorldSettings, world_to_envmap_matrix);
    REF_OFF(HIPRTCamera, inverse_view); REF_OFF(HIPRTCamera, inverse_projection); REF_OFF(HIPRTCamera, view_projection);
    REF_OFF(HIPRTCamera, do_jittering);
    REF_OFF(SimplifiedRendererMaterial, emissive_texture_used); REF_OFF(SimplifiedRendererMaterial, emission_strength);
    REF_OFF(SimplifiedRendererMaterial, base_color); REF_OFF(SimplifiedRendererMaterial, roughness);
    REF_OFF(SimplifiedRendererMaterial, oren_nayar_sigma); REF_OFF(SimplifiedRendererMaterial, metallic);
    REF_OFF(SimplifiedRendererMaterial, metallic_F90_falloff_exponent); REF_OFF(SimplifiedRendererMaterial, metallic_F82);
    REF_OFF(SimplifiedRendererMaterial, metallic_F90); REF_OFF(SimplifiedRendererMaterial, anisotropy);
    REF_OFF(SimplifiedRendererMaterial, anisotropy_rotation); REF_OFF(SimplifiedRendererMaterial, second_roughness_weight);
    REF_OFF(SimplifiedRendererMaterial, second_roughness); REF_OFF(SimplifiedRendererMaterial, specular);
    REF_OFF(SimplifiedRendererMaterial, specular_tint); REF_OFF(SimplifiedRendererMaterial, specular_color);
    REF_OFF(SimplifiedRendererMaterial, specular_darkening); REF_OFF(SimplifiedRendererMaterial, coat);
    REF_OFF(SimplifiedRendererMaterial, coat_medium_absorption); REF_OFF(SimplifiedRendererMaterial, coat_medium_thickness);
    REF_OFF(SimplifiedRendererMaterial, coat_roughness); REF_OFF(SimplifiedRendererMaterial, coat_roughening);
    REF_OFF(SimplifiedRendererMaterial, coat_darkening); REF_OFF(SimplifiedRendererMaterial, coat_anisotropy);
    REF_OFF(SimplifiedRendererMaterial, coat_anisotropy_rotation); REF_OFF(SimplifiedRendererMaterial, coat_ior);
    REF_OFF(SimplifiedRendererMaterial, sheen); REF_OFF(SimplifiedRendererMaterial, sheen_roughness);
    REF_OFF(SimplifiedRendererMaterial, sheen_color); REF_OFF(SimplifiedRendererMaterial, ior);
    REF_OFF(SimplifiedRendererMaterial, specular_transmission); REF_OFF(SimplifiedRendererMaterial, absorption_at_distance);
    REF_OFF(SimplifiedRendererMaterial, absorption_color); REF_OFF(SimplifiedRendererMaterial, dispersion_scale);
    REF_OFF(SimplifiedRendererMaterial, dispersion_abbe_number); REF_OFF(SimplifiedRendererMaterial, thin_walled);
    REF_OFF(SimplifiedRendererMaterial, thin_film); REF_OFF(SimplifiedRendererMaterial, thin_film_ior);
    REF_OFF(SimplifiedRendererMaterial, thin_film_thickness); REF_OFF(SimplifiedRendererMaterial, thin_film_kappa_3);
    REF_OFF(SimplifiedRendererMaterial, thin_film_hue_shift_degrees); REF_OFF(SimplifiedRendererMaterial, thin_film_base_ior_override);
    REF_OFF(SimplifiedRendererMaterial, thin_film_do_ior_override); REF_OFF(SimplifiedRendererMaterial, srgb);
    REF_OFF(SimplifiedRendererMaterial, alpha_opacity); REF_OFF(SimplifiedRendererMaterial, dielectric_priority);
    REF_OFF(SimplifiedRendererMaterial, energy_preservation_monte_carlo_samples);
    REF_OFF(SimplifiedRendererMaterial, enforce_strong_energy_conservation);
    REF_OFF(RendererMaterial, normal_map_texture_index); REF_OFF(RendererMaterial, emission_texture_index);
    REF_OFF(RendererMaterial, base_color_texture_index); REF_OFF(RendererMaterial, roughness_metallic_texture_index);
    REF_OFF(RendererMaterial, roughness_texture_index); REF_OFF(RendererMaterial, oren_sigma_texture_index);
    REF_OFF(RendererMaterial, metallic_texture_index); REF_OFF(RendererMaterial, specular_texture_index);
    REF_OFF(RendererMaterial, specular_tint_texture_index); REF_OFF(RendererMaterial, specular_color_texture_index);
    REF_OFF(RendererMaterial, anisotropic_texture_index); REF_OFF(RendererMaterial, anisotropic_rotation_texture_index);
    REF_OFF(RendererMaterial, coat_texture_index); REF_OFF(RendererMaterial, coat_roughness_texture_index);
    REF_OFF(RendererMaterial, coat_ior_texture_index); REF_OFF(RendererMaterial, sheen_texture_index);
    REF_OFF(RendererMaterial, sheen_roughness_texture_index); REF_OFF(RendererMaterial, sheen_color_texture_index);
    REF_OFF(RendererMaterial, specular_transmission_texture_index);
    REF_OFF(RayVolumeState, distance_in_volume); REF_OFF(RayVolumeState, interior_stack); REF_OFF(RayVolumeState, incident_mat_index);
    REF_OFF(RayVolumeState, outgoing_mat_index); REF_OFF(RayVolumeState, inside_material); REF_OFF(RayVolumeState, sampled_wavelength);
    REF_OFF(ReSTIRDIReservoir, UCW); REF_OFF(ReSTIRDIReservoir, M); REF_OFF(ReSTIRDIReservoir, weight_sum); REF_OFF(ReSTIRDIReservoir, sample);
    REF_OFF(ReSTIRDISample, point_on_light_source); REF_OFF(ReSTIRDISample, emissive_triangle_index);
    REF_OFF(ReSTIRDISample, target_function); REF_OFF(ReSTIRDISample, flags);
    REF_OFF(ReSTIRDIPresampledLight, point_on_light_source); REF_OFF(ReSTIRDIPresampledLight, light_source_normal);
    REF_OFF(ReSTIRDIPresampledLight, emissive_triangle_index); REF_OFF(ReSTIRDIPresampledLight, pdf);
    REF_OFF(ReSTIRDIPresampledLight, radiance); REF_OFF(ReSTIRDIPresampledLight, flags);
    return -1;
#pragma GCC diagnostic pop
}

const char* ref_variant()
{
    static std::string s;
    s = "LSS=" + std::to_string(DirectLightSamplingStrategy) + ";ESS=" + std::to_string(EnvmapSamplingStrategy)
        + ";GGX=" + std::to_string(PrincipledBSDFAnisotropicGGXSampleFunction) + ";MS=" + std::to_string(PrincipledBSDFGGXUseMultipleScattering)
        + ";DIFFUSE=" + std::to_string(PrincipledBSDFDiffuseLobe) + ";ENVMIS=" + std::to_string(EnvmapSamplingDoBSDFMIS)
        + ";PRESAMPLE=" + std::to_string(ReSTIR_DI_DoLightsPresampling) + ";BIAS=" + std::to_string(ReSTIR_DI_BiasCorrectionWeights)
        + ";LATER=" + std::to_string(ReSTIR_DI_LaterBouncesSamplingStrategy);
    return s.c_str();
}

int ref_num_threads() { return omp_get_max_threads(); }
void ref_set_num_threads(int n) { omp_set_num_threads(n); }

void* ref_create()
{
    RefCtx* c = new RefCtx();
    c->stop_noise_threshold_count.store(0);
    return c;
}

void ref_destroy(void* h) { delete reinterpret_cast<RefCtx*>(h); }

// Default-constructed structs of the reference, for the host mirror to compare its defaults against
void ref_default_render_settings(void* out) { HIPRTRenderSettings s; std::memcpy(out, &s, sizeof(s)); }
void ref_default_world_settings(void* out) { WorldSettings s; std::memcpy(out, &s, sizeof(s)); }
void ref_default_material(void* out) { RendererMaterial m; std::memcpy(out, &m, sizeof(m)); }
void ref_default_bsdfs_data(void* out) { BRDFsData m; std::memcpy(out, &m, sizeof(m)); }
void ref_default_volume_state(void* out) { RayVolumeState m; std::memset(out, 0, sizeof(m)); std::memcpy(out, &m, sizeof(m)); }
void ref_material_make_safe(void* mat) { RendererMaterial* m = reinterpret_cast<RendererMaterial*>(mat); m->make_safe(); m->precompute_properties(); }

void ref_sheen_ltc_table(float* out_32x32x3)
{
    std::memcpy(out_32x32x3, ltc_parameters_table_approximation.data(), 32 * 32 * 3 * sizeof(float));
}

int ref_load_hdr(const char* path, int channels, int flip, float* out, int capacity, int* w, int* h)
{
    Image32Bit img = Image32Bit::read_image_hdr(path, channels, flip != 0);
    *w = img.width; *h = img.height;
    if (img.width == 0) return -1;
    size_t n = (size_t)img.width * img.height * channels;
    if ((size_t)capacity < n) return -2;
    std::memcpy(out, img.data().data(), n * sizeof(float));
    return 0;
}

void ref_set_luts(void* h, const float* sheen, const float* conductor, const float* glossy, const float* glass, const float* glass_inv, const float* thin_glass)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    // src/Renderer/CPURenderer.cpp:93-132 (setup_brdfs_data) with the slices handed in as arrays
    c->sheen_ltc = make_image(sheen, 32, 32, 3);
    c->conductor_ess = make_image(conductor, GPUBakerConstants::GGX_CONDUCTOR_ESS_TEXTURE_SIZE_COS_THETA_O, GPUBakerConstants::GGX_CONDUCTOR_ESS_TEXTURE_SIZE_ROUGHNESS, 1);
    c->glossy_ess = make_image3d(glossy, GPUBakerConstants::GLOSSY_DIELECTRIC_TEXTURE_SIZE_COS_THETA_O, GPUBakerConstants::GLOSSY_DIELECTRIC_TEXTURE_SIZE_ROUGHNESS, GPUBakerConstants::GLOSSY_DIELECTRIC_TEXTURE_SIZE_IOR);
    c->glass_ess = make_image3d(glass, GPUBakerConstants::GGX_GLASS_ESS_TEXTURE_SIZE_COS_THETA_O, GPUBakerConstants::GGX_GLASS_ESS_TEXTURE_SIZE_ROUGHNESS, GPUBakerConstants::GGX_GLASS_ESS_TEXTURE_SIZE_IOR);
    c->glass_inv_ess = make_image3d(glass_inv, GPUBakerConstants::GGX_GLASS_ESS_TEXTURE_SIZE_COS_THETA_O, GPUBakerConstants::GGX_GLASS_ESS_TEXTURE_SIZE_ROUGHNESS, GPUBakerConstants::GGX_GLASS_ESS_TEXTURE_SIZE_IOR);
    c->thin_glass_ess = make_image3d(thin_glass, GPUBakerConstants::GGX_THIN_GLASS_ESS_TEXTURE_SIZE_COS_THETA_O, GPUBakerConstants::GGX_THIN_GLASS_ESS_TEXTURE_SIZE_ROUGHNESS, GPUBakerConstants::GGX_THIN_GLASS_ESS_TEXTURE_SIZE_IOR);

    c->rd.bsdfs_data.sheen_ltc_parameters_texture = &c->sheen_ltc;
    c->rd.bsdfs_data.GGX_conductor_Ess = &c->conductor_ess;
    c->rd.bsdfs_data.glossy_dielectric_Ess = &c->glossy_ess;
    c->rd.bsdfs_data.GGX_Ess_glass = &c->glass_ess;
    c->rd.bsdfs_data.GGX_Ess_glass_inverse = &c->glass_inv_ess;
    c->rd.bsdfs_data.GGX_Ess_thin_glass = &c->thin_glass_ess;
}

// 8-bit material textures (Scene::textures, src/Scene/SceneParser.h:80-131). Call before ref_set_scene.
void ref_add_texture(void* h, const unsigned char* texels, int w, int hgt, int channels)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    std::vector<unsigned char> v(texels, texels + (size_t)w * hgt * channels);
    c->textures.push_back(Image8Bit(v, w, hgt, channels));
    c->textures_dims.push_back(make_int2(w, hgt));
}

double ref_set_scene(void* h, const float* verts, int nverts, const int* indices, int ntris, const float* normals, const unsigned char* has_normals,
                     const float* texcoords, const int* material_indices, const void* materials, int nmats, const int* emissive, int nemissive)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    c->vertices.assign(reinterpret_cast<const float3*>(verts), reinterpret_cast<const float3*>(verts) + nverts);
    c->normals.assign(reinterpret_cast<const float3*>(normals), reinterpret_cast<const float3*>(normals) + nverts);
    c->texcoords.assign(reinterpret_cast<const float2*>(texcoords), reinterpret_cast<const float2*>(texcoords) + nverts);
    c->has_normals.assign(has_normals, has_normals + nverts);
    c->indices.assign(indices, indices + 3 * (size_t)ntris);
    c->material_indices.assign(material_indices, material_indices + ntris);
    c->materials.resize(nmats);
    std::memcpy((void*)c->materials.data(), materials, (size_t)nmats * sizeof(RendererMaterial));
    c->emissive.assign(emissive, emissive + nemissive);

    // src/Renderer/CPURenderer.cpp:134-200 (set_scene)
    HIPRTRenderData& rd = c->rd;
    rd.geom = nullptr;
    rd.buffers.materials_buffer = c->materials.data();
    rd.buffers.material_indices = c->material_indices.data();
    rd.buffers.has_vertex_normals = c->has_normals.data();
    rd.buffers.triangles_indices = c->indices.data();
    rd.buffers.vertices_positions = c->vertices.data();
    rd.buffers.vertex_normals = c->normals.data();
    rd.buffers.texcoords = c->texcoords.data();
    rd.buffers.material_textures = c->textures.data();
    rd.buffers.textures_dims = c->textures_dims.data();
    rd.buffers.emissive_triangles_count = nemissive;
    rd.buffers.emissive_triangles_indices = c->emissive.data();

    // Scene::get_triangles, src/Scene/SceneParser.h:118-130
    c->triangles.clear();
    c->triangles.reserve(ntris);
    for (size_t i = 0; i < 3 * (size_t)ntris; i += 3)
        c->triangles.push_back(Triangle(c->vertices[c->indices[i + 0]], c->vertices[c->indices[i + 1]], c->vertices[c->indices[i + 2]]));

    auto t0 = std::chrono::steady_clock::now();
    c->bvh = std::make_shared<BVH>(&c->triangles);
    rd.cpu_only.bvh = c->bvh.get();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void ref_set_envmap(void* h, const float* texels, int w, int hgt, int channels)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    // src/Renderer/CPURenderer.cpp:202-241 (set_envmap)
    if (w == 0 || hgt == 0)
    {
        c->rd.world_settings.ambient_light_type = AmbientLightType::UNIFORM;
        c->rd.world_settings.uniform_light_color = ColorRGB32F(0.5f, 0.5f, 0.5f);
        return;
    }
    c->envmap = make_image(texels, w, hgt, channels);
    if (EnvmapSamplingStrategy == ESS_BINARY_SEARCH)
    {
        c->envmap_cdf = c->envmap.compute_cdf();
        c->rd.world_settings.envmap_total_sum = c->envmap_cdf.back();
        c->rd.world_settings.envmap_cdf = c->envmap_cdf.data();
    }
    else
    {
        // Also done for ESS_NO_SAMPLING so that the alias-table builder can always be exercised
        float total_sum;
        c->envmap.compute_alias_table(c->alias_probas, c->alias_alias, &total_sum);
        c->rd.world_settings.envmap_total_sum = total_sum;
        c->rd.world_settings.alias_table_probas = c->alias_probas.data();
        c->rd.world_settings.alias_table_alias = c->alias_alias.data();
    }
    c->rd.world_settings.envmap = &c->envmap;
    c->rd.world_settings.envmap_width = w;
    c->rd.world_settings.envmap_height = hgt;
    c->rd.world_settings.ambient_light_type = AmbientLightType::ENVMAP;
}

int ref_get_alias_table(void* h, float* probas, int* alias, float* total_sum)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    std::memcpy(probas, c->alias_probas.data(), c->alias_probas.size() * sizeof(float));
    std::memcpy(alias, c->alias_alias.data(), c->alias_alias.size() * sizeof(int));
    *total_sum = c->rd.world_settings.envmap_total_sum;
    return (int)c->alias_probas.size();
}

void ref_set_camera(void* h, const void* hiprt_camera)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    std::memcpy(&c->rd.current_camera, hiprt_camera, sizeof(HIPRTCamera));
    c->rd.prev_camera = c->rd.current_camera;
}

// Raw access so that the Python side can poke run-time settings through the same struct mirrors it uses for the CUDA library.
void* ref_render_data(void* h) { return &reinterpret_cast<RefCtx*>(h)->rd; }
void* ref_render_settings(void* h) { return &reinterpret_cast<RefCtx*>(h)->rd.render_settings; }
void* ref_world_settings(void* h) { return &reinterpret_cast<RefCtx*>(h)->rd.world_settings; }
void* ref_bsdfs_data(void* h) { return &reinterpret_cast<RefCtx*>(h)->rd.bsdfs_data; }

void ref_resize(void* h, int w, int hgt)
{
    RefCtx* c = reinterpret_cast<RefCtx*>(h);
    // src/Renderer/CPURenderer.cpp:68-91 (constructor)
    size_t n = (size_t)w * hgt;
    c->res = make_int2(w, hgt);
    c->pixels.assign(n, ColorRGB32F(0.0f));
    c->pixel_active.assign(n, 0);
    c->denoiser_albedo.assign(n, ColorRGB32F(0.0f));
    c->denoiser_normals.assign(n, float3{0.0f, 0.0f, 0.0f});
    c->pixel_sample_count.assign(n, 0);
    c->pixel_converged_sample_count.assign(n, 0);
    c->pixel_squared_luminance.assign(n, 0.0f);
    c->initial_candidates_reservoirs.assign(n, ReSTIRDIReservoir());
    c->spatial_output_1.assign(n, ReSTIRDIReservoir());
    c->spatial_output_2.assign(n, ReSTIRDIReservoir());
    c->presampled_lights.assign(std::max(n, (size_t)c->rd.render_settings.restir_di_settings.light_presampling.number_of_subsets * c->rd.render_settings.restir_di_settings.light_presampling.subset_size), ReSTIRDIPresampledLight());
    c->output_reservoirs = c->spatial_output_1.data();
    c->g_buffer.resize(n);
    c->g_buffer_prev.resize(n);
    c->rng = Xorshift32Generator(42);
    c->odd_frame = false;

    HIPRTRenderData& rd = c->rd;
    rd.random_seed = 42;
    rd.render_settings.sample_number = 0;
    rd.render_settings.need_to_reset = true;
    rd.render_settings.denoiser_AOV_accumulation_counter = 0;
    rd.buffers.pixels = c->pixels.data();
    rd.aux_buffers.pixel_active = c->pixel_active.data();
    rd.aux_buffers.denoiser_albedo = c->denoiser_albedo.data();
    rd.aux_buffers.denoiser_normals = c->denoiser_normals.data();
    rd.aux_buffers.pixel_sample_count = c->pixel_sample_count.data();
    rd.aux_buffers.pixel_converged_sample_count = c->pixel_converged_sample_count.data();
    rd.aux_buffers.pixel_squared_luminance = c->pixel_squared_luminance.data();
    rd.aux_buffers.still_one_ray_active = &c->still_one_ray_active;
    rd.aux_buffers.stop_noise_threshold_converged_count = &c->stop_noise_threshold_count;

    auto bind = [](GBuffer& g, CPURendererGBuffer& s) {
        g.materials = s.materials.data();
        g.geometric_normals = s.geometric_normals.data();
        g.shading_normals = s.shading_normals.data();
        g.view_directions = s.view_directions.data();
        g.first_hits = s.first_hits.data();
        g.first_hit_prim_index = s.first_hit_prim_index.data();
        g.camera_ray_hit = s.cameray_ray_hit.data();
        g.ray_volume_states = s.ray_volume_states.data();
    };
    bind(rd.g_buffer, c->g_buffer);
    bind(rd.g_buffer_prev_frame, c->g_buffer_prev);

    ReSTIRDISettings& rs = rd.render_settings.restir_di_settings;
    rs.light_presampling.light_samples = c->presampled_lights.data();
    rs.initial_candidates.output_reservoirs = c->initial_candidates_reservoirs.data();
    rs.restir_output_reservoirs = c->spatial_output_1.data();
    rd.aux_buffers.restir_reservoir_buffer_1 = c->initial_candidates_reservoirs.data();
    rd.aux_buffers.restir_reservoir_buffer_2 = c->spatial_output_1.data();
    rd.aux_buffers.restir_reservoir_buffer_3 = c->spatial_output_2.data();
}

} // extern "C"

#include "ref_harness_passes.inl"
