/*
 * b200pt_types.h -- byte-compatible mirrors of the structs that cross the reference renderer's kernel
 * launch boundary (reference: src/HostDeviceCommon/*.h, src/Device/includes/{GBuffer,RayVolumeState}.h,
 * src/Device/includes/ReSTIR/DI/Reservoir.h, src/Device/kernel_parameters/ReSTIR/DI/LightPresamplingParameters.h).
 *
 * The reference passes `HIPRTRenderData` BY VALUE to every render kernel (src/Compiler/GPUKernel.cpp:169-178,
 * src/Renderer/GPURenderer.cpp:468,482); the C-ABI of b200pt.h takes a pointer to the same bytes and snapshots
 * them at call time. Layout facts below were measured by compiling the reference headers with g++ 13
 * (oracle/ref_harness.cpp: ref_sizeof / ref_offsetof) for the default layout-relevant kernel options
 * InteriorStackStrategy = ISS_WITH_PRIORITIES and NestedDielectricsStackSize = 3 (KernelOptions.h:144-152);
 * tests/test_abi_layout.py checks every offset of this header against the reference build.
 *
 * A maintainer compiling inside the reference tree defines B200PT_USE_REFERENCE_HEADERS and includes the
 * reference's own headers before this file: the mirrors are then skipped and the reference types are used.
 *
 * Pointer fields are DEVICE pointers when a struct is handed to b200pt_launch_*; fields documented
 * "b200pt_tex*" carry texture descriptors created with b200pt_tex2d_create / b200pt_tex3d_create where the
 * reference carried oroTextureObject_t or Image32Bit* (src/Device/includes/Texture.h:62,189).
 */
#ifndef B200PT_TYPES_H
#define B200PT_TYPES_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
#define B200PT_DEFAULT(...) = __VA_ARGS__
#else
#define B200PT_DEFAULT(...)
#endif

#ifndef B200PT_USE_REFERENCE_HEADERS

/* hiprtFloat2/3, hiprtInt2 (thirdparties/HIPRT-Fork/hiprt/hiprt_vec.h:30-58): unaligned PODs */
typedef struct { float x, y; } b200pt_float2;
typedef struct { float x, y, z; } b200pt_float3;
typedef struct { int x, y; } b200pt_int2;
/* float4x4 (src/HostDeviceCommon/Math.h:40-43): row-major, applied as M * column-vector */
typedef struct { float m[4][4]; } b200pt_float4x4;
/* ColorRGB32F (src/HostDeviceCommon/Color.h:64-110) */
typedef struct { float r, g, b; } ColorRGB32F;

/* ---- materials: src/HostDeviceCommon/Material.h:29-268 ---------------------------------------- */
enum { B200PT_NO_TEXTURE = -1, B200PT_CONSTANT_EMISSIVE_TEXTURE = -2 };

typedef struct SimplifiedRendererMaterial
{
    uint8_t emissive_texture_used B200PT_DEFAULT(0);             /*   0 */
    uint8_t _pad0[3] B200PT_DEFAULT({0});
    float emission_strength B200PT_DEFAULT(1.0f);                /*   4 */
    ColorRGB32F base_color B200PT_DEFAULT({1.0f, 1.0f, 1.0f}); /* 8 */
    float roughness B200PT_DEFAULT(0.3f);                        /*  20 */
    float oren_nayar_sigma B200PT_DEFAULT(0.34906585039886591538f);
    float metallic B200PT_DEFAULT(0.0f);                         /*  28 */
    float metallic_F90_falloff_exponent B200PT_DEFAULT(5.0f);
    ColorRGB32F metallic_F82 B200PT_DEFAULT({1.0f, 1.0f, 1.0f}); /* 36 */
    ColorRGB32F metallic_F90 B200PT_DEFAULT({1.0f, 1.0f, 1.0f}); /* 48 */
    float anisotropy B200PT_DEFAULT(0.0f);                       /*  60 */
    float anisotropy_rotation B200PT_DEFAULT(0.0f);
    float second_roughness_weight B200PT_DEFAULT(0.0f);
    float second_roughness B200PT_DEFAULT(0.5f);                 /*  72 */
    float specular B200PT_DEFAULT(1.0f);                         /*  76 */
    float specular_tint B200PT_DEFAULT(1.0f);
    ColorRGB32F specular_color B200PT_DEFAULT({1.0f, 1.0f, 1.0f}); /* 84 */
    float specular_darkening B200PT_DEFAULT(0.0f);               /*  96 */
    float coat B200PT_DEFAULT(0.0f);                             /* 100 */
    ColorRGB32F coat_medium_absorption B200PT_DEFAULT({1.0f, 1.0f, 1.0f}); /* 104 */
    float coat_medium_thickness B200PT_DEFAULT(5.0f);            /* 116 */
    float coat_roughness B200PT_DEFAULT(0.0f);
    float coat_roughening B200PT_DEFAULT(1.0f);
    float coat_darkening B200PT_DEFAULT(1.0f);
    float coat_anisotropy B200PT_DEFAULT(0.0f);
    float coat_anisotropy_rotation B200PT_DEFAULT(0.0f);
    float coat_ior B200PT_DEFAULT(1.5f);                         /* 140 */
    float sheen B200PT_DEFAULT(0.0f);                            /* 144 */
    float sheen_roughness B200PT_DEFAULT(0.5f);
    ColorRGB32F sheen_color B200PT_DEFAULT({1.0f, 1.0f, 1.0f}); /* 152 */
    float ior B200PT_DEFAULT(1.40f);                             /* 164 */
    float specular_transmission B200PT_DEFAULT(0.0f);            /* 168 */
    float absorption_at_distance B200PT_DEFAULT(1.0f);
    ColorRGB32F absorption_color B200PT_DEFAULT({1.0f, 1.0f, 1.0f}); /* 176 */
    float dispersion_scale B200PT_DEFAULT(0.0f);                 /* 188 */
    float dispersion_abbe_number B200PT_DEFAULT(20.0f);
    uint8_t thin_walled B200PT_DEFAULT(0);                       /* 196 */
    uint8_t _pad1[3] B200PT_DEFAULT({0});
    float thin_film B200PT_DEFAULT(0.0f);                        /* 200 */
    float thin_film_ior B200PT_DEFAULT(1.3f);
    float thin_film_thickness B200PT_DEFAULT(500.0f);
    float thin_film_kappa_3 B200PT_DEFAULT(0.0f);
    float thin_film_hue_shift_degrees B200PT_DEFAULT(0.0f);
    float thin_film_base_ior_override B200PT_DEFAULT(1.0f);      /* 220 */
    uint8_t thin_film_do_ior_override B200PT_DEFAULT(0);         /* 224 */
    uint8_t srgb B200PT_DEFAULT(1);                              /* 225 */
    uint8_t _pad2[2] B200PT_DEFAULT({0});
    float alpha_opacity B200PT_DEFAULT(1.0f);                    /* 228 */
    int dielectric_priority B200PT_DEFAULT(0);                   /* 232 */
    int energy_preservation_monte_carlo_samples B200PT_DEFAULT(12);
    uint8_t enforce_strong_energy_conservation B200PT_DEFAULT(0);/* 240 */
    uint8_t _pad3[3] B200PT_DEFAULT({0});
    ColorRGB32F emission B200PT_DEFAULT({0.0f, 0.0f, 0.0f}); /* 244 (private in the reference; get_emission() = emission * emission_strength) */
} SimplifiedRendererMaterial;

typedef struct RendererMaterial
{
    SimplifiedRendererMaterial base;                             /*   0 (C++ base class in the reference) */
    int normal_map_texture_index B200PT_DEFAULT(-1);             /* 256 */
    int emission_texture_index B200PT_DEFAULT(-1);
    int base_color_texture_index B200PT_DEFAULT(-1);
    int roughness_metallic_texture_index B200PT_DEFAULT(-1);
    int roughness_texture_index B200PT_DEFAULT(-1);
    int oren_sigma_texture_index B200PT_DEFAULT(-1);
    int metallic_texture_index B200PT_DEFAULT(-1);
    int specular_texture_index B200PT_DEFAULT(-1);
    int specular_tint_texture_index B200PT_DEFAULT(-1);
    int specular_color_texture_index B200PT_DEFAULT(-1);
    int anisotropic_texture_index B200PT_DEFAULT(-1);
    int anisotropic_rotation_texture_index B200PT_DEFAULT(-1);
    int coat_texture_index B200PT_DEFAULT(-1);
    int coat_roughness_texture_index B200PT_DEFAULT(-1);
    int coat_ior_texture_index B200PT_DEFAULT(-1);
    int sheen_texture_index B200PT_DEFAULT(-1);
    int sheen_roughness_texture_index B200PT_DEFAULT(-1);
    int sheen_color_texture_index B200PT_DEFAULT(-1);
    int specular_transmission_texture_index B200PT_DEFAULT(-1);  /* 328 */
} RendererMaterial;

/* ---- nested dielectrics: src/Device/includes/NestedDielectrics.h:145-288, RayVolumeState.h:11-32 ----
 * StackPriorityEntry is a 32-bit bit-field word. g++ lays the reference's (swapped-width) declaration
 *   unsigned priority:1; bool odd_parity:4; bool topmost:1; unsigned material_index:26
 * out LSB-first: bit 0 priority, bits 1-4 odd_parity (only 0/1 ever stored), bit 5 topmost, bits 6-31 material index.
 * Priorities therefore truncate to one bit -- a reference quirk that is part of the data contract. */
#define B200PT_NESTED_DIELECTRICS_STACK_SIZE 3
#define B200PT_MAX_MATERIAL_INDEX ((1 << 26) - 1)
#define B200PT_DIELECTRIC_PRIORITY_MAXIMUM 15
typedef struct InteriorStack
{
    uint32_t stack[B200PT_NESTED_DIELECTRICS_STACK_SIZE];        /* default entry: priority 0, odd_parity 1, topmost 1, index MAX */
    int stack_position B200PT_DEFAULT(0);
} InteriorStack;

typedef struct RayVolumeState
{
    float distance_in_volume B200PT_DEFAULT(0.0f);               /*  0 */
    InteriorStack interior_stack;                                /*  4 */
    int incident_mat_index B200PT_DEFAULT(-1);                   /* 20 */
    int outgoing_mat_index B200PT_DEFAULT(-1);                   /* 24 */
    uint8_t inside_material B200PT_DEFAULT(0);                   /* 28 */
    uint8_t _pad0[3] B200PT_DEFAULT({0});
    float sampled_wavelength B200PT_DEFAULT(0.0f);               /* 32 */
} RayVolumeState;

/* ---- ReSTIR DI: src/Device/includes/ReSTIR/DI/Reservoir.h:16-165, PresampledLight.h ------------ */
enum { RESTIR_DI_FLAGS_NONE = 0, RESTIR_DI_FLAGS_ENVMAP_SAMPLE = 1 << 0, RESTIR_DI_FLAGS_UNOCCLUDED = 1 << 1 };

typedef struct ReSTIRDISample
{
    int emissive_triangle_index B200PT_DEFAULT(-1);              /*  0 */
    b200pt_float3 point_on_light_source B200PT_DEFAULT({0, 0, 0}); /* 4: envmap-space direction for envmap samples */
    float target_function B200PT_DEFAULT(0.0f);                  /* 16 */
    uint8_t flags B200PT_DEFAULT(0);                             /* 20 */
    uint8_t _pad0[3] B200PT_DEFAULT({0});
} ReSTIRDISample;

typedef struct ReSTIRDIReservoir
{
    int M B200PT_DEFAULT(0);                                     /*  0 */
    float weight_sum B200PT_DEFAULT(0.0f);                       /*  4 */
    float UCW B200PT_DEFAULT(0.0f);                              /*  8 */
    ReSTIRDISample sample;                                       /* 12 */
} ReSTIRDIReservoir;

typedef struct ReSTIRDIPresampledLight
{
    int emissive_triangle_index B200PT_DEFAULT(-1);              /*  0 */
    b200pt_float3 point_on_light_source;                         /*  4 */
    b200pt_float3 light_source_normal;                           /* 16 */
    ColorRGB32F radiance;                                        /* 28 */
    float pdf B200PT_DEFAULT(0.0f);                              /* 40 */
    uint8_t flags B200PT_DEFAULT(0);                             /* 44 */
    uint8_t _pad0[3] B200PT_DEFAULT({0});
} ReSTIRDIPresampledLight;

/* ---- src/HostDeviceCommon/RenderData.h:32-173 ------------------------------------------------- */
typedef struct RenderBuffers
{
    ColorRGB32F* pixels B200PT_DEFAULT(NULL);                    /*  0: SUM of samples, not the average (RenderData.h:34-36) */
    int* triangles_indices B200PT_DEFAULT(NULL);                 /*  8: 3 ints per triangle */
    b200pt_float3* vertices_positions B200PT_DEFAULT(NULL);      /* 16 */
    unsigned char* has_vertex_normals B200PT_DEFAULT(NULL);      /* 24: per vertex */
    b200pt_float3* vertex_normals B200PT_DEFAULT(NULL);          /* 32 */
    b200pt_float2* texcoords B200PT_DEFAULT(NULL);               /* 40 */
    int* material_indices B200PT_DEFAULT(NULL);                  /* 48: per triangle */
    RendererMaterial* materials_buffer B200PT_DEFAULT(NULL);     /* 56 */
    int emissive_triangles_count B200PT_DEFAULT(0);              /* 64 */
    int _pad0 B200PT_DEFAULT(0);
    int* emissive_triangles_indices B200PT_DEFAULT(NULL);        /* 72 */
    void* material_textures B200PT_DEFAULT(NULL);                /* 80: device array of b200pt_tex2d descriptors */
    b200pt_int2* textures_dims B200PT_DEFAULT(NULL);             /* 88 */
} RenderBuffers;

/* src/HostDeviceCommon/BSDFsData.h:18-63 */
typedef enum GGXMaskingShadowingFlavor { HeightCorrelated = 0, HeightUncorrelated = 1 } GGXMaskingShadowingFlavor;
typedef struct BRDFsData
{
    uint8_t white_furnace_mode B200PT_DEFAULT(0);                /*  0 */
    uint8_t white_furnace_mode_turn_off_emissives B200PT_DEFAULT(1);
    uint8_t _pad0[6] B200PT_DEFAULT({0});
    void* sheen_ltc_parameters_texture B200PT_DEFAULT(NULL);     /*  8: b200pt_tex2d, 32x32 RGB */
    void* GGX_conductor_Ess B200PT_DEFAULT(NULL);                /* 16: b200pt_tex2d, 128x128 */
    void* glossy_dielectric_Ess B200PT_DEFAULT(NULL);            /* 24: b200pt_tex3d, 128x64x128 */
    void* GGX_Ess_glass B200PT_DEFAULT(NULL);                    /* 32: b200pt_tex3d, 256x16x128 */
    void* GGX_Ess_glass_inverse B200PT_DEFAULT(NULL);            /* 40 */
    void* GGX_Ess_thin_glass B200PT_DEFAULT(NULL);               /* 48: b200pt_tex3d, 32x32x96 */
    uint8_t use_hardware_tex_interpolation B200PT_DEFAULT(0);    /* 56 */
    uint8_t clearcoat_compensation_approximation B200PT_DEFAULT(1);
    uint8_t _pad1[2] B200PT_DEFAULT({0});
    int GGX_masking_shadowing B200PT_DEFAULT(HeightCorrelated);  /* 60 */
} BRDFsData;

typedef struct AuxiliaryBuffers
{
    unsigned char* pixel_active B200PT_DEFAULT(NULL);            /*  0 */
    b200pt_float3* denoiser_normals B200PT_DEFAULT(NULL);        /*  8 */
    ColorRGB32F* denoiser_albedo B200PT_DEFAULT(NULL);           /* 16 */
    int* pixel_sample_count B200PT_DEFAULT(NULL);                /* 24 */
    float* pixel_squared_luminance B200PT_DEFAULT(NULL);         /* 32 */
    int* pixel_converged_sample_count B200PT_DEFAULT(NULL);      /* 40 */
    unsigned char* still_one_ray_active B200PT_DEFAULT(NULL);    /* 48 */
    unsigned int* stop_noise_threshold_converged_count B200PT_DEFAULT(NULL); /* 56 */
    ReSTIRDIReservoir* restir_reservoir_buffer_1 B200PT_DEFAULT(NULL);       /* 64 */
    ReSTIRDIReservoir* restir_reservoir_buffer_2 B200PT_DEFAULT(NULL);
    ReSTIRDIReservoir* restir_reservoir_buffer_3 B200PT_DEFAULT(NULL);
} AuxiliaryBuffers;

/* src/Device/includes/GBuffer.h:17-34 (structure of arrays, one element per pixel) */
typedef struct GBuffer
{
    SimplifiedRendererMaterial* materials B200PT_DEFAULT(NULL);  /*  0 */
    int* first_hit_prim_index B200PT_DEFAULT(NULL);              /*  8 */
    b200pt_float3* shading_normals B200PT_DEFAULT(NULL);         /* 16 */
    b200pt_float3* geometric_normals B200PT_DEFAULT(NULL);       /* 24 */
    b200pt_float3* view_directions B200PT_DEFAULT(NULL);         /* 32 */
    b200pt_float3* first_hits B200PT_DEFAULT(NULL);              /* 40 */
    unsigned char* camera_ray_hit B200PT_DEFAULT(NULL);          /* 48 */
    RayVolumeState* ray_volume_states B200PT_DEFAULT(NULL);      /* 56 */
} GBuffer;

/* ---- src/HostDeviceCommon/ReSTIRDISettings.h:13-195 ------------------------------------------- */
typedef struct InitialCandidatesSettings
{
    int number_of_initial_light_candidates B200PT_DEFAULT(4);
    int number_of_initial_bsdf_candidates B200PT_DEFAULT(1);
    float envmap_candidate_probability B200PT_DEFAULT(0.25f);
    int _pad0 B200PT_DEFAULT(0);
    ReSTIRDIReservoir* output_reservoirs B200PT_DEFAULT(NULL);   /* 16 */
} InitialCandidatesSettings;

typedef struct TemporalPassSettings
{
    uint8_t do_temporal_reuse_pass B200PT_DEFAULT(1);            /*  0 */
    uint8_t use_permutation_sampling B200PT_DEFAULT(0);
    uint8_t _pad0[2] B200PT_DEFAULT({0});
    int permutation_sampling_random_bits B200PT_DEFAULT(42);     /*  4 */
    int max_neighbor_search_count B200PT_DEFAULT(8);
    int neighbor_search_radius B200PT_DEFAULT(4);
    uint8_t temporal_buffer_clear_requested B200PT_DEFAULT(0);   /* 16 */
    uint8_t _pad1[7] B200PT_DEFAULT({0});
    ReSTIRDIReservoir* input_reservoirs B200PT_DEFAULT(NULL);    /* 24 */
    ReSTIRDIReservoir* output_reservoirs B200PT_DEFAULT(NULL);   /* 32 */
} TemporalPassSettings;

typedef struct SpatialPassSettings
{
    uint8_t do_spatial_reuse_pass B200PT_DEFAULT(1);             /*  0 */
    uint8_t _pad0[3] B200PT_DEFAULT({0});
    int spatial_pass_index B200PT_DEFAULT(0);
    int number_of_passes B200PT_DEFAULT(2);
    int reuse_radius B200PT_DEFAULT(16);
    int reuse_neighbor_count B200PT_DEFAULT(2);                  /* 16 */
    uint8_t do_disocclusion_reuse_boost B200PT_DEFAULT(1);       /* 20 */
    uint8_t _pad1[3] B200PT_DEFAULT({0});
    int disocclusion_reuse_count B200PT_DEFAULT(5);              /* 24 */
    uint8_t debug_neighbor_location B200PT_DEFAULT(0);           /* 28 */
    uint8_t do_neighbor_rotation B200PT_DEFAULT(1);
    uint8_t allow_converged_neighbors_reuse B200PT_DEFAULT(0);
    uint8_t _pad2 B200PT_DEFAULT(0);
    float converged_neighbor_reuse_probability B200PT_DEFAULT(0.5f); /* 32 */
    uint8_t do_visibility_only_last_pass B200PT_DEFAULT(1);      /* 36 */
    uint8_t _pad3[3] B200PT_DEFAULT({0});
    int neighbor_visibility_count B200PT_DEFAULT(5);             /* 40 */
    int _pad4 B200PT_DEFAULT(0);
    ReSTIRDIReservoir* input_reservoirs B200PT_DEFAULT(NULL);    /* 48 */
    ReSTIRDIReservoir* output_reservoirs B200PT_DEFAULT(NULL);   /* 56 */
} SpatialPassSettings;

typedef struct LightPresamplingSettings
{
    int number_of_subsets B200PT_DEFAULT(128);
    int subset_size B200PT_DEFAULT(1024);
    int tile_size B200PT_DEFAULT(8);
    int _pad0 B200PT_DEFAULT(0);
    ReSTIRDIPresampledLight* light_samples B200PT_DEFAULT(NULL); /* 16 */
} LightPresamplingSettings;

typedef struct ReSTIRDISettings
{
    InitialCandidatesSettings initial_candidates;                /*   0 */
    TemporalPassSettings temporal_pass;                          /*  24 */
    SpatialPassSettings spatial_pass;                            /*  64 */
    LightPresamplingSettings light_presampling;                  /* 128 */
    uint8_t do_fused_spatiotemporal B200PT_DEFAULT(1);           /* 152 */
    uint8_t _pad0[3] B200PT_DEFAULT({0});
    int m_cap B200PT_DEFAULT(25);                                /* 156 */
    uint8_t use_confidence_weights B200PT_DEFAULT(1);            /* 160 */
    uint8_t use_normal_similarity_heuristic B200PT_DEFAULT(1);
    uint8_t _pad1[2] B200PT_DEFAULT({0});
    float normal_similarity_angle_degrees B200PT_DEFAULT(25.0f); /* 164 */
    float normal_similarity_angle_precomp B200PT_DEFAULT(0.906307787f);
    uint8_t use_plane_distance_heuristic B200PT_DEFAULT(1);      /* 172 */
    uint8_t _pad2[3] B200PT_DEFAULT({0});
    float plane_distance_threshold B200PT_DEFAULT(0.1f);         /* 176 */
    uint8_t use_roughness_similarity_heuristic B200PT_DEFAULT(0);/* 180 */
    uint8_t _pad3[3] B200PT_DEFAULT({0});
    float roughness_similarity_threshold B200PT_DEFAULT(0.25f);  /* 184 */
    uint8_t do_final_shading_visibility B200PT_DEFAULT(1);       /* 188 */
    uint8_t _pad4[3] B200PT_DEFAULT({0});
    ReSTIRDIReservoir* restir_output_reservoirs B200PT_DEFAULT(NULL); /* 192 */
} ReSTIRDISettings;

/* ---- src/HostDeviceCommon/RenderSettings.h:17-252 --------------------------------------------- */
typedef enum PathRussianRoulette { MAX_THROUGHPUT = 0, ARNOLD_2014 = 1 } PathRussianRoulette;
typedef struct RISSettings
{
    int number_of_light_candidates B200PT_DEFAULT(4);
    int number_of_bsdf_candidates B200PT_DEFAULT(1);
} RISSettings;

typedef struct HIPRTRenderSettings
{
    uint8_t need_to_reset B200PT_DEFAULT(1);                     /*  0 */
    uint8_t do_update_status_buffers B200PT_DEFAULT(0);
    uint8_t accumulate B200PT_DEFAULT(1);
    uint8_t _pad0 B200PT_DEFAULT(0);
    int denoiser_AOV_accumulation_counter B200PT_DEFAULT(0);     /*  4 */
    int sample_number B200PT_DEFAULT(0);
    int samples_per_frame B200PT_DEFAULT(1);
    int nb_bounces B200PT_DEFAULT(3);                            /* 16 */
    uint8_t use_russian_roulette B200PT_DEFAULT(1);              /* 20 */
    uint8_t _pad1[3] B200PT_DEFAULT({0});
    int russian_roulette_min_depth B200PT_DEFAULT(2);
    float russian_roulette_throughput_clamp B200PT_DEFAULT(10.0f);
    int path_russian_roulette_method B200PT_DEFAULT(MAX_THROUGHPUT); /* 32 */
    int freeze_random B200PT_DEFAULT(0);                         /* 36 */
    uint8_t display_NaNs B200PT_DEFAULT(0);                      /* 40 */
    uint8_t allow_render_low_resolution B200PT_DEFAULT(1);
    uint8_t wants_render_low_resolution B200PT_DEFAULT(0);
    uint8_t _pad2 B200PT_DEFAULT(0);
    int render_low_resolution_scaling B200PT_DEFAULT(2);         /* 44 */
    uint8_t enable_adaptive_sampling B200PT_DEFAULT(1);          /* 48 */
    uint8_t _pad3[3] B200PT_DEFAULT({0});
    int adaptive_sampling_min_samples B200PT_DEFAULT(64);
    float adaptive_sampling_noise_threshold B200PT_DEFAULT(0.3f);
    uint8_t enable_pixel_stop_noise_threshold B200PT_DEFAULT(1); /* 60 */
    uint8_t _pad4[3] B200PT_DEFAULT({0});
    float stop_pixel_percentage_converged B200PT_DEFAULT(90.0f);
    float stop_pixel_noise_threshold B200PT_DEFAULT(0.0f);
    float direct_contribution_clamp B200PT_DEFAULT(0.0f);        /* 72 */
    float envmap_contribution_clamp B200PT_DEFAULT(0.0f);
    float indirect_contribution_clamp B200PT_DEFAULT(15.0f);
    float minimum_light_contribution B200PT_DEFAULT(0.08f);
    int number_of_light_samples B200PT_DEFAULT(1);               /* 88 */
    uint8_t do_alpha_testing B200PT_DEFAULT(1);                  /* 92 */
    uint8_t _pad5[3] B200PT_DEFAULT({0});
    RISSettings ris_settings;                                    /* 96 */
    ReSTIRDISettings restir_di_settings;                         /* 104 */
} HIPRTRenderSettings;

/* ---- src/HostDeviceCommon/WorldSettings.h:11-52 ----------------------------------------------- */
typedef enum AmbientLightType { AMBIENT_NONE = 0, AMBIENT_UNIFORM = 1, AMBIENT_ENVMAP = 2 } AmbientLightType;
typedef struct WorldSettings
{
    int ambient_light_type B200PT_DEFAULT(AMBIENT_UNIFORM);      /*   0 */
    ColorRGB32F uniform_light_color B200PT_DEFAULT({0.5f, 0.5f, 0.5f});
    unsigned int envmap_width B200PT_DEFAULT(0);                 /*  16 */
    unsigned int envmap_height B200PT_DEFAULT(0);
    float envmap_intensity B200PT_DEFAULT(1.0f);                 /*  24 */
    int envmap_scale_background_intensity B200PT_DEFAULT(0);
    void* envmap B200PT_DEFAULT(NULL);                           /*  32: b200pt_tex2d (RGB(A) float32) */
    float envmap_total_sum B200PT_DEFAULT(0.0f);                 /*  40 */
    int _pad0 B200PT_DEFAULT(0);
    float* envmap_cdf B200PT_DEFAULT(NULL);                      /*  48 */
    int* alias_table_alias B200PT_DEFAULT(NULL);                 /*  56 */
    float* alias_table_probas B200PT_DEFAULT(NULL);              /*  64 */
    b200pt_float4x4 envmap_to_world_matrix B200PT_DEFAULT({{{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}}}); /* 72 */
    b200pt_float4x4 world_to_envmap_matrix B200PT_DEFAULT({{{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}}}); /* 136 */
} WorldSettings;

/* ---- src/HostDeviceCommon/HIPRTCamera.h:16-49 ------------------------------------------------- */
typedef struct HIPRTCamera
{
    b200pt_float4x4 inverse_view;                                /*   0 */
    b200pt_float4x4 inverse_projection;                          /*  64 */
    b200pt_float4x4 view_projection;                             /* 128 */
    uint8_t do_jittering B200PT_DEFAULT(1);                      /* 192 */
    uint8_t _pad0[3] B200PT_DEFAULT({0});
} HIPRTCamera;

/* hiprtGlobalStackBuffer (thirdparties/HIPRT-Fork/hiprt/hiprt_types.h:288-293), kept for layout only */
typedef struct b200pt_global_stack_buffer
{
    uint32_t stackSize B200PT_DEFAULT(0);
    uint32_t stackCount B200PT_DEFAULT(0);
    void* stackData B200PT_DEFAULT(NULL);
} b200pt_global_stack_buffer;

typedef struct HIPRTRenderData
{
    unsigned int random_seed B200PT_DEFAULT(42);                 /*    0 */
    int _pad0 B200PT_DEFAULT(0);
    void* geom B200PT_DEFAULT(NULL);                             /*    8: hiprtGeometry in the reference; here the device address returned by b200pt_bvh_device_handle */
    void* hiprt_function_table B200PT_DEFAULT(NULL);             /*   16: unused (the filter function is compiled in) */
    int global_traversal_stack_buffer_size B200PT_DEFAULT(32);   /*   24 */
    int _pad1 B200PT_DEFAULT(0);
    b200pt_global_stack_buffer global_traversal_stack_buffer;    /*   32: unused (stacks live in shared + local memory) */
    RenderBuffers buffers;                                       /*   48 */
    BRDFsData bsdfs_data;                                        /*  144 */
    AuxiliaryBuffers aux_buffers;                                /*  208 */
    GBuffer g_buffer;                                            /*  296 */
    GBuffer g_buffer_prev_frame;                                 /*  360 */
    HIPRTRenderSettings render_settings;                         /*  424 */
    WorldSettings world_settings;                                /*  728 */
    HIPRTCamera current_camera;                                  /*  928 */
    HIPRTCamera prev_camera;                                     /* 1124 */
    void* cpu_only_bvh B200PT_DEFAULT(NULL);                     /* 1320: CPUData, ignored */
} HIPRTRenderData;

/* src/Device/kernel_parameters/ReSTIR/DI/LightPresamplingParameters.h:16-49 */
typedef struct LightPresamplingParameters
{
    int number_of_subsets B200PT_DEFAULT(128);                   /*   0 */
    int subset_size B200PT_DEFAULT(1024);
    ReSTIRDIPresampledLight* out_light_samples B200PT_DEFAULT(NULL); /* 8 */
    int emissive_triangles_count B200PT_DEFAULT(0);              /*  16 */
    int _pad0 B200PT_DEFAULT(0);
    int* emissive_triangles_indices B200PT_DEFAULT(NULL);        /*  24 */
    int* triangles_indices B200PT_DEFAULT(NULL);
    b200pt_float3* vertices_positions B200PT_DEFAULT(NULL);
    int* material_indices B200PT_DEFAULT(NULL);
    RendererMaterial* materials B200PT_DEFAULT(NULL);            /*  56 */
    WorldSettings world_settings;                                /*  64 */
    uint8_t freeze_random B200PT_DEFAULT(0);                     /* 264 */
    uint8_t _pad1[3] B200PT_DEFAULT({0});
    int sample_number B200PT_DEFAULT(0);                         /* 268 */
    int random_seed B200PT_DEFAULT(42);                          /* 272 */
    float envmap_sampling_probability B200PT_DEFAULT(0.25f);     /* 276 */
} LightPresamplingParameters;

#endif /* !B200PT_USE_REFERENCE_HEADERS */

#ifdef __cplusplus
static_assert(sizeof(SimplifiedRendererMaterial) == 256, "SimplifiedRendererMaterial must be 256 bytes");
static_assert(sizeof(RendererMaterial) == 332, "RendererMaterial must be 332 bytes");
static_assert(sizeof(RayVolumeState) == 36, "RayVolumeState must be 36 bytes");
static_assert(sizeof(ReSTIRDIReservoir) == 36, "ReSTIRDIReservoir must be 36 bytes");
static_assert(sizeof(ReSTIRDIPresampledLight) == 48, "ReSTIRDIPresampledLight must be 48 bytes");
static_assert(sizeof(HIPRTRenderSettings) == 304, "HIPRTRenderSettings must be 304 bytes");
static_assert(sizeof(WorldSettings) == 200, "WorldSettings must be 200 bytes");
static_assert(sizeof(HIPRTCamera) == 196, "HIPRTCamera must be 196 bytes");
static_assert(sizeof(HIPRTRenderData) == 1328, "HIPRTRenderData must be 1328 bytes");
static_assert(sizeof(LightPresamplingParameters) == 280, "LightPresamplingParameters must be 280 bytes");
#endif

#endif /* B200PT_TYPES_H */
