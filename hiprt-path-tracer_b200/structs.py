"""ctypes mirrors of include/b200pt_types.h (themselves byte-compatible mirrors of the reference's boundary structs,
src/HostDeviceCommon/*.h). tests/test_abi_layout.py checks every offset against the reference build (oracle/_ref)."""
import ctypes as C

u8, i32, u32, f32, vp = C.c_uint8, C.c_int32, C.c_uint32, C.c_float, C.c_void_p


class Float3(C.Structure):
    _fields_ = [("x", f32), ("y", f32), ("z", f32)]


class ColorRGB32F(C.Structure):
    _fields_ = [("r", f32), ("g", f32), ("b", f32)]

    def set(self, r, g=None, b=None):
        self.r, self.g, self.b = (r, r, r) if g is None else (r, g, b)


class Float4x4(C.Structure):
    _fields_ = [("m", (f32 * 4) * 4)]

    def set(self, rows):
        for i in range(4):
            for j in range(4):
                self.m[i][j] = float(rows[i][j])


class Int2(C.Structure):
    _fields_ = [("x", i32), ("y", i32)]


class SimplifiedRendererMaterial(C.Structure):
    _fields_ = [
        ("emissive_texture_used", u8), ("_pad0", u8 * 3), ("emission_strength", f32), ("base_color", ColorRGB32F),
        ("roughness", f32), ("oren_nayar_sigma", f32), ("metallic", f32), ("metallic_F90_falloff_exponent", f32),
        ("metallic_F82", ColorRGB32F), ("metallic_F90", ColorRGB32F), ("anisotropy", f32), ("anisotropy_rotation", f32),
        ("second_roughness_weight", f32), ("second_roughness", f32), ("specular", f32), ("specular_tint", f32),
        ("specular_color", ColorRGB32F), ("specular_darkening", f32), ("coat", f32), ("coat_medium_absorption", ColorRGB32F),
        ("coat_medium_thickness", f32), ("coat_roughness", f32), ("coat_roughening", f32), ("coat_darkening", f32),
        ("coat_anisotropy", f32), ("coat_anisotropy_rotation", f32), ("coat_ior", f32), ("sheen", f32), ("sheen_roughness", f32),
        ("sheen_color", ColorRGB32F), ("ior", f32), ("specular_transmission", f32), ("absorption_at_distance", f32),
        ("absorption_color", ColorRGB32F), ("dispersion_scale", f32), ("dispersion_abbe_number", f32), ("thin_walled", u8),
        ("_pad1", u8 * 3), ("thin_film", f32), ("thin_film_ior", f32), ("thin_film_thickness", f32), ("thin_film_kappa_3", f32),
        ("thin_film_hue_shift_degrees", f32), ("thin_film_base_ior_override", f32), ("thin_film_do_ior_override", u8), ("srgb", u8),
        ("_pad2", u8 * 2), ("alpha_opacity", f32), ("dielectric_priority", i32), ("energy_preservation_monte_carlo_samples", i32),
        ("enforce_strong_energy_conservation", u8), ("_pad3", u8 * 3), ("emission", ColorRGB32F),
    ]


class RendererMaterial(C.Structure):
    _fields_ = [("base", SimplifiedRendererMaterial)] + [(n, i32) for n in (
        "normal_map_texture_index", "emission_texture_index", "base_color_texture_index", "roughness_metallic_texture_index",
        "roughness_texture_index", "oren_sigma_texture_index", "metallic_texture_index", "specular_texture_index",
        "specular_tint_texture_index", "specular_color_texture_index", "anisotropic_texture_index",
        "anisotropic_rotation_texture_index", "coat_texture_index", "coat_roughness_texture_index", "coat_ior_texture_index",
        "sheen_texture_index", "sheen_roughness_texture_index", "sheen_color_texture_index", "specular_transmission_texture_index")]


class InteriorStack(C.Structure):
    _fields_ = [("stack", u32 * 3), ("stack_position", i32)]


class RayVolumeState(C.Structure):
    _fields_ = [("distance_in_volume", f32), ("interior_stack", InteriorStack), ("incident_mat_index", i32), ("outgoing_mat_index", i32),
                ("inside_material", u8), ("_pad0", u8 * 3), ("sampled_wavelength", f32)]


class ReSTIRDISample(C.Structure):
    _fields_ = [("emissive_triangle_index", i32), ("point_on_light_source", Float3), ("target_function", f32), ("flags", u8), ("_pad0", u8 * 3)]


class ReSTIRDIReservoir(C.Structure):
    _fields_ = [("M", i32), ("weight_sum", f32), ("UCW", f32), ("sample", ReSTIRDISample)]


class ReSTIRDIPresampledLight(C.Structure):
    _fields_ = [("emissive_triangle_index", i32), ("point_on_light_source", Float3), ("light_source_normal", Float3), ("radiance", ColorRGB32F),
                ("pdf", f32), ("flags", u8), ("_pad0", u8 * 3)]


class RenderBuffers(C.Structure):
    _fields_ = [("pixels", vp), ("triangles_indices", vp), ("vertices_positions", vp), ("has_vertex_normals", vp), ("vertex_normals", vp),
                ("texcoords", vp), ("material_indices", vp), ("materials_buffer", vp), ("emissive_triangles_count", i32), ("_pad0", i32),
                ("emissive_triangles_indices", vp), ("material_textures", vp), ("textures_dims", vp)]


class BRDFsData(C.Structure):
    _fields_ = [("white_furnace_mode", u8), ("white_furnace_mode_turn_off_emissives", u8), ("_pad0", u8 * 6),
                ("sheen_ltc_parameters_texture", vp), ("GGX_conductor_Ess", vp), ("glossy_dielectric_Ess", vp), ("GGX_Ess_glass", vp),
                ("GGX_Ess_glass_inverse", vp), ("GGX_Ess_thin_glass", vp), ("use_hardware_tex_interpolation", u8),
                ("clearcoat_compensation_approximation", u8), ("_pad1", u8 * 2), ("GGX_masking_shadowing", i32)]


class AuxiliaryBuffers(C.Structure):
    _fields_ = [(n, vp) for n in ("pixel_active", "denoiser_normals", "denoiser_albedo", "pixel_sample_count", "pixel_squared_luminance",
                                  "pixel_converged_sample_count", "still_one_ray_active", "stop_noise_threshold_converged_count",
                                  "restir_reservoir_buffer_1", "restir_reservoir_buffer_2", "restir_reservoir_buffer_3")]


class GBuffer(C.Structure):
    _fields_ = [(n, vp) for n in ("materials", "first_hit_prim_index", "shading_normals", "geometric_normals", "view_directions", "first_hits",
                                  "camera_ray_hit", "ray_volume_states")]


class InitialCandidatesSettings(C.Structure):
    _fields_ = [("number_of_initial_light_candidates", i32), ("number_of_initial_bsdf_candidates", i32), ("envmap_candidate_probability", f32),
                ("_pad0", i32), ("output_reservoirs", vp)]


class TemporalPassSettings(C.Structure):
    _fields_ = [("do_temporal_reuse_pass", u8), ("use_permutation_sampling", u8), ("_pad0", u8 * 2), ("permutation_sampling_random_bits", i32),
                ("max_neighbor_search_count", i32), ("neighbor_search_radius", i32), ("temporal_buffer_clear_requested", u8), ("_pad1", u8 * 7),
                ("input_reservoirs", vp), ("output_reservoirs", vp)]


class SpatialPassSettings(C.Structure):
    _fields_ = [("do_spatial_reuse_pass", u8), ("_pad0", u8 * 3), ("spatial_pass_index", i32), ("number_of_passes", i32), ("reuse_radius", i32),
                ("reuse_neighbor_count", i32), ("do_disocclusion_reuse_boost", u8), ("_pad1", u8 * 3), ("disocclusion_reuse_count", i32),
                ("debug_neighbor_location", u8), ("do_neighbor_rotation", u8), ("allow_converged_neighbors_reuse", u8), ("_pad2", u8),
                ("converged_neighbor_reuse_probability", f32), ("do_visibility_only_last_pass", u8), ("_pad3", u8 * 3),
                ("neighbor_visibility_count", i32), ("_pad4", i32), ("input_reservoirs", vp), ("output_reservoirs", vp)]


class LightPresamplingSettings(C.Structure):
    _fields_ = [("number_of_subsets", i32), ("subset_size", i32), ("tile_size", i32), ("_pad0", i32), ("light_samples", vp)]


class ReSTIRDISettings(C.Structure):
    _fields_ = [("initial_candidates", InitialCandidatesSettings), ("temporal_pass", TemporalPassSettings), ("spatial_pass", SpatialPassSettings),
                ("light_presampling", LightPresamplingSettings), ("do_fused_spatiotemporal", u8), ("_pad0", u8 * 3), ("m_cap", i32),
                ("use_confidence_weights", u8), ("use_normal_similarity_heuristic", u8), ("_pad1", u8 * 2),
                ("normal_similarity_angle_degrees", f32), ("normal_similarity_angle_precomp", f32), ("use_plane_distance_heuristic", u8),
                ("_pad2", u8 * 3), ("plane_distance_threshold", f32), ("use_roughness_similarity_heuristic", u8), ("_pad3", u8 * 3),
                ("roughness_similarity_threshold", f32), ("do_final_shading_visibility", u8), ("_pad4", u8 * 3), ("restir_output_reservoirs", vp)]


class RISSettings(C.Structure):
    _fields_ = [("number_of_light_candidates", i32), ("number_of_bsdf_candidates", i32)]


class HIPRTRenderSettings(C.Structure):
    _fields_ = [("need_to_reset", u8), ("do_update_status_buffers", u8), ("accumulate", u8), ("_pad0", u8),
                ("denoiser_AOV_accumulation_counter", i32), ("sample_number", i32), ("samples_per_frame", i32), ("nb_bounces", i32),
                ("use_russian_roulette", u8), ("_pad1", u8 * 3), ("russian_roulette_min_depth", i32), ("russian_roulette_throughput_clamp", f32),
                ("path_russian_roulette_method", i32), ("freeze_random", i32), ("display_NaNs", u8), ("allow_render_low_resolution", u8),
                ("wants_render_low_resolution", u8), ("_pad2", u8), ("render_low_resolution_scaling", i32), ("enable_adaptive_sampling", u8),
                ("_pad3", u8 * 3), ("adaptive_sampling_min_samples", i32), ("adaptive_sampling_noise_threshold", f32),
                ("enable_pixel_stop_noise_threshold", u8), ("_pad4", u8 * 3), ("stop_pixel_percentage_converged", f32),
                ("stop_pixel_noise_threshold", f32), ("direct_contribution_clamp", f32), ("envmap_contribution_clamp", f32),
                ("indirect_contribution_clamp", f32), ("minimum_light_contribution", f32), ("number_of_light_samples", i32),
                ("do_alpha_testing", u8), ("_pad5", u8 * 3), ("ris_settings", RISSettings), ("restir_di_settings", ReSTIRDISettings)]


class WorldSettings(C.Structure):
    _fields_ = [("ambient_light_type", i32), ("uniform_light_color", ColorRGB32F), ("envmap_width", u32), ("envmap_height", u32),
                ("envmap_intensity", f32), ("envmap_scale_background_intensity", i32), ("envmap", vp), ("envmap_total_sum", f32), ("_pad0", i32),
                ("envmap_cdf", vp), ("alias_table_alias", vp), ("alias_table_probas", vp), ("envmap_to_world_matrix", Float4x4),
                ("world_to_envmap_matrix", Float4x4)]


class HIPRTCamera(C.Structure):
    _fields_ = [("inverse_view", Float4x4), ("inverse_projection", Float4x4), ("view_projection", Float4x4), ("do_jittering", u8), ("_pad0", u8 * 3)]


class GlobalStackBuffer(C.Structure):
    _fields_ = [("stackSize", u32), ("stackCount", u32), ("stackData", vp)]


class HIPRTRenderData(C.Structure):
    _fields_ = [("random_seed", u32), ("_pad0", i32), ("geom", vp), ("hiprt_function_table", vp), ("global_traversal_stack_buffer_size", i32),
                ("_pad1", i32), ("global_traversal_stack_buffer", GlobalStackBuffer), ("buffers", RenderBuffers), ("bsdfs_data", BRDFsData),
                ("aux_buffers", AuxiliaryBuffers), ("g_buffer", GBuffer), ("g_buffer_prev_frame", GBuffer), ("render_settings", HIPRTRenderSettings),
                ("world_settings", WorldSettings), ("current_camera", HIPRTCamera), ("prev_camera", HIPRTCamera), ("cpu_only_bvh", vp)]


class Options(C.Structure):
    """b200pt_options: compile-time kernel options of the reference (KernelOptions.h) selecting a prebuilt permutation."""
    _fields_ = [("direct_light_sampling_strategy", i32), ("envmap_sampling_strategy", i32), ("ggx_sample_function", i32),
                ("restir_di_later_bounces_strategy", i32)]


class Tile(C.Structure):
    _fields_ = [("x0", i32), ("y0", i32), ("x1", i32), ("y1", i32)]


class Counters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodes", C.c_uint64), ("triangles", C.c_uint64), ("stack_overflows", C.c_uint64),
                ("queue_rays", C.c_uint64), ("queue_nodes", C.c_uint64), ("queue_triangles", C.c_uint64)]


class BvhStats(C.Structure):
    _fields_ = [("node_count", C.c_uint), ("triangle_count", C.c_uint), ("max_depth", i32), ("built_on_device", i32), ("sah_cost", C.c_float),
                ("node_fill", C.c_float), ("build_seconds", C.c_double), ("device_build_seconds", C.c_double)]


LSS = {"none": 0, "uniform": 1, "bsdf": 2, "mis": 3, "ris": 4, "restir_di": 5}
ESS = {"none": 0, "binary_search": 1, "alias_table": 2}
AMBIENT_NONE, AMBIENT_UNIFORM, AMBIENT_ENVMAP = 0, 1, 2

ALL_STRUCTS = {c.__name__: c for c in (SimplifiedRendererMaterial, RendererMaterial, RayVolumeState, ReSTIRDISample, ReSTIRDIReservoir,
                                       ReSTIRDIPresampledLight, RenderBuffers, BRDFsData, AuxiliaryBuffers, GBuffer, InitialCandidatesSettings,
                                       TemporalPassSettings, SpatialPassSettings, LightPresamplingSettings, ReSTIRDISettings, HIPRTRenderSettings,
                                       WorldSettings, HIPRTCamera, HIPRTRenderData)}
