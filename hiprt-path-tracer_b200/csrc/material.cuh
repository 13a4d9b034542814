// Material fetch at a hit point and the nested-dielectrics interior stack.
// reference: src/Device/includes/Material.h:23-159 (get_intersection_material and friends),
//            src/Device/includes/NestedDielectrics.h:169-288 (InteriorStackImpl<ISS_WITH_PRIORITIES>),
//            src/HostDeviceCommon/Material.h:29-111 (is_emissive, get_alphas, get_thin_walled_roughness).
#pragma once

#include "texture.cuh"

namespace pt
{
typedef SimplifiedRendererMaterial Material;

PT_HD RGB rgb(const ColorRGB32F& c) { return RGB(c.r, c.g, c.b); }
PT_HD void set_rgb(ColorRGB32F& dst, RGB c) { dst.r = c.r; dst.g = c.g; dst.b = c.b; }

PT_DEV RGB material_emission(const Material& m) { return rgb(m.emission) * m.emission_strength; }

PT_DEV bool material_is_emissive(const Material& m)
{
    return !is_zero(m.emission.r * m.emission_strength) || !is_zero(m.emission.g * m.emission_strength)
        || !is_zero(m.emission.b * m.emission_strength) || m.emissive_texture_used;
}

constexpr float kRoughnessClamp = 1.0e-4f;

PT_DEV void get_alphas(float roughness, float anisotropy, float& alpha_x, float& alpha_y)
{
    float aspect = sqrtf(1.0f - 0.9f * anisotropy);
    alpha_x = maxf(kRoughnessClamp, roughness * roughness / aspect);
    alpha_y = maxf(kRoughnessClamp, roughness * roughness * aspect);
}

PT_DEV float get_thin_walled_roughness(bool thin_walled, float base_roughness, float relative_eta)
{
    if (!thin_walled)
        return base_roughness;
    float remapped = base_roughness * sqrtf(3.7f * (relative_eta - 1.0f) * square(relative_eta - 0.5f) / pow_3(relative_eta));
    return clampf(0.0f, 1.0f, remapped / 1.39f);
}

// ---- loads from the 332-byte, 4-byte-aligned RendererMaterial array ---------------------------
PT_DEV const unsigned char* material_ptr(const HIPRTRenderData& rd, int material_index)
{
    return reinterpret_cast<const unsigned char*>(rd.buffers.materials_buffer) + (size_t)material_index * sizeof(RendererMaterial);
}
PT_DEV float material_field_f(const HIPRTRenderData& rd, int material_index, size_t offset)
{
    return __ldg(reinterpret_cast<const float*>(material_ptr(rd, material_index) + offset));
}
PT_DEV int material_field_i(const HIPRTRenderData& rd, int material_index, size_t offset)
{
    return __ldg(reinterpret_cast<const int*>(material_ptr(rd, material_index) + offset));
}
PT_DEV float material_ior(const HIPRTRenderData& rd, int material_index) { return material_field_f(rd, material_index, offsetof(SimplifiedRendererMaterial, ior)); }
PT_DEV RGB material_buffer_emission(const HIPRTRenderData& rd, int material_index)
{
    size_t e = offsetof(SimplifiedRendererMaterial, emission);
    float s = material_field_f(rd, material_index, offsetof(SimplifiedRendererMaterial, emission_strength));
    return RGB(material_field_f(rd, material_index, e), material_field_f(rd, material_index, e + 4), material_field_f(rd, material_index, e + 8)) * s;
}

PT_DEV const b200pt_tex2d& material_texture(const HIPRTRenderData& rd, int texture_index)
{
    return reinterpret_cast<const b200pt_tex2d*>(rd.buffers.material_textures)[texture_index];
}

PT_DEV bool has_texture(int texture_index) { return texture_index != B200PT_NO_TEXTURE && texture_index != B200PT_CONSTANT_EMISSIVE_TEXTURE; }

PT_DEV void texture_override_f(const HIPRTRenderData& rd, float& value, float2 texcoords, int texture_index)
{
    if (has_texture(texture_index))
        value = sample_tex2d_rgba(material_texture(rd, texture_index), texcoords, false).r;
}
PT_DEV void texture_override_rgb(const HIPRTRenderData& rd, ColorRGB32F& value, float2 texcoords, int texture_index)
{
    if (has_texture(texture_index))
    {
        RGBA c = sample_tex2d_rgba(material_texture(rd, texture_index), texcoords, false);
        value.r = c.r; value.g = c.g; value.b = c.b;
    }
}

// Alpha of the base colour texture at a hit (get_hit_base_color_alpha, Material.h:23-45)
PT_DEV float base_color_alpha(const HIPRTRenderData& rd, int material_index, float2 texcoords)
{
    int tex = material_field_i(rd, material_index, offsetof(RendererMaterial, base_color_texture_index));
    if (tex == -1)
        return 1.0f;
    if (!has_texture(tex))
        return 1.0f;
    return sample_tex2d_rgba(material_texture(rd, tex), texcoords, true).a;
}

// Coat roughening of the base and second-metal roughness (Material.h:87-93) and the emissive-texture flag (:82)
PT_DEV void finish_intersection_material(Material& out, int emission_texture_index)
{
    out.emissive_texture_used = emission_texture_index > 0;   // reference quirk: index 0 does not count (Material.h:82)

    float coat = out.coat;
    float target_base = pow_1_4(minf(1.0f, pow_4(out.roughness) + 2.0f * pow_4(out.coat_roughness)));
    float roughened_base = lerpf(out.roughness, target_base, coat);
    out.roughness = lerpf(out.roughness, roughened_base, out.coat_roughening);

    float target_second = pow_1_4(minf(1.0f, pow_4(out.second_roughness) + 2.0f * pow_4(out.coat_roughness)));
    float roughened_second = lerpf(out.second_roughness, target_second, coat);
    out.second_roughness = lerpf(out.second_roughness, roughened_second, out.coat_roughening);
}

// get_intersection_material for a material WITHOUT texture overrides (none of the 18 parameter textures; the normal map is
// applied to the shading normal elsewhere): a function of the material buffer entry alone -- no texcoords, no hit. The
// wavefront keeps only the material INDEX of such a vertex in the path state and its kernels call this where they load the
// vertex (the inlined copy lets the compiler fetch just the fields a kernel reads, from a buffer that lives in L1/L2),
// instead of writing the 256-byte material to the slot and reading it back in every phase.
PT_DEV void plain_intersection_material(const HIPRTRenderData& rd, int material_index, Material& out)
{
    const uint32_t* src = reinterpret_cast<const uint32_t*>(material_ptr(rd, material_index));
    uint32_t* dst = reinterpret_cast<uint32_t*>(&out);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(Material) / 4); i++)
        dst[i] = __ldg(src + i);
    const int emission_texture_index = __ldg(reinterpret_cast<const int*>(src) + sizeof(Material) / 4 + 1);
    // The reference always round-trips the emission through '/ strength' and get_emission() multiplies again;
    // keep the division so that a zero strength yields the same NaN/Inf pattern as the oracle.
    RGB emission = material_emission(out) / out.emission_strength;
    set_rgb(out.emission, emission);
    if (rd.bsdfs_data.white_furnace_mode)
        set_rgb(out.base_color, RGB(1.0f));
    finish_intersection_material(out, emission_texture_index);
}

// get_intersection_material (Material.h:47-96). Returns true when the material has no texture override (see above).
PT_DEV_NOINLINE bool get_intersection_material(const HIPRTRenderData& rd, int material_index, float2 texcoords, Material& out)
{
    const uint32_t* src = reinterpret_cast<const uint32_t*>(material_ptr(rd, material_index));
    uint32_t* dst = reinterpret_cast<uint32_t*>(&out);

    int tex[19];
#pragma unroll
    for (int i = 0; i < 19; i++)
        tex[i] = __ldg(reinterpret_cast<const int*>(src) + sizeof(Material) / 4 + i);
    // order of the indices in RendererMaterial: 0 normal map, 1 emission, 2 base colour, 3 roughness-metallic, 4 roughness,
    // 5 oren sigma, 6 metallic, 7 specular, 8 specular tint, 9 specular colour, 10 anisotropic, 11 anisotropic rotation,
    // 12 coat, 13 coat roughness, 14 coat ior, 15 sheen, 16 sheen roughness, 17 sheen colour, 18 specular transmission
    bool any_texture = false;
#pragma unroll
    for (int i = 1; i < 19; i++)
        any_texture |= has_texture(tex[i]);
    if (!any_texture)
    {
        plain_intersection_material(rd, material_index, out);
        return true;
    }

#pragma unroll
    for (int i = 0; i < (int)(sizeof(Material) / 4); i++)
        dst[i] = __ldg(src + i);
    {
        // emission = get_emission() / emission_strength, overridden by the emission texture
        RGB emission = material_emission(out) / out.emission_strength;
        if (has_texture(tex[1]))
        {
            RGBA c = sample_tex2d_rgba(material_texture(rd, tex[1]), texcoords, false);
            emission = RGB(c.r, c.g, c.b);
        }
        set_rgb(out.emission, emission);
    }

    if (rd.bsdfs_data.white_furnace_mode)
        set_rgb(out.base_color, RGB(1.0f));
    else if (has_texture(tex[2]))
    {
        RGBA c = sample_tex2d_rgba(material_texture(rd, tex[2]), texcoords, true);
        set_rgb(out.base_color, RGB(c.r, c.g, c.b));
    }

    {
        if (tex[3] != B200PT_NO_TEXTURE)
        {
            RGBA c = sample_tex2d_rgba(material_texture(rd, tex[3]), texcoords, false);
            out.roughness = c.g;
            out.metallic = c.b;
        }
        else
        {
            texture_override_f(rd, out.metallic, texcoords, tex[6]);
            texture_override_f(rd, out.roughness, texcoords, tex[4]);
        }
        texture_override_f(rd, out.oren_nayar_sigma, texcoords, tex[5]);
        texture_override_f(rd, out.specular, texcoords, tex[7]);
        texture_override_f(rd, out.specular_tint, texcoords, tex[8]);
        texture_override_rgb(rd, out.specular_color, texcoords, tex[9]);
        texture_override_f(rd, out.anisotropy, texcoords, tex[10]);
        texture_override_f(rd, out.anisotropy_rotation, texcoords, tex[11]);
        texture_override_f(rd, out.coat, texcoords, tex[12]);
        texture_override_f(rd, out.coat_roughness, texcoords, tex[13]);
        texture_override_f(rd, out.coat_ior, texcoords, tex[14]);
        texture_override_f(rd, out.sheen, texcoords, tex[15]);
        texture_override_f(rd, out.sheen_roughness, texcoords, tex[16]);
        texture_override_rgb(rd, out.sheen_color, texcoords, tex[17]);
        texture_override_f(rd, out.specular_transmission, texcoords, tex[18]);
    }

    finish_intersection_material(out, tex[1]);
    return false;
}

// ---- interior stack (ISS_WITH_PRIORITIES, stack size 3) ---------------------------------------
// entry word: bit 0 priority (1 bit: reference quirk), bit 1 odd_parity, bit 5 topmost, bits 6..31 material index
namespace istack
{
PT_HD uint32_t priority(uint32_t e) { return e & 1u; }
PT_HD bool odd_parity(uint32_t e) { return (e >> 1) & 1u; }
PT_HD bool topmost(uint32_t e) { return (e >> 5) & 1u; }
PT_HD int material_index(uint32_t e) { return (int)(e >> 6); }
PT_HD uint32_t make(uint32_t prio, bool odd, bool top, int material) { return (prio & 1u) | ((odd ? 1u : 0u) << 1) | ((top ? 1u : 0u) << 5) | ((uint32_t)material << 6); }
PT_HD uint32_t set_topmost(uint32_t e, bool top) { return (e & ~(1u << 5)) | ((top ? 1u : 0u) << 5); }
constexpr uint32_t kDefaultEntry = 0u | (1u << 1) | (1u << 5) | ((uint32_t)B200PT_MAX_MATERIAL_INDEX << 6);
} // namespace istack

PT_HD void volume_state_init(RayVolumeState& vs)
{
    vs.distance_in_volume = 0.0f;
    for (int i = 0; i < B200PT_NESTED_DIELECTRICS_STACK_SIZE; i++)
        vs.interior_stack.stack[i] = istack::kDefaultEntry;
    vs.interior_stack.stack_position = 0;
    vs.incident_mat_index = -1;
    vs.outgoing_mat_index = -1;
    vs.inside_material = 0;
    vs._pad0[0] = vs._pad0[1] = vs._pad0[2] = 0;
    vs.sampled_wavelength = 0.0f;
}

// Returns true when the boundary must be skipped (NestedDielectrics.h:171-221)
PT_HD bool interior_stack_push(RayVolumeState& vs, int material_index, int material_priority)
{
    InteriorStack& st = vs.interior_stack;
    int last_entered = 0;
    for (last_entered = st.stack_position; last_entered >= 0; last_entered--)
    {
        uint32_t e = st.stack[last_entered];
        if (istack::material_index(e) != material_index && istack::topmost(e) && istack::odd_parity(e))
            break;
    }

    bool odd = true;
    int previous_same;
    for (previous_same = st.stack_position; previous_same >= 0; previous_same--)
    {
        uint32_t e = st.stack[previous_same];
        if (istack::material_index(e) == material_index)
        {
            st.stack[previous_same] = istack::set_topmost(e, false);
            odd = !istack::odd_parity(e);
            break;
        }
    }

    vs.inside_material = !odd;

    if (st.stack_position < B200PT_NESTED_DIELECTRICS_STACK_SIZE - 1)
        st.stack_position++;
    // material_index is stored in a 26-bit field, priority in a 1-bit field: both truncate like the bit-field does
    st.stack[st.stack_position] = istack::make((uint32_t)material_priority, odd, true, material_index & B200PT_MAX_MATERIAL_INDEX);

    // In the reference 'last_entered' can end at -1 when nothing qualifies: stack[-1] is then read (undefined
    // behaviour, it aliases distance_in_volume). Here that read is replaced by the default entry.
    uint32_t last_entry = last_entered >= 0 ? st.stack[last_entered] : istack::kDefaultEntry;
    if (material_priority < (int)istack::priority(last_entry))
        return true;

    if (odd)
    {
        vs.incident_mat_index = istack::material_index(last_entry);
        vs.outgoing_mat_index = material_index;
    }
    else
    {
        vs.incident_mat_index = material_index;
        vs.outgoing_mat_index = istack::material_index(last_entry);
    }
    return false;
}

// NestedDielectrics.h:223-253
PT_HD void interior_stack_pop(RayVolumeState& vs, bool inside_material)
{
    InteriorStack& st = vs.interior_stack;
    int stack_top_mat = istack::material_index(st.stack[st.stack_position]);
    if (st.stack_position > 0)
        st.stack_position--;

    if (inside_material)
    {
        int previous_same;
        for (previous_same = st.stack_position; previous_same >= 0; previous_same--)
            if (istack::material_index(st.stack[previous_same]) == stack_top_mat)
                break;

        if (previous_same >= 0)
            for (int i = previous_same + 1; i <= st.stack_position; i++)
                st.stack[i - 1] = st.stack[i];

        if (st.stack_position > 0)
            st.stack_position--;
    }

    for (int i = st.stack_position; i >= 0; i--)
    {
        if (istack::material_index(st.stack[i]) == stack_top_mat)
        {
            st.stack[i] = istack::set_topmost(st.stack[i], true);
            break;
        }
    }
}

PT_DEV bool is_inside_volume(const RayVolumeState& vs) { return vs.interior_stack.stack_position > 0; }
} // namespace pt
