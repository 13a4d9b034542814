// Texture / LUT fetches from plain linear device memory (no CUDA arrays, no texture objects).
//
// Implements the reference's *CPU* sampling rule, which is the one the oracle uses: nearest texel, repeat addressing,
// v flipped, x = int(u * (w - 1)) -- Image8Bit/Image32Bit::sample_rgba32f (src/Image/Image.cpp:161-195, 492-526) and
// Image32Bit3D::sample_rgba32f (:737-779) behind sample_texture_rgba / sample_texture_3D_rgb_32bits
// (src/Device/includes/Texture.h:30-76, 133-194). The reference's GPU path differs from its own CPU path here
// (tex2D point fetch for 2-D, manual bilinear for 3-D LUTs); the oracle, not HIPRT, is the spec.
#pragma once

#include "../../include/b200pt.h"
#include "vec.cuh"

namespace pt
{
PT_DEV float wrap_uv(float u)
{
    float w = u;
    if (w != 1.0f)
        w -= (float)(int)u;          // u -= static_cast<int>(u)
    return w < 0 ? 1.0f + w : w;
}

PT_DEV RGBA sample_tex2d_rgba(const b200pt_tex2d& tex, float2 uv, bool is_srgb)
{
    float u = wrap_uv(uv.x);
    float v = wrap_uv(uv.y);
    v = 1.0f - v;

    int x = (int)(u * (tex.width - 1));
    int y = (int)(v * (tex.height - 1));

    float c[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    size_t base = ((size_t)x + (size_t)y * tex.width) * tex.channels;
    if (tex.bytes_per_channel == 1)
    {
        const unsigned char* p = reinterpret_cast<const unsigned char*>(tex.texels) + base;
        for (int i = 0; i < tex.channels; i++)
            c[i] = __ldg(p + i) / 255.0f;
    }
    else
    {
        const float* p = reinterpret_cast<const float*>(tex.texels) + base;
        for (int i = 0; i < tex.channels; i++)
            c[i] = __ldg(p + i);
    }

    if (is_srgb)   // pow(rgba, 2.2f) on all four channels (Texture.h:72-73)
        return RGBA(powf(c[0], 2.2f), powf(c[1], 2.2f), powf(c[2], 2.2f), powf(c[3], 2.2f));
    return RGBA(c[0], c[1], c[2], c[3]);
}

// First channel of a 1- (or more-) channel float LUT
PT_DEV float sample_lut2d_r(const void* descriptor, float u_in, float v_in)
{
    const b200pt_tex2d* tex = reinterpret_cast<const b200pt_tex2d*>(descriptor);
    float u = wrap_uv(u_in);
    float v = 1.0f - wrap_uv(v_in);
    int w = __ldg(&tex->width), h = __ldg(&tex->height), ch = __ldg(&tex->channels);
    int x = (int)(u * (w - 1));
    int y = (int)(v * (h - 1));
    const float* p = reinterpret_cast<const float*>(tex->texels);
    return __ldg(p + ((size_t)x + (size_t)y * w) * ch);
}

PT_DEV RGB sample_lut2d_rgb(const void* descriptor, float u_in, float v_in)
{
    const b200pt_tex2d* tex = reinterpret_cast<const b200pt_tex2d*>(descriptor);
    float u = wrap_uv(u_in);
    float v = 1.0f - wrap_uv(v_in);
    int w = __ldg(&tex->width), h = __ldg(&tex->height), ch = __ldg(&tex->channels);
    int x = (int)(u * (w - 1));
    int y = (int)(v * (h - 1));
    const float* p = reinterpret_cast<const float*>(tex->texels) + ((size_t)x + (size_t)y * w) * ch;
    RGB out;
    out.r = __ldg(p);
    out.g = ch > 1 ? __ldg(p + 1) : 0.0f;
    out.b = ch > 2 ? __ldg(p + 2) : 0.0f;
    return out;
}

PT_DEV float sample_lut3d_r(const void* descriptor, float3 uvw)
{
    const b200pt_tex3d* tex = reinterpret_cast<const b200pt_tex3d*>(descriptor);
    float u = wrap_uv(uvw.x);
    float v = 1.0f - wrap_uv(uvw.y);
    float w = wrap_uv(uvw.z);
    int width = __ldg(&tex->width), height = __ldg(&tex->height), depth = __ldg(&tex->depth), ch = __ldg(&tex->channels);
    int x = (int)(u * (width - 1));
    int y = (int)(v * (height - 1));
    int z = (int)(w * (depth - 1));
    const float* p = tex->texels;
    return __ldg(p + ((size_t)z * width * height + (size_t)x + (size_t)y * width) * ch);
}
} // namespace pt
