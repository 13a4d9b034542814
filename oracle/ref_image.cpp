// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
//
// The reference's 8-bit texture loader (Image8Bit::read_image over stb_image, src/Image/Image.cpp:33-61), its constant-colour test
// (Image.cpp:250-271) and its texel fetch (Image.cpp:161-195), compiled from where they lie: the checker of the texture import of
// hiprt-path-tracer_b200/scene.py. Built by oracle/build_ref.py into oracle/_ref/libhippt_ref_image.so.
#include "Image/Image.h"

#include <cstring>

extern "C" int ref_read_image_8bit(const char* path, int output_channels, unsigned char* out, long long capacity, int* out_width, int* out_height, int* out_constant_threshold_5,
                                   float* out_rgba_at_uv00)
{
    Image8Bit image = Image8Bit::read_image(path, output_channels, false);
    if (image.width == 0 || image.height == 0)
        return -1;
    *out_width = image.width;
    *out_height = image.height;
    if (out_constant_threshold_5)
        *out_constant_threshold_5 = image.is_constant_color(5) ? 1 : 0;
    if (out_rgba_at_uv00)
    {
        ColorRGBA32F c = image.sample_rgba32f(make_float2(0, 0));
        out_rgba_at_uv00[0] = c.r; out_rgba_at_uv00[1] = c.g; out_rgba_at_uv00[2] = c.b; out_rgba_at_uv00[3] = c.a;
    }
    long long bytes = (long long)image.width * image.height * output_channels;
    if (out != nullptr)
    {
        if (bytes > capacity)
            return -2;
        std::memcpy(out, image.data().data(), (size_t)bytes);
    }
    return 0;
}
