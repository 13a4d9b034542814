// Small fp32 vector / colour algebra for the B200 path-tracing kernels.
//
// The expressions follow the evaluation ORDER of the reference's hippt:: helpers
// (reference: src/HostDeviceCommon/Math.h:61-235, thirdparties/HIPRT-Fork/hiprt/impl/Math.h:1078-1085,
// src/HostDeviceCommon/Color.h:64-110) so that, compiled with -fmad=false, results agree with the
// CPU oracle (g++ -ffp-contract=off) to the last bit wherever only + - * / sqrt are involved.
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define PT_DEV __device__ __forceinline__
#define PT_HD __host__ __device__ __forceinline__
// Large functions called from many sites: a real call keeps the megakernel's code size (and ptxas time) bounded.
#define PT_DEV_NOINLINE static __device__ __noinline__
// The BSDF entry points (principled eval / sample, glass eval, thin-film Fresnel). Out of line they take the 256-byte material
// by reference, which pins it (and the spills around the call) in local memory; PT_BSDF_INLINE lets the compiler keep the
// fields a kernel uses in registers at the price of code size.
#ifdef PT_BSDF_INLINE
#define PT_BSDF_FN static __device__ __forceinline__
#else
#define PT_BSDF_FN static __device__ __noinline__
#endif

namespace pt
{
constexpr float kPi = 3.14159265358979323846f;
constexpr float kTwoPi = 6.28318530717958647693f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInv2Pi = 0.15915494309189533577f;
constexpr float kTwoPiPi = 19.73920880217871723767f;
constexpr float kNearZero = 1.0e-10f;

// ---- float3 ---------------------------------------------------------------------------------
PT_HD float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
PT_HD float3 f3(float v) { return make_float3(v, v, v); }
PT_HD float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
PT_HD float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
PT_HD float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
PT_HD float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
PT_HD float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
PT_HD float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
PT_HD float3 operator/(float3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
PT_HD float3 operator/(float3 a, float3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
PT_HD float3& operator+=(float3& a, float3 b) { a = a + b; return a; }
PT_HD float3& operator*=(float3& a, float s) { a = a * s; return a; }
PT_HD float3& operator/=(float3& a, float s) { a = a / s; return a; }

PT_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PT_HD float3 cross(float3 a, float3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
PT_HD float length(float3 a) { return sqrtf(dot(a, a)); }
PT_HD float3 normalize(float3 a) { return a / sqrtf(dot(a, a)); }
PT_HD float3 abs3(float3 a) { return f3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }

// hippt::min / max on floats resolve to hiprt::min/max = fminf / fmaxf on the CPU path
// (thirdparties/HIPRT-Fork/hiprt/impl/Math.h:778-780): a NaN operand is dropped. Same here.
PT_HD float minf(float a, float b) { return fminf(a, b); }
PT_HD float maxf(float a, float b) { return fmaxf(a, b); }
PT_HD int mini(int a, int b) { return a < b ? a : b; }
PT_HD int maxi(int a, int b) { return a > b ? a : b; }
// hippt::clamp(min, max, val) = min(max_val, max(min_val, val))
PT_HD float clampf(float lo, float hi, float v) { return minf(hi, maxf(lo, v)); }
PT_HD float square(float x) { return x * x; }
PT_HD float pow_1_4(float x) { return sqrtf(sqrtf(x)); }
PT_HD float pow_3(float x) { return x * x * x; }
PT_HD float pow_4(float x) { float x2 = x * x; return x2 * x2; }
PT_HD float pow_5(float x) { float x2 = x * x; float x4 = x2 * x2; return x4 * x; }
PT_HD float pow_6(float x) { float x2 = x * x; float x4 = x2 * x2; return x4 * x2; }
PT_HD bool is_zero(float x) { return x < kNearZero && x > -kNearZero; }
PT_HD float lerpf(float a, float b, float t) { return (1.0f - t) * a + t * b; }
PT_HD float fractf(float a) { return a - floorf(a); }

// ---- colour ---------------------------------------------------------------------------------
struct RGB
{
    float r, g, b;
    PT_HD RGB() : r(0.0f), g(0.0f), b(0.0f) {}
    PT_HD explicit RGB(float v) : r(v), g(v), b(v) {}
    PT_HD RGB(float r_, float g_, float b_) : r(r_), g(g_), b(b_) {}
    PT_HD explicit RGB(float3 v) : r(v.x), g(v.y), b(v.z) {}

    PT_HD void operator+=(RGB o) { r += o.r; g += o.g; b += o.b; }
    PT_HD void operator*=(RGB o) { r *= o.r; g *= o.g; b *= o.b; }
    PT_HD void operator*=(float k) { r *= k; g *= k; b *= k; }
    PT_HD void operator/=(RGB o) { r /= o.r; g /= o.g; b /= o.b; }
    PT_HD void operator/=(float k) { r /= k; g /= k; b /= k; }
    PT_HD float luminance() const { return 0.3086f * r + 0.6094f * g + 0.0820f * b; }
    PT_HD void clamp(float lo, float hi) { r = clampf(lo, hi, r); g = clampf(lo, hi, g); b = clampf(lo, hi, b); }
    PT_HD bool has_NaN() const { return isnan(r) || isnan(g) || isnan(b); }
    PT_HD bool is_black() const { return !(r > 0.0f || g > 0.0f || b > 0.0f); }
    PT_HD bool is_white() const { return r == 1.0f && g == 1.0f && b == 1.0f; }
    PT_HD float max_component() const { return maxf(r, maxf(g, b)); }
    PT_HD float operator[](int i) const { return i == 0 ? r : (i == 1 ? g : b); }
};

PT_HD RGB operator+(RGB a, RGB b) { return RGB(a.r + b.r, a.g + b.g, a.b + b.b); }
PT_HD RGB operator-(RGB a, RGB b) { return RGB(a.r - b.r, a.g - b.g, a.b - b.b); }
PT_HD RGB operator-(RGB a) { return RGB(-a.r, -a.g, -a.b); }
PT_HD RGB operator*(RGB a, RGB b) { return RGB(a.r * b.r, a.g * b.g, a.b * b.b); }
PT_HD RGB operator*(float k, RGB c) { return RGB(c.r * k, c.g * k, c.b * k); }
PT_HD RGB operator*(RGB c, float k) { return RGB(c.r * k, c.g * k, c.b * k); }
PT_HD RGB operator/(RGB a, RGB b) { return RGB(a.r / b.r, a.g / b.g, a.b / b.b); }
PT_HD RGB operator/(float k, RGB c) { return RGB(k / c.r, k / c.g, k / c.b); }
PT_HD RGB operator/(RGB c, float k) { return RGB(c.r / k, c.g / k, c.b / k); }
PT_HD RGB sqrt(RGB c) { return RGB(sqrtf(c.r), sqrtf(c.g), sqrtf(c.b)); }
PT_HD RGB exp(RGB c) { return RGB(expf(c.r), expf(c.g), expf(c.b)); }
PT_HD RGB log(RGB c) { return RGB(logf(c.r), logf(c.g), logf(c.b)); }
PT_HD RGB pow(RGB c, float k) { return RGB(powf(c.r, k), powf(c.g, k), powf(c.b, k)); }
PT_HD RGB lerp(RGB a, RGB b, float t) { return (1.0f - t) * a + t * b; }
PT_HD RGB rgb_max(RGB a, RGB b) { return RGB(maxf(a.r, b.r), maxf(a.g, b.g), maxf(a.b, b.b)); }

struct RGBA
{
    float r, g, b, a;
    PT_HD RGBA() : r(0.0f), g(0.0f), b(0.0f), a(1.0f) {}
    PT_HD RGBA(float r_, float g_, float b_, float a_) : r(r_), g(g_), b(b_), a(a_) {}
};

// ---- matrices (row-major float m[4][4]; reference: src/HostDeviceCommon/Math.h:237-290) ------
struct Mat4 { float m[4][4]; };

PT_HD float3 matrix_X_point(const Mat4& m, float3 p)
{
    float xt = m.m[0][0] * p.x + m.m[0][1] * p.y + m.m[0][2] * p.z + m.m[0][3];
    float yt = m.m[1][0] * p.x + m.m[1][1] * p.y + m.m[1][2] * p.z + m.m[1][3];
    float zt = m.m[2][0] * p.x + m.m[2][1] * p.y + m.m[2][2] * p.z + m.m[2][3];
    float wt = m.m[3][0] * p.x + m.m[3][1] * p.y + m.m[3][2] * p.z + m.m[3][3];
    float inv_w = 1.0f;
    if (!is_zero(wt))
        inv_w = 1.0f / wt;
    return f3(xt * inv_w, yt * inv_w, zt * inv_w);
}

// NOTE: multiplies by the TRANSPOSE (columns), exactly like the reference's float4x4 overload
PT_HD float3 matrix_X_vec(const Mat4& m, float3 u)
{
    float xt = m.m[0][0] * u.x + m.m[1][0] * u.y + m.m[2][0] * u.z;
    float yt = m.m[0][1] * u.x + m.m[1][1] * u.y + m.m[2][1] * u.z;
    float zt = m.m[0][2] * u.x + m.m[1][2] * u.y + m.m[2][2] * u.z;
    float wt = m.m[0][3] * u.x + m.m[1][3] * u.y + m.m[2][3] * u.z;
    float inv_w = 1.0f;
    if (!is_zero(wt))
        inv_w = 1.0f / wt;
    return f3(xt * inv_w, yt * inv_w, zt * inv_w);
}

// ---- loads of the reference's unaligned 12-byte float3 / colour arrays -----------------------
PT_DEV float3 ld3(const void* base, size_t index)
{
    const float* p = reinterpret_cast<const float*>(base) + index * 3;
    return f3(__ldg(p), __ldg(p + 1), __ldg(p + 2));
}
PT_DEV void st3(void* base, size_t index, float3 v)
{
    float* p = reinterpret_cast<float*>(base) + index * 3;
    p[0] = v.x; p[1] = v.y; p[2] = v.z;
}
PT_DEV float3 ld3_rw(const void* base, size_t index)
{
    const float* p = reinterpret_cast<const float*>(base) + index * 3;
    return f3(p[0], p[1], p[2]);
}
PT_HD float3 to_f3(RGB c) { return f3(c.r, c.g, c.b); }
} // namespace pt
