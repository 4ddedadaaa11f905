// ORACLE (test infrastructure — see volrt_oracle.h).  A GLSL-450-as-C++ shim: just enough of the
// shading language for g++ to compile the REFERENCE's own shader text — /root/reference/shaders/frag.glsl and
// vert.glsl, read from where they lie (oracle/Makefile `ref` pipes them through sed, which only drops the
// `#version` line, into a translation unit that starts with this header) — into oracle/_ref/libvolrt_ref.so.
// No shader source is copied into the repo.  What the shim supplies:
//   * vec2 / vec3 / vec4 / mat4 with the constructors and operators those two files use, column-major mat4;
//   * min / max / abs / exp / normalize / distance / dot / length;
//   * sampler3D + texture(): the fixed-function sampler the reference creates (VulkanCore.cpp:676-710: LINEAR
//     mag/min, MIRRORED_REPEAT on U/V/W, normalised coordinates, one mip) over an RGBA8 UNORM image with the
//     layout of VulkanTexture.cpp:111-156 / VulkanCore.cpp:621-648 (tight, x fastest);
//   * the storage-qualifier keywords as macros (GLSL_SHIM_KEYWORDS): layout(...) -> nothing, uniform -> struct
//     (so `uniform Block {...} name;` declares a struct and one instance), in / out -> thread_local globals,
//     main -> shader_main.
// Where GLSL leaves the arithmetic to the implementation the shim takes one rounded binary32 operation per
// operator in the written order (the TU is compiled -ffp-contract=off), mat4*vec4 as the sum of scaled columns
// from column 0 up, dot as (x*x + y*y) + z*z, normalize(v) = v * (1 / sqrt(dot(v, v))), exp correctly rounded
// (double exp, rounded once), UNORM8 decode t / 255, trilinear weights in float in x, y, z order.
#ifndef VOLRT_GLSL_SHIM_H
#define VOLRT_GLSL_SHIM_H

#include <atomic>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace glsl {

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    vec2(float a, float b) : x(a), y(b) {}
};

struct vec4;
struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit vec3(const vec4& v);                       // vec3(vec4) drops w
    vec3& operator+=(const vec3& o) { x = x + o.x; y = y + o.y; z = z + o.z; return *this; }
    vec3& operator-=(const vec3& o) { x = x - o.x; y = y - o.y; z = z - o.z; return *this; }
    vec3& operator*=(const vec3& o) { x = x * o.x; y = y * o.y; z = z * o.z; return *this; }
    vec3& operator/=(const vec3& o) { x = x / o.x; y = y / o.y; z = z / o.z; return *this; }
    vec3& operator+=(float s) { x = x + s; y = y + s; z = z + s; return *this; }
    vec3& operator*=(float s) { x = x * s; y = y * s; z = z * s; return *this; }
};
struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(const vec3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
};
inline vec3::vec3(const vec4& v) : x(v.x), y(v.y), z(v.z) {}

inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(const vec3& a, const vec3& b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }

// column-major: m[c] is column c, m[c].x its row 0 (GLSL's indexing: MediaScroll[0].y = column 0, row 1)
struct mat4 {
    vec4 col[4];
    vec4& operator[](int c) { return col[c]; }
    const vec4& operator[](int c) const { return col[c]; }
};
inline vec4 operator*(const mat4& m, const vec4& v) {
    return ((m[0] * v.x + m[1] * v.y) + m[2] * v.z) + m[3] * v.w;
}
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r;
    for (int c = 0; c < 4; ++c) r[c] = a * b[c];
    return r;
}
inline mat4 mat4_from_floats(const float* f16) {
    mat4 m;
    for (int c = 0; c < 4; ++c) m[c] = vec4(f16[4 * c + 0], f16[4 * c + 1], f16[4 * c + 2], f16[4 * c + 3]);
    return m;
}

inline float min(float a, float b) { return std::fmin(a, b); }
inline float max(float a, float b) { return std::fmax(a, b); }
inline int min(int a, int b) { return b < a ? b : a; }
inline int max(int a, int b) { return a < b ? b : a; }
inline vec3 min(const vec3& a, const vec3& b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(const vec3& a, const vec3& b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float abs(float a) { return std::fabs(a); }
inline vec3 abs(const vec3& a) { return vec3(std::fabs(a.x), std::fabs(a.y), std::fabs(a.z)); }
inline float exp(float a) { return (float)std::exp((double)a); }
inline vec3 exp(const vec3& a) { return vec3(exp(a.x), exp(a.y), exp(a.z)); }
inline float dot(const vec3& a, const vec3& b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(const vec3& a) { return std::sqrt(dot(a, a)); }
inline float distance(const vec3& a, const vec3& b) { return length(a - b); }
inline vec3 normalize(const vec3& a) { const float inv = 1.0f / std::sqrt(dot(a, a)); return a * inv; }

// ---- the combined image sampler of binding 2
struct sampler3D {
    const uint8_t* texels = nullptr;      // R8G8B8A8_UNORM, texel (i,j,k) at 4*((k*h + j)*w + i)
    int w = 0, h = 0, d = 0;
};
// VK_SAMPLER_ADDRESS_MODE_MIRRORED_REPEAT as the Vulkan specification writes it:
//   i -> (size - 1) - mirror((i mod 2*size) - size),   mirror(n) = n >= 0 ? n : -(1 + n)
inline int mirrored_repeat(int i, int size) {
    int m = i % (2 * size);
    if (m < 0) m += 2 * size;
    const int n = m - size;
    return (size - 1) - (n >= 0 ? n : -(1 + n));
}
inline vec4 fetch_texel(const sampler3D& s, int i, int j, int k) {
    const uint8_t* t = s.texels + 4 * (((size_t)k * s.h + j) * s.w + i);
    return vec4((float)t[0] / 255.0f, (float)t[1] / 255.0f, (float)t[2] / 255.0f, (float)t[3] / 255.0f);
}
inline float lerp3(const float t[8], float fx, float fy, float fz) {
    const float wx0 = 1.0f - fx, wy0 = 1.0f - fy, wz0 = 1.0f - fz;
    const float a = (t[0] * wx0 + t[1] * fx) * wy0 + (t[2] * wx0 + t[3] * fx) * fy;
    const float b = (t[4] * wx0 + t[5] * fx) * wy0 + (t[6] * wx0 + t[7] * fx) * fy;
    return a * wz0 + b * fz;
}
// VK_FILTER_LINEAR, normalised coordinates: unnormalised u = s * size, then i0 = floor(u - 1/2), weight = frac(u - 1/2)
inline vec4 texture(const sampler3D& s, const vec3& P) {
    const float x = P.x * (float)s.w - 0.5f, y = P.y * (float)s.h - 0.5f, z = P.z * (float)s.d - 0.5f;
    const float x0 = std::floor(x), y0 = std::floor(y), z0 = std::floor(z);
    const float fx = x - x0, fy = y - y0, fz = z - z0;
    const int i0 = mirrored_repeat((int)x0, s.w), i1 = mirrored_repeat((int)x0 + 1, s.w);
    const int j0 = mirrored_repeat((int)y0, s.h), j1 = mirrored_repeat((int)y0 + 1, s.h);
    const int k0 = mirrored_repeat((int)z0, s.d), k1 = mirrored_repeat((int)z0 + 1, s.d);
    const vec4 c[8] = {fetch_texel(s, i0, j0, k0), fetch_texel(s, i1, j0, k0), fetch_texel(s, i0, j1, k0), fetch_texel(s, i1, j1, k0),
                       fetch_texel(s, i0, j0, k1), fetch_texel(s, i1, j0, k1), fetch_texel(s, i0, j1, k1), fetch_texel(s, i1, j1, k1)};
    float r[8], g[8], b[8], a[8];
    for (int q = 0; q < 8; ++q) { r[q] = c[q].x; g[q] = c[q].y; b[q] = c[q].z; a[q] = c[q].w; }
    return vec4(lerp3(r, fx, fy, fz), lerp3(g, fx, fy, fz), lerp3(b, fx, fy, fz), lerp3(a, fx, fy, fz));
}

// GLSL reads a decimal literal without suffix as float; C++ reads it as double and converts at the call.  Both
// routes give the same binary32 for the literals the shaders use:
static_assert((float)0.8 == 0.8f && (float)0.2 == 0.2f && (float)0.75 == 0.75f && (float)0.25 == 0.25f &&
              (float)0.7 == 0.7f && (float)0.3 == 0.3f && (float)1.0 == 1.0f, "double-rounded literal");

}  // namespace glsl

#endif  // VOLRT_GLSL_SHIM_H

// ---- storage qualifiers as macros; defined AFTER every standard header, immediately before the shader text
#ifdef GLSL_SHIM_KEYWORDS
#define layout(...)
#define uniform struct
#define in thread_local
#define out thread_local
#define main shader_main
#endif
#ifdef GLSL_SHIM_KEYWORDS_OFF
#undef layout
#undef uniform
#undef in
#undef out
#undef main
#endif
