// The handful of glm functions the reference's main() calls (source/Etna/TestMain.cpp:222-230), restated
// from their closed forms because glm is a system include of the reference and is not installed here.
// Conventions of source/Etna/Core/Vulkan/VulkanHeader.h:9-10: radians, depth 0..1 (=> perspectiveRH_ZO),
// right-handed lookAt.  Column-major: m[c][r], 16 contiguous floats like glm::mat4.
#pragma once
#include <cmath>
#include <cstring>

namespace glm {

struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};
static_assert(sizeof(vec3) == 12, "glm::vec3 is 12 bytes without GLM_FORCE_DEFAULT_ALIGNED_GENTYPES");

struct vec4 {
    float x, y, z, w;
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

struct mat4 {
    vec4 c[4];
    mat4() { std::memset(c, 0, sizeof c); }
    explicit mat4(float d) { std::memset(c, 0, sizeof c); c[0].x = c[1].y = c[2].z = c[3].w = d; }
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
    const float* data() const { return &c[0].x; }
    float* data() { return &c[0].x; }
};
static_assert(sizeof(mat4) == 64, "mat4 is 16 floats");

inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline vec3 normalize(vec3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }

// glm::rotate(m, angle, axis) = m * R(angle, axis)
inline mat4 rotate(const mat4& m, float angle, vec3 v) {
    const float c = std::cos(angle), s = std::sin(angle);
    const vec3 axis = normalize(v);
    const vec3 t = axis * (1.0f - c);
    float R[3][3];
    R[0][0] = c + t.x * axis.x;          R[0][1] = t.x * axis.y + s * axis.z;  R[0][2] = t.x * axis.z - s * axis.y;
    R[1][0] = t.y * axis.x - s * axis.z; R[1][1] = c + t.y * axis.y;           R[1][2] = t.y * axis.z + s * axis.x;
    R[2][0] = t.z * axis.x + s * axis.y; R[2][1] = t.z * axis.y - s * axis.x;  R[2][2] = c + t.z * axis.z;
    mat4 out;
    for (int col = 0; col < 3; ++col)
        for (int r = 0; r < 4; ++r)
            out[col][r] = m[0][r] * R[col][0] + m[1][r] * R[col][1] + m[2][r] * R[col][2];
    out[3] = m[3];
    return out;
}

inline mat4 lookAt(vec3 eye, vec3 center, vec3 up) {       // lookAtRH
    const vec3 f = normalize(center - eye), s = normalize(cross(f, up)), u = cross(s, f);
    mat4 m(1.0f);
    m[0][0] = s.x; m[1][0] = s.y; m[2][0] = s.z;
    m[0][1] = u.x; m[1][1] = u.y; m[2][1] = u.z;
    m[0][2] = -f.x; m[1][2] = -f.y; m[2][2] = -f.z;
    m[3][0] = -dot(s, eye); m[3][1] = -dot(u, eye); m[3][2] = dot(f, eye);
    return m;
}

inline mat4 perspective(float fovy, float aspect, float zNear, float zFar) {   // perspectiveRH_ZO
    const float t = std::tan(fovy / 2.0f);
    mat4 m;
    m[0][0] = 1.0f / (aspect * t);
    m[1][1] = 1.0f / t;
    m[2][2] = zFar / (zNear - zFar);
    m[2][3] = -1.0f;
    m[3][2] = -(zFar * zNear) / (zFar - zNear);
    return m;
}

inline mat4 inverse(const mat4& mm) {       // general 4x4 by cofactors, accumulated in double
    const float* a = mm.data();
    double inv[16], m[16];
    for (int i = 0; i < 16; ++i) m[i] = a[i];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    const double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    mat4 out;
    for (int i = 0; i < 16; ++i) out.data()[i] = (float)(inv[i] / det);
    return out;
}

}  // namespace glm
