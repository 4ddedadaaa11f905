// ORACLE (test infrastructure — see volrt_oracle.h).  Scene of the reference's
// main(): source/Etna/TestMain.cpp:222-245, with the glm functions it calls
// restated from their documented closed forms (glm is not vendored in the
// reference and not installed here — SURVEY.md §8(c)).  Conventions come from
// source/Etna/Core/Vulkan/VulkanHeader.h:9-10: radians, depth 0..1 (=> perspectiveRH_ZO,
// lookAtRH).  All arithmetic is binary32 like glm's; matrices are column-major m[c*4+r].
#include "volrt_oracle.h"
#include <cmath>
#include <cstring>

namespace {

struct V3 { float x, y, z; };
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline V3 normalize(V3 a) { float inv = 1.0f / std::sqrt(dot(a, a)); return {a.x * inv, a.y * inv, a.z * inv}; }

void identity(float* m) { std::memset(m, 0, 64); m[0] = m[5] = m[10] = m[15] = 1.0f; }

// glm::rotate(m, angle, axis) = m * R(angle, axis)   (TestMain.cpp:222,224)
void rotate(const float* m, float angle, V3 v, float* out) {
    const float c = std::cos(angle), s = std::sin(angle);
    V3 axis = normalize(v);
    V3 temp = {(1.0f - c) * axis.x, (1.0f - c) * axis.y, (1.0f - c) * axis.z};
    float R[3][3];
    R[0][0] = c + temp.x * axis.x;
    R[0][1] = temp.x * axis.y + s * axis.z;
    R[0][2] = temp.x * axis.z - s * axis.y;
    R[1][0] = temp.y * axis.x - s * axis.z;
    R[1][1] = c + temp.y * axis.y;
    R[1][2] = temp.y * axis.z + s * axis.x;
    R[2][0] = temp.z * axis.x + s * axis.y;
    R[2][1] = temp.z * axis.y - s * axis.x;
    R[2][2] = c + temp.z * axis.z;
    float res[16];
    for (int col = 0; col < 3; ++col)
        for (int r = 0; r < 4; ++r)
            res[col * 4 + r] = m[0 * 4 + r] * R[col][0] + m[1 * 4 + r] * R[col][1] + m[2 * 4 + r] * R[col][2];
    for (int r = 0; r < 4; ++r) res[12 + r] = m[12 + r];
    std::memcpy(out, res, 64);
}

// glm::lookAtRH (TestMain.cpp:225)
void look_at(V3 eye, V3 center, V3 up, float* m) {
    V3 f = normalize(sub(center, eye));
    V3 s = normalize(cross(f, up));
    V3 u = cross(s, f);
    identity(m);
    m[0 * 4 + 0] = s.x; m[1 * 4 + 0] = s.y; m[2 * 4 + 0] = s.z;
    m[0 * 4 + 1] = u.x; m[1 * 4 + 1] = u.y; m[2 * 4 + 1] = u.z;
    m[0 * 4 + 2] = -f.x; m[1 * 4 + 2] = -f.y; m[2 * 4 + 2] = -f.z;
    m[3 * 4 + 0] = -dot(s, eye);
    m[3 * 4 + 1] = -dot(u, eye);
    m[3 * 4 + 2] = dot(f, eye);
}

// glm::perspectiveRH_ZO (TestMain.cpp:226; GLM_FORCE_DEPTH_ZERO_TO_ONE)
void perspective(float fovy, float aspect, float zn, float zf, float* m) {
    const float t = std::tan(fovy / 2.0f);
    std::memset(m, 0, 64);
    m[0 * 4 + 0] = 1.0f / (aspect * t);
    m[1 * 4 + 1] = 1.0f / t;
    m[2 * 4 + 2] = zf / (zn - zf);
    m[2 * 4 + 3] = -1.0f;
    m[3 * 4 + 2] = -(zf * zn) / (zf - zn);
}

inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }

}  // namespace

extern "C" void vo_mat4_inverse(const float* m, float* out) {
    // general cofactor inverse (glm::inverse is the same expansion); done in double
    // and rounded once so that an independent implementation lands on the same floats.
    double a[16], inv[16];
    for (int i = 0; i < 16; ++i) a[i] = m[i];
    inv[0] = a[5] * a[10] * a[15] - a[5] * a[11] * a[14] - a[9] * a[6] * a[15] + a[9] * a[7] * a[14] + a[13] * a[6] * a[11] - a[13] * a[7] * a[10];
    inv[4] = -a[4] * a[10] * a[15] + a[4] * a[11] * a[14] + a[8] * a[6] * a[15] - a[8] * a[7] * a[14] - a[12] * a[6] * a[11] + a[12] * a[7] * a[10];
    inv[8] = a[4] * a[9] * a[15] - a[4] * a[11] * a[13] - a[8] * a[5] * a[15] + a[8] * a[7] * a[13] + a[12] * a[5] * a[11] - a[12] * a[7] * a[9];
    inv[12] = -a[4] * a[9] * a[14] + a[4] * a[10] * a[13] + a[8] * a[5] * a[14] - a[8] * a[6] * a[13] - a[12] * a[5] * a[10] + a[12] * a[6] * a[9];
    inv[1] = -a[1] * a[10] * a[15] + a[1] * a[11] * a[14] + a[9] * a[2] * a[15] - a[9] * a[3] * a[14] - a[13] * a[2] * a[11] + a[13] * a[3] * a[10];
    inv[5] = a[0] * a[10] * a[15] - a[0] * a[11] * a[14] - a[8] * a[2] * a[15] + a[8] * a[3] * a[14] + a[12] * a[2] * a[11] - a[12] * a[3] * a[10];
    inv[9] = -a[0] * a[9] * a[15] + a[0] * a[11] * a[13] + a[8] * a[1] * a[15] - a[8] * a[3] * a[13] - a[12] * a[1] * a[11] + a[12] * a[3] * a[9];
    inv[13] = a[0] * a[9] * a[14] - a[0] * a[10] * a[13] - a[8] * a[1] * a[14] + a[8] * a[2] * a[13] + a[12] * a[1] * a[10] - a[12] * a[2] * a[9];
    inv[2] = a[1] * a[6] * a[15] - a[1] * a[7] * a[14] - a[5] * a[2] * a[15] + a[5] * a[3] * a[14] + a[13] * a[2] * a[7] - a[13] * a[3] * a[6];
    inv[6] = -a[0] * a[6] * a[15] + a[0] * a[7] * a[14] + a[4] * a[2] * a[15] - a[4] * a[3] * a[14] - a[12] * a[2] * a[7] + a[12] * a[3] * a[6];
    inv[10] = a[0] * a[5] * a[15] - a[0] * a[7] * a[13] - a[4] * a[1] * a[15] + a[4] * a[3] * a[13] + a[12] * a[1] * a[7] - a[12] * a[3] * a[5];
    inv[14] = -a[0] * a[5] * a[14] + a[0] * a[6] * a[13] + a[4] * a[1] * a[14] - a[4] * a[2] * a[13] - a[12] * a[1] * a[6] + a[12] * a[2] * a[5];
    inv[3] = -a[1] * a[6] * a[11] + a[1] * a[7] * a[10] + a[5] * a[2] * a[11] - a[5] * a[3] * a[10] - a[9] * a[2] * a[7] + a[9] * a[3] * a[6];
    inv[7] = a[0] * a[6] * a[11] - a[0] * a[7] * a[10] - a[4] * a[2] * a[11] + a[4] * a[3] * a[10] + a[8] * a[2] * a[7] - a[8] * a[3] * a[6];
    inv[11] = -a[0] * a[5] * a[11] + a[0] * a[7] * a[9] + a[4] * a[1] * a[11] - a[4] * a[3] * a[9] - a[8] * a[1] * a[7] + a[8] * a[3] * a[5];
    inv[15] = a[0] * a[5] * a[10] - a[0] * a[6] * a[9] - a[4] * a[1] * a[10] + a[4] * a[2] * a[9] + a[8] * a[1] * a[6] - a[8] * a[2] * a[5];
    double det = a[0] * inv[0] + a[1] * inv[4] + a[2] * inv[8] + a[3] * inv[12];
    double r = 1.0 / det;
    for (int i = 0; i < 16; ++i) out[i] = (float)(inv[i] * r);
}

extern "C" void vo_scene_default(float phi_deg, float theta_deg, float frame_time,
                                 float* ubo0, float* ubo1_host) {
    float I[16], rot[16];
    identity(I);
    rotate(I, radians(phi_deg), {0.0f, 0.0f, 1.0f}, rot);                       // TestMain.cpp:222
    rotate(rot, radians(theta_deg), {0.0f, 1.0f, 0.0f}, ubo0 + 0);              // :224  Model
    look_at({3.0f, 3.0f, 3.0f}, {0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 1.0f}, ubo0 + 16);  // :225  View
    perspective(radians(45.0f), (float)1280 / (float)720, 0.1f, 10.0f, ubo0 + 32);   // :226  Projection
    ubo0[32 + 1 * 4 + 1] *= -1.0f;                                              // :228
    // GlobalShaderData as the C++ compiler lays it out (TestMain.cpp:34-39): 140 bytes,
    // mat4 @0, vec3 @64, mat4 @76 — NOT std140 (quirk Q2).
    std::memset(ubo1_host, 0, 140);
    vo_mat4_inverse(ubo0, ubo1_host);                                           // :230
    ubo1_host[16] = 3.0f; ubo1_host[17] = 3.0f; ubo1_host[18] = 3.0f;           // :242
    ubo1_host[19] = -frame_time;                                                // :233-238  MediaScroll[0][0]
}

extern "C" void vo_decode_ubo1(const void* bytes, uint32_t size, float* out36) {
    // std140: mat4 @0, vec3 @64 (+pad), mat4 @80; block = 144 B (frag.glsl:9-14).
    // The descriptor range is sizeof(GlobalShaderData)=140 (TestMain.cpp:150); bytes
    // beyond `size` read as zero (robust buffer access).
    unsigned char img[144];
    std::memset(img, 0, sizeof img);
    std::memcpy(img, bytes, size < 144u ? size : 144u);
    std::memcpy(out36, img, 144);
}

extern "C" void vo_params_default(vo_params* p) {
    std::memset(p, 0, sizeof *p);
    float host[35];
    vo_scene_default(0.0f, 0.0f, 0.0f, p->ubo0, host);
    vo_decode_ubo1(host, 140, p->ubo1);
    p->max_steps = 128;          // frag.glsl:30
    p->step_scale = 4.0f;        // frag.glsl:42
    p->density = 1.0f;           // frag.glsl:29
    p->sample_scale = 0.2f;      // frag.glsl:63
    const float cs[4] = {1.0f, 0.8f, 0.75f, 0.7f}, sw[4] = {0.0f, 0.2f, 0.25f, 0.3f};  // :66-69
    for (int i = 0; i < 4; ++i) { p->coord_scale[i] = cs[i]; p->scroll_weight[i] = sw[i]; }
    p->sun_color[0] = p->sun_color[1] = p->sun_color[2] = 1.0f;
    p->sun_dir[2] = 1.0f;
    p->fbm_octaves = 5; p->fbm_seed = 1337; p->fbm_frequency = 128.0f * 0.03f;
    p->fbm_gain = 0.5f; p->fbm_lacunarity = 2.0f; p->fbm_coverage = 0.45f;
    p->fbm_falloff0 = 0.35f; p->fbm_falloff1 = 0.5f;
}
