// ORACLE (test infrastructure).  Epilogue of the translation unit that holds the reference's vert.glsl compiled as C++
// (see ref_glsl_frag_api.h).  The shader's globals: osd (binding 0), inPosition / inTexCoord (in), fragPosition /
// fragTexCoord (out); gl_Position is declared by the prologue the Makefile writes.
#define GLSL_SHIM_KEYWORDS_OFF
#include "glsl_shim.h"      // (second inclusion: only switches the keyword macros off; the std headers it needs came in with the first)

namespace {
std::mutex g_vert_mu;
}

// ObjectShaderData (vert.glsl:4-9): three mat4 at 0 / 64 / 128, 192 bytes, identical in std140 and in C++
extern "C" int vref_vert(const void* ubo0_bytes, uint32_t ubo0_size, const float* in_positions, int n, float* out_gl_position,
                         float* out_frag_position) {
    if (!ubo0_bytes || !in_positions || n < 0) return 1;
    float u[48];
    std::memset(u, 0, sizeof u);
    std::memcpy(u, ubo0_bytes, ubo0_size < 192u ? ubo0_size : 192u);
    std::lock_guard<std::mutex> lock(g_vert_mu);
    osd.Model = glsl::mat4_from_floats(u);
    osd.View = glsl::mat4_from_floats(u + 16);
    osd.Projection = glsl::mat4_from_floats(u + 32);
    for (int i = 0; i < n; ++i) {
        inPosition = glsl::vec3(in_positions[3 * i], in_positions[3 * i + 1], in_positions[3 * i + 2]);
        inTexCoord = glsl::vec2(0.0f, 0.0f);
        shader_main();
        if (out_gl_position) { float* g = out_gl_position + 4 * i; g[0] = gl_Position.x; g[1] = gl_Position.y; g[2] = gl_Position.z; g[3] = gl_Position.w; }
        if (out_frag_position) { float* f = out_frag_position + 3 * i; f[0] = fragPosition.x; f[1] = fragPosition.y; f[2] = fragPosition.z; }
    }
    return 0;
}
