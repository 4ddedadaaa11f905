// ORACLE (test infrastructure).  Epilogue of the translation unit that holds the reference's frag.glsl compiled as C++
// (oracle/Makefile `ref`: glsl_shim.h + keywords, then the shader text from /root/reference/shaders/frag.glsl, then this
// file, all inside namespace glsl::frag).  C entry points over the shader's own globals: gsd (binding 1), texSampler
// (binding 2), fragPosition (location 0 in), outColor (location 0 out), shader_main() = the shader's main().
#define GLSL_SHIM_KEYWORDS_OFF
#include "glsl_shim.h"      // (second inclusion: only switches the keyword macros off; the std headers it needs came in with the first)

namespace {
std::mutex g_frag_mu;       // the uniforms are process globals, as in a bound descriptor set: one draw at a time

// std140 image of GlobalShaderData as frag.glsl:9-14 declares it: mat4 @0, vec3 @64 (4 pad bytes), mat4 @80; 144 bytes.
// Bytes beyond `size` read as zero (the reference binds a 140-byte range, TestMain.cpp:150).
void bind_uniforms(const void* ubo1_bytes, uint32_t size, const uint8_t* texels, int w, int h, int d) {
    float u[36];
    std::memset(u, 0, sizeof u);
    std::memcpy(u, ubo1_bytes, size < 144u ? size : 144u);
    gsd.WorldToLocal = glsl::mat4_from_floats(u);
    gsd.CameraPosition = glsl::vec3(u[16], u[17], u[18]);
    gsd.MediaScroll = glsl::mat4_from_floats(u + 20);
    texSampler.texels = texels; texSampler.w = w; texSampler.h = h; texSampler.d = d;
}
}  // namespace

// one fragment-shader invocation per entry of frag_positions[n][3]; entries with covered[i] == 0 get the render pass's
// clear colour (0,0,0,1) (VulkanRenderPass.cpp:17-19) and the shader is not run for them
extern "C" int vref_frag(const void* ubo1_bytes, uint32_t ubo1_size, const uint8_t* texels, int w, int h, int d,
                         const float* frag_positions, const int32_t* covered, int64_t n, float* out_rgba, int threads) {
    if (!ubo1_bytes || !texels || w <= 0 || h <= 0 || d <= 0 || !frag_positions || !out_rgba || n < 0) return 1;
    std::lock_guard<std::mutex> lock(g_frag_mu);
    bind_uniforms(ubo1_bytes, ubo1_size, texels, w, h, d);
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    std::atomic<int64_t> next{0};
    auto work = [&] {
        const int64_t chunk = 256;
        for (int64_t b; (b = next.fetch_add(chunk)) < n;) {
            const int64_t e = b + chunk < n ? b + chunk : n;
            for (int64_t i = b; i < e; ++i) {
                float* o = out_rgba + 4 * i;
                if (covered && !covered[i]) { o[0] = o[1] = o[2] = 0.0f; o[3] = 1.0f; continue; }
                fragPosition = glsl::vec3(frag_positions[3 * i], frag_positions[3 * i + 1], frag_positions[3 * i + 2]);
                shader_main();
                o[0] = outColor.x; o[1] = outColor.y; o[2] = outColor.z; o[3] = outColor.w;
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(work);
    work();
    for (auto& th : pool) th.join();
    return 0;
}

// the literals the shader was compiled with (frag.glsl:29-32), so a test can tell if the reference changes them
extern "C" void vref_frag_constants(float out4[4]) {
    out4[0] = density; out4[1] = (float)maxSteps; out4[2] = boxMin.x; out4[3] = boxMax.x;
}
