// ORACLE (test infrastructure).  The reference's draw, end to end on the CPU: its own vert.glsl and frag.glsl (compiled as
// C++ from /root/reference/shaders by oracle/Makefile, see glsl_shim.h) joined by a restatement of the fixed-function
// stages its pipeline configures.  Lives in oracle/_ref/libvolrt_ref.so next to the compiled shaders.
//   geometry     the cube's 8 vertices and 36 indices                  source/Etna/TestMain.cpp:94-112
//   draw         vkCmdDrawIndexed(36) with viewport (0,0,W,H), depth 0..1        TestMain.cpp:198-216
//   assembly     TRIANGLE_LIST                                          Core/Vulkan/VulkanPipeline.cpp:96
//   rasteriser   cull BACK, front = COUNTER_CLOCKWISE, fill              VulkanPipeline.cpp:107-108
//   depth        LESS against a buffer cleared to 1                      VulkanPipeline.cpp:130-132, VulkanRenderPass.cpp:48
//   blend off, clear colour (0,0,0,1)                                   VulkanPipeline.cpp:144, VulkanRenderPass.cpp:17-19
// Rasterisation rules follow the Vulkan specification: framebuffer coordinates x_f = (x_ndc + 1) W/2, y_f = (y_ndc + 1) H/2;
// facing from the sign of a = -1/2 sum(x_i y_j - x_j y_i) (positive = counter-clockwise = front here); a pixel is
// covered when its centre (px + 1/2, py + 1/2) is inside the triangle, ties by the top-left rule; attributes are
// interpolated perspective-correctly: f = sum(b_i f_i / w_i) / sum(b_i / w_i).  Interpolation runs in double (a hardware
// rasteriser uses fixed-point sub-pixel coordinates; no test depends on which side of an edge a tie falls).
// Triangles that cross the w = 0 plane are not clipped: vref_draw returns 2 for such a view (camera inside / too near).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

extern "C" int vref_vert(const void* ubo0_bytes, uint32_t ubo0_size, const float* in_positions, int n, float* out_gl_position,
                         float* out_frag_position);
extern "C" int vref_frag(const void* ubo1_bytes, uint32_t ubo1_size, const uint8_t* texels, int w, int h, int d,
                         const float* frag_positions, const int32_t* covered, int64_t n, float* out_rgba, int threads);

namespace {
const float kCube[8][3] = {{1, -1, -1}, {1, -1, 1}, {-1, -1, 1}, {-1, -1, -1}, {1, 1, -1}, {1, 1, 1}, {-1, 1, 1}, {-1, 1, -1}};
const uint16_t kIndices[36] = {0, 1, 2, 0, 2, 3, 3, 2, 6, 3, 6, 7, 0, 3, 7, 0, 7, 4, 4, 7, 6, 4, 6, 5, 1, 5, 6, 1, 6, 2, 4, 5, 1, 4, 1, 0};

struct V { double x, y, z, w; double f[3]; };   // framebuffer x, y, depth, clip w, fragPosition

inline double edge(const V& a, const V& b, double px, double py) { return (b.x - a.x) * (py - a.y) - (b.y - a.y) * (px - a.x); }
// top-left rule: an edge function of exactly 0 counts as inside only on a top edge (horizontal, interior below it on a
// y-down screen) or a left edge (interior at larger x).  For the orientation in which interior points have edge() > 0:
// d(edge)/d(px) = -dy, so a left edge has dy < 0; on a horizontal edge edge() = dx * (py - a.y), so a top edge has dx > 0.
inline bool owns_tie(const V& a, const V& b) {
    const double dx = b.x - a.x, dy = b.y - a.y;
    return (dy == 0.0 && dx > 0.0) || dy < 0.0;
}
}  // namespace

// out_rgba[H*W*4]; out_covered[H*W] and out_frag_position[H*W*3] may be null
extern "C" int vref_draw(const void* ubo0_bytes, uint32_t ubo0_size, const void* ubo1_bytes, uint32_t ubo1_size, const uint8_t* texels,
                         int w, int h, int d, int W, int H, float* out_rgba, int32_t* out_covered, float* out_frag_position, int threads) {
    if (!ubo0_bytes || !ubo1_bytes || !texels || W <= 0 || H <= 0 || !out_rgba) return 1;
    float gl[8][4], fp[8][3];
    if (vref_vert(ubo0_bytes, ubo0_size, &kCube[0][0], 8, &gl[0][0], &fp[0][0]) != 0) return 1;
    V v[8];
    for (int i = 0; i < 8; ++i) {
        if (!(gl[i][3] > 0.0f)) return 2;
        const double iw = 1.0 / (double)gl[i][3];
        v[i].x = ((double)gl[i][0] * iw + 1.0) * 0.5 * (double)W;
        v[i].y = ((double)gl[i][1] * iw + 1.0) * 0.5 * (double)H;
        v[i].z = (double)gl[i][2] * iw;
        v[i].w = (double)gl[i][3];
        for (int c = 0; c < 3; ++c) v[i].f[c] = (double)fp[i][c];
    }
    const size_t npx = (size_t)W * H;
    std::vector<int32_t> covered(npx, 0);
    std::vector<float> fpos(npx * 3, 0.0f);
    std::vector<double> depth(npx, 1.0);                                    // cleared to 1
    for (int t = 0; t < 12; ++t) {
        const V &a = v[kIndices[3 * t]], &b = v[kIndices[3 * t + 1]], &c = v[kIndices[3 * t + 2]];
        const double area2 = (a.x * b.y - b.x * a.y) + (b.x * c.y - c.x * b.y) + (c.x * a.y - a.x * c.y);   // sum(x_i y_j - x_j y_i)
        const double aspec = -0.5 * area2;
        if (!(aspec > 0.0)) continue;                                       // back-facing (or degenerate): culled
        // with a > 0 the edge functions below are negative inside; flip the winding so that inside is positive
        const V &p0 = a, &p1 = c, &p2 = b;
        const double tot = edge(p0, p1, p2.x, p2.y);
        if (!(tot > 0.0)) continue;
        const int x0 = (int)std::fmax(0.0, std::floor(std::fmin(std::fmin(a.x, b.x), c.x) - 0.5));
        const int x1 = (int)std::fmin((double)W - 1.0, std::ceil(std::fmax(std::fmax(a.x, b.x), c.x) - 0.5));
        const int y0 = (int)std::fmax(0.0, std::floor(std::fmin(std::fmin(a.y, b.y), c.y) - 0.5));
        const int y1 = (int)std::fmin((double)H - 1.0, std::ceil(std::fmax(std::fmax(a.y, b.y), c.y) - 0.5));
        for (int py = y0; py <= y1; ++py)
            for (int px = x0; px <= x1; ++px) {
                const double cx = (double)px + 0.5, cy = (double)py + 0.5;
                const double e0 = edge(p1, p2, cx, cy), e1 = edge(p2, p0, cx, cy), e2 = edge(p0, p1, cx, cy);   // weights of p0, p1, p2
                const bool in0 = e0 > 0.0 || (e0 == 0.0 && owns_tie(p1, p2));
                const bool in1 = e1 > 0.0 || (e1 == 0.0 && owns_tie(p2, p0));
                const bool in2 = e2 > 0.0 || (e2 == 0.0 && owns_tie(p0, p1));
                if (!(in0 && in1 && in2)) continue;
                const double b0 = e0 / tot, b1 = e1 / tot, b2 = e2 / tot;
                const double z = b0 * p0.z + b1 * p1.z + b2 * p2.z;         // depth is interpolated linearly in screen space
                const size_t i = (size_t)py * W + px;
                if (!(z >= 0.0 && z <= 1.0) || !(z < depth[i])) continue;  // depth clip 0..1, compare LESS
                depth[i] = z;
                const double q0 = b0 / p0.w, q1 = b1 / p1.w, q2 = b2 / p2.w, qs = q0 + q1 + q2;
                for (int k = 0; k < 3; ++k) fpos[3 * i + k] = (float)((q0 * p0.f[k] + q1 * p1.f[k] + q2 * p2.f[k]) / qs);
                covered[i] = 1;
            }
    }
    if (out_covered) std::memcpy(out_covered, covered.data(), npx * sizeof(int32_t));
    if (out_frag_position) std::memcpy(out_frag_position, fpos.data(), npx * 3 * sizeof(float));
    return vref_frag(ubo1_bytes, ubo1_size, texels, w, h, d, fpos.data(), covered.data(), (int64_t)npx, out_rgba, threads);
}
