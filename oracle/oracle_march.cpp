// ORACLE (test infrastructure — see volrt_oracle.h).  The per-pixel ray-march:
// a near-verbatim transliteration of /root/reference/shaders/frag.glsl:18-81 with
// ray generation restating shaders/vert.glsl:17-22 + the fixed-function state
// (SURVEY.md Appendix A.2), the sampler of VulkanCore.cpp:676-710 (A.4), the output
// conversions (A.5) and the build-defined extended integrator (A.6).
// Compiled with -ffp-contract=off: every line below is one rounded binary32 op per
// operator, in the order written — the CUDA kernel's per-ray set-up follows the same order.
#include "volrt_oracle.h"
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>
#include <algorithm>

namespace {

struct vec3 { float x, y, z; };
inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 vmin(vec3 a, vec3 b) { return {std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z)}; }
inline vec3 vmax(vec3 a, vec3 b) { return {std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z)}; }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
inline vec3 normalize(vec3 a) { float inv = 1.0f / std::sqrt(dot(a, a)); return a * inv; }

// column-major mat4 · (v, w): r_i = ((m0i·x + m1i·y) + m2i·z) + m3i·w
inline void mat_mul4(const float* m, float x, float y, float z, float w, float out[4]) {
    for (int i = 0; i < 4; ++i) out[i] = ((m[0 + i] * x + m[4 + i] * y) + m[8 + i] * z) + m[12 + i] * w;
}
inline vec3 xform_point(const float* m, vec3 p) { float r[4]; mat_mul4(m, p.x, p.y, p.z, 1.0f, r); return {r[0], r[1], r[2]}; }
inline vec3 xform_dir(const float* m, vec3 p) { float r[4]; mat_mul4(m, p.x, p.y, p.z, 0.0f, r); return {r[0], r[1], r[2]}; }

// ---------------------------------------------------------------- sampler (A.4)
inline int mirror(int i, int n) {
    int m = i % (2 * n);
    if (m < 0) m += 2 * n;
    return m < n ? m : 2 * n - 1 - m;
}
inline void texel(const vo_volume* v, int i, int j, int k, float out[4]) {
    size_t idx = ((size_t)k * v->h + j) * v->w + i;            // VulkanCore.cpp:625-636 (tight, x fastest)
    switch (v->format) {
        case VO_VOL_RGBA8: {
            const uint8_t* t = (const uint8_t*)v->texels + idx * 4;
            for (int c = 0; c < 4; ++c) out[c] = (float)t[c] / 255.0f;
            break;
        }
        case VO_VOL_R16F: out[0] = vo_half_to_float(((const uint16_t*)v->texels)[idx]); out[1] = out[2] = 0.0f; out[3] = 1.0f; break;
        case VO_VOL_R8: out[0] = (float)((const uint8_t*)v->texels)[idx] / 255.0f; out[1] = out[2] = 0.0f; out[3] = 1.0f; break;
        default: out[0] = ((const float*)v->texels)[idx]; out[1] = out[2] = 0.0f; out[3] = 1.0f; break;
    }
}
inline void sample(const vo_volume* v, vec3 u, float out[4]) {
    float x = u.x * (float)v->w - 0.5f, y = u.y * (float)v->h - 0.5f, z = u.z * (float)v->d - 0.5f;
    float fx0 = std::floor(x), fy0 = std::floor(y), fz0 = std::floor(z);
    float fx = x - fx0, fy = y - fy0, fz = z - fz0;
    int i0 = mirror((int)fx0, v->w), i1 = mirror((int)fx0 + 1, v->w);
    int j0 = mirror((int)fy0, v->h), j1 = mirror((int)fy0 + 1, v->h);
    int k0 = mirror((int)fz0, v->d), k1 = mirror((int)fz0 + 1, v->d);
    float t[8][4];
    texel(v, i0, j0, k0, t[0]); texel(v, i1, j0, k0, t[1]); texel(v, i0, j1, k0, t[2]); texel(v, i1, j1, k0, t[3]);
    texel(v, i0, j0, k1, t[4]); texel(v, i1, j0, k1, t[5]); texel(v, i0, j1, k1, t[6]); texel(v, i1, j1, k1, t[7]);
    float wx0 = 1.0f - fx, wy0 = 1.0f - fy, wz0 = 1.0f - fz;
    for (int c = 0; c < 4; ++c) {
        float a = (t[0][c] * wx0 + t[1][c] * fx) * wy0 + (t[2][c] * wx0 + t[3][c] * fx) * fy;
        float b = (t[4][c] * wx0 + t[5][c] * fx) * wy0 + (t[6][c] * wx0 + t[7][c] * fx) * fy;
        out[c] = a * wz0 + b * fz;
    }
}

// ---------------------------------------------------------------- densities
struct Ctx {
    const vo_volume* vol;
    const vo_params* p;
    vec3 cY, cZ, cW;       // frag.glsl:67-69 — MediaScroll rows .y/.z/.w of columns 0..2
};

inline float smoothstep(float e0, float e1, float x) {
    float t = std::fmin(std::fmax((x - e0) / (e1 - e0), 0.0f), 1.0f);
    return t * t * (3.0f - 2.0f * t);
}

// rho = max(0, .5 fbm + .5 - coverage) * (1 - smoothstep(f0, f1, maxabs(P - 1/2)))
inline float density_fbm(const vo_params* p, vec3 P) {
    float m = std::fmax(std::fmax(std::fabs(P.x - 0.5f), std::fabs(P.y - 0.5f)), std::fabs(P.z - 0.5f));
    float edge = 1.0f - smoothstep(p->fbm_falloff0, p->fbm_falloff1, m);
    if (edge <= 0.0f) return 0.0f;
    float f = vo_fbm_perlin(p->fbm_seed, P.x * p->fbm_frequency, P.y * p->fbm_frequency, P.z * p->fbm_frequency,
                            p->fbm_octaves, p->fbm_gain, p->fbm_lacunarity);
    return std::fmax(0.0f, (0.5f * f + 0.5f) - p->fbm_coverage) * edge;
}

inline float density(const Ctx& c, vec3 Pin) {
    const vo_params* p = c.p;
    float s[4];
    switch (p->density_mode) {
        case VO_DENSITY_RGBA_PRODUCT: {
            // frag.glsl:66-71
            sample(c.vol, Pin * p->coord_scale[0], s);                                             float sample1 = s[0];
            sample(c.vol, Pin * p->coord_scale[1] + c.cY * p->scroll_weight[1], s);                float sample2 = s[1];
            sample(c.vol, Pin * p->coord_scale[2] + c.cZ * p->scroll_weight[2], s);                float sample3 = s[2];
            sample(c.vol, Pin * p->coord_scale[3] + c.cW * p->scroll_weight[3], s);                float sample4 = s[3];
            return (sample1 * sample2) * (sample3 + sample4) * p->sample_scale;
        }
        case VO_DENSITY_SINGLE:
            sample(c.vol, Pin, s);
            return s[0] * p->sample_scale;
        default:
            return density_fbm(p, Pin);
    }
}

inline bool inside01(vec3 P) { return P.x >= 0.0f && P.x <= 1.0f && P.y >= 0.0f && P.y <= 1.0f && P.z >= 0.0f && P.z <= 1.0f; }

// ---------------------------------------------------------------- ray generation (A.2)
// What vert.glsl:17-22 + the rasteriser hand the fragment stage: is the pixel covered by a front face of the cube
// (cull back / depth LESS, VulkanPipeline.cpp:107-108,130-132), and the interpolated fragPosition = Model * (point of
// the face under the pixel centre).  false = the pixel keeps the clear colour.
bool ray_generate(const vo_params* p, const float* invPV, const float* invModel, int W, int H, int px, int py, vec3& fragPosition) {
    const float* Model = p->ubo0;
    float x = ((float)px + 0.5f) * (2.0f / (float)W) - 1.0f;
    float y = ((float)py + 0.5f) * (2.0f / (float)H) - 1.0f;
    float pn[4], pf[4];
    mat_mul4(invPV, x, y, 0.0f, 1.0f, pn);
    mat_mul4(invPV, x, y, 1.0f, 1.0f, pf);
    vec3 o_w = vec3{pn[0], pn[1], pn[2]} / pn[3];
    vec3 f_w = vec3{pf[0], pf[1], pf[2]} / pf[3];
    vec3 d_w = f_w - o_w;
    vec3 o_c = xform_point(invModel, o_w);
    vec3 d_c = xform_dir(invModel, d_w);
    const vec3 boxMin = {-1, -1, -1}, boxMax = {1, 1, 1};      // the cube of TestMain.cpp:94-112
    vec3 tMin = (boxMin - o_c) / d_c, tMax = (boxMax - o_c) / d_c;
    vec3 t1 = vmin(tMin, tMax), t2 = vmax(tMin, tMax);
    float t0 = std::fmax(std::fmax(t1.x, t1.y), t1.z), tF = std::fmin(std::fmin(t2.x, t2.y), t2.z);
    if (!(t0 <= tF && t0 >= 0.0f && t0 <= 1.0f)) return false;
    vec3 entry = o_c + d_c * t0;
    // fragPosition = vec3(Model * vec4(inPosition, 1))          vert.glsl:20
    fragPosition = xform_point(Model, entry);
    return true;
}

// ---------------------------------------------------------------- one pixel
// returns n (>= 0) for a covered pixel, -1 for the clear colour.
int shade_pixel(const Ctx& c, const float* invPV, const float* invModel, int W, int H, int px, int py,
                float out[4], uint64_t* early, uint64_t* light_samples) {
    const vo_params* p = c.p;
    const float* WorldToLocal = p->ubo1;
    const vec3 CameraPosition = {p->ubo1[16], p->ubo1[17], p->ubo1[18]};
    out[0] = out[1] = out[2] = 0.0f; out[3] = 1.0f;            // VulkanRenderPass.cpp:17-19 clear colour

    vec3 fragPosition;
    if (!ray_generate(p, invPV, invModel, W, H, px, py, fragPosition)) return -1;
    const vec3 boxMin = {-1, -1, -1}, boxMax = {1, 1, 1};      // frag.glsl:31-32
    {
        // --- frag.glsl:36-39
        vec3 cameraInBoxLocal = xform_point(WorldToLocal, CameraPosition);
        vec3 fragmentInBoxLocal = xform_point(WorldToLocal, fragPosition);
        vec3 rayDirection = normalize(fragmentInBoxLocal - cameraInBoxLocal);
        // IntersectAABB, frag.glsl:18-27
        vec3 tMin2 = (boxMin - cameraInBoxLocal) / rayDirection, tMax2 = (boxMax - cameraInBoxLocal) / rayDirection;
        vec3 u1 = vmin(tMin2, tMax2), u2 = vmax(tMin2, tMax2);
        float tNear = std::fmax(std::fmax(u1.x, u1.y), u1.z);
        float tFar = std::fmin(std::fmin(u2.x, u2.y), u2.z);

        // --- frag.glsl:42-46
        float stepSize = (1.0f / (float)p->max_steps) * p->step_scale;
        vec3 Pin = cameraInBoxLocal + rayDirection * tNear;
        vec3 Pout = cameraInBoxLocal + rayDirection * tFar;
        vec3 stepVec = stepSize * rayDirection;
        float q = length(Pin - Pout) / stepSize;
        int actualSteps = (q >= 0.0f) ? (q < 2147483520.0f ? (int)q : p->max_steps) : 0;   // int() truncates
        actualSteps = std::min(p->max_steps, actualSteps);

        // --- frag.glsl:49-54: to [0,1]^3
        Pin = Pin - boxMin;
        vec3 boxRange = {std::fabs(boxMax.x - boxMin.x), std::fabs(boxMax.y - boxMin.y), std::fabs(boxMax.z - boxMin.z)};
        Pin = Pin / boxRange;
        stepVec = stepVec / boxRange;

        if (p->sigma_t <= 0.0f) {
            // --- reference integrator, frag.glsl:55-80
            float accumDist = 0.0f;
            const float stop = p->early_term_eps > 0.0f ? -std::log(p->early_term_eps) : INFINITY;
            int i = 0;
            for (; i < actualSteps; ++i) {
                float currentSample = density(c, Pin);
                accumDist += currentSample;
                Pin = Pin + stepVec;
                if (accumDist * stepSize * p->density > stop) { ++i; ++*early; break; }
            }
            accumDist *= stepSize;
            // exp(): GLSL leaves its precision to the implementation; the oracle takes the correctly
            // rounded binary32 value (double exp, rounded once) so that an independent implementation
            // can land on the same bits.
            float color = 1.0f - (float)std::exp((double)(p->density * std::fmin(-accumDist, 0.0f)));
            out[0] = out[1] = out[2] = color; out[3] = 1.0f;
        } else {
            // --- compositing integrator (A.6)
            float T = 1.0f, C[3] = {0, 0, 0};
            vec3 sun = {p->sun_dir[0], p->sun_dir[1], p->sun_dir[2]};
            vec3 lightVec = (sun * p->light_step_len) / boxRange;
            for (int i = 0; i < actualSteps; ++i) {
                float rho = density(c, Pin);
                float a = 1.0f - std::exp(-p->sigma_t * rho * stepSize);
                float S = 1.0f;
                if (p->light_steps > 0) {
                    float sumL = 0.0f;
                    vec3 Pl = Pin;
                    for (int j = 1; j <= p->light_steps; ++j) {
                        Pl = Pl + lightVec;
                        if (inside01(Pl)) sumL += density(c, Pl);
                        ++*light_samples;
                    }
                    S = std::exp(-p->sigma_t * sumL * p->light_step_len);
                }
                float w = T * a;
                for (int k = 0; k < 3; ++k) C[k] += w * (p->sun_color[k] * S + p->ambient);
                T *= 1.0f - a;
                Pin = Pin + stepVec;
                if (p->early_term_eps > 0.0f && T < p->early_term_eps) { ++*early; break; }
            }
            out[0] = C[0]; out[1] = C[1]; out[2] = C[2]; out[3] = 1.0f;
        }
        return actualSteps;
    }
}

template <class F>
void parallel_for(int n, int threads, F f) {
    int nt = threads > 0 ? threads : vo_hw_threads();
    nt = std::max(1, std::min(nt, n));
    if (nt == 1) { for (int i = 0; i < n; ++i) f(i, 0); return; }
    std::atomic<int> next{0};
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t)
        pool.emplace_back([&, t] { for (int i; (i = next.fetch_add(1)) < n;) f(i, t); });
    for (auto& th : pool) th.join();
}

void mat_mul_d(const float* a, const float* b, double* out) {   // out = a·b, column-major
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += (double)a[k * 4 + r] * (double)b[c * 4 + k];
            out[c * 4 + r] = s;
        }
}
void inverse_d(const double* a, double* inv);

}  // namespace

// double-precision inverse of a double matrix, shared by the frame constants below
namespace {
void inverse_d(const double* m, double* out) {
    double a[4][8];
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) { a[r][c] = m[c * 4 + r]; a[r][4 + c] = (r == c); }
    for (int col = 0; col < 4; ++col) {   // Gauss-Jordan with partial pivoting
        int piv = col;
        for (int r = col + 1; r < 4; ++r) if (std::fabs(a[r][col]) > std::fabs(a[piv][col])) piv = r;
        if (piv != col) for (int c = 0; c < 8; ++c) std::swap(a[piv][c], a[col][c]);
        double d = 1.0 / a[col][col];
        for (int c = 0; c < 8; ++c) a[col][c] *= d;
        for (int r = 0; r < 4; ++r) if (r != col) {
            double f = a[r][col];
            if (f != 0.0) for (int c = 0; c < 8; ++c) a[r][c] -= f * a[col][c];
        }
    }
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out[c * 4 + r] = a[r][4 + c];
}
}  // namespace

extern "C" void vo_sample(const vo_volume* v, float u, float vv, float w, float out4[4]) { sample(v, {u, vv, w}, out4); }

// rows[n_rows]: the screen rows to shade, in the order their output is stored (W float4 each)
extern "C" int vo_render_rows(const vo_volume* v, const vo_params* p, int W, int H, const int32_t* rows, int n_rows,
                              float* out_rgba, int32_t* out_steps, uint64_t stats[4], int threads) {
    if (!p || W <= 0 || H <= 0 || n_rows < 0 || (n_rows && !rows) || p->max_steps <= 0) return 1;
    for (int r = 0; r < n_rows; ++r) if (rows[r] < 0 || rows[r] >= H) return 1;
    if (p->density_mode != VO_DENSITY_FBM && (!v || !v->texels)) return 2;
    // frame constants: inverse(Projection·View) and inverse(Model), formed in double and
    // rounded to binary32 once.
    double PV[16], PVi[16], Md[16], Mi[16];
    mat_mul_d(p->ubo0 + 32, p->ubo0 + 16, PV);
    inverse_d(PV, PVi);
    for (int i = 0; i < 16; ++i) Md[i] = p->ubo0[i];
    inverse_d(Md, Mi);
    float invPV[16], invModel[16];
    for (int i = 0; i < 16; ++i) { invPV[i] = (float)PVi[i]; invModel[i] = (float)Mi[i]; }

    Ctx c;
    c.vol = v; c.p = p;
    const float* MS = p->ubo1 + 20;                    // MediaScroll, column-major MS[col*4+row]
    c.cY = {MS[0 * 4 + 1], MS[1 * 4 + 1], MS[2 * 4 + 1]};
    c.cZ = {MS[0 * 4 + 2], MS[1 * 4 + 2], MS[2 * 4 + 2]};
    c.cW = {MS[0 * 4 + 3], MS[1 * 4 + 3], MS[2 * 4 + 3]};

    int nt = threads > 0 ? threads : vo_hw_threads();
    std::vector<uint64_t> acc((size_t)nt * 4, 0);
    parallel_for(n_rows, nt, [&](int r, int t) {
        int py = rows[r];
        uint64_t* a = &acc[(size_t)t * 4];
        for (int px = 0; px < W; ++px) {
            float* o = out_rgba + ((size_t)r * W + px) * 4;
            int n = shade_pixel(c, invPV, invModel, W, H, px, py, o, &a[2], &a[3]);
            if (out_steps) out_steps[(size_t)r * W + px] = n;
            if (n >= 0) { a[0] += 1; a[1] += (uint64_t)n; }
        }
    });
    if (stats) {
        for (int k = 0; k < 4; ++k) { stats[k] = 0; for (int t = 0; t < nt; ++t) stats[k] += acc[(size_t)t * 4 + k]; }
    }
    return 0;
}

// what the fragment stage receives for the pixels of `rows`: covered[r*W + px] (0/1) and fragPosition (world space);
// uncovered entries are zeroed.  Lets a test hand the very same inputs to the reference's frag.glsl compiled as C++
// (oracle/_ref, glsl_shim.h) and project them back through its vert.glsl.
extern "C" int vo_frag_positions(const vo_params* p, int W, int H, const int32_t* rows, int n_rows, float* out_xyz, int32_t* covered) {
    if (!p || W <= 0 || H <= 0 || n_rows < 0 || (n_rows && !rows) || !out_xyz || !covered) return 1;
    for (int r = 0; r < n_rows; ++r) if (rows[r] < 0 || rows[r] >= H) return 1;
    double PV[16], PVi[16], Md[16], Mi[16];
    mat_mul_d(p->ubo0 + 32, p->ubo0 + 16, PV);
    inverse_d(PV, PVi);
    for (int i = 0; i < 16; ++i) Md[i] = p->ubo0[i];
    inverse_d(Md, Mi);
    float invPV[16], invModel[16];
    for (int i = 0; i < 16; ++i) { invPV[i] = (float)PVi[i]; invModel[i] = (float)Mi[i]; }
    for (int r = 0; r < n_rows; ++r)
        for (int px = 0; px < W; ++px) {
            vec3 f = {0, 0, 0};
            const bool c = ray_generate(p, invPV, invModel, W, H, px, rows[r], f);
            const size_t i = (size_t)r * W + px;
            covered[i] = c ? 1 : 0;
            out_xyz[3 * i] = c ? f.x : 0.0f; out_xyz[3 * i + 1] = c ? f.y : 0.0f; out_xyz[3 * i + 2] = c ? f.z : 0.0f;
        }
    return 0;
}

extern "C" int vo_render(const vo_volume* v, const vo_params* p, int W, int H, int row0, int row1,
                         float* out_rgba, int32_t* out_steps, uint64_t stats[4], int threads) {
    if (row0 < 0 || row1 > H || row0 > row1) return 1;
    std::vector<int32_t> rows((size_t)(row1 - row0));
    for (int r = row0; r < row1; ++r) rows[(size_t)(r - row0)] = r;
    return vo_render_rows(v, p, W, H, rows.data(), row1 - row0, out_rgba, out_steps, stats, threads);
}

// ---------------------------------------------------------------- A.5 output conversion
extern "C" uint16_t vo_float_to_half(float f) {
    uint32_t x; std::memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t e = (x >> 23) & 0xffu, m = x & 0x7fffffu;
    if (e == 0xff) return (uint16_t)(sign | 0x7c00u | (m ? 0x200u | (m >> 13) : 0));
    int ex = (int)e - 127 + 15;
    if (ex >= 31) return (uint16_t)(sign | 0x7c00u);
    if (ex <= 0) {
        if (ex < -10) return (uint16_t)sign;
        m |= 0x800000u;
        int shift = 14 - ex;
        uint32_t h = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1))) ++h;
        return (uint16_t)(sign | h);
    }
    uint32_t h = ((uint32_t)ex << 10) | (m >> 13), rem = m & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) ++h;
    return (uint16_t)(sign | h);
}
extern "C" float vo_half_to_float(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1f, m = h & 0x3ffu, x;
    if (e == 0) {
        if (m == 0) x = sign;
        else { int s = 0; while (!(m & 0x400u)) { m <<= 1; ++s; } x = sign | ((uint32_t)(127 - 15 - s + 1) << 23) | ((m & 0x3ffu) << 13); }
    } else if (e == 31) x = sign | 0x7f800000u | (m << 13);
    else x = sign | ((e - 15 + 127) << 23) | (m << 13);
    float f; std::memcpy(&f, &x, 4); return f;
}

extern "C" void vo_encode(const float* rgba, int64_t n_px, int fmt, void* out) {
    for (int64_t i = 0; i < n_px * 4; ++i) {
        float c = rgba[i];
        if (fmt == VO_OUT_RGBA16F) { ((uint16_t*)out)[i] = vo_float_to_half(c); continue; }
        c = std::fmin(std::fmax(c, 0.0f), 1.0f);
        if (!(c == c)) c = 0.0f;
        if (fmt == VO_OUT_RGBA8_SRGB && (i & 3) != 3)
            c = c <= 0.0031308f ? 12.92f * c : 1.055f * std::pow(c, 1.0f / 2.4f) - 0.055f;
        ((uint8_t*)out)[i] = (uint8_t)std::nearbyintf(c * 255.0f);   // round to nearest even
    }
}

// C4/C5 smoke grid: the C2 density field at voxel centres, minus `threshold`, clamped, as fp16.
extern "C" void vo_bake_smoke_r16f(uint16_t* out, int n, const vo_params* p, float threshold, int threads) {
    parallel_for(n, threads, [&](int z, int) {
        for (int y = 0; y < n; ++y)
            for (int x = 0; x < n; ++x) {
                vec3 P = {((float)x + 0.5f) / (float)n, ((float)y + 0.5f) / (float)n, ((float)z + 0.5f) / (float)n};
                float rho = std::fmax(0.0f, density_fbm(p, P) - threshold);
                out[((size_t)z * n + y) * n + x] = vo_float_to_half(rho);
            }
    });
}
