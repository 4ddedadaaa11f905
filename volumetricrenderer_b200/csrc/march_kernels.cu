// libvolrt march kernels for sm_100a.
//
// One thread per pixel; a warp owns an 8x4-pixel tile (rays of a warp stay within ~2 voxels of
// each other, so their fetches share L1/TEX lines), a CTA of 4 warps owns 16x8 pixels.
// What each piece computes, with the reference line it replaces:
//   setup_ray      shaders/vert.glsl:17-22 + rasteriser coverage and
//                  shaders/frag.glsl:36-54 (camera/fragment to box space, IntersectAABB :18-27,
//                  step count :46, map to [0,1]^3 :49-54)
//   density_*      frag.glsl:66-71 (four fetches, (s1*s2)*(s3+s4)*scale) and the extension modes
//   march_kernel   frag.glsl:57-80 (step loop, Beer-Lambert) + A.6 compositing integrator
//   store_pixel    colour-target conversion (VulkanRenderer.cpp:75-80, VulkanSwapchain.cpp:181-191)
//
// Numerics: the per-ray set-up is written with explicitly rounded intrinsics (__fmul_rn ...)
// in the same operation order as the CPU oracle, because the step count n = int(|Pin-Pout|/step)
// is discontinuous: one ulp in the set-up can add or drop a whole sample.  The LINEAR variants keep
// that discipline through the whole loop and reproduce the oracle's binary32 values bit for bit;
// the TEX variants use the texture unit's fixed-point trilinear weights and FMA contraction.
#include "volrt_internal.h"
#include "noise_device.cuh"
#include "noise_tree_eval.cuh"
#include <math_constants.h>
#include <cuda.h>          // CUtensorMap (type only; the encoder is fetched at run time in volrt_api.cu)
#include <cstring>

namespace volrt {

struct f3 { float x, y, z; };

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float dvd(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ f3 add3(f3 a, f3 b) { return {add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)}; }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return {sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)}; }
__device__ __forceinline__ f3 mul3(f3 a, float s) { return {mul(a.x, s), mul(a.y, s), mul(a.z, s)}; }
__device__ __forceinline__ f3 div3(f3 a, float s) { return {dvd(a.x, s), dvd(a.y, s), dvd(a.z, s)}; }
__device__ __forceinline__ f3 div33(f3 a, f3 b) { return {dvd(a.x, b.x), dvd(a.y, b.y), dvd(a.z, b.z)}; }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }

// column-major mat4 * (x,y,z,w): r_i = ((m0i*x + m1i*y) + m2i*z) + m3i*w, each op rounded
__device__ __forceinline__ void mat_mul4(const float* __restrict__ m, float x, float y, float z, float w, float out[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
        out[i] = add(add(add(mul(m[i], x), mul(m[4 + i], y)), mul(m[8 + i], z)), mul(m[12 + i], w));
}
__device__ __forceinline__ f3 xform_point(const float* __restrict__ m, f3 p) {
    float r[4]; mat_mul4(m, p.x, p.y, p.z, 1.0f, r); return {r[0], r[1], r[2]};
}
__device__ __forceinline__ f3 xform_dir(const float* __restrict__ m, f3 p) {
    float r[4]; mat_mul4(m, p.x, p.y, p.z, 0.0f, r); return {r[0], r[1], r[2]};
}

// frag.glsl:18-27
__device__ __forceinline__ void intersect_aabb(f3 o, f3 d, float& tNear, float& tFar) {
    const f3 lo = {-1.0f, -1.0f, -1.0f}, hi = {1.0f, 1.0f, 1.0f};
    f3 tMin = div33(sub3(lo, o), d), tMax = div33(sub3(hi, o), d);
    f3 t1 = {fminf(tMin.x, tMax.x), fminf(tMin.y, tMax.y), fminf(tMin.z, tMax.z)};
    f3 t2 = {fmaxf(tMin.x, tMax.x), fmaxf(tMin.y, tMax.y), fmaxf(tMin.z, tMax.z)};
    tNear = fmaxf(fmaxf(t1.x, t1.y), t1.z);
    tFar = fminf(fminf(t2.x, t2.y), t2.z);
}

struct Ray {
    f3 pin;        // first sample, texture space [0,1]^3 (exactly on the entry face)
    f3 step;       // stepVec / boxRange
    int n;         // actualSteps
};

// returns false for pixels the rasteriser would not shade (they keep the clear colour)
__device__ __forceinline__ bool setup_ray(const MarchParams& P, const ViewConst* __restrict__ V, int px, int py,
                                          float step_size, Ray& r) {
    // outside the cube's screen rectangle (ViewConst::rect, from the projected corners): nothing to set up
    if (px < V->rect[0] || py < V->rect[1] || px >= V->rect[2] || py >= V->rect[3]) return false;
    // pixel centre -> NDC (viewport (0,0,W,H), TestMain.cpp:198-205)
    float x = sub(mul(add((float)px, 0.5f), dvd(2.0f, (float)P.width)), 1.0f);
    float y = sub(mul(add((float)py, 0.5f), dvd(2.0f, (float)P.height)), 1.0f);
    float pn[4], pf[4];
    mat_mul4(V->inv_pv, x, y, 0.0f, 1.0f, pn);      // near-plane point (depth 0..1 clip space)
    mat_mul4(V->inv_pv, x, y, 1.0f, 1.0f, pf);
    f3 o_w = div3({pn[0], pn[1], pn[2]}, pn[3]);
    f3 f_w = div3({pf[0], pf[1], pf[2]}, pf[3]);
    f3 d_w = sub3(f_w, o_w);
    f3 o_c = xform_point(V->inv_model, o_w);
    f3 d_c = xform_dir(V->inv_model, d_w);
    float t0, tF;
    intersect_aabb(o_c, d_c, t0, tF);
    // front face visible, in front of the near plane and before the far plane
    // (cull back, depth LESS: VulkanPipeline.cpp:107-108,130-132)
    if (!(t0 <= tF && t0 >= 0.0f && t0 <= 1.0f)) return false;
    f3 entry = add3(o_c, mul3(d_c, t0));
    f3 fragPosition = xform_point(V->model, entry);                                   // vert.glsl:20
    f3 cam = {V->cam[0], V->cam[1], V->cam[2]};
    f3 camL = xform_point(V->w2l, cam);                                               // frag.glsl:36
    f3 fragL = xform_point(V->w2l, fragPosition);                                     // :37
    f3 v = sub3(fragL, camL);
    float inv = dvd(1.0f, __fsqrt_rn(dot3(v, v)));
    f3 dir = mul3(v, inv);                                                            // :38
    float tNear, tFar;
    intersect_aabb(camL, dir, tNear, tFar);                                           // :39
    f3 Pin = add3(camL, mul3(dir, tNear));                                            // :43
    f3 Pout = add3(camL, mul3(dir, tFar));                                            // :44
    f3 stepVec = {mul(step_size, dir.x), mul(step_size, dir.y), mul(step_size, dir.z)};  // :45
    f3 dd = sub3(Pin, Pout);
    float q = dvd(__fsqrt_rn(dot3(dd, dd)), step_size);
    int n = (q >= 0.0f) ? __float2int_rz(q) : 0;                                      // :46 int() truncates
    r.n = min(P.max_steps, n);
    r.pin = {mul(add(Pin.x, 1.0f), 0.5f), mul(add(Pin.y, 1.0f), 0.5f), mul(add(Pin.z, 1.0f), 0.5f)};  // :49-53
    r.step = {mul(stepVec.x, 0.5f), mul(stepVec.y, 0.5f), mul(stepVec.z, 0.5f)};      // :54
    return true;
}

// ------------------------------------------------------------------ colour target
__device__ __forceinline__ float srgb_encode(float c) {
    return c <= 0.0031308f ? 12.92f * c : 1.055f * powf(c, 1.0f / 2.4f) - 0.055f;
}
__device__ __forceinline__ void store_pixel(const MarchParams& P, int view, int px, int py, float r, float g, float b, float a) {
    char* row = (char*)P.out + (size_t)view * P.view_stride + (size_t)py * P.row_pitch;
    switch (P.out_format) {
        case VR_OUT_RGBA8_UNORM:
        case VR_OUT_RGBA8_SRGB: {
            r = __saturatef(r); g = __saturatef(g); b = __saturatef(b); a = __saturatef(a);
            if (P.out_format == VR_OUT_RGBA8_SRGB) { r = srgb_encode(r); g = srgb_encode(g); b = srgb_encode(b); }
            uchar4 o;
            o.x = (unsigned char)__float2int_rn(r * 255.0f); o.y = (unsigned char)__float2int_rn(g * 255.0f);
            o.z = (unsigned char)__float2int_rn(b * 255.0f); o.w = (unsigned char)__float2int_rn(a * 255.0f);
            ((uchar4*)row)[px] = o;
            break;
        }
        case VR_OUT_RGBA16F: {
            __half2 lo = __floats2half2_rn(r, g), hi = __floats2half2_rn(b, a);
            uint2 o; o.x = *(unsigned int*)&lo; o.y = *(unsigned int*)&hi;
            ((uint2*)row)[px] = o;
            break;
        }
        default:
            ((float4*)row)[px] = make_float4(r, g, b, a);
    }
}

// ------------------------------------------------------------------ software sampler (LINEAR layout)
// SURVEY.md A.4 / VulkanCore.cpp:676-710: texel centre convention u*N - 0.5, linear filter,
// mirrored repeat.  Same operation order as the oracle => identical binary32 results.
__device__ __forceinline__ int wrap_index(int i, int n, int address_mode) {
    if ((unsigned)i < (unsigned)n) return i;
    if (address_mode == VR_ADDRESS_CLAMP_TO_EDGE) return i < 0 ? 0 : n - 1;
    int m = i % (2 * n);
    if (m < 0) m += 2 * n;
    return m < n ? m : 2 * n - 1 - m;
}

struct Taps {            // the 8 corner offsets and 3 fractions of one trilinear fetch
    int i0, i1, j0, j1, k0, k1;
    float fx, fy, fz;
};
__device__ __forceinline__ Taps make_taps(const MarchParams& P, f3 u) {
    float x = sub(mul(u.x, (float)P.vw), 0.5f), y = sub(mul(u.y, (float)P.vh), 0.5f), z = sub(mul(u.z, (float)P.vd), 0.5f);
    float x0 = floorf(x), y0 = floorf(y), z0 = floorf(z);
    Taps t;
    t.fx = sub(x, x0); t.fy = sub(y, y0); t.fz = sub(z, z0);
    int i = (int)x0, j = (int)y0, k = (int)z0;
    t.i0 = wrap_index(i, P.vw, P.address_mode); t.i1 = wrap_index(i + 1, P.vw, P.address_mode);
    t.j0 = wrap_index(j, P.vh, P.address_mode); t.j1 = wrap_index(j + 1, P.vh, P.address_mode);
    t.k0 = wrap_index(k, P.vd, P.address_mode); t.k1 = wrap_index(k + 1, P.vd, P.address_mode);
    return t;
}
__device__ __forceinline__ float trilerp(const Taps& t, const float c[8]) {
    float wx0 = sub(1.0f, t.fx), wy0 = sub(1.0f, t.fy), wz0 = sub(1.0f, t.fz);
    float a = add(mul(add(mul(c[0], wx0), mul(c[1], t.fx)), wy0), mul(add(mul(c[2], wx0), mul(c[3], t.fx)), t.fy));
    float b = add(mul(add(mul(c[4], wx0), mul(c[5], t.fx)), wy0), mul(add(mul(c[6], wx0), mul(c[7], t.fx)), t.fy));
    return add(mul(a, wz0), mul(b, t.fz));
}
// channel CH of an RGBA8 volume; unorm[] = { i / 255.0f } staged in shared memory
template <int CH>
__device__ __forceinline__ float sample_rgba8_linear(const MarchParams& P, const float* __restrict__ unorm, f3 u) {
    Taps t = make_taps(P, u);
    const unsigned char* __restrict__ base = (const unsigned char*)P.lin + CH;
    size_t r00 = ((size_t)t.k0 * P.vh + t.j0) * P.vw, r10 = ((size_t)t.k0 * P.vh + t.j1) * P.vw;
    size_t r01 = ((size_t)t.k1 * P.vh + t.j0) * P.vw, r11 = ((size_t)t.k1 * P.vh + t.j1) * P.vw;
    float c[8];
    c[0] = unorm[__ldg(base + 4 * (r00 + t.i0))]; c[1] = unorm[__ldg(base + 4 * (r00 + t.i1))];
    c[2] = unorm[__ldg(base + 4 * (r10 + t.i0))]; c[3] = unorm[__ldg(base + 4 * (r10 + t.i1))];
    c[4] = unorm[__ldg(base + 4 * (r01 + t.i0))]; c[5] = unorm[__ldg(base + 4 * (r01 + t.i1))];
    c[6] = unorm[__ldg(base + 4 * (r11 + t.i0))]; c[7] = unorm[__ldg(base + 4 * (r11 + t.i1))];
    return trilerp(t, c);
}
__device__ __forceinline__ float sample_single_linear(const MarchParams& P, const float* __restrict__ unorm, f3 u) {
    Taps t = make_taps(P, u);
    size_t r00 = ((size_t)t.k0 * P.vh + t.j0) * P.vw, r10 = ((size_t)t.k0 * P.vh + t.j1) * P.vw;
    size_t r01 = ((size_t)t.k1 * P.vh + t.j0) * P.vw, r11 = ((size_t)t.k1 * P.vh + t.j1) * P.vw;
    size_t idx[8] = {r00 + t.i0, r00 + t.i1, r10 + t.i0, r10 + t.i1, r01 + t.i0, r01 + t.i1, r11 + t.i0, r11 + t.i1};
    float c[8];
    if (P.vol_format == VR_VOL_R16F) {
        const __half* __restrict__ h = (const __half*)P.lin;
#pragma unroll
        for (int q = 0; q < 8; ++q) c[q] = __half2float(__ldg(h + idx[q]));
    } else {
        const unsigned char* __restrict__ b = (const unsigned char*)P.lin;
#pragma unroll
        for (int q = 0; q < 8; ++q) c[q] = unorm[__ldg(b + idx[q])];
    }
    return trilerp(t, c);
}

// ------------------------------------------------------------------ density functors
struct RayConst {           // per-ray constants of frag.glsl:67-69: cX * scroll_weight, rounded once
    f3 off1, off2, off3;
};

template <int VARIANT>
__device__ __forceinline__ float density_at(const MarchParams& P, const RayConst& rc, const float* __restrict__ unorm,
                                            const GradLut* lut, f3 p) {
    if (VARIANT == MARCH_RGBA_TEX) {
        // frag.glsl:66-71.  coord = Pin*scale_k + cK*weight_k
        const float x1 = p.x * P.coord_scale[0], y1 = p.y * P.coord_scale[0], z1 = p.z * P.coord_scale[0];
        const float x2 = fmaf(p.x, P.coord_scale[1], rc.off1.x), y2 = fmaf(p.y, P.coord_scale[1], rc.off1.y), z2 = fmaf(p.z, P.coord_scale[1], rc.off1.z);
        const float x3 = fmaf(p.x, P.coord_scale[2], rc.off2.x), y3 = fmaf(p.y, P.coord_scale[2], rc.off2.y), z3 = fmaf(p.z, P.coord_scale[2], rc.off2.z);
        const float x4 = fmaf(p.x, P.coord_scale[3], rc.off3.x), y4 = fmaf(p.y, P.coord_scale[3], rc.off3.y), z4 = fmaf(p.z, P.coord_scale[3], rc.off3.z);
        float s1, s2, s3, s4;
        if (P.planar) {
            // each fetch wants one channel: from a one-channel texture it moves a quarter of the bytes through L1/L2.
            // A channel that holds one value everywhere needs no fetch at all (a filter of equal texels returns the
            // texel): the reference's own bake leaves G at 255 throughout (TestMain.cpp:59-60,76 — SURVEY quirk Q1), so a
            // quarter of frag.glsl's fetches read a constant.  The kernel is bound by the fetch rate; the mask is uniform.
            const int cm = P.const_mask;
            s1 = (cm & 1) ? P.const_val[0] : tex3D<float>(P.tex_plane[0], x1, y1, z1);
            s2 = (cm & 2) ? P.const_val[1] : tex3D<float>(P.tex_plane[1], x2, y2, z2);
            s3 = (cm & 4) ? P.const_val[2] : tex3D<float>(P.tex_plane[2], x3, y3, z3);
            s4 = (cm & 8) ? P.const_val[3] : tex3D<float>(P.tex_plane[3], x4, y4, z4);
        } else {
            s1 = tex3D<float4>(P.tex, x1, y1, z1).x;
            s2 = tex3D<float4>(P.tex, x2, y2, z2).y;
            s3 = tex3D<float4>(P.tex, x3, y3, z3).z;
            s4 = tex3D<float4>(P.tex, x4, y4, z4).w;
        }
        return (s1 * s2) * (s3 + s4) * P.sample_scale;
    } else if (VARIANT == MARCH_RGBA_LINEAR) {
        f3 u1 = mul3(p, P.coord_scale[0]);
        f3 u2 = add3(mul3(p, P.coord_scale[1]), rc.off1);
        f3 u3 = add3(mul3(p, P.coord_scale[2]), rc.off2);
        f3 u4 = add3(mul3(p, P.coord_scale[3]), rc.off3);
        float s1 = sample_rgba8_linear<0>(P, unorm, u1);
        float s2 = sample_rgba8_linear<1>(P, unorm, u2);
        float s3 = sample_rgba8_linear<2>(P, unorm, u3);
        float s4 = sample_rgba8_linear<3>(P, unorm, u4);
        return mul(mul(mul(s1, s2), add(s3, s4)), P.sample_scale);
    } else if (VARIANT == MARCH_SINGLE_TEX) {
        return tex3D<float>(P.tex, p.x, p.y, p.z) * P.sample_scale;
    } else if (VARIANT == MARCH_SINGLE_LINEAR) {
        return mul(sample_single_linear(P, unorm, p), P.sample_scale);
    } else if (VARIANT == MARCH_TREE) {
        // the same box-edge fade as the fBm field, around whatever the tree computes; the program is warp-uniform
        const float m = fmaxf(fmaxf(fabsf(p.x - 0.5f), fabsf(p.y - 0.5f)), fabsf(p.z - 0.5f));
        const float t = __saturatef((m - P.fbm.falloff0) * P.fbm.inv_falloff);
        const float edge = 1.0f - t * t * (3.0f - 2.0f * t);
        if (edge <= 0.0f) return 0.0f;
        return fmaxf(0.0f, tree_eval(P.tree_prog, P.tree_n, P.tree_result, lut, P.tree_seed, p.x * P.tree_freq, p.y * P.tree_freq, p.z * P.tree_freq)) * edge;
    } else if (VARIANT == MARCH_FBM_CELLS) {
        return density_fbm_cells(GradLutFixed{}, P.fbm, P.cells, p.x, p.y, p.z);
    } else {
        return density_fbm(GradLutFixed{}, P.fbm, p.x, p.y, p.z);     // the table's place is checked by stage_tables()
    }
}
__host__ __device__ constexpr bool is_fbm(int variant) { return variant == MARCH_FBM || variant == MARCH_FBM_CELLS; }

__device__ __forceinline__ bool brick_occupied(const MarchParams& P, f3 p) {
    const BrickInfo& B = P.bricks;
    int bx = min(max((int)(p.x * B.inv_brick[0]), 0), B.bx - 1);
    int by = min(max((int)(p.y * B.inv_brick[1]), 0), B.by - 1);
    int bz = min(max((int)(p.z * B.inv_brick[2]), 0), B.bz - 1);
    return __ldg(B.occupied + ((size_t)bz * B.by + by) * B.bx + bx) != 0;
}
// 0 = the brick under p is occupied, else its Chebyshev distance (bricks) to the nearest occupied brick
__device__ __forceinline__ int brick_distance(const MarchParams& P, f3 p) {
    const BrickInfo& B = P.bricks;
    int bx = min(max((int)(p.x * B.inv_brick[0]), 0), B.bx - 1);
    int by = min(max((int)(p.y * B.inv_brick[1]), 0), B.by - 1);
    int bz = min(max((int)(p.z * B.inv_brick[2]), 0), B.bz - 1);
    return (int)__ldg(B.dist + ((size_t)bz * B.by + by) * B.bx + bx);
}
// The same lookup without the three float->int conversions (XU pipe: 41 % busy in the C5 march): x*inv + (2^23 - 1/2)
// rounds to an integer-valued float whose bit pattern is 0x4B000000 + round(x*inv - 1/2), i.e. the brick index except that a
// coordinate exactly on a brick face may name the brick below — which shares that face, so its distance bounds the same
// leap (the map is conservative on closed bricks).  Used by the hardware-filtered march only; the strict path and the
// counting build keep the truncation, so their skipped-step counts equal the oracle's.
__device__ __forceinline__ int brick_distance_fast(const MarchParams& P, f3 p) {
    const BrickInfo& B = P.bricks;
    constexpr float kM = 8388607.5f;
    // raw bit patterns 0x4B000000 + brick index, combined over the padded map (BrickInfo::dist_pad) and corrected once; the one
    // clamp of the linear index keeps a NaN position inside the map (it reads some brick's byte instead of faulting)
    const unsigned int rx = __float_as_uint(fmaf(p.x, B.inv_brick[0], kM));
    const unsigned int ry = __float_as_uint(fmaf(p.y, B.inv_brick[1], kM));
    const unsigned int rz = __float_as_uint(fmaf(p.z, B.inv_brick[2], kM));
    const unsigned int idx = (rz * (unsigned)(B.by + 2) + ry) * (unsigned)(B.bx + 2) + rx - B.pad_bias;
    return (int)__ldg(B.dist_pad + min(idx, B.pad_last));
}
// Empty-space leaping.  From a sample in a brick of distance d >= 2, every sample whose coordinates have moved by less
// than (d-1) bricks on every axis still lies in an empty brick, so m = floor(min_c((d-1)*brick_c - eps) / |step_c|)
// further steps can be skipped without a lookup.  eps (1e-4 texture units) absorbs the rounding of the running add.
struct Leap { float k, e; };
__device__ __forceinline__ Leap leap_setup(const MarchParams& P, f3 step) {
    const BrickInfo& B = P.bricks;
    const float a[3] = {fabsf(step.x), fabsf(step.y), fabsf(step.z)};
    Leap L = {1e30f, 0.0f};
#pragma unroll
    for (int c = 0; c < 3; ++c)
        if (a[c] > 1e-20f) {
            const float inv = 1.0f / a[c];
            L.k = fminf(L.k, inv / B.inv_brick[c]);
            L.e = fmaxf(L.e, 1e-4f * inv);
        }
    return L;
}
__device__ __forceinline__ int leap_steps(const Leap& L, int d, int left) {     // left = steps remaining after this one
    if (d < 2) return 0;
    const int m = (int)fmaf((float)(d - 1), L.k, -L.e);
    return max(0, min(m, left));
}
__device__ __forceinline__ float ex2_approx(float x) {       // MUFU.EX2: 2 ulp, no range fix-up (x <= 0 here)
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ bool inside01(f3 p) {
    return p.x >= 0.0f && p.x <= 1.0f && p.y >= 0.0f && p.y <= 1.0f && p.z >= 0.0f && p.z <= 1.0f;
}

// Sum of the densities along the sun light-march from p: samples at p + j*lv, j = 1..light_steps, until
// the march leaves the box (convex, so for good).  The fBm field takes the samples two at a time on the packed FP32
// instructions (noise_device.cuh: density_fbm2, same operations per point, same order of the sum).
#ifndef VOLRT_FBM_PAIRS
#define VOLRT_FBM_PAIRS 1
#endif
template <int VARIANT>
__device__ __forceinline__ float light_sum(const MarchParams& P, const RayConst& rc, const float* __restrict__ unorm,
                                           const GradLut* lut, f3 p, f3 lv, bool skip) {
    float sumL = 0.0f;
    f3 pl = p;
    if constexpr (is_fbm(VARIANT) && VOLRT_FBM_PAIRS) {
        for (int j = 1; j <= P.light_steps; j += 2) {
            const f3 pa = {pl.x + lv.x, pl.y + lv.y, pl.z + lv.z};
            if (!inside01(pa)) break;
            const f3 pb = {pa.x + lv.x, pa.y + lv.y, pa.z + lv.z};
            if (j < P.light_steps && inside01(pb)) {
                f2 r;
                if constexpr (VARIANT == MARCH_FBM_CELLS) r = density_fbm_cells2(GradLutFixed{}, P.fbm, P.cells, make_f2(pa.x, pb.x), make_f2(pa.y, pb.y), make_f2(pa.z, pb.z));
                else r = density_fbm2(GradLutFixed{}, P.fbm, make_f2(pa.x, pb.x), make_f2(pa.y, pb.y), make_f2(pa.z, pb.z));
                sumL += f2_a(r);
                sumL += f2_b(r);
            } else {
                sumL += density_at<VARIANT>(P, rc, unorm, lut, pa);
                break;
            }
            pl = pb;
        }
    } else {
        for (int j = 1; j <= P.light_steps; ++j) {
            pl = {pl.x + lv.x, pl.y + lv.y, pl.z + lv.z};
            if (!inside01(pl)) break;
            if (skip && !brick_occupied(P, pl)) continue;
            sumL += density_at<VARIANT>(P, rc, unorm, lut, pl);
        }
    }
    return sumL;
}

// q = n / d with m = ceil(2^32 / d) from the host (0 stands for d = 1): exact while n * d < 2^32
__device__ __forceinline__ unsigned int fast_div(unsigned int n, unsigned int m) { return m ? __umulhi(n, m) : n; }

// ------------------------------------------------------------------ the march
// one pixel: ray set-up, step loop, resolve, store  (frag.glsl:34-81 + A.6 extensions)
// seg: this thread's depth segment of the ray (0 when the launch has none).  A thick ray of the procedural cloud is a
// dependent chain of 128-256 steps x 7-9 fBm samples on one thread, and a 1/8 row band on one of eight GPUs holds a third of a
// wave of such rays: too few warps to keep an SM's schedulers fed.  Front-to-back compositing is associative —
// (C_a, T_a) then (C_b, T_b) is (C_a + T_a C_b, T_a T_b) — so a ray's step range is cut into P.segments contiguous pieces
// marched independently from T = 1 by different CTAs, and the last piece to finish composes them in order and stores the pixel.
// Unlike the lanes-per-ray split (march_composite_split_kernel: +22 % instructions at four lanes) a piece costs one extra ray
// set-up and nothing per step.  Grid x = tiles x pieces (tile_of_cta).  Not with early termination (a later piece cannot know
// the transmittance in front of it) and not in the counting build.  Measured on B200 (profiles/r02_depth_segments.txt): a
// 1/8 band of C2 0.267 -> 0.230 ms with 8 pieces, of C3 2.00 -> 1.86 ms with 4; whole frames lose ~2 % per doubling, so the
// automatic choice (march_pick_segments) only cuts launches that are less than two waves deep.
template <int VARIANT, bool COMPOSITE, bool STATS>
__device__ __forceinline__ void march_pixel(const MarchParams& P, const float* __restrict__ unorm, const GradLut& lut,
                                            int px, int py, int view, int seg = 0) {
    constexpr bool LINEAR = (VARIANT == MARCH_RGBA_LINEAR || VARIANT == MARCH_SINGLE_LINEAR);
    constexpr bool STRICT = LINEAR;                       // bit-for-bit with the oracle
    constexpr bool CAN_SKIP = (VARIANT == MARCH_SINGLE_TEX || VARIANT == MARCH_SINGLE_LINEAR);
    constexpr bool TEXTURED = (VARIANT == MARCH_RGBA_TEX || VARIANT == MARCH_SINGLE_TEX);
    const ViewConst* __restrict__ V = P.views ? P.views + view : &P.view0;

    // three quarters of a frame's pixels lie outside the cube's screen rectangle: they leave before anything is computed
    if (px < V->rect[0] || py < V->rect[1] || px >= V->rect[2] || py >= V->rect[3]) {
        if (seg == 0) store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f);            // VulkanRenderPass.cpp:17-19
        return;
    }
    const float step_size = mul(dvd(1.0f, (float)P.max_steps), P.step_scale);          // frag.glsl:42
    Ray ray;
    if (!setup_ray(P, V, px, py, step_size, ray)) {
        if (seg == 0) store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f);
        return;
    }
    RayConst rc;
    rc.off1 = {mul(V->scroll_y[0], P.scroll_weight[1]), mul(V->scroll_y[1], P.scroll_weight[1]), mul(V->scroll_y[2], P.scroll_weight[1])};
    rc.off2 = {mul(V->scroll_z[0], P.scroll_weight[2]), mul(V->scroll_z[1], P.scroll_weight[2]), mul(V->scroll_z[2], P.scroll_weight[2])};
    rc.off3 = {mul(V->scroll_w[0], P.scroll_weight[3]), mul(V->scroll_w[1], P.scroll_weight[3]), mul(V->scroll_w[2], P.scroll_weight[3])};

    f3 p = ray.pin;
    int n = ray.n;
    const bool segmented = COMPOSITE && !STATS && P.segments > 1;
    if (segmented) {
        // this piece: steps [n*seg/K, n*(seg+1)/K), first position Pin + i*stepVec (like the lanes-per-ray kernels).  The loops
        // below count from 0 to the piece's length, so the lanes of a warp start together whatever their rays' lengths
        // (n * seg / segments by the tile map's reciprocal of the inner count, which is the piece count here: two 64-bit
        // divisions were ~200 instructions per piece and warp)
        const int seg_begin = (int)fast_div((unsigned)(n * seg), P.map.div_inner);
        n = (int)fast_div((unsigned)(n * (seg + 1)), P.map.div_inner) - seg_begin;
        const float fi = (float)seg_begin;
        p = {fmaf(ray.step.x, fi, p.x), fmaf(ray.step.y, fi, p.y), fmaf(ray.step.z, fi, p.z)};
    }
    const bool skip = CAN_SKIP && P.skip_empty;
    Leap leap = {0.0f, 0.0f};
    if (skip) leap = leap_setup(P, ray.step);
    unsigned long long n_light = 0, n_skipped = 0;
    bool early = false;
    float cr, cg, cb;
    float T_piece = 1.0f;       // transmittance of this thread's piece of the ray (compositing march)

    if (!COMPOSITE) {
        // reference integrator, frag.glsl:55-80
        float acc = 0.0f;
        const bool et = P.early_term_eps > 0.0f;
        const float stop = et ? -logf(P.early_term_eps) : CUDART_INF_F;
        int i = 0;
        if (TEXTURED && !skip && !et) {
            // The texture path is latency-bound at the end of a frame (the long rays outlive the
            // short ones and too few fetches stay in flight to keep the TEX unit busy), so issue the
            // fetches of STEP_BATCH consecutive steps before consuming any of them.  Positions still come
            // from the running add of frag.glsl:74 and the sum keeps its order.
#ifndef VOLRT_STEP_BATCH
#define VOLRT_STEP_BATCH 4
#endif
            constexpr int STEP_BATCH = VOLRT_STEP_BATCH;
            for (; i + STEP_BATCH <= n; i += STEP_BATCH) {
                float d[STEP_BATCH];
#pragma unroll
                for (int k = 0; k < STEP_BATCH; ++k) {
                    d[k] = density_at<VARIANT>(P, rc, unorm, &lut, p);
                    p = add3(p, ray.step);
                }
#pragma unroll
                for (int k = 0; k < STEP_BATCH; ++k) acc += d[k];
            }
        }
        for (; i < n; ++i) {
            if (skip) {
                const int bd = brick_distance(P, p);
                if (bd) {
                    const int m = leap_steps(leap, bd, n - 1 - i);
                    if (STATS) n_skipped += (unsigned long long)(m + 1);
                    if (STRICT) { for (int j = 0; j <= m; ++j) p = add3(p, ray.step); }      // the oracle's running add, bit for bit
                    else { const float fm = (float)(m + 1); p = {fmaf(fm, ray.step.x, p.x), fmaf(fm, ray.step.y, p.y), fmaf(fm, ray.step.z, p.z)}; }
                    i += m;
                    continue;
                }
            }
            float d = density_at<VARIANT>(P, rc, unorm, &lut, p);
            acc = STRICT ? add(acc, d) : acc + d;
            p = add3(p, ray.step);                                                     // running add, :74
            if (et && mul(mul(acc, step_size), P.density) > stop) { early = true; break; }
        }
        acc = mul(acc, step_size);                                                     // :76
        const float e = mul(P.density, fminf(-acc, 0.0f));
        // STRICT: correctly rounded exp (double, rounded once) like the oracle; one per ray
        float c = STRICT ? sub(1.0f, (float)exp((double)e)) : 1.0f - __expf(e);            // :79
        cr = cg = cb = c;
    } else {
        // front-to-back compositing with a sun light-march
        float T = 1.0f;
        cr = cg = cb = 0.0f;
        const f3 lv = {P.sun_dir[0] * P.light_step_len * 0.5f, P.sun_dir[1] * P.light_step_len * 0.5f, P.sun_dir[2] * P.light_step_len * 0.5f};
        // Texture-filtered grids: a step is one dependent chain brick byte -> fetch -> exp, and on a grid far larger than L2
        // (C5: 2 GiB) a warp with ONE fetch in flight leaves the memory system idle (long_scoreboard 7 per issue, DRAM 45 %).
        // So the fetches of the next COMPOSITE_BATCH steps are issued together and consumed in order.  The brick map is
        // consulted once per batch; a batch member inside an empty brick fetches an exact 0 (that is what "empty" means)
        // and contributes nothing, and early termination still stops at the same step — the frame is unchanged.
        // Measured (tools/c5_probe.py, 16 views of C5, ms): batch 1 / 2 / 4 / 8 = 2.008 / 1.860 / 1.956 / 2.561 with
        // skipping + early termination, 4.57 / 4.58 / 5.89 / 6.24 with early termination alone: two in flight is all the
        // memory system rewards — deeper batches fetch past the terminating step and evict each other's lines.
#ifndef VOLRT_FAST_BRICK
#define VOLRT_FAST_BRICK 1
#endif
#ifndef VOLRT_COMPOSITE_BATCH
#define VOLRT_COMPOSITE_BATCH 2
#endif
        constexpr int CB = VOLRT_COMPOSITE_BATCH;
        constexpr bool BATCHED = TEXTURED && !STATS && CB > 1;
        int i0 = 0;
        if constexpr (BATCHED) {
            // T < eps_t ends the ray; without early termination eps_t = 0 and T, never negative, never is
            const float eps_t = P.early_term_eps > 0.0f ? P.early_term_eps : 0.0f;
            // exp(-sigma rho delta) = 2^(rho * kexp); the one-channel grid's sample scale is folded in (x = texel * kd), and
            // since kexp < 0 "rho > 0" reads "x < 0" whatever the sign of the scale
            const float kexp = -P.sigma_t * step_size * 1.4426950408889634f;
            const float kd = VARIANT == MARCH_SINGLE_TEX ? kexp * P.sample_scale : kexp;
            const bool lit = P.light_steps > 0;
            auto fetch = [&](f3 at) -> float {
                if constexpr (VARIANT == MARCH_SINGLE_TEX) return tex3D<float>(P.tex, at.x, at.y, at.z);
                else return density_at<VARIANT>(P, rc, unorm, &lut, at);
            };
            // one sample: T' = T * exp(-sigma rho delta) and its weight is what the ray lost, T - T' — the same numbers as
            // a = 1 - exp(..), w = T a, T *= 1 - a in five instructions instead of eleven (the march over a grid is issue-bound:
            // profiles/r02_final_c5.txt)
            auto composite = [&](float raw, f3 at) {
                const float x = raw * kd;
                if (x < 0.0f) {
                    const float Tn = T * ex2_approx(x);
                    const float w = T - Tn;
                    if (lit) {
                        const float S = expf(-P.sigma_t * light_sum<VARIANT>(P, rc, unorm, &lut, at, lv, skip) * P.light_step_len);
                        cr = fmaf(w, fmaf(P.sun_color[0], S, P.ambient), cr);
                        cg = fmaf(w, fmaf(P.sun_color[1], S, P.ambient), cg);
                        cb = fmaf(w, fmaf(P.sun_color[2], S, P.ambient), cb);
                    } else {
                        cr = fmaf(w, P.unlit[0], cr); cg = fmaf(w, P.unlit[1], cg); cb = fmaf(w, P.unlit[2], cb);
                    }
                    T = Tn;
                }
            };
            // whole batches only: no per-member bounds in the loop; the last n mod CB steps of a ray go through the loop below.
            // The brick byte of the position AFTER a batch is requested together with the batch's fetches, so the chain of an
            // iteration is one memory round trip, not lookup -> fetch (ncu, round 2: 16 % of the stall samples sat on that byte).
            // (A leap loop of its own in front of the fetches — so that lanes with empty bricks ahead catch up in cheap
            // iterations and the warp meets at the fetch batch — was measured: C5 6.02 -> 7.42 ms; the chain lookup -> leap ->
            // lookup of a lane running through empty space is better hidden behind its neighbours' fetches than run on its own.)
            auto lookup = [&](f3 at) -> int { return VOLRT_FAST_BRICK ? brick_distance_fast(P, at) : brick_distance(P, at); };
            int i = 0;
            int bd = (skip && CB <= n) ? lookup(p) : 0;
            while (i + CB <= n) {
                if (bd) {
                    const int m1 = leap_steps(leap, bd, n - 1 - i) + 1;
                    const float fm = (float)m1;
                    p = {fmaf(fm, ray.step.x, p.x), fmaf(fm, ray.step.y, p.y), fmaf(fm, ray.step.z, p.z)};
                    i += m1;
                    bd = lookup(p);
                    continue;
                }
                float d[CB];
                f3 q[CB + 1];
                q[0] = p;
#pragma unroll
                for (int k = 0; k < CB; ++k) { d[k] = fetch(q[k]); q[k + 1] = add3(q[k], ray.step); }
                if (skip) bd = lookup(q[CB]);
                bool stop = false;
#pragma unroll
                for (int k = 0; k < CB; ++k) {
                    if (!stop) {
                        composite(d[k], q[k]);
                        stop = T < eps_t;
                    }
                }
                p = q[CB];
                i += CB;
                if (stop) { early = true; i = n; }
            }
            i0 = i;
        }
        for (int i = i0; i < n; ++i) {
            float rho = 0.0f;
            if (skip) {
                const int bd = brick_distance(P, p);
                if (bd) {
                    const int m = leap_steps(leap, bd, n - 1 - i);
                    if (STATS) n_skipped += (unsigned long long)(m + 1);
                    if (STRICT) { for (int j = 0; j <= m; ++j) p = add3(p, ray.step); }
                    else { const float fm = (float)(m + 1); p = {fmaf(fm, ray.step.x, p.x), fmaf(fm, ray.step.y, p.y), fmaf(fm, ray.step.z, p.z)}; }
                    i += m;
                    continue;           // transmittance unchanged: nothing to test for early termination
                }
            }
            if constexpr (VARIANT == MARCH_TREE && !STATS) {
                rho = 0.0f;                         // evaluated below, in the light-march loop's iteration 0 (one interpreter call site)
            } else {
                rho = density_at<VARIANT>(P, rc, unorm, &lut, p);
            }
            float tree_sumL = 0.0f;
            if constexpr (VARIANT == MARCH_TREE && !STATS) {
                f3 q = p;
                for (int j = 0; j <= P.light_steps; ++j) {
                    const float d = density_at<VARIANT>(P, rc, unorm, &lut, q);
                    if (j == 0) { rho = d; if (!(rho > 0.0f)) break; }
                    else tree_sumL += d;
                    q = {q.x + lv.x, q.y + lv.y, q.z + lv.z};
                    if (!inside01(q)) break;
                }
            }
            if (rho > 0.0f) {                       // rho == 0 contributes exactly nothing
                float a = 1.0f - expf(-P.sigma_t * rho * step_size);
                float S = 1.0f;
                if constexpr (VARIANT == MARCH_TREE && !STATS) {
                    if (P.light_steps > 0) S = expf(-P.sigma_t * tree_sumL * P.light_step_len);
                } else if (!STATS && P.light_steps > 0) {
                    S = expf(-P.sigma_t * light_sum<VARIANT>(P, rc, unorm, &lut, p, lv, skip) * P.light_step_len);
                } else if (P.light_steps > 0) {
                    float sumL = 0.0f;
                    f3 pl = p;
                    for (int j = 1; j <= P.light_steps; ++j) {
                        pl = {pl.x + lv.x, pl.y + lv.y, pl.z + lv.z};
                        if (!inside01(pl)) break;   // the march leaves the box for good (box is convex)
                        if (skip && !brick_occupied(P, pl)) continue;
                        sumL += density_at<VARIANT>(P, rc, unorm, &lut, pl);
                        if (STATS) ++n_light;
                    }
                    S = expf(-P.sigma_t * sumL * P.light_step_len);
                }
                float w = T * a;
                cr = fmaf(w, fmaf(P.sun_color[0], S, P.ambient), cr);
                cg = fmaf(w, fmaf(P.sun_color[1], S, P.ambient), cg);
                cb = fmaf(w, fmaf(P.sun_color[2], S, P.ambient), cb);
                T *= 1.0f - a;
            }
            p = add3(p, ray.step);
            if (P.early_term_eps > 0.0f && T < P.early_term_eps) { early = true; break; }
        }
        T_piece = T;
    }
    if (segmented) {
        // hand the piece over; whoever arrives last composes the ray front to back and stores the pixel
        const size_t band_px = (size_t)(P.row_end - P.row_begin) * P.width;
        const size_t pix = (size_t)(py - P.row_begin) * P.width + px, v = (size_t)(view - P.view_begin);
        __stcg(P.seg_partials + (v * P.segments + seg) * band_px + pix, make_float4(cr, cg, cb, T_piece));
        __threadfence();
        unsigned int* arrived = P.seg_arrived + v * band_px + pix;
        if (atomicAdd(arrived, 1u) != (unsigned)(P.segments - 1)) return;
        __threadfence();
        *arrived = 0u;                                                                 // ready for the next launch
        float Tacc = 1.0f;
        cr = cg = cb = 0.0f;
        for (int k = 0; k < P.segments; ++k) {
            const float4 q = __ldcg(P.seg_partials + (v * P.segments + k) * band_px + pix);
            cr = fmaf(Tacc, q.x, cr); cg = fmaf(Tacc, q.y, cg); cb = fmaf(Tacc, q.z, cb);
            Tacc *= q.w;
        }
    }
    store_pixel(P, view, px, py, cr, cg, cb, 1.0f);                                    // :80
    if (STATS) {
        atomicAdd(&P.counters[0], 1ull);
        atomicAdd(&P.counters[1], (unsigned long long)n);
        if (early) atomicAdd(&P.counters[2], 1ull);
        if (n_light) atomicAdd(&P.counters[3], n_light);
        if (n_skipped) atomicAdd(&P.counters[4], n_skipped);
    }
}

// k-th index of 0..n-1 in centre-out order: m, m+1, m-1, m+2, ... (a bijection)
__device__ __forceinline__ int centre_out(int k, int n) {
    const int m = n >> 1;
    int c = (k & 1) ? m + ((k + 1) >> 1) : m - (k >> 1);
    return c >= n ? 0 : c;
}

// k-th index of 0..n-1 in order of distance from index c: c, c+1, c-1, c+2, ... and, once one side is used up, straight on
// along the other (a bijection; centre_out is the case c = n/2)
__device__ __forceinline__ int outward_from(int k, int n, int c) {
    const int left = c, right = n - 1 - c, m = min(left, right);
    if (k <= 2 * m) return (k & 1) ? c + ((k + 1) >> 1) : c - (k >> 1);
    const int rest = k - 2 * m - 1;
    return right > left ? c + m + 1 + rest : c - m - 1 - rest;
}

// Tile of this CTA.  CTAs are dispatched in linear order (blockIdx.x fastest); the SHORTER grid axis is made the fast one and
// both axes run outwards from the FRAME's centre, so the tiles over the object's middle — the expensive ones — are all
// dispatched first and the cheap silhouette and empty tiles fill the tail.  For a multi-GPU row band (a few tile rows,
// hundreds of columns) that is column by column from the middle, each column starting at the tile row nearest the frame's
// centre row (which for most bands is an edge of the band, not its middle).  With x always fastest and the band's own
// middle first, a third of the SM time of a 1/8 band was idle (profiles/r01_band8_before.txt).
// segs > 1 (depth segments, march_pixel): the grid's x extent is tiles x segs and the pieces of one tile are NEIGHBOURS in dispatch
// order (piece index fastest) — as separate z slices the pieces of a frame would be dispatched one whole frame after the other
// The map's numbers come from the host (set_tile_map): as runtime divisions this prologue was ~175 instructions per warp, 15 % of
// all the instructions of a 64-view batch, most of them in CTAs over empty tiles (ncu source page, round 2).
__device__ __forceinline__ void tile_of_cta(const MarchParams& P, int& tx, int& ty, int* inner = nullptr) {
    const TileMap& M = P.map;
    int kx = (int)blockIdx.x, ky = (int)blockIdx.y;
    if (M.inner > 1) {
        const int q = (int)fast_div((unsigned)kx, M.div_inner);
        if (inner) *inner = kx - q * M.inner;
        kx = q;
    }
    if (M.swap) {
        const int lin = kx + M.gx * ky;
        kx = (int)fast_div((unsigned)lin, M.div_gy);
        ky = lin - kx * M.gy;
    }
    if (M.linear) { tx = kx + M.tx0; ty = ky + M.ty0; return; }
    tx = centre_out(kx, M.gx) + M.tx0;
    ty = outward_from(ky, M.gy, M.cy) + M.ty0;
}
static unsigned int fast_div_magic(unsigned int d) { return d <= 1 ? 0u : (unsigned int)((0x100000000ull + d - 1) / d); }
// grid: tiles along x and y (before the inner index is folded into x); tile_h: rows of a tile
static void set_tile_map(MarchParams& p, dim3 grid, int inner, int tile_h, int tx0 = 0, int ty0 = 0) {
    TileMap& M = p.map;
    M.gx = (int)grid.x; M.gy = (int)grid.y;
    M.tx0 = tx0; M.ty0 = ty0;
    // interleaved view batches (view_group > 1: the volume is larger than L2) walk the tiles in plain row-major order: centre-out
    // alternates between two fronts that move apart, i.e. two working sets in L2, and a batch of a million CTAs has no tail to
    // shorten (C5 4.94 -> 4.80 ms)
    { const char* e = getenv("VOLRT_TILE_LINEAR"); M.linear = (p.view_group > 1 && !(e && atoi(e) == 0)) ? 1 : 0; }
    M.inner = inner > 1 ? inner : 1;
    M.div_inner = fast_div_magic((unsigned)M.inner);
    M.div_gy = fast_div_magic((unsigned)M.gy);
    // the reciprocal division is exact while n * d < 2^32: n < gx * gy, d = gy (frames up to 32768^2 are accepted; the ones
    // too large for that keep x as the fast axis)
    M.swap = M.gy < M.gx && (unsigned long long)M.gx * M.gy * M.gy < 0x100000000ull;
    const int cy = (p.height / 2 - p.row_begin) / tile_h - ty0;
    M.cy = cy < 0 ? 0 : cy > M.gy - 1 ? M.gy - 1 : cy;
}

// static shared memory of the march kernels: the gradient table first, so that it starts the CTA's shared window; the pair
// table of the cell-table march (noise_device.cuh: PairLut) at a fixed offset behind it
template <bool LINEAR, bool PAIR = false>
struct __align__(512) MarchSmem {
    GradLut lut;                       // FBM / TREE variants
    alignas(512) PairLut pair[1];      // MARCH_FBM_CELLS (PAIR; the other variants take the specialisation below)
    float unorm[LINEAR ? 256 : 1];     // LINEAR variants: i / 255
};
template <bool LINEAR>
struct __align__(256) MarchSmem<LINEAR, false> {
    GradLut lut;
    float unorm[LINEAR ? 256 : 1];
};
using MarchSmemPair = MarchSmem<false, true>;
static_assert(offsetof(MarchSmemPair, pair) == VOLRT_PAIR_LUT_OFFSET, "the cell-table march fetches its pair table at an immediate address");

// The fBm kernels fetch gradients at an IMMEDIATE shared address (noise_device.cuh: GradLutFixed): the first static __shared__
// object of a kernel sits at window address VOLRT_LUT_FIXED_ADDR on this toolchain and driver.  That is an observation, not
// a guarantee, so it is checked where a failure is an error code instead of a trap (a trap is sticky and takes the host
// application's CUDA context with it): this probe has the march kernels' shared layout and reports where the table landed.
__global__ void lut_address_probe_kernel(unsigned int* out) {
    __shared__ MarchSmem<false> smem;
    smem.lut.g[threadIdx.x & 15].x = 0.0f;             // keep the object
    if (threadIdx.x == 0) out[0] = (unsigned int)__cvta_generic_to_shared(&smem.lut);
}
cudaError_t march_probe_lut_address(unsigned int* d_out, cudaStream_t s) {
    lut_address_probe_kernel<<<1, 32, 0, s>>>(d_out);
    return cudaGetLastError();
}
unsigned int march_lut_fixed_address() { return (unsigned int)VOLRT_LUT_FIXED_ADDR; }

template <int VARIANT, class SMEM>
__device__ __forceinline__ void stage_tables(SMEM& smem) {
    constexpr bool LINEAR = (VARIANT == MARCH_RGBA_LINEAR || VARIANT == MARCH_SINGLE_LINEAR);
    if (LINEAR) for (int i = threadIdx.x; i < 256; i += blockDim.x) smem.unorm[i] = dvd((float)i, 255.0f);
    if (is_fbm(VARIANT)) {
        // the fBm march fetches gradients at an immediate shared address (noise_device.cuh: GradLutFixed)
        if ((unsigned)__cvta_generic_to_shared(&smem.lut) != (unsigned)VOLRT_LUT_FIXED_ADDR) __trap();
    }
    if (is_fbm(VARIANT) || VARIANT == MARCH_TREE) grad_lut_init(&smem.lut, threadIdx.x);
    if constexpr (VARIANT == MARCH_FBM_CELLS) pair_lut_init(&smem.pair[0], threadIdx.x, blockDim.x);
    if (LINEAR || is_fbm(VARIANT) || VARIANT == MARCH_TREE) __syncthreads();
}

// (a) direct mapping: one CTA of 4 warps per 16x8-pixel tile, a warp per 8x4 sub-tile.  Used where a step
// is cheap (texture fetches), i.e. the whole frame is ~100 us and a work queue would cost more than it saves.
// Tile order: CTAs are dispatched in blockIdx order, so visit tile rows and columns from the screen centre
// outwards — the object sits around the centre (the reference camera looks at the origin,
// TestMain.cpp:225), which makes the last CTAs silhouette-edge and empty tiles (a short tail).
#ifndef VOLRT_CTA_WX
#define VOLRT_CTA_WX 2          // warps per CTA along x (8 pixels each)
#endif
#ifndef VOLRT_CTA_WY
#define VOLRT_CTA_WY 2          // warps per CTA along y (4 pixels each)
#endif
#ifndef VOLRT_DIRECT_MIN_CTAS
#define VOLRT_DIRECT_MIN_CTAS 4     // register cap of the direct fBm kernel: 128.  With the octave loop unrolled (noise_device.cuh) ptxas
#endif                              // would take 223 registers (2 CTAs/SM, C2 1.96 ms); measured 3 / 4 / 5 / 6 CTAs: 1.797 / 1.72 / 1.76 / 1.76 ms

constexpr int kCtaWx = VOLRT_CTA_WX, kCtaWy = VOLRT_CTA_WY, kCtaThreads = 32 * kCtaWx * kCtaWy;
// Register caps (resident CTAs per SM asked of ptxas).  fBm: 128 registers (VOLRT_DIRECT_MIN_CTAS).  Texture marches, measured
// on B200: the compositing march over a grid far larger than L2 (C5) is latency-bound and wants warps — 8 / 10 / 12 CTAs
// (64 / 48 / 40 registers): 7.63 / 6.66 / 6.49 ms per 64 views; the reference integrator (C4) wants registers for its four-step
// fetch batches — 7 / 8 / 10 / 12 CTAs: 0.0925 / 0.0936 / 0.0975 / 0.0991 ms.  Left to itself ptxas took 103 registers for the
// compositing march once the depth-piece hand-over was added (46 before) and C5 doubled to 13.4 ms: hence explicit caps.
// FP32-filtered parity variants: 80 registers; the node-tree interpreter: 168 (its V[]/P[] live in local memory anyway).
#ifndef VOLRT_TEX_COMPOSITE_MIN_CTAS
#define VOLRT_TEX_COMPOSITE_MIN_CTAS 12
#endif
__host__ __device__ constexpr int direct_min_ctas(int variant, bool composite) {
    return is_fbm(variant) ? VOLRT_DIRECT_MIN_CTAS
         : (variant == MARCH_RGBA_TEX || variant == MARCH_SINGLE_TEX) ? (composite ? VOLRT_TEX_COMPOSITE_MIN_CTAS : 7)
         : variant == MARCH_TREE ? 3 : 6;
}
template <int VARIANT, bool COMPOSITE, bool STATS>
__global__ void __launch_bounds__(kCtaThreads, direct_min_ctas(VARIANT, COMPOSITE)) march_kernel(const __grid_constant__ MarchParams P) {
    constexpr bool LINEAR = (VARIANT == MARCH_RGBA_LINEAR || VARIANT == MARCH_SINGLE_LINEAR);
    __shared__ MarchSmem<LINEAR, VARIANT == MARCH_FBM_CELLS> smem;
    float* const unorm = smem.unorm;
    GradLut& lut = smem.lut;
    int tx, ty, seg = 0, inner = 0;
    int view = P.view_begin + blockIdx.z;
    tile_of_cta(P, tx, ty, &inner);
    if (COMPOSITE && !STATS && P.segments > 1) seg = inner;      // grid x = tiles x depth segments
    else if (P.view_group > 1) {                                 // grid x = tiles x views of a group
        view = P.view_begin + (int)blockIdx.z * P.view_group + inner;
        if (view >= P.view_begin + P.n_views) return;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px = tx * (8 * kCtaWx) + (warp % kCtaWx) * 8 + (lane & 7);
    const int py = P.row_begin + ty * (4 * kCtaWy) + (warp / kCtaWx) * 4 + (lane >> 3);
    {
        // a tile that misses the cube's screen rectangle altogether (three quarters of a frame's tiles) clears its pixels and
        // leaves before the tables are staged: CTA-uniform, ahead of the barrier in stage_tables
        const ViewConst* __restrict__ V = P.views ? P.views + view : &P.view0;
        const int x0 = tx * (8 * kCtaWx), y0 = P.row_begin + ty * (4 * kCtaWy);
        if (x0 >= V->rect[2] || x0 + 8 * kCtaWx <= V->rect[0] || y0 >= V->rect[3] || y0 + 4 * kCtaWy <= V->rect[1]) {
            if (seg == 0 && px < P.width && py < P.row_end) store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f);
            return;
        }
    }
    stage_tables<VARIANT>(smem);
    if (px >= P.width || py >= P.row_end) return;
    march_pixel<VARIANT, COMPOSITE, STATS>(P, unorm, lut, px, py, view, seg);
}

// (b) persistent warps: the grid is sized to the machine (SMs x resident CTAs) and every WARP pulls 8x4-pixel
// tiles off a global counter, centre-out.  A CTA of the direct mapping keeps its registers until its slowest
// warp retires — at the silhouette half of a CTA's warps exit at once and their slots idle (measured:
// 14 % of SM cycles idle, profiles/r01_march_fbm_lut.txt) — and a heavy CTA lives for a third of the
// frame, so the dispatch tail is long.  With warp-granular pulls no slot waits for a sibling.
// queue[0] = next tile, queue[1] = warps finished; the last warp out resets both for the next launch.
#ifndef VOLRT_PERSISTENT_MIN_CTAS
#define VOLRT_PERSISTENT_MIN_CTAS 9
#endif
template <int VARIANT, bool COMPOSITE, bool STATS>
__global__ void __launch_bounds__(128, VOLRT_PERSISTENT_MIN_CTAS) march_kernel_persistent(const __grid_constant__ MarchParams P) {
    constexpr bool LINEAR = (VARIANT == MARCH_RGBA_LINEAR || VARIANT == MARCH_SINGLE_LINEAR);
    __shared__ MarchSmem<LINEAR, VARIANT == MARCH_FBM_CELLS> smem;
    float* const unorm = smem.unorm;
    GradLut& lut = smem.lut;
    stage_tables<VARIANT>(smem);
    const int lane = threadIdx.x & 31;
    const int tiles_x = (P.width + 7) >> 3, tiles_y = (P.row_end - P.row_begin + 3) >> 2;
    const int per_view = tiles_x * tiles_y, total = per_view * P.n_views;
    // a pull claims `chunk` consecutive tiles: one same-address atomic costs a few cycles at the L2, and a
    // frame has tens of thousands of (mostly empty) tiles
    const int chunk = P.queue_chunk > 0 ? P.queue_chunk : 1;
    for (;;) {
        int t0 = 0;
        if (lane == 0) t0 = (int)atomicAdd(&P.queue[0], (unsigned)chunk);
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        if (t0 >= total) break;
        const int t1 = min(t0 + chunk, total);
        for (int t = t0; t < t1; ++t) {
            const int v = t / per_view, r = t - v * per_view;
            const int ky = r / tiles_x, kx = r - ky * tiles_x;
            const int px = centre_out(kx, tiles_x) * 8 + (lane & 7);
            const int py = P.row_begin + centre_out(ky, tiles_y) * 4 + (lane >> 3);
            if (px < P.width && py < P.row_end) march_pixel<VARIANT, COMPOSITE, STATS>(P, unorm, lut, px, py, P.view_begin + v);
            __syncwarp();
        }
    }
    if (lane == 0) {
        const unsigned int warps = gridDim.x * (blockDim.x >> 5);
        __threadfence();
        if (atomicAdd(&P.queue[1], 1u) == warps - 1) { P.queue[0] = 0u; P.queue[1] = 0u; __threadfence(); }
    }
}

// (c) ray-split texture march: S lanes share one ray (lane s takes steps s, s+S, s+2S, ...), a warp holds
// 32/S rays as a (32/S/2) x 2 pixel tile, and the partial optical depths meet in a shuffle reduction.
// Why: at 720p the default scene has ~6 100 warps of covered rays for ~5 900 warp slots — a single wave
// whose duration is the longest ray's dependent chain (110 steps x loaded TEX latency), with the
// SMs 29 % idle (profiles/r01_march_tex_c1_baseline.txt).  Splitting rays multiplies the warps and
// divides the chain.  The sum is re-associated and positions are Pin + i*stepVec instead of the running
// add (difference <= 1e-5 texture units, far below the 2^-8 texel resolution of the hardware
// filter), so this mapping serves the hardware-filtered variants only; parity for them is the
// north-star tolerance (2/255, PSNR 50 dB), the strict LINEAR variants keep one thread per ray.
template <int VARIANT, int S>
__global__ void __launch_bounds__(128) march_tex_split_kernel(const __grid_constant__ MarchParams P) {
    constexpr int RAYS = 32 / S;            // rays per warp
    constexpr int RW = RAYS / 2;            // warp tile = RW x 2 pixels
    int tx, ty;
    tile_of_cta(P, tx, ty);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = lane / S, sub = lane % S;
    // CTA tile = (2*RW) x 4 pixels: warps 0,1 side by side on rows 0-1, warps 2,3 on rows 2-3
    const int px = tx * (2 * RW) + (warp & 1) * RW + (r % RW);
    const int py = P.row_begin + ty * 4 + (warp >> 1) * 2 + (r / RW);
    const int view = P.view_begin + blockIdx.z;
    const bool in_frame = px < P.width && py < P.row_end;
    const ViewConst* __restrict__ V = P.views ? P.views + view : &P.view0;
    const float step_size = mul(dvd(1.0f, (float)P.max_steps), P.step_scale);
    Ray ray;
    const bool covered = in_frame && setup_ray(P, V, px, py, step_size, ray);
    float acc = 0.0f;
    if (covered) {
        RayConst rc;
        rc.off1 = {mul(V->scroll_y[0], P.scroll_weight[1]), mul(V->scroll_y[1], P.scroll_weight[1]), mul(V->scroll_y[2], P.scroll_weight[1])};
        rc.off2 = {mul(V->scroll_z[0], P.scroll_weight[2]), mul(V->scroll_z[1], P.scroll_weight[2]), mul(V->scroll_z[2], P.scroll_weight[2])};
        rc.off3 = {mul(V->scroll_w[0], P.scroll_weight[3]), mul(V->scroll_w[1], P.scroll_weight[3]), mul(V->scroll_w[2], P.scroll_weight[3])};
        const GradLut* no_lut = nullptr;
        const int n = ray.n;
        int i = sub;
#ifndef VOLRT_SPLIT_BATCH
#define VOLRT_SPLIT_BATCH 2
#endif
        constexpr int BATCH = VOLRT_SPLIT_BATCH;
        for (; i + (BATCH - 1) * S < n; i += BATCH * S) {
            float d[BATCH];
#pragma unroll
            for (int k = 0; k < BATCH; ++k) {
                const float fi = (float)(i + k * S);
                f3 p = {fmaf(ray.step.x, fi, ray.pin.x), fmaf(ray.step.y, fi, ray.pin.y), fmaf(ray.step.z, fi, ray.pin.z)};
                d[k] = density_at<VARIANT>(P, rc, nullptr, no_lut, p);
            }
#pragma unroll
            for (int k = 0; k < BATCH; ++k) acc += d[k];
        }
        for (; i < n; i += S) {
            const float fi = (float)i;
            f3 p = {fmaf(ray.step.x, fi, ray.pin.x), fmaf(ray.step.y, fi, ray.pin.y), fmaf(ray.step.z, fi, ray.pin.z)};
            acc += density_at<VARIANT>(P, rc, nullptr, no_lut, p);
        }
    }
#pragma unroll
    for (int o = 1; o < S; o <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (sub != 0 || !in_frame) return;
    if (!covered) { store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f); return; }
    acc *= step_size;                                                                  // frag.glsl:76
    const float c = 1.0f - __expf(P.density * fminf(-acc, 0.0f));                      // :79
    store_pixel(P, view, px, py, c, c, c, 1.0f);
}

// (d) ray-split compositing march: S lanes share one ray and take consecutive steps of a round
// (lane s -> step r*S + s), so the expensive part of a step — density plus the sun light-march — runs S
// steps at a time, and the order-dependent part is stitched with a segmented prefix product:
//     T_i = T_round * prod_{lanes before}(1 - a),   C += T_i * a_i * L_i,   T_round *= prod_{all lanes}(1 - a)
// Why: with one thread per ray the heaviest warps of the fBm frame execute ~0.5 M instructions each and share
// a scheduler with 8 other warps, so their dependent chain is as long as the whole frame: half a frame
// (a 2-GPU row band) took 80 % of the full frame's time.  Splitting divides that chain by S.
// Lanes of a ray sit at neighbouring steps, so they mostly agree on "is there density here" and the
// light-march branch stays coherent.  The sum and the products are re-associated and positions are
// Pin + i*stepVec: parity is tolerance-based (as it already is for every compositing variant).
#ifndef VOLRT_SPLIT_MIN_CTAS_OTHER
#define VOLRT_SPLIT_MIN_CTAS_OTHER 7   // every other density source (node trees keep V[]/P[] in local memory anyway; grids need few registers)
#endif
#ifndef VOLRT_SPLIT_MIN_CTAS
#define VOLRT_SPLIT_MIN_CTAS 5      // 96 registers: the two-point fBm light-march spills at 72 (7 CTAs/SM) — measured 4/5/6/7
#endif
template <int VARIANT, int S>
__global__ void __launch_bounds__(128, is_fbm(VARIANT) ? VOLRT_SPLIT_MIN_CTAS : VOLRT_SPLIT_MIN_CTAS_OTHER)
march_composite_split_kernel(const __grid_constant__ MarchParams P) {
    constexpr int RAYS = 32 / S, RW = RAYS / 2;
    constexpr bool CAN_SKIP = (VARIANT == MARCH_SINGLE_TEX || VARIANT == MARCH_SINGLE_LINEAR);
    constexpr bool LINEAR = (VARIANT == MARCH_RGBA_LINEAR || VARIANT == MARCH_SINGLE_LINEAR);
    __shared__ MarchSmem<LINEAR, VARIANT == MARCH_FBM_CELLS> smem;
    float* const unorm = smem.unorm;
    GradLut& lut = smem.lut;
    stage_tables<VARIANT>(smem);
    int tx, ty;
    tile_of_cta(P, tx, ty);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = lane / S, sub = lane % S;
    const int px = tx * (2 * RW) + (warp & 1) * RW + (r % RW);
    const int py = P.row_begin + ty * 4 + (warp >> 1) * 2 + (r / RW);
    const int view = P.view_begin + blockIdx.z;
    const bool in_frame = px < P.width && py < P.row_end;
    const ViewConst* __restrict__ V = P.views ? P.views + view : &P.view0;
    const float step_size = mul(dvd(1.0f, (float)P.max_steps), P.step_scale);
    Ray ray;
    const bool covered = in_frame && setup_ray(P, V, px, py, step_size, ray);
    const int n = covered ? ray.n : 0;
    int rounds = (n + S - 1) / S;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rounds = max(rounds, __shfl_xor_sync(0xffffffffu, rounds, o));
    if (rounds == 0) {
        if (sub == 0 && in_frame) store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f);
        return;
    }
    RayConst rc;
    rc.off1 = {mul(V->scroll_y[0], P.scroll_weight[1]), mul(V->scroll_y[1], P.scroll_weight[1]), mul(V->scroll_y[2], P.scroll_weight[1])};
    rc.off2 = {mul(V->scroll_z[0], P.scroll_weight[2]), mul(V->scroll_z[1], P.scroll_weight[2]), mul(V->scroll_z[2], P.scroll_weight[2])};
    rc.off3 = {mul(V->scroll_w[0], P.scroll_weight[3]), mul(V->scroll_w[1], P.scroll_weight[3]), mul(V->scroll_w[2], P.scroll_weight[3])};
    const bool skip = CAN_SKIP && P.skip_empty;
    const bool et = P.early_term_eps > 0.0f;
    const f3 lv = {P.sun_dir[0] * P.light_step_len * 0.5f, P.sun_dir[1] * P.light_step_len * 0.5f, P.sun_dir[2] * P.light_step_len * 0.5f};
    float T = 1.0f, cr = 0.0f, cg = 0.0f, cb = 0.0f;
    bool done = !covered;
    for (int round = 0; round < rounds; ++round) {
        const int i = round * S + sub;
        float a = 0.0f, Sl = 1.0f;
        if (!done && i < n) {
            const float fi = (float)i;
            const f3 p = {fmaf(ray.step.x, fi, ray.pin.x), fmaf(ray.step.y, fi, ray.pin.y), fmaf(ray.step.z, fi, ray.pin.z)};
            float rho = 0.0f;
            if constexpr (VARIANT == MARCH_TREE) {
                // one call site for the interpreter: the primary sample is iteration 0 of the light-march loop.  With two
                // inlined copies the kernel outgrew the instruction cache (`no_instruction` 1.9 stalls per issue).
                float sumL = 0.0f;
                f3 q = p;
                for (int j = 0; j <= P.light_steps; ++j) {
                    const float d = density_at<VARIANT>(P, rc, unorm, &lut, q);
                    if (j == 0) { rho = d; if (!(rho > 0.0f)) break; }
                    else sumL += d;
                    q = {q.x + lv.x, q.y + lv.y, q.z + lv.z};
                    if (!inside01(q)) break;
                }
                if (rho > 0.0f) {
                    a = 1.0f - expf(-P.sigma_t * rho * step_size);
                    if (P.light_steps > 0) Sl = expf(-P.sigma_t * sumL * P.light_step_len);
                }
            } else {
                if (!(skip && !brick_occupied(P, p))) rho = density_at<VARIANT>(P, rc, unorm, &lut, p);
                if (rho > 0.0f) {
                    a = 1.0f - expf(-P.sigma_t * rho * step_size);
                    if (P.light_steps > 0) Sl = expf(-P.sigma_t * light_sum<VARIANT>(P, rc, unorm, &lut, p, lv, skip) * P.light_step_len);
                }
            }
        }
        // inclusive prefix product of (1 - a) over the S lanes of the ray
        float incl = 1.0f - a;
#pragma unroll
        for (int o = 1; o < S; o <<= 1) {
            const float v = __shfl_up_sync(0xffffffffu, incl, o, S);
            if (sub >= o) incl *= v;
        }
        float excl = __shfl_up_sync(0xffffffffu, incl, 1, S);
        if (sub == 0) excl = 1.0f;
        const float total = __shfl_sync(0xffffffffu, incl, S - 1, S);
        const float Ti = T * excl;                        // transmittance in front of this lane's step
        if (!(et && Ti < P.early_term_eps)) {            // the reference loop would have stopped before this step
            const float w = Ti * a;
            cr = fmaf(w, fmaf(P.sun_color[0], Sl, P.ambient), cr);
            cg = fmaf(w, fmaf(P.sun_color[1], Sl, P.ambient), cg);
            cb = fmaf(w, fmaf(P.sun_color[2], Sl, P.ambient), cb);
        }
        T *= total;
        if (et && T < P.early_term_eps) done = true;     // same value in all lanes of the ray
    }
#pragma unroll
    for (int o = 1; o < S; o <<= 1) {
        cr += __shfl_xor_sync(0xffffffffu, cr, o);
        cg += __shfl_xor_sync(0xffffffffu, cg, o);
        cb += __shfl_xor_sync(0xffffffffu, cb, o);
    }
    if (sub != 0 || !in_frame) return;
    if (!covered) { store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f); return; }
    store_pixel(P, view, px, py, cr, cg, cb, 1.0f);
}

template <int VARIANT>
static cudaError_t launch_composite_split(const MarchParams& p_in, int split, int rows, int n_views, cudaStream_t s) {
    const int rw = 32 / split / 2;
    dim3 grid((p_in.width + 2 * rw - 1) / (2 * rw), (rows + 3) / 4, n_views);
    MarchParams p = p_in;
    set_tile_map(p, grid, 1, 4);
    switch (split) {
        case 2: march_composite_split_kernel<VARIANT, 2><<<grid, 128, 0, s>>>(p); break;
        case 4: march_composite_split_kernel<VARIANT, 4><<<grid, 128, 0, s>>>(p); break;
        case 8: march_composite_split_kernel<VARIANT, 8><<<grid, 128, 0, s>>>(p); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

template <int VARIANT>
static cudaError_t launch_split(const MarchParams& p_in, int split, int rows, int n_views, cudaStream_t s) {
    const int rw = 32 / split / 2;
    dim3 grid((p_in.width + 2 * rw - 1) / (2 * rw), (rows + 3) / 4, n_views);
    MarchParams p = p_in;
    set_tile_map(p, grid, 1, 4);
    switch (split) {
        case 2: march_tex_split_kernel<VARIANT, 2><<<grid, 128, 0, s>>>(p); break;
        case 4: march_tex_split_kernel<VARIANT, 4><<<grid, 128, 0, s>>>(p); break;
        case 8: march_tex_split_kernel<VARIANT, 8><<<grid, 128, 0, s>>>(p); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// (d') compositing march over a texture-filtered one-channel grid with RAY COMPACTION (north_star: "early-ray termination with
// warp-vote (__ballot_sync) compaction").  With early termination and empty-space leaping the rays of a warp end at very
// different times (C5: 25 of 32 lanes busy per issued instruction, profiles/r02_march_c5_viewgroup4.txt), and the kernel is
// issue-bound.  The refill scheme of round 1 (idle lanes take NEW pixels) put rays of unrelated depth into one warp and lost
// 2.5x to scattered fetches.  Here the unit is the CTA's own 16x8-pixel tile: every COMPACT_ROUND loop iterations the CTA
// counts its live rays; when a whole warp's worth of lanes has fallen idle, the live rays are ranked with __ballot_sync,
// written to shared memory and re-dealt to the lowest warps.  The rays of a warp still come from one screen tile and sit at
// neighbouring depths, so their fetches stay coherent; emptied warps only attend the CTA's barriers.  The loop body is the
// batched body of march_pixel (same leaps, same two fetches in flight, same early-termination test): the frame is identical.
// MEASURED, and NOT the default (VOLRT_COMPACT=1 selects it): C5 7.23 ms against 6.02 ms for the plain mapping, and the lanes
// busy per instruction do not move (25.0 of 32 either way, profiles/r02_c5_compaction.txt vs r02_march_c5_viewgroup4.txt).
// The idle lanes of this march are not finished rays: they are lanes that LEAP over empty bricks while their neighbours
// fetch (and lanes whose sample is exactly zero while the neighbours composite) — divergence inside an iteration, which
// re-dealing rays between warps cannot remove.  What the re-dealing costs is real: 2.4 barrier stalls per issue, the four
// warps of a CTA forced into step every 32 iterations, issue-active 73 -> 59 %.
#ifndef VOLRT_COMPACT_DEFAULT
#define VOLRT_COMPACT_DEFAULT 0
#endif
#ifndef VOLRT_COMPACT_ROUND
#define VOLRT_COMPACT_ROUND 32
#endif
struct CompactRay {                 // what a ray carries from lane to lane: 15 words
    f3 p, step;
    float T, cr, cg, cb, leap_k, leap_e;
    int i, n;
    unsigned int pix;               // py << 16 | px
};
template <int VARIANT>
__global__ void __launch_bounds__(128, 10) march_compact_kernel(const __grid_constant__ MarchParams P) {
    __shared__ CompactRay pool[128];
    __shared__ int warp_live[4];
    int tx, ty, vin = 0;
    int view = P.view_begin + blockIdx.z;
    tile_of_cta(P, tx, ty, &vin);
    if (P.view_group > 1) {
        view = P.view_begin + (int)blockIdx.z * P.view_group + vin;
        if (view >= P.view_begin + P.n_views) return;           // CTA-uniform
    }
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int px = tx * 16 + (warp & 1) * 8 + (lane & 7);
    const int py = P.row_begin + ty * 8 + (warp >> 1) * 4 + (lane >> 3);
    const ViewConst* __restrict__ V = P.views ? P.views + view : &P.view0;
    const float step_size = mul(dvd(1.0f, (float)P.max_steps), P.step_scale);
    const bool skip = P.skip_empty != 0;
    const bool et = P.early_term_eps > 0.0f;
    const RayConst rc = {};                                      // one-channel grids take no scroll offsets
    const float kd = -P.sigma_t * step_size * 1.4426950408889634f * P.sample_scale;
    const float eps_t = et ? P.early_term_eps : 0.0f;
    CompactRay r;
    r.T = 1.0f; r.cr = r.cg = r.cb = 0.0f; r.i = 0; r.n = 0; r.leap_k = r.leap_e = 0.0f;
    r.pix = ((unsigned)py << 16) | (unsigned)px;
    bool live = false;
    if (px < P.width && py < P.row_end) {
        Ray ray;
        if (setup_ray(P, V, px, py, step_size, ray)) {
            r.p = ray.pin; r.step = ray.step; r.n = ray.n;
            if (skip) { const Leap L = leap_setup(P, ray.step); r.leap_k = L.k; r.leap_e = L.e; }
            live = ray.n > 0;
            if (!live) store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f);
        } else {
            store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f);                      // VulkanRenderPass.cpp:17-19
        }
    }
    int slots = 128;                                             // lanes of the warps that still hold rays
    for (;;) {
        for (int round = 0; round < VOLRT_COMPACT_ROUND && live; ++round) {
            bool finished = false;
            int bd = 0;
            if (skip) bd = brick_distance_fast(P, r.p);
            if (bd) {
                const Leap L = {r.leap_k, r.leap_e};
                const int m1 = leap_steps(L, bd, r.n - 1 - r.i) + 1;
                const float fm = (float)m1;
                r.p = {fmaf(fm, r.step.x, r.p.x), fmaf(fm, r.step.y, r.p.y), fmaf(fm, r.step.z, r.p.z)};
                r.i += m1;
            } else {
                // the arithmetic of march_pixel's compositing loops, member for member (the frames are compared bit for bit):
                // whole batches on the 2^x form with the sample scale folded in, the last n mod CB steps on the exp form
                constexpr int CB = VOLRT_COMPOSITE_BATCH;
                if (r.i + CB <= r.n) {
                    float d[CB];
                    f3 q[CB + 1];
                    q[0] = r.p;
#pragma unroll
                    for (int k = 0; k < CB; ++k) { d[k] = tex3D<float>(P.tex, q[k].x, q[k].y, q[k].z); q[k + 1] = add3(q[k], r.step); }
#pragma unroll
                    for (int k = 0; k < CB; ++k) {
                        if (!finished) {
                            const float x = d[k] * kd;
                            if (x < 0.0f) {
                                const float Tn = r.T * ex2_approx(x);
                                const float w = r.T - Tn;
                                r.cr = fmaf(w, P.unlit[0], r.cr); r.cg = fmaf(w, P.unlit[1], r.cg); r.cb = fmaf(w, P.unlit[2], r.cb);
                                r.T = Tn;
                            }
                            finished = r.T < eps_t;
                        }
                    }
                    r.p = q[CB];
                    r.i += CB;
                } else {
                    const float rho = density_at<VARIANT>(P, rc, nullptr, nullptr, r.p);
                    if (rho > 0.0f) {
                        const float a = 1.0f - expf(-P.sigma_t * rho * step_size);
                        const float w = r.T * a;
                        r.cr = fmaf(w, fmaf(P.sun_color[0], 1.0f, P.ambient), r.cr);
                        r.cg = fmaf(w, fmaf(P.sun_color[1], 1.0f, P.ambient), r.cg);
                        r.cb = fmaf(w, fmaf(P.sun_color[2], 1.0f, P.ambient), r.cb);
                        r.T *= 1.0f - a;
                    }
                    r.p = add3(r.p, r.step);
                    r.i += 1;
                    if (et && r.T < P.early_term_eps) finished = true;
                }
            }
            if (finished || r.i >= r.n) {
                store_pixel(P, view, (int)(r.pix & 0xffffu), (int)(r.pix >> 16), r.cr, r.cg, r.cb, 1.0f);
                live = false;
            }
        }
        const int total = __syncthreads_count(live);
        if (total == 0) return;
        if (total <= slots - 32) {
            // a warp's worth of lanes is idle: rank the live rays (ballot within the warp, counts across warps) and re-deal them
            const unsigned int ballot = __ballot_sync(0xffffffffu, live);
            if (lane == 0) warp_live[warp] = __popc(ballot);
            __syncthreads();
            int base = 0;
            for (int w = 0; w < warp; ++w) base += warp_live[w];
            if (live) pool[base + __popc(ballot & ((1u << lane) - 1u))] = r;
            __syncthreads();
            live = tid < total;
            if (live) r = pool[tid];
            slots = (total + 31) & ~31;
            __syncthreads();                                     // pool and warp_live are rewritten by the next compaction
        }
    }
}
template <int VARIANT>
static cudaError_t launch_compact(const MarchParams& p_in, int rows, int n_views, cudaStream_t s) {
    MarchParams p = p_in;
    dim3 grid((p.width + 15) / 16, (rows + 7) / 8, n_views);
    set_tile_map(p, grid, p.view_group, 8);
    if (p.view_group > 1) {
        grid.x *= p.view_group;
        grid.z = (grid.z + p.view_group - 1) / p.view_group;
    }
    march_compact_kernel<VARIANT><<<grid, 128, 0, s>>>(p);
    return cudaGetLastError();
}

// (e) TMA-staged brick march (VR_LAYOUT_BRICK, one-channel fp16 grids, reference integrator).
// A CTA owns a 16x8-pixel tile and walks it through the volume in slabs of K steps.  For each slab it
// takes the bounding box (in voxels) of every tap its rays will touch, has ONE thread fetch that box with
// a single 3-D tiled TMA copy (cp.async.bulk.tensor.3d, straight out of the x-fastest volume — the
// tensor map does the gather, no re-bricked copy of the volume is kept) into shared memory, and
// filters from there in FP32 with the oracle's operation order (bit-exact, like the LINEAR layout).
// Slabs shrink (K/2) while the box does not fit; taps that would need mirrored addressing (index -1 or N,
// i.e. the first/last sample of a ray on a volume face) and anything else outside the staged box take the
// global-memory path of the LINEAR variant.
constexpr int BRICK_X = kBrickBox[0], BRICK_Y = kBrickBox[1], BRICK_Z = kBrickBox[2];   // voxels.  TMA wants the box's first byte 16-byte
                                                                   // aligned, so the x origin is rounded down to a multiple of 8 halves: usable x extent >= BRICK_X - 7
constexpr int BRICK_BOX_BYTES = BRICK_X * BRICK_Y * BRICK_Z * 2;
constexpr int BRICK_STAGES = VOLRT_BRICK_STAGES;
constexpr int BRICK_BOX_STRIDE = (BRICK_BOX_BYTES + 127) & ~127;   // a TMA destination is 128-byte aligned
constexpr int BRICK_BYTES = BRICK_BOX_STRIDE * BRICK_STAGES;
constexpr int BRICK_KMAX = 8;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Round 2's second attempt: the box shape is a build constant (volrt_internal.h: kBrickBox — rows padded against bank conflicts:
// -DVOLRT_BRICK_X=40 -DVOLRT_BRICK_Y=21 -DVOLRT_BRICK_Z=21) and so is the number of boxes in flight (VOLRT_BRICK_STAGES=2: the box
// of slab s+1 is requested — its AABB needs only the analytic tap positions — before slab s is filtered out of the other
// buffer, each buffer completing on its own mbarrier).  Both measured, neither faster: see kBrickBox.
struct BrickSlab { int i0, K, lo[3]; bool fits; };
__global__ void __launch_bounds__(128) march_brick_kernel(const __grid_constant__ MarchParams P,
                                                          const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(8) unsigned long long bar[BRICK_STAGES];
    __shared__ int red[8];            // lo xyz, hi xyz, n_max, (unused)
    const int tid = threadIdx.x;
    if (tid == 0) {
#pragma unroll
        for (int b = 0; b < BRICK_STAGES; ++b) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[b])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        red[6] = 0;
    }
    __syncthreads();
    int tx, ty;
    tile_of_cta(P, tx, ty);
    const int warp = tid >> 5, lane = tid & 31;
    const int px = tx * 16 + (warp & 1) * 8 + (lane & 7);
    const int py = P.row_begin + ty * 8 + (warp >> 1) * 4 + (lane >> 3);
    const int view = P.view_begin + blockIdx.z;
    const bool in_frame = px < P.width && py < P.row_end;
    const ViewConst* __restrict__ V = P.views ? P.views + view : &P.view0;
    const float step_size = mul(dvd(1.0f, (float)P.max_steps), P.step_scale);
    Ray ray;
    const bool covered = in_frame && setup_ray(P, V, px, py, step_size, ray);
    const int n = covered ? ray.n : 0;
    {
        const int wmax = __reduce_max_sync(0xffffffffu, n);
        if (lane == 0) atomicMax(&red[6], wmax);
    }
    __syncthreads();
    const int n_max = red[6];
    const float fw = (float)P.vw, fh = (float)P.vh, fd = (float)P.vd;
    f3 p = ray.pin;
    float acc = 0.0f;
    int K = BRICK_KMAX;
    // ---- the slab that starts at step i0: bounding box of its taps (analytic positions; 0.02-voxel slack against the running
    // add), halved until the box fits; CTA-uniform result.  Uses red[0..5] between two barriers of its own.
    auto plan = [&](int i0) -> BrickSlab {
        BrickSlab S;
        S.i0 = i0;
        for (;;) {
            if (tid < 6) red[tid] = tid < 3 ? 0x7fffffff : (int)0x80000000;
            __syncthreads();
            int l0 = 0x7fffffff, l1 = l0, l2 = l0, h0 = (int)0x80000000, h1 = h0, h2 = h0;
            if (i0 < n) {
                const float fa = (float)i0, fb = (float)(min(i0 + K, n) - 1);
                const float xa = fmaf(fmaf(ray.step.x, fa, ray.pin.x), fw, -0.5f), xb = fmaf(fmaf(ray.step.x, fb, ray.pin.x), fw, -0.5f);
                const float ya = fmaf(fmaf(ray.step.y, fa, ray.pin.y), fh, -0.5f), yb = fmaf(fmaf(ray.step.y, fb, ray.pin.y), fh, -0.5f);
                const float za = fmaf(fmaf(ray.step.z, fa, ray.pin.z), fd, -0.5f), zb = fmaf(fmaf(ray.step.z, fb, ray.pin.z), fd, -0.5f);
                l0 = (int)floorf(fminf(xa, xb) - 0.02f); h0 = (int)floorf(fmaxf(xa, xb) + 0.02f) + 1;
                l1 = (int)floorf(fminf(ya, yb) - 0.02f); h1 = (int)floorf(fmaxf(ya, yb) + 0.02f) + 1;
                l2 = (int)floorf(fminf(za, zb) - 0.02f); h2 = (int)floorf(fmaxf(za, zb) + 0.02f) + 1;
            }
            l0 = __reduce_min_sync(0xffffffffu, l0); l1 = __reduce_min_sync(0xffffffffu, l1); l2 = __reduce_min_sync(0xffffffffu, l2);
            h0 = __reduce_max_sync(0xffffffffu, h0); h1 = __reduce_max_sync(0xffffffffu, h1); h2 = __reduce_max_sync(0xffffffffu, h2);
            if (lane == 0 && l0 != 0x7fffffff) {
                atomicMin(&red[0], l0); atomicMin(&red[1], l1); atomicMin(&red[2], l2);
                atomicMax(&red[3], h0); atomicMax(&red[4], h1); atomicMax(&red[5], h2);
            }
            __syncthreads();
            int hi[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) { S.lo[a] = red[a]; hi[a] = red[3 + a]; }
            S.lo[0] &= ~7;                             // floor to a multiple of 8 (two's complement: also for negatives)
            S.fits = hi[0] - S.lo[0] < BRICK_X && hi[1] - S.lo[1] < BRICK_Y && hi[2] - S.lo[2] < BRICK_Z;
            __syncthreads();                       // red[] is rewritten by the next attempt / slab
            if (S.fits || K == 1) break;
            K >>= 1;
        }
        S.K = K;
        return S;
    };
    // ---- one TMA copy of the slab's box (zero-filled outside the volume) into buffer b, completing on bar[b]
    auto request = [&](const BrickSlab& S, int b) {
        if (S.fits && tid == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[b])), "r"(BRICK_BOX_BYTES) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(smem_u32(dyn_smem + b * BRICK_BOX_STRIDE)), "l"(&tmap), "r"(S.lo[0]), "r"(S.lo[1]), "r"(S.lo[2]), "r"(smem_u32(&bar[b])) : "memory");
        }
    };
    unsigned int phase_bits = 0;          // bit b: the parity bar[b] completes next
    int buf = 0;
    BrickSlab cur{};
    if (n_max > 0) { cur = plan(0); request(cur, 0); }
    for (int i0 = 0; i0 < n_max;) {
        // the next slab's box goes out before this one is consumed (its buffer was last read one iteration ago, behind a barrier)
        BrickSlab nxt{};
        const int i_next = i0 + cur.K;
        const bool more = i_next < n_max;
        if (BRICK_STAGES > 1 && more) { nxt = plan(i_next); request(nxt, buf ^ 1); }
        if (cur.fits) {
            unsigned int ok = 0;
            const unsigned int ph = (phase_bits >> buf) & 1u;
            while (!ok)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(ok) : "r"(smem_u32(&bar[buf])), "r"(ph) : "memory");
            phase_bits ^= 1u << buf;
        }
        // ---- K steps out of shared memory
        const __half* box = reinterpret_cast<const __half*>(dyn_smem + buf * BRICK_BOX_STRIDE);
        const int i1 = min(i0 + cur.K, n);
        for (int i = i0; i < i1; ++i) {
            const float x = sub(mul(p.x, fw), 0.5f), y = sub(mul(p.y, fh), 0.5f), z = sub(mul(p.z, fd), 0.5f);
            const float x0 = floorf(x), y0 = floorf(y), z0 = floorf(z);
            const int gi = (int)x0, gj = (int)y0, gk = (int)z0;
            const int li = gi - cur.lo[0], lj = gj - cur.lo[1], lk = gk - cur.lo[2];
            const bool staged = cur.fits && gi >= 0 && gi + 1 < P.vw && gj >= 0 && gj + 1 < P.vh && gk >= 0 && gk + 1 < P.vd &&
                                (unsigned)li < (unsigned)(BRICK_X - 1) && (unsigned)lj < (unsigned)(BRICK_Y - 1) && (unsigned)lk < (unsigned)(BRICK_Z - 1);
            float s;
            if (staged) {
                Taps t;
                t.fx = sub(x, x0); t.fy = sub(y, y0); t.fz = sub(z, z0);
                const __half* b = box + (lk * BRICK_Y + lj) * BRICK_X + li;
                float c[8];
                c[0] = __half2float(b[0]); c[1] = __half2float(b[1]);
                c[2] = __half2float(b[BRICK_X]); c[3] = __half2float(b[BRICK_X + 1]);
                c[4] = __half2float(b[BRICK_X * BRICK_Y]); c[5] = __half2float(b[BRICK_X * BRICK_Y + 1]);
                c[6] = __half2float(b[BRICK_X * BRICK_Y + BRICK_X]); c[7] = __half2float(b[BRICK_X * BRICK_Y + BRICK_X + 1]);
                s = trilerp(t, c);
            } else {
                s = sample_single_linear(P, nullptr, p);          // mirrored taps / outside the box: global memory
            }
            acc = add(acc, mul(s, P.sample_scale));
            p = add3(p, ray.step);                                  // running add, frag.glsl:74
        }
        i0 = i_next;
        __syncthreads();                           // everyone is done with this box before a copy lands in it again
        if (BRICK_STAGES > 1) { cur = nxt; buf ^= 1; }
        else if (more) { cur = plan(i0); request(cur, 0); }
    }
    if (!in_frame) return;
    if (!covered) { store_pixel(P, view, px, py, 0.0f, 0.0f, 0.0f, 1.0f); return; }
    acc = mul(acc, step_size);
    const float e = mul(P.density, fminf(-acc, 0.0f));
    const float c = sub(1.0f, (float)exp((double)e));
    store_pixel(P, view, px, py, c, c, c, 1.0f);
}

cudaError_t launch_march_brick(const MarchParams& p_in, const void* tensor_map, int n_views, cudaStream_t stream) {
    int rows = p_in.row_end - p_in.row_begin;
    if (rows <= 0 || n_views <= 0) return cudaSuccess;
    MarchParams p = p_in;
    p.n_views = n_views;
    // function attributes are per device: a process may hold contexts on several (vr_create_info.device_count)
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(march_brick_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BRICK_BYTES);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    CUtensorMap tm;
    memcpy(&tm, tensor_map, sizeof tm);
    dim3 grid((p.width + 8 * kCtaWx - 1) / (8 * kCtaWx), (rows + 4 * kCtaWy - 1) / (4 * kCtaWy), n_views);
    set_tile_map(p, grid, 1, 4 * kCtaWy);
    march_brick_kernel<<<grid, 128, BRICK_BYTES, stream>>>(p, tm);
    return cudaGetLastError();
}

// per-row sum of step counts (the cost model for row bands)
__global__ void __launch_bounds__(256) row_cost_kernel(const __grid_constant__ MarchParams P, int view, float* __restrict__ row_cost) {
    const int py = blockIdx.x;
    const ViewConst* __restrict__ V = P.views ? P.views + view : &P.view0;
    const float step_size = mul(dvd(1.0f, (float)P.max_steps), P.step_scale);
    float s = 0.0f;
    for (int px = threadIdx.x; px < P.width; px += blockDim.x) {
        Ray r;
        if (setup_ray(P, V, px, py, step_size, r)) s += (float)r.n + 2.0f;   // +2: set-up and store
    }
    __shared__ float red[8];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < 8; ++w) t += red[w];
        row_cost[py] = t + 0.05f * (float)P.width;     // uncovered pixels are cheap, not free
    }
}

// The clear colour (VulkanRenderPass.cpp:17-19) for every pixel of rows [row_begin, row_end) x n_views OUTSIDE the pixel block
// [x0, x1) x [y0, y1) that a restricted march launch covers.  x0 and x1 are multiples of 16, so a quad of four pixels is inside
// or outside as a whole; quads go out as 16-byte stores where the target allows.
__global__ void __launch_bounds__(256) clear_outside_kernel(const __grid_constant__ MarchParams P, int x0, int y0, int x1, int y1) {
    const int px = 4 * (int)(blockIdx.x * blockDim.x + threadIdx.x), py = P.row_begin + (int)blockIdx.y, view = P.view_begin + (int)blockIdx.z;
    if (px >= P.width) return;
    if (py >= y0 && py < y1 && px >= x0 && px < x1) return;
    const int bpp = (P.out_format == VR_OUT_RGBA8_UNORM || P.out_format == VR_OUT_RGBA8_SRGB) ? 4 : P.out_format == VR_OUT_RGBA16F ? 8 : 16;
    char* row = (char*)P.out + (size_t)view * P.view_stride + (size_t)py * P.row_pitch;
    char* at = row + (size_t)px * bpp;
    if (px + 4 <= P.width && ((size_t)at & 15) == 0) {
        uint4* q = (uint4*)at;
        if (bpp == 4) q[0] = make_uint4(0xFF000000u, 0xFF000000u, 0xFF000000u, 0xFF000000u);
        else if (bpp == 8) { q[0] = make_uint4(0u, 0x3C000000u, 0u, 0x3C000000u); q[1] = q[0]; }
        else { const uint4 one = make_uint4(0u, 0u, 0u, 0x3F800000u); q[0] = one; q[1] = one; q[2] = one; q[3] = one; }
        return;
    }
    for (int k = 0; k < 4 && px + k < P.width; ++k) store_pixel(P, view, px + k, py, 0.0f, 0.0f, 0.0f, 1.0f);
}
#ifndef VOLRT_RECT_GRID_MIN_SAVED
#define VOLRT_RECT_GRID_MIN_SAVED 100000       // CTAs a restricted launch must save before the extra clear launch is worth it
#endif

template <int VARIANT, bool COMPOSITE, bool STATS>
static cudaError_t launch_one(const MarchParams& p_in, dim3 grid, bool persistent, cudaStream_t s) {      // grid: tiles x tiles x views
    MarchParams p = p_in;
    if constexpr (VARIANT != MARCH_TREE) {      // a tree sample is hundreds of instructions: no queue variant is built for it
        if (persistent) {
            // resident CTAs per SM x SM count, from the occupancy calculator (registers decide)
            static int per_sm = 0;         // per instantiation; the same on every device of one architecture
            if (per_sm == 0) {
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, march_kernel_persistent<VARIANT, COMPOSITE, STATS>, 128, 0);
                if (per_sm <= 0) per_sm = 1;
            }
            const int resident = (p.sm_count > 0 ? p.sm_count : 148) * per_sm;
            const long long tiles = (long long)((p.width + 7) / 8) * ((p.row_end - p.row_begin + 3) / 4) * p.n_views;
            const long long want = (tiles + 3) / 4;
            march_kernel_persistent<VARIANT, COMPOSITE, STATS><<<(unsigned)(want < resident ? want : resident), 128, 0, s>>>(p);
            return cudaGetLastError();
        }
    }
    const bool segmented = COMPOSITE && !STATS && p.segments > 1;
    if (segmented) p.view_group = 1;
    // A batch of views: tiles over the union of the views' screen rectangles only (the cube fills a quarter of a frame; the CTAs
    // of the other three quarters did nothing but find that out and clear their pixels — 70 % of the 1 M CTAs of a 64-view
    // batch), the rest of the frames cleared by clear_outside_kernel.  One more launch: only where it saves many CTAs.
    int tx0 = 0, ty0 = 0;
    if (!STATS && !segmented && p.batch_rect[2] > p.batch_rect[0] && p.batch_rect[3] > p.batch_rect[1]) {
        constexpr int TW = 8 * kCtaWx, TH = 4 * kCtaWy;
        const int bx0 = p.batch_rect[0] / TW, bx1 = (p.batch_rect[2] + TW - 1) / TW;
        const int ry0 = p.batch_rect[1] - p.row_begin, ry1 = p.batch_rect[3] - p.row_begin;
        const int by0 = (ry0 > 0 ? ry0 : 0) / TH, by1 = ((ry1 < 0 ? 0 : ry1) + TH - 1) / TH;
        const int nx = (bx1 < (int)grid.x ? bx1 : (int)grid.x) - bx0, ny = (by1 < (int)grid.y ? by1 : (int)grid.y) - by0;
        const char* env = getenv("VOLRT_RECT_GRID_MIN_SAVED");         // tests: 0 = always, a huge number = never
        const long long min_saved = env ? atoll(env) : (long long)VOLRT_RECT_GRID_MIN_SAVED;
        if (nx > 0 && ny > 0 && ((long long)grid.x * grid.y - (long long)nx * ny) * grid.z >= min_saved) {
            tx0 = bx0; ty0 = by0;
            const dim3 cgrid((unsigned)((p.width + 1023) / 1024), (unsigned)(p.row_end - p.row_begin), grid.z);
            clear_outside_kernel<<<cgrid, 256, 0, s>>>(p, bx0 * TW, p.row_begin + by0 * TH, (bx0 + nx) * TW, p.row_begin + (by0 + ny) * TH);
            grid.x = (unsigned)nx; grid.y = (unsigned)ny;
            if (p.extra_launches) ++*p.extra_launches;
        }
    }
    set_tile_map(p, grid, segmented ? p.segments : p.view_group, 4 * kCtaWy, tx0, ty0);
    if (segmented) grid.x *= p.segments;             // grid x = tiles x depth segments
    else if (p.view_group > 1) {                     // grid x = tiles x views of a group, grid z = groups (march_kernel)
        grid.x *= p.view_group;
        grid.z = (grid.z + p.view_group - 1) / p.view_group;
    }
    march_kernel<VARIANT, COMPOSITE, STATS><<<grid, kCtaThreads, 0, s>>>(p);
    return cudaGetLastError();
}

template <int VARIANT>
static cudaError_t launch_v(const MarchParams& p, dim3 grid, bool composite, bool stats, bool persistent, cudaStream_t s) {
    if (composite) return stats ? launch_one<VARIANT, true, true>(p, grid, persistent, s) : launch_one<VARIANT, true, false>(p, grid, persistent, s);
    return stats ? launch_one<VARIANT, false, true>(p, grid, persistent, s) : launch_one<VARIANT, false, false>(p, grid, persistent, s);
}

// waves of the machine the covered rays of a launch make at one thread per ray: about a quarter of the pixels are covered
// (the reference's cube fills 21 % of its frame), an SM holds 32 warps of the cheap marches
static double launch_waves(const MarchParams& p, int rows, int n_views) {
    const double warps = (double)p.width * rows * n_views * 0.25 / 32.0;
    return warps / ((double)(p.sm_count > 0 ? p.sm_count : 148) * 32.0);
}

int march_pick_segments(const MarchParams& p, int variant, int n_views, bool stats) {
    const int rows = p.row_end - p.row_begin;
    if (rows <= 0 || n_views <= 0 || stats || !(p.sigma_t > 0.0f) || p.early_term_eps > 0.0f) return 1;
    if (p.schedule > 0 && p.queue) return 1;                     // persistent warps
    if (variant == MARCH_SINGLE_BRICK) return 1;
    int k = p.ray_segments;
    if (k < 0) {
        // auto: only where a step is expensive (procedural density: hundreds to thousands of instructions per step) and the
        // launch is too shallow for whole rays to fill the machine.  Measured (profiles/r02_depth_segments.txt): see march_pixel
        if (!(is_fbm(variant) || variant == MARCH_TREE)) return 1;
        const double waves = launch_waves(p, rows, n_views);
        k = waves >= 2.0 ? 1 : waves >= 0.8 ? 2 : waves >= 0.25 ? 4 : 8;
    }
    if (k > kMaxRaySegments) k = kMaxRaySegments;
    if (k > p.max_steps) k = p.max_steps;
    return k < 1 ? 1 : k;
}

cudaError_t launch_march(const MarchParams& p_in, int variant, int n_views, bool stats, cudaStream_t stream) {
    int rows = p_in.row_end - p_in.row_begin;
    if (rows <= 0 || n_views <= 0) return cudaSuccess;
    MarchParams p = p_in;
    p.n_views = n_views;
    for (int c = 0; c < 3; ++c) p.unlit[c] = p.sun_color[c] + p.ambient;
    dim3 grid((p.width + 8 * kCtaWx - 1) / (8 * kCtaWx), (rows + 4 * kCtaWy - 1) / (4 * kCtaWy), n_views);
    bool composite = p.sigma_t > 0.0f;
    // depth segments (march_pixel): one thread per ray and piece, grid x = tiles x pieces
    if (p.segments > 1) {
        if (!composite || stats || !p.seg_partials || !p.seg_arrived ||
            (size_t)rows * p.width * n_views * p.segments > p.seg_capacity) return cudaErrorInvalidValue;
        switch (variant) {
            case MARCH_RGBA_TEX: return launch_one<MARCH_RGBA_TEX, true, false>(p, grid, false, stream);
            case MARCH_RGBA_LINEAR: return launch_one<MARCH_RGBA_LINEAR, true, false>(p, grid, false, stream);
            case MARCH_SINGLE_TEX: return launch_one<MARCH_SINGLE_TEX, true, false>(p, grid, false, stream);
            case MARCH_SINGLE_LINEAR: return launch_one<MARCH_SINGLE_LINEAR, true, false>(p, grid, false, stream);
            case MARCH_FBM: return launch_one<MARCH_FBM, true, false>(p, grid, false, stream);
            case MARCH_FBM_CELLS: return launch_one<MARCH_FBM_CELLS, true, false>(p, grid, false, stream);
            case MARCH_TREE: return launch_one<MARCH_TREE, true, false>(p, grid, false, stream);
        }
        return cudaErrorInvalidValue;
    }
    p.segments = 1;
    // view batches over a volume that does not fit L2: interleave the views' tiles (MarchParams::view_group)
    int vg = 1;
    if (n_views > 1 && !stats) {
        static const int env_vg = [] { const char* e = getenv("VOLRT_VIEW_GROUP"); return e ? atoi(e) : -1; }();
        // auto: eight views' tiles interleaved when the volume cannot stay in the 126 MB L2 (C5: 64 views over a 2 GiB grid;
        // with centre-out tile order 1 / 2 / 4 / 8 / 16 views per group: 6.49 / 6.2 / 6.01 / 6.3 / 6.4 ms, with the row-major
        // order of set_tile_map 2 / 4 / 8 / 16: 5.06 / 4.80 / 4.72 / 4.96), otherwise whole views one after the other
        const double vol_bytes = (double)p.vw * p.vh * p.vd * (p.vol_format == VR_VOL_RGBA8_UNORM ? 4 : p.vol_format == VR_VOL_R16F ? 2 : 1);
        const bool grid_density = variant == MARCH_RGBA_TEX || variant == MARCH_SINGLE_TEX || variant == MARCH_RGBA_LINEAR || variant == MARCH_SINGLE_LINEAR;
        vg = env_vg >= 1 ? env_vg : (grid_density && vol_bytes > 96e6 && n_views >= 4) ? (n_views >= 8 ? 8 : 4) : 1;
        if (vg > n_views) vg = n_views;
        if (vg > 16) vg = 16;
    }
    p.view_group = vg;
    // schedule: 0 = direct tiles, 1 = persistent warps, -1 = auto
    bool persistent = p.schedule > 0;    // auto = direct: the queue did not pay off on B200 (measurements)
    if (!p.queue) persistent = false;
    // ray compaction (march_compact_kernel): early-terminating compositing march over a one-channel texture grid, no light-march
    if (composite && !stats && !persistent && variant == MARCH_SINGLE_TEX && p.early_term_eps > 0.0f && p.light_steps == 0) {
        const char* env = getenv("VOLRT_COMPACT");               // 1 = use it whatever the size, 0 = never; unset = the default below
        const int env_compact = env ? atoi(env) : -1;
        const bool use = env_compact >= 0 ? env_compact != 0 : VOLRT_COMPACT_DEFAULT != 0;
        if (use && (env_compact > 0 || launch_waves(p, rows, n_views) >= 3.0)) return launch_compact<MARCH_SINGLE_TEX>(p, rows, n_views, stream);
    }
    // ray split of the compositing integrator (any density source; not with device counters)
    if (composite && !stats && !persistent) {
        int split = p.ray_split_composite;
        if (split < 0) {
            // auto: by how many waves of the machine the covered rays (about a quarter of the pixels) make.
            // Measured on B200 with the two-point fBm light-march (ms, 1 / 2 / 4 lanes per ray): 1080p frame (3.4 waves)
            // 1.776 / 1.799 / 1.99; its half (1.7 waves) 0.97 / 0.935 / 1.0; its quarter 0.68 / 0.51 / 0.52; its eighth
            // 0.48 / 0.33 / 0.28; 4K frame 17.0 / 17.4 / -; its eighth (1.7 waves) 2.44 / 2.26 / -.
            const double waves = launch_waves(p, rows, n_views);
            split = waves >= 3.0 ? 1 : waves >= 0.8 ? 2 : 4;
            // a node-tree sample costs 167 registers with one thread per ray (3 CTAs/SM, issue 50 %): the split kernel's
            // register cap serves it better until the frame is many waves deep (C2 frame: 4.8 ms -> see DESIGN 5c)
            if (variant == MARCH_TREE) split = waves >= 10.0 ? 1 : waves >= 2.5 ? 2 : 4;
        }
        if (split == 2 || split == 4 || split == 8) {
            switch (variant) {
                case MARCH_RGBA_TEX: return launch_composite_split<MARCH_RGBA_TEX>(p, split, rows, n_views, stream);
                case MARCH_RGBA_LINEAR: return launch_composite_split<MARCH_RGBA_LINEAR>(p, split, rows, n_views, stream);
                case MARCH_SINGLE_TEX: return launch_composite_split<MARCH_SINGLE_TEX>(p, split, rows, n_views, stream);
                case MARCH_SINGLE_LINEAR: return launch_composite_split<MARCH_SINGLE_LINEAR>(p, split, rows, n_views, stream);
                case MARCH_FBM: return launch_composite_split<MARCH_FBM>(p, split, rows, n_views, stream);
                case MARCH_FBM_CELLS: return launch_composite_split<MARCH_FBM_CELLS>(p, split, rows, n_views, stream);
                case MARCH_TREE: return launch_composite_split<MARCH_TREE>(p, split, rows, n_views, stream);
            }
        }
    }
    // ray split (texture variants of the reference integrator only; no skipping / early-out / counters)
    if ((variant == MARCH_RGBA_TEX || variant == MARCH_SINGLE_TEX) && !composite && !stats && !persistent &&
        !(p.skip_empty && variant == MARCH_SINGLE_TEX) && !(p.early_term_eps > 0.0f)) {
        int split = p.ray_split;
        if (split < 0) {
            // auto: two lanes per ray while the covered rays (about a quarter of the pixels) would not fill two
            // waves of the machine; measured on B200 at 1280x720: 1 -> 0.099 ms, 2 -> 0.095, 4 -> 0.138, 8 -> 0.171
            // (more lanes per ray spread a warp's footprint along the ray and cost L1 data-stage wavefronts)
            split = launch_waves(p, rows, n_views) >= 2.5 ? 1 : 2;     // 2.5 waves of 32 warps = two waves of the 40 the texture march fits
        }
        if (split == 2 || split == 4 || split == 8)
            return variant == MARCH_RGBA_TEX ? launch_split<MARCH_RGBA_TEX>(p, split, rows, n_views, stream)
                                             : launch_split<MARCH_SINGLE_TEX>(p, split, rows, n_views, stream);
    }
    switch (variant) {
        case MARCH_RGBA_TEX: return launch_v<MARCH_RGBA_TEX>(p, grid, composite, stats, persistent, stream);
        case MARCH_RGBA_LINEAR: return launch_v<MARCH_RGBA_LINEAR>(p, grid, composite, stats, persistent, stream);
        case MARCH_SINGLE_TEX: return launch_v<MARCH_SINGLE_TEX>(p, grid, composite, stats, persistent, stream);
        case MARCH_SINGLE_LINEAR: return launch_v<MARCH_SINGLE_LINEAR>(p, grid, composite, stats, persistent, stream);
        case MARCH_FBM: return launch_v<MARCH_FBM>(p, grid, composite, stats, persistent, stream);
        case MARCH_FBM_CELLS: return launch_v<MARCH_FBM_CELLS>(p, grid, composite, stats, persistent, stream);
        case MARCH_TREE: return launch_v<MARCH_TREE>(p, grid, composite, stats, false, stream);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_row_cost(const MarchParams& p, int view, float* row_cost, cudaStream_t stream) {
    row_cost_kernel<<<p.height, 256, 0, stream>>>(p, view, row_cost);
    return cudaGetLastError();
}

}  // namespace volrt
