// Device-side coherent noise with FastNoise2 0.10.0 semantics (the library the reference bakes
// its volume with, vendor/noise @ 787b8c8a).  Paths: vendor/noise/include/FastNoise/Generators/.
//   perlin3    Perlin.inl:34-64
//   simplex3   Simplex.inl:56-125
//   cellular3  Cellular.inl:307-366 with the CellularDistance defaults (Cellular.h:16-17,88-90:
//              jitter 1, EuclideanSquared, Index0)
//   hash / gradient / fade / lerp   Utils.inl:229-247,294-340
//   fbm        Fractal.inl:17-39, bounding Fractal.h:27-38
// FMAs are contracted and 1/sqrt is rsqrtf (FastNoise2's own levels use rsqrtps / rsqrt14ps / a
// bit hack and are compiled -ffast-math), so like FastNoise2's ISA levels among themselves the
// results agree to ~1e-5, i.e. <= 1 LSB of the packed texel.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include "volrt_internal.h"

namespace volrt {

constexpr int kPrimeX = 501125321, kPrimeY = 1136930381, kPrimeZ = 1720413743;   // Utils.inl:9-11

__device__ __forceinline__ float grad_dot(int hash, float fx, float fy, float fz) {
    int h13 = hash & 13;
    float u = h13 < 8 ? fx : fy;
    float v = h13 == 12 ? fx : fz;
    v = h13 < 2 ? fy : v;
    return __int_as_float(__float_as_int(u) ^ (hash << 31)) + __int_as_float(__float_as_int(v) ^ ((hash & 2) << 30));
}
__device__ __forceinline__ int hash3(int seed_xz, int xy) {   // (seed ^ x ^ y ^ z) * C; (h >> 15) ^ h
    int h = (seed_xz ^ xy) * 0x27d4eb2d;
    return (h >> 15) ^ h;
}
__device__ __forceinline__ float quintic(float t) { return t * t * t * fmaf(t, fmaf(t, 6.0f, -15.0f), 10.0f); }
__device__ __forceinline__ float lerpf(float a, float b, float t) { return fmaf(t, b - a, a); }

__device__ __forceinline__ float perlin3(int seed, float x, float y, float z) {
    float xs = floorf(x), ys = floorf(y), zs = floorf(z);
    int x0 = __float2int_rn(xs) * kPrimeX, y0 = __float2int_rn(ys) * kPrimeY, z0 = __float2int_rn(zs) * kPrimeZ;
    int x1 = x0 + kPrimeX, y1 = y0 + kPrimeY, z1 = z0 + kPrimeZ;
    float xf0 = x - xs, yf0 = y - ys, zf0 = z - zs;
    float xf1 = xf0 - 1.0f, yf1 = yf0 - 1.0f, zf1 = zf0 - 1.0f;
    xs = quintic(xf0); ys = quintic(yf0); zs = quintic(zf0);
    int sz0 = seed ^ z0, sz1 = seed ^ z1;
    int xy00 = x0 ^ y0, xy10 = x1 ^ y0, xy01 = x0 ^ y1, xy11 = x1 ^ y1;
    float a = lerpf(lerpf(grad_dot(hash3(sz0, xy00), xf0, yf0, zf0), grad_dot(hash3(sz0, xy10), xf1, yf0, zf0), xs),
                    lerpf(grad_dot(hash3(sz0, xy01), xf0, yf1, zf0), grad_dot(hash3(sz0, xy11), xf1, yf1, zf0), xs), ys);
    float b = lerpf(lerpf(grad_dot(hash3(sz1, xy00), xf0, yf0, zf1), grad_dot(hash3(sz1, xy10), xf1, yf0, zf1), xs),
                    lerpf(grad_dot(hash3(sz1, xy01), xf0, yf1, zf1), grad_dot(hash3(sz1, xy11), xf1, yf1, zf1), xs), ys);
    return 0.964921414852142333984375f * lerpf(a, b, zs);
}

// ---- Perlin with the gradient taken from a 16-entry shared-memory table --------------------------
// The select/sign form above costs ~11 ALU-pipe instructions per lattice corner (LOP3 / ISETP / FSEL /
// SHF) and the ALU pipe issues one warp instruction every other cycle, which made the fBm march
// ALU-bound at 85 % (profiles/r01_march_fbm_baseline.txt).  GetGradientDot only looks at the low
// nibble of the hash, so the 16 possible gradients (cx,cy,cz in {-1,0,+1}) are tabulated once per
// CTA: dot = fx*cx + fy*cy + fz*cz is bit-identical to (u ^ h1) + (v ^ h2) (products by +-1 and
// the added zero are exact, so the one rounding is that of the same two-term sum).
// Layout: GradLut::g[n] = (cx, cy, cz, 0): one LDS.128 per lattice corner.  An earlier layout (LDS.64 for (cx,cy) +
// LDS.32 for cz at an 8-byte stride) was bank-conflict free but cost one more issue slot per corner; the kernel is
// issue-bound, not LSU-bound, so the single wider load wins although entries n and n+8 share banks
// (C2 2.04 -> 1.95 ms, C3 19.4 -> 18.7 ms).
// The table offset n*16 comes straight from the integer pipe of the FMA unit:
//   h = t*C;  h16 = t*(C<<4) = h<<4;  off = (h16 ^ (h16 >> 15)) & 0xF0  ==  (((h >> 15) ^ h) & 15) * 16
// (bits 4..7 of h16>>15 are bits 15..18 of h: the four bits the left shift drops are never looked at)
// the components (cx, cy, cz) of gradient n of the 16 (Utils.inl:229-247: what GetGradientDot multiplies by)
__host__ __device__ inline void grad_components(int n, float& cx, float& cy, float& cz) {
    const int h13 = n & 13;
    const int iu = h13 < 8 ? 0 : 1;
    int iv = h13 == 12 ? 0 : 2;
    iv = h13 < 2 ? 1 : iv;
    const float su = (n & 1) ? -1.0f : 1.0f, sv = (n & 2) ? -1.0f : 1.0f;
    cx = (iu == 0 ? su : 0.0f) + (iv == 0 ? sv : 0.0f);
    cy = (iu == 1 ? su : 0.0f) + (iv == 1 ? sv : 0.0f);
    cz = iv == 2 ? sv : 0.0f;
}
struct __align__(16) GradLut { float4 g[16]; };

__device__ __forceinline__ void grad_lut_init(GradLut* lut, int tid) {
    if (tid < 16) {
        const int n = tid;
        float cx, cy, cz;
        grad_components(n, cx, cy, cz);
        lut->g[n] = make_float4(cx, cy, cz, 0.0f);
    }
}

// A table the kernel guarantees to sit at shared-window address kLutFixedAddr (the first static __shared__ object of a
// non-cluster launch on sm_100a: the window starts with 1 KiB the system keeps).  The table base then is an immediate
// of the LDS and the per-corner address add disappears.  ptxas used to find that form itself (LDS [R + UR]), but in the
// two-point march kernel it demotes the base to a vector register and spends one IADD3 per fetch (16 of 214 per octave
// pair).  stage_tables() traps if the table is anywhere else, so a toolchain that lays shared memory out differently
// fails loudly instead of reading garbage.
#ifndef VOLRT_LUT_FIXED_ADDR
#define VOLRT_LUT_FIXED_ADDR 1024
#endif
struct GradLutFixed {};     // tag type: "the table is at VOLRT_LUT_FIXED_ADDR"

__device__ __forceinline__ unsigned int grad_lut_offset(int t) {
    const unsigned int h16 = (unsigned int)t * (0x27d4eb2du << 4);     // = h << 4 (mod 2^32)
    // (h16 >> 15 as the high word of h16 * 2^17 — an IMAD.HI on the FMA pipe instead of a shift on the half-rate ALU pipe — was
    // measured in round 1: C2 1.769 -> 1.936 ms; not kept)
    return (h16 ^ (h16 >> 15)) & 0xF0u;
}
__device__ __forceinline__ float4 grad_lut_fetch(const GradLut* lut, int t) {
    return *reinterpret_cast<const float4*>(reinterpret_cast<const char*>(lut->g) + grad_lut_offset(t));
}
__device__ __forceinline__ float4 grad_lut_fetch(GradLutFixed, int t) {
    float4 g;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4 + %5];" : "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w) : "r"(grad_lut_offset(t)), "n"(VOLRT_LUT_FIXED_ADDR));
    return g;
}
__device__ __forceinline__ float grad_dot_lut(const GradLut* lut, int t, float fx, float fy, float fz) {
    const float4 g = grad_lut_fetch(lut, t);
    return fmaf(fz, g.z, fmaf(fy, g.y, fx * g.x));
}
// lerp(d0, d1, s) over the two corners of an x edge, with d1 - d0 accumulated directly into the second dot product:
// 7 instructions instead of 8 (the difference is re-associated by a few ulp; FastNoise2's own SIMD levels contract
// and re-associate the same expression differently among themselves)
template <class LUT>
__device__ __forceinline__ float grad_edge_lut(LUT lut, int t0, int t1, float fx0, float fx1, float fy, float fz, float s) {
    const float4 g0 = grad_lut_fetch(lut, t0), g1 = grad_lut_fetch(lut, t1);
    const float d0 = fmaf(fz, g0.z, fmaf(fy, g0.y, fx0 * g0.x));
    const float e = fmaf(fz, g1.z, fmaf(fy, g1.y, fmaf(fx1, g1.x, -d0)));
    return fmaf(s, e, d0);
}

template <class LUT>
__device__ __forceinline__ float perlin3_lut(LUT lut, int seed, float x, float y, float z) {
    float xs = floorf(x), ys = floorf(y), zs = floorf(z);
    int x0 = __float2int_rn(xs) * kPrimeX, y0 = __float2int_rn(ys) * kPrimeY, z0 = __float2int_rn(zs) * kPrimeZ;
    int x1 = x0 + kPrimeX, y1 = y0 + kPrimeY, z1 = z0 + kPrimeZ;
    float xf0 = x - xs, yf0 = y - ys, zf0 = z - zs;
    float xf1 = xf0 - 1.0f, yf1 = yf0 - 1.0f, zf1 = zf0 - 1.0f;
    xs = quintic(xf0); ys = quintic(yf0); zs = quintic(zf0);
    int sz0 = seed ^ z0, sz1 = seed ^ z1;
    float a = lerpf(grad_edge_lut(lut, sz0 ^ x0 ^ y0, sz0 ^ x1 ^ y0, xf0, xf1, yf0, zf0, xs),
                    grad_edge_lut(lut, sz0 ^ x0 ^ y1, sz0 ^ x1 ^ y1, xf0, xf1, yf1, zf0, xs), ys);
    float b = lerpf(grad_edge_lut(lut, sz1 ^ x0 ^ y0, sz1 ^ x1 ^ y0, xf0, xf1, yf0, zf1, xs),
                    grad_edge_lut(lut, sz1 ^ x0 ^ y1, sz1 ^ x1 ^ y1, xf0, xf1, yf1, zf1, xs), ys);
    return lerpf(a, b, zs);           // WITHOUT Perlin.inl:63's 0.9649... output scale: the caller folds it into the octave amplitude
}
constexpr float kPerlinScale = 0.964921414852142333984375f;

__device__ __forceinline__ float simplex3(int seed, float x, float y, float z) {
    const float F3 = 1.0f / 3.0f, G3 = 1.0f / 2.0f;
    float s = F3 * (x + y + z);
    x += s; y += s; z += s;
    float x0 = floorf(x), y0 = floorf(y), z0 = floorf(z);
    float xi = x - x0, yi = y - y0, zi = z - z0;
    int i = __float2int_rn(x0) * kPrimeX, j = __float2int_rn(y0) * kPrimeY, k = __float2int_rn(z0) * kPrimeZ;
    bool x_ge_y = xi >= yi, y_ge_z = yi >= zi, x_ge_z = xi >= zi;
    float g = G3 * (xi + yi + zi);
    x0 = xi - g; y0 = yi - g; z0 = zi - g;
    bool i1 = x_ge_y && x_ge_z, j1 = y_ge_z && !x_ge_y, k1 = !x_ge_z && !y_ge_z;
    bool i2 = x_ge_y || x_ge_z, j2 = !x_ge_y || y_ge_z, k2 = x_ge_z && y_ge_z;   // k2 is used N-masked
    float x1 = (i1 ? x0 - 1.0f : x0) + G3, y1 = (j1 ? y0 - 1.0f : y0) + G3, z1 = (k1 ? z0 - 1.0f : z0) + G3;
    float x2 = (i2 ? x0 - 1.0f : x0) + G3 * 2, y2 = (j2 ? y0 - 1.0f : y0) + G3 * 2, z2 = (k2 ? z0 : z0 - 1.0f) + G3 * 2;
    float x3 = x0 + (G3 * 3 - 1), y3 = y0 + (G3 * 3 - 1), z3 = z0 + (G3 * 3 - 1);
    auto falloff = [](float a, float b, float c) {
        float t = fmaf(-a, a, fmaf(-b, b, fmaf(-c, c, 0.6f)));
        t = fmaxf(t, 0.0f);
        t *= t;
        return t * t;
    };
    float t0 = falloff(x0, y0, z0), t1 = falloff(x1, y1, z1), t2 = falloff(x2, y2, z2), t3 = falloff(x3, y3, z3);
    float n0 = grad_dot(hash3(seed ^ k, i ^ j), x0, y0, z0);
    float n1 = grad_dot(hash3(seed ^ (k1 ? k + kPrimeZ : k), (i1 ? i + kPrimeX : i) ^ (j1 ? j + kPrimeY : j)), x1, y1, z1);
    float n2 = grad_dot(hash3(seed ^ (k2 ? k : k + kPrimeZ), (i2 ? i + kPrimeX : i) ^ (j2 ? j + kPrimeY : j)), x2, y2, z2);
    float n3 = grad_dot(hash3(seed ^ (k + kPrimeZ), (i + kPrimeX) ^ (j + kPrimeY)), x3, y3, z3);
    return 32.69428253173828125f * fmaf(n0, t0, fmaf(n1, t1, fmaf(n2, t2, n3 * t3)));
}

// (float)(v & 0x3ff) - 511.5 without the int->float conversion (an XU-pipe instruction, 16 lanes/clk/SM, and the cellular
// walk needs 81 of them per point): with the 10 bits placed one position up in the mantissa of 2^22 (ulp 0.5) the float IS
// 4194304 + v, and 4194304 + 511.5 is representable, so one FADD gives the exact same value.  t = the hash shifted so that the
// wanted 10 bits sit at bits 1..10.
// The mask and the exponent pattern cannot both be immediates of one LOP3, and left to itself the compiler emits two;
// with the exponent pattern held in a register, (t & 0x7fe) | k is a single LOP3 (27 x 3 of them per point).
__device__ __forceinline__ float cell_jitter_component(int t) {
    unsigned k, r;
    asm("mov.b32 %0, 0x4A800000;" : "=r"(k));
    asm("lop3.b32 %0, %1, 0x7fe, %2, 0xEA;" : "=r"(r) : "r"(t), "r"(k));       // (a & b) | c
    return __uint_as_float(r) - 4194815.5f;
}

// MUFU.RSQ alone: rsqrtf() wraps it in a denormal fix-up (FSETP + 2 FMUL) that the squared length of a jitter vector
// (>= 0.75) never needs
__device__ __forceinline__ float rsqrt_fast(float v) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}

__device__ __forceinline__ float cellular3(int seed, float x, float y, float z) {
    const float jitter = 0.396144f;                   // kJitter3D * jitter modifier 1
    float best = CUDART_INF_F;                        // Index0 of the sorted distances = the minimum
    const int xb = __float2int_rn(x) - 1, yb = __float2int_rn(y) - 1, zb = __float2int_rn(z) - 1;
    // the three offsets and primed coordinates per axis, formed once (the running += of Cellular.inl gives the same values:
    // small integers and their sums are exact)
    float xcf[3], ycf[3], zcf[3];
    int xc[3], yc[3], zc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        xcf[i] = ((float)xb - x) + (float)i; ycf[i] = ((float)yb - y) + (float)i; zcf[i] = ((float)zb - z) + (float)i;
        xc[i] = (xb + i) * kPrimeX; yc[i] = (yb + i) * kPrimeY; zc[i] = (zb + i) * kPrimeZ;
    }
#pragma unroll
    for (int xi = 0; xi < 3; ++xi) {
#pragma unroll
        for (int yi = 0; yi < 3; ++yi) {
            const int sxy = seed ^ xc[xi] ^ yc[yi];
#pragma unroll
            for (int zi = 0; zi < 3; ++zi) {
                const int hash = (sxy ^ zc[zi]) * 0x27d4eb2d;   // HashPrimesHB, Utils.inl:304-312
                float xd = cell_jitter_component(hash << 1), yd = cell_jitter_component(hash >> 9), zd = cell_jitter_component(hash >> 19);
                const float invMag = jitter * rsqrt_fast(fmaf(xd, xd, fmaf(yd, yd, zd * zd)));
                xd = fmaf(xd, invMag, xcf[xi]); yd = fmaf(yd, invMag, ycf[yi]); zd = fmaf(zd, invMag, zcf[zi]);
                best = fminf(best, fmaf(zd, zd, fmaf(yd, yd, xd * xd)));
            }
        }
    }
    return best;
}

// (Unrolling this one-point routine of the HASH variant at five octaves was measured slower in round 1 — C2 1.733 -> 1.80 ms,
// spills under the 128-register cap; the cell-table variant below does unroll its one-point routine.)
template <int OCT, class LUT>
__device__ __forceinline__ float fbm_perlin_n(LUT lut, const FbmField& F, float x, float y, float z) {
    const int n = OCT > 0 ? OCT : F.octaves;
    int seed = F.seed;
    float amp = F.bounding * kPerlinScale;       // one multiply per fBm sample instead of one per octave
    float sum = perlin3_lut(lut, seed, x, y, z) * amp;
#pragma unroll
    for (int i = 1; i < n; ++i) {
        seed += 1;
        amp *= F.gain;
        x *= F.lacunarity; y *= F.lacunarity; z *= F.lacunarity;
        sum = fmaf(perlin3_lut(lut, seed, x, y, z), amp, sum);
    }
    return sum;
}
template <class LUT>
__device__ __forceinline__ float fbm_perlin(LUT lut, const FbmField& F, float x, float y, float z) {
    return fbm_perlin_n<0>(lut, F, x, y, z);
}

// ---- two sample points per thread on sm_100a's packed FP32 (FFMA2 / FMUL2 / FADD2) ---------------------
// The fBm march is issue-bound, not FMA-pipe bound (profiles/r01_march_fbm_lds128.txt: 88 % issue-active, FMA pipe
// 51 %).  Blackwell's fma/mul/add.f32x2 do two IEEE binary32 operations per instruction on a 64-bit register pair at
// the same FMA rate (tools/probes/ffma2_probe.cu: 126.6 FMA/clk/SM packed vs 123.7 scalar), i.e. half the issue
// slots.  Evaluating the lattice arithmetic of TWO points (A, B = two consecutive samples of the sun light-march) in
// the two halves packs everything that is not a hash, a table fetch or a gradient dot product (those read
// per-point LDS results): position scaling, fractions, the quintic fades, all seven lerps and the octave
// accumulation: 39 of ~117 instructions per octave and point become 20.  Each half performs exactly the operations
// of perlin3_lut/fbm_perlin in the same order with the same roundings, so the result is bit-identical to them.
struct f2 { unsigned long long v; };        // low half = point A, high half = point B, in one 64-bit register pair
__device__ __forceinline__ f2 make_f2(float a, float b) {
    f2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float f2_a(f2 v) { float a; asm("{ .reg .b32 t; mov.b64 {%0, t}, %1; }" : "=f"(a) : "l"(v.v)); return a; }
__device__ __forceinline__ float f2_b(f2 v) { float b; asm("{ .reg .b32 t; mov.b64 {t, %0}, %1; }" : "=f"(b) : "l"(v.v)); return b; }
__device__ __forceinline__ f2 splat2(float s) { return make_f2(s, s); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
    f2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d.v) : "l"(a.v), "l"(b.v), "l"(c.v));
    return d;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
    f2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d.v) : "l"(a.v), "l"(b.v));
    return d;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
    f2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d.v) : "l"(a.v), "l"(b.v));
    return d;
}
__device__ __forceinline__ f2 sub2(f2 a, f2 b) {
    f2 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d.v) : "l"(a.v), "l"(b.v));
    return d;
}
__device__ __forceinline__ f2 quintic2(f2 t) {        // t*t*t * fma(t, fma(t, 6, -15), 10), as quintic()
    const f2 p = fma2(t, fma2(t, splat2(6.0f), splat2(-15.0f)), splat2(10.0f));
    return mul2(mul2(mul2(t, t), t), p);
}
__device__ __forceinline__ f2 lerp2(f2 a, f2 b, f2 t) { return fma2(t, sub2(b, a), a); }

// the two corners of an x edge for both points; returns lerp(d0, d1, s) per point like grad_edge_lut
template <class LUT>
__device__ __forceinline__ f2 grad_edge_lut2(LUT lut, int tA0, int tA1, int tB0, int tB1,
                                             f2 fx0, f2 fx1, f2 fy, f2 fz, f2 s) {
    const float4 a0 = grad_lut_fetch(lut, tA0), a1 = grad_lut_fetch(lut, tA1);
    const float4 b0 = grad_lut_fetch(lut, tB0), b1 = grad_lut_fetch(lut, tB1);
    const float fyA = f2_a(fy), fyB = f2_b(fy), fzA = f2_a(fz), fzB = f2_b(fz);
    const float d0A = fmaf(fzA, a0.z, fmaf(fyA, a0.y, f2_a(fx0) * a0.x));
    const float d0B = fmaf(fzB, b0.z, fmaf(fyB, b0.y, f2_b(fx0) * b0.x));
    const float eA = fmaf(fzA, a1.z, fmaf(fyA, a1.y, fmaf(f2_a(fx1), a1.x, -d0A)));
    const float eB = fmaf(fzB, b1.z, fmaf(fyB, b1.y, fmaf(f2_b(fx1), b1.x, -d0B)));
    const f2 d0 = make_f2(d0A, d0B), e = make_f2(eA, eB);
    return fma2(s, e, d0);
}

template <class LUT>
__device__ __forceinline__ f2 perlin3_lut2(LUT lut, int seed, f2 x, f2 y, f2 z) {
    // The lattice coordinate as an integer without F2I (an XU-pipe instruction, 4 lanes/clk per scheduler): for an integral
    // |v| <= 2^22 the binary32 sum v + 1.5*2^23 is exact and its bit pattern is 0x4B400000 + v, so v * prime is one IMAD
    // on those bits with the constant -0x4B400000 * prime as addend.  Coordinates beyond 2^22 have no fraction bits
    // left for a noise lattice anyway.
    // (floor itself on the FMA pipe as rint(v - 1/2) was measured equal within noise in round 1 and slower on the cell-table
    // variant in round 2: profiles/r02_fbm_cells_experiments.txt)
    const f2 magic = splat2(12582912.0f);
    f2 xs = make_f2(floorf(f2_a(x)), floorf(f2_b(x))), ys = make_f2(floorf(f2_a(y)), floorf(f2_b(y))), zs = make_f2(floorf(f2_a(z)), floorf(f2_b(z)));
    const f2 xt = add2(xs, magic), yt = add2(ys, magic), zt = add2(zs, magic);
    constexpr unsigned int kBias = 0x4B400000u;
    constexpr int kx = (int)(0u - kBias * (unsigned int)kPrimeX), ky = (int)(0u - kBias * (unsigned int)kPrimeY), kz = (int)(0u - kBias * (unsigned int)kPrimeZ);
    const int ax0 = __float_as_int(f2_a(xt)) * kPrimeX + kx, ay0 = __float_as_int(f2_a(yt)) * kPrimeY + ky, az0 = __float_as_int(f2_a(zt)) * kPrimeZ + kz;
    const int bx0 = __float_as_int(f2_b(xt)) * kPrimeX + kx, by0 = __float_as_int(f2_b(yt)) * kPrimeY + ky, bz0 = __float_as_int(f2_b(zt)) * kPrimeZ + kz;
    const int ax1 = ax0 + kPrimeX, ay1 = ay0 + kPrimeY, az1 = az0 + kPrimeZ;
    const int bx1 = bx0 + kPrimeX, by1 = by0 + kPrimeY, bz1 = bz0 + kPrimeZ;
    const f2 xf0 = sub2(x, xs), yf0 = sub2(y, ys), zf0 = sub2(z, zs);
    const f2 m1 = splat2(-1.0f);
    const f2 xf1 = add2(xf0, m1), yf1 = add2(yf0, m1), zf1 = add2(zf0, m1);
    xs = quintic2(xf0); ys = quintic2(yf0); zs = quintic2(zf0);
    const int asz0 = seed ^ az0, asz1 = seed ^ az1, bsz0 = seed ^ bz0, bsz1 = seed ^ bz1;
    const f2 a = lerp2(grad_edge_lut2(lut, asz0 ^ ax0 ^ ay0, asz0 ^ ax1 ^ ay0, bsz0 ^ bx0 ^ by0, bsz0 ^ bx1 ^ by0, xf0, xf1, yf0, zf0, xs),
                       grad_edge_lut2(lut, asz0 ^ ax0 ^ ay1, asz0 ^ ax1 ^ ay1, bsz0 ^ bx0 ^ by1, bsz0 ^ bx1 ^ by1, xf0, xf1, yf1, zf0, xs), ys);
    const f2 b = lerp2(grad_edge_lut2(lut, asz1 ^ ax0 ^ ay0, asz1 ^ ax1 ^ ay0, bsz1 ^ bx0 ^ by0, bsz1 ^ bx1 ^ by0, xf0, xf1, yf0, zf1, xs),
                       grad_edge_lut2(lut, asz1 ^ ax0 ^ ay1, asz1 ^ ax1 ^ ay1, bsz1 ^ bx0 ^ by1, bsz1 ^ bx1 ^ by1, xf0, xf1, yf1, zf1, xs), ys);
    return lerp2(a, b, zs);
}

// OCT > 0: the octave count as a compile-time constant — the loop is unrolled, so the independent octaves (their positions
// are x * lacunarity^i) interleave in the instruction stream instead of each waiting for its own dependent chain.
// Same operations on the same values in the same order per octave and in the sum: the result does not change.
#ifndef VOLRT_FBM_UNROLL_OCTAVES
#define VOLRT_FBM_UNROLL_OCTAVES 5      // the C2 / C3 field; other counts take the loop.  C2 1.767 -> 1.72 ms, C3 17.16 -> 16.79 ms
#endif
template <int OCT, class LUT>
__device__ __forceinline__ f2 fbm_perlin2_n(LUT lut, const FbmField& F, f2 x, f2 y, f2 z) {
    const int n = OCT > 0 ? OCT : F.octaves;
    int seed = F.seed;
    float amp = F.bounding * kPerlinScale;
    const f2 lac = splat2(F.lacunarity);
    f2 sum = mul2(perlin3_lut2(lut, seed, x, y, z), splat2(amp));
#pragma unroll
    for (int i = 1; i < n; ++i) {
        seed += 1;
        amp *= F.gain;
        x = mul2(x, lac); y = mul2(y, lac); z = mul2(z, lac);
        sum = fma2(perlin3_lut2(lut, seed, x, y, z), splat2(amp), sum);
    }
    return sum;
}
template <class LUT>
__device__ __forceinline__ f2 fbm_perlin2(LUT lut, const FbmField& F, f2 x, f2 y, f2 z) {
    if (VOLRT_FBM_UNROLL_OCTAVES > 0 && F.octaves == VOLRT_FBM_UNROLL_OCTAVES) return fbm_perlin2_n<VOLRT_FBM_UNROLL_OCTAVES>(lut, F, x, y, z);
    return fbm_perlin2_n<0>(lut, F, x, y, z);
}

// ---- the same fBm with the lattice hashes read from per-cell tables (volrt_internal.h: FbmCells) ------------------------
// profiles/r01_march_fbm_pairs_unrolled.txt: the march is bound by instruction count, and of the ~100 instructions per
// octave and point 40 are the lattice hash: three IMADs by the primes, three adds, two seed XORs and per corner a
// three-input XOR, an IMAD, a shift and a mask.  None of it depends on anything but the integer cell coordinates, the
// octave and the seed — and the box holds only (frequency * lacunarity^o + 1)^3 cells per octave (C2: 5^3 ... 63^3).
// So vr_set_params evaluates the hashes once per cell (fbm_cell_table_kernel: 8 bytes per cell = the eight corners'
// gradient-table offsets, 2.4 MB for the five octaves of C2/C3, L1/L2 resident) and a sample fetches them with ONE
// 64-bit load: cell index in float (exact below 2^24: fma(zs, n^2, fma(ys, n, xs + K)), K = 1.5*2^23 + the index of
// cell (0,0,0)), whose bit pattern is 0x4B400000 + index, so the address is one IMAD.WIDE with an adjusted base.
// Everything after the hash — fractions, fades, gradient dot products, lerps, in the same order — is unchanged, and the
// table holds exactly what the hash would have produced: the value is bit-identical to perlin3_lut / perlin3_lut2.
// What a cell entry holds and how a sample uses it (round 2, third form — the first two are in the experiments log):
// the eight corners of a cell are taken as four PAIRS, (x0y0, x0y1) and (x1y0, x1y1) at z0 and at z1, and a pair fetches both
// its gradients with one index from a 256-entry shared-memory table (PairLut: pair (n_a, n_b) of the 16 gradients -> their
// components interleaved, (ax, bx, ay, by | az, bz)).  The dot products of the two x edges of a z plane then run side by side in
// the halves of the packed FP32 instructions: d0 = g.(fx0, fy, fz) for corners x0y0 | x0y1 is FMUL2 + 2 FFMA2 (fy differs between
// the halves: (fy0, fy1) is a register pair; fx0 and fz are broadcast operands), e' = d0 - g.(fx1, fy, fz) for x1y0 | x1y1 three
// more FFMA2 with negated operands, and the x lerp d0 - s e' one — seven packed instructions for four corners where the
// one-corner-at-a-time form took 28 scalar ones, four index extractions instead of eight, 24 LDS write-back cycles instead of
// 32.  Each half performs the operations of grad_edge_lut in the same order with the same roundings (negating both factors'
// sign and the accumulated value is exact), so the value is bit-identical to perlin3_lut.
// ncu source page of the C2 kernel before this (profiles/r02_c2_source_hot.txt): 70.6 instructions per octave and point in the
// light-march pairs, 26.7 of them the scalar corner FFMAs / FMULs, 8 PRMT, 8 LDS.
struct __align__(16) PairLut { float4 a[256]; float4 b[256]; };      // a[i] = (ax, bx, ay, by), b[i] = (az, bz, -, -); i = n_a * 16 + n_b
#ifndef VOLRT_PAIR_LUT_OFFSET
#define VOLRT_PAIR_LUT_OFFSET 512            // of MarchSmem::pair from the start of the CTA's static shared memory (march_kernels.cu asserts it)
#endif
__device__ __forceinline__ void pair_lut_init(PairLut* t, int tid, int nthreads) {
    for (int i = tid; i < 256; i += nthreads) {
        float ax, ay, az, bx, by, bz;
        grad_components(i >> 4, ax, ay, az);
        grad_components(i & 15, bx, by, bz);
        t->a[i] = make_float4(ax, bx, ay, by);
        t->b[i] = make_float4(az, bz, 0.0f, 0.0f);
    }
}
// byte offset of pair (ta, tb) in PairLut::a / ::b from the corners' lattice hashes (what a cell entry stores, 16 bits per pair)
__device__ __forceinline__ unsigned int pair_lut_offset(int ta, int tb) { return (grad_lut_offset(ta) << 4) | grad_lut_offset(tb); }

struct CellWords { unsigned int w[2]; };     // .w[0]: pairs of the z0 plane (x0 pair | x1 pair << 16), .w[1]: of the z1 plane
__device__ __forceinline__ CellWords fbm_cell_fetch(const FbmCellOctave& C, float xs, float ys, float zs) {
    const float t = fmaf(zs, C.n2f, fmaf(ys, C.nf, xs + C.kf));
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(C.base_adj + (unsigned long long)__float_as_uint(t) * 8ull));
    CellWords W;
    W.w[0] = v.x; W.w[1] = v.y;
    return W;
}
// (floor() on the FMA pipe instead of FRND — x + (2^23 - 1/2) rounds once to 2^23 + rint(x - 1/2) — was measured: 1.457 vs
// 1.441 ms at C2, profiles/r02_fbm_cells_experiments.txt; the XU pipe at 20 % is not what binds.)
__device__ __forceinline__ float cell_floor(float x) { return floorf(x); }
__device__ __forceinline__ f2 cell_floor2(f2 x) { return make_f2(floorf(f2_a(x)), floorf(f2_b(x))); }

struct PairGrad { f2 x, y, z; };             // components of the pair's two gradients, side by side
__device__ __forceinline__ PairGrad pair_lut_at(GradLutFixed, unsigned int off) {
    PairGrad g;
    asm("ld.shared.v2.b64 {%0, %1}, [%2 + %3];" : "=l"(g.x.v), "=l"(g.y.v) : "r"(off), "n"(VOLRT_LUT_FIXED_ADDR + VOLRT_PAIR_LUT_OFFSET));
    asm("ld.shared.b64 %0, [%1 + %2];" : "=l"(g.z.v) : "r"(off), "n"(VOLRT_LUT_FIXED_ADDR + VOLRT_PAIR_LUT_OFFSET + 4096));
    return g;
}
// one z plane of a cell: the x edges at y0 | y1 lerped along x, then along y
__device__ __forceinline__ float cell_plane(GradLutFixed lut, unsigned int w, float fx0, float fx1, f2 fy, float fz, float sx, float sy) {
    const PairGrad g0 = pair_lut_at(lut, w & 0xffffu), g1 = pair_lut_at(lut, w >> 16);
    f2 d0 = mul2(splat2(fx0), g0.x);
    d0 = fma2(fy, g0.y, d0);
    d0 = fma2(splat2(fz), g0.z, d0);
    const f2 nfy = make_f2(-f2_a(fy), -f2_b(fy));
    f2 e = fma2(splat2(-fx1), g1.x, d0);          // d0 - g1.(fx1, fy, fz): the negative of grad_edge_lut's e, to the bit
    e = fma2(nfy, g1.y, e);
    e = fma2(splat2(-fz), g1.z, e);
    const f2 r = fma2(splat2(-sx), e, d0);        // = fma(sx, e_of_grad_edge_lut, d0)
    return lerpf(f2_a(r), f2_b(r), sy);
}
template <int I = -1>
__device__ __forceinline__ float perlin3_cells(GradLutFixed lut, const FbmCellOctave& C, float x, float y, float z) {
    float xs = cell_floor(x), ys = cell_floor(y), zs = cell_floor(z);
    const CellWords W = fbm_cell_fetch(C, xs, ys, zs);
    const float xf0 = x - xs, yf0 = y - ys, zf0 = z - zs;
    const float xf1 = xf0 - 1.0f, zf1 = zf0 - 1.0f;
    xs = quintic(xf0); ys = quintic(yf0); zs = quintic(zf0);
    const f2 fy = add2(splat2(yf0), make_f2(0.0f, -1.0f));
    const float a = cell_plane(lut, W.w[0], xf0, xf1, fy, zf0, xs, ys);
    const float b = cell_plane(lut, W.w[1], xf0, xf1, fy, zf1, xs, ys);
    return lerpf(a, b, zs);
}
// two points: what does not go through the table (floor, fractions, fades, the z lerp) stays packed ACROSS the points
template <int I = -1>
__device__ __forceinline__ f2 perlin3_cells2(GradLutFixed lut, const FbmCellOctave& C, f2 x, f2 y, f2 z) {
    f2 xs = cell_floor2(x), ys = cell_floor2(y), zs = cell_floor2(z);
    const CellWords WA = fbm_cell_fetch(C, f2_a(xs), f2_a(ys), f2_a(zs));
    const CellWords WB = fbm_cell_fetch(C, f2_b(xs), f2_b(ys), f2_b(zs));
    const f2 xf0 = sub2(x, xs), yf0 = sub2(y, ys), zf0 = sub2(z, zs);
    const f2 m1 = splat2(-1.0f);
    const f2 xf1 = add2(xf0, m1), zf1 = add2(zf0, m1);
    xs = quintic2(xf0); ys = quintic2(yf0); zs = quintic2(zf0);
    // (fy0, fy0 - 1) per point as ONE packed add of the broadcast fraction and the pair (0, -1): t + 0 is t (fractions are never -0)
    const f2 c01 = make_f2(0.0f, -1.0f);
    const f2 fyA = add2(splat2(f2_a(yf0)), c01), fyB = add2(splat2(f2_b(yf0)), c01);
    const float a0 = cell_plane(lut, WA.w[0], f2_a(xf0), f2_a(xf1), fyA, f2_a(zf0), f2_a(xs), f2_a(ys));
    const float a1 = cell_plane(lut, WA.w[1], f2_a(xf0), f2_a(xf1), fyA, f2_a(zf1), f2_a(xs), f2_a(ys));
    const float b0 = cell_plane(lut, WB.w[0], f2_b(xf0), f2_b(xf1), fyB, f2_b(zf0), f2_b(xs), f2_b(ys));
    const float b1 = cell_plane(lut, WB.w[1], f2_b(xf0), f2_b(xf1), fyB, f2_b(zf1), f2_b(xs), f2_b(ys));
    return lerp2(make_f2(a0, b0), make_f2(a1, b1), zs);
}
// the compile-time octave count (C2 / C3: 5), octave by octave
template <int I, int OCT, class LUT>
__device__ __forceinline__ float fbm_cells_unrolled(LUT lut, const FbmField& F, const FbmCells& T, float x, float y, float z, float amp, float sum) {
    if constexpr (I >= OCT) return sum;
    else {
        const float p = perlin3_cells<I>(lut, T.oct[I], x, y, z);
        sum = I == 0 ? p * amp : fmaf(p, amp, sum);
        return fbm_cells_unrolled<I + 1, OCT>(lut, F, T, x * F.lacunarity, y * F.lacunarity, z * F.lacunarity, amp * F.gain, sum);
    }
}
template <int I, int OCT, class LUT>
__device__ __forceinline__ f2 fbm_cells_unrolled2(LUT lut, const FbmField& F, const FbmCells& T, f2 x, f2 y, f2 z, f2 lac, float amp, f2 sum) {
    if constexpr (I >= OCT) return sum;
    else {
        const f2 p = perlin3_cells2<I>(lut, T.oct[I], x, y, z);
        sum = I == 0 ? mul2(p, splat2(amp)) : fma2(p, splat2(amp), sum);
        return fbm_cells_unrolled2<I + 1, OCT>(lut, F, T, mul2(x, lac), mul2(y, lac), mul2(z, lac), lac, amp * F.gain, sum);
    }
}

template <int OCT, class LUT>
__device__ __forceinline__ float fbm_perlin_cells_n(LUT lut, const FbmField& F, const FbmCells& T, float x, float y, float z) {
    const int n = OCT > 0 ? OCT : F.octaves;
    float amp = F.bounding * kPerlinScale;
    if constexpr (OCT > 0) return fbm_cells_unrolled<0, OCT>(lut, F, T, x, y, z, amp, 0.0f);
    float sum = perlin3_cells(lut, T.oct[0], x, y, z) * amp;
#pragma unroll
    for (int i = 1; i < n; ++i) {
        amp *= F.gain;
        x *= F.lacunarity; y *= F.lacunarity; z *= F.lacunarity;
        sum = fmaf(perlin3_cells(lut, T.oct[i], x, y, z), amp, sum);
    }
    return sum;
}
template <int OCT, class LUT>
__device__ __forceinline__ f2 fbm_perlin_cells2_n(LUT lut, const FbmField& F, const FbmCells& T, f2 x, f2 y, f2 z) {
    const int n = OCT > 0 ? OCT : F.octaves;
    float amp = F.bounding * kPerlinScale;
    const f2 lac = splat2(F.lacunarity);
    if constexpr (OCT > 0) return fbm_cells_unrolled2<0, OCT>(lut, F, T, x, y, z, lac, amp, make_f2(0.0f, 0.0f));
    f2 sum = mul2(perlin3_cells2(lut, T.oct[0], x, y, z), splat2(amp));
#pragma unroll
    for (int i = 1; i < n; ++i) {
        amp *= F.gain;
        x = mul2(x, lac); y = mul2(y, lac); z = mul2(z, lac);
        sum = fma2(perlin3_cells2(lut, T.oct[i], x, y, z), splat2(amp), sum);
    }
    return sum;
}
// the one-point routine (the primary sample) unrolled too: C2 1.468 -> 1.441 ms on B200 (the hash variant lost by it: spills)
#ifndef VOLRT_FBM_CELLS_UNROLL_SCALAR
#define VOLRT_FBM_CELLS_UNROLL_SCALAR 5
#endif

__device__ __forceinline__ float fbm_edge(const FbmField& F, float ux, float uy, float uz) {
    const float m = fmaxf(fmaxf(fabsf(ux - 0.5f), fabsf(uy - 0.5f)), fabsf(uz - 0.5f));
    const float t = __saturatef((m - F.falloff0) * F.inv_falloff);
    return 1.0f - t * t * (3.0f - 2.0f * t);
}

// density_fbm at two points; a point whose edge factor is <= 0 gets exactly 0 like the scalar early-out
template <class LUT>
__device__ __forceinline__ f2 density_fbm2(LUT lut, const FbmField& F, f2 ux, f2 uy, f2 uz) {
    const float ea = fbm_edge(F, f2_a(ux), f2_a(uy), f2_a(uz)), eb = fbm_edge(F, f2_b(ux), f2_b(uy), f2_b(uz));
    if (ea <= 0.0f && eb <= 0.0f) return make_f2(0.0f, 0.0f);
    const f2 fr = splat2(F.frequency);
    const f2 f = fbm_perlin2(lut, F, mul2(ux, fr), mul2(uy, fr), mul2(uz, fr));
    const float fa = f2_a(f), fb = f2_b(f);
    return make_f2(ea <= 0.0f ? 0.0f : fmaxf(0.0f, (0.5f * fa + 0.5f) - F.coverage) * ea,
                   eb <= 0.0f ? 0.0f : fmaxf(0.0f, (0.5f * fb + 0.5f) - F.coverage) * eb);
}

// rho = max(0, .5 fbm + .5 - coverage) * (1 - smoothstep(f0, f1, maxabs(P - 1/2)))
template <class LUT>
__device__ __forceinline__ float density_fbm(LUT lut, const FbmField& F, float ux, float uy, float uz) {
    float m = fmaxf(fmaxf(fabsf(ux - 0.5f), fabsf(uy - 0.5f)), fabsf(uz - 0.5f));
    float t = __saturatef((m - F.falloff0) * F.inv_falloff);
    float edge = 1.0f - t * t * (3.0f - 2.0f * t);
    if (edge <= 0.0f) return 0.0f;
    float f = fbm_perlin(lut, F, ux * F.frequency, uy * F.frequency, uz * F.frequency);
    return fmaxf(0.0f, (0.5f * f + 0.5f) - F.coverage) * edge;
}

// The cell tables cover the unit box only.  vr_set_params enables them when falloff1 <= 1/2, so a positive edge fade
// already means maxabs(P - 1/2) < 1/2; what is left to keep out is a NaN coordinate (fmaxf drops NaNs, the fade of a NaN
// saturates to 0 and the edge would read 1): the maximum is taken NaN-propagating and anything not below 3/4 gets edge 0.
__device__ __forceinline__ float max_nan(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float fbm_edge_in_box(const FbmField& F, float ux, float uy, float uz) {
    const float m = max_nan(max_nan(fabsf(ux - 0.5f), fabsf(uy - 0.5f)), fabsf(uz - 0.5f));
    const float t = __saturatef((m - F.falloff0) * F.inv_falloff);
    const float e = 1.0f - t * t * (3.0f - 2.0f * t);
    return m < 0.75f ? e : 0.0f;
}
// both points must be inside the unit box (the light-march checks inside01 before it pairs two samples)
template <class LUT>
__device__ __forceinline__ f2 density_fbm_cells2(LUT lut, const FbmField& F, const FbmCells& T, f2 ux, f2 uy, f2 uz) {
    const float ea = fbm_edge(F, f2_a(ux), f2_a(uy), f2_a(uz)), eb = fbm_edge(F, f2_b(ux), f2_b(uy), f2_b(uz));
    if (ea <= 0.0f && eb <= 0.0f) return make_f2(0.0f, 0.0f);
    const f2 fr = splat2(F.frequency);
    const f2 x = mul2(ux, fr), y = mul2(uy, fr), z = mul2(uz, fr);
    const f2 f = (VOLRT_FBM_UNROLL_OCTAVES > 0 && F.octaves == VOLRT_FBM_UNROLL_OCTAVES) ? fbm_perlin_cells2_n<VOLRT_FBM_UNROLL_OCTAVES>(lut, F, T, x, y, z)
                                                                                         : fbm_perlin_cells2_n<0>(lut, F, T, x, y, z);
    const float fa = f2_a(f), fb = f2_b(f);
    return make_f2(ea <= 0.0f ? 0.0f : fmaxf(0.0f, (0.5f * fa + 0.5f) - F.coverage) * ea,
                   eb <= 0.0f ? 0.0f : fmaxf(0.0f, (0.5f * fb + 0.5f) - F.coverage) * eb);
}
template <class LUT>
__device__ __forceinline__ float density_fbm_cells(LUT lut, const FbmField& F, const FbmCells& T, float ux, float uy, float uz) {
    const float edge = fbm_edge_in_box(F, ux, uy, uz);
    if (edge <= 0.0f) return 0.0f;
    const float x = ux * F.frequency, y = uy * F.frequency, z = uz * F.frequency;
    const float f = (VOLRT_FBM_CELLS_UNROLL_SCALAR > 0 && F.octaves == VOLRT_FBM_CELLS_UNROLL_SCALAR) ? fbm_perlin_cells_n<VOLRT_FBM_CELLS_UNROLL_SCALAR>(lut, F, T, x, y, z)
                                                                                                       : fbm_perlin_cells_n<0>(lut, F, T, x, y, z);
    return fmaxf(0.0f, (0.5f * f + 0.5f) - F.coverage) * edge;
}

}  // namespace volrt
