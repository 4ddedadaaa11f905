// Device evaluation of FastNoise2 0.10.0 generator nodes beyond the three the reference's bake uses
// (SURVEY.md §8(f) rank 3).  Behaviour restated from vendor/noise/include/FastNoise/Generators/:
//   white3, value3                BasicGenerators.inl:20-33, Value.inl:30-57, Utils.inl:300-308 (GetValueCoord)
//   opensimplex2_3                Simplex.inl:309-372
//   opensimplex2s_3               Simplex.inl:443-524
//   checkerboard3                 BasicGenerators.inl:36-50
//   cellular_*                    Cellular.inl:95-160 (value), :307-366 + :439-474 (distance, return types), :520-580 (lookup)
//   calc_distance                 Utils.inl:329-372
//   warp_gradient3                DomainWarp.inl:69-122
//   fs_cos / fs_sin               FastSIMD/FunctionList.h:718-747
//   fs_log / fs_exp / fs_pow      FastSIMD/FunctionList.h:749-857
// 3-D only (the bake and the march are 3-D).  FMAs contract and 1/sqrt is rsqrtf as everywhere in noise_device.cuh.
#pragma once
#include "noise_device.cuh"

namespace volrt {

__device__ __forceinline__ float value_coord(int seed, int x, int y, int z) {       // GetValueCoord
    int h = seed ^ x ^ y ^ z;
    h *= h * 0x27d4eb2d;
    return (float)h * (1.0f / 2147483648.0f);          // (float)INT_MAX rounds to 2^31
}

__device__ __forceinline__ float white3(int seed, float x, float y, float z) {
    const int xi = __float_as_int(x), yi = __float_as_int(y), zi = __float_as_int(z);
    return value_coord(seed, (xi ^ (xi >> 16)) * kPrimeX, (yi ^ (yi >> 16)) * kPrimeY, (zi ^ (zi >> 16)) * kPrimeZ);
}

__device__ __forceinline__ float hermite(float t) { return t * t * fmaf(-t, 2.0f, 3.0f); }

__device__ __forceinline__ float value3(int seed, float x, float y, float z) {
    float xs = floorf(x), ys = floorf(y), zs = floorf(z);
    const int x0 = __float2int_rn(xs) * kPrimeX, y0 = __float2int_rn(ys) * kPrimeY, z0 = __float2int_rn(zs) * kPrimeZ;
    const int x1 = x0 + kPrimeX, y1 = y0 + kPrimeY, z1 = z0 + kPrimeZ;
    xs = hermite(x - xs); ys = hermite(y - ys); zs = hermite(z - zs);
    const float a = lerpf(lerpf(value_coord(seed, x0, y0, z0), value_coord(seed, x1, y0, z0), xs),
                          lerpf(value_coord(seed, x0, y1, z0), value_coord(seed, x1, y1, z0), xs), ys);
    const float b = lerpf(lerpf(value_coord(seed, x0, y0, z1), value_coord(seed, x1, y0, z1), xs),
                          lerpf(value_coord(seed, x0, y1, z1), value_coord(seed, x1, y1, z1), xs), ys);
    return lerpf(a, b, zs);
}

__device__ __forceinline__ float opensimplex2_3(int seed, float x, float y, float z) {
    const float f = (2.0f / 3.0f) * (x + y + z);
    float xr = f - x, yr = f - y, zr = f - z;
    float val = 0.0f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float v0x = rintf(xr), v0y = rintf(yr), v0z = rintf(zr);
        const float d0x = xr - v0x, d0y = yr - v0y, d0z = zr - v0z;
        const float sx = fabsf(d0x), sy = fabsf(d0y), sz = fabsf(d0z);
        const bool dirx = fmaxf(sy, sz) <= sx;
        const bool diry = (fmaxf(sz, sx) <= sy) && !dirx;
        const bool dirz = !(dirx || diry);
        const float v1x = dirx ? v0x + copysignf(1.0f, d0x) : v0x;
        const float v1y = diry ? v0y + copysignf(1.0f, d0y) : v0y;
        const float v1z = dirz ? v0z + copysignf(1.0f, d0z) : v0z;
        const float d1x = xr - v1x, d1y = yr - v1y, d1z = zr - v1z;
        const int h0x = __float2int_rn(v0x) * kPrimeX, h0y = __float2int_rn(v0y) * kPrimeY, h0z = __float2int_rn(v0z) * kPrimeZ;
        const int h1x = __float2int_rn(v1x) * kPrimeX, h1y = __float2int_rn(v1y) * kPrimeY, h1z = __float2int_rn(v1z) * kPrimeZ;
        float t0 = fmaf(-d0z, d0z, fmaf(-d0y, d0y, fmaf(-d0x, d0x, 0.6f)));
        float t1 = fmaf(-d1z, d1z, fmaf(-d1y, d1y, fmaf(-d1x, d1x, 0.6f)));
        t0 = fmaxf(t0, 0.0f); t1 = fmaxf(t1, 0.0f);
        t0 *= t0; t0 *= t0; t1 *= t1; t1 *= t1;
        const float g0 = grad_dot(hash3(seed ^ h0z, h0x ^ h0y), d0x, d0y, d0z);
        const float g1 = grad_dot(hash3(seed ^ h1z, h1x ^ h1y), d1x, d1y, d1z);
        val = fmaf(g0, t0, fmaf(g1, t1, val));
        xr += 0.5f; yr += 0.5f; zr += 0.5f;
        seed = ~seed;
    }
    return 32.69428253173828125f * val;
}

// OpenSimplex2S ("smooth"): two half-offset cubic lattices; on each, the cell's base corner or its opposite (whichever side
// of the plane x+y+z = 3/2 the point is on) plus, per axis, the corner across that axis or its mirror — four kernels of
// radius^2 3/4 per lattice.  One vertex is (sign pattern sx,sy,sz in {+1,-1}): p = the plane test of the pattern,
// chosen = p >= 0; an axis with sign +1 steps to the far corner when chosen, an axis with sign -1 when NOT chosen.
__device__ __forceinline__ float os2s_vertex(int seed, int bx, int by, int bz, float x, float y, float z, float falloff_base, float p,
                                             bool nx, bool ny, bool nz) {
    const bool chosen = p >= 0.0f;
    float t = fmaxf(chosen ? falloff_base + p : falloff_base, 0.0f);
    t *= t; t *= t;
    const bool fx = chosen != nx, fy = chosen != ny, fz = chosen != nz;       // far corner on this axis?
    const int h = hash3(seed ^ (fz ? bz + kPrimeZ : bz), (fx ? bx + kPrimeX : bx) ^ (fy ? by + kPrimeY : by));
    return t * grad_dot(h, fx ? x - 1.0f : x, fy ? y - 1.0f : y, fz ? z - 1.0f : z);
}
__device__ __forceinline__ float opensimplex2s_3(int seed, float x, float y, float z) {
    const float f = (2.0f / 3.0f) * (x + y + z);
    const float xr = f - x, yr = f - y, zr = f - z;
    const float xb = floorf(xr), yb = floorf(yr), zb = floorf(zr);
    float xi = xr - xb, yi = yr - yb, zi = zr - zb;
    int bx = __float2int_rn(xb) * kPrimeX, by = __float2int_rn(yb) * kPrimeY, bz = __float2int_rn(zb) * kPrimeZ;
    float val = 0.0f;
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        float a = fmaf(-xi, xi, fmaf(-yi, yi, fmaf(-zi, zi, 0.75f))) * 0.5f;
        val += os2s_vertex(seed, bx, by, bz, xi, yi, zi, a, zi + yi + xi - 1.5f, false, false, false);
        a -= 0.5f;
        val += os2s_vertex(seed, bx, by, bz, xi, yi, zi, a + xi, zi + yi - xi + -0.5f, true, false, false);
        val += os2s_vertex(seed, bx, by, bz, xi, yi, zi, a + yi, xi + -0.5f + (zi - yi), false, true, false);
        val += os2s_vertex(seed, bx, by, bz, xi, yi, zi, a + zi, xi + -0.5f - (zi - yi), false, false, true);
        if (pass == 1) break;
        const bool sx = xi >= 0.5f, sy = yi >= 0.5f, sz = zi >= 0.5f;       // the second lattice: shifted by half a cell
        if (sx) bx += kPrimeX;
        if (sy) by += kPrimeY;
        if (sz) bz += kPrimeZ;
        xi += sx ? -0.5f : 0.5f; yi += sy ? -0.5f : 0.5f; zi += sz ? -0.5f : 0.5f;
        seed = ~seed;
    }
    return 144.736422163332608f * val;
}

// +1 / -1 by the parity of round(x/size) ^ round(y/size) ^ round(z/size) (round to nearest even, like cvtps2dq).  The
// reference multiplies by an approximate reciprocal of the size (rcpps + one Newton step), so exactly on a cell border
// (a half-integer quotient) the two may land on different sides; everywhere else they agree.
__device__ __forceinline__ float checkerboard3(float size, float x, float y, float z) {
    const float m = 1.0f / size;
    const int v = __float2int_rn(x * m) ^ __float2int_rn(y * m) ^ __float2int_rn(z * m);
    return (v & 1) ? -1.0f : 1.0f;
}

// DistanceFunction: 0 Euclidean, 1 EuclideanSquared, 2 Manhattan, 3 Hybrid, 4 MaxAxis
__device__ __forceinline__ float calc_distance(int fn, float dx, float dy, float dz) {
    switch (fn) {
        default:
        case 0: { const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx)); return rsqrtf(d2) * d2; }
        case 1: return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
        case 2: return fabsf(dx) + (fabsf(dy) + fabsf(dz));
        case 3: return fmaf(dx, dx, fabsf(dx)) + fmaf(dy, dy, fabsf(dy)) + fmaf(dz, dz, fabsf(dz));
        case 4: return fmaxf(fabsf(dz), fmaxf(fabsf(dy), fabsf(dx)));
    }
}

// the 27-cell walk shared by the three cellular nodes; F(hash, xd, yd, zd) sees the jittered offset to each cell point
template <class F>
__device__ __forceinline__ void cellular_walk(int seed, float x, float y, float z, float jitter, F&& visit) {
    int xc = __float2int_rn(x) - 1, ycBase = __float2int_rn(y) - 1, zcBase = __float2int_rn(z) - 1;
    float xcf = (float)xc - x;
    const float ycfBase = (float)ycBase - y, zcfBase = (float)zcBase - z;
    xc *= kPrimeX; ycBase *= kPrimeY; zcBase *= kPrimeZ;
    for (int xi = 0; xi < 3; ++xi) {
        float ycf = ycfBase; int yc = ycBase;
        for (int yi = 0; yi < 3; ++yi) {
            float zcf = zcfBase; int zc = zcBase;
            const int sxy = seed ^ xc ^ yc;
#pragma unroll
            for (int zi = 0; zi < 3; ++zi) {
                const int hash = (sxy ^ zc) * 0x27d4eb2d;
                float xd = cell_jitter_component(hash << 1), yd = cell_jitter_component(hash >> 9), zd = cell_jitter_component(hash >> 19);
                const float invMag = jitter * rsqrt_fast(fmaf(xd, xd, fmaf(yd, yd, zd * zd)));
                xd = fmaf(xd, invMag, xcf); yd = fmaf(yd, invMag, ycf); zd = fmaf(zd, invMag, zcf);
                visit(hash, xd, yd, zd);
                zcf += 1.0f; zc += kPrimeZ;
            }
            ycf += 1.0f; yc += kPrimeY;
        }
        xcf += 1.0f; xc += kPrimeX;
    }
}

__device__ __forceinline__ float cellular_value3(int seed, float x, float y, float z, float jitter_mod, int dist_fn, int value_index) {
    float value[4] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F};
    float dist[4] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F};
    cellular_walk(seed, x, y, z, 0.396144f * jitter_mod, [&](int hash, float xd, float yd, float zd) {
        float nv = (float)(1.0 / 2147483647.0) * (float)hash;
        float nd = calc_distance(dist_fn, xd, yd, zd);
#pragma unroll
        for (int i = 0; i < 4; ++i) {               // sorted insertion up to value_index + 1 (Cellular.inl:139-154)
            const bool closer = nd < dist[i];
            const float ld = dist[i], lv = value[i];
            dist[i] = closer ? nd : dist[i];
            value[i] = closer ? nv : value[i];
            if (i > value_index) break;
            nd = closer ? ld : nd;
            nv = closer ? lv : nv;
        }
    });
    return value_index == 0 ? value[0] : value_index == 1 ? value[1] : value_index == 2 ? value[2] : value[3];
}

__device__ __forceinline__ float cellular_distance3(int seed, float x, float y, float z, float jitter_mod, int dist_fn, int idx0, int idx1, int ret) {
    float d[4] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F};
    cellular_walk(seed, x, y, z, 0.396144f * jitter_mod, [&](int, float xd, float yd, float zd) {
        const float nd = calc_distance(dist_fn, xd, yd, zd);
        d[3] = fmaxf(fminf(d[3], nd), d[2]);
        d[2] = fmaxf(fminf(d[2], nd), d[1]);
        d[1] = fmaxf(fminf(d[1], nd), d[0]);
        d[0] = fminf(d[0], nd);
    });
    auto pick = [&](int i) { return i == 0 ? d[0] : i == 1 ? d[1] : i == 2 ? d[2] : d[3]; };
    float a = pick(idx0), b = pick(idx1);
    if (dist_fn == 0) {                              // GetReturn: a second 1/sqrt for Euclidean, applied in place, so an
        a *= rsqrtf(a);                              // index named twice is scaled twice (Cellular.inl:441-445)
        if (idx1 == idx0) { a *= rsqrtf(a); b = a; }
        else b *= rsqrtf(b);
    }
    switch (ret) {
        default:
        case 0: return a;
        case 1: return a + b;
        case 2: return a - b;
        case 3: return a * b;
        case 4: return a * (1.0f / b);
    }
}

// nearest cell point (lookup position before the lookup frequency is applied)
__device__ __forceinline__ void cellular_lookup3(int seed, float x, float y, float z, float jitter_mod, int dist_fn, float out[3]) {
    float best = 3.402823466e+38f, cx = 0.0f, cy = 0.0f, cz = 0.0f;
    cellular_walk(seed, x, y, z, 0.396144f * jitter_mod, [&](int, float xd, float yd, float zd) {
        const float nd = calc_distance(dist_fn, xd, yd, zd);
        const bool closer = nd < best;
        best = fminf(nd, best);
        cx = closer ? xd + x : cx; cy = closer ? yd + y : cy; cz = closer ? zd + z : cz;
    });
    out[0] = cx; out[1] = cy; out[2] = cz;
}

// DomainWarpGradient::Warp: (x,y,z) is the position already multiplied by the warp frequency; out += warp * amp;
// returns the warp length
__device__ __forceinline__ float warp_gradient3(int seed, float amp, float x, float y, float z, float out[3]) {
    float xs = floorf(x), ys = floorf(y), zs = floorf(z);
    const int x0 = __float2int_rn(xs) * kPrimeX, y0 = __float2int_rn(ys) * kPrimeY, z0 = __float2int_rn(zs) * kPrimeZ;
    const int x1 = x0 + kPrimeX, y1 = y0 + kPrimeY, z1 = z0 + kPrimeZ;
    xs = hermite(x - xs); ys = hermite(y - ys); zs = hermite(z - zs);
    float cxv[8], cyv[8], czv[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int h = (seed ^ ((c & 1) ? x1 : x0) ^ ((c & 2) ? y1 : y0) ^ ((c & 4) ? z1 : z0)) * 0x27d4eb2d;
        cxv[c] = (float)(h & 0x3ff); cyv[c] = (float)((h >> 10) & 0x3ff); czv[c] = (float)((h >> 20) & 0x3ff);
    }
    auto tri = [&](const float* v) {
        return lerpf(lerpf(lerpf(v[0], v[1], xs), lerpf(v[2], v[3], xs), ys), lerpf(lerpf(v[4], v[5], xs), lerpf(v[6], v[7], xs), ys), zs);
    };
    const float norm = 1.0f / 511.5f;
    const float wx = (tri(cxv) - 511.5f) * norm, wy = (tri(cyv) - 511.5f) * norm, wz = (tri(czv) - 511.5f) * norm;
    out[0] = fmaf(wx, amp, out[0]); out[1] = fmaf(wy, amp, out[1]); out[2] = fmaf(wz, amp, out[2]);
    const float l2 = fmaf(wx, wx, fmaf(wy, wy, wz * wz));
    return l2 * rsqrtf(l2);
}

// FastSIMD's trigonometry is a quadratic in angle^2 after range reduction (FastSIMD/FunctionList.h:718-747), ~1e-3 away
// from the true cosine; SineWave is defined by it, so it is restated rather than replaced by sinf
__device__ __forceinline__ float fs_cos(float v) {
    v = fabsf(v);
    v -= floorf(v * 0.1591549f) * 6.283185f;
    const bool q1 = v >= 1.570796f, q3 = v >= 4.7123889f;
    float a = q1 ? 3.141593f - v : v;
    a = q3 ? 6.283185f - v : a;
    a *= a;
    a = fmaf(a, fmaf(a, 0.03679168f, -0.49558072f), 0.99940307f);
    return (q1 && !q3) ? -a : a;
}
__device__ __forceinline__ float fs_sin(float v) { return fs_cos(1.570796f - v); }

// PowFloat = FS_Pow_f32 = Exp_f32(pow * Log_f32(value)) (FunctionList.h:749-857).  The reference's Log_f32 multiplies
// its polynomial by e * -2.12e-4 where the routine it was adapted from adds that term, so for arguments in
// [0.71, 1.41) its log is x - x^2/2 and elsewhere it is off by up to ~1e-2: that IS the node's behaviour, restated as is.
__device__ __forceinline__ float fs_log(float x) {
    const bool valid = x > 0.0f;
    x = fmaxf(x, __int_as_float(0x00800000));
    int i = (int)(__float_as_uint(x) >> 23);
    x = __int_as_float((__float_as_int(x) & ~0x7f800000) | 0x3f000000);
    i -= 0x7f;
    float e = (float)i + 1.0f;
    const bool lt = x < 0.707106781186547524f;
    x = lt ? x + x : x;
    x -= 1.0f;
    e = lt ? e - 1.0f : e;
    float y = 7.0376836292E-2f;
    y = fmaf(y, x, -1.1514610310E-1f);
    y = fmaf(y, x, 1.1676998740E-1f);
    y = fmaf(y, x, -1.2420140846E-1f);
    y = fmaf(y, x, 1.4249322787E-1f);
    y = fmaf(y, x, -1.6668057665E-1f);
    y = fmaf(y, x, 2.0000714765E-1f);
    y = fmaf(y, x, -2.4999993993E-1f);
    y = fmaf(y, x, 3.3333331174E-1f);
    y *= x;
    const float xx = x * x;
    y *= xx;
    y *= e * -2.12194440e-4f;
    y -= xx * 0.5f;
    x += y;
    x += e * 0.693359375f;
    return valid ? x : 0.0f;
}
__device__ __forceinline__ float fs_exp(float x) {
    x = fminf(x, 88.3762626647949f);
    x = fmaxf(x, -88.3762626647949f);
    const float fx = floorf(fmaf(x, 1.44269504088896341f, 0.5f));
    x -= fx * 0.693359375f;
    x -= fx * -2.12194440e-4f;
    float y = 1.9875691500E-4f;
    y = fmaf(y, x, 1.3981999507E-3f);
    y = fmaf(y, x, 8.3334519073E-3f);
    y = fmaf(y, x, 4.1665795894E-2f);
    y = fmaf(y, x, 1.6666665459E-1f);
    y = fmaf(y, x, 5.0000001201E-1f);
    y *= x * x;
    y += x + 1.0f;
    return y * __int_as_float((__float2int_rn(fx) + 0x7f) << 23);
}
__device__ __forceinline__ float fs_pow(float v, float p) { return fs_exp(p * fs_log(v)); }

__device__ __forceinline__ float ping_pong(float t) {
    t -= rintf(t * 0.5f) * 2.0f;
    return t < 1.0f ? t : 2.0f - t;
}

}  // namespace volrt
