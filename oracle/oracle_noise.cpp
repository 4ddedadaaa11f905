// ORACLE (test infrastructure — see volrt_oracle.h).  Scalar restatement of the
// three FastNoise2 0.10.0 generators the reference bakes its volume from, the
// uniform-grid driver, the fBm wrapper and the normalise/pack loop of the app.
// Paths relative to /root/reference/vendor/noise/include/FastNoise/Generators/.
//
// Arithmetic choices: every FS_FMulAdd is evaluated UNFUSED (mul then add), which
// is what FastNoise2's SSE2/SSE4.1 levels do (FunctionList.h:586-597); InvSqrt is
// the exact 1/sqrt, where FastNoise2 uses rsqrtps / rsqrt14ps / a bit hack
// depending on the ISA (SSE.h:446, AVX512.h:387, Scalar.h:392).  FastNoise2's own
// SIMD levels differ from each other by <= 1 LSB of the packed byte for the same
// reason, so parity for noise is tolerance-based.
#include "volrt_oracle.h"
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>
#include <algorithm>
#include <atomic>

namespace {

constexpr int32_t PRIME_X = 501125321;   // Utils.inl:9-11
constexpr int32_t PRIME_Y = 1136930381;
constexpr int32_t PRIME_Z = 1720413743;

inline int32_t imul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }
inline int32_t iadd(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
inline int32_t ishl(int32_t a, int s) { return (int32_t)((uint32_t)a << s); }

// Utils.inl:294-302
inline int32_t hash_primes(int32_t seed, int32_t x, int32_t y, int32_t z) {
    int32_t h = imul(seed ^ x ^ y ^ z, 0x27d4eb2d);
    return (h >> 15) ^ h;
}
// Utils.inl:304-312
inline int32_t hash_primes_hb(int32_t seed, int32_t x, int32_t y, int32_t z) {
    return imul(seed ^ x ^ y ^ z, 0x27d4eb2d);
}
inline float xor_sign(float v, int32_t signbit) {
    uint32_t u; std::memcpy(&u, &v, 4); u ^= (uint32_t)signbit; std::memcpy(&v, &u, 4); return v;
}
// Utils.inl:229-247 (3-D, the non-AVX512 form; the AVX512 table holds the same 16 gradients)
inline float grad_dot(int32_t hash, float fx, float fy, float fz) {
    int32_t h13 = hash & 13;
    float u = h13 < 8 ? fx : fy;
    float v = h13 == 12 ? fx : fz;
    v = h13 < 2 ? fy : v;
    return xor_sign(u, ishl(hash, 31)) + xor_sign(v, ishl(hash & 2, 30));
}
inline float lerp(float a, float b, float t) { return t * (b - a) + a; }               // Utils.inl:324-328
inline float quintic(float t) { return t * t * t * (t * (t * 6.0f + -15.0f) + 10.0f); }  // Utils.inl:336-340
inline int32_t cvt_nearest(float x) { return (int32_t)std::nearbyintf(x); }             // _mm_cvtps_epi32

// Perlin.inl:34-64
float perlin3(int32_t seed, float x, float y, float z) {
    float xs = std::floor(x), ys = std::floor(y), zs = std::floor(z);
    int32_t x0 = imul(cvt_nearest(xs), PRIME_X), y0 = imul(cvt_nearest(ys), PRIME_Y), z0 = imul(cvt_nearest(zs), PRIME_Z);
    int32_t x1 = iadd(x0, PRIME_X), y1 = iadd(y0, PRIME_Y), z1 = iadd(z0, PRIME_Z);
    float xf0 = xs = x - xs, yf0 = ys = y - ys, zf0 = zs = z - zs;
    float xf1 = xf0 - 1.0f, yf1 = yf0 - 1.0f, zf1 = zf0 - 1.0f;
    xs = quintic(xs); ys = quintic(ys); zs = quintic(zs);
    return 0.964921414852142333984375f * lerp(
        lerp(lerp(grad_dot(hash_primes(seed, x0, y0, z0), xf0, yf0, zf0), grad_dot(hash_primes(seed, x1, y0, z0), xf1, yf0, zf0), xs),
             lerp(grad_dot(hash_primes(seed, x0, y1, z0), xf0, yf1, zf0), grad_dot(hash_primes(seed, x1, y1, z0), xf1, yf1, zf0), xs), ys),
        lerp(lerp(grad_dot(hash_primes(seed, x0, y0, z1), xf0, yf0, zf1), grad_dot(hash_primes(seed, x1, y0, z1), xf1, yf0, zf1), xs),
             lerp(grad_dot(hash_primes(seed, x0, y1, z1), xf0, yf1, zf1), grad_dot(hash_primes(seed, x1, y1, z1), xf1, yf1, zf1), xs), ys),
        zs);
}

// Simplex.inl:56-125
float simplex3(int32_t seed, float x, float y, float z) {
    const float F3 = 1.0f / 3.0f, G3 = 1.0f / 2.0f;
    float s = F3 * (x + y + z);
    x += s; y += s; z += s;
    float x0 = std::floor(x), y0 = std::floor(y), z0 = std::floor(z);
    float xi = x - x0, yi = y - y0, zi = z - z0;
    int32_t i = imul(cvt_nearest(x0), PRIME_X), j = imul(cvt_nearest(y0), PRIME_Y), k = imul(cvt_nearest(z0), PRIME_Z);
    bool x_ge_y = xi >= yi, y_ge_z = yi >= zi, x_ge_z = xi >= zi;
    float g = G3 * (xi + yi + zi);
    x0 = xi - g; y0 = yi - g; z0 = zi - g;
    bool i1 = x_ge_y && x_ge_z;
    bool j1 = y_ge_z && !x_ge_y;
    bool k1 = !x_ge_z && !y_ge_z;
    bool i2 = x_ge_y || x_ge_z;
    bool j2 = !x_ge_y || y_ge_z;
    bool k2 = x_ge_z && y_ge_z;  // used N-masked
    float x1 = (i1 ? x0 - 1.0f : x0) + G3, y1 = (j1 ? y0 - 1.0f : y0) + G3, z1 = (k1 ? z0 - 1.0f : z0) + G3;
    float x2 = (i2 ? x0 - 1.0f : x0) + G3 * 2, y2 = (j2 ? y0 - 1.0f : y0) + G3 * 2, z2 = (k2 ? z0 : z0 - 1.0f) + G3 * 2;
    float x3 = x0 + (G3 * 3 - 1), y3 = y0 + (G3 * 3 - 1), z3 = z0 + (G3 * 3 - 1);
    auto falloff = [](float a, float b, float c) {
        float t = -(a * a) + (-(b * b) + (-(c * c) + 0.6f));   // FNMulAdd chain, :104-107
        t = std::max(t, 0.0f);
        t *= t; t *= t;
        return t;
    };
    float t0 = falloff(x0, y0, z0), t1 = falloff(x1, y1, z1), t2 = falloff(x2, y2, z2), t3 = falloff(x3, y3, z3);
    float n0 = grad_dot(hash_primes(seed, i, j, k), x0, y0, z0);
    float n1 = grad_dot(hash_primes(seed, i1 ? iadd(i, PRIME_X) : i, j1 ? iadd(j, PRIME_Y) : j, k1 ? iadd(k, PRIME_Z) : k), x1, y1, z1);
    float n2 = grad_dot(hash_primes(seed, i2 ? iadd(i, PRIME_X) : i, j2 ? iadd(j, PRIME_Y) : j, k2 ? k : iadd(k, PRIME_Z)), x2, y2, z2);
    float n3 = grad_dot(hash_primes(seed, iadd(i, PRIME_X), iadd(j, PRIME_Y), iadd(k, PRIME_Z)), x3, y3, z3);
    return 32.69428253173828125f * (n0 * t0 + (n1 * t1 + (n2 * t2 + n3 * t3)));
}

// Cellular.inl:307-366 with the defaults of Cellular.h:16-17,88-90:
// jitter modifier 1, EuclideanSquared, return Index0 (no sqrt in GetReturn, :441-455).
float cellular_distance3(int32_t seed, float x, float y, float z) {
    const float jitter = 0.396144f * 1.0f;            // kJitter3D, Cellular.inl:14
    float distance[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
    int32_t xc = cvt_nearest(x) + -1, ycBase = cvt_nearest(y) + -1, zcBase = cvt_nearest(z) + -1;
    float xcf = (float)xc - x, ycfBase = (float)ycBase - y, zcfBase = (float)zcBase - z;
    xc = imul(xc, PRIME_X); ycBase = imul(ycBase, PRIME_Y); zcBase = imul(zcBase, PRIME_Z);
    for (int xi = 0; xi < 3; ++xi) {
        float ycf = ycfBase; int32_t yc = ycBase;
        for (int yi = 0; yi < 3; ++yi) {
            float zcf = zcfBase; int32_t zc = zcBase;
            for (int zi = 0; zi < 3; ++zi) {
                int32_t hash = hash_primes_hb(seed, xc, yc, zc);
                float xd = (float)(hash & 0x3ff) - (0x3ff / 2.0f);
                float yd = (float)((hash >> 10) & 0x3ff) - (0x3ff / 2.0f);
                float zd = (float)((hash >> 20) & 0x3ff) - (0x3ff / 2.0f);
                float invMag = jitter * (1.0f / std::sqrt(xd * xd + (yd * yd + zd * zd)));
                xd = xd * invMag + xcf; yd = yd * invMag + ycf; zd = zd * invMag + zcf;
                float nd = xd * xd; nd = yd * yd + nd; nd = zd * zd + nd;   // Utils.inl:356-362
                for (int i = 3; i > 0; --i) distance[i] = std::max(std::min(distance[i], nd), distance[i - 1]);
                distance[0] = std::min(distance[0], nd);
                zcf += 1.0f; zc = iadd(zc, PRIME_Z);
            }
            ycf += 1.0f; yc = iadd(yc, PRIME_Y);
        }
        xcf += 1.0f; xc = iadd(xc, PRIME_X);
    }
    return distance[0];
}

template <class F>
void parallel_rows(int n, int threads, F f) {
    int nt = threads > 0 ? threads : vo_hw_threads();
    nt = std::max(1, std::min(nt, n));
    if (nt == 1) { for (int i = 0; i < n; ++i) f(i); return; }
    std::atomic<int> next{0};
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t)
        pool.emplace_back([&] { for (int i; (i = next.fetch_add(1)) < n;) f(i); });
    for (auto& th : pool) th.join();
}

}  // namespace

extern "C" int vo_hw_threads(void) {
    unsigned n = std::thread::hardware_concurrency();
    return n ? (int)n : 1;
}

extern "C" float vo_noise(int gen, int seed, float x, float y, float z) {
    switch (gen) {
        case VO_NOISE_CELLULAR_DISTANCE: return cellular_distance3(seed, x, y, z);
        case VO_NOISE_PERLIN: return perlin3(seed, x, y, z);
        case VO_NOISE_SIMPLEX: return simplex3(seed, x, y, z);
    }
    return NAN;
}

// Fractal.inl:17-39 over Perlin, defaults Fractal.h:20-38 (weighted strength 0 ⇒ the
// per-octave amp factor Lerp(1, ., 0) is exactly 1).
extern "C" float vo_fbm_perlin(int seed, float x, float y, float z, int octaves, float gain, float lacunarity) {
    float g = std::fabs(gain), a = g, bound = 1.0f;     // CalculateFractalBounding, Fractal.h:27-38
    for (int i = 1; i < octaves; ++i) { bound += a; a *= g; }
    float amp = 1.0f / bound;
    float noise = perlin3(seed, x, y, z);
    float sum = noise * amp;
    for (int i = 1; i < octaves; ++i) {
        seed = iadd(seed, 1);                         // seed -= -1
        amp *= gain;
        x *= lacunarity; y *= lacunarity; z *= lacunarity;
        noise = perlin3(seed, x, y, z);
        sum += noise * amp;
    }
    return sum;
}

// Generator.inl:124-176: index i ↔ (x fastest, y, z); pos = float(start+idx)*frequency.
extern "C" void vo_noise_grid(int gen, float* out, int nx, int ny, int nz, float frequency, int seed,
                              float minmax[2], int threads) {
    std::vector<float> mins(nz, INFINITY), maxs(nz, -INFINITY);
    parallel_rows(nz, threads, [&](int z) {
        float mn = INFINITY, mx = -INFINITY;
        float zp = (float)z * frequency;
        for (int y = 0; y < ny; ++y) {
            float yp = (float)y * frequency;
            float* row = out + ((size_t)z * ny + y) * nx;
            for (int x = 0; x < nx; ++x) {
                float v = vo_noise(gen, seed, (float)x * frequency, yp, zp);
                row[x] = v; mn = std::min(mn, v); mx = std::max(mx, v);
            }
        }
        mins[z] = mn; maxs[z] = mx;
    });
    minmax[0] = *std::min_element(mins.begin(), mins.end());
    minmax[1] = *std::max_element(maxs.begin(), maxs.end());
}

// TestMain.cpp:64-92
extern "C" void vo_pack_rgba8(const float* c0, const float* c1, const float* c2, const float* c3,
                              const float* mm, int64_t n, uint8_t* out) {
    float inv1 = 1 / (mm[1] - mm[0]), inv2 = 1 / (mm[3] - mm[2]), inv3 = 1 / (mm[5] - mm[4]), inv4 = 1 / (mm[7] - mm[6]);
    for (int64_t i = 0; i < n; ++i) {
        float s1 = 1 - (c0[i] - mm[0]) * inv1;
        float s2 = 1 - (c1[i] - mm[2]) * inv2;
        float s3 = 1 - (c2[i] - mm[4]) * inv3;
        float s4 = 1 - (c3[i] - mm[6]) * inv4;
        s1 *= s1 * s1 * s1;                                   // :80
        out[i * 4 + 0] = (unsigned char)(s1 * 255.0f);         // :84-87 (truncation)
        out[i * 4 + 1] = (unsigned char)(s2 * 255.0f);
        out[i * 4 + 2] = (unsigned char)(s3 * 255.0f);
        out[i * 4 + 3] = (unsigned char)(s4 * 255.0f);
    }
}

// TestMain.cpp:43-92.  Quirk Q1: the second cellular call writes into buffer 1 again
// (:60), buffer 2 stays zero-initialised (:54), yet `sample2` is normalised with the
// min/max of call 2 and `sample1` (the call-2 field) with the min/max of call 1.
extern "C" void vo_bake_default(uint8_t* out, int size, int threads) {
    size_t n = (size_t)size * size * size;
    std::vector<float> b1(n), b2(n, 0.0f), b3(n), b4(n);
    float mm[8];
    vo_noise_grid(VO_NOISE_CELLULAR_DISTANCE, b1.data(), size, size, size, 0.01f, 1, mm + 0, threads);  // :59
    vo_noise_grid(VO_NOISE_CELLULAR_DISTANCE, b1.data(), size, size, size, 0.03f, 2, mm + 2, threads);  // :60
    vo_noise_grid(VO_NOISE_PERLIN, b3.data(), size, size, size, 0.19f, 3, mm + 4, threads);             // :61
    vo_noise_grid(VO_NOISE_SIMPLEX, b4.data(), size, size, size, 0.15f, 4, mm + 6, threads);            // :62
    vo_pack_rgba8(b1.data(), b2.data(), b3.data(), b4.data(), mm, (int64_t)n, out);
}
