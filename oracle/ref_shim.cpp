// ORACLE (test infrastructure).  Thin C entry points over the REFERENCE's own
// FastNoise2 (compiled from /root/reference/vendor/noise by oracle/Makefile into
// oracle/_ref/libvolrt_ref.so — no reference source is copied into this repo).
// Used to validate oracle_noise.cpp and the CUDA noise kernels, to generate
// tests/golden/, and as the "reference" CPU baseline of the volume bake.
//
// The calls mirror /root/reference/source/Etna/TestMain.cpp:43-62.
#include <FastNoise/FastNoise.h>
#include <cstdint>
#include <vector>

namespace {
FastSIMD::eLevel pick_level(int level) {
    // 0 = CPU max (what the app gets, TestMain.cpp:43-45), else a pinned level.
    switch (level) {
        case 1: return FastSIMD::Level_SSE2;
        case 2: return FastSIMD::Level_SSE41;
        case 3: return FastSIMD::Level_AVX2;
        case 4: return FastSIMD::Level_AVX512;
        default: return FastSIMD::Level_Null;
    }
}
FastNoise::SmartNode<FastNoise::Generator> make(int gen, int level) {
    FastSIMD::eLevel l = pick_level(level);
    switch (gen) {
        case 0: return FastNoise::New<FastNoise::CellularDistance>(l);
        case 1: return FastNoise::New<FastNoise::Perlin>(l);
        case 2: return FastNoise::New<FastNoise::Simplex>(l);
    }
    return {};
}
}  // namespace

extern "C" {

int vref_max_level(void) {
    switch (FastSIMD::CPUMaxSIMDLevel()) {
        case FastSIMD::Level_AVX512: return 4;
        case FastSIMD::Level_AVX2: return 3;
        case FastSIMD::Level_SSE41: case FastSIMD::Level_SSE42: return 2;
        default: return 1;
    }
}

// GenUniformGrid3D(out, 0,0,0, n,n,n, freq, seed) → minmax
int vref_grid(int gen, int level, float* out, int nx, int ny, int nz, float freq, int seed, float minmax[2]) {
    auto node = make(gen, level);
    if (!node) return 1;
    auto mm = node->GenUniformGrid3D(out, 0, 0, 0, nx, ny, nz, freq, seed);
    minmax[0] = mm.min; minmax[1] = mm.max;
    return 0;
}

// GenPositionArray3D at arbitrary points
int vref_points(int gen, int level, float* out, int n, const float* xs, const float* ys, const float* zs, int seed) {
    auto node = make(gen, level);
    if (!node) return 1;
    node->GenPositionArray3D(out, n, xs, ys, zs, 0.0f, 0.0f, 0.0f, seed);
    return 0;
}

// FractalFBm over Perlin at arbitrary points (Fractal.inl:11-40)
int vref_fbm_points(int level, float* out, int n, const float* xs, const float* ys, const float* zs, int seed,
                    int octaves, float gain, float lacunarity) {
    FastSIMD::eLevel l = pick_level(level);
    auto src = FastNoise::New<FastNoise::Perlin>(l);
    auto f = FastNoise::New<FastNoise::FractalFBm>(l);
    if (!src || !f) return 1;
    f->SetSource(src);
    f->SetOctaveCount(octaves);
    f->SetGain(gain);
    f->SetLacunarity(lacunarity);
    f->GenPositionArray3D(out, n, xs, ys, zs, 0.0f, 0.0f, 0.0f, seed);
    return 0;
}

// The whole start-up bake of the app (TestMain.cpp:43-92), single thread like the app.
int vref_bake_default(uint8_t* pixelData, int size, int level, float mm_out[8]) {
    auto fnCellular = make(0, level), fnPerlin = make(1, level), fnSimplex = make(2, level);
    if (!fnCellular || !fnPerlin || !fnSimplex) return 1;
    size_t n = (size_t)size * size * size;
    std::vector<float> n1(n), n2(n), n3(n), n4(n);
    auto o1 = fnCellular->GenUniformGrid3D(n1.data(), 0, 0, 0, size, size, size, 0.01f, 1);
    auto o2 = fnCellular->GenUniformGrid3D(n1.data(), 0, 0, 0, size, size, size, 0.03f, 2);
    auto o3 = fnPerlin->GenUniformGrid3D(n3.data(), 0, 0, 0, size, size, size, 0.19f, 3);
    auto o4 = fnSimplex->GenUniformGrid3D(n4.data(), 0, 0, 0, size, size, size, 0.15f, 4);
    float i1 = 1 / (o1.max - o1.min), i2 = 1 / (o2.max - o2.min), i3 = 1 / (o3.max - o3.min), i4 = 1 / (o4.max - o4.min);
    for (size_t i = 0; i < n; ++i) {
        float s1 = 1 - (n1[i] - o1.min) * i1, s2 = 1 - (n2[i] - o2.min) * i2;
        float s3 = 1 - (n3[i] - o3.min) * i3, s4 = 1 - (n4[i] - o4.min) * i4;
        s1 *= s1 * s1 * s1;
        pixelData[i * 4 + 0] = static_cast<unsigned char>(s1 * 255.0f);
        pixelData[i * 4 + 1] = static_cast<unsigned char>(s2 * 255.0f);
        pixelData[i * 4 + 2] = static_cast<unsigned char>(s3 * 255.0f);
        pixelData[i * 4 + 3] = static_cast<unsigned char>(s4 * 255.0f);
    }
    if (mm_out) { mm_out[0] = o1.min; mm_out[1] = o1.max; mm_out[2] = o2.min; mm_out[3] = o2.max;
                  mm_out[4] = o3.min; mm_out[5] = o3.max; mm_out[6] = o4.min; mm_out[7] = o4.max; }
    return 0;
}

}  // extern "C"
