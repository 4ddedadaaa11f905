// ORACLE (test infrastructure).  Thin C entry points over the REFERENCE's own
// FastNoise2 (compiled from /root/reference/vendor/noise by oracle/Makefile into
// oracle/_ref/libvolrt_ref.so — no reference source is copied into this repo).
// Used to validate oracle_noise.cpp and the CUDA noise kernels, to generate
// tests/golden/, and as the "reference" CPU baseline of the volume bake.
//
// The calls mirror /root/reference/source/Etna/TestMain.cpp:43-62.
#include <FastNoise/FastNoise.h>
#include <FastNoise/Metadata.h>
#include <cstdint>
#include <cstdio>
#include <string>
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

// ---- node trees: the reference's own decoder and generators on an encoded tree
// (FastNoise::NewFromEncodedNodeTree, FastNoise.h:49)
int vref_tree_grid(const char* encoded, int level, float* out, int x0, int y0, int z0, int nx, int ny, int nz, float freq, int seed, float minmax[2]) {
    auto node = FastNoise::NewFromEncodedNodeTree(encoded, pick_level(level));
    if (!node) return 1;
    auto mm = node->GenUniformGrid3D(out, x0, y0, z0, nx, ny, nz, freq, seed);
    minmax[0] = mm.min; minmax[1] = mm.max;
    return 0;
}
int vref_tree_points(const char* encoded, int level, float* out, int n, const float* xs, const float* ys, const float* zs, int seed) {
    auto node = FastNoise::NewFromEncodedNodeTree(encoded, pick_level(level));
    if (!node) return 1;
    node->GenPositionArray3D(out, n, xs, ys, zs, 0.0f, 0.0f, 0.0f, seed);
    return 0;
}
// every registered node class with its serialised members, as JSON lines: the wire format of an encoded tree is
// {u16 id, i32 per variable, child per node lookup, (u8 is_node, child | f32) per hybrid} (Metadata.cpp:199-300)
int vref_metadata_dump(char* buf, int cap) {
    std::string o;
    char tmp[256];
    for (const FastNoise::Metadata* m : FastNoise::Metadata::GetAll()) {
        std::snprintf(tmp, sizeof tmp, "{\"id\": %d, \"name\": \"%s\", \"variables\": [", (int)m->id, m->name);
        o += tmp;
        bool first = true;
        for (const auto& v : m->memberVariables) {
            const char* ty = v.type == FastNoise::Metadata::MemberVariable::EFloat ? "float" : v.type == FastNoise::Metadata::MemberVariable::EInt ? "int" : "enum";
            std::snprintf(tmp, sizeof tmp, "%s{\"name\": \"%s\", \"dim\": %d, \"type\": \"%s\", \"default_f\": %.9g, \"default_i\": %d, \"enum\": [",
                          first ? "" : ", ", v.name, v.dimensionIdx, ty, (double)v.valueDefault.f, v.valueDefault.i);
            o += tmp;
            for (size_t i = 0; i < v.enumNames.size(); ++i) { o += i ? ", \"" : "\""; o += v.enumNames[i]; o += "\""; }
            o += "]}";
            first = false;
        }
        o += "], \"nodes\": [";
        first = true;
        for (const auto& n : m->memberNodeLookups) { std::snprintf(tmp, sizeof tmp, "%s{\"name\": \"%s\", \"dim\": %d}", first ? "" : ", ", n.name, n.dimensionIdx); o += tmp; first = false; }
        o += "], \"hybrids\": [";
        first = true;
        for (const auto& h : m->memberHybrids) { std::snprintf(tmp, sizeof tmp, "%s{\"name\": \"%s\", \"dim\": %d, \"default\": %.9g}", first ? "" : ", ", h.name, h.dimensionIdx, (double)h.valueDefault); o += tmp; first = false; }
        o += "]}\n";
    }
    if ((int)o.size() + 1 > cap) return -(int)o.size() - 1;
    std::snprintf(buf, cap, "%s", o.c_str());
    return (int)o.size();
}

}  // extern "C"
