/*
 * volrt_oracle.h — CPU ORACLE for the volumetric ray-march hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (`volumetricrenderer_b200/`,
 * `include/volrt.h`, `host/`) may include, link or call this.  Only `tests/`,
 * `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
 * legs use it, and only as the checker / the timed CPU baseline.
 *
 * It restates, in plain scalar C++ (compiled -O2 -ffp-contract=off), the
 * arithmetic of the reference (paths relative to /root/reference):
 *   shaders/frag.glsl:18-81          ray set-up, slab test, step loop, Beer-Lambert
 *   shaders/vert.glsl:17-22          fragPosition / coverage (ray generation)
 *   source/Etna/TestMain.cpp:21-39   UBO structs (incl. the non-std140 C++ layout)
 *   source/Etna/TestMain.cpp:43-92   volume bake
 *   source/Etna/TestMain.cpp:222-245 default scene matrices
 *   source/Etna/Core/Vulkan/VulkanCore.cpp:676-710   sampler state (linear, mirrored repeat)
 *   source/Etna/Core/Vulkan/VulkanTexture.cpp:111-156 RGBA8 UNORM, tight x-fastest layout
 *   vendor/noise/include/FastNoise/Generators/{Cellular,Perlin,Simplex,Fractal,Utils,Generator}.inl
 *
 * PARITY STATUS: pinned to the reference's own sources on both halves of the path.
 *   Ray-march: /root/reference/shaders/frag.glsl and vert.glsl are compiled AS C++ (oracle/glsl_shim.h supplies
 *   vec/mat types, the built-ins and the fixed-function sampler; oracle/Makefile `ref` pipes the shader text from
 *   where it lies into g++) into oracle/_ref/libvolrt_ref.so.  tests/test_oracle_march.py holds oracle_march.cpp
 *   to that build bit for bit on binary32 outColor (full config-1 frame + rotated / ragged / scrolled scenes), and
 *   tests/golden/glsl_shaders_golden.npz (made by tests/golden/make_golden_glsl.py from the same build) does so
 *   where /root/reference is absent.  Not executed: the reference's Vulkan pipeline itself (no graphics stack on
 *   either box — unavailable); its fixed-function stages are restated in oracle/ref_raster.cpp from the pipeline
 *   state the reference sets, and glm (a system include of the reference, not vendored) from its closed forms.
 *   Noise / bake: the reference's FastNoise2 compiled from /root/reference/vendor/noise (same library file) and
 *   fixtures generated from it (tests/golden/, made by tests/golden/make_golden.py).
 *   The build-defined extensions (configs 2-5: light-march, fBm density, fp16 grids, skipping) have no reference
 *   counterpart; their parity is GPU vs this oracle's statement of the build's own definition.
 */
#ifndef VOLRT_ORACLE_H
#define VOLRT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* volume formats / density modes / output encodings (oracle's own numbering) */
enum { VO_VOL_RGBA8 = 0, VO_VOL_R16F = 1, VO_VOL_R8 = 2, VO_VOL_R32F = 3 };
enum { VO_DENSITY_RGBA_PRODUCT = 0, VO_DENSITY_SINGLE = 1, VO_DENSITY_FBM = 2 };
enum { VO_OUT_RGBA8_UNORM = 0, VO_OUT_RGBA8_SRGB = 1, VO_OUT_RGBA16F = 2 };
enum { VO_NOISE_CELLULAR_DISTANCE = 0, VO_NOISE_PERLIN = 1, VO_NOISE_SIMPLEX = 2 };

typedef struct vo_volume {
    const void* texels;   /* tightly packed, x fastest (VulkanCore.cpp:621-648) */
    int32_t w, h, d;
    int32_t format;       /* VO_VOL_* */
} vo_volume;

typedef struct vo_params {
    float ubo0[48];          /* Model, View, Projection — column-major (vert.glsl:4-9)       */
    float ubo1[36];          /* std140 image as frag.glsl:9-14 reads it: W2L@0 Cam@16 MS@20  */
    int32_t max_steps;       /* frag.glsl:30  (128)                                          */
    float step_scale;        /* frag.glsl:42  (4)                                            */
    float density;           /* frag.glsl:29  (1)                                            */
    float sample_scale;      /* frag.glsl:63  (0.2)                                          */
    float coord_scale[4];    /* frag.glsl:66-69 (1, .8, .75, .7)                             */
    float scroll_weight[4];  /* frag.glsl:66-69 (0, .2, .25, .3)                             */
    /* ---- build-defined extensions; all 0 = reference ---- */
    int32_t density_mode;    /* VO_DENSITY_*                                                 */
    int32_t light_steps;
    float light_step_len;    /* box units                                                    */
    float sun_dir[3];        /* box space, normalised by the caller                          */
    float sun_color[3];
    float ambient;
    float sigma_t;           /* > 0 selects compositing mode                                 */
    float early_term_eps;    /* 0 = off                                                      */
    /* procedural fBm density */
    int32_t fbm_octaves;
    int32_t fbm_seed;
    float fbm_frequency;     /* noise position = P_tex * fbm_frequency                       */
    float fbm_gain;
    float fbm_lacunarity;
    float fbm_coverage;      /* rho = max(0, .5 fbm + .5 - coverage) * edge falloff          */
    float fbm_falloff0, fbm_falloff1;
} vo_params;

/* ---- scene (TestMain.cpp:222-245, glm closed forms restated) ---- */
/* ubo0: 192 B as the host writes them; ubo1_host: the 140-byte C++ struct image
 * (MediaScroll at byte 76 — quirk Q2), frame_time goes to MediaScroll[0][0]. */
void vo_scene_default(float phi_deg, float theta_deg, float frame_time,
                      float* ubo0_48, float* ubo1_host_35);
/* decode the bytes a host hands over into what the shader sees through std140
 * (bytes >= size read as 0). */
void vo_decode_ubo1(const void* bytes, uint32_t size, float* ubo1_std140_36);
void vo_params_default(vo_params* p);   /* reference literals + default scene at phi=theta=0 */
void vo_mat4_inverse(const float* m16, float* out16);  /* general 4x4, computed in double */

/* ---- sampler T(u) ---- */
void vo_sample(const vo_volume* v, float u, float vv, float w, float out4[4]);

/* ---- the march.  out_rgba: W*(row1-row0) float4, linear, row-major from row0.
 * out_steps (nullable): per-pixel step count n (−1 for uncovered).
 * stats[0]=covered px, [1]=steps, [2]=rays terminated early, [3]=light samples.
 * threads: 0 = all cores (rows are distributed dynamically). */
int vo_render(const vo_volume* v, const vo_params* p, int W, int H, int row0, int row1,
              float* out_rgba, int32_t* out_steps, uint64_t stats[4], int threads);

/* same for an arbitrary list of screen rows (the bounded, strided sample bench.py times) */
int vo_render_rows(const vo_volume* v, const vo_params* p, int W, int H, const int32_t* rows, int n_rows,
                   float* out_rgba, int32_t* out_steps, uint64_t stats[4], int threads);

/* the inputs of the fragment stage for the pixels of `rows` (SURVEY.md A.2 = vert.glsl:17-22 + rasteriser):
 * covered[n_rows*W] and fragPosition xyz[n_rows*W*3] (world space; zero where not covered) */
int vo_frag_positions(const vo_params* p, int W, int H, const int32_t* rows, int n_rows, float* out_xyz, int32_t* covered);

/* ---- output conversion ---- */
void vo_encode(const float* rgba, int64_t n_px, int fmt, void* out);

/* ---- noise (FastNoise2 restated, scalar, unfused) ---- */
float vo_noise(int gen, int seed, float x, float y, float z);
float vo_fbm_perlin(int seed, float x, float y, float z, int octaves, float gain, float lacunarity);
/* Generator.inl:124-176 — out[i] for x fastest; pos = float(idx)*frequency; minmax[2] */
void vo_noise_grid(int gen, float* out, int nx, int ny, int nz, float frequency, int seed,
                   float minmax[2], int threads);
/* TestMain.cpp:64-92 normalise + pack (ch = float planes, mm = 4 × {min,max}) */
void vo_pack_rgba8(const float* ch0, const float* ch1, const float* ch2, const float* ch3,
                   const float* mm8, int64_t n, uint8_t* out);
/* TestMain.cpp:43-92 complete, incl. quirk Q1 (second cellular bake overwrites buffer 1) */
void vo_bake_default(uint8_t* out_rgba8, int size, int threads);
/* C2 density field sampled on an n^3 grid -> fp16, minus `threshold` clamped at 0 */
void vo_bake_smoke_r16f(uint16_t* out, int n, const vo_params* p, float threshold, int threads);

uint16_t vo_float_to_half(float f);
float vo_half_to_float(uint16_t h);
int vo_hw_threads(void);

#ifdef __cplusplus
}
#endif
#endif
