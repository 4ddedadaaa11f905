// libvolrt internals: what crosses from the host side of the C ABI (volrt_api.cu) to the
// kernels (march_kernels.cu, noise_kernels.cu).  Not installed; the public surface is
// include/volrt.h only.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/volrt.h"

namespace volrt {

// Per-view constants, decoded from the two UBOs at vr_set_params time.  Column-major mat4s.
// inv_pv / inv_model are formed in double and rounded once (the oracle does the same).
struct ViewConst {
    float inv_pv[16];       // inverse(Projection * View)           vert.glsl:19
    float inv_model[16];    // inverse(Model)
    float model[16];        // Model                                 vert.glsl:20
    float w2l[16];          // gsd.WorldToLocal                      frag.glsl:11
    float cam[3];           // gsd.CameraPosition                    frag.glsl:12
    float scroll_y[3];      // (MS[0].y, MS[1].y, MS[2].y)           frag.glsl:67
    float scroll_z[3];      // (MS[0].z, ...)                        frag.glsl:68
    float scroll_w[3];      // (MS[0].w, ...)                        frag.glsl:69
    // Pixels [x0,x1) x [y0,y1) that can see the cube: the bounding rectangle of its eight projected corners (+2 px), or
    // the whole frame when a corner is not in front of the camera.  Four fifths of the reference's frame lie outside it
    // and keep the clear colour without a ray set-up.
    int rect[4];            // x0, y0, x1, y1
};

struct BrickInfo {          // min-max grid for empty-space skipping (8^3 voxel bricks)
    const uint8_t* occupied;   // 1 byte per brick: 0 = density upper bound is exactly 0
    const uint8_t* dist;       // 1 byte per brick: 0 = occupied, else Chebyshev distance (bricks, <= 255) to the nearest occupied brick
    int bx, by, bz;            // bricks per axis
    float inv_brick[3];        // bricks per unit of texture coordinate
    // the distance map again with one brick of border on every side (the border repeats the nearest brick, which is what clamping
    // the index did), for the march's lookup: brick_distance_fast forms the index from the raw bit patterns of three float
    // additions, takes pad_bias off once and clamps the LINEAR index once instead of three coordinates twice
    const uint8_t* dist_pad;   // (bx+2) x (by+2) x (bz+2)
    unsigned int pad_bias, pad_last;
};

// procedural density: FastNoise2 FractalFBm over Perlin, remapped
struct TreeInstr;
struct FbmField {
    int octaves, seed;
    float frequency, gain, lacunarity, coverage, falloff0, falloff1, bounding, threshold;
    float inv_falloff;      // 1 / (falloff1 - falloff0): a multiply per sample instead of a divide
};

// Lattice-cell tables of the fBm field (noise_device.cuh: perlin3_cells).  Per octave, one uint2 per lattice CELL of the
// box: its eight corners as four pairs, each a 16-bit offset into the shared-memory pair table (x0y0 | x0y1 and x1y0 | x1y1 at z0
// in .x, at z1 in .y), i.e. FastNoise2's HashPrimes + the low nibble GetGradientDot looks at (Utils.inl:229-247,295-302), evaluated once per cell
// at vr_set_params instead of eight times per sample per octave.  Cells cover lattice coordinates -1 .. n-2 per axis.
constexpr int kFbmCellMaxOctaves = 8;
constexpr int kFbmCellBytes = 8;     // four 16-bit offsets into the shared-memory pair table (noise_device.cuh: PairLut)
struct FbmCellOctave {
    unsigned long long base_adj;   // device address of cell (-1,-1,-1)'s entry, minus kFbmCellBytes * 0x4B400000 (the index arrives as float bits)
    float nf, n2f, kf;             // n, n*n, and 1.5*2^23 + n*n + n + 1 (magic + the offset of lattice coordinate (0,0,0))
    float pad;
};
struct FbmCells {
    FbmCellOctave oct[kFbmCellMaxOctaves];
};

// blockIdx -> tile, with the divisions as multiplications by host-made reciprocals (q = umulhi(n, m), exact while n * d < 2^32;
// m = 0 stands for d = 1)
struct TileMap {
    int gx, gy;                 // tiles of the launch along x and y (grid x = gx * inner)
    int inner;                  // depth segments or views of a group folded into grid x (fastest in dispatch order); 1 = none
    unsigned int div_inner, div_gy;
    int swap;                   // gy < gx: the linear index runs down the columns
    int cy;                     // tile row nearest the frame's centre row
    int tx0, ty0;               // first tile of the launch (a launch restricted to the views' screen rectangle: batch_rect)
    int linear;                 // tiles in plain row-major order instead of centre-out (view batches over a volume larger than L2)
};
struct MarchParams {
    // target
    int width, height;
    int row_begin, row_end;
    int view_begin;             // blockIdx.z + view_begin = view index
    int n_views;                // views in this launch (set by launch_march)
    int schedule;               // -1 auto, 0 direct tiles, 1 persistent warps pulling tiles off `queue`
    unsigned int* queue;        // [0] next tile, [1] warps finished (self-resetting)
    int queue_chunk;            // tiles claimed per pull
    int ray_split;              // lanes per ray of the texture march: -1 auto, 1 none, 2/4/8
    int ray_split_composite;    // lanes per ray of the compositing march: -1 auto, 1 none, 2/4/8
    void* out;                  // view 0 / row 0 of the destination frame
    size_t row_pitch, view_stride;
    int out_format;             // vr_output_format
    const ViewConst* views;     // device array (multi-view batches); nullptr = the single view below
    ViewConst view0;            // single view, carried in the kernel parameters (no H2D copy per frame)
    // reference literals (frag.glsl:29-32,42,63,66-69)
    int max_steps;
    float step_scale, density, sample_scale;
    float coord_scale[4], scroll_weight[4];
    // volume
    cudaTextureObject_t tex;    // TEXTURE layout
    cudaTextureObject_t tex_plane[4];   // RGBA8 volume as four R8 textures (planar != 0)
    int planar;
    int const_mask;             // bit k: channel k of the RGBA8 volume has one value everywhere -> no fetch, const_val[k]
    float const_val[4];
    const void* lin;            // LINEAR layout (x fastest)
    int vw, vh, vd;
    int vol_format;             // vr_volume_format
    int address_mode;
    // extensions
    int light_steps;
    float light_step_len, sun_dir[3], sun_color[3], ambient, sigma_t, early_term_eps;
    int skip_empty;
    float unlit[3];             // sun_color + ambient: what a sample scatters when there is no light-march (filled by launch_march)
    BrickInfo bricks;
    FbmField fbm;
    FbmCells cells;             // MARCH_FBM_CELLS only
    // VR_DENSITY_TREE: the compiled program (device copy owned by the context), evaluated at P * tree_freq
    const TreeInstr* tree_prog;
    int tree_n, tree_result, tree_seed;
    float tree_freq;
    // device counters (VR_RENDER_STATS): covered, steps, early, light samples, skipped
    unsigned long long* counters;
    int sm_count;               // multiprocessors of the context's device (launch heuristics)
    // depth segments of the compositing march (march_kernels.cu: march_pixel): a ray's steps are cut into `segments`
    // contiguous pieces marched by different CTAs (grid x = tiles x segments); each writes its (colour, transmittance) to
    // seg_partials and the last one to arrive composes them front to back and stores the pixel.  segments <= 1: off.
    int segments;               // set by the caller from march_pick_segments (0 / 1 = off)
    int ray_segments;           // request: -1 auto, 1 off, 2..8
    float4* seg_partials;       // [view][segment][row - row_begin][x], context scratch
    unsigned int* seg_arrived;  // [view][row - row_begin][x], zero between launches (the composing thread resets it)
    size_t seg_capacity;        // pixel-segments the scratch holds
    // view batches: the same screen tile of `view_group` consecutive views is dispatched back to back (grid x = tiles x
    // view_group, grid z = groups), so that CTAs resident together look at the same region of a volume larger than L2
    int view_group;             // 0 / 1 = views are whole z slices
    // union of the screen rectangles (ViewConst::rect) of the views of this launch, from the caller (x1 <= x0: not known).  A
    // batch of many views launches tiles over this rectangle only and clears the rest of the frames with a copy-like kernel
    int batch_rect[4];
    int* extra_launches;        // HOST pointer or null: launch_march adds the kernels it launched besides the march itself
    // how a CTA finds its tile (march_kernels.cu: tile_of_cta), filled by the launch functions (set_tile_map): every warp of
    // every CTA runs this, and three quarters of a frame's CTAs are empty tiles that do nothing else
    TileMap map;
};
constexpr int kMaxRaySegments = 8;

enum MarchVariant {
    MARCH_RGBA_TEX = 0,      // reference integrator, 4 HW-filtered fetches/step
    MARCH_RGBA_LINEAR = 1,   // reference integrator, FP32 software trilinear from global memory
    MARCH_SINGLE_TEX = 2,    // one-channel grid through the texture unit
    MARCH_SINGLE_LINEAR = 3,
    MARCH_FBM = 4,           // procedural density
    MARCH_SINGLE_BRICK = 5,  // one-channel fp16 grid, boxes staged to shared memory by TMA
    MARCH_TREE = 6,          // procedural density from a compiled FastNoise2 node tree (noise_tree_eval.cuh)
    MARCH_FBM_CELLS = 7      // MARCH_FBM with the lattice hashes read from per-cell tables (MarchParams::cells)
};

constexpr size_t kGatherMailboxBytes = 256;    // behind the context's frames: counters of the device-side gather protocol (volrt_api.cu)

// launches one march over rows [row_begin,row_end) x views; returns cudaGetLastError()
cudaError_t launch_march(const MarchParams& p, int variant, int n_views, bool stats, cudaStream_t stream);
// where the march kernels' gradient table lands in the shared window (one word written to d_out), and where the fBm kernels
// expect it: the host checks the two agree before the first procedural render (an error code instead of a device trap)
cudaError_t march_probe_lut_address(unsigned int* d_out, cudaStream_t s);
unsigned int march_lut_fixed_address();
// how many depth segments (MarchParams::segments) this launch should use: 1 = none.  The caller sizes seg_partials / seg_arrived
// for rows x width x views x segments before it launches.
int march_pick_segments(const MarchParams& p, int variant, int n_views, bool stats);

// TMA-staged brick march (VR_LAYOUT_BRICK): tensor_map = 128-byte CUtensorMap over the x-fastest fp16 volume
cudaError_t launch_march_brick(const MarchParams& p, const void* tensor_map, int n_views, cudaStream_t stream);
// The box a CTA of the TMA-staged march pulls per slab.  Round 1's 32 x 24 x 24 halves (16-word rows, 384-word planes) puts every
// other row and every plane on the same banks: 70 % of the shared-memory wavefronts were conflicts.  Rows of 40 halves = 20 words
// (consecutive rows start 20 banks apart: 20 m mod 32 runs through 8 windows of 4 banks) with 21 rows per plane (window 5 lk + lj
// mod 8) cut the wavefronts 109 M -> 41 M and the conflicts 77 M -> 10.5 M per C4 frame — and the frame takes 1.55 ms instead of
// 1.50: the kernel is bound by its 0.62 G warp instructions (software trilinear, ~75 per step against one TEX), not by the
// shared-memory pipe.  Two boxes in flight (VOLRT_BRICK_STAGES=2, 3 CTAs per SM instead of 6): 1.81 ms.  Texture unit: 0.252 ms.
// profiles/r02_march_brick_c4.txt.  Default = the fastest measured.
#ifndef VOLRT_BRICK_X
#define VOLRT_BRICK_X 32
#define VOLRT_BRICK_Y 24
#define VOLRT_BRICK_Z 24
#endif
#ifndef VOLRT_BRICK_STAGES
#define VOLRT_BRICK_STAGES 1        // boxes in flight per CTA (each on its own mbarrier)
#endif
constexpr int kBrickBox[3] = {VOLRT_BRICK_X, VOLRT_BRICK_Y, VOLRT_BRICK_Z};

// analytic per-row cost (sum of chord steps) for band balancing; cost[height] on device
cudaError_t launch_row_cost(const MarchParams& p, int view, float* row_cost, cudaStream_t stream);

// ---- noise / bake (noise_kernels.cu) ----
struct NoiseField { uint32_t generator; float frequency; int32_t seed; };
// fills out[n^3] (x fastest) and atomically folds min/max into minmax_bits[2] (ordered-int floats)
cudaError_t launch_noise_fill(const NoiseField& f, int n, float* out, unsigned int* minmax_bits, cudaStream_t s);
cudaError_t launch_noise_minmax_only(const NoiseField& f, int n, unsigned int* minmax_bits, cudaStream_t s);
struct PackChannel { const float* field; float mn, mx; uint32_t power; };
cudaError_t launch_noise_pack(const PackChannel ch[4], int64_t n, uint8_t* out_rgba8, cudaStream_t s);
cudaError_t launch_smoke_fill(const FbmField& f, int n, __half* out, cudaStream_t s);
// fBm lattice-cell tables: entries[n^3] for one octave (seed already offset by the octave index), cells -1 .. n-2 per axis
cudaError_t launch_fbm_cell_table(int seed, int n, void* entries, cudaStream_t s);      // n^3 entries of kFbmCellBytes
cudaError_t launch_occupancy(const void* lin, int vol_format, int w, int h, int d, int address_mode,
                             uint8_t* occupied, int bx, int by, int bz, cudaStream_t s);
cudaError_t launch_deinterleave_rgba8(const uint8_t* rgba, uint8_t* planes4, size_t voxels, unsigned int* mm8, cudaStream_t s);   // mm8 may be null
cudaError_t launch_brick_distance(const uint8_t* occupied, uint8_t* dist, uint8_t* tmp, int bx, int by, int bz, cudaStream_t s);
// pad[(bz+2)(by+2)(bx+2)] = dist with a border that repeats the nearest brick
cudaError_t launch_brick_pad(const uint8_t* dist, uint8_t* pad, int bx, int by, int bz, cudaStream_t s);
// ---- peak_kernels.cu: measured roofline denominators
cudaError_t launch_fma_peak(float* sink, int blocks, int iters, cudaStream_t s);
cudaError_t launch_l2_sweep(const void* buf, size_t bytes, unsigned int* sink, int sms, cudaStream_t s);
cudaError_t launch_tex_peak(cudaTextureObject_t tex, float* sink, int blocks, int iters, cudaStream_t s);
cudaError_t launch_tex_stream(cudaTextureObject_t tex, int nx, int ny, int nz, float* sink, cudaStream_t s);
// ---- min-max pyramid (noise_kernels.cu; layout in include/volrt.h)
// level 0 from the texels: out[brick][channel][{min,max}]
cudaError_t launch_pyramid_base(const void* lin, int vol_format, int w, int h, int d, float* out, int bx, int by, int bz, cudaStream_t s);
// one reduction step: (bx,by,bz) entries -> ceil(/2) entries
cudaError_t launch_pyramid_reduce(const float* in, float* out, int channels, int bx, int by, int bz, cudaStream_t s);
// occupancy from level 0: occupied = some channel has min != 0 or max != 0 (NaN counts as occupied)
cudaError_t launch_occupancy_from_pyramid(const float* level0, int channels, uint8_t* occupied, int nbricks, cudaStream_t s);

// ---- FastNoise2 node trees (noise_tree.cu): an encoded tree is compiled on the host into a straight-line program
// over value registers V[] and position registers P[][3]; fractal octaves are unrolled, domain modifiers become
// position instructions, so the device needs no recursion and every thread runs the same instruction stream.
enum TreeOp : uint16_t {
    T_CONST = 0, T_WHITE, T_VALUE, T_PERLIN, T_SIMPLEX, T_OPENSIMPLEX2, T_SINE, T_POSITION_OUTPUT, T_DISTANCE_TO_POINT,
    T_CELL_VALUE, T_CELL_DISTANCE, T_CELL_LOOKUP_POS,
    T_ADD, T_SUB, T_MUL, T_DIV, T_MIN, T_MAX, T_MIN_SMOOTH, T_MAX_SMOOTH, T_FADE, T_POW_FLOAT, T_POW_INT, T_REMAP, T_TERRACE,
    T_CONVERT_RGBA8, T_COPY,
    T_FBM_FIRST, T_FBM_NEXT, T_RIDGED_FIRST, T_RIDGED_NEXT, T_PINGPONG_FIRST, T_PINGPONG_NEXT,
    T_POS_COPY, T_POS_SCALE, T_POS_AXIS_SCALE, T_POS_OFFSET, T_POS_ROTATE, T_POS_WARP, T_WARP_AMP_NEXT,
    T_FRACTAL_LEAF,
    T_OPENSIMPLEX2S, T_CHECKERBOARD,
    T_OP_COUNT
};
struct TreeInstr {
    uint16_t op;
    uint8_t dst;          // value register written (or position register for T_POS_* / T_CELL_LOOKUP_POS)
    uint8_t pos;          // position register read
    uint8_t a, b, c, d;   // value-register operands
    int32_t seed_off;     // added to the caller's seed
    int32_t i0, i1, i2, i3;
    float f[9];
};
constexpr int kTreeMaxInstr = 512, kTreeMaxValues = 64, kTreeMaxPositions = 16;
// out[nz][ny][nx] = tree((x0+ix)*freq, (y0+iy)*freq, (z0+iz)*freq) with seed; min/max folded into minmax_bits[2]
cudaError_t launch_tree_grid(const TreeInstr* d_prog, int n_instr, int result_reg, int x0, int y0, int z0, int nx, int ny, int nz,
                             float frequency, int seed, float* out, unsigned int* minmax_bits, cudaStream_t s);
}  // namespace volrt
struct vr_noise_tree {           // the opaque handle of include/volrt.h
    std::vector<volrt::TreeInstr> prog;
    int result_reg = 0, n_nodes = 0, n_values = 0, n_positions = 0;
    std::string encoded;
};
namespace volrt {

// ---- volume files (volume_file.cu; host code)
struct Sha256 {
    uint32_t h[8];
    uint64_t total;
    uint8_t buf[64];
    size_t fill;
    void reset();
    void update(const void* data, size_t n);
    void finish(uint8_t out[32]);
private:
    void block(const uint8_t* p);
};
constexpr uint32_t kVolumeFileHeaderBytes = 128;
struct VolumeFileHeader {
    uint32_t width, height, depth, format, address_mode, pyramid_levels;
    uint64_t texel_offset, texel_bytes, pyramid_offset, pyramid_bytes;
    uint8_t sha256[32];
    void encode(uint8_t out[kVolumeFileHeaderBytes]) const;
    const char* decode(const uint8_t in[kVolumeFileHeaderBytes], uint64_t file_size);   // nullptr = ok, else what is wrong
};
const char* read_volume_header(FILE* f, VolumeFileHeader* h);
size_t pyramid_floats(uint32_t w, uint32_t h, uint32_t d, uint32_t format, uint32_t* levels);

void minmax_init(unsigned int* bits2_host);
float minmax_decode(unsigned int bits);

}  // namespace volrt
