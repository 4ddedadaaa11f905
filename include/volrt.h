/*
 * volrt.h — C ABI of libvolrt: the B200-native (sm_100a CUDA) drop-in for the one
 * data-parallel hot path of Raspy-Py/VolumetricRenderer — the per-pixel volumetric
 * ray-march of shaders/frag.glsl and the noise-volume bake that feeds it.
 *
 * The reference has no plugin / FFI layer: its hot path is reached through Vulkan
 * object construction and three call sites in one main().  Every export below names
 * the reference call site it replaces (paths relative to the reference root).
 *
 * Conventions
 *   - extern "C", fixed-width types, plain pointers + sizes; no C++/torch types, no exceptions.
 *   - every call returns vr_status; the message of the last failure is kept per context
 *     (vr_last_error).  The reference's Error(...) macro throws and terminates
 *     (source/Etna/Core/Utils.h:22-29) — that is deliberately not carried over.
 *   - the library owns all device memory; the caller owns every pointer it passes and may free
 *     it on return; out-handles are written only on VR_OK.
 *   - no global state; a context is not internally locked (one thread at a time per context).
 *   - every struct starts with struct_size (= sizeof the struct the caller compiled against).
 *   - there is NO CPU fallback: without a CUDA device vr_create fails with VR_ERR_CUDA.
 */
#ifndef VOLRT_H
#define VOLRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define VR_API __declspec(dllexport)
#else
#define VR_API __attribute__((visibility("default")))
#endif

#define VR_VERSION_MAJOR 0
#define VR_VERSION_MINOR 2
#define VR_MAX_DEVICES 8

typedef struct vr_context vr_context;
typedef int32_t vr_status;

enum {
    VR_OK = 0,
    VR_ERR_INVALID_ARGUMENT = 1,
    VR_ERR_CUDA = 2,
    VR_ERR_OUT_OF_MEMORY = 3,
    VR_ERR_UNSUPPORTED = 4,
    VR_ERR_BAD_STATE = 5,
    VR_ERR_IO = 6                /* volume files: open/read/write failed, bad magic, hash mismatch */
};

/* colour-target encodings.  The reference renders into the swapchain format, which prefers
 * R8G8B8A8_SRGB and otherwise takes the platform's first format
 * (source/Etna/Core/Vulkan/VulkanSwapchain.cpp:181-191; target: VulkanRenderer.cpp:75-80). */
typedef enum vr_output_format {
    VR_OUT_RGBA8_UNORM = 0,
    VR_OUT_RGBA8_SRGB = 1,
    VR_OUT_RGBA16F = 2,
    VR_OUT_RGBA32F = 3          /* raw shader output, for parity work */
} vr_output_format;

/* volume texel formats.  Reference: VK_FORMAT_R8G8B8A8_UNORM (VulkanTexture.cpp:116). */
typedef enum vr_volume_format {
    VR_VOL_RGBA8_UNORM = 0,
    VR_VOL_R16F = 1,
    VR_VOL_R8_UNORM = 2
} vr_volume_format;

/* sampler addressing.  Reference: MIRRORED_REPEAT on U,V,W (VulkanCore.cpp:683-685). */
typedef enum vr_address_mode {
    VR_ADDRESS_MIRRORED_REPEAT = 0,
    VR_ADDRESS_CLAMP_TO_EDGE = 1
} vr_address_mode;

/* where the volume lives and who filters it */
typedef enum vr_layout {
    VR_LAYOUT_AUTO = 0,      /* TEXTURE */
    VR_LAYOUT_TEXTURE = 1,   /* cudaArray + texture unit: hardware trilinear (fixed-point weights) */
    VR_LAYOUT_LINEAR = 2,    /* x-fastest global memory, trilinear in FP32 in the kernel (bit-exact
                                against the CPU oracle) */
    VR_LAYOUT_BRICK = 3      /* R16F only (width % 8 == 0): the march stages 32x24x24-voxel boxes of the volume
                                in shared memory with TMA (cp.async.bulk.tensor.3d through a tiled tensor map
                                over the x-fastest volume) and filters in FP32 — bit-exact like LINEAR.
                                Serves the reference integrator; extensions run the LINEAR kernel */
} vr_layout;

typedef enum vr_density_mode {
    VR_DENSITY_RGBA_PRODUCT = 0,   /* frag.glsl:66-71: (s1*s2)*(s3+s4)*scale, four fetches  */
    VR_DENSITY_SINGLE = 1,         /* T(P).r * sample_scale, one fetch (smoke grids)        */
    VR_DENSITY_FBM = 2,            /* in-kernel FastNoise2-semantics fBm over Perlin        */
    VR_DENSITY_TREE = 3            /* a compiled FastNoise2 node tree, evaluated per sample */
} vr_density_mode;

typedef enum vr_memory_kind {
    VR_MEM_NONE = 0,     /* render into the context's own device frame only */
    VR_MEM_HOST = 1,
    VR_MEM_DEVICE = 2    /* device pointer on the context's device, or peer-mapped */
} vr_memory_kind;

enum {
    VR_RENDER_ASYNC = 1u << 0,     /* return after enqueueing; complete with vr_wait.  With VR_MEM_HOST the
                                      context keeps two frames in flight (the reference's
                                      MAX_FRAMES_IN_FLIGHT = 2, VulkanRenderer.cpp:13): the march of the
                                      next frame overlaps this frame's device-to-host copy          */
    VR_RENDER_DIRECT = 1u << 1,    /* VR_MEM_DEVICE only: the march kernel stores its pixels straight
                                      into dst (which may be another GPU's memory mapped through
                                      NVLink peer access / CUDA IPC) instead of the context frame */
    VR_RENDER_STATS = 1u << 2,     /* count covered rays / steps on the device (slower variant) */
    VR_RENDER_NO_MARCH = 1u << 3   /* launch nothing; only copy the context frame (e.g. the one the
                                      peers have just gathered into) to dst.  The view range may name any of
                                      the context's max_views frames; with VR_RENDER_ASYNC + VR_MEM_HOST the
                                      copy runs on the copy stream and completes at vr_wait                 */
};

/* vr_create_info.flags: how pixels are handed to warps.  Default (0) = one CTA per 16x8-pixel tile, tiles dispatched from
 * the frame's centre outwards (measured fastest for every workload on B200).  PERSISTENT = warps pulling 8x4 tiles off a
 * device-side queue (one vr_render in flight per context); kept for comparison. */
enum {
    VR_CREATE_SCHEDULE_DIRECT = 1u << 0,
    VR_CREATE_SCHEDULE_PERSISTENT = 1u << 1
};

/* replaces renderer.Init() + AddRenderPass("BasePass", ...) + render-target creation:
 * source/Etna/TestMain.cpp:121-122,156-166; Core/Vulkan/VulkanRenderer.cpp:17-118,232-256 */
typedef struct vr_create_info {
    uint32_t struct_size;
    int32_t device;              /* CUDA ordinal of this context's GPU (device_count <= 1)   */
    uint32_t width, height;      /* reference: 1280 x 720 (VulkanContext.cpp:24)             */
    uint32_t format;             /* vr_output_format                                         */
    uint32_t max_views;          /* frames held by the context (0 -> 1)                      */
    uint32_t flags;
    /* ---- several GPUs behind ONE context, in ONE process (read only when struct_size covers them; a caller compiled
     * against the 28-byte struct of version 0.1 keeps working).  The reference host is one process, one thread, one main()
     * (source/Etna/TestMain.cpp:41,173-256) and picks one physical device (Core/Vulkan/VulkanDevice.cpp:42-49); with
     * device_count = N > 1 the same five calls drive N GPUs of one box: the frame lives on devices[0], the volume is
     * replicated over NVLink at upload, and vr_render cuts the frame into cost-balanced row bands (or a view batch into view
     * shards), one per GPU, whose march kernels store their pixels straight into devices[0]'s frame through peer access;
     * the cut follows the measured band times from frame to frame.  Needs peer access between devices[0] and the others. */
    uint32_t device_count;       /* 0 / 1: the single `device` above                          */
    int32_t devices[VR_MAX_DEVICES];
} vr_create_info;

/* replaces vkc::Texture3D texture(pixelData.data(), VkExtent3D(size,size,size)):
 * TestMain.cpp:128; VulkanTexture.cpp:111-156; sampler VulkanCore.cpp:676-710 */
typedef struct vr_volume_desc {
    uint32_t struct_size;
    uint32_t width, height, depth;
    uint32_t format;             /* vr_volume_format  */
    uint32_t address_mode;       /* vr_address_mode   */
    uint32_t layout;             /* vr_layout         */
} vr_volume_desc;

/* one channel recipe of the bake, TestMain.cpp:59-62,75-80:
 *   s = 1 - (field - min) / (max - min);  s <- s^power;  byte = (uint8)(s * 255)   (truncation)
 * where {min,max} are those of the `norm_*` field — which in the reference is NOT always the
 * stored field: the second cellular call overwrites buffer 1 and buffer 2 stays zero (quirk Q1). */
#define VR_NOISE_CELLULAR_DISTANCE 0u
#define VR_NOISE_PERLIN 1u
#define VR_NOISE_SIMPLEX 2u
#define VR_NOISE_ZERO 3u
#define VR_NOISE_TREE 4u           /* the compiled FastNoise2 node tree in vr_noise_desc.tree[k]; the channel's frequency and seed apply */
#define VR_NOISE_SAME 0xffffffffu
typedef struct vr_noise_channel {
    uint32_t generator;          /* VR_NOISE_*: the field whose values are packed             */
    float frequency;
    int32_t seed;
    uint32_t norm_generator;     /* field whose min/max normalises; VR_NOISE_SAME = the above */
    float norm_frequency;
    int32_t norm_seed;
    uint32_t power;              /* 1 or 4 in the reference                                   */
} vr_noise_channel;

typedef struct vr_noise_tree vr_noise_tree;     /* "FastNoise2 node trees" below */
/* replaces the start-up bake, TestMain.cpp:43-92 (FastNoise2 GenUniformGrid3D x4 + pack) */
typedef struct vr_noise_desc {
    uint32_t struct_size;
    uint32_t size;               /* cube edge; reference 128                                  */
    uint32_t format;             /* VR_VOL_RGBA8_UNORM: the 4 channel recipes below;
                                    VR_VOL_R16F: the fBm smoke field rho(P) of vr_params.fbm_*
                                    at voxel centres, minus `threshold`, clamped at 0          */
    uint32_t layout;             /* vr_layout                                                 */
    uint32_t address_mode;       /* vr_address_mode                                           */
    float threshold;
    vr_noise_channel channel[4];
    const vr_noise_tree* tree[4]; /* per channel, for generator VR_NOISE_TREE; read only when struct_size covers it */
} vr_noise_desc;

/* replaces objectUniformBuffer.Update(&osd, frame) / globalUniformBuffer.Update(&gsd, frame):
 * TestMain.cpp:248-249; Core/Vulkan/VulkanUniformBuffer.h:58-61.  The UBO bytes are copied at
 * call time and decoded with the std140 offsets the SHADER uses (frag.glsl:9-14, vert.glsl:4-9);
 * bytes past *_size read as zero.  Handing over the reference's C++ structs verbatim (192 B and
 * 140 B, TestMain.cpp:27-39) therefore reproduces its MediaScroll misalignment exactly. */
typedef struct vr_params {
    uint32_t struct_size;
    uint32_t ubo0_size;          /* 192                                                        */
    const void* ubo0;            /* ObjectShaderData: Model, View, Projection                  */
    const void* ubo1;            /* GlobalShaderData: WorldToLocal, CameraPosition, MediaScroll */
    uint32_t ubo1_size;          /* 140 (C++ sizeof) or 144 (std140)                           */
    uint32_t max_steps;          /* frag.glsl:30   128                                         */
    float step_scale;            /* frag.glsl:42   4                                           */
    float density;               /* frag.glsl:29   1                                           */
    float sample_scale;          /* frag.glsl:63   0.2                                         */
    float coord_scale[4];        /* frag.glsl:66-69  1, .8, .75, .7                            */
    float scroll_weight[4];      /* frag.glsl:66-69  0, .2, .25, .3                            */
    /* ---- extensions beyond the reference; all zero = frag.glsl ---- */
    uint32_t density_mode;       /* vr_density_mode                                            */
    uint32_t light_steps;        /* sun light-march samples per primary step                   */
    float light_step_len;        /* box units                                                  */
    float sun_dir[3];            /* box space, unit length                                     */
    float sun_color[3];
    float ambient;
    float sigma_t;               /* > 0: front-to-back compositing instead of optical depth    */
    float early_term_eps;        /* > 0: stop a ray once transmittance < eps                   */
    uint32_t skip_empty;         /* use the min-max brick grid to jump over empty space        */
    int32_t fbm_octaves;         /* VR_DENSITY_FBM: FastNoise2 FractalFBm over Perlin          */
    int32_t fbm_seed;
    float fbm_frequency;
    float fbm_gain;
    float fbm_lacunarity;
    float fbm_coverage;
    float fbm_falloff0, fbm_falloff1;
    /* multi-view batch: view k uses view_ubos + k*336 bytes = {ubo0[192], ubo1 std140[144]};
     * view_count 0/1 with view_ubos NULL = the single view above */
    uint32_t view_count;
    const void* view_ubos;
    /* VR_DENSITY_TREE: rho(P) = max(0, tree(P * tree_frequency; tree_seed)) * edge(P), P in texture
     * coordinates, edge = the fbm_falloff0/1 fade above.  The program is copied at vr_set_params; the
     * handle may be freed afterwards.  A caller built against the struct without these three fields
     * (struct_size = offset of density_tree) is accepted for every other mode. */
    const vr_noise_tree* density_tree;
    float tree_frequency;
    int32_t tree_seed;
} vr_params;

/* replaces EnqueueRenderPass + BeginFrame/EndFrame -> vkCmdDrawIndexed + submit:
 * TestMain.cpp:194-217,251,255; VulkanRenderer.cpp:174-230 */
typedef struct vr_target {
    uint32_t struct_size;
    uint32_t memory_kind;        /* vr_memory_kind                                             */
    void* dst;                   /* view-major, row-major from row 0 of the FULL frame         */
    size_t row_pitch;            /* bytes; 0 = tightly packed                                  */
    uint32_t row_begin, row_end; /* band of rows to render (0,0 = all rows)                    */
    uint32_t view_begin, view_end; /* views to render (0,0 = all)                              */
    uint32_t flags;              /* VR_RENDER_*                                                */
    void* stream;                /* cudaStream_t; NULL = the context's own stream              */
} vr_target;

typedef struct vr_stats {
    uint32_t struct_size;
    float gpu_ms;                /* device time (CUDA events) of the last vr_render's kernels, or of the last upload / bake */
    uint64_t rays;               /* pixels shaded or cleared                                   */
    uint64_t covered;            /* VR_RENDER_STATS only from here on                          */
    uint64_t steps;
    uint64_t early_terminated;
    uint64_t light_samples;
    uint64_t skipped_steps;
    uint32_t kernel_launches;    /* kernels launched by the last vr_render                     */
    uint32_t constant_channels;  /* bit k: channel k of the resident RGBA8 volume holds one value everywhere; the
                                    hardware-filtered march does not fetch such a channel (the reference's default bake
                                    leaves G constant, TestMain.cpp:59-60,76)                    */
} vr_stats;

VR_API uint32_t vr_version(void);
VR_API vr_status vr_params_default(vr_params* p);           /* reference literals, extensions off */
VR_API vr_status vr_create(const vr_create_info* info, vr_context** out);
VR_API void vr_destroy(vr_context* ctx);                    /* replaces renderer.Shutdown(), TestMain.cpp:260 */
VR_API const char* vr_last_error(const vr_context* ctx);    /* ctx may be NULL: last vr_create failure */
VR_API vr_status vr_upload_volume(vr_context* ctx, const vr_volume_desc* desc, const void* texels);
VR_API vr_status vr_generate_volume(vr_context* ctx, const vr_noise_desc* desc, const vr_params* fbm);
VR_API vr_status vr_download_volume(vr_context* ctx, void* texels, size_t bytes);  /* x-fastest copy back */
VR_API vr_status vr_noise_desc_default(vr_noise_desc* d);   /* the reference's bake incl. quirk Q1 */
VR_API vr_status vr_set_params(vr_context* ctx, const vr_params* p);
VR_API vr_status vr_render(vr_context* ctx, const vr_target* target);
VR_API vr_status vr_wait(vr_context* ctx);
VR_API vr_status vr_get_stats(vr_context* ctx, vr_stats* out);
/* the context's device frame (view 0, row 0) and its layout, for zero-copy consumers */
VR_API vr_status vr_frame_info(vr_context* ctx, void** dev_ptr, size_t* row_pitch, size_t* view_stride);
/* row bands of equal estimated cost for `parts` GPUs: bounds[parts+1] */
VR_API vr_status vr_row_bands(vr_context* ctx, uint32_t parts, uint32_t* bounds);
/* the device half of vr_row_bands: cost[height] = per-row sum of step counts from the ray set-up alone */
VR_API vr_status vr_row_cost(vr_context* ctx, float* cost);
/* the host half of vr_row_bands (no device needed): cut `height` rows with per-row cost[] into `parts`
 * contiguous bands of nearly equal cost; bounds[0] = 0 ... bounds[parts] = height */
VR_API vr_status vr_split_rows(const float* cost, uint32_t height, uint32_t parts, uint32_t* bounds);
/* NVLink peer gather across processes: rank 0 exports its frame, the others map it and render
 * their band with VR_RENDER_DIRECT into the mapped pointer. handle = 64 bytes (cudaIpcMemHandle_t) */
VR_API vr_status vr_frame_export(vr_context* ctx, void* handle64);
VR_API vr_status vr_frame_import(vr_context* ctx, const void* handle64, void** mapped);
VR_API vr_status vr_frame_unmap(vr_context* ctx, void* mapped);
/* Completion of that gather on the device, without a host barrier per frame.  GPU 0's context holds `frames`
 * (= its max_views) frame slots followed by a mailbox of counters; frame0 = NULL on GPU 0 itself, the mapped pointer on the
 * others (same width, height and format everywhere).  Per frame k into slot s, every rank (GPU 0 included) enqueues
 *   vr_gather_acquire(ctx, frame0, frames, s, r)  r = how many earlier frames used slot s: the render stream waits until
 *                                                 GPU 0 has copied that many out of it;
 *   vr_render(VR_RENDER_DIRECT | VR_RENDER_ASYNC) into frame0 + s * view_stride;
 *   vr_gather_signal(ctx, frame0, frames, s)      "my band of this frame has landed" (system-scope release);
 * and GPU 0 alone
 *   vr_gather_collect(ctx, s, ranks * (r + 1), dst_host, row_pitch)   its copy stream waits for all the bands, copies slot s
 *                                                 to page-locked host memory and releases the slot.
 * Nothing blocks the host; vr_wait on GPU 0 completes the copies.  A wait that is not satisfied within ~10 s traps. */
VR_API vr_status vr_gather_acquire(vr_context* ctx, void* frame0, uint32_t frames, uint32_t slot, uint32_t releases);
VR_API vr_status vr_gather_signal(vr_context* ctx, void* frame0, uint32_t frames, uint32_t slot);
VR_API vr_status vr_gather_collect(vr_context* ctx, uint32_t slot, uint32_t arrivals, void* dst_host, size_t row_pitch);
/* measurement helpers: evict the 126 MB L2 between timed frames (writes a 256 MiB scratch buffer on
 * the context's stream and reads it back, so that L2 is left without dirty lines); page-locked host memory so that vr_render's D2H copy and
 * vr_upload_volume's H2D copy run at full PCIe rate (the reference maps its UBOs and stages
 * its texture through host-visible Vulkan memory: VulkanUniformBuffer.h:30-45, VulkanTexture.cpp:121-128) */
VR_API vr_status vr_flush_l2(vr_context* ctx);
/* roofline denominators measured on this device: peaks4 = { FP32-FMA TFLOP/s, trilinear RGBA8 texture
 * fetches G/s (0 without a TEXTURE-layout RGBA8 volume), SM count, max SM clock MHz } */
VR_API vr_status vr_measure_peaks(vr_context* ctx, float* peaks4);
/* GB/s at which the texture unit streams the resident one-channel TEXTURE-layout volume out of DRAM when every texel is fetched
 * exactly once (trilinear fetches on a 2-texel lattice, L2 flushed): the ceiling of a grid march whose fetches are all first
 * touches — what the "fraction of HBM bandwidth" of the 512^3 smoke workload is bounded by.  No counterpart in the reference. */
VR_API vr_status vr_measure_tex_stream(vr_context* ctx, float* gbs);
VR_API void* vr_host_alloc(size_t bytes);
VR_API void vr_host_free(void* p);

/* ---- volume files.  The reference keeps its volume in memory only (TestMain.cpp:51-92) and its
 * one loader is stb_image for 2-D pictures (VulkanTexture.cpp:14-55); this is the step in front of vr_upload_volume
 * for volumes that are shared between runs (a 1024^3 fp16 grid is 2 GiB and 29 ms of bake, or minutes on the CPU).
 * Layout, little-endian:   [0,128)  header          magic "VOLRTVOL", version 1, dims, format, address mode, section table
 *                          texels   width*height*depth*texel_bytes, x fastest, tightly packed (what vr_upload_volume takes)
 *                          pyramid  optional; min-max mip pyramid over 8^3-texel bricks (below)
 * The header carries the SHA-256 of the texel section.
 * Min-max pyramid: level 0 has one entry per 8^3 brick (ceil(dim/8) per axis), entry = channels x {min, max} as float32
 * of the decoded texel values (UNORM8 -> v/255, fp16 -> float), taken over the brick plus a one-texel apron (clamped at
 * the volume faces) — everything a trilinear fetch inside the brick can touch.  Level l+1 halves each axis (ceil) and
 * reduces the up to 2x2x2 children.  Levels follow each other, x fastest, until a level is 1x1x1.  A NaN texel makes its
 * bricks' max NaN on every level.  channels = 4 for RGBA8, else 1. */
typedef struct vr_volume_file_info {
    uint32_t struct_size;
    uint32_t width, height, depth;
    uint32_t format;             /* vr_volume_format  */
    uint32_t address_mode;       /* vr_address_mode   */
    uint32_t pyramid_levels;     /* 0 = the file has no pyramid section */
    uint32_t reserved;
    uint64_t texel_bytes;
    uint64_t pyramid_bytes;
    uint8_t  sha256[32];         /* of the texel section */
} vr_volume_file_info;
#define VR_FILE_VERIFY        1u     /* reading: recompute the SHA-256 of the texels and compare with the header */
#define VR_FILE_WITH_PYRAMID  2u     /* vr_save_volume_file: append the min-max pyramid (computed on the device) */
#define VR_FILE_TRUST_PYRAMID 4u     /* vr_upload_volume_file: take the brick occupancy from the file's pyramid level 0
                                        instead of scanning the texels again */
/* host only (no device needed) */
VR_API vr_status vr_sha256(const void* data, size_t bytes, uint8_t out32[32]);
VR_API vr_status vr_volume_file_write(const char* path, const vr_volume_desc* desc, const void* texels,
                                      const float* pyramid, uint32_t pyramid_levels);   /* pyramid may be NULL */
VR_API vr_status vr_volume_file_info_read(const char* path, vr_volume_file_info* info);
VR_API vr_status vr_volume_file_read(const char* path, void* texels, size_t texel_bytes, float* pyramid, size_t pyramid_bytes,
                                     uint32_t flags);                                    /* either destination may be NULL */
VR_API size_t vr_pyramid_floats(uint32_t width, uint32_t height, uint32_t depth, uint32_t format, uint32_t* levels);
/* device side: file -> pinned staging (two buffers) -> H2D, overlapped; then as vr_upload_volume */
VR_API vr_status vr_upload_volume_file(vr_context* ctx, const char* path, uint32_t layout, uint32_t flags);
VR_API vr_status vr_save_volume_file(vr_context* ctx, const char* path, uint32_t flags);
/* min-max pyramid of the resident volume, computed on the device; out holds vr_pyramid_floats() floats */
VR_API vr_status vr_volume_pyramid(vr_context* ctx, float* out, size_t out_bytes);

/* ---- FastNoise2 node trees.  The reference builds its generators in code
 * (FastNoise::New<...>, TestMain.cpp:43-45); FastNoise2's portable form of a generator graph is the encoded node tree
 * string (FastNoise::NewFromEncodedNodeTree, vendor/noise/include/FastNoise/FastNoise.h:49, the strings its NoiseTool
 * exports).  vr_noise_tree_compile decodes such a string on the host (no device needed) into a straight-line
 * program; vr_noise_tree_grid evaluates it on the device with GenUniformGrid3D semantics (Generator.inl:124-176):
 * out[z][y][x] = tree((x0+x)*frequency, (y0+y)*frequency, (z0+z)*frequency; seed), min/max returned like
 * FastNoise::OutputMinMax.  Built (40 of the reference's 42 node classes): Constant, White, Checkerboard, SineWave,
 * PositionOutput, DistanceToPoint, Value, Perlin, Simplex, OpenSimplex2, OpenSimplex2S, CellularValue/Distance/Lookup,
 * FractalFBm/Ridged/PingPong, DomainWarpGradient, DomainWarpFractalProgressive/Independant,
 * DomainScale/Offset/Rotate/AxisScale, SeedOffset, Remap, ConvertRGBA8, Terrace,
 * Add/Subtract/Multiply/Divide/Min/Max/MinSmooth/MaxSmooth/Fade/PowFloat/PowInt, GeneratorCache; 3-D positions only.
 * AddDimension / RemoveDimension (they change the dimension count) -> VR_ERR_UNSUPPORTED with the node's name in err.
 * Checkerboard divides by the size exactly; the reference multiplies by an approximate reciprocal, so a point exactly
 * on a cell border may fall on the other side there. */
VR_API vr_status vr_noise_tree_compile(const char* encoded_node_tree, vr_noise_tree** out, char* err, size_t err_cap);
VR_API void vr_noise_tree_free(vr_noise_tree* tree);
VR_API vr_status vr_noise_tree_info(const vr_noise_tree* tree, uint32_t* nodes, uint32_t* instructions, uint32_t* value_regs,
                                    uint32_t* position_regs);
/* out_host may be NULL (min/max only); it holds nx*ny*nz floats, x fastest */
VR_API vr_status vr_noise_tree_grid(vr_context* ctx, const vr_noise_tree* tree, int32_t x0, int32_t y0, int32_t z0, uint32_t nx,
                                    uint32_t ny, uint32_t nz, float frequency, int32_t seed, float* out_host, float minmax[2]);

/* ---- allocation-free variant of vr_noise_tree_grid.  No counterpart in the reference; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same computation, same arguments and results as vr_noise_tree_grid, but it needs no context and calls no cudaMalloc /
 * cudaFree / cudaMallocHost or any allocating library routine: the program upload buffer, the two min/max words and the output
 * plane live in one file-scope device array of libvolrt.so (csrc/noalloc.cu, ~1.03 MiB, held by every GPU that loads the
 * library until the process exits).  Limit: nx*ny*nz <= VR_STATIC_GRID_MAX_POINTS (64^3); larger grids are refused with
 * VR_ERR_INVALID_ARGUMENT and a message naming the limit.  `device` = CUDA ordinal; `stream` = cudaStream_t, NULL = the default
 * stream; the call returns after the results have reached out_host (may be NULL) and minmax.  All calls on a device share the
 * array and it keeps the last call's contents (each call re-initialises what must start clean); calls are serialised by a
 * mutex inside the library and complete before they return — work that overlapped on the GPU would race on the array.
 * err (may be NULL) receives the message of a failure, as in vr_noise_tree_compile. */
#define VR_STATIC_GRID_MAX_POINTS 262144u
VR_API vr_status vr_noise_tree_grid_static(int32_t device, const vr_noise_tree* tree, int32_t x0, int32_t y0, int32_t z0, uint32_t nx,
                                           uint32_t ny, uint32_t nz, float frequency, int32_t seed, float* out_host, float minmax[2],
                                           void* stream, char* err, size_t err_cap);

#ifdef __cplusplus
}
#endif
#endif
