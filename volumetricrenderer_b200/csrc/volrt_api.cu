// libvolrt: the C ABI of include/volrt.h over the sm_100a kernels.  Host-side only logic here:
// context + resource ownership, UBO decoding (std140), volume upload into a cudaArray / linear
// buffer, launch configuration, row-band cost model, CUDA-IPC export of the frame for the
// NVLink peer gather.  There is no CPU rendering path in this library.
#include "volrt_internal.h"
#include <cuda.h>              // CUtensorMap + enums (types only: cuTensorMapEncodeTiled is fetched at run time)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstddef>
#include <cstring>
#include <condition_variable>
#include <mutex>
#include <new>
#include <thread>
#include <unistd.h>
#include <string>
#include <vector>

using namespace volrt;

struct vr_context {
    int device = 0;
    uint32_t width = 0, height = 0, format = 0, max_views = 1;
    size_t px_bytes = 4, row_pitch = 0, view_stride = 0;
    void* frame = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // frames in flight (VR_RENDER_ASYNC into host memory): the march of frame k+1 overlaps the D2H copy of
    // frame k, like the reference's MAX_FRAMES_IN_FLIGHT = 2 (VulkanRenderer.cpp:13).  frame2 = second
    // device frame, allocated by the first asynchronous host render.
    cudaStream_t copy_stream = nullptr;
    void* frame2 = nullptr;
    cudaEvent_t ev_render_done[2] = {nullptr, nullptr}, ev_copy_done[2] = {nullptr, nullptr};
    bool copy_pending[2] = {false, false};
    int slot = 0;
    cudaEvent_t ev_params = nullptr;   // the H2D copy of the pinned view block has completed
    bool params_pending = false;
    bool views_on_device = false;
    bool foreign_stream_pending = false;   // last render ran on a caller stream (ev1 marks its end)
    ViewConst view0;
    bool timed = false;
    // volume
    bool has_volume = false;
    vr_volume_desc vdesc{};
    void* lin = nullptr;            // x-fastest copy (all layouts: source of the array, occupancy, download)
    size_t lin_bytes = 0;
    cudaArray_t array = nullptr;
    cudaTextureObject_t tex = 0;
    // RGBA8 volumes are also kept as four one-channel textures: frag.glsl reads ONE channel from each of its four fetches
    cudaArray_t plane_array[4] = {};
    cudaTextureObject_t plane_tex[4] = {};
    unsigned char* plane_lin = nullptr;    // de-interleave scratch (4 planes)
    bool planar = false;
    uint32_t const_mask = 0;           // channels of the RGBA8 volume that hold a single value (bit k), and that value as the sampler returns it
    float const_val[4] = {0, 0, 0, 0};
    unsigned int* d_chan_mm = nullptr; // 8 words: per-channel min / max of the last RGBA8 volume
    uint8_t* occupied = nullptr;
    uint8_t* brick_dist = nullptr;         // 2 planes: the distance map and the scratch of its passes
    alignas(64) unsigned char tensor_map[128] = {};   // BRICK layout: CUtensorMap over `lin`
    bool has_tensor_map = false;
    BrickInfo bricks{};
    // bake scratch, kept between vr_generate_volume calls of the same size
    float* bake_plane[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t bake_plane_bytes = 0;
    unsigned int* d_mm = nullptr;
    // vr_upload_volume_file: ring of page-locked staging buffers + their copy-done events, kept for the next file
    static constexpr int kFileRing = 6;
    static constexpr size_t kFileChunk = (size_t)16 << 20;
    unsigned char* file_stage[kFileRing] = {};
    cudaEvent_t file_done[kFileRing] = {};
    TreeInstr* d_march_tree = nullptr;      // VR_DENSITY_TREE: the program the march interprets (copied at vr_set_params)
    int march_tree_n = 0, march_tree_result = 0;
    TreeInstr* d_tree_prog = nullptr;       // vr_noise_tree_grid: program upload buffer and a grow-only output plane
    float* d_tree_out = nullptr;
    size_t tree_out_bytes = 0;
    // fBm lattice-cell tables (volrt_internal.h: FbmCells), rebuilt by vr_set_params when the field's lattice changes
    unsigned char* d_fbm_cells = nullptr;  // the tables of every octave, back to back (kFbmCellBytes per cell)
    size_t fbm_cells_bytes = 0;
    FbmCells fbm_cells{};
    bool fbm_cells_ok = false;
    int fbm_cells_key_oct = 0, fbm_cells_key_seed = 0;
    float fbm_cells_key_freq = 0.0f, fbm_cells_key_lac = 0.0f;
    const float* trusted_level0 = nullptr; // device copy of a file's pyramid level 0 (VR_FILE_TRUST_PYRAMID), for one finish_volume
    uint32_t fill_launches = 0;            // kernels the current upload / bake has launched so far
    // params
    bool has_params = false;
    vr_params params{};
    ViewConst* d_views = nullptr;
    ViewConst* h_views = nullptr;   // pinned staging
    uint32_t view_count = 1, view_capacity = 0;
    unsigned long long* d_counters = nullptr;
    unsigned long long* h_counters = nullptr;   // pinned
    float* d_row_cost = nullptr;
    unsigned int* d_queue = nullptr;   // tile queue of the persistent-warp schedule (self-resetting)
    int schedule = -1;                 // -1 auto, 0 direct, 1 persistent
    int queue_chunk = 1;
    int ray_split = -1;                // lanes per ray (texture march): -1 auto
    int ray_split_composite = -1;      // lanes per ray (compositing march): -1 auto
    int lut_addr_ok = -1;              // -1 not probed yet; fBm kernels fetch their gradient table at an immediate shared address
    int ray_segments = -1;             // depth segments per ray (compositing march): -1 auto
    int sm_count = 0;
    float4* d_seg_partials = nullptr;  // depth-segment scratch (MarchParams::seg_partials), grow-only
    unsigned int* d_seg_arrived = nullptr;
    size_t seg_capacity = 0, seg_arrived_px = 0;
    void* flush_buf = nullptr;      // 256 MiB scratch, allocated by the first vr_flush_l2
    vr_stats stats{};
    std::string err;
    // ---- several GPUs behind this context (vr_create_info.device_count > 1): sub-contexts of devices[1..]; the frame,
    // the band plan and everything the caller sees live here, on devices[0]
    std::vector<vr_context*> peers;
    vr_context* leader = nullptr;            // set in a sub-context
    vr_params plan_params{};                 // the parameters the plan was made for
    std::vector<float> plan_cost;            // analytic per-row cost (vr_row_cost) the bands are cut from
    std::vector<uint32_t> plan_bounds;       // row bands of the last single-view frame, one per device
    std::vector<uint32_t> plan_views;        // view shards of the last view batch
    uint32_t plan_age = 0;                   // frames since plan_cost was computed
    bool plan_valid = false;
    bool group_timed = false;                // the last vr_render went through the group path (stats: max over the devices)
    cudaEvent_t ev_consumed = nullptr;       // the frame of the last group render has been copied out / handed over
};

static thread_local std::string g_create_error;

static vr_status fail(vr_context* ctx, vr_status st, const char* what, cudaError_t e = cudaSuccess) {
    std::string m = what;
    if (e != cudaSuccess) { m += ": "; m += cudaGetErrorName(e); m += " — "; m += cudaGetErrorString(e); }
    if (ctx) ctx->err = m; else g_create_error = m;
    return st;
}
#define VR_CUDA(ctx, call)                                                                        \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? VR_ERR_OUT_OF_MEMORY : VR_ERR_CUDA, #call, e_); \
    } while (0)

static size_t out_px_bytes(uint32_t fmt) {
    switch (fmt) {
        case VR_OUT_RGBA8_UNORM: case VR_OUT_RGBA8_SRGB: return 4;
        case VR_OUT_RGBA16F: return 8;
        case VR_OUT_RGBA32F: return 16;
    }
    return 0;
}
static size_t texel_bytes(uint32_t fmt) {
    switch (fmt) {
        case VR_VOL_RGBA8_UNORM: return 4;
        case VR_VOL_R16F: return 2;
        case VR_VOL_R8_UNORM: return 1;
    }
    return 0;
}

extern "C" uint32_t vr_version(void) { return (VR_VERSION_MAJOR << 16) | VR_VERSION_MINOR; }

extern "C" const char* vr_last_error(const vr_context* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

extern "C" vr_status vr_params_default(vr_params* p) {
    if (!p) return VR_ERR_INVALID_ARGUMENT;
    std::memset(p, 0, sizeof *p);
    p->struct_size = sizeof *p;
    p->ubo0_size = 192;
    p->ubo1_size = 140;
    p->max_steps = 128;            // frag.glsl:30
    p->step_scale = 4.0f;          // frag.glsl:42
    p->density = 1.0f;             // frag.glsl:29
    p->sample_scale = 0.2f;        // frag.glsl:63
    const float cs[4] = {1.0f, 0.8f, 0.75f, 0.7f}, sw[4] = {0.0f, 0.2f, 0.25f, 0.3f};   // frag.glsl:66-69
    for (int i = 0; i < 4; ++i) { p->coord_scale[i] = cs[i]; p->scroll_weight[i] = sw[i]; }
    p->sun_dir[2] = 1.0f;
    p->sun_color[0] = p->sun_color[1] = p->sun_color[2] = 1.0f;
    p->fbm_octaves = 5; p->fbm_seed = 1337; p->fbm_frequency = 128.0f * 0.03f;
    p->fbm_gain = 0.5f; p->fbm_lacunarity = 2.0f; p->fbm_coverage = 0.45f;
    p->fbm_falloff0 = 0.35f; p->fbm_falloff1 = 0.5f;
    p->view_count = 1;
    p->tree_frequency = 1.0f; p->tree_seed = 1337;
    return VR_OK;
}

extern "C" vr_status vr_create(const vr_create_info* info, vr_context** out) {
    if (!info || !out || info->struct_size < offsetof(vr_create_info, device_count))      // the 28-byte struct of version 0.1 is accepted
        return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: null argument or struct_size too small");
    if (info->width == 0 || info->height == 0 || info->width > 32768 || info->height > 32768)
        return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: width/height out of range");
    if (out_px_bytes(info->format) == 0) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: unknown output format");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, VR_ERR_CUDA, "vr_create: no CUDA device (libvolrt has no CPU path)", e);
    int device0 = info->device;
    if (info->struct_size >= offsetof(vr_create_info, devices) + sizeof(info->devices) && info->device_count > 1) {
        if (info->device_count > VR_MAX_DEVICES) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: device_count exceeds VR_MAX_DEVICES");
        for (uint32_t a = 0; a < info->device_count; ++a) {
            if (info->devices[a] < 0 || info->devices[a] >= ndev) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: devices[] ordinal out of range");
            for (uint32_t b = 0; b < a; ++b)
                if (info->devices[a] == info->devices[b]) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: devices[] names a device twice");
        }
        device0 = info->devices[0];
    }
    if (device0 < 0 || device0 >= ndev) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: device ordinal out of range");
    VR_CUDA(nullptr, cudaSetDevice(device0));
    vr_context* c = new (std::nothrow) vr_context();
    if (!c) return fail(nullptr, VR_ERR_OUT_OF_MEMORY, "vr_create: host allocation failed");
    c->device = device0;
    c->width = info->width; c->height = info->height; c->format = info->format;
    c->max_views = info->max_views ? info->max_views : 1;
    c->px_bytes = out_px_bytes(info->format);
    c->row_pitch = c->px_bytes * c->width;
    c->view_stride = c->row_pitch * c->height;
    c->stats.struct_size = sizeof(vr_stats);
    auto bail = [&](const char* what, cudaError_t ce) {
        vr_status st = fail(nullptr, ce == cudaErrorMemoryAllocation ? VR_ERR_OUT_OF_MEMORY : VR_ERR_CUDA, what, ce);
        vr_destroy(c);
        return st;
    };
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    if ((e = cudaEventCreate(&c->ev0)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&c->ev1)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate(copy)", e);
    for (int i = 0; i < 2; ++i) {
        if ((e = cudaEventCreateWithFlags(&c->ev_render_done[i], cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
        if ((e = cudaEventCreateWithFlags(&c->ev_copy_done[i], cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    }
    if ((e = cudaEventCreateWithFlags(&c->ev_params, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    // the frames, followed by the gather mailbox (kGatherMailboxBytes, zeroed): arrival / release counters per frame slot
    if ((e = cudaMalloc(&c->frame, c->view_stride * c->max_views + kGatherMailboxBytes)) != cudaSuccess) return bail("cudaMalloc(frame)", e);
    if ((e = cudaMemset((char*)c->frame + c->view_stride * c->max_views, 0, kGatherMailboxBytes)) != cudaSuccess) return bail("cudaMemset(mailbox)", e);
    c->view_capacity = c->max_views;
    if ((e = cudaMalloc(&c->d_views, sizeof(ViewConst) * c->view_capacity)) != cudaSuccess) return bail("cudaMalloc(views)", e);
    if ((e = cudaMallocHost(&c->h_views, sizeof(ViewConst) * c->view_capacity)) != cudaSuccess) return bail("cudaMallocHost(views)", e);
    if ((e = cudaMalloc(&c->d_counters, 8 * sizeof(unsigned long long))) != cudaSuccess) return bail("cudaMalloc(counters)", e);
    if ((e = cudaMallocHost(&c->h_counters, 8 * sizeof(unsigned long long))) != cudaSuccess) return bail("cudaMallocHost(counters)", e);
    if ((e = cudaMalloc(&c->d_row_cost, sizeof(float) * c->height)) != cudaSuccess) return bail("cudaMalloc(row_cost)", e);
    if ((e = cudaMalloc(&c->d_queue, 2 * sizeof(unsigned int))) != cudaSuccess) return bail("cudaMalloc(queue)", e);
    if ((e = cudaMemset(c->d_queue, 0, 2 * sizeof(unsigned int))) != cudaSuccess) return bail("cudaMemset(queue)", e);
    c->schedule = (info->flags & VR_CREATE_SCHEDULE_DIRECT) ? 0 : (info->flags & VR_CREATE_SCHEDULE_PERSISTENT) ? 1 : -1;
    if (const char* ev = std::getenv("VOLRT_SCHEDULE")) {            // experiment knob: direct | persistent | auto
        if (!std::strcmp(ev, "direct")) c->schedule = 0;
        else if (!std::strcmp(ev, "persistent")) c->schedule = 1;
        else if (!std::strcmp(ev, "auto")) c->schedule = -1;
    }
    if (const char* ev = std::getenv("VOLRT_RAY_SPLIT")) c->ray_split = std::atoi(ev);
    if (const char* ev = std::getenv("VOLRT_RAY_SPLIT_COMPOSITE")) c->ray_split_composite = std::atoi(ev);
    if (const char* ev = std::getenv("VOLRT_RAY_SEGMENTS")) c->ray_segments = std::atoi(ev);
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, c->device);
    if (const char* ev = std::getenv("VOLRT_QUEUE_CHUNK")) { int v = std::atoi(ev); if (v > 0 && v <= 1024) c->queue_chunk = v; }
    // several GPUs behind this context: one sub-context per further device, peer access both ways with devices[0]
    const bool has_group = info->struct_size >= offsetof(vr_create_info, devices) + sizeof(info->devices) && info->device_count > 1;
    if (has_group) {
        if ((e = cudaEventCreateWithFlags(&c->ev_consumed, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
        for (uint32_t k = 1; k < info->device_count; ++k) {
            vr_create_info sub = {};
            sub.struct_size = (uint32_t)offsetof(vr_create_info, device_count);
            sub.device = info->devices[k]; sub.width = info->width; sub.height = info->height; sub.format = info->format;
            sub.max_views = 1; sub.flags = info->flags;
            vr_context* pc = nullptr;
            const vr_status st = vr_create(&sub, &pc);
            if (st != VR_OK) { vr_destroy(c); return st; }           // the message is in g_create_error already
            pc->leader = c;
            c->peers.push_back(pc);
            int can = 0, can_back = 0;
            cudaDeviceCanAccessPeer(&can, pc->device, c->device);
            cudaDeviceCanAccessPeer(&can_back, c->device, pc->device);
            if (!can || !can_back) { vr_destroy(c); return fail(nullptr, VR_ERR_UNSUPPORTED, "vr_create: no peer access between devices[0] and another device of the group"); }
            cudaSetDevice(pc->device);
            e = cudaDeviceEnablePeerAccess(c->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return bail("cudaDeviceEnablePeerAccess", e);
            cudaSetDevice(c->device);
            e = cudaDeviceEnablePeerAccess(pc->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return bail("cudaDeviceEnablePeerAccess", e);
            cudaGetLastError();
        }
        cudaSetDevice(c->device);
    }
    *out = c;
    return VR_OK;
}

static vr_status replicate_volume(vr_context* c);      // group contexts: copy the leader's volume to the other devices

static void release_volume(vr_context* c) {
    if (c->tex) { cudaDestroyTextureObject(c->tex); c->tex = 0; }
    if (c->array) { cudaFreeArray(c->array); c->array = nullptr; }
    for (int k = 0; k < 4; ++k) {
        if (c->plane_tex[k]) { cudaDestroyTextureObject(c->plane_tex[k]); c->plane_tex[k] = 0; }
        if (c->plane_array[k]) { cudaFreeArray(c->plane_array[k]); c->plane_array[k] = nullptr; }
    }
    if (c->plane_lin) { cudaFree(c->plane_lin); c->plane_lin = nullptr; }
    c->planar = false;
    if (c->lin) { cudaFree(c->lin); c->lin = nullptr; }
    if (c->occupied) { cudaFree(c->occupied); c->occupied = nullptr; }
    if (c->brick_dist) { cudaFree(c->brick_dist); c->brick_dist = nullptr; }
    c->lin_bytes = 0;
    c->has_volume = false;
}

extern "C" void vr_destroy(vr_context* c) {
    if (!c) return;
    for (vr_context* pc : c->peers) { if (pc->stream) { cudaSetDevice(pc->device); cudaStreamSynchronize(pc->stream); } }
    for (vr_context* pc : c->peers) vr_destroy(pc);
    c->peers.clear();
    cudaSetDevice(c->device);
    if (c->ev_consumed) cudaEventDestroy(c->ev_consumed);
    if (c->stream) cudaStreamSynchronize(c->stream);
    release_volume(c);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    if (c->frame) cudaFree(c->frame);
    if (c->frame2) cudaFree(c->frame2);
    if (c->d_chan_mm) cudaFree(c->d_chan_mm);
    for (int i = 0; i < 2; ++i) {
        if (c->ev_render_done[i]) cudaEventDestroy(c->ev_render_done[i]);
        if (c->ev_copy_done[i]) cudaEventDestroy(c->ev_copy_done[i]);
    }
    if (c->ev_params) cudaEventDestroy(c->ev_params);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->d_views) cudaFree(c->d_views);
    if (c->h_views) cudaFreeHost(c->h_views);
    if (c->d_counters) cudaFree(c->d_counters);
    if (c->h_counters) cudaFreeHost(c->h_counters);
    if (c->d_row_cost) cudaFree(c->d_row_cost);
    if (c->d_queue) cudaFree(c->d_queue);
    if (c->d_tree_prog) cudaFree(c->d_tree_prog);
    if (c->d_march_tree) cudaFree(c->d_march_tree);
    if (c->d_tree_out) cudaFree(c->d_tree_out);
    if (c->d_fbm_cells) cudaFree(c->d_fbm_cells);
    if (c->d_seg_partials) cudaFree(c->d_seg_partials);
    if (c->d_seg_arrived) cudaFree(c->d_seg_arrived);
    for (int i = 0; i < vr_context::kFileRing; ++i) {
        if (c->file_stage[i]) cudaFreeHost(c->file_stage[i]);
        if (c->file_done[i]) cudaEventDestroy(c->file_done[i]);
    }
    for (float* pl : c->bake_plane) if (pl) cudaFree(pl);
    if (c->d_mm) cudaFree(c->d_mm);
    if (c->flush_buf) cudaFree(c->flush_buf);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

// build the texture object (TEXTURE layout) and the occupancy grid from c->lin
static vr_status finish_volume(vr_context* c, const vr_volume_desc& d) {
    const uint32_t layout = d.layout == VR_LAYOUT_AUTO ? (uint32_t)VR_LAYOUT_TEXTURE : d.layout;
    bool range_pending = false;
    c->const_mask = 0;
    if (layout == VR_LAYOUT_TEXTURE) {
        cudaChannelFormatDesc cd;
        switch (d.format) {
            case VR_VOL_RGBA8_UNORM: cd = cudaCreateChannelDesc<uchar4>(); break;
            case VR_VOL_R16F: cd = cudaCreateChannelDescHalf(); break;
            default: cd = cudaCreateChannelDesc<unsigned char>(); break;
        }
        cudaExtent ext = make_cudaExtent(d.width, d.height, d.depth);
        if (!c->array) VR_CUDA(c, cudaMalloc3DArray(&c->array, &cd, ext));
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr(c->lin, d.width * texel_bytes(d.format), d.width, d.height);
        cp.dstArray = c->array;
        cp.extent = ext;
        cp.kind = cudaMemcpyDeviceToDevice;
        VR_CUDA(c, cudaMemcpy3DAsync(&cp, c->stream));
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = c->array;
        cudaTextureDesc td = {};
        // VulkanCore.cpp:680-694: LINEAR mag/min, MIRRORED_REPEAT x3, normalised coordinates, one mip
        cudaTextureAddressMode am = d.address_mode == VR_ADDRESS_CLAMP_TO_EDGE ? cudaAddressModeClamp : cudaAddressModeMirror;
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = am;
        td.filterMode = cudaFilterModeLinear;
        td.readMode = d.format == VR_VOL_R16F ? cudaReadModeElementType : cudaReadModeNormalizedFloat;
        td.normalizedCoords = 1;
        if (!c->tex) VR_CUDA(c, cudaCreateTextureObject(&c->tex, &rd, &td, nullptr));
        // planar copies of an RGBA8 volume (decision 10); VOLRT_PLANAR=0 keeps the interleaved texture only
        static const bool want_planar = [] { const char* e = getenv("VOLRT_PLANAR"); return !e || atoi(e) != 0; }();
        c->planar = false;
        if (d.format == VR_VOL_RGBA8_UNORM && want_planar) {
            const size_t vox = (size_t)d.width * d.height * d.depth;
            if (!c->plane_lin) VR_CUDA(c, cudaMalloc(&c->plane_lin, 4 * vox));
            static const bool want_const = [] { const char* e = getenv("VOLRT_CONST_CHANNELS"); return !e || atoi(e) != 0; }();
            if (want_const) {
                // min words start at 0xFFFFFFFF, max words at 0; the de-interleave kernel folds the texels' bytes into them
                if (!c->d_chan_mm) VR_CUDA(c, cudaMalloc(&c->d_chan_mm, 8 * sizeof(unsigned int)));
                VR_CUDA(c, cudaMemsetAsync(c->d_chan_mm, 0xFF, 4 * sizeof(unsigned int), c->stream));
                VR_CUDA(c, cudaMemsetAsync(c->d_chan_mm + 4, 0, 4 * sizeof(unsigned int), c->stream));
                range_pending = true;
            }
            VR_CUDA(c, launch_deinterleave_rgba8((const uint8_t*)c->lin, c->plane_lin, vox, want_const ? c->d_chan_mm : nullptr, c->stream));
            const cudaChannelFormatDesc cd1 = cudaCreateChannelDesc<unsigned char>();
            for (int k = 0; k < 4; ++k) {
                if (!c->plane_array[k]) VR_CUDA(c, cudaMalloc3DArray(&c->plane_array[k], &cd1, ext));
                cudaMemcpy3DParms pp = {};
                pp.srcPtr = make_cudaPitchedPtr(c->plane_lin + (size_t)k * vox, d.width, d.width, d.height);
                pp.dstArray = c->plane_array[k];
                pp.extent = ext;
                pp.kind = cudaMemcpyDeviceToDevice;
                VR_CUDA(c, cudaMemcpy3DAsync(&pp, c->stream));
                cudaResourceDesc rdk = {};
                rdk.resType = cudaResourceTypeArray;
                rdk.res.array.array = c->plane_array[k];
                if (!c->plane_tex[k]) VR_CUDA(c, cudaCreateTextureObject(&c->plane_tex[k], &rdk, &td, nullptr));
            }
            c->planar = true;
        }
    }
    c->has_tensor_map = false;
    if (layout == VR_LAYOUT_BRICK) {
        // 3-D tiled tensor map over the x-fastest fp16 volume; the march kernel pulls 24^3-voxel boxes through it
        typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        VR_CUDA(c, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
        if (!fn || qr != cudaDriverEntryPointSuccess) return fail(c, VR_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled is not available in this driver");
        const cuuint64_t dims[3] = {d.width, d.height, d.depth};
        const cuuint64_t strides[2] = {(cuuint64_t)d.width * 2, (cuuint64_t)d.width * d.height * 2};
        const cuuint32_t boxdim[3] = {(cuuint32_t)kBrickBox[0], (cuuint32_t)kBrickBox[1], (cuuint32_t)kBrickBox[2]};
        const cuuint32_t estr[3] = {1, 1, 1};
        CUresult r = ((EncodeTiled)fn)((CUtensorMap*)c->tensor_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, c->lin, dims, strides, boxdim, estr,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(c, VR_ERR_CUDA, "cuTensorMapEncodeTiled failed");
        c->has_tensor_map = true;
    }
    // min-max bricks (8^3 voxels + 1-texel apron) for empty-space skipping
    BrickInfo& B = c->bricks;
    B.bx = (d.width + 7) / 8; B.by = (d.height + 7) / 8; B.bz = (d.depth + 7) / 8;
    B.inv_brick[0] = (float)d.width / 8.0f; B.inv_brick[1] = (float)d.height / 8.0f; B.inv_brick[2] = (float)d.depth / 8.0f;
    if (!c->occupied) VR_CUDA(c, cudaMalloc(&c->occupied, (size_t)B.bx * B.by * B.bz));
    B.occupied = c->occupied;
    const size_t nbricks = (size_t)B.bx * B.by * B.bz;
    const size_t npad = (size_t)(B.bx + 2) * (B.by + 2) * (B.bz + 2);
    if (!c->brick_dist) VR_CUDA(c, cudaMalloc(&c->brick_dist, 2 * nbricks + npad));      // distance map, pass scratch, padded copy
    B.dist = c->brick_dist;
    B.dist_pad = c->brick_dist + 2 * nbricks;
    {
        const unsigned long long cell = (unsigned long long)(B.by + 2) * (B.bx + 2) + (B.bx + 2) + 1;
        B.pad_bias = (unsigned int)((0x4B000000ull - 1ull) * cell);       // mod 2^32, like the index arithmetic it corrects
        B.pad_last = (unsigned int)(npad - 1);
    }
    if (c->trusted_level0) VR_CUDA(c, launch_occupancy_from_pyramid(c->trusted_level0, d.format == VR_VOL_RGBA8_UNORM ? 4 : 1, c->occupied, (int)nbricks, c->stream));
    else VR_CUDA(c, launch_occupancy(c->lin, d.format, d.width, d.height, d.depth, d.address_mode, c->occupied, B.bx, B.by, B.bz, c->stream));
    VR_CUDA(c, launch_brick_distance(c->occupied, c->brick_dist, c->brick_dist + nbricks, B.bx, B.by, B.bz, c->stream));
    VR_CUDA(c, launch_brick_pad(c->brick_dist, c->brick_dist + 2 * nbricks, B.bx, B.by, B.bz, c->stream));
    VR_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    c->foreign_stream_pending = false;
    c->stats.rays = 0;
    c->stats.kernel_launches = c->fill_launches + 5;       // + the occupancy kernel, the 3 distance passes and the padded copy
    c->fill_launches = 0;
    std::memset(c->h_counters, 0, 8 * sizeof(unsigned long long));
    unsigned int mm[8];
    if (range_pending) VR_CUDA(c, cudaMemcpyAsync(mm, c->d_chan_mm, sizeof mm, cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    if (range_pending)
        for (int k = 0; k < 4; ++k)
            if (mm[k] == mm[4 + k]) { c->const_mask |= 1u << k; c->const_val[k] = (float)mm[k] / 255.0f; }
    c->stats.constant_channels = c->const_mask;
    c->vdesc = d;
    c->vdesc.layout = layout;
    c->has_volume = true;
    return VR_OK;
}

// storage for a volume of this description: kept when the previous volume had the same shape (an interactive
// host re-uploads or re-bakes every so often; allocation and texture-object creation cost far more than the fill)
static vr_status prepare_volume(vr_context* c, const vr_volume_desc& d) {
    const uint32_t layout = d.layout == VR_LAYOUT_AUTO ? (uint32_t)VR_LAYOUT_TEXTURE : d.layout;
    const vr_volume_desc& o = c->vdesc;
    const bool same = c->has_volume && o.width == d.width && o.height == d.height && o.depth == d.depth && o.format == d.format &&
                      o.address_mode == d.address_mode && o.layout == layout;
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    // a march enqueued on a caller's stream, or a frame copy on the copy stream, may still be reading the storage about
    // to be rewritten
    if (c->foreign_stream_pending) { VR_CUDA(c, cudaEventSynchronize(c->ev1)); c->foreign_stream_pending = false; }
    VR_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    if (same) c->has_volume = false;                         // storage stays; has_volume returns with finish_volume
    else {
        release_volume(c);
        c->lin_bytes = (size_t)d.width * d.height * d.depth * texel_bytes(d.format);
        VR_CUDA(c, cudaMalloc(&c->lin, c->lin_bytes));
    }
    c->fill_launches = 0;
    VR_CUDA(c, cudaEventRecord(c->ev0, c->stream));          // vr_get_stats().gpu_ms of an upload / bake: fill + texture + occupancy
    return VR_OK;
}

static vr_status check_volume_desc(vr_context* c, const vr_volume_desc* d) {
    if (!d || d->struct_size < sizeof(vr_volume_desc)) return fail(c, VR_ERR_INVALID_ARGUMENT, "volume desc: null or struct_size too small");
    if (d->width == 0 || d->height == 0 || d->depth == 0 || d->width > 4096 || d->height > 4096 || d->depth > 4096)
        return fail(c, VR_ERR_INVALID_ARGUMENT, "volume desc: dimensions must be in 1..4096");
    if (texel_bytes(d->format) == 0) return fail(c, VR_ERR_INVALID_ARGUMENT, "volume desc: unknown format");
    if (d->address_mode > VR_ADDRESS_CLAMP_TO_EDGE) return fail(c, VR_ERR_INVALID_ARGUMENT, "volume desc: unknown address mode");
    if (d->layout > VR_LAYOUT_BRICK) return fail(c, VR_ERR_INVALID_ARGUMENT, "volume desc: unknown layout");
    if (d->layout == VR_LAYOUT_BRICK && (d->format != VR_VOL_R16F || d->width % 8 != 0))
        return fail(c, VR_ERR_UNSUPPORTED, "volume desc: VR_LAYOUT_BRICK needs an R16F volume whose width is a multiple of 8 (TMA: 16-byte row pitch)");
    return VR_OK;
}

extern "C" vr_status vr_upload_volume(vr_context* c, const vr_volume_desc* d, const void* texels) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!texels) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_upload_volume: texels is null");
    vr_status st = check_volume_desc(c, d);
    if (st != VR_OK) return st;
    VR_CUDA(c, cudaSetDevice(c->device));
    st = prepare_volume(c, *d);
    if (st != VR_OK) return st;
    // synchronous like the reference's staging copy + vkQueueWaitIdle (VulkanCore.cpp:529-541)
    VR_CUDA(c, cudaMemcpyAsync(c->lin, texels, c->lin_bytes, cudaMemcpyHostToDevice, c->stream));
    st = finish_volume(c, *d);
    if (st != VR_OK) { release_volume(c); return st; }
    return replicate_volume(c);
}

// group context: the leader's resident volume, copied to every other device over NVLink (peer copy of the x-fastest texels), where
// the texture objects, the occupancy grid and the distance map are built like after an upload
static vr_status replicate_volume(vr_context* c) {
    if (c->peers.empty() || !c->has_volume) return VR_OK;
    vr_volume_desc d = c->vdesc;
    d.struct_size = sizeof d;
    vr_status st = VR_OK;
    for (vr_context* pc : c->peers) {
        cudaError_t e = cudaSetDevice(pc->device);
        if (e == cudaSuccess) st = prepare_volume(pc, d);
        if (e == cudaSuccess && st == VR_OK) e = cudaMemcpyPeerAsync(pc->lin, pc->device, c->lin, c->device, c->lin_bytes, pc->stream);
        if (e == cudaSuccess && st == VR_OK) st = finish_volume(pc, d);
        if (e != cudaSuccess || st != VR_OK) {
            release_volume(pc);
            cudaSetDevice(c->device);
            if (e != cudaSuccess) return fail(c, e == cudaErrorMemoryAllocation ? VR_ERR_OUT_OF_MEMORY : VR_ERR_CUDA, "replicating the volume to a device of the group", e);
            return fail(c, st, ("replicating the volume to a device of the group: " + pc->err).c_str());
        }
    }
    VR_CUDA(c, cudaSetDevice(c->device));
    return VR_OK;
}

extern "C" vr_status vr_download_volume(vr_context* c, void* texels, size_t bytes) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!c->has_volume) return fail(c, VR_ERR_BAD_STATE, "vr_download_volume: no volume");
    if (!texels || bytes < c->lin_bytes) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_download_volume: buffer too small");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaMemcpyAsync(texels, c->lin, c->lin_bytes, cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    return VR_OK;
}

// ------------------------------------------------------------------ min-max pyramid and volume files
// device pyramid of c->lin into a fresh device buffer (caller frees)
static vr_status build_pyramid(vr_context* c, float** d_out, size_t* floats, uint32_t* levels) {
    const vr_volume_desc& d = c->vdesc;
    *floats = pyramid_floats(d.width, d.height, d.depth, d.format, levels);
    VR_CUDA(c, cudaMalloc(d_out, *floats * sizeof(float)));
    const int ch = d.format == VR_VOL_RGBA8_UNORM ? 4 : 1;
    int bx = (d.width + 7) / 8, by = (d.height + 7) / 8, bz = (d.depth + 7) / 8;
    float* level = *d_out;
    cudaError_t e = launch_pyramid_base(c->lin, d.format, d.width, d.height, d.depth, level, bx, by, bz, c->stream);
    while (e == cudaSuccess && !(bx == 1 && by == 1 && bz == 1)) {
        float* next = level + (size_t)bx * by * bz * ch * 2;
        e = launch_pyramid_reduce(level, next, ch, bx, by, bz, c->stream);
        level = next;
        bx = (bx + 1) / 2; by = (by + 1) / 2; bz = (bz + 1) / 2;
    }
    if (e != cudaSuccess) { cudaFree(*d_out); *d_out = nullptr; return fail(c, VR_ERR_CUDA, "pyramid kernels", e); }
    return VR_OK;
}

extern "C" vr_status vr_volume_pyramid(vr_context* c, float* out, size_t out_bytes) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!c->has_volume) return fail(c, VR_ERR_BAD_STATE, "vr_volume_pyramid: no volume");
    VR_CUDA(c, cudaSetDevice(c->device));
    float* d_pyr = nullptr;
    size_t floats = 0;
    uint32_t levels = 0;
    if (!out || out_bytes < pyramid_floats(c->vdesc.width, c->vdesc.height, c->vdesc.depth, c->vdesc.format, nullptr) * sizeof(float))
        return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_volume_pyramid: buffer too small (vr_pyramid_floats)");
    vr_status st = build_pyramid(c, &d_pyr, &floats, &levels);
    if (st != VR_OK) return st;
    cudaError_t e = cudaMemcpyAsync(out, d_pyr, floats * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_pyr);
    if (e != cudaSuccess) return fail(c, VR_ERR_CUDA, "vr_volume_pyramid", e);
    return VR_OK;
}

extern "C" vr_status vr_save_volume_file(vr_context* c, const char* path, uint32_t flags) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!path) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_save_volume_file: path is null");
    if (!c->has_volume) return fail(c, VR_ERR_BAD_STATE, "vr_save_volume_file: no volume");
    VR_CUDA(c, cudaSetDevice(c->device));
    std::vector<unsigned char> texels;
    std::vector<float> pyr;
    uint32_t levels = 0;
    try {
        texels.resize(c->lin_bytes);
        if (flags & VR_FILE_WITH_PYRAMID) pyr.resize(pyramid_floats(c->vdesc.width, c->vdesc.height, c->vdesc.depth, c->vdesc.format, &levels));
    } catch (const std::bad_alloc&) { return fail(c, VR_ERR_OUT_OF_MEMORY, "vr_save_volume_file: host memory"); }
    vr_status st = vr_download_volume(c, texels.data(), texels.size());
    if (st == VR_OK && !pyr.empty()) st = vr_volume_pyramid(c, pyr.data(), pyr.size() * sizeof(float));
    if (st != VR_OK) return st;
    st = vr_volume_file_write(path, &c->vdesc, texels.data(), pyr.empty() ? nullptr : pyr.data(), levels);
    if (st != VR_OK) return fail(c, st, "vr_save_volume_file: cannot write the file");
    return VR_OK;
}

// file -> ring of page-locked staging buffers -> device; reads, hashing and H2D copies overlap
extern "C" vr_status vr_upload_volume_file(vr_context* c, const char* path, uint32_t layout, uint32_t flags) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!path) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_upload_volume_file: path is null");
    std::FILE* f = std::fopen(path, "rb");
    if (!f) return fail(c, VR_ERR_IO, "vr_upload_volume_file: cannot open the file");
    VolumeFileHeader h{};
    if (const char* err = read_volume_header(f, &h)) { std::fclose(f); return fail(c, VR_ERR_IO, (std::string("vr_upload_volume_file: ") + err).c_str()); }
    vr_volume_desc vd = {sizeof(vr_volume_desc), h.width, h.height, h.depth, h.format, h.address_mode, layout};
    vr_status st = check_volume_desc(c, &vd);
    if (st != VR_OK) { std::fclose(f); return st; }
    if ((flags & VR_FILE_TRUST_PYRAMID) && !h.pyramid_levels) { std::fclose(f); return fail(c, VR_ERR_BAD_STATE, "vr_upload_volume_file: the file has no pyramid to trust"); }
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) { std::fclose(f); return fail(c, VR_ERR_CUDA, "cudaSetDevice", e); }
    st = prepare_volume(c, vd);
    if (st != VR_OK) { std::fclose(f); return st; }
    // kReaders threads pread() 16 MiB chunks into a ring of page-locked buffers and queue the H2D copies themselves;
    // a buffer is reused once the copy that last read it has finished (event), and the SHA-256 (inherently
    // sequential) is taken chunk by chunk in file order by whichever thread holds that chunk.
    constexpr int kReaders = 4, kRing = vr_context::kFileRing;
    const size_t chunk = vr_context::kFileChunk;
    const size_t n_chunks = (size_t)((h.texel_bytes + chunk - 1) / chunk);
    unsigned char** stage = c->file_stage;
    cudaEvent_t* done = c->file_done;
    float* d_level0 = nullptr;
    const char* what = nullptr;
    Sha256 sha;
    sha.reset();
    const int ring = (int)(n_chunks < (size_t)kRing ? n_chunks : (size_t)kRing);
    for (int i = 0; i < ring && e == cudaSuccess; ++i) {
        if (!stage[i]) e = cudaMallocHost(&stage[i], chunk);
        if (e == cudaSuccess && !done[i]) e = cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) {
        std::mutex mu;
        std::condition_variable cv;
        size_t next_chunk = 0, hash_turn = 0;
        size_t issued[kRing];                      // issued[b] = number of copies queued from buffer b so far
        for (size_t& v : issued) v = 0;
        cudaError_t first_err = cudaSuccess;
        const char* first_what = nullptr;
        const int fd = fileno(f);
        const bool verify = (flags & VR_FILE_VERIFY) != 0;
        auto worker = [&]() {
            cudaSetDevice(c->device);
            for (;;) {
                size_t k;
                {
                    std::unique_lock<std::mutex> lk(mu);
                    if (first_err != cudaSuccess || first_what || next_chunk >= n_chunks) return;
                    k = next_chunk++;
                }
                const int b = (int)(k % ring);
                const size_t off = k * chunk, n = h.texel_bytes - off < chunk ? (size_t)(h.texel_bytes - off) : chunk;
                {   // wait until chunk k - ring has been queued from this buffer, then until its copy has finished
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return issued[b] == k / ring || first_err != cudaSuccess || first_what; });
                    if (first_err != cudaSuccess || first_what) return;
                }
                cudaError_t le = k >= (size_t)ring ? cudaEventSynchronize(done[b]) : cudaSuccess;
                const char* lw = nullptr;
                if (le == cudaSuccess) {
                    size_t got = 0;
                    while (got < n) {
                        const ssize_t rd = pread(fd, stage[b] + got, n - got, (off_t)(h.texel_offset + off + got));
                        if (rd <= 0) { lw = "short read (texels)"; break; }
                        got += (size_t)rd;
                    }
                }
                if (le == cudaSuccess && !lw && verify) {
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return hash_turn == k || first_err != cudaSuccess || first_what; });
                    if (hash_turn == k) {
                        lk.unlock();
                        sha.update(stage[b], n);
                        lk.lock();
                        ++hash_turn;
                        cv.notify_all();
                    }
                }
                if (le == cudaSuccess && !lw) le = cudaMemcpyAsync((char*)c->lin + off, stage[b], n, cudaMemcpyHostToDevice, c->stream);
                if (le == cudaSuccess && !lw) le = cudaEventRecord(done[b], c->stream);
                {
                    std::unique_lock<std::mutex> lk(mu);
                    if (le != cudaSuccess && first_err == cudaSuccess) first_err = le;
                    if (lw && !first_what) first_what = lw;
                    ++issued[b];
                    cv.notify_all();
                }
            }
        };
        std::vector<std::thread> pool;
        const int nthreads = (int)(n_chunks < (size_t)kReaders ? n_chunks : (size_t)kReaders);
        for (int t = 1; t < nthreads; ++t) pool.emplace_back(worker);
        worker();
        for (auto& t : pool) t.join();
        e = first_err;
        what = first_what;
    }
    if (e == cudaSuccess && !what && (flags & VR_FILE_VERIFY)) {
        uint8_t got[32];
        sha.finish(got);
        if (std::memcmp(got, h.sha256, 32) != 0) what = "SHA-256 of the texels does not match the header";
    }
    if (e == cudaSuccess && !what && (flags & VR_FILE_TRUST_PYRAMID)) {
        const size_t l0 = (size_t)((h.width + 7) / 8) * ((h.height + 7) / 8) * ((h.depth + 7) / 8) * (h.format == VR_VOL_RGBA8_UNORM ? 4 : 1) * 2 * sizeof(float);
        std::vector<unsigned char> host(l0);
        if (fseeko(f, (off_t)h.pyramid_offset, SEEK_SET) != 0 || std::fread(host.data(), 1, l0, f) != l0) what = "short read (pyramid)";
        else {
            e = cudaMalloc(&d_level0, l0);
            if (e == cudaSuccess) e = cudaMemcpyAsync(d_level0, host.data(), l0, cudaMemcpyHostToDevice, c->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);    // `host` goes out of scope
        }
    }
    std::fclose(f);
    if (e == cudaSuccess && !what) {
        c->trusted_level0 = d_level0;
        st = finish_volume(c, vd);
        c->trusted_level0 = nullptr;
    }
    cudaStreamSynchronize(c->stream);
    if (d_level0) cudaFree(d_level0);
    if (what) { release_volume(c); return fail(c, VR_ERR_IO, (std::string("vr_upload_volume_file: ") + what).c_str()); }
    if (e != cudaSuccess) { release_volume(c); return fail(c, e == cudaErrorMemoryAllocation ? VR_ERR_OUT_OF_MEMORY : VR_ERR_CUDA, "vr_upload_volume_file", e); }
    if (st != VR_OK) { release_volume(c); return st; }
    return replicate_volume(c);
}

// ------------------------------------------------------------------ volume bake on the device
extern "C" vr_status vr_noise_desc_default(vr_noise_desc* d) {
    if (!d) return VR_ERR_INVALID_ARGUMENT;
    std::memset(d, 0, sizeof *d);
    d->struct_size = sizeof *d;
    d->size = 128;                                   // TestMain.cpp:51
    d->format = VR_VOL_RGBA8_UNORM;
    d->layout = VR_LAYOUT_TEXTURE;
    d->address_mode = VR_ADDRESS_MIRRORED_REPEAT;
    // TestMain.cpp:59-62 with quirk Q1: buffer 1 ends up holding the (0.03, seed 2) cellular field
    // but is normalised with the (0.01, seed 1) field's range; buffer 2 is never written (zeros)
    // and is normalised with the (0.03, seed 2) range.
    d->channel[0] = {VR_NOISE_CELLULAR_DISTANCE, 0.03f, 2, VR_NOISE_CELLULAR_DISTANCE, 0.01f, 1, 4};   // :80 s1^4
    d->channel[1] = {VR_NOISE_ZERO, 0.0f, 0, VR_NOISE_CELLULAR_DISTANCE, 0.03f, 2, 1};
    d->channel[2] = {VR_NOISE_PERLIN, 0.19f, 3, VR_NOISE_SAME, 0.0f, 0, 1};
    d->channel[3] = {VR_NOISE_SIMPLEX, 0.15f, 4, VR_NOISE_SAME, 0.0f, 0, 1};
    return VR_OK;
}

static float fbm_bounding(int octaves, float gain) {    // Fractal.h:27-38
    float g = std::fabs(gain), amp = g, sum = 1.0f;
    for (int i = 1; i < octaves; ++i) { sum += amp; amp *= g; }
    return 1.0f / sum;
}

extern "C" vr_status vr_generate_volume(vr_context* c, const vr_noise_desc* d, const vr_params* fbm) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!d || d->struct_size < offsetof(vr_noise_desc, tree)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_generate_volume: null desc or struct_size too small");   // the tree pointers are optional
    vr_volume_desc vd = {sizeof(vr_volume_desc), d->size, d->size, d->size, d->format, d->address_mode, d->layout};
    vr_status st = check_volume_desc(c, &vd);
    if (st != VR_OK) return st;
    if (d->format == VR_VOL_R8_UNORM) return fail(c, VR_ERR_UNSUPPORTED, "vr_generate_volume: RGBA8 (channel recipes) or R16F (fBm smoke) only");
    // every argument is checked BEFORE the resident volume's storage is touched: a bad call leaves the context as it was
    if (d->format == VR_VOL_R16F) {
        if (fbm && (fbm->fbm_octaves < 1 || fbm->fbm_octaves > 16)) return fail(c, VR_ERR_INVALID_ARGUMENT, "fbm_octaves must be in 1..16");
    } else {
        const bool trees = d->struct_size >= offsetof(vr_noise_desc, tree) + sizeof(d->tree);
        for (int k = 0; k < 4; ++k) {
            const vr_noise_channel& ch = d->channel[k];
            const vr_noise_tree* tr = (ch.generator == VR_NOISE_TREE && trees) ? d->tree[k] : nullptr;
            if (ch.generator > VR_NOISE_TREE || (ch.generator == VR_NOISE_TREE && (!tr || tr->prog.empty())) ||
                (ch.norm_generator > VR_NOISE_ZERO && ch.norm_generator != VR_NOISE_SAME) || ch.power < 1 || ch.power > 8)
                return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_generate_volume: bad channel recipe (a VR_NOISE_TREE channel needs desc.tree[k] and normalises with VR_NOISE_SAME or a built-in field)");
            if (ch.norm_generator == VR_NOISE_ZERO || (ch.norm_generator == VR_NOISE_SAME && ch.generator == VR_NOISE_ZERO))
                return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_generate_volume: a channel needs a non-zero field to normalise with");
        }
    }
    VR_CUDA(c, cudaSetDevice(c->device));
    st = prepare_volume(c, vd);
    if (st != VR_OK) return st;
    const int n = (int)d->size;
    const size_t vox = (size_t)n * n * n;
    if (d->format == VR_VOL_R16F) {
        vr_params def;
        vr_params_default(&def);
        const vr_params* f = fbm ? fbm : &def;
        if (f->fbm_octaves < 1 || f->fbm_octaves > 16) { release_volume(c); return fail(c, VR_ERR_INVALID_ARGUMENT, "fbm_octaves must be in 1..16"); }
        FbmField ff = {f->fbm_octaves, f->fbm_seed, f->fbm_frequency, f->fbm_gain, f->fbm_lacunarity, f->fbm_coverage,
                       f->fbm_falloff0, f->fbm_falloff1, fbm_bounding(f->fbm_octaves, f->fbm_gain), d->threshold,
                       1.0f / (f->fbm_falloff1 - f->fbm_falloff0)};
        VR_CUDA(c, launch_smoke_fill(ff, n, (__half*)c->lin, c->stream));
        c->fill_launches = 1;
    } else {
        // each distinct field is generated once; min/max folded on the device
        struct Gen { NoiseField f; float* buf; unsigned int* mm; bool keep; const vr_noise_tree* tree; };
        std::vector<Gen> gens;
        auto find = [&](uint32_t g, float fr, int32_t s, const vr_noise_tree* tr) -> int {
            for (size_t i = 0; i < gens.size(); ++i)
                if (gens[i].f.generator == g && gens[i].f.frequency == fr && gens[i].f.seed == s && gens[i].tree == tr) return (int)i;
            return -1;
        };
        const bool has_trees = d->struct_size >= offsetof(vr_noise_desc, tree) + sizeof(d->tree);
        int field_of[4], norm_of[4];
        for (int k = 0; k < 4; ++k) {
            const vr_noise_channel& ch = d->channel[k];
            const vr_noise_tree* tr = (ch.generator == VR_NOISE_TREE && has_trees) ? d->tree[k] : nullptr;
            const bool bad_gen = ch.generator > VR_NOISE_TREE || (ch.generator == VR_NOISE_TREE && (!tr || tr->prog.empty()));
            if (bad_gen || (ch.norm_generator > VR_NOISE_ZERO && ch.norm_generator != VR_NOISE_SAME) || ch.power < 1 || ch.power > 8) {
                release_volume(c);
                return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_generate_volume: bad channel recipe (a VR_NOISE_TREE channel needs desc.tree[k] and normalises with VR_NOISE_SAME or a built-in field)");
            }
            field_of[k] = -1;
            if (ch.generator != VR_NOISE_ZERO) {
                field_of[k] = find(ch.generator, ch.frequency, ch.seed, tr);
                if (field_of[k] < 0) { gens.push_back({{ch.generator, ch.frequency, ch.seed}, nullptr, nullptr, true, tr}); field_of[k] = (int)gens.size() - 1; }
                gens[field_of[k]].keep = true;
            }
            if (ch.norm_generator == VR_NOISE_SAME) norm_of[k] = field_of[k];
            else if (ch.norm_generator == VR_NOISE_ZERO) norm_of[k] = -1;
            else {
                norm_of[k] = find(ch.norm_generator, ch.norm_frequency, ch.norm_seed, nullptr);
                if (norm_of[k] < 0) { gens.push_back({{ch.norm_generator, ch.norm_frequency, ch.norm_seed}, nullptr, nullptr, false, nullptr}); norm_of[k] = (int)gens.size() - 1; }
            }
            if (norm_of[k] < 0) { release_volume(c); return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_generate_volume: a channel needs a non-zero field to normalise with"); }
        }
        // scratch planes (one fp32 plane per kept field, at most 4) and the min/max words live in the context
        std::vector<unsigned int> h_mm(gens.size() * 2);
        for (size_t i = 0; i < gens.size(); ++i) minmax_init(&h_mm[i * 2]);
        cudaError_t e = cudaSuccess;
        if (!c->d_mm) e = cudaMalloc(&c->d_mm, 16 * sizeof(unsigned int));     // <= 8 distinct fields (4 channels x {field, norm})
        unsigned int* d_mm = c->d_mm;
        if (c->bake_plane_bytes != vox * sizeof(float)) {
            for (float*& pl : c->bake_plane) { if (pl) cudaFree(pl); pl = nullptr; }
            c->bake_plane_bytes = vox * sizeof(float);
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_mm, h_mm.data(), h_mm.size() * sizeof(unsigned int), cudaMemcpyHostToDevice, c->stream);
        int plane = 0;
        for (size_t i = 0; i < gens.size() && e == cudaSuccess; ++i) {
            gens[i].mm = d_mm + i * 2;
            if (gens[i].keep) {
                if (!c->bake_plane[plane]) e = cudaMalloc(&c->bake_plane[plane], c->bake_plane_bytes);
                gens[i].buf = c->bake_plane[plane++];
                if (e == cudaSuccess && gens[i].tree) {
                    // a node tree: the same GenUniformGrid3D walk, interpreted (noise_tree.cu); the program buffer is reused in stream order
                    if (!c->d_tree_prog) e = cudaMalloc(&c->d_tree_prog, kTreeMaxInstr * sizeof(TreeInstr));
                    if (e == cudaSuccess) e = cudaMemcpyAsync(c->d_tree_prog, gens[i].tree->prog.data(), gens[i].tree->prog.size() * sizeof(TreeInstr), cudaMemcpyHostToDevice, c->stream);
                    if (e == cudaSuccess) e = launch_tree_grid(c->d_tree_prog, (int)gens[i].tree->prog.size(), gens[i].tree->result_reg, 0, 0, 0, n, n, n,
                                                               gens[i].f.frequency, gens[i].f.seed, gens[i].buf, gens[i].mm, c->stream);
                } else if (e == cudaSuccess) e = launch_noise_fill(gens[i].f, n, gens[i].buf, gens[i].mm, c->stream);
            } else {
                e = launch_noise_minmax_only(gens[i].f, n, gens[i].mm, c->stream);
            }
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(h_mm.data(), d_mm, h_mm.size() * sizeof(unsigned int), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e == cudaSuccess) {
            PackChannel pc[4];
            for (int k = 0; k < 4; ++k) {
                pc[k].field = field_of[k] >= 0 ? gens[field_of[k]].buf : nullptr;
                pc[k].mn = minmax_decode(h_mm[norm_of[k] * 2]);
                pc[k].mx = minmax_decode(h_mm[norm_of[k] * 2 + 1]);
                pc[k].power = d->channel[k].power;
            }
            e = launch_noise_pack(pc, (int64_t)vox, (uint8_t*)c->lin, c->stream);
            c->fill_launches = (uint32_t)gens.size() + 1;
        }
        if (c->bake_plane_bytes > ((size_t)1 << 30)) {         // very large bakes give their scratch back (cudaFree waits for the pack)
            for (float*& pl : c->bake_plane) { if (pl) cudaFree(pl); pl = nullptr; }
            c->bake_plane_bytes = 0;
        }
        if (e != cudaSuccess) { release_volume(c); return fail(c, e == cudaErrorMemoryAllocation ? VR_ERR_OUT_OF_MEMORY : VR_ERR_CUDA, "vr_generate_volume", e); }
    }
    st = finish_volume(c, vd);
    if (st != VR_OK) { release_volume(c); return st; }
    return replicate_volume(c);
}

// ------------------------------------------------------------------ FastNoise2 node trees on the device
extern "C" vr_status vr_noise_tree_grid(vr_context* c, const vr_noise_tree* t, int32_t x0, int32_t y0, int32_t z0, uint32_t nx, uint32_t ny,
                                        uint32_t nz, float frequency, int32_t seed, float* out_host, float minmax[2]) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!t || t->prog.empty()) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_noise_tree_grid: no compiled tree");
    if (nx == 0 || ny == 0 || nz == 0 || nx > 4096 || ny > 4096 || nz > 4096) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_noise_tree_grid: grid dimensions must be in 1..4096");
    VR_CUDA(c, cudaSetDevice(c->device));
    const size_t bytes = (size_t)nx * ny * nz * sizeof(float);
    if (!c->d_tree_prog) VR_CUDA(c, cudaMalloc(&c->d_tree_prog, kTreeMaxInstr * sizeof(TreeInstr)));
    if (!c->d_mm) VR_CUDA(c, cudaMalloc(&c->d_mm, 16 * sizeof(unsigned int)));
    if (c->tree_out_bytes < bytes) {
        VR_CUDA(c, cudaStreamSynchronize(c->stream));
        if (c->d_tree_out) { cudaFree(c->d_tree_out); c->d_tree_out = nullptr; c->tree_out_bytes = 0; }
        cudaError_t e = cudaMalloc(&c->d_tree_out, bytes);
        if (e != cudaSuccess) return fail(c, e == cudaErrorMemoryAllocation ? VR_ERR_OUT_OF_MEMORY : VR_ERR_CUDA, "vr_noise_tree_grid: output plane", e);
        c->tree_out_bytes = bytes;
    }
    unsigned int h_mm[2];
    minmax_init(h_mm);
    VR_CUDA(c, cudaMemcpyAsync(c->d_tree_prog, t->prog.data(), t->prog.size() * sizeof(TreeInstr), cudaMemcpyHostToDevice, c->stream));
    VR_CUDA(c, cudaMemcpyAsync(c->d_mm, h_mm, sizeof h_mm, cudaMemcpyHostToDevice, c->stream));
    VR_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    VR_CUDA(c, launch_tree_grid(c->d_tree_prog, (int)t->prog.size(), t->result_reg, x0, y0, z0, (int)nx, (int)ny, (int)nz, frequency, seed,
                                c->d_tree_out, c->d_mm, c->stream));
    VR_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    c->foreign_stream_pending = false;
    c->stats.rays = 0;
    c->stats.kernel_launches = 1;
    std::memset(c->h_counters, 0, 8 * sizeof(unsigned long long));
    VR_CUDA(c, cudaMemcpyAsync(h_mm, c->d_mm, sizeof h_mm, cudaMemcpyDeviceToHost, c->stream));
    if (out_host) VR_CUDA(c, cudaMemcpyAsync(out_host, c->d_tree_out, bytes, cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    if (minmax) { minmax[0] = minmax_decode(h_mm[0]); minmax[1] = minmax_decode(h_mm[1]); }
    return VR_OK;
}

// ------------------------------------------------------------------ params
static void mat_mul_d(const float* a, const float* b, double* out) {      // out = a*b, column-major
    for (int col = 0; col < 4; ++col)
        for (int r = 0; r < 4; ++r) {
            double s = 0.0;
            for (int k = 0; k < 4; ++k) s += (double)a[k * 4 + r] * (double)b[col * 4 + k];
            out[col * 4 + r] = s;
        }
}
static bool inverse_d(const double* m, double* inv) {                      // cofactor expansion in double
    double c[16];
    c[0] = m[5] * (m[10] * m[15] - m[11] * m[14]) - m[9] * (m[6] * m[15] - m[7] * m[14]) + m[13] * (m[6] * m[11] - m[7] * m[10]);
    c[1] = -(m[1] * (m[10] * m[15] - m[11] * m[14]) - m[9] * (m[2] * m[15] - m[3] * m[14]) + m[13] * (m[2] * m[11] - m[3] * m[10]));
    c[2] = m[1] * (m[6] * m[15] - m[7] * m[14]) - m[5] * (m[2] * m[15] - m[3] * m[14]) + m[13] * (m[2] * m[7] - m[3] * m[6]);
    c[3] = -(m[1] * (m[6] * m[11] - m[7] * m[10]) - m[5] * (m[2] * m[11] - m[3] * m[10]) + m[9] * (m[2] * m[7] - m[3] * m[6]));
    c[4] = -(m[4] * (m[10] * m[15] - m[11] * m[14]) - m[8] * (m[6] * m[15] - m[7] * m[14]) + m[12] * (m[6] * m[11] - m[7] * m[10]));
    c[5] = m[0] * (m[10] * m[15] - m[11] * m[14]) - m[8] * (m[2] * m[15] - m[3] * m[14]) + m[12] * (m[2] * m[11] - m[3] * m[10]);
    c[6] = -(m[0] * (m[6] * m[15] - m[7] * m[14]) - m[4] * (m[2] * m[15] - m[3] * m[14]) + m[12] * (m[2] * m[7] - m[3] * m[6]));
    c[7] = m[0] * (m[6] * m[11] - m[7] * m[10]) - m[4] * (m[2] * m[11] - m[3] * m[10]) + m[8] * (m[2] * m[7] - m[3] * m[6]);
    c[8] = m[4] * (m[9] * m[15] - m[11] * m[13]) - m[8] * (m[5] * m[15] - m[7] * m[13]) + m[12] * (m[5] * m[11] - m[7] * m[9]);
    c[9] = -(m[0] * (m[9] * m[15] - m[11] * m[13]) - m[8] * (m[1] * m[15] - m[3] * m[13]) + m[12] * (m[1] * m[11] - m[3] * m[9]));
    c[10] = m[0] * (m[5] * m[15] - m[7] * m[13]) - m[4] * (m[1] * m[15] - m[3] * m[13]) + m[12] * (m[1] * m[7] - m[3] * m[5]);
    c[11] = -(m[0] * (m[5] * m[11] - m[7] * m[9]) - m[4] * (m[1] * m[11] - m[3] * m[9]) + m[8] * (m[1] * m[7] - m[3] * m[5]));
    c[12] = -(m[4] * (m[9] * m[14] - m[10] * m[13]) - m[8] * (m[5] * m[14] - m[6] * m[13]) + m[12] * (m[5] * m[10] - m[6] * m[9]));
    c[13] = m[0] * (m[9] * m[14] - m[10] * m[13]) - m[8] * (m[1] * m[14] - m[2] * m[13]) + m[12] * (m[1] * m[10] - m[2] * m[9]);
    c[14] = -(m[0] * (m[5] * m[14] - m[6] * m[13]) - m[4] * (m[1] * m[14] - m[2] * m[13]) + m[12] * (m[1] * m[6] - m[2] * m[5]));
    c[15] = m[0] * (m[5] * m[10] - m[6] * m[9]) - m[4] * (m[1] * m[10] - m[2] * m[9]) + m[8] * (m[1] * m[6] - m[2] * m[5]);
    double det = m[0] * c[0] + m[4] * c[1] + m[8] * c[2] + m[12] * c[3];
    if (det == 0.0 || !(det == det)) return false;
    double r = 1.0 / det;
    for (int i = 0; i < 16; ++i) inv[i] = c[i] * r;
    return true;
}

// ubo0: ObjectShaderData (vert.glsl:4-9), ubo1: GlobalShaderData as std140 reads it (frag.glsl:9-14)
static bool decode_view(const unsigned char* ubo0, uint32_t n0, const unsigned char* ubo1, uint32_t n1, uint32_t W, uint32_t H, ViewConst* v) {
    float u0[48], u1[36];
    std::memset(u0, 0, sizeof u0); std::memset(u1, 0, sizeof u1);
    std::memcpy(u0, ubo0, n0 < 192u ? n0 : 192u);      // bytes past the descriptor range read as zero
    std::memcpy(u1, ubo1, n1 < 144u ? n1 : 144u);
    const float *Model = u0, *View = u0 + 16, *Proj = u0 + 32;
    double pv[16], pvi[16], md[16], mi[16];
    mat_mul_d(Proj, View, pv);
    for (int i = 0; i < 16; ++i) md[i] = Model[i];
    if (!inverse_d(pv, pvi) || !inverse_d(md, mi)) return false;
    for (int i = 0; i < 16; ++i) {
        v->inv_pv[i] = (float)pvi[i]; v->inv_model[i] = (float)mi[i];
        v->model[i] = Model[i]; v->w2l[i] = u1[i];
    }
    for (int i = 0; i < 3; ++i) v->cam[i] = u1[16 + i];
    // screen rectangle of the cube [-1,1]^3 (TestMain.cpp:94-112) under Projection * View * Model, in the pixel convention of
    // setup_ray (centre of pixel px at NDC (px + 1/2) * 2/W - 1).  The silhouette of a convex body whose corners are all in
    // front of the camera is the hull of the projected corners, so their bounding rectangle holds every covered pixel.
    v->rect[0] = 0; v->rect[1] = 0; v->rect[2] = (int)W; v->rect[3] = (int)H;
    {
        double lo[2] = {1e300, 1e300}, hi[2] = {-1e300, -1e300};
        bool all_front = true;
        for (int k = 0; k < 8 && all_front; ++k) {
            const double cx = (k & 1) ? 1.0 : -1.0, cy = (k & 2) ? 1.0 : -1.0, cz = (k & 4) ? 1.0 : -1.0;
            double w4[4], c4[4];
            for (int r = 0; r < 4; ++r) w4[r] = md[r] * cx + md[4 + r] * cy + md[8 + r] * cz + md[12 + r];
            for (int r = 0; r < 4; ++r) c4[r] = pv[r] * w4[0] + pv[4 + r] * w4[1] + pv[8 + r] * w4[2] + pv[12 + r] * w4[3];
            if (!(c4[3] > 1e-6) || !(w4[3] == w4[3])) { all_front = false; break; }
            const double px = (c4[0] / c4[3] + 1.0) * 0.5 * (double)W - 0.5, py = (c4[1] / c4[3] + 1.0) * 0.5 * (double)H - 0.5;
            if (!(px == px) || !(py == py)) { all_front = false; break; }
            lo[0] = px < lo[0] ? px : lo[0]; hi[0] = px > hi[0] ? px : hi[0];
            lo[1] = py < lo[1] ? py : lo[1]; hi[1] = py > hi[1] ? py : hi[1];
        }
        if (all_front) {
            auto clampi = [](double x, int a, int b) { return x < (double)a ? a : x > (double)b ? b : (int)x; };
            v->rect[0] = clampi(std::floor(lo[0]) - 2.0, 0, (int)W); v->rect[1] = clampi(std::floor(lo[1]) - 2.0, 0, (int)H);
            v->rect[2] = clampi(std::ceil(hi[0]) + 3.0, 0, (int)W); v->rect[3] = clampi(std::ceil(hi[1]) + 3.0, 0, (int)H);
        }
    }
    const float* MS = u1 + 20;                          // std140: MediaScroll at byte 80
    for (int col = 0; col < 3; ++col) {
        v->scroll_y[col] = MS[col * 4 + 1];             // (MS[0].y, MS[1].y, MS[2].y)  frag.glsl:67
        v->scroll_z[col] = MS[col * 4 + 2];
        v->scroll_w[col] = MS[col * 4 + 3];
    }
    return true;
}

// The fBm field's lattice hashes, once per cell (noise_device.cuh: perlin3_cells).  Eligible when every octave's lattice
// over the unit box is small enough for the float-encoded cell index (n^3 + 1.5*2^23 < 2^24) and the tables fit a budget;
// anything else marches with the per-sample hash (MARCH_FBM).  Rebuilt only when octaves / seed / frequency / lacunarity change.
static vr_status prepare_fbm_cells(vr_context* c, const vr_params* p) {
    const char* env = std::getenv("VOLRT_FBM_CELLS");        // 0 = march with the per-sample hash (A/B runs, tests)
    const bool want = !env || std::atoi(env) != 0;
    c->fbm_cells_ok = false;
    if (!want || p->fbm_octaves < 1 || p->fbm_octaves > kFbmCellMaxOctaves) return VR_OK;
    if (!(p->fbm_frequency > 0.0f) || !(p->fbm_lacunarity > 0.0f) || !(p->fbm_frequency < 1e6f) || !(p->fbm_lacunarity < 1e3f)) return VR_OK;
    // the noise is only evaluated where the edge fade is positive, i.e. maxabs(P - 1/2) < falloff1: inside the unit box
    if (!(p->fbm_falloff1 <= 0.5f) || !(p->fbm_falloff1 > p->fbm_falloff0)) return VR_OK;
    int n[kFbmCellMaxOctaves];
    size_t total = 0;
    double top = (double)p->fbm_frequency * (1.0 + 1e-5);
    for (int o = 0; o < p->fbm_octaves; ++o) {
        const double cells = std::floor(top) + 3.0;                 // lattice coordinates -1 .. floor(top) + 1
        if (cells > 160.0) return VR_OK;                            // 160^3 + 1.5*2^23 < 2^24
        n[o] = (int)cells;
        total += (size_t)n[o] * n[o] * n[o];
        top *= (double)p->fbm_lacunarity * (1.0 + 1e-5);
    }
    if (total * (size_t)kFbmCellBytes > ((size_t)256 << 20)) return VR_OK;
    const size_t need = total * (size_t)kFbmCellBytes;
    const bool same = c->d_fbm_cells && c->fbm_cells_key_oct == p->fbm_octaves && c->fbm_cells_key_seed == p->fbm_seed &&
                      c->fbm_cells_key_freq == p->fbm_frequency && c->fbm_cells_key_lac == p->fbm_lacunarity;
    if (!same) {
        if (c->foreign_stream_pending) { VR_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev1, 0)); c->foreign_stream_pending = false; }
        if (c->fbm_cells_bytes < need) {
            VR_CUDA(c, cudaStreamSynchronize(c->stream));
            if (c->d_fbm_cells) { cudaFree(c->d_fbm_cells); c->d_fbm_cells = nullptr; c->fbm_cells_bytes = 0; }
            VR_CUDA(c, cudaMalloc(&c->d_fbm_cells, need));
            c->fbm_cells_bytes = need;
        }
        unsigned char* at = c->d_fbm_cells;
        for (int o = 0; o < p->fbm_octaves; ++o) {
            VR_CUDA(c, launch_fbm_cell_table(p->fbm_seed + o, n[o], at, c->stream));     // Fractal.inl: seed + octave index
            FbmCellOctave& C = c->fbm_cells.oct[o];
            C.base_adj = (unsigned long long)(uintptr_t)at - (unsigned long long)kFbmCellBytes * 0x4B400000ull;
            C.nf = (float)n[o];
            C.n2f = (float)(n[o] * n[o]);
            C.kf = 12582912.0f + (float)(n[o] * n[o] + n[o] + 1);
            C.pad = 0.0f;
            at += (size_t)n[o] * n[o] * n[o] * (size_t)kFbmCellBytes;
        }
        VR_CUDA(c, cudaEventRecord(c->ev_params, c->stream));      // a render on a caller's stream waits for the tables
        c->fbm_cells_key_oct = p->fbm_octaves; c->fbm_cells_key_seed = p->fbm_seed;
        c->fbm_cells_key_freq = p->fbm_frequency; c->fbm_cells_key_lac = p->fbm_lacunarity;
    }
    c->fbm_cells_ok = true;
    return VR_OK;
}

// a sub-context of a group needs max_views >= the batch: its view block is grown on demand
static vr_status vr_set_params_peer(vr_context* pc, const vr_params* p) {
    const uint32_t views = p->view_count ? p->view_count : 1;
    if (views > pc->max_views) {
        cudaError_t e = cudaSetDevice(pc->device);
        if (e == cudaSuccess) e = cudaStreamSynchronize(pc->stream);
        if (e != cudaSuccess) return fail(pc, VR_ERR_CUDA, "vr_set_params (group)", e);
        if (pc->d_views) { cudaFree(pc->d_views); pc->d_views = nullptr; }
        if (pc->h_views) { cudaFreeHost(pc->h_views); pc->h_views = nullptr; }
        if ((e = cudaMalloc(&pc->d_views, sizeof(ViewConst) * views)) != cudaSuccess) return fail(pc, VR_ERR_OUT_OF_MEMORY, "cudaMalloc(views)", e);
        if ((e = cudaMallocHost(&pc->h_views, sizeof(ViewConst) * views)) != cudaSuccess) return fail(pc, VR_ERR_OUT_OF_MEMORY, "cudaMallocHost(views)", e);
        pc->view_capacity = views;
        pc->max_views = views;             // its own frames are never used (it stores into the leader's), only the view block is
    }
    return vr_set_params(pc, p);
}

extern "C" vr_status vr_set_params(vr_context* c, const vr_params* p) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    // a caller built before the VR_DENSITY_TREE fields were appended passes the shorter struct
    if (!p || p->struct_size < offsetof(vr_params, density_tree)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: null or struct_size too small");
    vr_params full;
    if (p->struct_size < sizeof(vr_params)) {
        std::memset(&full, 0, sizeof full);
        std::memcpy(&full, p, offsetof(vr_params, density_tree));
        p = &full;
    }
    if (p->max_steps == 0 || p->max_steps > (1u << 20)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: max_steps out of range");
    if (!(p->step_scale > 0.0f)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: step_scale must be > 0");
    if (p->density_mode > VR_DENSITY_TREE) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: unknown density_mode");
    if (p->density_mode == VR_DENSITY_TREE) {
        if (!p->density_tree || p->density_tree->prog.empty()) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: VR_DENSITY_TREE needs density_tree (vr_noise_tree_compile)");
        if (!(p->tree_frequency > 0.0f)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: tree_frequency must be > 0");
    }
    if (p->density_mode == VR_DENSITY_FBM && (p->fbm_octaves < 1 || p->fbm_octaves > 16)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: fbm_octaves must be in 1..16");
    if (p->light_steps > 0 && !(p->sigma_t > 0.0f)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: light_steps needs sigma_t > 0 (compositing mode)");
    if (p->light_steps > 1024) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: light_steps out of range");
    uint32_t views = p->view_count ? p->view_count : 1;
    if (views > c->max_views) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: view_count exceeds vr_create_info.max_views");
    if (views > 1 && !p->view_ubos) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: view_count > 1 needs view_ubos");
    if (!p->view_ubos && (!p->ubo0 || !p->ubo1)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: ubo0/ubo1 are null");
    VR_CUDA(c, cudaSetDevice(c->device));
    if (views == 1 && !p->view_ubos) {
        // the reference's case: one view per frame.  Its constants travel in the kernel parameters.
        if (!decode_view((const unsigned char*)p->ubo0, p->ubo0_size, (const unsigned char*)p->ubo1, p->ubo1_size, c->width, c->height, &c->view0))
            return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: Model or Projection*View is singular");
    } else {
        // the pinned staging block may still be in flight from the previous batch
        if (c->params_pending) { VR_CUDA(c, cudaEventSynchronize(c->ev_params)); c->params_pending = false; }
        for (uint32_t k = 0; k < views; ++k) {
            const unsigned char* b = (const unsigned char*)p->view_ubos + (size_t)k * 336;
            if (!decode_view(b, 192, b + 192, 144, c->width, c->height, &c->h_views[k]))
                return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_set_params: Model or Projection*View is singular");
        }
        // d_views may still be read by a march that was enqueued on a caller's stream
        if (c->foreign_stream_pending) { VR_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev1, 0)); c->foreign_stream_pending = false; }
        VR_CUDA(c, cudaMemcpyAsync(c->d_views, c->h_views, sizeof(ViewConst) * views, cudaMemcpyHostToDevice, c->stream));
        VR_CUDA(c, cudaEventRecord(c->ev_params, c->stream));
        c->params_pending = true;
        c->view0 = c->h_views[0];
    }
    if (p->density_mode == VR_DENSITY_TREE) {
        // the march reads its own copy of the program: the caller may free or recompile the handle right away
        const vr_noise_tree* t = p->density_tree;
        if (!c->d_march_tree) VR_CUDA(c, cudaMalloc(&c->d_march_tree, kTreeMaxInstr * sizeof(TreeInstr)));
        if (c->foreign_stream_pending) { VR_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev1, 0)); c->foreign_stream_pending = false; }
        VR_CUDA(c, cudaMemcpyAsync(c->d_march_tree, t->prog.data(), t->prog.size() * sizeof(TreeInstr), cudaMemcpyHostToDevice, c->stream));
        VR_CUDA(c, cudaStreamSynchronize(c->stream));          // t->prog is pageable and the caller's
        c->march_tree_n = (int)t->prog.size();
        c->march_tree_result = t->result_reg;
    }
    if (p->density_mode == VR_DENSITY_FBM) {
        vr_status st = prepare_fbm_cells(c, p);
        if (st != VR_OK) return st;
    }
    c->views_on_device = !(views == 1 && !p->view_ubos);
    c->params = *p;
    c->params.density_tree = nullptr;
    c->params.ubo0 = c->params.ubo1 = c->params.view_ubos = nullptr;   // bytes were copied; do not keep caller pointers
    c->view_count = views;
    c->has_params = true;
    // group context: every device decodes the same bytes (and builds its own fBm cell tables / tree copy); a peer holds one
    // view's constants at a time, so view batches keep their whole view block (sharded at vr_render)
    for (vr_context* pc : c->peers) {
        vr_params q = *p;
        q.struct_size = sizeof q;
        const vr_status st = vr_set_params_peer(pc, &q);
        if (st != VR_OK) { cudaSetDevice(c->device); return fail(c, st, ("vr_set_params on a device of the group: " + pc->err).c_str()); }
    }
    if (!c->peers.empty()) {
        VR_CUDA(c, cudaSetDevice(c->device));
        // a new camera keeps the band plan (the cut then follows the measured times); other changes start it over
        const vr_params& o = c->plan_params;
        if (!c->plan_valid || o.max_steps != p->max_steps || o.density_mode != p->density_mode || o.light_steps != p->light_steps ||
            o.sigma_t != p->sigma_t || o.step_scale != p->step_scale || o.view_count != p->view_count) c->plan_valid = false;
        c->plan_params = c->params;
    }
    return VR_OK;
}

// ------------------------------------------------------------------ render
static vr_status fill_march_params(vr_context* c, MarchParams& m, int& variant) {
    const vr_params& p = c->params;
    std::memset(&m, 0, sizeof m);
    m.width = (int)c->width; m.height = (int)c->height;
    m.row_begin = 0; m.row_end = (int)c->height;
    m.out = c->frame; m.row_pitch = c->row_pitch; m.view_stride = c->view_stride;
    m.out_format = (int)c->format;
    m.views = c->views_on_device ? c->d_views : nullptr;
    m.view0 = c->view0;
    m.max_steps = (int)p.max_steps;
    m.step_scale = p.step_scale; m.density = p.density; m.sample_scale = p.sample_scale;
    for (int i = 0; i < 4; ++i) { m.coord_scale[i] = p.coord_scale[i]; m.scroll_weight[i] = p.scroll_weight[i]; }
    m.light_steps = (int)p.light_steps; m.light_step_len = p.light_step_len;
    for (int i = 0; i < 3; ++i) { m.sun_dir[i] = p.sun_dir[i]; m.sun_color[i] = p.sun_color[i]; }
    m.ambient = p.ambient; m.sigma_t = p.sigma_t; m.early_term_eps = p.early_term_eps;
    m.fbm = {p.fbm_octaves, p.fbm_seed, p.fbm_frequency, p.fbm_gain, p.fbm_lacunarity, p.fbm_coverage,
             p.fbm_falloff0, p.fbm_falloff1, fbm_bounding(p.fbm_octaves, p.fbm_gain), 0.0f, 1.0f / (p.fbm_falloff1 - p.fbm_falloff0)};
    m.counters = c->d_counters;
    m.queue = c->d_queue;
    m.schedule = c->schedule;
    m.queue_chunk = c->queue_chunk;
    m.ray_split = c->ray_split;
    m.ray_split_composite = c->ray_split_composite;
    m.ray_segments = c->ray_segments;
    m.sm_count = c->sm_count;
    if (p.density_mode == VR_DENSITY_FBM) {
        if (c->lut_addr_ok < 0) {            // once per context: is the gradient table where the fBm kernels fetch it?
            unsigned int got = 0;
            cudaError_t e = march_probe_lut_address((unsigned int*)c->d_counters + 14, c->stream);
            if (e == cudaSuccess) e = cudaMemcpyAsync(&got, (unsigned int*)c->d_counters + 14, sizeof got, cudaMemcpyDeviceToHost, c->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
            if (e != cudaSuccess) return fail(c, VR_ERR_CUDA, "vr_render: gradient-table address probe", e);
            c->lut_addr_ok = got == march_lut_fixed_address() ? 1 : 0;
        }
        if (!c->lut_addr_ok)
            return fail(c, VR_ERR_UNSUPPORTED, "vr_render: this toolchain / driver places static shared memory elsewhere than the fBm kernels "
                                                "assume (rebuild libvolrt with -DVOLRT_LUT_FIXED_ADDR=<address>); VR_DENSITY_TREE renders the same field");
        variant = MARCH_FBM;
        if (c->fbm_cells_ok) { m.cells = c->fbm_cells; variant = MARCH_FBM_CELLS; }
        return VR_OK;
    }
    if (p.density_mode == VR_DENSITY_TREE) {
        m.tree_prog = c->d_march_tree; m.tree_n = c->march_tree_n; m.tree_result = c->march_tree_result;
        m.tree_seed = p.tree_seed; m.tree_freq = p.tree_frequency;
        if (!(p.fbm_falloff1 > p.fbm_falloff0)) { m.fbm.falloff0 = 0.0f; m.fbm.inv_falloff = 0.0f; }   // no fade: edge == 1
        variant = MARCH_TREE;
        return VR_OK;
    }
    if (!c->has_volume) return fail(c, VR_ERR_BAD_STATE, "vr_render: no volume uploaded (vr_upload_volume / vr_generate_volume)");
    const vr_volume_desc& d = c->vdesc;
    m.tex = c->tex; m.lin = c->lin;
    m.planar = c->planar ? 1 : 0;
    m.const_mask = c->planar ? (int)c->const_mask : 0;
    for (int k = 0; k < 4; ++k) m.const_val[k] = c->const_val[k];
    for (int k = 0; k < 4; ++k) m.tex_plane[k] = c->plane_tex[k];
    m.vw = (int)d.width; m.vh = (int)d.height; m.vd = (int)d.depth;
    m.vol_format = (int)d.format; m.address_mode = (int)d.address_mode;
    m.bricks = c->bricks;
    m.skip_empty = p.skip_empty ? 1 : 0;
    const bool tex = d.layout == VR_LAYOUT_TEXTURE;
    if (p.density_mode == VR_DENSITY_RGBA_PRODUCT) {
        if (d.format != VR_VOL_RGBA8_UNORM) return fail(c, VR_ERR_BAD_STATE, "vr_render: VR_DENSITY_RGBA_PRODUCT needs an RGBA8 volume");
        variant = tex ? MARCH_RGBA_TEX : MARCH_RGBA_LINEAR;
    } else {
        if (d.format == VR_VOL_RGBA8_UNORM) return fail(c, VR_ERR_BAD_STATE, "vr_render: VR_DENSITY_SINGLE needs a one-channel volume (R16F / R8)");
        variant = tex ? MARCH_SINGLE_TEX : MARCH_SINGLE_LINEAR;
        // TMA-staged boxes serve the reference integrator; the extensions fall back to the (bit-identical) LINEAR kernel
        if (d.layout == VR_LAYOUT_BRICK && c->has_tensor_map && !(p.sigma_t > 0.0f) && !p.skip_empty && !(p.early_term_eps > 0.0f))
            variant = MARCH_SINGLE_BRICK;
    }
    return VR_OK;
}

// the march of rows [r0,r1) x views [v0,v1) with m (already pointing at its destination) on stream s of context c's device:
// depth-segment scratch, events around the launch
static vr_status enqueue_march(vr_context* c, MarchParams& m, int variant, uint32_t r0, uint32_t r1, uint32_t v0, uint32_t v1, bool stats,
                               bool no_march, cudaStream_t s, int* n_launches) {
    if (variant == MARCH_SINGLE_BRICK && stats) variant = MARCH_SINGLE_LINEAR;     // device counters live in the generic kernel
    if (!no_march && r1 > r0 && v1 > v0) {
        // depth segments: the scratch (one float4 per pixel and piece, one arrival counter per pixel) is grown here, once
        int segs = march_pick_segments(m, variant, (int)(v1 - v0), stats);
        const size_t px = (size_t)(r1 - r0) * c->width * (v1 - v0);
        while (segs > 1 && px * segs * sizeof(float4) > ((size_t)2 << 30)) --segs;
        if (segs > 1) {
            if (c->seg_capacity < px * segs || c->seg_arrived_px < px) {
                VR_CUDA(c, cudaStreamSynchronize(c->stream));
                if (c->foreign_stream_pending) { VR_CUDA(c, cudaEventSynchronize(c->ev1)); c->foreign_stream_pending = false; }
                if (c->d_seg_partials) { cudaFree(c->d_seg_partials); c->d_seg_partials = nullptr; c->seg_capacity = 0; }
                if (c->d_seg_arrived) { cudaFree(c->d_seg_arrived); c->d_seg_arrived = nullptr; c->seg_arrived_px = 0; }
                // room for a taller band of the same frame: the cut of a group moves from frame to frame
                const size_t want_px = px + px / 4;
                VR_CUDA(c, cudaMalloc(&c->d_seg_partials, want_px * segs * sizeof(float4)));
                c->seg_capacity = want_px * segs;
                VR_CUDA(c, cudaMalloc(&c->d_seg_arrived, want_px * sizeof(unsigned int)));
                VR_CUDA(c, cudaMemsetAsync(c->d_seg_arrived, 0, want_px * sizeof(unsigned int), s));
                c->seg_arrived_px = want_px;
            }
            m.segments = segs;
            m.seg_partials = c->d_seg_partials; m.seg_arrived = c->d_seg_arrived; m.seg_capacity = c->seg_capacity;
        }
    }
    // a batch of views: the union of their screen rectangles (launch_one restricts the grid to it when that saves enough CTAs)
    if (v1 - v0 > 1 && c->views_on_device && c->h_views) {
        int u[4] = {(int)c->width, (int)c->height, 0, 0};
        for (uint32_t v = v0; v < v1; ++v) {
            const int* r = c->h_views[v].rect;
            if (r[2] <= r[0] || r[3] <= r[1]) continue;
            u[0] = r[0] < u[0] ? r[0] : u[0]; u[1] = r[1] < u[1] ? r[1] : u[1];
            u[2] = r[2] > u[2] ? r[2] : u[2]; u[3] = r[3] > u[3] ? r[3] : u[3];
        }
        if (u[2] > u[0] && u[3] > u[1]) for (int k = 0; k < 4; ++k) m.batch_rect[k] = u[k];
    }
    VR_CUDA(c, cudaEventRecord(c->ev0, s));
    if (!no_march) {
        if (variant == MARCH_SINGLE_BRICK) VR_CUDA(c, launch_march_brick(m, c->tensor_map, (int)(v1 - v0), s));
        else { m.extra_launches = n_launches; VR_CUDA(c, launch_march(m, variant, (int)(v1 - v0), stats, s)); m.extra_launches = nullptr; }
        if (r1 > r0 && v1 > v0) ++*n_launches;
    }
    VR_CUDA(c, cudaEventRecord(c->ev1, s));
    return VR_OK;
}

// ---- the frame of a group context (vr_create_info.device_count > 1), cut over its devices.
// One view: contiguous row bands of equal estimated cost.  The analytic cost (row_cost_kernel: primary steps per row) does
// not know how many steps carry a light-march, so from the second frame on every band's model cost is rescaled by the device
// time its GPU measured for it and the rows are cut again — the cut follows the scene from frame to frame.  A view batch:
// contiguous view shards.  Every device marches on its own stream straight into `frame` on devices[0] (peer access); the
// leader's stream waits for their completion events, so whatever follows on it (the copy to the caller) sees the whole frame.
static vr_status render_group(vr_context* c, char* frame, int slot, bool pipelined, cudaStream_t s, int* n_launches) {
    const uint32_t n = (uint32_t)c->peers.size() + 1, H = c->height;
    auto ctx_of = [&](uint32_t d) { return d == 0 ? c : c->peers[d - 1]; };
    std::vector<uint32_t> lo(n, 0), hi(n, 0);
    const bool by_views = c->view_count > 1;
    if (by_views) {
        const uint32_t base = c->view_count / n, extra = c->view_count % n;
        uint32_t b = 0;
        for (uint32_t d = 0; d < n; ++d) { lo[d] = b; b += base + (d < extra ? 1 : 0); hi[d] = b; }
    } else {
        if (!c->plan_valid || c->plan_cost.size() != H || c->plan_bounds.size() != n + 1 || c->plan_age >= 256) {
            c->plan_cost.assign(H, 0.0f);
            vr_status st = vr_row_cost(c, c->plan_cost.data());
            if (st != VR_OK) return st;
            c->plan_bounds.assign(n + 1, 0);
            st = vr_split_rows(c->plan_cost.data(), H, n, c->plan_bounds.data());
            if (st != VR_OK) return fail(c, st, "vr_render: cutting the row bands");
            c->plan_valid = true;
            c->plan_age = 0;
        } else {
            // feedback: the device times of the previous frame's bands, if they are all in
            std::vector<float> ms(n, 0.0f);
            bool all = true;
            for (uint32_t d = 0; d < n && all; ++d) {
                vr_context* x = ctx_of(d);
                if (c->plan_bounds[d + 1] <= c->plan_bounds[d]) continue;
                cudaSetDevice(x->device);
                if (!x->timed || cudaEventQuery(x->ev1) != cudaSuccess || cudaEventElapsedTime(&ms[d], x->ev0, x->ev1) != cudaSuccess) all = false;
            }
            cudaGetLastError();
            cudaSetDevice(c->device);
            if (all) {
                float mx = 0.0f, mn = 1e30f;
                for (uint32_t d = 0; d < n; ++d) if (ms[d] > 0.0f) { mx = std::fmax(mx, ms[d]); mn = std::fmin(mn, ms[d]); }
                if (mx > 1.03f * mn) {
                    std::vector<float> cost(c->plan_cost);
                    for (uint32_t d = 0; d < n; ++d) {
                        double sum = 0.0;
                        for (uint32_t y = c->plan_bounds[d]; y < c->plan_bounds[d + 1]; ++y) sum += cost[y];
                        if (sum > 0.0 && ms[d] > 0.0f)
                            for (uint32_t y = c->plan_bounds[d]; y < c->plan_bounds[d + 1]; ++y) cost[y] = (float)(cost[y] * ms[d] / sum);
                    }
                    std::vector<uint32_t> nb(n + 1, 0);
                    if (vr_split_rows(cost.data(), H, n, nb.data()) == VR_OK) { c->plan_bounds = nb; c->plan_cost = cost; }
                }
            }
            ++c->plan_age;
        }
        for (uint32_t d = 0; d < n; ++d) { lo[d] = c->plan_bounds[d]; hi[d] = c->plan_bounds[d + 1]; }
    }
    for (uint32_t d = 0; d < n; ++d) {
        vr_context* x = ctx_of(d);
        x->timed = false;
        if (hi[d] <= lo[d]) continue;
        VR_CUDA(c, cudaSetDevice(x->device));
        cudaStream_t xs = d == 0 ? s : x->stream;
        MarchParams m;
        int variant = 0;
        vr_status st = fill_march_params(x, m, variant);
        if (st != VR_OK) { cudaSetDevice(c->device); return d == 0 ? st : fail(c, st, ("vr_render on a device of the group: " + x->err).c_str()); }
        m.out = frame; m.row_pitch = c->row_pitch; m.view_stride = c->view_stride;
        uint32_t r0 = 0, r1 = H, v0 = 0, v1 = 1;
        if (by_views) { v0 = lo[d]; v1 = hi[d]; } else { r0 = lo[d]; r1 = hi[d]; }
        m.row_begin = (int)r0; m.row_end = (int)r1; m.view_begin = (int)v0;
        if (d > 0) {
            // the frame slot must have left for the caller (copy stream / the leader's stream) before a peer writes it again
            if (pipelined) { if (c->copy_pending[slot]) VR_CUDA(c, cudaStreamWaitEvent(xs, c->ev_copy_done[slot], 0)); }
            else VR_CUDA(c, cudaStreamWaitEvent(xs, c->ev_consumed, 0));
        }
        st = enqueue_march(x, m, variant, r0, r1, v0, v1, false, false, xs, n_launches);
        if (st != VR_OK) { cudaSetDevice(c->device); return d == 0 ? st : fail(c, st, ("vr_render on a device of the group: " + x->err).c_str()); }
        x->timed = true;
    }
    VR_CUDA(c, cudaSetDevice(c->device));
    for (vr_context* pc : c->peers) if (pc->timed) VR_CUDA(c, cudaStreamWaitEvent(s, pc->ev1, 0));
    if (!c->timed) { VR_CUDA(c, cudaEventRecord(c->ev0, s)); VR_CUDA(c, cudaEventRecord(c->ev1, s)); }   // the leader's band was empty
    c->group_timed = true;
    return VR_OK;
}

extern "C" vr_status vr_render(vr_context* c, const vr_target* t) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!c->has_params) return fail(c, VR_ERR_BAD_STATE, "vr_render: vr_set_params has not been called");
    vr_target def = {};
    def.struct_size = sizeof(vr_target);
    if (!t) t = &def;
    if (t->struct_size < sizeof(vr_target)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_render: target struct_size too small");
    if (t->memory_kind > VR_MEM_DEVICE) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_render: unknown memory_kind");
    if (t->memory_kind != VR_MEM_NONE && !t->dst) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_render: dst is null");
    const bool direct = (t->flags & VR_RENDER_DIRECT) != 0;
    if (direct && t->memory_kind != VR_MEM_DEVICE) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_render: VR_RENDER_DIRECT needs VR_MEM_DEVICE");
    uint32_t r0 = t->row_begin, r1 = t->row_end;
    if (r0 == 0 && r1 == 0) r1 = c->height;
    if (r0 > r1 || r1 > c->height) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_render: row band out of range");
    const bool no_march = (t->flags & VR_RENDER_NO_MARCH) != 0;
    uint32_t v0 = t->view_begin, v1 = t->view_end;
    if (v0 == 0 && v1 == 0) v1 = c->view_count;
    // a copy-only call may name any frame the context holds (the multi-GPU gather double-buffers GPU 0's frames)
    if (v0 > v1 || v1 > (no_march ? c->max_views : c->view_count)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_render: view range out of range");
    const size_t pitch = t->row_pitch ? t->row_pitch : c->row_pitch;
    if (pitch < c->row_pitch) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_render: row_pitch smaller than a row");

    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = t->stream ? (cudaStream_t)t->stream : c->stream;
    MarchParams m;
    int variant = 0;
    vr_status st = fill_march_params(c, m, variant);
    if (st != VR_OK) return st;
    m.row_begin = (int)r0; m.row_end = (int)r1; m.view_begin = (int)v0;
    if (direct) { m.out = t->dst; m.row_pitch = pitch; m.view_stride = pitch * c->height; }
    const bool stats = (t->flags & VR_RENDER_STATS) != 0;
    if (stats) VR_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), s));
    if (t->stream && (c->views_on_device || variant == MARCH_FBM_CELLS)) {   // the view block / the fBm cell tables were written on the context stream
        VR_CUDA(c, cudaStreamWaitEvent(s, c->ev_params, 0));
    }
    // asynchronous render into host memory: alternate between two device frames and hand the D2H copy to
    // the copy stream, so that the next frame's march overlaps it
    const bool pipelined = (t->flags & VR_RENDER_ASYNC) && t->memory_kind == VR_MEM_HOST && !direct && !t->stream && !no_march;
    // copy-only + async: the frame(s) named by the view range leave on the copy stream, behind whatever the render
    // stream holds now; the caller keeps the frame untouched until vr_wait (multi-GPU: the peers write the other frame)
    const bool copy_async = (t->flags & VR_RENDER_ASYNC) && t->memory_kind == VR_MEM_HOST && !direct && !t->stream && no_march;
    char* frame = (char*)c->frame;
    int slot = 0;
    if (pipelined) {
        if (!c->frame2) VR_CUDA(c, cudaMalloc(&c->frame2, c->view_stride * c->max_views));
        slot = c->slot;
        c->slot ^= 1;
        if (slot == 1) frame = (char*)c->frame2;
        if (c->copy_pending[slot]) VR_CUDA(c, cudaStreamWaitEvent(s, c->ev_copy_done[slot], 0));   // frame slot is free again
        m.out = frame;
    } else if (!no_march && !direct && c->copy_pending[0]) {
        // a pipelined frame may still be leaving slot 0 on the copy stream: this render writes the same memory
        VR_CUDA(c, cudaStreamWaitEvent(s, c->ev_copy_done[0], 0));
    }
    c->group_timed = false;
    // several GPUs behind this context: whole-frame renders into the context's own frame are cut over the devices
    const bool group = !c->peers.empty() && !no_march && !direct && !t->stream && !stats && r0 == 0 && r1 == c->height && v0 == 0 &&
                       v1 == c->view_count;
    int n_launches = 0;
    if (group) {
        st = render_group(c, frame, slot, pipelined, s, &n_launches);
        if (st != VR_OK) return st;
    } else {
        st = enqueue_march(c, m, variant, r0, r1, v0, v1, stats, no_march, s, &n_launches);
        if (st != VR_OK) return st;
    }
    c->timed = true;
    c->foreign_stream_pending = t->stream != nullptr;
    c->stats.kernel_launches = (uint32_t)n_launches;
    c->stats.rays = (uint64_t)(r1 - r0) * c->width * (v1 - v0);
    if (stats) VR_CUDA(c, cudaMemcpyAsync(c->h_counters, c->d_counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    else std::memset(c->h_counters, 0, 8 * sizeof(unsigned long long));

    if (t->memory_kind != VR_MEM_NONE && !direct && r1 > r0) {
        cudaStream_t cs = s;
        if (pipelined) {
            VR_CUDA(c, cudaEventRecord(c->ev_render_done[slot], s));
            VR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_render_done[slot], 0));
            cs = c->copy_stream;
        } else if (copy_async) {
            VR_CUDA(c, cudaEventRecord(c->ev_render_done[0], s));
            VR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_render_done[0], 0));
            cs = c->copy_stream;
        }
        for (uint32_t v = v0; v < v1; ++v) {
            const char* src = frame + (size_t)v * c->view_stride + (size_t)r0 * c->row_pitch;
            char* dst = (char*)t->dst + (size_t)v * pitch * c->height + (size_t)r0 * pitch;
            VR_CUDA(c, cudaMemcpy2DAsync(dst, pitch, src, c->row_pitch, c->row_pitch, r1 - r0,
                                         t->memory_kind == VR_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, cs));
        }
        if (pipelined) { VR_CUDA(c, cudaEventRecord(c->ev_copy_done[slot], c->copy_stream)); c->copy_pending[slot] = true; }
        if (c->ev_consumed && !pipelined) VR_CUDA(c, cudaEventRecord(c->ev_consumed, cs));   // the peers may write this frame again after it
    }
    if (!(t->flags & VR_RENDER_ASYNC)) VR_CUDA(c, cudaStreamSynchronize(s));
    return VR_OK;
}

extern "C" vr_status vr_wait(vr_context* c) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    for (vr_context* pc : c->peers) {
        VR_CUDA(c, cudaSetDevice(pc->device));
        VR_CUDA(c, cudaStreamSynchronize(pc->stream));
    }
    VR_CUDA(c, cudaSetDevice(c->device));
    if (c->timed) VR_CUDA(c, cudaEventSynchronize(c->ev1));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    c->copy_pending[0] = c->copy_pending[1] = false;
    return VR_OK;
}

extern "C" vr_status vr_get_stats(vr_context* c, vr_stats* out) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!out || out->struct_size < sizeof(vr_stats)) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_get_stats: null or struct_size too small");
    VR_CUDA(c, cudaSetDevice(c->device));
    if (c->timed) {
        VR_CUDA(c, cudaEventSynchronize(c->ev1));
        VR_CUDA(c, cudaEventElapsedTime(&c->stats.gpu_ms, c->ev0, c->ev1));
    }
    if (c->group_timed) {          // a frame cut over the group's devices: the slowest band is the frame's device time
        for (vr_context* pc : c->peers) {
            if (!pc->timed) continue;
            float ms = 0.0f;
            VR_CUDA(c, cudaSetDevice(pc->device));
            VR_CUDA(c, cudaEventSynchronize(pc->ev1));
            VR_CUDA(c, cudaEventElapsedTime(&ms, pc->ev0, pc->ev1));
            if (ms > c->stats.gpu_ms) c->stats.gpu_ms = ms;
        }
        VR_CUDA(c, cudaSetDevice(c->device));
    }
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    c->stats.covered = c->h_counters[0];
    c->stats.steps = c->h_counters[1];
    c->stats.early_terminated = c->h_counters[2];
    c->stats.light_samples = c->h_counters[3];
    c->stats.skipped_steps = c->h_counters[4];
    *out = c->stats;
    out->struct_size = sizeof(vr_stats);
    return VR_OK;
}

extern "C" vr_status vr_frame_info(vr_context* c, void** dev_ptr, size_t* row_pitch, size_t* view_stride) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (dev_ptr) *dev_ptr = c->frame;
    if (row_pitch) *row_pitch = c->row_pitch;
    if (view_stride) *view_stride = c->view_stride;
    return VR_OK;
}

// analytic per-row cost (sum of the rows' step counts from the ray set-up alone); cost[height]
extern "C" vr_status vr_row_cost(vr_context* c, float* cost) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!cost) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_row_cost: null output");
    if (!c->has_params) return fail(c, VR_ERR_BAD_STATE, "vr_row_cost: vr_set_params has not been called");
    VR_CUDA(c, cudaSetDevice(c->device));
    MarchParams m;
    int variant = 0;
    // the cost model only needs the ray set-up, so it works without a volume as well
    vr_status st = VR_OK;
    if (!c->has_volume && c->params.density_mode != VR_DENSITY_FBM) {
        vr_params saved = c->params;
        c->params.density_mode = VR_DENSITY_FBM;
        st = fill_march_params(c, m, variant);
        c->params = saved;
    } else st = fill_march_params(c, m, variant);
    if (st != VR_OK) return st;
    VR_CUDA(c, launch_row_cost(m, 0, c->d_row_cost, c->stream));
    VR_CUDA(c, cudaMemcpyAsync(cost, c->d_row_cost, sizeof(float) * c->height, cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    return VR_OK;
}

// contiguous row bands of (nearly) equal estimated cost; bounds[0]=0 ... bounds[parts]=height
extern "C" vr_status vr_row_bands(vr_context* c, uint32_t parts, uint32_t* bounds) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!bounds || parts == 0 || parts > 64) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_row_bands: parts must be in 1..64");
    std::vector<float> cost(c->height);
    vr_status st = vr_row_cost(c, cost.data());
    if (st != VR_OK) return st;
    return vr_split_rows(cost.data(), c->height, parts, bounds);
}

extern "C" vr_status vr_split_rows(const float* cost, uint32_t height, uint32_t parts, uint32_t* bounds) {
    if (!cost || !bounds || parts == 0 || parts > 64 || height == 0) return VR_ERR_INVALID_ARGUMENT;
    double total = 0.0;
    for (uint32_t y = 0; y < height; ++y) total += cost[y] > 0.0f ? cost[y] : 0.0f;
    bounds[0] = 0;
    double acc = 0.0;
    uint32_t k = 1;
    for (uint32_t y = 0; y < height && k < parts; ++y) {
        acc += cost[y] > 0.0f ? cost[y] : 0.0f;
        while (k < parts && acc >= total * (double)k / (double)parts && total > 0.0) bounds[k++] = y + 1;
    }
    if (total <= 0.0) for (; k < parts; ++k) bounds[k] = (uint32_t)((uint64_t)height * k / parts);   // no cost: equal heights
    while (k <= parts) bounds[k++] = height;
    return VR_OK;
}

// ------------------------------------------------------------------ NVLink peer gather across processes
extern "C" vr_status vr_frame_export(vr_context* c, void* handle64) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!handle64) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_frame_export: null handle");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    VR_CUDA(c, cudaIpcGetMemHandle(&h, c->frame));
    std::memcpy(handle64, &h, 64);
    return VR_OK;
}
extern "C" vr_status vr_frame_import(vr_context* c, const void* handle64, void** mapped) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!handle64 || !mapped) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_frame_import: null argument");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    void* p = nullptr;
    VR_CUDA(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *mapped = p;
    return VR_OK;
}
extern "C" vr_status vr_frame_unmap(vr_context* c, void* mapped) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaIpcCloseMemHandle(mapped));
    return VR_OK;
}

// ------------------------------------------------------------------ device-side completion of the gather
// The peer stores fuse the gather into the march; what remained per frame was a host barrier (every band has landed)
// before GPU 0 may copy the frame out, and another before a frame slot may be written again.  Both become counters in
// GPU 0's memory, next to the frames (same IPC mapping):
//   arrived[slot]   += 1 by every rank after its band of a frame (system-scope release behind the march's stores);
//   released[slot]  += 1 by GPU 0 after the frame has left for the host.
// GPU 0's copy stream spins on arrived[slot] (one thread), copies, releases; a rank's render stream spins on
// released[slot] before it marches into that slot again.  The host only enqueues.  Spins are bounded (~10 s) and trap.
namespace {
struct GatherMailbox { unsigned int arrived[8]; unsigned int released[8]; };
static_assert(sizeof(GatherMailbox) <= volrt::kGatherMailboxBytes, "mailbox size");
constexpr uint32_t kGatherSlots = 8;

__global__ void gather_bump_kernel(unsigned int* counter) {
    __threadfence_system();
    atomicAdd_system(counter, 1u);
}
__global__ void gather_wait_kernel(const unsigned int* counter, unsigned int target) {
    const long long t0 = clock64();
    for (;;) {
        unsigned int v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
        if ((int)(v - target) >= 0) return;
        __nanosleep(100);
        if (clock64() - t0 > 20000000000ll) __trap();        // ~10 s: a rank died or the protocol was misused — fail loudly
    }
}
GatherMailbox* mailbox_of(vr_context* c, void* frame0, uint32_t frames) {
    char* base = frame0 ? (char*)frame0 : (char*)c->frame;
    return (GatherMailbox*)(base + c->view_stride * (frame0 ? frames : c->max_views));
}
}  // namespace

extern "C" vr_status vr_gather_acquire(vr_context* c, void* frame0, uint32_t frames, uint32_t slot, uint32_t releases) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (slot >= kGatherSlots || (frame0 && (frames == 0 || slot >= frames)) || (!frame0 && slot >= c->max_views))
        return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_gather_acquire: slot out of range");
    if (releases == 0) return VR_OK;                          // first use of the slot: nothing to wait for
    VR_CUDA(c, cudaSetDevice(c->device));
    gather_wait_kernel<<<1, 1, 0, c->stream>>>(&mailbox_of(c, frame0, frames)->released[slot], releases);
    VR_CUDA(c, cudaGetLastError());
    return VR_OK;
}
extern "C" vr_status vr_gather_signal(vr_context* c, void* frame0, uint32_t frames, uint32_t slot) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (slot >= kGatherSlots || (frame0 && (frames == 0 || slot >= frames)) || (!frame0 && slot >= c->max_views))
        return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_gather_signal: slot out of range");
    VR_CUDA(c, cudaSetDevice(c->device));
    gather_bump_kernel<<<1, 1, 0, c->stream>>>(&mailbox_of(c, frame0, frames)->arrived[slot]);
    VR_CUDA(c, cudaGetLastError());
    return VR_OK;
}
extern "C" vr_status vr_gather_collect(vr_context* c, uint32_t slot, uint32_t arrivals, void* dst_host, size_t row_pitch) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (slot >= kGatherSlots || slot >= c->max_views) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_gather_collect: slot out of range");
    if (!dst_host) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_gather_collect: dst is null");
    const size_t pitch = row_pitch ? row_pitch : c->row_pitch;
    if (pitch < c->row_pitch) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_gather_collect: row_pitch smaller than a row");
    VR_CUDA(c, cudaSetDevice(c->device));
    GatherMailbox* mb = mailbox_of(c, nullptr, 0);
    gather_wait_kernel<<<1, 1, 0, c->copy_stream>>>(&mb->arrived[slot], arrivals);
    VR_CUDA(c, cudaGetLastError());
    VR_CUDA(c, cudaMemcpy2DAsync(dst_host, pitch, (const char*)c->frame + (size_t)slot * c->view_stride, c->row_pitch, c->row_pitch,
                                 c->height, cudaMemcpyDeviceToHost, c->copy_stream));
    gather_bump_kernel<<<1, 1, 0, c->copy_stream>>>(&mb->released[slot]);
    VR_CUDA(c, cudaGetLastError());
    return VR_OK;
}

// ------------------------------------------------------------------ measurement helpers
extern "C" vr_status vr_flush_l2(vr_context* c) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    for (vr_context* pc : c->peers) {                    // a group context: every device's L2
        const vr_status st = vr_flush_l2(pc);
        if (st != VR_OK) return fail(c, st, ("vr_flush_l2 on a device of the group: " + pc->err).c_str());
    }
    const size_t bytes = (size_t)256 << 20;             // 2x the 126 MB L2
    VR_CUDA(c, cudaSetDevice(c->device));
    if (!c->flush_buf) VR_CUDA(c, cudaMalloc(&c->flush_buf, bytes));
    VR_CUDA(c, cudaMemsetAsync(c->flush_buf, 0xA5, bytes, c->stream));
    // ... and read it back, so that what the write leaves in L2 is clean (peak_kernels.cu: l2_sweep_kernel)
    const char* env = std::getenv("VOLRT_FLUSH_READ");
    if (!env || std::atoi(env) != 0) VR_CUDA(c, launch_l2_sweep(c->flush_buf, bytes, (unsigned int*)c->d_counters + 15, c->sm_count, c->stream));
    return VR_OK;
}
extern "C" void* vr_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
extern "C" void vr_host_free(void* p) { if (p) cudaFreeHost(p); }

// GB/s at which the texture path streams the resident one-channel TEXTURE-layout volume out of DRAM when every texel is read
// exactly once (tex_stream_kernel; L2 flushed before every repetition): the ceiling of a march whose fetches are all first
// touches (C4 at 128 steps)
extern "C" vr_status vr_measure_tex_stream(vr_context* c, float* gbs) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!gbs) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_measure_tex_stream: null output");
    if (!c->has_volume || !c->tex || c->vdesc.format == VR_VOL_RGBA8_UNORM || c->vdesc.layout != VR_LAYOUT_TEXTURE)
        return fail(c, VR_ERR_BAD_STATE, "vr_measure_tex_stream: needs a resident one-channel volume in the TEXTURE layout");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaEvent_t e0, e1;
    VR_CUDA(c, cudaEventCreate(&e0));
    VR_CUDA(c, cudaEventCreate(&e1));
    float best = 0.0f;
    const double bytes = (double)c->lin_bytes;
    for (int rep = 0; rep < 6; ++rep) {
        vr_status st = vr_flush_l2(c);
        cudaError_t e = st == VR_OK ? cudaEventRecord(e0, c->stream) : cudaErrorUnknown;
        if (e == cudaSuccess) e = launch_tex_stream(c->tex, (int)c->vdesc.width, (int)c->vdesc.height, (int)c->vdesc.depth, (float*)c->d_counters, c->stream);
        if (e == cudaSuccess) e = cudaEventRecord(e1, c->stream);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return fail(c, VR_ERR_CUDA, "tex_stream", e); }
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        const float g = (float)(bytes / (ms * 1e-3) / 1e9);
        if (rep >= 1 && g > best) best = g;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *gbs = best;
    return VR_OK;
}

// peaks[0] = FP32 FMA TFLOP/s (FMA = 2 flop); peaks[1] = trilinear RGBA8 texture fetches, G/s
// (needs a TEXTURE-layout volume, else 0); peaks[2] = SM count; peaks[3] = SM clock MHz (max).
extern "C" vr_status vr_measure_peaks(vr_context* c, float* peaks4) {
    if (!c) return VR_ERR_INVALID_ARGUMENT;
    if (!peaks4) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_measure_peaks: null output");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaDeviceProp prop;
    VR_CUDA(c, cudaGetDeviceProperties(&prop, c->device));
    int clock_khz = 0;
    cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, c->device);
    const int sms = prop.multiProcessorCount;
    float* sink = (float*)c->d_counters;
    cudaEvent_t e0, e1;
    VR_CUDA(c, cudaEventCreate(&e0));
    VR_CUDA(c, cudaEventCreate(&e1));
    float best_fma = 0.0f, best_tex = 0.0f;
    const int blocks = sms * 8, iters = 8192;
    for (int rep = 0; rep < 6; ++rep) {               // first reps warm the clocks up
        cudaEventRecord(e0, c->stream);
        cudaError_t e = launch_fma_peak(sink, blocks, iters, c->stream);
        cudaEventRecord(e1, c->stream);
        if (e != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return fail(c, VR_ERR_CUDA, "fma_peak", e); }
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        double flop = 2.0 * 16.0 * (double)iters * 256.0 * (double)blocks;
        float tf = (float)(flop / (ms * 1e-3) / 1e12);
        if (rep >= 2 && tf > best_fma) best_fma = tf;
    }
    // the fetch peak is measured on the context's RGBA8 texture when it has one, else on a scratch 64^3 one
    cudaTextureObject_t ptex = (c->tex && c->vdesc.format == VR_VOL_RGBA8_UNORM) ? c->tex : 0;
    cudaArray_t scratch = nullptr;
    if (!ptex) {
        const int n = 64;
        std::vector<unsigned char> host((size_t)n * n * n * 4);
        for (size_t i = 0; i < host.size(); ++i) host[i] = (unsigned char)(i * 2654435761u >> 24);
        cudaChannelFormatDesc cd = cudaCreateChannelDesc<uchar4>();
        cudaExtent ext = make_cudaExtent(n, n, n);
        if (cudaMalloc3DArray(&scratch, &cd, ext) == cudaSuccess) {
            cudaMemcpy3DParms cp = {};
            cp.srcPtr = make_cudaPitchedPtr(host.data(), n * 4, n, n);
            cp.dstArray = scratch; cp.extent = ext; cp.kind = cudaMemcpyHostToDevice;
            cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = scratch;
            cudaTextureDesc td = {};
            td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeMirror;
            td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeNormalizedFloat; td.normalizedCoords = 1;
            if (cudaMemcpy3D(&cp) != cudaSuccess || cudaCreateTextureObject(&ptex, &rd, &td, nullptr) != cudaSuccess) ptex = 0;
        }
        cudaGetLastError();
    }
    if (ptex) {
        const int tblocks = sms * 16, titers = 2048;
        for (int rep = 0; rep < 5; ++rep) {
            cudaEventRecord(e0, c->stream);
            cudaError_t e = launch_tex_peak(ptex, sink, tblocks, titers, c->stream);
            cudaEventRecord(e1, c->stream);
            if (e != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return fail(c, VR_ERR_CUDA, "tex_peak", e); }
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, e0, e1);
            float g = (float)(4.0 * (double)titers * 128.0 * (double)tblocks / (ms * 1e-3) / 1e9);
            if (rep >= 1 && g > best_tex) best_tex = g;
        }
    }
    if (scratch) { if (ptex) cudaDestroyTextureObject(ptex); cudaFreeArray(scratch); }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    peaks4[0] = best_fma; peaks4[1] = best_tex; peaks4[2] = (float)sms; peaks4[3] = (float)clock_khz / 1000.0f;
    return VR_OK;
}
