// vr_noise_tree_grid_static: vr_noise_tree_grid for callers that may not allocate (code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The existing entry point keeps its scratch in the context and grows it with cudaMalloc on demand (program
// upload buffer, the two min/max words, the output plane); here the same three pieces live in ONE file-scope device array
// of a fixed, modest size, so the call allocates nothing — no cudaMalloc / cudaFree / cudaMallocHost, no context.
//
// Every GPU that loads libvolrt.so holds the array (kStaticScratchBytes, ~1.03 MiB) until the process exits.  All calls on
// a device share it and it keeps the previous call's contents: the min/max words are re-initialised by every call, the
// program and the output plane are overwritten up to the size the call uses.  Calls are serialised by a mutex and complete
// (stream synchronised) before they return, so two calls never overlap on the GPU; if that were relaxed — an asynchronous
// variant, or the mutex removed — overlapping calls would race on the array.
#include "volrt_internal.h"
#include <cstring>
#include <mutex>

using namespace volrt;

namespace {

constexpr size_t kProgBytes = (size_t)kTreeMaxInstr * sizeof(TreeInstr);        // 512 x 64 B = 32 KiB
constexpr size_t kMinMaxOffset = kProgBytes;                                    // 2 words, in a 256-byte slot of their own
constexpr size_t kPlaneOffset = kProgBytes + 256;
constexpr size_t kStaticScratchBytes = kPlaneOffset + (size_t)VR_STATIC_GRID_MAX_POINTS * sizeof(float);
static_assert(kProgBytes % 256 == 0, "the sections of the scratch array stay 256-byte aligned");

// In host code this name is a host-side stand-in, NOT a device address: the address comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_addr[kMaxDevices] = {};      // per device, resolved once with that device current

vr_status fail_msg(char* err, size_t cap, vr_status st, const char* what, cudaError_t e = cudaSuccess) {
    if (err && cap) {
        if (e != cudaSuccess) snprintf(err, cap, "%s: %s — %s", what, cudaGetErrorName(e), cudaGetErrorString(e));
        else snprintf(err, cap, "%s", what);
    }
    return st;
}

}  // namespace

extern "C" vr_status vr_noise_tree_grid_static(int32_t device, const vr_noise_tree* t, int32_t x0, int32_t y0, int32_t z0, uint32_t nx,
                                               uint32_t ny, uint32_t nz, float frequency, int32_t seed, float* out_host, float minmax[2],
                                               void* stream, char* err, size_t err_cap) {
    if (!t || t->prog.empty()) return fail_msg(err, err_cap, VR_ERR_INVALID_ARGUMENT, "vr_noise_tree_grid_static: no compiled tree");
    if (t->prog.size() > (size_t)kTreeMaxInstr) return fail_msg(err, err_cap, VR_ERR_INVALID_ARGUMENT, "vr_noise_tree_grid_static: program longer than the static buffer");
    if (nx == 0 || ny == 0 || nz == 0) return fail_msg(err, err_cap, VR_ERR_INVALID_ARGUMENT, "vr_noise_tree_grid_static: empty grid");
    const unsigned long long points = (unsigned long long)nx * ny * nz;
    if (nx > 4096 || ny > 4096 || nz > 4096 || points > VR_STATIC_GRID_MAX_POINTS)
        return fail_msg(err, err_cap, VR_ERR_INVALID_ARGUMENT,
                        "vr_noise_tree_grid_static: grid exceeds VR_STATIC_GRID_MAX_POINTS (262144 points = 64^3, the size of the static scratch); "
                        "use vr_noise_tree_grid for larger grids");
    if (device < 0 || device >= kMaxDevices) return fail_msg(err, err_cap, VR_ERR_INVALID_ARGUMENT, "vr_noise_tree_grid_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_mu);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail_msg(err, err_cap, VR_ERR_CUDA, "vr_noise_tree_grid_static: cudaSetDevice (libvolrt has no CPU path)", e);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);       // with `device` current: each device has its own copy
        if (e != cudaSuccess || !p) return fail_msg(err, err_cap, VR_ERR_CUDA, "vr_noise_tree_grid_static: cudaGetSymbolAddress", e);
        g_scratch_addr[device] = (unsigned char*)p;
    }
    unsigned char* base = g_scratch_addr[device];
    TreeInstr* d_prog = (TreeInstr*)base;
    unsigned int* d_mm = (unsigned int*)(base + kMinMaxOffset);
    float* d_out = (float*)(base + kPlaneOffset);
    cudaStream_t s = (cudaStream_t)stream;
    unsigned int h_mm[2];
    minmax_init(h_mm);                                         // the words hold the last call's range: start over
    const size_t bytes = (size_t)points * sizeof(float);
    e = cudaMemcpyAsync(d_prog, t->prog.data(), t->prog.size() * sizeof(TreeInstr), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_mm, h_mm, sizeof h_mm, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = launch_tree_grid(d_prog, (int)t->prog.size(), t->result_reg, x0, y0, z0, (int)nx, (int)ny, (int)nz, frequency, seed, d_out, d_mm, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_mm, d_mm, sizeof h_mm, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && out_host) e = cudaMemcpyAsync(out_host, d_out, bytes, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return fail_msg(err, err_cap, VR_ERR_CUDA, "vr_noise_tree_grid_static", e);
    if (minmax) { minmax[0] = minmax_decode(h_mm[0]); minmax[1] = minmax_decode(h_mm[1]); }
    return VR_OK;
}
