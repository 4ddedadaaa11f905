// Roofline denominators that MEASURED_PEAKS.json does not hold: the march is neither HBM- nor
// tensor-bound, so the FP32-FMA issue peak and the texture-unit trilinear fetch peak are measured on
// the box, in the same process as the numbers they normalise.
#include "volrt_internal.h"

namespace volrt {

// 16 independent FMA chains per thread: no dependent-issue stalls, no memory traffic.
__global__ void __launch_bounds__(256) fma_peak_kernel(float* __restrict__ sink, int iters, float a, float b) {
    float acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = (float)(threadIdx.x + k);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) acc[k] = fmaf(acc[k], a, b);
    }
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += acc[k];
    if (s == 12345.678f) sink[0] = s;        // never true; keeps the chains alive
}

// Trilinear fetch rate of an L2/L1-resident RGBA8 volume with warp-coherent coordinates
// (an 8x4 pixel tile stepping along nearly parallel rays), 4 independent fetches per iteration.
__global__ void __launch_bounds__(128) tex_peak_kernel(cudaTextureObject_t tex, float* __restrict__ sink, int iters) {
    const int lane = threadIdx.x & 31;
    float x = (float)(lane & 7) * (1.0f / 1024.0f) + (float)(blockIdx.x % 97) * (1.0f / 97.0f);
    float y = (float)(lane >> 3) * (1.0f / 1024.0f) + (float)(threadIdx.x >> 5) * 0.11f;
    float z = 0.0f, s = 0.0f;
    for (int i = 0; i < iters; ++i) {
        float4 a = tex3D<float4>(tex, x, y, z);
        float4 b = tex3D<float4>(tex, x * 0.8f, y * 0.8f, z * 0.8f);
        float4 c = tex3D<float4>(tex, x * 0.75f, y * 0.75f, z * 0.75f);
        float4 d = tex3D<float4>(tex, x * 0.7f, y * 0.7f, z * 0.7f);
        s += a.x * b.y * (c.z + d.w);
        x += 0.0021f; y += 0.0017f; z += 0.0023f;
    }
    if (s == 12345.678f) sink[0] = s;
}

// What the texture path can stream out of DRAM: every texel of a one-channel 3-D texture read exactly once, by trilinear
// fetches placed between texels 2i and 2i+1 on every axis (one fetch filters a 2x2x2 block: 1/8 fetch per texel, so the
// fetch rate — 16.8 M fetches for 512^3 — is not the limit; at the unit's peak they take ~30 us, the 268 MB take 41 us at
// the measured HBM rate).  SHAPE = how a warp's 32 fetches are laid out, PER_THREAD = independent fetches per thread along z.
// This is the ceiling of the C4 march at 128 steps, where nearly every fetch is a first touch: DESIGN.md §7.
template <int PER_THREAD, int SHAPE>
__global__ void __launch_bounds__(256) tex_stream_kernel(cudaTextureObject_t tex, int nx, int ny, int nz, float* __restrict__ sink) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int bx, by, bz;                                                            // fetch coordinates (units of 2 texels)
    if (SHAPE == 0) {            // warp = 8x2x2 fetches, CTA = 16x8x2
        bx = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
        by = blockIdx.y * 8 + ((warp >> 1) & 3) * 2 + ((lane >> 3) & 1);
        bz = (blockIdx.z * 2 + (lane >> 4)) * PER_THREAD;
    } else if (SHAPE == 1) {     // warp = 32x1x1 fetches, CTA = 32x8x1
        bx = blockIdx.x * 32 + lane;
        by = blockIdx.y * 8 + warp;
        bz = blockIdx.z * PER_THREAD;
    } else {                     // warp = 4x4x2 fetches, CTA = 8x8x4
        bx = blockIdx.x * 8 + (warp & 1) * 4 + (lane & 3);
        by = blockIdx.y * 8 + ((warp >> 1) & 1) * 4 + ((lane >> 2) & 3);
        bz = (blockIdx.z * 4 + (warp >> 2) * 2 + (lane >> 4)) * PER_THREAD;
    }
    float s = 0.0f;
    if (2 * bx < nx && 2 * by < ny) {
        float v[PER_THREAD];
#pragma unroll
        for (int k = 0; k < PER_THREAD; ++k) {
            const int z = bz + k;
            v[k] = 2 * z < nz ? tex3D<float>(tex, (float)(2 * bx + 1) / (float)nx, (float)(2 * by + 1) / (float)ny, (float)(2 * z + 1) / (float)nz) : 0.0f;
        }
#pragma unroll
        for (int k = 0; k < PER_THREAD; ++k) s += v[k];
    }
    if (s == 12345.678f) sink[0] = s;
}
// Measured on B200 over the 512^3 fp16 grid (268 MB, L2 flushed), GB/s: warp = 8x2x2 fetches with 1 / 4 / 8 per thread along z
// 3449 / 3547 / 3542; warp = 32x1x1 with 1 / 4: 3451 / 3591; warp = 4x4x2 with 1 / 4: 3544 / 3644.  Whatever the walk, the
// texture path streams first-touch texels out of DRAM at 3.45-3.65 TB/s behind a write-only L2 flush and at 4.30-4.37 TB/s = 66 % of the
// 6551 GB/s a copy reaches behind a flush that leaves L2 clean (l2_sweep_kernel below).
cudaError_t launch_tex_stream(cudaTextureObject_t tex, int nx, int ny, int nz, float* sink, cudaStream_t s) {
    const int fx = nx / 2, fy = ny / 2, fz = nz / 2;
    tex_stream_kernel<4, 2><<<dim3((fx + 7) / 8, (fy + 7) / 8, (fz + 15) / 16), 256, 0, s>>>(tex, nx, ny, nz, sink);
    return cudaGetLastError();
}

// Second half of vr_flush_l2: read the buffer the memset has just written.  The write alone leaves L2 full of DIRTY lines, and
// the kernel that is then timed pays for their write-back while it streams its own data in (a 512^3 fp16 march: 92 us against the
// 80 us ncu measures for it behind a clean flush).  After this sweep L2 holds clean lines of an unrelated buffer: cold and quiet.
__global__ void __launch_bounds__(256) l2_sweep_kernel(const uint4* __restrict__ buf, size_t n16, unsigned int* __restrict__ sink) {
    unsigned int acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = __ldcg(buf + i);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    if (acc == 0x12345679u) *sink = acc;          // never (the buffer holds 0xA5 bytes): keeps the loads
}
cudaError_t launch_l2_sweep(const void* buf, size_t bytes, unsigned int* sink, int sms, cudaStream_t s) {
    l2_sweep_kernel<<<(sms > 0 ? sms : 148) * 8, 256, 0, s>>>((const uint4*)buf, bytes / 16, sink);
    return cudaGetLastError();
}

cudaError_t launch_fma_peak(float* sink, int blocks, int iters, cudaStream_t s) {
    fma_peak_kernel<<<blocks, 256, 0, s>>>(sink, iters, 0.999f, 0.001f);
    return cudaGetLastError();
}
cudaError_t launch_tex_peak(cudaTextureObject_t tex, float* sink, int blocks, int iters, cudaStream_t s) {
    tex_peak_kernel<<<blocks, 128, 0, s>>>(tex, sink, iters);
    return cudaGetLastError();
}

}  // namespace volrt
