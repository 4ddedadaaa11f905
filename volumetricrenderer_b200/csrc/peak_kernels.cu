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

cudaError_t launch_fma_peak(float* sink, int blocks, int iters, cudaStream_t s) {
    fma_peak_kernel<<<blocks, 256, 0, s>>>(sink, iters, 0.999f, 0.001f);
    return cudaGetLastError();
}
cudaError_t launch_tex_peak(cudaTextureObject_t tex, float* sink, int blocks, int iters, cudaStream_t s) {
    tex_peak_kernel<<<blocks, 128, 0, s>>>(tex, sink, iters);
    return cudaGetLastError();
}

}  // namespace volrt
