// Does sm_100a's packed FP32 (fma.rn.f32x2 -> FFMA2) free issue slots?  Measures, per SM and clock:
//   A  scalar FFMA chains                       (the FP32 peak: 128 FMA/clk/SM)
//   B  FFMA2 chains                             (same FMA count in half the instructions?)
//   C  scalar FFMA + as many LOP3/IADD as FFMA  (issue-bound mix)
//   D  FFMA2 + the same integer work per FMA    (does the mix now run at the FMA rate?)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ffma2_probe ffma2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a);
    rb = *reinterpret_cast<unsigned long long*>(&b);
    rc = *reinterpret_cast<unsigned long long*>(&c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}

template <int MODE>
__global__ void __launch_bounds__(256) k(float* sink, int iters, float a, float b, int m) {
    float acc[16];
    int h[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = (float)(threadIdx.x + i);
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = threadIdx.x * (i + 3);
    const float2 a2 = {a, a}, b2 = {b, b};
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
        } else {
#pragma unroll
            for (int i = 0; i < 16; i += 2) {
                float2 v = fma2(make_float2(acc[i], acc[i + 1]), a2, b2);
                acc[i] = v.x; acc[i + 1] = v.y;
            }
        }
        if (MODE >= 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {         // 2 ALU-pipe instructions per i -> 16 per iteration
                h[i] = (h[i] ^ m) + i;
                h[i] = (h[i] >> 3) ^ h[i];
            }
        }
    }
    float s = 0.0f;
    int hs = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) hs ^= h[i];
    if (s == 12345.678f || hs == 0x12345678) sink[0] = s + hs;
}

template <int MODE>
static void run(const char* name, float* sink, int sms, float mhz) {
    const int iters = 20000, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(sink, 100, 0.999f, 0.001f, 0x5bd1e995);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(sink, iters, 0.999f, 0.001f, 0x5bd1e995);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fma = (double)blocks * 256 * 16 * iters;
    const double clk = ms * 1e-3 * mhz * 1e6;
    printf("%-28s %8.3f ms  %6.2f TFLOP/s  %6.1f FMA/clk/SM\n", name, ms, 2 * fma / ms * 1e-9, fma / clk / sms);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    float* sink;
    cudaMalloc(&sink, 4);
    const float mhz = khz / 1000.0f;
    printf("%s, %d SMs, %.0f MHz\n", p.name, p.multiProcessorCount, mhz);
    run<0>("A scalar FFMA", sink, p.multiProcessorCount, mhz);
    run<1>("B FFMA2", sink, p.multiProcessorCount, mhz);
    run<2>("C FFMA + 1 int op per FMA", sink, p.multiProcessorCount, mhz);
    run<3>("D FFMA2 + 1 int op per FMA", sink, p.multiProcessorCount, mhz);
    return 0;
}
