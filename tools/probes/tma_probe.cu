// minimal check of the 3-D tiled TMA load used by march_brick_kernel (build: nvcc -arch=sm_100a -o tma_probe tma_probe.cu)
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
constexpr int BX = 32, BY = 24, BZ = 24, BYTES = BX * BY * BZ * 2;
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(const __grid_constant__ CUtensorMap tmap, int x, int y, int z, float* out) {
    extern __shared__ __align__(128) unsigned char dyn[];
    __shared__ __align__(8) unsigned long long bar;
    __half* box = (__half*)dyn;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(BYTES) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(s32(box)), "l"(&tmap), "r"(x), "r"(y), "r"(z), "r"(s32(&bar)) : "memory");
    }
    unsigned ok = 0;
    while (!ok) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(s32(&bar)), "r"(0u) : "memory");
    float s = 0;
    for (int i = threadIdx.x; i < BX * BY * BZ; i += blockDim.x) s += __half2float(box[i]);
    atomicAdd(out, s);
    if (threadIdx.x == 0) out[1] = __half2float(box[(1 * BY + 2) * BX + 3]);
}
int main() {
    const int N = 64;
    std::vector<__half> h((size_t)N * N * N);
    for (int zz = 0; zz < N; ++zz) for (int yy = 0; yy < N; ++yy) for (int xx = 0; xx < N; ++xx) h[((size_t)zz * N + yy) * N + xx] = __float2half((float)(xx + 2 * yy + 3 * zz) / 64.f);
    __half* d; cudaMalloc(&d, h.size() * 2); cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
    printf("entry %p qr %d\n", fn, (int)qr);
    CUtensorMap tm;
    cuuint64_t dims[3] = {N, N, N}, strides[2] = {N * 2, (cuuint64_t)N * N * 2};
    cuuint32_t box[3] = {BX, BY, BZ}, es[3] = {1, 1, 1};
    CUresult r = ((Enc)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d\n", (int)r);
    float* out; cudaMalloc(&out, 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, BYTES);
    int cs[3][3] = {{8, 7, 9}, {-8, 50, 60}, {56, -3, 63}};
    for (auto& c : cs) {
        cudaMemset(out, 0, 8);
        k<<<1, 128, BYTES>>>(tm, c[0], c[1], c[2], out);
        cudaError_t e = cudaDeviceSynchronize();
        float ho[2]; cudaMemcpy(ho, out, 8, cudaMemcpyDeviceToHost);
        double want = 0; for (int zz = 0; zz < BZ; ++zz) for (int yy = 0; yy < BY; ++yy) for (int xx = 0; xx < BX; ++xx) { int gx = c[0] + xx, gy = c[1] + yy, gz = c[2] + zz; if (gx >= 0 && gx < N && gy >= 0 && gy < N && gz >= 0 && gz < N) want += __half2float(h[((size_t)gz * N + gy) * N + gx]); }
        printf("coords %d %d %d: %s sum %.1f want %.1f elem %.4f\n", c[0], c[1], c[2], cudaGetErrorString(e), ho[0], want, ho[1]);
    }
    return 0;
}
