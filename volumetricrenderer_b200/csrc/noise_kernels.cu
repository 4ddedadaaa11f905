// libvolrt volume-bake kernels for sm_100a: the device-side replacement of the reference's
// start-up bake (source/Etna/TestMain.cpp:43-92), i.e. FastNoise2's uniform-grid driver
// (vendor/noise/include/FastNoise/Generators/Generator.inl:124-176) over the three generators of
// noise_device.cuh, the global min/max it returns (Generator.h:43-61), and the app's
// normalise / power / truncate-to-byte loop (TestMain.cpp:64-92).
//
// Layout: x fastest, then y, then z (VulkanCore.cpp:621-648), one voxel per thread, so a warp
// writes 128 contiguous bytes of a float plane / of the packed RGBA8 volume.
// The bake is ALU-bound (cellular: 27 hashed cells per voxel) and runs once per volume.
#include "noise_device.cuh"

namespace volrt {

// ---- order-preserving float <-> uint map so that min/max fold with integer atomics
__host__ __device__ __forceinline__ unsigned int ordered_bits(float f) {
#ifdef __CUDA_ARCH__
    unsigned int b = __float_as_uint(f);
#else
    unsigned int b; memcpy(&b, &f, 4);
#endif
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
void minmax_init(unsigned int* bits2) { bits2[0] = 0xffffffffu; bits2[1] = 0u; }
float minmax_decode(unsigned int o) {
    unsigned int b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    float f; memcpy(&f, &b, 4); return f;
}

template <int GEN>
__device__ __forceinline__ float eval_noise(int seed, float x, float y, float z) {
    if (GEN == VR_NOISE_CELLULAR_DISTANCE) return cellular3(seed, x, y, z);
    if (GEN == VR_NOISE_PERLIN) return perlin3(seed, x, y, z);
    return simplex3(seed, x, y, z);
}

// Generator.inl:148-167: pos = float(idx) * frequency (int -> float, then ONE multiply), x fastest.
template <int GEN, bool STORE>
__global__ void __launch_bounds__(256) noise_fill_kernel(int n, float frequency, int seed, float* __restrict__ out,
                                                         unsigned int* __restrict__ minmax_bits) {
    const size_t total = (size_t)n * n * n;
    float mn = CUDART_INF_F, mx = -CUDART_INF_F;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int x = (int)(i % n), y = (int)((i / n) % n), z = (int)(i / ((size_t)n * n));
        float v = eval_noise<GEN>(seed, __fmul_rn((float)x, frequency), __fmul_rn((float)y, frequency), __fmul_rn((float)z, frequency));
        if (STORE) out[i] = v;
        mn = fminf(mn, v); mx = fmaxf(mx, v);
    }
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ float smn[8], smx[8];
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
        if (mn <= mx) {
            atomicMin(&minmax_bits[0], ordered_bits(mn));
            atomicMax(&minmax_bits[1], ordered_bits(mx));
        }
    }
}

static int fill_grid(int n) {
    size_t total = (size_t)n * n * n;
    size_t blocks = (total + 255) / 256;
    const size_t cap = 148u * 8u * 16u;          // grid-stride beyond 16 waves of 8 CTAs/SM
    return (int)(blocks < cap ? blocks : cap);
}

template <bool STORE>
static cudaError_t fill_dispatch(const NoiseField& f, int n, float* out, unsigned int* mm, cudaStream_t s) {
    int g = fill_grid(n);
    switch (f.generator) {
        case VR_NOISE_CELLULAR_DISTANCE: noise_fill_kernel<VR_NOISE_CELLULAR_DISTANCE, STORE><<<g, 256, 0, s>>>(n, f.frequency, f.seed, out, mm); break;
        case VR_NOISE_PERLIN: noise_fill_kernel<VR_NOISE_PERLIN, STORE><<<g, 256, 0, s>>>(n, f.frequency, f.seed, out, mm); break;
        case VR_NOISE_SIMPLEX: noise_fill_kernel<VR_NOISE_SIMPLEX, STORE><<<g, 256, 0, s>>>(n, f.frequency, f.seed, out, mm); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
cudaError_t launch_noise_fill(const NoiseField& f, int n, float* out, unsigned int* mm, cudaStream_t s) {
    return fill_dispatch<true>(f, n, out, mm, s);
}
cudaError_t launch_noise_minmax_only(const NoiseField& f, int n, unsigned int* mm, cudaStream_t s) {
    return fill_dispatch<false>(f, n, nullptr, mm, s);
}

// TestMain.cpp:64-92.  The app is compiled -O0 (CMakeLists.txt:6-7) with no contraction, so
// every operator is one rounded binary32 op:  s = 1 - (v - min) * (1 / (max - min));
// s1 *= s1*s1*s1;  byte = (unsigned char)(s * 255.0f)  (truncation toward zero).
struct PackArgs { PackChannel ch[4]; float inv[4]; };

__device__ __forceinline__ unsigned int pack_one(float v, float mn, float inv, unsigned int power) {
    float s = __fsub_rn(1.0f, __fmul_rn(__fsub_rn(v, mn), inv));
    if (power == 4) s = __fmul_rn(s, __fmul_rn(__fmul_rn(s, s), s));       // :80  s *= s*s*s
    else { float b = s; for (unsigned int k = 1; k < power; ++k) s = __fmul_rn(s, b); }
    return (unsigned int)__float2int_rz(__fmul_rn(s, 255.0f)) & 0xffu;      // :84-87
}

__global__ void __launch_bounds__(256) noise_pack_kernel(const __grid_constant__ PackArgs A, long long n, uchar4* __restrict__ out) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        unsigned int b[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float v = A.ch[k].field ? __ldg(A.ch[k].field + i) : 0.0f;      // :54 zero-initialised buffer
            b[k] = pack_one(v, A.ch[k].mn, A.inv[k], A.ch[k].power);
        }
        out[i] = make_uchar4((unsigned char)b[0], (unsigned char)b[1], (unsigned char)b[2], (unsigned char)b[3]);
    }
}

cudaError_t launch_noise_pack(const PackChannel ch[4], int64_t n, uint8_t* out_rgba8, cudaStream_t s) {
    PackArgs a;
    for (int k = 0; k < 4; ++k) {
        a.ch[k] = ch[k];
        volatile float d = ch[k].mx - ch[k].mn;     // binary32, like `1 / (max - min)` at :70-73
        a.inv[k] = 1.0f / d;
    }
    size_t blocks = ((size_t)n + 255) / 256;
    const size_t cap = 148u * 8u * 16u;
    noise_pack_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, s>>>(a, (long long)n, (uchar4*)out_rgba8);
    return cudaGetLastError();
}

// C4/C5 smoke grid: the C2 density field at voxel centres, minus threshold, clamped at 0, as fp16
__global__ void __launch_bounds__(256) smoke_fill_kernel(const __grid_constant__ FbmField F, int n, __half* __restrict__ out) {
    __shared__ GradLut lut;
    grad_lut_init(&lut, threadIdx.x);
    __syncthreads();
    const size_t total = (size_t)n * n * n;
    const float fn = (float)n;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int x = (int)(i % n), y = (int)((i / n) % n), z = (int)(i / ((size_t)n * n));
        float ux = __fdiv_rn((float)x + 0.5f, fn), uy = __fdiv_rn((float)y + 0.5f, fn), uz = __fdiv_rn((float)z + 0.5f, fn);
        float rho = fmaxf(0.0f, density_fbm(&lut, F, ux, uy, uz) - F.threshold);
        out[i] = __float2half_rn(rho);
    }
}
cudaError_t launch_smoke_fill(const FbmField& f, int n, __half* out, cudaStream_t s) {
    smoke_fill_kernel<<<fill_grid(n), 256, 0, s>>>(f, n, out);
    return cudaGetLastError();
}

// fBm lattice-cell table of one octave (volrt_internal.h: FbmCells; noise_device.cuh: perlin3_cells): entry (ix,iy,iz), stored at
// ((iz+1)*n + iy+1)*n + ix+1 for lattice coordinates -1 .. n-2, holds the gradient-table offsets of the cell's eight corners —
// what perlin3_lut derives per sample from HashPrimes (Utils.inl:295-302) — as four 16-bit offsets into the pair table
// (noise_device.cuh: PairLut), the z0 plane's two corner pairs in .x and the z1 plane's in .y.
__global__ void __launch_bounds__(256) fbm_cell_table_kernel(int seed, int n, void* __restrict__ entries) {
    const int total = n * n * n, idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int ix = idx % n - 1, iy = (idx / n) % n - 1, iz = idx / (n * n) - 1;
    const int x0 = ix * kPrimeX, y0 = iy * kPrimeY, z0 = iz * kPrimeZ;
    const int x1 = x0 + kPrimeX, y1 = y0 + kPrimeY, z1 = z0 + kPrimeZ;
    const int sz0 = seed ^ z0, sz1 = seed ^ z1;
    const int t[8] = {sz0 ^ x0 ^ y0, sz0 ^ x1 ^ y0, sz0 ^ x0 ^ y1, sz0 ^ x1 ^ y1, sz1 ^ x0 ^ y0, sz1 ^ x1 ^ y0, sz1 ^ x0 ^ y1, sz1 ^ x1 ^ y1};
    uint2 e;
    e.x = pair_lut_offset(t[0], t[2]) | pair_lut_offset(t[1], t[3]) << 16;      // z0: corners x0y0 | x0y1, then x1y0 | x1y1
    e.y = pair_lut_offset(t[4], t[6]) | pair_lut_offset(t[5], t[7]) << 16;      // z1
    reinterpret_cast<uint2*>(entries)[idx] = e;
}
cudaError_t launch_fbm_cell_table(int seed, int n, void* entries, cudaStream_t s) {
    const int total = n * n * n;
    fbm_cell_table_kernel<<<(total + 255) / 256, 256, 0, s>>>(seed, n, entries);
    return cudaGetLastError();
}

// Occupancy of 8^3-voxel bricks: a trilinear fetch at a position inside brick b touches texels
// 8b-1 .. 8b+8 per axis, so the brick is "empty" iff all of those are exactly zero.  Index -1
// and N wrap to 0 and N-1 in both addressing modes, so clamping the apron is exact.
// One CTA per row of bricks along x (fixed iy, iz): its 10x10 (y,z) texel rows are read once, whole and
// coalesced (16 bytes per lane when the row pitch allows), and every non-zero texel marks its own brick plus the
// x-neighbour whose apron it sits in.  The y/z aprons make the read 100/64 of the volume; the excess hits L2.
__device__ __forceinline__ void occ_mark(unsigned char* s, int bx, int x0, unsigned m, int T) {
    // m: non-zero mask of T consecutive texels starting at x0 (T divides 8 or is a multiple of 8, x0 % T == 0)
    const int groups = T >= 8 ? T / 8 : 1, gw = T >= 8 ? 8 : T;
    for (int g = 0; g < groups; ++g) {
        const unsigned gm = T >= 8 ? (m >> (8 * g)) & 0xffu : m;
        if (!gm) continue;
        const int gx = x0 + 8 * g, brick = gx >> 3, off = gx & 7;
        s[brick] = 1;
        if ((gm & 1u) && off == 0 && brick > 0) s[brick - 1] = 1;
        if ((gm & (1u << (gw - 1))) && off + gw == 8 && brick + 1 < bx) s[brick + 1] = 1;
    }
}
__device__ __forceinline__ unsigned occ_mask16(uint4 q, int vol_format) {
    const unsigned wv[4] = {q.x, q.y, q.z, q.w};
    unsigned m = 0;
    if (vol_format == VR_VOL_R16F) {
#pragma unroll
        for (int i = 0; i < 4; ++i) m |= ((wv[i] & 0x7fffu) ? 1u : 0u) << (2 * i) | ((wv[i] & 0x7fff0000u) ? 1u : 0u) << (2 * i + 1);
    } else if (vol_format == VR_VOL_R8_UNORM) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int k = 0; k < 4; ++k) m |= ((wv[i] >> (8 * k)) & 0xffu ? 1u : 0u) << (4 * i + k);
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) m |= (wv[i] ? 1u : 0u) << i;
    }
    return m;
}
__global__ void __launch_bounds__(256) occupancy_kernel(const void* __restrict__ lin, int vol_format, int w, int h, int d,
                                                        unsigned char* __restrict__ occupied, int bx, int by) {
    __shared__ unsigned char s[512];                       // bx <= 4096 / 8
    const int iy = blockIdx.x % by, iz = blockIdx.x / by;
    for (int i = threadIdx.x; i < bx; i += blockDim.x) s[i] = 0;
    __syncthreads();
    const int tb = vol_format == VR_VOL_R16F ? 2 : vol_format == VR_VOL_R8_UNORM ? 1 : 4;
    const int T = 16 / tb;
    const bool vec = ((size_t)w * tb) % 16 == 0;          // every row starts 16-byte aligned (lin comes from cudaMalloc)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int row = warp; row < 100; row += 8) {
        const int yy = min(max(iy * 8 - 1 + row % 10, 0), h - 1);
        const int zz = min(max(iz * 8 - 1 + row / 10, 0), d - 1);
        const char* base = (const char*)lin + ((size_t)zz * h + yy) * w * tb;
        if (vec) {
            const int nvec = w / T;
            for (int v = lane; v < nvec; v += 32) {
                const unsigned m = occ_mask16(__ldg((const uint4*)base + v), vol_format);
                if (m) occ_mark(s, bx, v * T, m, T);
            }
        } else {
            for (int x = lane; x < w; x += 32) {
                unsigned nz;
                if (vol_format == VR_VOL_R16F) nz = (((const unsigned short*)base)[x] & 0x7fffu) != 0;
                else if (vol_format == VR_VOL_R8_UNORM) nz = ((const unsigned char*)base)[x] != 0;
                else nz = ((const unsigned int*)base)[x] != 0;
                if (nz) occ_mark(s, bx, x, 1u, 1);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < bx; i += blockDim.x) occupied[(size_t)blockIdx.x * bx + i] = s[i];
}
cudaError_t launch_occupancy(const void* lin, int vol_format, int w, int h, int d, int /*address_mode*/,
                             uint8_t* occupied, int bx, int by, int bz, cudaStream_t s) {
    occupancy_kernel<<<by * bz, 256, 0, s>>>(lin, vol_format, w, h, d, occupied, bx, by);
    return cudaGetLastError();
}

// ---- brick distance map: dist[b] = 0 for an occupied brick, else the Chebyshev distance (in bricks, capped at 255)
// to the nearest occupied one.  A ray sample in a brick of distance d can advance while no coordinate leaves the
// (2d-1)^3 block of empty bricks around it, so the march leaps over empty space with one byte lookup per leap.
// min-max is separable: min_q max(|dx|,|dy|,|dz|) = min_z' max(|dz|, min_y' max(|dy|, min_x' |dx|)), one pass per axis;
// a pass only looks at offsets smaller than its current best, so it costs O(distance) per brick.
template <int AXIS, bool FIRST>
__global__ void __launch_bounds__(256) brick_distance_pass_kernel(const unsigned char* __restrict__ in, unsigned char* __restrict__ out,
                                                                  int bx, int by, int bz) {
    const int total = bx * by * bz, idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int ix = idx % bx, iy = (idx / bx) % by, iz = idx / (bx * by);
    const int n = AXIS == 0 ? bx : AXIS == 1 ? by : bz, pos = AXIS == 0 ? ix : AXIS == 1 ? iy : iz;
    const int stride = AXIS == 0 ? 1 : AXIS == 1 ? bx : bx * by;
    auto val = [&](int q) -> int {
        const int v = in[idx + (q - pos) * stride];
        return FIRST ? (v ? 0 : 255) : v;
    };
    int best = val(pos);
    for (int k = 1; k < best; ++k) {
        if (pos - k >= 0) best = min(best, max(k, val(pos - k)));
        if (pos + k < n) best = min(best, max(k, val(pos + k)));
    }
    out[idx] = (unsigned char)best;
}
cudaError_t launch_brick_distance(const uint8_t* occupied, uint8_t* dist, uint8_t* tmp, int bx, int by, int bz, cudaStream_t s) {
    const int total = bx * by * bz, grid = (total + 255) / 256;
    brick_distance_pass_kernel<0, true><<<grid, 256, 0, s>>>(occupied, dist, bx, by, bz);
    brick_distance_pass_kernel<1, false><<<grid, 256, 0, s>>>(dist, tmp, bx, by, bz);
    brick_distance_pass_kernel<2, false><<<grid, 256, 0, s>>>(tmp, dist, bx, by, bz);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) brick_pad_kernel(const unsigned char* __restrict__ dist, unsigned char* __restrict__ pad, int bx, int by, int bz) {
    const int px = bx + 2, py = by + 2, total = px * py * (bz + 2), idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int ix = min(max(idx % px - 1, 0), bx - 1), iy = min(max((idx / px) % py - 1, 0), by - 1), iz = min(max(idx / (px * py) - 1, 0), bz - 1);
    pad[idx] = dist[((size_t)iz * by + iy) * bx + ix];
}
cudaError_t launch_brick_pad(const uint8_t* dist, uint8_t* pad, int bx, int by, int bz, cudaStream_t s) {
    const int total = (bx + 2) * (by + 2) * (bz + 2);
    brick_pad_kernel<<<(total + 255) / 256, 256, 0, s>>>(dist, pad, bx, by, bz);
    return cudaGetLastError();
}

// ---- min-max pyramid (include/volrt.h): level 0 = per 8^3 brick incl. the one-texel apron, then 2x2x2 reductions.
// Same traversal as the occupancy kernel: one CTA per brick row, whole texel rows read coalesced; the 8 lanes that
// share a brick fold their values with REDUX on order-preserving bits, and three lanes per group (the leader, and
// the two texels that sit in a neighbour's apron) touch shared memory.  NaN texels count for the max only (as the
// canonical quiet NaN, which is the largest ordered value), so a NaN anywhere keeps its bricks "occupied".
__device__ __forceinline__ float decode_texel(const void* row, int vol_format, int x, int c) {
    if (vol_format == VR_VOL_R16F) return __half2float(((const __half*)row)[x]);
    if (vol_format == VR_VOL_R8_UNORM) return __fdiv_rn((float)((const unsigned char*)row)[x], 255.0f);
    return __fdiv_rn((float)((const unsigned char*)row)[4 * x + c], 255.0f);
}
template <int CH>
__global__ void __launch_bounds__(256) pyramid_base_kernel(const void* __restrict__ lin, int vol_format, int w, int h, int d,
                                                           float* __restrict__ out, int bx, int by) {
    __shared__ unsigned int smin[512 * CH], smax[512 * CH];
    const int iy = blockIdx.x % by, iz = blockIdx.x / by;
    for (int i = threadIdx.x; i < bx * CH; i += blockDim.x) { smin[i] = 0xffffffffu; smax[i] = 0u; }
    __syncthreads();
    const int tb = vol_format == VR_VOL_R16F ? 2 : vol_format == VR_VOL_R8_UNORM ? 1 : 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned int group = 0xffu << (lane & 24);                 // the 8 lanes of one brick
    for (int row = warp; row < 100; row += 8) {
        const int yy = min(max(iy * 8 - 1 + row % 10, 0), h - 1);
        const int zz = min(max(iz * 8 - 1 + row / 10, 0), d - 1);
        const char* base = (const char*)lin + ((size_t)zz * h + yy) * w * tb;
        for (int x0 = 0; x0 < w; x0 += 32) {
            const int x = x0 + lane, brick = x >> 3;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                unsigned int lo = 0xffffffffu, hi = 0u;
                if (x < w) {
                    const float v = decode_texel(base, vol_format, x, c);
                    if (v != v) hi = ordered_bits(__int_as_float(0x7fc00000));
                    else lo = hi = ordered_bits(v);
                }
                const unsigned int gmin = __reduce_min_sync(group, lo), gmax = __reduce_max_sync(group, hi);
                if (x < w) {
                    if ((lane & 7) == 0) {
                        atomicMin(&smin[brick * CH + c], gmin); atomicMax(&smax[brick * CH + c], gmax);
                        if (brick > 0) { atomicMin(&smin[(brick - 1) * CH + c], lo); atomicMax(&smax[(brick - 1) * CH + c], hi); }
                    } else if ((lane & 7) == 7 && brick + 1 < bx) {
                        atomicMin(&smin[(brick + 1) * CH + c], lo); atomicMax(&smax[(brick + 1) * CH + c], hi);
                    }
                }
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < bx * CH; i += blockDim.x) {
        const unsigned int a = smin[i], b = smax[i];
        const unsigned int ab = (a & 0x80000000u) ? (a & 0x7fffffffu) : ~a, bb = (b & 0x80000000u) ? (b & 0x7fffffffu) : ~b;
        float2 mm = make_float2(a == 0xffffffffu ? CUDART_INF_F : __uint_as_float(ab), __uint_as_float(bb));
        reinterpret_cast<float2*>(out)[(size_t)blockIdx.x * bx * CH + i] = mm;
    }
}
cudaError_t launch_pyramid_base(const void* lin, int vol_format, int w, int h, int d, float* out, int bx, int by, int bz, cudaStream_t s) {
    if (vol_format == VR_VOL_RGBA8_UNORM) pyramid_base_kernel<4><<<by * bz, 256, 0, s>>>(lin, vol_format, w, h, d, out, bx, by);
    else pyramid_base_kernel<1><<<by * bz, 256, 0, s>>>(lin, vol_format, w, h, d, out, bx, by);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) pyramid_reduce_kernel(const float2* __restrict__ in, float2* __restrict__ out, int ch,
                                                             int bx, int by, int bz, int ox, int oy, int oz) {
    const int total = ox * oy * oz * ch, idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int c = idx % ch, e = idx / ch, x = e % ox, y = (e / ox) % oy, z = e / (ox * oy);
    unsigned int lo = 0xffffffffu, hi = 0u;
    for (int dz = 0; dz < 2; ++dz)
        for (int dy = 0; dy < 2; ++dy)
            for (int dx = 0; dx < 2; ++dx) {
                const int sx = 2 * x + dx, sy = 2 * y + dy, sz = 2 * z + dz;
                if (sx >= bx || sy >= by || sz >= bz) continue;
                const float2 v = in[(((size_t)sz * by + sy) * bx + sx) * ch + c];
                lo = min(lo, ordered_bits(v.x));
                hi = max(hi, ordered_bits(v.y != v.y ? __int_as_float(0x7fc00000) : v.y));
            }
    const unsigned int ab = (lo & 0x80000000u) ? (lo & 0x7fffffffu) : ~lo, bb = (hi & 0x80000000u) ? (hi & 0x7fffffffu) : ~hi;
    out[idx] = make_float2(__uint_as_float(ab), __uint_as_float(bb));
}
cudaError_t launch_pyramid_reduce(const float* in, float* out, int channels, int bx, int by, int bz, cudaStream_t s) {
    const int ox = (bx + 1) / 2, oy = (by + 1) / 2, oz = (bz + 1) / 2, total = ox * oy * oz * channels;
    pyramid_reduce_kernel<<<(total + 255) / 256, 256, 0, s>>>((const float2*)in, (float2*)out, channels, bx, by, bz, ox, oy, oz);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) occupancy_from_pyramid_kernel(const float2* __restrict__ level0, int ch, unsigned char* __restrict__ occupied, int nbricks) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbricks) return;
    int any = 0;
    for (int c = 0; c < ch; ++c) { const float2 v = level0[(size_t)b * ch + c]; any |= (v.x != 0.0f) || (v.y != 0.0f); }
    occupied[b] = any ? 1 : 0;
}
cudaError_t launch_occupancy_from_pyramid(const float* level0, int channels, uint8_t* occupied, int nbricks, cudaStream_t s) {
    occupancy_from_pyramid_kernel<<<(nbricks + 255) / 256, 256, 0, s>>>((const float2*)level0, channels, occupied, nbricks);
    return cudaGetLastError();
}

// RGBA8 (x fastest) -> four tightly packed R8 planes; 4 voxels per thread so that every plane store is a 32-bit word.
// mm (may be null): per-channel min (mm[0..3], initialised >= 255) and max (mm[4..7], initialised 0) of the volume, folded in
// here because this kernel reads every texel anyway (as a separate pass it cost the 128^3 bake 0.05 ms): byte-wise SIMD
// min/max on the packed texels (VMNMX), one atomic pair per channel and CTA.
__global__ void __launch_bounds__(256) deinterleave_rgba8_kernel(const uint4* __restrict__ rgba, unsigned char* __restrict__ planes, size_t voxels,
                                                                 unsigned int* __restrict__ mm) {
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int mn = 0xffffffffu, mx = 0u;
    // the word stores below land on planes + c*voxels + 4q: 4-byte aligned for every plane only when voxels % 4 == 0
    // (a 5^3 or 10x13x27 volume would raise a sticky misaligned-address fault); other volumes go texel by texel
    if ((voxels & 3) == 0 && q * 4 + 3 < voxels) {
        const uint4 v = __ldg(rgba + q);
        mn = __vminu4(__vminu4(v.x, v.y), __vminu4(v.z, v.w));
        mx = __vmaxu4(__vmaxu4(v.x, v.y), __vmaxu4(v.z, v.w));
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const unsigned sel = 0x0040u + 0x0011u * c;              // byte c of the first word, byte c of the second
            const unsigned w = (__byte_perm(v.x, v.y, sel) & 0xffffu) | (__byte_perm(v.z, v.w, sel) << 16);
            reinterpret_cast<unsigned*>(planes + (size_t)c * voxels)[q] = w;
        }
    } else {
        const size_t end = (voxels & 3) == 0 ? voxels : (q * 4 + 4 < voxels ? q * 4 + 4 : voxels);
        for (size_t i = q * 4; i < end; ++i) {
            const unsigned int t = reinterpret_cast<const unsigned int*>(rgba)[i];
            mn = __vminu4(mn, t);
            mx = __vmaxu4(mx, t);
            for (int c = 0; c < 4; ++c) planes[(size_t)c * voxels + i] = reinterpret_cast<const unsigned char*>(rgba)[i * 4 + c];
        }
    }
    if (mm) {
        __shared__ unsigned int s_mn[8], s_mx[8];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn = __vminu4(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = __vmaxu4(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if ((threadIdx.x & 31) == 0) { s_mn[threadIdx.x >> 5] = mn; s_mx[threadIdx.x >> 5] = mx; }
        __syncthreads();
        if (threadIdx.x < 4) {
            const int c = threadIdx.x;
            unsigned int lo = 255u, hi = 0u;
            for (int w = 0; w < 8; ++w) { lo = min(lo, (s_mn[w] >> (8 * c)) & 0xffu); hi = max(hi, (s_mx[w] >> (8 * c)) & 0xffu); }
            atomicMin(mm + c, lo);
            atomicMax(mm + 4 + c, hi);
        }
    }
}
cudaError_t launch_deinterleave_rgba8(const uint8_t* rgba, uint8_t* planes4, size_t voxels, unsigned int* mm8, cudaStream_t s) {
    const size_t quads = (voxels + 3) / 4;
    deinterleave_rgba8_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, s>>>((const uint4*)rgba, planes4, voxels, mm8);
    return cudaGetLastError();
}

}  // namespace volrt
