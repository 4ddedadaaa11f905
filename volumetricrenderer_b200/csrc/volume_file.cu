// Volume files: header + texels (+ min-max pyramid), SHA-256 of the texel section.  Host code only — the
// device side (staged upload, pyramid kernels) is in volrt_api.cu / noise_kernels.cu.  Format: include/volrt.h.
#include <cstdio>
#if defined(__x86_64__)
#include <cpuid.h>
#include <immintrin.h>
#endif
#include <cstring>
#include <string>
#include <vector>

#include "volrt_internal.h"

namespace volrt {

// ---------------------------------------------------------------- SHA-256 (FIPS 180-4), streaming
static const uint32_t kK[64] = {
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u, 0xd807aa98u, 0x12835b01u,
    0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u, 0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu,
    0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, 0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u,
    0x06ca6351u, 0x14292967u, 0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
    0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u, 0x19a4c116u, 0x1e376c08u,
    0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u, 0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u,
    0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};

static inline uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

void Sha256::reset() {
    static const uint32_t h0[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au, 0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};
    std::memcpy(h, h0, sizeof h);
    total = 0;
    fill = 0;
}

// SHA-NI path (x86 sha extensions, chosen at run time): 2-round sha256rnds2 steps with the message schedule kept in
// four xmm registers; state is held as (ABEF, CDGH) as the instruction wants it.
#if defined(__x86_64__) && defined(__GNUC__)
#define VOLRT_SHA_NI 1
__attribute__((target("sha,sse4.1,ssse3"))) static void sha256_blocks_ni(uint32_t h[8], const uint8_t* p, size_t blocks) {
    const __m128i bswap = _mm_set_epi64x(0x0c0d0e0f08090a0bULL, 0x0405060700010203ULL);
    __m128i t = _mm_loadu_si128((const __m128i*)&h[0]), s1 = _mm_loadu_si128((const __m128i*)&h[4]);
    t = _mm_shuffle_epi32(t, 0xB1);                 // CDAB
    s1 = _mm_shuffle_epi32(s1, 0x1B);               // EFGH
    __m128i s0 = _mm_alignr_epi8(t, s1, 8);         // ABEF
    s1 = _mm_blend_epi16(s1, t, 0xF0);              // CDGH
    for (; blocks; --blocks, p += 64) {
        const __m128i save0 = s0, save1 = s1;
        __m128i m[4];
        for (int i = 0; i < 4; ++i) m[i] = _mm_shuffle_epi8(_mm_loadu_si128((const __m128i*)(p + 16 * i)), bswap);
        for (int r = 0; r < 16; ++r) {
            __m128i msg = _mm_add_epi32(m[r & 3], _mm_loadu_si128((const __m128i*)&kK[4 * r]));
            s1 = _mm_sha256rnds2_epu32(s1, s0, msg);
            msg = _mm_shuffle_epi32(msg, 0x0E);
            s0 = _mm_sha256rnds2_epu32(s0, s1, msg);
            if (r < 12) {                           // schedule words 16+4r .. 19+4r into the slot just consumed
                __m128i x = _mm_sha256msg1_epu32(m[r & 3], m[(r + 1) & 3]);
                x = _mm_add_epi32(x, _mm_alignr_epi8(m[(r + 3) & 3], m[(r + 2) & 3], 4));
                m[r & 3] = _mm_sha256msg2_epu32(x, m[(r + 3) & 3]);
            }
        }
        s0 = _mm_add_epi32(s0, save0);
        s1 = _mm_add_epi32(s1, save1);
    }
    t = _mm_shuffle_epi32(s0, 0x1B);                // FEBA
    s1 = _mm_shuffle_epi32(s1, 0xB1);               // DCHG
    s0 = _mm_blend_epi16(t, s1, 0xF0);              // DCBA
    s1 = _mm_alignr_epi8(s1, t, 8);                 // HGFE
    _mm_storeu_si128((__m128i*)&h[0], s0);
    _mm_storeu_si128((__m128i*)&h[4], s1);
}
static bool have_sha_ni() {
    static const int v = [] { unsigned a, b, c, d; return __get_cpuid_count(7, 0, &a, &b, &c, &d) ? (int)((b >> 29) & 1) : 0; }();
    return v != 0;
}
#endif

void Sha256::block(const uint8_t* p) {
#ifdef VOLRT_SHA_NI
    if (have_sha_ni()) { sha256_blocks_ni(h, p, 1); return; }
#endif
    uint32_t w[64];
    for (int i = 0; i < 16; ++i) w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
    for (int i = 16; i < 64; ++i) {
        const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3);
        const uint32_t s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
        w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int i = 0; i < 64; ++i) {
        const uint32_t S1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25), ch = (e & f) ^ (~e & g);
        const uint32_t t1 = hh + S1 + ch + kK[i] + w[i];
        const uint32_t S0 = rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22), mj = (a & b) ^ (a & c) ^ (b & c);
        const uint32_t t2 = S0 + mj;
        hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

void Sha256::update(const void* data, size_t n) {
    const uint8_t* p = (const uint8_t*)data;
    total += n;
    if (fill) {
        const size_t take = n < 64 - fill ? n : 64 - fill;
        std::memcpy(buf + fill, p, take);
        fill += take; p += take; n -= take;
        if (fill == 64) { block(buf); fill = 0; }
    }
#ifdef VOLRT_SHA_NI
    if (n >= 64 && have_sha_ni()) { const size_t nb = n / 64; sha256_blocks_ni(h, p, nb); p += nb * 64; n -= nb * 64; }
#endif
    for (; n >= 64; p += 64, n -= 64) block(p);
    if (n) { std::memcpy(buf, p, n); fill = n; }
}

void Sha256::finish(uint8_t out[32]) {
    const uint64_t bits = total * 8;
    uint8_t pad[72] = {0x80};
    const size_t padn = (fill < 56 ? 56 : 120) - fill;
    for (int i = 0; i < 8; ++i) pad[padn + i] = (uint8_t)(bits >> (56 - 8 * i));
    update(pad, padn + 8);
    for (int i = 0; i < 8; ++i) { out[4 * i] = (uint8_t)(h[i] >> 24); out[4 * i + 1] = (uint8_t)(h[i] >> 16); out[4 * i + 2] = (uint8_t)(h[i] >> 8); out[4 * i + 3] = (uint8_t)h[i]; }
}

// ---------------------------------------------------------------- pyramid geometry
size_t pyramid_floats(uint32_t w, uint32_t h, uint32_t d, uint32_t format, uint32_t* levels) {
    const size_t ch = format == VR_VOL_RGBA8_UNORM ? 4 : 1;
    size_t bx = (w + 7) / 8, by = (h + 7) / 8, bz = (d + 7) / 8, total = 0;
    uint32_t n = 0;
    for (;;) {
        total += bx * by * bz * ch * 2;
        ++n;
        if (bx == 1 && by == 1 && bz == 1) break;
        bx = (bx + 1) / 2; by = (by + 1) / 2; bz = (bz + 1) / 2;
    }
    if (levels) *levels = n;
    return total;
}

// ---------------------------------------------------------------- header
static const char kMagic[8] = {'V', 'O', 'L', 'R', 'T', 'V', 'O', 'L'};

static void put32(uint8_t* p, uint32_t v) { for (int i = 0; i < 4; ++i) p[i] = (uint8_t)(v >> (8 * i)); }
static void put64(uint8_t* p, uint64_t v) { for (int i = 0; i < 8; ++i) p[i] = (uint8_t)(v >> (8 * i)); }
static uint32_t get32(const uint8_t* p) { uint32_t v = 0; for (int i = 0; i < 4; ++i) v |= (uint32_t)p[i] << (8 * i); return v; }
static uint64_t get64(const uint8_t* p) { uint64_t v = 0; for (int i = 0; i < 8; ++i) v |= (uint64_t)p[i] << (8 * i); return v; }

void VolumeFileHeader::encode(uint8_t out[kVolumeFileHeaderBytes]) const {
    std::memset(out, 0, kVolumeFileHeaderBytes);
    std::memcpy(out, kMagic, 8);
    put32(out + 8, 1);                           // version
    put32(out + 12, kVolumeFileHeaderBytes);
    put32(out + 16, width); put32(out + 20, height); put32(out + 24, depth);
    put32(out + 28, format); put32(out + 32, address_mode); put32(out + 36, pyramid_levels);
    put64(out + 40, texel_offset); put64(out + 48, texel_bytes);
    put64(out + 56, pyramid_offset); put64(out + 64, pyramid_bytes);
    std::memcpy(out + 72, sha256, 32);
}

const char* VolumeFileHeader::decode(const uint8_t in[kVolumeFileHeaderBytes], uint64_t file_size) {
    if (std::memcmp(in, kMagic, 8) != 0) return "not a volrt volume file (bad magic)";
    if (get32(in + 8) != 1) return "unsupported volume file version";
    if (get32(in + 12) != kVolumeFileHeaderBytes) return "bad header size";
    width = get32(in + 16); height = get32(in + 20); depth = get32(in + 24);
    format = get32(in + 28); address_mode = get32(in + 32); pyramid_levels = get32(in + 36);
    texel_offset = get64(in + 40); texel_bytes = get64(in + 48);
    pyramid_offset = get64(in + 56); pyramid_bytes = get64(in + 64);
    std::memcpy(sha256, in + 72, 32);
    const uint64_t tb = format == VR_VOL_RGBA8_UNORM ? 4 : format == VR_VOL_R16F ? 2 : format == VR_VOL_R8_UNORM ? 1 : 0;
    if (tb == 0) return "unknown texel format";
    if (width == 0 || height == 0 || depth == 0 || width > 4096 || height > 4096 || depth > 4096) return "dimensions out of range";
    if (address_mode > VR_ADDRESS_CLAMP_TO_EDGE) return "unknown address mode";
    if (texel_bytes != (uint64_t)width * height * depth * tb) return "texel section size does not match the dimensions";
    if (texel_offset < kVolumeFileHeaderBytes || texel_offset + texel_bytes > file_size) return "texel section outside the file (truncated?)";
    if (pyramid_levels) {
        uint32_t lv = 0;
        const uint64_t want = pyramid_floats(width, height, depth, format, &lv) * sizeof(float);
        if (lv != pyramid_levels || pyramid_bytes != want) return "pyramid section does not match the dimensions";
        if (pyramid_offset < texel_offset + texel_bytes || pyramid_offset + pyramid_bytes > file_size) return "pyramid section outside the file (truncated?)";
    } else if (pyramid_bytes) return "pyramid bytes without levels";
    return nullptr;
}

const char* read_volume_header(std::FILE* f, VolumeFileHeader* h) {
    if (std::fseek(f, 0, SEEK_END) != 0) return "seek failed";
    const long long size = ftello(f);
    if (size < (long long)kVolumeFileHeaderBytes) return "file shorter than a header";
    if (std::fseek(f, 0, SEEK_SET) != 0) return "seek failed";
    uint8_t raw[kVolumeFileHeaderBytes];
    if (std::fread(raw, 1, sizeof raw, f) != sizeof raw) return "short read (header)";
    return h->decode(raw, (uint64_t)size);
}

}  // namespace volrt

using namespace volrt;

// ---------------------------------------------------------------- C ABI, host-only entries
extern "C" vr_status vr_sha256(const void* data, size_t bytes, uint8_t out32[32]) {
    if ((!data && bytes) || !out32) return VR_ERR_INVALID_ARGUMENT;
    Sha256 s;
    s.reset();
    s.update(data, bytes);
    s.finish(out32);
    return VR_OK;
}

extern "C" size_t vr_pyramid_floats(uint32_t w, uint32_t h, uint32_t d, uint32_t format, uint32_t* levels) {
    if (w == 0 || h == 0 || d == 0) { if (levels) *levels = 0; return 0; }
    return pyramid_floats(w, h, d, format, levels);
}

extern "C" vr_status vr_volume_file_write(const char* path, const vr_volume_desc* d, const void* texels, const float* pyramid, uint32_t pyramid_levels) {
    if (!path || !d || d->struct_size < sizeof(vr_volume_desc) || !texels) return VR_ERR_INVALID_ARGUMENT;
    VolumeFileHeader h{};
    h.width = d->width; h.height = d->height; h.depth = d->depth; h.format = d->format; h.address_mode = d->address_mode;
    const uint64_t tb = d->format == VR_VOL_RGBA8_UNORM ? 4 : d->format == VR_VOL_R16F ? 2 : d->format == VR_VOL_R8_UNORM ? 1 : 0;
    if (tb == 0 || d->width == 0 || d->height == 0 || d->depth == 0 || d->width > 4096 || d->height > 4096 || d->depth > 4096 ||
        d->address_mode > VR_ADDRESS_CLAMP_TO_EDGE)
        return VR_ERR_INVALID_ARGUMENT;
    h.texel_offset = kVolumeFileHeaderBytes;
    h.texel_bytes = (uint64_t)d->width * d->height * d->depth * tb;
    if (pyramid) {
        uint32_t lv = 0;
        h.pyramid_bytes = pyramid_floats(d->width, d->height, d->depth, d->format, &lv) * sizeof(float);
        if (pyramid_levels != lv) return VR_ERR_INVALID_ARGUMENT;
        h.pyramid_levels = lv;
        h.pyramid_offset = h.texel_offset + h.texel_bytes;
    }
    Sha256 s;
    s.reset();
    s.update(texels, h.texel_bytes);
    s.finish(h.sha256);
    uint8_t raw[kVolumeFileHeaderBytes];
    h.encode(raw);
    std::FILE* f = std::fopen(path, "wb");
    if (!f) return VR_ERR_IO;
    bool ok = std::fwrite(raw, 1, sizeof raw, f) == sizeof raw && std::fwrite(texels, 1, h.texel_bytes, f) == h.texel_bytes;
    if (ok && pyramid) ok = std::fwrite(pyramid, 1, h.pyramid_bytes, f) == h.pyramid_bytes;
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) std::remove(path);
    return ok ? VR_OK : VR_ERR_IO;
}

extern "C" vr_status vr_volume_file_info_read(const char* path, vr_volume_file_info* info) {
    if (!path || !info || info->struct_size < sizeof(vr_volume_file_info)) return VR_ERR_INVALID_ARGUMENT;
    std::FILE* f = std::fopen(path, "rb");
    if (!f) return VR_ERR_IO;
    VolumeFileHeader h{};
    const char* err = read_volume_header(f, &h);
    std::fclose(f);
    if (err) return VR_ERR_IO;
    info->width = h.width; info->height = h.height; info->depth = h.depth; info->format = h.format; info->address_mode = h.address_mode;
    info->pyramid_levels = h.pyramid_levels; info->reserved = 0; info->texel_bytes = h.texel_bytes; info->pyramid_bytes = h.pyramid_bytes;
    std::memcpy(info->sha256, h.sha256, 32);
    return VR_OK;
}

extern "C" vr_status vr_volume_file_read(const char* path, void* texels, size_t texel_bytes, float* pyramid, size_t pyramid_bytes, uint32_t flags) {
    if (!path) return VR_ERR_INVALID_ARGUMENT;
    std::FILE* f = std::fopen(path, "rb");
    if (!f) return VR_ERR_IO;
    VolumeFileHeader h{};
    vr_status st = VR_OK;
    if (read_volume_header(f, &h)) st = VR_ERR_IO;
    if (st == VR_OK && texels) {
        if (texel_bytes < h.texel_bytes) st = VR_ERR_INVALID_ARGUMENT;
        else if (std::fseek(f, (long)h.texel_offset, SEEK_SET) != 0 || std::fread(texels, 1, h.texel_bytes, f) != h.texel_bytes) st = VR_ERR_IO;
        else if (flags & VR_FILE_VERIFY) {
            uint8_t got[32];
            vr_sha256(texels, h.texel_bytes, got);
            if (std::memcmp(got, h.sha256, 32) != 0) st = VR_ERR_IO;
        }
    }
    if (st == VR_OK && pyramid) {
        if (!h.pyramid_levels) st = VR_ERR_BAD_STATE;
        else if (pyramid_bytes < h.pyramid_bytes) st = VR_ERR_INVALID_ARGUMENT;
        else if (fseeko(f, (off_t)h.pyramid_offset, SEEK_SET) != 0 || std::fread(pyramid, 1, h.pyramid_bytes, f) != h.pyramid_bytes) st = VR_ERR_IO;
    }
    std::fclose(f);
    return st;
}
