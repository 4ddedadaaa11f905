"""ctypes binding of the CPU ORACLE (oracle/libvolrt_oracle.so) and, when built,
of the reference's own FastNoise2 (oracle/_ref/libvolrt_ref.so).

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py; never by volumetricrenderer_b200.
See oracle/volrt_oracle.h for what each function restates (reference file:line).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(HERE, "libvolrt_oracle.so")
_REF_SO = os.path.join(HERE, "_ref", "libvolrt_ref.so")

VOL_RGBA8, VOL_R16F, VOL_R8, VOL_R32F = 0, 1, 2, 3
DENSITY_RGBA_PRODUCT, DENSITY_SINGLE, DENSITY_FBM = 0, 1, 2
OUT_RGBA8_UNORM, OUT_RGBA8_SRGB, OUT_RGBA16F = 0, 1, 2
NOISE_CELLULAR, NOISE_PERLIN, NOISE_SIMPLEX = 0, 1, 2
LEVEL_AUTO, LEVEL_SSE2, LEVEL_SSE41, LEVEL_AVX2, LEVEL_AVX512 = 0, 1, 2, 3, 4


class VoVolume(C.Structure):
    _fields_ = [("texels", C.c_void_p), ("w", C.c_int32), ("h", C.c_int32), ("d", C.c_int32),
                ("format", C.c_int32)]


class VoParams(C.Structure):
    _fields_ = [
        ("ubo0", C.c_float * 48), ("ubo1", C.c_float * 36),
        ("max_steps", C.c_int32), ("step_scale", C.c_float), ("density", C.c_float),
        ("sample_scale", C.c_float), ("coord_scale", C.c_float * 4), ("scroll_weight", C.c_float * 4),
        ("density_mode", C.c_int32), ("light_steps", C.c_int32), ("light_step_len", C.c_float),
        ("sun_dir", C.c_float * 3), ("sun_color", C.c_float * 3), ("ambient", C.c_float),
        ("sigma_t", C.c_float), ("early_term_eps", C.c_float),
        ("fbm_octaves", C.c_int32), ("fbm_seed", C.c_int32), ("fbm_frequency", C.c_float),
        ("fbm_gain", C.c_float), ("fbm_lacunarity", C.c_float), ("fbm_coverage", C.c_float),
        ("fbm_falloff0", C.c_float), ("fbm_falloff1", C.c_float),
    ]


def build(ref=True):
    """Compile the oracle (always) and oracle/_ref (only where /root/reference exists)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "all"])
    if ref and os.path.isdir("/root/reference/vendor/noise"):
        subprocess.check_call(["make", "-s", "-C", HERE, "-j8", "ref"])


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_ORACLE_SO):
            build(ref=False)
        L = C.CDLL(_ORACLE_SO)
        fp, ip, u8p = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
        L.vo_scene_default.argtypes = [C.c_float, C.c_float, C.c_float, fp, fp]
        L.vo_decode_ubo1.argtypes = [C.c_void_p, C.c_uint32, fp]
        L.vo_params_default.argtypes = [C.POINTER(VoParams)]
        L.vo_mat4_inverse.argtypes = [fp, fp]
        L.vo_sample.argtypes = [C.POINTER(VoVolume), C.c_float, C.c_float, C.c_float, fp]
        L.vo_render.argtypes = [C.POINTER(VoVolume), C.POINTER(VoParams), C.c_int, C.c_int, C.c_int, C.c_int,
                                fp, ip, C.POINTER(C.c_uint64), C.c_int]
        L.vo_render.restype = C.c_int
        L.vo_render_rows.argtypes = [C.POINTER(VoVolume), C.POINTER(VoParams), C.c_int, C.c_int, ip, C.c_int,
                                     fp, ip, C.POINTER(C.c_uint64), C.c_int]
        L.vo_render_rows.restype = C.c_int
        L.vo_frag_positions.argtypes = [C.POINTER(VoParams), C.c_int, C.c_int, ip, C.c_int, fp, ip]
        L.vo_frag_positions.restype = C.c_int
        L.vo_encode.argtypes = [fp, C.c_int64, C.c_int, C.c_void_p]
        L.vo_noise.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_float]
        L.vo_noise.restype = C.c_float
        L.vo_fbm_perlin.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float]
        L.vo_fbm_perlin.restype = C.c_float
        L.vo_noise_grid.argtypes = [C.c_int, fp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, fp, C.c_int]
        L.vo_pack_rgba8.argtypes = [fp, fp, fp, fp, fp, C.c_int64, u8p]
        L.vo_bake_default.argtypes = [u8p, C.c_int, C.c_int]
        L.vo_bake_smoke_r16f.argtypes = [C.POINTER(C.c_uint16), C.c_int, C.POINTER(VoParams), C.c_float, C.c_int]
        L.vo_float_to_half.argtypes = [C.c_float]
        L.vo_float_to_half.restype = C.c_uint16
        L.vo_half_to_float.argtypes = [C.c_uint16]
        L.vo_half_to_float.restype = C.c_float
        L.vo_hw_threads.restype = C.c_int
        _lib = L
    return _lib


def have_ref():
    return os.path.exists(_REF_SO)


def ref():
    """The reference's FastNoise2 (oracle/_ref). Raises if it was never built."""
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libvolrt_ref.so missing: run `make -C oracle ref` where /root/reference exists")
        R = C.CDLL(_REF_SO)
        fp = C.POINTER(C.c_float)
        R.vref_max_level.restype = C.c_int
        R.vref_grid.argtypes = [C.c_int, C.c_int, fp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, fp]
        R.vref_points.argtypes = [C.c_int, C.c_int, fp, C.c_int, fp, fp, fp, C.c_int]
        R.vref_fbm_points.argtypes = [C.c_int, fp, C.c_int, fp, fp, fp, C.c_int, C.c_int, C.c_float, C.c_float]
        R.vref_bake_default.argtypes = [C.POINTER(C.c_uint8), C.c_int, C.c_int, fp]
        R.vref_tree_grid.argtypes = [C.c_char_p, C.c_int, fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, fp]
        R.vref_metadata_dump.argtypes = [C.c_char_p, C.c_int]
        if hasattr(R, "vref_frag"):          # the reference's shaders compiled as C++ (glsl_shim.h)
            ip = C.POINTER(C.c_int32)
            R.vref_frag.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint8), C.c_int, C.c_int, C.c_int, fp, ip, C.c_int64, fp, C.c_int]
            R.vref_frag.restype = C.c_int
            R.vref_frag_constants.argtypes = [fp]
            R.vref_vert.argtypes = [C.c_void_p, C.c_uint32, fp, C.c_int, fp, fp]
            R.vref_vert.restype = C.c_int
            R.vref_draw.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint8), C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, fp, ip, fp, C.c_int]
            R.vref_draw.restype = C.c_int
        _ref = R
    return _ref


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


# ------------------------------------------------------------------ scene / params
def scene_default(phi_deg=0.0, theta_deg=0.0, frame_time=0.0):
    """-> (ubo0: 192 bytes, ubo1: the 140-byte C++ struct image) as numpy float32 arrays."""
    u0 = np.zeros(48, np.float32)
    u1 = np.zeros(35, np.float32)
    lib().vo_scene_default(phi_deg, theta_deg, frame_time, _fp(u0), _fp(u1))
    return u0, u1


def decode_ubo1(raw_bytes):
    out = np.zeros(36, np.float32)
    buf = (C.c_char * len(raw_bytes)).from_buffer_copy(raw_bytes)
    lib().vo_decode_ubo1(C.cast(buf, C.c_void_p), len(raw_bytes), _fp(out))
    return out


def params_default(**kw):
    p = VoParams()
    lib().vo_params_default(C.byref(p))
    set_params(p, **kw)
    return p


def set_params(p, **kw):
    for k, v in kw.items():
        cur = getattr(p, k)
        if hasattr(cur, "__len__"):
            v = list(np.asarray(v, np.float32).ravel())
            assert len(v) == len(cur), k
            for i, x in enumerate(v):
                cur[i] = float(x)
        else:
            setattr(p, k, v)
    return p


def set_scene(p, ubo0, ubo1_raw):
    """ubo0: 48 floats; ubo1_raw: bytes/array as the host writes them (140 or 144 B)."""
    u0 = np.asarray(ubo0, np.float32).ravel()
    for i in range(48):
        p.ubo0[i] = float(u0[i])
    dec = decode_ubo1(np.asarray(ubo1_raw, np.float32).tobytes())
    for i in range(36):
        p.ubo1[i] = float(dec[i])
    return p


def mat4_inverse(m):
    m = np.ascontiguousarray(m, np.float32).ravel()
    out = np.zeros(16, np.float32)
    lib().vo_mat4_inverse(_fp(m), _fp(out))
    return out


# ------------------------------------------------------------------ volume / render
_FMT_DTYPE = {VOL_RGBA8: np.uint8, VOL_R16F: np.uint16, VOL_R8: np.uint8, VOL_R32F: np.float32}


def make_volume(texels, fmt):
    """texels: numpy array shaped (d,h,w[,4]); returns (VoVolume, keepalive)."""
    t = np.ascontiguousarray(texels)
    if fmt == VOL_R16F and t.dtype == np.float16:
        t = t.view(np.uint16)
    assert t.dtype == _FMT_DTYPE[fmt], (t.dtype, fmt)
    d, h, w = t.shape[:3]
    v = VoVolume(t.ctypes.data, w, h, d, fmt)
    return v, t


def sample(vol, u, v, w):
    out = np.zeros(4, np.float32)
    lib().vo_sample(C.byref(vol), u, v, w, _fp(out))
    return out


def render(vol, p, W, H, rows=None, threads=0, want_steps=False):
    """-> (rgba float32 (rows,W,4), stats dict[, steps int32 (rows,W)])"""
    r0, r1 = rows if rows is not None else (0, H)
    out = np.zeros((r1 - r0, W, 4), np.float32)
    steps = np.zeros((r1 - r0, W), np.int32) if want_steps else None
    st = (C.c_uint64 * 4)()
    rc = lib().vo_render(C.byref(vol) if vol is not None else None, C.byref(p), W, H, r0, r1, _fp(out),
                         steps.ctypes.data_as(C.POINTER(C.c_int32)) if want_steps else None, st, threads)
    if rc != 0:
        raise ValueError("vo_render rc=%d" % rc)
    stats = {"covered": int(st[0]), "steps": int(st[1]), "early": int(st[2]), "light_samples": int(st[3])}
    return (out, stats, steps) if want_steps else (out, stats)


def render_rows(vol, p, W, H, rows, threads=0):
    """shade only the listed screen rows -> (rgba float32 (len(rows),W,4), stats dict)"""
    rows = np.ascontiguousarray(rows, np.int32)
    out = np.zeros((len(rows), W, 4), np.float32)
    st = (C.c_uint64 * 4)()
    rc = lib().vo_render_rows(C.byref(vol) if vol is not None else None, C.byref(p), W, H,
                              rows.ctypes.data_as(C.POINTER(C.c_int32)), len(rows), _fp(out), None, st, threads)
    if rc != 0:
        raise ValueError("vo_render_rows rc=%d" % rc)
    return out, {"covered": int(st[0]), "steps": int(st[1]), "early": int(st[2]), "light_samples": int(st[3])}


def frag_positions(p, W, H, rows=None):
    """The fragment stage's inputs as the oracle generates them for the listed rows (default: all):
    -> (fragPosition float32 (len(rows), W, 3), covered int32 (len(rows), W))"""
    rows = np.arange(H, dtype=np.int32) if rows is None else np.ascontiguousarray(rows, np.int32)
    xyz = np.zeros((len(rows), W, 3), np.float32)
    cov = np.zeros((len(rows), W), np.int32)
    ip = C.POINTER(C.c_int32)
    rc = lib().vo_frag_positions(C.byref(p), W, H, rows.ctypes.data_as(ip), len(rows), _fp(xyz), cov.ctypes.data_as(ip))
    if rc != 0:
        raise ValueError("vo_frag_positions rc=%d" % rc)
    return xyz, cov


# ------------------------------------------------------------------ the reference's own shaders, compiled as C++ (oracle/_ref)
def have_ref_glsl():
    return have_ref() and hasattr(ref(), "vref_frag")


def ref_frag(ubo1_raw, texels_rgba8, frag_pos, covered=None, threads=0):
    """/root/reference/shaders/frag.glsl main(), one invocation per fragPosition.  ubo1_raw: the bytes the host
    uploads for binding 1 (140 or 144; decoded with std140 offsets, bytes past the end read as zero); texels_rgba8:
    uint8 (d, h, w, 4); frag_pos: float32 (..., 3); covered: int32 (...) or None (all).  -> outColor float32 (..., 4)"""
    t = np.ascontiguousarray(texels_rgba8, np.uint8)
    d, h, w = t.shape[:3]
    fpos = np.ascontiguousarray(frag_pos, np.float32)
    n = fpos.size // 3
    cov = None if covered is None else np.ascontiguousarray(covered, np.int32)
    out = np.zeros(fpos.shape[:-1] + (4,), np.float32)
    raw = np.asarray(ubo1_raw, np.float32).tobytes() if not isinstance(ubo1_raw, (bytes, bytearray)) else bytes(ubo1_raw)
    buf = (C.c_char * len(raw)).from_buffer_copy(raw)
    rc = ref().vref_frag(C.cast(buf, C.c_void_p), len(raw), t.ctypes.data_as(C.POINTER(C.c_uint8)), w, h, d, _fp(fpos),
                         cov.ctypes.data_as(C.POINTER(C.c_int32)) if cov is not None else None, n, _fp(out), threads)
    if rc != 0:
        raise ValueError("vref_frag rc=%d" % rc)
    return out


def ref_frag_constants():
    """(density, maxSteps, boxMin.x, boxMax.x) as compiled from frag.glsl:29-32"""
    out = np.zeros(4, np.float32)
    ref().vref_frag_constants(_fp(out))
    return out


def ref_vert(ubo0, positions):
    """/root/reference/shaders/vert.glsl main() per inPosition -> (gl_Position (n,4), fragPosition (n,3))"""
    raw = np.asarray(ubo0, np.float32).tobytes()
    buf = (C.c_char * len(raw)).from_buffer_copy(raw)
    pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
    gl = np.zeros((len(pos), 4), np.float32)
    fr = np.zeros((len(pos), 3), np.float32)
    rc = ref().vref_vert(C.cast(buf, C.c_void_p), len(raw), _fp(pos), len(pos), _fp(gl), _fp(fr))
    if rc != 0:
        raise ValueError("vref_vert rc=%d" % rc)
    return gl, fr


def ref_draw(ubo0, ubo1_raw, texels_rgba8, W, H, threads=0):
    """The reference's whole draw on the CPU: its vert.glsl -> fixed-function rasteriser (oracle/ref_raster.cpp) -> its
    frag.glsl.  -> (rgba float32 (H,W,4), covered int32 (H,W), fragPosition float32 (H,W,3))"""
    t = np.ascontiguousarray(texels_rgba8, np.uint8)
    d, h, w = t.shape[:3]
    r0 = np.asarray(ubo0, np.float32).tobytes()
    r1 = np.asarray(ubo1_raw, np.float32).tobytes() if not isinstance(ubo1_raw, (bytes, bytearray)) else bytes(ubo1_raw)
    b0 = (C.c_char * len(r0)).from_buffer_copy(r0)
    b1 = (C.c_char * len(r1)).from_buffer_copy(r1)
    out = np.zeros((H, W, 4), np.float32)
    cov = np.zeros((H, W), np.int32)
    fpos = np.zeros((H, W, 3), np.float32)
    rc = ref().vref_draw(C.cast(b0, C.c_void_p), len(r0), C.cast(b1, C.c_void_p), len(r1), t.ctypes.data_as(C.POINTER(C.c_uint8)),
                         w, h, d, W, H, _fp(out), cov.ctypes.data_as(C.POINTER(C.c_int32)), _fp(fpos), threads)
    if rc != 0:
        raise ValueError("vref_draw rc=%d" % rc)
    return out, cov, fpos


def encode(rgba, fmt):
    rgba = np.ascontiguousarray(rgba, np.float32)
    n = rgba.size // 4
    out = np.zeros(rgba.shape, np.uint16 if fmt == OUT_RGBA16F else np.uint8)
    lib().vo_encode(_fp(rgba), n, fmt, out.ctypes.data)
    return out


# ------------------------------------------------------------------ noise
def noise(gen, seed, x, y, z):
    return float(lib().vo_noise(gen, seed, x, y, z))


def noise_points(gen, seed, xs, ys, zs):
    L = lib()
    return np.array([L.vo_noise(gen, seed, float(a), float(b), float(c)) for a, b, c in zip(xs, ys, zs)], np.float32)


def fbm_points(seed, xs, ys, zs, octaves=5, gain=0.5, lacunarity=2.0):
    L = lib()
    return np.array([L.vo_fbm_perlin(seed, float(a), float(b), float(c), octaves, gain, lacunarity)
                     for a, b, c in zip(xs, ys, zs)], np.float32)


def noise_grid(gen, n, freq, seed, threads=0):
    nx, ny, nz = (n, n, n) if np.isscalar(n) else n
    out = np.zeros((nz, ny, nx), np.float32)
    mm = np.zeros(2, np.float32)
    lib().vo_noise_grid(gen, _fp(out), nx, ny, nz, freq, seed, _fp(mm), threads)
    return out, mm


def bake_default(size=128, threads=0):
    out = np.zeros((size, size, size, 4), np.uint8)
    lib().vo_bake_default(out.ctypes.data_as(C.POINTER(C.c_uint8)), size, threads)
    return out


def bake_smoke_r16f(n, p, threshold=0.0, threads=0):
    out = np.zeros((n, n, n), np.uint16)
    lib().vo_bake_smoke_r16f(out.ctypes.data_as(C.POINTER(C.c_uint16)), n, C.byref(p), threshold, threads)
    return out


# ------------------------------------------------------------------ reference FastNoise2
def ref_grid(gen, n, freq, seed, level=LEVEL_AUTO):
    nx, ny, nz = (n, n, n) if np.isscalar(n) else n
    out = np.zeros((nz, ny, nx), np.float32)
    mm = np.zeros(2, np.float32)
    if ref().vref_grid(gen, level, _fp(out), nx, ny, nz, freq, seed, _fp(mm)) != 0:
        raise RuntimeError("FastNoise2 level %d unavailable on this CPU" % level)
    return out, mm


def ref_points(gen, seed, xs, ys, zs, level=LEVEL_AUTO):
    xs, ys, zs = (np.ascontiguousarray(a, np.float32) for a in (xs, ys, zs))
    out = np.zeros(len(xs), np.float32)
    if ref().vref_points(gen, level, _fp(out), len(xs), _fp(xs), _fp(ys), _fp(zs), seed) != 0:
        raise RuntimeError("FastNoise2 level %d unavailable on this CPU" % level)
    return out


def ref_fbm_points(seed, xs, ys, zs, octaves=5, gain=0.5, lacunarity=2.0, level=LEVEL_AUTO):
    xs, ys, zs = (np.ascontiguousarray(a, np.float32) for a in (xs, ys, zs))
    out = np.zeros(len(xs), np.float32)
    if ref().vref_fbm_points(level, _fp(out), len(xs), _fp(xs), _fp(ys), _fp(zs), seed, octaves, gain, lacunarity) != 0:
        raise RuntimeError("FastNoise2 level %d unavailable on this CPU" % level)
    return out


def ref_bake_default(size=128, level=LEVEL_AUTO):
    out = np.zeros((size, size, size, 4), np.uint8)
    mm = np.zeros(8, np.float32)
    if ref().vref_bake_default(out.ctypes.data_as(C.POINTER(C.c_uint8)), size, level, _fp(mm)) != 0:
        raise RuntimeError("FastNoise2 level %d unavailable on this CPU" % level)
    return out, mm


# ------------------------------------------------------------------ min-max pyramid (build-defined, include/volrt.h)
def minmax_pyramid(texels, fmt):
    """numpy restatement of the min-max pyramid definition in include/volrt.h ("volume files"): level 0 = per
    8^3 brick over the brick plus a one-texel apron clamped at the faces, entry = channels x (min, max) of the
    decoded values; level l+1 reduces up to 2x2x2 children.  NaN texels count for the max only.
    texels: (d,h,w) float16/uint16-bits/uint8 or (d,h,w,4) uint8 -> list of (bz,by,bx,channels,2) float32 arrays.
    The reference has no such structure (it has no empty-space skipping); parity for it is self-referential."""
    t = np.asarray(texels)
    if fmt == VOL_R16F:
        v = (t.view(np.float16) if t.dtype == np.uint16 else t).astype(np.float32)[..., None]
    elif t.ndim == 4:
        v = t.astype(np.float32) / np.float32(255.0)
    else:
        v = (t.astype(np.float32) / np.float32(255.0))[..., None]
    d, h, w, ch = v.shape
    nb = [(n + 7) // 8 for n in (d, h, w)]
    lo = np.where(np.isnan(v), np.float32(np.inf), v)                       # NaN ignored by the min
    hi = np.where(np.isnan(v), np.float32(np.inf), v)
    nan = np.isnan(v)
    level = np.zeros((nb[0], nb[1], nb[2], ch, 2), np.float32)
    for bz in range(nb[0]):
        z0, z1 = max(bz * 8 - 1, 0), min(bz * 8 + 9, d)
        for by in range(nb[1]):
            y0, y1 = max(by * 8 - 1, 0), min(by * 8 + 9, h)
            for bx in range(nb[2]):
                x0, x1 = max(bx * 8 - 1, 0), min(bx * 8 + 9, w)
                blk_lo = lo[z0:z1, y0:y1, x0:x1].reshape(-1, ch)
                blk_hi = hi[z0:z1, y0:y1, x0:x1].reshape(-1, ch)
                blk_nan = nan[z0:z1, y0:y1, x0:x1].reshape(-1, ch)
                level[bz, by, bx, :, 0] = blk_lo.min(axis=0)
                mx = np.where(blk_nan, -np.float32(np.inf), blk_hi).max(axis=0)
                level[bz, by, bx, :, 1] = np.where(blk_nan.any(axis=0), np.float32(np.nan), mx)
    out = [level]
    while out[-1].shape[:3] != (1, 1, 1):
        p = out[-1]
        sz, sy, sx = p.shape[:3]
        nz, ny, nx = (sz + 1) // 2, (sy + 1) // 2, (sx + 1) // 2
        q = np.zeros((nz, ny, nx, ch, 2), np.float32)
        for z in range(nz):
            for y in range(ny):
                for x in range(nx):
                    blk = p[2 * z:2 * z + 2, 2 * y:2 * y + 2, 2 * x:2 * x + 2].reshape(-1, ch, 2)
                    q[z, y, x, :, 0] = blk[..., 0].min(axis=0)
                    mx = blk[..., 1]
                    q[z, y, x, :, 1] = np.where(np.isnan(mx).any(axis=0), np.float32(np.nan), np.nanmax(np.where(np.isnan(mx), -np.inf, mx), axis=0))
        out.append(q)
    return out


# ------------------------------------------------------------------ reference FastNoise2 on encoded node trees
def ref_tree_grid(encoded, shape, freq, seed, start=(0, 0, 0), level=LEVEL_AUTO):
    """the reference's own decoder + generators (FastNoise::NewFromEncodedNodeTree -> GenUniformGrid3D);
    shape = (nz, ny, nx) -> (grid float32, minmax float32[2])"""
    nz, ny, nx = shape
    out = np.zeros(shape, np.float32)
    mm = np.zeros(2, np.float32)
    rc = ref().vref_tree_grid(encoded.encode(), level, _fp(out), start[0], start[1], start[2], nx, ny, nz, freq, seed, _fp(mm))
    if rc != 0:
        raise RuntimeError("FastNoise2 could not build the tree (or SIMD level %d is unavailable)" % level)
    return out, mm


def ref_metadata():
    import json
    buf = C.create_string_buffer(1 << 18)
    n = ref().vref_metadata_dump(buf, len(buf))
    if n < 0:
        raise RuntimeError("metadata dump needs %d bytes" % -n)
    return [json.loads(l) for l in buf.value.decode().splitlines()]
