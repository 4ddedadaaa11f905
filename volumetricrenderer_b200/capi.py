"""ctypes mirror of include/volrt.h — the C ABI of libvolrt.so (hand-written sm_100a CUDA).

This module only declares types and loads the library.  There is NO CPU path behind it:
if libvolrt.so is missing, `load()` raises; if there is no CUDA device, `vr_create` fails with
VR_ERR_CUDA.  Build the library with `python -c "import __graft_entry__ as g; g.build()"` or
`make -C volumetricrenderer_b200/csrc`.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VOLRT_LIB") or os.path.join(HERE, "libvolrt.so")   # VOLRT_LIB: A/B builds

VR_OK, VR_ERR_INVALID_ARGUMENT, VR_ERR_CUDA, VR_ERR_OUT_OF_MEMORY, VR_ERR_UNSUPPORTED, VR_ERR_BAD_STATE, VR_ERR_IO = range(7)
STATUS_NAMES = ["VR_OK", "VR_ERR_INVALID_ARGUMENT", "VR_ERR_CUDA", "VR_ERR_OUT_OF_MEMORY", "VR_ERR_UNSUPPORTED",
                "VR_ERR_BAD_STATE", "VR_ERR_IO"]

OUT_RGBA8_UNORM, OUT_RGBA8_SRGB, OUT_RGBA16F, OUT_RGBA32F = 0, 1, 2, 3
VOL_RGBA8_UNORM, VOL_R16F, VOL_R8_UNORM = 0, 1, 2
ADDRESS_MIRRORED_REPEAT, ADDRESS_CLAMP_TO_EDGE = 0, 1
LAYOUT_AUTO, LAYOUT_TEXTURE, LAYOUT_LINEAR, LAYOUT_BRICK = 0, 1, 2, 3
DENSITY_RGBA_PRODUCT, DENSITY_SINGLE, DENSITY_FBM, DENSITY_TREE = 0, 1, 2, 3
MEM_NONE, MEM_HOST, MEM_DEVICE = 0, 1, 2
RENDER_ASYNC, RENDER_DIRECT, RENDER_STATS, RENDER_NO_MARCH = 1, 2, 4, 8
NOISE_CELLULAR_DISTANCE, NOISE_PERLIN, NOISE_SIMPLEX, NOISE_ZERO, NOISE_SAME = 0, 1, 2, 3, 0xFFFFFFFF
NOISE_TREE = 4


MAX_DEVICES = 8


class vr_create_info(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("width", C.c_uint32), ("height", C.c_uint32),
                ("format", C.c_uint32), ("max_views", C.c_uint32), ("flags", C.c_uint32),
                ("device_count", C.c_uint32), ("devices", C.c_int32 * MAX_DEVICES)]


class vr_volume_desc(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("depth", C.c_uint32),
                ("format", C.c_uint32), ("address_mode", C.c_uint32), ("layout", C.c_uint32)]


class vr_volume_file_info(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("depth", C.c_uint32),
                ("format", C.c_uint32), ("address_mode", C.c_uint32), ("pyramid_levels", C.c_uint32), ("reserved", C.c_uint32),
                ("texel_bytes", C.c_uint64), ("pyramid_bytes", C.c_uint64), ("sha256", C.c_uint8 * 32)]


FILE_VERIFY, FILE_WITH_PYRAMID, FILE_TRUST_PYRAMID = 1, 2, 4


class vr_noise_channel(C.Structure):
    _fields_ = [("generator", C.c_uint32), ("frequency", C.c_float), ("seed", C.c_int32),
                ("norm_generator", C.c_uint32), ("norm_frequency", C.c_float), ("norm_seed", C.c_int32),
                ("power", C.c_uint32)]


class vr_noise_desc(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("size", C.c_uint32), ("format", C.c_uint32), ("layout", C.c_uint32),
                ("address_mode", C.c_uint32), ("threshold", C.c_float), ("channel", vr_noise_channel * 4),
                ("tree", C.c_void_p * 4)]


class vr_params(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("ubo0_size", C.c_uint32), ("ubo0", C.c_void_p), ("ubo1", C.c_void_p),
        ("ubo1_size", C.c_uint32), ("max_steps", C.c_uint32), ("step_scale", C.c_float), ("density", C.c_float),
        ("sample_scale", C.c_float), ("coord_scale", C.c_float * 4), ("scroll_weight", C.c_float * 4),
        ("density_mode", C.c_uint32), ("light_steps", C.c_uint32), ("light_step_len", C.c_float),
        ("sun_dir", C.c_float * 3), ("sun_color", C.c_float * 3), ("ambient", C.c_float), ("sigma_t", C.c_float),
        ("early_term_eps", C.c_float), ("skip_empty", C.c_uint32), ("fbm_octaves", C.c_int32),
        ("fbm_seed", C.c_int32), ("fbm_frequency", C.c_float), ("fbm_gain", C.c_float),
        ("fbm_lacunarity", C.c_float), ("fbm_coverage", C.c_float), ("fbm_falloff0", C.c_float),
        ("fbm_falloff1", C.c_float), ("view_count", C.c_uint32), ("view_ubos", C.c_void_p),
        ("density_tree", C.c_void_p), ("tree_frequency", C.c_float), ("tree_seed", C.c_int32),
    ]


class vr_target(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("memory_kind", C.c_uint32), ("dst", C.c_void_p),
                ("row_pitch", C.c_size_t), ("row_begin", C.c_uint32), ("row_end", C.c_uint32),
                ("view_begin", C.c_uint32), ("view_end", C.c_uint32), ("flags", C.c_uint32), ("stream", C.c_void_p)]


class vr_stats(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("gpu_ms", C.c_float), ("rays", C.c_uint64), ("covered", C.c_uint64),
                ("steps", C.c_uint64), ("early_terminated", C.c_uint64), ("light_samples", C.c_uint64),
                ("skipped_steps", C.c_uint64), ("kernel_launches", C.c_uint32), ("constant_channels", C.c_uint32)]


# every symbol include/volrt.h declares: name -> (restype, argtypes)
_VP = C.c_void_p
EXPORTS = {
    "vr_version": (C.c_uint32, []),
    "vr_params_default": (C.c_int32, [C.POINTER(vr_params)]),
    "vr_create": (C.c_int32, [C.POINTER(vr_create_info), C.POINTER(_VP)]),
    "vr_destroy": (None, [_VP]),
    "vr_last_error": (C.c_char_p, [_VP]),
    "vr_upload_volume": (C.c_int32, [_VP, C.POINTER(vr_volume_desc), _VP]),
    "vr_generate_volume": (C.c_int32, [_VP, C.POINTER(vr_noise_desc), C.POINTER(vr_params)]),
    "vr_download_volume": (C.c_int32, [_VP, _VP, C.c_size_t]),
    "vr_noise_desc_default": (C.c_int32, [C.POINTER(vr_noise_desc)]),
    "vr_set_params": (C.c_int32, [_VP, C.POINTER(vr_params)]),
    "vr_render": (C.c_int32, [_VP, C.POINTER(vr_target)]),
    "vr_wait": (C.c_int32, [_VP]),
    "vr_get_stats": (C.c_int32, [_VP, C.POINTER(vr_stats)]),
    "vr_frame_info": (C.c_int32, [_VP, C.POINTER(_VP), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "vr_row_bands": (C.c_int32, [_VP, C.c_uint32, C.POINTER(C.c_uint32)]),
    "vr_row_cost": (C.c_int32, [_VP, C.POINTER(C.c_float)]),
    "vr_split_rows": (C.c_int32, [C.POINTER(C.c_float), C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]),
    "vr_frame_export": (C.c_int32, [_VP, _VP]),
    "vr_frame_import": (C.c_int32, [_VP, _VP, C.POINTER(_VP)]),
    "vr_frame_unmap": (C.c_int32, [_VP, _VP]),
    "vr_gather_acquire": (C.c_int32, [_VP, _VP, C.c_uint32, C.c_uint32, C.c_uint32]),
    "vr_gather_signal": (C.c_int32, [_VP, _VP, C.c_uint32, C.c_uint32]),
    "vr_gather_collect": (C.c_int32, [_VP, C.c_uint32, C.c_uint32, _VP, C.c_size_t]),
    "vr_flush_l2": (C.c_int32, [_VP]),
    "vr_measure_peaks": (C.c_int32, [_VP, C.POINTER(C.c_float)]),
    "vr_measure_tex_stream": (C.c_int32, [_VP, C.POINTER(C.c_float)]),
    "vr_host_alloc": (_VP, [C.c_size_t]),
    "vr_host_free": (None, [_VP]),
    "vr_sha256": (C.c_int32, [_VP, C.c_size_t, C.POINTER(C.c_uint8)]),
    "vr_volume_file_write": (C.c_int32, [C.c_char_p, C.POINTER(vr_volume_desc), _VP, C.POINTER(C.c_float), C.c_uint32]),
    "vr_volume_file_info_read": (C.c_int32, [C.c_char_p, C.POINTER(vr_volume_file_info)]),
    "vr_volume_file_read": (C.c_int32, [C.c_char_p, _VP, C.c_size_t, C.POINTER(C.c_float), C.c_size_t, C.c_uint32]),
    "vr_pyramid_floats": (C.c_size_t, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]),
    "vr_upload_volume_file": (C.c_int32, [_VP, C.c_char_p, C.c_uint32, C.c_uint32]),
    "vr_save_volume_file": (C.c_int32, [_VP, C.c_char_p, C.c_uint32]),
    "vr_volume_pyramid": (C.c_int32, [_VP, C.POINTER(C.c_float), C.c_size_t]),
    "vr_noise_tree_compile": (C.c_int32, [C.c_char_p, C.POINTER(_VP), C.c_char_p, C.c_size_t]),
    "vr_noise_tree_free": (None, [_VP]),
    "vr_noise_tree_info": (C.c_int32, [_VP, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "vr_noise_tree_grid": (C.c_int32, [_VP, _VP, C.c_int32, C.c_int32, C.c_int32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_int32,
                                       C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "vr_noise_tree_grid_static": (C.c_int32, [C.c_int32, _VP, C.c_int32, C.c_int32, C.c_int32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float,
                                              C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_float), _VP, C.c_char_p, C.c_size_t]),
}
STATIC_GRID_MAX_POINTS = 262144

_lib = None


def load():
    """dlopen libvolrt.so and type every export.  Raises if the library was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libvolrt.so is not built (%s): run __graft_entry__.build() — libvolrt has no "
                               "fallback path" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            f = getattr(L, name)          # AttributeError if the .so lacks a declared symbol
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib
