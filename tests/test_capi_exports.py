"""The C-ABI library loads and exports every symbol include/volrt.h declares (no compute without a GPU)."""
import ctypes as C
import os
import re

import pytest

from volumetricrenderer_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "volrt.h")).read()
    return sorted(set(re.findall(r"^VR_API [^;(]*?\b(vr_\w+)\(", text, flags=re.M)))


def test_header_and_binding_agree():
    assert _declared() == sorted(capi.EXPORTS)


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    for name in _declared():
        assert hasattr(lib, name), name
    assert lib.vr_version() == (0 << 16) | 2


def test_struct_sizes_match_the_header_layout():
    # natural alignment, as a C compiler lays the structs of include/volrt.h out
    assert C.sizeof(capi.vr_create_info) == 28 + 4 + 4 * capi.MAX_DEVICES      # the 28-byte struct of version 0.1 + the device group
    assert C.sizeof(capi.vr_volume_desc) == 28
    assert C.sizeof(capi.vr_noise_channel) == 28
    assert C.sizeof(capi.vr_noise_desc) == 24 + 4 * 28 + 4 * 8          # channel recipes + the per-channel node-tree pointers
    assert C.sizeof(capi.vr_target) == 56
    assert C.sizeof(capi.vr_stats) == 64
    p = capi.vr_params()
    assert capi.load().vr_params_default(C.byref(p)) == capi.VR_OK
    assert p.struct_size == C.sizeof(capi.vr_params)      # the library compiled the same layout
    assert (p.max_steps, p.step_scale, p.density) == (128, 4.0, 1.0)           # frag.glsl:29-30,42
    assert [round(x, 4) for x in p.coord_scale] == [1.0, 0.8, 0.75, 0.7]       # frag.glsl:66-69
    assert [round(x, 4) for x in p.scroll_weight] == [0.0, 0.2, 0.25, 0.3]
    assert abs(p.sample_scale - 0.2) < 1e-7 and p.ubo0_size == 192 and p.ubo1_size == 140


def test_default_noise_recipe_encodes_quirk_q1():
    d = capi.vr_noise_desc()
    assert capi.load().vr_noise_desc_default(C.byref(d)) == capi.VR_OK
    assert d.struct_size == C.sizeof(capi.vr_noise_desc) and d.size == 128
    ch = d.channel
    # TestMain.cpp:59-60: both cellular bakes land in buffer 1; buffer 2 stays zero
    assert (ch[0].generator, round(ch[0].frequency, 4), ch[0].seed) == (capi.NOISE_CELLULAR_DISTANCE, 0.03, 2)
    assert (ch[0].norm_generator, round(ch[0].norm_frequency, 4), ch[0].norm_seed, ch[0].power) == (0, 0.01, 1, 4)
    assert ch[1].generator == capi.NOISE_ZERO and (round(ch[1].norm_frequency, 4), ch[1].norm_seed) == (0.03, 2)
    assert (ch[2].generator, ch[2].seed, ch[3].generator, ch[3].seed) == (capi.NOISE_PERLIN, 3, capi.NOISE_SIMPLEX, 4)


def test_no_cpu_fallback_without_a_device():
    """without a CUDA device vr_create must fail loudly (there is no CPU path in the product)"""
    lib = capi.load()
    ctx = C.c_void_p()
    info = capi.vr_create_info(C.sizeof(capi.vr_create_info), 0, 64, 64, 0, 1, 0)
    st = lib.vr_create(C.byref(info), C.byref(ctx))
    if st != capi.VR_OK:                      # a caller compiled against version 0.1's 28-byte struct gets the same answer
        old = capi.vr_create_info(28, 0, 64, 64, 0, 1, 0)
        assert lib.vr_create(C.byref(old), C.byref(ctx)) == st
    if st == capi.VR_OK:                      # a GPU is present: fine, clean up
        lib.vr_destroy(ctx)
        pytest.skip("CUDA device present")
    assert st == capi.VR_ERR_CUDA and not ctx.value
    assert b"no CUDA device" in lib.vr_last_error(None)
    # argument validation works without a device
    assert lib.vr_create(None, C.byref(ctx)) == capi.VR_ERR_INVALID_ARGUMENT
    assert lib.vr_set_params(None, None) == capi.VR_ERR_INVALID_ARGUMENT
    assert lib.vr_render(None, None) == capi.VR_ERR_INVALID_ARGUMENT


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "volumetricrenderer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "volrt_oracle" not in src and "from oracle" not in src and "import oracle" not in src, f
