"""Volume files (include/volrt.h "volume files"): host-only entry points of libvolrt.
No device needed: header/sections/SHA-256 round trips, corrupt and truncated files, the pyramid's geometry."""
import ctypes as C
import hashlib
import os

import numpy as np
import pytest

from volumetricrenderer_b200 import capi, volfile


def test_sha256_matches_hashlib_at_block_boundaries():
    rng = np.random.default_rng(7)
    assert volfile.sha256(b"") == hashlib.sha256(b"").digest()
    assert volfile.sha256(b"abc").hex() == "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"     # FIPS 180-4 B.1
    for n in (1, 55, 56, 57, 63, 64, 65, 119, 120, 127, 128, 1 << 16, (1 << 20) + 3):
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert volfile.sha256(b) == hashlib.sha256(b).digest(), n


@pytest.mark.parametrize("fmt,shape", [(capi.VOL_RGBA8_UNORM, (5, 6, 7, 4)), (capi.VOL_R16F, (9, 16, 24)), (capi.VOL_R8_UNORM, (1, 1, 1))])
def test_round_trip_and_header(tmp_path, fmt, shape):
    rng = np.random.default_rng(sum(shape))
    t = rng.integers(0, 256, shape, dtype=np.uint8) if fmt != capi.VOL_R16F else rng.random(shape).astype(np.float16)
    path = str(tmp_path / "v.vol")
    n = capi.load().vr_pyramid_floats(shape[2], shape[1], shape[0], fmt, None)
    pyr = rng.random(n).astype(np.float32)
    volfile.write(path, t, fmt, capi.ADDRESS_CLAMP_TO_EDGE, pyramid=pyr)
    i = volfile.info(path)
    assert (i["depth"], i["height"], i["width"]) == tuple(shape[:3]) and i["format"] == fmt
    assert i["address_mode"] == capi.ADDRESS_CLAMP_TO_EDGE and i["texel_bytes"] == t.nbytes
    assert i["sha256"] == hashlib.sha256(t.tobytes()).digest()
    assert i["pyramid_levels"] == len(volfile.pyramid_shapes(shape[2], shape[1], shape[0], fmt)) and i["pyramid_bytes"] == 4 * n
    back, _, pyr2 = volfile.read(path, verify=True, want_pyramid=True)
    assert back.dtype == t.dtype and np.array_equal(back.view(np.uint8), t.view(np.uint8)) and np.array_equal(pyr, pyr2)
    assert os.path.getsize(path) == 128 + t.nbytes + 4 * n
    raw = open(path, "rb").read()
    assert raw[:8] == b"VOLRTVOL" and int.from_bytes(raw[8:12], "little") == 1 and raw[128:128 + t.nbytes] == t.tobytes()


def test_corrupt_truncated_and_foreign_files(tmp_path):
    t = np.arange(4 * 8 * 8, dtype=np.uint8).reshape(4, 8, 8)
    path = str(tmp_path / "v.vol")
    volfile.write(path, t, capi.VOL_R8_UNORM)
    raw = bytearray(open(path, "rb").read())
    # one flipped texel bit: readable without verification, rejected with it
    bad = bytearray(raw); bad[128 + 17] ^= 0x10
    open(path, "wb").write(bad)
    got, _ = volfile.read(path, verify=False)
    assert got[0, 2, 1] == (t[0, 2, 1] ^ 0x10)
    with pytest.raises(volfile.VolumeFileError):
        volfile.read(path, verify=True)
    # truncated texel section
    open(path, "wb").write(raw[:-5])
    with pytest.raises(volfile.VolumeFileError):
        volfile.info(path)
    # not a volume file / shorter than a header / missing
    open(path, "wb").write(b"P6\n1 1\n255\n" + bytes(200))
    with pytest.raises(volfile.VolumeFileError):
        volfile.info(path)
    open(path, "wb").write(raw[:100])
    with pytest.raises(volfile.VolumeFileError):
        volfile.info(path)
    with pytest.raises(volfile.VolumeFileError):
        volfile.info(str(tmp_path / "nope.vol"))
    # dims that disagree with the section size
    bad = bytearray(raw); bad[16:20] = (9).to_bytes(4, "little")
    open(path, "wb").write(bad)
    with pytest.raises(volfile.VolumeFileError):
        volfile.info(path)
    # asking for a pyramid the file does not have
    open(path, "wb").write(raw)
    with pytest.raises(volfile.VolumeFileError):
        volfile.read(path, want_pyramid=True)


def test_pyramid_geometry():
    lv = C.c_uint32(0)
    lib = capi.load()
    assert lib.vr_pyramid_floats(128, 128, 128, capi.VOL_RGBA8_UNORM, C.byref(lv)) == 8 * (16 ** 3 + 8 ** 3 + 4 ** 3 + 2 ** 3 + 1) and lv.value == 5
    assert lib.vr_pyramid_floats(1, 1, 1, capi.VOL_R16F, C.byref(lv)) == 2 and lv.value == 1
    assert lib.vr_pyramid_floats(17, 9, 20, capi.VOL_R16F, C.byref(lv)) == 2 * (3 * 2 * 3 + 2 * 1 * 2 + 1) and lv.value == 3
    assert [s[:3] for s in volfile.pyramid_shapes(20, 9, 17, capi.VOL_R16F)] == [(3, 2, 3), (2, 1, 2), (1, 1, 1)]


def test_oracle_pyramid_is_the_brute_force_definition(oracle):
    rng = np.random.default_rng(3)
    t = (rng.random((11, 18, 26)) * (rng.random((11, 18, 26)) > 0.7)).astype(np.float16)
    lv = oracle.minmax_pyramid(t, oracle.VOL_R16F)
    f = t.astype(np.float32)
    # brick (1,2,3) of level 0: texels z 7..10 (clamped), y 15..17, x 23..25
    assert lv[0][1, 2, 3, 0, 0] == f[7:11, 15:18, 23:26].min() and lv[0][1, 2, 3, 0, 1] == f[7:11, 15:18, 23:26].max()
    assert lv[0][0, 0, 0, 0, 1] == f[0:9, 0:9, 0:9].max()
    assert lv[-1].shape == (1, 1, 1, 1, 2) and lv[-1][0, 0, 0, 0, 0] == f.min() and lv[-1][0, 0, 0, 0, 1] == f.max()
