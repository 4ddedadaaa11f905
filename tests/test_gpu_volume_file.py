"""Volume files and the min-max pyramid on the device, through the C ABI."""
import numpy as np
import pytest

import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene, volfile
from util import smoke_volume

pytestmark = pytest.mark.gpu


def _eq_nan(a, b):
    return np.array_equal(a.view(np.uint32) & 0x7fffffff, b.view(np.uint32) & 0x7fffffff) if np.isnan(a).any() or np.isnan(b).any() \
        else np.array_equal(a, b)


@pytest.mark.parametrize("fmt,shape", [(capi.VOL_R16F, (40, 24, 64)), (capi.VOL_R16F, (17, 9, 29)), (capi.VOL_R8_UNORM, (9, 31, 50)),
                                        (capi.VOL_RGBA8_UNORM, (24, 16, 36)), (capi.VOL_RGBA8_UNORM, (10, 13, 27)), (capi.VOL_R16F, (1, 1, 1))])
def test_device_pyramid_equals_the_oracle(oracle, fmt, shape):
    rng = np.random.default_rng(sum(shape) + fmt)
    mask = rng.random(shape) > 0.8
    if fmt == capi.VOL_R16F:
        t = ((rng.random(shape) - 0.3) * mask).astype(np.float16)          # negative values too
        if t.size > 100:
            t[shape[0] // 2, shape[1] // 2, 7] = np.nan
        ofmt = oracle.VOL_R16F
    elif fmt == capi.VOL_R8_UNORM:
        t = (rng.integers(0, 256, shape) * mask).astype(np.uint8)
        ofmt = oracle.VOL_R8
    else:
        t = (rng.integers(0, 256, shape + (4,)) * mask[..., None]).astype(np.uint8)
        ofmt = oracle.VOL_RGBA8
    ref = oracle.minmax_pyramid(t, ofmt)
    with vr.Renderer(32, 32) as r:
        r.upload_volume(t, fmt, capi.LAYOUT_LINEAR)
        got = r.volume_pyramid()
    assert len(got) == len(ref)
    for a, b in zip(got, ref):
        assert a.shape == b.shape
        assert _eq_nan(a, b)


def test_save_load_round_trip_renders_the_same_frame(tmp_path):
    W, H = 320, 180
    vol = smoke_volume(48, 5)
    vol[vol < np.float16(0.25)] = 0                     # leave empty space for the occupancy / distance map
    u0, u1 = scene.default_scene(20.0, 10.0)
    kw = dict(density_mode=capi.DENSITY_SINGLE, sample_scale=0.5, max_steps=192, skip_empty=1)
    path = str(tmp_path / "smoke.vol")
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_R16F, capi.LAYOUT_LINEAR)
        r.set_params(u0, u1, **kw)
        a = r.render()[0].copy()
        r.save_volume_file(path, with_pyramid=True)
        pyr = r.volume_pyramid()
    i = volfile.info(path)
    assert i["pyramid_levels"] == len(pyr) and i["sha256"] == volfile.sha256(vol)
    back, _, flat = volfile.read(path, verify=True, want_pyramid=True)
    assert np.array_equal(back.view(np.uint16), vol.view(np.uint16))
    assert np.array_equal(np.concatenate([p.reshape(-1) for p in pyr]), flat)
    for flags in (capi.FILE_VERIFY, 0, capi.FILE_VERIFY | capi.FILE_TRUST_PYRAMID):
        with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
            r.upload_volume_file(path, capi.LAYOUT_LINEAR, flags)
            r.set_params(u0, u1, **kw)
            b = r.render(flags=capi.RENDER_STATS)[0]
            assert r.stats()["skipped_steps"] > 0
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), flags
            assert np.array_equal(r.download_volume().view(np.uint16), vol.view(np.uint16))


def test_upload_rejects_bad_files(tmp_path):
    vol = smoke_volume(16, 1)
    path = str(tmp_path / "v.vol")
    volfile.write(path, vol, capi.VOL_R16F)
    raw = bytearray(open(path, "rb").read())
    raw[128 + 100] ^= 1
    open(path, "wb").write(raw)
    with vr.Renderer(32, 32) as r:
        with pytest.raises(vr.VolrtError) as e:
            r.upload_volume_file(path, flags=capi.FILE_VERIFY)
        assert e.value.status == capi.VR_ERR_IO and "SHA-256" in str(e.value)
        with pytest.raises(vr.VolrtError) as e:
            r.upload_volume_file(path, flags=capi.FILE_TRUST_PYRAMID)       # the file has no pyramid
        assert e.value.status == capi.VR_ERR_BAD_STATE
        with pytest.raises(vr.VolrtError):
            r.upload_volume_file(str(tmp_path / "missing.vol"))
        r.upload_volume_file(path, flags=0)                                  # unverified: loads
        with pytest.raises(vr.VolrtError):
            r.upload_volume_file(path, layout=capi.LAYOUT_BRICK + 7)


def test_chunked_upload_of_a_file_larger_than_the_staging_buffers(tmp_path):
    """8 chunks through the 6-deep ring of 16 MiB staging buffers (120 MiB volume), SHA verified on the way"""
    rng = np.random.default_rng(11)
    t = rng.integers(0, 256, (480, 256, 256, 4), dtype=np.uint8)
    path = str(tmp_path / "big.vol")
    volfile.write(path, t, capi.VOL_RGBA8_UNORM)
    with vr.Renderer(32, 32) as r:
        r.upload_volume_file(path, capi.LAYOUT_TEXTURE, capi.FILE_VERIFY)
        assert np.array_equal(r.download_volume(), t)
