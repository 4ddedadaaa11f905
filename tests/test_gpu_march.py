"""GPU parity: libvolrt's march kernels (through the C ABI) vs the CPU oracle, same inputs.

Bars (BASELINE.json north_star): per-pixel max abs error <= 2/255 and PSNR >= 50 dB in 8-bit for the
hardware-filtered TEXTURE layout; the FP32-filtered LINEAR layout follows the oracle's operation
order and must match its binary32 output BIT FOR BIT.
"""
import numpy as np
import pytest

import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
from util import psnr8, max_abs8, oracle_params, smoke_volume

pytestmark = pytest.mark.gpu

TOL8 = 2          # /255, stated by north_star
PSNR_MIN = 50.0   # dB, stated by north_star


def _oracle_frame(oracle, vol, fmt, p, W, H, rows=None):
    ovol, keep = oracle.make_volume(vol, fmt)
    img, st, steps = oracle.render(ovol, p, W, H, rows=rows, want_steps=True)
    return img, st, steps


def test_default_scene_linear_is_bit_exact(oracle, default_volume):
    """config 1: the reference's default scene, 1280x720, FP32 software trilinear."""
    W, H = 1280, 720
    ubo0, ubo1 = scene.default_scene()
    p = oracle_params(oracle, ubo0, ubo1)
    ref, st, _ = _oracle_frame(oracle, default_volume, oracle.VOL_RGBA8, p, W, H)
    assert abs(st["covered"] - 196251) <= 4 and abs(st["steps"] - 7439290) < 2000
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1)
        img = r.render(flags=capi.RENDER_STATS)[0]
        s = r.stats()
    assert s["covered"] == st["covered"] and s["steps"] == st["steps"]
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))


def test_default_scene_against_the_reference_shaders_compiled_here(oracle, default_volume):
    """config 1 against the REFERENCE ITSELF, no restatement in between: /root/reference/shaders/frag.glsl compiled as C++
    (oracle/_ref, built by `make -C oracle ref`; oracle/glsl_shim.h) on the reference's own FastNoise2 bake.  LINEAR layout:
    binary32 outColor bit for bit.  TEXTURE layout: north-star tolerance.  And the reference's whole draw (vert.glsl ->
    rasteriser -> frag.glsl, oracle/ref_raster.cpp) within the tolerance too."""
    if not oracle.have_ref_glsl():
        pytest.skip("oracle/_ref with the compiled shaders did not travel to this box")
    W, H = 1280, 720
    ubo0, ubo1 = scene.default_scene()
    p = oracle_params(oracle, ubo0, ubo1)
    fpos, cov = oracle.frag_positions(p, W, H)
    theirs = oracle.ref_frag(ubo1, default_volume, fpos, cov)                    # frag.glsl main() per covered pixel
    drawn, dcov, _ = oracle.ref_draw(ubo0, ubo1, default_volume, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1)
        strict = r.render()[0].copy()
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        fast = r.render()[0].copy()
    assert np.array_equal(strict.view(np.uint32), theirs.view(np.uint32))
    e8 = lambda a: oracle.encode(a, oracle.OUT_RGBA8_UNORM)
    assert max_abs8(e8(fast), e8(theirs)) <= TOL8 and psnr8(e8(fast), e8(theirs)) >= PSNR_MIN
    assert (dcov != cov).sum() <= 4
    assert max_abs8(e8(strict), e8(drawn)) <= TOL8 and psnr8(e8(strict), e8(drawn)) >= PSNR_MIN
    assert max_abs8(e8(fast), e8(drawn)) <= TOL8 and psnr8(e8(fast), e8(drawn)) >= PSNR_MIN


@pytest.mark.parametrize("fmt,ofmt", [(capi.OUT_RGBA8_UNORM, 0), (capi.OUT_RGBA8_SRGB, 1)])
def test_default_scene_texture_within_tolerance(oracle, default_volume, fmt, ofmt):
    """config 1 through the texture unit (fixed-point trilinear weights), both target encodings (Q3)."""
    W, H = 1280, 720
    ubo0, ubo1 = scene.default_scene()
    p = oracle_params(oracle, ubo0, ubo1)
    ref, _, _ = _oracle_frame(oracle, default_volume, oracle.VOL_RGBA8, p, W, H)
    ref8 = oracle.encode(ref, ofmt)
    with vr.Renderer(W, H, fmt) as r:
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        r.set_params(ubo0, ubo1)
        img = r.render()[0]
    assert img[..., 3].min() == 255
    assert max_abs8(img, ref8) <= TOL8
    assert psnr8(img, ref8) >= PSNR_MIN
    assert ref8[..., 0].max() > 40            # the cloud is actually there


@pytest.mark.parametrize("phi,theta", [(30.0, 0.0), (-77.0, 41.0), (180.0, 90.0), (12.3, -160.0)])
def test_rotated_model_random_volume(oracle, phi, theta):
    """A/D/W/S orbit (TestMain.cpp:177-184), non-cubic random volume, 16F target, frame time != 0."""
    W, H = 640, 360
    rng = np.random.default_rng(int(abs(phi) * 10))
    vol = rng.integers(0, 256, size=(24, 40, 56, 4), dtype=np.uint8)
    ubo0, ubo1 = scene.default_scene(phi, theta, 2.5)
    p = oracle_params(oracle, ubo0, ubo1)
    ref, st, _ = _oracle_frame(oracle, vol, oracle.VOL_RGBA8, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1)
        lin = r.render()[0]
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        tex = r.render()[0]
    assert st["covered"] > 1000
    assert np.array_equal(lin.view(np.uint32), ref.view(np.uint32))
    assert np.abs(tex - ref).max() <= 2.0 / 255.0
    with vr.Renderer(W, H, capi.OUT_RGBA16F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1)
        h = r.render()[0]
    assert np.array_equal(h.view(np.uint16), oracle.encode(ref, oracle.OUT_RGBA16F))


def test_media_scroll_std140_decoding(oracle):
    """A host that fills a real std140 block (144 B) animates the fetches; the reference's 140-byte
    C++ struct does not (quirk Q2) — both decode like the shader would read them."""
    W, H = 320, 180
    rng = np.random.default_rng(5)
    vol = rng.integers(0, 256, size=(32, 32, 32, 4), dtype=np.uint8)
    ubo0, ubo1_host = scene.default_scene(10.0, 5.0, 3.0)
    std140 = np.zeros(36, np.float32)
    std140[:19] = ubo1_host[:19]
    std140[20:36] = rng.uniform(-1, 1, 16).astype(np.float32)
    outs = []
    for u1 in (ubo1_host, std140):
        p = oracle_params(oracle, ubo0, u1)
        ref, _, _ = _oracle_frame(oracle, vol, oracle.VOL_RGBA8, p, W, H)
        with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
            r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
            r.set_params(ubo0, u1)
            img = r.render()[0]
        assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
        outs.append(img)
    assert not np.array_equal(outs[0], outs[1])
    # frame time lands in std140 padding: the image is time-invariant with the C++ layout
    ubo0b, ubo1b = scene.default_scene(10.0, 5.0, 77.0)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0b, ubo1b)
        assert np.array_equal(r.render()[0], outs[0])


def test_max_steps_and_literals(oracle, default_volume):
    """runtime maxSteps / step factor / scales (frag.glsl:29-32,42,63 as vr_params fields)."""
    W, H = 480, 270
    ubo0, ubo1 = scene.default_scene(33.0, 12.0)
    kw = dict(max_steps=256, step_scale=3.0, density=1.7, sample_scale=0.35)
    p = oracle_params(oracle, ubo0, ubo1, **kw)
    oracle.set_params(p, coord_scale=[1.0, 0.9, 0.6, 0.5])
    ref, st, _ = _oracle_frame(oracle, default_volume, oracle.VOL_RGBA8, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1, coord_scale=[1.0, 0.9, 0.6, 0.5], **kw)
        img = r.render(flags=capi.RENDER_STATS)[0]
        assert r.stats()["steps"] == st["steps"]
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        tex = r.render()[0]
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
    assert np.abs(tex - ref).max() <= 2.0 / 255.0


def test_row_bands_and_views(oracle):
    """row-band rendering (the multi-GPU partition) and a multi-view batch equal the full frames."""
    W, H = 320, 200
    rng = np.random.default_rng(11)
    vol = rng.integers(0, 256, size=(16, 16, 16, 4), dtype=np.uint8)
    views = scene.orbit_views(4)
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, max_views=4) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(view_ubos=views)
        full = r.render().copy()
        banded = np.zeros_like(full)
        bounds = r.row_bands(3)
        assert bounds[0] == 0 and bounds[-1] == H and all(b1 >= b0 for b0, b1 in zip(bounds, bounds[1:]))
        for b0, b1 in zip(bounds, bounds[1:]):
            if b1 > b0:
                r.render(out=banded, rows=(b0, b1))
        assert np.array_equal(full, banded)
        for k in range(4):
            u0 = views[k, :192].view(np.float32)
            u1 = views[k, 192:332].view(np.float32)
            p = oracle_params(oracle, u0, u1)
            ref, _, _ = _oracle_frame(oracle, vol, oracle.VOL_RGBA8, p, W, H)
            assert np.array_equal(full[k], oracle.encode(ref, oracle.OUT_RGBA8_UNORM))


@pytest.mark.parametrize("layout", [capi.LAYOUT_LINEAR, capi.LAYOUT_TEXTURE, capi.LAYOUT_BRICK])
def test_single_channel_fp16_grid(oracle, layout):
    """config 4 shape at test size: one-channel fp16 grid, one fetch per step."""
    W, H = 480, 270
    vol = smoke_volume(48, 3)
    ubo0, ubo1 = scene.default_scene(25.0, 15.0)
    p = oracle_params(oracle, ubo0, ubo1, density_mode=oracle.DENSITY_SINGLE, sample_scale=0.5, max_steps=192)
    ref, st, _ = _oracle_frame(oracle, vol.view(np.uint16), oracle.VOL_R16F, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_R16F, layout)
        r.set_params(ubo0, ubo1, density_mode=capi.DENSITY_SINGLE, sample_scale=0.5, max_steps=192)
        img = r.render()[0]
        r.set_params(ubo0, ubo1, density_mode=capi.DENSITY_SINGLE, sample_scale=0.5, max_steps=192, skip_empty=1)
        skipped = r.render(flags=capi.RENDER_STATS)[0]
        s = r.stats()
    assert ref[..., 0].max() > 0.3
    if layout in (capi.LAYOUT_LINEAR, capi.LAYOUT_BRICK):      # FP32 filtering in the oracle's order: bit-exact
        assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
    else:
        assert np.abs(img - ref).max() <= 2.0 / 255.0
    # empty-space skipping only ever skips samples that are exactly zero
    assert s["skipped_steps"] > 0
    # same samples; the ray-split schedule of the texture path re-associates the sum and takes positions as
    # Pin + i*stepVec (a 1e-5 shift can move a hardware filter weight by one 1/256 step)
    assert np.abs(skipped - img).max() <= (2e-4 if layout == capi.LAYOUT_TEXTURE else 1e-6)


def test_compositing_light_march_and_early_termination(oracle):
    """A.6 extensions vs the oracle's restatement of them (self-referential parity)."""
    W, H = 384, 216
    vol = smoke_volume(40, 9)
    ubo0, ubo1 = scene.default_scene(-20.0, 30.0)
    sun = np.array([1.0, 1.0, 2.0], np.float32)
    sun /= np.linalg.norm(sun)
    kw = dict(sample_scale=1.0, max_steps=128, sigma_t=8.0, light_steps=6, light_step_len=2 * 4.0 / 128,
              ambient=0.15)
    p = oracle_params(oracle, ubo0, ubo1, density_mode=oracle.DENSITY_SINGLE, **kw)
    oracle.set_params(p, sun_dir=sun, sun_color=[1.0, 0.95, 0.85])
    ref, st, _ = _oracle_frame(oracle, vol.view(np.uint16), oracle.VOL_R16F, p, W, H)
    oracle.set_params(p, early_term_eps=1.0 / 64)
    ref_et, st_et, _ = _oracle_frame(oracle, vol.view(np.uint16), oracle.VOL_R16F, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_R16F, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1, density_mode=capi.DENSITY_SINGLE, sun_dir=sun, sun_color=[1.0, 0.95, 0.85], **kw)
        img = r.render()[0]
        r.set_params(ubo0, ubo1, density_mode=capi.DENSITY_SINGLE, sun_dir=sun, sun_color=[1.0, 0.95, 0.85],
                     early_term_eps=1.0 / 64, **kw)
        img_et = r.render(flags=capi.RENDER_STATS)[0]
        s = r.stats()
    assert ref[..., :3].max() > 0.3
    assert np.abs(img - ref).max() <= 1e-4            # expf / FMA contraction vs libm, unfused
    assert np.abs(img_et - ref_et).max() <= 1e-4
    assert st_et["early"] > 0 and abs(s["early_terminated"] - st_et["early"]) <= max(2, st_et["early"] // 100)


def test_procedural_fbm_cloud(oracle):
    """config 2 shape at test size: in-kernel FastNoise2-semantics fBm density + sun light-march."""
    W, H = 320, 180
    ubo0, ubo1 = scene.default_scene()
    sun = np.array([1.0, 1.0, 2.0], np.float32)
    sun /= np.linalg.norm(sun)
    kw = dict(max_steps=64, sigma_t=8.0, light_steps=4, light_step_len=2 * 4.0 / 64, ambient=0.15)
    p = oracle_params(oracle, ubo0, ubo1, density_mode=oracle.DENSITY_FBM, **kw)
    oracle.set_params(p, sun_dir=sun, sun_color=[1.0, 0.95, 0.85])
    ref, st = oracle.render(None, p, W, H)
    ref8 = oracle.encode(ref, oracle.OUT_RGBA8_UNORM)
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM) as r:
        r.set_params(ubo0, ubo1, density_mode=capi.DENSITY_FBM, sun_dir=sun, sun_color=[1.0, 0.95, 0.85], **kw)
        img = r.render()[0]
    assert ref8[..., 0].max() > 60
    assert max_abs8(img, ref8) <= TOL8
    assert psnr8(img, ref8) >= PSNR_MIN


@pytest.mark.parametrize("eye,center,fov", [((1.9, 1.7, 1.6), (0.0, 0.0, 0.0), 45.0),      # the cube overflows the frame
                                            ((1.25, 0.3, 0.2), (0.0, 0.1, 0.0), 70.0),     # corners behind the near plane
                                            ((6.0, -5.0, 4.0), (1.5, 0.0, 0.5), 30.0),     # small, off-centre, partly off-screen
                                            ((0.4, 0.3, 0.2), (0.0, 0.0, 0.0), 45.0)])     # camera inside the cube: nothing drawn
def test_screen_rectangle_is_conservative(oracle, eye, center, fov):
    """The ray set-up is skipped outside the bounding rectangle of the cube's projected corners (whole frame when a corner
    is not in front of the camera).  Cameras the reference's orbit never reaches: the strict path must still equal the
    oracle bit for bit, i.e. no covered pixel was culled."""
    W, H = 320, 200
    rng = np.random.default_rng(11)
    vol = rng.integers(0, 256, size=(16, 16, 16, 4), dtype=np.uint8)
    ubo0, ubo1 = scene.default_scene(15.0, -20.0)
    ubo0[16:32] = scene.look_at(eye, center, (0.0, 0.0, 1.0))
    proj = scene.perspective(scene.radians(fov), np.float32(W) / np.float32(H), 0.1, 20.0)
    proj[5] *= np.float32(-1.0)
    ubo0[32:48] = proj
    ubo1[16:19] = eye
    p = oracle_params(oracle, ubo0, ubo1)
    ref, st, _ = _oracle_frame(oracle, vol, oracle.VOL_RGBA8, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1)
        img = r.render(flags=capi.RENDER_STATS)[0]
        s = r.stats()
    assert s["covered"] == st["covered"] and s["steps"] == st["steps"]
    assert (st["covered"] == 0) == (max(abs(c) for c in eye) < 1.0)
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))


def test_constant_channels_are_not_fetched(oracle, default_volume):
    """A channel that holds one value everywhere is replaced by that value in the hardware-filtered march.  The reference's own
    bake has one (G = 255, TestMain.cpp:59-60,76); arbitrary constants (B = 230, A = 64) must reproduce the oracle as well."""
    W, H = 320, 180
    ubo0, ubo1 = scene.default_scene(20.0, -10.0)
    rng = np.random.default_rng(3)
    vol = rng.integers(0, 256, size=(24, 40, 32, 4), dtype=np.uint8)
    vol[..., 2] = 230
    vol[..., 3] = 64
    p = oracle_params(oracle, ubo0, ubo1)
    ovol, _keep = oracle.make_volume(vol, oracle.VOL_RGBA8)
    ref8 = oracle.encode(oracle.render(ovol, p, W, H)[0], oracle.OUT_RGBA8_UNORM)
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM) as r:
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        assert r.stats()["constant_channels"] == 0b0010
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        assert r.stats()["constant_channels"] == 0b1100
        r.set_params(ubo0, ubo1)
        img = r.render()[0]
        assert ref8[..., 0].max() > 15
        assert max_abs8(img, ref8) <= TOL8 and psnr8(img, ref8) >= PSNR_MIN
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)        # the strict path fetches everything
        assert r.stats()["constant_channels"] == 0


def test_copy_only_async_of_a_second_frame(default_volume):
    """What the multi-GPU gather pipelines on GPU 0: a march stores straight into frame slot 1 of the context
    (VR_RENDER_DIRECT), then a copy-only asynchronous vr_render hands that slot to the copy stream; vr_wait completes it."""
    W, H = 256, 144
    ubo0, ubo1 = scene.default_scene(10.0, 5.0)
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, max_views=2) as r:
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        r.set_params(ubo0, ubo1)
        ref = r.render()[0].copy()
        ptr, pitch, vstride = r.frame_info()
        assert vstride == pitch * H
        host = r.new_frame_buffer(views=2, pinned=True)
        host[:] = 7
        r.render_device(ptr + vstride, rows=(0, H), flags=capi.RENDER_DIRECT | capi.RENDER_ASYNC)
        got = r.render(out=host, views=(1, 2), flags=capi.RENDER_NO_MARCH | capi.RENDER_ASYNC)
        r.wait()
        assert np.array_equal(got[0], ref) and (host[0] == 7).all()
        with pytest.raises(Exception):                       # a marching call may only name views that have parameters
            r.render(out=host, views=(1, 2))


def test_device_side_gather_protocol_single_rank(default_volume):
    """vr_gather_acquire / _signal / _collect with one rank: six frames through two slots with no host synchronisation in
    between (slot reuse waits for the release on the device), every delivered frame equal to an ordinary render."""
    W, H = 256, 144
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, max_views=2) as r:
        r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        ptr, pitch, vstride = r.frame_info()
        host = r.new_frame_buffer(views=6, pinned=True)
        refs = []
        for i in range(6):
            r.set_params(*scene.default_scene(10.0 * i, 5.0))
            refs.append(r.render()[0].copy())
        for i in range(6):
            slot = i & 1
            r.set_params(*scene.default_scene(10.0 * i, 5.0))
            r.gather_acquire(None, 2, slot, i // 2)
            r.render_device(ptr + slot * vstride, rows=(0, H // 2), flags=capi.RENDER_DIRECT | capi.RENDER_ASYNC)
            r.render_device(ptr + slot * vstride, rows=(H // 2, H), flags=capi.RENDER_DIRECT | capi.RENDER_ASYNC)
            r.gather_signal(None, 2, slot)
            r.gather_collect(slot, i // 2 + 1, host[i])
        r.wait()
        for i in range(6):
            assert np.array_equal(host[i], refs[i]), i
        with pytest.raises(Exception):
            r.gather_signal(None, 2, 2)                      # only two slots


@pytest.mark.parametrize("light_steps", [6, 5, 1])
def test_fbm_light_march_pairs_equal_the_scalar_loop(monkeypatch, light_steps):
    """The fBm light-march evaluates its samples two at a time on packed FP32 (FFMA2); each half does the
    scalar routine's operations in the same order, so one thread per ray must give the very same binary32 frame as
    the counting build (VR_RENDER_STATS), which keeps the scalar loop — for even and odd sample counts."""
    monkeypatch.setenv("VOLRT_RAY_SPLIT_COMPOSITE", "1")
    monkeypatch.setenv("VOLRT_RAY_SEGMENTS", "1")          # one thread per WHOLE ray: depth pieces re-associate the compositing
    W, H = 320, 180
    ubo0, ubo1 = scene.default_scene(25.0, -15.0)
    sun = np.array([1.0, 1.0, 2.0], np.float32)
    sun /= np.linalg.norm(sun)
    kw = dict(max_steps=48, sigma_t=8.0, light_steps=light_steps, light_step_len=2 * 4.0 / 48, ambient=0.15)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.set_params(ubo0, ubo1, density_mode=capi.DENSITY_FBM, sun_dir=sun, sun_color=[1.0, 0.95, 0.85], **kw)
        packed = r.render()[0].copy()
        scalar = r.render(flags=capi.RENDER_STATS)[0].copy()
        s = r.stats()
    assert s["light_samples"] > 0 and packed[..., :3].max() > 0.2
    assert np.array_equal(packed.view(np.uint32), scalar.view(np.uint32))


@pytest.mark.parametrize("octaves,freq,lac,falloff1,eligible", [
    (5, 128 * 0.03, 2.0, 0.5, True),        # the C2 / C3 field
    (3, 7.3, 1.7, 0.45, True),              # ragged lattice sizes, the three-octave loop
    (8, 1.1, 2.0, 0.5, True),               # the most octaves the tables take (lattice up to 143^3)
    (9, 0.5, 2.0, 0.5, False),              # too many octaves -> per-sample hash
    (5, 128 * 0.03, 2.0, 0.6, False),       # fade reaches beyond the box -> per-sample hash
    (2, 120.0, 2.0, 0.5, False),            # lattice larger than the float-encoded index allows -> per-sample hash
])
def test_fbm_cell_tables_equal_the_per_sample_hash(monkeypatch, octaves, freq, lac, falloff1, eligible):
    """MARCH_FBM_CELLS reads, per lattice cell, the eight gradient-table offsets the hash would have produced; everything after
    the hash is the same arithmetic, so the frame is bit-identical to the hash variant's — whichever of the two the field
    is eligible for (ineligible fields must simply keep working)."""
    W, H = 320, 180
    u0, u1 = scene.default_scene(15.0, 25.0)
    SUN = np.array([1.0, 1.0, 2.0], np.float32) / np.float32(np.sqrt(6.0))
    kw = dict(density_mode=2, max_steps=64, sigma_t=8.0, light_steps=5, light_step_len=0.09, ambient=0.15, sun_dir=SUN,
              sun_color=[1.0, 0.95, 0.85], fbm_octaves=octaves, fbm_frequency=freq, fbm_lacunarity=lac, fbm_falloff1=falloff1, fbm_seed=77)
    frames = {}
    for cells in ("1", "0"):
        monkeypatch.setenv("VOLRT_FBM_CELLS", cells)
        for split in (1, 2):                                  # one thread per ray, and two lanes per ray
            monkeypatch.setenv("VOLRT_RAY_SPLIT_COMPOSITE", str(split))
            with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
                r.set_params(u0, u1, **kw)
                frames[cells, split] = r.render()[0].copy()
    for split in (1, 2):
        assert np.array_equal(frames["1", split].view(np.uint32), frames["0", split].view(np.uint32)), (split, eligible)
    assert frames["1", 1][..., :3].max() > 0.05


@pytest.mark.parametrize("mode", ["fbm", "grid"])
def test_depth_segments_compose_to_the_whole_ray(monkeypatch, oracle, mode):
    """The compositing march may cut a ray's steps into K contiguous pieces marched by different CTAs and compose them with
    (C_a + T_a C_b, T_a T_b).  Every K gives the frame of K = 1 up to re-association and the pieces' first positions being
    Pin + i*stepVec rather than a running sum (1e-4 in float, 1/80 of the 2/255 bar), which in turn is the
    oracle's frame; a second frame through the same context reuses the arrival counters the first one reset."""
    W, H = 333, 187
    u0, u1 = scene.default_scene(35.0, -20.0)
    sun = np.array([1.0, 1.0, 2.0], np.float32) / np.float32(np.sqrt(6.0))
    kw = dict(max_steps=96, sigma_t=8.0, light_steps=3, light_step_len=0.08, ambient=0.15, sun_dir=sun, sun_color=[1.0, 0.95, 0.85])
    if mode == "fbm":
        kw.update(density_mode=capi.DENSITY_FBM)
        vol, ovol = None, None
    else:
        kw.update(density_mode=capi.DENSITY_SINGLE, sample_scale=1.0)
        vol = smoke_volume(24, 3)
        ovol, _k = oracle.make_volume(vol.view(np.uint16), oracle.VOL_R16F)
    scal = {k: v for k, v in kw.items() if not hasattr(v, "__len__")}
    p = oracle_params(oracle, u0, u1, **scal)
    oracle.set_params(p, sun_dir=sun, sun_color=[1.0, 0.95, 0.85])
    ref, _ = oracle.render(ovol, p, W, H)
    monkeypatch.setenv("VOLRT_RAY_SPLIT_COMPOSITE", "1")
    frames = {}
    for k in (1, 2, 3, 8):
        monkeypatch.setenv("VOLRT_RAY_SEGMENTS", str(k))
        with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
            if vol is not None:
                r.upload_volume(vol, capi.VOL_R16F, capi.LAYOUT_LINEAR)
            r.set_params(u0, u1, **kw)
            a = r.render()[0].copy()
            b = r.render(rows=(40, 131))[0].copy()                 # a band: other scratch shape, counters must be clean
            c = r.render()[0].copy()
        assert np.array_equal(a, c) and np.array_equal(a[40:131], b)
        frames[k] = a
    assert frames[1][..., :3].max() > 0.1
    for k in (2, 3, 8):
        assert np.abs(frames[k] - frames[1]).max() <= 1e-4, k
    assert np.abs(frames[8] - ref).max() <= (2e-4 if mode == "grid" else 2.0 / 255.0)


def test_known_answers_on_gpu():
    """closed forms that do not depend on the oracle: constant volume, axis-aligned centre ray."""
    W, H = 256, 256
    vol = np.full((8, 8, 8, 4), 255, np.uint8)
    # camera on the +X axis looking at the origin; square aspect
    ubo0 = np.zeros(48, np.float32)
    ubo0[0:16] = scene.identity()
    ubo0[16:32] = scene.look_at((4.0, 0.0, 0.0), (0.0, 0.0, 0.0), (0.0, 0.0, 1.0))
    proj = scene.perspective(scene.radians(60.0), 1.0, 0.1, 10.0)
    proj[5] *= -1
    ubo0[32:48] = proj
    ubo1 = np.zeros(36, np.float32)
    ubo1[0:16] = scene.identity()
    ubo1[16:19] = (4.0, 0.0, 0.0)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        r.set_params(ubo0, ubo1)
        img = r.render()[0]
    # rho = (1*1)*(1+1)*0.2 = 0.4 per step; chord 2 / stepSize 0.03125 = 64 steps on the axis
    c = img[H // 2, W // 2, 0]
    n = np.round(-np.log1p(-c) / (0.4 * 0.03125))
    assert n in (63.0, 64.0)
    assert abs(c - (1 - np.exp(-0.4 * n * 0.03125))) < 1e-5
    assert np.all(img[..., 3] == 1.0)
    assert np.all(img[0, 0, :3] == 0.0)           # clear colour outside the silhouette
    assert np.array_equal(img[..., 0], img[..., 1]) and np.array_equal(img[..., 0], img[..., 2])


@pytest.mark.parametrize("steps,phi,theta", [(96, 0.0, 0.0), (384, 40.0, -25.0), (700, 90.0, 90.0)])
def test_tma_brick_march_is_bit_exact(oracle, steps, phi, theta):
    """VR_LAYOUT_BRICK: boxes staged by TMA; sparse stepping (boxes shrink to fit), voxel-rate stepping, and an
    axis-aligned view; non-cubic volume whose faces exercise the mirrored-tap fallback."""
    W, H = 400, 240
    rng = np.random.default_rng(steps)
    vol = (rng.random((40, 56, 64), np.float32) ** 3).astype(np.float16)        # (d, h, w), w % 8 == 0
    ubo0, ubo1 = scene.default_scene(phi, theta)
    kw = dict(density_mode=1, sample_scale=0.7, max_steps=steps)
    p = oracle_params(oracle, ubo0, ubo1, density_mode=oracle.DENSITY_SINGLE, sample_scale=0.7, max_steps=steps)
    ref, st, _ = _oracle_frame(oracle, vol.view(np.uint16), oracle.VOL_R16F, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_R16F, capi.LAYOUT_BRICK)
        r.set_params(ubo0, ubo1, **kw)
        img = r.render()[0]
    assert st["covered"] > 5000 and ref[..., 0].max() > 0.2
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        with pytest.raises(vr.VolrtError):
            r.upload_volume(np.zeros((8, 8, 12), np.float16), capi.VOL_R16F, capi.LAYOUT_BRICK)    # width % 8 != 0
