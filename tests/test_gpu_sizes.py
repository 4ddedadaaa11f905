"""Edge sizes and BASELINE.json's full sizes, through the C ABI.

Small/ragged frames check every kernel mapping's bounds handling against the oracle; the full-size
configurations (1080p / 4K procedural cloud, 512^3 fp16 grid) are checked on a strided sample of rows
against the oracle plus size-independent properties: the union of row bands is the frame, rendering is
idempotent, and a view batch equals its views rendered one by one."""
import numpy as np
import pytest

import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
from util import psnr8, max_abs8, oracle_params, smoke_volume

pytestmark = pytest.mark.gpu
SUN = np.array([1.0, 1.0, 2.0], np.float32) / np.float32(np.sqrt(6.0))


@pytest.mark.parametrize("W,H", [(1, 1), (7, 3), (33, 17), (331, 173), (16, 8), (130, 66)])
def test_ragged_frames_every_mapping(oracle, W, H):
    rng = np.random.default_rng(W * 1000 + H)
    vol = rng.integers(0, 256, size=(12, 20, 16, 4), dtype=np.uint8)
    u0, u1 = scene.default_scene(17.0, -8.0)
    p = oracle_params(oracle, u0, u1)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    ref, st = oracle.render(ov, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(u0, u1)
        assert np.array_equal(r.render()[0].view(np.uint32), ref.view(np.uint32))          # direct mapping, strict
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        assert np.abs(r.render()[0] - ref).max() <= 2.0 / 255.0                             # ray-split texture mapping
    # one-channel grid: TMA brick mapping (bit-exact) and the ray-split compositing mapping
    g = smoke_volume(16, 1)
    ps = oracle_params(oracle, u0, u1, density_mode=oracle.DENSITY_SINGLE, sample_scale=0.5)
    og, _k2 = oracle.make_volume(g.view(np.uint16), oracle.VOL_R16F)
    refs, _ = oracle.render(og, ps, W, H)
    pc = oracle_params(oracle, u0, u1, density_mode=oracle.DENSITY_SINGLE, sample_scale=1.0, sigma_t=6.0, light_steps=3,
                       light_step_len=0.06, ambient=0.1)
    oracle.set_params(pc, sun_dir=SUN, sun_color=[1.0, 0.9, 0.8])
    refc, _ = oracle.render(og, pc, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(g, capi.VOL_R16F, capi.LAYOUT_BRICK)
        r.set_params(u0, u1, density_mode=capi.DENSITY_SINGLE, sample_scale=0.5)
        assert np.array_equal(r.render()[0].view(np.uint32), refs.view(np.uint32))
        r.set_params(u0, u1, density_mode=capi.DENSITY_SINGLE, sample_scale=1.0, sigma_t=6.0, light_steps=3, light_step_len=0.06,
                     ambient=0.1, sun_dir=SUN, sun_color=[1.0, 0.9, 0.8])
        assert np.abs(r.render()[0] - refc).max() <= 1e-4


@pytest.mark.parametrize("shape", [(5, 5, 5), (27, 13, 10), (3, 3, 3), (1, 1, 1), (7, 1, 2)])
def test_rgba8_texture_layout_with_a_voxel_count_not_divisible_by_four(oracle, shape):
    """The planar copies of an RGBA8 volume are written as 32-bit words per plane; plane k starts at k * voxels bytes, which is
    word aligned only when voxels % 4 == 0.  Odd volumes must take the byte path (a misaligned store is a sticky fault)."""
    W, H = 96, 54
    rng = np.random.default_rng(shape[0] * 100 + shape[1] * 10 + shape[2])
    vol = rng.integers(0, 256, size=shape + (4,), dtype=np.uint8)
    vol[..., 1] = 200                                   # a constant channel: exercises the per-channel min/max of the byte path
    u0, u1 = scene.default_scene(11.0, 23.0)
    p = oracle_params(oracle, u0, u1)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    ref, _ = oracle.render(ov, p, W, H)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
        assert r.stats()["constant_channels"] & 2
        r.set_params(u0, u1)
        a = r.render()[0].copy()
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)     # same shape again: storage is reused
        b = r.render()[0].copy()
    assert np.abs(a - ref).max() <= 2.0 / 255.0
    assert np.array_equal(a, b)


def test_r8_volume_and_all_encodings_on_split_kernels(oracle):
    W, H = 200, 120
    rng = np.random.default_rng(3)
    v8 = rng.integers(0, 256, size=(24, 24, 24), dtype=np.uint8)
    u0, u1 = scene.default_scene(-30.0, 12.0)
    p = oracle_params(oracle, u0, u1, density_mode=oracle.DENSITY_SINGLE, sample_scale=0.3)
    ov, _k = oracle.make_volume(v8, oracle.VOL_R8)
    ref, _ = oracle.render(ov, p, W, H)
    for fmt, ofmt in ((capi.OUT_RGBA8_UNORM, oracle.OUT_RGBA8_UNORM), (capi.OUT_RGBA8_SRGB, oracle.OUT_RGBA8_SRGB)):
        with vr.Renderer(W, H, fmt) as r:
            r.upload_volume(v8, capi.VOL_R8_UNORM, capi.LAYOUT_LINEAR)
            r.set_params(u0, u1, density_mode=capi.DENSITY_SINGLE, sample_scale=0.3)
            assert np.array_equal(r.render()[0], oracle.encode(ref, ofmt))
            r.upload_volume(v8, capi.VOL_R8_UNORM, capi.LAYOUT_TEXTURE)
            assert max_abs8(r.render()[0], oracle.encode(ref, ofmt)) <= 2
    with vr.Renderer(W, H, capi.OUT_RGBA16F) as r:
        r.upload_volume(v8, capi.VOL_R8_UNORM, capi.LAYOUT_TEXTURE)
        r.set_params(u0, u1, density_mode=capi.DENSITY_SINGLE, sample_scale=0.3)
        h = r.render()[0].astype(np.float32)
        assert np.abs(h - ref).max() <= 2.0 / 255.0 + 1e-3


def test_clamp_address_mode_texture_vs_linear():
    """CLAMP_TO_EDGE is not in the reference (its sampler mirrors); the two GPU samplers must agree with each other"""
    W, H = 160, 90
    rng = np.random.default_rng(8)
    vol = rng.integers(0, 256, size=(8, 8, 8, 4), dtype=np.uint8)
    u0, u1 = scene.default_scene(5.0, 5.0)
    std140 = np.zeros(36, np.float32)
    std140[:19] = u1[:19]
    std140[20:32] = 3.0                     # 0.8 P + 0.6 etc.: fetches 2-4 cross the border
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR, capi.ADDRESS_CLAMP_TO_EDGE)
        r.set_params(u0, std140)
        a = r.render()[0].copy()
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE, capi.ADDRESS_CLAMP_TO_EDGE)
        b = r.render()[0].copy()
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR, capi.ADDRESS_MIRRORED_REPEAT)
        c = r.render()[0].copy()
    assert np.abs(a - b).max() <= 2.0 / 255.0 and a[..., 0].max() > 0.05
    assert np.abs(a - c).max() > 1e-3       # the address mode matters for these coordinates


def _c2_kw(max_steps, light):
    return dict(density_mode=2, max_steps=max_steps, sigma_t=8.0, light_steps=light, light_step_len=2 * 4.0 / max_steps,
                ambient=0.15, sun_dir=SUN, sun_color=[1.0, 0.95, 0.85])


@pytest.mark.parametrize("W,H,steps,light", [(1920, 1080, 128, 6), (3840, 2160, 256, 8)])
def test_full_size_procedural_cloud(oracle, W, H, steps, light):
    """BASELINE.json configs[1] and configs[2] at full size: oracle parity on a strided row sample (north-star
    tolerance), bands == frame, idempotence."""
    u0, u1 = scene.default_scene()
    kw = _c2_kw(steps, light)
    rows = np.arange(H // 12, H, H // 6, dtype=np.int32)[:5]
    scal = {k: v for k, v in kw.items() if not hasattr(v, "__len__")}
    p = oracle_params(oracle, u0, u1, **scal)
    oracle.set_params(p, sun_dir=SUN, sun_color=[1.0, 0.95, 0.85])
    ref, st = oracle.render_rows(None, p, W, H, rows)
    ref8 = oracle.encode(ref, oracle.OUT_RGBA8_UNORM)
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM) as r:
        r.set_params(u0, u1, **kw)
        img = r.render()[0].copy()
        again = r.render()[0].copy()
        bands = np.zeros_like(img)[None]
        b = r.row_bands(5)
        for b0, b1 in zip(b, b[1:]):
            if b1 > b0:
                r.render(out=bands, rows=(b0, b1))
    assert np.array_equal(img, again)
    # a band of a different height may pick a different lanes-per-ray mapping: equal within re-association noise
    assert max_abs8(img, bands[0]) <= 1
    assert st["covered"] > 1000 and ref8[..., 0].max() > 60
    assert max_abs8(img[rows], ref8) <= 2
    assert psnr8(img[rows], ref8) >= 50.0
    assert img[..., 3].min() == 255


def test_full_size_fp16_grid_512(oracle):
    """BASELINE.json configs[3]: 512^3 fp16 grid at 1080p; the device-baked grid is downloaded so that both sides
    sample the same bytes; texture path within tolerance, TMA-brick path bit-exact, on a strided row sample."""
    W, H, n = 1920, 1080, 512
    u0, u1 = scene.default_scene()
    rows = np.array([300, 540, 800], np.int32)
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.generate_volume(size=n, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE)
        vol = r.download_volume()
        r.set_params(u0, u1, density_mode=capi.DENSITY_SINGLE, max_steps=512, sample_scale=4.0)
        tex = r.render()[0][rows].copy()
        r.upload_volume(vol, capi.VOL_R16F, capi.LAYOUT_BRICK)
        brick = r.render()[0][rows].copy()
    p = oracle_params(oracle, u0, u1, density_mode=oracle.DENSITY_SINGLE, max_steps=512, sample_scale=4.0)
    ov, _k = oracle.make_volume(vol.view(np.uint16), oracle.VOL_R16F)
    ref, st = oracle.render_rows(ov, p, W, H, rows)
    assert ref[..., 0].max() > 0.3
    assert np.array_equal(brick.view(np.uint32), ref.view(np.uint32))
    assert np.abs(tex - ref).max() <= 2.0 / 255.0


def test_view_batch_equals_single_views(oracle):
    """C5 shape at test size: skipping + early termination over a view batch == each view alone"""
    W, H = 320, 180
    g = smoke_volume(64, 5)
    g[g < np.float16(0.4)] = 0
    views = scene.orbit_views(6)
    kw = dict(density_mode=1, max_steps=256, sample_scale=1.0, sigma_t=30.0, early_term_eps=1.0 / 512, skip_empty=1,
              ambient=1.0, sun_color=[0.0, 0.0, 0.0])
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, max_views=6) as r:
        r.upload_volume(g, capi.VOL_R16F, capi.LAYOUT_TEXTURE)
        r.set_params(view_ubos=views, **kw)
        batch = r.render().copy()
        some = r.render(views=(2, 5)).copy()
        r.render(flags=capi.RENDER_STATS)
        s = r.stats()
    assert np.array_equal(batch[2:5], some)
    assert s["skipped_steps"] > 0 and s["early_terminated"] > 0
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM) as r:
        r.upload_volume(g, capi.VOL_R16F, capi.LAYOUT_TEXTURE)
        for k in (0, 3, 5):
            r.set_params(views[k, :192].view(np.float32), views[k, 192:332].view(np.float32), **kw)
            assert np.array_equal(r.render()[0], batch[k])
    # and against the oracle (which neither skips nor splits rays)
    scal = {k: v for k, v in kw.items() if not hasattr(v, "__len__") and k != "skip_empty"}
    p = oracle_params(oracle, views[1, :192].view(np.float32), views[1, 192:332].view(np.float32), **scal)
    oracle.set_params(p, sun_color=[0.0, 0.0, 0.0])
    ov, _k = oracle.make_volume(g.view(np.uint16), oracle.VOL_R16F)
    ref, _ = oracle.render(ov, p, W, H)
    assert max_abs8(batch[1], oracle.encode(ref, oracle.OUT_RGBA8_UNORM)) <= 2


@pytest.mark.parametrize("fmt,shape", [(capi.VOL_R16F, (40, 24, 64)), (capi.VOL_R16F, (17, 23, 29)),
                                        (capi.VOL_R8_UNORM, (32, 40, 48)), (capi.VOL_R8_UNORM, (9, 31, 50)),
                                        (capi.VOL_RGBA8_UNORM, (24, 16, 36)), (capi.VOL_RGBA8_UNORM, (10, 13, 27))])
def test_occupancy_grid_aprons(fmt, shape):
    """Empty-space skipping must not change a single sample: isolated non-zero texels on and next to brick borders
    (x, y, z = 7, 8 mod 8, first and last texel) light up the neighbouring bricks through the 1-texel apron.  Row
    pitches that are and are not multiples of 16 bytes take the vector and the scalar path of the occupancy kernel."""
    d, h, w = shape
    rng = np.random.default_rng(d * 10000 + h * 100 + w)
    ch = 4 if fmt == capi.VOL_RGBA8_UNORM else 1
    vol = np.zeros((d, h, w, ch), np.float32)
    pts = [(0, 0, 0), (d - 1, h - 1, w - 1), (7, 7, 7), (8, 8, 8), (7, 8, 15), (d // 2, 7, 16), (8, h // 2, 23), (d - 1, 0, 8)]
    pts += [tuple(int(rng.integers(0, n)) for n in (d, h, w)) for _ in range(6)]
    for z, y, x in pts:
        vol[min(z, d - 1), min(y, h - 1), min(x, w - 1), :] = 1.0
    if fmt == capi.VOL_R16F:
        texels = vol[..., 0].astype(np.float16)
    elif fmt == capi.VOL_R8_UNORM:
        texels = (vol[..., 0] * 255).astype(np.uint8)
    else:
        texels = (vol * 255).astype(np.uint8)
    W, H = 320, 200
    u0, u1 = scene.default_scene(31.0, 12.0)
    mode = capi.DENSITY_SINGLE if ch == 1 else 0
    with vr.Renderer(W, H, capi.OUT_RGBA32F) as r:
        r.upload_volume(texels, fmt, capi.LAYOUT_LINEAR)
        r.set_params(u0, u1, density_mode=mode, sample_scale=0.25, max_steps=400, density=40.0)
        plain = r.render()[0].copy()
        r.set_params(u0, u1, density_mode=mode, sample_scale=0.25, max_steps=400, density=40.0, skip_empty=1)
        skipped = r.render(flags=capi.RENDER_STATS)[0]
        s = r.stats()
    assert plain[..., 0].max() > 0.0
    if ch == 1:
        assert s["skipped_steps"] > 0
        assert np.array_equal(plain.view(np.uint32), skipped.view(np.uint32))


def test_c5_machinery_properties_at_256():
    """BASELINE config 5 at a size that runs in a second: a thresholded fp16 smoke grid, compositing with early termination,
    several views.  Size-independent properties of the texture-filtered path: (1) empty-space skipping does not change the
    frame; (2) the march with two steps' fetches in flight and the brick index taken from the float's bits equals the
    counting build, which walks one step at a time with the truncating index; (3) early termination only ever removes
    light, by less than its threshold per channel.  (1) and (2) hold to 3e-3 (under half of 2/255) rather than to the bit: a
    leap, and the several-lanes-per-ray kernel that serves frames without early termination, place a sample with one FMA
    where the plain loop adds step by step — an ulp the 8-bit filter weights may or may not see, amplified by sigma_t = 400."""
    W, H, NV = 480, 270, 4
    views = scene.orbit_views(64)[::16][:NV]
    kw = dict(density_mode=1, max_steps=256, sample_scale=1.0, sigma_t=400.0, ambient=1.0, sun_color=[0.0, 0.0, 0.0])
    with vr.Renderer(W, H, capi.OUT_RGBA32F, max_views=NV) as r:
        r.generate_volume(size=256, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=0.07)
        frames = {}
        for skip in (0, 1):
            for eps in (0.0, 1.0 / 512):
                r.set_params(view_ubos=views, skip_empty=skip, early_term_eps=eps, **kw)
                fast = r.render().copy()
                counted = r.render(flags=capi.RENDER_STATS).copy()
                st = r.stats()
                assert np.abs(fast - counted).max() <= 3e-3, (skip, eps)
                assert st["covered"] > 0 and (st["skipped_steps"] > 0) == bool(skip)
                assert (st["early_terminated"] > 0) == (eps > 0)
                frames[(skip, eps)] = fast
        for eps in (0.0, 1.0 / 512):
            assert np.abs(frames[(0, eps)] - frames[(1, eps)]).max() <= 3e-3
        full, cut = frames[(1, 0.0)][..., :3], frames[(1, 1.0 / 512)][..., :3]
        assert (cut <= full + 1e-7).all() and (full - cut).max() <= 1.0 / 512 + 1e-6 and full.max() > 0.5


def test_c5_at_its_stated_size_against_the_oracle(oracle):
    """BASELINE.json configs[4] as benchmarked: 1024^3 fp16 grid (2 GiB, device-baked, thresholded at 0.07), sigma_t = 400,
    512 steps, empty-space skipping + early termination at 1/512, 1080p, three of the 64 orbit views — a strided sample of rows
    of each against the oracle marching the very same texels (downloaded from the device): <= 2/255 and PSNR >= 50 dB.
    The oracle neither skips nor batches its fetches; it stops at the same transmittance."""
    import bench
    W, H, _desc, kw = bench.WORKLOADS["c5"]
    pick = [0, 23, 47]
    views = scene.orbit_views(bench.C5_VIEWS)[pick]
    rows = np.arange(H // 16, H, H // 8, dtype=np.int32)[:8]
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, max_views=len(pick)) as r:
        r.generate_volume(size=bench.C5_GRID, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=bench.C5_THRESHOLD)
        r.set_params(view_ubos=views, **kw)
        frames = r.render().copy()
        r.render(flags=capi.RENDER_STATS)
        st = r.stats()
        grid = r.download_volume()                                   # 2 GiB: the texels the device marched
    assert grid.shape == (bench.C5_GRID,) * 3 and st["skipped_steps"] > 0.3 * st["steps"] and st["early_terminated"] > 0.2 * st["covered"]
    ov, _keep = oracle.make_volume(grid.view(np.uint16), oracle.VOL_R16F)
    scal = {k: v for k, v in kw.items() if not hasattr(v, "__len__") and k != "skip_empty"}
    worst, psnrs = 0, []
    for k in range(len(pick)):
        u0 = np.frombuffer(views[k, :192].tobytes(), np.float32)
        u1 = np.frombuffer(views[k, 192:192 + 144].tobytes(), np.float32)
        p = oracle_params(oracle, u0, u1, **scal)
        oracle.set_params(p, sun_color=kw["sun_color"])
        ref, _ = oracle.render_rows(ov, p, W, H, rows)
        ref8 = oracle.encode(ref, oracle.OUT_RGBA8_UNORM)
        got = frames[k][rows]
        assert ref8[..., :3].max() > 100                                  # dense smoke: the rows are not empty
        worst = max(worst, max_abs8(got, ref8))
        psnrs.append(psnr8(got, ref8))
    assert worst <= 2 and min(psnrs) >= 50.0, (worst, psnrs)


def test_ray_compaction_kernel_gives_the_same_frame(monkeypatch):
    """VOLRT_COMPACT=1: the early-terminating grid march with ballot-ranked re-dealing of live rays inside a CTA's tile
    (march_compact_kernel; measured slower than the plain mapping, so not the default).  Same loop body, same positions:
    the frames are identical."""
    W, H, NV = 480, 270, 4
    views = scene.orbit_views(64)[::16][:NV]
    kw = dict(density_mode=1, max_steps=256, sample_scale=1.0, sigma_t=400.0, ambient=1.0, sun_color=[0.0, 0.0, 0.0], skip_empty=1,
              early_term_eps=1.0 / 512)
    frames = {}
    monkeypatch.setenv("VOLRT_RAY_SPLIT_COMPOSITE", "1")      # one thread per ray in the plain mapping too (a frame this small would get two lanes)
    with vr.Renderer(W, H, capi.OUT_RGBA32F, max_views=NV) as r:
        r.generate_volume(size=256, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=0.07)
        r.set_params(view_ubos=views, **kw)
        for knob in ("0", "1"):
            monkeypatch.setenv("VOLRT_COMPACT", knob)
            frames[knob] = r.render().copy()
    assert frames["0"][..., :3].max() > 0.5
    assert np.array_equal(frames["0"].view(np.uint32), frames["1"].view(np.uint32))


@pytest.mark.parametrize("fmt", ["OUT_RGBA8_UNORM", "OUT_RGBA16F", "OUT_RGBA32F"])
def test_view_batch_over_its_screen_rectangle_only_gives_the_same_frames(monkeypatch, fmt):
    """A batch of views launches tiles over the union of the views' screen rectangles and clears the rest of the frames with
    clear_outside_kernel (launch_one; by default only where that saves >= 100 000 CTAs — forced here).  Every pixel of every
    view, in all three target formats and for an odd frame width, equals the full-grid launch bit for bit; so does a row band."""
    W, H, NV = 482, 270, 4
    views = scene.orbit_views(64)[::16][:NV]
    kw = dict(density_mode=1, max_steps=256, sample_scale=1.0, sigma_t=400.0, ambient=1.0, sun_color=[0.0, 0.0, 0.0], skip_empty=1,
              early_term_eps=1.0 / 512)
    monkeypatch.setenv("VOLRT_RAY_SPLIT_COMPOSITE", "1")
    frames = {}
    with vr.Renderer(W, H, getattr(capi, fmt), max_views=NV) as r:
        r.generate_volume(size=128, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=0.07)
        r.set_params(view_ubos=views, **kw)
        for knob in ("1000000000", "0"):
            monkeypatch.setenv("VOLRT_RECT_GRID_MIN_SAVED", knob)
            frames[knob] = r.render().copy()
            launches = r.stats()["kernel_launches"]
            assert launches == (2 if knob == "0" else 1)
            r.render(rows=(40, 201))                       # a band that cuts the rectangle: only its rows may change
            frames[knob + "band"] = r.render(flags=capi.RENDER_NO_MARCH).copy()
    full, rect = frames["1000000000"], frames["0"]
    assert (full[..., :3].reshape(NV, -1).max(axis=1) > 0).all()
    assert np.array_equal(full.view(np.uint8), rect.view(np.uint8))
    assert np.array_equal(frames["1000000000band"].view(np.uint8), frames["0band"].view(np.uint8))
