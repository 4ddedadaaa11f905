"""Several GPUs behind ONE context in ONE process (vr_create_info.device_count): the reference host is a single process with a
single main() (source/Etna/TestMain.cpp:41,173-256), so that is the shape a drop-in must offer.  The frame of a group context
must be the frame of a single-GPU context: its row bands / view shards are rendered by different devices through peer
stores into devices[0]'s frame, each band possibly with another lanes-per-ray / depth-piece mapping (re-association only:
<= 1/255).  Needs >= 2 GPUs with peer access; skipped otherwise (the driver's 8-GPU scaling run exercises it through
bench.py --single-process)."""
import ctypes as C

import numpy as np
import pytest

import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
from util import max_abs8

pytestmark = pytest.mark.gpu
SUN = np.array([1.0, 1.0, 2.0], np.float32) / np.float32(np.sqrt(6.0))


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs_two = pytest.mark.skipif(_gpu_count() < 2, reason="needs two GPUs")


def test_group_argument_checks():
    lib = capi.load()
    ctx = C.c_void_p()
    info = capi.vr_create_info(C.sizeof(capi.vr_create_info), 0, 64, 64, 0, 1, 0)
    info.device_count = 2
    info.devices[0], info.devices[1] = 0, 0
    assert lib.vr_create(C.byref(info), C.byref(ctx)) == capi.VR_ERR_INVALID_ARGUMENT and b"twice" in lib.vr_last_error(None)
    info.devices[1] = 999
    assert lib.vr_create(C.byref(info), C.byref(ctx)) == capi.VR_ERR_INVALID_ARGUMENT
    info.device_count = 99
    assert lib.vr_create(C.byref(info), C.byref(ctx)) == capi.VR_ERR_INVALID_ARGUMENT
    old = capi.vr_create_info(28, 0, 64, 64, 0, 1, 0)          # version 0.1's struct: still a single-GPU context
    assert lib.vr_create(C.byref(old), C.byref(ctx)) == capi.VR_OK
    lib.vr_destroy(ctx)


@needs_two
@pytest.mark.parametrize("workload", ["c3", "c1", "smoke"])
def test_group_frame_equals_the_single_gpu_frame(default_volume, workload):
    n = min(_gpu_count(), 8)
    u0, u1 = scene.default_scene(20.0, -10.0)
    if workload == "c3":
        W, H = 3840, 2160
        kw = dict(density_mode=2, max_steps=256, sigma_t=8.0, light_steps=8, light_step_len=2 * 4.0 / 256, ambient=0.15, sun_dir=SUN,
                  sun_color=[1.0, 0.95, 0.85])
        setup = lambda r: None
    elif workload == "c1":
        W, H, kw = 1280, 720, {}
        setup = lambda r: r.upload_volume(default_volume, capi.VOL_RGBA8_UNORM, capi.LAYOUT_TEXTURE)
    else:
        W, H = 640, 360
        kw = dict(density_mode=1, max_steps=256, sample_scale=2.0, sigma_t=20.0, early_term_eps=1.0 / 256, skip_empty=1, ambient=1.0,
                  sun_color=[0.0, 0.0, 0.0])
        setup = lambda r: r.generate_volume(size=128, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=0.07)
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM) as r:
        setup(r)
        r.set_params(u0, u1, **kw)
        alone = r.render()[0].copy()
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, devices=list(range(n))) as g:
        setup(g)
        g.set_params(u0, u1, **kw)
        frames = [g.render()[0].copy() for _ in range(4)]            # the cut moves with the measured band times
        assert g.stats()["kernel_launches"] >= 2                    # more than one device marched
        # a new camera through the same context, two frames in flight with host read-back
        u0b, u1b = scene.default_scene(40.0, 15.0)
        g.set_params(u0b, u1b, **kw)
        bufs = [g.new_frame_buffer(pinned=True), g.new_frame_buffer(pinned=True)]
        for i in range(4):
            g.render(out=bufs[i & 1], flags=capi.RENDER_ASYNC)
        g.wait()
        moved = [b[0].copy() for b in bufs]
    with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM) as r:
        setup(r)
        r.set_params(u0b, u1b, **kw)
        alone_b = r.render()[0].copy()
    assert alone[..., :3].max() > 20
    for f in frames:
        assert max_abs8(f, alone) <= 1
    for f in moved:
        assert max_abs8(f, alone_b) <= 1


@needs_two
def test_group_view_batch_is_sharded_by_views():
    n = min(_gpu_count(), 8)
    W, H, views = 320, 180, 5
    vu = scene.orbit_views(16)[:views]
    kw = dict(density_mode=1, max_steps=128, sample_scale=2.0, sigma_t=20.0, early_term_eps=1.0 / 256, skip_empty=1, ambient=1.0,
              sun_color=[0.0, 0.0, 0.0])
    out = []
    for devices in (None, list(range(n))):
        with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, max_views=views, devices=devices) as r:
            r.generate_volume(size=64, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=0.05)
            r.set_params(view_ubos=vu, **kw)
            out.append(r.render().copy())
    assert out[0].shape == (views, H, W, 4) and out[0][..., :3].max() > 20
    assert np.array_equal(out[0], out[1])                            # whole views, same kernels: identical
