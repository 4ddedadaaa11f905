"""debug: where does the LINEAR layout differ from the oracle (float32 target)?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
from oracle import oracle

def run(W, H, vol, phi, theta, ft):
    ubo0, ubo1 = scene.default_scene(phi, theta, ft)
    p = oracle.params_default(); oracle.set_scene(p, ubo0, ubo1)
    ovol, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    ref, st, steps = oracle.render(ovol, p, W, H, want_steps=True)
    with vr.Renderer(W, H, capi.OUT_RGBA32F, device=0) as r:
        r.upload_volume(vol, capi.VOL_RGBA8_UNORM, capi.LAYOUT_LINEAR)
        r.set_params(ubo0, ubo1)
        img = r.render(flags=capi.RENDER_STATS)[0]
        s = r.stats()
    d = img[..., 0] != ref[..., 0]
    print("scene", phi, theta, "covered", st["covered"], s["covered"], "steps", st["steps"], s["steps"], "diff px", int(d.sum()))
    ys, xs = np.nonzero(d)
    for y, x in list(zip(ys, xs))[:10]:
        a, b = img[y, x, 0], ref[y, x, 0]
        print("  px", x, y, "gpu %.9g ref %.9g ulps %d n=%d" % (a, b, int(a.view(np.int32)) - int(b.view(np.int32)), steps[y, x]))

rng = np.random.default_rng(7)
vol = rng.integers(0, 256, size=(32, 32, 32, 4), dtype=np.uint8)
run(320, 180, vol, 20.0, -10.0, 0.25)
run(320, 180, vol, 0.0, 0.0, 0.0)
run(640, 360, vol, -77.0, 41.0, 0.0)
