import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
vol = (np.random.default_rng(0).random((48, 48, 48), np.float32)).astype(np.float16)
u0, u1 = scene.default_scene()
with vr.Renderer(160, 96, capi.OUT_RGBA32F) as r:
    r.upload_volume(vol, capi.VOL_R16F, capi.LAYOUT_BRICK)
    r.set_params(u0, u1, density_mode=1, max_steps=64)
    img = r.render()[0]
    print("ok", img[..., 0].max())
