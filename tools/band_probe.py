"""time row bands of one frame on one GPU: does half the frame cost half the time?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
W, H, desc, kw = bench.WORKLOADS[wl]
u0, u1 = scene.default_scene()
r = vr.Renderer(W, H, capi.OUT_RGBA8_UNORM)
bench.setup_volume(r, wl, "texture")
r.set_params(u0, u1, **kw)
b2 = r.row_bands(2); b4 = r.row_bands(4); b8 = r.row_bands(8)
def t(rows, reps=5):
    ms = []
    for _ in range(reps + 2):
        r.flush_l2(); r.render(rows=rows, to_host=False); ms.append(r.stats()["gpu_ms"])
    return float(np.median(ms[2:]))
full = t((0, H))
print(wl, "full", round(full, 3))
for name, b in (("2", b2), ("4", b4), ("8", b8)):
    ts = [t((b[i], b[i + 1])) for i in range(len(b) - 1)]
    print(" bands/%s %s -> ms %s  sum %.3f  max %.3f  ideal %.3f" % (name, b, [round(x, 3) for x in ts], sum(ts), max(ts), full / len(ts)))
