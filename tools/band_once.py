"""render one 1/8 row band of a workload a few times (for an ncu capture of a multi-GPU-sized launch on one GPU)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
W, H, desc, kw = bench.WORKLOADS[wl]
u0, u1 = scene.default_scene()
r = vr.Renderer(W, H, capi.OUT_RGBA8_UNORM)
bench.setup_volume(r, wl, "texture")
r.set_params(u0, u1, **kw)
b = r.row_bands(8)
rows = (b[3], b[4])
for _ in range(6):
    r.flush_l2(); r.render(rows=rows, to_host=False)
print(wl, "rows", rows, "gpu_ms", r.stats()["gpu_ms"])
