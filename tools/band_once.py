"""render one 1/N row band of a workload a few times (what one of N GPUs runs; also for an ncu capture of such a launch on one GPU)
usage: python tools/band_once.py [workload] [N=8] [band index = N/2 - 1]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8
k = int(sys.argv[3]) if len(sys.argv) > 3 else N // 2 - 1
W, H, desc, kw = bench.WORKLOADS[wl]
u0, u1 = scene.default_scene()
r = vr.Renderer(W, H, capi.OUT_RGBA8_UNORM)
bench.setup_volume(r, wl, "texture")
r.set_params(u0, u1, **kw)
b = r.row_bands(N)
rows = (b[k], b[k + 1])
ms = []
for _ in range(8):
    r.flush_l2(); r.render(rows=rows, to_host=False); ms.append(r.stats()["gpu_ms"])
print(wl, "1/%d band" % N, "rows", rows, "segments", os.environ.get("VOLRT_RAY_SEGMENTS", "auto"), "lanes", os.environ.get("VOLRT_RAY_SPLIT_COMPOSITE", "auto"),
      "gpu_ms %.4f" % min(ms[2:]))
