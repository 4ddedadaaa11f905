"""C5 anatomy on one GPU: the 1024^3 fp16 grid, 16 of the 64 orbit views, with empty-space skipping and early
termination switched on and off — what each of them buys, and what a step costs.
usage: python tools/c5_probe.py [views]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
import bench

NV = int(sys.argv[1]) if len(sys.argv) > 1 else 16
W, H, desc, kw = bench.WORKLOADS["c5"]
views = scene.orbit_views(bench.C5_VIEWS)[:: bench.C5_VIEWS // NV][:NV]
r = vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, max_views=NV)
grid = int(os.environ.get("VOLRT_C5_GRID", bench.C5_GRID))
r.generate_volume(size=grid, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=bench.C5_THRESHOLD)


def t(**over):
    k = dict(kw)
    k.update(over)
    r.set_params(view_ubos=views, **k)
    r.render(to_host=False, flags=capi.RENDER_STATS)
    st = r.stats()
    ms = []
    for _ in range(6):
        r.flush_l2()
        r.render(to_host=False)
        ms.append(r.stats()["gpu_ms"])
    m = float(np.median(ms[1:]))
    fetched = st["steps"] - st["skipped_steps"]
    print("%-34s %8.3f ms  %7.1f Mrays/s  steps %.3g  skipped %.3g  early %d  -> %.1f Gfetch/s (upper bound)" % (
        " ".join("%s=%s" % kv for kv in over.items()) or "c5 as defined", m, W * H * NV / m / 1e3, st["steps"],
        st["skipped_steps"], st["early_terminated"], fetched / m / 1e6))


t()
t(skip_empty=0)
t(early_term_eps=0.0)
t(skip_empty=0, early_term_eps=0.0)
