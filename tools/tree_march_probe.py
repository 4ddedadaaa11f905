"""VR_DENSITY_TREE at the C2 frame (1920x1080, 128 steps, 6 light samples): the fBm cloud written as a node tree, the same
with a domain warp, and a cellular field, beside the built-in fBm kernel"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, scene
from volumetricrenderer_b200.noisetree import N, CompiledTree
import bench

W, H, desc, kw = bench.WORKLOADS["c2"]
u0, u1 = scene.default_scene()


def cloud(src):
    return N("Max", RHS=0.0, LHS=N("Subtract", RHS=0.45, LHS=N("Remap", FromMin=-1, FromMax=1, ToMin=0, ToMax=1, Source=src)))


fbm = N("FractalFBm", Octaves=5, Source=N("Perlin"))
TREES = {
    "fbm5_perlin": cloud(fbm),
    "fbm5_perlin_warped": cloud(N("DomainWarpGradient", WarpFrequency=0.7, WarpAmplitude=0.6, Source=fbm)),
    "fbm5_simplex": cloud(N("FractalFBm", Octaves=5, Source=N("Simplex"))),
    "ridged4_opensimplex2": cloud(N("FractalRidged", Octaves=4, Source=N("OpenSimplex2"))),
    "cellular_distance": N("Subtract", LHS=0.35, RHS=N("CellularDistance")),
}
with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM) as r:
    def timed(**extra):
        k = dict(kw); k.update(extra)
        r.set_params(u0, u1, **k)
        ts = []
        for i in range(6):
            r.flush_l2(); r.render(to_host=False)
            if i >= 2:
                ts.append(r.stats()["gpu_ms"])
        return float(np.mean(ts))
    base = timed()
    print("built-in fBm kernel           %8.3f ms" % base)
    d = r.default_params()
    for name, t in TREES.items():
        ct = CompiledTree(t.encode())
        ms = timed(density_tree=ct, density_mode=capi.DENSITY_TREE, tree_frequency=d.fbm_frequency, tree_seed=d.fbm_seed)
        print("%-28s  %8.3f ms  (%d instructions, x%.2f the fBm kernel)" % (name, ms, ct.info()["instructions"], ms / base))
        ct.close()
