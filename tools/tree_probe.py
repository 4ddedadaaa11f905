"""Node-tree evaluation rate: libvolrt on the device vs the reference's FastNoise2 on one host thread (its best SIMD
level), same encoded trees, 128^3 grid (the reference's bake size), device time from vr_get_stats."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200.noisetree import CompiledTree
from oracle import oracle
from noise_trees import TREES

n = 128
names = ["perlin", "opensimplex2", "cell_distance_default", "cell_value_idx2_manhattan", "fbm_perlin5", "ridged_simplex",
         "warp_progressive", "blend_fade_pow", "cloud_density"]
with vr.Renderer(64, 64) as r:
    for name in names:
        tree, freq, seed, start = TREES[name]
        enc = tree.encode()
        ct = CompiledTree(enc)
        ms = []
        for _ in range(5):
            r.noise_tree_grid(ct, (n, n, n), freq, seed, start, want_grid=False)
            ms.append(r.stats()["gpu_ms"])
        g = min(ms)
        line = "%-28s %3d instr  device %7.3f ms = %8.1f Mpoints/s" % (name, ct.info()["instructions"], g, n ** 3 / g / 1e3)
        if oracle.have_ref():
            ts = []
            for _ in range(3):
                t0 = time.perf_counter(); oracle.ref_tree_grid(enc, (n, n, n), freq, seed, start); ts.append(time.perf_counter() - t0)
            c = min(ts) * 1e3
            line += " | FastNoise2 1 thread (SIMD level %d) %8.2f ms = %7.1f Mpoints/s | x%.0f" % (oracle.ref().vref_max_level(), c, n ** 3 / c / 1e3, c / g)
        print(line, flush=True)
        ct.close()
