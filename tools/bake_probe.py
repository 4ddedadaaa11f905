import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi
from oracle import oracle
with vr.Renderer(64, 64) as r:
    for n in (128, 256, 512):
        ts = []
        for _ in range(6):
            t0 = time.perf_counter(); r.generate_volume(size=n); ts.append(time.perf_counter() - t0)
        print("rgba8 bake %d^3: %.3f ms (best of 6), %.1f Mpoints/s" % (n, min(ts) * 1e3, 4 * n ** 3 / min(ts) / 1e6))
    for n in (512, 1024):
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); r.generate_volume(size=n, format=capi.VOL_R16F); ts.append(time.perf_counter() - t0)
        print("fp16 smoke %d^3: %.3f ms, %.1f Mvoxels/s" % (n, min(ts) * 1e3, n ** 3 / min(ts) / 1e6))
if oracle.have_ref():
    t0 = time.perf_counter(); oracle.ref_bake_default(128); t = time.perf_counter() - t0
    print("FastNoise2 reference bake 128^3 (1 thread, max SIMD level %d): %.1f ms, %.1f Mpoints/s" % (oracle.ref().vref_max_level(), t * 1e3, 4 * 128 ** 3 / t / 1e6))
