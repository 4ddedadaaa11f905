"""Generate tests/golden/fastnoise2_golden.npz from the REFERENCE's own FastNoise2.

Run in the build container only (needs /root/reference → oracle/_ref/libvolrt_ref.so, built by
`make -C oracle ref`):   python tests/golden/make_golden.py
The fixtures pin the oracle's noise restatement (oracle/oracle_noise.cpp) and libvolrt's noise
kernels to the reference's outputs; they travel to the GPU box, the reference does not.

Every call mirrors /root/reference/source/Etna/TestMain.cpp:43-62 (generators, frequencies, seeds)
at SIMD level SSE2 (bit-identical to SSE4.1; the lowest level an x86-64 build has,
vendor/noise/include/FastSIMD/FastSIMD_Config.h:17) and AVX2.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fastnoise2_golden.npz")
RECIPES = {  # TestMain.cpp:59-62
    "cellular_a": (oracle.NOISE_CELLULAR, 0.01, 1),
    "cellular_b": (oracle.NOISE_CELLULAR, 0.03, 2),
    "perlin": (oracle.NOISE_PERLIN, 0.19, 3),
    "simplex": (oracle.NOISE_SIMPLEX, 0.15, 4),
}
LEVELS = {"sse2": oracle.LEVEL_SSE2, "avx2": oracle.LEVEL_AVX2}


def main():
    if not oracle.have_ref():
        raise SystemExit("oracle/_ref/libvolrt_ref.so missing: make -C oracle ref (needs /root/reference)")
    d = {}
    rng = np.random.default_rng(20261104)
    pts = rng.uniform(-40.0, 40.0, size=(3, 512)).astype(np.float32)
    d["points_xyz"] = pts
    for lname, lvl in LEVELS.items():
        for name, (gen, freq, seed) in RECIPES.items():
            # non-cubic grid: pins the index order (x fastest) of Generator.inl:124-176
            g, mm = oracle.ref_grid(gen, (24, 10, 6), freq, seed, lvl)
            d["grid_%s_%s" % (name, lname)] = g
            d["grid_mm_%s_%s" % (name, lname)] = mm
            d["points_%s_%s" % (name, lname)] = oracle.ref_points(gen, seed, pts[0], pts[1], pts[2], lvl)
        d["fbm5_%s" % lname] = oracle.ref_fbm_points(1337, pts[0] * 0.25, pts[1] * 0.25, pts[2] * 0.25, 5, 0.5, 2.0, lvl)
        vol, mm = oracle.ref_bake_default(128, lvl)
        d["bake_mm_%s" % lname] = mm
        d["bake_sha256_%s" % lname] = np.frombuffer(hashlib.sha256(vol.tobytes()).digest(), np.uint8)
        d["bake_mean_%s" % lname] = vol.reshape(-1, 4).mean(axis=0)
        d["bake_crop_%s" % lname] = vol[40:56, 60:76, 100:116].copy()        # 16^3 texels
        d["bake_hist_r_%s" % lname] = np.bincount(vol[..., 0].ravel(), minlength=256).astype(np.int64)
    np.savez_compressed(OUT, **d)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
