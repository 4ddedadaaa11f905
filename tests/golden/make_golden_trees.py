"""Generate tests/golden/fastnoise2_trees.npz and fastnoise2_metadata.json from the REFERENCE's own FastNoise2.

Run in the build container only (needs /root/reference -> oracle/_ref/libvolrt_ref.so, `make -C oracle ref`):
    python tests/golden/make_golden_trees.py
For every tree of tests/noise_trees.py: the encoded string (written by volumetricrenderer_b200.noisetree, decoded by
the reference's NewFromEncodedNodeTree), the reference's GenUniformGrid3D output at AVX2 — the level the reference's
author tunes for — its min/max, and `spread` = the largest difference between the reference's own SIMD levels
(SSE2, AVX2, AVX-512 when the CPU has it), which is the natural tolerance scale: FastNoise2 is compiled -ffast-math
with per-level 1/sqrt, reciprocal, sin and pow approximations.
The metadata JSON is vref_metadata_dump(): ids and serialised members of every node class, as the reference registers them.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))
from oracle import oracle  # noqa: E402
from noise_trees import TREES, SHAPE  # noqa: E402


def main():
    if not oracle.have_ref():
        raise SystemExit("oracle/_ref/libvolrt_ref.so missing: make -C oracle ref (needs /root/reference)")
    json.dump(oracle.ref_metadata(), open(os.path.join(HERE, "fastnoise2_metadata.json"), "w"), indent=0)
    levels = [oracle.LEVEL_SSE2, oracle.LEVEL_AVX2] + ([oracle.LEVEL_AVX512] if oracle.ref().vref_max_level() >= 4 else [])
    d = {"shape": np.array(SHAPE, np.int32), "names": np.array(list(TREES))}
    for name, (tree, freq, seed, start) in TREES.items():
        enc = tree.encode()
        grids = [oracle.ref_tree_grid(enc, SHAPE, freq, seed, start, level=l) for l in levels]
        ref, mm = grids[1]
        if name == "convert_rgba8":
            spread = max(float(np.abs(((g.view(np.uint32) & 0xff).astype(np.int64) - (ref.view(np.uint32) & 0xff).astype(np.int64))).max()) for g, _ in grids)
        else:
            spread = max(float(np.abs(g - ref).max()) for g, _ in grids)
        d[name + "/encoded"] = np.array(enc)
        d[name + "/grid"] = ref
        d[name + "/minmax"] = mm
        d[name + "/spread"] = np.float32(spread)
        d[name + "/params"] = np.array([freq, seed, start[0], start[1], start[2]], np.float64)
        print("%-32s spread %.2e" % (name, spread))
    np.savez_compressed(os.path.join(HERE, "fastnoise2_trees.npz"), **d)
    print("wrote", os.path.getsize(os.path.join(HERE, "fastnoise2_trees.npz")), "bytes")


if __name__ == "__main__":
    main()
