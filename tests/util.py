"""helpers shared by the parity tests"""
import numpy as np


def psnr8(a, b):
    d = a.astype(np.float64) - b.astype(np.float64)
    mse = float((d * d).mean())
    return 99.0 if mse == 0 else 10.0 * np.log10(255.0 * 255.0 / mse)


def max_abs8(a, b):
    return int(np.abs(a.astype(np.int32) - b.astype(np.int32)).max())


def oracle_params(oracle, ubo0, ubo1, **kw):
    p = oracle.params_default(**kw)
    oracle.set_scene(p, ubo0, ubo1)
    return p


def smoke_volume(n, seed=0):
    """cheap synthetic one-channel 'smoke' (blobs, ~half empty) as float16, shaped (n,n,n)"""
    rng = np.random.default_rng(seed)
    g = np.linspace(0, 1, n, dtype=np.float32)
    z, y, x = np.meshgrid(g, g, g, indexing="ij")
    v = np.zeros((n, n, n), np.float32)
    for _ in range(6):
        c = rng.uniform(0.2, 0.8, 3)
        r = rng.uniform(0.08, 0.2)
        v += np.maximum(0.0, 1.0 - ((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) / (r * r))
    return v.astype(np.float16)
