"""Generate tests/golden/glsl_shaders_golden.npz from the REFERENCE's own shaders.

Run in the build container only (needs /root/reference -> oracle/_ref/libvolrt_ref.so, built by `make -C oracle ref`,
which compiles /root/reference/shaders/frag.glsl and vert.glsl as C++ behind oracle/glsl_shim.h):
    python tests/golden/make_golden_glsl.py
The fixture pins the oracle's ray-march restatement (oracle/oracle_march.cpp) to what the reference's shader text computes;
it travels to machines that have no /root/reference.  Per scene it holds the inputs (UBO bytes as a host would upload
them, the RGBA8 volume, the per-pixel fragPosition / coverage handed to the fragment stage) and the outputs: frag.glsl's
outColor per pixel, vert.glsl's gl_Position / fragPosition for the cube's corners, and the frame of the whole draw
(vert.glsl -> fixed-function rasteriser of oracle/ref_raster.cpp -> frag.glsl).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "glsl_shaders_golden.npz")
W, H = 96, 54
CUBE = np.array([[1, -1, -1], [1, -1, 1], [-1, -1, 1], [-1, -1, -1], [1, 1, -1], [1, 1, 1], [-1, 1, 1], [-1, 1, -1]], np.float32)
# name -> (phi, theta, frame_time, volume shape (d,h,w), volume seed, MediaScroll y/z/w rows of columns 0..2 or None)
SCENES = {
    "default_pose": (0.0, 0.0, 0.0, (16, 16, 16), 1, None),
    "rotated_time": (20.0, -10.0, 0.25, (16, 16, 16), 2, None),
    "steep_ragged_volume": (77.0, 33.0, 1.5, (9, 20, 12), 3, None),
    "scrolled_std140": (-35.0, 12.0, 0.0, (16, 16, 16), 4, [[0.3, -0.7, 1.9], [0.11, 0.5, -2.25], [4.0, 0.05, -0.6]]),
}


def scene_bytes(phi, theta, frame_time, scroll):
    """ubo0 (192 B) and ubo1 as uploaded: the reference host's 140-byte C++ struct (TestMain.cpp:34-39), or — for the
    scrolled scene — a 144-byte std140 image whose MediaScroll has the .y/.z/.w the shader reads set"""
    u0, u1 = oracle.scene_default(phi, theta, frame_time)
    if scroll is None:
        return u0, u1
    s = np.zeros(36, np.float32)
    s[:19] = u1[:19]                        # WorldToLocal @0, CameraPosition @64 line up in both layouts
    for col in range(3):
        for row, vals in zip((1, 2, 3), scroll):
            s[20 + 4 * col + row] = vals[col]
    s[20] = -frame_time
    return u0, s


def main():
    if not oracle.have_ref_glsl():
        raise SystemExit("oracle/_ref/libvolrt_ref.so lacks the compiled shaders: make -C oracle ref (needs /root/reference)")
    d = {"frag_constants": oracle.ref_frag_constants(), "size": np.array([W, H], np.int32)}
    for name, (phi, theta, ft, shape, seed, scroll) in SCENES.items():
        rng = np.random.default_rng(seed)
        vol = rng.integers(0, 256, size=shape + (4,), dtype=np.uint8)
        u0, u1 = scene_bytes(phi, theta, ft, scroll)
        p = oracle.params_default()
        oracle.set_scene(p, u0, u1)
        fpos, cov = oracle.frag_positions(p, W, H)
        d[name + "/ubo0"] = u0
        d[name + "/ubo1"] = u1
        d[name + "/volume"] = vol
        d[name + "/frag_position"] = fpos
        d[name + "/covered"] = cov.astype(np.uint8)
        d[name + "/out_color"] = oracle.ref_frag(u1, vol, fpos, cov)
        gl, fr = oracle.ref_vert(u0, CUBE)
        d[name + "/vert_gl_position"] = gl
        d[name + "/vert_frag_position"] = fr
        img, c2, _ = oracle.ref_draw(u0, u1, vol, W, H)
        d[name + "/draw_rgba"] = img
        d[name + "/draw_covered"] = c2.astype(np.uint8)
    np.savez_compressed(OUT, **d)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
