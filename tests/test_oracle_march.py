"""Tests that pin the CPU oracle's restatement of shaders/frag.glsl (oracle/oracle_march.cpp).

Two kinds.  (1) Closed forms derived from the shader text itself — none of these expectations comes out of the oracle.
(2) At the end of the file: the oracle against the reference's own shaders COMPILED AS C++ (oracle/_ref), bit for bit.
The reference holds no golden vector for the ray-march and its Vulkan pipeline cannot run here (no graphics stack).
  frag.glsl:42     stepSize = 4/128
  frag.glsl:46     n = min(128, int(|Pin-Pout| / stepSize))
  frag.glsl:66-71  rho = (s1*s2)*(s3+s4)*0.2
  frag.glsl:76-80  colour = 1 - exp(-sum(rho)*stepSize), alpha 1
  VulkanCore.cpp:676-710   linear filter, mirrored repeat, texel centres at (i+0.5)/N
  VulkanRenderPass.cpp:17-19   clear colour (0,0,0,1)
"""
import os

import numpy as np
import pytest

from volumetricrenderer_b200 import scene
from util import oracle_params


def _axis_scene(fov_deg=60.0, eye=(4.0, 0.0, 0.0)):
    ubo0 = np.zeros(48, np.float32)
    ubo0[0:16] = scene.identity()
    ubo0[16:32] = scene.look_at(eye, (0.0, 0.0, 0.0), (0.0, 0.0, 1.0))
    proj = scene.perspective(scene.radians(fov_deg), 1.0, 0.1, 10.0)
    proj[5] *= -1
    ubo0[32:48] = proj
    ubo1 = np.zeros(36, np.float32)
    ubo1[0:16] = scene.identity()
    ubo1[16:19] = eye
    return ubo0, ubo1


def test_constant_volume_closed_form(oracle):
    W = H = 129                                   # odd: the centre pixel's ray is the +X axis
    vol = np.full((4, 4, 4, 4), 255, np.uint8)
    ubo0, ubo1 = _axis_scene()
    p = oracle_params(oracle, ubo0, ubo1)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    img, st, steps = oracle.render(ov, p, W, H, want_steps=True)
    c = H // 2
    assert steps[c, c] == 64                      # chord 2 / (4/128)
    rho = (1.0 * 1.0) * (1.0 + 1.0) * 0.2
    assert abs(img[c, c, 0] - (1.0 - np.exp(-rho * 64 * 0.03125))) < 2e-7
    # every covered pixel: colour is the closed form of ITS step count
    cov = steps >= 0
    want = 1.0 - np.exp(-rho * steps[cov].astype(np.float64) * 0.03125)
    assert np.abs(img[..., 0][cov] - want).max() < 5e-7
    assert np.all(img[..., 3] == 1.0)
    assert np.all(img[~cov][:, :3] == 0.0)        # clear colour
    assert 0.2 < cov.mean() < 0.5 and st["covered"] == int(cov.sum()) and st["steps"] == int(steps[cov].sum())
    # grey: the three colour lanes are the same accumulator (frag.glsl:55,80)
    assert np.array_equal(img[..., 0], img[..., 1]) and np.array_equal(img[..., 0], img[..., 2])


def test_step_count_is_truncated_chord_and_capped(oracle):
    W = H = 65
    vol = np.full((2, 2, 2, 4), 255, np.uint8)
    ubo0, ubo1 = _axis_scene()
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    for max_steps, step_scale, want in [(128, 4.0, 64), (256, 4.0, 128), (128, 2.0, 128), (32, 4.0, 16), (16, 0.5, 16)]:
        p = oracle_params(oracle, ubo0, ubo1, max_steps=max_steps, step_scale=step_scale)
        _, _, steps = oracle.render(ov, p, W, H, want_steps=True)
        assert steps[H // 2, W // 2] == want      # min(maxSteps, int(2 / ((1/maxSteps)*scale)))
        assert steps.max() <= max_steps


def test_camera_inside_or_missing_box_draws_nothing(oracle):
    vol = np.full((2, 2, 2, 4), 255, np.uint8)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    # camera inside the cube: only back faces remain and they are culled (VulkanPipeline.cpp:107)
    ubo0, ubo1 = _axis_scene(eye=(0.5, 0.0, 0.0))
    img, st = oracle.render(ov, oracle_params(oracle, ubo0, ubo1), 33, 33)
    assert st["covered"] == 0 and np.all(img[..., :3] == 0) and np.all(img[..., 3] == 1)
    # cube beyond the far plane (10)
    ubo0, ubo1 = _axis_scene(eye=(14.0, 0.0, 0.0))
    _, st = oracle.render(ov, oracle_params(oracle, ubo0, ubo1), 33, 33)
    assert st["covered"] == 0


def test_sampler_texel_centres_filtering_and_mirrored_repeat(oracle):
    rng = np.random.default_rng(1)
    N = 8
    vol = rng.integers(0, 256, size=(N, N, N, 4), dtype=np.uint8)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    t = lambda i, j, k: vol[k % N, j % N, i % N].astype(np.float32) / np.float32(255)   # noqa: E731
    # at a texel centre the filter returns the texel
    for (i, j, k) in [(0, 0, 0), (3, 5, 7), (7, 7, 7)]:
        s = oracle.sample(ov, (i + .5) / N, (j + .5) / N, (k + .5) / N)
        assert np.allclose(s, t(i, j, k), atol=1e-7)
    # half-way between two texels along x: the mean
    s = oracle.sample(ov, 3.0 / N, 2.5 / N, 4.5 / N)
    assert np.allclose(s, 0.5 * (t(2, 2, 4) + t(3, 2, 4)), atol=1e-6)
    # u = 0 and u = 1 sit on the mirror: both taps are the edge texel
    assert np.allclose(oracle.sample(ov, 0.0, 2.5 / N, 4.5 / N), t(0, 2, 4), atol=1e-7)
    assert np.allclose(oracle.sample(ov, 1.0, 2.5 / N, 4.5 / N), t(7, 2, 4), atol=1e-7)
    # MIRRORED_REPEAT: T(u) = T(-u) = T(2 - u) = T(u + 2)   (SURVEY.md A.4 cases 1.5, -0.25)
    for u in (0.3, 0.77, 0.25):
        a = oracle.sample(ov, u, 0.4, 0.6)
        for um in (-u, 2.0 - u, u + 2.0, u - 2.0):
            assert np.allclose(oracle.sample(ov, um, 0.4, 0.6), a, atol=2e-6)
    assert np.allclose(oracle.sample(ov, 1.5, 0.4, 0.6), oracle.sample(ov, 0.5, 0.4, 0.6), atol=2e-6)
    assert np.allclose(oracle.sample(ov, 0.1, -0.25, 0.6), oracle.sample(ov, 0.1, 0.25, 0.6), atol=2e-6)


def test_single_channel_formats(oracle):
    v16 = np.array([[[0.0, 1.0], [0.5, 0.25]], [[2.0, 0.125], [0.0, 0.0]]], np.float16)
    ov, _k = oracle.make_volume(v16.view(np.uint16), oracle.VOL_R16F)
    assert oracle.sample(ov, 0.75, 0.25, 0.25)[0] == 1.0
    assert oracle.sample(ov, 0.5, 0.25, 0.25)[0] == 0.5
    v8 = np.array([[[0, 255], [51, 102]], [[0, 0], [0, 0]]], np.uint8)
    ov, _k = oracle.make_volume(v8, oracle.VOL_R8)
    assert abs(oracle.sample(ov, 0.25, 0.75, 0.25)[0] - 0.2) < 1e-7


def test_ubo1_std140_decoding_quirk_q2(oracle):
    """TestMain.cpp:34-39 packs MediaScroll at byte 76, frag.glsl:9-14 reads it at byte 80."""
    host = np.zeros(35, np.float32)
    host[19:35] = np.arange(100, 116, dtype=np.float32)     # C++ MediaScroll flat elements 0..15
    dec = oracle.decode_ubo1(host.tobytes())
    ms = dec[20:36]                                         # what the shader sees
    assert ms[:15].tolist() == list(range(101, 116))        # shifted by one float
    assert ms[15] == 0.0                                    # beyond the 140-byte descriptor range
    u0, u1 = oracle.scene_default(0.0, 0.0, 123.0)
    assert u1[19] == -123.0 and oracle.decode_ubo1(u1.tobytes())[20:36].max() == 0.0   # time never reaches the shader


def test_image_is_time_invariant_but_scroll_moves_it(oracle):
    rng = np.random.default_rng(2)
    vol = rng.integers(0, 256, size=(8, 8, 8, 4), dtype=np.uint8)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    u0, u1a = scene.default_scene(0, 0, 0.0)
    _, u1b = scene.default_scene(0, 0, 9.0)
    a, _ = oracle.render(ov, oracle_params(oracle, u0, u1a), 96, 54)
    b, _ = oracle.render(ov, oracle_params(oracle, u0, u1b), 96, 54)
    assert np.array_equal(a, b)
    std140 = np.zeros(36, np.float32)
    std140[:19] = u1a[:19]
    std140[20 + 1] = 0.37                                   # MediaScroll[0].y -> moves fetch 2 only
    c, _ = oracle.render(ov, oracle_params(oracle, u0, std140), 96, 54)
    assert not np.array_equal(a, c)


def test_scene_matrices(oracle):
    """glm closed forms: the Python host scene and the oracle's C scene agree to a few ulp."""
    for phi, theta, t in [(0, 0, 0), (30, -12.5, 1.5), (-170, 88, 0.25)]:
        a0, a1 = scene.default_scene(phi, theta, t)
        b0, b1 = oracle.scene_default(phi, theta, t)
        assert np.abs(a0 - b0).max() <= 5e-7           # numpy vs libm cos/sin: a few ulp
        assert np.abs(a1 - b1).max() <= 5e-7
    u0, u1 = scene.default_scene(40.0, 10.0)
    M, V, P = (u0[i * 16:(i + 1) * 16].reshape(4, 4).T for i in range(3))
    assert np.allclose(V @ np.array([3, 3, 3, 1.0]), [0, 0, 0, 1], atol=1e-6)          # View * eye = 0
    clip = P @ V @ np.array([0, 0, 0, 1.0])
    assert np.allclose(clip[:2] / clip[3], 0, atol=1e-6)                                # centre -> NDC (0,0)
    assert P[1, 1] < 0 and P[3, 2] == -1                                                # Y flip, RH
    W2L = u1[:16].reshape(4, 4).T
    assert np.allclose(W2L @ M, np.eye(4), atol=1e-6)
    assert np.allclose(oracle.mat4_inverse(u0[:16]).reshape(4, 4).T @ M, np.eye(4), atol=1e-6)


def test_rotating_the_model_equals_orbiting_the_camera(oracle):
    """WorldToLocal = inverse(Model) (TestMain.cpp:230): spinning the cube about Z by phi shows the
    same picture as a static cube seen from a camera spun by -phi — checked through chord lengths."""
    vol = np.full((2, 2, 2, 4), 255, np.uint8)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    u0, u1 = scene.default_scene(90.0, 0.0)      # cube has 4-fold symmetry about Z: 90 deg = identity picture
    v0, v1 = scene.default_scene(0.0, 0.0)
    a, sa, na = oracle.render(ov, oracle_params(oracle, u0, u1), 160, 90, want_steps=True)
    b, sb, nb = oracle.render(ov, oracle_params(oracle, v0, v1), 160, 90, want_steps=True)
    assert abs(sa["covered"] - sb["covered"]) <= 4
    both = (na >= 0) & (nb >= 0)
    assert np.abs(na[both] - nb[both]).max() <= 1
    assert np.abs(a - b).max() < 0.01


def test_output_encodings(oracle):
    x = np.array([[0.0, 0.5, 1.0, 1.0], [0.0031308, 0.2, 1.5, 0.25], [-0.5, 0.00196078, 0.00196079 * 1.5, 1.0]], np.float32)
    u = oracle.encode(x, oracle.OUT_RGBA8_UNORM)
    assert u.tolist() == [[0, 128, 255, 255], [1, 51, 255, 64], [0, 0, 1, 255]]        # RNE, clamp
    s = oracle.encode(x, oracle.OUT_RGBA8_SRGB)
    want = lambda c: int(np.rint(255 * (12.92 * c if c <= 0.0031308 else 1.055 * c ** (1 / 2.4) - 0.055)))  # noqa: E731
    assert s[0].tolist() == [0, want(0.5), 255, 255] and s[1, 1] == want(0.2) and s[1, 3] == 64   # alpha stays linear
    h = oracle.encode(x, oracle.OUT_RGBA16F)
    assert np.array_equal(h.view(np.float16), x.astype(np.float16))
    r = np.random.default_rng(0).normal(0, 100, 4096).astype(np.float32)
    assert np.array_equal(oracle.encode(r, oracle.OUT_RGBA16F).view(np.float16), r.astype(np.float16))


def test_extended_integrator_reduces_to_the_reference(oracle):
    """compositing with no light, white sun, no ambient == 1 - exp(-sigma_t * sum(rho) * dt)."""
    rng = np.random.default_rng(4)
    vol = rng.integers(0, 256, size=(8, 8, 8, 4), dtype=np.uint8)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    u0, u1 = scene.default_scene(15.0, 5.0)
    ref, _ = oracle.render(ov, oracle_params(oracle, u0, u1, density=3.0), 96, 54)
    comp, _ = oracle.render(ov, oracle_params(oracle, u0, u1, sigma_t=3.0), 96, 54)
    assert np.abs(ref - comp).max() < 1e-5
    # early termination never changes a pixel by more than eps
    full, _ = oracle.render(ov, oracle_params(oracle, u0, u1, sigma_t=40.0), 96, 54)
    et, st = oracle.render(ov, oracle_params(oracle, u0, u1, sigma_t=40.0, early_term_eps=1 / 256), 96, 54)
    assert st["early"] > 0 and np.abs(full - et).max() <= 1 / 256 + 1e-6


def test_render_rows_is_a_slice_of_the_frame(oracle):
    rng = np.random.default_rng(6)
    vol = rng.integers(0, 256, size=(8, 8, 8, 4), dtype=np.uint8)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    u0, u1 = scene.default_scene()
    p = oracle_params(oracle, u0, u1)
    full, st = oracle.render(ov, p, 128, 72)
    band, _ = oracle.render(ov, p, 128, 72, rows=(20, 50), threads=1)
    assert np.array_equal(full[20:50], band)
    rows = np.array([5, 33, 34, 70], np.int32)
    some, _ = oracle.render_rows(ov, p, 128, 72, rows, threads=3)
    assert np.array_equal(full[rows], some)
    empty, st0 = oracle.render(ov, p, 128, 72, rows=(10, 10))
    assert empty.shape == (0, 128, 4) and st0["covered"] == 0


# ---------------------------------------------------------------------------------------------------------------------
# The oracle held to the REFERENCE'S OWN SHADER TEXT.  oracle/Makefile `ref` compiles /root/reference/shaders/frag.glsl and
# vert.glsl — from where they lie, as C++ behind oracle/glsl_shim.h — into oracle/_ref/libvolrt_ref.so (it travels to the
# GPU box; /root/reference does not).  tests/golden/glsl_shaders_golden.npz holds outputs of that build for machines
# without it (script: tests/golden/make_golden_glsl.py).  Bit-exact on binary32 outColor: the oracle restates the same
# operations in the same order, and both sides take GLSL's implementation-defined pieces (rounding per operator, mat*vec
# order, exp, the fixed-function sampler) the same way — see the header of glsl_shim.h.
GOLDEN_GLSL = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "glsl_shaders_golden.npz")
CUBE = np.array([[1, -1, -1], [1, -1, 1], [-1, -1, 1], [-1, -1, -1], [1, 1, -1], [1, 1, 1], [-1, 1, 1], [-1, 1, -1]], np.float32)


def _needs_compiled_shaders(oracle):
    if not oracle.have_ref_glsl():
        pytest.skip("oracle/_ref with the compiled shaders is not built (needs /root/reference: make -C oracle ref)")


def _std140_with_scroll(u1_host, scroll_rows):
    """144-byte std140 image of GlobalShaderData whose MediaScroll columns 0..2 carry the given .y / .z / .w"""
    s = np.zeros(36, np.float32)
    s[:19] = u1_host[:19]
    for col in range(3):
        for row, vals in zip((1, 2, 3), scroll_rows):
            s[20 + 4 * col + row] = vals[col]
    return s


def test_compiled_shader_literals_are_the_oracle_defaults(oracle):
    _needs_compiled_shaders(oracle)
    density, max_steps, box_min, box_max = oracle.ref_frag_constants()          # frag.glsl:29-32 as compiled
    p = oracle.params_default()
    assert (density, max_steps, box_min, box_max) == (p.density, p.max_steps, -1.0, 1.0)
    assert (p.step_scale, round(p.sample_scale, 6)) == (4.0, 0.2)


def test_oracle_equals_compiled_frag_glsl(oracle):
    """the full config-1 frame (1280x720, the reference's own bake) and five more scenes: binary32 outColor, bit for bit"""
    _needs_compiled_shaders(oracle)
    vol_default, _mm = oracle.ref_bake_default(128)                              # the reference's FastNoise2, TestMain.cpp:43-92
    rng = np.random.default_rng(11)
    scenes = [
        # W, H, phi, theta, frame_time, volume, scroll (None = the host's 140-byte struct, quirk Q2)
        (1280, 720, 0.0, 0.0, 0.0, vol_default, None),
        (640, 360, 20.0, -10.0, 0.25, rng.integers(0, 256, size=(32, 32, 32, 4), dtype=np.uint8), None),
        (640, 360, 77.0, 33.0, 3.0, rng.integers(0, 256, size=(9, 20, 12, 4), dtype=np.uint8), None),
        (640, 360, -35.0, 12.0, 0.0, rng.integers(0, 256, size=(16, 16, 16, 4), dtype=np.uint8), [[0.3, -0.7, 1.9], [0.11, 0.5, -2.25], [4.0, 0.05, -0.6]]),
        (333, 187, 181.0, -64.0, 0.0, vol_default[40:72, 8:40, 64:96].copy(), [[-1.5, 2.5, 0.25], [0.0, 0.0, 0.0], [9.0, -9.0, 0.5]]),
        (200, 120, 45.0, 89.0, 7.0, rng.integers(0, 2, size=(5, 5, 5, 4), dtype=np.uint8) * 255, None),
    ]
    for W, H, phi, theta, ft, vol, scroll in scenes:
        u0, u1 = oracle.scene_default(phi, theta, ft)
        if scroll is not None:
            u1 = _std140_with_scroll(u1, scroll)
        p = oracle_params(oracle, u0, u1)
        ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
        mine, st, steps = oracle.render(ov, p, W, H, want_steps=True)
        fpos, cov = oracle.frag_positions(p, W, H)
        assert np.array_equal(cov == 1, steps >= 0) and st["covered"] == int(cov.sum()) > 0.1 * W * H
        theirs = oracle.ref_frag(u1, vol, fpos, cov)
        assert np.array_equal(mine.view(np.uint32), theirs.view(np.uint32)), (W, H, phi, theta)
        assert theirs[..., 0].max() > 0.02                                         # not a frame of zeros


def test_vert_glsl_projects_the_oracles_fragments_onto_their_pixel_centres(oracle):
    """SURVEY A.2 (the oracle's ray generation) against the reference's vertex shader: the point of the cube the oracle
    hands the fragment stage, pushed through vert.glsl, lands on the pixel centre it was generated for, and comes
    back as the same fragPosition"""
    _needs_compiled_shaders(oracle)
    W, H = 320, 180
    for phi, theta in ((0.0, 0.0), (20.0, -10.0), (123.0, 45.0)):
        u0, u1 = oracle.scene_default(phi, theta)
        p = oracle_params(oracle, u0, u1)
        fpos, cov = oracle.frag_positions(p, W, H)
        ys, xs = np.nonzero(cov)
        inv_model = oracle.mat4_inverse(u0[:16]).reshape(4, 4).T.astype(np.float64)       # column-major -> rows
        world = np.concatenate([fpos[ys, xs].astype(np.float64), np.ones((len(ys), 1))], axis=1)
        local = (inv_model @ world.T).T[:, :3]
        assert np.abs(np.abs(local).max(axis=1) - 1.0).max() < 2e-4                       # on a face of the cube [-1,1]^3
        gl, fr = oracle.ref_vert(u0, local.astype(np.float32))
        ndc = gl[:, :2] / gl[:, 3:4]
        want = np.stack([(xs + 0.5) * 2.0 / W - 1.0, (ys + 0.5) * 2.0 / H - 1.0], axis=1)
        assert np.abs(ndc - want).max() < 0.02 * 2.0 / W                                   # within 2 % of a pixel
        assert np.abs(fr - fpos[ys, xs]).max() < 2e-4
        depth = gl[:, 2] / gl[:, 3]
        assert depth.min() > 0.0 and depth.max() < 1.0                                     # between the near and far planes


def test_reference_draw_pipeline_vs_oracle(oracle):
    """vert.glsl -> the fixed-function stages (oracle/ref_raster.cpp) -> frag.glsl, against the oracle's frame: the same
    pixels are shaded (silhouette ties aside) and the frames agree within the north-star tolerance; they are not bit-equal
    because the rasteriser interpolates fragPosition from the triangle's corners while the oracle un-projects the pixel"""
    _needs_compiled_shaders(oracle)
    W, H = 640, 360
    rng = np.random.default_rng(5)
    vol = rng.integers(0, 256, size=(24, 24, 24, 4), dtype=np.uint8)
    for phi, theta in ((0.0, 0.0), (20.0, -10.0), (77.0, 33.0), (200.0, -80.0)):
        u0, u1 = oracle.scene_default(phi, theta, 0.5)
        p = oracle_params(oracle, u0, u1)
        ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
        mine, st = oracle.render(ov, p, W, H)
        fpos, cov = oracle.frag_positions(p, W, H)
        theirs, rcov, rfpos = oracle.ref_draw(u0, u1, vol, W, H)
        differ = cov != rcov
        assert differ.sum() <= 1e-4 * cov.sum() + 2
        if differ.any():                                  # only on the silhouette: a neighbour is uncovered
            ys, xs = np.nonzero(differ)
            for y, x in zip(ys, xs):
                assert (cov[max(y - 1, 0):y + 2, max(x - 1, 0):x + 2] == 0).any()
        both = (cov == 1) & (rcov == 1)
        assert np.abs(fpos[both] - rfpos[both]).max() < 3e-4
        err = np.abs(mine - theirs).max(axis=-1)
        assert err.max() <= 2.0 / 255.0
        assert (err > 1e-5).mean() < 1e-3                 # a few rays whose truncated step count flips by one
        a8, b8 = oracle.encode(mine, oracle.OUT_RGBA8_UNORM), oracle.encode(theirs, oracle.OUT_RGBA8_UNORM)
        from util import psnr8
        assert psnr8(a8, b8) >= 60.0


def test_oracle_against_the_golden_outputs_of_the_compiled_shaders(oracle):
    """runs everywhere: the committed outputs of the reference's compiled shaders (make_golden_glsl.py)"""
    g = np.load(GOLDEN_GLSL)
    W, H = (int(v) for v in g["size"])
    names = sorted({k.split("/")[0] for k in g.files if "/" in k})
    assert len(names) >= 4
    p0 = oracle.params_default()
    assert tuple(g["frag_constants"]) == (p0.density, p0.max_steps, -1.0, 1.0)
    for name in names:
        u0, u1, vol = g[name + "/ubo0"], g[name + "/ubo1"], g[name + "/volume"]
        p = oracle_params(oracle, u0, u1)
        ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
        mine, st = oracle.render(ov, p, W, H)
        fpos, cov = oracle.frag_positions(p, W, H)
        assert np.array_equal(cov, g[name + "/covered"]) and np.array_equal(fpos.view(np.uint32), g[name + "/frag_position"].view(np.uint32))
        assert np.array_equal(mine.view(np.uint32), g[name + "/out_color"].view(np.uint32)), name      # frag.glsl, bit for bit
        # vert.glsl on the cube's corners against the oracle's own matrices (float64 product, rounded)
        M, V, P = (u0[i:i + 16].reshape(4, 4).T.astype(np.float64) for i in (0, 16, 32))
        corners = np.concatenate([CUBE.astype(np.float64), np.ones((8, 1))], axis=1)
        assert np.abs((P @ V @ M @ corners.T).T - g[name + "/vert_gl_position"]).max() < 2e-5
        assert np.abs((M @ corners.T).T[:, :3] - g[name + "/vert_frag_position"]).max() < 1e-6
        # the whole draw
        err = np.abs(mine - g[name + "/draw_rgba"]).max(axis=-1)
        assert (cov != g[name + "/draw_covered"]).sum() <= 2 and err.max() <= 2.0 / 255.0 and (err > 1e-5).mean() < 2e-3
