"""The C++ host above the C ABI (host/): builds against include/volrt.h + libvolrt.so only, fails loudly
without a device, and on a GPU replays the reference's main() loop headless (TestMain.cpp:173-256)."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "host", "volrt_headless")


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host")])


def test_host_builds_and_does_not_touch_the_oracle():
    _build()
    assert os.path.exists(EXE)
    for f in os.listdir(os.path.join(ROOT, "host")):
        if f.endswith((".cpp", ".hpp", "Makefile")):
            src = open(os.path.join(ROOT, "host", f)).read()
            assert "volrt_oracle" not in src and "oracle/" not in src
    ldd = subprocess.run(["ldd", EXE], capture_output=True, text=True).stdout
    assert "libvolrt.so" in ldd and "oracle" not in ldd


def test_host_struct_layout_is_the_references():
    """GlobalShaderData must be the 140-byte C++ image (MediaScroll at 76) for quirk Q2 to reproduce;
    headless_main.cpp static_asserts it — building is the check; here: the header mirrors the names."""
    hpp = open(os.path.join(ROOT, "host", "volrt_host.hpp")).read()
    for name in ("class Renderer", "class Texture3D", "class UniformBuffer", "void Init(", "void Shutdown(", "void AddRenderPass(",
                 "void EnqueueRenderPass(", "void BeginFrame(", "void EndFrame(", "GetSwapchainExtent", "GetCurrentFrame",
                 "GetFramesCount", "void Update(const T* data, uint32_t index"):
        assert name in hpp, name
    main = open(os.path.join(ROOT, "host", "headless_main.cpp")).read()
    assert "sizeof(GlobalShaderData) == 140" in main and "sizeof(ObjectShaderData) == 192" in main


def test_host_fails_loudly_without_a_device():
    _build()
    r = subprocess.run([EXE, "--frames", "1"], capture_output=True, text=True)
    if r.returncode == 0:
        pytest.skip("CUDA device present")
    assert r.returncode == 1 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_headless_loop_matches_the_python_host(tmp_path, oracle):
    """same key trace through the C++ host and through the Python host -> same frames (sum of R); and the frame the C++ host
    wrote to disk is the ORACLE's frame of that scene (and the compiled reference shaders' where oracle/_ref is present)"""
    import volumetricrenderer_b200 as vr
    from volumetricrenderer_b200 import capi, scene
    _build()
    r = subprocess.run([EXE, "--frames", "12", "--keys", "D*6,W*6", "--every", "1", "--out", str(tmp_path)],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    sums = {int(m.group(1)): int(m.group(2)) for m in re.finditer(r"frame\s+(\d+) key . phi\s+\S+ theta\s+\S+ sum\(R\) (\d+)", r.stdout)}
    assert len(sums) == 12
    phi = theta = 0.0
    with vr.Renderer(1280, 720, capi.OUT_RGBA8_UNORM) as rr:
        rr.generate_volume()
        for f in range(12):
            if f < 6:
                phi += np.float32(100.0) * np.float32(0.016)
            else:
                theta -= np.float32(100.0) * np.float32(0.016)
            u0, u1 = scene.default_scene(float(phi), float(theta), 0.1 * f)
            rr.set_params(u0, u1)
            img = rr.render()[0]
            s = int(img[..., 0].astype(np.int64).sum())
            assert abs(s - sums[f]) <= 2e-4 * max(s, 1), (f, s, sums[f])     # libm vs numpy sin/cos: a few ulp in Model
        baked = rr.download_volume()                                         # the bake is deterministic: the C++ host marched these texels
    # the PPM of the last frame is the frame
    ppm = open(os.path.join(str(tmp_path), "frame_0011.ppm"), "rb").read()
    assert ppm.startswith(b"P6\n1280 720\n255\n")
    px = np.frombuffer(ppm[len(b"P6\n1280 720\n255\n"):], np.uint8).reshape(720, 1280, 3)
    assert int(px[..., 0].astype(np.int64).sum()) == sums[11]
    # ... and it is the right frame: the oracle (and the reference's own shaders compiled as C++) on the scene of frame 11
    from util import oracle_params, psnr8, max_abs8
    p = oracle_params(oracle, u0, u1)
    ov, _keep = oracle.make_volume(baked, oracle.VOL_RGBA8)
    ref, st = oracle.render(ov, p, 1280, 720)
    ref8 = oracle.encode(ref, oracle.OUT_RGBA8_UNORM)[..., :3]
    assert st["covered"] > 150000 and ref8.max() > 40
    assert max_abs8(px, ref8) <= 2 and psnr8(px, ref8) >= 50.0
    if oracle.have_ref_glsl():
        drawn, _cov, _fp = oracle.ref_draw(u0, u1, baked, 1280, 720)
        d8 = oracle.encode(drawn, oracle.OUT_RGBA8_UNORM)[..., :3]
        assert max_abs8(px, d8) <= 2 and psnr8(px, d8) >= 50.0


@pytest.mark.gpu
def test_media_scroll_never_moves_in_the_reference_and_can_be_made_to():
    """The reference writes -frameTime into MediaScroll[0][0] only (TestMain.cpp:233-238).  With its 140-byte struct that float
    lands in std140 padding (quirk Q2); with the shader's own layout (--std140) it lands in MediaScroll[0].x, which only a
    commented-out line reads (frag.glsl:65).  Either way the frames do not depend on time.  A host that also fills .y/.z/.w
    in the std140 layout gets the scroll the shader implements (frag.glsl:67-69)."""
    _build()

    def sums(*extra):
        r = subprocess.run([EXE, "--frames", "4", "--keys", "-", "--every", "1", "--dt", "0.5", *extra], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stderr
        return [int(m.group(1)) for m in re.finditer(r"sum\(R\) (\d+)", r.stdout)]
    q2 = sums()
    fixed = sums("--std140")
    moving = sums("--std140", "--scroll-yzw")
    assert len(q2) == 4 and len(set(q2)) == 1
    assert fixed == q2
    assert moving[0] == q2[0] and len(set(moving)) == 4        # frame 0 has frameTime 0; later frames scroll
