"""volumetricrenderer_b200 — host side of libvolrt, the B200-native (sm_100a CUDA) drop-in for
VolumetricRenderer's per-pixel volumetric ray-march (shaders/frag.glsl) and its noise-volume bake.

    capi      ctypes mirror of include/volrt.h (the C ABI; loads ./libvolrt.so)
    renderer  Renderer: the reference's renderer surface for this path, one C-ABI call per method
    scene     the two uniform blocks the reference's main() writes each frame (TestMain.cpp:222-245)
    csrc/     CUDA kernels + the C ABI implementation

There is no CPU path in this package.  The CPU oracle lives in /oracle and is test infrastructure.
"""
from . import capi, scene  # noqa: F401
from .renderer import Renderer, VolrtError, PinnedArray  # noqa: F401
