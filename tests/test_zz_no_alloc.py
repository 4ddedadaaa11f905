"""vr_noise_tree_grid_static: the allocation-free variant of vr_noise_tree_grid (include/volrt.h; no counterpart in the
reference).  Its scratch is a file-scope device array that every call shares and
that keeps the previous call's contents, so the test makes two calls in a row — the second one smaller and with another tree,
so that stale contents (the longer program, the larger plane, the wider min/max range) would show — and holds each to the
existing entry point.  Both run the same kernel on the same program, so the tolerance is zero: bit-identical grids, identical
min/max.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, noisetree
from volumetricrenderer_b200.noisetree import N

pytestmark = pytest.mark.gpu

CALLS = [
    # tree, shape (nz, ny, nx), frequency, seed, start
    (N("FractalFBm", Octaves=5, Source=N("Perlin")), (64, 64, 64), 0.16, 1337, (0, 0, 0)),              # the full static plane
    (N("Add", LHS=N("Simplex"), RHS=N("Multiply", LHS=N("CellularDistance"), RHS=0.25)), (7, 11, 13), 0.11, 5, (-3, 2, 9)),
]


def test_static_entry_point_equals_the_allocating_one_twice_in_a_row():
    with vr.Renderer(32, 32) as r:
        refs = [r.noise_tree_grid(t, shape, f, seed, start) for t, shape, f, seed, start in CALLS]
    for rep in range(2):                                   # and the pair again: the array now holds call 2's leftovers
        for (t, shape, f, seed, start), (ref, (rmn, rmx)) in zip(CALLS, refs):
            got, (mn, mx) = noisetree.grid_static(t, shape, f, seed, start)
            assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), (rep, shape)
            assert (mn, mx) == (rmn, rmx) == (float(ref.min()), float(ref.max()))
    # min/max only (no output pointer), on a caller stream = the default stream
    _, (mn, mx) = noisetree.grid_static(CALLS[1][0], CALLS[1][1], CALLS[1][2], CALLS[1][3], CALLS[1][4], want_grid=False)
    assert (mn, mx) == refs[1][1]


def test_static_entry_point_refuses_what_does_not_fit():
    lib = capi.load()
    t = noisetree.CompiledTree(N("Perlin").encode())
    err = C.create_string_buffer(384)
    mm = (C.c_float * 2)()
    try:
        for nx, ny, nz in ((65, 64, 64), (4096, 4096, 1), (8192, 1, 1)):
            st = lib.vr_noise_tree_grid_static(0, t.handle, 0, 0, 0, nx, ny, nz, 0.1, 1, None, mm, None, err, len(err))
            assert st == capi.VR_ERR_INVALID_ARGUMENT and b"VR_STATIC_GRID_MAX_POINTS" in err.value
        assert lib.vr_noise_tree_grid_static(0, t.handle, 0, 0, 0, 0, 4, 4, 0.1, 1, None, mm, None, err, len(err)) == capi.VR_ERR_INVALID_ARGUMENT
        assert lib.vr_noise_tree_grid_static(0, None, 0, 0, 0, 4, 4, 4, 0.1, 1, None, mm, None, err, len(err)) == capi.VR_ERR_INVALID_ARGUMENT
        assert lib.vr_noise_tree_grid_static(99, t.handle, 0, 0, 0, 4, 4, 4, 0.1, 1, None, mm, None, err, len(err)) != capi.VR_OK
        # the largest grid that fits still works after the refusals
        st = lib.vr_noise_tree_grid_static(0, t.handle, 0, 0, 0, 64, 64, 64, 0.1, 1, None, mm, None, err, len(err))
        assert st == capi.VR_OK and mm[0] < 0.0 < mm[1]
    finally:
        t.close()
