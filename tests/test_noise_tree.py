"""FastNoise2 encoded node trees, host side: the wire-format table, the Python encoder and
libvolrt's decoder/compiler agree with what the reference reports and accepts.  No device needed."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from volumetricrenderer_b200 import capi, noisetree
from volumetricrenderer_b200.noisetree import N, CompiledTree
from noise_trees import TREES, SHAPE

HERE = os.path.dirname(os.path.abspath(__file__))
UNSUPPORTED = {"AddDimension", "RemoveDimension"}


def _metadata():
    return json.load(open(os.path.join(HERE, "golden", "fastnoise2_metadata.json")))


def _key(m):
    return m["name"].replace(" ", "") + ("" if m["dim"] < 0 else "XYZW"[m["dim"]])


def test_python_schema_is_the_reference_metadata():
    meta = _metadata()
    assert len(meta) == len(noisetree.SCHEMA) == 42
    for m in meta:
        nid, vs, ls, hs = noisetree.SCHEMA[m["name"]]
        assert nid == m["id"]
        assert [(v[0], v[1]) for v in vs] == [(_key(v), v["type"][0]) for v in m["variables"]]
        assert ls == [_key(x) for x in m["nodes"]]
        assert [h[0] for h in hs] == [_key(h) for h in m["hybrids"]]


def test_library_decoder_knows_every_node_class_of_the_reference():
    """a default-constructed node of every class (children = Perlin) parses to the end of the stream: the C++ table has
    the right member counts for all 42 classes; the ones libvolrt does not build are refused BY NAME"""
    for m in _metadata():
        name = m["name"]
        _id, vs, ls, hs = noisetree.SCHEMA[name]
        kw = {l: (N("DomainWarpGradient", Source=N("Perlin")) if l == "DomainWarpSource" else N("Perlin")) for l in ls}
        enc = N(name, **kw).encode()
        if name in UNSUPPORTED:
            with pytest.raises(ValueError) as e:
                CompiledTree(enc)
            assert "VR_ERR_UNSUPPORTED" in str(e.value) and name in str(e.value)
        else:
            info = CompiledTree(enc).info()
            assert info["nodes"] == 1 + len(ls) + (1 if "DomainWarpSource" in ls else 0) and info["instructions"] >= 1


def test_known_encoded_string():
    # the FastNoise2 wiki's example shape: FractalFBm(5 octaves, lacunarity 2, gain 0.5, weighted strength 0) over a coherent source
    enc = N("FractalFBm", Octaves=5, Source=N("Simplex")).encode()
    assert enc == "DQAFAAAAAAAAQAgAAAAAAD8AAAAAAA=="
    # constant gain over a lattice generator: one looping instruction; with a gain that is itself a node, one per octave step
    assert CompiledTree(enc).info() == {"nodes": 2, "instructions": 1, "value_regs": 1, "position_regs": 1}
    unrolled = CompiledTree(N("FractalFBm", Octaves=5, Source=N("Simplex"), Gain=N("Constant", Value=0.5)).encode()).info()
    assert unrolled["nodes"] == 3 and unrolled["instructions"] == 17 and unrolled["position_regs"] == 2


def test_fixture_strings_are_what_the_encoder_writes_and_compile():
    g = np.load(os.path.join(HERE, "golden", "fastnoise2_trees.npz"))
    assert list(g["names"]) == list(TREES)
    for name, (tree, freq, seed, start) in TREES.items():
        assert str(g[name + "/encoded"]) == tree.encode(), name
        info = CompiledTree(tree.encode()).info()
        assert 1 <= info["instructions"] <= 512 and info["value_regs"] <= 64 and info["position_regs"] <= 16
    # a node object used twice is written once and referenced: 5 node records, 4 distinct nodes
    assert CompiledTree(TREES["shared_subtree"][0].encode()).info()["nodes"] == 4


def test_malformed_streams_are_rejected():
    good = N("FractalFBm", Octaves=3, Source=N("Perlin")).encode()
    import base64
    raw = base64.b64decode(good)
    for bad in (raw[:-3], raw[:5], b"\xff\xff\x05\x00", b"\x63\x00", raw[:2] + b"\x40\x00\x00\x00" + raw[6:]):
        with pytest.raises(ValueError):
            CompiledTree(base64.b64encode(bad).decode())
    with pytest.raises(ValueError):
        CompiledTree("not base64 !!")
    with pytest.raises(ValueError) as e:                       # unrolled octaves overflow the register file: refused, not truncated
        g = N("Constant", Value=0.5)                            # a gain that is a node keeps the inner fractal unrolled too
        CompiledTree(N("FractalFBm", Octaves=16, Source=N("FractalFBm", Octaves=16, Gain=g, Source=N("Perlin"))).encode())
    assert "VR_ERR_UNSUPPORTED" in str(e.value)
    # with constant gains the inner fractal is one looping instruction and the same nesting fits
    assert CompiledTree(N("FractalFBm", Octaves=16, Source=N("FractalFBm", Octaves=16, Source=N("Perlin"))).encode()).info()["instructions"] < 64
    with pytest.raises(KeyError):
        N("Perlin", Frequency=2.0)
    with pytest.raises(ValueError):
        N("FractalFBm")                                        # Source is a required node


def test_reference_accepts_the_encoder_output(oracle):
    """when the reference's FastNoise2 is built here: its own decoder takes every string and reproduces the fixture"""
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    g = np.load(os.path.join(HERE, "golden", "fastnoise2_trees.npz"))
    for name, (tree, freq, seed, start) in TREES.items():
        ref, mm = oracle.ref_tree_grid(tree.encode(), SHAPE, freq, seed, start, level=oracle.LEVEL_AVX2)
        assert np.array_equal(ref.view(np.uint32), g[name + "/grid"].view(np.uint32)), name
    meta = oracle.ref_metadata()
    assert meta == _metadata()
