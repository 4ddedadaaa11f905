"""FastNoise2 node trees: a small builder that writes the reference's encoded-node-tree strings
(FastNoise::NewFromEncodedNodeTree, vendor/noise/include/FastNoise/FastNoise.h:49) and the binding of
libvolrt's compiler / device evaluator for them (include/volrt.h, "FastNoise2 node trees").

Wire format (restated from the behaviour of vendor/noise/src/FastNoise/Metadata.cpp:51-183): base64 of, per node in
pre-order, {u16 id, 4 bytes per member variable (f32 / i32 / enum index), every node lookup as a nested node, every
hybrid as u8 0 + f32 constant or u8 1 + nested node}.  A node object used twice is written once; later uses are
{u16 0xFFFF, u16 index in post-order}.  SCHEMA is the member table the reference itself reports
(tests/golden/fastnoise2_metadata.json, dumped by oracle/ref_shim.cpp: vref_metadata_dump).

    from volumetricrenderer_b200.noisetree import N
    tree = N("FractalFBm", Octaves=5, Source=N("Perlin"), Gain=0.5)
    enc = tree.encode()                          # the string FastNoise2's NoiseTool would export
    grid, (mn, mx) = renderer.noise_tree_grid(enc, (128, 128, 128), frequency=0.03, seed=1337)
"""
import base64
import ctypes as C
import struct

import numpy as np

from . import capi

# name: (id, [(variable, type f|i|e, default)], [node lookups], [(hybrid, default)])
SCHEMA = {
    'Constant': (0, [('Value', 'f', 1)], [], []),
    'White': (1, [], [], []),
    'Checkerboard': (2, [('Size', 'f', 1)], [], []),
    'SineWave': (3, [('Scale', 'f', 1)], [], []),
    'PositionOutput': (4, [('MultiplierX', 'f', 0), ('MultiplierY', 'f', 0), ('MultiplierZ', 'f', 0), ('MultiplierW', 'f', 0), ('OffsetX', 'f', 0), ('OffsetY', 'f', 0), ('OffsetZ', 'f', 0), ('OffsetW', 'f', 0)], [], []),
    'DistanceToPoint': (5, [('DistanceFunction', 'e', 0), ('PointX', 'f', 0), ('PointY', 'f', 0), ('PointZ', 'f', 0), ('PointW', 'f', 0)], [], []),
    'Value': (6, [], [], []),
    'Perlin': (7, [], [], []),
    'Simplex': (8, [], [], []),
    'OpenSimplex2': (9, [], [], []),
    'CellularValue': (10, [('DistanceFunction', 'e', 1), ('ValueIndex', 'i', 0)], [], [('JitterModifier', 1)]),
    'CellularDistance': (11, [('DistanceFunction', 'e', 1), ('DistanceIndex0', 'i', 0), ('DistanceIndex1', 'i', 1), ('ReturnType', 'e', 0)], [], [('JitterModifier', 1)]),
    'CellularLookup': (12, [('DistanceFunction', 'e', 1), ('LookupFrequency', 'f', 0.100000001)], ['Lookup'], [('JitterModifier', 1)]),
    'FractalFBm': (13, [('Octaves', 'i', 3), ('Lacunarity', 'f', 2)], ['Source'], [('Gain', 0.5), ('WeightedStrength', 0)]),
    'FractalPingPong': (14, [('Octaves', 'i', 3), ('Lacunarity', 'f', 2)], ['Source'], [('Gain', 0.5), ('WeightedStrength', 0), ('PingPongStrength', 2)]),
    'FractalRidged': (15, [('Octaves', 'i', 3), ('Lacunarity', 'f', 2)], ['Source'], [('Gain', 0.5), ('WeightedStrength', 0)]),
    'DomainWarpGradient': (16, [('WarpFrequency', 'f', 0.5)], ['Source'], [('WarpAmplitude', 1)]),
    'DomainWarpFractalProgressive': (17, [('Octaves', 'i', 3), ('Lacunarity', 'f', 2)], ['DomainWarpSource'], [('Gain', 0.5), ('WeightedStrength', 0)]),
    'DomainWarpFractalIndependant': (18, [('Octaves', 'i', 3), ('Lacunarity', 'f', 2)], ['DomainWarpSource'], [('Gain', 0.5), ('WeightedStrength', 0)]),
    'DomainScale': (19, [('Scale', 'f', 1)], ['Source'], []),
    'DomainOffset': (20, [], ['Source'], [('OffsetX', 0), ('OffsetY', 0), ('OffsetZ', 0), ('OffsetW', 0)]),
    'DomainRotate': (21, [('Yaw', 'f', 0), ('Pitch', 'f', 0), ('Roll', 'f', 0)], ['Source'], []),
    'SeedOffset': (22, [('SeedOffset', 'i', 1)], ['Source'], []),
    'Remap': (23, [('FromMin', 'f', -1), ('FromMax', 'f', 1), ('ToMin', 'f', 0), ('ToMax', 'f', 1)], ['Source'], []),
    'ConvertRGBA8': (24, [('Min', 'f', -1), ('Max', 'f', 1)], ['Source'], []),
    'Add': (25, [], ['LHS'], [('RHS', 0)]),
    'Subtract': (26, [], [], [('LHS', 0), ('RHS', 0)]),
    'Multiply': (27, [], ['LHS'], [('RHS', 0)]),
    'Divide': (28, [], [], [('LHS', 0), ('RHS', 0)]),
    'Min': (29, [], ['LHS'], [('RHS', 0)]),
    'Max': (30, [], ['LHS'], [('RHS', 0)]),
    'MinSmooth': (31, [], ['LHS'], [('RHS', 0), ('Smoothness', 0.100000001)]),
    'MaxSmooth': (32, [], ['LHS'], [('RHS', 0), ('Smoothness', 0.100000001)]),
    'Fade': (33, [], ['A', 'B'], [('Fade', 0.5)]),
    'Terrace': (34, [('Multiplier', 'f', 1), ('Smoothness', 'f', 0)], ['Source'], []),
    'PowFloat': (35, [], [], [('Value', 2), ('Pow', 2)]),
    'PowInt': (36, [('Pow', 'i', 2)], ['Value'], []),
    'DomainAxisScale': (37, [('ScaleX', 'f', 1), ('ScaleY', 'f', 1), ('ScaleZ', 'f', 1), ('ScaleW', 'f', 1)], ['Source'], []),
    'AddDimension': (38, [], ['Source'], [('NewDimensionPosition', 0)]),
    'RemoveDimension': (39, [('RemoveDimension', 'e', 1)], ['Source'], []),
    'GeneratorCache': (40, [], ['Source'], []),
    'OpenSimplex2S': (41, [], [], []),
}


class N:
    """one node; keyword arguments name its members (spaces of the reference's member names removed, per-dimension
    members suffixed X/Y/Z/W); unnamed members keep the reference's defaults"""

    def __init__(self, name, **members):
        if name not in SCHEMA:
            raise KeyError("unknown FastNoise2 node %r" % name)
        self.name, self.members = name, members
        _id, vs, ls, hs = SCHEMA[name]
        known = {v[0] for v in vs} | set(ls) | {h[0] for h in hs}
        bad = set(members) - known
        if bad:
            raise KeyError("%s has no member(s) %s; it has %s" % (name, sorted(bad), sorted(known)))
        for l in ls:
            if not isinstance(members.get(l), N):
                raise ValueError("%s.%s must be a node" % (name, l))

    def _write(self, out, order):
        if id(self) in order:
            out += struct.pack("<HH", 0xFFFF, order[id(self)])
            return
        nid, vs, ls, hs = SCHEMA[self.name]
        out += struct.pack("<H", nid)
        for vname, ty, default in vs:
            v = self.members.get(vname, default)
            out += struct.pack("<f", float(v)) if ty == "f" else struct.pack("<i", int(v))
        for l in ls:
            self.members[l]._write(out, order)
        for hname, default in hs:
            v = self.members.get(hname, default)
            if isinstance(v, N):
                out += b"\x01"
                v._write(out, order)
            else:
                out += b"\x00" + struct.pack("<f", float(v))
        order[id(self)] = len(order)

    def encode(self):
        out = bytearray()
        self._write(out, {})
        return base64.b64encode(bytes(out)).decode()


class CompiledTree:
    """vr_noise_tree handle (host-only: compiling needs no device)"""

    def __init__(self, encoded):
        self._lib = capi.load()
        self.encoded = encoded if isinstance(encoded, str) else encoded.encode()
        h = C.c_void_p()
        err = C.create_string_buffer(256)
        st = self._lib.vr_noise_tree_compile(self.encoded.encode(), C.byref(h), err, len(err))
        if st != capi.VR_OK:
            raise ValueError("%s: %s" % (capi.STATUS_NAMES[st], err.value.decode()))
        self.handle = h

    def info(self):
        v = [C.c_uint32() for _ in range(4)]
        self._lib.vr_noise_tree_info(self.handle, *[C.byref(x) for x in v])
        return dict(zip(("nodes", "instructions", "value_regs", "position_regs"), [x.value for x in v]))

    def close(self):
        if self.handle:
            self._lib.vr_noise_tree_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def grid_static(tree, shape, frequency, seed, start=(0, 0, 0), device=0, stream=None, want_grid=True):
    """vr_noise_tree_grid_static: the allocation-free, context-free variant of Renderer.noise_tree_grid (include/volrt.h).
    tree: encoded string, N or CompiledTree; shape = (nz, ny, nx) with nz*ny*nx <= capi.STATIC_GRID_MAX_POINTS."""
    import numpy as np
    lib = capi.load()
    own = None
    if isinstance(tree, N):
        tree = tree.encode()
    if not isinstance(tree, CompiledTree):
        tree = own = CompiledTree(tree)
    nz, ny, nx = shape
    out = np.empty(shape, np.float32) if want_grid else None
    mm = (C.c_float * 2)()
    err = C.create_string_buffer(384)
    try:
        st = lib.vr_noise_tree_grid_static(device, tree.handle, start[0], start[1], start[2], nx, ny, nz, frequency, seed,
                                           out.ctypes.data_as(C.POINTER(C.c_float)) if want_grid else None, mm, stream, err, len(err))
    finally:
        if own is not None:
            own.close()
    if st != capi.VR_OK:
        from .renderer import VolrtError
        raise VolrtError(st, err.value.decode())
    return out, (mm[0], mm[1])
