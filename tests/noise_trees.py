"""The node trees the rank-3 parity tests run: every node libvolrt builds, alone and
combined, with the members that change behaviour set away from their defaults.  Shared by the fixture generator
(tests/golden/make_golden.py) and the tests; order and names are part of the fixture."""
from volumetricrenderer_b200.noisetree import N

# name -> (tree, frequency, seed, (x0, y0, z0))
_shared = N("Simplex")
TREES = {
    "constant": (N("Constant", Value=0.37), 0.1, 1, (0, 0, 0)),
    "white": (N("White"), 0.1, 5, (3, -2, 7)),
    "value": (N("Value"), 0.23, 11, (-5, 2, 9)),
    "perlin": (N("Perlin"), 0.19, 3, (0, 0, 0)),
    "simplex": (N("Simplex"), 0.15, 4, (-7, 0, 3)),
    "opensimplex2": (N("OpenSimplex2"), 0.21, 9, (2, -9, 5)),
    "sinewave": (N("SineWave", Scale=1.7), 0.31, 0, (1, 2, 3)),
    "position_output": (N("PositionOutput", MultiplierX=0.5, MultiplierY=-1.25, MultiplierZ=2.0, OffsetX=3.0, OffsetY=0.5, OffsetZ=-1.0), 0.2, 0, (-3, 1, 0)),
    "distance_to_point_euclid": (N("DistanceToPoint", DistanceFunction=0, PointX=0.7, PointY=-0.2, PointZ=1.1), 0.2, 0, (-4, -4, -4)),
    "distance_to_point_hybrid": (N("DistanceToPoint", DistanceFunction=3, PointX=0.7, PointY=-0.2, PointZ=1.1), 0.2, 0, (-4, -4, -4)),
    "cell_value_default": (N("CellularValue"), 0.13, 21, (0, 0, 0)),
    "cell_value_idx2_manhattan": (N("CellularValue", DistanceFunction=2, ValueIndex=2, JitterModifier=0.7), 0.17, 22, (-6, 3, 1)),
    "cell_distance_default": (N("CellularDistance"), 0.03, 2, (0, 0, 0)),
    "cell_distance_euclid_sub": (N("CellularDistance", DistanceFunction=0, DistanceIndex0=1, DistanceIndex1=0, ReturnType=2), 0.11, 8, (5, 5, -5)),
    "cell_distance_maxaxis_mul": (N("CellularDistance", DistanceFunction=4, DistanceIndex0=0, DistanceIndex1=2, ReturnType=3, JitterModifier=1.3), 0.09, 8, (0, -8, 4)),
    "cell_distance_hybrid_add": (N("CellularDistance", DistanceFunction=3, DistanceIndex0=0, DistanceIndex1=3, ReturnType=1), 0.12, 30, (1, 1, 1)),
    "cell_distance_div_jitter_node": (N("CellularDistance", DistanceFunction=1, DistanceIndex0=0, DistanceIndex1=1, ReturnType=4,
                                        JitterModifier=N("Remap", Source=N("Perlin"), FromMin=-1, FromMax=1, ToMin=0.5, ToMax=1.0)), 0.1, 31, (0, 0, 0)),
    "cell_lookup": (N("CellularLookup", Lookup=N("Perlin"), LookupFrequency=0.35, DistanceFunction=1), 0.14, 40, (-2, 6, 0)),
    "fbm_perlin5": (N("FractalFBm", Octaves=5, Source=N("Perlin")), 0.03 * 128 / 24, 1337, (0, 0, 0)),
    "fbm_weighted_gain_node": (N("FractalFBm", Octaves=4, Lacunarity=2.3, Source=N("OpenSimplex2"), Gain=N("Remap", Source=N("Value"), FromMin=-1, FromMax=1, ToMin=0.3, ToMax=0.7),
                                 WeightedStrength=0.6), 0.08, 77, (3, 3, 3)),
    "ridged_simplex": (N("FractalRidged", Octaves=4, Lacunarity=1.9, Source=N("Simplex"), Gain=0.6, WeightedStrength=0.4), 0.07, 12, (0, -3, 8)),
    "pingpong_value": (N("FractalPingPong", Octaves=3, Source=N("Value"), Gain=0.55, WeightedStrength=0.2, PingPongStrength=2.5), 0.12, 13, (2, 2, -2)),
    "warp_gradient": (N("DomainWarpGradient", WarpFrequency=0.6, WarpAmplitude=1.5, Source=N("Perlin")), 0.11, 50, (0, 0, 0)),
    "warp_progressive": (N("DomainWarpFractalProgressive", Octaves=3, Lacunarity=2.1, Gain=0.5, WeightedStrength=0.3,
                           DomainWarpSource=N("DomainWarpGradient", WarpFrequency=0.4, WarpAmplitude=2.0, Source=N("Simplex"))), 0.09, 51, (-4, 0, 4)),
    "warp_independant": (N("DomainWarpFractalIndependant", Octaves=3, Lacunarity=1.8, Gain=0.6,
                           DomainWarpSource=N("DomainWarpGradient", WarpFrequency=0.5, WarpAmplitude=1.2, Source=N("CellularDistance"))), 0.09, 52, (0, 5, -5)),
    "domain_scale_offset": (N("DomainOffset", OffsetX=0.4, OffsetY=N("Perlin"), OffsetZ=-1.5, Source=N("DomainScale", Scale=1.7, Source=N("Simplex"))), 0.13, 60, (1, -1, 1)),
    "domain_rotate_axis_scale": (N("DomainRotate", Yaw=0.6, Pitch=-0.3, Roll=1.1, Source=N("DomainAxisScale", ScaleX=1.0, ScaleY=2.5, ScaleZ=0.5, Source=N("Perlin"))), 0.16, 61, (0, 0, 0)),
    "seed_offset_cache": (N("SeedOffset", SeedOffset=17, Source=N("GeneratorCache", Source=N("Perlin"))), 0.19, 3, (0, 0, 0)),
    "remap_terrace": (N("Terrace", Multiplier=4.0, Smoothness=0.3, Source=N("Remap", Source=N("Perlin"), FromMin=-1, FromMax=1, ToMin=0, ToMax=2)), 0.1, 62, (0, 0, 0)),
    "terrace_hard": (N("Terrace", Multiplier=3.0, Smoothness=0.0, Source=N("Simplex")), 0.1, 63, (0, 0, 0)),
    "blend_arith": (N("Add", LHS=N("Multiply", LHS=N("Perlin"), RHS=N("Value")), RHS=N("Subtract", LHS=N("Simplex"), RHS=0.25)), 0.14, 64, (0, 0, 0)),
    "blend_divide_minmax": (N("Max", LHS=N("Min", LHS=N("Perlin"), RHS=N("Simplex")), RHS=N("Divide", LHS=N("Value"), RHS=3.0)), 0.14, 65, (2, 0, -1)),
    "blend_smooth": (N("MinSmooth", LHS=N("Perlin"), RHS=N("MaxSmooth", LHS=N("Simplex"), RHS=-0.2, Smoothness=0.3), Smoothness=0.25), 0.12, 66, (0, 0, 0)),
    "blend_fade_pow": (N("Fade", A=N("PowInt", Pow=3, Value=N("Perlin")), B=N("PowFloat", Value=N("Remap", Source=N("Simplex"), FromMin=-1, FromMax=1, ToMin=0.1, ToMax=1.1), Pow=1.7),
                         Fade=N("Value")), 0.12, 67, (0, 0, 0)),
    "shared_subtree": (N("Add", LHS=N("DomainScale", Scale=2.0, Source=_shared), RHS=N("Multiply", LHS=_shared, RHS=0.5)), 0.1, 68, (0, 0, 0)),
    "convert_rgba8": (N("ConvertRGBA8", Min=-0.8, Max=0.9, Source=N("Perlin")), 0.17, 69, (0, 0, 0)),
    "cloud_density": (N("Max", RHS=0.0, LHS=N("Subtract", RHS=0.45, LHS=N("Remap", FromMin=-1, FromMax=1, ToMin=0, ToMax=1,
                        Source=N("DomainWarpGradient", WarpFrequency=0.7, WarpAmplitude=0.6, Source=N("FractalFBm", Octaves=5, Source=N("Perlin")))))), 0.16, 1337, (0, 0, 0)),
    # added after the first fixture: the smooth simplex variant, alone / under a looping fractal / under an unrolled one, and
    # the checkerboard (size and frequency chosen so that no grid point sits on a cell border, where the reference's
    # approximate reciprocal decides the side)
    "opensimplex2s": (N("OpenSimplex2S"), 0.21, 9, (2, -9, 5)),
    "fbm_opensimplex2s": (N("FractalFBm", Octaves=4, Lacunarity=2.1, Source=N("OpenSimplex2S"), Gain=0.55), 0.09, 78, (-3, 4, 1)),
    "ridged_opensimplex2s_gain_node": (N("FractalRidged", Octaves=3, Source=N("OpenSimplex2S"), Gain=N("Remap", Source=N("Value"), FromMin=-1, FromMax=1, ToMin=0.4, ToMax=0.6)), 0.11, 79, (0, 0, 0)),
    "checkerboard": (N("Checkerboard", Size=0.9), 0.07, 0, (-5, 3, 0)),
    "checkerboard_blend": (N("Multiply", LHS=N("Perlin"), RHS=N("Checkerboard", Size=2.3)), 0.37, 70, (1, -6, 2)),
}
SHAPE = (9, 10, 12)        # (nz, ny, nx)
