// The node-tree interpreter (one thread, one position): shared by the grid kernel of noise_tree.cu and the
// VR_DENSITY_TREE march of march_kernels.cu.  The program format is volrt_internal.h: TreeInstr.
#pragma once
#include "volrt_internal.h"
#include "noise_nodes.cuh"

namespace volrt {

// ------------------------------------------------------------------------------------------------ interpreter
// one of the lattice generators without members (what a fractal is usually built over); the Perlin gradients come out
// of the shared-memory table of noise_device.cuh
__device__ __forceinline__ float tree_leaf(int gen, const GradLut* lut, int s, float x, float y, float z) {
    switch (gen) {
        case T_VALUE: return value3(s, x, y, z);
        case T_SIMPLEX: return simplex3(s, x, y, z);
        case T_OPENSIMPLEX2: return opensimplex2_3(s, x, y, z);
        case T_OPENSIMPLEX2S: return opensimplex2s_3(s, x, y, z);
        default: return perlin3_lut(lut, s, x, y, z) * kPerlinScale;
    }
}

// lut: a GradLut in shared memory, filled by grad_lut_init before the first call
__device__ __forceinline__ float tree_eval(const TreeInstr* __restrict__ prog, int n_instr, int result_reg, const GradLut* lut, int seed,
                                           float x, float y, float z) {
    float V[kTreeMaxValues];
    float P[kTreeMaxPositions][3];
    P[0][0] = x; P[0][1] = y; P[0][2] = z;
    for (int pc = 0; pc < n_instr; ++pc) {
        const TreeInstr& in = prog[pc];
        const int s = seed + in.seed_off;
        const float px = P[in.pos][0], py = P[in.pos][1], pz = P[in.pos][2];
        switch (in.op) {
            case T_CONST: V[in.dst] = in.f[0]; break;
            case T_WHITE: V[in.dst] = white3(s, px, py, pz); break;
            // a lattice generator alone, or a whole fractal over one with constant gain / weighted strength (the compiler
            // fuses those, noise_tree.cu: emit_fractal): octaves run here with sum, amplitude and position in registers
            // instead of one interpreted instruction per octave step.  Same operations in the same order as the
            // T_*_FIRST / T_*_NEXT instructions below (Fractal.inl), so fused and unfused programs agree bit for bit.
            // i0 = generator, i1 = octaves, i2 = 0 fBm | 1 ridged | 2 ping-pong; f = bounding, lacunarity, gain, weighted
            // strength, ping-pong strength
            case T_VALUE: case T_PERLIN: case T_SIMPLEX: case T_OPENSIMPLEX2: case T_OPENSIMPLEX2S: case T_FRACTAL_LEAF: {
                const bool fr = in.op == T_FRACTAL_LEAF;
                const int gen = fr ? in.i0 : (int)in.op, oct = fr ? in.i1 : 1, kind = fr ? in.i2 : 0;
                const float lac = in.f[1], gain = in.f[2], ws = in.f[3], pps = in.f[4];
                float fx = px, fy = py, fz = pz, amp = fr ? in.f[0] : 1.0f;
                float n = tree_leaf(gen, lut, s, fx, fy, fz);
                float sh = kind == 0 ? n : kind == 1 ? fmaf(fabsf(n), -2.0f, 1.0f) : ping_pong((n + 1.0f) * pps);
                float sum = sh * amp;
                for (int i = 1; i < oct; ++i) {
                    fx *= lac; fy *= lac; fz *= lac;
                    const float w = kind == 0 ? (n + 1.0f) * 0.5f : kind == 1 ? 1.0f - fabsf(n) : (sh + 1.0f) * 0.5f;
                    amp *= lerpf(1.0f, w, ws);
                    amp *= gain;
                    n = tree_leaf(gen, lut, s + i, fx, fy, fz);
                    sh = kind == 0 ? n : kind == 1 ? fmaf(fabsf(n), -2.0f, 1.0f) : ping_pong((n + 1.0f) * pps);
                    sum = fmaf(sh, amp, sum);
                }
                V[in.dst] = sum;
                break;
            }
            case T_CHECKERBOARD: V[in.dst] = checkerboard3(in.f[0], px, py, pz); break;
            case T_SINE: V[in.dst] = (fs_sin(px * in.f[0]) * fs_sin(py * in.f[0])) * fs_sin(pz * in.f[0]); break;
            case T_POSITION_OUTPUT: V[in.dst] = ((px + in.f[3]) * in.f[0] + (py + in.f[4]) * in.f[1]) + (pz + in.f[5]) * in.f[2]; break;
            case T_DISTANCE_TO_POINT: V[in.dst] = calc_distance(in.i0, px - in.f[0], py - in.f[1], pz - in.f[2]); break;
            case T_CELL_VALUE: V[in.dst] = cellular_value3(s, px, py, pz, V[in.a], in.i0, in.i1); break;
            case T_CELL_DISTANCE:
                // the node's defaults (what the reference's bake uses, TestMain.cpp:43) keep only the nearest squared distance
                if (in.i0 == 1 && in.i1 == 0 && in.i3 == 0 && V[in.a] == 1.0f) V[in.dst] = cellular3(s, px, py, pz);
                else V[in.dst] = cellular_distance3(s, px, py, pz, V[in.a], in.i0, in.i1, in.i2, in.i3);
                break;
            case T_CELL_LOOKUP_POS: {
                float c[3];
                cellular_lookup3(s, px, py, pz, V[in.a], in.i0, c);
                P[in.dst][0] = c[0] * in.f[0]; P[in.dst][1] = c[1] * in.f[0]; P[in.dst][2] = c[2] * in.f[0];
                break;
            }
            case T_ADD: V[in.dst] = V[in.a] + V[in.b]; break;
            case T_SUB: V[in.dst] = V[in.a] - V[in.b]; break;
            case T_MUL: V[in.dst] = V[in.a] * V[in.b]; break;
            case T_DIV: V[in.dst] = V[in.a] / V[in.b]; break;
            case T_MIN: V[in.dst] = fminf(V[in.a], V[in.b]); break;
            case T_MAX: V[in.dst] = fmaxf(V[in.a], V[in.b]); break;
            case T_MIN_SMOOTH: case T_MAX_SMOOTH: {                                      // Blends.inl (MinSmooth / MaxSmooth)
                const float sg = in.op == T_MAX_SMOOTH ? -1.0f : 1.0f;
                const float a = sg * V[in.a], b = sg * V[in.b];
                const float sm = fmaxf(1.175494351e-38f, fabsf(V[in.c]));
                float h = fmaxf(sm - fabsf(a - b), 0.0f);
                h *= 1.0f / sm;
                V[in.dst] = sg * fmaf(-(1.0f / 6.0f), h * h * h * sm, fminf(a, b));
                break;
            }
            case T_FADE: { const float fd = fabsf(V[in.c]); V[in.dst] = fmaf(V[in.a], 1.0f - fd, V[in.b] * fd); break; }
            case T_POW_FLOAT: V[in.dst] = fs_pow(V[in.a], V[in.b]); break;
            case T_POW_INT: { const float v = V[in.a]; float p = v * v; for (int i = 2; i < in.i0; ++i) p *= v; V[in.dst] = p; break; }
            case T_REMAP: V[in.dst] = in.f[3] + ((V[in.a] - in.f[0]) / in.f[1] * in.f[2]); break;
            case T_TERRACE: {
                const float src = V[in.a] * in.f[0];
                float r = rintf(src);
                if (in.f[1] != 0.0f) {
                    float diff = r - src;
                    const bool neg = diff < 0.0f;
                    diff = fminf((0.5f - fabsf(diff)) * in.f[2], 0.5f);
                    r += neg ? 0.5f - diff : diff - 0.5f;
                }
                V[in.dst] = r * in.f[3];
                break;
            }
            case T_CONVERT_RGBA8: {
                float v = fmaxf(fminf(V[in.a], in.f[1]), in.f[0]);
                const int b = __float2int_rn((v - in.f[0]) * in.f[2]);
                V[in.dst] = __int_as_float((int)(255u << 24) | b | (b << 8) | (b << 16));
                break;
            }
            case T_COPY: V[in.dst] = V[in.a]; break;
            // fractals (Fractal.inl): a = this octave's source value, b = amplitude register, c = gain, d = weighted strength,
            // i0 = ping-pong strength register, i1 = the previous octave's source value
            case T_FBM_FIRST: V[in.b] = in.f[0]; V[in.dst] = V[in.a] * in.f[0]; break;
            case T_FBM_NEXT: {
                float amp = V[in.b] * lerpf(1.0f, (V[in.i1] + 1.0f) * 0.5f, V[in.d]);
                amp *= V[in.c];
                V[in.b] = amp;
                V[in.dst] = fmaf(V[in.a], amp, V[in.dst]);
                break;
            }
            case T_RIDGED_FIRST: V[in.b] = in.f[0]; V[in.dst] = fmaf(fabsf(V[in.a]), -2.0f, 1.0f) * in.f[0]; break;
            case T_RIDGED_NEXT: {
                float amp = V[in.b] * lerpf(1.0f, 1.0f - fabsf(V[in.i1]), V[in.d]);
                amp *= V[in.c];
                V[in.b] = amp;
                V[in.dst] = fmaf(fmaf(fabsf(V[in.a]), -2.0f, 1.0f), amp, V[in.dst]);
                break;
            }
            case T_PINGPONG_FIRST: V[in.b] = in.f[0]; V[in.dst] = ping_pong((V[in.a] + 1.0f) * V[in.i0]) * in.f[0]; break;
            case T_PINGPONG_NEXT: {
                const float prev = ping_pong((V[in.i1] + 1.0f) * V[in.i0]);
                float amp = V[in.b] * lerpf(1.0f, (prev + 1.0f) * 0.5f, V[in.d]);
                amp *= V[in.c];
                V[in.b] = amp;
                V[in.dst] = fmaf(ping_pong((V[in.a] + 1.0f) * V[in.i0]), amp, V[in.dst]);
                break;
            }
            case T_POS_COPY: P[in.dst][0] = px; P[in.dst][1] = py; P[in.dst][2] = pz; break;
            case T_POS_SCALE: P[in.dst][0] = px * in.f[0]; P[in.dst][1] = py * in.f[0]; P[in.dst][2] = pz * in.f[0]; break;
            case T_POS_AXIS_SCALE: P[in.dst][0] = px * in.f[0]; P[in.dst][1] = py * in.f[1]; P[in.dst][2] = pz * in.f[2]; break;
            case T_POS_OFFSET: P[in.dst][0] = px + V[in.a]; P[in.dst][1] = py + V[in.b]; P[in.dst][2] = pz + V[in.c]; break;
            case T_POS_ROTATE:
                P[in.dst][0] = fmaf(px, in.f[0], fmaf(py, in.f[1], pz * in.f[2]));
                P[in.dst][1] = fmaf(px, in.f[3], fmaf(py, in.f[4], pz * in.f[5]));
                P[in.dst][2] = fmaf(px, in.f[6], fmaf(py, in.f[7], pz * in.f[8]));
                break;
            case T_POS_WARP: V[in.b] = warp_gradient3(s, V[in.a], px * in.f[0], py * in.f[0], pz * in.f[0], P[in.dst]); break;
            case T_WARP_AMP_NEXT: {
                float amp = V[in.dst] * lerpf(1.0f, 1.0f - V[in.a], V[in.b]);
                V[in.dst] = amp * V[in.c];
                break;
            }
            default: break;
        }
    }
    return V[result_reg];
}

}  // namespace volrt
