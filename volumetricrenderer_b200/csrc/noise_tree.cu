// FastNoise2 encoded node trees on the device.
//
// Host: decode the reference's wire format (FastNoise::NewFromEncodedNodeTree, FastNoise.h:49; format restated from
// the behaviour of vendor/noise/src/FastNoise/Metadata.cpp:199-300: base64 of {u16 node id | 0xFFFF + u16 back
// reference, i32/f32 per member variable, child per node lookup, (u8 is_node, child | f32) per hybrid}; node ids
// and member order as the reference registers them — tests/golden/fastnoise2_metadata.json is the dump the
// reference itself prints) and compile the tree into a straight-line program (volrt_internal.h: TreeInstr).
// Device: one thread per grid point interprets that program; every thread runs the same instruction, so the
// switch is warp-uniform.  Generators: noise_device.cuh / noise_nodes.cuh.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>

#include "noise_nodes.cuh"
#include "noise_tree_eval.cuh"

namespace volrt {

// ------------------------------------------------------------------------------------------------ wire format
namespace {

struct VarSpec { char type; };                     // 'f' float, 'i' int, 'e' enum
struct NodeSpec {
    const char* name;
    const char* vars;                              // one char per member variable, in wire order
    int lookups;                                   // node lookups
    int hybrids;                                   // hybrids
};
// index = node id on the wire (the order the reference registers its classes in)
const NodeSpec kNodes[] = {
    /* 0*/ {"Constant", "f", 0, 0},
    /* 1*/ {"White", "", 0, 0},
    /* 2*/ {"Checkerboard", "f", 0, 0},
    /* 3*/ {"SineWave", "f", 0, 0},
    /* 4*/ {"PositionOutput", "ffffffff", 0, 0},
    /* 5*/ {"DistanceToPoint", "effff", 0, 0},
    /* 6*/ {"Value", "", 0, 0},
    /* 7*/ {"Perlin", "", 0, 0},
    /* 8*/ {"Simplex", "", 0, 0},
    /* 9*/ {"OpenSimplex2", "", 0, 0},
    /*10*/ {"CellularValue", "ei", 0, 1},
    /*11*/ {"CellularDistance", "eiie", 0, 1},
    /*12*/ {"CellularLookup", "ef", 1, 1},
    /*13*/ {"FractalFBm", "if", 1, 2},
    /*14*/ {"FractalPingPong", "if", 1, 3},
    /*15*/ {"FractalRidged", "if", 1, 2},
    /*16*/ {"DomainWarpGradient", "f", 1, 1},
    /*17*/ {"DomainWarpFractalProgressive", "if", 1, 2},
    /*18*/ {"DomainWarpFractalIndependant", "if", 1, 2},
    /*19*/ {"DomainScale", "f", 1, 0},
    /*20*/ {"DomainOffset", "", 1, 4},
    /*21*/ {"DomainRotate", "fff", 1, 0},
    /*22*/ {"SeedOffset", "i", 1, 0},
    /*23*/ {"Remap", "ffff", 1, 0},
    /*24*/ {"ConvertRGBA8", "ff", 1, 0},
    /*25*/ {"Add", "", 1, 1},
    /*26*/ {"Subtract", "", 0, 2},
    /*27*/ {"Multiply", "", 1, 1},
    /*28*/ {"Divide", "", 0, 2},
    /*29*/ {"Min", "", 1, 1},
    /*30*/ {"Max", "", 1, 1},
    /*31*/ {"MinSmooth", "", 1, 2},
    /*32*/ {"MaxSmooth", "", 1, 2},
    /*33*/ {"Fade", "", 2, 1},
    /*34*/ {"Terrace", "ff", 1, 0},
    /*35*/ {"PowFloat", "", 0, 2},
    /*36*/ {"PowInt", "i", 1, 0},
    /*37*/ {"DomainAxisScale", "ffff", 1, 0},
    /*38*/ {"AddDimension", "", 1, 1},
    /*39*/ {"RemoveDimension", "e", 1, 0},
    /*40*/ {"GeneratorCache", "", 1, 0},
    /*41*/ {"OpenSimplex2S", "", 0, 0},
};
constexpr int kNodeCount = (int)(sizeof kNodes / sizeof kNodes[0]);

struct Hybrid { int node = -1; float value = 0.0f; };      // node = index into Tree::nodes, or -1 for the constant
struct Node {
    int id = -1;
    std::vector<int32_t> vars;                              // raw 4-byte values
    std::vector<int> lookups;
    std::vector<Hybrid> hybrids;
    float varf(size_t i) const { float f; std::memcpy(&f, &vars[i], 4); return f; }
};
struct Tree { std::vector<Node> nodes; int root = -1; };

std::vector<uint8_t> base64_decode(const char* s) {
    std::vector<uint8_t> out;
    uint32_t acc = 0;
    int bits = 0;
    for (; *s; ++s) {
        const char ch = *s;
        int v;
        if (ch >= 'A' && ch <= 'Z') v = ch - 'A';
        else if (ch >= 'a' && ch <= 'z') v = ch - 'a' + 26;
        else if (ch >= '0' && ch <= '9') v = ch - '0' + 52;
        else if (ch == '+') v = 62;
        else if (ch == '/') v = 63;
        else if (ch == '=') break;
        else throw std::runtime_error("encoded tree: character outside the base64 alphabet");
        acc = (acc << 6) | (uint32_t)v;
        bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
    }
    return out;
}

struct Reader {
    const std::vector<uint8_t>& d;
    size_t i = 0;
    template <class T> T get() {
        if (d.size() < i + sizeof(T)) throw std::runtime_error("encoded tree: truncated");
        T v;
        std::memcpy(&v, d.data() + i, sizeof(T));
        i += sizeof(T);
        return v;
    }
};

// returns the index of the parsed node; `refs` is the reference's post-order list that 0xFFFF entries point into
int parse_node(Reader& r, Tree& t, std::vector<int>& refs, int depth) {
    if (depth > 64) throw std::runtime_error("encoded tree: nesting deeper than 64");
    const uint16_t id = r.get<uint16_t>();
    if (id == 0xFFFF) {
        const uint16_t ref = r.get<uint16_t>();
        if (ref >= refs.size()) throw std::runtime_error("encoded tree: back reference to a node that does not exist yet");
        return refs[ref];
    }
    if (id >= kNodeCount) throw std::runtime_error("encoded tree: unknown node id " + std::to_string(id));
    const NodeSpec& sp = kNodes[id];
    Node n;
    n.id = id;
    for (const char* v = sp.vars; *v; ++v) n.vars.push_back(r.get<int32_t>());
    for (int k = 0; k < sp.lookups; ++k) n.lookups.push_back(parse_node(r, t, refs, depth + 1));
    for (int k = 0; k < sp.hybrids; ++k) {
        const uint8_t is_node = r.get<uint8_t>();
        if (is_node > 1) throw std::runtime_error("encoded tree: bad hybrid tag");
        Hybrid h;
        if (is_node) h.node = parse_node(r, t, refs, depth + 1);
        else h.value = r.get<float>();
        n.hybrids.push_back(h);
    }
    if (t.nodes.size() >= 1024) throw std::runtime_error("encoded tree: more than 1024 nodes");
    t.nodes.push_back(std::move(n));
    refs.push_back((int)t.nodes.size() - 1);
    return (int)t.nodes.size() - 1;
}

// ------------------------------------------------------------------------------------------------ compiler
struct Compiler {
    const Tree& t;
    std::vector<TreeInstr> prog;
    int n_values = 0, n_positions = 1;              // P[0] is the input position
    bool fuse = std::getenv("VOLRT_TREE_NO_FUSE") == nullptr;      // the unfused form stays reachable for tests

    int new_value() {
        if (n_values >= kTreeMaxValues) throw std::runtime_error("tree needs more than 64 value registers (octaves are unrolled)");
        return n_values++;
    }
    int new_position() {
        if (n_positions >= kTreeMaxPositions) throw std::runtime_error("tree needs more than 16 position registers");
        return n_positions++;
    }
    TreeInstr& push(TreeOp op, int dst, int pos, int seed_off) {
        if ((int)prog.size() >= kTreeMaxInstr) throw std::runtime_error("compiled tree is longer than 512 instructions");
        TreeInstr in;
        std::memset(&in, 0, sizeof in);
        in.op = op; in.dst = (uint8_t)dst; in.pos = (uint8_t)pos; in.seed_off = seed_off;
        prog.push_back(in);
        return prog.back();
    }
    int constant(float v) { const int r = new_value(); push(T_CONST, r, 0, 0).f[0] = v; return r; }
    int hybrid(const Hybrid& h, int pos, int seed_off) { return h.node >= 0 ? emit(h.node, pos, seed_off) : constant(h.value); }
    [[noreturn]] void unsupported(const Node& n, const char* why) {
        throw std::runtime_error(std::string("node ") + kNodes[n.id].name + ": " + why);
    }
    static float fractal_bounding(int octaves, float gain_const) {          // Fractal.h:27-38
        const float g = std::fabs(gain_const);
        float amp = g, sum = 1.0f;
        for (int i = 1; i < octaves; ++i) { sum += amp; amp *= g; }
        return 1.0f / sum;
    }

    // value of node `ni` at position register `pos` with seed offset `so` -> value register
    int emit(int ni, int pos, int so) {
        const Node& n = t.nodes[ni];
        switch (n.id) {
            case 0: return constant(n.varf(0));
            case 1: case 6: case 7: case 8: case 9: case 41: {
                static const TreeOp ops[10] = {T_CONST, T_WHITE, T_CONST, T_CONST, T_CONST, T_CONST, T_VALUE, T_PERLIN, T_SIMPLEX, T_OPENSIMPLEX2};
                const int r = new_value();
                push(n.id == 41 ? T_OPENSIMPLEX2S : ops[n.id], r, pos, so);
                return r;
            }
            case 2: {
                if (!(n.varf(0) != 0.0f)) unsupported(n, "size must not be 0");
                const int r = new_value();
                push(T_CHECKERBOARD, r, pos, so).f[0] = n.varf(0);
                return r;
            }
            case 3: { const int r = new_value(); push(T_SINE, r, pos, so).f[0] = 1.0f / n.varf(0); return r; }
            case 4: {
                const int r = new_value();
                TreeInstr& in = push(T_POSITION_OUTPUT, r, pos, so);
                for (int c = 0; c < 3; ++c) { in.f[c] = n.varf(c); in.f[3 + c] = n.varf(4 + c); }
                return r;
            }
            case 5: {
                const int r = new_value();
                TreeInstr& in = push(T_DISTANCE_TO_POINT, r, pos, so);
                in.i0 = n.vars[0];
                for (int c = 0; c < 3; ++c) in.f[c] = n.varf(1 + c);
                return r;
            }
            case 10: case 11: {
                const int jit = hybrid(n.hybrids[0], pos, so);
                const int r = new_value();
                TreeInstr& in = push(n.id == 10 ? T_CELL_VALUE : T_CELL_DISTANCE, r, pos, so);
                in.a = (uint8_t)jit;
                in.i0 = n.vars[0];
                auto clamp3 = [](int v) { return v < 0 ? 0 : v > 3 ? 3 : v; };       // Cellular.h:42,83-84
                if (n.id == 10) in.i1 = clamp3(n.vars[1]);
                else { in.i1 = clamp3(n.vars[1]); in.i2 = clamp3(n.vars[2]); in.i3 = n.vars[3]; }
                if (in.i0 < 0 || in.i0 > 4) unsupported(n, "distance function out of range");
                return r;
            }
            case 12: {
                const int jit = hybrid(n.hybrids[0], pos, so);
                const int p = new_position();
                TreeInstr& in = push(T_CELL_LOOKUP_POS, p, pos, so);
                in.a = (uint8_t)jit; in.i0 = n.vars[0]; in.f[0] = n.varf(1);
                return emit(n.lookups[0], p, so + 1);                                  // seed - (-1), Cellular.inl:580
            }
            case 13: case 14: case 15: return emit_fractal(n, pos, so);
            case 16: {
                const int amp = hybrid(n.hybrids[0], pos, so);
                const int p = new_position();
                push(T_POS_COPY, p, pos, 0);
                const int strength = new_value();
                TreeInstr& in = push(T_POS_WARP, p, pos, so);
                in.a = (uint8_t)amp; in.b = (uint8_t)strength; in.f[0] = n.varf(0);
                return emit(n.lookups[0], p, so);
            }
            case 17: case 18: return emit_warp_fractal(n, pos, so);
            case 19: { const int p = new_position(); push(T_POS_SCALE, p, pos, 0).f[0] = n.varf(0); return emit(n.lookups[0], p, so); }
            case 20: {
                int off[3];
                for (int c = 0; c < 3; ++c) off[c] = hybrid(n.hybrids[c], pos, so);
                const int p = new_position();
                TreeInstr& in = push(T_POS_OFFSET, p, pos, 0);
                in.a = (uint8_t)off[0]; in.b = (uint8_t)off[1]; in.c = (uint8_t)off[2];
                return emit(n.lookups[0], p, so);
            }
            case 21: {                                                                  // Modifiers.h:80-116
                const float yc = std::cos(n.varf(0)), ys = std::sin(n.varf(0)), pc = std::cos(n.varf(1)), ps = std::sin(n.varf(1));
                const float rc = std::cos(n.varf(2)), rs = std::sin(n.varf(2));
                const int p = new_position();
                TreeInstr& in = push(T_POS_ROTATE, p, pos, 0);
                in.f[0] = yc * pc; in.f[1] = yc * ps * rs - ys * rc; in.f[2] = yc * ps * rc + ys * rs;
                in.f[3] = ys * pc; in.f[4] = ys * ps * rs + yc * rc; in.f[5] = ys * ps * rc - yc * rs;
                in.f[6] = -ps;     in.f[7] = pc * rs;                in.f[8] = pc * rc;
                return emit(n.lookups[0], p, so);
            }
            case 22: return emit(n.lookups[0], pos, so + n.vars[0]);
            case 23: {
                const int src = emit(n.lookups[0], pos, so);
                const int r = new_value();
                TreeInstr& in = push(T_REMAP, r, pos, so);
                in.a = (uint8_t)src;
                in.f[0] = n.varf(0); in.f[1] = n.varf(1) - n.varf(0); in.f[2] = n.varf(3) - n.varf(2); in.f[3] = n.varf(2);
                return r;
            }
            case 24: {
                const int src = emit(n.lookups[0], pos, so);
                const int r = new_value();
                TreeInstr& in = push(T_CONVERT_RGBA8, r, pos, so);
                in.a = (uint8_t)src; in.f[0] = n.varf(0); in.f[1] = n.varf(1); in.f[2] = 255.0f / (n.varf(1) - n.varf(0));
                return r;
            }
            case 25: case 27: case 29: case 30: {                                       // LHS node, RHS hybrid
                const int l = emit(n.lookups[0], pos, so), rh = hybrid(n.hybrids[0], pos, so);
                const int r = new_value();
                TreeInstr& in = push(n.id == 25 ? T_ADD : n.id == 27 ? T_MUL : n.id == 29 ? T_MIN : T_MAX, r, pos, so);
                in.a = (uint8_t)l; in.b = (uint8_t)rh;
                return r;
            }
            case 26: case 28: case 35: {                                                // two hybrids
                const int l = hybrid(n.hybrids[0], pos, so), rh = hybrid(n.hybrids[1], pos, so);
                const int r = new_value();
                TreeInstr& in = push(n.id == 26 ? T_SUB : n.id == 28 ? T_DIV : T_POW_FLOAT, r, pos, so);
                in.a = (uint8_t)l; in.b = (uint8_t)rh;
                return r;
            }
            case 31: case 32: {
                const int l = emit(n.lookups[0], pos, so), rh = hybrid(n.hybrids[0], pos, so), sm = hybrid(n.hybrids[1], pos, so);
                const int r = new_value();
                TreeInstr& in = push(n.id == 31 ? T_MIN_SMOOTH : T_MAX_SMOOTH, r, pos, so);
                in.a = (uint8_t)l; in.b = (uint8_t)rh; in.c = (uint8_t)sm;
                return r;
            }
            case 33: {
                const int fade = hybrid(n.hybrids[0], pos, so);                          // Blends.inl: fade, then A, then B
                const int a = emit(n.lookups[0], pos, so), b = emit(n.lookups[1], pos, so);
                const int r = new_value();
                TreeInstr& in = push(T_FADE, r, pos, so);
                in.a = (uint8_t)a; in.b = (uint8_t)b; in.c = (uint8_t)fade;
                return r;
            }
            case 34: {
                const int src = emit(n.lookups[0], pos, so);
                const int r = new_value();
                TreeInstr& in = push(T_TERRACE, r, pos, so);
                in.a = (uint8_t)src;
                in.f[0] = n.varf(0); in.f[1] = n.varf(1); in.f[2] = n.varf(1) != 0.0f ? 1.0f + 1.0f / n.varf(1) : 0.0f; in.f[3] = 1.0f / n.varf(0);
                return r;
            }
            case 36: {
                const int src = emit(n.lookups[0], pos, so);
                const int r = new_value();
                TreeInstr& in = push(T_POW_INT, r, pos, so);
                in.a = (uint8_t)src; in.i0 = n.vars[0];
                return r;
            }
            case 37: {
                const int p = new_position();
                TreeInstr& in = push(T_POS_AXIS_SCALE, p, pos, 0);
                for (int c = 0; c < 3; ++c) in.f[c] = n.varf(c);
                return emit(n.lookups[0], p, so);
            }
            case 40: return emit(n.lookups[0], pos, so);                                 // GeneratorCache: same value
            case 38: case 39: unsupported(n, "changes the dimension count; libvolrt evaluates 3-D positions only");
            default: unsupported(n, "not built");
        }
    }

    int emit_fractal(const Node& n, int pos, int so) {
        const int octaves = n.vars[0];
        const float lac = n.varf(1);
        if (octaves < 1 || octaves > 16) unsupported(n, "octaves must be in 1..16");
        // the usual case — constant gain and weighted strength over Value / Perlin / Simplex / OpenSimplex2(S) — is one
        // instruction that loops over the octaves in registers (noise_tree_eval.cuh)
        const Node& src = t.nodes[n.lookups[0]];
        const bool leaf = (src.id >= 6 && src.id <= 9) || src.id == 41;
        const bool constants = n.hybrids[0].node < 0 && n.hybrids[1].node < 0 && (n.id != 14 || n.hybrids[2].node < 0);
        if (fuse && leaf && constants) {
            static const TreeOp gen[4] = {T_VALUE, T_PERLIN, T_SIMPLEX, T_OPENSIMPLEX2};
            const int r = new_value();
            TreeInstr& in = push(T_FRACTAL_LEAF, r, pos, so);
            in.i0 = src.id == 41 ? T_OPENSIMPLEX2S : gen[src.id - 6]; in.i1 = octaves; in.i2 = n.id == 13 ? 0 : n.id == 15 ? 1 : 2;
            in.f[0] = fractal_bounding(octaves, n.hybrids[0].value); in.f[1] = lac; in.f[2] = n.hybrids[0].value;
            in.f[3] = n.hybrids[1].value; in.f[4] = n.id == 14 ? n.hybrids[2].value : 0.0f;
            return r;
        }
        const int gain = hybrid(n.hybrids[0], pos, so), ws = hybrid(n.hybrids[1], pos, so);
        const int strength = n.id == 14 ? hybrid(n.hybrids[2], pos, so) : 0;
        // bounding uses the constant part of the gain (1 when the gain is a generator, Fractal.h:12,27-38)
        const float bounding = fractal_bounding(octaves, n.hybrids[0].node >= 0 ? 1.0f : n.hybrids[0].value);
        const TreeOp first = n.id == 13 ? T_FBM_FIRST : n.id == 15 ? T_RIDGED_FIRST : T_PINGPONG_FIRST;
        const TreeOp next = n.id == 13 ? T_FBM_NEXT : n.id == 15 ? T_RIDGED_NEXT : T_PINGPONG_NEXT;
        const int p = new_position();
        push(T_POS_COPY, p, pos, 0);
        int noise = emit(n.lookups[0], p, so);
        const int sum = new_value(), amp = new_value();
        {
            TreeInstr& in = push(first, sum, pos, so);
            in.a = (uint8_t)noise; in.b = (uint8_t)amp; in.f[0] = bounding; in.i0 = strength;
        }
        for (int i = 1; i < octaves; ++i) {
            push(T_POS_SCALE, p, p, 0).f[0] = lac;                                      // pos *= lacunarity, in place
            const int prev = noise;
            noise = emit(n.lookups[0], p, so + i);
            TreeInstr& in = push(next, sum, pos, so);
            in.a = (uint8_t)noise; in.b = (uint8_t)amp; in.c = (uint8_t)gain; in.d = (uint8_t)ws; in.i0 = strength; in.i1 = prev;
        }
        return sum;
    }

    int emit_warp_fractal(const Node& n, int pos, int so) {
        const int octaves = n.vars[0];
        const float lac = n.varf(1);
        if (octaves < 1 || octaves > 16) unsupported(n, "octaves must be in 1..16");
        const Node& w = t.nodes[n.lookups[0]];
        if (w.id != 16) unsupported(n, "the domain warp source must be a DomainWarpGradient node");
        const float bounding = fractal_bounding(octaves, n.hybrids[0].node >= 0 ? 1.0f : n.hybrids[0].value);
        // DomainWarpFractal.inl:15-35 — amp, weighted strength, gain are sampled at the unwarped position
        const int warp_amp = hybrid(w.hybrids[0], pos, so);
        const int amp = new_value();
        {
            const int b = constant(bounding);
            TreeInstr& in = push(T_MUL, amp, pos, so);
            in.a = (uint8_t)b; in.b = (uint8_t)warp_amp;
        }
        const int ws = hybrid(n.hybrids[1], pos, so), gain = hybrid(n.hybrids[0], pos, so);
        const int p = new_position();
        push(T_POS_COPY, p, pos, 0);
        const int strength = new_value();
        float freq = w.varf(0);
        const bool progressive = n.id == 17;
        for (int i = 0; i < octaves; ++i) {
            if (i) {
                freq *= lac;
                TreeInstr& in = push(T_WARP_AMP_NEXT, amp, pos, so);
                in.a = (uint8_t)strength; in.b = (uint8_t)ws; in.c = (uint8_t)gain;
            }
            TreeInstr& in = push(T_POS_WARP, p, progressive ? p : pos, so + i);       // noise position: warped so far | original
            in.a = (uint8_t)amp; in.b = (uint8_t)strength; in.f[0] = freq;
        }
        return emit(w.lookups[0], p, so);
    }
};

}  // namespace

// the min/max fold of noise_kernels.cu, on order-preserving bits
__device__ __forceinline__ unsigned int tree_ordered_bits(float f) {
    const unsigned int b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// Generator.inl:124-176 (GenUniformGrid3D): pos = float(start + idx) * frequency, x fastest
__global__ void __launch_bounds__(128) tree_grid_kernel(const TreeInstr* __restrict__ prog, int n_instr, int result_reg, int x0, int y0, int z0,
                                                        int nx, int ny, int nz, float frequency, int seed, float* __restrict__ out,
                                                        unsigned int* __restrict__ minmax_bits) {
    __shared__ GradLut lut;
    grad_lut_init(&lut, threadIdx.x);
    __syncthreads();
    const size_t total = (size_t)nx * ny * nz;
    float mn = CUDART_INF_F, mx = -CUDART_INF_F;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int ix = (int)(i % nx), iy = (int)((i / nx) % ny), iz = (int)(i / ((size_t)nx * ny));
        const float v = tree_eval(prog, n_instr, result_reg, &lut, seed, __fmul_rn((float)(x0 + ix), frequency), __fmul_rn((float)(y0 + iy), frequency),
                                  __fmul_rn((float)(z0 + iz), frequency));
        out[i] = v;
        mn = fminf(mn, v); mx = fmaxf(mx, v);
    }
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    // one pair of atomics per CTA: 65 k warps hammering the same two words cost more than a whole Perlin pass
    __shared__ float smn[4], smx[4];
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 4; ++w) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
        if (mn <= mx) {
            atomicMin(&minmax_bits[0], tree_ordered_bits(mn));
            atomicMax(&minmax_bits[1], tree_ordered_bits(mx));
        }
    }
}

cudaError_t launch_tree_grid(const TreeInstr* d_prog, int n_instr, int result_reg, int x0, int y0, int z0, int nx, int ny, int nz,
                             float frequency, int seed, float* out, unsigned int* minmax_bits, cudaStream_t s) {
    const size_t total = (size_t)nx * ny * nz;
    size_t blocks = (total + 127) / 128;
    const size_t cap = 148u * 9u * 4u;              // 4 waves of the 9 CTAs an SM holds at 56 registers; grid-stride beyond
    if (blocks > cap) blocks = cap;
    if (blocks == 0) return cudaSuccess;
    tree_grid_kernel<<<(unsigned)blocks, 128, 0, s>>>(d_prog, n_instr, result_reg, x0, y0, z0, nx, ny, nz, frequency, seed, out, minmax_bits);
    return cudaGetLastError();
}

}  // namespace volrt

using namespace volrt;

// ------------------------------------------------------------------------------------------------ C ABI (host-only half)
static void copy_error(char* err, size_t cap, const char* what) {
    if (err && cap) { std::strncpy(err, what, cap - 1); err[cap - 1] = 0; }
}

extern "C" vr_status vr_noise_tree_compile(const char* encoded, vr_noise_tree** out, char* err, size_t err_cap) {
    if (!encoded || !out) { copy_error(err, err_cap, "null argument"); return VR_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    try {
        const std::vector<uint8_t> bytes = base64_decode(encoded);
        Tree t;
        std::vector<int> refs;
        Reader r{bytes};
        t.root = parse_node(r, t, refs, 0);
        Compiler c{t};
        const int result = c.emit(t.root, 0, 0);
        std::unique_ptr<vr_noise_tree> h(new vr_noise_tree);
        h->prog = std::move(c.prog);
        h->result_reg = result;
        h->n_nodes = (int)t.nodes.size();
        h->n_values = c.n_values;
        h->n_positions = c.n_positions;
        h->encoded = encoded;
        *out = h.release();
        return VR_OK;
    } catch (const std::runtime_error& e) {
        copy_error(err, err_cap, e.what());
        const bool unsupported = std::strncmp(e.what(), "node ", 5) == 0 || std::strncmp(e.what(), "tree needs", 10) == 0 ||
                                 std::strncmp(e.what(), "compiled tree", 13) == 0;
        return unsupported ? VR_ERR_UNSUPPORTED : VR_ERR_INVALID_ARGUMENT;
    } catch (const std::bad_alloc&) {
        copy_error(err, err_cap, "out of host memory");
        return VR_ERR_OUT_OF_MEMORY;
    }
}

extern "C" void vr_noise_tree_free(vr_noise_tree* t) { delete t; }

extern "C" vr_status vr_noise_tree_info(const vr_noise_tree* t, uint32_t* nodes, uint32_t* instructions, uint32_t* value_regs, uint32_t* position_regs) {
    if (!t) return VR_ERR_INVALID_ARGUMENT;
    if (nodes) *nodes = (uint32_t)t->n_nodes;
    if (instructions) *instructions = (uint32_t)t->prog.size();
    if (value_regs) *value_regs = (uint32_t)t->n_values;
    if (position_regs) *position_regs = (uint32_t)t->n_positions;
    return VR_OK;
}
