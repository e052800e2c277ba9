// Marching-cubes mesh extraction for sm_100a (B200).
//
// Replaces the `measure.marching_cubes(tsdf_vol, level=0)` call of TSDFVolume.get_mesh / get_point_cloud
// (third_party/tsdf_fusion_python/fusion.py:317,338) and the in-tree post-processing after it (fusion.py:339-348:
// nearest-voxel colour gather, b/g/r unpack, voxel -> world transform).  The arithmetic contract is the one restated in
// oracle/mc_oracle.c (scikit-image itself is not available anywhere on the box: parity unpinned).
//
// Pipeline (all HBM-bound streaming / compaction, no tensor cores):
//   1. mc_sign_kernel    one pass over the TSDF plane (4 B/voxel): one bit per voxel (v > level) and min/max.  The bit
//                        volume is N/8 bytes (16 MB for 512^3) and stays L2-resident for everything that follows.
//   2. mc_count_kernel   one thread per group of 32 z-consecutive voxels: bit-parallel edge masks (which voxels own a
//                        vertex on their +axis0/+axis1/+axis2 edge or a cube-centre vertex) and the triangle count of
//                        the group's active cubes.  TSDF values are only touched for cubes with an ambiguous face.
//   3. scan kernels      exclusive prefix sums of the per-group (vertex, triangle) counts.
//   4. mc_emit_kernel    re-walks the active cubes and writes triangles; a vertex index is recovered from the owner
//                        group's prefix + popcounts of its masks, so no per-voxel index volume exists.  Vertices come out
//                        sorted by (owner voxel in scan order, axis) and triangles in cube scan order: deterministic.
//   5. mc_vertex_kernel  one thread per vertex: position, accumulated normal, colour, world transform.
//
// z-slab shards pass one halo plane below / above the slab (`halo_lo` / `halo_hi`); keys carry global voxel indices so
// per-slab meshes can be stitched exactly (hive_b200/sharded.py).
#include "hv_common.cuh"

#include <float.h>
#include <string.h>

#define HV_MC_TABLE static __device__ const
#include "hv_mc_tables.h"

namespace hv {

struct McGrid {
    const float* tsdf;
    const float* halo_lo;  // [X][Y] plane just below local z = 0, or null
    const float* halo_hi;  // [X][Y] plane just above local z = Z - 1, or null
    const float* color;
    const float* chalo_lo;
    const float* chalo_hi;
    int X, Y, Z, pitch;
    int hl;              // 1 if halo_lo is present: "zb" coordinates are local z + hl
    int Zb;              // Z + hl + hh: planes visible to this shard
    int G;               // groups of 32 zb per column
    int own_lo, own_hi;  // zb range of the voxels / cubes this shard owns
    long long zoff;      // global z of zb = 0
    long long Zg;        // global Z
    double level;
    // scratch
    unsigned int* bits;   // [X*Y*G]
    uint4* gmask;         // [X*Y*G] owner masks: axis0, axis1, axis2, centre
    uint2* gcount;        // [X*Y*G] (vertices, triangles)
    uint2* goff;          // [X*Y*G] exclusive prefix of gcount
    uint2* bsum;          // scratch of the scan
    unsigned int* active; // [n_groups] compacted ids of the groups that own a vertex or a triangle (any order)
    unsigned int* n_active;
    unsigned long long* totals;  // [0] vertices [1] triangles [2] min key [3] max key (ordered-int float)
    long long n_groups;
};

__device__ __forceinline__ float mc_fetch(const McGrid& g, const float* vol, const float* lo, const float* hi, int x,
                                          int y, int zb) {
    const int z = zb - g.hl;
    const size_t col = (size_t)x * g.Y + y;
    if (z < 0) return __ldg(lo + col);
    if (z >= g.Z) return __ldg(hi + col);
    return __ldg(vol + col * (size_t)g.pitch + z);
}

__device__ __forceinline__ int float_order_key(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}

__global__ void mc_init_kernel(unsigned long long* totals, unsigned int* n_active) {
    if (threadIdx.x == 0) {
        totals[0] = 0ull;
        totals[1] = 0ull;
        totals[2] = 0x7fffffffull;  // min key
        totals[3] = 0x80000000ull;  // max key
        *n_active = 0u;
    }
}

// ---- 1. sign bits + min/max ----------------------------------------------------------------------------------------
// A warp covers 128 consecutive zb of one column (each lane a float4 -> 4 sign bits, 8 lanes -> one 32-bit word) and
// strides over (column, segment) tiles; whole-volume shards (no low halo, 16-byte aligned columns) take the LDG.128 path.
struct McSignGeom {
    int segs;        // 128-zb segments per column
    int n_tiles;     // columns * segs
    int vec;         // 1: z == zb and columns are float4-aligned
};

// Sign nibble of four values + running min/max.
__device__ __forceinline__ unsigned int mc_sign4(const float (&v)[4], float level_f, bool level_exact, double level,
                                                 float& vmin, float& vmax) {
    unsigned int nib = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const bool pos = level_exact ? (v[k] > level_f) : ((double)v[k] > level);  // NaN (padding) -> 0
        nib |= pos ? (1u << k) : 0u;
    }
    vmin = fminf(vmin, fminf(fminf(v[0], v[1]), fminf(v[2], v[3])));  // fminf / fmaxf ignore NaN
    vmax = fmaxf(vmax, fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3])));
    return nib;
}

__device__ __forceinline__ void mc_store_word(const McGrid& g, unsigned int nib, int lane, size_t col, int seg) {
    unsigned int word = nib << (4 * (lane & 7));
    word |= __shfl_xor_sync(0xffffffffu, word, 1);
    word |= __shfl_xor_sync(0xffffffffu, word, 2);
    word |= __shfl_xor_sync(0xffffffffu, word, 4);
    const int gq = seg * 4 + (lane >> 3);
    if ((lane & 7) == 0 && gq < g.G) g.bits[col * g.G + gq] = word;
}

// A warp takes whole columns (grid-stride) and walks their 128-zb segments: no tile decode, and in the vector path two
// segments (2 x 512 B per warp) are in flight together.  r01: the tile-indexed version spent 78 warp-instructions per
// 128 voxels, a third of them in the integer division of the tile decode, and ran at 63 % issue / 55 % DRAM.
__global__ void __launch_bounds__(256) mc_sign_kernel(const McGrid g, const McSignGeom sg) {
    const int lane = threadIdx.x & 31;
    const int warp0 = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int n_warps = (int)((gridDim.x * blockDim.x) >> 5);
    float vmin = INFINITY, vmax = -INFINITY;
    const float level_f = (float)g.level;  // exact compare needs double only when the level is not a float
    const bool level_exact = (double)level_f == g.level;
    const int n_cols = g.X * g.Y;
    for (int col = warp0; col < n_cols; col += n_warps) {
        if (sg.vec) {
            const float* base = g.tsdf + (size_t)col * g.pitch + lane * 4;
            int seg = 0;
            for (; seg + 1 < sg.segs; seg += 2) {
                const int za = seg * 128 + lane * 4, zb = za + 128;
                float a[4] = {NAN, NAN, NAN, NAN}, b[4] = {NAN, NAN, NAN, NAN};
                const bool fa = za + 3 < g.Z, fb = zb + 3 < g.Z;
                if (fa) {
                    const float4 q = __ldcs(reinterpret_cast<const float4*>(base + seg * 128));
                    a[0] = q.x; a[1] = q.y; a[2] = q.z; a[3] = q.w;
                }
                if (fb) {
                    const float4 q = __ldcs(reinterpret_cast<const float4*>(base + seg * 128 + 128));
                    b[0] = q.x; b[1] = q.y; b[2] = q.z; b[3] = q.w;
                }
                if (!fa) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (za + k < g.Zb) a[k] = mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, col / g.Y, col % g.Y, za + k);
                }
                if (!fb) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (zb + k < g.Zb) b[k] = mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, col / g.Y, col % g.Y, zb + k);
                }
                mc_store_word(g, mc_sign4(a, level_f, level_exact, g.level, vmin, vmax), lane, (size_t)col, seg);
                mc_store_word(g, mc_sign4(b, level_f, level_exact, g.level, vmin, vmax), lane, (size_t)col, seg + 1);
            }
            if (seg < sg.segs) {
                const int za = seg * 128 + lane * 4;
                float a[4] = {NAN, NAN, NAN, NAN};
                if (za + 3 < g.Z) {
                    const float4 q = __ldcs(reinterpret_cast<const float4*>(base + seg * 128));
                    a[0] = q.x; a[1] = q.y; a[2] = q.z; a[3] = q.w;
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (za + k < g.Zb) a[k] = mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, col / g.Y, col % g.Y, za + k);
                }
                mc_store_word(g, mc_sign4(a, level_f, level_exact, g.level, vmin, vmax), lane, (size_t)col, seg);
            }
        } else {
            const int x = col / g.Y, y = col - x * g.Y;
            for (int seg = 0; seg < sg.segs; ++seg) {
                const int zb0 = seg * 128 + lane * 4;
                float v[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    v[k] = (zb0 + k < g.Zb) ? mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, x, y, zb0 + k) : NAN;
                mc_store_word(g, mc_sign4(v, level_f, level_exact, g.level, vmin, vmax), lane, (size_t)col, seg);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        vmin = fminf(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
        vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    }
    if (lane == 0) {
        if (vmin <= vmax) {  // at least one real value seen
            atomicMin(reinterpret_cast<int*>(g.totals + 2), float_order_key(vmin));
            atomicMax(reinterpret_cast<int*>(g.totals + 3), float_order_key(vmax));
        }
    }
}

// ---- cube helpers -----------------------------------------------------------------------------------------------------
struct CubeWords {
    unsigned int w[8];    // per corner: the word whose bit k is that corner of cube (x, y, 32*gq + k)
    unsigned int active;  // cubes of this group that exist, are owned and are crossed by the surface
    unsigned int own;     // owned voxels of this group
    unsigned int b00, b01, b10, u00;
};

__device__ __forceinline__ unsigned int range_mask(int lo, int hi, int base) {
    // bits k with lo <= base + k < hi
    const int a = max(lo - base, 0), b = min(hi - base, 32);
    if (a >= b) return 0u;
    const unsigned int upto_b = (b >= 32) ? 0xffffffffu : ((1u << b) - 1u);
    return upto_b & ~((1u << a) - 1u);
}

__device__ __forceinline__ void load_pair(const McGrid& g, int x, int y, int gq, bool in, unsigned int& b, unsigned int& u) {
    b = 0u;
    u = 0u;
    if (!in) return;
    const size_t gi = ((size_t)x * g.Y + y) * g.G + gq;
    b = g.bits[gi];
    const unsigned int nxt = (gq + 1 < g.G) ? g.bits[gi + 1] : 0u;
    u = (b >> 1) | (nxt << 31);
}

__device__ __forceinline__ CubeWords load_cube_words(const McGrid& g, int x, int y, int gq) {
    CubeWords c;
    const bool x1 = x + 1 < g.X, y1 = y + 1 < g.Y;
    unsigned int b00, b01, b10, b11, u00, u01, u10, u11;
    load_pair(g, x, y, gq, true, b00, u00);
    load_pair(g, x, y + 1, gq, y1, b01, u01);
    load_pair(g, x + 1, y, gq, x1, b10, u10);
    load_pair(g, x + 1, y + 1, gq, x1 && y1, b11, u11);
    // corner i offsets (axis0, axis1, axis2): see kMcCornerOff
    c.w[0] = b00; c.w[1] = u00; c.w[2] = u01; c.w[3] = b01;
    c.w[4] = b10; c.w[5] = u10; c.w[6] = u11; c.w[7] = b11;
    c.b00 = b00; c.b01 = b01; c.b10 = b10; c.u00 = u00;
    const int base = gq * 32;
    c.own = range_mask(g.own_lo, g.own_hi, base);
    const unsigned int all8 = b00 & b01 & b10 & b11 & u00 & u01 & u10 & u11;
    const unsigned int any8 = b00 | b01 | b10 | b11 | u00 | u01 | u10 | u11;
    c.active = (x1 && y1) ? (any8 & ~all8 & c.own & range_mask(0, g.Zb - 1, base)) : 0u;
    return c;
}

__device__ __forceinline__ int cube_index(const CubeWords& c, int k) {
    int idx = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) idx |= (int)((c.w[i] >> k) & 1u) << i;
    return idx;
}

__device__ __forceinline__ void load_cube_values(const McGrid& g, int x, int y, int zb, double v[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float f = mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, x + kMcCornerOff[3 * i], y + kMcCornerOff[3 * i + 1],
                                 zb + kMcCornerOff[3 * i + 2]);
        v[i] = (double)f - g.level;
    }
}

// Lewiner's test_interior (same statement as oracle/mc_oracle.c:interior_says_tunnel): slab form
// for cases 4 / 10, reference-edge form for 6 / 7 / 12 / 13.5.  `word`: kMcInterior descriptor.  Returns true when the
// tube tiling is to be used.  Plain double arithmetic; the TU is compiled with -fmad=false, so no contraction.
__device__ __noinline__ bool cube_interior_tunnel(unsigned int word, const double v[8]) {
    const bool inverted = (word >> 2) & 1u, tunnel_if_test = (word >> 3) & 1u;
    const bool ret_pos = !inverted, ret_neg = inverted;  // Lewiner's "return s > 0" / "return s < 0"
    int c[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = (int)((word >> (4 + 3 * i)) & 7u);
    double a0 = 0, a1 = 0, b0 = 0, b1 = 0, c0 = 0, c1 = 0, d0 = 0, d1 = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {  // select by compare: keeps v[] in registers
        const double vi = v[i];
        if (c[0] == i) a0 = vi;
        if (c[1] == i) a1 = vi;
        if (c[2] == i) b0 = vi;
        if (c[3] == i) b1 = vi;
        if (c[4] == i) c0 = vi;
        if (c[5] == i) c1 = vi;
        if (c[6] == i) d0 = vi;
        if (c[7] == i) d1 = vi;
    }
    double t, At, Bt, Ct, Dt;
    bool result;
    if ((word & 3u) == 1u) {
        const double dA = a1 - a0, dB = b1 - b0, dC = c1 - c0, dD = d1 - d0;
        const double qa = dA * dC - dB * dD;
        const double qb = c0 * dA + a0 * dC - d0 * dB - b0 * dD;
        t = -qb / (2 * qa);
        if (t < 0 || t > 1) return tunnel_if_test ? ret_pos : !ret_pos;
        At = a0 + dA * t;
        Bt = b0 + dB * t;
        Ct = c0 + dC * t;
        Dt = d0 + dD * t;
    } else {
        t = a0 / (a0 - a1);
        At = 0;
        Bt = b0 + (b1 - b0) * t;
        Ct = c0 + (c1 - c0) * t;
        Dt = d0 + (d1 - d0) * t;
    }
    const int test = (At >= 0 ? 1 : 0) | (Bt >= 0 ? 2 : 0) | (Ct >= 0 ? 4 : 0) | (Dt >= 0 ? 8 : 0);
    if (test == 5) result = (At * Ct - Bt * Dt < (double)FLT_EPSILON) ? ret_pos : ret_neg;
    else if (test == 10) result = (At * Ct - Bt * Dt >= (double)FLT_EPSILON) ? ret_pos : ret_neg;
    else if (test == 7 || test == 11 || test >= 13) result = ret_neg;
    else result = ret_pos;
    return tunnel_if_test ? result : !result;
}

// Tiling of one cube: each ambiguous face is resolved with Lewiner's face test, p =
// "the positive corners are joined" = (pos - neg > -FLT_EPSILON); where his decision tree evaluates the interior
// test, a tunnel selects the tube tiling stored after the 2^namb plain ones.
__device__ __forceinline__ int cube_variant(int idx, const double v[8]) {
    int p_bits = 0;
    const int namb = kMcAmbCount[idx];
    const unsigned int faces = kMcAmbFaces[idx];
    for (int k = 0; k < namb; ++k) {
        const int f = (faces >> (3 * k)) & 7;
        int a = kMcFaceCorner[4 * f], b = kMcFaceCorner[4 * f + 1], c = kMcFaceCorner[4 * f + 2], d = kMcFaceCorner[4 * f + 3];
        if (!((idx >> a) & 1)) {
            const int t = a; a = b; b = c; c = d; d = t;
        }
        const double pos = __dmul_rn(v[a], v[c]), neg = __dmul_rn(v[b], v[d]);
        if (__dsub_rn(pos, neg) > -(double)FLT_EPSILON) p_bits |= 1 << k;
    }
    int variant = kMcVariantBase[idx] + p_bits;
    const int ib = kMcInteriorBase[idx];
    if (ib != 255) {
        const unsigned long long mask = kMcInteriorMask[idx];
        if ((mask >> p_bits) & 1ull) {
            const int rank = __popcll(mask & ((1ull << p_bits) - 1ull));
            if (cube_interior_tunnel(kMcInterior[ib + rank], v)) variant = kMcVariantBase[idx] + (1 << namb) + rank;
        }
    }
    return variant;
}

// The asymptotic decider needs the eight corner values in float64 (16 registers): out of line, so that the common
// unambiguous path of the count / emit kernels keeps a small register footprint (count: 64 -> 40 registers, 4 -> 6 CTAs/SM).
// (scalars by value: a reference to the kernel-parameter struct would make every thread spill the struct to its stack)
__device__ __noinline__ int cube_variant_ambiguous(const float* tsdf, const float* halo_lo, const float* halo_hi, int Y,
                                                   int pitch, int Z, int hl, double level, int idx, int x, int y, int zb) {
    double v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int cx = x + kMcCornerOff[3 * i], cy = y + kMcCornerOff[3 * i + 1], cz = zb + kMcCornerOff[3 * i + 2] - hl;
        const size_t col = (size_t)cx * Y + cy;
        const float f = cz < 0 ? __ldg(halo_lo + col) : (cz >= Z ? __ldg(halo_hi + col) : __ldg(tsdf + col * (size_t)pitch + cz));
        v[i] = (double)f - level;
    }
    return cube_variant(idx, v);
}

// count kernel: one thread per group, nearly all of them with nothing ambiguous -> occupancy matters, call out of line
__device__ __forceinline__ int cube_variant_at_ool(const McGrid& g, int idx, int x, int y, int zb) {
    if (kMcAmbCount[idx] == 0 && kMcInteriorBase[idx] == 255) return kMcVariantBase[idx];  // (case 4: no face, one interior test)
    return cube_variant_ambiguous(g.tsdf, g.halo_lo, g.halo_hi, g.Y, g.pitch, g.Z, g.hl, g.level, idx, x, y, zb);
}

// emit kernel: works on surface cubes only; inlined (the call cost 30 us at 512^3 there)
__device__ __forceinline__ int cube_variant_at(const McGrid& g, int idx, int x, int y, int zb) {
    if (kMcAmbCount[idx] == 0 && kMcInteriorBase[idx] == 255) return kMcVariantBase[idx];
    double v[8];
    load_cube_values(g, x, y, zb, v);
    return cube_variant(idx, v);
}

// ---- 2. per-group owner masks and counts ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) mc_count_kernel(const McGrid g) {
    const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gi >= g.n_groups) return;
    int x, y, gq;
    if (gi < 0x100000000ll) {  // 32-bit decode (a 64-bit division is ~100 instructions; this kernel has 4.2 M threads at 512^3)
        const unsigned int g32 = (unsigned int)gi, col = g32 / (unsigned int)g.G;
        gq = (int)(g32 - col * (unsigned int)g.G);
        x = (int)(col / (unsigned int)g.Y);
        y = (int)(col - (unsigned int)x * (unsigned int)g.Y);
    } else {
        const long long col = gi / g.G;
        gq = (int)(gi - col * g.G);
        x = (int)(col / g.Y);
        y = (int)(col - (long long)x * g.Y);
    }
    const CubeWords c = load_cube_words(g, x, y, gq);
    const int base = gq * 32;
    uint4 m;
    m.x = (x + 1 < g.X) ? ((c.b00 ^ c.b10) & c.own) : 0u;
    m.y = (y + 1 < g.Y) ? ((c.b00 ^ c.b01) & c.own) : 0u;
    m.z = (c.b00 ^ c.u00) & c.own & range_mask(0, g.Zb - 1, base);
    m.w = 0u;
    unsigned int ntri = 0;
    for (unsigned int rest = c.active; rest; rest &= rest - 1) {
        const int k = __ffs(rest) - 1;
        const int idx = cube_index(c, k);
        const int variant = cube_variant_at_ool(g, idx, x, y, base + k);
        ntri += kMcVarNtri[variant];
        if ((kMcVarMult[variant] >> 48) & 15ull) m.w |= 1u << k;
    }
    g.gmask[gi] = m;
    const unsigned int nv = __popc(m.x) + __popc(m.y) + __popc(m.z) + __popc(m.w);
    g.gcount[gi] = make_uint2(nv, ntri);
    // warp-aggregated append to the work list of mc_emit_kernel
    const bool has_work = (nv | ntri) != 0u;
    const unsigned int vote = __ballot_sync(__activemask(), has_work);
    if (has_work) {
        const int lane = threadIdx.x & 31;
        const int leader = __ffs(vote) - 1;
        unsigned int base0 = 0;
        if (lane == leader) base0 = atomicAdd(g.n_active, (unsigned int)__popc(vote));
        base0 = __shfl_sync(vote, base0, leader);
        g.active[base0 + __popc(vote & ((1u << lane) - 1u))] = (unsigned int)gi;
    }
}

// ---- 4. emit triangles and vertex keys ------------------------------------------------------------------------------
struct McEmit {
    int32_t* faces;                  // [T][3] local vertex indices, or null
    unsigned long long* face_keys;   // [T][3] global vertex keys, or null
    unsigned long long* vert_keys;   // [V]
    unsigned long long* face_cubes;  // [T] global linear index of the cube that produced the face, or null
};

__device__ __forceinline__ unsigned long long vertex_key(const McGrid& g, int x, int y, int zb, int axis) {
    const unsigned long long lin = ((unsigned long long)x * g.Y + y) * (unsigned long long)g.Zg + (unsigned long long)(zb + g.zoff);
    return lin * 4ull + (unsigned long long)axis;
}

__device__ __forceinline__ int vertex_index(const McGrid& g, int x, int y, int zb, int axis) {
    const size_t og = ((size_t)x * g.Y + y) * g.G + (zb >> 5);
    const int bit = zb & 31;
    const uint4 m = g.gmask[og];
    const unsigned int low = (1u << bit) - 1u;
    int r = (int)g.goff[og].x + __popc(m.x & low) + __popc(m.y & low) + __popc(m.z & low) + __popc(m.w & low);
    if (axis > 0) r += (m.x >> bit) & 1u;
    if (axis > 1) r += (m.y >> bit) & 1u;
    if (axis > 2) r += (m.z >> bit) & 1u;
    return r;
}

// Writes triangle `j` of the cube (x, y, zb) with tiling `variant` as face `ti`: three independent look-ups (edge ->
// owning voxel -> that voxel's group mask and offset), unrolled so that all of their loads are in flight together.
__device__ __forceinline__ void emit_triangle(const McGrid& g, const McEmit& out, int x, int y, int zb, int variant, int j,
                                              size_t ti) {
    const uint8_t* tri = kMcTris + 3 * ((size_t)kMcVarTriOff[variant] + (size_t)j);
    int ox[3], oy[3], oz[3], axis[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int e = tri[c];
        ox[c] = oy[c] = oz[c] = 0;
        axis[c] = 3;
        if (e < 12) {
            ox[c] = kMcEdgeOwner[4 * e], oy[c] = kMcEdgeOwner[4 * e + 1], oz[c] = kMcEdgeOwner[4 * e + 2];
            axis[c] = kMcEdgeOwner[4 * e + 3];
        }
    }
    if (out.face_keys) {
#pragma unroll
        for (int c = 0; c < 3; ++c) out.face_keys[3 * ti + c] = vertex_key(g, x + ox[c], y + oy[c], zb + oz[c], axis[c]);
        if (out.face_cubes) out.face_cubes[ti] = vertex_key(g, x, y, zb, 0) >> 2;
    } else {
        int v[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) v[c] = vertex_index(g, x + ox[c], y + oy[c], zb + oz[c], axis[c]);
#pragma unroll
        for (int c = 0; c < 3; ++c) out.faces[3 * ti + c] = v[c];
    }
}

// A block takes kEmitThreads entries of the active-group list.  Phase 1, one thread per group: the keys of the vertices
// the group owns, and one 8-byte record per *triangle* of its surface cubes (phase-1 slot, voxel, tiling, triangle number,
// face index) pushed into a shared-memory list.  Phase 2, one thread per record: one triangle = three independent
// vertex look-ups.  (r01 history: one warp per group / one lane per cube left 75-90 % of the lanes idle in the triangle
// loop; one thread per cube record still walked up to 12 triangles x 3 dependent look-ups serially -- 47 % of the
// kernel's stall samples were long-scoreboard waits inside that loop.)
constexpr int kEmitThreads = 128;
constexpr int kEmitCap = 2048;  // triangle records per round (16 KB); triangles beyond it are written by phase 1 directly
static_assert(HV_MC_NUM_VARIANTS <= 1024 && kEmitThreads <= 128, "record packing: 7 bits slot, 5 voxel, 10 tiling, 4+ triangle");

__global__ void __launch_bounds__(kEmitThreads) mc_emit_kernel(const McGrid g, const McEmit out) {
    __shared__ uint2 s_tri[kEmitCap];
    __shared__ int s_x[kEmitThreads], s_y[kEmitThreads], s_z0[kEmitThreads];
    __shared__ unsigned int s_n;
    const unsigned int n_active = *g.n_active;
    for (unsigned int base = blockIdx.x * kEmitThreads; base < n_active; base += gridDim.x * kEmitThreads) {
        if (threadIdx.x == 0) s_n = 0u;
        __syncthreads();
        const unsigned int a = base + threadIdx.x;
        if (a < n_active) {
            const unsigned int gi = g.active[a];
            const uint2 cnt = g.gcount[gi];
            const uint2 off = g.goff[gi];
            const unsigned int col = gi / (unsigned int)g.G;
            const int gq = (int)(gi - col * (unsigned int)g.G);
            const int x = (int)(col / (unsigned int)g.Y), y = (int)(col - (unsigned int)x * (unsigned int)g.Y);
            s_x[threadIdx.x] = x;
            s_y[threadIdx.x] = y;
            s_z0[threadIdx.x] = gq * 32;
            if (cnt.x) {  // vertex keys, order inside the group: (voxel, axis)
                const uint4 m = g.gmask[gi];
                size_t vi = off.x;
                for (unsigned int rest = m.x | m.y | m.z | m.w; rest; rest &= rest - 1) {
                    const int k = __ffs(rest) - 1;
                    const int zb = gq * 32 + k;
                    if ((m.x >> k) & 1u) out.vert_keys[vi++] = vertex_key(g, x, y, zb, 0);
                    if ((m.y >> k) & 1u) out.vert_keys[vi++] = vertex_key(g, x, y, zb, 1);
                    if ((m.z >> k) & 1u) out.vert_keys[vi++] = vertex_key(g, x, y, zb, 2);
                    if ((m.w >> k) & 1u) out.vert_keys[vi++] = vertex_key(g, x, y, zb, 3);
                }
            }
            if (cnt.y) {
                const CubeWords c = load_cube_words(g, x, y, gq);
                unsigned int toff = off.y;
                for (unsigned int rest = c.active; rest; rest &= rest - 1) {
                    const int k = __ffs(rest) - 1;
                    const int variant = cube_variant_at(g, cube_index(c, k), x, y, gq * 32 + k);
                    const unsigned int ntri = kMcVarNtri[variant];
                    if (ntri == 0u) continue;
                    const unsigned int slot = atomicAdd(&s_n, ntri);
                    const unsigned int head = threadIdx.x | ((unsigned int)k << 7) | ((unsigned int)variant << 12);
                    for (unsigned int j = 0; j < ntri; ++j) {
                        if (slot + j < (unsigned int)kEmitCap)
                            s_tri[slot + j] = make_uint2(head | (j << 22), toff + j);
                        else
                            emit_triangle(g, out, x, y, gq * 32 + k, variant, (int)j, toff + j);
                    }
                    toff += ntri;
                }
            }
        }
        __syncthreads();
        const unsigned int n = min(s_n, (unsigned int)kEmitCap);
        for (unsigned int i = threadIdx.x; i < n; i += kEmitThreads) {
            const uint2 r = s_tri[i];
            const unsigned int t = r.x & 127u;
            emit_triangle(g, out, s_x[t], s_y[t], s_z0[t] + (int)((r.x >> 7) & 31u), (int)((r.x >> 12) & 1023u),
                          (int)(r.x >> 22), r.y);
        }
        __syncthreads();
    }
}

// ---- 5. vertices: position, normal, colour ----------------------------------------------------------------------------
struct McVerts {
    const unsigned long long* vert_keys;
    long long n_verts;
    float* verts_world;  // [V][3]
    float* verts_voxel;  // [V][3] or null
    float* normals;      // [V][3]
    uint8_t* colors;     // [V][3] r, g, b
    float voxel_size;
    float origin[3];
};

// Adds `mult` times this cube's interpolated down-hill gradient at edge `e` (12 = centre) to the float32 accumulator.
__device__ __forceinline__ void add_cube_gradient(const double v[8], int e, int mult, float acc[3]) {
    if (mult == 0) return;
    // weights 1 / (FLT_EPSILON + |v|): only the edge's two end corners are needed for e < 12 (all 8 for the centre);
    // computing all eight per cube cost 32 double divisions per vertex (r01 ncu: mc_vertex_kernel 100 us)
    double w[8];
    int lo = 0, hi = 0, clo = 0, chi = 0;
    if (e < 12) {
        lo = kMcEdgeCorner[2 * e], hi = kMcEdgeCorner[2 * e + 1];
        clo = kMcCornerOff[3 * lo] * 4 + kMcCornerOff[3 * lo + 1] * 2 + kMcCornerOff[3 * lo + 2];
        chi = kMcCornerOff[3 * hi] * 4 + kMcCornerOff[3 * hi + 1] * 2 + kMcCornerOff[3 * hi + 2];
        w[0] = 1.0 / ((double)FLT_EPSILON + fabs(v[lo]));
        w[1] = 1.0 / ((double)FLT_EPSILON + fabs(v[hi]));
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = 1.0 / ((double)FLT_EPSILON + fabs(v[i]));
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int bitk = 4 >> k;  // offset bit of axis k in the d0*4 + d1*2 + d2 code
        double gk;
        if (e < 12) {
            const double g_lo = v[kMcCornerAt[clo & ~bitk]] - v[kMcCornerAt[clo | bitk]];
            const double g_hi = v[kMcCornerAt[chi & ~bitk]] - v[kMcCornerAt[chi | bitk]];
            gk = __dadd_rn(__dmul_rn(g_lo, w[0]), __dmul_rn(g_hi, w[1]));
        } else {
            gk = 0.0;
            for (int i = 0; i < 8; ++i) {
                const int ci = kMcCornerOff[3 * i] * 4 + kMcCornerOff[3 * i + 1] * 2 + kMcCornerOff[3 * i + 2];
                const double gi = v[kMcCornerAt[ci & ~bitk]] - v[kMcCornerAt[ci | bitk]];
                gk = __dadd_rn(gk, __dmul_rn(w[i], gi));
            }
        }
        const float gf = (float)gk;
        for (int j = 0; j < mult; ++j) acc[k] = __fadd_rn(acc[k], gf);
    }
}

__global__ void __launch_bounds__(128, 8) mc_vertex_kernel(const McGrid g, const McVerts out) {
    const long long vi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (vi >= out.n_verts) return;
    const unsigned long long key = out.vert_keys[vi];
    const int axis = (int)(key & 3ull);
    const unsigned long long lin = key >> 2;
    // key -> (x, y, zg): 32-bit divisions whenever the linear index fits (volumes below 2^32 voxels); the four 64-bit
    // divisions this replaces were ~400 instructions per vertex
    long long zg;
    int x, y;
    if (lin < 0x100000000ull) {
        const unsigned int l32 = (unsigned int)lin, c32 = l32 / (unsigned int)g.Zg;
        zg = (long long)(l32 - c32 * (unsigned int)g.Zg);
        x = (int)(c32 / (unsigned int)g.Y);
        y = (int)(c32 - (unsigned int)x * (unsigned int)g.Y);
    } else {
        const unsigned long long col = lin / (unsigned long long)g.Zg;
        zg = (long long)(lin - col * (unsigned long long)g.Zg);
        x = (int)(col / (unsigned long long)g.Y);
        y = (int)(col - (unsigned long long)x * (unsigned long long)g.Y);
    }
    const int zb = (int)(zg - g.zoff);
    const int p[3] = {x, y, zb};
    const int dim[3] = {g.X, g.Y, g.Zb};

    double pos[3] = {(double)x, (double)y, (double)zg};
    float acc[3] = {0.0f, 0.0f, 0.0f};
    double v[8];
    if (axis == 3) {
        load_cube_values(g, x, y, zb, v);
        int idx = 0;
        for (int i = 0; i < 8; ++i) idx |= (v[i] > 0.0 ? 1 : 0) << i;
        const int variant = cube_variant(idx, v);
        double f[3] = {0.0, 0.0, 0.0}, ff = 0.0;
        for (int i = 0; i < 8; ++i) {
            const double w = 1.0 / ((double)FLT_EPSILON + fabs(v[i]));
            for (int k = 0; k < 3; ++k) f[k] = __dadd_rn(f[k], __dmul_rn((double)kMcCornerOff[3 * i + k], w));
            ff = __dadd_rn(ff, w);
        }
        for (int k = 0; k < 3; ++k) pos[k] = __dadd_rn(pos[k], f[k] / ff);
        add_cube_gradient(v, 12, (int)((kMcVarMult[variant] >> 48) & 15ull), acc);
    } else {
        int q[3] = {x, y, zb};
        q[axis] += 1;
        const double v_lo = (double)mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, x, y, zb) - g.level;
        const double v_hi = (double)mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, q[0], q[1], q[2]) - g.level;
        const double w_lo = 1.0 / ((double)FLT_EPSILON + fabs(v_lo)), w_hi = 1.0 / ((double)FLT_EPSILON + fabs(v_hi));
        pos[axis] = __dadd_rn(pos[axis], w_hi / __dadd_rn(w_lo, w_hi));
        // the (up to four) cubes around the edge, in scan order (axis0 outermost): larger offsets first
        const int b = (axis == 0) ? 1 : 0, c = (axis == 2) ? 1 : 2;
        for (int db = 1; db >= 0; --db) {
            for (int dc = 1; dc >= 0; --dc) {
                int base[3] = {p[0], p[1], p[2]};
                base[b] -= db;
                base[c] -= dc;
                if (base[b] < 0 || base[c] < 0 || base[b] + 1 >= dim[b] || base[c] + 1 >= dim[c]) continue;
                int d[3] = {0, 0, 0};
                d[b] = db;
                d[c] = dc;
                const int e = kMcEdgeAt[axis * 8 + d[0] * 4 + d[1] * 2 + d[2]];
                load_cube_values(g, base[0], base[1], base[2], v);
                int idx = 0;
                for (int i = 0; i < 8; ++i) idx |= (v[i] > 0.0 ? 1 : 0) << i;
                const int variant = cube_variant(idx, v);
                add_cube_gradient(v, e, (int)((kMcVarMult[variant] >> (4 * e)) & 15ull), acc);
            }
        }
    }
    const float pv[3] = {(float)pos[0], (float)pos[1], (float)pos[2]};
    const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn((double)acc[0], (double)acc[0]), __dmul_rn((double)acc[1], (double)acc[1])),
                                      __dmul_rn((double)acc[2], (double)acc[2])));
    if (len > 0.0)
        for (int k = 0; k < 3; ++k) acc[k] = (float)((double)acc[k] / len);
    // colour of the nearest voxel (fusion.py:339,343-348), half-to-even like np.round
    const int i0 = (int)rintf(pv[0]), i1 = (int)rintf(pv[1]);
    const int i2 = (int)((long long)rintf(pv[2]) - g.zoff);
    const unsigned int packed = (unsigned int)mc_fetch(g, g.color, g.chalo_lo, g.chalo_hi, i0, i1, i2);
    for (int k = 0; k < 3; ++k) {
        out.verts_world[3 * vi + k] = __fadd_rn(__fmul_rn(pv[k], out.voxel_size), out.origin[k]);  // fusion.py:340
        if (out.verts_voxel) out.verts_voxel[3 * vi + k] = pv[k];
        out.normals[3 * vi + k] = acc[k];
    }
    out.colors[3 * vi + 0] = (uint8_t)(packed & 255u);
    out.colors[3 * vi + 1] = (uint8_t)((packed >> 8) & 255u);
    out.colors[3 * vi + 2] = (uint8_t)((packed >> 16) & 255u);
}

}  // namespace hv

// ---- C ABI ---------------------------------------------------------------------------------------------------------------
namespace {

struct McLayout {
    long long n_groups;
    int G, Zb, n_scan_blocks;
    size_t off_bits, off_gmask, off_gcount, off_goff, off_bsum, off_active, off_nactive, total;
};

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

int mc_layout(const int64_t* h_dims, bool lo, bool hi, McLayout* L) {
    HV_REQUIRE(h_dims, "marching cubes: null dims");
    const int64_t X = h_dims[0], Y = h_dims[1], Z = h_dims[2];
    HV_REQUIRE(X >= 1 && Y >= 1 && Z >= 1 && X < (1ll << 30) && Y < (1ll << 30) && Z < (1ll << 30),
               "marching cubes: bad dims %lld x %lld x %lld", (long long)X, (long long)Y, (long long)Z);
    L->Zb = (int)Z + (lo ? 1 : 0) + (hi ? 1 : 0);
    L->G = (L->Zb + 31) / 32;
    L->n_groups = (long long)X * Y * L->G;
    HV_REQUIRE(L->n_groups < (1ll << 32) && L->Zb < (1 << 22), "marching cubes: volume too large");
    L->n_scan_blocks = (int)hv::scan_scratch_pairs(L->n_groups);
    size_t o = 0;
    L->off_gmask = o;  // 16-byte elements first
    o = align_up(o + (size_t)L->n_groups * sizeof(uint4));
    L->off_gcount = o;
    o = align_up(o + (size_t)L->n_groups * sizeof(uint2));
    L->off_goff = o;
    o = align_up(o + (size_t)L->n_groups * sizeof(uint2));
    L->off_bits = o;
    o = align_up(o + (size_t)L->n_groups * sizeof(unsigned int));
    L->off_bsum = o;
    o = align_up(o + (size_t)L->n_scan_blocks * sizeof(uint2));
    L->off_active = o;
    o = align_up(o + (size_t)L->n_groups * sizeof(unsigned int));
    L->off_nactive = o;
    o = align_up(o + sizeof(unsigned int));
    L->total = o;
    return HV_OK;
}

int mc_grid(hv::McGrid* g, const float* tsdf, const float* halo_lo, const float* halo_hi, const int64_t* h_dims,
            int64_t pitch, double level, void* scratch, int64_t scratch_bytes, McLayout* L) {
    int rc = mc_layout(h_dims, halo_lo != nullptr, halo_hi != nullptr, L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(tsdf && scratch, "marching cubes: null pointer");
    HV_REQUIRE(pitch >= h_dims[2], "marching cubes: pitch %lld < Z", (long long)pitch);
    HV_REQUIRE(((uintptr_t)scratch & 255) == 0, "marching cubes: scratch must be 256-byte aligned");
    if ((size_t)scratch_bytes < L->total) {
        hv::set_error("marching cubes: scratch of %lld bytes, need %lld", (long long)scratch_bytes, (long long)L->total);
        return HV_E_CAPACITY;
    }
    char* s = static_cast<char*>(scratch);
    g->tsdf = tsdf;
    g->halo_lo = halo_lo;
    g->halo_hi = halo_hi;
    g->color = g->chalo_lo = g->chalo_hi = nullptr;
    g->X = (int)h_dims[0];
    g->Y = (int)h_dims[1];
    g->Z = (int)h_dims[2];
    g->pitch = (int)pitch;
    g->hl = halo_lo ? 1 : 0;
    g->Zb = L->Zb;
    g->G = L->G;
    g->own_lo = g->hl;
    g->own_hi = g->hl + g->Z;
    g->zoff = -(long long)g->hl;
    g->Zg = g->Z;
    g->level = level;
    g->bits = reinterpret_cast<unsigned int*>(s + L->off_bits);
    g->gmask = reinterpret_cast<uint4*>(s + L->off_gmask);
    g->gcount = reinterpret_cast<uint2*>(s + L->off_gcount);
    g->goff = reinterpret_cast<uint2*>(s + L->off_goff);
    g->bsum = reinterpret_cast<uint2*>(s + L->off_bsum);
    g->active = reinterpret_cast<unsigned int*>(s + L->off_active);
    g->n_active = reinterpret_cast<unsigned int*>(s + L->off_nactive);
    g->totals = nullptr;
    g->n_groups = L->n_groups;
    return HV_OK;
}

}  // namespace

extern "C" int hv_mc_scratch_bytes(const int64_t* h_dims, int has_halo_lo, int has_halo_hi, int64_t* h_bytes) {
    McLayout L;
    int rc = mc_layout(h_dims, has_halo_lo != 0, has_halo_hi != 0, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(h_bytes, "hv_mc_scratch_bytes: null output");
    *h_bytes = (int64_t)L.total;
    return HV_OK;
}

extern "C" int hv_mc_count(const float* tsdf, const float* halo_lo, const float* halo_hi, const int64_t* h_dims,
                           int64_t pitch, double level, void* scratch, int64_t scratch_bytes,
                           unsigned long long* totals, void* stream) {
    hv::McGrid g;
    McLayout L;
    int rc = mc_grid(&g, tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(totals, "hv_mc_count: null totals");
    g.totals = totals;
    cudaStream_t st = hv::as_stream(stream);
    hv::mc_init_kernel<<<1, 32, 0, st>>>(totals, g.n_active);  // (a pageable H2D of the four counters would block the host)
    const int sms = hv::device_info().sm_count;
    hv::McSignGeom sg;
    sg.segs = (g.Zb + 127) / 128;
    const long long n_tiles = (long long)g.X * g.Y * sg.segs;
    HV_REQUIRE(n_tiles < (1ll << 31), "hv_mc_count: too many columns");
    sg.n_tiles = (int)n_tiles;
    sg.vec = (!halo_lo && (pitch & 3) == 0 && ((uintptr_t)tsdf & 15) == 0) ? 1 : 0;
    const long long want = ((long long)g.X * g.Y + 7) / 8;  // 8 warps per block, a warp per column (grid-stride)
    const long long cap = (long long)sms * 8;
    const int sign_blocks = (int)(want < cap ? (want > 0 ? want : 1) : cap);
    hv::mc_sign_kernel<<<sign_blocks, 256, 0, st>>>(g, sg);
    const unsigned int group_blocks = (unsigned int)((g.n_groups + 255) / 256);
    hv::mc_count_kernel<<<group_blocks, 256, 0, st>>>(g);
    hv::exclusive_scan_u2(g.gcount, g.goff, g.bsum, g.n_groups, totals, st);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_mc_emit(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                          const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                          double level, void* scratch, int64_t scratch_bytes, int64_t z_global_offset, int64_t z_global_dim,
                          float voxel_size, const float* h_origin, int64_t n_verts, int64_t n_tris, float* verts_world,
                          float* verts_voxel, float* normals, uint8_t* colors, int32_t* faces,
                          unsigned long long* face_keys, unsigned long long* face_cubes, unsigned long long* vert_keys,
                          void* stream) {
    hv::McGrid g;
    McLayout L;
    int rc = mc_grid(&g, tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(n_verts >= 0 && n_tris >= 0 && n_verts < (1ll << 31) && n_tris < (1ll << 31), "hv_mc_emit: bad counts");
    if (n_verts == 0 && n_tris == 0) return HV_OK;  // a shard the surface does not cross
    HV_REQUIRE(color && h_origin && vert_keys && verts_world && normals && colors, "hv_mc_emit: null pointer");
    HV_REQUIRE((faces != nullptr) != (face_keys != nullptr), "hv_mc_emit: pass exactly one of faces / face_keys");
    HV_REQUIRE((halo_lo != nullptr) == (color_halo_lo != nullptr) && (halo_hi != nullptr) == (color_halo_hi != nullptr),
               "hv_mc_emit: tsdf and colour halos must come in pairs");
    HV_REQUIRE(faces == nullptr || (!halo_lo && !halo_hi), "hv_mc_emit: a shard with halos must emit face_keys");
    HV_REQUIRE(z_global_dim >= z_global_offset + g.Z, "hv_mc_emit: slab exceeds the global z extent");
    g.color = color;
    g.chalo_lo = color_halo_lo;
    g.chalo_hi = color_halo_hi;
    g.zoff = z_global_offset - g.hl;
    g.Zg = z_global_dim;
    cudaStream_t st = hv::as_stream(stream);
    hv::McEmit em{faces, face_keys, vert_keys, face_keys ? face_cubes : nullptr};
    // one thread per active group; the list length lives on the device, so size the grid by what the outputs imply
    long long emit_groups = n_verts + n_tris;  // >= number of active groups
    if (emit_groups > g.n_groups) emit_groups = g.n_groups;
    const long long max_blocks = (long long)hv::device_info().sm_count * 8;
    long long emit_blocks = (emit_groups + hv::kEmitThreads - 1) / hv::kEmitThreads;
    if (emit_blocks > max_blocks) emit_blocks = max_blocks;
    if (emit_blocks < 1) emit_blocks = 1;
    hv::mc_emit_kernel<<<(unsigned int)emit_blocks, hv::kEmitThreads, 0, st>>>(g, em);
    if (n_verts > 0) {
        hv::McVerts mv;
        mv.vert_keys = vert_keys;
        mv.n_verts = n_verts;
        mv.verts_world = verts_world;
        mv.verts_voxel = verts_voxel;
        mv.normals = normals;
        mv.colors = colors;
        mv.voxel_size = voxel_size;
        for (int k = 0; k < 3; ++k) mv.origin[k] = h_origin[k];
        hv::mc_vertex_kernel<<<(unsigned int)((n_verts + 127) / 128), 128, 0, st>>>(g, mv);
    }
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

// Count + emit in one call: the only host round trip (the output sizes are data dependent) is a stream synchronise on
// four counters copied into the caller's pinned `h_totals`; the caller provides output buffers with a capacity and gets
// HV_E_CAPACITY (with h_totals filled, scratch still valid for a plain hv_mc_emit) if the surface is larger.
extern "C" int hv_mc_extract(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                             const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                             double level, void* scratch, int64_t scratch_bytes, int64_t z_global_offset,
                             int64_t z_global_dim, float voxel_size, const float* h_origin, int64_t cap_verts,
                             int64_t cap_tris, float* verts_world, float* verts_voxel, float* normals, uint8_t* colors,
                             int32_t* faces, unsigned long long* face_keys, unsigned long long* face_cubes,
                             unsigned long long* vert_keys, unsigned long long* totals, unsigned long long* h_totals,
                             void* stream) {
    HV_REQUIRE(h_totals, "hv_mc_extract: null h_totals");
    int rc = hv_mc_count(tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, totals, stream);
    if (rc != HV_OK) return rc;
    cudaStream_t st = hv::as_stream(stream);
    HV_CUDA(cudaMemcpyAsync(h_totals, totals, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    HV_CUDA(cudaStreamSynchronize(st));
    const int64_t n_verts = (int64_t)h_totals[0], n_tris = (int64_t)h_totals[1];
    auto key_to_float = [](unsigned long long k) {
        int i = (int)(unsigned int)k;
        i = i >= 0 ? i : i ^ 0x7fffffff;
        float f;
        memcpy(&f, &i, sizeof(f));
        return f;
    };
    const double vmin = key_to_float(h_totals[2]), vmax = key_to_float(h_totals[3]);
    if (level < vmin || level > vmax) {
        hv::set_error("Surface level must be within volume data range.");
        return HV_E_LEVEL_RANGE;
    }
    if (n_verts == 0) {
        hv::set_error("No surface found at the given iso value.");
        return HV_E_NOSURFACE;
    }
    if (n_verts > cap_verts || n_tris > cap_tris) {
        hv::set_error("hv_mc_extract: %lld vertices / %lld triangles exceed the capacity %lld / %lld", (long long)n_verts,
                      (long long)n_tris, (long long)cap_verts, (long long)cap_tris);
        return HV_E_CAPACITY;
    }
    return hv_mc_emit(tsdf, halo_lo, halo_hi, color, color_halo_lo, color_halo_hi, h_dims, pitch, level, scratch,
                      scratch_bytes, z_global_offset, z_global_dim, voxel_size, h_origin, n_verts, n_tris, verts_world,
                      verts_voxel, normals, colors, faces, face_keys, face_cubes, vert_keys, stream);
}
