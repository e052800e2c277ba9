// Marching-cubes mesh extraction for sm_100a (B200).
//
// Replaces the `measure.marching_cubes(tsdf_vol, level=0)` call of TSDFVolume.get_mesh / get_point_cloud
// (third_party/tsdf_fusion_python/fusion.py:317,338) and the in-tree post-processing after it (fusion.py:339-348:
// nearest-voxel colour gather, b/g/r unpack, voxel -> world transform).  The arithmetic contract is the one restated in
// oracle/mc_oracle.c (scikit-image itself is not available anywhere on the box: parity unpinned).
//
// Pipeline (HBM-bound streaming + sparse compaction, no tensor cores), DESIGN.md 4.2.  The global order of the 32-voxel
// z-groups (column-major, z fastest) is the order of vertices and triangles; a tile = 1024 consecutive groups.
//   1. mc_sign_kernel    the one pass over the TSDF plane (4 B/voxel): one sign bit per voxel (N/8 bytes, L2 resident),
//                        a warp vote per word; per-block min / max for the skimage level-range check.
//   2. mc_detect_kernel  a thread per four groups: does any sign change come near?  Ordered compaction of those groups
//                        into the tile's list; one work item per 32 listed groups.
//   3. mc_count_kernel   a warp per work item, no barriers: lane = listed group (masks of the voxels owning a vertex on
//                        their +axis0 / +axis1 / +axis2 edge, surface-cube mask), then lane = surface cube (warp scans
//                        spread the cubes, 32 at a time whatever the shape of the surface).  Every tiling decision is made
//                        once: cubes that need the face / interior tests are queued and evaluated 32 at a time (the only
//                        step that waits on TSDF values).  Written: a 16-byte record {first vertex, mx, my, mz} in place
//                        for every group that owns a vertex (`gtab`, sparse: the index look-up of a triangle corner is
//                        ONE load plus two hot small-table entries), one work record per surface cube (with its tiling
//                        and batch-local triangle offset) and per edge vertex, the batch totals.
//   4. mc_scan_kernel    batch totals -> batch bases within the tile (a warp per tile), tile totals -> tile bases (last
//                        block), grand totals, min / max; resets the counters for the next extraction.
//   5. mc_emit_kernel    one thread per cube record (its triangles: the vertex indices of the cube's edges are looked up
//                        first, all loads independent; its centre vertex) and one thread per vertex record (position,
//                        normal accumulated over the <= 4 cubes around the edge in scan order, colour, world transform
//                        -- the edge's 2x3x3 neighbourhood in registers, nothing indexed dynamically), the two kinds of
//                        block interleaved in one grid.
// Launches 2-5 are programmatic dependents (their set-up overlaps the predecessor's tail) and read the totals on the
// device, so all five can be queued back to back; the host reads the totals once, afterwards.
// Vertex order: edge vertices sorted by (owner voxel in scan order, axis), then the cube-centre vertices by cube;
// triangles in cube scan order.  Deterministic; no dedup pass because each lattice edge has exactly one owner.
//
// z-slab shards pass one halo plane below / above the slab (`halo_lo` / `halo_hi`); keys carry global voxel indices so
// per-slab meshes can be stitched exactly (hive_b200/sharded.py).
#include "hv_common.cuh"

#include <float.h>
#include <stdlib.h>
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
    unsigned int* bits;          // [n_groups] sign words
    uint4* gtab;                 // [n_groups], written only where the group owns a vertex: {batch-local first vertex | batch of the tile << 12, mx, my, mz}
    unsigned int* glist;         // [n_tiles][1024] ids of the groups near the surface, per tile, ascending
    unsigned int* tile_cnt;      // [n_tiles] entries of glist used
    unsigned short* cvar;        // [n_groups][32] tiling of a cube, written only for cubes whose tiling needs the face / interior tests
    int2* sign_minmax;           // [sign_blocks] min / max key (ordered-int float) of the voxels a block of the sign kernel saw
    int sign_blocks;
    unsigned int* work;          // [n_batches] work items of the count kernel: tile * 32 + batch of the tile, any order
    uint4* batch_tot;            // [n_batches] {edge vertices, triangles, centre vertices, cubes} of a batch
    uint4* batch_base;           // [n_batches] {edge vertices, triangles, centre vertices} of the tile before this batch of 32 listed groups
    uint4* tile_tot;             // [n_tiles] {edge vertices, triangles, centre vertices, cubes} of every tile of 32 batches
    uint4* tile_base;            // [n_tiles] exclusive prefix of tile_tot (x, y, z)
    unsigned int* ctrl;          // [0] scan blocks done, [3] work items, [4..5] cursors of the record arrays (vertices | cubes << 32), [6] work ticket
    // work records (caller-sized: hv_mc_work_bytes)
    uint4* cube_rec;             // {group, voxel | tiling << 5 | batch of the tile << 16, batch-local first triangle, batch-local centre vertex}
    unsigned long long* vert_rec;  // pack_vertex_record()
    long long cap_cubes, cap_vrec;
    unsigned long long* totals;  // [0] vertices [1] triangles [2] min key [3] max key (ordered-int float) [4] edge vertices
                                 // [5] surface cubes [6] 1 if the records did not fit
    long long n_groups, n_batches;
    int n_tiles;
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

// ---- cube helpers -----------------------------------------------------------------------------------------------------
struct CubeWords {
    unsigned int w[8];    // per corner: the word whose bit k is that corner of cube (x, y, 32*gq + k)
    unsigned int active;  // cubes of this group that exist, are owned and are crossed by the surface
    unsigned int mx, my, mz;  // voxels of the group owning a vertex on their +axis0 / +axis1 / +axis2 edge
};

__device__ __forceinline__ unsigned int range_mask(int lo, int hi, int base) {
    // bits k with lo <= base + k < hi
    const int a = max(lo - base, 0), b = min(hi - base, 32);
    if (a >= b) return 0u;
    const unsigned int upto_b = (b >= 32) ? 0xffffffffu : ((1u << b) - 1u);
    return upto_b & ~((1u << a) - 1u);
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

// ---- 1. sign bits + min/max ----------------------------------------------------------------------------------------
// A warp takes whole columns (grid-stride).  Lane l reads the voxels zb = 32 w + l of sixteen sign words w at a time (sixteen
// coalesced 128-byte loads in flight), and a word is one warp vote -- no nibble packing, no shuffles.  (r01: the float4 /
// shuffle form spent 79 warp-instructions per 128 voxels and sat at 73 % issue utilisation with DRAM at 66 %.)
constexpr int kSignBatch = 16;

__global__ void __launch_bounds__(256) mc_sign_kernel(const __grid_constant__ McGrid g) {
    const int lane = threadIdx.x & 31;
    const int warp0 = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int n_warps = (int)((gridDim.x * blockDim.x) >> 5);
    float vmin = INFINITY, vmax = -INFINITY;
    const float level_f = (float)g.level;  // exact compare needs double only when the level is not a float
    const bool level_exact = (double)level_f == g.level;
    const int n_cols = g.X * g.Y;
    for (int col = warp0; col < n_cols; col += n_warps) {
        const float* base = g.tsdf + (size_t)col * g.pitch + lane;
        unsigned int* dst = g.bits + (size_t)col * g.G;
        for (int w0 = 0; w0 < g.G; w0 += kSignBatch) {
            float v[kSignBatch];
            if (g.hl == 0 && (w0 + kSignBatch) * 32 <= g.Z) {
#pragma unroll
                for (int k = 0; k < kSignBatch; ++k) v[k] = __ldcs(base + (w0 + k) * 32);
            } else {
#pragma unroll
                for (int k = 0; k < kSignBatch; ++k) {
                    const int zb = (w0 + k) * 32 + lane, z = zb - g.hl;
                    v[k] = NAN;  // padding: compares false, ignored by fminf / fmaxf
                    if (zb < g.Zb) v[k] = z < 0 ? __ldg(g.halo_lo + col) : (z >= g.Z ? __ldg(g.halo_hi + col) : __ldcs(g.tsdf + (size_t)col * g.pitch + z));
                }
            }
            unsigned int mine = 0u;
#pragma unroll
            for (int k = 0; k < kSignBatch; ++k) {
                const bool pos = level_exact ? (v[k] > level_f) : ((double)v[k] > g.level);
                const unsigned int word = __ballot_sync(0xffffffffu, pos);
                if (lane == k) mine = word;
                vmin = fminf(vmin, v[k]);
                vmax = fmaxf(vmax, v[k]);
            }
            if (lane < kSignBatch && w0 + lane < g.G) dst[w0 + lane] = mine;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        vmin = fminf(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
        vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    }
    __shared__ float s_min[8], s_max[8];
    if (lane == 0) {
        s_min[threadIdx.x >> 5] = vmin;
        s_max[threadIdx.x >> 5] = vmax;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) {
            vmin = fminf(vmin, s_min[w]);
            vmax = fmaxf(vmax, s_max[w]);
        }
        // (no voxel seen: min > max, keys stay at their neutral values)
        g.sign_minmax[blockIdx.x] = vmin <= vmax ? make_int2(float_order_key(vmin), float_order_key(vmax)) : make_int2(0x7fffffff, (int)0x80000000);
    }
}

// ---- group helpers ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned int decode_group(const McGrid& g, unsigned int gid, int& x, int& y, int& gq) {
    const unsigned int col = gid / (unsigned int)g.G;
    gq = (int)(gid - col * (unsigned int)g.G);
    x = (int)(col / (unsigned int)g.Y);
    y = (int)(col - (unsigned int)x * (unsigned int)g.Y);
    return col;
}

__device__ __forceinline__ void load_pair(const McGrid& g, size_t gi, int gq, bool in, unsigned int& b, unsigned int& u) {
    b = 0u;
    u = 0u;
    if (!in) return;
    b = __ldg(g.bits + gi);
    const unsigned int nxt = (gq + 1 < g.G) ? __ldg(g.bits + gi + 1) : 0u;
    u = (b >> 1) | (nxt << 31);
}

// Sign words around group (x, y, gq) from the bit volume, the masks of the vertices its voxels own and of its surface cubes.
__device__ __forceinline__ CubeWords load_cube_words(const McGrid& g, int x, int y, int gq) {
    CubeWords w;
    const bool x1 = x + 1 < g.X, y1 = y + 1 < g.Y;
    const size_t gi = ((size_t)x * g.Y + y) * g.G + gq;
    const size_t dx = (size_t)g.Y * g.G;
    unsigned int b00, b01, b10, b11, u00, u01, u10, u11;
    load_pair(g, gi, gq, true, b00, u00);
    load_pair(g, gi + g.G, gq, y1, b01, u01);
    load_pair(g, gi + dx, gq, x1, b10, u10);
    load_pair(g, gi + dx + g.G, gq, x1 && y1, b11, u11);
    // corner i offsets (axis0, axis1, axis2): see kMcCornerOff
    w.w[0] = b00; w.w[1] = u00; w.w[2] = u01; w.w[3] = b01;
    w.w[4] = b10; w.w[5] = u10; w.w[6] = u11; w.w[7] = b11;
    const int base = gq * 32;
    const unsigned int own = range_mask(g.own_lo, g.own_hi, base), zin = range_mask(0, g.Zb - 1, base);
    w.mx = x1 ? ((b00 ^ b10) & own) : 0u;
    w.my = y1 ? ((b00 ^ b01) & own) : 0u;
    w.mz = (b00 ^ u00) & own & zin;
    const unsigned int all8 = b00 & b01 & b10 & b11 & u00 & u01 & u10 & u11;
    const unsigned int any8 = b00 | b01 | b10 | b11 | u00 | u01 | u10 | u11;
    w.active = (x1 && y1) ? (any8 & ~all8 & own & zin) : 0u;
    return w;
}

// ---- warp-level expansion helpers ----------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned int warp_excl_scan(unsigned int v, unsigned int& total) {
    const int lane = threadIdx.x & 31;
    unsigned int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}

// Largest lane s with off[s] <= q, `off` being an exclusive scan over the lanes: the lane whose items cover ordinal q.
__device__ __forceinline__ int warp_owner(unsigned int off, unsigned int q) {
    int s = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
        const unsigned int v = __shfl_sync(0xffffffffu, off, s + step);
        if (v <= q) s += step;
    }
    return s;
}

// Position of the r-th (0-based) set bit of `mask` (r < popc(mask)).
__device__ __forceinline__ int select_bit(unsigned int mask, int r) {
    int pos = 0;
#pragma unroll
    for (int w = 16; w > 0; w >>= 1) {
        const int cnt = __popc((mask >> pos) & ((1u << w) - 1u));
        if (r >= cnt) {
            r -= cnt;
            pos += w;
        }
    }
    return pos;
}

// ---- 2a. detect: which groups does the surface come near? ---------------------------------------------------------------------
// A tile = 1024 consecutive groups (the global group order is the vertex / triangle order).  One thread looks at four
// consecutive groups (one 16-byte load per neighbouring column when the column length allows it) and the block compacts the
// ids of the groups with a sign change in reach -- in order -- into the tile's slice of `glist`.  With a floor in the scene
// every column has one such group and 30 that have none: the kernels after this one only ever see the former, 32 to a warp.
constexpr int kTileGroups = 1024;
constexpr int kTileBatches = kTileGroups / 32;
constexpr int kDetectThreads = kTileGroups / 4;

__device__ __forceinline__ bool group_near_surface(const unsigned int (&b)[4], const unsigned int (&n)[4]) {
    unsigned int all8 = 0xffffffffu, any8 = 0u;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const unsigned int u = (b[i] >> 1) | (n[i] << 31);
        all8 &= b[i] & u;
        any8 |= b[i] | u;
    }
    return (any8 & ~all8) != 0u;  // (columns / words beyond the volume read as zero: conservative)
}

__global__ void __launch_bounds__(kDetectThreads) mc_detect_kernel(const __grid_constant__ McGrid g) {
    __shared__ unsigned int s_warp[kDetectThreads / 32];
    cudaGridDependencySynchronize();  // launched as a programmatic dependent of the sign kernel (launch latency hidden)
    const unsigned int tile = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long gid0 = (long long)tile * kTileGroups + 4 * (long long)threadIdx.x;
    unsigned int flags = 0u;
    if (gid0 < g.n_groups) {
        const int G = g.G;
        if ((G & 3) == 0) {  // the four groups share a column and a 16-byte word
            const unsigned int col = (unsigned int)gid0 / (unsigned int)G;
            const int gq = (int)((unsigned int)gid0 - col * (unsigned int)G);
            const unsigned int x = col / (unsigned int)g.Y, y = col - x * (unsigned int)g.Y;
            const bool x1 = (int)x + 1 < g.X, y1 = (int)y + 1 < g.Y;
            const unsigned int* p[4];
            p[0] = g.bits + gid0;
            p[1] = y1 ? p[0] + G : nullptr;
            p[2] = x1 ? p[0] + (size_t)g.Y * G : nullptr;
            p[3] = (x1 && y1) ? p[2] + G : nullptr;
            unsigned int w[4][5];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                uint4 q = make_uint4(0u, 0u, 0u, 0u);
                unsigned int nx = 0u;
                if (p[i]) {
                    q = __ldg(reinterpret_cast<const uint4*>(p[i]));
                    if (gq + 4 < G) nx = __ldg(p[i] + 4);
                }
                w[i][0] = q.x; w[i][1] = q.y; w[i][2] = q.z; w[i][3] = q.w; w[i][4] = nx;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned int b[4] = {w[0][k], w[1][k], w[2][k], w[3][k]};
                const unsigned int n[4] = {w[0][k + 1], w[1][k + 1], w[2][k + 1], w[3][k + 1]};
                if (group_near_surface(b, n)) flags |= 1u << k;
            }
        } else {
            for (int k = 0; k < 4 && gid0 + k < g.n_groups; ++k) {
                int x, y, gq;
                decode_group(g, (unsigned int)(gid0 + k), x, y, gq);
                const CubeWords c = load_cube_words(g, x, y, gq);
                unsigned int all8 = 0xffffffffu, any8 = 0u;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    all8 &= c.w[i];
                    any8 |= c.w[i];
                }
                if (any8 & ~all8) flags |= 1u << k;
            }
        }
    }
    const unsigned int cnt = __popc(flags);
    unsigned int total;
    const unsigned int ex = warp_excl_scan(cnt, total);
    if (lane == 31) s_warp[warp] = ex + cnt;
    __syncthreads();
    unsigned int carry = 0, sum = 0;
#pragma unroll
    for (int w = 0; w < kDetectThreads / 32; ++w) {
        const unsigned int v = s_warp[w];
        if (w < warp) carry += v;
        sum += v;
    }
    unsigned int* dst = g.glist + (size_t)tile * kTileGroups + carry + ex;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if ((flags >> k) & 1u) *dst++ = (unsigned int)(gid0 + k);
    if (threadIdx.x == 0) g.tile_cnt[tile] = sum;
    // one work item per batch of 32 listed groups (their order in `work` does not matter)
    if (warp == 0 && sum) {
        const unsigned int nb = (sum + 31u) / 32u;
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(g.ctrl + 3, nb);
        base = __shfl_sync(0xffffffffu, base, 0);
        if ((unsigned int)lane < nb) g.work[base + lane] = tile * kTileBatches + (unsigned int)lane;
    }
}

// ---- 2b. count: per-group masks, work records, scan ------------------------------------------------------------------------
// The unit of work is a batch of 32 consecutive entries of a tile's group list, one per lane: sign words of the four
// columns around the group, owner masks, surface-cube mask.  (Batches of 32 consecutive groups -- two columns -- cost
// 167 us at 512^3: with a floor in the scene every one of the 131 k batches had one or two lanes with work.)  The cubes
// of a batch are then spread over the lanes with a warp scan -- 32 cubes at a time whatever the
// shape of the surface (one thread per group walked up to 32 cubes serially, one thread per column up to Z: 315 us and
// 370 us at 512^3 with a wall parallel to z in the scene) -- so every tiling decision (face and interior tests included)
// is made ONCE, lane-parallel, and lands in the cube's work record together with the cube's batch-local triangle offset.
// A tile has at most 32 batches (16 warps x 2); the batch totals are scanned in the tile, the last tile to finish scans the tile
// totals: a consumer adds two hot small-table entries per look-up.  (A decoupled look-back over tiles of 256 groups was
// measured at 520 us: with ~600 tiles in flight every tile walks back ~20 windows.)  Records go to a bump-allocated range
// per batch: their order in the record arrays is arbitrary, the indices they carry are not.
constexpr int kCountThreads = 128;
constexpr int kCubeChunk = 256;
static_assert(HV_MC_NUM_VARIANTS <= 1024, "cube words keep the tiling in 10 bits");

// vertex record: group (31 bits) | voxel << 31 (5) | axis << 36 (2) | batch-local vertex index << 38 (12: <= 32 * 96) |
// batch of the tile << 50 (5)
__device__ __forceinline__ unsigned long long pack_vertex_record(unsigned int gid, int k, int axis, unsigned int local, int slot) {
    return (unsigned long long)gid | ((unsigned long long)k << 31) | ((unsigned long long)axis << 36) |
           ((unsigned long long)local << 38) | ((unsigned long long)slot << 50);
}

__global__ void __launch_bounds__(kCountThreads, 8) mc_count_kernel(const __grid_constant__ McGrid g) {
    __shared__ unsigned int s_cube[kCountThreads / 32][kCubeChunk];
    __shared__ unsigned short s_queue[kCountThreads / 32][kCubeChunk];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned int lt = (1u << lane) - 1u;
    cudaGridDependencySynchronize();
    const unsigned int n_work = g.ctrl[3];
    const unsigned int n_warps = gridDim.x * (kCountThreads / 32);
    // the first item of a warp is its own index, the following ones come from a ticket counter -- requested before the
    // current item is processed, so the round trip is never waited for
    unsigned int wi = blockIdx.x * (kCountThreads / 32) + warp;
#pragma unroll 1
    while (wi < n_work) {
        unsigned int ticket = 0;
        if (lane == 0) ticket = atomicAdd(g.ctrl + 6, 1u);
        const unsigned int item = __ldg(g.work + wi);
        const unsigned int tile = item / kTileBatches;
        const int slot = (int)(item % kTileBatches);
        const unsigned int n_listed = __ldg(g.tile_cnt + tile);
        const unsigned int li = (unsigned int)slot * 32u + (unsigned int)lane;
        uint4 tot = make_uint4(0u, 0u, 0u, 0u);
        unsigned int gid = 0u;
        int x = 0, y = 0, gq = 0;
        CubeWords c;
#pragma unroll
        for (int i = 0; i < 8; ++i) c.w[i] = 0u;
        c.active = c.mx = c.my = c.mz = 0u;
        if (li < n_listed) {
            gid = __ldg(g.glist + (size_t)tile * kTileGroups + li);
            decode_group(g, gid, x, y, gq);
            c = load_cube_words(g, x, y, gq);
        }
        if (__ballot_sync(0xffffffffu, (c.active | c.mx | c.my | c.mz) != 0u)) {
            // ---- edge vertices: the group's record in place, one work record per vertex
            const unsigned int nv = __popc(c.mx) + __popc(c.my) + __popc(c.mz);
            unsigned int NV;
            const unsigned int vpre = warp_excl_scan(nv, NV);
            if (nv) g.gtab[gid] = make_uint4(vpre | ((unsigned int)slot << 12), c.mx, c.my, c.mz);
            unsigned int C;
            const unsigned int coff = warp_excl_scan(__popc(c.active), C);
            // space in both record arrays with one atomic: (vertex cursor | cube cursor << 32)
            unsigned long long cursors = 0ull;
            if (lane == 0)
                cursors = atomicAdd(reinterpret_cast<unsigned long long*>(g.ctrl + 4), (unsigned long long)NV | ((unsigned long long)C << 32));
            cursors = __shfl_sync(0xffffffffu, cursors, 0);
            const unsigned int vbase = (unsigned int)cursors, cbase = (unsigned int)(cursors >> 32);
            if (NV) {
                if ((long long)vbase + NV <= g.cap_vrec) {
                    unsigned long long* dst = g.vert_rec + vbase;
                    unsigned int o = vpre;
                    for (unsigned int rest = c.mx | c.my | c.mz; rest; rest &= rest - 1) {
                        const int k = __ffs(rest) - 1;
                        if ((c.mx >> k) & 1u) { dst[o] = pack_vertex_record(gid, k, 0, o, slot); ++o; }
                        if ((c.my >> k) & 1u) { dst[o] = pack_vertex_record(gid, k, 1, o, slot); ++o; }
                        if ((c.mz >> k) & 1u) { dst[o] = pack_vertex_record(gid, k, 2, o, slot); ++o; }
                    }
                }
            }
            // ---- surface cubes, one lane per cube
            const bool write_ok = (long long)cbase + C <= g.cap_cubes;
            unsigned int tri_run = 0, cen_run = 0;
            // 256 cubes at a time, three sweeps: (A) configuration index of every cube, tiling where the tables alone
            // decide, the others queued; (B) the queued cubes, one per lane -- the only step that waits on TSDF values, and
            // it waits once per 32 ambiguous cubes instead of once per 32 cubes (r02: a dense batch in a noisy region had an
            // ambiguous cube in every one of its 32 rounds: 64 us, with the other 15 warps of the tile at the barrier);
            // (C) scans + records.
            unsigned int* s_var = s_cube[warp];
            unsigned short* s_q = s_queue[warp];
            for (unsigned int ch0 = 0; ch0 < C; ch0 += kCubeChunk) {
                const unsigned int ch1 = min(C, ch0 + kCubeChunk);
                unsigned int n_queued = 0;
                for (unsigned int c0 = ch0; c0 < ch1; c0 += 32u) {  // (A)
                    const unsigned int q = c0 + (unsigned int)lane;
                    const bool cv = q < ch1;
                    const int s = warp_owner(coff, cv ? q : ch1 - 1u);
                    const unsigned int r = q - __shfl_sync(0xffffffffu, coff, s);
                    const unsigned int am = __shfl_sync(0xffffffffu, c.active, s);
                    unsigned int w[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) w[i] = __shfl_sync(0xffffffffu, c.w[i], s);
                    bool amb = false;
                    if (cv) {
                        const int k = select_bit(am, (int)r);
                        int idx = 0;
#pragma unroll
                        for (int i = 0; i < 8; ++i) idx |= (int)((w[i] >> k) & 1u) << i;
                        amb = kMcAmbCount[idx] != 0 || kMcInteriorBase[idx] != 255;
                        // tiling (10 bits; the configuration index for now if it is still open) | voxel << 10 | lane of the group << 15
                        s_var[q - ch0] = (amb ? (unsigned int)idx : (unsigned int)kMcVariantBase[idx]) | ((unsigned int)k << 10) | ((unsigned int)s << 15);
                    }
                    const unsigned int ab = __ballot_sync(0xffffffffu, amb);
                    if (amb) s_q[n_queued + __popc(ab & lt)] = (unsigned short)(q - ch0);
                    n_queued += __popc(ab);
                }
                __syncwarp();
                for (unsigned int i0 = 0; i0 < n_queued; i0 += 32u) {  // (B)
                    const bool qv = i0 + (unsigned int)lane < n_queued;
                    const unsigned int e = qv ? s_q[i0 + lane] : 0u;
                    const unsigned int word = s_var[e];
                    const int s = (int)(word >> 15), k = (int)((word >> 10) & 31u);
                    const int cx = __shfl_sync(0xffffffffu, x, s), cy = __shfl_sync(0xffffffffu, y, s);
                    const int cgq = __shfl_sync(0xffffffffu, gq, s);
                    const unsigned int qgid = __shfl_sync(0xffffffffu, gid, s);
                    if (qv) {
                        const int variant = cube_variant_ambiguous(g.tsdf, g.halo_lo, g.halo_hi, g.Y, g.pitch, g.Z, g.hl, g.level,
                                                                   (int)(word & 1023u), cx, cy, cgq * 32 + k);
                        s_var[e] = (word & ~1023u) | (unsigned int)variant;
                        g.cvar[(size_t)qgid * 32 + k] = (unsigned short)variant;  // the vertex kernel looks it up
                    }
                }
                __syncwarp();
                for (unsigned int c0 = ch0; c0 < ch1; c0 += 32u) {  // (C)
                    const unsigned int q = c0 + (unsigned int)lane;
                    const bool cv = q < ch1;
                    const unsigned int word = cv ? s_var[q - ch0] : 0u;
                    const int variant = (int)(word & 1023u);
                    const unsigned int cgid = __shfl_sync(0xffffffffu, gid, (int)(word >> 15));
                    const unsigned int ntri = cv ? kMcVarNtri[variant] : 0u;
                    const bool hasc = cv && ((kMcVarMult[variant] >> 48) & 15ull) != 0ull;
                    unsigned int T;
                    const unsigned int toff = warp_excl_scan(ntri, T);
                    const unsigned int cb = __ballot_sync(0xffffffffu, hasc);
                    if (cv && write_ok)
                        g.cube_rec[cbase + q] = make_uint4(cgid, ((word >> 10) & 31u) | ((unsigned int)variant << 5) | ((unsigned int)slot << 16),
                                                           tri_run + toff, cen_run + __popc(cb & lt));
                    tri_run += T;
                    cen_run += __popc(cb);
                }
                __syncwarp();
            }
            tot = make_uint4(NV, tri_run, cen_run, C);
        }
        if (lane == 0) g.batch_tot[item] = tot;
        wi = n_warps + __shfl_sync(0xffffffffu, ticket, 0);
    }
}

// ---- 2c. scan: batch totals within their tile, tile totals over the volume ----------------------------------------------------
// One warp per tile (its <= 32 batch totals); the last block to finish scans the tile totals: a consumer adds two hot
// small-table entries per look-up.  (A decoupled look-back over tiles of 256 groups inside the count kernel was measured at
// 520 us -- with ~600 tiles in flight every tile walks back ~20 windows -- and a tile-wide barrier in the count kernel at
// 135 us: most warps of a tile have no batch, all of them wait for the one that has a dense one.)
constexpr int kScanThreads2 = 256;

__global__ void __launch_bounds__(kScanThreads2) mc_scan_kernel(const __grid_constant__ McGrid g) {
    __shared__ unsigned long long s_wa[kScanThreads2 / 32], s_wb[kScanThreads2 / 32];
    __shared__ unsigned int s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long tile = (long long)blockIdx.x * (kScanThreads2 / 32) + warp;
    cudaGridDependencySynchronize();
    if (tile < g.n_tiles) {
        const unsigned int nb = (__ldg(g.tile_cnt + tile) + 31u) / 32u;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if ((unsigned int)lane < nb) v = __ldg(g.batch_tot + tile * kTileBatches + lane);
        unsigned int sx, sy, sz, sw;
        const unsigned int ex = warp_excl_scan(v.x, sx), ey = warp_excl_scan(v.y, sy), ez = warp_excl_scan(v.z, sz);
        warp_excl_scan(v.w, sw);
        if ((unsigned int)lane < nb) g.batch_base[tile * kTileBatches + lane] = make_uint4(ex, ey, ez, 0u);
        if (lane == 0) g.tile_tot[tile] = make_uint4(sx, sy, sz, sw);
    }
    // the last block to get here turns the tile totals into bases
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(g.ctrl, 1u) == gridDim.x - 1u) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // every thread takes a run of consecutive tiles, 16 at a time with all loads in flight together (a block-wide scan per
    // 256 tiles cost a load round trip per round: 12.5 us at 4096 tiles)
    const int n_tiles = g.n_tiles;
    const int per = (n_tiles + kScanThreads2 - 1) / kScanThreads2;
    const int i0 = min(n_tiles, (int)threadIdx.x * per), i1 = min(n_tiles, i0 + per);
    unsigned long long a = 0, bsum = 0;  // (edge vertices | triangles << 32), (centre vertices | cubes << 32): all < 2^31
    for (int j0 = i0; j0 < i1; j0 += 16) {
        uint4 v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = (j0 + j < i1) ? __ldcg(g.tile_tot + j0 + j) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            a += (unsigned long long)v[j].x | ((unsigned long long)v[j].y << 32);
            bsum += (unsigned long long)v[j].z | ((unsigned long long)v[j].w << 32);
        }
    }
    unsigned long long ia = a, ib = bsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long ua = __shfl_up_sync(0xffffffffu, ia, o), ub = __shfl_up_sync(0xffffffffu, ib, o);
        if (lane >= o) {
            ia += ua;
            ib += ub;
        }
    }
    if (lane == 31) {
        s_wa[warp] = ia;
        s_wb[warp] = ib;
    }
    __syncthreads();
    unsigned long long ea = ia - a, eb = ib - bsum, carry_a = 0, carry_b = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads2 / 32; ++w) {
        const unsigned long long xa = s_wa[w], xb = s_wb[w];
        if (w < warp) {
            ea += xa;
            eb += xb;
        }
        carry_a += xa;
        carry_b += xb;
    }
    for (int j0 = i0; j0 < i1; j0 += 16) {
        uint4 v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = (j0 + j < i1) ? __ldcg(g.tile_tot + j0 + j) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (j0 + j < i1) g.tile_base[j0 + j] = make_uint4((unsigned int)ea, (unsigned int)(ea >> 32), (unsigned int)eb, 0u);
            ea += (unsigned long long)v[j].x | ((unsigned long long)v[j].y << 32);
            eb += (unsigned long long)v[j].z | ((unsigned long long)v[j].w << 32);
        }
    }
    if (threadIdx.x == 0) {
        const unsigned long long n_edge = carry_a & 0xffffffffull, n_tri = carry_a >> 32, n_centre = carry_b & 0xffffffffull, n_cube = carry_b >> 32;
        g.totals[0] = n_edge + n_centre;  // edge vertices, then the cube-centre vertices
        g.totals[1] = n_tri;
        g.totals[4] = n_edge;
        g.totals[5] = n_cube;
        g.totals[6] = ((long long)n_cube > g.cap_cubes || (long long)n_edge > g.cap_vrec) ? 1ull : 0ull;  // records incomplete
    }
    // min / max of the volume (skimage's level-range check) from the sign kernel's per-block values
    int kmin = 0x7fffffff, kmax = (int)0x80000000;
    for (int i = threadIdx.x; i < g.sign_blocks; i += kScanThreads2) {
        const int2 v = __ldcg(g.sign_minmax + i);
        kmin = min(kmin, v.x);
        kmax = max(kmax, v.y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
        kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
    }
    __shared__ int s_kmin[kScanThreads2 / 32], s_kmax[kScanThreads2 / 32];
    if (lane == 0) {
        s_kmin[warp] = kmin;
        s_kmax[warp] = kmax;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kScanThreads2 / 32; ++w) {
            kmin = min(kmin, s_kmin[w]);
            kmax = max(kmax, s_kmax[w]);
        }
        g.totals[2] = (unsigned long long)(unsigned int)kmin;
        g.totals[3] = (unsigned long long)(unsigned int)kmax;
        // every counter back to zero for the next extraction (hv_mc_scratch_init zeroed them for the first)
        for (int i = 0; i < 8; ++i) g.ctrl[i] = 0u;
    }
}

// ---- 3. / 4. emit ----------------------------------------------------------------------------------------------------------
struct McEmit {
    int32_t* faces;                  // [T][3] local vertex indices, or null
    unsigned long long* face_keys;   // [T][3] global vertex keys, or null
    unsigned long long* vert_keys;   // [V], or null
    unsigned long long* face_cubes;  // [T] global linear index of the cube that produced the face, or null
    float* verts_world;              // [V][3]
    float* verts_voxel;              // [V][3] or null
    float* normals;                  // [V][3]
    uint8_t* colors;                 // [V][3] r, g, b
    long long cap_verts, cap_tris;   // capacity of the buffers above: anything beyond is counted, not written
    float voxel_size;
    float origin[3];
    // placement of a z-slab's mesh in the mesh of the whole volume (hv_mc_column_shifts), or null: per column, what to add
    // to the slab's own vertex / triangle / centre-vertex numbering, and the global index of the first vertex the slab
    // above owns in that column (the seam vertices a slab's top cubes refer to)
    const int32_t* vshift;
    const int32_t* tshift;
    const int32_t* cshift;
    const int32_t* vnext;
    unsigned int n_edge_global;      // edge vertices of the whole volume (centre vertices come after them)
    int32_t* vert_dst;               // [V] index of every vertex of this slab in the whole mesh
    int32_t* tri_dst;                // [T] ... of every triangle
    // mc_emit_kernel<.., DIRECT>: rows are written AT their whole-mesh index -- verts / normals / colors / faces are the whole
    // mesh's arrays (on this GPU or, mapped through NVLink, on a peer's), of this many rows
    long long gcap_verts, gcap_tris;
};

__device__ __forceinline__ unsigned long long vertex_key(const McGrid& g, int x, int y, int zb, int axis) {
    const unsigned long long lin = ((unsigned long long)x * g.Y + y) * (unsigned long long)g.Zg + (unsigned long long)(zb + g.zoff);
    return lin * 4ull + (unsigned long long)axis;
}

// Index of the edge vertex owned by voxel (xo, yo, zo) on `axis`: the tile's and the batch's base + the group's first vertex,
// plus the vertices of lower voxels, plus lower axes of its own -- one 16-byte load and two hot small-table entries.
__device__ __forceinline__ unsigned int edge_vertex_index(const McGrid& g, int xo, int yo, int zo, int axis) {
    const unsigned int og = ((unsigned int)xo * (unsigned int)g.Y + (unsigned int)yo) * (unsigned int)g.G + (unsigned int)(zo >> 5);  // < 2^31
    const int bit = zo & 31;
    const uint4 m = __ldg(g.gtab + og);
    const unsigned int ot = og / kTileGroups;
    const unsigned int base = __ldg(&g.tile_base[ot].x) + __ldg(&g.batch_base[ot * kTileBatches + (m.x >> 12)].x);
    const unsigned int low = (1u << bit) - 1u;
    unsigned int r = base + (m.x & 4095u) + __popc(m.y & low) + __popc(m.z & low) + __popc(m.w & low);
    if (axis > 0) r += (m.y >> bit) & 1u;
    if (axis > 1) r += (m.z >> bit) & 1u;
    return r;
}

// kMcEdgeOwner as a compile-time constant (5 bits per edge: d0 | d1 << 1 | d2 << 2 | axis << 3), so that the unrolled
// per-edge look-ups of mc_tri_kernel have static offsets.
constexpr unsigned long long kEdgeOwnerPacked =
    16ull | (12ull << 5) | (18ull << 10) | (8ull << 15) | (17ull << 20) | (13ull << 25) | (19ull << 30) | (9ull << 35) |
    (0ull << 40) | (4ull << 45) | (6ull << 50) | (2ull << 55);

// Triangle `j` of the cube (x, y, zb) with tiling `variant`, written as face `ti`.
__device__ __forceinline__ void emit_triangle(const McGrid& g, const McEmit& out, int x, int y, int zb, int variant, int j,
                                              size_t ti, unsigned int centre_index) {
    if ((long long)ti >= out.cap_tris) return;
    const uint8_t* tri = kMcTris + 3 * ((size_t)kMcVarTriOff[variant] + (size_t)j);
    int ox[3], oy[3], oz[3], axis[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const int e = tri[q];
        ox[q] = oy[q] = oz[q] = 0;
        axis[q] = 3;
        if (e < 12) {
            ox[q] = kMcEdgeOwner[4 * e], oy[q] = kMcEdgeOwner[4 * e + 1], oz[q] = kMcEdgeOwner[4 * e + 2];
            axis[q] = kMcEdgeOwner[4 * e + 3];
        }
    }
    if (out.face_keys) {
#pragma unroll
        for (int q = 0; q < 3; ++q) out.face_keys[3 * ti + q] = vertex_key(g, x + ox[q], y + oy[q], zb + oz[q], axis[q]);
        if (out.face_cubes) out.face_cubes[ti] = vertex_key(g, x, y, zb, 0) >> 2;
        return;
    }
    unsigned int v[3];
#pragma unroll
    for (int q = 0; q < 3; ++q)
        v[q] = axis[q] < 3 ? edge_vertex_index(g, x + ox[q], y + oy[q], zb + oz[q], axis[q]) : centre_index;
#pragma unroll
    for (int q = 0; q < 3; ++q) out.faces[3 * ti + q] = (int32_t)v[q];
}

// Cube-local frame of an edge vertex: corner i of a cube sits at (ia, ib, ic) along (edge axis, lower other axis, upper
// other axis); kCornerPos[axis] holds ia * 4 + ib * 2 + ic for i = 0..7, 3 bits each.
// corner offsets (axis0, axis1, axis2): 0:(0,0,0) 1:(0,0,1) 2:(0,1,1) 3:(0,1,0) 4:(1,0,0) 5:(1,0,1) 6:(1,1,1) 7:(1,1,0)
// axis 0: (ia, ib, ic) = (o0, o1, o2); axis 1: (o1, o0, o2); axis 2: (o2, o0, o1)   [tools/gen_mc_tables.py CORNERS]
constexpr unsigned int kCornerPos[3] = {0xdec4c8u, 0xdda948u, 0x7f2360u};

// sign bits at the eight cube-local positions (bit ia * 4 + ib * 2 + ic) -> Lewiner configuration index, per edge axis
struct LocToIndex {
    uint8_t t[3][256];
};
constexpr LocToIndex make_loc_to_index() {
    LocToIndex r{};
    for (int axis = 0; axis < 3; ++axis)
        for (int loc = 0; loc < 256; ++loc) {
            int idx = 0;
            for (int i = 0; i < 8; ++i) idx |= ((loc >> ((kCornerPos[axis] >> (3 * i)) & 7u)) & 1) << i;
            r.t[axis][loc] = (uint8_t)idx;
        }
    return r;
}
static __device__ const LocToIndex kLocToIndex = make_loc_to_index();

// cube-local position -> corner id (the inverse of kCornerPos), 3 bits each
constexpr unsigned int inverse_corner_pos(unsigned int cpos) {
    unsigned int inv = 0;
    for (unsigned int i = 0; i < 8; ++i) inv |= i << (3 * ((cpos >> (3 * i)) & 7u));
    return inv;
}
constexpr unsigned int kPosCorner0 = inverse_corner_pos(kCornerPos[0]), kPosCorner1 = inverse_corner_pos(kCornerPos[1]),
                       kPosCorner2 = inverse_corner_pos(kCornerPos[2]);

// out of line: the decision tree indexes its eight values dynamically (they live on the stack either way)
__device__ __noinline__ int cube_variant_values(int idx, const double* v) { return cube_variant(idx, v); }

// One edge vertex (fusion.py:338-348 for one vertex): position on the edge, normal accumulated over the (up to four) cubes
// around the edge in scan order, colour of the nearest voxel, world transform.  The edge's 2 x 3 x 3 neighbourhood lives in
// registers in an edge-relative frame (a = along the edge, b < c the other two axes): one base pointer and three strides,
// nothing is indexed dynamically, and the tilings of the surrounding cubes -- face and interior tests included -- are decided
// from those registers.  (r02: fetching the corners of every ambiguous cube again was 25 % of this kernel's instructions on
// the room scene, the per-value coordinate permutation another 25 %.)
// WHOLE: a whole volume (no halo planes): direct loads, and the tiling of an ambiguous cube is read from `cvar` (the count
// kernel decided it) instead of being derived again.  LEVEL0: level == 0, the value every caller in HIVE uses: x - 0.0 == x.
template <bool WHOLE, bool LEVEL0>
__device__ __forceinline__ void emit_edge_vertex(const McGrid& g, const McEmit& out, int x, int y, int zb, int axis, size_t vi) {
    const int pb = axis == 0 ? y : x, pc = axis == 2 ? y : zb;
    const int db_dim = axis == 0 ? g.Y : g.X, dc_dim = axis == 2 ? g.Y : g.Zb;
    bool okb[3], okc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        okb[i] = pb + i - 1 >= 0 && pb + i - 1 < db_dim;
        okc[i] = pc + i - 1 >= 0 && pc + i - 1 < dc_dim;
    }
    float n[2][3][3];
    if (WHOLE) {  // zb == z, every neighbour is in the plane array
        const ptrdiff_t sx = (ptrdiff_t)g.Y * g.pitch, sy = g.pitch;
        const ptrdiff_t sa = axis == 0 ? sx : (axis == 1 ? sy : 1), sb = axis == 0 ? sy : sx, sc = axis == 2 ? sy : 1;
        const float* p0 = g.tsdf + ((size_t)x * g.Y + y) * (size_t)g.pitch + zb;
#pragma unroll
        for (int ia = 0; ia < 2; ++ia)
#pragma unroll
            for (int ib = 0; ib < 3; ++ib)
#pragma unroll
                for (int ic = 0; ic < 3; ++ic)
                    n[ia][ib][ic] = (okb[ib] && okc[ic]) ? __ldg(p0 + ia * sa + (ib - 1) * sb + (ic - 1) * sc) : 0.0f;
    } else {
#pragma unroll
        for (int ia = 0; ia < 2; ++ia)
#pragma unroll
            for (int ib = 0; ib < 3; ++ib)
#pragma unroll
                for (int ic = 0; ic < 3; ++ic) {
                    const int qb = pb + ib - 1, qc = pc + ic - 1;
                    float f = 0.0f;
                    if (okb[ib] && okc[ic]) {
                        // (x, y, zb) of the neighbour: the three coordinates are (a, b, c) permuted by `axis`
                        const int qx = axis == 0 ? x + ia : qb;
                        const int qy = axis == 1 ? y + ia : (axis == 0 ? qb : qc);
                        const int qz = axis == 2 ? zb + ia : qc;
                        f = mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, qx, qy, qz);
                    }
                    n[ia][ib][ic] = f;
                }
    }
    const double level = LEVEL0 ? 0.0 : g.level;
#define HV_DV(f) (LEVEL0 ? (double)(f) : ((double)(f) - level))
    // sign bit of every neighbour, position ia * 9 + ib * 3 + ic
    unsigned int sgn = 0u;
#pragma unroll
    for (int ia = 0; ia < 2; ++ia)
#pragma unroll
        for (int ib = 0; ib < 3; ++ib)
#pragma unroll
            for (int ic = 0; ic < 3; ++ic)
                sgn |= ((LEVEL0 ? (n[ia][ib][ic] > 0.0f) : (HV_DV(n[ia][ib][ic]) > 0.0)) ? 1u : 0u) << (ia * 9 + ib * 3 + ic);
    const double v_lo = HV_DV(n[0][1][1]), v_hi = HV_DV(n[1][1][1]);
    const double w_lo = 1.0 / ((double)FLT_EPSILON + fabs(v_lo)), w_hi = 1.0 / ((double)FLT_EPSILON + fabs(v_hi));
    const double frac = w_hi / __dadd_rn(w_lo, w_hi);
    const unsigned int pos_corner = axis == 0 ? kPosCorner0 : (axis == 1 ? kPosCorner1 : kPosCorner2);
    float acc_a = 0.0f, acc_b = 0.0f, acc_c = 0.0f;
    // the cubes around the edge in scan order (axis0 outermost): larger offsets first; b is the lower axis
#pragma unroll
    for (int db = 1; db >= 0; --db) {
#pragma unroll
        for (int dc = 1; dc >= 0; --dc) {
            if (!okb[1 - db] || !okc[1 - dc] || !okb[2 - db] || !okc[2 - dc]) continue;  // the cube does not exist
            // sign bits of the cube's corners at cube-local (ia, ib, ic)
            unsigned int loc = 0;
#pragma unroll
            for (int ia = 0; ia < 2; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib)
#pragma unroll
                    for (int ic = 0; ic < 2; ++ic)
                        loc |= ((sgn >> (ia * 9 + (ib + 1 - db) * 3 + (ic + 1 - dc))) & 1u) << (ia * 4 + ib * 2 + ic);
            const int idx = kLocToIndex.t[axis][loc];
            int variant = kMcVariantBase[idx];
            if (kMcAmbCount[idx] != 0 || kMcInteriorBase[idx] != 255) {
              if (WHOLE) {
                // base voxel of the cube: (a, b, c) = (edge position, pb - db, pc - dc) permuted back by `axis`
                const int bb = pb - db, bc = pc - dc;
                const int cx = axis == 0 ? x : bb, cy = axis == 1 ? y : (axis == 0 ? bb : bc), cz = axis == 2 ? zb : bc;
                variant = __ldg(g.cvar + (((size_t)cx * g.Y + cy) * g.G + (cz >> 5)) * 32 + (cz & 31));
              } else {
                double v[8];
#pragma unroll
                for (int ia = 0; ia < 2; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib)
#pragma unroll
                        for (int ic = 0; ic < 2; ++ic)
                            v[(pos_corner >> (3 * (ia * 4 + ib * 2 + ic))) & 7u] = HV_DV(n[ia][ib + 1 - db][ic + 1 - dc]);
                variant = cube_variant_values(idx, v);
              }
            }
            // the edge inside this cube: lower corner at cube-local (0, db, dc)
            const int d0 = axis == 0 ? 0 : db, d1 = axis == 1 ? 0 : (axis == 0 ? db : dc), d2 = axis == 2 ? 0 : dc;
            const int e = kMcEdgeAt[axis * 8 + d0 * 4 + d1 * 2 + d2];
            const int mult = (int)((kMcVarMult[variant] >> (4 * e)) & 15ull);
            if (mult == 0) continue;
            // down-hill in-cube finite differences at the edge's two end corners, weighted like the position
            const double ga = v_lo - v_hi;
            const double gb_lo = HV_DV(n[0][1 - db][1]) - HV_DV(n[0][2 - db][1]);
            const double gb_hi = HV_DV(n[1][1 - db][1]) - HV_DV(n[1][2 - db][1]);
            const double gc_lo = HV_DV(n[0][1][1 - dc]) - HV_DV(n[0][1][2 - dc]);
            const double gc_hi = HV_DV(n[1][1][1 - dc]) - HV_DV(n[1][1][2 - dc]);
            const float fa = (float)__dadd_rn(__dmul_rn(ga, w_lo), __dmul_rn(ga, w_hi));
            const float fb = (float)__dadd_rn(__dmul_rn(gb_lo, w_lo), __dmul_rn(gb_hi, w_hi));
            const float fc = (float)__dadd_rn(__dmul_rn(gc_lo, w_lo), __dmul_rn(gc_hi, w_hi));
            for (int j = 0; j < mult; ++j) {
                acc_a = __fadd_rn(acc_a, fa);
                acc_b = __fadd_rn(acc_b, fb);
                acc_c = __fadd_rn(acc_c, fc);
            }
        }
    }
    float acc[3];
    acc[0] = axis == 0 ? acc_a : acc_b;                     // axis 0 is a for axis 0, else b
    acc[1] = axis == 1 ? acc_a : (axis == 0 ? acc_b : acc_c);  // axis 1: a for axis 1, b for axis 0, c for axis 2
    acc[2] = axis == 2 ? acc_a : acc_c;
    const long long zg = (long long)zb + g.zoff;
    double pos[3] = {(double)x, (double)y, (double)zg};
    if (axis == 0) pos[0] = __dadd_rn(pos[0], frac);
    else if (axis == 1) pos[1] = __dadd_rn(pos[1], frac);
    else pos[2] = __dadd_rn(pos[2], frac);
    const float pv[3] = {(float)pos[0], (float)pos[1], (float)pos[2]};
    const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn((double)acc[0], (double)acc[0]), __dmul_rn((double)acc[1], (double)acc[1])),
                                      __dmul_rn((double)acc[2], (double)acc[2])));
    if (len > 0.0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) acc[k] = (float)((double)acc[k] / len);
    }
    // colour of the nearest voxel (fusion.py:339,343-348), half-to-even like np.round
    const int i0 = (int)rintf(pv[0]), i1 = (int)rintf(pv[1]);
    const int i2 = (int)((long long)rintf(pv[2]) - g.zoff);
    const unsigned int packed = WHOLE ? (unsigned int)__ldg(g.color + ((size_t)i0 * g.Y + i1) * (size_t)g.pitch + i2)
                                      : (unsigned int)mc_fetch(g, g.color, g.chalo_lo, g.chalo_hi, i0, i1, i2);
#undef HV_DV
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        out.verts_world[3 * vi + k] = __fadd_rn(__fmul_rn(pv[k], out.voxel_size), out.origin[k]);  // fusion.py:340
        if (out.verts_voxel) out.verts_voxel[3 * vi + k] = pv[k];
        out.normals[3 * vi + k] = acc[k];
    }
    out.colors[3 * vi + 0] = (uint8_t)(packed & 255u);
    out.colors[3 * vi + 1] = (uint8_t)((packed >> 8) & 255u);
    out.colors[3 * vi + 2] = (uint8_t)((packed >> 16) & 255u);
    if (out.vert_keys) out.vert_keys[vi] = vertex_key(g, x, y, zb, axis);
}

// Adds `mult` times this cube's interpolated down-hill gradient at the centre vertex to the float32 accumulator.
__device__ __noinline__ void emit_centre_vertex(const McGrid& g, const McEmit& out, int x, int y, int zb, size_t vi) {
    double v[8];
    load_cube_values(g, x, y, zb, v);
    int idx = 0;
    for (int i = 0; i < 8; ++i) idx |= (v[i] > 0.0 ? 1 : 0) << i;
    const int variant = cube_variant(idx, v);
    const int mult = (int)((kMcVarMult[variant] >> 48) & 15ull);
    double w[8], f[3] = {0.0, 0.0, 0.0}, ff = 0.0;
    for (int i = 0; i < 8; ++i) {
        w[i] = 1.0 / ((double)FLT_EPSILON + fabs(v[i]));
        for (int k = 0; k < 3; ++k) f[k] = __dadd_rn(f[k], __dmul_rn((double)kMcCornerOff[3 * i + k], w[i]));
        ff = __dadd_rn(ff, w[i]);
    }
    const long long zg = (long long)zb + g.zoff;
    double pos[3] = {(double)x, (double)y, (double)zg};
    float acc[3] = {0.0f, 0.0f, 0.0f};
    for (int k = 0; k < 3; ++k) {
        pos[k] = __dadd_rn(pos[k], f[k] / ff);
        const int bitk = 4 >> k;  // offset bit of axis k in the d0*4 + d1*2 + d2 code
        double gk = 0.0;
        for (int i = 0; i < 8; ++i) {
            const int ci = kMcCornerOff[3 * i] * 4 + kMcCornerOff[3 * i + 1] * 2 + kMcCornerOff[3 * i + 2];
            const double gi = v[kMcCornerAt[ci & ~bitk]] - v[kMcCornerAt[ci | bitk]];
            gk = __dadd_rn(gk, __dmul_rn(w[i], gi));
        }
        const float gf = (float)gk;
        for (int j = 0; j < mult; ++j) acc[k] = __fadd_rn(acc[k], gf);
    }
    const float pv[3] = {(float)pos[0], (float)pos[1], (float)pos[2]};
    const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn((double)acc[0], (double)acc[0]), __dmul_rn((double)acc[1], (double)acc[1])),
                                      __dmul_rn((double)acc[2], (double)acc[2])));
    if (len > 0.0)
        for (int k = 0; k < 3; ++k) acc[k] = (float)((double)acc[k] / len);
    const int i0 = (int)rintf(pv[0]), i1 = (int)rintf(pv[1]);
    const int i2 = (int)((long long)rintf(pv[2]) - g.zoff);
    const unsigned int packed = (unsigned int)mc_fetch(g, g.color, g.chalo_lo, g.chalo_hi, i0, i1, i2);
    for (int k = 0; k < 3; ++k) {
        out.verts_world[3 * vi + k] = __fadd_rn(__fmul_rn(pv[k], out.voxel_size), out.origin[k]);
        if (out.verts_voxel) out.verts_voxel[3 * vi + k] = pv[k];
        out.normals[3 * vi + k] = acc[k];
    }
    out.colors[3 * vi + 0] = (uint8_t)(packed & 255u);
    out.colors[3 * vi + 1] = (uint8_t)((packed >> 8) & 255u);
    out.colors[3 * vi + 2] = (uint8_t)((packed >> 16) & 255u);
    if (out.vert_keys) out.vert_keys[vi] = vertex_key(g, x, y, zb, 3);
}

constexpr int kEmitThreads = 256;

// ---- 3. triangles (+ cube-centre vertices): one thread per cube record ---------------------------------------------------------
// The vertex indices of the cube's (up to 12) edges are looked up first, all loads independent; its triangles then pick
// from that array.  (One thread walking its triangles with three dependent look-ups each was measured at 66 us.)
template <bool DIRECT>
__device__ __forceinline__ void emit_cubes(const McGrid& g, const McEmit& out, unsigned int block, unsigned int n_blocks) {
    __shared__ unsigned int s_vidx[13][kEmitThreads];  // [edge][thread]: dynamic indexing without local memory
    const unsigned int n_cubes = (unsigned int)g.totals[5], n_edge = (unsigned int)g.totals[4];
    for (unsigned int i = block * kEmitThreads + threadIdx.x; i < n_cubes; i += n_blocks * kEmitThreads) {
        const uint4 rec = __ldg(g.cube_rec + i);
        int x, y, gq;
        decode_group(g, rec.x, x, y, gq);
        const size_t tile = rec.x / kTileGroups;
        const uint4 tb = __ldg(g.tile_base + tile), bb = __ldg(g.batch_base + tile * kTileBatches + (rec.y >> 16));
        const uint4 base = make_uint4(tb.x + bb.x, tb.y + bb.y, tb.z + bb.z, 0u);
        const int k = (int)(rec.y & 31u), variant = (int)((rec.y >> 5) & 2047u), zb = gq * 32 + k;
        const int ntri = kMcVarNtri[variant];
        const unsigned long long mult = kMcVarMult[variant];
        const unsigned int centre_index = n_edge + base.z + rec.w;
        const size_t t0 = (size_t)base.y + rec.z;
        const size_t col = (size_t)x * g.Y + y;
        unsigned int centre_global = centre_index;
        if (out.vshift) centre_global = out.n_edge_global + (base.z + rec.w) + (unsigned int)__ldg(out.cshift + col);
        if (out.face_keys) {
            for (int j = 0; j < ntri; ++j) emit_triangle(g, out, x, y, zb, variant, j, t0 + j, centre_index);
        } else {
            unsigned int vidx[12];
#pragma unroll
            for (int e = 0; e < 12; ++e) {
                constexpr unsigned long long P = kEdgeOwnerPacked;
                const int code = (int)((P >> (5 * e)) & 31ull);
                vidx[e] = 0u;
                if ((mult >> (4 * e)) & 15ull) {
                    const int xo = x + (code & 1), yo = y + ((code >> 1) & 1), zo = zb + ((code >> 2) & 1), axis = code >> 3;
                    if (!out.vshift) {
                        vidx[e] = edge_vertex_index(g, xo, yo, zo, axis);
                    } else {
                        const size_t ocol = (size_t)xo * g.Y + yo;
                        if (zo < g.own_hi) {
                            vidx[e] = edge_vertex_index(g, xo, yo, zo, axis) + (unsigned int)__ldg(out.vshift + ocol);
                        } else {
                            // a seam vertex: on an axis-0 / axis-1 edge of the first voxel the slab above owns in this
                            // column, i.e. that slab's first (axis 0) or first / second (axis 1) vertex of the column
                            unsigned int r = (unsigned int)__ldg(out.vnext + ocol);
                            if (axis == 1 && xo + 1 < g.X) {
                                const unsigned int b0 = __ldg(g.bits + ocol * g.G + (zo >> 5)) >> (zo & 31);
                                const unsigned int b1 = __ldg(g.bits + (ocol + g.Y) * g.G + (zo >> 5)) >> (zo & 31);
                                r += (b0 ^ b1) & 1u;
                            }
                            vidx[e] = r;
                        }
                    }
                }
            }
#pragma unroll
            for (int e = 0; e < 12; ++e) s_vidx[e][threadIdx.x] = vidx[e];
            s_vidx[12][threadIdx.x] = centre_global;
            const uint8_t* tri = kMcTris + 3 * (size_t)kMcVarTriOff[variant];
            for (int j = 0; j < ntri; ++j) {
                const size_t ti = t0 + j;
                if ((long long)ti >= out.cap_tris) break;
                const unsigned int a = s_vidx[tri[3 * j]][threadIdx.x], b = s_vidx[tri[3 * j + 1]][threadIdx.x],
                                   c = s_vidx[tri[3 * j + 2]][threadIdx.x];
                size_t row = ti;
                if (DIRECT) {
                    row = ti + (size_t)__ldg(out.tshift + col);
                    if ((long long)row >= out.gcap_tris) break;
                }
                out.faces[3 * row] = (int32_t)a;
                out.faces[3 * row + 1] = (int32_t)b;
                out.faces[3 * row + 2] = (int32_t)c;
                if (!DIRECT && out.tri_dst) out.tri_dst[ti] = (int32_t)(ti + (size_t)__ldg(out.tshift + col));
            }
        }
        if (((mult >> 48) & 15ull) && (long long)centre_index < out.cap_verts) {
            if (!DIRECT) {
                emit_centre_vertex(g, out, x, y, zb, centre_index);
                if (out.vert_dst) out.vert_dst[centre_index] = (int32_t)centre_global;
            } else if ((long long)centre_global < out.gcap_verts) {
                emit_centre_vertex(g, out, x, y, zb, centre_global);
            }
        }
    }
}

// ---- 4. edge vertices: one thread per vertex record ----------------------------------------------------------------------------------
template <bool WHOLE, bool LEVEL0, bool DIRECT>
__device__ __forceinline__ void emit_vertices(const McGrid& g, const McEmit& out, unsigned int block, unsigned int n_blocks) {
    const unsigned int n_edge = (unsigned int)g.totals[4];
    for (unsigned int i = block * kEmitThreads + threadIdx.x; i < n_edge; i += n_blocks * kEmitThreads) {
        const unsigned long long rec = __ldg(g.vert_rec + i);
        const unsigned int gid = (unsigned int)(rec & 0x7fffffffull);
        const int k = (int)((rec >> 31) & 31ull), axis = (int)((rec >> 36) & 3ull);
        int x, y, gq;
        decode_group(g, gid, x, y, gq);
        const size_t tile = gid / kTileGroups;
        const size_t vi = (size_t)__ldg(&g.tile_base[tile].x) + __ldg(&g.batch_base[tile * kTileBatches + (unsigned int)((rec >> 50) & 31ull)].x) +
                          (unsigned int)((rec >> 38) & 4095ull);
        if ((long long)vi < out.cap_verts) {
            if (!DIRECT) {
                emit_edge_vertex<WHOLE, LEVEL0>(g, out, x, y, gq * 32 + k, axis, vi);
                if (out.vert_dst) out.vert_dst[vi] = (int32_t)(vi + (size_t)__ldg(out.vshift + (size_t)x * g.Y + y));
            } else {
                const size_t row = vi + (size_t)__ldg(out.vshift + (size_t)x * g.Y + y);
                if ((long long)row < out.gcap_verts) emit_edge_vertex<WHOLE, LEVEL0>(g, out, x, y, gq * 32 + k, axis, row);
            }
        }
    }
}

// One launch for both: triangles and vertices do not depend on each other and stall on different things (index look-ups in
// L2 / arithmetic), so their blocks are interleaved in the grid in the ratio n_tri_blocks : n_vert_blocks.  (Two launches:
// 33 + 40 us; the triangle kernel on a forked side stream gained 6 us of that -- the event hand-over costs what it hides.)
template <bool WHOLE, bool LEVEL0, bool DIRECT>
__global__ void __launch_bounds__(kEmitThreads, (WHOLE && LEVEL0) ? 4 : 3)
    mc_emit_kernel(const __grid_constant__ McGrid g, const __grid_constant__ McEmit out, const unsigned int n_tri_blocks,
                   const unsigned int n_vert_blocks) {
    cudaGridDependencySynchronize();
    if (g.totals[6]) return;  // the records are incomplete: the host re-runs the count with a larger work buffer
    // (32-bit: the host keeps the grid below 2^16 blocks, so b * n_tri_blocks < 2^32; the 64-bit divisions were 7 % of the
    //  kernel's instructions)
    const unsigned int b = blockIdx.x, n = n_tri_blocks + n_vert_blocks;
    const unsigned int t_before = b * n_tri_blocks / n, t_after = (b + 1u) * n_tri_blocks / n;
    if (t_after > t_before) emit_cubes<DIRECT>(g, out, t_before, n_tri_blocks);
    else emit_vertices<WHOLE, LEVEL0, DIRECT>(g, out, b - t_before, n_vert_blocks);
}

// ---- z-slab placement (hive_b200/sharded.py): where does a slab's mesh go in the mesh of the whole volume? --------------------
// The whole-volume order is column-major, so the slabs interleave column by column: every slab counts its edge vertices,
// triangles and centre vertices per column (from its work records), the counts of all slabs are all-gathered, and every
// slab derives its per-column shifts itself.  Nothing is sorted and nothing is searched.
__global__ void __launch_bounds__(256) mc_column_counts_kernel(const __grid_constant__ McGrid g, int32_t* __restrict__ counts,
                                                                const long long XY) {
    if (g.totals[6]) return;
    const unsigned int n_edge = (unsigned int)g.totals[4], n_cubes = (unsigned int)g.totals[5];
    const unsigned int n = n_edge > n_cubes ? n_edge : n_cubes;
    for (unsigned int i = blockIdx.x * 256u + threadIdx.x; i < n; i += gridDim.x * 256u) {
        if (i < n_edge) {
            const unsigned int gid = (unsigned int)(__ldg(g.vert_rec + i) & 0x7fffffffull);
            atomicAdd(counts + gid / (unsigned int)g.G, 1);
        }
        if (i < n_cubes) {
            const uint4 rec = __ldg(g.cube_rec + i);
            const int variant = (int)((rec.y >> 5) & 2047u);
            const unsigned int col = rec.x / (unsigned int)g.G;
            atomicAdd(counts + XY + col, (int)kMcVarNtri[variant]);
            if ((kMcVarMult[variant] >> 48) & 15ull) atomicAdd(counts + 2 * XY + col, 1);
        }
    }
}

// all[w * slab_stride + k * XY + c]: count k (0 edge vertices, 1 triangles, 2 centre vertices) of slab w in column c
// (PEERS: slab w's counts are read where they are -- in the memory of the GPU that holds slab w, mapped into this process
// (hv_peer_open); the loads travel over NVLink and nothing is gathered first.  The slabs' tail words are copied out for the host.)
constexpr int kMaxSlabPtrs = 16;
struct SlabCounts {
    const int32_t* base[kMaxSlabPtrs];
};

template <bool PEERS>
__global__ void __launch_bounds__(256) mc_shift_prepare_kernel(const int32_t* __restrict__ all, const long long slab_stride,
                                                                const SlabCounts peers, const int W, const int rank,
                                                                const long long XY, uint2* __restrict__ pa, uint2* __restrict__ pb,
                                                                uint2* __restrict__ pc, int32_t* __restrict__ below,
                                                                int32_t* __restrict__ tails, const int tail_words) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    if (PEERS && blockIdx.x == 0 && (int)threadIdx.x < W * tail_words) {
        const int w = (int)threadIdx.x / tail_words, j = (int)threadIdx.x - w * tail_words;
        tails[threadIdx.x] = peers.base[w][3 * XY + j];
    }
    if (c >= XY) return;
    unsigned int tot[3] = {0u, 0u, 0u}, low[3] = {0u, 0u, 0u}, own[3] = {0u, 0u, 0u};
    for (int w = 0; w < W; ++w)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const unsigned int v = (unsigned int)(PEERS ? peers.base[w][k * XY + c] : __ldg(all + (long long)w * slab_stride + k * XY + c));
            tot[k] += v;
            if (w < rank) low[k] += v;
            if (w == rank) own[k] = v;
        }
    pa[c] = make_uint2(tot[0], tot[1]);
    pb[c] = make_uint2(tot[2], own[0]);
    pc[c] = make_uint2(own[1], own[2]);
#pragma unroll
    for (int k = 0; k < 3; ++k) below[k * XY + c] = (int32_t)low[k];
    below[3 * XY + c] = (int32_t)own[0];
}

// sa / sb / sc: exclusive scans of pa / pb / pc.  shifts[0..3][c] = vshift, tshift, cshift, vnext (holds `below` on entry).
__global__ void __launch_bounds__(256) mc_shift_finish_kernel(const uint2* __restrict__ sa, const uint2* __restrict__ sb,
                                                               const uint2* __restrict__ sc, const long long XY,
                                                               int32_t* __restrict__ shifts) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    if (c >= XY) return;
    const uint2 a = sa[c], b = sb[c], cc = sc[c];
    const unsigned int low_v = (unsigned int)shifts[c], low_t = (unsigned int)shifts[XY + c], low_c = (unsigned int)shifts[2 * XY + c];
    const unsigned int own_v = (unsigned int)shifts[3 * XY + c];
    shifts[c] = (int32_t)(a.x + low_v - b.y);            // global start of my vertices in the column - my own start
    shifts[XY + c] = (int32_t)(a.y + low_t - cc.x);
    shifts[2 * XY + c] = (int32_t)(b.x + low_c - cc.y);
    shifts[3 * XY + c] = (int32_t)(a.x + low_v + own_v);  // first vertex of the slab above in this column
}

// dst[index[i]] = src[i] for rows of `words` 32-bit words (or of 3 bytes when words == 0)
__global__ void __launch_bounds__(256) scatter_rows_kernel(const void* __restrict__ src, const int32_t* __restrict__ index,
                                                            const long long n, const int words, void* __restrict__ dst) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (words > 0) {
        const long long row = i / words;
        const int k = (int)(i - row * words);
        if (row < n) static_cast<unsigned int*>(dst)[(long long)index[row] * words + k] = static_cast<const unsigned int*>(src)[i];
    } else if (i < n) {
        const unsigned char* s3 = static_cast<const unsigned char*>(src) + 3 * i;
        unsigned char* d3 = static_cast<unsigned char*>(dst) + 3 * (long long)index[i];
        d3[0] = s3[0];
        d3[1] = s3[1];
        d3[2] = s3[2];
    }
}

}  // namespace hv

// ---- C ABI ---------------------------------------------------------------------------------------------------------------
namespace {

struct McLayout {
    long long n_groups, n_batches, n_tiles;
    int G, Zb;
    size_t off_bits, off_gtab, off_glist, off_tile_cnt, off_work, off_cvar, off_minmax, off_batch_tot, off_batch_base, off_tile_tot, off_tile_base, off_ctrl, total;
};

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }
constexpr int kMaxSignBlocks = 8192;

// Launch as a programmatic dependent of the kernel before it in the stream: the grid may be set up while its predecessor
// drains; the kernel itself waits (cudaGridDependencySynchronize, its first statement) before it reads anything.
template <typename... KArgs, typename... Args>
cudaError_t launch_dependent(void (*kernel)(KArgs...), unsigned int grid, unsigned int block, cudaStream_t st, Args... args) {
    static const bool plain = getenv("HV_MC_NO_PDL") != nullptr;  // A/B switch
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = plain ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

int mc_layout(const int64_t* h_dims, bool lo, bool hi, McLayout* L) {
    HV_REQUIRE(h_dims, "marching cubes: null dims");
    const int64_t X = h_dims[0], Y = h_dims[1], Z = h_dims[2];
    HV_REQUIRE(X >= 1 && Y >= 1 && Z >= 1 && X < (1ll << 30) && Y < (1ll << 30) && Z < (1ll << 30),
               "marching cubes: bad dims %lld x %lld x %lld", (long long)X, (long long)Y, (long long)Z);
    L->Zb = (int)Z + (lo ? 1 : 0) + (hi ? 1 : 0);
    L->G = (L->Zb + 31) / 32;
    L->n_groups = (long long)X * Y * L->G;
    HV_REQUIRE(L->n_groups < (1ll << 31) && L->Zb < (1 << 22), "marching cubes: volume too large");
    L->n_tiles = (L->n_groups + hv::kTileGroups - 1) / hv::kTileGroups;
    L->n_batches = L->n_tiles * hv::kTileBatches;
    size_t o = 0;
    L->off_gtab = o;  // 16-byte elements first
    o = align_up(o + (size_t)L->n_groups * sizeof(uint4));
    L->off_batch_base = o;
    o = align_up(o + (size_t)L->n_batches * sizeof(uint4));
    L->off_batch_tot = o;
    o = align_up(o + (size_t)L->n_batches * sizeof(uint4));
    L->off_tile_tot = o;
    o = align_up(o + (size_t)L->n_tiles * sizeof(uint4));
    L->off_tile_base = o;
    o = align_up(o + (size_t)L->n_tiles * sizeof(uint4));
    L->off_bits = o;
    o = align_up(o + (size_t)L->n_groups * sizeof(unsigned int));
    L->off_glist = o;
    o = align_up(o + (size_t)L->n_tiles * hv::kTileGroups * sizeof(unsigned int));
    L->off_tile_cnt = o;
    o = align_up(o + (size_t)L->n_tiles * sizeof(unsigned int));
    L->off_work = o;
    o = align_up(o + (size_t)L->n_batches * sizeof(unsigned int));
    L->off_cvar = o;
    o = align_up(o + (size_t)L->n_groups * 32 * sizeof(unsigned short));
    L->off_minmax = o;
    o = align_up(o + (size_t)kMaxSignBlocks * sizeof(int2));
    L->off_ctrl = o;
    o = align_up(o + 8 * sizeof(unsigned int));
    L->total = o;
    return HV_OK;
}

size_t work_bytes(int64_t rec_verts, int64_t rec_cubes) {
    return align_up((size_t)rec_cubes * sizeof(uint4)) + align_up((size_t)rec_verts * sizeof(unsigned long long));
}

int mc_grid(hv::McGrid* g, const float* tsdf, const float* halo_lo, const float* halo_hi, const int64_t* h_dims,
            int64_t pitch, double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts,
            int64_t rec_cubes, McLayout* L) {
    int rc = mc_layout(h_dims, halo_lo != nullptr, halo_hi != nullptr, L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(tsdf && scratch, "marching cubes: null pointer");
    HV_REQUIRE(pitch >= h_dims[2], "marching cubes: pitch %lld < Z", (long long)pitch);
    HV_REQUIRE(((uintptr_t)scratch & 255) == 0, "marching cubes: scratch must be 256-byte aligned");
    HV_REQUIRE(rec_verts >= 0 && rec_cubes >= 0 && rec_verts < (1ll << 31) && rec_cubes < (1ll << 31) &&
                   (work != nullptr || (rec_verts == 0 && rec_cubes == 0)) && ((uintptr_t)work & 255) == 0,
               "marching cubes: bad work buffer");
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
    g->gtab = reinterpret_cast<uint4*>(s + L->off_gtab);
    g->glist = reinterpret_cast<unsigned int*>(s + L->off_glist);
    g->tile_cnt = reinterpret_cast<unsigned int*>(s + L->off_tile_cnt);
    g->work = reinterpret_cast<unsigned int*>(s + L->off_work);
    g->cvar = reinterpret_cast<unsigned short*>(s + L->off_cvar);
    g->sign_minmax = reinterpret_cast<int2*>(s + L->off_minmax);
    g->sign_blocks = 0;
    g->batch_tot = reinterpret_cast<uint4*>(s + L->off_batch_tot);
    g->batch_base = reinterpret_cast<uint4*>(s + L->off_batch_base);
    g->tile_tot = reinterpret_cast<uint4*>(s + L->off_tile_tot);
    g->tile_base = reinterpret_cast<uint4*>(s + L->off_tile_base);
    g->ctrl = reinterpret_cast<unsigned int*>(s + L->off_ctrl);
    g->cube_rec = reinterpret_cast<uint4*>(work);
    g->vert_rec = reinterpret_cast<unsigned long long*>(static_cast<char*>(work) + align_up((size_t)rec_cubes * sizeof(uint4)));
    g->cap_cubes = rec_cubes;
    g->cap_vrec = rec_verts;
    g->totals = nullptr;
    g->n_groups = L->n_groups;
    g->n_batches = L->n_batches;
    g->n_tiles = (int)L->n_tiles;
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

extern "C" int hv_mc_scratch_init(const int64_t* h_dims, int has_halo_lo, int has_halo_hi, void* scratch, int64_t scratch_bytes,
                                   void* stream) {
    McLayout L;
    int rc = mc_layout(h_dims, has_halo_lo != 0, has_halo_hi != 0, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(scratch && ((uintptr_t)scratch & 255) == 0 && (size_t)scratch_bytes >= L.total, "hv_mc_scratch_init: bad scratch");
    HV_CUDA(cudaMemsetAsync(static_cast<char*>(scratch) + L.off_ctrl, 0, 8 * sizeof(unsigned int), hv::as_stream(stream)));
    return HV_OK;
}

extern "C" int hv_mc_work_bytes(int64_t rec_verts, int64_t rec_cubes, int64_t* h_bytes) {
    HV_REQUIRE(h_bytes && rec_verts >= 0 && rec_cubes >= 0 && rec_verts < (1ll << 31) && rec_cubes < (1ll << 31),
               "hv_mc_work_bytes: bad arguments");
    *h_bytes = (int64_t)work_bytes(rec_verts, rec_cubes);
    return HV_OK;
}

extern "C" int hv_mc_count(const float* tsdf, const float* halo_lo, const float* halo_hi, const int64_t* h_dims,
                           int64_t pitch, double level, void* scratch, int64_t scratch_bytes, void* work,
                           int64_t rec_verts, int64_t rec_cubes, unsigned long long* totals, void* stream) {
    hv::McGrid g;
    McLayout L;
    int rc = mc_grid(&g, tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, work, rec_verts, rec_cubes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(totals, "hv_mc_count: null totals");
    g.totals = totals;
    cudaStream_t st = hv::as_stream(stream);
    const int sms = hv::device_info().sm_count;
    static int sign_res = 0;  // resident CTAs per SM of the sign kernel: the grid is exactly one wave
    if (sign_res == 0 &&
        (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&sign_res, hv::mc_sign_kernel, 256, 0) != cudaSuccess || sign_res < 1))
        sign_res = 4;
    const long long want = ((long long)g.X * g.Y + 7) / 8;  // 8 warps per block, a warp per column (grid-stride)
    const long long cap = (long long)sms * sign_res;
    int sign_blocks = (int)(want < cap ? (want > 0 ? want : 1) : cap);
    if (sign_blocks > kMaxSignBlocks) sign_blocks = kMaxSignBlocks;
    g.sign_blocks = sign_blocks;
    hv::mc_sign_kernel<<<sign_blocks, 256, 0, st>>>(g);
    HV_CUDA(launch_dependent(hv::mc_detect_kernel, (unsigned int)L.n_tiles, hv::kDetectThreads, st, g));
    {
        // one warp per batch of 32 listed groups; how many there are is only known on the device
        static int count_res = 0;
        if (count_res == 0 && (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&count_res, hv::mc_count_kernel, hv::kCountThreads, 0) !=
                                   cudaSuccess || count_res < 1))
            count_res = 4;
        const long long want = (L.n_batches + hv::kCountThreads / 32 - 1) / (hv::kCountThreads / 32);
        const long long cap = (long long)sms * count_res;
        HV_CUDA(launch_dependent(hv::mc_count_kernel, (unsigned int)(want < cap ? want : cap), hv::kCountThreads, st, g));
    }
    HV_CUDA(launch_dependent(hv::mc_scan_kernel, (unsigned int)((L.n_tiles + hv::kScanThreads2 / 32 - 1) / (hv::kScanThreads2 / 32)),
                             hv::kScanThreads2, st, g));
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

namespace {
struct Placement {
    const int32_t* col_shifts;  // [4][X*Y] from hv_mc_column_shifts
    int64_t n_edge_global;
    int32_t* vert_dst;
    int32_t* tri_dst;
    int64_t whole_verts, whole_tris;  // > 0: direct placement, the output arrays are the whole mesh's, of this many rows
};
}  // namespace

static int mc_emit_impl(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                        const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                        double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts,
                        int64_t rec_cubes, int64_t z_global_offset, int64_t z_global_dim, float voxel_size,
                        const float* h_origin, int64_t n_verts, int64_t n_tris, float* verts_world, float* verts_voxel,
                        float* normals, uint8_t* colors, int32_t* faces, unsigned long long* face_keys,
                        unsigned long long* face_cubes, unsigned long long* vert_keys, unsigned long long* totals,
                        const Placement* place, void* stream) {
    hv::McGrid g;
    McLayout L;
    int rc = mc_grid(&g, tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, work, rec_verts, rec_cubes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(n_verts >= 0 && n_tris >= 0 && n_verts < (1ll << 31) && n_tris < (1ll << 31), "hv_mc_emit: bad counts");
    if (n_verts == 0 && n_tris == 0) return HV_OK;  // a shard the surface does not cross / no capacity at all
    HV_REQUIRE(color && h_origin && verts_world && normals && colors && totals, "hv_mc_emit: null pointer");
    HV_REQUIRE((faces != nullptr) != (face_keys != nullptr), "hv_mc_emit: pass exactly one of faces / face_keys");
    HV_REQUIRE((halo_lo != nullptr) == (color_halo_lo != nullptr) && (halo_hi != nullptr) == (color_halo_hi != nullptr),
               "hv_mc_emit: tsdf and colour halos must come in pairs");
    HV_REQUIRE(faces == nullptr || place != nullptr || (!halo_lo && !halo_hi),
               "hv_mc_emit: a shard with halos must emit face_keys (or be placed: hv_mc_emit_placed)");
    const bool direct = place && place->whole_verts > 0;
    HV_REQUIRE(place == nullptr || (faces && place->col_shifts && place->n_edge_global >= 0 && place->n_edge_global < (1ll << 31)),
               "hv_mc_emit_placed: bad placement");
    HV_REQUIRE(place == nullptr || (direct ? (!place->vert_dst && !place->tri_dst && place->whole_verts >= n_verts &&
                                              place->whole_tris >= n_tris && place->whole_verts < (1ll << 31) &&
                                              place->whole_tris < (1ll << 31))
                                           : (place->vert_dst && place->tri_dst)),
               "hv_mc_emit_placed: bad placement (row indices or whole-mesh sizes)");
    HV_REQUIRE(z_global_dim >= z_global_offset + g.Z, "hv_mc_emit: slab exceeds the global z extent");
    g.color = color;
    g.chalo_lo = color_halo_lo;
    g.chalo_hi = color_halo_hi;
    g.zoff = z_global_offset - g.hl;
    g.Zg = z_global_dim;
    g.totals = totals;
    cudaStream_t st = hv::as_stream(stream);
    hv::McEmit em;
    em.faces = faces;
    em.face_keys = face_keys;
    em.vert_keys = vert_keys;
    em.face_cubes = face_keys ? face_cubes : nullptr;
    em.verts_world = verts_world;
    em.verts_voxel = verts_voxel;
    em.normals = normals;
    em.colors = colors;
    em.cap_verts = n_verts;
    em.cap_tris = n_tris;
    em.voxel_size = voxel_size;
    for (int k = 0; k < 3; ++k) em.origin[k] = h_origin[k];
    em.vshift = em.tshift = em.cshift = em.vnext = nullptr;
    em.n_edge_global = 0u;
    em.vert_dst = em.tri_dst = nullptr;
    em.gcap_verts = em.gcap_tris = 0;
    if (place) {
        const long long XY = (long long)g.X * g.Y;
        em.vshift = place->col_shifts;
        em.tshift = place->col_shifts + XY;
        em.cshift = place->col_shifts + 2 * XY;
        em.vnext = place->col_shifts + 3 * XY;
        em.n_edge_global = (unsigned int)place->n_edge_global;
        em.vert_dst = place->vert_dst;
        em.tri_dst = place->tri_dst;
        em.gcap_verts = place->whole_verts;
        em.gcap_tris = place->whole_tris;
    }
    const int sms = hv::device_info().sm_count;
    // one thread per record that can exist: a surface cube has a triangle, an edge vertex is a vertex
    const long long max_cubes = rec_cubes < n_tris ? rec_cubes : n_tris, max_verts = rec_verts < n_verts ? rec_verts : n_verts;
    // (grid-stride: a capacity far above the surface would otherwise launch tens of thousands of CTAs that only read the totals)
    const long long grid_cap = (long long)sms * 48;
    const long long tb = max_cubes > 0 ? (max_cubes + hv::kEmitThreads - 1) / hv::kEmitThreads : 0;
    const long long vb = max_verts > 0 ? (max_verts + hv::kEmitThreads - 1) / hv::kEmitThreads : 0;
    // (both roles grid-stride, so the block counts can be scaled down together when the capacity is far above the surface)
    long long nt = tb, nv = vb;
    if (nt + nv > grid_cap) {
        nt = tb * grid_cap / (tb + vb);
        nv = vb * grid_cap / (tb + vb);
        if (tb > 0 && nt == 0) nt = 1;
        if (vb > 0 && nv == 0) nv = 1;
    }
    if (nt + nv > 0) {
        const unsigned int grid = (unsigned int)(nt + nv);
        const bool whole = !halo_lo && !halo_hi, level0 = level == 0.0;
        // (a dependent launch is harmless when something else precedes it in the stream: it then just waits for that)
        const unsigned int unt = (unsigned int)nt, unv = (unsigned int)nv;
        if (direct && level0) HV_CUDA(launch_dependent(hv::mc_emit_kernel<false, true, true>, grid, hv::kEmitThreads, st, g, em, unt, unv));
        else if (direct) HV_CUDA(launch_dependent(hv::mc_emit_kernel<false, false, true>, grid, hv::kEmitThreads, st, g, em, unt, unv));
        else if (whole && level0) HV_CUDA(launch_dependent(hv::mc_emit_kernel<true, true, false>, grid, hv::kEmitThreads, st, g, em, unt, unv));
        else if (whole) HV_CUDA(launch_dependent(hv::mc_emit_kernel<true, false, false>, grid, hv::kEmitThreads, st, g, em, unt, unv));
        else if (level0) HV_CUDA(launch_dependent(hv::mc_emit_kernel<false, true, false>, grid, hv::kEmitThreads, st, g, em, unt, unv));
        else HV_CUDA(launch_dependent(hv::mc_emit_kernel<false, false, false>, grid, hv::kEmitThreads, st, g, em, unt, unv));
    }
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_mc_emit(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                          const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                          double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts,
                          int64_t rec_cubes, int64_t z_global_offset, int64_t z_global_dim, float voxel_size,
                          const float* h_origin, int64_t n_verts, int64_t n_tris, float* verts_world, float* verts_voxel,
                          float* normals, uint8_t* colors, int32_t* faces, unsigned long long* face_keys,
                          unsigned long long* face_cubes, unsigned long long* vert_keys, unsigned long long* totals,
                          void* stream) {
    return mc_emit_impl(tsdf, halo_lo, halo_hi, color, color_halo_lo, color_halo_hi, h_dims, pitch, level, scratch, scratch_bytes,
                        work, rec_verts, rec_cubes, z_global_offset, z_global_dim, voxel_size, h_origin, n_verts, n_tris,
                        verts_world, verts_voxel, normals, colors, faces, face_keys, face_cubes, vert_keys, totals, nullptr, stream);
}

extern "C" int hv_mc_emit_placed(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                                 const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                                 double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts,
                                 int64_t rec_cubes, int64_t z_global_offset, int64_t z_global_dim, float voxel_size,
                                 const float* h_origin, int64_t n_verts, int64_t n_tris, float* verts_world, float* normals,
                                 uint8_t* colors, int32_t* faces, const int32_t* col_shifts, int64_t n_edge_global,
                                 int32_t* vert_dst, int32_t* tri_dst, unsigned long long* totals, void* stream) {
    Placement pl = {col_shifts, n_edge_global, vert_dst, tri_dst, 0, 0};
    return mc_emit_impl(tsdf, halo_lo, halo_hi, color, color_halo_lo, color_halo_hi, h_dims, pitch, level, scratch, scratch_bytes,
                        work, rec_verts, rec_cubes, z_global_offset, z_global_dim, voxel_size, h_origin, n_verts, n_tris,
                        verts_world, nullptr, normals, colors, faces, nullptr, nullptr, nullptr, totals, &pl, stream);
}

// The slab's rows written AT their whole-mesh positions: verts_world / normals / colors (whole_verts rows) and faces
// (whole_tris rows) are the whole mesh's arrays -- on this GPU, or on another one whose memory is mapped into this
// process (hv_peer_open): the stores then travel over NVLink and nothing is gathered or scattered afterwards.
extern "C" int hv_mc_emit_direct(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                                 const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                                 double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts,
                                 int64_t rec_cubes, int64_t z_global_offset, int64_t z_global_dim, float voxel_size,
                                 const float* h_origin, int64_t n_verts, int64_t n_tris, const int32_t* col_shifts,
                                 int64_t n_edge_global, int64_t whole_verts, int64_t whole_tris, float* verts_world,
                                 float* normals, uint8_t* colors, int32_t* faces, unsigned long long* totals, void* stream) {
    HV_REQUIRE(whole_verts > 0 && whole_tris >= 0, "hv_mc_emit_direct: bad whole-mesh sizes");
    Placement pl = {col_shifts, n_edge_global, nullptr, nullptr, whole_verts, whole_tris};
    return mc_emit_impl(tsdf, halo_lo, halo_hi, color, color_halo_lo, color_halo_hi, h_dims, pitch, level, scratch, scratch_bytes,
                        work, rec_verts, rec_cubes, z_global_offset, z_global_dim, voxel_size, h_origin, n_verts, n_tris,
                        verts_world, nullptr, normals, colors, faces, nullptr, nullptr, nullptr, totals, &pl, stream);
}

extern "C" int hv_mc_column_counts(const float* tsdf, const float* halo_lo, const float* halo_hi, const int64_t* h_dims,
                                   int64_t pitch, double level, void* scratch, int64_t scratch_bytes, void* work,
                                   int64_t rec_verts, int64_t rec_cubes, unsigned long long* totals, int32_t* counts,
                                   void* stream) {
    hv::McGrid g;
    McLayout L;
    int rc = mc_grid(&g, tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, work, rec_verts, rec_cubes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(totals && counts, "hv_mc_column_counts: null pointer");
    g.totals = totals;
    cudaStream_t st = hv::as_stream(stream);
    const long long XY = (long long)g.X * g.Y;
    HV_CUDA(cudaMemsetAsync(counts, 0, (size_t)3 * XY * sizeof(int32_t), st));
    const long long n = rec_verts > rec_cubes ? rec_verts : rec_cubes;
    if (n > 0) {
        const long long want = (n + 255) / 256, cap = (long long)hv::device_info().sm_count * 16;
        hv::mc_column_counts_kernel<<<(unsigned int)(want < cap ? want : cap), 256, 0, st>>>(g, counts, XY);
    }
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

namespace {
struct ShiftLayout {
    size_t off_pairs[6], off_bsum, total;
};
void shift_layout(int64_t XY, ShiftLayout* L) {
    size_t o = 0;
    for (int i = 0; i < 6; ++i) {
        L->off_pairs[i] = o;
        o = align_up(o + (size_t)XY * sizeof(uint2));
    }
    L->off_bsum = o;
    o = align_up(o + (size_t)hv::scan_scratch_pairs(XY) * sizeof(uint2));
    L->total = o;
}
}  // namespace

extern "C" int hv_mc_shift_scratch_bytes(int64_t n_columns, int64_t* h_bytes) {
    HV_REQUIRE(h_bytes && n_columns > 0 && n_columns < (1ll << 31), "hv_mc_shift_scratch_bytes: bad arguments");
    ShiftLayout L;
    shift_layout(n_columns, &L);
    *h_bytes = (int64_t)L.total;
    return HV_OK;
}

static int mc_column_shifts_impl(const int32_t* all_counts, int64_t slab_stride, const void* const* h_slab_ptrs, int32_t* tails,
                                 int tail_words, int world_size, int rank, int64_t n_columns, int32_t* shifts,
                                 unsigned long long* g_totals, void* scratch, int64_t scratch_bytes, void* stream);

extern "C" int hv_mc_column_shifts(const int32_t* all_counts, int64_t slab_stride, int world_size, int rank, int64_t n_columns, int32_t* shifts,
                                   unsigned long long* g_totals, void* scratch, int64_t scratch_bytes, void* stream) {
    HV_REQUIRE(all_counts && slab_stride >= 3 * n_columns, "hv_mc_column_shifts: bad arguments");
    return mc_column_shifts_impl(all_counts, slab_stride, nullptr, nullptr, 0, world_size, rank, n_columns, shifts, g_totals, scratch,
                                 scratch_bytes, stream);
}

// hv_mc_column_shifts reading every slab's counts where they are: h_slab_ptrs (HOST array of world_size DEVICE addresses, at
// most 16) -- slab w's [3][X*Y] counts followed by tail_words words of per-slab data -- typically the memory of the GPU holding
// slab w, mapped with hv_peer_open.  `tails` (DEVICE, [world_size][tail_words]) receives the tail words of all slabs.
extern "C" int hv_mc_column_shifts_peers(const void* const* h_slab_ptrs, int world_size, int rank, int64_t n_columns, int tail_words,
                                         int32_t* shifts, unsigned long long* g_totals, int32_t* tails, void* scratch,
                                         int64_t scratch_bytes, void* stream) {
    HV_REQUIRE(h_slab_ptrs && tails && world_size >= 1 && world_size <= hv::kMaxSlabPtrs && tail_words >= 0 &&
                   world_size * tail_words <= 256, "hv_mc_column_shifts_peers: bad arguments (at most 16 slabs)");
    for (int w = 0; w < world_size; ++w) HV_REQUIRE(h_slab_ptrs[w], "hv_mc_column_shifts_peers: null slab pointer");
    return mc_column_shifts_impl(nullptr, 0, h_slab_ptrs, tails, tail_words, world_size, rank, n_columns, shifts, g_totals, scratch,
                                 scratch_bytes, stream);
}

static int mc_column_shifts_impl(const int32_t* all_counts, int64_t slab_stride, const void* const* h_slab_ptrs, int32_t* tails,
                                 int tail_words, int world_size, int rank, int64_t n_columns, int32_t* shifts,
                                 unsigned long long* g_totals, void* scratch, int64_t scratch_bytes, void* stream) {
    HV_REQUIRE(shifts && g_totals && scratch && world_size >= 1 && rank >= 0 && rank < world_size &&
                   n_columns > 0 && n_columns < (1ll << 31) && ((uintptr_t)scratch & 255) == 0,
               "hv_mc_column_shifts: bad arguments");
    ShiftLayout L;
    shift_layout(n_columns, &L);
    HV_REQUIRE((size_t)scratch_bytes >= L.total, "hv_mc_column_shifts: scratch of %lld bytes, need %lld", (long long)scratch_bytes,
               (long long)L.total);
    char* s = static_cast<char*>(scratch);
    uint2* p[6];
    for (int i = 0; i < 6; ++i) p[i] = reinterpret_cast<uint2*>(s + L.off_pairs[i]);
    uint2* bsum = reinterpret_cast<uint2*>(s + L.off_bsum);
    cudaStream_t st = hv::as_stream(stream);
    const unsigned int blocks = (unsigned int)((n_columns + 255) / 256);
    hv::SlabCounts sc;
    for (int w = 0; w < hv::kMaxSlabPtrs; ++w) sc.base[w] = (h_slab_ptrs && w < world_size) ? static_cast<const int32_t*>(h_slab_ptrs[w]) : nullptr;
    if (h_slab_ptrs)
        hv::mc_shift_prepare_kernel<true><<<blocks, 256, 0, st>>>(nullptr, 0, sc, world_size, rank, n_columns, p[0], p[1], p[2], shifts,
                                                                   tails, tail_words);
    else
        hv::mc_shift_prepare_kernel<false><<<blocks, 256, 0, st>>>(all_counts, slab_stride, sc, world_size, rank, n_columns, p[0], p[1],
                                                                    p[2], shifts, nullptr, 0);
    // g_totals: [0] edge vertices, [1] triangles, [2] centre vertices of the whole volume ([3] scratch)
    hv::exclusive_scan_u2(p[0], p[3], bsum, n_columns, g_totals, st);
    hv::exclusive_scan_u2(p[1], p[4], bsum, n_columns, g_totals + 2, st);
    hv::exclusive_scan_u2(p[2], p[5], bsum, n_columns, nullptr, st);
    hv::mc_shift_finish_kernel<<<blocks, 256, 0, st>>>(p[3], p[4], p[5], n_columns, shifts);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_scatter_rows(const void* src, const int32_t* index, int64_t n, int row_bytes, void* dst, void* stream) {
    HV_REQUIRE(n >= 0 && (row_bytes == 3 || (row_bytes > 0 && row_bytes % 4 == 0)), "hv_scatter_rows: rows of 3 or 4k bytes");
    if (n == 0) return HV_OK;
    HV_REQUIRE(src && index && dst, "hv_scatter_rows: null pointer");
    const int words = row_bytes == 3 ? 0 : row_bytes / 4;
    const long long threads = words ? n * words : n;
    hv::scatter_rows_kernel<<<(unsigned int)((threads + 255) / 256), 256, 0, hv::as_stream(stream)>>>(src, index, n, words, dst);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

// Count + emit in one call: the only host round trip (the output sizes are data dependent) is a stream synchronise on
// the counters copied into the caller's pinned `h_totals`; the caller provides output buffers with a capacity and gets
// HV_E_CAPACITY (with h_totals filled, scratch still valid) if the surface -- or its work records -- is larger.
extern "C" int hv_mc_extract(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                             const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                             double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts,
                             int64_t rec_cubes, int64_t z_global_offset, int64_t z_global_dim, float voxel_size,
                             const float* h_origin, int64_t cap_verts, int64_t cap_tris, float* verts_world,
                             float* verts_voxel, float* normals, uint8_t* colors, int32_t* faces,
                             unsigned long long* face_keys, unsigned long long* face_cubes, unsigned long long* vert_keys,
                             unsigned long long* totals, unsigned long long* h_totals, void* stream) {
    HV_REQUIRE(h_totals, "hv_mc_extract: null h_totals");
    int rc = hv_mc_count(tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, work, rec_verts, rec_cubes,
                         totals, stream);
    if (rc != HV_OK) return rc;
    cudaStream_t st = hv::as_stream(stream);
    HV_CUDA(cudaMemcpyAsync(h_totals, totals, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
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
    if (n_verts > cap_verts || n_tris > cap_tris || h_totals[6]) {
        hv::set_error("hv_mc_extract: %lld vertices / %lld triangles (%lld edge vertices / %lld cubes to record) exceed the "
                      "capacity %lld / %lld (%lld / %lld)", (long long)n_verts, (long long)n_tris, (long long)h_totals[4],
                      (long long)h_totals[5], (long long)cap_verts, (long long)cap_tris, (long long)rec_verts, (long long)rec_cubes);
        return HV_E_CAPACITY;
    }
    return hv_mc_emit(tsdf, halo_lo, halo_hi, color, color_halo_lo, color_halo_hi, h_dims, pitch, level, scratch,
                      scratch_bytes, work, rec_verts, rec_cubes, z_global_offset, z_global_dim, voxel_size, h_origin, n_verts,
                      n_tris, verts_world, verts_voxel, normals, colors, faces, face_keys, face_cubes, vert_keys, totals, stream);
}
