// Marching-cubes mesh extraction for sm_100a (B200).
//
// Replaces the `measure.marching_cubes(tsdf_vol, level=0)` call of TSDFVolume.get_mesh / get_point_cloud
// (third_party/tsdf_fusion_python/fusion.py:317,338) and the in-tree post-processing after it (fusion.py:339-348:
// nearest-voxel colour gather, b/g/r unpack, voxel -> world transform).  The arithmetic contract is the one restated in
// oracle/mc_oracle.c (scikit-image itself is not available anywhere on the box: parity unpinned).
//
// Pipeline (HBM-bound streaming + sparse compaction, no tensor cores), DESIGN.md 4.2:
//   1. mc_classify_kernel  the one pass over the TSDF plane (4 B/voxel).  A CTA sweeps a strip of TY columns along
//                          axis 0; the sign bits of planes x and x + 1 sit in shared memory, so the per-group vertex /
//                          triangle counts are taken from the bits the same CTA has just produced.  Writes the bit volume
//                          (N/8 bytes), one packed count word per 32-voxel group, and the totals of every run (= one
//                          plane of one strip, a contiguous range of the global group order); min / max for the
//                          skimage level-range check.
//   2. mc_run_scan_kernel  exclusive scan of the run totals (X * ceil(Y/TY) entries), one CTA.
//   3. mc_emit_kernel      sweeps the same strips: bits of three planes + count words of two planes in shared memory,
//                          block scans turn them into the index base of every group in reach, so a triangle corner's
//                          vertex index is run base + shared-memory prefix + popcounts -- no index volume, no global
//                          look-ups.  One thread per triangle record, one thread per vertex record (position, normal
//                          accumulated over the <= 4 cubes around the edge in scan order, colour, world transform) with
//                          the edge's 2x3x3 neighbourhood in registers.
// Vertex order: edge vertices sorted by (owner voxel in scan order, axis), then the cube-centre vertices by cube;
// triangles in cube scan order.  Deterministic; no dedup pass because each lattice edge has exactly one owner.
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
    // sweep geometry: a strip = TY columns (axis 1), a run = one plane (axis 0) of one strip
    int TY, S, GP;        // columns per strip, strips per plane, G + 2 (row pitch of the shared bit tables)
    // scratch
    unsigned int* bits;   // [X*Y*G]
    unsigned int* gcnt;   // [X*Y*G] edge vertices | centre vertices << 8 | triangles << 16 of the group
    uint4* run_tot;       // [X*S] (edge vertices, triangles, centre vertices, 0) of every run
    uint4* run_base;      // [X*S] exclusive prefix of run_tot
    unsigned long long* totals;  // [0] vertices [1] triangles [2] min key [3] max key (ordered-int float) [4] edge vertices
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

__global__ void mc_init_kernel(unsigned long long* totals) {
    if (threadIdx.x == 0) {
        totals[0] = 0ull;
        totals[1] = 0ull;
        totals[2] = 0x7fffffffull;  // min key
        totals[3] = 0x80000000ull;  // max key
        totals[4] = 0ull;
    }
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

// ---- shared-memory bit tables ---------------------------------------------------------------------------------------------
// One table = the sign words of one plane (axis 0) for TY + 2 columns of a strip (its own columns, then up to two columns of
// the next strip), GP = G + 2 words per column; words past the last group, columns past Y and planes past X read as zero.
constexpr int kSweepThreads = 256;

// Sign words around group (column c of the tables, word gq) of plane x: p0 = table of plane x, p1 = of plane x + 1.
__device__ __forceinline__ CubeWords cube_words(const McGrid& g, const unsigned int* p0, const unsigned int* p1, int x, int y,
                                                int c, int gq) {
    CubeWords w;
    const int i = c * g.GP + gq;
    const unsigned int b00 = p0[i], b01 = p0[i + g.GP], b10 = p1[i], b11 = p1[i + g.GP];
    const unsigned int u00 = (b00 >> 1) | (p0[i + 1] << 31), u01 = (b01 >> 1) | (p0[i + g.GP + 1] << 31);
    const unsigned int u10 = (b10 >> 1) | (p1[i + 1] << 31), u11 = (b11 >> 1) | (p1[i + g.GP + 1] << 31);
    // corner i offsets (axis0, axis1, axis2): see kMcCornerOff
    w.w[0] = b00; w.w[1] = u00; w.w[2] = u01; w.w[3] = b01;
    w.w[4] = b10; w.w[5] = u10; w.w[6] = u11; w.w[7] = b11;
    const bool x1 = x + 1 < g.X, y1 = y + 1 < g.Y;
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

__device__ __forceinline__ unsigned int warp_sum(unsigned int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- 1. classify: sign bits, per-group counts, run totals, min / max -------------------------------------------------------
// Loads the sign words of plane x (columns y0 .. y0 + ny of the strip; column ny belongs to the next strip) into `tab`;
// `owned`: this CTA is the one that publishes the plane's bits (and takes its values into min / max).
__device__ __forceinline__ void classify_load_plane(const McGrid& g, unsigned int* tab, int x, int y0, int ny, bool owned,
                                                    int vec, float level_f, bool level_exact, float& vmin, float& vmax) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int segs = (g.Zb + 127) >> 7;
    const int tiles = (ny + 1) * segs;
    for (int tb = warp; tb < tiles; tb += 4 * (kSweepThreads / 32)) {
        float v[4][4];
        int cc[4], sg[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {  // four 512-byte column segments of this warp in flight together
            const int t = tb + u * (kSweepThreads / 32);
            v[u][0] = v[u][1] = v[u][2] = v[u][3] = NAN;
            cc[u] = -1;
            if (t < tiles) {
                const int c = t / segs, seg = t - c * segs, y = y0 + c;
                cc[u] = c;
                sg[u] = seg;
                if (x < g.X && y < g.Y) {
                    const int z0 = seg * 128 + lane * 4;
                    const size_t col = (size_t)x * g.Y + y;
                    if (vec && z0 + 3 < g.Z) {
                        const float4* p = reinterpret_cast<const float4*>(g.tsdf + col * (size_t)g.pitch + z0);
                        const float4 q = (owned && c < ny) ? __ldcs(p) : __ldg(p);
                        v[u][0] = q.x; v[u][1] = q.y; v[u][2] = q.z; v[u][3] = q.w;
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (z0 + k < g.Zb) v[u][k] = mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, x, y, z0 + k);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (cc[u] < 0) continue;  // warp-uniform
            float lmin = INFINITY, lmax = -INFINITY;
            const unsigned int nib = mc_sign4(v[u], level_f, level_exact, g.level, lmin, lmax);
            unsigned int word = nib << (4 * (lane & 7));
            word |= __shfl_xor_sync(0xffffffffu, word, 1);
            word |= __shfl_xor_sync(0xffffffffu, word, 2);
            word |= __shfl_xor_sync(0xffffffffu, word, 4);
            const int gq = sg[u] * 4 + (lane >> 3);
            if ((lane & 7) == 0 && gq < g.G) {
                tab[cc[u] * g.GP + gq] = word;
                if (owned && cc[u] < ny) g.bits[((size_t)x * g.Y + (y0 + cc[u])) * g.G + gq] = word;
            }
            if (owned && cc[u] < ny) {
                vmin = fminf(vmin, lmin);
                vmax = fmaxf(vmax, lmax);
            }
        }
    }
}

__global__ void __launch_bounds__(kSweepThreads) mc_classify_kernel(const McGrid g, const int CX, const int vec) {
    extern __shared__ unsigned int s_mem[];
    __shared__ unsigned int s_tot[3];
    const int tabw = (g.TY + 2) * g.GP;
    const int strip = blockIdx.x % g.S, chunk = blockIdx.x / g.S;
    const int y0 = strip * g.TY, ny = min(g.TY, g.Y - y0);
    const int xa = chunk * CX, xb = min(g.X, xa + CX);
    if (xa >= xb) return;
    const float level_f = (float)g.level;  // exact compare needs double only when the level is not a float
    const bool level_exact = (double)level_f == g.level;
    float vmin = INFINITY, vmax = -INFINITY;
    for (int i = threadIdx.x; i < 2 * tabw; i += kSweepThreads) s_mem[i] = 0u;
    if (threadIdx.x < 3) s_tot[threadIdx.x] = 0u;
    __syncthreads();
    unsigned int* cur = s_mem;
    unsigned int* nxt = s_mem + tabw;
    classify_load_plane(g, cur, xa, y0, ny, true, vec, level_f, level_exact, vmin, vmax);
    for (int x = xa; x < xb; ++x) {
        classify_load_plane(g, nxt, x + 1, y0, ny, x + 1 < xb, vec, level_f, level_exact, vmin, vmax);
        __syncthreads();
        unsigned int a_nv = 0, a_nt = 0, a_nc = 0;
        for (int t = threadIdx.x; t < ny * g.G; t += kSweepThreads) {
            const int c = t / g.G, gq = t - c * g.G, y = y0 + c;
            const CubeWords w = cube_words(g, cur, nxt, x, y, c, gq);
            const unsigned int nv = __popc(w.mx) + __popc(w.my) + __popc(w.mz);
            unsigned int nt = 0, nc = 0;
            for (unsigned int rest = w.active; rest; rest &= rest - 1) {
                const int k = __ffs(rest) - 1;
                const int variant = cube_variant_at_ool(g, cube_index(w, k), x, y, gq * 32 + k);
                nt += kMcVarNtri[variant];
                nc += ((kMcVarMult[variant] >> 48) & 15ull) ? 1u : 0u;
            }
            g.gcnt[((size_t)x * g.Y + y) * g.G + gq] = nv | (nc << 8) | (nt << 16);
            a_nv += nv;
            a_nt += nt;
            a_nc += nc;
        }
        a_nv = warp_sum(a_nv);
        a_nt = warp_sum(a_nt);
        a_nc = warp_sum(a_nc);
        if ((threadIdx.x & 31) == 0 && (a_nv | a_nt | a_nc)) {
            atomicAdd(&s_tot[0], a_nv);
            atomicAdd(&s_tot[1], a_nt);
            atomicAdd(&s_tot[2], a_nc);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            g.run_tot[(size_t)x * g.S + strip] = make_uint4(s_tot[0], s_tot[1], s_tot[2], 0u);
            s_tot[0] = s_tot[1] = s_tot[2] = 0u;
        }
        unsigned int* t = cur;
        cur = nxt;
        nxt = t;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        vmin = fminf(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
        vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    }
    if ((threadIdx.x & 31) == 0 && vmin <= vmax) {  // at least one real value seen
        atomicMin(reinterpret_cast<int*>(g.totals + 2), float_order_key(vmin));
        atomicMax(reinterpret_cast<int*>(g.totals + 3), float_order_key(vmax));
    }
}

// ---- 2. exclusive scan of the run totals (one CTA) ----------------------------------------------------------------------------
constexpr int kRunScanThreads = 1024;

__global__ void __launch_bounds__(kRunScanThreads) mc_run_scan_kernel(const uint4* __restrict__ tot, uint4* __restrict__ base,
                                                                      int n, unsigned long long* totals) {
    __shared__ unsigned long long s_w[3][kRunScanThreads / 32];
    const int per = (n + kRunScanThreads - 1) / kRunScanThreads;
    const int i0 = threadIdx.x * per, i1 = min(n, i0 + per);
    unsigned long long a = 0, b = 0, c = 0;
    for (int i = i0; i < i1; ++i) {
        const uint4 v = tot[i];
        a += v.x;
        b += v.y;
        c += v.z;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long ia = a, ib = b, ic = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long ua = __shfl_up_sync(0xffffffffu, ia, o), ub = __shfl_up_sync(0xffffffffu, ib, o),
                                 uc = __shfl_up_sync(0xffffffffu, ic, o);
        if (lane >= o) {
            ia += ua;
            ib += ub;
            ic += uc;
        }
    }
    if (lane == 31) {
        s_w[0][warp] = ia;
        s_w[1][warp] = ib;
        s_w[2][warp] = ic;
    }
    __syncthreads();
    unsigned long long ca = 0, cb = 0, cc = 0, ta = 0, tb = 0, tc = 0;
    for (int w = 0; w < kRunScanThreads / 32; ++w) {
        if (w < warp) {
            ca += s_w[0][w];
            cb += s_w[1][w];
            cc += s_w[2][w];
        }
        ta += s_w[0][w];
        tb += s_w[1][w];
        tc += s_w[2][w];
    }
    unsigned long long ea = ca + ia - a, eb = cb + ib - b, ec = cc + ic - c;  // exclusive prefix of this thread's range
    for (int i = i0; i < i1; ++i) {
        const uint4 v = tot[i];
        base[i] = make_uint4((unsigned int)ea, (unsigned int)eb, (unsigned int)ec, 0u);
        ea += v.x;
        eb += v.y;
        ec += v.z;
    }
    if (threadIdx.x == 0) {
        totals[0] = ta + tc;  // vertices: edge vertices, then the cube-centre vertices
        totals[1] = tb;
        totals[4] = ta;
    }
}

// ---- 3. emit -----------------------------------------------------------------------------------------------------------------
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
    int CX;
};

__device__ __forceinline__ unsigned long long vertex_key(const McGrid& g, int x, int y, int zb, int axis) {
    const unsigned long long lin = ((unsigned long long)x * g.Y + y) * (unsigned long long)g.Zg + (unsigned long long)(zb + g.zoff);
    return lin * 4ull + (unsigned long long)axis;
}

// Exclusive block scan of one value per thread, with a running carry kept in shared memory (s_carry); `s_warp`: 8 words.
__device__ __forceinline__ unsigned int block_scan_carry(unsigned int v, unsigned int* s_warp, unsigned int* s_carry) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned int before = *s_carry, total = 0;
#pragma unroll
    for (int w = 0; w < kSweepThreads / 32; ++w) {
        const unsigned int sw = s_warp[w];
        if (w < warp) before += sw;
        total += sw;
    }
    __syncthreads();
    if (threadIdx.x == 0) *s_carry += total;
    return before + inc - v;
}

// Tables the emit kernel keeps in shared memory.
struct EmitTables {
    unsigned int* bits[3];  // planes x, x + 1, x + 2 (rotating): [(TY + 2) * GP]
    unsigned int* vpre[2];  // planes x, x + 1: global index of the first edge vertex of every group, [(TY + 1) * G]
    unsigned int* tpre;     // own groups of plane x: run-local index of the group's first triangle, [TY * G]
    unsigned int* cpre;     //   ... of its first centre vertex
    unsigned int* cnt;      //   ... its packed counts
    unsigned int* mw;       //   ... mask of its cubes that have a centre vertex
    uint2* rec;             // record list of one round
    unsigned int* s_warp;   // 8 words of scan scratch
    unsigned int* s_misc;   // [0] carry, [1] own total of a scan, [2..4] run totals (edge vertices, triangles, centres)
};

constexpr int kRecCap = 2048;  // records per round (16 KB)

__device__ __forceinline__ void emit_load_bits(const McGrid& g, unsigned int* tab, int p, int y0) {
    const int rows = g.TY + 2;
    for (int i = threadIdx.x; i < rows * g.GP; i += kSweepThreads) {
        const int c = i / g.GP, gq = i - c * g.GP, y = y0 + c;
        tab[i] = (p < g.X && y < g.Y && gq < g.G) ? __ldg(g.bits + ((size_t)p * g.Y + y) * g.G + gq) : 0u;
    }
}

// Global index of the first edge vertex of every group of plane p in reach: the strip's own columns continue the run
// (p, strip); column ny is the first column of the run (p, strip + 1).
__device__ __forceinline__ void emit_build_vpre(const McGrid& g, const EmitTables& T, unsigned int* vpre, int p, int strip,
                                                int y0, int ny) {
    const int n_own = ny * g.G, n_all = (y0 + ny < g.Y) ? n_own + g.G : n_own;
    if (p >= g.X) return;  // never referenced
    const unsigned int base_own = g.run_base[(size_t)p * g.S + strip].x;
    const unsigned int base_next = (y0 + ny < g.Y) ? g.run_base[(size_t)p * g.S + strip + 1].x : 0u;
    if (threadIdx.x == 0) T.s_misc[0] = 0u;
    __syncthreads();
    for (int t0 = 0; t0 < n_all; t0 += kSweepThreads) {
        const int t = t0 + threadIdx.x;
        unsigned int v = 0;
        if (t < n_all) {
            const int c = t / g.G, gq = t - c * g.G;
            v = __ldg(g.gcnt + ((size_t)p * g.Y + (y0 + c)) * g.G + gq) & 255u;
        }
        const unsigned int ex = block_scan_carry(v, T.s_warp, T.s_misc);
        if (t == n_own) T.s_misc[1] = ex;  // total of the own columns
        if (t < n_all) vpre[t] = ex;       // rebased below
    }
    __syncthreads();
    const unsigned int own_total = T.s_misc[1];
    for (int t = threadIdx.x; t < n_all; t += kSweepThreads)
        vpre[t] = (t < n_own) ? base_own + vpre[t] : base_next + (vpre[t] - own_total);
}

// Run-local prefixes of the triangles and centre vertices of the own groups of plane x; run totals into s_misc[2..4].
__device__ __forceinline__ void emit_build_own(const McGrid& g, const EmitTables& T, int x, int y0, int ny) {
    const int n_own = ny * g.G;
    unsigned int nv_total = 0;
    for (int comp = 0; comp < 2; ++comp) {  // 0: triangles, 1: centre vertices
        if (threadIdx.x == 0) T.s_misc[0] = 0u;
        __syncthreads();
        for (int t0 = 0; t0 < n_own; t0 += kSweepThreads) {
            const int t = t0 + threadIdx.x;
            unsigned int word = 0;
            if (t < n_own) {
                const int c = t / g.G, gq = t - c * g.G;
                word = __ldg(g.gcnt + ((size_t)x * g.Y + (y0 + c)) * g.G + gq);
                if (comp == 0) T.cnt[t] = word;
            }
            const unsigned int v = comp == 0 ? (word >> 16) : ((word >> 8) & 255u);
            const unsigned int ex = block_scan_carry(v, T.s_warp, T.s_misc);
            if (t < n_own) (comp == 0 ? T.tpre : T.cpre)[t] = ex;
            if (comp == 0) nv_total += word & 255u;
        }
        __syncthreads();
        if (threadIdx.x == 0) T.s_misc[comp == 0 ? 3 : 4] = T.s_misc[0];
    }
    nv_total = warp_sum(nv_total);
    if (threadIdx.x == 0) T.s_misc[2] = 0u;
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && nv_total) atomicAdd(&T.s_misc[2], nv_total);
    __syncthreads();
}

// Index of the edge vertex owned by voxel (column co of the tables, z = zo) of the plane whose tables are (bp, bq = the
// plane after it), on `axis`: the group's base from `vpre`, plus the vertices of lower voxels, plus lower axes of its own.
__device__ __forceinline__ unsigned int edge_vertex_index(const McGrid& g, const unsigned int* bp, const unsigned int* bq,
                                                          const unsigned int* vpre, int xo, int yo, int co, int zo, int axis) {
    const int og = zo >> 5, bit = zo & 31;
    const int i = co * g.GP + og;
    const unsigned int b00 = bp[i], u00 = (b00 >> 1) | (bp[i + 1] << 31), b01 = bp[i + g.GP], b10 = bq[i];
    const unsigned int own = range_mask(g.own_lo, g.own_hi, og * 32), zin = range_mask(0, g.Zb - 1, og * 32);
    const unsigned int mx = (xo + 1 < g.X) ? ((b00 ^ b10) & own) : 0u;
    const unsigned int my = (yo + 1 < g.Y) ? ((b00 ^ b01) & own) : 0u;
    const unsigned int mz = (b00 ^ u00) & own & zin;
    const unsigned int low = (1u << bit) - 1u;
    unsigned int r = vpre[co * g.G + og] + __popc(mx & low) + __popc(my & low) + __popc(mz & low);
    if (axis > 0) r += (mx >> bit) & 1u;
    if (axis > 1) r += (my >> bit) & 1u;
    return r;
}

// One triangle record: rec.x = column | word << 5 | voxel << 13 | tiling << 18 | triangle << 28, rec.y = run-local index.
__device__ __forceinline__ void emit_triangle(const McGrid& g, const McEmit& out, const EmitTables& T, int s0, int x, int y0,
                                              uint2 rec, unsigned int tri_base, unsigned int centre_base) {
    const int c = (int)(rec.x & 31u), gq = (int)((rec.x >> 5) & 255u), k = (int)((rec.x >> 13) & 31u);
    const int variant = (int)((rec.x >> 18) & 1023u), j = (int)(rec.x >> 28);
    const int y = y0 + c, zb = gq * 32 + k;
    const size_t ti = (size_t)tri_base + rec.y;
    if ((long long)ti >= out.cap_tris) return;
    const uint8_t* tri = kMcTris + 3 * ((size_t)kMcVarTriOff[variant] + (size_t)j);
    if (out.face_keys) {
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int e = tri[q];
            unsigned long long key;
            if (e < 12)
                key = vertex_key(g, x + kMcEdgeOwner[4 * e], y + kMcEdgeOwner[4 * e + 1], zb + kMcEdgeOwner[4 * e + 2],
                                 kMcEdgeOwner[4 * e + 3]);
            else
                key = vertex_key(g, x, y, zb, 3);
            out.face_keys[3 * ti + q] = key;
        }
        if (out.face_cubes) out.face_cubes[ti] = vertex_key(g, x, y, zb, 0) >> 2;
        return;
    }
    unsigned int v[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const int e = tri[q];
        if (e < 12) {
            const int ox = kMcEdgeOwner[4 * e], oy = kMcEdgeOwner[4 * e + 1], oz = kMcEdgeOwner[4 * e + 2];
            const unsigned int* bp = T.bits[(s0 + ox) % 3];
            const unsigned int* bq = T.bits[(s0 + ox + 1) % 3];
            v[q] = edge_vertex_index(g, bp, bq, T.vpre[(x + ox) & 1], x + ox, y + oy, c + oy, zb + oz, kMcEdgeOwner[4 * e + 3]);
        } else {
            const int t = c * g.G + gq;
            v[q] = centre_base + T.cpre[t] + __popc(T.mw[t] & ((1u << k) - 1u));
        }
    }
#pragma unroll
    for (int q = 0; q < 3; ++q) out.faces[3 * ti + q] = (int32_t)v[q];
}

// Bits (i0 | i1 << 1 | ... ) -> Lewiner configuration index: corner i of the cube sits at cube-local (ia, ib, ic) along
// (edge axis, lower other axis, upper other axis); kCornerPos[axis] holds ia * 4 + ib * 2 + ic for i = 0..7, 3 bits each.
__device__ __forceinline__ unsigned int corner_positions(int axis) {
    // corner offsets (axis0, axis1, axis2): 0:(0,0,0) 1:(0,0,1) 2:(0,1,1) 3:(0,1,0) 4:(1,0,0) 5:(1,0,1) 6:(1,1,1) 7:(1,1,0)
    // axis 0: (ia, ib, ic) = (o0, o1, o2); axis 1: (o1, o0, o2); axis 2: (o2, o0, o1)   [tools/gen_mc_tables.py CORNERS]
    return axis == 0 ? 0xdec4c8u : (axis == 1 ? 0xdda948u : 0x7f2360u);
}

// One edge vertex (fusion.py:338-348 for one vertex): position on the edge, normal accumulated over the (up to four) cubes
// around the edge in scan order, colour of the nearest voxel, world transform.  The edge's 2 x 3 x 3 neighbourhood lives in
// registers in an edge-relative frame (a = along the edge, b < c the other two axes), so nothing is indexed dynamically.
__device__ __forceinline__ void emit_edge_vertex(const McGrid& g, const McEmit& out, int x, int y, int zb, int axis, size_t vi) {
    const int bax = axis == 0 ? 1 : 0, cax = axis == 2 ? 1 : 2;
    const int pb = bax == 0 ? x : y, pc = cax == 1 ? y : zb;
    const int db_dim = bax == 0 ? g.X : g.Y, dc_dim = cax == 1 ? g.Y : g.Zb;
    float n[2][3][3];
#pragma unroll
    for (int ia = 0; ia < 2; ++ia)
#pragma unroll
        for (int ib = 0; ib < 3; ++ib)
#pragma unroll
            for (int ic = 0; ic < 3; ++ic) {
                const int qb = pb + ib - 1, qc = pc + ic - 1;
                float f = 0.0f;
                if (qb >= 0 && qb < db_dim && qc >= 0 && qc < dc_dim) {
                    // (x, y, zb) of the neighbour: the three coordinates are (a, b, c) permuted by `axis`
                    const int qx = axis == 0 ? x + ia : qb;                      // bax == 0 whenever axis != 0
                    const int qy = axis == 1 ? y + ia : (axis == 0 ? qb : qc);    // axis 0: b = y; axis 2: c = y
                    const int qz = axis == 2 ? zb + ia : qc;                     // cax == 2 whenever axis != 2
                    f = mc_fetch(g, g.tsdf, g.halo_lo, g.halo_hi, qx, qy, qz);
                }
                n[ia][ib][ic] = f;
            }
    const double level = g.level;
    const double v_lo = (double)n[0][1][1] - level, v_hi = (double)n[1][1][1] - level;
    const double w_lo = 1.0 / ((double)FLT_EPSILON + fabs(v_lo)), w_hi = 1.0 / ((double)FLT_EPSILON + fabs(v_hi));
    const double frac = w_hi / __dadd_rn(w_lo, w_hi);
    const unsigned int cpos = corner_positions(axis);
    float acc_a = 0.0f, acc_b = 0.0f, acc_c = 0.0f;
    // the cubes around the edge in scan order (axis0 outermost): larger offsets first; b is the lower axis
#pragma unroll
    for (int db = 1; db >= 0; --db) {
#pragma unroll
        for (int dc = 1; dc >= 0; --dc) {
            const int bb = pb - db, bc = pc - dc;  // base of the cube along b, c
            if (bb < 0 || bc < 0 || bb + 1 >= db_dim || bc + 1 >= dc_dim) continue;
            // sign bits of the cube's corners at cube-local (ia, ib, ic)
            unsigned int loc = 0;
#pragma unroll
            for (int ia = 0; ia < 2; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib)
#pragma unroll
                    for (int ic = 0; ic < 2; ++ic)
                        loc |= (((double)n[ia][ib + 1 - db][ic + 1 - dc] - level) > 0.0 ? 1u : 0u) << (ia * 4 + ib * 2 + ic);
            int idx = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) idx |= (int)((loc >> ((cpos >> (3 * i)) & 7u)) & 1u) << i;
            int variant;
            if (kMcAmbCount[idx] == 0 && kMcInteriorBase[idx] == 255) {
                variant = kMcVariantBase[idx];
            } else {
                const int cx = axis == 0 ? x : bb;
                const int cy = axis == 1 ? y : (axis == 0 ? bb : bc);
                const int cz = axis == 2 ? zb : bc;
                variant = cube_variant_ambiguous(g.tsdf, g.halo_lo, g.halo_hi, g.Y, g.pitch, g.Z, g.hl, g.level, idx, cx, cy, cz);
            }
            // the edge inside this cube: lower corner at cube-local (0, db, dc)
            const int d0 = axis == 0 ? 0 : db, d1 = axis == 1 ? 0 : (axis == 0 ? db : dc), d2 = axis == 2 ? 0 : dc;
            const int e = kMcEdgeAt[axis * 8 + d0 * 4 + d1 * 2 + d2];
            const int mult = (int)((kMcVarMult[variant] >> (4 * e)) & 15ull);
            if (mult == 0) continue;
            // down-hill in-cube finite differences at the edge's two end corners, weighted like the position
            const double ga = v_lo - v_hi;
            const double gb_lo = ((double)n[0][1 - db][1] - level) - ((double)n[0][2 - db][1] - level);
            const double gb_hi = ((double)n[1][1 - db][1] - level) - ((double)n[1][2 - db][1] - level);
            const double gc_lo = ((double)n[0][1][1 - dc] - level) - ((double)n[0][1][2 - dc] - level);
            const double gc_hi = ((double)n[1][1][1 - dc] - level) - ((double)n[1][1][2 - dc] - level);
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
    const unsigned int packed = (unsigned int)mc_fetch(g, g.color, g.chalo_lo, g.chalo_hi, i0, i1, i2);
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

__global__ void __launch_bounds__(kSweepThreads, 2) mc_emit_kernel(const McGrid g, const McEmit out) {
    extern __shared__ unsigned int s_mem[];
    EmitTables T;
    {
        const int tabw = (g.TY + 2) * g.GP, vw = (g.TY + 1) * g.G, ow = g.TY * g.G;
        unsigned int* p = s_mem;
        for (int i = 0; i < 3; ++i, p += tabw) T.bits[i] = p;
        for (int i = 0; i < 2; ++i, p += vw) T.vpre[i] = p;
        T.tpre = p; p += ow;
        T.cpre = p; p += ow;
        T.cnt = p; p += ow;
        T.mw = p; p += ow;
        T.s_warp = p; p += 8;
        T.s_misc = p; p += 8;
        T.rec = reinterpret_cast<uint2*>(p + ((reinterpret_cast<uintptr_t>(p) & 4) ? 1 : 0));
    }
    const int strip = blockIdx.x % g.S, chunk = blockIdx.x / g.S;
    const int y0 = strip * g.TY, ny = min(g.TY, g.Y - y0);
    const int xa = chunk * out.CX, xb = min(g.X, xa + out.CX);
    const int n_own = ny * g.G;
    const unsigned int n_edge_total = (unsigned int)g.totals[4];
    int held_bits[3] = {-1, -1, -1}, held_v[2] = {-1, -1};
    for (int x = xa; x < xb; ++x) {
        const uint4 rt = g.run_tot[(size_t)x * g.S + strip];
        if ((rt.x | rt.y | rt.z) == 0u) continue;  // nothing owned by this run (uniform over the block)
        __syncthreads();                           // the previous plane's readers are done with the tables
        for (int d = 0; d < 3; ++d)
            if (held_bits[(x + d) % 3] != x + d) {
                emit_load_bits(g, T.bits[(x + d) % 3], x + d, y0);
                held_bits[(x + d) % 3] = x + d;
            }
        for (int d = 0; d < 2; ++d)
            if (held_v[(x + d) & 1] != x + d) {
                emit_build_vpre(g, T, T.vpre[(x + d) & 1], x + d, strip, y0, ny);
                held_v[(x + d) & 1] = x + d;
            }
        emit_build_own(g, T, x, y0, ny);
        __syncthreads();
        const uint4 rb = g.run_base[(size_t)x * g.S + strip];
        const unsigned int run_nv = T.s_misc[2], run_nt = T.s_misc[3], run_nc = T.s_misc[4];
        const int s0 = x % 3;
        const unsigned int* B0 = T.bits[s0];
        const unsigned int* B1 = T.bits[(s0 + 1) % 3];
        // ---- triangles: one record per triangle, then one thread per record
        for (unsigned int lo = 0; lo < run_nt; lo += kRecCap) {
            const unsigned int hi = lo + kRecCap;
            for (int t = threadIdx.x; t < n_own; t += kSweepThreads) {
                const unsigned int nt = T.cnt[t] >> 16;
                unsigned int o = T.tpre[t];
                if (nt == 0u || o + nt <= lo || o >= hi) continue;
                const int c = t / g.G, gq = t - c * g.G;
                const CubeWords w = cube_words(g, B0, B1, x, y0 + c, c, gq);
                unsigned int mw = 0u;
                for (unsigned int rest = w.active; rest; rest &= rest - 1) {
                    const int k = __ffs(rest) - 1;
                    const int variant = cube_variant_at(g, cube_index(w, k), x, y0 + c, gq * 32 + k);
                    const unsigned int ntri = kMcVarNtri[variant];
                    if ((kMcVarMult[variant] >> 48) & 15ull) mw |= 1u << k;
                    const unsigned int head = (unsigned int)c | ((unsigned int)gq << 5) | ((unsigned int)k << 13) |
                                              ((unsigned int)variant << 18);
                    for (unsigned int j = 0; j < ntri; ++j)
                        if (o + j >= lo && o + j < hi) T.rec[o + j - lo] = make_uint2(head | (j << 28), o + j);
                    o += ntri;
                }
                T.mw[t] = mw;
            }
            __syncthreads();
            const unsigned int n = min(run_nt - lo, (unsigned int)kRecCap);
            for (unsigned int i = threadIdx.x; i < n; i += kSweepThreads)
                emit_triangle(g, out, T, s0, x, y0, T.rec[i], rb.y, n_edge_total + rb.z);
            __syncthreads();
        }
        // ---- edge vertices
        for (unsigned int lo = 0; lo < run_nv; lo += kRecCap) {
            const unsigned int hi = lo + kRecCap;
            for (int t = threadIdx.x; t < n_own; t += kSweepThreads) {
                const unsigned int nv = T.cnt[t] & 255u;
                unsigned int o = T.vpre[x & 1][t] - rb.x;  // run-local
                if (nv == 0u || o + nv <= lo || o >= hi) continue;
                const int c = t / g.G, gq = t - c * g.G;
                const CubeWords w = cube_words(g, B0, B1, x, y0 + c, c, gq);
                const unsigned int head = (unsigned int)c | ((unsigned int)gq << 5);
                for (unsigned int rest = w.mx | w.my | w.mz; rest; rest &= rest - 1) {
                    const int k = __ffs(rest) - 1;
#pragma unroll
                    for (int axis = 0; axis < 3; ++axis) {
                        const unsigned int m = axis == 0 ? w.mx : (axis == 1 ? w.my : w.mz);
                        if ((m >> k) & 1u) {
                            if (o >= lo && o < hi) T.rec[o - lo] = make_uint2(head | ((unsigned int)k << 13) | ((unsigned int)axis << 18), o);
                            ++o;
                        }
                    }
                }
            }
            __syncthreads();
            const unsigned int n = min(run_nv - lo, (unsigned int)kRecCap);
            for (unsigned int i = threadIdx.x; i < n; i += kSweepThreads) {
                const uint2 r = T.rec[i];
                const size_t vi = (size_t)rb.x + r.y;
                if ((long long)vi < out.cap_verts)
                    emit_edge_vertex(g, out, x, y0 + (int)(r.x & 31u), (int)((r.x >> 5) & 255u) * 32 + (int)((r.x >> 13) & 31u),
                                     (int)((r.x >> 18) & 3u), vi);
            }
            __syncthreads();
        }
        // ---- cube-centre vertices (tilings 6.1.2, 7.3, 10.2, 12.2, 13.3, 13.4): after all edge vertices
        if (run_nc) {
            for (int t = threadIdx.x; t < n_own; t += kSweepThreads) {
                unsigned int m = ((T.cnt[t] >> 8) & 255u) ? T.mw[t] : 0u;
                unsigned int o = T.cpre[t];
                const int c = t / g.G, gq = t - c * g.G;
                for (; m; m &= m - 1, ++o) {
                    const size_t vi = (size_t)n_edge_total + rb.z + o;
                    if ((long long)vi < out.cap_verts) emit_centre_vertex(g, out, x, y0 + c, gq * 32 + (__ffs(m) - 1), vi);
                }
            }
        }
    }
}

}  // namespace hv

// ---- C ABI ---------------------------------------------------------------------------------------------------------------
namespace {

struct McLayout {
    long long n_groups, n_runs;
    int G, Zb, TY, S, GP;
    size_t off_bits, off_gcnt, off_run_tot, off_run_base, total;
};

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

int mc_layout(const int64_t* h_dims, bool lo, bool hi, McLayout* L) {
    HV_REQUIRE(h_dims, "marching cubes: null dims");
    const int64_t X = h_dims[0], Y = h_dims[1], Z = h_dims[2];
    HV_REQUIRE(X >= 1 && Y >= 1 && Z >= 1 && X < (1ll << 30) && Y < (1ll << 30) && Z < (1ll << 30),
               "marching cubes: bad dims %lld x %lld x %lld", (long long)X, (long long)Y, (long long)Z);
    L->Zb = (int)Z + (lo ? 1 : 0) + (hi ? 1 : 0);
    L->G = (L->Zb + 31) / 32;
    HV_REQUIRE(L->G <= 256, "marching cubes: more than 8192 voxels along axis 2 per shard (%d): cut the volume into z-slabs", L->Zb);
    L->n_groups = (long long)X * Y * L->G;
    HV_REQUIRE(L->n_groups < (1ll << 32), "marching cubes: volume too large");
    int ty = 512 / L->G;  // about two rounds of the 256 threads per plane of a strip
    ty = ty < 1 ? 1 : (ty > 30 ? 30 : ty);
    if (ty > Y) ty = (int)Y;
    L->TY = ty;
    L->S = (int)((Y + ty - 1) / ty);
    L->GP = L->G + 2;
    L->n_runs = (long long)X * L->S;
    HV_REQUIRE(L->n_runs < (1ll << 31), "marching cubes: too many runs");
    size_t o = 0;
    L->off_run_tot = o;
    o = align_up(o + (size_t)L->n_runs * sizeof(uint4));
    L->off_run_base = o;
    o = align_up(o + (size_t)L->n_runs * sizeof(uint4));
    L->off_bits = o;
    o = align_up(o + (size_t)L->n_groups * sizeof(unsigned int));
    L->off_gcnt = o;
    o = align_up(o + (size_t)L->n_groups * sizeof(unsigned int));
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
    g->TY = L->TY;
    g->S = L->S;
    g->GP = L->GP;
    g->bits = reinterpret_cast<unsigned int*>(s + L->off_bits);
    g->gcnt = reinterpret_cast<unsigned int*>(s + L->off_gcnt);
    g->run_tot = reinterpret_cast<uint4*>(s + L->off_run_tot);
    g->run_base = reinterpret_cast<uint4*>(s + L->off_run_base);
    g->totals = nullptr;
    g->n_groups = L->n_groups;
    return HV_OK;
}

// planes per chunk so that about `target` CTAs exist
int planes_per_chunk(int X, int S, long long target) {
    long long chunks = (target + S - 1) / S;
    if (chunks < 1) chunks = 1;
    if (chunks > X) chunks = X;
    return (int)((X + chunks - 1) / chunks);
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
    hv::mc_init_kernel<<<1, 32, 0, st>>>(totals);  // (a pageable H2D of the counters would block the host)
    const int sms = hv::device_info().sm_count;
    const int vec = (!halo_lo && (pitch & 3) == 0 && ((uintptr_t)tsdf & 15) == 0) ? 1 : 0;
    const int CX = planes_per_chunk(g.X, g.S, (long long)sms * 4);
    const int chunks = (g.X + CX - 1) / CX;
    const size_t smem = (size_t)2 * (g.TY + 2) * g.GP * sizeof(unsigned int);
    hv::mc_classify_kernel<<<(unsigned int)(g.S * chunks), hv::kSweepThreads, smem, st>>>(g, CX, vec);
    hv::mc_run_scan_kernel<<<1, hv::kRunScanThreads, 0, st>>>(g.run_tot, g.run_base, (int)L.n_runs, totals);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_mc_emit(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                          const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                          double level, void* scratch, int64_t scratch_bytes, int64_t z_global_offset, int64_t z_global_dim,
                          float voxel_size, const float* h_origin, int64_t n_verts, int64_t n_tris, float* verts_world,
                          float* verts_voxel, float* normals, uint8_t* colors, int32_t* faces,
                          unsigned long long* face_keys, unsigned long long* face_cubes, unsigned long long* vert_keys,
                          unsigned long long* totals, void* stream) {
    hv::McGrid g;
    McLayout L;
    int rc = mc_grid(&g, tsdf, halo_lo, halo_hi, h_dims, pitch, level, scratch, scratch_bytes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(n_verts >= 0 && n_tris >= 0 && n_verts < (1ll << 31) && n_tris < (1ll << 31), "hv_mc_emit: bad counts");
    if (n_verts == 0 && n_tris == 0) return HV_OK;  // a shard the surface does not cross / no capacity at all
    HV_REQUIRE(color && h_origin && verts_world && normals && colors && totals, "hv_mc_emit: null pointer");
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
    const int sms = hv::device_info().sm_count;
    em.CX = planes_per_chunk(g.X, g.S, (long long)sms * 16);
    if (em.CX < 4 && g.X >= 4) em.CX = 4;
    const int chunks = (g.X + em.CX - 1) / em.CX;
    const size_t words = (size_t)3 * (g.TY + 2) * g.GP + (size_t)2 * (g.TY + 1) * g.G + (size_t)4 * g.TY * g.G + 16 + 2;
    const size_t smem = words * sizeof(unsigned int) + (size_t)hv::kRecCap * sizeof(uint2);
    HV_REQUIRE(smem <= 48 * 1024, "hv_mc_emit: %lld bytes of shared memory", (long long)smem);
    hv::mc_emit_kernel<<<(unsigned int)(g.S * chunks), hv::kSweepThreads, smem, st>>>(g, em);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

// Count + emit in one call: the only host round trip (the output sizes are data dependent) is a stream synchronise on
// the counters copied into the caller's pinned `h_totals`; the caller provides output buffers with a capacity and gets
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
                      verts_voxel, normals, colors, faces, face_keys, face_cubes, vert_keys, totals, stream);
}
