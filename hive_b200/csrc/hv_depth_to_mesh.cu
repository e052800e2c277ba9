// Per-frame depth-to-mesh for sm_100a (B200): back-projection + lattice triangulation + face filter, batched over frames.
//
// Replaces, for HIVE's dynamic foreground meshes (hive/pipeline.py:386-409):
//   point_cloud_from_depth / image2world        hive/geometric.py:107-126,183-206
//   Pipeline._triangulate_faces                 hive/pipeline.py:651-667   (scipy.spatial.Delaunay on pixel coordinates)
//   Pipeline._filter_faces                      hive/pipeline.py:670-694   (edge <= max_pixel_distance px, |d depth| <= max)
//
// The Delaunay triangulation of a subset of the integer lattice, after the 2-pixel edge filter, has a closed form
// (validated there against scipy): only two triangle shapes survive --
//   * unit right triangles: a 2x2 quad with 4 valid corners gives two (split along one diagonal), with exactly 3 valid
//     corners gives one;
//   * "bridge" triangles (sqrt2, sqrt2, 2) around an ABSENT pixel whose 4-neighbours N, E, S, W are present: 3 present
//     give one, 4 present give two.
// Co-circular points make Qhull's choice of diagonal arbitrary; here it is a rule (the diagonal with the smaller depth
// difference, ties -> main / horizontal) or, for parity tests, a caller-provided per-pixel bitmap.
//
// One thread per pixel; vertex ids are row-major ranks of the valid pixels (np.nonzero order, geometric.py:121),
// obtained from per-32-pixel valid-bit words + a prefix over the words, so no per-pixel id map is materialised.
// Faces are compacted with warp scans in pixel order: the output is deterministic.
#include "hv_common.cuh"

namespace hv {

struct D2mGrid {
    const float* depth;    // [B][H][W]
    const uint8_t* mask;   // [B][H][W] or null (all true)
    const uint8_t* diag;   // [B][H][W] or null; bit 0: quad owned by the pixel uses the anti-diagonal, bit 1: hole at the
                           // pixel is split N-S
    int B, H, W;
    int wpr;               // 32-pixel words per image row: ceil(W / 32); word (b, v, c) holds pixels u = 32 c + lane
    int words_per_frame;   // H * wpr
    long long n_words;
    float max_depth_distance;
    int bridges;
    int verts_only;        // skip triangulation (point_cloud_from_depth)
    uint8_t* slots;        // [B][H][W] which of a pixel's eight candidate faces exist (count -> emit), written where a word row has faces
    unsigned int* wbits;   // [n_words] valid bits
    uint2* wcount;         // [n_words] (vertices, faces)
    uint2* woff;           // [n_words] exclusive prefix
    uint2* bsum;
};

// The faces a pixel (u, v) can own, in the order they are emitted, as corner triples over the six pixels 0 = (u, v),
// 1 = (u+1, v), 2 = (u, v+1), 3 = (u+1, v+1), 4 = (u, v-1), 5 = (u-1, v); winding with negative signed area in (u, v):
//   quad:  slot 0 (A, D, B)   1 (A, C, D)   [main diagonal]      2 (A, C, B)   3 (B, C, D)   [anti-diagonal]
//   hole:  slot 4 (W, E, N)   5 (W, S, E)   [split W-E]          6 (N, S, E)   7 (N, W, S)   [split N-S]
// A quad / hole with three of its four vertices gives the one triangle of these that avoids the missing one.  The count
// kernel decides which slots exist (validity, diagonal, depth filter) and leaves the slot mask -- one byte -- for the emit
// kernel, which therefore needs neither the neighbourhood depths nor the tests again.  (r01: both kernels enumerated the
// faces; the emit kernel spent ~680 warp-instructions per word row.)
struct PixelFaces {
    int n;
    unsigned int slots;
};
// corner ids of the quad slots 0..3 and the hole slots 4..7, 3 bits per corner, 9 per slot
constexpr unsigned long long kQuadCorners = (0ull | 3ull << 3 | 1ull << 6) | (0ull | 2ull << 3 | 3ull << 6) << 9 |
                                            (0ull | 2ull << 3 | 1ull << 6) << 18 | (1ull | 2ull << 3 | 3ull << 6) << 27;
constexpr unsigned long long kHoleCorners = (5ull | 1ull << 3 | 4ull << 6) | (5ull | 2ull << 3 | 1ull << 6) << 9 |
                                            (4ull | 2ull << 3 | 1ull << 6) << 18 | (4ull | 5ull << 3 | 2ull << 6) << 27;

// Depth of pixel (u, v) if it is a vertex (inside the image, mask set, depth > 0: geometric.py:117), else 0.
__device__ __forceinline__ float d2m_vdepth(const D2mGrid& g, int b, int u, int v) {
    if (u < 0 || v < 0 || u >= g.W || v >= g.H) return 0.0f;
    const size_t i = ((size_t)b * g.H + v) * g.W + u;
    if (g.mask && !g.mask[i]) return 0.0f;
    const float d = __ldg(g.depth + i);
    return d > 0.0f ? d : 0.0f;
}

// The neighbourhood values a pixel's faces depend on.  A warp owns one 32-pixel word column and walks kD2mRows image
// rows downwards, keeping three rows in registers: every row is loaded once per warp (each lane its own pixel, coalesced;
// lanes 0 and 1 the halo pixels right and left of the word), horizontal neighbours come by shuffle, the row below
// becomes the current row of the next step.  (r01: one thread per pixel re-loading up to eight neighbours -- mask byte
// + depth each, behind a 64-bit index and an integer division -- made d2m_count_kernel 18x slower than its HBM time.)
constexpr int kD2mRows = 8;

struct Row {
    float c;               // vertex depth of the lane's pixel, or 0
    float halo_r, halo_l;  // warp-uniform: pixels 32 cw + 32 and 32 cw - 1 of the row
    unsigned int bits;     // warp-uniform: vertices of the word
};

// The loads of a row and what is derived from them are separate steps, so that a kernel can issue row v + 2 before it
// works on row v: with load-then-use in one call every row cost a full memory round trip per warp (r01 ncu of the
// one-step version: 12 long-scoreboard stall cycles per issued instruction in d2m_count_kernel).
struct RawRow {
    float d, dh;
    unsigned int m, mh;
};

__device__ __forceinline__ RawRow load_raw(const D2mGrid& g, int b, int v, int cw, int lane) {
    RawRow r;
    // the four loads (own depth + mask, halo depth + mask) are independent and issued together
    const int u = cw * 32 + lane;
    const int hu = lane == 0 ? cw * 32 + 32 : cw * 32 - 1;
    const bool row_in = v >= 0 && v < g.H;
    const bool in = row_in && u < g.W, hin = row_in && lane < 2 && hu >= 0 && hu < g.W;
    const size_t row = ((size_t)b * g.H + (row_in ? v : 0)) * g.W;
    r.d = 0.0f;
    r.dh = 0.0f;
    r.m = 1u;
    r.mh = 1u;
    if (in) r.d = __ldg(g.depth + row + u);
    if (hin) r.dh = __ldg(g.depth + row + hu);
    if (g.mask) {
        if (in) r.m = __ldg(g.mask + row + u);
        if (hin) r.mh = __ldg(g.mask + row + hu);
    }
    return r;
}

__device__ __forceinline__ Row finish_row(const RawRow& q) {
    Row r;
    r.c = (q.m != 0u && q.d > 0.0f) ? q.d : 0.0f;
    const float halo = (q.mh != 0u && q.dh > 0.0f) ? q.dh : 0.0f;
    r.halo_r = __shfl_sync(0xffffffffu, halo, 0);
    r.halo_l = __shfl_sync(0xffffffffu, halo, 1);
    r.bits = __ballot_sync(0xffffffffu, r.c > 0.0f);
    return r;
}

__device__ __forceinline__ Row load_row(const D2mGrid& g, int b, int v, int cw, int lane) {
    return finish_row(load_raw(g, b, v, cw, lane));
}

struct Hood {
    float c, e, w, s, se, n;  // centre, (u+1, v), (u-1, v), (u, v+1), (u+1, v+1), (u, v-1): vertex depth or 0
};

// warp-uniform: no vertex in rows v, v+1 of the word and its halo => no pixel of the word owns a face
__device__ __forceinline__ bool hood_empty(const Row& cur, const Row& next) {
    return (cur.bits | next.bits) == 0u && !(cur.halo_r > 0.0f) && !(cur.halo_l > 0.0f) && !(next.halo_r > 0.0f);
}

__device__ __forceinline__ Hood make_hood(const Row& prev, const Row& cur, const Row& next, int lane) {
    Hood h;
    h.c = cur.c;
    h.s = next.c;
    h.n = prev.c;
    const float e = __shfl_down_sync(0xffffffffu, cur.c, 1), se = __shfl_down_sync(0xffffffffu, next.c, 1),
                w = __shfl_up_sync(0xffffffffu, cur.c, 1);
    h.e = lane == 31 ? cur.halo_r : e;
    h.se = lane == 31 ? next.halo_r : se;
    h.w = lane == 0 ? cur.halo_l : w;
    return h;
}

__device__ __forceinline__ bool depth_ok(const D2mGrid& g, float a, float b, float c) {
    // pipeline.py:686-690: all three edges, float32 differences against the float32 threshold
    return fabsf(__fsub_rn(a, b)) <= g.max_depth_distance && fabsf(__fsub_rn(c, b)) <= g.max_depth_distance &&
           fabsf(__fsub_rn(a, c)) <= g.max_depth_distance;
}

__device__ __forceinline__ void push_face(const D2mGrid& g, PixelFaces& f, int slot, float da, float db, float dc) {
    if (!depth_ok(g, da, db, dc)) return;
    f.slots |= 1u << slot;
    ++f.n;
}

// Faces owned by pixel (u, v): the quad whose top-left corner it is, and -- if the pixel itself is absent -- the hole
// around it.
__device__ __forceinline__ void pixel_faces(const D2mGrid& g, int b, int u, int v, const Hood& h, PixelFaces& f) {
    f.n = 0;
    f.slots = 0u;
    if (u >= g.W) return;
    const unsigned int ov = g.diag ? g.diag[((size_t)b * g.H + v) * g.W + u] : 0u;
    const float da = h.c, db = h.e, dc = h.s, dd = h.se;  // A = (u, v), B = (u+1, v), C = (u, v+1), D = (u+1, v+1)
    const bool va = da > 0.0f, vb = db > 0.0f, vc = dc > 0.0f, vd = dd > 0.0f;
    const int nq = (int)va + (int)vb + (int)vc + (int)vd;
    if (nq == 4) {
        const bool anti = g.diag ? (ov & 1u) : !(fabsf(__fsub_rn(da, dd)) <= fabsf(__fsub_rn(db, dc)));
        if (!anti) {
            push_face(g, f, 0, da, dd, db);
            push_face(g, f, 1, da, dc, dd);
        } else {
            push_face(g, f, 2, da, dc, db);
            push_face(g, f, 3, db, dc, dd);
        }
    } else if (nq == 3) {
        if (!vd) push_face(g, f, 2, da, dc, db);
        else if (!va) push_face(g, f, 3, db, dc, dd);
        else if (!vb) push_face(g, f, 1, da, dc, dd);
        else push_face(g, f, 0, da, dd, db);
    }
    if (!va && g.bridges) {
        const float dn = h.n, de = h.e, ds = h.s, dw = h.w;
        const bool vn = dn > 0.0f, ve = de > 0.0f, vs = ds > 0.0f, vw = dw > 0.0f;
        const int nh = (int)vn + (int)ve + (int)vs + (int)vw;
        if (nh == 4) {
            const bool vertical = g.diag ? ((ov >> 1) & 1u) : !(fabsf(__fsub_rn(dw, de)) <= fabsf(__fsub_rn(dn, ds)));
            if (!vertical) {
                push_face(g, f, 4, dw, de, dn);
                push_face(g, f, 5, dw, ds, de);
            } else {
                push_face(g, f, 6, dn, ds, de);
                push_face(g, f, 7, dn, dw, ds);
            }
        } else if (nh == 3) {
            if (!vn) push_face(g, f, 5, dw, ds, de);
            else if (!vs) push_face(g, f, 4, dw, de, dn);
            else if (!ve) push_face(g, f, 7, dn, dw, ds);
            else push_face(g, f, 6, dn, ds, de);
        }
    }
}

// One warp per word column cw and kD2mRows rows; the grid is (words of a row / warps per block, rows / kD2mRows, B), so
// no thread divides (r01: a 64-bit division + a 32-bit one per thread were a third of the kernel).
constexpr int kD2mThreads = 128;

__global__ void __launch_bounds__(kD2mThreads) d2m_count_kernel(const D2mGrid g) {
    const int lane = threadIdx.x & 31;
    const int cw = (int)blockIdx.x * (kD2mThreads / 32) + (int)(threadIdx.x >> 5);
    if (cw >= g.wpr) return;
    const int b = (int)blockIdx.z;
    const int v0 = (int)blockIdx.y * kD2mRows, v1 = min(v0 + kD2mRows, g.H);
    const bool faces = !g.verts_only;
    {
        // Most 32 x 8 pixel blocks of a foreground frame hold no vertex at all: look at the block and its one-pixel rim
        // first (mask bytes when there is a mask, depths otherwise; all ten rows requested together) and leave with
        // zeros.  (r01: the row-by-row walk spent ~70 warp-instructions per empty word row, 80 % of the rows.)
        bool any = false;
        if (g.mask && (g.W & 3) == 0 && (reinterpret_cast<uintptr_t>(g.mask) & 3) == 0) {
            // ten 4-byte words per row (pixels 32 cw - 4 .. 32 cw + 35: the word and, generously, its rim), ten rows:
            // 100 words, one per lane in four rounds
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k = i * 32 + lane, r = k / 10, c = k - r * 10;
                const int v = v0 - 1 + r, uu = cw * 32 - 4 + 4 * c;
                if (k < 100 && v >= 0 && v < g.H && v <= v1 && uu >= 0 && uu < g.W)
                    any |= __ldg(reinterpret_cast<const unsigned int*>(g.mask + ((size_t)b * g.H + v) * g.W + uu)) != 0u;
            }
        } else {
            const int u = cw * 32 + lane, hu = lane == 0 ? cw * 32 + 32 : cw * 32 - 1;
            const bool hin = lane < 2 && hu >= 0 && hu < g.W;
#pragma unroll
            for (int r = -1; r <= kD2mRows; ++r) {
                const int v = v0 + r;
                if (v < 0 || v >= g.H || v > v1) continue;
                const size_t row = ((size_t)b * g.H + v) * g.W;
                if (g.mask) {
                    if (u < g.W) any |= __ldg(g.mask + row + u) != 0;
                    if (hin) any |= __ldg(g.mask + row + hu) != 0;
                } else {
                    if (u < g.W) any |= __ldg(g.depth + row + u) > 0.0f;
                    if (hin) any |= __ldg(g.depth + row + hu) > 0.0f;
                }
            }
        }
        if (!__any_sync(0xffffffffu, any)) {
            if (lane < v1 - v0) {
                const size_t w = (size_t)b * g.words_per_frame + (size_t)(v0 + lane) * g.wpr + cw;
                g.wbits[w] = 0u;
                g.wcount[w] = make_uint2(0u, 0u);
            }
            return;
        }
    }
    const RawRow q0 = load_raw(g, b, v0 - 1, cw, lane), q1 = load_raw(g, b, v0, cw, lane);
    RawRow raw_next = load_raw(g, b, v0 + 1, cw, lane);
    Row prev = finish_row(q0), cur = finish_row(q1);
    size_t w = (size_t)b * g.words_per_frame + (size_t)v0 * g.wpr + cw;
    for (int v = v0; v < v1; ++v) {
        const RawRow raw_next2 = load_raw(g, b, v + 2 <= v1 ? v + 2 : -1, cw, lane);  // in flight while row v is processed
        const Row next = finish_row(raw_next);
        raw_next = raw_next2;
        unsigned int nf = 0;
        if (faces && !hood_empty(cur, next)) {
            const Hood h = make_hood(prev, cur, next, lane);
            PixelFaces f;
            pixel_faces(g, b, cw * 32 + lane, v, h, f);
            nf = (unsigned int)f.n;
            if (cw * 32 + lane < g.W) g.slots[((size_t)b * g.H + v) * g.W + cw * 32 + lane] = (uint8_t)f.slots;
            nf = __reduce_add_sync(0xffffffffu, nf);
        }
        if (lane == 0) {
            g.wbits[w] = cur.bits;
            g.wcount[w] = make_uint2(__popc(cur.bits), nf);
        }
        w += g.wpr;
        prev = cur;
        cur = next;
    }
}

struct D2mOut {
    const double* cams;  // [B][12]: R row-major (9), t (3) -- world-to-camera, as hive/geometric.py passes them
    double Kinv[9];
    double* verts;        // [V][3]
    float* verts32;       // HV_D2M_VERTS_F32: the same rounded to float32 (what a glTF export keeps), verts is null then
    int32_t* faces;       // [F][3] frame-local vertex ids
    int32_t* points2d;    // [V][2] (u, v) or null
    float* vdepth;        // [V] or null
    long long* frame_off; // [B+1][2] (vertex, face) start of every frame in the concatenated outputs
    unsigned long long n_verts, n_faces;  // the totals, or (d2m_emit_kernel<true>) the capacity of the output arrays
    const unsigned long long* d_totals;   // d2m_emit_kernel<true>: where hv_d2m_count left the totals (DEVICE)
};

// CAPPED: the host has not read the totals yet -- the outputs hold out.n_verts / out.n_faces rows (a guess from earlier
// batches), rows past that are dropped (the host sees totals > capacity afterwards and repeats the call exactly sized).
template <bool CAPPED>
__global__ void __launch_bounds__(kD2mThreads) d2m_emit_kernel(const D2mGrid g, const D2mOut out) {
    __shared__ unsigned int s_rank[6][kD2mThreads];
    const int lane = threadIdx.x & 31;
    const int cw = (int)blockIdx.x * (kD2mThreads / 32) + (int)(threadIdx.x >> 5);
    if (cw >= g.wpr) return;
    const int b = (int)blockIdx.z;
    const int v0 = (int)blockIdx.y * kD2mRows, v1 = min(v0 + kD2mRows, g.H);
    const size_t wframe = (size_t)b * g.words_per_frame;
    const size_t w0 = wframe + (size_t)v0 * g.wpr + cw;
    if (v0 == 0 && cw == 0 && lane == 0) {
        const uint2 off = g.woff[wframe];
        out.frame_off[2 * b] = off.x;
        out.frame_off[2 * b + 1] = off.y;
        if (b == g.B - 1) {
            out.frame_off[2 * g.B] = (long long)(CAPPED ? out.d_totals[0] : out.n_verts);
            out.frame_off[2 * g.B + 1] = (long long)(CAPPED ? out.d_totals[1] : out.n_faces);
        }
    }
    // counts (lane r holds row v0 + r), prefix offsets and vertex bits (lane l holds row v0 - 1 + l) of the warp's words and
    // the frame's first offset: all requested together -- most word columns are empty and leave right after
    uint2 my_cnt = make_uint2(0u, 0u), my_off = make_uint2(0u, 0u);
    unsigned int my_bits = 0u;
    if (lane < v1 - v0) my_cnt = g.wcount[w0 + (size_t)lane * g.wpr];
    {
        const int row = v0 - 1 + lane;
        if (lane < (v1 - v0) + 2 && row >= 0 && row < g.H) {
            my_off = g.woff[wframe + (size_t)row * g.wpr + cw];
            my_bits = g.wbits[wframe + (size_t)row * g.wpr + cw];
        }
    }
    const unsigned int vbase = g.woff[wframe].x;
    const unsigned int has_v = __ballot_sync(0xffffffffu, my_cnt.x != 0u), has_f = __ballot_sync(0xffffffffu, my_cnt.y != 0u);
    if ((has_v | has_f) == 0u) return;
    const int u = cw * 32 + lane;
    const unsigned int lt = (1u << lane) - 1u;
    // slot byte and depth of the lane's pixel in every row that has something, all requested before the first is used
    // (r02 ncu: with one load per row inside the loop the kernel sat at 44 % issue utilisation, 12 stall cycles per
    // instruction on those loads)
    unsigned int sl_row[kD2mRows];
    float d_row[kD2mRows];
    double cam[12];  // R row-major, t of this frame: once per warp, not once per row
#pragma unroll
    for (int k = 0; k < 12; ++k) cam[k] = __ldg(out.cams + 12 * (size_t)b + k);
#pragma unroll
    for (int r = 0; r < kD2mRows; ++r) {
        const size_t px = ((size_t)b * g.H + (v0 + r)) * g.W + u;
        sl_row[r] = (((has_f >> r) & 1u) && u < g.W) ? g.slots[px] : 0u;
        d_row[r] = (((has_v >> r) & 1u) && u < g.W) ? __ldg(g.depth + px) : 0.0f;
    }
#pragma unroll
    for (int r = 0; r < kD2mRows; ++r) {
        if (!(((has_v | has_f) >> r) & 1u)) continue;
        const int v = v0 + r;
        const uint2 off = make_uint2(__shfl_sync(0xffffffffu, my_off.x, r + 1), __shfl_sync(0xffffffffu, my_off.y, r + 1));
        const unsigned int bits = __shfl_sync(0xffffffffu, my_bits, r + 1);
        if ((has_f >> r) & 1u) {
            // the slot mask the count kernel left for this pixel: nothing is tested again
            const unsigned int sl = sl_row[r];
            const unsigned int nf = __popc(sl);
            unsigned int inc = nf;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int up = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += up;
            }
            // frame-local ranks (row-major rank among the frame's vertices, np.nonzero order, geometric.py:121) of the six
            // pixels a face can touch, from the vertex bits of the word in the three rows and the ranks of their first
            // pixels.  A corner is always a vertex, so rank(u - 1) = rank(u) - 1 and rank(u + 1) = rank(u) + valid(u).
            const unsigned int bits_prev = __shfl_sync(0xffffffffu, my_bits, r), bits_next = __shfl_sync(0xffffffffu, my_bits, r + 2);
            const unsigned int first_prev = __shfl_sync(0xffffffffu, my_off.x, r) - vbase, first_cur = off.x - vbase;
            const unsigned int first_next = __shfl_sync(0xffffffffu, my_off.x, r + 2) - vbase;
            if (sl) {
                const unsigned int rc = first_cur + __popc(bits & lt), rs = first_next + __popc(bits_next & lt);
                // the six ranks go through shared memory so that a corner is one indexed load (as a 6-way select chain
                // they were 27 % of this kernel's instructions)
                s_rank[0][threadIdx.x] = rc;
                s_rank[1][threadIdx.x] = rc + ((bits >> lane) & 1u);
                s_rank[2][threadIdx.x] = rs;
                s_rank[3][threadIdx.x] = rs + ((bits_next >> lane) & 1u);
                s_rank[4][threadIdx.x] = first_prev + __popc(bits_prev & lt);
                s_rank[5][threadIdx.x] = rc - 1u;
                size_t fi = (size_t)off.y + inc - nf;
                for (unsigned int rest = sl; rest; rest &= rest - 1u, ++fi) {
                    if (CAPPED && fi >= out.n_faces) break;
                    const int slot = __ffs(rest) - 1;
                    const unsigned int code = (unsigned int)((slot < 4 ? kQuadCorners >> (9 * slot) : kHoleCorners >> (9 * (slot - 4))) & 511ull);
#pragma unroll
                    for (int c = 0; c < 3; ++c) out.faces[3 * fi + c] = (int32_t)s_rank[(code >> (3 * c)) & 7u][threadIdx.x];
                }
            }
        }
        if ((bits >> lane) & 1u) {  // image2world, hive/geometric.py:200-204 (float64)
            const float d = d_row[r];
            const size_t vi = off.x + __popc(bits & lt);
            if (CAPPED && vi >= out.n_verts) continue;
            const double du = (double)u, dv = (double)v, dd = (double)d;
            double X[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double pc = fma(out.Kinv[3 * k + 2], 1.0, fma(out.Kinv[3 * k + 1], dv, __dmul_rn(out.Kinv[3 * k], du)));
                X[k] = __dsub_rn(__dmul_rn(dd, pc), cam[9 + k]);
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {  // R^T @ X
                const double w = fma(cam[6 + k], X[2], fma(cam[3 + k], X[1], __dmul_rn(cam[k], X[0])));
                if (out.verts32) out.verts32[3 * vi + k] = __double2float_rn(w);
                else out.verts[3 * vi + k] = w;
            }
            if (out.points2d) {
                out.points2d[2 * vi] = u;
                out.points2d[2 * vi + 1] = v;
            }
            if (out.vdepth) out.vdepth[vi] = d;
        }
    }
}

// ---- elementwise projection helpers (hive/geometric.py:155-206) and the generic face filter (pipeline.py:670-694) --------
__global__ void __launch_bounds__(256) image2world_kernel(const double* __restrict__ pts, const double* __restrict__ depth,
                                                          long long n, const double* __restrict__ cam /* Kinv 9, R 9, t 3 */,
                                                          double scale, double* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double u = __dmul_rn(pts[2 * i], scale), v = __dmul_rn(pts[2 * i + 1], scale), d = depth[i];
    const double *Kinv = cam, *R = cam + 9, *t = cam + 18;
    double X[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double pc = fma(Kinv[3 * k + 2], 1.0, fma(Kinv[3 * k + 1], v, __dmul_rn(Kinv[3 * k], u)));
        X[k] = __dsub_rn(__dmul_rn(d, pc), t[k]);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) out[3 * i + k] = fma(R[6 + k], X[2], fma(R[3 + k], X[1], __dmul_rn(R[k], X[0])));
}

__global__ void __launch_bounds__(256) world2image_kernel(const double* __restrict__ pts, long long n,
                                                          const double* __restrict__ cam /* K 9, R 9, t 3 */, double scale,
                                                          int round_half_even, double* __restrict__ uv,
                                                          double* __restrict__ depth) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *K = cam, *R = cam + 9, *t = cam + 18;
    const double p0 = pts[3 * i], p1 = pts[3 * i + 1], p2 = pts[3 * i + 2];
    double c[3], q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
        c[k] = __dadd_rn(fma(R[3 * k + 2], p2, fma(R[3 * k + 1], p1, __dmul_rn(R[3 * k], p0))), t[k]);
#pragma unroll
    for (int k = 0; k < 3; ++k) q[k] = fma(K[3 * k + 2], c[2], fma(K[3 * k + 1], c[1], __dmul_rn(K[3 * k], c[0])));
    double a = (q[0] / q[2]) / scale, b = (q[1] / q[2]) / scale;
    if (round_half_even) {
        a = rint(a);
        b = rint(b);
    }
    uv[2 * i] = a;
    uv[2 * i + 1] = b;
    depth[i] = q[2];
}

struct FilterArgs {
    const int32_t* pts;   // [N][2]
    const void* depth;    // [N] float32 or float64
    const int32_t* faces; // [F][3]
    long long n_faces;
    double max_pixel_distance, max_depth_distance;
    unsigned int* wbits;
    uint2* wcount;
    uint2* woff;
    int32_t* out;
};

template <typename T>
__device__ __forceinline__ bool face_ok(const FilterArgs& a, long long f) {
    const int i0 = a.faces[3 * f], i1 = a.faces[3 * f + 1], i2 = a.faces[3 * f + 2];
    const int ea[3] = {i0, i2, i0}, eb[3] = {i1, i1, i2};  // pipeline.py:686: edges (0,1), (2,1), (0,2)
    const T* depth = static_cast<const T*>(a.depth);
    const T max_dd = (T)a.max_depth_distance;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const long long dx = (long long)a.pts[2 * ea[k]] - a.pts[2 * eb[k]], dy = (long long)a.pts[2 * ea[k] + 1] - a.pts[2 * eb[k] + 1];
        ok = ok && (sqrt((double)(dx * dx + dy * dy)) <= a.max_pixel_distance);
        const T dd = depth[ea[k]] - depth[eb[k]];
        ok = ok && ((dd < (T)0 ? -dd : dd) <= max_dd);
    }
    return ok;
}

template <typename T, bool kEmit>
__global__ void __launch_bounds__(256) filter_faces_kernel(const FilterArgs a) {
    const int lane = threadIdx.x & 31;
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long w = f >> 5;
    if (w * 32 >= a.n_faces) return;
    const bool keep = f < a.n_faces && face_ok<T>(a, f);
    const unsigned int bits = __ballot_sync(0xffffffffu, keep);
    if (!kEmit) {
        if (lane == 0) a.wcount[w] = make_uint2(__popc(bits), 0u);
    } else if (keep) {
        const size_t o = a.woff[w].x + __popc(bits & ((1u << lane) - 1u));
        a.out[3 * o] = a.faces[3 * f];
        a.out[3 * o + 1] = a.faces[3 * f + 1];
        a.out[3 * o + 2] = a.faces[3 * f + 2];
    }
}

}  // namespace hv

// ---- C ABI ---------------------------------------------------------------------------------------------------------------
namespace {

struct D2mLayout {
    int wpr, words_per_frame;
    long long n_words;
    size_t off_bits, off_count, off_off, off_bsum, off_slots, total;
};

size_t d2m_align(size_t v) { return (v + 255) & ~(size_t)255; }

int d2m_layout(int B, int H, int W, D2mLayout* L) {
    HV_REQUIRE(B >= 1 && H >= 1 && W >= 1 && (long long)H * W < (1ll << 30) && B <= 65535 && H <= 65535,
               "depth-to-mesh: bad batch %d x %d x %d", B, H, W);
    L->wpr = (W + 31) / 32;
    L->words_per_frame = H * L->wpr;
    L->n_words = (long long)B * L->words_per_frame;
    size_t o = 0;
    L->off_count = o;
    o = d2m_align(o + (size_t)L->n_words * sizeof(uint2));
    L->off_off = o;
    o = d2m_align(o + (size_t)L->n_words * sizeof(uint2));
    L->off_bits = o;
    o = d2m_align(o + (size_t)L->n_words * sizeof(unsigned int));
    L->off_bsum = o;
    o = d2m_align(o + (size_t)hv::scan_scratch_pairs(L->n_words) * sizeof(uint2));
    L->off_slots = o;
    o = d2m_align(o + (size_t)B * H * W);
    L->total = o;
    return HV_OK;
}

int d2m_grid(hv::D2mGrid* g, const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
             float max_depth_distance, int flags, void* scratch, int64_t scratch_bytes, D2mLayout* L) {
    int rc = d2m_layout(B, H, W, L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(depth && scratch, "depth-to-mesh: null pointer");
    HV_REQUIRE(((uintptr_t)scratch & 255) == 0, "depth-to-mesh: scratch must be 256-byte aligned");
    HV_REQUIRE(max_depth_distance >= 0.0f, "depth-to-mesh: max_depth_distance must be >= 0");
    if ((size_t)scratch_bytes < L->total) {
        hv::set_error("depth-to-mesh: scratch of %lld bytes, need %lld", (long long)scratch_bytes, (long long)L->total);
        return HV_E_CAPACITY;
    }
    char* s = static_cast<char*>(scratch);
    g->depth = depth;
    g->mask = mask;
    g->diag = diag;
    g->B = B;
    g->H = H;
    g->W = W;
    g->wpr = L->wpr;
    g->words_per_frame = L->words_per_frame;
    g->n_words = L->n_words;
    g->max_depth_distance = max_depth_distance;
    g->bridges = (flags & HV_D2M_BRIDGES) ? 1 : 0;
    g->verts_only = (flags & HV_D2M_VERTS_ONLY) ? 1 : 0;
    g->wbits = reinterpret_cast<unsigned int*>(s + L->off_bits);
    g->wcount = reinterpret_cast<uint2*>(s + L->off_count);
    g->woff = reinterpret_cast<uint2*>(s + L->off_off);
    g->bsum = reinterpret_cast<uint2*>(s + L->off_bsum);
    g->slots = reinterpret_cast<uint8_t*>(s + L->off_slots);
    return HV_OK;
}

}  // namespace

extern "C" int hv_d2m_scratch_bytes(int B, int H, int W, int64_t* h_bytes) {
    D2mLayout L;
    int rc = d2m_layout(B, H, W, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(h_bytes, "hv_d2m_scratch_bytes: null output");
    *h_bytes = (int64_t)L.total;
    return HV_OK;
}

extern "C" int hv_d2m_count(const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
                            float max_depth_distance, int flags, void* scratch, int64_t scratch_bytes,
                            unsigned long long* totals, void* stream) {
    hv::D2mGrid g;
    D2mLayout L;
    int rc = d2m_grid(&g, depth, mask, diag, B, H, W, max_depth_distance, flags, scratch, scratch_bytes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(totals, "hv_d2m_count: null totals");
    cudaStream_t st = hv::as_stream(stream);
    const dim3 grid((unsigned int)((g.wpr + hv::kD2mThreads / 32 - 1) / (hv::kD2mThreads / 32)), (unsigned int)((H + hv::kD2mRows - 1) / hv::kD2mRows), (unsigned int)B);
    hv::d2m_count_kernel<<<grid, hv::kD2mThreads, 0, st>>>(g);
    hv::exclusive_scan_u2(g.wcount, g.woff, g.bsum, g.n_words, totals, st);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

static int d2m_emit_impl(const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
                         float max_depth_distance, int flags, const double* h_kinv, const double* cams, void* scratch,
                         int64_t scratch_bytes, int64_t n_verts, int64_t n_faces, double* verts, int32_t* faces,
                         int32_t* points2d, float* vdepth, long long* frame_offsets, const unsigned long long* d_totals,
                         void* stream) {
    hv::D2mGrid g;
    D2mLayout L;
    int rc = d2m_grid(&g, depth, mask, diag, B, H, W, max_depth_distance, flags, scratch, scratch_bytes, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(h_kinv && cams && verts && frame_offsets && (faces || n_faces == 0), "hv_d2m_emit: null pointer");
    HV_REQUIRE(n_verts >= 0 && n_faces >= 0, "hv_d2m_emit: bad counts");
    hv::D2mOut out;
    out.cams = cams;
    for (int k = 0; k < 9; ++k) out.Kinv[k] = h_kinv[k];
    out.verts = (flags & HV_D2M_VERTS_F32) ? nullptr : verts;
    out.verts32 = (flags & HV_D2M_VERTS_F32) ? reinterpret_cast<float*>(verts) : nullptr;
    out.faces = faces;
    out.points2d = points2d;
    out.vdepth = vdepth;
    out.frame_off = frame_offsets;
    out.n_verts = (unsigned long long)n_verts;
    out.n_faces = (unsigned long long)n_faces;
    out.d_totals = d_totals;
    const dim3 grid((unsigned int)((g.wpr + hv::kD2mThreads / 32 - 1) / (hv::kD2mThreads / 32)), (unsigned int)((H + hv::kD2mRows - 1) / hv::kD2mRows), (unsigned int)B);
    if (d_totals) hv::d2m_emit_kernel<true><<<grid, hv::kD2mThreads, 0, hv::as_stream(stream)>>>(g, out);
    else hv::d2m_emit_kernel<false><<<grid, hv::kD2mThreads, 0, hv::as_stream(stream)>>>(g, out);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_d2m_emit(const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
                           float max_depth_distance, int flags, const double* h_kinv, const double* cams, void* scratch,
                           int64_t scratch_bytes, int64_t n_verts, int64_t n_faces, double* verts, int32_t* faces,
                           int32_t* points2d, float* vdepth, long long* frame_offsets, void* stream) {
    return d2m_emit_impl(depth, mask, diag, B, H, W, max_depth_distance, flags, h_kinv, cams, scratch, scratch_bytes, n_verts,
                         n_faces, verts, faces, points2d, vdepth, frame_offsets, nullptr, stream);
}

// hv_d2m_emit without the host having read the totals: the outputs hold cap_verts / cap_faces rows, `totals` is the DEVICE
// pair hv_d2m_count wrote on the same stream.  Rows beyond the capacity are dropped; frame_offsets still end in the true
// totals, so the caller sees the overflow when it reads them (and repeats with hv_d2m_emit, exactly sized).
extern "C" int hv_d2m_emit_capped(const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
                                  float max_depth_distance, int flags, const double* h_kinv, const double* cams,
                                  void* scratch, int64_t scratch_bytes, int64_t cap_verts, int64_t cap_faces, double* verts,
                                  int32_t* faces, int32_t* points2d, float* vdepth, long long* frame_offsets,
                                  const unsigned long long* totals, void* stream) {
    HV_REQUIRE(totals, "hv_d2m_emit_capped: null totals");
    return d2m_emit_impl(depth, mask, diag, B, H, W, max_depth_distance, flags, h_kinv, cams, scratch, scratch_bytes, cap_verts,
                         cap_faces, verts, faces, points2d, vdepth, frame_offsets, totals, stream);
}

extern "C" int hv_image2world(const double* points2d, const double* depth, int64_t n, const double* cam, double scale,
                              double* out, void* stream) {
    if (n == 0) return HV_OK;
    HV_REQUIRE(points2d && depth && cam && out && n > 0, "hv_image2world: bad arguments");
    hv::image2world_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, hv::as_stream(stream)>>>(points2d, depth, n, cam, scale, out);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_world2image(const double* points, int64_t n, const double* cam, double scale, int round_half_even,
                              double* uv, double* depth, void* stream) {
    if (n == 0) return HV_OK;
    HV_REQUIRE(points && cam && uv && depth && n > 0, "hv_world2image: bad arguments");
    HV_REQUIRE(scale != 0.0, "hv_world2image: scale_factor must be non-zero");
    hv::world2image_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, hv::as_stream(stream)>>>(points, n, cam, scale,
                                                                                              round_half_even, uv, depth);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_filter_faces_scratch_bytes(int64_t n_faces, int64_t* h_bytes) {
    HV_REQUIRE(n_faces >= 0 && h_bytes, "hv_filter_faces_scratch_bytes: bad arguments");
    const long long words = (n_faces + 31) / 32;
    *h_bytes = (int64_t)(d2m_align((size_t)words * 8) * 2 + d2m_align((size_t)hv::scan_tiles(words) * 8 + 8));
    return HV_OK;
}

/* phase 0: count (totals[0] = kept faces); phase 1: write the kept faces, order preserved, to `out`. */
extern "C" int hv_filter_faces(const int32_t* points2d, const void* depth, int depth_is_f64, const int32_t* faces,
                               int64_t n_faces, double max_pixel_distance, double max_depth_distance, void* scratch,
                               int64_t scratch_bytes, int phase, unsigned long long* totals, int32_t* out, void* stream) {
    if (n_faces == 0) return HV_OK;
    HV_REQUIRE(points2d && depth && faces && scratch && n_faces > 0, "hv_filter_faces: bad arguments");
    int64_t need = 0;
    hv_filter_faces_scratch_bytes(n_faces, &need);
    if (scratch_bytes < need) {
        hv::set_error("hv_filter_faces: scratch of %lld bytes, need %lld", (long long)scratch_bytes, (long long)need);
        return HV_E_CAPACITY;
    }
    const long long words = (n_faces + 31) / 32;
    char* s = static_cast<char*>(scratch);
    hv::FilterArgs a;
    a.pts = points2d;
    a.depth = depth;
    a.faces = faces;
    a.n_faces = n_faces;
    a.max_pixel_distance = max_pixel_distance;
    a.max_depth_distance = max_depth_distance;
    a.wbits = nullptr;
    a.wcount = reinterpret_cast<uint2*>(s);
    a.woff = reinterpret_cast<uint2*>(s + d2m_align((size_t)words * 8));
    uint2* bsum = reinterpret_cast<uint2*>(s + 2 * d2m_align((size_t)words * 8));
    a.out = out;
    cudaStream_t st = hv::as_stream(stream);
    const unsigned int blocks = (unsigned int)((words * 32 + 255) / 256);
    if (phase == 0) {
        HV_REQUIRE(totals, "hv_filter_faces: null totals");
        if (depth_is_f64) hv::filter_faces_kernel<double, false><<<blocks, 256, 0, st>>>(a);
        else hv::filter_faces_kernel<float, false><<<blocks, 256, 0, st>>>(a);
        hv::exclusive_scan_u2(a.wcount, a.woff, bsum, words, totals, st);
    } else {
        HV_REQUIRE(out, "hv_filter_faces: null output");
        if (depth_is_f64) hv::filter_faces_kernel<double, true><<<blocks, 256, 0, st>>>(a);
        else hv::filter_faces_kernel<float, true><<<blocks, 256, 0, st>>>(a);
    }
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}
