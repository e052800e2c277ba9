// Connected components of a triangle mesh over its face-adjacency graph.
//
// Replaces the body of Pipeline._cleanup_with_connected_components (hive/pipeline.py:741-779), which builds a
// trimesh.Trimesh and calls trimesh.graph.connected_components(mesh.face_adjacency, min_len=min_components):
//   * two faces are adjacent iff they share an (undirected) edge that exactly two faces use;
//   * nodes of the graph are the faces that have at least one adjacent face -- a face without neighbours is in no
//     component and is always dropped;
//   * components with fewer than `min_len` faces are discarded; for an object only the largest survivor is kept
//     (first in label order on a tie, i.e. the one containing the lowest face index), for the background all of them.
// trimesh is not vendored and not installed (parity unpinned); the statement above is what oracle/d2m_oracle.py restates
// with scipy.sparse.csgraph.
//
// Kernels: every face edge is inserted into an open-addressing hash table keyed by its (min vertex, max vertex) pair --
// the entry counts the uses of the edge and remembers the first two faces -> entries with exactly two uses hook their
// faces into a lock-free union-find (atomicCAS on the larger root) -> flatten -> component sizes -> selection -> keep
// mask.  No sort (r01 sorted the 3F edge keys with torch.sort); which of two faces lands in which slot of an entry is a
// race, the components are not.  Everything is a streaming pass over 3F edges, the table, or F faces.
#include "hv_common.cuh"

namespace hv {

constexpr unsigned long long kCcEmpty = ~0ull;

__device__ __forceinline__ unsigned long long cc_hash(unsigned long long k) {  // splitmix64 finaliser
    k ^= k >> 30;
    k *= 0xbf58476d1ce4e5b9ull;
    k ^= k >> 27;
    k *= 0x94d049bb133111ebull;
    k ^= k >> 31;
    return k;
}

// One thread per (face, edge): find or claim the edge's entry, count the use, remember the face if it is one of the first two.
__global__ void __launch_bounds__(256) cc_insert_kernel(const int32_t* __restrict__ faces, long long n_faces, long long n_verts,
                                                        unsigned long long* keys, unsigned int* cnt, unsigned int* fa,
                                                        unsigned int* fb, unsigned long long mask) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * n_faces) return;
    const long long f = i / 3;
    const int k = (int)(i - 3 * f);
    const long long a = faces[3 * f + k], b = faces[3 * f + (k == 2 ? 0 : k + 1)];
    const unsigned long long key = (unsigned long long)((a < b ? a : b) * n_verts + (a < b ? b : a));
    unsigned long long h = cc_hash(key) & mask;
    for (;;) {
        const unsigned long long old = atomicCAS(keys + h, kCcEmpty, key);
        if (old == kCcEmpty || old == key) {
            const unsigned int c = atomicAdd(cnt + h, 1u);
            if (c == 0u) fa[h] = (unsigned int)f;
            else if (c == 1u) fb[h] = (unsigned int)f;
            return;
        }
        h = (h + 1) & mask;
    }
}

__device__ __forceinline__ unsigned int cc_find(volatile unsigned int* parent, unsigned int x) {
    unsigned int p = parent[x];
    while (p != x) {
        x = p;
        p = parent[x];
    }
    return x;
}

__device__ __forceinline__ void cc_union(unsigned int* parent, unsigned int a, unsigned int b) {
    for (;;) {
        a = cc_find(parent, a);
        b = cc_find(parent, b);
        if (a == b) return;
        if (a < b) {
            const unsigned int t = a;
            a = b;
            b = t;
        }
        // hook the larger root under the smaller one; retry if somebody re-parented it first
        const unsigned int old = atomicCAS(parent + a, a, b);
        if (old == a) return;
    }
}

__global__ void __launch_bounds__(256) cc_init_kernel(unsigned int* __restrict__ parent, unsigned int* __restrict__ has_adj,
                                                      unsigned int* __restrict__ size, long long n_faces,
                                                      unsigned long long* best) {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f == 0) *best = 0ull;
    if (f >= n_faces) return;
    parent[f] = (unsigned int)f;
    has_adj[f] = 0u;
    size[f] = 0u;
}

// One thread per table entry: an edge used by exactly two faces links them (three or more uses link nothing).
__global__ void __launch_bounds__(256) cc_hook_kernel(const unsigned int* __restrict__ cnt, const unsigned int* __restrict__ fa,
                                                      const unsigned int* __restrict__ fb, unsigned long long n_slots,
                                                      unsigned int* parent, unsigned int* has_adj) {
    const unsigned long long h = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= n_slots || cnt[h] != 2u) return;
    const unsigned int a = fa[h], b = fb[h];
    has_adj[a] = 1u;
    has_adj[b] = 1u;
    if (a != b) cc_union(parent, a, b);
}

__global__ void __launch_bounds__(256) cc_flatten_kernel(unsigned int* parent, const unsigned int* __restrict__ has_adj,
                                                         unsigned int* size, long long n_faces) {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_faces) return;
    const unsigned int root = cc_find(parent, (unsigned int)f);
    parent[f] = root;  // roots keep parent[root] == root, so concurrent flattening stays consistent
    if (has_adj[f]) atomicAdd(size + root, 1u);
}

// Largest component (>= min_len); ties go to the smallest root = the component containing the lowest face index.
__global__ void __launch_bounds__(256) cc_best_kernel(const unsigned int* __restrict__ parent,
                                                      const unsigned int* __restrict__ size, long long n_faces,
                                                      unsigned int min_len, unsigned long long* best) {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_faces) return;
    if (parent[f] != (unsigned int)f) return;
    const unsigned int s = size[f];
    if (s == 0u || s < min_len) return;
    atomicMax(best, ((unsigned long long)s << 32) | (unsigned long long)(0xffffffffu - (unsigned int)f));
}

__global__ void __launch_bounds__(256) cc_mask_kernel(const unsigned int* __restrict__ parent,
                                                      const unsigned int* __restrict__ has_adj,
                                                      const unsigned int* __restrict__ size, long long n_faces,
                                                      unsigned int min_len, int largest_only,
                                                      const unsigned long long* __restrict__ best,
                                                      uint8_t* __restrict__ keep, unsigned int* __restrict__ labels) {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_faces) return;
    const unsigned int root = parent[f];
    bool k = has_adj[f] != 0u && size[root] >= min_len && size[root] > 0u;
    if (largest_only) {
        const unsigned long long b = *best;
        k = k && b != 0ull && root == 0xffffffffu - (unsigned int)(b & 0xffffffffull);
    }
    keep[f] = k ? 1 : 0;
    if (labels) labels[f] = has_adj[f] ? root : 0xffffffffu;
}

// ---- vertices no face refers to (trimesh.Trimesh(vertices, faces) drops them when it is built, pipeline.py:758) -----------
__global__ void __launch_bounds__(256) cc_mark_kernel(const int32_t* __restrict__ faces, long long n3, uint2* __restrict__ flags) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n3) flags[faces[i]] = make_uint2(1u, 0u);
}

__global__ void __launch_bounds__(256) cc_gather_vertices_kernel(const uint2* __restrict__ flags, const uint2* __restrict__ offs,
                                                                 long long n_verts, const double* __restrict__ verts_in,
                                                                 double* __restrict__ verts_out) {
    const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n_verts || flags[v].x == 0u) return;
    const long long o = offs[v].x;
    verts_out[3 * o] = verts_in[3 * v];
    verts_out[3 * o + 1] = verts_in[3 * v + 1];
    verts_out[3 * o + 2] = verts_in[3 * v + 2];
}

__global__ void __launch_bounds__(256) cc_remap_kernel(const int32_t* __restrict__ faces, long long n3,
                                                       const uint2* __restrict__ offs, int32_t* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n3) out[i] = (int32_t)offs[faces[i]].x;
}

__global__ void __launch_bounds__(256) cc_face_flags_kernel(const uint8_t* __restrict__ keep, long long n_faces,
                                                            uint2* __restrict__ flags) {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f < n_faces) flags[f] = make_uint2(keep[f] ? 1u : 0u, 0u);
}

// kept faces, in their original order, renumbered to the compacted vertices
__global__ void __launch_bounds__(256) cc_face_gather_kernel(const int32_t* __restrict__ faces, const uint8_t* __restrict__ keep,
                                                             long long n_faces, const uint2* __restrict__ foffs,
                                                             const uint2* __restrict__ voffs, int32_t* __restrict__ out) {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_faces || !keep[f]) return;
    const long long o = 3 * (long long)foffs[f].x;
#pragma unroll
    for (int c = 0; c < 3; ++c) out[o + c] = (int32_t)voffs[faces[3 * f + c]].x;
}

}  // namespace hv

namespace {
struct CcLayout {
    unsigned long long slots;
    size_t off_parent, off_keys, off_cnt, off_fa, off_fb, total;
};
void cc_layout(int64_t n_faces, CcLayout* L) {
    unsigned long long slots = 1024;
    while (slots < 6ull * (unsigned long long)n_faces) slots <<= 1;  // load factor <= 0.5
    L->slots = slots;
    size_t o = 256;  // [0] best
    L->off_parent = o;
    o += 3 * (size_t)n_faces * sizeof(unsigned int);
    o = (o + 255) & ~(size_t)255;
    L->off_keys = o;
    o += slots * sizeof(unsigned long long);
    L->off_cnt = o;
    o += slots * sizeof(unsigned int);
    L->off_fa = o;
    o += slots * sizeof(unsigned int);
    L->off_fb = o;
    o += slots * sizeof(unsigned int);
    L->total = o;
}
}  // namespace

extern "C" int hv_cc_scratch_bytes(int64_t n_faces, int64_t* h_bytes) {
    HV_REQUIRE(h_bytes && n_faces > 0 && n_faces < (1ll << 31) / 3, "hv_cc_scratch_bytes: bad arguments");
    CcLayout L;
    cc_layout(n_faces, &L);
    *h_bytes = (int64_t)L.total;
    return HV_OK;
}

extern "C" int hv_cc_mask(const int32_t* faces, int64_t n_faces, int64_t n_verts, int min_len, int largest_only, void* scratch,
                          int64_t scratch_bytes, uint8_t* keep, uint32_t* labels, void* stream) {
    HV_REQUIRE(faces && keep && scratch && n_faces > 0 && n_verts > 0 && min_len >= 0, "hv_cc_mask: bad arguments");
    HV_REQUIRE(n_verts < (1ll << 31) && n_faces < (1ll << 31) / 3, "hv_cc_mask: mesh too large");
    CcLayout L;
    cc_layout(n_faces, &L);
    HV_REQUIRE((size_t)scratch_bytes >= L.total && ((uintptr_t)scratch & 255) == 0, "hv_cc_mask: scratch too small or unaligned");
    char* s = static_cast<char*>(scratch);
    unsigned long long* best = reinterpret_cast<unsigned long long*>(s);
    unsigned int* parent = reinterpret_cast<unsigned int*>(s + L.off_parent);
    unsigned int* has_adj = parent + n_faces;
    unsigned int* size = has_adj + n_faces;
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(s + L.off_keys);
    unsigned int* cnt = reinterpret_cast<unsigned int*>(s + L.off_cnt);
    unsigned int* fa = reinterpret_cast<unsigned int*>(s + L.off_fa);
    unsigned int* fb = reinterpret_cast<unsigned int*>(s + L.off_fb);
    cudaStream_t st = hv::as_stream(stream);
    HV_CUDA(cudaMemsetAsync(keys, 0xff, L.slots * sizeof(unsigned long long), st));
    HV_CUDA(cudaMemsetAsync(cnt, 0, L.slots * sizeof(unsigned int), st));
    const unsigned int fblk = (unsigned int)((n_faces + 255) / 256);
    const unsigned int eblk = (unsigned int)((3 * n_faces + 255) / 256);
    hv::cc_init_kernel<<<fblk, 256, 0, st>>>(parent, has_adj, size, n_faces, best);
    hv::cc_insert_kernel<<<eblk, 256, 0, st>>>(faces, n_faces, n_verts, keys, cnt, fa, fb, L.slots - 1);
    hv::cc_hook_kernel<<<(unsigned int)((L.slots + 255) / 256), 256, 0, st>>>(cnt, fa, fb, L.slots, parent, has_adj);
    hv::cc_flatten_kernel<<<fblk, 256, 0, st>>>(parent, has_adj, size, n_faces);
    if (largest_only) hv::cc_best_kernel<<<fblk, 256, 0, st>>>(parent, size, n_faces, (unsigned int)min_len, best);
    hv::cc_mask_kernel<<<fblk, 256, 0, st>>>(parent, has_adj, size, n_faces, (unsigned int)min_len, largest_only, best, keep,
                                             labels);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_cc_compact_scratch_bytes(int64_t n_verts, int64_t* h_bytes) {
    HV_REQUIRE(h_bytes && n_verts > 0 && n_verts < (1ll << 31), "hv_cc_compact_scratch_bytes: bad arguments");
    *h_bytes = (int64_t)((2 * (size_t)n_verts + (size_t)hv::scan_scratch_pairs(n_verts)) * sizeof(uint2) + 1024);
    return HV_OK;
}

// verts_out (capacity n_verts rows) receives, in their original order, the vertices that `faces_all` refer to; faces_out the
// `faces_keep` renumbered accordingly; d_count[0] (DEVICE) how many vertices survived.
extern "C" int hv_cc_compact_vertices(const int32_t* faces_all, int64_t n_faces_all, const int32_t* faces_keep,
                                      int64_t n_faces_keep, const double* verts_in, int64_t n_verts, double* verts_out,
                                      int32_t* faces_out, unsigned long long* d_count, void* scratch, int64_t scratch_bytes,
                                      void* stream) {
    HV_REQUIRE(faces_all && verts_in && verts_out && d_count && scratch && n_faces_all > 0 && n_verts > 0 && n_faces_keep >= 0 &&
                   (n_faces_keep == 0 || (faces_keep && faces_out)), "hv_cc_compact_vertices: bad arguments");
    int64_t need = 0;
    int rc = hv_cc_compact_scratch_bytes(n_verts, &need);
    if (rc != HV_OK) return rc;
    HV_REQUIRE(scratch_bytes >= need && ((uintptr_t)scratch & 255) == 0, "hv_cc_compact_vertices: scratch too small or unaligned");
    // (the scan moves its pairs as 16-byte vectors: every array starts on a 256-byte boundary)
    uint2* flags = reinterpret_cast<uint2*>(scratch);
    uint2* offs = reinterpret_cast<uint2*>((reinterpret_cast<uintptr_t>(flags + n_verts) + 255) & ~(uintptr_t)255);
    uint2* bsum = reinterpret_cast<uint2*>((reinterpret_cast<uintptr_t>(offs + n_verts) + 255) & ~(uintptr_t)255);
    cudaStream_t st = hv::as_stream(stream);
    HV_CUDA(cudaMemsetAsync(flags, 0, (size_t)n_verts * sizeof(uint2), st));
    hv::cc_mark_kernel<<<(unsigned int)((3 * n_faces_all + 255) / 256), 256, 0, st>>>(faces_all, 3 * n_faces_all, flags);
    hv::exclusive_scan_u2(flags, offs, bsum, n_verts, d_count, st);
    hv::cc_gather_vertices_kernel<<<(unsigned int)((n_verts + 255) / 256), 256, 0, st>>>(flags, offs, n_verts, verts_in, verts_out);
    if (n_faces_keep > 0)
        hv::cc_remap_kernel<<<(unsigned int)((3 * n_faces_keep + 255) / 256), 256, 0, st>>>(faces_keep, 3 * n_faces_keep, offs, faces_out);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

// ---- the whole of Pipeline._cleanup_with_connected_components on device arrays (no host hop between its steps) ------------
namespace {
struct CleanupLayout {
    size_t off_cc, off_keep, off_vflags, off_voffs, off_fflags, off_foffs, off_bsum, total;
};
size_t a256(size_t v) { return (v + 255) & ~(size_t)255; }
int cleanup_layout(int64_t n_faces, int64_t n_verts, CleanupLayout* L) {
    int64_t cc = 0;
    int rc = hv_cc_scratch_bytes(n_faces, &cc);
    if (rc != HV_OK) return rc;
    const int64_t n_max = n_faces > n_verts ? n_faces : n_verts;
    size_t o = 0;
    L->off_cc = o;
    o = a256(o + (size_t)cc);
    L->off_keep = o;
    o = a256(o + (size_t)n_faces);
    L->off_vflags = o;
    o = a256(o + (size_t)n_verts * sizeof(uint2));
    L->off_voffs = o;
    o = a256(o + (size_t)n_verts * sizeof(uint2));
    L->off_fflags = o;
    o = a256(o + (size_t)n_faces * sizeof(uint2));
    L->off_foffs = o;
    o = a256(o + (size_t)n_faces * sizeof(uint2));
    L->off_bsum = o;
    o = a256(o + (size_t)hv::scan_scratch_pairs(n_max) * sizeof(uint2));
    L->total = o;
    return HV_OK;
}
}  // namespace

extern "C" int hv_cc_cleanup_scratch_bytes(int64_t n_faces, int64_t n_verts, int64_t* h_bytes) {
    HV_REQUIRE(h_bytes && n_faces > 0 && n_verts > 0 && n_verts < (1ll << 31), "hv_cc_cleanup_scratch_bytes: bad arguments");
    CleanupLayout L;
    int rc = cleanup_layout(n_faces, n_verts, &L);
    if (rc != HV_OK) return rc;
    *h_bytes = (int64_t)L.total;
    return HV_OK;
}

// verts_out (capacity n_verts rows) / faces_out (capacity n_faces rows); d_counts (DEVICE, 4 x uint64): [0] = vertices kept,
// [2] = faces kept.
extern "C" int hv_cc_cleanup(const int32_t* faces, int64_t n_faces, const double* verts_in, int64_t n_verts, int min_len,
                             int largest_only, double* verts_out, int32_t* faces_out, unsigned long long* d_counts,
                             void* scratch, int64_t scratch_bytes, void* stream) {
    HV_REQUIRE(faces && verts_in && verts_out && faces_out && d_counts && scratch && n_faces > 0 && n_verts > 0,
               "hv_cc_cleanup: bad arguments");
    CleanupLayout L;
    int rc = cleanup_layout(n_faces, n_verts, &L);
    if (rc != HV_OK) return rc;
    HV_REQUIRE((size_t)scratch_bytes >= L.total && ((uintptr_t)scratch & 255) == 0, "hv_cc_cleanup: scratch too small or unaligned");
    char* s = static_cast<char*>(scratch);
    uint8_t* keep = reinterpret_cast<uint8_t*>(s + L.off_keep);
    uint2* vflags = reinterpret_cast<uint2*>(s + L.off_vflags);
    uint2* voffs = reinterpret_cast<uint2*>(s + L.off_voffs);
    uint2* fflags = reinterpret_cast<uint2*>(s + L.off_fflags);
    uint2* foffs = reinterpret_cast<uint2*>(s + L.off_foffs);
    uint2* bsum = reinterpret_cast<uint2*>(s + L.off_bsum);
    int64_t cc_bytes = 0;
    hv_cc_scratch_bytes(n_faces, &cc_bytes);
    rc = hv_cc_mask(faces, n_faces, n_verts, min_len, largest_only, s + L.off_cc, cc_bytes, keep, nullptr, stream);
    if (rc != HV_OK) return rc;
    cudaStream_t st = hv::as_stream(stream);
    const unsigned int fblk = (unsigned int)((n_faces + 255) / 256), vblk = (unsigned int)((n_verts + 255) / 256);
    HV_CUDA(cudaMemsetAsync(vflags, 0, (size_t)n_verts * sizeof(uint2), st));
    hv::cc_mark_kernel<<<(unsigned int)((3 * n_faces + 255) / 256), 256, 0, st>>>(faces, 3 * n_faces, vflags);
    hv::exclusive_scan_u2(vflags, voffs, bsum, n_verts, d_counts, st);
    hv::cc_gather_vertices_kernel<<<vblk, 256, 0, st>>>(vflags, voffs, n_verts, verts_in, verts_out);
    hv::cc_face_flags_kernel<<<fblk, 256, 0, st>>>(keep, n_faces, fflags);
    hv::exclusive_scan_u2(fflags, foffs, bsum, n_faces, d_counts + 2, st);
    hv::cc_face_gather_kernel<<<fblk, 256, 0, st>>>(faces, keep, n_faces, foffs, voffs, faces_out);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}
