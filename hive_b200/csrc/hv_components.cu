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
// Kernels: edge keys (3 per face) -> [sorted by the caller: radix sort is plumbing] -> adjacent pairs from runs of exactly
// two equal keys, hooked into a lock-free union-find (atomicCAS on the larger root) -> flatten -> component sizes ->
// selection -> keep mask.  Everything is a streaming pass over 3F keys or F faces.
#include "hv_common.cuh"

namespace hv {

__global__ void __launch_bounds__(256) cc_edge_keys_kernel(const int32_t* __restrict__ faces, long long n_faces,
                                                           long long n_verts, unsigned long long* __restrict__ keys) {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_faces) return;
    const long long a = faces[3 * f], b = faces[3 * f + 1], c = faces[3 * f + 2];
    const long long e0a = a < b ? a : b, e0b = a < b ? b : a;
    const long long e1a = b < c ? b : c, e1b = b < c ? c : b;
    const long long e2a = c < a ? c : a, e2b = c < a ? a : c;
    // (edge, face) packed so that one sort of plain 64-bit keys groups the uses of an edge: edge id in the high bits
    // would need 2*log2(V) + log2(F) bits, so the face id travels separately: keys[3f + k] = edge, value = slot index
    keys[3 * f + 0] = (unsigned long long)(e0a * n_verts + e0b);
    keys[3 * f + 1] = (unsigned long long)(e1a * n_verts + e1b);
    keys[3 * f + 2] = (unsigned long long)(e2a * n_verts + e2b);
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

// sorted_keys[i] ascending, slot[i] = index into the unsorted key array (face = slot / 3).
__global__ void __launch_bounds__(256) cc_hook_kernel(const unsigned long long* __restrict__ sorted_keys,
                                                      const long long* __restrict__ slot, long long n,
                                                      unsigned int* parent, unsigned int* has_adj) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= n) return;
    const unsigned long long k = sorted_keys[i];
    if (sorted_keys[i + 1] != k) return;
    if (i > 0 && sorted_keys[i - 1] == k) return;      // not the start of the run
    if (i + 2 < n && sorted_keys[i + 2] == k) return;  // an edge used by three or more faces links nothing
    const unsigned int fa = (unsigned int)(slot[i] / 3), fb = (unsigned int)(slot[i + 1] / 3);
    has_adj[fa] = 1u;
    has_adj[fb] = 1u;
    if (fa != fb) cc_union(parent, fa, fb);
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

}  // namespace hv

extern "C" int hv_cc_edge_keys(const int32_t* faces, int64_t n_faces, int64_t n_verts, unsigned long long* keys,
                               void* stream) {
    HV_REQUIRE(faces && keys && n_faces > 0 && n_verts > 0, "hv_cc_edge_keys: bad arguments");
    HV_REQUIRE(n_verts < (1ll << 31) && n_faces < (1ll << 31) / 3, "hv_cc_edge_keys: mesh too large");
    hv::cc_edge_keys_kernel<<<(unsigned)((n_faces + 255) / 256), 256, 0, hv::as_stream(stream)>>>(faces, n_faces, n_verts,
                                                                                                 keys);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_cc_scratch_bytes(int64_t n_faces, int64_t* h_bytes) {
    HV_REQUIRE(h_bytes && n_faces > 0, "hv_cc_scratch_bytes: bad arguments");
    *h_bytes = 3 * n_faces * (int64_t)sizeof(unsigned int) + 256;
    return HV_OK;
}

extern "C" int hv_cc_label(const unsigned long long* sorted_keys, const int64_t* slot, int64_t n_faces, int min_len,
                           int largest_only, void* scratch, int64_t scratch_bytes, uint8_t* keep, uint32_t* labels,
                           void* stream) {
    HV_REQUIRE(sorted_keys && slot && keep && scratch && n_faces > 0 && min_len >= 0, "hv_cc_label: bad arguments");
    int64_t need = 0;
    hv_cc_scratch_bytes(n_faces, &need);
    HV_REQUIRE(scratch_bytes >= need && ((uintptr_t)scratch & 7) == 0, "hv_cc_label: scratch too small or unaligned");
    unsigned long long* best = reinterpret_cast<unsigned long long*>(scratch);
    unsigned int* parent = reinterpret_cast<unsigned int*>(static_cast<char*>(scratch) + 256);
    unsigned int* has_adj = parent + n_faces;
    unsigned int* size = has_adj + n_faces;
    cudaStream_t st = hv::as_stream(stream);
    const unsigned int fb = (unsigned int)((n_faces + 255) / 256);
    const unsigned int eb = (unsigned int)((3 * n_faces + 255) / 256);
    hv::cc_init_kernel<<<fb, 256, 0, st>>>(parent, has_adj, size, n_faces, best);
    hv::cc_hook_kernel<<<eb, 256, 0, st>>>(sorted_keys, reinterpret_cast<const long long*>(slot), 3 * n_faces, parent,
                                           has_adj);
    hv::cc_flatten_kernel<<<fb, 256, 0, st>>>(parent, has_adj, size, n_faces);
    if (largest_only) hv::cc_best_kernel<<<fb, 256, 0, st>>>(parent, size, n_faces, (unsigned int)min_len, best);
    hv::cc_mask_kernel<<<fb, 256, 0, st>>>(parent, has_adj, size, n_faces, (unsigned int)min_len, largest_only, best, keep,
                                           labels);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}
