/*
 * hive_b200 -- C ABI of the B200-native geometric hot path of HIVE.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch / CUDA types in any signature
 * (a stream is passed as `void*` holding a cudaStream_t; NULL = the legacy default stream).  Every pointer is a
 * DEVICE pointer unless its name starts with `h_`.  The library keeps no persistent device memory: volumes, frame
 * buffers and scratch are owned by the caller (the Python host keeps them in torch tensors).
 *
 * Every entry point returns 0 on success or a negative HV_E_* code; `hv_last_error` gives the text of the last
 * failure on the calling thread.  File:line citations name the reference interface each entry point replaces
 * (paths relative to the HIVE tree; "fusion.py" is third_party/tsdf_fusion_python/fusion.py).
 */
#ifndef HIVE_B200_H
#define HIVE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HV_VERSION 100

#define HV_OK 0
#define HV_E_INVALID (-1)     /* bad argument (shape, alignment, null pointer)            -> AssertionError / ValueError */
#define HV_E_CUDA (-2)        /* CUDA runtime error                                       -> RuntimeError */
#define HV_E_NOSURFACE (-3)   /* marching cubes produced no vertex                        -> RuntimeError (skimage contract) */
#define HV_E_LEVEL_RANGE (-4) /* iso level outside [min, max] of the volume               -> ValueError  (skimage contract) */
#define HV_E_CAPACITY (-5)    /* caller-provided output / scratch buffer too small        -> RuntimeError */

int hv_version(void);
/* Copies the build id (hash of the sources and flags the library was compiled from, hive_b200/build.py:source_hash)
 * into buf; returns its length. */
int hv_build_id(char* buf, int n);
/* Copies the calling thread's last error text into buf (NUL-terminated, at most n bytes); returns its length. */
int hv_last_error(char* buf, int n);
/* Number of SMs of the current device (grid sizing is in multiples of it); <0 on error. */
int hv_sm_count(void);

/* ------------------------------------------------------------------------------------------------------------
 * Frame staging.  Replaces the per-call `cuda.InOut(color_im)`, `cuda.InOut(depth_im)` uploads (and their
 * pointless downloads) of fusion.py:251-252 and the colour fold of fusion.py:231-232.
 *
 * A frame is stored on the device as H*W 8-byte texels {float depth; uint32 b<<16|g<<8|r} so that one 64-bit
 * gather per voxel fetches both the depth and the folded colour.
 * ---------------------------------------------------------------------------------------------------------- */

/* depth: (H,W) float32 metres; rgb: (H,W,3) uint8; texels: H*W*8 bytes, 8-byte aligned. All device. */
int hv_pack_frame(const float* depth, const uint8_t* rgb, int H, int W, void* texels, void* stream);

/* Host-buffer form: async H2D of h_depth / h_rgb (pinned memory recommended) into the device staging buffers
 * d_depth (H*W*4 B) / d_rgb (H*W*3 B), then hv_pack_frame. */
int hv_upload_frame(const float* h_depth, const uint8_t* h_rgb, int H, int W, float* d_depth, uint8_t* d_rgb,
                    void* texels, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * TSDF integration.  Replaces TSDFVolume.integrate, fusion.py:217-303 (CPU semantics; the PyCUDA
 * kernel fusion.py:76-149 is the code path it supersedes).
 *
 * tsdf / weight / color : float32 planes laid out [X][Y][pitch] (z fastest, fusion.py:54-57), `pitch` >= Z floats per
 *                         column, a multiple of 4, base pointers 16-byte aligned.  Only z in [0, Z) is touched.
 * wx, wy, wz            : world coordinate of voxel index i on each axis, computed by the reference formula
 *                         float32(float64(origin) + voxel_size * i) (fusion.py:183-188) and then widened to float64
 *                         (the kernel projects in float64); wz holds `pitch` entries, 16-byte aligned.
 *                         For a z-slab shard the caller passes the slab's slice of the global table.
 * h_world2cam           : HOST, 12 doubles, rows 0..2 of inv(cam_pose) (fusion.py:263).
 * h_intr                : HOST, 4 floats fx, fy, cx, cy (float32 like fusion.py:196-198).
 * texels                : frame from hv_pack_frame.
 * trunc                 : truncation margin, 5 * voxel_size (fusion.py:39).   obs_weight: fusion.py:217.
 * flags                 : HV_INTEGRATE_INTEGER_WEIGHTS = the caller guarantees that obs_weight and every weight already
 *                         stored in the volume are non-negative integers below 2^20 (true for a volume only ever fed
 *                         integer obs_weights, e.g. the reference's constant 1.0).  It selects a shorter, still
 *                         bit-exact, division sequence; 0 is always safe.
 * updated               : optional device uint64 accumulator; the kernel adds the number of voxels it updated
 *                         (valid_pts.sum(), fusion.py:279).  NULL to skip.
 * ---------------------------------------------------------------------------------------------------------- */
#define HV_INTEGRATE_INTEGER_WEIGHTS 1
/* Bits 8..15 of `flags` pick a kernel explicitly (profiling and A/B tests; every kernel gives identical results);
 * 0 = the library's default. */
#define HV_INTEGRATE_KERNEL_V3 1         /* straight-line loop over tiles (profiles/r01_integrate_v3b.md) */
#define HV_INTEGRATE_KERNEL_PIPE 2       /* column walker: projection of tile i+1 overlapped with the loads of tile i */
#define HV_INTEGRATE_KERNEL_PIPE_CULL 3  /* + segment-level frustum culling (switched off by the kernel itself when the
                                            whole volume is inside the frustum) */
#define HV_INTEGRATE_KERNEL_PIPE_CULL_OCC4 4 /* the same at 64 registers / 4 CTAs per SM (A/B: spills vs occupancy) */
#define HV_INTEGRATE_KERNEL_DEFAULT HV_INTEGRATE_KERNEL_PIPE_CULL
#define HV_INTEGRATE_KERNEL(v) ((v) << 8)
int hv_integrate(float* tsdf, float* weight, float* color, const int64_t* h_dims, int64_t pitch,
                 const double* wx, const double* wy, const double* wz, const double* h_world2cam, const float* h_intr,
                 const void* texels, int H, int W, double trunc, double obs_weight, int flags,
                 unsigned long long* updated, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Marching cubes.  Replaces `measure.marching_cubes(tsdf_vol, level=0)` inside TSDFVolume.get_mesh /
 * get_point_cloud (fusion.py:317,338 -- scikit-image's Lewiner implementation, a third-party dependency) and the
 * in-tree post-processing fusion.py:339-348 (nearest-voxel colour gather, b/g/r unpack, voxel -> world transform).
 *
 * Two calls with a host read of the counters in between (the output size is data dependent):
 *   hv_mc_count  classifies the volume into `scratch`, writes one work record per surface cube and per edge vertex
 *                into `work`, and writes totals (DEVICE, 8 x uint64): [0] = #vertices, [1] = #triangles, [2] / [3] =
 *                min / max of the volume as order-preserving int32 keys of the floats (key = bits >= 0 ? bits :
 *                bits ^ 0x7fffffff) for skimage's "level within data range" check (ValueError, relied on by
 *                scripts/experiments.py:167), [4] = #vertices that lie on lattice edges, [5] = #surface cubes,
 *                [6] = 1 if the records did not fit (rec_verts < [4] or rec_cubes < [5]: call hv_mc_count again with a
 *                larger work buffer before hv_mc_emit; [5] <= [1] and [4] <= [0], so rec_cubes = triangle capacity and
 *                rec_verts = vertex capacity is always enough).
 *   hv_mc_emit   writes the mesh: at most n_verts vertices / n_tris triangles (the capacity of the output buffers;
 *                `totals` as left by hv_mc_count says how many there are -- a caller that reuses a capacity from an
 *                earlier extraction can launch both calls back to back and read the totals afterwards; nothing is
 *                written when [6] is set).  Vertices on
 *                lattice edges come first, sorted by (owner voxel in C scan order, edge axis), then the vertices
 *                inside cubes (Lewiner's c-vertex) by cube; triangles by cube scan order; right-hand normals of the
 *                faces point towards increasing values (what skimage returns for gradient_direction='descent').
 *
 * tsdf / color : float32 planes [X][Y][pitch] (z fastest), h_dims = {X, Y, Z}.
 * halo_lo / halo_hi (and the colour pair): optional [X][Y] planes just below z = 0 / above z = Z-1 of a z-slab
 *                shard (NULL for a whole volume).  A shard owns the cubes whose base voxel lies in its slab.
 * scratch      : device buffer of hv_mc_scratch_bytes() bytes, 256-byte aligned, passed to hv_mc_scratch_init once;
 *                must be kept between the two calls.
 * work         : device buffer of hv_mc_work_bytes(rec_verts, rec_cubes) bytes, 256-byte aligned; kept likewise, and
 *                passed with the same rec_verts / rec_cubes to both calls.
 * z_global_offset / z_global_dim : position of the slab in the global volume (0 / Z for a whole volume); vertex keys
 *                are ((x*Y + y)*z_global_dim + z_global)*4 + axis, axis 3 = the extra vertex inside a cube.
 * faces / face_keys : exactly one is non-NULL: int32 [T][3] indices into this call's vertices (whole volume only), or
 *                uint64 [T][3] global vertex keys (shards: seam vertices belong to the neighbouring shard).
 * face_cubes   : optional, with face_keys: uint64 [T] global linear index ((x*Y + y)*z_global_dim + z) of the cube each
 *                face came from -- the merge key that restores the whole-volume face order when shards are stitched.
 * verts_voxel  : optional [V][3] positions in voxel units (what skimage itself returns); NULL to skip.
 * vert_keys    : optional uint64 [V] key of every vertex (shards: what face_keys refer to); NULL to skip.
 * ---------------------------------------------------------------------------------------------------------- */
int hv_mc_scratch_bytes(const int64_t* h_dims, int has_halo_lo, int has_halo_hi, int64_t* h_bytes);
/* once per scratch buffer, before its first hv_mc_count: zeroes the counters the kernels keep (and reset themselves) */
int hv_mc_scratch_init(const int64_t* h_dims, int has_halo_lo, int has_halo_hi, void* scratch, int64_t scratch_bytes, void* stream);
int hv_mc_work_bytes(int64_t rec_verts, int64_t rec_cubes, int64_t* h_bytes);
int hv_mc_count(const float* tsdf, const float* halo_lo, const float* halo_hi, const int64_t* h_dims, int64_t pitch,
                double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts, int64_t rec_cubes,
                unsigned long long* totals, void* stream);
int hv_mc_emit(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
               const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
               double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts, int64_t rec_cubes,
               int64_t z_global_offset, int64_t z_global_dim,
               float voxel_size, const float* h_origin, int64_t n_verts, int64_t n_tris, float* verts_world,
               float* verts_voxel, float* normals, uint8_t* colors, int32_t* faces, unsigned long long* face_keys,
               unsigned long long* face_cubes, unsigned long long* vert_keys, unsigned long long* totals, void* stream);

/* ---- z-slab shards: placing a slab's mesh in the mesh of the whole volume (hive_b200/sharded.py; nothing in the reference:
 * SURVEY.md 8(e)).  The whole-volume order is column-major, so the slabs interleave column by column.
 *   hv_mc_column_counts  after hv_mc_count: counts[k][c] (DEVICE int32 [3][X*Y]) = edge vertices (k = 0), triangles (1), centre
 *                        vertices (2) this slab owns in column c = x*Y + y.
 *   hv_mc_column_shifts  all_counts (DEVICE int32, slab w's [3][X*Y] counts at w * slab_stride: the all-gathered counts, a
 *                        few words of other per-slab data may follow each) -> shifts (DEVICE int32
 *                        [4][X*Y]: what to add per column to the slab's own vertex / triangle / centre-vertex numbering, and
 *                        the global index of the first vertex of the slab above in the column) and g_totals (DEVICE
 *                        4 x uint64: [0] edge vertices, [1] triangles, [2] centre vertices of the whole volume).
 *   hv_mc_emit_placed    hv_mc_emit for a slab (halo planes allowed) with faces as indices into the WHOLE mesh; vert_dst [V]
 *                        / tri_dst [T] receive the index of every vertex / triangle of the slab in the whole mesh.
 *   hv_scatter_rows      dst[index[i]] = src[i] for rows of row_bytes = 3 or a multiple of 4 (assembling the whole mesh). */
int hv_mc_column_counts(const float* tsdf, const float* halo_lo, const float* halo_hi, const int64_t* h_dims, int64_t pitch,
                        double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts, int64_t rec_cubes,
                        unsigned long long* totals, int32_t* counts, void* stream);
int hv_mc_shift_scratch_bytes(int64_t n_columns, int64_t* h_bytes);
int hv_mc_column_shifts(const int32_t* all_counts, int64_t slab_stride, int world_size, int rank, int64_t n_columns, int32_t* shifts,
                        unsigned long long* g_totals, void* scratch, int64_t scratch_bytes, void* stream);
int hv_mc_emit_placed(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                      const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                      double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts, int64_t rec_cubes,
                      int64_t z_global_offset, int64_t z_global_dim, float voxel_size, const float* h_origin, int64_t n_verts,
                      int64_t n_tris, float* verts_world, float* normals, uint8_t* colors, int32_t* faces,
                      const int32_t* col_shifts, int64_t n_edge_global, int32_t* vert_dst, int32_t* tri_dst,
                      unsigned long long* totals, void* stream);
int hv_scatter_rows(const void* src, const int32_t* index, int64_t n, int row_bytes, void* dst, void* stream);

/* hv_mc_emit_direct: hv_mc_emit_placed with every row written AT its whole-mesh index -- verts_world / normals / colors
 * (whole_verts rows) and faces (whole_tris rows) are the whole mesh's arrays, on this GPU or on a peer GPU whose buffer
 * is mapped into this process (hv_peer_open): the slab's stores then go over NVLink straight into the final mesh and
 * nothing is gathered or scattered afterwards.  n_verts / n_tris are still this slab's own counts. */
int hv_mc_emit_direct(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                      const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                      double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts, int64_t rec_cubes,
                      int64_t z_global_offset, int64_t z_global_dim, float voxel_size, const float* h_origin, int64_t n_verts,
                      int64_t n_tris, const int32_t* col_shifts, int64_t n_edge_global, int64_t whole_verts,
                      int64_t whole_tris, float* verts_world, float* normals, uint8_t* colors, int32_t* faces,
                      unsigned long long* totals, void* stream);

/* Device buffers the other ranks of the node can store into (one process per GPU; nothing in the reference).
 *   hv_peer_alloc  cudaMalloc of `bytes` + its CUDA IPC handle (HV_PEER_HANDLE_BYTES bytes, HOST) -- send the handle to the
 *                  other ranks by any means; free with hv_peer_free once every rank has closed it.
 *   hv_peer_open   maps the buffer behind a handle made by ANOTHER process into this one (enables peer access between the
 *                  two GPUs); the pointer is usable by kernels and copies of this process.  hv_peer_close unmaps it.
 *   hv_copy_async  cudaMemcpyAsync between any two of: device, mapped peer, pinned host addresses. */
#define HV_PEER_HANDLE_BYTES 64
int hv_peer_alloc(int64_t bytes, void** h_ptr, void* h_handle);
int hv_peer_free(void* ptr);
int hv_peer_open(const void* h_handle, void** h_ptr);
int hv_peer_close(void* ptr);
int hv_copy_async(void* dst, const void* src, int64_t bytes, void* stream);

/* Signals and seam planes between the GPUs of a node (hive_b200/sharded.py, no counterpart in the reference): a flag is a 32-bit
 * epoch in the receiver's (mapped) memory.
 *   hv_peer_signal  stores `epoch` into n <= 16 flags (HOST array of DEVICE addresses, typically in peers' memory).  Stream
 *                   order is the fence: everything earlier kernels of the stream wrote -- into peer memory too -- is visible
 *                   to whoever sees the flag.
 *   hv_peer_wait    a one-warp kernel that returns when each of the n <= 32 consecutive flags at `flags` (this GPU's memory)
 *                   has reached `epoch`; after ~30 s it gives up and leaves 1 + the flag's index in *err (DEVICE uint32).
 *   hv_copy_planes  dst_a[c] = a[c * pitch + z], dst_b[c] = b[c * pitch + z] for n_cols columns: one z-plane of the tsdf and
 *                   colour volumes, e.g. into a neighbour's halo inbox.
 *   hv_mc_column_shifts_peers  hv_mc_column_shifts reading every slab's counts where they are (h_slab_ptrs: HOST array of
 *                   world_size <= 16 DEVICE addresses, slab w's [3][X*Y] counts + tail_words words) -- no all-gather;
 *                   tails (DEVICE [world_size][tail_words]) receives every slab's tail words. */
int hv_peer_signal(void* const* h_flags, int n, unsigned int epoch, void* stream);
int hv_peer_wait(const void* flags, int n, unsigned int epoch, void* err, void* stream);
int hv_copy_planes(const float* a, const float* b, int64_t pitch, int64_t z, int64_t n_cols, float* dst_a, float* dst_b,
                   void* stream);
int hv_mc_column_shifts_peers(const void* const* h_slab_ptrs, int world_size, int rank, int64_t n_columns, int tail_words,
                              int32_t* shifts, unsigned long long* g_totals, int32_t* tails, void* scratch,
                              int64_t scratch_bytes, void* stream);

/* hv_mc_count + hv_mc_emit in one call with caller-provided capacity (cap_verts vertices, cap_tris triangles in every
 * output buffer).  The single host round trip -- the counters -- lands in `h_totals` (HOST, 8 x uint64, pinned memory
 * recommended; same meaning as `totals`).  Returns HV_E_LEVEL_RANGE / HV_E_NOSURFACE like skimage would raise
 * (fusion.py:338), or HV_E_CAPACITY when the surface (or its work records) exceeds the capacity: h_totals is filled, so
 * the caller can allocate exactly and finish with hv_mc_count (if [6] is set) + hv_mc_emit. */
int hv_mc_extract(const float* tsdf, const float* halo_lo, const float* halo_hi, const float* color,
                  const float* color_halo_lo, const float* color_halo_hi, const int64_t* h_dims, int64_t pitch,
                  double level, void* scratch, int64_t scratch_bytes, void* work, int64_t rec_verts, int64_t rec_cubes,
                  int64_t z_global_offset, int64_t z_global_dim,
                  float voxel_size, const float* h_origin, int64_t cap_verts, int64_t cap_tris, float* verts_world,
                  float* verts_voxel, float* normals, uint8_t* colors, int32_t* faces, unsigned long long* face_keys,
                  unsigned long long* face_cubes, unsigned long long* vert_keys, unsigned long long* totals,
                  unsigned long long* h_totals, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Depth-to-mesh, batched over frames.  Replaces, for the per-frame foreground meshes of hive/pipeline.py:386-409,
 *   point_cloud_from_depth / image2world   hive/geometric.py:107-126, 183-206
 *   Pipeline._triangulate_faces            hive/pipeline.py:651-667 (scipy.spatial.Delaunay on the pixel lattice)
 *   Pipeline._filter_faces                 hive/pipeline.py:670-694
 * using the closed form of "lattice Delaunay + 2-pixel edge filter" (SURVEY.md App. D).
 *
 * depth : [B][H][W] float32 (0 = invalid).   mask : [B][H][W] uint8 (non-zero = keep) or NULL.
 * diag  : NULL, or [B][H][W] uint8 overriding the otherwise arbitrary diagonal of co-circular points: bit 0 = the 2x2
 *         quad whose top-left pixel this is splits along the anti-diagonal, bit 1 = the hole at this pixel splits N-S.
 * flags : HV_D2M_BRIDGES = also emit the (sqrt2, sqrt2, 2) triangles over 1-pixel holes, i.e. max_pixel_distance in
 *         [2, sqrt5) (HIVE's default 2, hive/options.py:274); without it, max_pixel_distance in [sqrt2, 2).
 * max_depth_distance : pipeline.py:689, compared in float32 like the reference.
 * Two calls (output sizes are data dependent): hv_d2m_count writes totals[0] = #vertices, totals[1] = #faces of the
 * whole batch; hv_d2m_emit writes the concatenated outputs.  Vertex ids in `faces` are frame-local ranks of the valid
 * pixels in row-major order (np.nonzero order); winding has negative signed area in (u, v), like the reference's
 * reversed Delaunay simplices.  frame_offsets: [B+1][2] int64 (vertex, face) start of each frame (device memory, or
 * page-locked host memory: 16 bytes per frame, written once).
 * h_kinv : HOST, 9 doubles, inv(K) computed in K's own dtype (geometric.py:202).  cams: DEVICE, [B][12] doubles, the
 * world-to-camera R (row-major) and t that geometric.py takes.  points2d ([V][2] int32 u,v) and vdepth ([V] float32)
 * are optional.
 * ---------------------------------------------------------------------------------------------------------- */
#define HV_D2M_BRIDGES 1
#define HV_D2M_VERTS_ONLY 2 /* point_cloud_from_depth: back-projection only, no faces */
#define HV_D2M_VERTS_F32 4  /* hv_d2m_emit*: `verts` is float32 [V][3] -- the float64 result rounded to nearest (what a glTF
                               export keeps; the reference's arrays are float64) -- for callers that pay for the bytes */
int hv_d2m_scratch_bytes(int B, int H, int W, int64_t* h_bytes);
int hv_d2m_count(const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
                 float max_depth_distance, int flags, void* scratch, int64_t scratch_bytes, unsigned long long* totals,
                 void* stream);
int hv_d2m_emit(const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
                float max_depth_distance, int flags, const double* h_kinv, const double* cams, void* scratch,
                int64_t scratch_bytes, int64_t n_verts, int64_t n_faces, double* verts, int32_t* faces, int32_t* points2d,
                float* vdepth, long long* frame_offsets, void* stream);
/* hv_d2m_emit before the host has read the totals (no round trip between count and emit): the outputs hold cap_verts /
 * cap_faces rows -- a guess, e.g. from the previous batch -- and `totals` is the DEVICE pair hv_d2m_count wrote on the same
 * stream.  Rows past the capacity are dropped; frame_offsets[B] still carries the true totals, so the caller sees an
 * overflow when it reads them and repeats with hv_d2m_emit, exactly sized (the scratch stays valid). */
int hv_d2m_emit_capped(const float* depth, const uint8_t* mask, const uint8_t* diag, int B, int H, int W,
                       float max_depth_distance, int flags, const double* h_kinv, const double* cams, void* scratch,
                       int64_t scratch_bytes, int64_t cap_verts, int64_t cap_faces, double* verts, int32_t* faces,
                       int32_t* points2d, float* vdepth, long long* frame_offsets, const unsigned long long* totals,
                       void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Elementwise projections, float64 like the reference.  `cam` is a DEVICE array of 21 doubles: a 3x3 matrix
 * (row-major), then the world-to-camera R (row-major) and t.
 *   hv_image2world  hive/geometric.py:183-206: the matrix is inv(K); out[i] = R^T (depth[i] * Kinv [u*s, v*s, 1] - t)
 *   hv_world2image  hive/geometric.py:155-180: the matrix is K; c = K (R p + t); depth = c_z; uv = c_xy / c_z / s,
 *                   rounded half-to-even when the reference would (integer dtype).
 * ---------------------------------------------------------------------------------------------------------- */
int hv_image2world(const double* points2d, const double* depth, int64_t n, const double* cam, double scale, double* out,
                   void* stream);
int hv_world2image(const double* points, int64_t n, const double* cam, double scale, int round_half_even, double* uv,
                   double* depth, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Generic face filter, Pipeline._filter_faces (hive/pipeline.py:670-694) for arbitrary faces: keeps a face iff its
 * three edges (0,1), (2,1), (0,2) are <= max_pixel_distance long in the image and differ by <= max_depth_distance
 * in depth.  Order preserving.  Two phases around a host read of totals[0] (phase 0 counts, phase 1 writes).
 * ---------------------------------------------------------------------------------------------------------- */
int hv_filter_faces_scratch_bytes(int64_t n_faces, int64_t* h_bytes);
int hv_filter_faces(const int32_t* points2d, const void* depth, int depth_is_f64, const int32_t* faces, int64_t n_faces,
                    double max_pixel_distance, double max_depth_distance, void* scratch, int64_t scratch_bytes, int phase,
                    unsigned long long* totals, int32_t* out, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Frame preparation either side of the kernels above.
 *
 * hv_dilate_mask replaces hive.image_processing.dilate_mask (hive/image_processing.py:32-48): cv2.dilate with the
 * default 3x3 rectangle (hive/options.py:248), `iterations` times, then astype(bool) -- i.e. out = 1 where any
 * mask pixel within Chebyshev distance `iterations` is non-zero.  Batched over B frames.  mask/out: (B,H,W) uint8.
 *
 * hv_pack_frame_masked / hv_upload_frame_masked replace the per-frame masking of the TSDF driver,
 * hive/fusion.py:118-121 (`mask = dilate_mask(mask); depth_im[mask > 0] = 0.0`), fused into the texel pack:
 * texel.depth = 0 where the dilated mask is set.
 *
 * hv_depth_max: np.max(depth_im) of get_view_frustum (fusion.py:364) for B frames of n pixels; keys_scratch = B ints.
 * ---------------------------------------------------------------------------------------------------------- */
int hv_dilate_scratch_bytes(int B, int H, int W, int64_t* h_bytes);
int hv_dilate_mask(const uint8_t* mask, int B, int H, int W, int iterations, void* scratch, int64_t scratch_bytes,
                   uint8_t* out, void* stream);
/* hv_dilate_mask_static: hv_dilate_mask for callers that cannot allocate device memory (CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.  The
 * two bit-planes live in a 2 MiB array inside the library (hv_dilate_static_max_bytes; held by every GPU that loads the
 * library until exit), so the call allocates nothing and takes no scratch argument.  2 * B * H * ceil(W/32) * 4 bytes must
 * fit, e.g. four 1920x1080 masks; larger batches return HV_E_CAPACITY (use hv_dilate_mask).  All calls on a device share
 * the array: calls that overlap on the GPU (different streams, no ordering between them) race; calls on one stream, or
 * ordered by events, are safe.  Nothing in it has to start at zero -- each call overwrites what it reads. */
int hv_dilate_static_max_bytes(int64_t* h_bytes);
int hv_dilate_mask_static(const uint8_t* mask, int B, int H, int W, int iterations, uint8_t* out, void* stream);
int hv_pack_frame_masked(const float* depth, const uint8_t* rgb, const uint8_t* mask, int H, int W, int iterations,
                         void* scratch, int64_t scratch_bytes, void* texels, void* stream);
int hv_upload_frame_masked(const float* h_depth, const uint8_t* h_rgb, const uint8_t* h_mask, int H, int W,
                           int iterations, float* d_depth, uint8_t* d_rgb, uint8_t* d_mask, void* scratch,
                           int64_t scratch_bytes, void* texels, void* stream);
int hv_depth_max(const float* depth, int B, int64_t n, int* keys_scratch, float* out, void* stream);

/* Key-frame selection, the inner loop of HiveDataset.select_key_frames (hive/io.py:1149-1175): for one candidate frame
 * (depth (H,W) f32, mask (H,W) u8 with 0 = static pixel, cam = Kinv 9 | R 9 | t 3 float64, world-to-camera) and n_keys
 * key frames (key_cams = n_keys x (K 9 | R 9 | t 3) float64): box[k] = {min_u, min_v, max_u, max_v} of the candidate's
 * static points projected (rounded half-even to int, as world2image does) inside key frame k's image, count[k] = how
 * many landed inside.  All device pointers. */
int hv_keyframe_overlap(const float* depth, const uint8_t* mask, int H, int W, const double* cam, const double* key_cams,
                        int n_keys, int* box, unsigned int* count, void* stream);

/* Vertex colours of the TSDF mesh to the renderer's colour space, hive/pipeline.py:281-282
 * (`vertex_colors[:, :3] = (255 * np.power(vertex_colors / 255, 2.2)).astype(int)`): a function of one byte, applied
 * in place through a 256-entry table the host builds with that expression.  bytes: n device bytes, 16-byte aligned;
 * lut: 256 device bytes. */
int hv_byte_lut(uint8_t* bytes, int64_t n, const uint8_t* lut, void* stream);

/* np.count_nonzero per mask of a batch (the 1 % coverage rule, hive/pipeline.py:377): counts = DEVICE uint64 [B]. */
int hv_count_nonzero(const uint8_t* masks, int B, int64_t n, unsigned long long* counts, void* stream);
/* box (DEVICE int64 [4]) = min u, min v, max u, max v over uv rows holding integers (hive/pipeline.py:797-800). */
int hv_uv_bbox(const double* uv, int64_t n, int64_t* box, void* stream);
/* Texture atlas of a frame's objects, the pixel and UV work of Pipeline._pack_textures (hive/pipeline.py:810-868); the
 * placement (a few integers per crop) is the host's.  src: uint8 pixels the crops are read from -- the crops back to back,
 * or the frame itself; desc[t] = {byte offset of the crop's first pixel in src, x0, y0, w, h, first row of the crop's UVs
 * in `uv`, row pitch of the crop in src in pixels (w, or the frame's width), u shift, v shift} (int64 x 9); rows
 * r = 0..n_rows-1 hold crops row_first[r] .. row_first[r+1]-1 (in x order) and start at canvas row row_y0[r] (both
 * n_rows+1 entries).  atlas: (canvas_h, canvas_w, 3) uint8, every byte written (black where no crop lies).  uv: (n_uv, 2)
 * float64 pixel coordinates; uv_out[i] = (((u - u shift) + x0) / canvas_w, 1 - ((v - v shift) + y0) / canvas_h)
 * (pipeline.py:806, 837, 857, 865-866).  All pointers device. */
int hv_atlas_pack(const uint8_t* src, const int64_t* desc, int n_tex, const int64_t* row_first, const int64_t* row_y0,
                  int n_rows, int canvas_h, int canvas_w, uint8_t* atlas, const double* uv, int64_t n_uv, double* uv_out,
                  void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Connected components over the face-adjacency graph: the body of Pipeline._cleanup_with_connected_components
 * (hive/pipeline.py:741-779 -> trimesh.graph.connected_components(mesh.face_adjacency, min_len)).
 *   hv_cc_mask: keep[f] = 1 iff face f (int32 [F][3], DEVICE) survives: faces adjacent across an edge that exactly two
 *   faces use form the graph (edges found through a hash table in `scratch`: hv_cc_scratch_bytes, 256-byte aligned;
 *   nothing is sorted); components below min_len faces are dropped; largest_only keeps the largest survivor, lowest
 *   face index on a tie; labels (optional, F uint32) = component root per face, 0xffffffff for faces without any
 *   neighbour.
 * ---------------------------------------------------------------------------------------------------------- */
int hv_cc_scratch_bytes(int64_t n_faces, int64_t* h_bytes);
int hv_cc_mask(const int32_t* faces, int64_t n_faces, int64_t n_verts, int min_len, int largest_only, void* scratch,
               int64_t scratch_bytes, uint8_t* keep, uint32_t* labels, void* stream);
/* The vertices `faces_all` refer to, in their original order, into verts_out (float64 [n_verts][3] capacity), `faces_keep`
 * renumbered accordingly into faces_out; d_count[0..1] (DEVICE, 2 x uint64): [0] = vertices kept.  What building
 * trimesh.Trimesh(vertices, faces) does to unreferenced vertices before the components are taken (pipeline.py:758). */
/* hv_cc_mask + the compaction above in one call on device arrays (no host hop between the steps): verts_out / faces_out have
 * the capacity of the inputs; d_counts (DEVICE, 4 x uint64): [0] = vertices kept, [2] = faces kept. */
int hv_cc_cleanup_scratch_bytes(int64_t n_faces, int64_t n_verts, int64_t* h_bytes);
int hv_cc_cleanup(const int32_t* faces, int64_t n_faces, const double* verts_in, int64_t n_verts, int min_len, int largest_only,
                  double* verts_out, int32_t* faces_out, unsigned long long* d_counts, void* scratch, int64_t scratch_bytes,
                  void* stream);
int hv_cc_compact_scratch_bytes(int64_t n_verts, int64_t* h_bytes);
int hv_cc_compact_vertices(const int32_t* faces_all, int64_t n_faces_all, const int32_t* faces_keep, int64_t n_faces_keep,
                           const double* verts_in, int64_t n_verts, double* verts_out, int32_t* faces_out,
                           unsigned long long* d_count, void* scratch, int64_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * PLY writers (host pointers, host threads; no CUDA).  Replace meshwrite / pcwrite, fusion.py:378-438: ASCII output
 * byte-identical to the reference's; binary != 0 writes binary_little_endian with the same elements instead.
 * verts/norms/xyz: (n,3) float64 (the reference formats Python floats, i.e. doubles); faces (n,3) int32; colours uint8.
 * ---------------------------------------------------------------------------------------------------------- */
int hv_meshwrite_ply(const char* filename, const double* verts, const int32_t* faces, const double* norms,
                     const uint8_t* colors, int64_t n_verts, int64_t n_faces, int binary);
int hv_pcwrite_ply(const char* filename, const double* xyz, const uint8_t* rgb, int64_t n, int binary);

#ifdef __cplusplus
}
#endif
#endif /* HIVE_B200_H */
