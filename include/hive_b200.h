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
int hv_integrate(float* tsdf, float* weight, float* color, const int64_t* h_dims, int64_t pitch,
                 const double* wx, const double* wy, const double* wz, const double* h_world2cam, const float* h_intr,
                 const void* texels, int H, int W, double trunc, double obs_weight, int flags,
                 unsigned long long* updated, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HIVE_B200_H */
