// Per-frame image preparation that sits either side of the TSDF / depth-to-mesh kernels:
//
//   * dilate_mask: hive/image_processing.py:32-48 -- cv2.dilate(mask, 3x3 rect, iterations=n) followed by astype(bool).
//     n iterations of a 3x3 rectangle are one dilation with a (2n+1)^2 square (OpenCV itself folds the iterations into
//     the kernel size for rectangular elements); the border is ignored (morphologyDefaultBorderValue).  So
//         out[v][u] = any(mask[v-n..v+n][u-n..u+n] != 0),   n = 0 -> mask != 0.
//     Done on bit-planes: one ballot per 32 pixels, a horizontal pass of 2n+1 funnel shifts per word, a vertical OR of
//     2n+1 words.  A 640x480 mask is 9600 words: everything stays in L1/L2.
//   * masked frame upload: hive/fusion.py:118-121 (`depth_im[dilate_mask(mask) > 0] = 0`) fused into the texel pack of
//     hv_pack_frame, so the dilated mask never exists as an image and the depth frame is never rewritten.
//   * depth_max: the np.max(depth_im) of get_view_frustum (fusion.py:364) for a batch of frames.
#include "hv_common.cuh"
#include <limits.h>

namespace hv {

// bits[(b*H + v)*Ww + w] bit l  <=>  mask[b][v][32*w + l] != 0 ; bits past W are zero.  One warp per word.
__global__ void __launch_bounds__(256) mask_bits_kernel(const uint8_t* __restrict__ mask, long long rows, int W, int Ww,
                                                        uint32_t* __restrict__ bits) {
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= rows * Ww) return;
    const long long row = warp / Ww;
    const int w = (int)(warp - row * Ww);
    const int u = w * 32 + lane;
    const bool on = u < W && mask[row * W + u] != 0;
    const uint32_t word = __ballot_sync(0xffffffffu, on);
    if (lane == 0) bits[warp] = word;
}

// 32 bits of a row starting at bit offset `off` (may be negative or run past the row: zeros there).
__device__ __forceinline__ uint32_t row_window(const uint32_t* __restrict__ row, int Ww, int off) {
    const int i = off >> 5;  // arithmetic shift: floor
    const int sh = off & 31;
    const uint32_t lo = (i >= 0 && i < Ww) ? row[i] : 0u;
    const uint32_t hi = (i + 1 >= 0 && i + 1 < Ww) ? row[i + 1] : 0u;
    return __funnelshift_r(lo, hi, sh);
}

// Horizontal dilation by n on each side: hb word = OR over s in [-n, n] of the row shifted by s.
__global__ void __launch_bounds__(256) mask_hdilate_kernel(const uint32_t* __restrict__ bits, long long rows, int W,
                                                           int Ww, int n, uint32_t* __restrict__ hbits) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * Ww) return;
    const long long row = idx / Ww;
    const int w = (int)(idx - row * Ww);
    const uint32_t* r = bits + row * Ww;
    uint32_t acc = 0;
    for (int s = -n; s <= n; ++s) acc |= row_window(r, Ww, w * 32 + s);
    const int tail = W - w * 32;  // keep bits past W clear
    if (tail < 32) acc &= (tail <= 0) ? 0u : ((1u << tail) - 1u);
    hbits[idx] = acc;
}

__device__ __forceinline__ uint32_t vdilate_word(const uint32_t* __restrict__ hbits, int H, int Ww, int v, int w, int n) {
    const int v0 = max(0, v - n), v1 = min(H - 1, v + n);
    uint32_t acc = 0;
    for (int r = v0; r <= v1; ++r) acc |= hbits[(long long)r * Ww + w];
    return acc;
}

// Vertical pass + expansion to a uint8 0/1 image.
__global__ void __launch_bounds__(256) mask_vdilate_kernel(const uint32_t* __restrict__ hbits, int B, int H, int W, int Ww,
                                                           int n, uint8_t* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)B * H * W;
    if (i >= total) return;
    const long long row = i / W;
    const int u = (int)(i - row * W);
    const int b = (int)(row / H);
    const int v = (int)(row - (long long)b * H);
    const uint32_t word = vdilate_word(hbits + (long long)b * H * Ww, H, Ww, v, u >> 5, n);
    out[i] = (word >> (u & 31)) & 1u;
}

// hv_pack_frame with `depth = 0` wherever the dilated mask is set (hive/fusion.py:121).
__global__ void __launch_bounds__(256) pack_frame_masked_kernel(const float* __restrict__ depth,
                                                                const uint8_t* __restrict__ rgb,
                                                                const uint32_t* __restrict__ hbits, int H, int W, int Ww,
                                                                int n, uint2* __restrict__ texels) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= H * W) return;
    const int v = i / W, u = i - v * W;
    const uint32_t word = vdilate_word(hbits, H, Ww, v, u >> 5, n);
    const bool masked = (word >> (u & 31)) & 1u;
    const unsigned int r = rgb[3 * i + 0], g = rgb[3 * i + 1], b = rgb[3 * i + 2];
    texels[i] = make_uint2(masked ? 0u : __float_as_uint(depth[i]), (b << 16) | (g << 8) | r);
}

// Per-frame maximum of a (B, n) float32 array, NaN-free input assumed (depth maps); result as ordered-int atomicMax.
__global__ void __launch_bounds__(256) depth_max_kernel(const float* __restrict__ depth, long long n,
                                                        int* __restrict__ out_keys) {
    const int b = blockIdx.y;
    const float* d = depth + (long long)b * n;
    float m = -INFINITY;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        m = fmaxf(m, d[i]);
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) {
        const int bitsv = __float_as_int(m);
        atomicMax(out_keys + b, bitsv >= 0 ? bitsv : bitsv ^ 0x7fffffff);  // order-preserving float -> int key
    }
}

__global__ void depth_max_finish_kernel(const int* __restrict__ keys, int B, float* __restrict__ out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int k = keys[b];
    out[b] = __int_as_float(k >= 0 ? k : k ^ 0x7fffffff);
}

static inline int words_per_row(int W) { return (W + 31) / 32; }

static int launch_bits_and_hdilate(const uint8_t* mask, long long rows, int W, int n, uint32_t* raw, uint32_t* hb,
                                   cudaStream_t st) {
    const int Ww = words_per_row(W);
    const long long words = rows * Ww;
    mask_bits_kernel<<<(unsigned)((words * 32 + 255) / 256), 256, 0, st>>>(mask, rows, W, Ww, raw);
    HV_CUDA(cudaGetLastError());
    mask_hdilate_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(raw, rows, W, Ww, n, hb);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

}  // namespace hv

extern "C" int hv_dilate_scratch_bytes(int B, int H, int W, int64_t* bytes) {
    HV_REQUIRE(bytes && B > 0 && H > 0 && W > 0, "hv_dilate_scratch_bytes: bad arguments");
    *bytes = 2ll * B * H * hv::words_per_row(W) * (int64_t)sizeof(uint32_t);
    return HV_OK;
}

extern "C" int hv_dilate_mask(const uint8_t* mask, int B, int H, int W, int iterations, void* scratch,
                              int64_t scratch_bytes, uint8_t* out, void* stream) {
    HV_REQUIRE(mask && out && scratch, "hv_dilate_mask: null pointer");
    HV_REQUIRE(B > 0 && H > 0 && W > 0 && iterations >= 0, "hv_dilate_mask: bad arguments");
    HV_REQUIRE((long long)B * H * W < (1ll << 40), "hv_dilate_mask: batch too large");
    int64_t need = 0;
    hv_dilate_scratch_bytes(B, H, W, &need);
    HV_REQUIRE(scratch_bytes >= need && ((uintptr_t)scratch & 3) == 0, "hv_dilate_mask: scratch too small or unaligned");
    const int Ww = hv::words_per_row(W);
    uint32_t* raw = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* hb = raw + (long long)B * H * Ww;
    cudaStream_t st = hv::as_stream(stream);
    const int rc = hv::launch_bits_and_hdilate(mask, (long long)B * H, W, iterations, raw, hb, st);
    if (rc != HV_OK) return rc;
    const long long total = (long long)B * H * W;
    hv::mask_vdilate_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(hb, B, H, W, Ww, iterations, out);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

// ---- hv_dilate_mask without caller scratch ---------------------------------------------------------------------------------
// The two bit-planes of hv_dilate_mask live in a fixed array in device memory, for callers that cannot allocate (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the 2 MiB until exit.
namespace hv {

constexpr long long kDilateStaticWords = 512 * 1024;  // 2 MiB: both planes of e.g. four 1920x1080 masks
static __device__ __align__(256) uint32_t g_dilate_static[kDilateStaticWords];

constexpr int kMaxDevices = 64;
static uint32_t* g_dilate_static_ptr[kMaxDevices];  // device address of the array, per device (0: not asked yet)

// In host code `g_dilate_static` is only a handle for the runtime, not a device address: the address the kernels get comes
// from cudaGetSymbolAddress, asked once per device with that device current.  (Two threads racing here store the same value.)
static int dilate_static_address(uint32_t** out) {
    int dev = -1;
    HV_CUDA(cudaGetDevice(&dev));
    HV_REQUIRE(dev >= 0 && dev < kMaxDevices, "hv_dilate_mask_static: device ordinal %d out of range", dev);
    uint32_t* p = g_dilate_static_ptr[dev];
    if (!p) {
        HV_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_dilate_static));
        HV_REQUIRE(p && ((uintptr_t)p & 255) == 0, "hv_dilate_mask_static: bad symbol address");
        g_dilate_static_ptr[dev] = p;
    }
    *out = p;
    return HV_OK;
}

}  // namespace hv

extern "C" int hv_dilate_static_max_bytes(int64_t* bytes) {
    HV_REQUIRE(bytes, "hv_dilate_static_max_bytes: null pointer");
    *bytes = hv::kDilateStaticWords * (int64_t)sizeof(uint32_t);
    return HV_OK;
}

extern "C" int hv_dilate_mask_static(const uint8_t* mask, int B, int H, int W, int iterations, uint8_t* out, void* stream) {
    HV_REQUIRE(mask && out, "hv_dilate_mask_static: null pointer");
    HV_REQUIRE(B > 0 && H > 0 && W > 0 && iterations >= 0, "hv_dilate_mask_static: bad arguments");
    const int Ww = hv::words_per_row(W);
    const long long plane = (long long)B * H * Ww;
    if (2 * plane > hv::kDilateStaticWords) {
        hv::set_error("hv_dilate_mask_static: %d x %d x %d needs %lld bytes of scratch, the built-in array holds %lld; "
                      "use hv_dilate_mask with caller scratch", B, H, W, 2 * plane * 4, hv::kDilateStaticWords * 4);
        return HV_E_CAPACITY;
    }
    uint32_t* raw = nullptr;
    const int ra = hv::dilate_static_address(&raw);
    if (ra != HV_OK) return ra;
    uint32_t* hb = raw + plane;
    // nothing to clear: both planes are written in full (all B*H*Ww words) before the next kernel reads them
    cudaStream_t st = hv::as_stream(stream);
    const int rc = hv::launch_bits_and_hdilate(mask, (long long)B * H, W, iterations, raw, hb, st);
    if (rc != HV_OK) return rc;
    const long long total = (long long)B * H * W;
    hv::mask_vdilate_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(hb, B, H, W, Ww, iterations, out);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_pack_frame_masked(const float* depth, const uint8_t* rgb, const uint8_t* mask, int H, int W,
                                    int iterations, void* scratch, int64_t scratch_bytes, void* texels, void* stream) {
    HV_REQUIRE(depth && rgb && mask && texels && scratch, "hv_pack_frame_masked: null pointer");
    HV_REQUIRE(H > 0 && W > 0 && (long long)H * W < (1ll << 30) && iterations >= 0, "hv_pack_frame_masked: bad arguments");
    HV_REQUIRE(((uintptr_t)texels & 7) == 0, "hv_pack_frame_masked: texels must be 8-byte aligned");
    int64_t need = 0;
    hv_dilate_scratch_bytes(1, H, W, &need);
    HV_REQUIRE(scratch_bytes >= need && ((uintptr_t)scratch & 3) == 0,
               "hv_pack_frame_masked: scratch too small or unaligned");
    const int Ww = hv::words_per_row(W);
    uint32_t* raw = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* hb = raw + (long long)H * Ww;
    cudaStream_t st = hv::as_stream(stream);
    const int rc = hv::launch_bits_and_hdilate(mask, H, W, iterations, raw, hb, st);
    if (rc != HV_OK) return rc;
    const int n = H * W;
    hv::pack_frame_masked_kernel<<<(n + 255) / 256, 256, 0, st>>>(depth, rgb, hb, H, W, Ww, iterations,
                                                                 reinterpret_cast<uint2*>(texels));
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_upload_frame_masked(const float* h_depth, const uint8_t* h_rgb, const uint8_t* h_mask, int H, int W,
                                      int iterations, float* d_depth, uint8_t* d_rgb, uint8_t* d_mask, void* scratch,
                                      int64_t scratch_bytes, void* texels, void* stream) {
    HV_REQUIRE(h_depth && h_rgb && h_mask && d_depth && d_rgb && d_mask, "hv_upload_frame_masked: null pointer");
    HV_REQUIRE(H > 0 && W > 0, "hv_upload_frame_masked: bad frame size %d x %d", H, W);
    const size_t n = (size_t)H * W;
    cudaStream_t st = hv::as_stream(stream);
    HV_CUDA(cudaMemcpyAsync(d_depth, h_depth, n * sizeof(float), cudaMemcpyHostToDevice, st));
    HV_CUDA(cudaMemcpyAsync(d_rgb, h_rgb, n * 3, cudaMemcpyHostToDevice, st));
    HV_CUDA(cudaMemcpyAsync(d_mask, h_mask, n, cudaMemcpyHostToDevice, st));
    return hv_pack_frame_masked(d_depth, d_rgb, d_mask, H, W, iterations, scratch, scratch_bytes, texels, stream);
}

extern "C" int hv_depth_max(const float* depth, int B, int64_t n, int* keys_scratch, float* out, void* stream) {
    HV_REQUIRE(depth && keys_scratch && out && B > 0 && n > 0, "hv_depth_max: bad arguments");
    HV_REQUIRE(B <= 65535, "hv_depth_max: at most 65535 frames per call");
    cudaStream_t st = hv::as_stream(stream);
    HV_CUDA(cudaMemsetAsync(keys_scratch, 0x80, (size_t)B * sizeof(int), st));  // 0x80808080: below every finite key
    const int bx = (int)((n + 256 * 8 - 1) / (256 * 8));
    hv::depth_max_kernel<<<dim3(bx < 1 ? 1 : (bx > 64 ? 64 : bx), B), 256, 0, st>>>(depth, n, keys_scratch);
    HV_CUDA(cudaGetLastError());
    hv::depth_max_finish_kernel<<<(B + 127) / 128, 128, 0, st>>>(keys_scratch, B, out);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

// ---- key-frame selection (hive/io.py:1117-1189) ---------------------------------------------------------------------------
// For one candidate frame and K current key frames: back-project every static pixel (mask == 0, depth > 0;
// point_cloud_from_depth, hive/geometric.py:107-126), project it into each key frame (world2image with the default int32
// rounding, geometric.py:155-180), and reduce the bounding box of the projections that land inside the image
// (io.py:1163-1175).  The reference materialises the point cloud and K projected copies of it per candidate; here nothing
// is materialised: one thread per pixel, the key frames are an inner loop, boxes are reduced per warp and merged with
// atomics.  box[k] = {min_u, min_v, max_u, max_v}, count[k] = number of visible points.
namespace hv {

__global__ void __launch_bounds__(256) keyframe_overlap_kernel(const float* __restrict__ depth, const uint8_t* __restrict__ mask,
                                                               int H, int W, const double* __restrict__ cam /* Kinv 9, R 9, t 3 */,
                                                               const double* __restrict__ key_cams /* K x (K 9, R 9, t 3) */,
                                                               int n_keys, int* __restrict__ box, unsigned int* __restrict__ count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool in_image = i < H * W;
    const float d32 = in_image ? depth[i] : 0.0f;
    const bool valid = in_image && mask[i] == 0 && d32 > 0.0f;
    if (__ballot_sync(0xffffffffu, valid) == 0u) return;
    double P[3] = {0.0, 0.0, 0.0};
    if (valid) {
        const int v = i / W, u = i - v * W;
        const double *Kinv = cam, *R = cam + 9, *t = cam + 18;
        const double d = (double)d32, ud = (double)u, vd = (double)v;
        double X[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {  // same operation order as image2world_kernel (hv_depth_to_mesh.cu)
            const double pc = fma(Kinv[3 * k + 2], 1.0, fma(Kinv[3 * k + 1], vd, __dmul_rn(Kinv[3 * k], ud)));
            X[k] = __dsub_rn(__dmul_rn(d, pc), t[k]);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) P[k] = fma(R[6 + k], X[2], fma(R[3 + k], X[1], __dmul_rn(R[k], X[0])));
    }
    for (int kf = 0; kf < n_keys; ++kf) {
        const double *K = key_cams + 21 * kf, *R = K + 9, *t = K + 18;
        int pu = 0, pv = 0;
        bool vis = false;
        if (valid) {
            double c[3], q[3];
#pragma unroll
            for (int k = 0; k < 3; ++k)  // same operation order as world2image_kernel
                c[k] = __dadd_rn(fma(R[3 * k + 2], P[2], fma(R[3 * k + 1], P[1], __dmul_rn(R[3 * k], P[0]))), t[k]);
#pragma unroll
            for (int k = 0; k < 3; ++k) q[k] = fma(K[3 * k + 2], c[2], fma(K[3 * k + 1], c[1], __dmul_rn(K[3 * k], c[0])));
            const double a = rint(q[0] / q[2]), b = rint(q[1] / q[2]);
            // astype(int32) of the rounded value, then 0 <= u < width, 0 <= v < height (io.py:1163-1167)
            vis = a >= 0.0 && a < (double)W && b >= 0.0 && b < (double)H;
            pu = vis ? (int)a : 0;
            pv = vis ? (int)b : 0;
        }
        const unsigned int m = __ballot_sync(0xffffffffu, vis);
        if (m == 0u) continue;
        int lo_u = vis ? pu : INT_MAX, lo_v = vis ? pv : INT_MAX, hi_u = vis ? pu : INT_MIN, hi_v = vis ? pv : INT_MIN;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo_u = min(lo_u, __shfl_xor_sync(0xffffffffu, lo_u, o));
            lo_v = min(lo_v, __shfl_xor_sync(0xffffffffu, lo_v, o));
            hi_u = max(hi_u, __shfl_xor_sync(0xffffffffu, hi_u, o));
            hi_v = max(hi_v, __shfl_xor_sync(0xffffffffu, hi_v, o));
        }
        if (lane == 0) {
            atomicMin(box + 4 * kf + 0, lo_u);
            atomicMin(box + 4 * kf + 1, lo_v);
            atomicMax(box + 4 * kf + 2, hi_u);
            atomicMax(box + 4 * kf + 3, hi_v);
            atomicAdd(count + kf, (unsigned int)__popc(m));
        }
    }
}

__global__ void keyframe_box_init_kernel(int* box, unsigned int* count, int n_keys) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_keys) return;
    box[4 * k + 0] = INT_MAX;
    box[4 * k + 1] = INT_MAX;
    box[4 * k + 2] = INT_MIN;
    box[4 * k + 3] = INT_MIN;
    count[k] = 0u;
}

// ---- vertex colours through a 256-entry table (hive/pipeline.py:281-282) --------------------------------------------------
// The reference converts the TSDF mesh's vertex colours for the renderer with
//     vertex_colors[:, :3] = (255 * np.power(vertex_colors / 255, 2.2)).astype(int)
// a pure function of one byte.  The table is built on the host by that very numpy expression (so it is the reference's
// arithmetic by construction) and applied in place here: 16 bytes per thread, table in shared memory.
__global__ void __launch_bounds__(256) byte_lut_kernel(uint8_t* __restrict__ bytes, long long n,
                                                       const uint8_t* __restrict__ lut) {
    __shared__ uint8_t s_lut[256];
    s_lut[threadIdx.x] = lut[threadIdx.x];
    __syncthreads();
    const long long n16 = n >> 4;
    uint4* v = reinterpret_cast<uint4*>(bytes);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (long long)gridDim.x * blockDim.x) {
        uint4 w = v[i];
        unsigned int* c = &w.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned int x = c[k];
            c[k] = (unsigned int)s_lut[x & 255u] | ((unsigned int)s_lut[(x >> 8) & 255u] << 8) |
                   ((unsigned int)s_lut[(x >> 16) & 255u] << 16) | ((unsigned int)s_lut[x >> 24] << 24);
        }
        v[i] = w;
    }
    // ragged tail (< 16 bytes), by the first threads of block 0
    if (blockIdx.x == 0) {
        const long long i = (n16 << 4) + threadIdx.x;
        if (threadIdx.x < 16 && i < n) bytes[i] = s_lut[bytes[i]];
    }
}

}  // namespace hv

extern "C" int hv_keyframe_overlap(const float* depth, const uint8_t* mask, int H, int W, const double* cam,
                                   const double* key_cams, int n_keys, int* box, unsigned int* count, void* stream) {
    HV_REQUIRE(depth && mask && cam && key_cams && box && count, "hv_keyframe_overlap: null pointer");
    HV_REQUIRE(H > 0 && W > 0 && (long long)H * W < (1ll << 30) && n_keys > 0, "hv_keyframe_overlap: bad arguments");
    cudaStream_t st = hv::as_stream(stream);
    hv::keyframe_box_init_kernel<<<(n_keys + 127) / 128, 128, 0, st>>>(box, count, n_keys);
    hv::keyframe_overlap_kernel<<<(H * W + 255) / 256, 256, 0, st>>>(depth, mask, H, W, cam, key_cams, n_keys, box, count);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_byte_lut(uint8_t* bytes, int64_t n, const uint8_t* lut, void* stream) {
    HV_REQUIRE(bytes && lut && n >= 0, "hv_byte_lut: bad arguments");
    HV_REQUIRE(((uintptr_t)bytes & 15) == 0, "hv_byte_lut: bytes must be 16-byte aligned");
    if (n == 0) return HV_OK;
    const long long blocks = (n / 16 + 255) / 256;
    const int cap = hv::device_info().sm_count * 8;
    hv::byte_lut_kernel<<<(unsigned)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks)), 256, 0, hv::as_stream(stream)>>>(
        bytes, n, lut);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

// ---- coverage of (dilated) object masks: np.count_nonzero(mask) of hive/pipeline.py:377, per frame of a batch ---------------
namespace hv {
__global__ void __launch_bounds__(256) count_nonzero_kernel(const uint8_t* __restrict__ m, long long n, unsigned long long* out) {
    const uint8_t* row = m + (size_t)blockIdx.y * n;
    unsigned int c = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) c += row[i] != 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out + blockIdx.y, (unsigned long long)c);
}
}  // namespace hv

extern "C" int hv_count_nonzero(const uint8_t* masks, int B, int64_t n, unsigned long long* counts, void* stream) {
    HV_REQUIRE(masks && counts && B > 0 && n > 0, "hv_count_nonzero: bad arguments");
    cudaStream_t st = hv::as_stream(stream);
    HV_CUDA(cudaMemsetAsync(counts, 0, (size_t)B * sizeof(unsigned long long), st));
    const long long want = (n + 255) / 256;
    const dim3 grid((unsigned int)(want < 64 ? want : 64), (unsigned int)B);
    hv::count_nonzero_kernel<<<grid, 256, 0, st>>>(masks, n, counts);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}
