// Texture atlas packing: the pixel and UV work of Pipeline._pack_textures (hive/pipeline.py:810-868).
//
// The reference lays the per-object texture crops of a frame out in `n_rows` rows, left to right, rows top to bottom, on
// a black canvas (row height = tallest crop of the row, canvas width = widest row), shifts every object's UVs by its
// crop's position and normalises them: u / canvas_w, 1 - v / canvas_h.  The layout is a few integers per crop and is
// decided on the host (hive_b200/pipeline.py); the bytes move here.
//
// Atlas kernel: one thread per canvas pixel.  It finds its row (<= a handful) and its crop (binary search over the x
// offsets of the row), then copies the crop's pixel or writes black -- every canvas byte is written exactly once, no
// memset + overdraw.  UV kernel: one thread per vertex, binary search over the per-crop vertex ranges, float64 with the
// reference's operation order (add offset, divide, subtract from one).
#include "hv_common.cuh"

#include <limits.h>

namespace hv {

// desc[t] = {src byte offset, x0, y0, w, h, first uv row, src row pitch in pixels (crop taken straight out of a larger image),
//            u shift, v shift (subtracted from the uv before the crop's place is added: the crop's origin in that image)}
constexpr int kAtlasDescWords = 9;

__global__ void __launch_bounds__(256) atlas_pixels_kernel(const uint8_t* __restrict__ src,
                                                           const long long* __restrict__ desc,
                                                           const long long* __restrict__ row_first,
                                                           const long long* __restrict__ row_y0, int n_rows, int canvas_h,
                                                           int canvas_w, uint8_t* __restrict__ atlas) {
    const long long n = (long long)canvas_h * canvas_w;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (long long)gridDim.x * blockDim.x) {
        const int y = (int)(p / canvas_w), x = (int)(p - (long long)y * canvas_w);
        int r = 0;
        while (r + 1 < n_rows && row_y0[r + 1] <= y) ++r;
        // last crop of row r that starts at or left of x
        long long lo = row_first[r], hi = row_first[r + 1];
        while (hi - lo > 1) {
            const long long mid = (lo + hi) >> 1;
            if (desc[mid * kAtlasDescWords + 1] <= x) lo = mid;
            else hi = mid;
        }
        const long long* d = desc + lo * kAtlasDescWords;
        const int dx = x - (int)d[1], dy = y - (int)d[2];
        uint8_t c0 = 0, c1 = 0, c2 = 0;
        if (dx >= 0 && dx < (int)d[3] && dy >= 0 && dy < (int)d[4]) {
            const uint8_t* s = src + d[0] + ((long long)dy * d[6] + dx) * 3;
            c0 = s[0];
            c1 = s[1];
            c2 = s[2];
        }
        uint8_t* o = atlas + p * 3;
        o[0] = c0;
        o[1] = c1;
        o[2] = c2;
    }
}

__global__ void __launch_bounds__(256) atlas_uv_kernel(const double* __restrict__ uv, long long n_uv,
                                                       const long long* __restrict__ desc, int n_tex, int canvas_h,
                                                       int canvas_w, double* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_uv) return;
    int lo = 0, hi = n_tex;  // last crop whose first uv row is <= i (crops without vertices share a start: take the last)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (desc[(long long)mid * kAtlasDescWords + 5] <= i) lo = mid;
        else hi = mid;
    }
    const long long* d = desc + (long long)lo * kAtlasDescWords;
    // uv -= crop origin (pipeline.py:806, integers: exact), then uvs_atlas[k] += (moving_w, moving_h) (pipeline.py:837, 857)
    const double u = __dadd_rn(__dsub_rn(uv[2 * i], (double)d[7]), (double)d[1]);
    const double v = __dadd_rn(__dsub_rn(uv[2 * i + 1], (double)d[8]), (double)d[2]);
    out[2 * i] = __ddiv_rn(u, (double)canvas_w);                              // pipeline.py:865
    out[2 * i + 1] = __dsub_rn(1.0, __ddiv_rn(v, (double)canvas_h));          // pipeline.py:866
}

// box = {min u, min v, max u, max v} of uv rows holding integers (world2image with rounding), pipeline.py:797-800
__global__ void __launch_bounds__(256) uv_bbox_kernel(const double* __restrict__ uv, long long n, long long* box) {
    long long lo_u = LLONG_MAX, lo_v = LLONG_MAX, hi_u = LLONG_MIN, hi_v = LLONG_MIN;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long u = (long long)rint(uv[2 * i]), v = (long long)rint(uv[2 * i + 1]);
        lo_u = min(lo_u, u);
        hi_u = max(hi_u, u);
        lo_v = min(lo_v, v);
        hi_v = max(hi_v, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo_u = min(lo_u, __shfl_xor_sync(0xffffffffu, lo_u, o));
        lo_v = min(lo_v, __shfl_xor_sync(0xffffffffu, lo_v, o));
        hi_u = max(hi_u, __shfl_xor_sync(0xffffffffu, hi_u, o));
        hi_v = max(hi_v, __shfl_xor_sync(0xffffffffu, hi_v, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(box, lo_u);
        atomicMin(box + 1, lo_v);
        atomicMax(box + 2, hi_u);
        atomicMax(box + 3, hi_v);
    }
}

__global__ void uv_bbox_init_kernel(long long* box) {
    box[0] = box[1] = LLONG_MAX;
    box[2] = box[3] = LLONG_MIN;
}

}  // namespace hv

extern "C" int hv_uv_bbox(const double* uv, int64_t n, int64_t* box, void* stream) {
    HV_REQUIRE(uv && box && n > 0, "hv_uv_bbox: bad arguments");
    cudaStream_t st = hv::as_stream(stream);
    hv::uv_bbox_init_kernel<<<1, 1, 0, st>>>(reinterpret_cast<long long*>(box));
    const long long want = (n + 255) / 256;
    hv::uv_bbox_kernel<<<(unsigned int)(want < 256 ? want : 256), 256, 0, st>>>(uv, n, reinterpret_cast<long long*>(box));
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_atlas_pack(const uint8_t* src, const int64_t* desc, int n_tex, const int64_t* row_first,
                             const int64_t* row_y0, int n_rows, int canvas_h, int canvas_w, uint8_t* atlas,
                             const double* uv, int64_t n_uv, double* uv_out, void* stream) {
    HV_REQUIRE(desc && row_first && row_y0 && n_tex > 0 && n_rows > 0, "hv_atlas_pack: bad layout arguments");
    HV_REQUIRE(canvas_h >= 0 && canvas_w >= 0 && n_uv >= 0, "hv_atlas_pack: negative size");
    cudaStream_t st = hv::as_stream(stream);
    const long long n_px = (long long)canvas_h * canvas_w;
    if (n_px > 0) {
        HV_REQUIRE(src && atlas, "hv_atlas_pack: null pixel pointer");
        const long long blocks = (n_px + 255) / 256;
        const long long cap = (long long)hv::device_info().sm_count * 16;
        hv::atlas_pixels_kernel<<<(unsigned)(blocks > cap ? cap : blocks), 256, 0, st>>>(
            src, reinterpret_cast<const long long*>(desc), reinterpret_cast<const long long*>(row_first),
            reinterpret_cast<const long long*>(row_y0), n_rows, canvas_h, canvas_w, atlas);
        HV_CUDA(cudaGetLastError());
    }
    if (n_uv > 0) {
        HV_REQUIRE(uv && uv_out, "hv_atlas_pack: null uv pointer");
        hv::atlas_uv_kernel<<<(unsigned)((n_uv + 255) / 256), 256, 0, st>>>(uv, n_uv, reinterpret_cast<const long long*>(desc),
                                                                           n_tex, canvas_h, canvas_w, uv_out);
        HV_CUDA(cudaGetLastError());
    }
    return HV_OK;
}
