// Microbenchmark (not part of the product): what does the B200 memory system give for the integrate kernel's access
// pattern with NO arithmetic?  Three float planes [X][Y][Z], every float4 read, incremented and written back.
//   mode 0: flat streaming (thread i -> float4 i, grid-stride)                     -- the copy-like ceiling
//   mode 1: v3 tiling   (warp = 128 z of one column, 8 warps = 8 adjacent y, seg fastest over CTAs, persistent)
//   mode 2: column walk (warp walks all segs of its column, persistent over column groups)
//   mode 3: column walk, 2 tiles in flight per warp (loads of seg s+1 issued before the stores of seg s)
// nvcc -O3 -arch=sm_100a -o rmw_pattern rmw_pattern.cu && ./rmw_pattern
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ldcs(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void stcs(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }
__device__ __forceinline__ float4 inc(float4 v) { return make_float4(v.x + 1.f, v.y + 1.f, v.z + 1.f, v.w + 1.f); }

template <int kMode>
__global__ void __launch_bounds__(256) rmw(float* a, float* b, float* c, int X, int Y, int Z) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int segs = Z / 128, ytiles = Y / 8;
    if (kMode == 0) {
        const size_t n4 = (size_t)X * Y * Z / 4;
        for (size_t i = blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
            const float4 va = ldcs(a + 4 * i), vb = ldcs(b + 4 * i), vc = ldcs(c + 4 * i);
            stcs(a + 4 * i, inc(va)); stcs(b + 4 * i, inc(vb)); stcs(c + 4 * i, inc(vc));
        }
    } else if (kMode == 1) {
        const int n_tiles = X * ytiles * segs;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int t = tile / segs, seg = tile - t * segs, x = t / ytiles, y = (t - x * ytiles) * 8 + warp;
            const size_t off = ((size_t)(x * Y + y) * Z) + (seg * 32 + lane) * 4;
            const float4 va = ldcs(a + off), vb = ldcs(b + off), vc = ldcs(c + off);
            stcs(a + off, inc(va)); stcs(b + off, inc(vb)); stcs(c + off, inc(vc));
        }
    } else {
        const int n_groups = X * ytiles;
        for (int g = blockIdx.x; g < n_groups; g += gridDim.x) {
            const int x = g / ytiles, y = (g - x * ytiles) * 8 + warp;
            const size_t col = (size_t)(x * Y + y) * Z + lane * 4;
            if (kMode == 2) {
                for (int seg = 0; seg < segs; ++seg) {
                    const size_t off = col + seg * 128;
                    const float4 va = ldcs(a + off), vb = ldcs(b + off), vc = ldcs(c + off);
                    stcs(a + off, inc(va)); stcs(b + off, inc(vb)); stcs(c + off, inc(vc));
                }
            } else {
                float4 va = ldcs(a + col), vb = ldcs(b + col), vc = ldcs(c + col);
                for (int seg = 0; seg < segs; ++seg) {
                    const size_t off = col + seg * 128;
                    float4 na = va, nb = vb, nc = vc;
                    if (seg + 1 < segs) { na = ldcs(a + off + 128); nb = ldcs(b + off + 128); nc = ldcs(c + off + 128); }
                    stcs(a + off, inc(va)); stcs(b + off, inc(vb)); stcs(c + off, inc(vc));
                    va = na; vb = nb; vc = nc;
                }
            }
        }
    }
}

template <int kMode>
static void run(const char* name, float* a, float* b, float* c, int X, int Y, int Z, int ctas_per_sm) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int grid = sms * ctas_per_sm;
    if (kMode != 0) grid -= 1;  // coprime with segs * ytiles
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) rmw<kMode><<<grid, 256>>>(a, b, c, X, Y, Z);
    cudaEventRecord(e0);
    const int reps = 20;
    for (int i = 0; i < reps; ++i) rmw<kMode><<<grid, 256>>>(a, b, c, X, Y, Z);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= reps;
    const double bytes = 24.0 * X * Y * Z;
    printf("%-28s ctas/SM %d  %.4f ms  %.0f GB/s  (%s)\n", name, ctas_per_sm, ms, bytes / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const int X = 512, Y = 512, Z = 512;
    const size_t n = (size_t)X * Y * Z;
    float *a, *b, *c;
    cudaMalloc(&a, n * 4); cudaMalloc(&b, n * 4); cudaMalloc(&c, n * 4);
    cudaMemset(a, 0, n * 4); cudaMemset(b, 0, n * 4); cudaMemset(c, 0, n * 4);
    for (int cps = 2; cps <= 8; cps += (cps < 4 ? 1 : 2)) {
        run<0>("flat streaming", a, b, c, X, Y, Z, cps);
        run<1>("v3 tiling", a, b, c, X, Y, Z, cps);
        run<2>("column walk", a, b, c, X, Y, Z, cps);
        run<3>("column walk, 2 in flight", a, b, c, X, Y, Z, cps);
    }
    return 0;
}
