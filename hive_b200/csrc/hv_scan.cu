// Exclusive prefix sum of (uint32, uint32) count pairs -- the compaction backbone of marching cubes (vertex / triangle
// counts per voxel group) and depth-to-mesh (vertex / face counts per 32-pixel word).  Three small kernels: per-tile
// reduce, one block scanning the tile sums (64-bit carries), per-tile scan + carry-in.  Reads each pair twice and
// writes it once: negligible next to the kernels that produce the counts.
#include "hv_common.cuh"

namespace hv {

__device__ __forceinline__ uint2 add2(uint2 a, uint2 b) { return make_uint2(a.x + b.x, a.y + b.y); }

__device__ uint2 block_exclusive_scan(uint2 v, uint2& total) {
    __shared__ uint2 warp_sums[kScanThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint2 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int ux = __shfl_up_sync(0xffffffffu, inc.x, o), uy = __shfl_up_sync(0xffffffffu, inc.y, o);
        if (lane >= o) inc = add2(inc, make_uint2(ux, uy));
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint2 carry = make_uint2(0, 0), tot = make_uint2(0, 0);
    for (int w = 0; w < kScanThreads / 32; ++w) {
        if (w < warp) carry = add2(carry, warp_sums[w]);
        tot = add2(tot, warp_sums[w]);
    }
    __syncthreads();
    total = tot;
    return make_uint2(carry.x + inc.x - v.x, carry.y + inc.y - v.y);
}

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint2* __restrict__ counts, uint2* __restrict__ bsum,
                                                                   long long n) {
    const long long base = (long long)blockIdx.x * kScanTile + (long long)threadIdx.x * kScanItems;
    uint2 s = make_uint2(0, 0);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i)
        if (base + i < n) s = add2(s, counts[base + i]);
    uint2 total;
    block_exclusive_scan(s, total);
    if (threadIdx.x == 0) bsum[blockIdx.x] = total;
}

// one block: exclusive scan of the tile sums in place; 64-bit grand totals
__global__ void __launch_bounds__(kScanThreads) scan_top_kernel(uint2* bsum, int n_blocks, unsigned long long* totals) {
    __shared__ unsigned long long carry_a, carry_b;
    if (threadIdx.x == 0) carry_a = carry_b = 0ull;
    __syncthreads();
    for (int start = 0; start < n_blocks; start += kScanThreads) {
        const int i = start + threadIdx.x;
        const uint2 v = (i < n_blocks) ? bsum[i] : make_uint2(0, 0);
        uint2 total;
        const uint2 ex = block_exclusive_scan(v, total);
        const unsigned long long ca = carry_a, cb = carry_b;
        if (i < n_blocks) bsum[i] = make_uint2((unsigned int)(ca + ex.x), (unsigned int)(cb + ex.y));
        __syncthreads();
        if (threadIdx.x == 0) {
            carry_a = ca + total.x;
            carry_b = cb + total.y;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0 && totals) {
        totals[0] = carry_a;
        totals[1] = carry_b;
    }
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint2* __restrict__ counts, uint2* __restrict__ offs,
                                                                  const uint2* __restrict__ bsum, long long n) {
    const long long base = (long long)blockIdx.x * kScanTile + (long long)threadIdx.x * kScanItems;
    uint2 item[kScanItems];
    uint2 s = make_uint2(0, 0);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        item[i] = (base + i < n) ? counts[base + i] : make_uint2(0, 0);
        s = add2(s, item[i]);
    }
    uint2 total;
    uint2 run = add2(block_exclusive_scan(s, total), bsum[blockIdx.x]);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        if (base + i < n) offs[base + i] = run;
        run = add2(run, item[i]);
    }
}

long long scan_tiles(long long n) { return (n + kScanTile - 1) / kScanTile; }

void exclusive_scan_u2(const uint2* counts, uint2* offs, uint2* bsum, long long n, unsigned long long* totals,
                       cudaStream_t st) {
    const int tiles = (int)scan_tiles(n);
    if (tiles <= 0) return;
    scan_reduce_kernel<<<tiles, kScanThreads, 0, st>>>(counts, bsum, n);
    scan_top_kernel<<<1, kScanThreads, 0, st>>>(bsum, tiles, totals);
    scan_apply_kernel<<<tiles, kScanThreads, 0, st>>>(counts, offs, bsum, n);
}

}  // namespace hv
