// Exclusive prefix sum of (uint32, uint32) count pairs -- the compaction backbone of marching cubes (vertex / triangle
// counts per voxel group) and depth-to-mesh (vertex / face counts per 32-pixel word).  One pass with decoupled look-back
// (scan_onepass_kernel below); the three-kernel form it replaced (reduce / top / apply) is kept for A/B and as the
// reference in tests of the scan (HV_SCAN_THREE_PASS=1 in the environment selects it).
#include "hv_common.cuh"
#include <stdlib.h>

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
long long scan_scratch_pairs(long long n) { return scan_tiles(n) + 1; }  // one status word per tile + the ticket

// ---- single pass: decoupled look-back ------------------------------------------------------------------------------------
// r01 launch list: the three-kernel form cost 12 + 8 + 25 us per marching-cubes extraction at 512^3 (4.2 M pairs), the
// middle kernel a single block.  One kernel now: tiles take a ticket (so every predecessor of a running tile is itself
// running or done), publish {flag, x, y} as ONE 64-bit word -- flag 1 = tile aggregate, 2 = inclusive prefix, 31 bits
// per component (all totals are < 2^31: hv_mc_emit / hv_d2m_emit require it) -- and warp 0 looks back 32 tiles at a time.
constexpr unsigned long long kScanAgg = 1ull << 62, kScanInc = 2ull << 62;
__device__ __forceinline__ unsigned long long scan_pack(unsigned long long flag, uint2 v) {
    return flag | ((unsigned long long)(v.y & 0x7fffffffu) << 31) | (unsigned long long)(v.x & 0x7fffffffu);
}

__global__ void __launch_bounds__(kScanThreads) scan_onepass_kernel(const uint2* __restrict__ counts, uint2* __restrict__ offs,
                                                                    long long n, int tiles, volatile unsigned long long* state,
                                                                    unsigned int* ticket, unsigned long long* totals) {
    __shared__ unsigned int s_tile;
    __shared__ uint2 s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const long long base = (long long)tile * kScanTile + (long long)threadIdx.x * kScanItems;
    uint2 item[kScanItems];
    uint2 s = make_uint2(0, 0);
    const bool full = base + kScanItems <= n;  // the thread's 64 bytes as four 16-byte accesses (arrays are 256-byte aligned)
    if (full) {
        const uint4* src = reinterpret_cast<const uint4*>(counts + base);
#pragma unroll
        for (int i = 0; i < kScanItems / 2; ++i) {
            const uint4 q = __ldg(src + i);
            item[2 * i] = make_uint2(q.x, q.y);
            item[2 * i + 1] = make_uint2(q.z, q.w);
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) item[i] = (base + i < n) ? counts[base + i] : make_uint2(0, 0);
    }
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) s = add2(s, item[i]);
    uint2 total;
    const uint2 ex = block_exclusive_scan(s, total);
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        uint2 acc = make_uint2(0, 0);
        if (tile > 0) {
            if (lane == 0) state[tile] = scan_pack(kScanAgg, total);
            for (int p = tile - 1;; p -= 32) {
                const int idx = p - lane;
                unsigned long long w = kScanInc;  // tiles before the first: an inclusive prefix of zero
                if (idx >= 0) {
                    do {
                        w = state[idx];
                    } while ((w >> 62) == 0ull);
                }
                const unsigned int inc_mask = __ballot_sync(0xffffffffu, (w >> 62) == 2ull);
                const int first = inc_mask ? (__ffs(inc_mask) - 1) : 32;
                unsigned int vx = (lane <= first) ? (unsigned int)(w & 0x7fffffffull) : 0u;
                unsigned int vy = (lane <= first) ? (unsigned int)((w >> 31) & 0x7fffffffull) : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    vx += __shfl_xor_sync(0xffffffffu, vx, o);
                    vy += __shfl_xor_sync(0xffffffffu, vy, o);
                }
                acc = add2(acc, make_uint2(vx, vy));
                if (inc_mask) break;
            }
        }
        if (lane == 0) {
            state[tile] = scan_pack(kScanInc, add2(acc, total));
            s_prefix = acc;
            if (tile == tiles - 1 && totals) {
                totals[0] = (unsigned long long)acc.x + total.x;
                totals[1] = (unsigned long long)acc.y + total.y;
            }
        }
    }
    __syncthreads();
    uint2 run = add2(ex, s_prefix);
    if (full) {
        uint4* dst = reinterpret_cast<uint4*>(offs + base);
#pragma unroll
        for (int i = 0; i < kScanItems / 2; ++i) {
            const uint2 a = run;
            const uint2 b = add2(run, item[2 * i]);
            dst[i] = make_uint4(a.x, a.y, b.x, b.y);
            run = add2(b, item[2 * i + 1]);
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) {
            if (base + i < n) offs[base + i] = run;
            run = add2(run, item[i]);
        }
    }
}

// `bsum`: scratch of scan_scratch_pairs(n) pairs (8 bytes each).
void exclusive_scan_u2(const uint2* counts, uint2* offs, uint2* bsum, long long n, unsigned long long* totals,
                       cudaStream_t st) {
    const int tiles = (int)scan_tiles(n);
    if (tiles <= 0) return;
    static const bool three_pass = getenv("HV_SCAN_THREE_PASS") != nullptr;
    if (three_pass) {
        scan_reduce_kernel<<<tiles, kScanThreads, 0, st>>>(counts, bsum, n);
        scan_top_kernel<<<1, kScanThreads, 0, st>>>(bsum, tiles, totals);
        scan_apply_kernel<<<tiles, kScanThreads, 0, st>>>(counts, offs, bsum, n);
        return;
    }
    cudaMemsetAsync(bsum, 0, (size_t)(tiles + 1) * sizeof(uint2), st);
    unsigned long long* state = reinterpret_cast<unsigned long long*>(bsum);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(state + tiles);
    scan_onepass_kernel<<<tiles, kScanThreads, 0, st>>>(counts, offs, n, tiles, state, ticket, totals);
}

}  // namespace hv
