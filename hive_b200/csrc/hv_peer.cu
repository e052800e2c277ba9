// Device memory that the other ranks of one node can store into (hive_b200/sharded.py: the whole-volume mesh is assembled
// in rank 0's HBM by the emit kernels of all slabs, each writing its rows over NVLink -- no gather, no scatter).
//
// One process per GPU, so a buffer crosses processes as a CUDA IPC handle: the owner allocates with cudaMalloc (IPC handles
// cover whole cudaMalloc allocations; a block of torch's caching allocator would export its neighbours too), the 64-byte
// handle travels through whatever the ranks already talk over, and every other rank maps it once.  Opening the handle
// enables peer access between the two devices.  The reference is single-GPU: nothing there corresponds to this file.
#include "hv_common.cuh"
#include <string.h>

static_assert(sizeof(cudaIpcMemHandle_t) == HV_PEER_HANDLE_BYTES, "HV_PEER_HANDLE_BYTES must be the size of a CUDA IPC handle");

extern "C" int hv_peer_alloc(int64_t bytes, void** ptr, void* handle) {
    HV_REQUIRE(ptr && handle && bytes > 0, "hv_peer_alloc: bad arguments");
    void* p = nullptr;
    HV_CUDA(cudaMalloc(&p, (size_t)bytes));
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return hv::cuda_fail(e, "cudaIpcGetMemHandle");
    }
    memcpy(handle, &h, sizeof(h));
    *ptr = p;
    return HV_OK;
}

extern "C" int hv_peer_free(void* ptr) {
    if (ptr) HV_CUDA(cudaFree(ptr));
    return HV_OK;
}

extern "C" int hv_peer_open(const void* handle, void** ptr) {
    HV_REQUIRE(handle && ptr, "hv_peer_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    HV_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *ptr = p;
    return HV_OK;
}

extern "C" int hv_peer_close(void* ptr) {
    if (ptr) HV_CUDA(cudaIpcCloseMemHandle(ptr));
    return HV_OK;
}

// Plain asynchronous copy between two device (or mapped peer / pinned host) addresses, for buffers torch does not own.
extern "C" int hv_copy_async(void* dst, const void* src, int64_t bytes, void* stream) {
    HV_REQUIRE(bytes >= 0 && (bytes == 0 || (dst && src)), "hv_copy_async: bad arguments");
    if (bytes) HV_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, hv::as_stream(stream)));
    return HV_OK;
}

// ---- signals between the GPUs of a node: a 32-bit epoch per flag, stored into the peer's memory -------------------------
// Kernels of one stream run in order and a kernel's stores -- those into a peer's memory included -- are performed before the
// next kernel starts, so "data, then flag" needs no fence inside a kernel: the data is written by the kernels before
// hv_peer_signal, the flag by the one-warp kernel of hv_peer_signal itself.  The receiver's hv_peer_wait spins (system-scope
// acquire loads) on flags in its OWN memory until every one has reached the epoch -- a peer may still be busy with work queued
// before its signal, so the wait is long -- and gives up after ~30 s (and reports which
// flag through `err`) instead of hanging the GPU.
namespace hv {

constexpr int kMaxPeers = 16;
struct FlagPtrs {
    unsigned int* p[kMaxPeers];
};

__global__ void peer_signal_kernel(const FlagPtrs flags, const int n, const unsigned int epoch) {
    const int i = threadIdx.x;
    if (i < n && flags.p[i]) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flags.p[i]), "r"(epoch) : "memory");
}

__global__ void peer_wait_kernel(const unsigned int* flags, const int n, const unsigned int epoch, const long long timeout_clocks,
                                 unsigned int* err) {
    const int i = threadIdx.x;
    if (i >= n) return;
    const long long t0 = clock64();
    for (;;) {
        unsigned int v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + i) : "memory");
        if ((int)(v - epoch) >= 0) return;
        if (clock64() - t0 > timeout_clocks) {
            atomicMax(err, (unsigned int)(i + 1));
            return;
        }
        __nanosleep(200);
    }
}

// dst_a[c] = a[c * pitch + z], dst_b[c] = b[c * pitch + z]: one z-plane of two [X][Y][pitch] volumes (the seam planes)
__global__ void __launch_bounds__(256) copy_planes_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                           const long long pitch, const long long z, const long long n_cols,
                                                           float* __restrict__ dst_a, float* __restrict__ dst_b) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    if (c >= n_cols) return;
    dst_a[c] = __ldg(a + c * pitch + z);
    dst_b[c] = __ldg(b + c * pitch + z);
}

}  // namespace hv

extern "C" int hv_peer_signal(void* const* h_flags, int n, unsigned int epoch, void* stream) {
    HV_REQUIRE(h_flags && n >= 1 && n <= hv::kMaxPeers, "hv_peer_signal: 1..16 flags");
    hv::FlagPtrs f;
    for (int i = 0; i < hv::kMaxPeers; ++i) f.p[i] = i < n ? static_cast<unsigned int*>(h_flags[i]) : nullptr;
    hv::peer_signal_kernel<<<1, 32, 0, hv::as_stream(stream)>>>(f, n, epoch);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_peer_wait(const void* flags, int n, unsigned int epoch, void* err, void* stream) {
    HV_REQUIRE(flags && err && n >= 1 && n <= 32, "hv_peer_wait: 1..32 flags");
    hv::peer_wait_kernel<<<1, 32, 0, hv::as_stream(stream)>>>(static_cast<const unsigned int*>(flags), n, epoch,
                                                             60000000000ll /* ~30 s at 1.9 GHz */, static_cast<unsigned int*>(err));
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}

extern "C" int hv_copy_planes(const float* a, const float* b, int64_t pitch, int64_t z, int64_t n_cols, float* dst_a, float* dst_b,
                              void* stream) {
    HV_REQUIRE(a && b && dst_a && dst_b && pitch > 0 && z >= 0 && z < pitch && n_cols > 0, "hv_copy_planes: bad arguments");
    hv::copy_planes_kernel<<<(unsigned int)((n_cols + 255) / 256), 256, 0, hv::as_stream(stream)>>>(a, b, pitch, z, n_cols, dst_a, dst_b);
    HV_CUDA(cudaGetLastError());
    return HV_OK;
}
