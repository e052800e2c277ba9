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
