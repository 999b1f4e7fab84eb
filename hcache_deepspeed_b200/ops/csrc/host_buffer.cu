// Host staging buffers with explicit allocation flags (role of the reference's fast host buffer,
// deepspeed/inference/v2/ragged/csrc/fast_host_buffer.cu: cudaHostAlloc(Portable | Mapped | WriteCombined)).
//
// A batch-metadata staging buffer is written once by the host and read once by the copy engine (or, when mapped, by
// the GPU directly), which is the access pattern write-combined memory is for: stores bypass the CPU caches, so the
// DMA read finds the data in DRAM without snooping. Reads BY THE HOST from write-combined memory are uncached and slow,
// so the flag is a choice of the caller; the default here is plain portable pinned memory.
#include <cuda_runtime.h>

#include "hds_b200.h"
#include "hds_host.cuh"

using namespace hds;

extern "C" int hds_host_buffer_alloc(int64_t bytes, int mapped, int write_combined, void** out) {
    HDS_REQUIRE(out != nullptr && bytes > 0, "hds_host_buffer_alloc: bytes > 0 and a result pointer are required");
    unsigned int flags = cudaHostAllocPortable;
    if (mapped) flags |= cudaHostAllocMapped;
    if (write_combined) flags |= cudaHostAllocWriteCombined;
    void* p = nullptr;
    HDS_CHECK_CUDA(cudaHostAlloc(&p, static_cast<size_t>(bytes), flags));
    *out = p;
    return HDS_OK;
}

// Device-side address of a MAPPED buffer (kernels may read the staging area in place, no copy at all).
extern "C" int hds_host_buffer_device_ptr(void* host_ptr, void** out) {
    HDS_REQUIRE(host_ptr != nullptr && out != nullptr, "hds_host_buffer_device_ptr: null argument");
    HDS_CHECK_CUDA(cudaHostGetDevicePointer(out, host_ptr, 0));
    return HDS_OK;
}

extern "C" int hds_host_buffer_free(void* host_ptr) {
    if (host_ptr == nullptr) return HDS_OK;
    HDS_CHECK_CUDA(cudaFreeHost(host_ptr));
    return HDS_OK;
}
