// Host-side helpers shared by the launchers: error plumbing, SM count, TMA descriptor encoding.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <mutex>
#include <utility>

namespace hds {

// Every extern "C" entry point returns 0 on success or a negative HDS error / positive cudaError_t.
enum : int { HDS_OK = 0, HDS_ERR_BAD_ARG = -1, HDS_ERR_DRIVER = -2, HDS_ERR_UNSUPPORTED = -3, HDS_ERR_TOO_LARGE = -4 };

inline thread_local char g_last_error[512] = {0};
inline void set_error(const char* msg) {
    strncpy(g_last_error, msg, sizeof(g_last_error) - 1);
    g_last_error[sizeof(g_last_error) - 1] = 0;
}

#define HDS_CHECK_CUDA(expr)                                                                         \
    do {                                                                                             \
        cudaError_t _e = (expr);                                                                     \
        if (_e != cudaSuccess) {                                                                     \
            char _b[400];                                                                            \
            snprintf(_b, sizeof(_b), "%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            ::hds::set_error(_b);                                                                    \
            return static_cast<int>(_e);                                                             \
        }                                                                                            \
    } while (0)

#define HDS_REQUIRE(cond, msg)               \
    do {                                     \
        if (!(cond)) {                       \
            ::hds::set_error(msg);           \
            return ::hds::HDS_ERR_BAD_ARG;   \
        }                                    \
    } while (0)

// Every kernel of the library is launched with the programmatic-stream-serialization attribute (PDL): it may begin
// while its predecessor drains and synchronises with `griddepcontrol.wait` (hds_ptx.cuh). HDS_PDL=0 turns it off.
inline bool pdl_enabled() {
    static const int on = getenv("HDS_PDL") ? atoi(getenv("HDS_PDL")) : 1;
    return on != 0;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                 Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

inline int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (cached[dev] == 0) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

// cuTensorMapEncodeTiled is a driver entry point; fetch it through the runtime so the library
// does not link libcuda (the build host has no driver installed).
inline PFN_cuTensorMapEncodeTiled_v12000 tensor_map_encoder() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    });
    return fn;
}

// 2D row-major tensor [outer, inner] of 16-bit elements, row pitch in bytes; SWIZZLE_128B boxes
// of [box_outer, box_inner] (box_inner * 2 bytes must be <= 128). OOB reads are zero-filled.
inline int make_tmap_2d_16b(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                            uint32_t box_inner, uint32_t box_outer, bool bf16, bool swizzle128 = true) {
    auto enc = tensor_map_encoder();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
        return HDS_ERR_DRIVER;
    }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (pitch_bytes & 15)) {
        set_error("TMA operand must be 16-byte aligned with a 16-byte-multiple row pitch");
        return HDS_ERR_BAD_ARG;
    }
    cuuint64_t gdim[2] = {inner, outer};
    cuuint64_t gstride[1] = {pitch_bytes};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                     const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char b[200];
        snprintf(b, sizeof(b), "cuTensorMapEncodeTiled failed: CUresult %d (inner=%llu outer=%llu pitch=%llu box=%ux%u)",
                 (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)pitch_bytes, box_inner,
                 box_outer);
        set_error(b);
        return HDS_ERR_DRIVER;
    }
    return HDS_OK;
}

// 2D tensor of 8-bit elements (fp8 operands): box_inner is in BYTES (= elements), 128 for SWIZZLE_128B
inline int make_tmap_2d_8b(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                           uint32_t box_inner, uint32_t box_outer) {
    auto enc = tensor_map_encoder();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
        return HDS_ERR_DRIVER;
    }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (pitch_bytes & 15)) {
        set_error("TMA operand must be 16-byte aligned with a 16-byte-multiple row pitch");
        return HDS_ERR_BAD_ARG;
    }
    cuuint64_t gdim[2] = {inner, outer};
    cuuint64_t gstride[1] = {pitch_bytes};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (8-bit) failed");
        return HDS_ERR_DRIVER;
    }
    return HDS_OK;
}

// 2D tensor of PACKED 6-bit elements (e3m2 / e2m3 weights): global memory holds 16 values in 12 bytes, no gaps; the TMA
// unpacks every 16-value group into a 16-byte slot of shared memory (12 bytes + 4 bytes of padding), which is the layout
// tcgen05.mma.kind::f8f6f4 reads for 6-bit operands — so HBM carries 0.75 byte per weight while a shared-memory tile row
// stays 128 bytes for 128 elements, exactly like an fp8 tile. Rules of the format (cuTensorMapEncodeTiled): inner
// extent a multiple of 128 elements, box inner extent exactly 128, row pitch a multiple of 32 bytes, base 32-byte aligned.
// Bytes the TMA reports to the mbarrier per 128-element row of a 16U6_ALIGN16B box: the 96 packed bytes it moves, not
// the 128-byte shared-memory footprint of the unpacked row (HDS_F6_TX_PER_ROW overrides, for experiments).
inline uint32_t f6_tx_bytes_per_row() {
    static const int v = getenv("HDS_F6_TX_PER_ROW") ? atoi(getenv("HDS_F6_TX_PER_ROW")) : 96;
    return static_cast<uint32_t>(v);
}
inline int make_tmap_2d_6b(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                           uint32_t box_outer) {
    auto enc = tensor_map_encoder();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
        return HDS_ERR_DRIVER;
    }
    if ((reinterpret_cast<uintptr_t>(base) & 31) || (pitch_bytes & 31) || (inner & 127)) {
        set_error("6-bit TMA operand: base 32-byte aligned, row pitch a multiple of 32 bytes, K a multiple of 128");
        return HDS_ERR_BAD_ARG;
    }
    cuuint64_t gdim[2] = {inner, outer};
    cuuint64_t gstride[1] = {pitch_bytes};
    cuuint32_t box[2] = {128, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_16U6_ALIGN16B, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char b[160];
        snprintf(b, sizeof(b), "cuTensorMapEncodeTiled (6-bit packed) failed: CUresult %d", (int)r);
        set_error(b);
        return HDS_ERR_DRIVER;
    }
    return HDS_OK;
}

// 3D tensor [d2, d1, d0] of 16-bit elements (d0 innermost), strides in bytes for d1 and d2.
inline int make_tmap_3d_16b(CUtensorMap* map, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                            uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box0, uint32_t box1,
                            uint32_t box2, bool bf16) {
    auto enc = tensor_map_encoder();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
        return HDS_ERR_DRIVER;
    }
    cuuint64_t gdim[3] = {d0, d1, d2};
    cuuint64_t gstride[2] = {stride1_bytes, stride2_bytes};
    cuuint32_t box[3] = {box0, box1, box2};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3,
                     const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char b[160];
        snprintf(b, sizeof(b), "cuTensorMapEncodeTiled(3d) failed: CUresult %d", (int)r);
        set_error(b);
        return HDS_ERR_DRIVER;
    }
    return HDS_OK;
}

// ALiBi slopes armed by hds_attention_next_alibi for the next paged-attention launch of this thread (defined in
// attention_splitkv.cu); returns them once, then nullptr.
const float* attn_take_alibi();
int attn_take_window();   // same for hds_attention_next_window (0 = no sliding window)

}  // namespace hds
