// Group-wise quantizers: integer (int8 / int4, symmetric or asymmetric) and minifloat (FP8 e4m3, FP6 e3m2, FP4 e2m1,
// FP12 e4m7, all-finite encodings scaled per group), plus their dequantizers and a row-selective dequantizer.
// One CTA per group: block-wide max(/min) reduction, then a vectorised quantise-and-pack pass. Values are packed
// little-endian at their natural bit width (4-bit: 2 per byte, 6-bit: 4 per 3 bytes, 12-bit: 2 per 3 bytes).
// Role in the reference: csrc/quantization/{quantize,dequantize,quantize_intX}.cu, csrc/fp_quantizer/fp_quantize.cu.
#include "hds_host.cuh"
#include "hds_ptx.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

constexpr int kQThreads = 256;

__device__ __forceinline__ float block_reduce_max(float v, float* sh) {
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = sh[0];
    for (int w = 1; w < kQThreads / 32; ++w) r = fmaxf(r, sh[w]);
    __syncthreads();
    return r;
}

// ---- integer -------------------------------------------------------------------------------------------------
template <int DT>
__global__ void __launch_bounds__(kQThreads)
quantize_int_kernel(const void* __restrict__ src, int8_t* __restrict__ dst, float* __restrict__ params, int group_size,
                    int bits, int symmetric) {
    __shared__ float sh[kQThreads / 32];
    const int64_t base = static_cast<int64_t>(blockIdx.x) * group_size;
    float mx = -INFINITY, mn = INFINITY;
    for (int i = threadIdx.x; i < group_size; i += kQThreads) {
        const float v = load_any<DT>(src, base + i);
        mx = fmaxf(mx, v);
        mn = fminf(mn, v);
    }
    mx = block_reduce_max(mx, sh);
    mn = -block_reduce_max(-mn, sh);
    const float qmax = static_cast<float>((1 << (bits - 1)) - 1), qmin = -static_cast<float>(1 << (bits - 1));
    float scale, offset;
    if (symmetric) {
        const float a = fmaxf(fabsf(mx), fabsf(mn));
        scale = a > 0.0f ? a / qmax : 1.0f;
        offset = 0.0f;
    } else {
        scale = mx > mn ? (mx - mn) / static_cast<float>((1 << bits) - 1) : 1.0f;
        offset = mn - qmin * scale;  // value represented by code qmin is mn
    }
    if (threadIdx.x == 0) {
        params[2 * blockIdx.x] = scale;
        params[2 * blockIdx.x + 1] = offset;
    }
    const float inv = 1.0f / scale;
    if (bits == 8) {
        for (int i = threadIdx.x; i < group_size; i += kQThreads) {
            const float q = rintf((load_any<DT>(src, base + i) - offset) * inv);
            dst[base + i] = static_cast<int8_t>(fminf(fmaxf(q, qmin), qmax));
        }
    } else {  // 4 bits: two codes per byte, low nibble first
        for (int i = threadIdx.x; i < group_size / 2; i += kQThreads) {
            const float q0 = fminf(fmaxf(rintf((load_any<DT>(src, base + 2 * i) - offset) * inv), qmin), qmax);
            const float q1 = fminf(fmaxf(rintf((load_any<DT>(src, base + 2 * i + 1) - offset) * inv), qmin), qmax);
            dst[base / 2 + i] = static_cast<int8_t>((static_cast<int>(q0) & 0xf) | ((static_cast<int>(q1) & 0xf) << 4));
        }
    }
}

template <int DT>
__global__ void __launch_bounds__(kQThreads)
dequantize_int_kernel(const int8_t* __restrict__ src, const float* __restrict__ params, void* __restrict__ dst,
                      int group_size, int bits) {
    const int64_t base = static_cast<int64_t>(blockIdx.x) * group_size;
    const float scale = params[2 * blockIdx.x], offset = params[2 * blockIdx.x + 1];
    for (int i = threadIdx.x; i < group_size; i += kQThreads) {
        int q;
        if (bits == 8) {
            q = src[base + i];
        } else {
            const int b = src[(base + i) >> 1];
            q = (i & 1) ? (b >> 4) : ((b << 28) >> 28);  // sign-extend the nibble
            if (i & 1) q = (q << 28) >> 28;
        }
        const float v = static_cast<float>(q) * scale + offset;
        if constexpr (DT == 2) reinterpret_cast<float*>(dst)[base + i] = v;
        else store_scalar<DT == 0>(dst, base + i, v);
    }
}

// ---- minifloat ----------------------------------------------------------------------------------------------
struct FpFormat {
    int ebits, mbits, bias;
    float max_value;
};
__host__ __device__ inline FpFormat fp_format(int q_bits) {
    switch (q_bits) {
        case 8: return {4, 3, 7, 480.0f};
        case 6: return {3, 2, 3, 28.0f};
        case 4: return {2, 1, 1, 6.0f};
        default: return {4, 7, 7, 510.0f};  // 12
    }
}

__device__ __forceinline__ uint32_t fp_encode(float y, const FpFormat f) {
    const uint32_t sign = y < 0.0f ? 1u : 0u;
    float a = fminf(fabsf(y), f.max_value);
    const float min_normal = ldexpf(1.0f, 1 - f.bias);
    uint32_t code;
    if (a < min_normal) {
        code = static_cast<uint32_t>(rintf(ldexpf(a, f.bias - 1 + f.mbits)));  // subnormal step 2^(1-bias-m)
    } else {
        int e;
        const float fr = frexpf(a, &e);  // a = fr * 2^e, fr in [0.5, 1)
        int eu = e - 1;
        uint32_t mant = static_cast<uint32_t>(rintf((fr * 2.0f - 1.0f) * static_cast<float>(1 << f.mbits)));
        if (mant == (1u << f.mbits)) {
            mant = 0;
            ++eu;
        }
        code = (static_cast<uint32_t>(eu + f.bias) << f.mbits) | mant;
        const uint32_t max_code = (1u << (f.ebits + f.mbits)) - 1;
        code = code > max_code ? max_code : code;
    }
    return (sign << (f.ebits + f.mbits)) | code;
}

__device__ __forceinline__ float fp_decode(uint32_t code, const FpFormat f) {
    const uint32_t sign = (code >> (f.ebits + f.mbits)) & 1u;
    const uint32_t e = (code >> f.mbits) & ((1u << f.ebits) - 1);
    const uint32_t m = code & ((1u << f.mbits) - 1);
    const float v = e == 0 ? ldexpf(static_cast<float>(m), 1 - f.bias - f.mbits)
                           : ldexpf(1.0f + static_cast<float>(m) / static_cast<float>(1 << f.mbits), static_cast<int>(e) - f.bias);
    return sign ? -v : v;
}

__host__ __device__ inline int64_t fp_packed_bytes(int64_t n, int q_bits) { return (n * q_bits + 7) / 8; }

__device__ __forceinline__ void fp_store_codes(uint8_t* dst, int64_t i, int q_bits, const uint32_t* c) {
    // called per packing unit starting at element i
    if (q_bits == 8) dst[i] = static_cast<uint8_t>(c[0]);
    else if (q_bits == 4) dst[i >> 1] = static_cast<uint8_t>(c[0] | (c[1] << 4));
    else if (q_bits == 6) {
        const uint32_t w = c[0] | (c[1] << 6) | (c[2] << 12) | (c[3] << 18);
        uint8_t* p = dst + (i >> 2) * 3;
        p[0] = w & 0xff; p[1] = (w >> 8) & 0xff; p[2] = (w >> 16) & 0xff;
    } else {
        const uint32_t w = c[0] | (c[1] << 12);
        uint8_t* p = dst + (i >> 1) * 3;
        p[0] = w & 0xff; p[1] = (w >> 8) & 0xff; p[2] = (w >> 16) & 0xff;
    }
}

__device__ __forceinline__ uint32_t fp_load_code(const uint8_t* src, int64_t i, int q_bits) {
    if (q_bits == 8) return src[i];
    if (q_bits == 4) return (src[i >> 1] >> ((i & 1) * 4)) & 0xf;
    if (q_bits == 6) {
        const uint8_t* p = src + (i >> 2) * 3;
        const uint32_t w = p[0] | (p[1] << 8) | (p[2] << 16);
        return (w >> ((i & 3) * 6)) & 0x3f;
    }
    const uint8_t* p = src + (i >> 1) * 3;
    const uint32_t w = p[0] | (p[1] << 8) | (p[2] << 16);
    return (w >> ((i & 1) * 12)) & 0xfff;
}

template <int DT>
__global__ void __launch_bounds__(kQThreads)
fp_quantize_kernel(const void* __restrict__ src, uint8_t* __restrict__ dst, float* __restrict__ scales, int group_size,
                   int q_bits) {
    __shared__ float sh[kQThreads / 32];
    const FpFormat f = fp_format(q_bits);
    const int64_t base = static_cast<int64_t>(blockIdx.x) * group_size;
    float mx = 0.0f;
    for (int i = threadIdx.x; i < group_size; i += kQThreads) mx = fmaxf(mx, fabsf(load_any<DT>(src, base + i)));
    mx = block_reduce_max(mx, sh);
    const float scale = mx > 0.0f ? mx / f.max_value : 1.0f;
    if (threadIdx.x == 0) scales[blockIdx.x] = scale;
    const float inv = 1.0f / scale;
    const int unit = q_bits == 8 ? 1 : (q_bits == 6 ? 4 : 2);
    uint8_t* out = dst + fp_packed_bytes(base, q_bits);
    for (int u = threadIdx.x; u < group_size / unit; u += kQThreads) {
        uint32_t c[4];
        for (int j = 0; j < unit; ++j) c[j] = fp_encode(load_any<DT>(src, base + u * unit + j) * inv, f);
        fp_store_codes(out, static_cast<int64_t>(u) * unit, q_bits, c);
    }
}

template <int DT>
__global__ void __launch_bounds__(kQThreads)
fp_dequantize_kernel(const uint8_t* __restrict__ src, const float* __restrict__ scales, void* __restrict__ dst,
                     const int32_t* __restrict__ group_idx, int group_size, int q_bits) {
    const FpFormat f = fp_format(q_bits);
    const int64_t g = group_idx ? group_idx[blockIdx.x] : blockIdx.x;  // selective: output row i <- group group_idx[i]
    const uint8_t* in = src + fp_packed_bytes(g * group_size, q_bits);
    const float scale = scales[g];
    const int64_t obase = static_cast<int64_t>(blockIdx.x) * group_size;
    for (int i = threadIdx.x; i < group_size; i += kQThreads) {
        const float v = fp_decode(fp_load_code(in, i, q_bits), f) * scale;
        if constexpr (DT == 2) reinterpret_cast<float*>(dst)[obase + i] = v;
        else store_scalar<DT == 0>(dst, obase + i, v);
    }
}

}  // namespace hds

using namespace hds;

#define HDS_DT_SWITCH(dt, CALL)                \
    do {                                       \
        if (dt == 0) { CALL(0); }              \
        else if (dt == 1) { CALL(1); }         \
        else { CALL(2); }                      \
    } while (0)

extern "C" int hds_quantize_int(const void* src, int dtype, int64_t n_groups, int group_size, int bits, int symmetric,
                                void* dst, float* params, void* stream) {
    HDS_REQUIRE(bits == 4 || bits == 8, "hds_quantize_int: 4 or 8 bits");
    HDS_REQUIRE(dtype >= 0 && dtype <= 2 && group_size > 0 && group_size % 2 == 0, "hds_quantize_int: bad dtype / group size");
    if (n_groups == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
#define CALL(D) quantize_int_kernel<D><<<static_cast<unsigned>(n_groups), kQThreads, 0, s>>>(src, static_cast<int8_t*>(dst), params, group_size, bits, symmetric)
    HDS_DT_SWITCH(dtype, CALL);
#undef CALL
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_dequantize_int(const void* src, const float* params, int64_t n_groups, int group_size, int bits,
                                  void* dst, int dtype, void* stream) {
    HDS_REQUIRE(bits == 4 || bits == 8, "hds_dequantize_int: 4 or 8 bits");
    if (n_groups == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
#define CALL(D) dequantize_int_kernel<D><<<static_cast<unsigned>(n_groups), kQThreads, 0, s>>>(static_cast<const int8_t*>(src), params, dst, group_size, bits)
    HDS_DT_SWITCH(dtype, CALL);
#undef CALL
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_fp_quantize(const void* src, int dtype, int64_t n_groups, int group_size, int q_bits, void* dst,
                               float* scales, void* stream) {
    HDS_REQUIRE(q_bits == 8 || q_bits == 6 || q_bits == 4 || q_bits == 12, "hds_fp_quantize: q_bits must be 4, 6, 8 or 12");
    HDS_REQUIRE(group_size % 4 == 0, "hds_fp_quantize: group size must be a multiple of 4");
    if (n_groups == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
#define CALL(D) fp_quantize_kernel<D><<<static_cast<unsigned>(n_groups), kQThreads, 0, s>>>(src, static_cast<uint8_t*>(dst), scales, group_size, q_bits)
    HDS_DT_SWITCH(dtype, CALL);
#undef CALL
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_fp_dequantize(const void* src, const float* scales, const int32_t* group_idx, int64_t n_out_groups,
                                 int group_size, int q_bits, void* dst, int dtype, void* stream) {
    HDS_REQUIRE(q_bits == 8 || q_bits == 6 || q_bits == 4 || q_bits == 12, "hds_fp_dequantize: q_bits must be 4, 6, 8 or 12");
    if (n_out_groups == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
#define CALL(D) fp_dequantize_kernel<D><<<static_cast<unsigned>(n_out_groups), kQThreads, 0, s>>>(static_cast<const uint8_t*>(src), scales, dst, group_idx, group_size, q_bits)
    HDS_DT_SWITCH(dtype, CALL);
#undef CALL
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}
