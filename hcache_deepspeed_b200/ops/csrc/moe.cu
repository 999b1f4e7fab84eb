// Mixture-of-experts routing kernels for the ragged forward (tensor-parallel experts, no expert parallelism):
//
//   top-k gating  : softmax over the expert logits of each token, k rounds of arg-max (ties -> lower index), optional
//                   renormalisation of the k scores, one atomic per (token, choice) to claim a row inside the expert
//   offsets       : per-expert row ranges padded to the GEMM's 128-row tile (so no A tile straddles two experts);
//                   re-zeroes the counters for the next layer
//   scatter       : copies each token's activation row to its k expert rows, records the row index per choice
//   gather        : out[t] += sum_j score[t, j] * y[row(t, j)]   (fp32, straight into the pending residual delta)
//
// The expert GEMMs in between are ONE grouped tcgen05 launch each (gemm_core.cuh, GemmArgs::group_pad_off).
// Role in the reference: ragged_ops/top_k_gating/top_k_gating_cuda.cu:14-87, moe_scatter/moe_scatter_cuda.cu:22-150,
// moe_gather/moe_gather_cuda.cu:21-112, cutlass_ops/moe_gemm (external library).
#include "hds_host.cuh"
#include "hds_ptx.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

constexpr int kGateThreads = 128;
constexpr int kMoeMaxK = 8;
constexpr int kMoeRowPad = 128;

template <int DT>
__global__ void __launch_bounds__(kGateThreads)
moe_topk_gating_kernel(const void* __restrict__ logits, int64_t ld, int E, int k, int normalize,
                       float* __restrict__ scores, int32_t* __restrict__ expert_ids, int32_t* __restrict__ offsets,
                       int32_t* __restrict__ counts) {
    extern __shared__ float probs[];  // [E]
    __shared__ float red_v[kGateThreads / 32];
    __shared__ int red_i[kGateThreads / 32];
    __shared__ float sel_v[kMoeMaxK];
    __shared__ int sel_i[kMoeMaxK];
    pdl_enter();
    const int t = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float mx = -INFINITY;
    for (int e = tid; e < E; e += kGateThreads) {
        const float v = load_any<DT>(logits, static_cast<int64_t>(t) * ld + e);
        probs[e] = v;
        mx = fmaxf(mx, v);
    }
    mx = warp_max(mx);
    if (lane == 0) red_v[warp] = mx;
    __syncthreads();
    mx = red_v[0];
#pragma unroll
    for (int w = 1; w < kGateThreads / 32; ++w) mx = fmaxf(mx, red_v[w]);
    __syncthreads();
    float sum = 0.0f;
    for (int e = tid; e < E; e += kGateThreads) {
        const float p = __expf(probs[e] - mx);
        probs[e] = p;
        sum += p;
    }
    sum = warp_sum(sum);
    if (lane == 0) red_v[warp] = sum;
    __syncthreads();
    sum = 0.0f;
#pragma unroll
    for (int w = 0; w < kGateThreads / 32; ++w) sum += red_v[w];
    const float inv = 1.0f / sum;
    __syncthreads();
    for (int j = 0; j < k; ++j) {
        float bv = -1.0f;
        int bi = 0x7fffffff;
        for (int e = tid; e < E; e += kGateThreads) {
            const float p = probs[e];
            if (p > bv || (p == bv && e < bi)) {
                bv = p;
                bi = e;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) {
                bv = ov;
                bi = oi;
            }
        }
        if (lane == 0) {
            red_v[warp] = bv;
            red_i[warp] = bi;
        }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < kGateThreads / 32; ++w) {
                if (red_v[w] > bv || (red_v[w] == bv && red_i[w] < bi)) {
                    bv = red_v[w];
                    bi = red_i[w];
                }
            }
            sel_v[j] = bv * inv;
            sel_i[j] = bi;
            probs[bi] = -1.0f;  // taken
        }
        __syncthreads();
    }
    if (tid < k) {
        float s = sel_v[tid];
        if (normalize) {
            float tot = 0.0f;
            for (int j = 0; j < k; ++j) tot += sel_v[j];
            s /= tot;
        }
        const int e = sel_i[tid];
        scores[static_cast<int64_t>(t) * k + tid] = s;
        expert_ids[static_cast<int64_t>(t) * k + tid] = e;
        offsets[static_cast<int64_t>(t) * k + tid] = atomicAdd(&counts[e], 1);
    }
}

__global__ void moe_offsets_kernel(int32_t* __restrict__ counts, int32_t* __restrict__ pad_off, int E) {
    pdl_enter();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        int acc = 0;
        pad_off[0] = 0;
        for (int e = 0; e < E; ++e) {
            acc += (counts[e] + kMoeRowPad - 1) / kMoeRowPad * kMoeRowPad;
            pad_off[e + 1] = acc;
            counts[e] = 0;
        }
    }
}

__global__ void moe_scatter_kernel(const void* __restrict__ x, int64_t ldx, void* __restrict__ xs, int64_t lds,
                                   const int32_t* __restrict__ expert_ids, const int32_t* __restrict__ offsets,
                                   const int32_t* __restrict__ pad_off, int32_t* __restrict__ mapped_slots, int H, int k) {
    pdl_enter();
    const int t = blockIdx.x, j = blockIdx.y;
    const int64_t c = static_cast<int64_t>(t) * k + j;
    const int slot = pad_off[expert_ids[c]] + offsets[c];
    if (threadIdx.x == 0) mapped_slots[c] = slot;
    const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(x) + static_cast<int64_t>(t) * ldx * 2);
    uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<char*>(xs) + static_cast<int64_t>(slot) * lds * 2);
    for (int i = threadIdx.x; i < (H >> 3); i += blockDim.x) dst[i] = src[i];
}

template <bool kBf16>
__global__ void moe_gather_kernel(const void* __restrict__ y, int64_t ldy, const int32_t* __restrict__ mapped_slots,
                                  const float* __restrict__ scores, float* __restrict__ out, int64_t ldo, int H, int k) {
    pdl_enter();
    const int t = blockIdx.x;
    for (int v = threadIdx.x; v < (H >> 3); v += blockDim.x) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = 0; j < k; ++j) {
            const int slot = mapped_slots[static_cast<int64_t>(t) * k + j];
            const float s = scores[static_cast<int64_t>(t) * k + j];
            const uint4 q = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(y) +
                                                            (static_cast<int64_t>(slot) * ldy + v * 8) * 2);
            acc[0] += s * unpack_lo<kBf16>(q.x); acc[1] += s * unpack_hi<kBf16>(q.x);
            acc[2] += s * unpack_lo<kBf16>(q.y); acc[3] += s * unpack_hi<kBf16>(q.y);
            acc[4] += s * unpack_lo<kBf16>(q.z); acc[5] += s * unpack_hi<kBf16>(q.z);
            acc[6] += s * unpack_lo<kBf16>(q.w); acc[7] += s * unpack_hi<kBf16>(q.w);
        }
        float4* o = reinterpret_cast<float4*>(out + static_cast<int64_t>(t) * ldo + v * 8);
        float4 a = o[0], b = o[1];
        a.x += acc[0]; a.y += acc[1]; a.z += acc[2]; a.w += acc[3];
        b.x += acc[4]; b.y += acc[5]; b.z += acc[6]; b.w += acc[7];
        o[0] = a;
        o[1] = b;
    }
}

}  // namespace hds

using namespace hds;

extern "C" int hds_moe_topk_gating(const void* logits, int64_t ld_logits, int T, int E, int k, int normalize,
                                   int logits_dtype, float* scores, int32_t* expert_ids, int32_t* offsets,
                                   int32_t* counts, void* stream) {
    HDS_REQUIRE(E >= 1 && E <= 4096, "hds_moe_topk_gating: 1..4096 experts");
    HDS_REQUIRE(k >= 1 && k <= kMoeMaxK && k <= E, "hds_moe_topk_gating: top_k must be in 1..8 and <= experts");
    HDS_REQUIRE(logits_dtype >= 0 && logits_dtype <= 2, "hds_moe_topk_gating: logits dtype 0 bf16 / 1 fp16 / 2 fp32");
    if (T == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t smem = static_cast<size_t>(E) * sizeof(float);
    cudaError_t rc;
    if (logits_dtype == 0)
        rc = launch_kernel(moe_topk_gating_kernel<0>, dim3(T), dim3(kGateThreads), smem, s, logits, ld_logits, E, k,
                           normalize, scores, expert_ids, offsets, counts);
    else if (logits_dtype == 1)
        rc = launch_kernel(moe_topk_gating_kernel<1>, dim3(T), dim3(kGateThreads), smem, s, logits, ld_logits, E, k,
                           normalize, scores, expert_ids, offsets, counts);
    else
        rc = launch_kernel(moe_topk_gating_kernel<2>, dim3(T), dim3(kGateThreads), smem, s, logits, ld_logits, E, k,
                           normalize, scores, expert_ids, offsets, counts);
    HDS_CHECK_CUDA(rc);
    return HDS_OK;
}

extern "C" int hds_moe_scatter(const void* x, int64_t ldx, void* xs, int64_t ld_xs, const int32_t* expert_ids,
                               const int32_t* offsets, int32_t* counts, int32_t* pad_off, int32_t* mapped_slots, int T,
                               int E, int k, int H, void* stream) {
    HDS_REQUIRE(H % 8 == 0 && ldx % 8 == 0 && ld_xs % 8 == 0, "hds_moe_scatter: H and strides must be multiples of 8");
    if (T == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    HDS_CHECK_CUDA(launch_kernel(moe_offsets_kernel, dim3(1), dim3(32), 0, s, counts, pad_off, E));
    HDS_CHECK_CUDA(launch_kernel(moe_scatter_kernel, dim3(T, k), dim3(H / 8 >= 256 ? 256 : 128), 0, s, x, ldx, xs, ld_xs,
                                 expert_ids, offsets, static_cast<const int32_t*>(pad_off), mapped_slots, H, k));
    return HDS_OK;
}

extern "C" int hds_moe_gather(const void* y, int64_t ldy, const int32_t* mapped_slots, const float* scores, float* out,
                              int64_t ldo, int T, int k, int H, int dtype, void* stream) {
    HDS_REQUIRE(H % 8 == 0 && ldy % 8 == 0 && ldo % 4 == 0, "hds_moe_gather: H / strides must be vector aligned");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_moe_gather: dtype must be 0 (bf16) or 1 (fp16)");
    if (T == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const dim3 block(H / 8 >= 256 ? 256 : 128);
    if (dtype == 0)
        HDS_CHECK_CUDA(launch_kernel(moe_gather_kernel<true>, dim3(T), block, 0, s, y, ldy, mapped_slots, scores, out, ldo, H, k));
    else
        HDS_CHECK_CUDA(launch_kernel(moe_gather_kernel<false>, dim3(T), block, 0, s, y, ldy, mapped_slots, scores, out, ldo, H, k));
    return HDS_OK;
}

extern "C" int64_t hds_moe_padded_rows(int T, int E, int k) {
    // every expert's range is rounded up to the 128-row GEMM tile
    return (static_cast<int64_t>(T) * k + static_cast<int64_t>(E) * (kMoeRowPad - 1)) / kMoeRowPad * kMoeRowPad + kMoeRowPad;
}
