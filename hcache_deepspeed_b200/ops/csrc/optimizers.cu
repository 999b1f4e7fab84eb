// Multi-tensor fused optimizers for the training runtime: Adam / AdamW and Lion over lists of tensors in one launch
// per <= 24 tensors (a (tensor, chunk) table rides in the kernel arguments, so there is no per-tensor launch and no
// device-side metadata to keep alive). Parameters and gradients may be fp32 / bf16 / fp16; moments are fp32.
// Role in the reference: csrc/adam/multi_tensor_adam.cu + multi_tensor_apply.cuh, csrc/lion/multi_tensor_lion.cu.
#include "hds_host.cuh"
#include "hds_ptx.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

constexpr int kMtaTensors = 24;
constexpr int kMtaBlocks = 480;
constexpr int kMtaChunk = 2048 * 16;
constexpr int kMtaThreads = 512;

struct MtaTable {
    void* p[kMtaTensors];
    const void* g[kMtaTensors];
    float* m[kMtaTensors];
    float* v[kMtaTensors];  // unused by Lion
    int64_t n[kMtaTensors];
    int32_t block_chunk[kMtaBlocks];
    uint8_t block_tensor[kMtaBlocks];
};

struct OptArgs {
    float lr, beta1, beta2, eps, weight_decay, bc1, bc2, grad_scale;
    int adamw;  // 1: decoupled weight decay, 0: L2 added to the gradient
};

template <int DT>
__device__ __forceinline__ float ld_elem(const void* base, int64_t i) { return load_any<DT>(base, i); }
template <int DT>
__device__ __forceinline__ void st_elem(void* base, int64_t i, float x) {
    if constexpr (DT == 2) reinterpret_cast<float*>(base)[i] = x;
    else store_scalar<DT == 0>(base, i, x);
}

template <int PDT, int GDT, bool kLion>
__global__ void __launch_bounds__(kMtaThreads) multi_tensor_opt_kernel(const MtaTable tab, const OptArgs a) {
    const int t = tab.block_tensor[blockIdx.x];
    const int64_t start = static_cast<int64_t>(tab.block_chunk[blockIdx.x]) * kMtaChunk;
    const int64_t n = tab.n[t];
    const int64_t end = start + kMtaChunk < n ? start + kMtaChunk : n;
    void* p = tab.p[t];
    const void* g = tab.g[t];
    float* m = tab.m[t];
    float* v = tab.v[t];
    for (int64_t i = start + threadIdx.x; i < end; i += kMtaThreads) {
        float w = ld_elem<PDT>(p, i);
        float grad = ld_elem<GDT>(g, i) * a.grad_scale;
        if constexpr (kLion) {
            // Lion: update = sign(beta1 * m + (1 - beta1) * g);  m = beta2 * m + (1 - beta2) * g
            const float mi = m[i];
            const float c = a.beta1 * mi + (1.0f - a.beta1) * grad;
            const float upd = c > 0.0f ? 1.0f : (c < 0.0f ? -1.0f : 0.0f);
            w = w * (1.0f - a.lr * a.weight_decay) - a.lr * upd;
            m[i] = a.beta2 * mi + (1.0f - a.beta2) * grad;
        } else {
            if (!a.adamw && a.weight_decay != 0.0f) grad += a.weight_decay * w;
            const float mi = a.beta1 * m[i] + (1.0f - a.beta1) * grad;
            const float vi = a.beta2 * v[i] + (1.0f - a.beta2) * grad * grad;
            m[i] = mi;
            v[i] = vi;
            const float denom = sqrtf(vi / a.bc2) + a.eps;
            float upd = (mi / a.bc1) / denom;
            if (a.adamw && a.weight_decay != 0.0f) upd += a.weight_decay * w;
            w -= a.lr * upd;
        }
        st_elem<PDT>(p, i, w);
    }
}

template <bool kLion>
static int launch_mta(const MtaTable& tab, const OptArgs& a, int blocks, int pdt, int gdt, cudaStream_t s) {
#define HDS_MTA(P, G)                                                                                         \
    if (pdt == P && gdt == G) {                                                                               \
        multi_tensor_opt_kernel<P, G, kLion><<<blocks, kMtaThreads, 0, s>>>(tab, a);                           \
        HDS_CHECK_CUDA(cudaGetLastError());                                                                   \
        return HDS_OK;                                                                                        \
    }
    HDS_MTA(2, 2) HDS_MTA(2, 0) HDS_MTA(2, 1) HDS_MTA(0, 0) HDS_MTA(1, 1) HDS_MTA(0, 2) HDS_MTA(1, 2)
#undef HDS_MTA
    set_error("multi-tensor optimizer: unsupported (param, grad) dtype pair");
    return HDS_ERR_UNSUPPORTED;
}

static int run_mta(bool lion, int n_tensors, void* const* params, const void* const* grads, float* const* exp_avg,
                   float* const* exp_avg_sq, const int64_t* numels, int param_dtype, int grad_dtype, const OptArgs& a,
                   cudaStream_t s) {
    HDS_REQUIRE(n_tensors >= 0, "multi-tensor optimizer: negative tensor count");
    MtaTable tab;
    memset(&tab, 0, sizeof(tab));
    int nt = 0, nb = 0;
    auto flush = [&]() -> int {
        if (nb == 0) return HDS_OK;
        int rc = lion ? launch_mta<true>(tab, a, nb, param_dtype, grad_dtype, s)
                      : launch_mta<false>(tab, a, nb, param_dtype, grad_dtype, s);
        nt = nb = 0;
        return rc;
    };
    for (int i = 0; i < n_tensors; ++i) {
        if (numels[i] <= 0) continue;
        const int64_t chunks = (numels[i] + kMtaChunk - 1) / kMtaChunk;
        int64_t c = 0;
        while (c < chunks) {
            if (nt == kMtaTensors || nb == kMtaBlocks) {
                int rc = flush();
                if (rc) return rc;
            }
            tab.p[nt] = params[i];
            tab.g[nt] = grads[i];
            tab.m[nt] = exp_avg[i];
            tab.v[nt] = exp_avg_sq ? exp_avg_sq[i] : nullptr;
            tab.n[nt] = numels[i];
            while (c < chunks && nb < kMtaBlocks) {
                tab.block_tensor[nb] = static_cast<uint8_t>(nt);
                tab.block_chunk[nb] = static_cast<int32_t>(c);
                ++nb;
                ++c;
            }
            ++nt;
        }
    }
    return flush();
}

}  // namespace hds

using namespace hds;

extern "C" int hds_multi_tensor_adam(int n_tensors, void* const* params, const void* const* grads,
                                     float* const* exp_avg, float* const* exp_avg_sq, const int64_t* numels,
                                     int param_dtype, int grad_dtype, float lr, float beta1, float beta2, float eps,
                                     float weight_decay, int step, int adamw_mode, int bias_correction,
                                     float grad_scale, void* stream) {
    HDS_REQUIRE(step >= 1, "hds_multi_tensor_adam: step counts from 1");
    OptArgs a;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
    a.bc1 = bias_correction ? 1.0f - powf(beta1, static_cast<float>(step)) : 1.0f;
    a.bc2 = bias_correction ? 1.0f - powf(beta2, static_cast<float>(step)) : 1.0f;
    a.grad_scale = grad_scale;
    a.adamw = adamw_mode;
    return run_mta(false, n_tensors, params, grads, exp_avg, exp_avg_sq, numels, param_dtype, grad_dtype, a,
                   static_cast<cudaStream_t>(stream));
}

extern "C" int hds_multi_tensor_lion(int n_tensors, void* const* params, const void* const* grads,
                                     float* const* exp_avg, const int64_t* numels, int param_dtype, int grad_dtype,
                                     float lr, float beta1, float beta2, float weight_decay, float grad_scale,
                                     void* stream) {
    OptArgs a;
    memset(&a, 0, sizeof(a));
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.weight_decay = weight_decay; a.grad_scale = grad_scale;
    a.bc1 = a.bc2 = 1.0f;
    return run_mta(true, n_tensors, params, grads, exp_avg, nullptr, numels, param_dtype, grad_dtype, a,
                   static_cast<cudaStream_t>(stream));
}
