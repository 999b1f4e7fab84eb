// Host-side optimizers for ZeRO-Offload: Adam / AdamW, Adagrad and Lion over fp32 master parameters that live in
// (pinned) host memory, with an optional fused write-back of the updated values as bf16 / fp16 into a second
// buffer (the copy that is sent back to the GPU). Work is split over a small thread pool; the inner loops are
// written so the compiler vectorises them and are cloned for AVX-512 / AVX2 / baseline at load time.
// Role in the reference: csrc/adam/cpu_adam_impl.cpp, csrc/adagrad/cpu_adagrad.cpp, csrc/lion/cpu_lion_impl.cpp
// (+ csrc/includes/simd.h).
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <algorithm>
#include <thread>
#include <vector>
#include "../../../include/hds_host.h"

#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
#define HDS_CLONES __attribute__((target_clones("avx512f", "avx2", "default")))
#else
#define HDS_CLONES
#endif

namespace {

inline uint16_t f32_to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);  // quiet NaN
    u += 0x7fffu + ((u >> 16) & 1u);
    return static_cast<uint16_t>(u >> 16);
}

inline uint16_t f32_to_fp16(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    const int32_t exp = static_cast<int32_t>((x >> 23) & 0xff) - 127 + 15;
    uint32_t mant = x & 0x7fffffu;
    if (((x >> 23) & 0xff) == 0xff) return static_cast<uint16_t>(sign | 0x7c00u | (mant ? 0x200u : 0));
    if (exp >= 31) return static_cast<uint16_t>(sign | 0x7c00u);
    if (exp <= 0) {
        if (exp < -10) return static_cast<uint16_t>(sign);
        mant |= 0x800000u;
        const int shift = 14 - exp;
        uint32_t h = mant >> shift;
        const uint32_t rem = mant & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1))) ++h;
        return static_cast<uint16_t>(sign | h);
    }
    uint32_t h = (static_cast<uint32_t>(exp) << 10) | (mant >> 13);
    const uint32_t rem = mant & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) ++h;
    return static_cast<uint16_t>(sign | h);
}

inline void write_back(void* dst, int dst_dtype, int64_t lo, int64_t hi, const float* p) {
    if (dst == nullptr) return;
    uint16_t* d = static_cast<uint16_t*>(dst);
    if (dst_dtype == 0) for (int64_t i = lo; i < hi; ++i) d[i] = f32_to_bf16(p[i]);
    else if (dst_dtype == 1) for (int64_t i = lo; i < hi; ++i) d[i] = f32_to_fp16(p[i]);
    else memcpy(static_cast<float*>(dst) + lo, p + lo, static_cast<size_t>(hi - lo) * 4);
}

HDS_CLONES void adam_range(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                           float* __restrict__ v, int64_t lo, int64_t hi, float lr, float b1, float b2, float eps,
                           float wd, float bc1, float bc2_sqrt, int adamw) {
    const float step_size = lr / bc1;
    for (int64_t i = lo; i < hi; ++i) {
        float grad = g[i];
        float w = p[i];
        if (!adamw && wd != 0.0f) grad += wd * w;
        const float mi = b1 * m[i] + (1.0f - b1) * grad;
        const float vi = b2 * v[i] + (1.0f - b2) * grad * grad;
        m[i] = mi;
        v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        if (adamw && wd != 0.0f) w -= lr * wd * w;
        p[i] = w - step_size * (mi / denom);
    }
}

HDS_CLONES void adagrad_range(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ s, int64_t lo,
                              int64_t hi, float lr, float eps, float wd) {
    for (int64_t i = lo; i < hi; ++i) {
        float grad = g[i];
        if (wd != 0.0f) grad += wd * p[i];
        const float si = s[i] + grad * grad;
        s[i] = si;
        p[i] -= lr * grad / (sqrtf(si) + eps);
    }
}

HDS_CLONES void lion_range(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, int64_t lo,
                           int64_t hi, float lr, float b1, float b2, float wd) {
    for (int64_t i = lo; i < hi; ++i) {
        const float grad = g[i], mi = m[i];
        const float c = b1 * mi + (1.0f - b1) * grad;
        const float upd = c > 0.0f ? 1.0f : (c < 0.0f ? -1.0f : 0.0f);
        p[i] = p[i] * (1.0f - lr * wd) - lr * upd;
        m[i] = b2 * mi + (1.0f - b2) * grad;
    }
}

template <typename F>
void parallel_ranges(int64_t n, int threads, F fn) {
    const int64_t min_per_thread = 1 << 16;
    int t = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    t = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(t, (n + min_per_thread - 1) / min_per_thread)));
    if (t == 1) {
        fn(0, n);
        return;
    }
    const int64_t per = ((n + t - 1) / t + 15) & ~int64_t(15);
    std::vector<std::thread> pool;
    for (int i = 0; i < t; ++i) {
        const int64_t lo = i * per, hi = std::min<int64_t>(n, lo + per);
        if (lo >= hi) break;
        pool.emplace_back([=] { fn(lo, hi); });
    }
    for (auto& th : pool) th.join();
}

}  // namespace

extern "C" int hds_cpu_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                 float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                                 int adamw_mode, int bias_correction, void* copy_out, int copy_dtype, int threads) {
    if (n < 0 || step < 1 || !params || !grads || !exp_avg || !exp_avg_sq) return -1;
    const float bc1 = bias_correction ? 1.0f - powf(beta1, static_cast<float>(step)) : 1.0f;
    const float bc2 = bias_correction ? sqrtf(1.0f - powf(beta2, static_cast<float>(step))) : 1.0f;
    parallel_ranges(n, threads, [=](int64_t lo, int64_t hi) {
        adam_range(params, grads, exp_avg, exp_avg_sq, lo, hi, lr, beta1, beta2, eps, weight_decay, bc1, bc2, adamw_mode);
        write_back(copy_out, copy_dtype, lo, hi, params);
    });
    return 0;
}

extern "C" int hds_cpu_adagrad_step(float* params, const float* grads, float* exp_avg_sq, int64_t n, float lr,
                                    float eps, float weight_decay, void* copy_out, int copy_dtype, int threads) {
    if (n < 0 || !params || !grads || !exp_avg_sq) return -1;
    parallel_ranges(n, threads, [=](int64_t lo, int64_t hi) {
        adagrad_range(params, grads, exp_avg_sq, lo, hi, lr, eps, weight_decay);
        write_back(copy_out, copy_dtype, lo, hi, params);
    });
    return 0;
}

extern "C" int hds_cpu_lion_step(float* params, const float* grads, float* exp_avg, int64_t n, float lr, float beta1,
                                 float beta2, float weight_decay, void* copy_out, int copy_dtype, int threads) {
    if (n < 0 || !params || !grads || !exp_avg) return -1;
    parallel_ranges(n, threads, [=](int64_t lo, int64_t hi) {
        lion_range(params, grads, exp_avg, lo, hi, lr, beta1, beta2, weight_decay);
        write_back(copy_out, copy_dtype, lo, hi, params);
    });
    return 0;
}
