// Bandwidth-bound sm_100a kernels of the ragged forward: RMSNorm / LayerNorm (with fused residual
// add, optional fp32 partial-sum input that is consumed-and-zeroed), RoPE + paged KV append,
// ragged embedding, last-token gather, arg-max sampling, KV block gather/scatter for KV offload.
//
// Reference counterparts (behaviour only; these are re-designed, not translated):
//   core_ops/cuda_rms_norm/rms_norm_cuda.cu:18-164, core_ops/cuda_layer_norm/layer_norm_cuda.cu:30-325,
//   ragged_ops/linear_blocked_kv_rotary/blocked_kv_rotary_cuda.cu:25-194, ragged_ops/embed/embed_cuda.cu:18-80,
//   ragged_ops/logits_gather/logits_gather_cuda.cu:18-50  (all under deepspeed/inference/v2/kernels/).
#include "hds_ptx.cuh"
#include "hds_host.cuh"
#include "tp_sync_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

constexpr int kNormThreads = 256;

__device__ __forceinline__ float block_sum_256(float v, float* red) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) red[w] = v;
    __syncthreads();
    float t = (l < (kNormThreads / 32)) ? red[l] : 0.0f;
    t = warp_sum(t);
    __syncthreads();
    return t;
}

// One CTA per row; the row lives in registers (VPT uint4 = 8*VPT elements per thread).
//   x_in      : [rows, cols] input (the residual stream when kHasDelta)
//   delta     : optional [rows, cols] added to x_in first; 16-bit or fp32 (kDeltaF32). fp32 deltas are the
//               split-K / tensor-parallel partial sums produced by hds_gemm_accum; with zero_delta they
//               are cleared after being read so the next accumulating GEMM starts from zero.
//   res_out   : optional, receives x_in + delta (pre-norm) or the normalised value (post_norm)
//   out       : normalised output;  row_idx: optional gather of input rows (last-token gather)
template <int VPT, bool kBf16, bool kLayerNorm, bool kDeltaF32>
__global__ void __launch_bounds__(kNormThreads)
norm_kernel(const void* __restrict__ x_in, void* delta, void* res_out, void* __restrict__ out,
            const void* __restrict__ gamma, const void* __restrict__ beta, const int32_t* __restrict__ row_idx,
            int cols, float eps, int zero_delta, int post_norm, int64_t ld_in, int64_t ld_out) {
    __shared__ float red[kNormThreads / 32];
    pdl_enter();
    const int r_out = blockIdx.x;
    const int r_in = row_idx ? row_idx[r_out] : r_out;
    const int nvec = cols >> 3;
    const char* xrow = reinterpret_cast<const char*>(x_in) + static_cast<int64_t>(r_in) * ld_in * 2;
    float v[VPT][8];
    float sum = 0.0f, sumsq = 0.0f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int vi = threadIdx.x + i * kNormThreads;
        if (vi < nvec) {
            uint4 q = *reinterpret_cast<const uint4*>(xrow + vi * 16);
            v[i][0] = unpack_lo<kBf16>(q.x); v[i][1] = unpack_hi<kBf16>(q.x);
            v[i][2] = unpack_lo<kBf16>(q.y); v[i][3] = unpack_hi<kBf16>(q.y);
            v[i][4] = unpack_lo<kBf16>(q.z); v[i][5] = unpack_hi<kBf16>(q.z);
            v[i][6] = unpack_lo<kBf16>(q.w); v[i][7] = unpack_hi<kBf16>(q.w);
            if (delta != nullptr) {
                if (kDeltaF32) {
                    float4* dp = reinterpret_cast<float4*>(reinterpret_cast<float*>(delta) +
                                                           static_cast<int64_t>(r_in) * cols + vi * 8);
                    float4 d0 = dp[0], d1 = dp[1];
                    v[i][0] += d0.x; v[i][1] += d0.y; v[i][2] += d0.z; v[i][3] += d0.w;
                    v[i][4] += d1.x; v[i][5] += d1.y; v[i][6] += d1.z; v[i][7] += d1.w;
                    if (zero_delta) {
                        dp[0] = make_float4(0.f, 0.f, 0.f, 0.f);
                        dp[1] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                } else {
                    uint4 d = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(delta) +
                                                             (static_cast<int64_t>(r_in) * cols + vi * 8) * 2);
                    v[i][0] += unpack_lo<kBf16>(d.x); v[i][1] += unpack_hi<kBf16>(d.x);
                    v[i][2] += unpack_lo<kBf16>(d.y); v[i][3] += unpack_hi<kBf16>(d.y);
                    v[i][4] += unpack_lo<kBf16>(d.z); v[i][5] += unpack_hi<kBf16>(d.z);
                    v[i][6] += unpack_lo<kBf16>(d.w); v[i][7] += unpack_hi<kBf16>(d.w);
                }
                // the residual stream is stored in the 16-bit dtype: round before normalising so the
                // statistics match what later layers read back
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    uint32_t p = pack2<kBf16>(v[i][j], v[i][j + 1]);
                    v[i][j] = unpack_lo<kBf16>(p);
                    v[i][j + 1] = unpack_hi<kBf16>(p);
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                sum += v[i][j];
                sumsq += v[i][j] * v[i][j];
            }
        }
    }
    if (res_out != nullptr && !post_norm) {
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int vi = threadIdx.x + i * kNormThreads;
            if (vi < nvec) {
                uint4 q;
                q.x = pack2<kBf16>(v[i][0], v[i][1]); q.y = pack2<kBf16>(v[i][2], v[i][3]);
                q.z = pack2<kBf16>(v[i][4], v[i][5]); q.w = pack2<kBf16>(v[i][6], v[i][7]);
                *reinterpret_cast<uint4*>(reinterpret_cast<char*>(res_out) +
                                          (static_cast<int64_t>(r_out) * ld_out + vi * 8) * 2) = q;
            }
        }
    }
    float mean = 0.0f, rstd;
    if (kLayerNorm) {
        sum = block_sum_256(sum, red);
        mean = sum / cols;
        float var = 0.0f;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int vi = threadIdx.x + i * kNormThreads;
            if (vi < nvec) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float d = v[i][j] - mean;
                    var += d * d;
                }
            }
        }
        var = block_sum_256(var, red);
        rstd = rsqrtf(var / cols + eps);
    } else {
        sumsq = block_sum_256(sumsq, red);
        rstd = rsqrtf(sumsq / cols + eps);
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int vi = threadIdx.x + i * kNormThreads;
        if (vi < nvec) {
            uint4 g = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(gamma) + vi * 16);
            float gg[8] = {unpack_lo<kBf16>(g.x), unpack_hi<kBf16>(g.x), unpack_lo<kBf16>(g.y), unpack_hi<kBf16>(g.y),
                           unpack_lo<kBf16>(g.z), unpack_hi<kBf16>(g.z), unpack_lo<kBf16>(g.w), unpack_hi<kBf16>(g.w)};
            float bb[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (kLayerNorm && beta != nullptr) {
                uint4 b = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(beta) + vi * 16);
                bb[0] = unpack_lo<kBf16>(b.x); bb[1] = unpack_hi<kBf16>(b.x);
                bb[2] = unpack_lo<kBf16>(b.y); bb[3] = unpack_hi<kBf16>(b.y);
                bb[4] = unpack_lo<kBf16>(b.z); bb[5] = unpack_hi<kBf16>(b.z);
                bb[6] = unpack_lo<kBf16>(b.w); bb[7] = unpack_hi<kBf16>(b.w);
            }
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * gg[j] + bb[j];
            uint4 q;
            q.x = pack2<kBf16>(o[0], o[1]); q.y = pack2<kBf16>(o[2], o[3]);
            q.z = pack2<kBf16>(o[4], o[5]); q.w = pack2<kBf16>(o[6], o[7]);
            const int64_t off = (static_cast<int64_t>(r_out) * ld_out + vi * 8) * 2;
            *reinterpret_cast<uint4*>(reinterpret_cast<char*>(out) + off) = q;
            if (res_out != nullptr && post_norm) *reinterpret_cast<uint4*>(reinterpret_cast<char*>(res_out) + off) = q;
        }
    }
}

template <bool kBf16, bool kLayerNorm, bool kDeltaF32>
static int launch_norm_vpt(int rows, int cols, const void* x, void* delta, void* res_out, void* out, const void* gamma,
                           const void* beta, const int32_t* row_idx, float eps, int zero_delta, int post_norm,
                           int64_t ld_in, int64_t ld_out, cudaStream_t s) {
    const int nvec = cols / 8;
    const int vpt = (nvec + kNormThreads - 1) / kNormThreads;
#define HDS_NORM_LAUNCH(V)                                                                                        \
    launch_kernel(norm_kernel<V, kBf16, kLayerNorm, kDeltaF32>, dim3(rows), dim3(kNormThreads), 0, s, x, delta,    \
                  res_out, out, gamma, beta, row_idx, cols, eps, zero_delta, post_norm, ld_in, ld_out)
    if (vpt <= 1) HDS_NORM_LAUNCH(1);
    else if (vpt <= 2) HDS_NORM_LAUNCH(2);
    else if (vpt <= 4) HDS_NORM_LAUNCH(4);
    else if (vpt <= 8) HDS_NORM_LAUNCH(8);
    else {
        set_error("norm: cols > 16384 unsupported");
        return HDS_ERR_UNSUPPORTED;
    }
#undef HDS_NORM_LAUNCH
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

// ---------------------------------------------------------------------------------------------
// Tensor-parallel norm: the collective that follows a row-parallel GEMM, fused with residual add + normalisation.
//
// Every rank's GEMM left its partial sums in its OWN buffer (symmetric memory) and released them with a flag
// (tp_signal_tail). This kernel acquires all ranks' flags and then runs in one of two shapes:
//   * all-reduce (scatter == 0, decode-size batches): every rank sums all ranks' partials for ALL rows — one pass of
//     `multimem.ld_reduce` (the switch adds the ranks' copies) or plain 16-byte peer loads in rank order — adds them to
//     its replica of the residual stream and normalises locally. One flag exchange per reduction, nothing else.
//   * reduce-scatter + all-gather (scatter == 1, larger batches): a rank reduces only the rows it owns
//     (n-times fewer NVLink bytes), keeps the residual stream of those rows, and writes the normalised rows into EVERY
//     rank's activation buffer (`multimem.st`, or one store per peer); the last CTA then signals, and the consuming
//     GEMM's TMA producer waits for all ranks' signals before its first activation load (all-gather -> GEMM).
// fp32 partial sums (accumulating decode GEMMs) live in two alternating buffers; the kernel clears the one NOT being
// read so the next accumulating GEMM starts from zero. 16-bit partials (store epilogues of the prefill GEMMs) need no
// clearing. Replaces NCCL all-reduce + residual add + norm (llama_v2/model.py:159-170 in the reference).
// ---------------------------------------------------------------------------------------------
struct TpReduce {
    const char* src[kTpMaxRanks];  // every rank's partial-sum buffer through this rank's mappings, rank order
    const char* src_mc;            // NVLS alias of that buffer (multimem.ld_reduce), or null
    char* out_peer[kTpMaxRanks];   // scatter: `out` on every rank (this rank's mappings)
    char* out_mc;                  // scatter: NVLS alias of `out` (multimem.st), or null
    float* zero_buf;               // local fp32 buffer to clear ([rows_total, cols]), may be null
    int n;                         // ranks
    int scatter;
    int row0, n_rows;              // rows [row0, row0 + n_rows) are normalised by this launch
    int rows_total;                // rows of the whole batch (extent of zero_buf)
    TpSync sync;
};

template <int THREADS>
__device__ __forceinline__ float block_sum_n(float v, float* red) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) red[w] = v;
    __syncthreads();
    float t = (l < (THREADS / 32)) ? red[l] : 0.0f;
    t = warp_sum(t);
    __syncthreads();
    return t;
}

template <bool kBf16>
__device__ __forceinline__ void add_h8(float (&v)[8], const uint4& d) {
    v[0] += unpack_lo<kBf16>(d.x); v[1] += unpack_hi<kBf16>(d.x);
    v[2] += unpack_lo<kBf16>(d.y); v[3] += unpack_hi<kBf16>(d.y);
    v[4] += unpack_lo<kBf16>(d.z); v[5] += unpack_hi<kBf16>(d.z);
    v[6] += unpack_lo<kBf16>(d.w); v[7] += unpack_hi<kBf16>(d.w);
}

// One CTA per row, THREADS x VPT 16-byte vectors per row (VPT is 1 up to hidden size 8192: with a remote round trip of
// a few microseconds per load the row must be fetched by ONE wave of loads, all issued before the first is consumed).
template <int THREADS, int VPT, bool kBf16, bool kLayerNorm, bool kPartial16>
__global__ void __launch_bounds__(THREADS)
norm_tp_kernel(const void* __restrict__ x_in, void* res_out, void* __restrict__ out, const void* __restrict__ gamma,
               const void* __restrict__ beta, int cols, float eps, int64_t ld_in, int64_t ld_out, const TpReduce tp) {
    __shared__ float red[32];
    pdl_enter();
    const int nvec = cols >> 3;
    const bool live = static_cast<int>(blockIdx.x) < tp.n_rows;
    const int r = tp.row0 + blockIdx.x;
    // the local operands first: their latency hides behind the flag wait
    uint4 xq[VPT];
    if (live) {
        const char* xrow = reinterpret_cast<const char*>(x_in) + static_cast<int64_t>(r) * ld_in * 2;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int vi = threadIdx.x + i * THREADS;
            if (vi < nvec) xq[i] = *reinterpret_cast<const uint4*>(xrow + vi * 16);
        }
    }
    if (tp.sync.wait) tp_wait_cta(tp.sync);
    if (tp.zero_buf != nullptr) {
        // all-reduce shape: CTA r clears row r; scatter shape: the CTAs of this launch share all rows_total rows
        const int step = tp.scatter ? static_cast<int>(gridDim.x) : tp.rows_total;
        for (int rr = tp.scatter ? static_cast<int>(blockIdx.x) : r; rr < tp.rows_total; rr += step) {
            float4* zp = reinterpret_cast<float4*>(tp.zero_buf + static_cast<int64_t>(rr) * cols);
            for (int vi = threadIdx.x; vi < 2 * nvec; vi += THREADS) zp[vi] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    float v[VPT][8];
    float sum = 0.0f, sumsq = 0.0f;
    if (live) {
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int vi = threadIdx.x + i * THREADS;
            if (vi < nvec) {
                v[i][0] = unpack_lo<kBf16>(xq[i].x); v[i][1] = unpack_hi<kBf16>(xq[i].x);
                v[i][2] = unpack_lo<kBf16>(xq[i].y); v[i][3] = unpack_hi<kBf16>(xq[i].y);
                v[i][4] = unpack_lo<kBf16>(xq[i].z); v[i][5] = unpack_hi<kBf16>(xq[i].z);
                v[i][6] = unpack_lo<kBf16>(xq[i].w); v[i][7] = unpack_hi<kBf16>(xq[i].w);
                const int64_t e = static_cast<int64_t>(r) * cols + vi * 8;  // element offset in a partial-sum buffer
                if constexpr (kPartial16) {
                    if (tp.src_mc != nullptr) {
                        const uint4 d = multimem_ld_reduce_v4_h2<kBf16>(tp.src_mc + e * 2);
                        add_h8<kBf16>(v[i], d);
                    } else {
                        uint4 d[kTpMaxRanks];
#pragma unroll
                        for (int s = 0; s < kTpMaxRanks; ++s)
                            if (s < tp.n) d[s] = ld_relaxed_sys_v4(tp.src[s] + e * 2);
#pragma unroll
                        for (int s = 0; s < kTpMaxRanks; ++s)
                            if (s < tp.n) add_h8<kBf16>(v[i], d[s]);
                    }
                } else {
                    if (tp.src_mc != nullptr) {
                        const float4 d0 = multimem_ld_reduce_v4_f32(tp.src_mc + e * 4);
                        const float4 d1 = multimem_ld_reduce_v4_f32(tp.src_mc + e * 4 + 16);
                        v[i][0] += d0.x; v[i][1] += d0.y; v[i][2] += d0.z; v[i][3] += d0.w;
                        v[i][4] += d1.x; v[i][5] += d1.y; v[i][6] += d1.z; v[i][7] += d1.w;
                    } else {
#pragma unroll
                        for (int s0 = 0; s0 < kTpMaxRanks; s0 += 4) {   // four ranks' loads in flight at a time
                            if (s0 >= tp.n) break;
                            uint4 a[4], b[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (s0 + j < tp.n) {
                                    a[j] = ld_relaxed_sys_v4(tp.src[s0 + j] + e * 4);
                                    b[j] = ld_relaxed_sys_v4(tp.src[s0 + j] + e * 4 + 16);
                                }
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (s0 + j < tp.n) {
                                    v[i][0] += __uint_as_float(a[j].x); v[i][1] += __uint_as_float(a[j].y);
                                    v[i][2] += __uint_as_float(a[j].z); v[i][3] += __uint_as_float(a[j].w);
                                    v[i][4] += __uint_as_float(b[j].x); v[i][5] += __uint_as_float(b[j].y);
                                    v[i][6] += __uint_as_float(b[j].z); v[i][7] += __uint_as_float(b[j].w);
                                }
                        }
                    }
                }
                // the residual stream is stored in the 16-bit dtype: round before normalising
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    const uint32_t p = pack2<kBf16>(v[i][j], v[i][j + 1]);
                    v[i][j] = unpack_lo<kBf16>(p);
                    v[i][j + 1] = unpack_hi<kBf16>(p);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    sum += v[i][j];
                    sumsq += v[i][j] * v[i][j];
                }
                if (res_out != nullptr) {
                    uint4 o;
                    o.x = pack2<kBf16>(v[i][0], v[i][1]); o.y = pack2<kBf16>(v[i][2], v[i][3]);
                    o.z = pack2<kBf16>(v[i][4], v[i][5]); o.w = pack2<kBf16>(v[i][6], v[i][7]);
                    *reinterpret_cast<uint4*>(reinterpret_cast<char*>(res_out) + (static_cast<int64_t>(r) * ld_out + vi * 8) * 2) = o;
                }
            }
        }
    }
    float mean = 0.0f, rstd;
    if (kLayerNorm) {
        sum = block_sum_n<THREADS>(sum, red);
        mean = sum / cols;
        float var = 0.0f;
        if (live) {
#pragma unroll
            for (int i = 0; i < VPT; ++i) {
                const int vi = threadIdx.x + i * THREADS;
                if (vi < nvec) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float d = v[i][j] - mean;
                        var += d * d;
                    }
                }
            }
        }
        var = block_sum_n<THREADS>(var, red);
        rstd = rsqrtf(var / cols + eps);
    } else {
        sumsq = block_sum_n<THREADS>(sumsq, red);
        rstd = rsqrtf(sumsq / cols + eps);
    }
    if (live) {
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int vi = threadIdx.x + i * THREADS;
            if (vi < nvec) {
                const uint4 g = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(gamma) + vi * 16);
                const float gg[8] = {unpack_lo<kBf16>(g.x), unpack_hi<kBf16>(g.x), unpack_lo<kBf16>(g.y), unpack_hi<kBf16>(g.y),
                                     unpack_lo<kBf16>(g.z), unpack_hi<kBf16>(g.z), unpack_lo<kBf16>(g.w), unpack_hi<kBf16>(g.w)};
                float bb[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                if (kLayerNorm && beta != nullptr) {
                    const uint4 b = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(beta) + vi * 16);
                    bb[0] = unpack_lo<kBf16>(b.x); bb[1] = unpack_hi<kBf16>(b.x);
                    bb[2] = unpack_lo<kBf16>(b.y); bb[3] = unpack_hi<kBf16>(b.y);
                    bb[4] = unpack_lo<kBf16>(b.z); bb[5] = unpack_hi<kBf16>(b.z);
                    bb[6] = unpack_lo<kBf16>(b.w); bb[7] = unpack_hi<kBf16>(b.w);
                }
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * gg[j] + bb[j];
                uint4 q;
                q.x = pack2<kBf16>(o[0], o[1]); q.y = pack2<kBf16>(o[2], o[3]);
                q.z = pack2<kBf16>(o[4], o[5]); q.w = pack2<kBf16>(o[6], o[7]);
                const int64_t off = (static_cast<int64_t>(r) * ld_out + vi * 8) * 2;
                if (!tp.scatter) {
                    *reinterpret_cast<uint4*>(reinterpret_cast<char*>(out) + off) = q;
                } else if (tp.out_mc != nullptr) {
                    multimem_st_v4(tp.out_mc + off, q);
                } else {
                    for (int s = 0; s < tp.n; ++s) st_v4(tp.out_peer[s] + off, q);
                }
            }
        }
    }
    if (tp.sync.signal) tp_signal_tail<true>(tp.sync, gridDim.x);
}

template <bool kBf16, bool kLayerNorm, bool kPartial16>
static int launch_norm_tp_vpt(int grid, int cols, const void* x, void* res_out, void* out, const void* gamma, const void* beta,
                              float eps, int64_t ld_in, int64_t ld_out, const TpReduce& tp, cudaStream_t s) {
    const int nvec = cols / 8;
#define HDS_NORM_TP_LAUNCH(TH, V)                                                                                     \
    launch_kernel(norm_tp_kernel<TH, V, kBf16, kLayerNorm, kPartial16>, dim3(grid), dim3(TH), 0, s, x, res_out, out, \
                  gamma, beta, cols, eps, ld_in, ld_out, tp)
    if (nvec <= 256) HDS_NORM_TP_LAUNCH(256, 1);
    else if (nvec <= 512) HDS_NORM_TP_LAUNCH(512, 1);
    else if (nvec <= 1024) HDS_NORM_TP_LAUNCH(1024, 1);
    else if (nvec <= 2048) HDS_NORM_TP_LAUNCH(1024, 2);
    else {
        set_error("norm_tp: cols > 16384 unsupported");
        return HDS_ERR_UNSUPPORTED;
    }
#undef HDS_NORM_TP_LAUNCH
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

// ---------------------------------------------------------------------------------------------
// RoPE + paged KV append (unfused path: arbitrary head_dim % 8 == 0, partial rotary dims)
// grid (tokens, ceil(vectors per token / 256)); each thread owns one 8-element vector of one head (full warps: with one
// 16-lane block per (token, head) a 4160-token prefill launched 400k blocks and ran 13x over its memory time).
// ---------------------------------------------------------------------------------------------
template <bool kBf16>
__global__ void rope_kv_append_kernel(const void* __restrict__ qkv, int64_t ld_qkv, void* __restrict__ q_out,
                                      void* __restrict__ kv_cache, const int32_t* __restrict__ tok_pos,
                                      const int32_t* __restrict__ tok_slot, const float* __restrict__ inv_freq,
                                      int n_q, int n_kv, int d, int rotary_dim, int block_size, int do_rope) {
    pdl_enter();
    const int t = blockIdx.x;
    const int vec_per_head = d >> 3;
    const int idx = blockIdx.y * blockDim.x + threadIdx.x;   // vector index over all heads of this token
    const int head = idx / vec_per_head;
    const int vi = idx - head * vec_per_head;
    if (head >= n_q + 2 * n_kv) return;
    const bool is_q = head < n_q;
    const bool is_k = !is_q && head < n_q + n_kv;
    const int slot = tok_slot[t];
    if (!is_q && slot < 0) return;
    const char* src = reinterpret_cast<const char*>(qkv) + (static_cast<int64_t>(t) * ld_qkv + head * d) * 2;
    uint4 own = *reinterpret_cast<const uint4*>(src + vi * 16);
    float x[8] = {unpack_lo<kBf16>(own.x), unpack_hi<kBf16>(own.x), unpack_lo<kBf16>(own.y), unpack_hi<kBf16>(own.y),
                  unpack_lo<kBf16>(own.z), unpack_hi<kBf16>(own.z), unpack_lo<kBf16>(own.w), unpack_hi<kBf16>(own.w)};
    const int e0 = vi * 8;
    if (do_rope && (is_q || is_k) && e0 < rotary_dim) {
        const int half = rotary_dim >> 1;
        const bool low = e0 < half;
        const int pe0 = low ? e0 + half : e0 - half;
        uint4 pr = *reinterpret_cast<const uint4*>(src + pe0 * 2);
        float y[8] = {unpack_lo<kBf16>(pr.x), unpack_hi<kBf16>(pr.x), unpack_lo<kBf16>(pr.y), unpack_hi<kBf16>(pr.y),
                      unpack_lo<kBf16>(pr.z), unpack_hi<kBf16>(pr.z), unpack_lo<kBf16>(pr.w), unpack_hi<kBf16>(pr.w)};
        const float pos = static_cast<float>(tok_pos[t]);
        const int f0 = low ? e0 : e0 - half;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float s, c;
            rope_sincos(pos * inv_freq[f0 + j], &s, &c);
            x[j] = low ? (x[j] * c - y[j] * s) : (x[j] * c + y[j] * s);
        }
    }
    uint4 o;
    o.x = pack2<kBf16>(x[0], x[1]); o.y = pack2<kBf16>(x[2], x[3]);
    o.z = pack2<kBf16>(x[4], x[5]); o.w = pack2<kBf16>(x[6], x[7]);
    if (is_q) {
        *reinterpret_cast<uint4*>(reinterpret_cast<char*>(q_out) +
                                  ((static_cast<int64_t>(t) * n_q + head) * d + e0) * 2) = o;
    } else {
        const int kvh = is_k ? head - n_q : head - n_q - n_kv;
        const int64_t blk = slot / block_size, off = slot % block_size;
        const int64_t e = (((blk * 2 + (is_k ? 0 : 1)) * n_kv + kvh) * block_size + off) * d + e0;
        *reinterpret_cast<uint4*>(reinterpret_cast<char*>(kv_cache) + e * 2) = o;
    }
}

// ---------------------------------------------------------------------------------------------
// ragged embedding: out[t] = W[ids[t]] (+ P[tok_pos[t] + pos_offset]); ids < 0 -> zeros
// ---------------------------------------------------------------------------------------------
template <bool kBf16>
__global__ void embed_kernel(const int32_t* __restrict__ ids, const void* __restrict__ W, const void* __restrict__ P,
                             const int32_t* __restrict__ tok_pos, int pos_offset, int max_pos, void* __restrict__ out,
                             int H, int vocab) {
    pdl_enter();
    const int t = blockIdx.x;
    const int id = ids[t];
    const int nvec = H >> 3;
    const bool valid = id >= 0 && id < vocab;
    const char* wrow = reinterpret_cast<const char*>(W) + static_cast<int64_t>(valid ? id : 0) * H * 2;
    const char* prow = nullptr;
    if (P != nullptr) {
        int p = tok_pos[t] + pos_offset;
        p = p < 0 ? 0 : (p >= max_pos ? max_pos - 1 : p);
        prow = reinterpret_cast<const char*>(P) + static_cast<int64_t>(p) * H * 2;
    }
    for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
        uint4 q = valid ? ld_nc_v4(wrow + vi * 16) : make_uint4(0, 0, 0, 0);
        if (prow != nullptr) {
            uint4 p = ld_nc_v4(prow + vi * 16);
            q.x = pack2<kBf16>(unpack_lo<kBf16>(q.x) + unpack_lo<kBf16>(p.x), unpack_hi<kBf16>(q.x) + unpack_hi<kBf16>(p.x));
            q.y = pack2<kBf16>(unpack_lo<kBf16>(q.y) + unpack_lo<kBf16>(p.y), unpack_hi<kBf16>(q.y) + unpack_hi<kBf16>(p.y));
            q.z = pack2<kBf16>(unpack_lo<kBf16>(q.z) + unpack_lo<kBf16>(p.z), unpack_hi<kBf16>(q.z) + unpack_hi<kBf16>(p.z));
            q.w = pack2<kBf16>(unpack_lo<kBf16>(q.w) + unpack_lo<kBf16>(p.w), unpack_hi<kBf16>(q.w) + unpack_hi<kBf16>(p.w));
        }
        *reinterpret_cast<uint4*>(reinterpret_cast<char*>(out) + (static_cast<int64_t>(t) * H + vi * 8) * 2) = q;
    }
}

// rows gather: out[i] = x[row_idx[i]]
__global__ void gather_rows_kernel(const void* __restrict__ x, const int32_t* __restrict__ row_idx,
                                   void* __restrict__ out, int H) {
    pdl_enter();
    const int i = blockIdx.x;
    const int r = row_idx[i];
    const int nvec = H >> 3;
    for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
        uint4 q = ld_nc_v4(reinterpret_cast<const char*>(x) + (static_cast<int64_t>(r) * H + vi * 8) * 2);
        *reinterpret_cast<uint4*>(reinterpret_cast<char*>(out) + (static_cast<int64_t>(i) * H + vi * 8) * 2) = q;
    }
}

// ---------------------------------------------------------------------------------------------
// greedy sampling: arg-max over the vocabulary, one CTA per sequence (ties -> lowest index)
// ---------------------------------------------------------------------------------------------
template <int DT>
__global__ void __launch_bounds__(1024)
argmax_kernel(const void* __restrict__ logits, int64_t ld, int V, int32_t* __restrict__ out_ids,
              float* __restrict__ out_vals) {
    __shared__ float sv[32];
    __shared__ int si[32];
    pdl_enter();
    const int s = blockIdx.x;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    bool vec = false;
    if constexpr (DT != 2) {
        // 16-bit logits: 8 per 16-byte load (a 32000-wide row is 4 loads per thread instead of 32 dependent scalar ones)
        const char* row = reinterpret_cast<const char*>(logits) + static_cast<int64_t>(s) * ld * 2;
        vec = (V & 7) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0;
        if (vec) {
            for (int i = threadIdx.x * 8; i < V; i += blockDim.x * 8) {
                const uint4 p = *reinterpret_cast<const uint4*>(row + i * 2);
                const uint32_t w[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float lo = unpack_lo<DT == 0>(w[j]), hi = unpack_hi<DT == 0>(w[j]);
                    if (lo > best || (lo == best && i + 2 * j < bi)) {
                        best = lo;
                        bi = i + 2 * j;
                    }
                    if (hi > best || (hi == best && i + 2 * j + 1 < bi)) {
                        best = hi;
                        bi = i + 2 * j + 1;
                    }
                }
            }
        }
    }
    if (!vec) {
        for (int i = threadIdx.x; i < V; i += blockDim.x) {
            const float v = load_any<DT>(logits, static_cast<int64_t>(s) * ld + i);
            if (v > best || (v == best && i < bi)) {
                best = v;
                bi = i;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) {
            best = ov;
            bi = oi;
        }
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) {
        sv[w] = best;
        si[w] = bi;
    }
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        best = l < nw ? sv[l] : -INFINITY;
        bi = l < nw ? si[l] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) {
                best = ov;
                bi = oi;
            }
        }
        if (l == 0) {
            out_ids[s] = bi;
            if (out_vals) out_vals[s] = best;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// KV block gather / scatter between the paged cache (all layers) and a dense staging buffer:
// the device half of plain KV offload. One CTA copies one (layer, block) pair of K+V.
//   cache  : [layers][blocks][2][n_kv][bs][d]      staging: [n_sel][layers][2][n_kv][bs][d]
// ---------------------------------------------------------------------------------------------
__global__ void kv_blocks_copy_kernel(char* __restrict__ cache, char* __restrict__ staging,
                                      const int32_t* __restrict__ block_ids, int64_t block_bytes,
                                      int64_t layer_stride_bytes, int n_layers, int to_staging) {
    const int sel = blockIdx.x;
    const int layer = blockIdx.y;
    const int64_t blk = block_ids[sel];
    char* c = cache + layer * layer_stride_bytes + blk * block_bytes;
    char* s = staging + (static_cast<int64_t>(sel) * n_layers + layer) * block_bytes;
    const int64_t nvec = block_bytes >> 4;
    for (int64_t i = threadIdx.x; i < nvec; i += blockDim.x) {
        if (to_staging)
            reinterpret_cast<uint4*>(s)[i] = reinterpret_cast<const uint4*>(c)[i];
        else
            reinterpret_cast<uint4*>(c)[i] = reinterpret_cast<const uint4*>(s)[i];
    }
}

// ---------------------------------------------------------------------------------------------
// P2P row gather for the tensor-parallel restore: dst[r] = peer_of(r)[r] for every row not owned by this rank.
// One launch pulls all peers' slices over NVLink (16-byte loads from peer-mapped symmetric memory).
// ---------------------------------------------------------------------------------------------
struct PeerPtrs {
    const char* p[8];
};
__global__ void p2p_gather_rows_kernel(char* __restrict__ dst, const PeerPtrs peers, int rows_per_peer, int n_rows,
                                       int64_t row_bytes, int my_rank, const TpSync sync) {
    if (sync.wait) tp_wait_cta(sync);  // every rank has published "my rows of this slot have landed"
    const int r = blockIdx.x;
    const int owner = r / rows_per_peer;
    if (r >= n_rows || owner == my_rank) return;
    const char* src = peers.p[owner] + static_cast<int64_t>(r) * row_bytes;
    uint4* d = reinterpret_cast<uint4*>(dst + static_cast<int64_t>(r) * row_bytes);
    const int nvec = static_cast<int>(row_bytes >> 4);
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) d[i] = ld_relaxed_sys_v4(src + i * 16);
}

// fp32 -> 16-bit convert (used to materialise accumulated fp32 partial sums), optional zeroing
template <bool kBf16>
__global__ void cvt_f32_kernel(float* __restrict__ src, void* __restrict__ dst, int64_t n, int zero_src) {
    pdl_enter();
    const int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (i + 3 < n) {
        float4 v = *reinterpret_cast<float4*>(src + i);
        uint2 o;
        o.x = pack2<kBf16>(v.x, v.y);
        o.y = pack2<kBf16>(v.z, v.w);
        *reinterpret_cast<uint2*>(reinterpret_cast<char*>(dst) + i * 2) = o;
        if (zero_src) *reinterpret_cast<float4*>(src + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
        for (int64_t j = i; j < n; ++j) {
            store_scalar<kBf16>(dst, j, src[j]);
            if (zero_src) src[j] = 0.f;
        }
    }
}

}  // namespace hds

using namespace hds;

extern "C" int hds_norm(const void* x, int64_t ld_in, void* delta, int delta_is_f32, int zero_delta, void* res_out,
                        void* out, int64_t ld_out, const void* gamma, const void* beta, const int32_t* row_idx,
                        int rows, int cols, float eps, int layer_norm, int post_norm, int dtype, void* stream) {
    HDS_REQUIRE(rows >= 0 && cols > 0 && cols % 8 == 0, "hds_norm: cols must be a positive multiple of 8");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_norm: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(delta == nullptr || row_idx == nullptr, "hds_norm: delta and row gather are mutually exclusive");
    if (rows == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool f32 = delta != nullptr && delta_is_f32;
#define HDS_NORM_DISPATCH(BF, LN, F32)                                                                            \
    return launch_norm_vpt<BF, LN, F32>(rows, cols, x, delta, res_out, out, gamma, beta, row_idx, eps, zero_delta, \
                                        post_norm, ld_in, ld_out, s)
    if (dtype == 0) {
        if (layer_norm) { if (f32) HDS_NORM_DISPATCH(true, true, true); else HDS_NORM_DISPATCH(true, true, false); }
        else            { if (f32) HDS_NORM_DISPATCH(true, false, true); else HDS_NORM_DISPATCH(true, false, false); }
    } else {
        if (layer_norm) { if (f32) HDS_NORM_DISPATCH(false, true, true); else HDS_NORM_DISPATCH(false, true, false); }
        else            { if (f32) HDS_NORM_DISPATCH(false, false, true); else HDS_NORM_DISPATCH(false, false, false); }
    }
#undef HDS_NORM_DISPATCH
}

extern "C" int hds_norm_tp(int ctx, int wait, int signal, const void* x, int64_t ld_in, const void* const* src_ptrs,
                           const void* src_mc, int partial_16bit, float* zero_buf, void* res_out, void* out, int64_t ld_out,
                           void* const* out_ptrs, void* out_mc, int row0, int n_rows, int rows_total, const void* gamma,
                           const void* beta, int cols, float eps, int layer_norm, int dtype, void* stream) {
    HDS_REQUIRE(cols > 0 && cols % 8 == 0, "hds_norm_tp: cols must be a positive multiple of 8");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_norm_tp: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(row0 >= 0 && n_rows >= 0 && row0 + n_rows <= rows_total, "hds_norm_tp: row range outside the batch");
    TpReduce tp;
    memset(&tp, 0, sizeof(tp));
    if (!tp_get_context(ctx, &tp.sync)) return HDS_ERR_BAD_ARG;
    tp.sync.wait = wait ? 1 : 0;
    tp.sync.signal = signal ? 1 : 0;
    tp.n = tp.sync.n_ranks;
    for (int i = 0; i < tp.n; ++i) tp.src[i] = reinterpret_cast<const char*>(src_ptrs[i]);
    tp.src_mc = reinterpret_cast<const char*>(src_mc);
    tp.scatter = out_ptrs != nullptr || out_mc != nullptr;
    if (out_ptrs != nullptr)
        for (int i = 0; i < tp.n; ++i) tp.out_peer[i] = reinterpret_cast<char*>(out_ptrs[i]);
    tp.out_mc = reinterpret_cast<char*>(out_mc);
    tp.zero_buf = zero_buf;
    tp.row0 = row0;
    tp.n_rows = n_rows;
    tp.rows_total = rows_total;
    // a rank that owns no row of a tiny batch still takes part in the flag exchange (and clears its buffer)
    const int grid = n_rows > 0 ? n_rows : 1;
    if (n_rows == 0 && !wait && !signal && zero_buf == nullptr) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HDS_NORM_TP_DISPATCH(BF, LN, P16) \
    return launch_norm_tp_vpt<BF, LN, P16>(grid, cols, x, res_out, out, gamma, beta, eps, ld_in, ld_out, tp, s)
    if (dtype == 0) {
        if (layer_norm) { if (partial_16bit) HDS_NORM_TP_DISPATCH(true, true, true); else HDS_NORM_TP_DISPATCH(true, true, false); }
        else            { if (partial_16bit) HDS_NORM_TP_DISPATCH(true, false, true); else HDS_NORM_TP_DISPATCH(true, false, false); }
    } else {
        if (layer_norm) { if (partial_16bit) HDS_NORM_TP_DISPATCH(false, true, true); else HDS_NORM_TP_DISPATCH(false, true, false); }
        else            { if (partial_16bit) HDS_NORM_TP_DISPATCH(false, false, true); else HDS_NORM_TP_DISPATCH(false, false, false); }
    }
#undef HDS_NORM_TP_DISPATCH
}

extern "C" int hds_rope_kv_append(const void* qkv, int64_t ld_qkv, void* q_out, void* kv_cache_layer,
                                  const int32_t* tok_pos, const int32_t* tok_slot, const float* inv_freq, int T,
                                  int n_q_heads, int n_kv_heads, int head_dim, int rotary_dim, int block_size,
                                  int do_rope, int dtype, void* stream) {
    HDS_REQUIRE(head_dim % 8 == 0 && head_dim <= 512, "hds_rope_kv_append: head_dim must be a multiple of 8 (<=512)");
    HDS_REQUIRE(!do_rope || (rotary_dim % 16 == 0 && rotary_dim <= head_dim && inv_freq != nullptr),
                "hds_rope_kv_append: rotary_dim must be a multiple of 16 and <= head_dim");
    HDS_REQUIRE(ld_qkv % 8 == 0, "hds_rope_kv_append: ld_qkv must be a multiple of 8");
    if (T == 0) return HDS_OK;
    const int vectors = (n_q_heads + 2 * n_kv_heads) * (head_dim / 8);
    const int threads = vectors < 256 ? ((vectors + 31) / 32) * 32 : 256;
    dim3 grid(T, (vectors + threads - 1) / threads);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype == 0)
        launch_kernel(rope_kv_append_kernel<true>, grid, dim3(threads), 0, s, qkv, ld_qkv, q_out, kv_cache_layer, tok_pos, tok_slot,
                                                             inv_freq, n_q_heads, n_kv_heads, head_dim, rotary_dim,
                                                             block_size, do_rope);
    else
        launch_kernel(rope_kv_append_kernel<false>, grid, dim3(threads), 0, s, qkv, ld_qkv, q_out, kv_cache_layer, tok_pos, tok_slot,
                                                              inv_freq, n_q_heads, n_kv_heads, head_dim, rotary_dim,
                                                              block_size, do_rope);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_embed(const int32_t* ids, const void* W, const void* P, const int32_t* tok_pos, int pos_offset,
                         int max_pos, void* out, int T, int H, int vocab, int dtype, void* stream) {
    HDS_REQUIRE(H % 8 == 0, "hds_embed: hidden size must be a multiple of 8");
    if (T == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int threads = H / 8 >= 256 ? 256 : 128;
    if (dtype == 0)
        launch_kernel(embed_kernel<true>, dim3(T), dim3(threads), 0, s, ids, W, P, tok_pos, pos_offset, max_pos, out, H, vocab);
    else
        launch_kernel(embed_kernel<false>, dim3(T), dim3(threads), 0, s, ids, W, P, tok_pos, pos_offset, max_pos, out, H, vocab);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_gather_rows(const void* x, const int32_t* row_idx, void* out, int n_rows, int H, void* stream) {
    HDS_REQUIRE(H % 8 == 0, "hds_gather_rows: hidden size must be a multiple of 8");
    if (n_rows == 0) return HDS_OK;
    launch_kernel(gather_rows_kernel, dim3(n_rows), dim3(256), 0, static_cast<cudaStream_t>(stream), x, row_idx, out, H);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_argmax(const void* logits, int64_t ld, int n_rows, int V, int32_t* out_ids, float* out_vals,
                          int dtype, void* stream) {
    HDS_REQUIRE(dtype >= 0 && dtype <= 2, "hds_argmax: dtype 0 (bf16) / 1 (fp16) / 2 (fp32)");
    if (n_rows == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype == 0) launch_kernel(argmax_kernel<0>, dim3(n_rows), dim3(1024), 0, s, logits, ld, V, out_ids, out_vals);
    else if (dtype == 1) launch_kernel(argmax_kernel<1>, dim3(n_rows), dim3(1024), 0, s, logits, ld, V, out_ids, out_vals);
    else launch_kernel(argmax_kernel<2>, dim3(n_rows), dim3(1024), 0, s, logits, ld, V, out_ids, out_vals);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_kv_blocks_copy(void* cache, void* staging, const int32_t* block_ids, int n_sel, int n_layers,
                                  int64_t block_bytes, int64_t layer_stride_bytes, int to_staging, void* stream) {
    HDS_REQUIRE(block_bytes % 16 == 0, "hds_kv_blocks_copy: block bytes must be a multiple of 16");
    if (n_sel == 0) return HDS_OK;
    dim3 grid(n_sel, n_layers);
    kv_blocks_copy_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<char*>(cache), reinterpret_cast<char*>(staging), block_ids, block_bytes, layer_stride_bytes,
        n_layers, to_staging);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_p2p_gather_rows(void* dst, const void* const* peer_ptrs, int n_peers, int rows_per_peer, int n_rows,
                                   int64_t row_bytes, int my_rank, void* stream) {
    HDS_REQUIRE(n_peers >= 1 && n_peers <= 8, "hds_p2p_gather_rows: 1..8 peers");
    HDS_REQUIRE(row_bytes % 16 == 0 && rows_per_peer > 0, "hds_p2p_gather_rows: rows must be 16-byte multiples");
    if (n_rows == 0) return HDS_OK;
    PeerPtrs pp;
    memset(&pp, 0, sizeof(pp));
    for (int i = 0; i < n_peers; ++i) pp.p[i] = reinterpret_cast<const char*>(peer_ptrs[i]);
    TpSync sync = tp_take_armed();
    sync.signal = 0;
    p2p_gather_rows_kernel<<<n_rows, 256, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<char*>(dst), pp,
                                                                                  rows_per_peer, n_rows, row_bytes, my_rank, sync);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

extern "C" int hds_cvt_f32(float* src, void* dst, int64_t n, int zero_src, int dtype, void* stream) {
    if (n == 0) return HDS_OK;
    const int threads = 256;
    const int64_t blocks = ((n + 3) / 4 + threads - 1) / threads;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype == 0) launch_kernel(cvt_f32_kernel<true>, dim3(static_cast<unsigned>(blocks)), dim3(threads), 0, s, src, dst, n, zero_src);
    else launch_kernel(cvt_f32_kernel<false>, dim3(static_cast<unsigned>(blocks)), dim3(threads), 0, s, src, dst, n, zero_src);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}
