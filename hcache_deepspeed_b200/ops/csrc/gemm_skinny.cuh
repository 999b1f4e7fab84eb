// Weight-streaming tcgen05 GEMM for decode-size M (<= 64 tokens):  D[M,N] = X[M,K] * W[N,K]^T
//
// The operands are swapped relative to gemm_core.cuh: the 128 rows of the MMA are 128 output FEATURES (a
// 128 x 64 weight tile, K-major, straight from the [N, K] checkpoint layout) and the MMA's N side is the token
// tile (NT = 16 / 32 / 64 tokens). A 128 x NT x 16 tcgen05.mma costs NT/128 of the tensor-pipe time of the
// token-major 128 x 128 x 16 instruction for the same 4 KB of weights, so the tensor pipe stops being the limit
// of a skinny GEMM and the kernel is a pure HBM stream. To keep enough bytes in flight without a persistent
// schedule each CTA owns ONE (feature tile, K split) unit and 2-3 CTAs are resident per SM (4-5 stages of
// 16 KB weights each); prologue / epilogue of one CTA overlap the streaming of its neighbours.
//
//   warp 0   : TMA producer         warp 1 : TMEM alloc + MMA issuer        warps 2-5 : epilogue
//   accumulator: TMEM lanes = features, columns = tokens  ->  one epilogue thread owns one output feature
//
// Epilogues: store (+bias +act), gated (interleaved gate/up rows of W -> act(g)*u; the partner feature lives
// in the neighbouring lane), fp32 accumulate into the pending residual delta (split-K needs no tickets there).
// Split-K for store/gated goes through a feature-major fp32 scratch [N][NT] with one ticket per (tile, warp).
//
// Role in the reference: the cuBLAS GEMMs of a decode step (core_ops/blas_kernels/blas.h:31-136) plus the
// bias / gated activation kernels that follow them.
#pragma once
#include "gemm_core.cuh"

namespace hds {

constexpr int kSkinnyBF = 128;  // features per CTA (MMA M)

__host__ __device__ constexpr int skinny_stages(int nt) { return nt == 32 ? 5 : 4; }
__host__ __device__ constexpr int skinny_occupancy(int nt) { return nt == 16 ? 3 : 2; }
__host__ __device__ constexpr size_t skinny_smem_bytes(int nt) {
    return static_cast<size_t>(skinny_stages(nt)) * (kSkinnyBF * kGemmBK * 2 + nt * kGemmBK * 2) + 256 + 1024;
}

struct SkinnyArgs {
    int M, N, K;
    int splits, kb_per_split;
    int act;
    void* out;          // EPI_STORE / EPI_GATED
    int64_t ldo;
    const void* bias;   // [N] I/O dtype, may be null
    float* accum;       // EPI_RED_F32: [M, ld_accum] fp32, accumulated with red.global.add
    int64_t ld_accum;
    float* ws;          // split-K scratch [N][NT] fp32 (store / gated), zero between calls
    int32_t* tickets;   // [n_tiles * 4], zero between calls
    const float* a_scale;  // fp8 operands only: per-token scales [M]
    const float* w_scale;  // fp8 operands only: per-feature scales [N]
    uint32_t idesc_or;     // fp8 kernels: operand-format bits for 6-bit (E3M2) weights, else 0
    uint32_t w_tx_bytes;   // bytes one weight-tile TMA load reports to the mbarrier (0 = the tile's shared-memory size)
    TpSync sync;           // tensor parallel: the producer acquires the peers' flags before its first activation load
                           // (sync.wait); the last CTA to retire releases the output to the peers (sync.signal)
};

template <int NT>
__device__ __forceinline__ void skinny_tmem_load(uint32_t taddr, float (&v)[NT]) {
    if constexpr (NT == 16) {
        uint32_t r[16];
        tmem_ld_32x16(taddr, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
    } else {
#pragma unroll
        for (int c = 0; c < NT; c += 32) {
            uint32_t r[32];
            tmem_ld_32x32(taddr + c, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) v[c + j] = __uint_as_float(r[j]);
        }
    }
}

// Final store of one feature column (this thread) for all live tokens. For the gated form the whole warp must call
// (shuffles): even lanes hold gate features, odd lanes the matching up features.
template <int NT, bool kBf16, bool kGated>
__device__ __forceinline__ void skinny_finalize(const SkinnyArgs& a, int n, uint32_t lane, float (&v)[NT]) {
    const bool live = n < a.N;
    const float b = (a.bias != nullptr && live) ? load_scalar<kBf16>(a.bias, n) : 0.0f;
#pragma unroll
    for (int t = 0; t < NT; ++t) v[t] += b;
    if constexpr (kGated) {
        // even lanes hold gate features (activated), odd lanes the matching up features (left as they are)
        apply_act_vec<NT>(v, (lane & 1) ? static_cast<int>(ACT_IDENTITY) : a.act);
#pragma unroll
        for (int t = 0; t < NT; ++t) {
            const float partner = __shfl_xor_sync(0xffffffffu, v[t], 1);
            if (live && !(lane & 1) && t < a.M)
                store_scalar<kBf16>(a.out, static_cast<int64_t>(t) * a.ldo + (n >> 1), v[t] * partner);
        }
    } else {
        if (!live) return;
        apply_act_vec<NT>(v, a.act);
#pragma unroll
        for (int t = 0; t < NT; ++t)
            if (t < a.M) store_scalar<kBf16>(a.out, static_cast<int64_t>(t) * a.ldo + n, v[t]);
    }
}

// kFp8: both operands are E4M3 codes (a 128-byte row of a tile then holds 128 K-elements instead of 64, i.e. half the
// weight bytes per MAC); the accumulator is scaled by w_scale[feature] * a_scale[token] right after it leaves TMEM, so
// split-K partial sums and all three epilogues see real-valued numbers. kBf16 then names the OUTPUT / bias dtype.
template <int NT, int EPI, bool kBf16, bool kFp8 = false>
__global__ void __launch_bounds__(kGemmThreads, skinny_occupancy(NT))
gemm_skinny_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                   const SkinnyArgs args) {
    constexpr int STAGES = skinny_stages(NT);
    constexpr uint32_t W_BYTES = kSkinnyBF * kGemmBK * 2;
    constexpr uint32_t X_BYTES = NT * kGemmBK * 2;
    constexpr uint32_t TMEM_COLS = NT < 32 ? 32 : NT;
    constexpr uint32_t IDESC = kFp8 ? make_idesc_f8(kSkinnyBF, NT) : make_idesc_f16(kSkinnyBF, NT, kBf16, false, false);
    constexpr int KB = kFp8 ? 2 * kGemmBK : kGemmBK;  // K-elements per 128-byte tile row
    static_assert(NT == 16 || NT == 32 || NT == 64, "token tile must be 16, 32 or 64");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_w = smem;
    uint8_t* smem_x = smem + STAGES * W_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * (W_BYTES + X_BYTES));
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull_bar + 1);

    const uint32_t warp = warp_id();
    const uint32_t lane = lane_id();
    const int n_tiles = (args.N + kSkinnyBF - 1) / kSkinnyBF;
    const int kb_total = (args.K + KB - 1) / KB;
    const int n_t = static_cast<int>(blockIdx.x) % n_tiles;
    const int sp = static_cast<int>(blockIdx.x) / n_tiles;
    const int kb0 = sp * args.kb_per_split;
    const int kb1 = min(kb_total, kb0 + args.kb_per_split);
    const uint32_t w_tx = args.w_tx_bytes ? args.w_tx_bytes : W_BYTES;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_w);
        prefetch_tmap(&tmap_x);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        mbar_init(tfull_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<TMEM_COLS>(tmem_slot);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    griddep_launch_dependents();
    if (warp == 0) {
        if (lane == 0) {
            // Weights are never written by a kernel: the first STAGES weight tiles are requested BEFORE waiting for
            // the previous kernel (PDL), so HBM keeps streaming across the kernel boundary; the activation tiles
            // (produced by the predecessor) follow after the wait and complete the same stage barriers.
            const int pre = min(STAGES, kb1 - kb0);
            for (int i = 0; i < pre; ++i) {
                mbar_expect_tx(&full_bar[i], w_tx + X_BYTES);
                tma_load_2d(smem_w + i * W_BYTES, &tmap_w, &full_bar[i], (kb0 + i) * KB, n_t * kSkinnyBF,
                            kEvictFirst);
            }
            griddep_wait();
            if (args.sync.wait) {  // tensor parallel: the activations were all-gathered into this buffer by the peers
                tp_wait_thread(args.sync);
                fence_proxy_async_global();
            }
            for (int i = 0; i < pre; ++i)
                tma_load_2d(smem_x + i * X_BYTES, &tmap_x, &full_bar[i], (kb0 + i) * KB, 0, kEvictLast);
            uint32_t stage = 0, phase = 1;  // the ring has been filled once
            for (int kb = kb0 + pre; kb < kb1; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                mbar_expect_tx(&full_bar[stage], w_tx + X_BYTES);
                tma_load_2d(smem_x + stage * X_BYTES, &tmap_x, &full_bar[stage], kb * KB, 0, kEvictLast);
                tma_load_2d(smem_w + stage * W_BYTES, &tmap_w, &full_bar[stage], kb * KB, n_t * kSkinnyBF,
                            kEvictFirst);
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint64_t dw = make_desc_kmajor_sw128(smem_u32(smem_w + stage * W_BYTES));
                const uint64_t dx = make_desc_kmajor_sw128(smem_u32(smem_x + stage * X_BYTES));
#pragma unroll
                for (int k = 0; k < 4; ++k) {  // 32 bytes of K per instruction: 16 (16-bit) or 32 (fp8) elements
                    if constexpr (kFp8) umma_f8_ss(tmem_base, dw + 2 * k, dx + 2 * k, IDESC | args.idesc_or, (kb > kb0 || k > 0) ? 1u : 0u);
                    else umma_f16_ss(tmem_base, dw + 2 * k, dx + 2 * k, IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[stage]);
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(tfull_bar);
        }
    } else {
        const uint32_t quad = warp & 3;
        const int n = n_t * kSkinnyBF + static_cast<int>(quad * 32 + lane);
        griddep_wait();  // this warp writes global memory the previous kernel may still be reading
        mbar_wait(tfull_bar, 0);
        tc_fence_after();
        float v[NT];
        skinny_tmem_load<NT>(tmem_base + ((quad * 32u) << 16), v);
        if constexpr (kFp8) {
            const float sw = n < args.N ? __ldg(args.w_scale + n) : 0.0f;
#pragma unroll
            for (int t = 0; t < NT; ++t) v[t] *= sw * (t < args.M ? __ldg(args.a_scale + t) : 0.0f);
        }
        if constexpr (EPI == EPI_RED_F32) {
            if (n < args.N) {
                const float b = (args.bias != nullptr && sp == 0) ? load_scalar<kBf16>(args.bias, n) : 0.0f;
#pragma unroll
                for (int t = 0; t < NT; ++t)
                    if (t < args.M) atomicAdd(args.accum + static_cast<int64_t>(t) * args.ld_accum + n, v[t] + b);
            }
        } else {
            constexpr bool kGated = EPI == EPI_GATED;
            if (args.splits == 1) {
                skinny_finalize<NT, kBf16, kGated>(args, n, lane, v);
            } else if (n_t * kSkinnyBF + static_cast<int>(quad * 32) < args.N) {  // warp-uniform
                float* slot = args.ws + static_cast<int64_t>(n) * NT;
                if (n < args.N) {
#pragma unroll
                    for (int t = 0; t < NT; t += 4) red_add_v4_f32(slot + t, v[t], v[t + 1], v[t + 2], v[t + 3]);
                }
                __threadfence();
                __syncwarp();
                int last = 0;
                if (lane == 0) last = (atomicAdd(&args.tickets[n_t * 4 + quad], 1) == args.splits - 1) ? 1 : 0;
                last = __shfl_sync(0xffffffffu, last, 0);
                if (last) {
                    __threadfence();
                    if (n < args.N) {
#pragma unroll
                        for (int t = 0; t < NT; t += 4) {
                            const float4 q = __ldcg(reinterpret_cast<const float4*>(slot + t));
                            v[t] = q.x; v[t + 1] = q.y; v[t + 2] = q.z; v[t + 3] = q.w;
                            __stcg(reinterpret_cast<float4*>(slot + t), make_float4(0.f, 0.f, 0.f, 0.f));
                        }
                    }
                    skinny_finalize<NT, kBf16, kGated>(args, n, lane, v);
                    __syncwarp();
                    if (lane == 0) args.tickets[n_t * 4 + quad] = 0;
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
    if (args.sync.signal) tp_signal_tail(args.sync, gridDim.x);
}

}  // namespace hds
