// CTA-pair GEMM (tcgen05 `cta_group::2`): two SMs of one TPC compute one 256 x BN output tile.
//
//   * a cluster of 2 CTAs; CTA r stages ITS 128 rows of the A tile and ITS HALF (BN/2 rows of W) of the B tile, so each
//     SM fills 32 KB of shared memory per 128-byte k-block instead of 48 KB (2/3 of the L2 -> SM traffic per FLOP) and
//     6 stages fit instead of 4
//   * both producers issue `cp.async.bulk.tensor ... .cta_group::2`, whose completion bytes land on the LEADER's
//     (rank 0) "full" mbarrier; the leader's elected thread issues `tcgen05.mma.cta_group::2` (M = 256): every SM's
//     tensor core multiplies its own A rows with both halves of B, reading the peer's half out of the peer's smem
//   * `tcgen05.commit ... .multicast::cluster` releases the stage in BOTH CTAs and publishes the accumulator to both
//     epilogues; each CTA's 4 epilogue warps drain their own 128 TMEM lanes and arrive (remotely) on the leader's
//     "accumulator empty" barrier. Accumulators are double buffered (2 x BN of the 512 TMEM columns).
//   * operand kinds: bf16 / fp16 (`kind::f16`, k-block = 64 elements) and E4M3 (`kind::f8f6f4`, k-block = 128 elements,
//     per-row / per-channel scales in the epilogue) share the kernel: a stage is 128 bytes of K either way.
//
// Operands must live in THIS GPU's memory: every host entry checks the operand pointers (cudaPointerGetAttributes) and
// refuses anything else, and the peer-sharded restore (hds_restore_kv_proj_peers) stays on the single-CTA kernel.
//
// Single-CTA kernels (gemm_core.cuh, gemm_fp8.cu) remain for M <= 128, split-K, grouped and fused-epilogue variants.
// Reference role: the cuBLAS / CUTLASS GEMMs behind `blas_kernels` and `cutlass_ops` (inference/v2/kernels).
#include <stdlib.h>
#include "gemm_core.cuh"
#include "hds_host.cuh"
#include "tp_sync_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared::cluster address -> leader CTA

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same smem offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}\n" ::"r"(smem_u32(bar)),
        "r"(cta)
        : "memory");
}
// TMA tile load of one CTA of the pair; completion bytes are credited to the LEADER's mbarrier
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* tmap, uint64_t* bar, int32_t c0, int32_t c1,
                                                uint64_t hint) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "l"(hint)
        : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_result) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "n"(kCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
template <bool kFp8>
__device__ __forceinline__ void umma_ss_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
    if constexpr (kFp8) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
            "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
            "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}
// arrive (once the MMAs issued so far have completed) on the barrier at this offset in every CTA of `mask`
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"(mask)
                 : "memory");
}

constexpr uint32_t k2smStageA = 128 * 128;  // 128 rows x 128 B of K per CTA
__host__ __device__ constexpr uint32_t k2sm_stage_b(int bn) { return (bn / 2) * 128; }
__host__ __device__ constexpr int k2sm_stages(int bn) { return (200 * 1024) / (k2smStageA + k2sm_stage_b(bn)); }
__host__ __device__ constexpr size_t k2sm_smem_bytes(int bn) {
    return static_cast<size_t>(k2sm_stages(bn)) * (k2smStageA + k2sm_stage_b(bn)) + 512 + 1024;
}

struct Gemm2smArgs {
    GemmArgs g;            // M, N, K, out, ldo, bias, act
    const float* a_scale;  // fp8 only: [M]
    const float* w_scale;  // fp8 only: [N]
};

template <int BN, bool kFp8, bool kBf16, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_2sm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w, const Gemm2smArgs fa) {
    constexpr int STAGES = k2sm_stages(BN);
    constexpr uint32_t A_BYTES = k2smStageA, W_BYTES = k2sm_stage_b(BN);
    constexpr uint32_t TMEM_COLS = gemm_tmem_cols(BN);
    constexpr int KB_ELEMS = kFp8 ? 128 : 64;            // elements of K in one 128-byte k-block
    constexpr int MMA_K = kFp8 ? 32 : 16;                // elements of K per instruction (32 bytes either way)
    constexpr uint32_t IDESC = kFp8 ? make_idesc_f8(256, BN) : make_idesc_f16(256, BN, kBf16, false, false);
    constexpr bool kGated = EPI == EPI_GATED;
    constexpr bool kRope = EPI == EPI_KV_ROPE || EPI == EPI_QKV_ROPE;  // RoPE + paged-KV epilogue (head_dim 128)
    static_assert(!(kRope && kFp8), "the RoPE epilogues take 16-bit operands");
    const GemmArgs& args = fa.g;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_w = smem + STAGES * A_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + W_BYTES));
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const uint32_t warp = warp_id(), lane = lane_id();
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int m_tiles = (args.M + 255) / 256;
    const int n_tiles = (args.N + BN - 1) / BN;
    const int kb_total = (args.K + KB_ELEMS - 1) / KB_ELEMS;
    const int total_work = m_tiles * n_tiles;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_w);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], 8);  // 4 epilogue warps x 2 CTAs (only the leader's copy is waited on)
        }
        fence_barrier_init();
    }
    __syncwarp();
    cluster_sync_all();  // the peer's barriers exist before anything is signalled across the pair
    if (warp == 1) tmem_alloc_2sm<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    griddep_launch_dependents();
    griddep_wait();

    if (warp == 0) {
        if (lane == 0) {
            if (args.sync.wait) {  // the activation operand was all-gathered into THIS GPU's buffer by the peers
                tp_wait_thread(args.sync);
                fence_proxy_async_global();
            }
            uint32_t stage = 0, phase = 0;
            for (int w = pair; w < total_work; w += n_pairs) {
                const int m_t = w % m_tiles, n_t = w / m_tiles;
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * (A_BYTES + (args.w_tx_bytes ? args.w_tx_bytes : W_BYTES)));
                    tma_load_2d_2sm(smem_a + stage * A_BYTES, &tmap_a, &full_bar[stage], kb * KB_ELEMS,
                                    m_t * 256 + static_cast<int>(rank) * 128, kEvictLast);
                    tma_load_2d_2sm(smem_w + stage * W_BYTES, &tmap_w, &full_bar[stage], kb * KB_ELEMS,
                                    n_t * BN + static_cast<int>(rank) * (BN / 2), m_tiles > 1 ? kEvictNormal : kEvictFirst);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0 && lane == 0) {
            uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
            for (int w = pair; w < total_work; w += n_pairs) {
                mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint64_t da = make_desc_kmajor_sw128(smem_u32(smem_a + stage * A_BYTES));
                    const uint64_t dw = make_desc_kmajor_sw128(smem_u32(smem_w + stage * W_BYTES));
#pragma unroll
                    for (int k = 0; k < KB_ELEMS / MMA_K; ++k)  // 32 bytes of K per instruction inside the swizzle atom
                        umma_ss_2sm<kFp8>(d_tmem, da + 2 * k, dw + 2 * k, IDESC | args.idesc_or, (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit_2sm(&empty_bar[stage], 0b11);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit_2sm(&tfull_bar[acc], 0b11);
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else {
        const uint32_t quad = warp & 3;
        uint32_t acc = 0, acc_phase = 0;
        for (int w = pair; w < total_work; w += n_pairs) {
            const int m_t = w % m_tiles, n_t = w / m_tiles;
            mbar_wait(&tfull_bar[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + acc * BN;
            const int row = m_t * 256 + static_cast<int>(rank) * 128 + quad * 32 + lane;
            float sa = 1.0f;
            if constexpr (kFp8) sa = row < args.M ? fa.a_scale[row] : 0.0f;
            if constexpr (kRope) epilogue_kv_rope<BN, 128, kBf16, EPI == EPI_QKV_ROPE>(args, taddr, row, n_t);
#pragma unroll 1
            for (int c = 0; c < (kRope ? 0 : BN); c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(taddr + c, r);
                tmem_ld_wait();
                const int n = n_t * BN + c;
                if (row < args.M && n < args.N) {
                    float v[32];
                    if constexpr (kFp8) {
                        if (n + 32 <= args.N && (reinterpret_cast<uintptr_t>(fa.w_scale + n) & 15) == 0) {
                            const float4* ws = reinterpret_cast<const float4*>(fa.w_scale + n);
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 s = __ldg(ws + j);
                                v[4 * j + 0] = __uint_as_float(r[4 * j + 0]) * sa * s.x;
                                v[4 * j + 1] = __uint_as_float(r[4 * j + 1]) * sa * s.y;
                                v[4 * j + 2] = __uint_as_float(r[4 * j + 2]) * sa * s.z;
                                v[4 * j + 3] = __uint_as_float(r[4 * j + 3]) * sa * s.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                v[j] = __uint_as_float(r[j]) * sa * (n + j < args.N ? __ldg(fa.w_scale + n + j) : 0.0f);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    }
                    if (kGated) gated_chunk<kBf16>(args, row, n, v);
                    else store_chunk<kBf16>(args, row, n, v);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc], 0);
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1;
            }
        }
    }
    // the leader's MMAs read the peer's shared memory and both CTAs' TMEM: nobody leaves before everyone is done
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2sm<TMEM_COLS>(tmem_base);
    }
    if (args.sync.signal) tp_signal_tail(args.sync, gridDim.x);
}

template <int BN, bool kFp8, bool kBf16, int EPI>
static int launch_2sm_inst(const CUtensorMap& ta, const CUtensorMap& tw, const Gemm2smArgs& fa, int grid, cudaStream_t stream) {
    auto kern = gemm_2sm_kernel<BN, kFp8, kBf16, EPI>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(k2sm_smem_bytes(BN))));
        configured[dev & 63] = true;
    }
    HDS_CHECK_CUDA(launch_kernel(kern, dim3(grid), dim3(kGemmThreads), k2sm_smem_bytes(BN), stream, ta, tw, fa));
    return HDS_OK;
}

template <bool kFp8, bool kBf16>
static int launch_2sm(int bn, bool gated, const CUtensorMap& ta, const CUtensorMap& tw, const Gemm2smArgs& fa, int grid,
                      cudaStream_t s) {
    if (bn == 256)
        return gated ? launch_2sm_inst<256, kFp8, kBf16, EPI_GATED>(ta, tw, fa, grid, s)
                     : launch_2sm_inst<256, kFp8, kBf16, EPI_STORE>(ta, tw, fa, grid, s);
    return gated ? launch_2sm_inst<128, kFp8, kBf16, EPI_GATED>(ta, tw, fa, grid, s)
                 : launch_2sm_inst<128, kFp8, kBf16, EPI_STORE>(ta, tw, fa, grid, s);
}

// true iff `p` is ordinary device memory owned by the current GPU (not a peer mapping, not host / managed memory)
bool is_local_device_ptr(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    int dev = -1;
    cudaGetDevice(&dev);
    return at.type == cudaMemoryTypeDevice && at.device == dev;
}

// Shared host path: operands are 16-bit (a_scale == nullptr) or E4M3 with row / channel scales.
int gemm_2sm_dispatch(const void* A, int64_t lda_bytes, const float* a_scale, const void* W, int64_t ldw_bytes,
                      const float* w_scale, int M, int N, int K, bool fp8, bool bf16, int epilogue, int act, const void* bias,
                      void* out, int64_t ldo, int bn, cudaStream_t stream, int w_fmt) {
    if (bn <= 0) {
        const int64_t m_tiles = (M + 255) / 256;
        bn = m_tiles * ((N + 255) / 256) * 2 >= sm_count() ? 256 : 128;
    }
    if (bn != 128 && bn != 256) {
        set_error("CTA-pair GEMM: bn must be 128 or 256");
        return HDS_ERR_BAD_ARG;
    }
    if (!is_local_device_ptr(A) || !is_local_device_ptr(W)) {
        set_error("CTA-pair GEMM: operands must be this GPU's own device memory");
        return HDS_ERR_UNSUPPORTED;
    }
    Gemm2smArgs fa;
    memset(&fa, 0, sizeof(fa));
    fa.g.M = M;
    fa.g.N = N;
    fa.g.K = K;
    fa.g.act = act;
    fa.g.out = out;
    fa.g.ldo = ldo;
    fa.g.bias = bias;
    fa.a_scale = a_scale;
    fa.w_scale = w_scale;
    fa.g.sync = tp_take_armed();
    CUtensorMap ta, tw;
    int rc;
    if (fp8) {
        rc = make_tmap_2d_8b(&ta, A, K, M, lda_bytes, 128, 128);
        if (rc) return rc;
        if (w_fmt == 1) {   // E3M2 weights, packed 6-bit in global memory; the weights are the MMA's B operand
            rc = make_tmap_2d_6b(&tw, W, K, N, ldw_bytes, bn / 2);
            fa.g.idesc_or = 4u << 10;
            fa.g.w_tx_bytes = static_cast<uint32_t>(bn / 2) * f6_tx_bytes_per_row();
        } else {
            rc = make_tmap_2d_8b(&tw, W, K, N, ldw_bytes, 128, bn / 2);
        }
    } else {
        rc = make_tmap_2d_16b(&ta, A, K, M, lda_bytes, 64, 128, bf16);
        if (rc) return rc;
        rc = make_tmap_2d_16b(&tw, W, K, N, ldw_bytes, 64, bn / 2, bf16);
    }
    if (rc) return rc;
    const int work = ((M + 255) / 256) * ((N + bn - 1) / bn);
    const int pairs = work < sm_count() / 2 ? work : sm_count() / 2;
    const int grid = 2 * pairs;
    const bool gated = epilogue == EPI_GATED;
    if (fp8) return bf16 ? launch_2sm<true, true>(bn, gated, ta, tw, fa, grid, stream)
                         : launch_2sm<true, false>(bn, gated, ta, tw, fa, grid, stream);
    return bf16 ? launch_2sm<false, true>(bn, gated, ta, tw, fa, grid, stream)
                : launch_2sm<false, false>(bn, gated, ta, tw, fa, grid, stream);
}

// QKV / KV projection with the RoPE + paged-KV epilogue on the pair kernel. `g` carries the epilogue fields
// (bias, q_out, kv_cache, tok_pos, tok_slot, inv_freq, head counts, block_size, do_rope) and M, N, K; head_dim is 128.
int gemm_2sm_dispatch_rope(const GemmArgs& g, const void* A, int64_t lda_bytes, const void* W, int64_t ldw_bytes, bool bf16,
                           bool has_q, cudaStream_t stream) {
    if (!is_local_device_ptr(A) || !is_local_device_ptr(W)) {
        set_error("CTA-pair GEMM: operands must be this GPU's own device memory");
        return HDS_ERR_UNSUPPORTED;
    }
    Gemm2smArgs fa;
    memset(&fa, 0, sizeof(fa));
    fa.g = g;
    CUtensorMap ta, tw;
    int rc = make_tmap_2d_16b(&ta, A, g.K, g.M, lda_bytes, 64, 128, bf16);
    if (rc) return rc;
    rc = make_tmap_2d_16b(&tw, W, g.K, g.N, ldw_bytes, 64, 128, bf16);
    if (rc) return rc;
    const int work = ((g.M + 255) / 256) * ((g.N + 255) / 256);
    const int pairs = work < sm_count() / 2 ? work : sm_count() / 2;
    if (has_q)
        return bf16 ? launch_2sm_inst<256, false, true, EPI_QKV_ROPE>(ta, tw, fa, 2 * pairs, stream)
                    : launch_2sm_inst<256, false, false, EPI_QKV_ROPE>(ta, tw, fa, 2 * pairs, stream);
    return bf16 ? launch_2sm_inst<256, false, true, EPI_KV_ROPE>(ta, tw, fa, 2 * pairs, stream)
                : launch_2sm_inst<256, false, false, EPI_KV_ROPE>(ta, tw, fa, 2 * pairs, stream);
}

}  // namespace hds

using namespace hds;

extern "C" int hds_gemm_2sm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int dtype,
                            int epilogue, int act, const void* bias, void* out, int64_t ldo, int bn, void* stream) {
    HDS_REQUIRE(epilogue == EPI_STORE || epilogue == EPI_GATED, "hds_gemm_2sm: epilogue must be 0 (store) or 1 (gated)");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_gemm_2sm: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(epilogue != EPI_GATED || (N % 2 == 0), "hds_gemm_2sm: gated epilogue needs an even N");
    HDS_REQUIRE(M > 0 && N > 0 && K > 0 && K % 8 == 0, "hds_gemm_2sm: K must be a positive multiple of 8");
    return gemm_2sm_dispatch(A, lda * 2, nullptr, W, ldw * 2, nullptr, M, N, K, false, dtype == 0, epilogue, act, bias, out,
                             ldo, bn, static_cast<cudaStream_t>(stream), 0);
}
