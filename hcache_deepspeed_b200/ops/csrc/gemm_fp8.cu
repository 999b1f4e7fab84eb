// FP8 (E4M3) GEMM on the 5th-generation tensor cores:  out[M,N] = act((Aq * a_scale) (Wq * w_scale)^T + bias)
//
//   * `tcgen05.mma.kind::f8f6f4`, 128 x BN x 32 per instruction, fp32 accumulation in TMEM
//   * activations quantised per TOKEN (row-wise amax / 448) by hds_quantize_fp8_rows just before the GEMM, weights per
//     OUTPUT CHANNEL once at load time; both scales are applied in the epilogue (acc * a_scale[m] * w_scale[n])
//   * same persistent warp-specialised pipeline as gemm_core.cuh (TMA producer / MMA issuer / 4 epilogue warps,
//     double-buffered accumulator); a k-block is 128 BYTES of K = 128 elements, i.e. half the HBM traffic per MAC
//
// This is the Blackwell replacement for the reference's weight-only FP6 path, which is an sm_80 mma.sync kernel whose
// body is compiled out on sm_90+ (core_ops/cuda_linear/include/kernel_matmul.cuh:21-272, linear_kernels_cuda.cu:51-71).
#include <cuda_fp8.h>
#include <stdlib.h>
#include "gemm_core.cuh"
#include "gemm_skinny.cuh"
#include "hds_host.cuh"
#include "tp_sync_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

constexpr int kF8BK = 128;  // elements (= bytes) of K per pipeline stage
constexpr float kE4M3Max = 448.0f;

__host__ __device__ constexpr int f8_stages(int bn) { return (200 * 1024) / (128 * kF8BK + bn * kF8BK); }
__host__ __device__ constexpr size_t f8_smem_bytes(int bn) {
    return static_cast<size_t>(f8_stages(bn)) * (128 * kF8BK + bn * kF8BK) + 512 + 1024;
}

// ---- row-wise quantiser: one CTA per row, 16-byte loads / 8-byte stores ---------------------------------------
template <bool kBf16>
__global__ void __launch_bounds__(256) quantize_fp8_rows_kernel(const void* __restrict__ x, int64_t ldx,
                                                                uint8_t* __restrict__ q, int64_t ldq,
                                                                float* __restrict__ scales, int K, const TpSync sync) {
    __shared__ float red[8];
    pdl_enter();
    if (sync.wait) tp_wait_cta(sync);  // tensor parallel: the rows were all-gathered into this buffer by the peers
    const int r = blockIdx.x;
    const char* row = reinterpret_cast<const char*>(x) + static_cast<int64_t>(r) * ldx * 2;
    uint8_t* qrow = q + static_cast<int64_t>(r) * ldq;
    // vector path: 8 elements per step (rows 16-byte aligned, K a multiple of 8); the second pass re-reads the row from L1/L2
    const bool vec = (K & 7) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0 && (reinterpret_cast<uintptr_t>(qrow) & 7) == 0;
    float mx = 0.0f;
    if (vec) {
        for (int i = threadIdx.x * 8; i < K; i += blockDim.x * 8) {
            const uint4 p = *reinterpret_cast<const uint4*>(row + i * 2);
            const uint32_t w[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) mx = fmaxf(mx, fmaxf(fabsf(unpack_lo<kBf16>(w[j])), fabsf(unpack_hi<kBf16>(w[j]))));
        }
    } else {
        for (int i = threadIdx.x; i < K; i += blockDim.x) mx = fmaxf(mx, fabsf(load_scalar<kBf16>(x, static_cast<int64_t>(r) * ldx + i)));
    }
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    mx = red[0];
    for (int w = 1; w < static_cast<int>(blockDim.x >> 5); ++w) mx = fmaxf(mx, red[w]);
    const float scale = mx > 0.0f ? mx / kE4M3Max : 1.0f;
    if (threadIdx.x == 0) scales[r] = scale;
    const float inv = 1.0f / scale;
    if (vec) {
        for (int i = threadIdx.x * 8; i < K; i += blockDim.x * 8) {
            const uint4 p = *reinterpret_cast<const uint4*>(row + i * 2);
            const uint32_t w[4] = {p.x, p.y, p.z, p.w};
            uint32_t o[2] = {0u, 0u};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t lo = __nv_cvt_float_to_fp8(unpack_lo<kBf16>(w[j]) * inv, __NV_SATFINITE, __NV_E4M3);
                const uint32_t hi = __nv_cvt_float_to_fp8(unpack_hi<kBf16>(w[j]) * inv, __NV_SATFINITE, __NV_E4M3);
                o[j >> 1] |= (lo | (hi << 8)) << ((j & 1) * 16);
            }
            *reinterpret_cast<uint2*>(qrow + i) = make_uint2(o[0], o[1]);
        }
    } else {
        for (int i = threadIdx.x; i < K; i += blockDim.x) {
            const float v = load_scalar<kBf16>(x, static_cast<int64_t>(r) * ldx + i) * inv;
            qrow[i] = static_cast<uint8_t>(__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3));
        }
    }
}

struct F8Args {
    GemmArgs g;              // M, N, K, out, ldo, bias, act (store / gated epilogue helpers of gemm_core.cuh)
    const float* a_scale;    // [M]
    const float* w_scale;    // [N]
};

template <int BN, bool kBf16, bool kGated>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_fp8_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w, const F8Args fa) {
    constexpr int STAGES = f8_stages(BN);
    constexpr uint32_t A_BYTES = 128 * kF8BK;
    constexpr uint32_t W_BYTES = BN * kF8BK;
    constexpr uint32_t TMEM_COLS = gemm_tmem_cols(BN);
    constexpr uint32_t IDESC = make_idesc_f8(128, BN);
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
    const int m_tiles = (args.M + 127) / 128;
    const int n_tiles = (args.N + BN - 1) / BN;
    const int kb_total = (args.K + kF8BK - 1) / kF8BK;
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
            mbar_init(&tempty_bar[i], 4);
        }
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
    griddep_wait();

    if (warp == 0) {
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
                const int m_t = w % m_tiles, n_t = w / m_tiles;
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], A_BYTES + (args.w_tx_bytes ? args.w_tx_bytes : W_BYTES));
                    tma_load_2d(smem_a + stage * A_BYTES, &tmap_a, &full_bar[stage], kb * kF8BK, m_t * 128, kEvictLast);
                    tma_load_2d(smem_w + stage * W_BYTES, &tmap_w, &full_bar[stage], kb * kF8BK, n_t * BN,
                                m_tiles > 1 ? kEvictNormal : kEvictFirst);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
            for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
                mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint64_t da = make_desc_kmajor_sw128(smem_u32(smem_a + stage * A_BYTES));
                    const uint64_t dw = make_desc_kmajor_sw128(smem_u32(smem_w + stage * W_BYTES));
#pragma unroll
                    for (int k = 0; k < kF8BK / 32; ++k)  // 32 elements = 32 B per MMA inside the 128 B swizzle atom
                        umma_f8_ss(d_tmem, da + 2 * k, dw + 2 * k, IDESC | args.idesc_or, (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&empty_bar[stage]);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(&tfull_bar[acc]);
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else {
        const uint32_t quad = warp & 3;
        uint32_t acc = 0, acc_phase = 0;
        for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
            const int m_t = w % m_tiles, n_t = w / m_tiles;
            mbar_wait(&tfull_bar[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + acc * BN;
            const int row = m_t * 128 + quad * 32 + lane;
            const float sa = row < args.M ? fa.a_scale[row] : 0.0f;
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(taddr + c, r);
                tmem_ld_wait();
                const int n = n_t * BN + c;
                if (row < args.M && n < args.N) {
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        v[j] = __uint_as_float(r[j]) * sa * (n + j < args.N ? __ldg(fa.w_scale + n + j) : 0.0f);
                    if (kGated) gated_chunk<kBf16>(args, row, n, v);
                    else store_chunk<kBf16>(args, row, n, v);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[acc]);
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1;
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

namespace hds {

template <int BN, bool kBf16, bool kGated>
static int launch_fp8_inst(const CUtensorMap& ta, const CUtensorMap& tw, const F8Args& fa, int grid, cudaStream_t stream) {
    auto kern = gemm_fp8_kernel<BN, kBf16, kGated>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(f8_smem_bytes(BN))));
        configured[dev & 63] = true;
    }
    HDS_CHECK_CUDA(launch_kernel(kern, dim3(grid), dim3(kGemmThreads), f8_smem_bytes(BN), stream, ta, tw, fa));
    return HDS_OK;
}

template <bool kBf16, bool kGated>
static int launch_fp8_bn(int bn, const CUtensorMap& ta, const CUtensorMap& tw, const F8Args& fa, int grid, cudaStream_t s) {
    if (bn == 256) return launch_fp8_inst<256, kBf16, kGated>(ta, tw, fa, grid, s);
    if (bn == 128) return launch_fp8_inst<128, kBf16, kGated>(ta, tw, fa, grid, s);
    return launch_fp8_inst<64, kBf16, kGated>(ta, tw, fa, grid, s);
}

}  // namespace hds

namespace hds {
// gemm_2sm.cu: CTA-pair (cta_group::2) kernel shared by the 16-bit and fp8 operand kinds
int gemm_2sm_dispatch(const void* A, int64_t lda_bytes, const float* a_scale, const void* W, int64_t ldw_bytes,
                      const float* w_scale, int M, int N, int K, bool fp8, bool bf16, int epilogue, int act, const void* bias,
                      void* out, int64_t ldo, int bn, cudaStream_t stream, int w_fmt = 0);

// Weight format of the NEXT hds_gemm_fp8 / hds_gemm_fp8_skinny call of this host thread (consumed by it):
// 0 = E4M3 bytes (default), 1 = E3M2 packed 6-bit.
static thread_local int g_next_w_fmt = 0;
static int take_w_fmt() {
    const int f = g_next_w_fmt;
    g_next_w_fmt = 0;
    return f;
}
}  // namespace hds

using namespace hds;

extern "C" int hds_quantize_fp8_rows(const void* x, int64_t ldx, int rows, int K, int dtype, void* q, int64_t ldq,
                                     float* scales, void* stream) {
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_quantize_fp8_rows: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(rows >= 0 && K > 0, "hds_quantize_fp8_rows: bad shape");
    if (rows == 0) return HDS_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    TpSync sync = tp_take_armed();
    sync.signal = 0;
    if (dtype == 0)
        HDS_CHECK_CUDA(launch_kernel(quantize_fp8_rows_kernel<true>, dim3(rows), dim3(K >= 8192 ? 256 : 128), 0, s, x, ldx,
                                     reinterpret_cast<uint8_t*>(q), ldq, scales, K, sync));
    else
        HDS_CHECK_CUDA(launch_kernel(quantize_fp8_rows_kernel<false>, dim3(rows), dim3(K >= 8192 ? 256 : 128), 0, s, x, ldx,
                                     reinterpret_cast<uint8_t*>(q), ldq, scales, K, sync));
    return HDS_OK;
}

extern "C" int hds_fp8_next_weight_format(int fmt) {
    HDS_REQUIRE(fmt == 0 || fmt == 1, "hds_fp8_next_weight_format: 0 (E4M3) or 1 (E3M2 packed 6-bit)");
    hds::g_next_w_fmt = fmt;
    return HDS_OK;
}

extern "C" int hds_gemm_fp8(const void* Aq, int64_t lda, const float* a_scale, const void* Wq, int64_t ldw,
                            const float* w_scale, int M, int N, int K, int out_dtype, int epilogue, int act,
                            const void* bias, void* out, int64_t ldo, int bn, void* stream) {
    HDS_REQUIRE(epilogue == EPI_STORE || epilogue == EPI_GATED, "hds_gemm_fp8: epilogue must be 0 (store) or 1 (gated)");
    HDS_REQUIRE(out_dtype == 0 || out_dtype == 1, "hds_gemm_fp8: out dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(epilogue != EPI_GATED || (N % 2 == 0), "hds_gemm_fp8: gated epilogue needs an even N");
    HDS_REQUIRE(M > 0 && N > 0 && K > 0 && K % 16 == 0, "hds_gemm_fp8: K must be a positive multiple of 16");
    const int w_fmt = take_w_fmt();
    HDS_REQUIRE(w_fmt == 0 || K % 128 == 0, "hds_gemm_fp8: 6-bit weights need K to be a multiple of 128");
    // bn: 0 = auto (CTA-pair kernel above one m-tile), 64 / 128 / 256 = single-CTA tile, 1128 / 1256 = CTA-pair tile
    if ((bn == 0 && M > 128) || bn > 1000)
        return gemm_2sm_dispatch(Aq, lda, a_scale, Wq, ldw, w_scale, M, N, K, true, out_dtype == 0, epilogue, act, bias, out,
                                 ldo, bn > 1000 ? bn - 1000 : 0, static_cast<cudaStream_t>(stream), w_fmt);
    if (bn <= 0) {
        // widest tile that still gives every SM work; decode-size problems take the narrow tile for more CTAs
        const int64_t m_tiles = (M + 127) / 128;
        bn = m_tiles * ((N + 255) / 256) >= sm_count() ? 256 : m_tiles * ((N + 127) / 128) >= sm_count() ? 128 : 64;
    }
    HDS_REQUIRE(bn == 64 || bn == 128 || bn == 256, "hds_gemm_fp8: bn must be 64, 128 or 256");
    F8Args fa;
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
    int rc = make_tmap_2d_8b(&ta, Aq, K, M, lda, kF8BK, 128);
    if (rc) return rc;
    if (w_fmt == 1) {   // E3M2 weights (packed 6-bit): the weights are the MMA's B operand in this kernel
        rc = make_tmap_2d_6b(&tw, Wq, K, N, ldw, bn);
        fa.g.idesc_or = 4u << 10;
        fa.g.w_tx_bytes = static_cast<uint32_t>(bn) * f6_tx_bytes_per_row();
    } else {
        rc = make_tmap_2d_8b(&tw, Wq, K, N, ldw, kF8BK, bn);
    }
    if (rc) return rc;
    const int work = ((M + 127) / 128) * ((N + bn - 1) / bn);
    const int grid = work < sm_count() ? work : sm_count();
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (out_dtype == 0)
        return epilogue == EPI_STORE ? launch_fp8_bn<true, false>(bn, ta, tw, fa, grid, s)
                                     : launch_fp8_bn<true, true>(bn, ta, tw, fa, grid, s);
    return epilogue == EPI_STORE ? launch_fp8_bn<false, false>(bn, ta, tw, fa, grid, s)
                                 : launch_fp8_bn<false, true>(bn, ta, tw, fa, grid, s);
}

// ---- decode-size M: the feature-major weight-streaming kernel (gemm_skinny.cuh) with E4M3 operands --------------
namespace hds {

template <int NT, int EPI, bool kBf16>
static int launch_skinny_fp8_inst(const CUtensorMap& tw, const CUtensorMap& tx, const SkinnyArgs& a, int grid, cudaStream_t stream) {
    auto kern = gemm_skinny_kernel<NT, EPI, kBf16, true>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(skinny_smem_bytes(NT))));
        configured[dev & 63] = true;
    }
    HDS_CHECK_CUDA(launch_kernel(kern, dim3(grid), dim3(kGemmThreads), skinny_smem_bytes(NT), stream, tw, tx, a));
    return HDS_OK;
}

template <int EPI, bool kBf16>
static int launch_skinny_fp8_nt(int nt, const CUtensorMap& tw, const CUtensorMap& tx, const SkinnyArgs& a, int grid,
                                cudaStream_t s) {
    if (nt == 16) return launch_skinny_fp8_inst<16, EPI, kBf16>(tw, tx, a, grid, s);
    if (nt == 32) return launch_skinny_fp8_inst<32, EPI, kBf16>(tw, tx, a, grid, s);
    return launch_skinny_fp8_inst<64, EPI, kBf16>(tw, tx, a, grid, s);
}

}  // namespace hds

extern "C" int64_t hds_gemm_fp8_skinny_workspace_bytes(int M, int N) {
    const int nt = M <= 16 ? 16 : M <= 32 ? 32 : 64;
    const int64_t n_tiles = (N + hds::kSkinnyBF - 1) / hds::kSkinnyBF;
    return ((n_tiles * hds::kSkinnyBF * nt * 4 + 255) & ~int64_t(255)) + n_tiles * 16;
}

extern "C" int hds_gemm_fp8_skinny(const void* Aq, int64_t lda, const float* a_scale, const void* Wq, int64_t ldw,
                                   const float* w_scale, int M, int N, int K, int out_dtype, int epilogue, int act,
                                   const void* bias, void* out, int64_t ldo, float* accum, int64_t ld_accum, void* ws,
                                   int64_t ws_bytes, int want_splits, void* stream) {
    using namespace hds;
    HDS_REQUIRE(epilogue == EPI_STORE || epilogue == EPI_GATED || epilogue == EPI_RED_F32,
                "hds_gemm_fp8_skinny: epilogue must be 0 (store), 1 (gated) or 2 (fp32 accumulate)");
    HDS_REQUIRE(out_dtype == 0 || out_dtype == 1, "hds_gemm_fp8_skinny: out dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(M > 0 && M <= 64 && N > 0 && K > 0 && K % 16 == 0, "hds_gemm_fp8_skinny: 1 <= M <= 64, K a multiple of 16");
    HDS_REQUIRE(epilogue != EPI_GATED || N % 2 == 0, "hds_gemm_fp8_skinny: gated epilogue needs an even N");
    HDS_REQUIRE(epilogue != EPI_RED_F32 || accum != nullptr, "hds_gemm_fp8_skinny: accum required");
    const int nt = M <= 16 ? 16 : M <= 32 ? 32 : 64;
    const int n_tiles = (N + kSkinnyBF - 1) / kSkinnyBF;
    const int kb_total = (K + 127) / 128;
    const int64_t ws_data = (static_cast<int64_t>(n_tiles) * kSkinnyBF * nt * 4 + 255) & ~int64_t(255);
    const bool can_split = epilogue == EPI_RED_F32 || (ws != nullptr && ws_bytes >= ws_data + static_cast<int64_t>(n_tiles) * 16);
    int splits = 1;
    if (can_split) {
        if (want_splits > 0) {
            splits = want_splits;
        } else {
            // as launch_skinny (gemm.cu): measured (profiles/kernel_bench_decode_gemm.jsonl, BIG_SHAPES run): fill the resident slots once; past one full
            // wave of tiles, split until the waves are >= 80 % full (448 tiles of a 57344-wide MLP-1: 5.25 -> 5.66 TB/s at 4 splits)
            const int slots = sm_count() * skinny_occupancy(nt);
            splits = slots / n_tiles;
            if (n_tiles > slots) {
                splits = 1;
                for (int s = 1; s <= 8; ++s) {
                    const int64_t ctas = static_cast<int64_t>(n_tiles) * s;
                    const int64_t waves = (ctas + slots - 1) / slots;
                    if (ctas * 100 >= waves * slots * 80) {
                        splits = s;
                        break;
                    }
                }
            }
            if (splits > kb_total / 4) splits = kb_total / 4;
            if (splits > 8) splits = 8;
        }
        if (splits < 1) splits = 1;
        if (splits > kb_total) splits = kb_total;
    }
    const int per = (kb_total + splits - 1) / splits;
    splits = (kb_total + per - 1) / per;
    SkinnyArgs a;
    memset(&a, 0, sizeof(a));
    a.M = M; a.N = N; a.K = K;
    a.splits = splits;
    a.kb_per_split = per;
    a.act = act;
    a.out = out; a.ldo = ldo; a.bias = bias;
    a.accum = accum; a.ld_accum = ld_accum;
    a.a_scale = a_scale; a.w_scale = w_scale;
    a.sync = tp_take_armed();
    if (splits > 1 && epilogue != EPI_RED_F32) {
        a.ws = reinterpret_cast<float*>(ws);
        a.tickets = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(ws) + ws_data);
    }
    CUtensorMap tw, tx;
    const int w_fmt = take_w_fmt();
    HDS_REQUIRE(w_fmt == 0 || K % 128 == 0, "hds_gemm_fp8_skinny: 6-bit weights need K to be a multiple of 128");
    int rc;
    if (w_fmt == 1) {   // E3M2 weights (packed 6-bit): the weights are the MMA's A operand in the feature-major kernel
        rc = make_tmap_2d_6b(&tw, Wq, K, N, ldw, kSkinnyBF);
        a.idesc_or = 4u << 7;
        a.w_tx_bytes = static_cast<uint32_t>(kSkinnyBF) * f6_tx_bytes_per_row();
    } else {
        rc = make_tmap_2d_8b(&tw, Wq, K, N, ldw, 128, kSkinnyBF);
    }
    if (rc) return rc;
    rc = make_tmap_2d_8b(&tx, Aq, K, M, lda, 128, nt);
    if (rc) return rc;
    const int grid = n_tiles * splits;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (out_dtype == 0) {
        if (epilogue == EPI_STORE) return launch_skinny_fp8_nt<EPI_STORE, true>(nt, tw, tx, a, grid, s);
        if (epilogue == EPI_GATED) return launch_skinny_fp8_nt<EPI_GATED, true>(nt, tw, tx, a, grid, s);
        return launch_skinny_fp8_nt<EPI_RED_F32, true>(nt, tw, tx, a, grid, s);
    }
    if (epilogue == EPI_STORE) return launch_skinny_fp8_nt<EPI_STORE, false>(nt, tw, tx, a, grid, s);
    if (epilogue == EPI_GATED) return launch_skinny_fp8_nt<EPI_GATED, false>(nt, tw, tx, a, grid, s);
    return launch_skinny_fp8_nt<EPI_RED_F32, false>(nt, tw, tx, a, grid, s);
}
