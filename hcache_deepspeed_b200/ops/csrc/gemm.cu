// Launchers for the tcgen05 GEMM family (see gemm_core.cuh for the kernel).
#include "gemm_core.cuh"
#include "hds_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

template <int BN, int EPI, bool kBf16>
static int launch_gemm_inst(const CUtensorMap& ta, const CUtensorMap& tw, const GemmArgs& args, int grid,
                            cudaStream_t stream) {
    auto kern = gemm_tn_kernel<BN, EPI, kBf16>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(gemm_smem_bytes(BN))));
        configured[dev & 63] = true;
    }
    kern<<<grid, kGemmThreads, gemm_smem_bytes(BN), stream>>>(ta, tw, args);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

template <int EPI, bool kBf16>
static int launch_gemm_bn(int bn, const CUtensorMap& ta, const CUtensorMap& tw, const GemmArgs& args, int grid,
                          cudaStream_t stream) {
    if constexpr (EPI == EPI_KV_ROPE || EPI == EPI_QKV_ROPE) {
        switch (bn) {
            case 64: return launch_gemm_inst<64, EPI, kBf16>(ta, tw, args, grid, stream);
            case 128: return launch_gemm_inst<128, EPI, kBf16>(ta, tw, args, grid, stream);
            default: set_error("fused RoPE/KV epilogue supports head_dim 64 or 128"); return HDS_ERR_UNSUPPORTED;
        }
    } else {
        switch (bn) {
            case 32: return launch_gemm_inst<32, EPI, kBf16>(ta, tw, args, grid, stream);
            case 64: return launch_gemm_inst<64, EPI, kBf16>(ta, tw, args, grid, stream);
            case 128: return launch_gemm_inst<128, EPI, kBf16>(ta, tw, args, grid, stream);
            case 256: return launch_gemm_inst<256, EPI, kBf16>(ta, tw, args, grid, stream);
            default: set_error("tile N must be one of 32/64/128/256"); return HDS_ERR_BAD_ARG;
        }
    }
}

// Tile-width heuristic: cost ~ waves * BN (each CTA streams BN*K weight bytes per tile); ties go to
// the wider tile (fewer activation re-reads).
static int pick_bn(int M, int N, int sms, bool allow_256) {
    const int m_tiles = (M + kGemmBM - 1) / kGemmBM;
    const int cands[4] = {256, 128, 64, 32};
    int best = 128;
    long best_cost = -1;
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        if (bn == 256 && (!allow_256 || m_tiles < 2)) continue;
        if (bn > 32 && bn / 2 >= N) continue;
        const long tiles = static_cast<long>(m_tiles) * ((N + bn - 1) / bn);
        const long waves = (tiles + sms - 1) / sms;
        const long cost = waves * bn;
        if (best_cost < 0 || cost < best_cost) {
            best_cost = cost;
            best = bn;
        }
    }
    return best;
}

struct Operands {
    CUtensorMap ta, tw;
};

static int prepare(GemmArgs& g, Operands& ops, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N,
                   int K, bool bf16, int bn, int splits) {
    HDS_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem");
    HDS_REQUIRE(K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0, "gemm: K, lda, ldw must be multiples of 8 elements");
    g.M = M;
    g.N = N;
    g.K = K;
    const int kb_total = (K + kGemmBK - 1) / kGemmBK;
    if (splits < 1) splits = 1;
    if (splits > kb_total) splits = kb_total;
    int per = (kb_total + splits - 1) / splits;
    splits = (kb_total + per - 1) / per;
    g.splits = splits;
    g.kb_per_split = per;
    // A box covers only the rows that exist (rounded to 8) when there is a single m-tile: small-M
    // (decode) GEMMs then move T*128 B instead of 16 KB of zero fill per k-block.
    int box_rows = kGemmBM;
    if (M < kGemmBM) box_rows = ((M + 7) / 8) * 8;
    g.a_tx_bytes = static_cast<uint32_t>(box_rows) * kGemmBK * 2;
    int rc = make_tmap_2d_16b(&ops.ta, A, static_cast<uint64_t>(K), static_cast<uint64_t>(M),
                              static_cast<uint64_t>(lda) * 2, kGemmBK, box_rows, bf16);
    if (rc) return rc;
    rc = make_tmap_2d_16b(&ops.tw, W, static_cast<uint64_t>(K), static_cast<uint64_t>(N),
                          static_cast<uint64_t>(ldw) * 2, kGemmBK, bn, bf16);
    return rc;
}

static int grid_for(const GemmArgs& g, int bn) {
    const long work = static_cast<long>((g.M + kGemmBM - 1) / kGemmBM) * ((g.N + bn - 1) / bn) * g.splits;
    const int sms = sm_count();
    return static_cast<int>(work < sms ? work : sms);
}

}  // namespace hds

using namespace hds;

extern "C" const char* hds_last_error(void) { return g_last_error; }

extern "C" int hds_gemm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int dtype,
                        int epilogue, int act, const void* bias, void* out, int64_t ldo, int bn, void* stream) {
    HDS_REQUIRE(epilogue == EPI_STORE || epilogue == EPI_GATED, "hds_gemm: epilogue must be 0 (store) or 1 (gated)");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_gemm: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(epilogue != EPI_GATED || (N % 2 == 0), "hds_gemm: gated epilogue needs an even N");
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    Operands ops;
    if (bn <= 0) bn = pick_bn(M, N, sm_count(), true);
    int rc = prepare(g, ops, A, lda, W, ldw, M, N, K, dtype == 0, bn, 1);
    if (rc) return rc;
    g.act = act;
    g.out = out;
    g.ldo = ldo;
    g.bias = bias;
    const int grid = grid_for(g, bn);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype == 0) {
        return epilogue == EPI_STORE ? launch_gemm_bn<EPI_STORE, true>(bn, ops.ta, ops.tw, g, grid, s)
                                     : launch_gemm_bn<EPI_GATED, true>(bn, ops.ta, ops.tw, g, grid, s);
    }
    return epilogue == EPI_STORE ? launch_gemm_bn<EPI_STORE, false>(bn, ops.ta, ops.tw, g, grid, s)
                                 : launch_gemm_bn<EPI_GATED, false>(bn, ops.ta, ops.tw, g, grid, s);
}

extern "C" int hds_gemm_accum(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int dtype,
                              const void* bias, float* const* accum_ptrs, int n_targets, int red_mode,
                              int64_t ld_accum, int bn, int splits, void* stream) {
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_gemm_accum: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(n_targets >= 1 && n_targets <= kMaxPeers, "hds_gemm_accum: 1..8 accumulation targets");
    HDS_REQUIRE(red_mode >= 0 && red_mode <= 2, "hds_gemm_accum: red_mode 0 (local) / 1 (p2p) / 2 (multicast)");
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    Operands ops;
    const int sms = sm_count();
    if (bn <= 0) bn = M > kGemmBM ? pick_bn(M, N, sms, true) : (N >= 2048 ? 64 : 32);
    if (splits <= 0) {
        const long tiles = static_cast<long>((M + kGemmBM - 1) / kGemmBM) * ((N + bn - 1) / bn);
        splits = static_cast<int>(sms / (tiles > 0 ? tiles : 1));
        const int kb_total = (K + kGemmBK - 1) / kGemmBK;
        if (splits > kb_total / 4) splits = kb_total / 4;  // keep >= 4 k-blocks (256 of K) per split
        if (splits < 1) splits = 1;
    }
    int rc = prepare(g, ops, A, lda, W, ldw, M, N, K, dtype == 0, bn, splits);
    if (rc) return rc;
    g.bias = bias;
    for (int i = 0; i < n_targets; ++i) g.accum[i] = accum_ptrs[i];
    g.n_targets = n_targets;
    g.red_mode = red_mode;
    g.ld_accum = ld_accum;
    const int grid = grid_for(g, bn);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    return dtype == 0 ? launch_gemm_bn<EPI_RED_F32, true>(bn, ops.ta, ops.tw, g, grid, s)
                      : launch_gemm_bn<EPI_RED_F32, false>(bn, ops.ta, ops.tw, g, grid, s);
}

static int gemm_rope_common(bool has_q, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int K,
                            int dtype, const void* bias, void* q_out, void* kv_cache_layer, const int32_t* tok_pos,
                            const int32_t* tok_slot, const float* inv_freq, int n_q_heads, int n_kv_heads,
                            int head_dim, int block_size, int do_rope, void* stream) {
    HDS_REQUIRE(dtype == 0 || dtype == 1, "gemm_rope: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(head_dim == 64 || head_dim == 128, "gemm_rope: fused epilogue supports head_dim 64 / 128");
    HDS_REQUIRE(n_kv_heads > 0 && block_size > 0, "gemm_rope: bad head/block geometry");
    HDS_REQUIRE(!do_rope || inv_freq != nullptr, "gemm_rope: inv_freq required when do_rope");
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    Operands ops;
    const int N = ((has_q ? n_q_heads : 0) + 2 * n_kv_heads) * head_dim;
    int rc = prepare(g, ops, A, lda, W, ldw, M, N, K, dtype == 0, head_dim, 1);
    if (rc) return rc;
    g.bias = bias;
    g.q_out = q_out;
    g.kv_cache = kv_cache_layer;
    g.tok_pos = tok_pos;
    g.tok_slot = tok_slot;
    g.inv_freq = inv_freq;
    g.n_q_heads = has_q ? n_q_heads : 0;
    g.n_kv_heads = n_kv_heads;
    g.block_size = block_size;
    g.do_rope = do_rope;
    const int grid = grid_for(g, head_dim);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (has_q)
        return dtype == 0 ? launch_gemm_bn<EPI_QKV_ROPE, true>(head_dim, ops.ta, ops.tw, g, grid, s)
                          : launch_gemm_bn<EPI_QKV_ROPE, false>(head_dim, ops.ta, ops.tw, g, grid, s);
    return dtype == 0 ? launch_gemm_bn<EPI_KV_ROPE, true>(head_dim, ops.ta, ops.tw, g, grid, s)
                      : launch_gemm_bn<EPI_KV_ROPE, false>(head_dim, ops.ta, ops.tw, g, grid, s);
}

extern "C" int hds_restore_kv_proj(const void* latents, int64_t ld_latents, const void* w_kv, int64_t ldw, int T,
                                   int H, int dtype, const void* bias_kv, void* kv_cache_layer,
                                   const int32_t* tok_pos, const int32_t* tok_slot, const float* inv_freq,
                                   int n_kv_heads, int head_dim, int block_size, int do_rope, void* stream) {
    return gemm_rope_common(false, latents, ld_latents, w_kv, ldw, T, H, dtype, bias_kv, nullptr, kv_cache_layer,
                            tok_pos, tok_slot, inv_freq, 0, n_kv_heads, head_dim, block_size, do_rope, stream);
}

extern "C" int hds_qkv_rope_proj(const void* x, int64_t ldx, const void* w_qkv, int64_t ldw, int T, int H, int dtype,
                                 const void* bias_qkv, void* q_out, void* kv_cache_layer, const int32_t* tok_pos,
                                 const int32_t* tok_slot, const float* inv_freq, int n_q_heads, int n_kv_heads,
                                 int head_dim, int block_size, int do_rope, void* stream) {
    HDS_REQUIRE(q_out != nullptr && n_q_heads > 0, "hds_qkv_rope_proj: q_out and n_q_heads required");
    return gemm_rope_common(true, x, ldx, w_qkv, ldw, T, H, dtype, bias_qkv, q_out, kv_cache_layer, tok_pos,
                            tok_slot, inv_freq, n_q_heads, n_kv_heads, head_dim, block_size, do_rope, stream);
}
