// Launchers for the tcgen05 GEMM family (see gemm_core.cuh for the kernel).
#include <stdlib.h>
#include "gemm_core.cuh"
#include "gemm_skinny.cuh"
#include "hds_host.cuh"
#include "tp_sync_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

static thread_local const TmapSet* g_peer_maps = nullptr;  // set by the peer-sharded entry point around its launch
static thread_local int g_peer_rows = 0, g_peer_count = 0;

template <int BN, int EPI, bool kBf16, int HD = 128, int AM = 128>
static int launch_gemm_inst(const CUtensorMap& ta, const CUtensorMap& tw, const GemmArgs& args, int grid,
                            cudaStream_t stream) {
    static const TmapSet kNoPeers = {};
    auto kern = gemm_tn_kernel<BN, EPI, kBf16, HD, AM>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(gemm_smem_bytes(BN, AM))));
        configured[dev & 63] = true;
    }
    HDS_CHECK_CUDA(launch_kernel(kern, dim3(grid), dim3(kGemmThreads), gemm_smem_bytes(BN, AM), stream, ta, tw, args,
                                 g_peer_maps ? *g_peer_maps : kNoPeers));
    return HDS_OK;
}

// fused RoPE/KV epilogues: tile = BN / HD whole heads
template <int EPI, bool kBf16>
static int launch_gemm_rope(int bn, int hd, const CUtensorMap& ta, const CUtensorMap& tw, const GemmArgs& args,
                            int grid, cudaStream_t stream) {
    if (hd == 128 && bn == 128) return launch_gemm_inst<128, EPI, kBf16, 128>(ta, tw, args, grid, stream);
    if (hd == 128 && bn == 256) return launch_gemm_inst<256, EPI, kBf16, 128>(ta, tw, args, grid, stream);
    if (hd == 64 && bn == 64) return launch_gemm_inst<64, EPI, kBf16, 64>(ta, tw, args, grid, stream);
    if (hd == 64 && bn == 128) return launch_gemm_inst<128, EPI, kBf16, 64>(ta, tw, args, grid, stream);
    if (hd == 64 && bn == 256) return launch_gemm_inst<256, EPI, kBf16, 64>(ta, tw, args, grid, stream);
    set_error("fused RoPE/KV epilogue supports head_dim 64 or 128");
    return HDS_ERR_UNSUPPORTED;
}

template <int EPI, bool kBf16>
static int launch_gemm_bn(int bn, const CUtensorMap& ta, const CUtensorMap& tw, const GemmArgs& args, int grid,
                          cudaStream_t stream) {
    // decode-size M: compact A stages (16 or 64 rows) and a deeper weight pipeline
    if (args.M <= 64 && (bn == 64 || bn == 128)) {
        if (args.M <= 16) {
            return bn == 64 ? launch_gemm_inst<64, EPI, kBf16, 128, 16>(ta, tw, args, grid, stream)
                            : launch_gemm_inst<128, EPI, kBf16, 128, 16>(ta, tw, args, grid, stream);
        }
        return bn == 64 ? launch_gemm_inst<64, EPI, kBf16, 128, 64>(ta, tw, args, grid, stream)
                        : launch_gemm_inst<128, EPI, kBf16, 128, 64>(ta, tw, args, grid, stream);
    }
    {
        switch (bn) {
            case 32: return launch_gemm_inst<32, EPI, kBf16>(ta, tw, args, grid, stream);
            case 64: return launch_gemm_inst<64, EPI, kBf16>(ta, tw, args, grid, stream);
            case 128: return launch_gemm_inst<128, EPI, kBf16>(ta, tw, args, grid, stream);
            case 256: return launch_gemm_inst<256, EPI, kBf16>(ta, tw, args, grid, stream);
            default: set_error("tile N must be one of 32/64/128/256"); return HDS_ERR_BAD_ARG;
        }
    }
}

// Tile / split heuristic. A CTA streams (bn + a_rows) * 128 B per k-block; the critical path is
// waves * k-blocks-per-unit of that, plus a per-unit epilogue term. Splitting K only pays when the
// un-split tile count leaves SMs idle (decode-size M); ties go to wider tiles and fewer splits.
struct TileCfg {
    int bn, splits;
};
static TileCfg pick_config(int M, int N, int K, int sms, bool allow_split, int force_bn) {
    const int m_tiles = (M + kGemmBM - 1) / kGemmBM;
    const int kb_total = (K + kGemmBK - 1) / kGemmBK;
    const int a_rows = M < kGemmBM ? ((M + 7) / 8) * 8 : kGemmBM;
    const int cands[4] = {256, 128, 64, 32};
    TileCfg best{128, 1};
    double best_cost = -1.0;
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        if (force_bn > 0 && bn != force_bn) continue;
        if (force_bn <= 0) {
            if (bn == 256 && m_tiles < 2) continue;
            if (bn == 32 && N > 64) continue;  // too little in flight per SM to stream well
            if (bn > 32 && bn / 2 >= N) continue;
        }
        const long tiles = static_cast<long>(m_tiles) * ((N + bn - 1) / bn);
        // measured on B200: splitting K only wins when the un-split tiles cover less than half of the SMs
        // (e.g. N=4096: 32 tiles -> 4 splits is 1.6x faster; N=12288: 96 tiles -> any split is slower)
        int max_splits = 1;
        if (allow_split && m_tiles == 1 && tiles * 2 <= sms) {
            max_splits = static_cast<int>(sms / tiles);
            if (max_splits > kb_total / 4) max_splits = kb_total / 4;
            if (max_splits > 16) max_splits = 16;
            if (max_splits < 1) max_splits = 1;
        }
        for (int sp = 1; sp <= max_splits; ++sp) {
            const int per = (kb_total + sp - 1) / sp;
            const int eff_sp = (kb_total + per - 1) / per;
            const long units = tiles * eff_sp;
            const long waves = (units + sms - 1) / sms;
            // per-unit fixed cost ~ 6 k-blocks' worth of time (pipeline hand-off + epilogue); split units
            // also pay the fp32 partial-sum traffic
            const double unit_cost = per * (bn + a_rows) + 6.0 * bn + (eff_sp > 1 ? 2.0 * bn : 0.0);
            const double cost = waves * unit_cost;
            if (best_cost < 0 || cost < best_cost * 0.97) {
                best_cost = cost;
                best = TileCfg{bn, eff_sp};
            }
        }
    }
    return best;
}

struct Operands {
    CUtensorMap ta, tw;
};

// ---- prefill-size M: CTA-pair kernel (gemm_2sm.cu) ------------------------------------------------------------
int gemm_2sm_dispatch(const void* A, int64_t lda_bytes, const float* a_scale, const void* W, int64_t ldw_bytes,
                      const float* w_scale, int M, int N, int K, bool fp8, bool bf16, int epilogue, int act, const void* bias,
                      void* out, int64_t ldo, int bn, cudaStream_t stream, int w_fmt = 0);
int gemm_2sm_dispatch_rope(const GemmArgs& g, const void* A, int64_t lda_bytes, const void* W, int64_t ldw_bytes, bool bf16,
                           bool has_q, cudaStream_t stream);
bool is_local_device_ptr(const void* p);
static bool pair_enabled() {
    static const int on = getenv("HDS_GEMM_2SM") ? atoi(getenv("HDS_GEMM_2SM")) : 1;
    return on != 0;
}

// ---- decode-size M: feature-major weight-streaming kernel (gemm_skinny.cuh) -------------------------------
static inline int skinny_nt(int M) { return M <= 16 ? 16 : M <= 32 ? 32 : 64; }
static bool skinny_enabled() {
    static const int on = getenv("HDS_GEMM_SKINNY") ? atoi(getenv("HDS_GEMM_SKINNY")) : 1;
    return on != 0;
}

template <int NT, int EPI, bool kBf16>
static int launch_skinny_inst(const CUtensorMap& tw, const CUtensorMap& tx, const SkinnyArgs& a, int grid,
                              cudaStream_t stream) {
    auto kern = gemm_skinny_kernel<NT, EPI, kBf16>;
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
static int launch_skinny_nt(int nt, const CUtensorMap& tw, const CUtensorMap& tx, const SkinnyArgs& a, int grid,
                            cudaStream_t stream) {
    if (nt == 16) return launch_skinny_inst<16, EPI, kBf16>(tw, tx, a, grid, stream);
    if (nt == 32) return launch_skinny_inst<32, EPI, kBf16>(tw, tx, a, grid, stream);
    return launch_skinny_inst<64, EPI, kBf16>(tw, tx, a, grid, stream);
}

// One CTA per (128-feature tile, K split); splits fill the resident slots (SMs x CTAs per SM) once.
static int launch_skinny(int epi, bool bf16, const void* X, int64_t ldx, const void* W, int64_t ldw, int M, int N,
                         int K, int act, const void* bias, void* out, int64_t ldo, float* accum, int64_t ld_accum,
                         void* ws, int64_t ws_bytes, int want_splits, cudaStream_t stream) {
    HDS_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem");
    HDS_REQUIRE(K % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "gemm: K, lda, ldw must be multiples of 8 elements");
    const int nt = skinny_nt(M);
    const int n_tiles = (N + kSkinnyBF - 1) / kSkinnyBF;
    const int kb_total = (K + kGemmBK - 1) / kGemmBK;
    const int64_t ws_data = (static_cast<int64_t>(n_tiles) * kSkinnyBF * nt * 4 + 255) & ~int64_t(255);
    const bool can_split = epi == EPI_RED_F32 || (ws != nullptr && ws_bytes >= ws_data + static_cast<int64_t>(n_tiles) * 16);
    int splits = 1;
    if (can_split) {
        if (want_splits > 0) {
            splits = want_splits;
        } else {
            // measured (profiles/kernel_bench_decode_gemm.jsonl, BIG_SHAPES run): fill the resident slots once; past one full
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
    a.sync = tp_take_armed();
    if (splits > 1 && epi != EPI_RED_F32) {
        a.ws = reinterpret_cast<float*>(ws);
        a.tickets = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(ws) + ws_data);
    }
    CUtensorMap tw, tx;
    int rc = make_tmap_2d_16b(&tw, W, static_cast<uint64_t>(K), static_cast<uint64_t>(N),
                              static_cast<uint64_t>(ldw) * 2, kGemmBK, kSkinnyBF, bf16);
    if (rc) return rc;
    rc = make_tmap_2d_16b(&tx, X, static_cast<uint64_t>(K), static_cast<uint64_t>(M), static_cast<uint64_t>(ldx) * 2,
                          kGemmBK, nt, bf16);
    if (rc) return rc;
    const int grid = n_tiles * splits;
    if (bf16) {
        if (epi == EPI_STORE) return launch_skinny_nt<EPI_STORE, true>(nt, tw, tx, a, grid, stream);
        if (epi == EPI_GATED) return launch_skinny_nt<EPI_GATED, true>(nt, tw, tx, a, grid, stream);
        return launch_skinny_nt<EPI_RED_F32, true>(nt, tw, tx, a, grid, stream);
    }
    if (epi == EPI_STORE) return launch_skinny_nt<EPI_STORE, false>(nt, tw, tx, a, grid, stream);
    if (epi == EPI_GATED) return launch_skinny_nt<EPI_GATED, false>(nt, tw, tx, a, grid, stream);
    return launch_skinny_nt<EPI_RED_F32, false>(nt, tw, tx, a, grid, stream);
}

static int prepare(GemmArgs& g, Operands& ops, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N,
                   int K, bool bf16, int bn, int splits) {
    HDS_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem");
    HDS_REQUIRE(K % 8 == 0 && lda % 8 == 0 && (ldw < 0 || ldw % 8 == 0),
                "gemm: K, lda, ldw must be multiples of 8 elements");
    HDS_REQUIRE(ldw >= 0 || bn >= 32, "gemm: tiled weights need a tile width >= 32");
    g.M = M;
    g.N = N;
    g.K = K;
    static const int debug_skip = getenv("HDS_GEMM_DEBUG") ? atoi(getenv("HDS_GEMM_DEBUG")) : 0;
    g.debug_skip = debug_skip;
    g.a_peers = g_peer_count;
    g.a_rows_per_peer = g_peer_rows;
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
    if (ldw < 0) {
        // pre-tiled weights: [ceil(N/32)][ceil(K/64)][32][64] — one TMA instruction per k-block fetches BN/32
        // contiguous 4 KB chunks (one per 32-row strip)
        g.w_tiled = 1;
        const uint64_t strips = (static_cast<uint64_t>(N) + 31) / 32;
        rc = make_tmap_3d_16b(&ops.tw, W, 64, static_cast<uint64_t>(kb_total) * 32, strips, 128,
                              static_cast<uint64_t>(kb_total) * 4096, 64, 32, static_cast<uint32_t>(bn / 32), bf16);
    } else {
        rc = make_tmap_2d_16b(&ops.tw, W, static_cast<uint64_t>(K), static_cast<uint64_t>(N),
                              static_cast<uint64_t>(ldw) * 2, kGemmBK, bn, bf16);
    }
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

static inline int64_t align256(int64_t x) { return (x + 255) & ~int64_t(255); }

extern "C" int64_t hds_gemm_splitk_workspace_bytes(int M, int N) {
    // fp32 partial sums [M, N] + tickets for up to N/32 tiles x 4 quadrants
    // (the feature-major kernel for M <= 64 pads rows to its token tile and N to 128)
    const int64_t rows = M <= 64 ? skinny_nt(M) : M;
    const int64_t cols = (static_cast<int64_t>(N) + 127) / 128 * 128;
    return align256(rows * cols * 4) + align256((static_cast<int64_t>(N) / 32 + 1) * 4 * 4);
}

extern "C" int hds_gemm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int dtype,
                        int epilogue, int act, const void* bias, void* out, int64_t ldo, int bn, int splits,
                        void* splitk_ws, int64_t splitk_ws_bytes, void* stream) {
    HDS_REQUIRE(epilogue == EPI_STORE || epilogue == EPI_GATED, "hds_gemm: epilogue must be 0 (store) or 1 (gated)");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_gemm: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(epilogue != EPI_GATED || (N % 2 == 0), "hds_gemm: gated epilogue needs an even N");
    // measured (profiles/kernel_bench.jsonl): the feature-major kernel wins for M <= 16 everywhere and for the gated
    // MLP-1 up to 64 tokens (1.3-1.5x); plain stores at 32-64 tokens stay on the token-major kernel
    // (a tensor-parallel rank's MLP-1 is narrow — N = 2752 for Llama-2-7B at TP=8: with 17-64 tokens the token-major kernel is
    // 1.3x faster there, gpurun_out/tp_rank_gemm_probe.jsonl — so the wide-token feature-major form needs N >= 8192)
    if (M <= 64 && ldw >= 0 &&
        (bn == -1 || (bn == 0 && skinny_enabled() && (M <= 16 || (epilogue == EPI_GATED && N >= 8192)))))
        return launch_skinny(epilogue, dtype == 0, A, lda, W, ldw, M, N, K, act, bias, out, ldo, nullptr, 0, splitk_ws,
                             splitk_ws_bytes, splits, static_cast<cudaStream_t>(stream));
    // measured (profiles/kernel_bench_gemm_2sm.txt): with at least one full wave of 256 x 256 pair tiles the
    // cta_group::2 kernel is 1.10-1.20x the single-CTA kernel and level with cuBLAS
    if (bn == 0 && M > 128 && ldw >= 0 && pair_enabled() && lda % 8 == 0 && ldw % 8 == 0 && K % 8 == 0 &&
        static_cast<int64_t>((M + 255) / 256) * ((N + 255) / 256) * 2 >= sm_count() && is_local_device_ptr(A) &&
        is_local_device_ptr(W))
        return hds::gemm_2sm_dispatch(A, lda * 2, nullptr, W, ldw * 2, nullptr, M, N, K, false, dtype == 0, epilogue, act, bias, out,
                                 ldo, 256, static_cast<cudaStream_t>(stream));
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    Operands ops;
    const bool can_split = splitk_ws != nullptr && M <= kGemmBM &&
                           splitk_ws_bytes >= hds_gemm_splitk_workspace_bytes(M, N);
    if (splits <= 0 || bn <= 0) {
        const TileCfg cfg = pick_config(M, N, K, sm_count(), can_split && splits <= 0, bn);
        if (bn <= 0) bn = cfg.bn;
        if (splits <= 0) splits = cfg.splits;
    }
    if (!can_split) splits = 1;
    int rc = prepare(g, ops, A, lda, W, ldw, M, N, K, dtype == 0, bn, splits);
    if (rc) return rc;
    g.act = act;
    g.out = out;
    g.ldo = ldo;
    g.bias = bias;
    g.sync = tp_take_armed();
    if (g.splits > 1) {
        g.splitk_ws = reinterpret_cast<float*>(splitk_ws);
        g.splitk_tickets = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(splitk_ws) +
                                                      align256(static_cast<int64_t>(M) * N * 4));
    }
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
    if (M <= 64 && ldw >= 0 && red_mode == 0 && n_targets == 1 && (bn == -1 || (bn == 0 && skinny_enabled() && M <= 16)))
        return launch_skinny(EPI_RED_F32, dtype == 0, A, lda, W, ldw, M, N, K, 0, bias, nullptr, 0, accum_ptrs[0],
                             ld_accum, nullptr, 0, splits, static_cast<cudaStream_t>(stream));
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    Operands ops;
    if (splits <= 0 || bn <= 0) {
        const TileCfg cfg = pick_config(M, N, K, sm_count(), splits <= 0, bn);
        if (bn <= 0) bn = cfg.bn;
        if (splits <= 0) splits = cfg.splits;
    }
    int rc = prepare(g, ops, A, lda, W, ldw, M, N, K, dtype == 0, bn, splits);
    if (rc) return rc;
    g.bias = bias;
    for (int i = 0; i < n_targets; ++i) g.accum[i] = accum_ptrs[i];
    g.n_targets = n_targets;
    g.red_mode = red_mode;
    g.ld_accum = ld_accum;
    g.sync = tp_take_armed();
    const int grid = grid_for(g, bn);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    return dtype == 0 ? launch_gemm_bn<EPI_RED_F32, true>(bn, ops.ta, ops.tw, g, grid, s)
                      : launch_gemm_bn<EPI_RED_F32, false>(bn, ops.ta, ops.tw, g, grid, s);
}

/* Grouped GEMM over experts: rows [pad_off[e], pad_off[e+1]) of A use W[e*N : (e+1)*N]. */
extern "C" int hds_moe_gemm(const void* A, int64_t lda, const void* W_stacked, int64_t ldw, const int32_t* pad_off,
                            int n_experts, int max_rows, int N, int K, int dtype, int epilogue, int act, void* out,
                            int64_t ldo, void* stream) {
    HDS_REQUIRE(epilogue == EPI_STORE || epilogue == EPI_GATED, "hds_moe_gemm: epilogue must be 0 (store) or 1 (gated)");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "hds_moe_gemm: dtype must be 0 (bf16) or 1 (fp16)");
    HDS_REQUIRE(n_experts >= 1 && pad_off != nullptr && max_rows % kGemmBM == 0,
                "hds_moe_gemm: needs expert offsets and a row capacity that is a multiple of 128");
    HDS_REQUIRE(epilogue != EPI_GATED || (N % 2 == 0), "hds_moe_gemm: gated epilogue needs an even N");
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    Operands ops;
    const int bn = N >= 256 ? 256 : 128;
    int rc = prepare(g, ops, A, lda, W_stacked, ldw, max_rows, N, K, dtype == 0, bn, 1);
    if (rc) return rc;
    // the weight map spans every expert: n_experts * N rows
    rc = make_tmap_2d_16b(&ops.tw, W_stacked, static_cast<uint64_t>(K), static_cast<uint64_t>(N) * n_experts,
                          static_cast<uint64_t>(ldw) * 2, kGemmBK, bn, dtype == 0);
    if (rc) return rc;
    g.act = act;
    g.out = out;
    g.ldo = ldo;
    g.group_pad_off = pad_off;
    g.n_groups = n_experts;
    g.sync = tp_take_armed();
    const int grid = grid_for(g, bn);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype == 0) {
        return epilogue == EPI_STORE ? (bn == 128 ? launch_gemm_inst<128, EPI_STORE, true>(ops.ta, ops.tw, g, grid, s)
                                                  : launch_gemm_inst<256, EPI_STORE, true>(ops.ta, ops.tw, g, grid, s))
                                     : (bn == 128 ? launch_gemm_inst<128, EPI_GATED, true>(ops.ta, ops.tw, g, grid, s)
                                                  : launch_gemm_inst<256, EPI_GATED, true>(ops.ta, ops.tw, g, grid, s));
    }
    return epilogue == EPI_STORE ? (bn == 128 ? launch_gemm_inst<128, EPI_STORE, false>(ops.ta, ops.tw, g, grid, s)
                                              : launch_gemm_inst<256, EPI_STORE, false>(ops.ta, ops.tw, g, grid, s))
                                 : (bn == 128 ? launch_gemm_inst<128, EPI_GATED, false>(ops.ta, ops.tw, g, grid, s)
                                              : launch_gemm_inst<256, EPI_GATED, false>(ops.ta, ops.tw, g, grid, s));
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
    // two (head_dim 128) or four (head_dim 64) heads per 256-wide tile once there is enough work to
    // fill the machine with the wider tiles; otherwise one head per tile
    const long m_tiles = (M + kGemmBM - 1) / kGemmBM;
    int bn = head_dim;
    if (m_tiles * ((N + 255) / 256) >= sm_count()) bn = 256;
    // prefill-size problems with 128-wide heads: the CTA-pair kernel (same epilogue, M = 256 tiles over two SMs)
    const bool pair = head_dim == 128 && M > 128 && ldw >= 0 && g_peer_maps == nullptr && pair_enabled() &&
                      lda % 8 == 0 && ldw % 8 == 0 && K % 8 == 0 &&
                      static_cast<int64_t>((M + 255) / 256) * ((N + 255) / 256) * 2 >= sm_count() &&
                      is_local_device_ptr(A) && is_local_device_ptr(W);
    int rc = 0;
    if (pair) {
        g.M = M;
        g.N = N;
        g.K = K;
    } else {
        rc = prepare(g, ops, A, lda, W, ldw, M, N, K, dtype == 0, bn, 1);
    }
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
    g.sync = tp_take_armed();
    if (pair)
        return gemm_2sm_dispatch_rope(g, A, lda * 2, W, ldw * 2, dtype == 0, has_q, static_cast<cudaStream_t>(stream));
    const int grid = grid_for(g, bn);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (has_q)
        return dtype == 0 ? launch_gemm_rope<EPI_QKV_ROPE, true>(bn, head_dim, ops.ta, ops.tw, g, grid, s)
                          : launch_gemm_rope<EPI_QKV_ROPE, false>(bn, head_dim, ops.ta, ops.tw, g, grid, s);
    return dtype == 0 ? launch_gemm_rope<EPI_KV_ROPE, true>(bn, head_dim, ops.ta, ops.tw, g, grid, s)
                      : launch_gemm_rope<EPI_KV_ROPE, false>(bn, head_dim, ops.ta, ops.tw, g, grid, s);
}

extern "C" int hds_restore_kv_proj(const void* latents, int64_t ld_latents, const void* w_kv, int64_t ldw, int T,
                                   int H, int dtype, const void* bias_kv, void* kv_cache_layer,
                                   const int32_t* tok_pos, const int32_t* tok_slot, const float* inv_freq,
                                   int n_kv_heads, int head_dim, int block_size, int do_rope, void* stream) {
    return gemm_rope_common(false, latents, ld_latents, w_kv, ldw, T, H, dtype, bias_kv, nullptr, kv_cache_layer,
                            tok_pos, tok_slot, inv_freq, 0, n_kv_heads, head_dim, block_size, do_rope, stream);
}

extern "C" int hds_restore_kv_proj_peers(const void* const* latent_ptrs, int n_peers, int rows_per_peer,
                                         int64_t ld_latents, const void* w_kv, int64_t ldw, int T, int H, int dtype,
                                         const void* bias_kv, void* kv_cache_layer, const int32_t* tok_pos,
                                         const int32_t* tok_slot, const float* inv_freq, int n_kv_heads, int head_dim,
                                         int block_size, int do_rope, void* stream) {
    HDS_REQUIRE(n_peers >= 1 && n_peers <= kMaxPeers, "restore_kv_proj_peers: 1..8 peers");
    HDS_REQUIRE(rows_per_peer > 0 && rows_per_peer % kGemmBM == 0, "restore_kv_proj_peers: rows_per_peer must be a multiple of 128");
    HDS_REQUIRE(static_cast<int64_t>(n_peers) * rows_per_peer >= T, "restore_kv_proj_peers: peers do not cover T rows");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "restore_kv_proj_peers: dtype must be 0 (bf16) or 1 (fp16)");
    TmapSet set;
    memset(&set, 0, sizeof(set));
    for (int p = 0; p < n_peers; ++p) {
        const int64_t first = static_cast<int64_t>(p) * rows_per_peer;
        int64_t rows = T - first;
        if (rows > rows_per_peer) rows = rows_per_peer;
        if (rows <= 0) rows = 1;  // a peer without rows is never addressed; keep the descriptor valid
        const uint32_t box_rows = T < kGemmBM ? static_cast<uint32_t>(((T + 7) / 8) * 8) : kGemmBM;  // as prepare()
        int rc = make_tmap_2d_16b(&set.m[p], latent_ptrs[p], static_cast<uint64_t>(H), static_cast<uint64_t>(rows),
                                  static_cast<uint64_t>(ld_latents) * 2, kGemmBK, box_rows, dtype == 0);
        if (rc) return rc;
    }
    g_peer_maps = &set;
    g_peer_rows = rows_per_peer;
    g_peer_count = n_peers;
    const int rc = gemm_rope_common(false, latent_ptrs[0], ld_latents, w_kv, ldw, T, H, dtype, bias_kv, nullptr,
                                    kv_cache_layer, tok_pos, tok_slot, inv_freq, 0, n_kv_heads, head_dim, block_size,
                                    do_rope, stream);
    g_peer_maps = nullptr;
    g_peer_count = 0;
    return rc;
}

extern "C" int hds_qkv_rope_proj(const void* x, int64_t ldx, const void* w_qkv, int64_t ldw, int T, int H, int dtype,
                                 const void* bias_qkv, void* q_out, void* kv_cache_layer, const int32_t* tok_pos,
                                 const int32_t* tok_slot, const float* inv_freq, int n_q_heads, int n_kv_heads,
                                 int head_dim, int block_size, int do_rope, void* stream) {
    HDS_REQUIRE(q_out != nullptr && n_q_heads > 0, "hds_qkv_rope_proj: q_out and n_q_heads required");
    return gemm_rope_common(true, x, ldx, w_qkv, ldw, T, H, dtype, bias_qkv, q_out, kv_cache_layer, tok_pos,
                            tok_slot, inv_freq, n_q_heads, n_kv_heads, head_dim, block_size, do_rope, stream);
}
