// Split-KV ragged paged attention for work items with few query rows (decode and short
// continuations): memory-bound, so the design goal is HBM streaming rate, not FLOPs.
//
//   * K/V pages of one KV head are gathered by TMA (cp.async.bulk.tensor, SWIZZLE_128B) into a 3-stage
//     shared-memory ring of 128-token tiles; one elected thread walks the block table.
//   * The 128B swizzle stores 16-byte chunk c of row j at chunk (c ^ (j & 7)), so "one thread per KV
//     token" reads of K and "eight lanes per KV token" reads of V are bank-conflict free while every
//     lane uses the same (compile-time) logical chunk index — q can stay broadcast in shared memory.
//   * KV range is split over gridDim.z CTAs; partial (O, m, l) go to a scratch buffer and the last CTA
//     to finish (atomic ticket) merges them — one launch, no separate combine kernel.
//
// Stands in for the external paged flash kernel the reference links
// (deepspeed/inference/v2/kernels/ragged_ops/blocked_flash/blocked_flash.cpp:21-101, flash.h:74), which
// refuses sm_100 at runtime (blocked_flash.cpp:31-33).
#include "hds_ptx.cuh"
#include "hds_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

// W consumer warps + 1 TMA producer warp; every consumer warp owns a 16-token strip, so a pipeline stage holds
// 16 * W KV tokens. W = 8: 128-token tiles, 192 KB of stages, one CTA per SM. W = 4: 64-token tiles, 96 KB, TWO CTAs
// per SM — same warps and bytes in flight per SM, but twice as many CTA slots, which removes the wave-quantisation
// loss when (items x kv heads x splits) is not a multiple of the SM count (e.g. 8 sequences x 32 heads = 256 CTAs).
constexpr int kAttnStages = 3;
__host__ __device__ constexpr int attn_threads(int w) { return (w + 1) * 32; }
__host__ __device__ constexpr int attn_tok(int w) { return 16 * w; }

struct AttnArgs {
    const void* q;
    void* out;
    const int32_t* block_tables;
    const hds_attn_item* items;
    int max_blocks_per_seq;
    int n_q_heads, n_kv_heads, block_size;
    int n_splits;
    float scale_log2;  // softmax scale * log2(e)
    int32_t* counters;  // [n_items][n_kv_heads]
    float* ws_ml;       // [n_items][n_kv_heads][n_splits][R][2]
    float* ws_o;        // [n_items][n_kv_heads][n_splits][R][D]
    int consumer_warps; // 8 (one CTA per SM) or 4 (two CTAs per SM); host-side choice, see attn_pick_warps
    const float* alibi; // [n_q_heads] ALiBi slopes (natural-log units) or nullptr: score += slope * (key pos - query pos)
    int window;         // sliding-window attention: a query at position p sees keys (p - window, p]; 0 = unlimited
};

template <int D, int W>
__host__ __device__ constexpr size_t attn_stage_bytes() {
    return static_cast<size_t>(2) * attn_tok(W) * D * 2;  // K + V
}
template <int D, int R, int W>
__host__ __device__ constexpr size_t attn_smem_bytes() {
    return kAttnStages * attn_stage_bytes<D, W>() + static_cast<size_t>(R) * D * 4 /*q*/ + 512 + 1024;
}

__device__ __forceinline__ float ex2f(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Warp-independent formulation: consumer warp w owns tokens [16w, 16w+16) of every 128-token tile for BOTH
// the q.K^T and the P.V step and keeps its own running (max, sum, accumulator); the eight partial states are
// merged once at the end. Nothing in the main loop synchronises the CTA: warps wait on the stage's TMA
// mbarrier and release the stage through a second mbarrier that the producer warp waits on.
template <int D, int R, bool kBf16, int W>
__global__ void __launch_bounds__(attn_threads(W), W == 8 ? 1 : 2)
paged_attn_splitkv_kernel(const __grid_constant__ CUtensorMap tmap_kv, const AttnArgs a) {
    constexpr int kAttnConsumerWarps = W;
    constexpr int kAttnThreads = attn_threads(W);
    constexpr int kAttnTok = attn_tok(W);
    constexpr int PANELS = D / 64;                 // 128-byte column panels per row
    constexpr int PANEL_BYTES = kAttnTok * 128;    // one panel of a 128-token tile
    constexpr int TILE_BYTES = PANELS * PANEL_BYTES;
    constexpr int CH = D / 16;                     // 16-byte K chunks per lane (two lanes share a token)
    constexpr int EPL = D / 32;                    // output dims per lane in the PV step

    pdl_enter();  // PDL: the next kernel may be scheduled; this one needs its predecessor's Q and KV writes
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* stages = smem;
    float* q_s = reinterpret_cast<float*>(smem + kAttnStages * attn_stage_bytes<D, W>());  // [R][D], pre-scaled
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(q_s + R * D);
    uint64_t* empty_bar = full_bar + kAttnStages;
    int* flag_s = reinterpret_cast<int*>(empty_bar + kAttnStages);
    float* ml_fin = reinterpret_cast<float*>(flag_s + 2);  // [R][2] merged (m, l) of this CTA

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int item_idx = blockIdx.x, kvh = blockIdx.y, split = blockIdx.z;
    const hds_attn_item item = a.items[item_idx];
    const int G = a.n_q_heads / a.n_kv_heads;
    const int rows = item.n_q * G;  // <= R (host guarantees)
    const int kv_len = item.pos0 + item.n_q;
    const int tiles_total = (kv_len + kAttnTok - 1) / kAttnTok;
    // sliding window: tiles that lie wholly before the first query's window are never loaded
    const int tile_first = a.window > 0 ? max(0, item.pos0 - a.window + 1) / kAttnTok : 0;
    const int tiles_live = tiles_total - tile_first;
    const int tps = (tiles_live + a.n_splits - 1) / a.n_splits;
    const int eff_splits = (tiles_live + tps - 1) / tps;
    if (split >= eff_splits) return;
    const int t0 = tile_first + split * tps;
    const int t1 = min(tiles_total, t0 + tps);
    const int bs = a.block_size;

    if (tid == 0) {
        prefetch_tmap(&tmap_kv);
        for (int i = 0; i < kAttnStages; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], kAttnConsumerWarps);
        }
        fence_barrier_init();
    }
    for (int e = tid; e < R * D; e += kAttnThreads) {
        const int r = e / D, dd = e % D;
        float v = 0.0f;
        if (r < rows) {
            const int qi = r / G, g = r % G;
            v = load_scalar<kBf16>(a.q, (static_cast<int64_t>(item.q_row0 + qi) * a.n_q_heads + kvh * G + g) * D + dd) *
                a.scale_log2;
        }
        q_s[e] = v;
    }
    __syncthreads();

    float m_run[R], l_run[R], acc[R][EPL];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        m_run[r] = -INFINITY;
        l_run[r] = 0.0f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) acc[r][i] = 0.0f;
    }

    if (warp == kAttnConsumerWarps) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const int32_t* table = a.block_tables + static_cast<int64_t>(item.seq) * a.max_blocks_per_seq;
            uint32_t stage = 0, phase = 0;
            for (int ti = t0; ti < t1; ++ti) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* kdst = stages + stage * attn_stage_bytes<D, W>();
                uint8_t* vdst = kdst + TILE_BYTES;
                const int tok0 = ti * kAttnTok;
                const int n_tok = min(kAttnTok, kv_len - tok0);
                const int box = bs < kAttnTok ? bs : kAttnTok;   // rows per TMA box (host builds the map with it)
                const int nb = (n_tok + box - 1) / box;
                mbar_expect_tx(&full_bar[stage], static_cast<uint32_t>(nb) * box * D * 2 * 2);
                for (int b = 0; b < nb; ++b) {
                    const int tok = tok0 + b * box;
                    const int blk = table[tok / bs];
                    const int in_blk = tok % bs;
                    const int krow = ((blk * 2 + 0) * a.n_kv_heads + kvh) * bs + in_blk;
                    const int vrow = ((blk * 2 + 1) * a.n_kv_heads + kvh) * bs + in_blk;
#pragma unroll
                    for (int p = 0; p < PANELS; ++p) {
                        tma_load_2d(kdst + p * PANEL_BYTES + b * box * 128, &tmap_kv, &full_bar[stage], p * 64, krow,
                                    kEvictFirst);
                        tma_load_2d(vdst + p * PANEL_BYTES + b * box * 128, &tmap_kv, &full_bar[stage], p * 64, vrow,
                                    kEvictFirst);
                    }
                }
                if (++stage == kAttnStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else {
        // ===================== consumers =====================
        const int j = lane & 15, hh = lane >> 4;   // token slot inside the warp's 16-token strip, dim half
        const int rt = warp * 16 + j;              // row of this lane's token inside the tile
        float qreg[R == 1 ? CH * 8 : 1];
        if (R == 1) {
#pragma unroll
            for (int i = 0; i < CH * 8; ++i) qreg[i] = q_s[hh * CH * 8 + i];
        }
        // ALiBi slope of each row's head, in log2 units (0 without ALiBi: the fused multiply-add below is then a no-op)
        float slope[R];
#pragma unroll
        for (int r = 0; r < R; ++r) slope[r] = a.alibi ? a.alibi[kvh * G + r % G] * 1.4426950408889634f : 0.0f;
        uint32_t stage = 0, phase = 0;
        for (int ti = t0; ti < t1; ++ti) {
            const uint8_t* ktile = stages + stage * attn_stage_bytes<D, W>();
            const uint8_t* vtile = ktile + TILE_BYTES;
            const int tok0 = ti * kAttnTok;
            const int n_tok = min(kAttnTok, kv_len - tok0);
            const int n_mine = min(16, max(0, n_tok - warp * 16));  // valid tokens of this warp's strip (uniform)
            mbar_wait(&full_bar[stage], phase);
            if (n_mine > 0) {
                // ---- s = q . k for my token (half of the head dim per lane, combined by one shuffle) ----
                float part[R];
#pragma unroll
                for (int r = 0; r < R; ++r) part[r] = 0.0f;
                if (j < n_mine) {
#pragma unroll
                    for (int c = 0; c < CH; ++c) {
                        const int cg = hh * CH + c;
                        const int panel = cg >> 3, cc = cg & 7;
                        const uint4 kk = *reinterpret_cast<const uint4*>(ktile + panel * PANEL_BYTES + rt * 128 +
                                                                        ((cc ^ (rt & 7)) << 4));
                        float kf[8] = {unpack_lo<kBf16>(kk.x), unpack_hi<kBf16>(kk.x), unpack_lo<kBf16>(kk.y),
                                       unpack_hi<kBf16>(kk.y), unpack_lo<kBf16>(kk.z), unpack_hi<kBf16>(kk.z),
                                       unpack_lo<kBf16>(kk.w), unpack_hi<kBf16>(kk.w)};
                        if (R == 1) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) part[0] = fmaf(qreg[c * 8 + i], kf[i], part[0]);
                        } else {
#pragma unroll
                            for (int r = 0; r < R; ++r) {
                                const float4 qa = *reinterpret_cast<const float4*>(q_s + r * D + cg * 8);
                                const float4 qb = *reinterpret_cast<const float4*>(q_s + r * D + cg * 8 + 4);
                                part[r] += qa.x * kf[0] + qa.y * kf[1] + qa.z * kf[2] + qa.w * kf[3] + qb.x * kf[4] +
                                           qb.y * kf[5] + qb.z * kf[6] + qb.w * kf[7];
                            }
                        }
                    }
                }
                float p[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float sc = part[r] + __shfl_xor_sync(0xffffffffu, part[r], 16);
                    const int q_pos = item.pos0 + r / G;
                    const bool ok = (j < n_mine) && (tok0 + rt <= q_pos) && (r < rows) &&
                                    (a.window <= 0 || tok0 + rt > q_pos - a.window);
                    sc = ok ? fmaf(slope[r], static_cast<float>(tok0 + rt - q_pos), sc) : -INFINITY;
                    float mx = sc;
#pragma unroll
                    for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                    const float m_new = fmaxf(m_run[r], mx);
                    const float scale = (m_run[r] == -INFINITY) ? 0.0f : ex2f(m_run[r] - m_new);
                    const float pv = (m_new == -INFINITY) ? 0.0f : ex2f(sc - m_new);
                    float ps = pv;
#pragma unroll
                    for (int o = 8; o > 0; o >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, o);
                    l_run[r] = l_run[r] * scale + ps;
                    m_run[r] = m_new;
                    p[r] = pv;
#pragma unroll
                    for (int i = 0; i < EPL; ++i) acc[r][i] *= scale;
                }
                // ---- acc += p_j * v_j : every lane owns EPL output dims, tokens of the strip in turn ----
                const int boff = lane * EPL * 2;                      // logical byte offset inside the V row
                const int vpanel = boff >> 7, vcc = (boff & 127) >> 4, vin = boff & 15;
                for (int jj = 0; jj < n_mine; ++jj) {
                    const int rv = warp * 16 + jj;
                    const uint8_t* vp = vtile + vpanel * PANEL_BYTES + rv * 128 + ((vcc ^ (rv & 7)) << 4) + vin;
                    float vf[EPL];
                    if (EPL == 4) {
                        const uint2 vv = *reinterpret_cast<const uint2*>(vp);
                        vf[0] = unpack_lo<kBf16>(vv.x); vf[1] = unpack_hi<kBf16>(vv.x);
                        vf[2 % EPL] = unpack_lo<kBf16>(vv.y); vf[3 % EPL] = unpack_hi<kBf16>(vv.y);
                    } else {
                        const uint32_t vv = *reinterpret_cast<const uint32_t*>(vp);
                        vf[0] = unpack_lo<kBf16>(vv); vf[1] = unpack_hi<kBf16>(vv);
                    }
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const float pj = __shfl_sync(0xffffffffu, p[r], jj);
#pragma unroll
                        for (int i = 0; i < EPL; ++i) acc[r][i] = fmaf(pj, vf[i], acc[r][i]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            if (++stage == kAttnStages) {
                stage = 0;
                phase ^= 1;
            }
        }
    }
    __syncthreads();  // every consumer is done with the ring: reuse it to merge the eight warp states

    float* mg_ml = reinterpret_cast<float*>(stages);                       // [warps][R][2]
    float* mg_o = mg_ml + kAttnConsumerWarps * R * 2;                      // [warps][R][D]
    if (warp < kAttnConsumerWarps) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (lane == 0) {
                mg_ml[(warp * R + r) * 2 + 0] = m_run[r];
                mg_ml[(warp * R + r) * 2 + 1] = l_run[r];
            }
#pragma unroll
            for (int i = 0; i < EPL; ++i) mg_o[(warp * R + r) * D + lane * EPL + i] = acc[r][i];
        }
    }
    __syncthreads();
    const int64_t ws_base = (static_cast<int64_t>(item_idx) * a.n_kv_heads + kvh) * a.n_splits;
    for (int e = tid; e < rows * D; e += kAttnThreads) {
        const int r = e / D, dd = e % D;
        float mx = -INFINITY;
#pragma unroll
        for (int w = 0; w < kAttnConsumerWarps; ++w) mx = fmaxf(mx, mg_ml[(w * R + r) * 2]);
        float o = 0.0f, l = 0.0f;
#pragma unroll
        for (int w = 0; w < kAttnConsumerWarps; ++w) {
            const float mw = mg_ml[(w * R + r) * 2];
            const float wt = (mw == -INFINITY) ? 0.0f : ex2f(mw - mx);
            l = fmaf(wt, mg_ml[(w * R + r) * 2 + 1], l);
            o = fmaf(wt, mg_o[(w * R + r) * D + dd], o);
        }
        if (dd == 0) {
            ml_fin[r * 2 + 0] = mx;
            ml_fin[r * 2 + 1] = l;
        }
        if (eff_splits == 1) {
            const int qi = r / G, gg = r % G;
            store_scalar<kBf16>(a.out, (static_cast<int64_t>(item.q_row0 + qi) * a.n_q_heads + kvh * G + gg) * D + dd,
                                o / l);
        } else {
            a.ws_o[((ws_base + split) * R + r) * D + dd] = o;
        }
    }
    if (eff_splits == 1) return;
    __syncthreads();
    if (tid < rows) {
        a.ws_ml[((ws_base + split) * R + tid) * 2 + 0] = ml_fin[tid * 2 + 0];
        a.ws_ml[((ws_base + split) * R + tid) * 2 + 1] = ml_fin[tid * 2 + 1];
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const int prev = atomicAdd(&a.counters[item_idx * a.n_kv_heads + kvh], 1);
        *flag_s = (prev == eff_splits - 1) ? 1 : 0;
    }
    __syncthreads();
    if (*flag_s == 0) return;
    __threadfence();
    // ---- last CTA of this (item, kv head): merge the splits ----
    for (int e = tid; e < rows * D; e += kAttnThreads) {
        const int r = e / D, dd = e % D;
        float mx = -INFINITY;
        for (int s = 0; s < eff_splits; ++s) mx = fmaxf(mx, __ldcg(&a.ws_ml[((ws_base + s) * R + r) * 2]));
        float o = 0.0f, l = 0.0f;
        for (int s = 0; s < eff_splits; ++s) {
            const float ms = __ldcg(&a.ws_ml[((ws_base + s) * R + r) * 2]);
            const float w = (ms == -INFINITY) ? 0.0f : ex2f(ms - mx);
            l += w * __ldcg(&a.ws_ml[((ws_base + s) * R + r) * 2 + 1]);
            o += w * __ldcg(&a.ws_o[((ws_base + s) * R + r) * D + dd]);
        }
        const int qi = r / G, gg = r % G;
        store_scalar<kBf16>(a.out, (static_cast<int64_t>(item.q_row0 + qi) * a.n_q_heads + kvh * G + gg) * D + dd,
                            o / l);
    }
    if (tid == 0) a.counters[item_idx * a.n_kv_heads + kvh] = 0;
}

static inline int64_t align256(int64_t x) { return (x + 255) & ~int64_t(255); }

template <int D, int R, bool kBf16, int W>
static int launch_splitkv_w(const CUtensorMap& tm, const AttnArgs& a, int n_items, cudaStream_t s) {
    constexpr int kAttnThreads = attn_threads(W);
    auto kern = paged_attn_splitkv_kernel<D, R, kBf16, W>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(attn_smem_bytes<D, R, W>())));
        configured[dev & 63] = true;
    }
    dim3 grid(n_items, a.n_kv_heads, a.n_splits);
    HDS_CHECK_CUDA(launch_kernel(kern, grid, dim3(kAttnThreads), attn_smem_bytes<D, R, W>(), s, tm, a));
    return HDS_OK;
}

template <int D, int R, bool kBf16>
static int launch_splitkv(const CUtensorMap& tm, const AttnArgs& a, int n_items, cudaStream_t s) {
    return a.consumer_warps == 4 ? launch_splitkv_w<D, R, kBf16, 4>(tm, a, n_items, s)
                                 : launch_splitkv_w<D, R, kBf16, 8>(tm, a, n_items, s);
}

template <int D, bool kBf16>
static int dispatch_rows(int rmax, const CUtensorMap& tm, const AttnArgs& a, int n_items, cudaStream_t s) {
    if (rmax <= 1) return launch_splitkv<D, 1, kBf16>(tm, a, n_items, s);
    if (rmax <= 2) return launch_splitkv<D, 2, kBf16>(tm, a, n_items, s);
    if (rmax <= 4) return launch_splitkv<D, 4, kBf16>(tm, a, n_items, s);
    return launch_splitkv<D, 8, kBf16>(tm, a, n_items, s);
}

static inline int rows_bucket(int r) { return r <= 1 ? 1 : r <= 2 ? 2 : r <= 4 ? 4 : 8; }

// Measured on B200 (profiles/kernel_bench_attn_warps.txt): with more CTAs than SMs the two-CTAs-per-SM variant is
// 2-8 % faster (8 seqs x 32 heads x 4K ctx: 101 -> 93 us; 16 seqs: 196 -> 181 us) because a second CTA fills the SM
// while the first drains; with fewer CTAs than SMs the bigger tile wins (each CTA is alone on its SM anyway).
static inline int attn_pick_warps(int64_t ctas) {
    static const int forced = getenv("HDS_ATTN_WARPS") ? atoi(getenv("HDS_ATTN_WARPS")) : 0;
    if (forced == 4 || forced == 8) return forced;
    return ctas > sm_count() ? 4 : 8;
}

}  // namespace hds

// ALiBi slopes for the NEXT paged-attention launch of this thread (either kernel family); consumed by that launch.
static thread_local const float* g_next_alibi = nullptr;
static thread_local int g_next_window = 0;
const float* hds::attn_take_alibi() {
    const float* p = g_next_alibi;
    g_next_alibi = nullptr;
    return p;
}
int hds::attn_take_window() {
    const int w = g_next_window;
    g_next_window = 0;
    return w;
}

using namespace hds;

extern "C" int hds_attention_next_alibi(const float* slopes) {
    g_next_alibi = slopes;
    return HDS_OK;
}

extern "C" int hds_attention_next_window(int window) {
    HDS_REQUIRE(window >= 0, "hds_attention_next_window: window >= 0 (0 = unlimited)");
    g_next_window = window;
    return HDS_OK;
}

extern "C" int64_t hds_paged_attention_splitkv_workspace_bytes(int n_items, int n_q_heads, int n_kv_heads,
                                                               int head_dim, int max_splits, int rows_per_item) {
    (void)n_q_heads;
    const int R = rows_bucket(rows_per_item);
    const int64_t units = static_cast<int64_t>(n_items) * n_kv_heads;
    return align256(units * 4) + align256(units * max_splits * R * 2 * 4) + align256(units * max_splits * R * head_dim * 4);
}

extern "C" int hds_paged_attention_splitkv(const void* q, void* out, const void* kv_cache_layer,
                                                int64_t num_blocks, const int32_t* block_tables,
                                                int max_blocks_per_seq, const hds_attn_item* items, int n_items,
                                                int rows_per_item, int n_q_heads, int n_kv_heads, int head_dim,
                                                int block_size, int n_splits, float scale, void* workspace,
                                                int64_t workspace_bytes, int dtype, void* stream) {
    const float* alibi = attn_take_alibi();
    const int window = attn_take_window();
    HDS_REQUIRE(head_dim == 64 || head_dim == 128, "paged_attention_splitkv: head_dim must be 64 or 128");
    HDS_REQUIRE(block_size == 16 || block_size == 32 || block_size == 64 || block_size == 128,
                "paged_attention_splitkv: block_size must be 16/32/64/128");
    HDS_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0, "paged_attention_splitkv: n_q_heads % n_kv_heads != 0");
    HDS_REQUIRE(rows_per_item >= 1 && rows_per_item <= 8, "paged_attention_splitkv: at most 8 query rows per item");
    HDS_REQUIRE(n_splits >= 1 && n_splits <= 64, "paged_attention_splitkv: n_splits in [1,64]");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "paged_attention_splitkv: dtype must be 0 (bf16) or 1 (fp16)");
    if (n_items == 0) return HDS_OK;
    const int64_t need =
        hds_paged_attention_splitkv_workspace_bytes(n_items, n_q_heads, n_kv_heads, head_dim, n_splits, rows_per_item);
    if (n_splits > 1) {
        HDS_REQUIRE(workspace != nullptr && workspace_bytes >= need, "paged_attention_splitkv: workspace too small");
    }
    const int R = rows_bucket(rows_per_item);
    AttnArgs a;
    memset(&a, 0, sizeof(a));
    a.q = q;
    a.out = out;
    a.block_tables = block_tables;
    a.items = items;
    a.max_blocks_per_seq = max_blocks_per_seq;
    a.n_q_heads = n_q_heads;
    a.n_kv_heads = n_kv_heads;
    a.block_size = block_size;
    a.n_splits = n_splits;
    a.scale_log2 = scale * 1.4426950408889634f;
    a.alibi = alibi;
    a.window = window;
    const int64_t units = static_cast<int64_t>(n_items) * n_kv_heads;
    char* ws = reinterpret_cast<char*>(workspace);
    a.counters = reinterpret_cast<int32_t*>(ws);
    a.ws_ml = reinterpret_cast<float*>(ws + align256(units * 4));
    a.ws_o = reinterpret_cast<float*>(ws + align256(units * 4) + align256(units * n_splits * R * 2 * 4));
    CUtensorMap tm;
    const uint64_t rows = static_cast<uint64_t>(num_blocks) * 2 * n_kv_heads * block_size;
    a.consumer_warps = attn_pick_warps(units * n_splits);
    const int tile_tok = attn_tok(a.consumer_warps);
    int rc = make_tmap_2d_16b(&tm, kv_cache_layer, head_dim, rows, static_cast<uint64_t>(head_dim) * 2, 64,
                              block_size < tile_tok ? block_size : tile_tok, dtype == 0);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (head_dim == 128)
        return dtype == 0 ? dispatch_rows<128, true>(R, tm, a, n_items, s) : dispatch_rows<128, false>(R, tm, a, n_items, s);
    return dtype == 0 ? dispatch_rows<64, true>(R, tm, a, n_items, s) : dispatch_rows<64, false>(R, tm, a, n_items, s);
}

// =================================================================================================
// hds_paged_attention_splitkv_static — the same split-KV attention, for callers that must not allocate
// =================================================================================================
// Some users of this library cannot hand it a scratch buffer because they have no way to obtain one:
// a serving process that gave every free byte of HBM to the KV cache at start-up (RESERVE mode does
// exactly that) has nothing left to cudaMalloc from; a C or C++ host that embeds the library without
// PyTorch has no caching allocator to borrow from; and latency-critical loops cannot afford cudaMalloc
// at all, because it synchronises the whole device. For them the partial (O, m, l) results and the
// per-(item, kv head) arrival counters live in the fixed array below instead of caller memory.
//
// Cost and limits, so nobody is surprised:
//   * the array is part of the library image: every GPU that loads libhds_b200.so holds
//     HDS_STATIC_SCRATCH_BYTES (8 MiB) of device memory until the process exits, used or not;
//   * a problem whose scratch need (hds_paged_attention_splitkv_workspace_bytes) exceeds that size is
//     refused with HDS_ERR_TOO_LARGE (-4) and a message saying how much was needed — lower n_splits or
//     n_items, or use the caller-scratch entry point;
//   * every call on a device shares the one array and finds whatever the previous call left in it. The
//     arrival counters are therefore cleared on the caller's stream at the start of each call (the layout
//     depends on n_items, so an earlier call's partial sums may sit where this call's counters go);
//     nothing else needs a defined starting value;
//   * because the array is shared, two calls that overlap in time on the same GPU (different streams, or
//     different host threads) race on it and produce garbage. Serialise them on one stream;
//   * the first call on each device looks up the array's device address (see below). Make that first
//     call during start-up, before any stream capture is begun on that device.
//
// In host code the identifier g_splitkv_static_scratch is only a stand-in for the symbol: taking its
// address there yields a host pointer that kernels must never see. The device address is fetched once
// per device, with that device current, through cudaGetSymbolAddress and passed to the kernel.
#define HDS_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
static __device__ __align__(256) unsigned char g_splitkv_static_scratch[HDS_STATIC_SCRATCH_BYTES];

namespace hds {
static int static_scratch_address(char** out) {
    static void* addr[64] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    HDS_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error("hds_paged_attention_splitkv_static: device ordinal out of range");
        return HDS_ERR_BAD_ARG;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (addr[dev] == nullptr) {
        void* p = nullptr;
        HDS_CHECK_CUDA(cudaGetSymbolAddress(&p, g_splitkv_static_scratch));  // device `dev` is current
        addr[dev] = p;
    }
    *out = static_cast<char*>(addr[dev]);
    return HDS_OK;
}
}  // namespace hds

extern "C" int64_t hds_paged_attention_splitkv_static_capacity(void) { return HDS_STATIC_SCRATCH_BYTES; }

extern "C" int hds_paged_attention_splitkv_static(const void* q, void* out, const void* kv_cache_layer,
                                                  int64_t num_blocks, const int32_t* block_tables,
                                                  int max_blocks_per_seq, const hds_attn_item* items, int n_items,
                                                  int rows_per_item, int n_q_heads, int n_kv_heads, int head_dim,
                                                  int block_size, int n_splits, float scale, int dtype, void* stream) {
    HDS_REQUIRE(head_dim == 64 || head_dim == 128, "paged_attention_splitkv_static: head_dim must be 64 or 128");
    HDS_REQUIRE(block_size == 16 || block_size == 32 || block_size == 64 || block_size == 128,
                "paged_attention_splitkv_static: block_size must be 16/32/64/128");
    HDS_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0,
                "paged_attention_splitkv_static: n_q_heads % n_kv_heads != 0");
    HDS_REQUIRE(rows_per_item >= 1 && rows_per_item <= 8, "paged_attention_splitkv_static: at most 8 query rows per item");
    HDS_REQUIRE(n_splits >= 1 && n_splits <= 64, "paged_attention_splitkv_static: n_splits in [1,64]");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "paged_attention_splitkv_static: dtype must be 0 (bf16) or 1 (fp16)");
    if (n_items == 0) return HDS_OK;
    const int64_t need =
        hds_paged_attention_splitkv_workspace_bytes(n_items, n_q_heads, n_kv_heads, head_dim, n_splits, rows_per_item);
    if (need > HDS_STATIC_SCRATCH_BYTES) {
        char b[256];
        snprintf(b, sizeof(b),
                 "paged_attention_splitkv_static: this problem needs %lld bytes of scratch but the built-in array holds "
                 "%d; reduce n_splits / n_items or call hds_paged_attention_splitkv with your own scratch",
                 static_cast<long long>(need), HDS_STATIC_SCRATCH_BYTES);
        set_error(b);
        return HDS_ERR_TOO_LARGE;
    }
    char* ws = nullptr;
    int rc = static_scratch_address(&ws);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int R = rows_bucket(rows_per_item);
    const int64_t units = static_cast<int64_t>(n_items) * n_kv_heads;
    // the counters must read zero when the kernel starts; the previous call may have left partial sums here
    HDS_CHECK_CUDA(cudaMemsetAsync(ws, 0, static_cast<size_t>(align256(units * 4)), s));
    AttnArgs a;
    memset(&a, 0, sizeof(a));
    a.q = q;
    a.out = out;
    a.block_tables = block_tables;
    a.items = items;
    a.max_blocks_per_seq = max_blocks_per_seq;
    a.n_q_heads = n_q_heads;
    a.n_kv_heads = n_kv_heads;
    a.block_size = block_size;
    a.n_splits = n_splits;
    a.scale_log2 = scale * 1.4426950408889634f;
    a.counters = reinterpret_cast<int32_t*>(ws);
    a.ws_ml = reinterpret_cast<float*>(ws + align256(units * 4));
    a.ws_o = reinterpret_cast<float*>(ws + align256(units * 4) + align256(units * n_splits * R * 2 * 4));
    CUtensorMap tm;
    const uint64_t rows = static_cast<uint64_t>(num_blocks) * 2 * n_kv_heads * block_size;
    rc = make_tmap_2d_16b(&tm, kv_cache_layer, head_dim, rows, static_cast<uint64_t>(head_dim) * 2, 64, block_size,
                          dtype == 0);
    if (rc) return rc;
    if (head_dim == 128)
        return dtype == 0 ? dispatch_rows<128, true>(R, tm, a, n_items, s) : dispatch_rows<128, false>(R, tm, a, n_items, s);
    return dtype == 0 ? dispatch_rows<64, true>(R, tm, a, n_items, s) : dispatch_rows<64, false>(R, tm, a, n_items, s);
}
