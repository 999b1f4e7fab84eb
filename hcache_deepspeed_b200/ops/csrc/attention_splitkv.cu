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

constexpr int kAttnThreads = 256;
constexpr int kAttnTok = 128;  // KV tokens per pipeline stage
constexpr int kAttnStages = 3;

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
};

template <int D>
__host__ __device__ constexpr size_t attn_stage_bytes() {
    return static_cast<size_t>(2) * kAttnTok * D * 2;  // K + V
}
template <int D, int R>
__host__ __device__ constexpr size_t attn_smem_bytes() {
    return kAttnStages * attn_stage_bytes<D>() + static_cast<size_t>(R) * D * 4 /*q*/ +
           2ull * R * kAttnTok * 4 /*partial scores*/ + static_cast<size_t>(R) * kAttnTok * 4 /*p*/ + 512 + 1024;
}

template <int D, int R, bool kBf16>
__global__ void __launch_bounds__(kAttnThreads, 1)
paged_attn_splitkv_kernel(const __grid_constant__ CUtensorMap tmap_kv, const AttnArgs a) {
    constexpr int PANELS = D / 64;                 // 128-byte column panels per row
    constexpr int PANEL_BYTES = kAttnTok * 128;    // one panel of a 128-token tile
    constexpr int TILE_BYTES = PANELS * PANEL_BYTES;
    constexpr int CH_QK = D / 16;                  // 16-byte chunks per thread in the QK phase (2 threads/token)
    constexpr int CV = D / 8;                      // chunks per V row
    constexpr int NG = kAttnThreads / CV;          // token groups in the PV phase

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* stages = smem;
    float* q_s = reinterpret_cast<float*>(smem + kAttnStages * attn_stage_bytes<D>());
    float* sc_s = q_s + R * D;                     // [2][R][128]
    float* p_s = sc_s + 2 * R * kAttnTok;          // [R][128]
    float* m_s = p_s + R * kAttnTok;               // [R]
    float* l_s = m_s + R;                          // [R]
    float* scale_s = l_s + R;                      // [R]
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(scale_s + R + 2);
    full_bar = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(full_bar) + 7) & ~uintptr_t(7));
    int* flag_s = reinterpret_cast<int*>(full_bar + kAttnStages);

    const int tid = threadIdx.x;
    const int item_idx = blockIdx.x, kvh = blockIdx.y, split = blockIdx.z;
    const hds_attn_item item = a.items[item_idx];
    const int G = a.n_q_heads / a.n_kv_heads;
    const int rows = item.n_q * G;  // <= R (host guarantees)
    const int kv_len = item.pos0 + item.n_q;
    const int tiles_total = (kv_len + kAttnTok - 1) / kAttnTok;
    const int tps = (tiles_total + a.n_splits - 1) / a.n_splits;
    const int eff_splits = (tiles_total + tps - 1) / tps;
    if (split >= eff_splits) return;
    const int t0 = split * tps;
    const int t1 = min(tiles_total, t0 + tps);
    const int bs = a.block_size;
    const int bps = kAttnTok / bs;  // blocks per stage
    const int32_t* table = a.block_tables + static_cast<int64_t>(item.seq) * a.max_blocks_per_seq;

    if (tid == 0) {
        prefetch_tmap(&tmap_kv);
        for (int i = 0; i < kAttnStages; ++i) mbar_init(&full_bar[i], 1);
        fence_barrier_init();
    }
    // q -> shared (fp32, pre-scaled so that softmax is exp2)
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
    if (tid < R) {
        m_s[tid] = -INFINITY;
        l_s[tid] = 0.0f;
    }
    __syncthreads();

    auto issue_tile = [&](int ti) {
        const int stage = (ti - t0) % kAttnStages;
        uint8_t* kdst = stages + stage * attn_stage_bytes<D>();
        uint8_t* vdst = kdst + TILE_BYTES;
        const int tok0 = ti * kAttnTok;
        const int n_tok = min(kAttnTok, kv_len - tok0);
        const int nb = (n_tok + bs - 1) / bs;
        mbar_expect_tx(&full_bar[stage], static_cast<uint32_t>(nb) * bs * D * 2 * 2);
        for (int b = 0; b < nb; ++b) {
            const int blk = table[tok0 / bs + b];
            const int krow = ((blk * 2 + 0) * a.n_kv_heads + kvh) * bs;
            const int vrow = ((blk * 2 + 1) * a.n_kv_heads + kvh) * bs;
#pragma unroll
            for (int p = 0; p < PANELS; ++p) {
                tma_load_2d(kdst + p * PANEL_BYTES + b * bs * 128, &tmap_kv, &full_bar[stage], p * 64, krow,
                            kEvictFirst);
                tma_load_2d(vdst + p * PANEL_BYTES + b * bs * 128, &tmap_kv, &full_bar[stage], p * 64, vrow,
                            kEvictFirst);
            }
        }
    };
    if (tid == 0) {
        for (int ti = t0; ti < t1 && ti < t0 + kAttnStages; ++ti) issue_tile(ti);
    }

    float acc[R][8];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[r][i] = 0.0f;

    const int warp = tid >> 5, lane = tid & 31;
    for (int ti = t0; ti < t1; ++ti) {
        const int it = ti - t0;
        const int stage = it % kAttnStages;
        const uint32_t phase = (it / kAttnStages) & 1;
        const uint8_t* ktile = stages + stage * attn_stage_bytes<D>();
        const uint8_t* vtile = ktile + TILE_BYTES;
        const int tok0 = ti * kAttnTok;
        const int n_tok = min(kAttnTok, kv_len - tok0);
        mbar_wait(&full_bar[stage], phase);

        // ---- S = q . K^T : thread (j, hh) covers half of token j's head dimension ----
        {
            const int j = tid & (kAttnTok - 1), hh = tid >> 7;
            float part[R];
#pragma unroll
            for (int r = 0; r < R; ++r) part[r] = 0.0f;
            if (j < n_tok) {
#pragma unroll
                for (int c = 0; c < CH_QK; ++c) {
                    const int cg = hh * CH_QK + c;
                    const int panel = cg >> 3, cc = cg & 7;
                    const uint4 kk = *reinterpret_cast<const uint4*>(ktile + panel * PANEL_BYTES + j * 128 +
                                                                    ((cc ^ (j & 7)) << 4));
                    const float k0 = unpack_lo<kBf16>(kk.x), k1 = unpack_hi<kBf16>(kk.x);
                    const float k2 = unpack_lo<kBf16>(kk.y), k3 = unpack_hi<kBf16>(kk.y);
                    const float k4 = unpack_lo<kBf16>(kk.z), k5 = unpack_hi<kBf16>(kk.z);
                    const float k6 = unpack_lo<kBf16>(kk.w), k7 = unpack_hi<kBf16>(kk.w);
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const float4 qa = *reinterpret_cast<const float4*>(q_s + r * D + cg * 8);
                        const float4 qb = *reinterpret_cast<const float4*>(q_s + r * D + cg * 8 + 4);
                        part[r] += qa.x * k0 + qa.y * k1 + qa.z * k2 + qa.w * k3 + qb.x * k4 + qb.y * k5 + qb.z * k6 +
                                   qb.w * k7;
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) sc_s[(hh * R + r) * kAttnTok + j] = part[r];
        }
        __syncthreads();
        // every thread is past the previous tile's PV: its stage can be refilled
        if (tid == 0 && it >= 1 && ti - 1 + kAttnStages < t1) issue_tile(ti - 1 + kAttnStages);

        // ---- online softmax: warp r owns query row r ----
        if (warp < R && warp < rows) {
            const int r = warp;
            const int q_pos = item.pos0 + r / G;
            float s[4];
            float tmax = -INFINITY;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int j = lane + 32 * i;
                const bool ok = j < n_tok && (tok0 + j) <= q_pos;
                s[i] = ok ? sc_s[r * kAttnTok + j] + sc_s[(R + r) * kAttnTok + j] : -INFINITY;
                tmax = fmaxf(tmax, s[i]);
            }
            tmax = warp_max(tmax);
            const float m_old = m_s[r];
            const float m_new = fmaxf(m_old, tmax);
            float psum = 0.0f;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float p = (m_new == -INFINITY) ? 0.0f : exp2f(s[i] - m_new);
                p_s[r * kAttnTok + lane + 32 * i] = p;
                psum += p;
            }
            psum = warp_sum(psum);
            if (lane == 0) {
                const float sc = (m_old == -INFINITY) ? 0.0f : exp2f(m_old - m_new);
                scale_s[r] = sc;
                m_s[r] = m_new;
                l_s[r] = l_s[r] * sc + psum;
            }
        }
        __syncthreads();

        // ---- O += P . V : thread (cv, g) owns 8 output dims for every NG-th token ----
        {
            const int cv = tid % CV, g = tid / CV;
            const int panel = cv >> 3, cc = cv & 7;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float sc = scale_s[r];
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[r][i] *= sc;
            }
            for (int j = g; j < n_tok; j += NG) {
                const uint4 vv =
                    *reinterpret_cast<const uint4*>(vtile + panel * PANEL_BYTES + j * 128 + ((cc ^ (j & 7)) << 4));
                const float v0 = unpack_lo<kBf16>(vv.x), v1 = unpack_hi<kBf16>(vv.x);
                const float v2 = unpack_lo<kBf16>(vv.y), v3 = unpack_hi<kBf16>(vv.y);
                const float v4 = unpack_lo<kBf16>(vv.z), v5 = unpack_hi<kBf16>(vv.z);
                const float v6 = unpack_lo<kBf16>(vv.w), v7 = unpack_hi<kBf16>(vv.w);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float p = p_s[r * kAttnTok + j];
                    acc[r][0] += p * v0; acc[r][1] += p * v1; acc[r][2] += p * v2; acc[r][3] += p * v3;
                    acc[r][4] += p * v4; acc[r][5] += p * v5; acc[r][6] += p * v6; acc[r][7] += p * v7;
                }
            }
        }
    }
    __syncthreads();  // all PV reads of the ring are done; reuse it for the cross-group reduction

    // ---- reduce the NG partial accumulators through shared memory ----
    float* red = reinterpret_cast<float*>(stages);  // [NG][R][D]
    {
        const int cv = tid % CV, g = tid / CV;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float4* dst = reinterpret_cast<float4*>(red + (g * R + r) * D + cv * 8);
            dst[0] = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
            dst[1] = make_float4(acc[r][4], acc[r][5], acc[r][6], acc[r][7]);
        }
    }
    __syncthreads();
    const int64_t ws_base = (static_cast<int64_t>(item_idx) * a.n_kv_heads + kvh) * a.n_splits;
    for (int e = tid; e < rows * D; e += kAttnThreads) {
        const int r = e / D, dd = e % D;
        float o = 0.0f;
#pragma unroll 4
        for (int g = 0; g < NG; ++g) o += red[(g * R + r) * D + dd];
        if (eff_splits == 1) {
            const int qi = r / G, gg = r % G;
            store_scalar<kBf16>(a.out, (static_cast<int64_t>(item.q_row0 + qi) * a.n_q_heads + kvh * G + gg) * D + dd,
                                o / l_s[r]);
        } else {
            a.ws_o[((ws_base + split) * R + r) * D + dd] = o;
        }
    }
    if (eff_splits == 1) return;
    if (tid < rows) {
        a.ws_ml[((ws_base + split) * R + tid) * 2 + 0] = m_s[tid];
        a.ws_ml[((ws_base + split) * R + tid) * 2 + 1] = l_s[tid];
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
            const float w = (ms == -INFINITY) ? 0.0f : exp2f(ms - mx);
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

template <int D, int R, bool kBf16>
static int launch_splitkv(const CUtensorMap& tm, const AttnArgs& a, int n_items, cudaStream_t s) {
    auto kern = paged_attn_splitkv_kernel<D, R, kBf16>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(attn_smem_bytes<D, R>())));
        configured[dev & 63] = true;
    }
    dim3 grid(n_items, a.n_kv_heads, a.n_splits);
    kern<<<grid, kAttnThreads, attn_smem_bytes<D, R>(), s>>>(tm, a);
    HDS_CHECK_CUDA(cudaGetLastError());
    return HDS_OK;
}

template <int D, bool kBf16>
static int dispatch_rows(int rmax, const CUtensorMap& tm, const AttnArgs& a, int n_items, cudaStream_t s) {
    if (rmax <= 1) return launch_splitkv<D, 1, kBf16>(tm, a, n_items, s);
    if (rmax <= 2) return launch_splitkv<D, 2, kBf16>(tm, a, n_items, s);
    if (rmax <= 4) return launch_splitkv<D, 4, kBf16>(tm, a, n_items, s);
    return launch_splitkv<D, 8, kBf16>(tm, a, n_items, s);
}

static inline int rows_bucket(int r) { return r <= 1 ? 1 : r <= 2 ? 2 : r <= 4 ? 4 : 8; }

}  // namespace hds

using namespace hds;

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
    const int64_t units = static_cast<int64_t>(n_items) * n_kv_heads;
    char* ws = reinterpret_cast<char*>(workspace);
    a.counters = reinterpret_cast<int32_t*>(ws);
    a.ws_ml = reinterpret_cast<float*>(ws + align256(units * 4));
    a.ws_o = reinterpret_cast<float*>(ws + align256(units * 4) + align256(units * n_splits * R * 2 * 4));
    CUtensorMap tm;
    const uint64_t rows = static_cast<uint64_t>(num_blocks) * 2 * n_kv_heads * block_size;
    int rc = make_tmap_2d_16b(&tm, kv_cache_layer, head_dim, rows, static_cast<uint64_t>(head_dim) * 2, 64, block_size,
                              dtype == 0);
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
