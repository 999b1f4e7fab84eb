// tcgen05 ragged paged attention for work items of up to 128 query tokens (prefill, chunked
// continuation after an HCache restore). One CTA = one (item, query head).
//
//   warp 0    : TMA producer — Q tile once, then 128-token K/V tiles gathered page by page from the
//               paged cache through the block table (2-stage ring, SWIZZLE_128B)
//   warp 1    : TMEM allocator + tcgen05.mma issuer:  S = Q K^T (K-major x K-major),
//               O_tile = P V (P K-major from shared memory, V MN-major straight from its page layout)
//   warps 2-9 : softmax — TWO threads per query row (warps w and w+4 share a TMEM lane quadrant and split the tile's
//               128 key columns 64/64, exchanging only the row max through shared memory), so every scheduler has two
//               warps to interleave: ONE tcgen05.ld pass pulls this thread's 64 scores into
//               registers and releases S at once (so Q K^T of the next tile runs under this tile's exponentials),
//               causal mask, online max/sum, P -> shared (UMMA K-major swizzled layout).
//   O accumulates in TMEM across all KV tiles (PV issued with accumulate); it is rescaled in place only when a
//   row's running max grows by more than 2^8 (lazy rescale: probabilities are computed against a slightly stale
//   max, which is exact after the final division by the matching row sum)
//
// Head-packed mode (kPacked, grouped-query decode): one CTA = one (item, KV head); the 128 MMA rows are the
// (token, query head of the group) pairs of the item — G heads x up to 128/G tokens, fetched by ONE 3-D TMA box from
// q[T][heads][D] — so a decode step of a GQA model (Llama-2-70B: 8 query heads per KV head) streams every K/V page ONCE
// for the whole group and does the 8x larger arithmetic on the tensor cores instead of on FFMA (the split-KV SIMT kernel
// is compute-bound there: 35 ms per 64-session iteration against a 9 ms HBM floor). TMEM lane quadrants without a live
// row skip the softmax arithmetic and only keep the barrier protocol going.
//
// Stands in for the external paged flash kernel of the reference
// (deepspeed/inference/v2/kernels/ragged_ops/blocked_flash/blocked_flash.cpp:21-101; atoms:
//  ragged_ops/includes/attention_atom.h:11-39).
#include "hds_ptx.cuh"
#include "hds_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

constexpr int kPfThreads = 320;   // TMA warp + MMA warp + 8 softmax warps (two per TMEM lane quadrant)
constexpr int kPfTok = 128;
constexpr int kPfStages = 2;

struct PrefillArgs {
    void* out;
    const int32_t* block_tables;
    const hds_attn_item* items;
    int max_blocks_per_seq;
    int n_q_heads, n_kv_heads, block_size;
    float scale_log2;
    int n_splits;       // KV range split over gridDim.z CTAs
    int32_t* tickets;   // [n_items][n_q_heads][4 quadrants][2 halves], zero between calls
    float* ws_ml;       // [n_items][n_q_heads][n_splits][128][2]
    float* ws_o;        // [n_items][n_q_heads][n_splits][128][D]
    const float* alibi; // [n_q_heads] ALiBi slopes (natural-log units) or nullptr
    int window;         // sliding-window attention: a query at position p sees keys (p - window, p]; 0 = unlimited
};

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int D>
__host__ __device__ constexpr size_t pf_smem_bytes() {
    // Q [128 x D] + P [128 x 128] + stages x (K + V) [128 x D] (16-bit) + barriers + alignment slack
    return static_cast<size_t>(128) * D * 2 + 128 * 128 * 2 + static_cast<size_t>(kPfStages) * 2 * kPfTok * D * 2 + 256 +
           2 * 2 * 128 * 4 /*row-max exchange*/ + 1024;
}

template <int D, bool kBf16, bool kPacked = false>
__global__ void __launch_bounds__(kPfThreads, 1)
paged_attn_prefill_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                          const PrefillArgs a) {
    constexpr int PANELS = D / 64;
    constexpr int PANEL_BYTES = kPfTok * 128;  // [128 rows x 128 B]
    constexpr int TILE_BYTES = PANELS * PANEL_BYTES;
    constexpr uint32_t TMEM_COLS = 256;  // S: 128 columns, O_tile: D (<=128) columns
    constexpr uint32_t IDESC_QK = make_idesc_f16(128, kPfTok, kBf16, false, false);
    constexpr uint32_t IDESC_PV = make_idesc_f16(128, D, kBf16, false, true);  // B = V is MN-major

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_sm = smem;                         // PANELS panels of [128 x 64]
    uint8_t* p_sm = q_sm + TILE_BYTES;            // 2 panels of [128 rows x 64 tokens]
    uint8_t* kv_sm = p_sm + 2 * PANEL_BYTES;      // stages x (K tile, V tile)
    uint64_t* bars = reinterpret_cast<uint64_t*>(kv_sm + kPfStages * 2 * TILE_BYTES);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;
    uint64_t* kv_empty = kv_full + kPfStages;
    uint64_t* s_full = kv_empty + kPfStages;
    uint64_t* p_full = s_full + 1;
    uint64_t* o_full = p_full + 1;   // PV of a tile has completed: O is stable, the P buffer is free
    uint64_t* s_empty = o_full + 1;  // all four softmax warps hold S in registers: the S accumulator is free
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + 1);
    float* xch = reinterpret_cast<float*>(tmem_slot + 2);  // [2 (tile parity)][2 (half)][128 rows] row-max exchange

    const uint32_t warp = warp_id(), lane = lane_id();
    pdl_enter();  // PDL: the next kernel may be scheduled; this one needs its predecessor's Q and KV writes
    // heaviest work first: the host lists a sequence's query tiles in position order, so later items see more keys;
    // walking them backwards leaves the short ones to fill the tail of the schedule
    const int item_idx = static_cast<int>(gridDim.x - 1 - blockIdx.x);
    const hds_attn_item item = a.items[item_idx];
    const int head = blockIdx.y;                       // query head, or (kPacked) KV head
    const int G = a.n_q_heads / a.n_kv_heads;
    const int kvh = kPacked ? head : head / G;
    const int n_live_rows = kPacked ? item.n_q * G : item.n_q;   // rows of the 128-row tile that are ever stored
    const int kv_len = item.pos0 + item.n_q;
    const int tiles_total = (kv_len + kPfTok - 1) / kPfTok;
    // sliding window: key tiles wholly before the FIRST query's window are skipped (later rows see even less of them)
    const int tile_first = a.window > 0 ? max(0, item.pos0 - a.window + 1) / kPfTok : 0;
    const int tiles_live = tiles_total - tile_first;
    const int tps = (tiles_live + a.n_splits - 1) / a.n_splits;
    const int eff_splits = (tiles_live + tps - 1) / tps;
    const int split = blockIdx.z;
    if (split >= eff_splits) return;
    const int t0 = tile_first + split * tps;
    const int n_tiles = min(tiles_total, t0 + tps) - t0;  // tiles of this split: absolute index t0 + ti
    const int bs = a.block_size;
    const int bps = kPfTok / bs;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_q);
        prefetch_tmap(&tmap_kv);
        mbar_init(q_full, 1);
        for (int i = 0; i < kPfStages; ++i) {
            mbar_init(&kv_full[i], 1);
            mbar_init(&kv_empty[i], 1);
        }
        mbar_init(s_full, 1);
        mbar_init(p_full, 8);
        mbar_init(o_full, 1);
        mbar_init(s_empty, 8);
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
    const uint32_t tmem_s = tmem_base;
    const uint32_t tmem_o = tmem_base + 128;

    if (warp == 0) {
        if (lane == 0) {
            // ---- Q: rows q_row0 .. q_row0+127 of head `head` (rows past n_q are never stored) ----
            mbar_expect_tx(q_full, TILE_BYTES);
#pragma unroll
            for (int p = 0; p < PANELS; ++p) {
                if constexpr (kPacked)   // box [128/G tokens][G heads][64 dims]: tile row = token * G + head in group
                    tma_load_3d(q_sm + p * PANEL_BYTES, &tmap_q, q_full, p * 64, kvh * G, item.q_row0, kEvictFirst);
                else
                    tma_load_2d(q_sm + p * PANEL_BYTES, &tmap_q, q_full, head * D + p * 64, item.q_row0, kEvictFirst);
            }
            // ---- K/V tiles ----
            const int32_t* table = a.block_tables + static_cast<int64_t>(item.seq) * a.max_blocks_per_seq;
            const int last_blk_idx = (kv_len - 1) / bs;
            uint32_t stage = 0, phase = 0;
            for (int ti = 0; ti < n_tiles; ++ti) {
                mbar_wait(&kv_empty[stage], phase ^ 1);
                uint8_t* kdst = kv_sm + stage * 2 * TILE_BYTES;
                uint8_t* vdst = kdst + TILE_BYTES;
                mbar_expect_tx(&kv_full[stage], 2 * TILE_BYTES);
                for (int b = 0; b < bps; ++b) {
                    // past the end of the sequence re-load its last page: masked out later, but keeps
                    // every V value that meets a zero probability finite (0 * NaN would poison the row)
                    const int bi = min((t0 + ti) * bps + b, last_blk_idx);
                    const int blk = table[bi];
                    const int krow = ((blk * 2 + 0) * a.n_kv_heads + kvh) * bs;
                    const int vrow = ((blk * 2 + 1) * a.n_kv_heads + kvh) * bs;
#pragma unroll
                    for (int p = 0; p < PANELS; ++p) {
                        tma_load_2d(kdst + p * PANEL_BYTES + b * bs * 128, &tmap_kv, &kv_full[stage], p * 64, krow,
                                    kEvictNormal);
                        tma_load_2d(vdst + p * PANEL_BYTES + b * bs * 128, &tmap_kv, &kv_full[stage], p * 64, vrow,
                                    kEvictNormal);
                    }
                }
                if (++stage == kPfStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            auto issue_qk = [&](uint32_t stage) {
                const uint8_t* ktile = kv_sm + stage * 2 * TILE_BYTES;
#pragma unroll
                for (int p = 0; p < PANELS; ++p) {
                    const uint64_t dq = make_desc_kmajor_sw128(smem_u32(q_sm + p * PANEL_BYTES));
                    const uint64_t dk = make_desc_kmajor_sw128(smem_u32(ktile + p * PANEL_BYTES));
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16_ss(tmem_s, dq + 2 * k, dk + 2 * k, IDESC_QK, (p > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(s_full);
            };
            mbar_wait(q_full, 0);
            mbar_wait(&kv_full[0], 0);
            tc_fence_after();
            issue_qk(0);
            uint32_t stage = 0, phase = 0;
            for (int ti = 0; ti < n_tiles; ++ti) {
                // ---- S for tile ti+1 as soon as the softmax warps have pulled S(ti) into registers ----
                const uint32_t nstage = (stage + 1 == kPfStages) ? 0 : stage + 1;
                const uint32_t nphase = (stage + 1 == kPfStages) ? phase ^ 1 : phase;
                if (ti + 1 < n_tiles) {
                    mbar_wait(s_empty, ti & 1);
                    mbar_wait(&kv_full[nstage], nphase);
                    tc_fence_after();
                    issue_qk(nstage);
                }
                // ---- O += P V for tile ti ----
                mbar_wait(p_full, ti & 1);
                tc_fence_after();
                const uint8_t* vtile = kv_sm + stage * 2 * TILE_BYTES + TILE_BYTES;
#pragma unroll
                for (int kk = 0; kk < kPfTok / 16; ++kk) {
                    // A = P: K-major; 16 tokens = 32 B inside the 64-token panel, 4 steps per panel
                    const uint64_t dp =
                        make_desc_kmajor_sw128(smem_u32(p_sm + (kk >> 2) * PANEL_BYTES)) + 2 * (kk & 3);
                    // B = V: MN-major; 16 tokens = 16 rows of 128 B; panels of 64 dims are PANEL_BYTES apart
                    const uint64_t dv = make_desc_mnmajor_sw128(smem_u32(vtile + kk * 16 * 128), PANEL_BYTES);
                    umma_f16_ss(tmem_o, dp, dv, IDESC_PV, (ti > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(o_full);
                umma_commit(&kv_empty[stage]);
                stage = nstage;
                phase = nphase;
            }
        }
    } else {
        // ===================== softmax / output (warps 2..5, one thread per query row) =====================
        const uint32_t quad = warp & 3;              // TMEM lane quadrant this warp may touch
        const uint32_t half = (warp - 2) >> 2;       // which 64 key columns (and which half of the output dims)
        const int r = quad * 32 + lane;              // row of the 128-row tile
        const int r_tok = kPacked ? r / G : r;       // query token inside the item
        const int r_head = kPacked ? kvh * G + (r - r_tok * G) : head;
        const int q_pos = item.pos0 + r_tok;
        const bool quad_live = static_cast<int>(quad * 32) < n_live_rows;   // warp-uniform, same for both halves
        const uint32_t lane_addr = (quad * 32u) << 16;
        constexpr int HC = kPfTok / 2;               // key columns per thread
        constexpr int HD2 = D / 2;                   // output dims per thread
        float m_run = -INFINITY, l_run = 0.0f;       // l_run: this thread's columns only
        uint8_t* p_row = p_sm + half * PANEL_BYTES + r * 128;   // half h owns P panel h (64 tokens = 128 B per row)
        constexpr float kRescaleThreshold = 8.0f;    // log2 units: probabilities stay below 2^8 against a stale max
        auto pair_sync = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory"); };
        for (int ti = 0; ti < n_tiles; ++ti) {
            const int tok0 = (t0 + ti) * kPfTok + half * HC;
            // warp-uniform: the tile reaches past the first query (causal edge) or starts before the last query's window
            const bool need_mask = (t0 + ti) * kPfTok + kPfTok - 1 > item.pos0 ||
                                   (a.window > 0 && (t0 + ti) * kPfTok <= item.pos0 + item.n_q - 1 - a.window);
            const int win_lo = a.window > 0 ? q_pos - a.window : -(1 << 30);   // keys at or before this position are hidden
            mbar_wait(s_full, ti & 1);
            tc_fence_after();
            if (!quad_live) {
                // no live row in this lane quadrant (head-packed decode): keep the barrier protocol going, skip the math
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(s_empty);
                if (ti > 0) {
                    mbar_wait(o_full, (ti - 1) & 1);
                    tc_fence_after();
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_full);
                continue;
            }
            // ---- my 64 scores -> registers, then hand the S accumulator back to the MMA warp ----
            uint32_t sr[HC];
#pragma unroll
            for (int c = 0; c < HC; c += 32)
                tmem_ld_32x32(tmem_s + lane_addr + half * HC + c, *reinterpret_cast<uint32_t(*)[32]>(&sr[c]));
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_empty);
            if (a.alibi != nullptr) {
                // ALiBi: raw score += slope * (key - query position), expressed in raw-score units so that the scaling below
                // stays one multiply-add per element (kernel-uniform branch; slopes differ per row only in packed mode)
                const float sl = a.alibi[r_head] * 1.4426950408889634f / a.scale_log2;
                const float d0 = static_cast<float>(tok0 - q_pos);
#pragma unroll
                for (int j = 0; j < HC; ++j)
                    sr[j] = __float_as_uint(fmaf(sl, d0 + static_cast<float>(j), __uint_as_float(sr[j])));
            }
            float tmax = -INFINITY;
            if (need_mask) {
#pragma unroll
                for (int j = 0; j < HC; ++j) {
                    const int kv_pos = tok0 + j;
                    const float sv = (kv_pos <= q_pos && kv_pos < kv_len && kv_pos > win_lo) ? __uint_as_float(sr[j]) : -INFINITY;
                    sr[j] = __float_as_uint(sv);
                    tmax = fmaxf(tmax, sv);
                }
            } else {
#pragma unroll
                for (int j = 0; j < HC; ++j) tmax = fmaxf(tmax, __uint_as_float(sr[j]));
            }
            // row max over both halves
            float* x = xch + (ti & 1) * 256;
            x[half * 128 + r] = tmax;
            pair_sync();
            tmax = fmaxf(tmax, x[(half ^ 1) * 128 + r]);
            // lazy rescale: move the reference max only when it grew by more than the threshold
            const float m_cand = fmaxf(m_run, tmax * a.scale_log2);
            const bool grow = (m_cand > m_run + kRescaleThreshold) || (m_run == -INFINITY && m_cand != -INFINITY);
            const float m_new = grow ? m_cand : m_run;
            const bool fresh = (m_run == -INFINITY);   // nothing accumulated for this row yet
            const float alpha = (!grow || fresh) ? (grow ? 0.0f : 1.0f) : ex2_approx(m_run - m_new);
            const float m_sub = (m_new == -INFINITY) ? 0.0f : m_new;  // fully masked row: exp2(-inf - 0) = 0
            float ps0 = 0.0f, ps1 = 0.0f;
            uint32_t packed[HC / 2];
#pragma unroll
            for (int j = 0; j < HC; j += 2) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(sr[j]), a.scale_log2, -m_sub));
                const float p1 = ex2_approx(fmaf(__uint_as_float(sr[j + 1]), a.scale_log2, -m_sub));
                ps0 += p0;
                ps1 += p1;
                packed[j >> 1] = pack2<kBf16>(p0, p1);
            }
            l_run = l_run * alpha + (ps0 + ps1);
            m_run = m_new;
            // PV of the previous tile must be complete before P is overwritten or O is rescaled
            if (ti > 0) {
                mbar_wait(o_full, (ti - 1) & 1);
                tc_fence_after();
                // (alpha may underflow to exactly 0 when the maximum jumps by > 126 log2 units, e.g. under a steep ALiBi
                // slope: the stale accumulation must then still be wiped, so "fresh" is tracked separately)
                const bool rescale = grow && !fresh;
                if (__any_sync(0xffffffffu, rescale)) {
                    const float f = rescale ? alpha : 1.0f;
#pragma unroll
                    for (int c = 0; c < HD2; c += 32) {
                        uint32_t orr[32];
                        tmem_ld_32x32(tmem_o + lane_addr + half * HD2 + c, orr);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) orr[j] = __float_as_uint(__uint_as_float(orr[j]) * f);
                        tmem_st_32x32(tmem_o + lane_addr + half * HD2 + c, orr);
                    }
                    tmem_st_wait();
                }
            }
            // probabilities -> shared memory (A operand of the PV MMA): 8 tokens per 16-byte chunk, swizzled by row
#pragma unroll
            for (int ch = 0; ch < HC / 8; ++ch) {
                const uint4 v = make_uint4(packed[4 * ch], packed[4 * ch + 1], packed[4 * ch + 2], packed[4 * ch + 3]);
                *reinterpret_cast<uint4*>(p_row + ((ch ^ (r & 7)) << 4)) = v;
            }
            // make P (and a rescaled O) visible to the tensor core, then signal the MMA warp
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
        }
        // ---- total row sum over both halves, then my half of the final O row from TMEM ----
        if (quad_live) {
            float* x = xch + (n_tiles & 1) * 256;
            x[half * 128 + r] = l_run;
            pair_sync();
            l_run += x[(half ^ 1) * 128 + r];
        }
        mbar_wait(o_full, (n_tiles - 1) & 1);
        tc_fence_after();
        float o_acc[HD2];
#pragma unroll
        for (int c = 0; c < HD2; c += 32) {
            uint32_t orr[32];
            tmem_ld_32x32(tmem_o + lane_addr + half * HD2 + c, orr);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) o_acc[c + j] = __uint_as_float(orr[j]);
        }
        tc_fence_before();
        const bool row_valid = r < n_live_rows;
        char* dst = reinterpret_cast<char*>(a.out) +
                    ((static_cast<int64_t>(item.q_row0 + r_tok) * a.n_q_heads + r_head) * D + half * HD2) * 2;
        if (eff_splits == 1) {
            if (row_valid) {
                const float inv = 1.0f / l_run;
#pragma unroll
                for (int c = 0; c < HD2; c += 8) {
                    uint4 v;
                    v.x = pack2<kBf16>(o_acc[c + 0] * inv, o_acc[c + 1] * inv);
                    v.y = pack2<kBf16>(o_acc[c + 2] * inv, o_acc[c + 3] * inv);
                    v.z = pack2<kBf16>(o_acc[c + 4] * inv, o_acc[c + 5] * inv);
                    v.w = pack2<kBf16>(o_acc[c + 6] * inv, o_acc[c + 7] * inv);
                    st_v4(dst + c * 2, v);
                }
            }
        } else if (quad_live) {  // warp-uniform: this quadrant owns valid rows
            // ---- split-KV: park the partial (O, m, l); the last split of this (quadrant, half) merges its dims ----
            const int64_t unit = static_cast<int64_t>(item_idx) * gridDim.y + head;
            const int64_t ws_row = (unit * a.n_splits + split) * 128 + r;
            if (row_valid) {
                float4* wo = reinterpret_cast<float4*>(a.ws_o + ws_row * D + half * HD2);
#pragma unroll
                for (int c = 0; c < HD2; c += 4) __stcg(wo + (c >> 2), make_float4(o_acc[c], o_acc[c + 1], o_acc[c + 2], o_acc[c + 3]));
                if (half == 0) __stcg(reinterpret_cast<float2*>(a.ws_ml + ws_row * 2), make_float2(m_run, l_run));
            }
            __threadfence();
            pair_sync();  // both halves of the rows (O dims and (m, l)) are published before either takes a ticket
            int last = 0;
            if (lane == 0) last = (atomicAdd(&a.tickets[(unit * 4 + quad) * 2 + half], 1) == eff_splits - 1) ? 1 : 0;
            last = __shfl_sync(0xffffffffu, last, 0);
            if (last) {
                __threadfence();
                if (row_valid) {
                    float mx = -INFINITY;
                    for (int sp = 0; sp < eff_splits; ++sp)
                        mx = fmaxf(mx, __ldcg(a.ws_ml + ((unit * a.n_splits + sp) * 128 + r) * 2));
                    float l_tot = 0.0f;
#pragma unroll
                    for (int i = 0; i < HD2; ++i) o_acc[i] = 0.0f;
                    for (int sp = 0; sp < eff_splits; ++sp) {
                        const int64_t row_sp = (unit * a.n_splits + sp) * 128 + r;
                        const float2 ml = __ldcg(reinterpret_cast<const float2*>(a.ws_ml + row_sp * 2));
                        const float w = (ml.x == -INFINITY) ? 0.0f : ex2_approx(ml.x - mx);
                        l_tot = fmaf(w, ml.y, l_tot);
                        const float4* wo = reinterpret_cast<const float4*>(a.ws_o + row_sp * D + half * HD2);
#pragma unroll
                        for (int c = 0; c < HD2; c += 4) {
                            const float4 t = __ldcg(wo + (c >> 2));
                            o_acc[c] = fmaf(w, t.x, o_acc[c]);
                            o_acc[c + 1] = fmaf(w, t.y, o_acc[c + 1]);
                            o_acc[c + 2] = fmaf(w, t.z, o_acc[c + 2]);
                            o_acc[c + 3] = fmaf(w, t.w, o_acc[c + 3]);
                        }
                    }
                    const float inv = 1.0f / l_tot;
#pragma unroll
                    for (int c = 0; c < HD2; c += 8) {
                        uint4 v;
                        v.x = pack2<kBf16>(o_acc[c + 0] * inv, o_acc[c + 1] * inv);
                        v.y = pack2<kBf16>(o_acc[c + 2] * inv, o_acc[c + 3] * inv);
                        v.z = pack2<kBf16>(o_acc[c + 4] * inv, o_acc[c + 5] * inv);
                        v.w = pack2<kBf16>(o_acc[c + 6] * inv, o_acc[c + 7] * inv);
                        st_v4(dst + c * 2, v);
                    }
                }
                __syncwarp();
                if (lane == 0) a.tickets[(unit * 4 + quad) * 2 + half] = 0;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

}  // namespace hds
#include "attention_prefill64.cuh"
namespace hds {

// 64-key tiles, two CTAs per SM (attention_prefill64.cuh): head_dim 128, pages of at most 64 tokens
template <bool kBf16>
static int launch_prefill64(const CUtensorMap& tq, const CUtensorMap& tkv, const PrefillArgs& a, int n_items, cudaStream_t s) {
    auto kern = paged_attn_prefill64_kernel<kBf16>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kP64SmemBytes)));
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        configured[dev & 63] = true;
    }
    dim3 grid(n_items, a.n_q_heads, a.n_splits);
    HDS_CHECK_CUDA(launch_kernel(kern, grid, dim3(kP64Threads), kP64SmemBytes, s, tq, tkv, a));
    return HDS_OK;
}

// Measured (tools/prefill64_probe.sh, 4224 causal tokens x 32 heads x d128): 292 us (515 TFLOP/s) against 302 us (498) for the
// 128-key kernel, and 53.7 ms against 51.8 ms for a whole fresh 4160-token Llama-2-7B prefill: two co-resident CTAs do not buy
// the overlap that was hoped for, so the 128-key kernel stays the default and this shape is opt-in (HDS_PREFILL_TILE=64).
static bool prefill64_enabled() {
    static const int tile = getenv("HDS_PREFILL_TILE") ? atoi(getenv("HDS_PREFILL_TILE")) : 128;
    return tile == 64;
}

template <int D, bool kBf16, bool kPacked = false>
static int launch_prefill(const CUtensorMap& tq, const CUtensorMap& tkv, const PrefillArgs& a, int n_items,
                          cudaStream_t s) {
    auto kern = paged_attn_prefill_kernel<D, kBf16, kPacked>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        HDS_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(pf_smem_bytes<D>())));
        configured[dev & 63] = true;
    }
    dim3 grid(n_items, kPacked ? a.n_kv_heads : a.n_q_heads, a.n_splits);
    HDS_CHECK_CUDA(launch_kernel(kern, grid, dim3(kPfThreads), pf_smem_bytes<D>(), s, tq, tkv, a));
    return HDS_OK;
}

}  // namespace hds

using namespace hds;

static inline int64_t pf_align256(int64_t x) { return (x + 255) & ~int64_t(255); }

extern "C" int64_t hds_paged_attention_prefill_workspace_bytes(int n_items, int n_q_heads, int head_dim, int n_splits) {
    if (n_splits <= 1) return 0;
    const int64_t units = static_cast<int64_t>(n_items) * n_q_heads;
    return pf_align256(units * 8 * 4) + pf_align256(units * n_splits * 128 * 2 * 4) +
           pf_align256(units * n_splits * 128 * head_dim * 4);
}

static int attention_tc_impl(bool packed, const void* q, void* out, const void* kv_cache_layer, int64_t num_blocks,
                             const int32_t* block_tables, int max_blocks_per_seq, const hds_attn_item* items, int n_items,
                             int n_q_rows, int n_q_heads, int n_kv_heads, int head_dim, int block_size, float scale,
                             int n_splits, void* workspace, int64_t workspace_bytes, int dtype, void* stream) {
    HDS_REQUIRE(head_dim == 64 || head_dim == 128, "paged_attention_prefill: head_dim must be 64 or 128");
    HDS_REQUIRE(block_size == 16 || block_size == 32 || block_size == 64 || block_size == 128,
                "paged_attention_prefill: block_size must be 16/32/64/128");
    HDS_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0, "paged_attention_prefill: n_q_heads % n_kv_heads != 0");
    HDS_REQUIRE(dtype == 0 || dtype == 1, "paged_attention_prefill: dtype must be 0 (bf16) or 1 (fp16)");
    const float* alibi = attn_take_alibi();
    const int window = attn_take_window();
    if (n_items == 0) return HDS_OK;
    PrefillArgs a;
    memset(&a, 0, sizeof(a));
    a.alibi = alibi;
    a.window = window;
    a.out = out;
    a.block_tables = block_tables;
    a.items = items;
    a.max_blocks_per_seq = max_blocks_per_seq;
    a.n_q_heads = n_q_heads;
    a.n_kv_heads = n_kv_heads;
    a.block_size = block_size;
    a.scale_log2 = scale * 1.4426950408889634f;
    if (n_splits < 1) n_splits = 1;
    HDS_REQUIRE(n_splits <= 64, "paged_attention_prefill: n_splits in [1,64]");
    a.n_splits = n_splits;
    const int grid_heads = packed ? n_kv_heads : n_q_heads;
    const int G = n_q_heads / n_kv_heads;
    HDS_REQUIRE(!packed || (head_dim == 128 && G >= 2 && G <= 128 && 128 % G == 0),
                "paged_attention_packed: head_dim 128 and a group size that divides 128");
    if (n_splits > 1) {
        const int64_t need = hds_paged_attention_prefill_workspace_bytes(n_items, grid_heads, head_dim, n_splits);
        HDS_REQUIRE(workspace != nullptr && workspace_bytes >= need, "paged_attention_prefill: workspace too small");
        const int64_t units = static_cast<int64_t>(n_items) * grid_heads;
        char* ws = reinterpret_cast<char*>(workspace);
        a.tickets = reinterpret_cast<int32_t*>(ws);
        a.ws_ml = reinterpret_cast<float*>(ws + pf_align256(units * 8 * 4));
        a.ws_o = reinterpret_cast<float*>(ws + pf_align256(units * 8 * 4) + pf_align256(units * n_splits * 128 * 2 * 4));
    }
    CUtensorMap tq, tkv;
    int rc;
    if (packed)   // q[T][heads][D]: one box = [128/G tokens][G heads][64 dims] -> 128 tile rows ordered (token, head)
        rc = make_tmap_3d_16b(&tq, q, static_cast<uint64_t>(head_dim), static_cast<uint64_t>(n_q_heads),
                              static_cast<uint64_t>(n_q_rows), static_cast<uint64_t>(head_dim) * 2,
                              static_cast<uint64_t>(n_q_heads) * head_dim * 2, 64, static_cast<uint32_t>(G),
                              static_cast<uint32_t>(128 / G), dtype == 0);
    else
        rc = make_tmap_2d_16b(&tq, q, static_cast<uint64_t>(n_q_heads) * head_dim, static_cast<uint64_t>(n_q_rows),
                              static_cast<uint64_t>(n_q_heads) * head_dim * 2, 64, 128, dtype == 0);
    if (rc) return rc;
    const uint64_t rows = static_cast<uint64_t>(num_blocks) * 2 * n_kv_heads * block_size;
    rc = make_tmap_2d_16b(&tkv, kv_cache_layer, head_dim, rows, static_cast<uint64_t>(head_dim) * 2, 64, block_size,
                          dtype == 0);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (packed)
        return dtype == 0 ? launch_prefill<128, true, true>(tq, tkv, a, n_items, s)
                          : launch_prefill<128, false, true>(tq, tkv, a, n_items, s);
    if (head_dim == 128 && block_size <= 64 && alibi == nullptr && window == 0 && prefill64_enabled())   // no ALiBi / window there
        return dtype == 0 ? launch_prefill64<true>(tq, tkv, a, n_items, s) : launch_prefill64<false>(tq, tkv, a, n_items, s);
    if (head_dim == 128)
        return dtype == 0 ? launch_prefill<128, true>(tq, tkv, a, n_items, s)
                          : launch_prefill<128, false>(tq, tkv, a, n_items, s);
    return dtype == 0 ? launch_prefill<64, true>(tq, tkv, a, n_items, s)
                      : launch_prefill<64, false>(tq, tkv, a, n_items, s);
}

extern "C" int hds_paged_attention_prefill(const void* q, void* out, const void* kv_cache_layer, int64_t num_blocks,
                                           const int32_t* block_tables, int max_blocks_per_seq,
                                           const hds_attn_item* items, int n_items, int n_q_rows, int n_q_heads,
                                           int n_kv_heads, int head_dim, int block_size, float scale, int n_splits,
                                           void* workspace, int64_t workspace_bytes, int dtype, void* stream) {
    return attention_tc_impl(false, q, out, kv_cache_layer, num_blocks, block_tables, max_blocks_per_seq, items, n_items,
                             n_q_rows, n_q_heads, n_kv_heads, head_dim, block_size, scale, n_splits, workspace,
                             workspace_bytes, dtype, stream);
}

extern "C" int hds_paged_attention_packed(const void* q, void* out, const void* kv_cache_layer, int64_t num_blocks,
                                          const int32_t* block_tables, int max_blocks_per_seq,
                                          const hds_attn_item* items, int n_items, int n_q_rows, int n_q_heads,
                                          int n_kv_heads, int head_dim, int block_size, float scale, int n_splits,
                                          void* workspace, int64_t workspace_bytes, int dtype, void* stream) {
    return attention_tc_impl(true, q, out, kv_cache_layer, num_blocks, block_tables, max_blocks_per_seq, items, n_items,
                             n_q_rows, n_q_heads, n_kv_heads, head_dim, block_size, scale, n_splits, workspace,
                             workspace_bytes, dtype, stream);
}
