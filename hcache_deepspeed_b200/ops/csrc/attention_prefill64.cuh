// tcgen05 ragged paged attention, 64-key tiles, TWO CTAs per SM (head_dim 128, KV pages of 16 / 32 / 64 tokens).
//
// Why a second shape of the same kernel: with one CTA per SM (attention_prefill.cu) a single query tile is in flight,
// so the tensor pipe idles while the softmax warps work on S and the softmax warps idle while P V runs — ncu showed the
// tensor pipe 23 % active (profiles/ncu_prefill.txt). Halving the key tile halves every buffer (K/V stages 2 x 32 KB,
// P 16 KB, S 64 TMEM columns), which lets TWO independent CTAs share an SM (2 x 112.3 KB of shared memory, 2 x 256 TMEM
// columns): one CTA's MMAs run under the other's exponentials without any extra synchronisation in the kernel.
//
//   warp 0    : TMA producer (Q once, then 64-key K/V tiles page by page through the block table, 2-stage ring)
//   warp 1    : TMEM allocator + tcgen05.mma issuer: S = Q K^T (128 x 64 x 128), O += P V (128 x 128 x 64)
//   warps 2-5 : softmax, ONE thread per query row (64 scores in registers): causal mask, online max / sum with lazy
//               rescale of O in TMEM, P -> shared memory (UMMA K-major swizzled panel)
// Split-KV, the work-item list and the workspace layout are those of attention_prefill.cu (PrefillArgs).
#pragma once
#include "hds_ptx.cuh"

namespace hds {

constexpr int kP64Threads = 192;
constexpr int kP64Tok = 64;
constexpr int kP64Stages = 2;
constexpr int kP64D = 128;
constexpr int kP64QPanel = 128 * 128;                      // [128 query rows x 64 dims]
constexpr int kP64KvPanel = kP64Tok * 128;                 // [64 keys x 64 dims]
constexpr int kP64QTile = 2 * kP64QPanel;                  // 32 KB
constexpr int kP64KvTile = 2 * kP64KvPanel;                // 16 KB
constexpr int kP64PBytes = 128 * 128;                      // [128 rows x 64 keys] = 16 KB
constexpr size_t kP64SmemBytes = kP64QTile + kP64PBytes + kP64Stages * 2 * kP64KvTile + 256;   // 114944 B: two fit in 227 KB

template <bool kBf16>
__global__ void __launch_bounds__(kP64Threads, 2)
paged_attn_prefill64_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                            const PrefillArgs a) {
    constexpr int D = kP64D;
    constexpr uint32_t TMEM_COLS = 256;  // S: 64 columns, O: 128 columns
    constexpr uint32_t IDESC_QK = make_idesc_f16(128, kP64Tok, kBf16, false, false);
    constexpr uint32_t IDESC_PV = make_idesc_f16(128, D, kBf16, false, true);  // B = V is MN-major

    extern __shared__ __align__(1024) uint8_t smem64[];
    uint8_t* q_sm = smem64;
    uint8_t* p_sm = q_sm + kP64QTile;
    uint8_t* kv_sm = p_sm + kP64PBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(kv_sm + kP64Stages * 2 * kP64KvTile);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;
    uint64_t* kv_empty = kv_full + kP64Stages;
    uint64_t* s_full = kv_empty + kP64Stages;
    uint64_t* p_full = s_full + 1;
    uint64_t* o_full = p_full + 1;
    uint64_t* s_empty = o_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + 1);

    const uint32_t warp = warp_id(), lane = lane_id();
    if (smem_u32(smem64) & 1023u) __trap();   // the swizzled tiles need a 1024-byte aligned base and there is no slack to re-align
    pdl_enter();
    const int item_idx = static_cast<int>(gridDim.x - 1 - blockIdx.x);   // heaviest items first (see attention_prefill.cu)
    const hds_attn_item item = a.items[item_idx];
    const int head = blockIdx.y;
    const int kvh = head / (a.n_q_heads / a.n_kv_heads);
    const int kv_len = item.pos0 + item.n_q;
    const int tiles_total = (kv_len + kP64Tok - 1) / kP64Tok;
    const int tps = (tiles_total + a.n_splits - 1) / a.n_splits;
    const int eff_splits = (tiles_total + tps - 1) / tps;
    const int split = blockIdx.z;
    if (split >= eff_splits) return;
    const int t0 = split * tps;
    const int n_tiles = min(tiles_total, t0 + tps) - t0;
    const int bs = a.block_size;
    const int bps = kP64Tok / bs;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_q);
        prefetch_tmap(&tmap_kv);
        mbar_init(q_full, 1);
        for (int i = 0; i < kP64Stages; ++i) {
            mbar_init(&kv_full[i], 1);
            mbar_init(&kv_empty[i], 1);
        }
        mbar_init(s_full, 1);
        mbar_init(p_full, 4);
        mbar_init(o_full, 1);
        mbar_init(s_empty, 4);
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
    const uint32_t tmem_o = tmem_base + 64;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(q_full, kP64QTile);
#pragma unroll
            for (int p = 0; p < 2; ++p)
                tma_load_2d(q_sm + p * kP64QPanel, &tmap_q, q_full, head * D + p * 64, item.q_row0, kEvictFirst);
            const int32_t* table = a.block_tables + static_cast<int64_t>(item.seq) * a.max_blocks_per_seq;
            const int last_blk_idx = (kv_len - 1) / bs;
            uint32_t stage = 0, phase = 0;
            for (int ti = 0; ti < n_tiles; ++ti) {
                mbar_wait(&kv_empty[stage], phase ^ 1);
                uint8_t* kdst = kv_sm + stage * 2 * kP64KvTile;
                uint8_t* vdst = kdst + kP64KvTile;
                mbar_expect_tx(&kv_full[stage], 2 * kP64KvTile);
                for (int b = 0; b < bps; ++b) {
                    // past the end of the sequence re-load its last page (masked later; keeps V finite under zero weights)
                    const int bi = min((t0 + ti) * bps + b, last_blk_idx);
                    const int blk = table[bi];
                    const int krow = ((blk * 2 + 0) * a.n_kv_heads + kvh) * bs;
                    const int vrow = ((blk * 2 + 1) * a.n_kv_heads + kvh) * bs;
#pragma unroll
                    for (int p = 0; p < 2; ++p) {
                        tma_load_2d(kdst + p * kP64KvPanel + b * bs * 128, &tmap_kv, &kv_full[stage], p * 64, krow, kEvictNormal);
                        tma_load_2d(vdst + p * kP64KvPanel + b * bs * 128, &tmap_kv, &kv_full[stage], p * 64, vrow, kEvictNormal);
                    }
                }
                if (++stage == kP64Stages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            auto issue_qk = [&](uint32_t stage) {
                const uint8_t* ktile = kv_sm + stage * 2 * kP64KvTile;
#pragma unroll
                for (int p = 0; p < 2; ++p) {
                    const uint64_t dq = make_desc_kmajor_sw128(smem_u32(q_sm + p * kP64QPanel));
                    const uint64_t dk = make_desc_kmajor_sw128(smem_u32(ktile + p * kP64KvPanel));
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
                const uint32_t nstage = (stage + 1 == kP64Stages) ? 0 : stage + 1;
                const uint32_t nphase = (stage + 1 == kP64Stages) ? phase ^ 1 : phase;
                if (ti + 1 < n_tiles) {   // S of the next tile as soon as the softmax warps hold this one in registers
                    mbar_wait(s_empty, ti & 1);
                    mbar_wait(&kv_full[nstage], nphase);
                    tc_fence_after();
                    issue_qk(nstage);
                }
                mbar_wait(p_full, ti & 1);
                tc_fence_after();
                const uint8_t* vtile = kv_sm + stage * 2 * kP64KvTile + kP64KvTile;
#pragma unroll
                for (int kk = 0; kk < kP64Tok / 16; ++kk) {
                    // A = P (K-major, one 64-key panel: 16 keys = 32 B); B = V (MN-major: 16 keys = 16 rows of 128 B; the two
                    // 64-dim panels are kP64KvPanel apart)
                    const uint64_t dp = make_desc_kmajor_sw128(smem_u32(p_sm)) + 2 * kk;
                    const uint64_t dv = make_desc_mnmajor_sw128(smem_u32(vtile + kk * 16 * 128), kP64KvPanel);
                    umma_f16_ss(tmem_o, dp, dv, IDESC_PV, (ti > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(o_full);
                umma_commit(&kv_empty[stage]);
                stage = nstage;
                phase = nphase;
            }
        }
    } else {
        // ===================== softmax / output: warps 2..5, one thread per query row =====================
        const uint32_t quad = warp & 3;
        const int r = quad * 32 + lane;
        const int q_pos = item.pos0 + r;
        const uint32_t lane_addr = (quad * 32u) << 16;
        float m_run = -INFINITY, l_run = 0.0f;
        uint8_t* p_row = p_sm + r * 128;
        constexpr float kRescaleThreshold = 8.0f;
        for (int ti = 0; ti < n_tiles; ++ti) {
            const int tok0 = (t0 + ti) * kP64Tok;
            const bool need_mask = tok0 + kP64Tok - 1 > item.pos0;  // warp-uniform
            mbar_wait(s_full, ti & 1);
            tc_fence_after();
            uint32_t sr[kP64Tok];
            tmem_ld_32x32(tmem_s + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&sr[0]));
            tmem_ld_32x32(tmem_s + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[32]));
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_empty);
            float tmax = -INFINITY;
            if (need_mask) {
#pragma unroll
                for (int j = 0; j < kP64Tok; ++j) {
                    const int kv_pos = tok0 + j;
                    const float sv = (kv_pos <= q_pos && kv_pos < kv_len) ? __uint_as_float(sr[j]) : -INFINITY;
                    sr[j] = __float_as_uint(sv);
                    tmax = fmaxf(tmax, sv);
                }
            } else {
#pragma unroll
                for (int j = 0; j < kP64Tok; ++j) tmax = fmaxf(tmax, __uint_as_float(sr[j]));
            }
            const float m_cand = fmaxf(m_run, tmax * a.scale_log2);
            const bool grow = (m_cand > m_run + kRescaleThreshold) || (m_run == -INFINITY && m_cand != -INFINITY);
            const float m_new = grow ? m_cand : m_run;
            const bool fresh = (m_run == -INFINITY);   // nothing accumulated for this row yet
            const float alpha = (!grow || fresh) ? (grow ? 0.0f : 1.0f) : ex2_approx(m_run - m_new);
            const float m_sub = (m_new == -INFINITY) ? 0.0f : m_new;
            float ps0 = 0.0f, ps1 = 0.0f;
            uint32_t packed[kP64Tok / 2];
#pragma unroll
            for (int j = 0; j < kP64Tok; j += 2) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(sr[j]), a.scale_log2, -m_sub));
                const float p1 = ex2_approx(fmaf(__uint_as_float(sr[j + 1]), a.scale_log2, -m_sub));
                ps0 += p0;
                ps1 += p1;
                packed[j >> 1] = pack2<kBf16>(p0, p1);
            }
            l_run = l_run * alpha + (ps0 + ps1);
            m_run = m_new;
            if (ti > 0) {   // P V of the previous tile must be complete before P is overwritten or O is rescaled
                mbar_wait(o_full, (ti - 1) & 1);
                tc_fence_after();
                // (alpha may underflow to exactly 0 when the maximum jumps by > 126 log2 units, e.g. under a steep ALiBi
                // slope: the stale accumulation must then still be wiped, so "fresh" is tracked separately)
                const bool rescale = grow && !fresh;
                if (__any_sync(0xffffffffu, rescale)) {
                    const float f = rescale ? alpha : 1.0f;
#pragma unroll
                    for (int c = 0; c < D; c += 32) {
                        uint32_t orr[32];
                        tmem_ld_32x32(tmem_o + lane_addr + c, orr);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) orr[j] = __float_as_uint(__uint_as_float(orr[j]) * f);
                        tmem_st_32x32(tmem_o + lane_addr + c, orr);
                    }
                    tmem_st_wait();
                }
            }
#pragma unroll
            for (int ch = 0; ch < kP64Tok / 8; ++ch) {   // 8 keys per 16-byte chunk, chunk index swizzled by row
                const uint4 v = make_uint4(packed[4 * ch], packed[4 * ch + 1], packed[4 * ch + 2], packed[4 * ch + 3]);
                *reinterpret_cast<uint4*>(p_row + ((ch ^ (r & 7)) << 4)) = v;
            }
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
        }
        mbar_wait(o_full, (n_tiles - 1) & 1);
        tc_fence_after();
        const bool row_valid = r < item.n_q;
        char* dst = reinterpret_cast<char*>(a.out) + ((static_cast<int64_t>(item.q_row0 + r) * a.n_q_heads + head) * D) * 2;
        if (eff_splits == 1) {
            const float inv = 1.0f / l_run;
#pragma unroll
            for (int c = 0; c < D; c += 32) {
                uint32_t orr[32];
                tmem_ld_32x32(tmem_o + lane_addr + c, orr);
                tmem_ld_wait();
                if (row_valid) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        uint4 v;
                        v.x = pack2<kBf16>(__uint_as_float(orr[j + 0]) * inv, __uint_as_float(orr[j + 1]) * inv);
                        v.y = pack2<kBf16>(__uint_as_float(orr[j + 2]) * inv, __uint_as_float(orr[j + 3]) * inv);
                        v.z = pack2<kBf16>(__uint_as_float(orr[j + 4]) * inv, __uint_as_float(orr[j + 5]) * inv);
                        v.w = pack2<kBf16>(__uint_as_float(orr[j + 6]) * inv, __uint_as_float(orr[j + 7]) * inv);
                        st_v4(dst + (c + j) * 2, v);
                    }
                }
            }
            tc_fence_before();
        } else {
            // ---- split-KV: park the partial (O, m, l); the last split of this quadrant merges and writes the rows ----
            const int64_t unit = static_cast<int64_t>(item_idx) * gridDim.y + head;
            const int64_t ws_row = (unit * a.n_splits + split) * 128 + r;
            const bool quad_live = static_cast<int>(quad * 32) < item.n_q;   // warp-uniform
#pragma unroll
            for (int c = 0; c < D; c += 32) {
                uint32_t orr[32];
                tmem_ld_32x32(tmem_o + lane_addr + c, orr);
                tmem_ld_wait();
                if (row_valid) {
                    float4* wo = reinterpret_cast<float4*>(a.ws_o + ws_row * D + c);
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        __stcg(wo + (j >> 2), make_float4(__uint_as_float(orr[j]), __uint_as_float(orr[j + 1]),
                                                          __uint_as_float(orr[j + 2]), __uint_as_float(orr[j + 3])));
                }
            }
            tc_fence_before();
            if (quad_live) {
                if (row_valid) __stcg(reinterpret_cast<float2*>(a.ws_ml + ws_row * 2), make_float2(m_run, l_run));
                __threadfence();
                __syncwarp();
                int last = 0;
                if (lane == 0) last = (atomicAdd(&a.tickets[(unit * 4 + quad) * 2], 1) == eff_splits - 1) ? 1 : 0;
                last = __shfl_sync(0xffffffffu, last, 0);
                if (last) {
                    __threadfence();
                    if (row_valid) {
                        float mx = -INFINITY;
                        for (int sp = 0; sp < eff_splits; ++sp)
                            mx = fmaxf(mx, __ldcg(a.ws_ml + ((unit * a.n_splits + sp) * 128 + r) * 2));
                        float l_tot = 0.0f;
                        for (int sp = 0; sp < eff_splits; ++sp) {
                            const float2 ml = __ldcg(reinterpret_cast<const float2*>(a.ws_ml + ((unit * a.n_splits + sp) * 128 + r) * 2));
                            l_tot = fmaf((ml.x == -INFINITY) ? 0.0f : ex2_approx(ml.x - mx), ml.y, l_tot);
                        }
                        const float inv = 1.0f / l_tot;
#pragma unroll 1
                        for (int c = 0; c < D; c += 32) {
                            float acc[32];
#pragma unroll
                            for (int j = 0; j < 32; ++j) acc[j] = 0.0f;
                            for (int sp = 0; sp < eff_splits; ++sp) {
                                const int64_t row_sp = (unit * a.n_splits + sp) * 128 + r;
                                const float mv = __ldcg(a.ws_ml + row_sp * 2);
                                const float w = (mv == -INFINITY) ? 0.0f : ex2_approx(mv - mx);
                                const float4* wo = reinterpret_cast<const float4*>(a.ws_o + row_sp * D + c);
#pragma unroll
                                for (int j = 0; j < 32; j += 4) {
                                    const float4 t = __ldcg(wo + (j >> 2));
                                    acc[j] = fmaf(w, t.x, acc[j]);
                                    acc[j + 1] = fmaf(w, t.y, acc[j + 1]);
                                    acc[j + 2] = fmaf(w, t.z, acc[j + 2]);
                                    acc[j + 3] = fmaf(w, t.w, acc[j + 3]);
                                }
                            }
#pragma unroll
                            for (int j = 0; j < 32; j += 8) {
                                uint4 v;
                                v.x = pack2<kBf16>(acc[j + 0] * inv, acc[j + 1] * inv);
                                v.y = pack2<kBf16>(acc[j + 2] * inv, acc[j + 3] * inv);
                                v.z = pack2<kBf16>(acc[j + 4] * inv, acc[j + 5] * inv);
                                v.w = pack2<kBf16>(acc[j + 6] * inv, acc[j + 7] * inv);
                                st_v4(dst + (c + j) * 2, v);
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) a.tickets[(unit * 4 + quad) * 2] = 0;
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
}

}  // namespace hds
