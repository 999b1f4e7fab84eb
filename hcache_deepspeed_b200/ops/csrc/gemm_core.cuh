// Persistent warp-specialised tcgen05 GEMM for sm_100a:  D[M,N] = A[M,K] * W[N,K]^T  (fp32 accumulate)
//
//   warp 0   : TMA producer  (cp.async.bulk.tensor, SWIZZLE_128B boxes, STAGES-deep mbarrier ring)
//   warp 1   : TMEM allocator + single-thread tcgen05.mma issuer (128 x BN x 16 per instruction)
//   warps 2-5: epilogue (tcgen05.ld from TMEM, fused epilogue, global stores / reductions)
//
// The accumulator is double buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps
// the main loop of tile i+1. Tiles are (m, n, k-split) triples walked persistently by
// gridDim.x <= #SM CTAs. Tokens sit on the M side (TMEM lanes): one epilogue thread owns one
// token row, which is what makes the RoPE + paged-KV epilogue a register-only operation.
//
// Replaces the reference's cuBLAS call + separate epilogue kernels
// (deepspeed/inference/v2/kernels/core_ops/blas_kernels/blas.h:31-136,
//  core_ops/bias_activations/bias_activation_cuda.cu, core_ops/gated_activations/
//  gated_activation_kernels_cuda.cu:44-88) and, for the KV epilogues,
//  ragged_ops/linear_blocked_kv_rotary/blocked_kv_rotary_cuda.cu:25-194.
#pragma once
#include "hds_ptx.cuh"

namespace hds {

enum GemmEpilogue : int {
    EPI_STORE = 0,      // out = act(acc + bias)
    EPI_GATED = 1,      // out[:, j] = act(acc[:, 2j] + b[2j]) * (acc[:, 2j+1] + b[2j+1])   (interleaved gate/up)
    EPI_RED_F32 = 2,    // accum[m, n] += acc (+bias on split 0): split-K and/or tensor-parallel partial sums
    EPI_KV_ROPE = 3,    // HCache restore: K (RoPE) and V written straight into paged KV blocks
    EPI_QKV_ROPE = 4,   // forward QKV: Q (RoPE) to a dense buffer, K (RoPE) / V into paged KV blocks
};

constexpr int kMaxPeers = 8;
constexpr int kGemmBM = 128;
constexpr int kGemmBK = 64;
constexpr int kGemmThreads = 192;

// Pipeline depth: as many (A, W) stages as fit in ~200 KB, at most 16. `am` is the number of A rows kept in
// shared memory: 128 normally; 16 / 64 for decode-size M, where the TMA box only delivers the live rows and
// the tensor core reads the remaining rows of its 128-row operand from whatever follows in shared memory
// (those rows only feed accumulator rows that are never stored). Shrinking A leaves room for 2x the weight
// bytes in flight, which is what a weight-streaming GEMM needs to cover HBM latency.
__host__ __device__ constexpr int gemm_stages(int bn, int am = 128) {
    return (200 * 1024) / (am * 128 + bn * 128) > 16 ? 16 : (200 * 1024) / (am * 128 + bn * 128);
}
__host__ __device__ constexpr uint32_t gemm_tmem_cols(int bn) {
    return 2 * bn <= 32 ? 32 : 2 * bn <= 64 ? 64 : 2 * bn <= 128 ? 128 : 2 * bn <= 256 ? 256 : 512;
}
__host__ __device__ constexpr size_t gemm_smem_bytes(int bn, int am = 128) {
    // stages + barriers (2*stages + 4 mbarriers, TMEM slot) + 1 KB alignment slack; with am < 128 the last A
    // stage is read 16 KB deep, which stays inside the W region that follows (>= 16 KB for every config used)
    return static_cast<size_t>(gemm_stages(bn, am)) * (am * kGemmBK * 2 + bn * kGemmBK * 2) + 512 + 1024;
}

// Activation operand spread over tensor-parallel peers: map i describes rows [i * rows_per_peer, ...) as they sit
// in rank i's buffer (a peer mapping of symmetric memory). The TMA producer picks the map from the tile's row
// range, so the all-gather of the operand happens inside the GEMM: tiles stream straight out of peer HBM over
// NVLink / NVSwitch into shared memory, no gathered copy is ever materialised.
struct TmapSet {
    CUtensorMap m[kMaxPeers];
};

struct GemmArgs {
    int M, N, K;
    int a_peers;           // > 0: A rows are sharded over this many peers (TmapSet), rows_per_peer multiple of 128
    int a_rows_per_peer;
    int splits;            // split-K factor (>=1, no empty split)
    int kb_per_split;      // k-blocks (of 64) per split
    uint32_t a_tx_bytes;   // bytes delivered by one A box (box rows * 128)
    int debug_skip;        // experiments only (HDS_GEMM_DEBUG): bit0 = no A loads, bit1 = no MMA issue (results invalid)
    int w_tiled;           // W is stored pre-tiled [N/32][K/64][32][64]: every TMA box is one contiguous 4 KB chunk
    int act;               // Activation
    // EPI_STORE / EPI_GATED
    void* out;
    int64_t ldo;           // elements
    const void* bias;      // [N] in the I/O dtype, may be null
    // EPI_RED_F32
    float* accum[kMaxPeers];  // accum[0] = local (or multicast) target; others = peer mappings
    int n_targets;            // number of valid accum pointers (1 for local / multicast)
    int red_mode;             // 0 = red.gpu  1 = red.sys to every accum[i]  2 = multimem.red on accum[0]
    int64_t ld_accum;
    // split-K for EPI_STORE / EPI_GATED (single m-tile only): fp32 partial sums + one ticket per
    // (n-tile, 32-row quadrant); the last split to arrive applies the epilogue and re-zeroes both
    float* splitk_ws;         // [M, N] fp32, zero between calls
    int32_t* splitk_tickets;  // [n_tiles * 4], zero between calls
    // EPI_KV_ROPE / EPI_QKV_ROPE
    void* kv_cache;           // this layer's cache: [blocks][2][n_kv_heads][block_size][head_dim]
    void* q_out;              // [M, n_q_heads * head_dim]
    const int32_t* tok_pos;   // [M] absolute position of each token in its sequence
    const int32_t* tok_slot;  // [M] block_id * block_size + in-block offset; < 0 = skip token
    const float* inv_freq;    // [rotary_dim / 2]
    int n_q_heads, n_kv_heads, block_size, do_rope;
    // grouped (mixture-of-experts) mode: A rows [group_pad_off[e], group_pad_off[e+1]) belong to expert e (ranges are
    // multiples of 128 so no tile straddles two experts); W stacks the experts' [N, K] matrices along its rows.
    // M is the row capacity of A / out; tiles at or beyond group_pad_off[n_groups] are skipped.
    const int32_t* group_pad_off;
    int n_groups;
    // tensor-parallel flag protocol (hds_ptx.cuh): `sync.wait` — the TMA producer acquires every rank's latest signal
    // before its first activation load (the operand was all-gathered into this GPU's buffer by the peers' norm kernels);
    // `sync.signal` — the last CTA to retire releases this kernel's output to the peers.
    TpSync sync;
    // kind::f8f6f4 kernels: bits OR-ed into the instruction descriptor to select a non-E4M3 weight format
    // (E3M2 = code 4 in the operand-format field of whichever side the weights are on)
    uint32_t idesc_or;
    uint32_t w_tx_bytes;   // bytes one weight-tile TMA load reports to the mbarrier (0 = the tile's shared-memory size)
};

// Expert that owns the 128-row tile starting at row0 (-1: past the last expert's rows).
__device__ __forceinline__ int gemm_group_of(const GemmArgs& a, int row0) {
    const int32_t* off = a.group_pad_off;
    if (row0 >= off[a.n_groups]) return -1;
    int lo = 0, hi = a.n_groups;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] <= row0) lo = mid;
        else hi = mid;
    }
    return lo;
}

template <bool kBf16>
__device__ __forceinline__ void store_row32(void* dst, const float (&v)[32]) {
    uint4* p = reinterpret_cast<uint4*>(dst);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        uint4 q;
        q.x = pack2<kBf16>(v[8 * i + 0], v[8 * i + 1]);
        q.y = pack2<kBf16>(v[8 * i + 2], v[8 * i + 3]);
        q.z = pack2<kBf16>(v[8 * i + 4], v[8 * i + 5]);
        q.w = pack2<kBf16>(v[8 * i + 6], v[8 * i + 7]);
        st_v4(p + i, q);
    }
}

// ---------------------------------------------------------------------------------------------
// Epilogues. `taddr` addresses this warp's 32 lanes at column 0 of the accumulator stage; `row`
// is the global token row owned by this thread (may be >= M: then only the collective tcgen05.ld
// is executed and nothing is written).
// ---------------------------------------------------------------------------------------------
// Apply the STORE epilogue to 32 consecutive accumulator columns [n, n+32) of one row.
template <bool kBf16>
__device__ __forceinline__ void store_chunk(const GemmArgs& a, int row, int n, float (&v)[32]) {
    if (a.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (n + j < a.N) v[j] += load_scalar<kBf16>(a.bias, n + j);
    }
    apply_act_vec<32>(v, a.act);
    const int64_t off = static_cast<int64_t>(row) * a.ldo + n;
    char* dst = reinterpret_cast<char*>(a.out) + off * 2;
    if (n + 32 <= a.N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        store_row32<kBf16>(dst, v);
    } else {
        for (int j = 0; j < 32 && n + j < a.N; ++j) store_scalar<kBf16>(a.out, off + j, v[j]);
    }
}

// GATED epilogue: N counts interleaved (gate, up) columns; the output has N/2 columns.
template <bool kBf16>
__device__ __forceinline__ void gated_chunk(const GemmArgs& a, int row, int n, float (&v)[32]) {
    float o[16], u[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        o[j] = v[2 * j];
        u[j] = v[2 * j + 1];
        if (a.bias != nullptr && n + 2 * j + 1 < a.N) {
            o[j] += load_scalar<kBf16>(a.bias, n + 2 * j);
            u[j] += load_scalar<kBf16>(a.bias, n + 2 * j + 1);
        }
    }
    apply_act_vec<16>(o, a.act);
#pragma unroll
    for (int j = 0; j < 16; ++j) o[j] *= u[j];
    const int64_t off = static_cast<int64_t>(row) * a.ldo + (n >> 1);
    char* dst = reinterpret_cast<char*>(a.out) + off * 2;
    if (n + 32 <= a.N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        uint4* p = reinterpret_cast<uint4*>(dst);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            uint4 q;
            q.x = pack2<kBf16>(o[8 * i + 0], o[8 * i + 1]);
            q.y = pack2<kBf16>(o[8 * i + 2], o[8 * i + 3]);
            q.z = pack2<kBf16>(o[8 * i + 4], o[8 * i + 5]);
            q.w = pack2<kBf16>(o[8 * i + 6], o[8 * i + 7]);
            st_v4(p + i, q);
        }
    } else {
        for (int j = 0; j < 16 && n + 2 * j + 1 < a.N; ++j) store_scalar<kBf16>(a.out, off + j, o[j]);
    }
}

// STORE / GATED epilogue of one tile. With split-K (a.splits > 1, single m-tile) every split adds its
// partial into the fp32 workspace; each warp takes a ticket for its 32-row quadrant and the last one
// reads the finished sums back (L2), applies the epilogue and leaves workspace + ticket zeroed.
template <int BN, bool kBf16, bool kGated>
__device__ __forceinline__ void epilogue_store(const GemmArgs& a, uint32_t taddr, int row, int n0, int n_t,
                                               uint32_t quad, uint32_t lane) {
    const bool split = a.splits > 1;
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c, r);
        tmem_ld_wait();
        const int n = n0 + c;
        if (row < a.M && n < a.N) {
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            if (!split) {
                if (kGated) gated_chunk<kBf16>(a, row, n, v);
                else store_chunk<kBf16>(a, row, n, v);
            } else {
                float* dst = a.splitk_ws + static_cast<int64_t>(row) * a.N + n;
                if (n + 32 <= a.N && (a.N & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) red_add_v4_f32(dst + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
                } else {
                    for (int j = 0; j < 32 && n + j < a.N; ++j) atomicAdd(dst + j, v[j]);
                }
            }
        }
    }
    if (!split) return;
    if (static_cast<int>(quad * 32) >= a.M) return;  // no valid row in this quadrant (warp-uniform)
    __threadfence();
    __syncwarp();
    int last = 0;
    if (lane == 0) last = (atomicAdd(&a.splitk_tickets[n_t * 4 + quad], 1) == a.splits - 1) ? 1 : 0;
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    __threadfence();
    if (row < a.M) {
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
            const int n = n0 + c;
            if (n >= a.N) break;
            float* src = a.splitk_ws + static_cast<int64_t>(row) * a.N + n;
            float v[32];
            if (n + 32 <= a.N && (a.N & 3) == 0) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const float4 t = __ldcg(reinterpret_cast<const float4*>(src + j));
                    v[j] = t.x; v[j + 1] = t.y; v[j + 2] = t.z; v[j + 3] = t.w;
                    __stcg(reinterpret_cast<float4*>(src + j), make_float4(0.f, 0.f, 0.f, 0.f));
                }
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    v[j] = (n + j < a.N) ? __ldcg(src + j) : 0.0f;
                    if (n + j < a.N) __stcg(src + j, 0.0f);
                }
            }
            if (kGated) gated_chunk<kBf16>(a, row, n, v);
            else store_chunk<kBf16>(a, row, n, v);
        }
    }
    __syncwarp();
    if (lane == 0) a.splitk_tickets[n_t * 4 + quad] = 0;
}

template <int BN, bool kBf16>
__device__ __forceinline__ void epilogue_red(const GemmArgs& a, uint32_t taddr, int row, int n0, bool first_split) {
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c, r);
        tmem_ld_wait();
        const int n = n0 + c;
        if (row < a.M && n < a.N) {
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            if (a.bias != nullptr && first_split) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (n + j < a.N) v[j] += load_scalar<kBf16>(a.bias, n + j);
            }
            const int64_t off = static_cast<int64_t>(row) * a.ld_accum + n;
            const bool vec = (n + 32 <= a.N) && ((off & 3) == 0);
            if (a.red_mode == 2) {
                float* dst = a.accum[0] + off;
                if (vec) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) multimem_red_add_v4_f32(dst + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
                } else {
                    for (int j = 0; j < 32 && n + j < a.N; ++j)
                        asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(dst + j), "f"(v[j])
                                     : "memory");
                }
            } else if (a.red_mode == 1) {
                for (int t = 0; t < a.n_targets; ++t) {
                    float* dst = a.accum[t] + off;
                    if (vec) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            red_add_v4_f32_sys(dst + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
                    } else {
                        for (int j = 0; j < 32 && n + j < a.N; ++j)
                            asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(dst + j), "f"(v[j])
                                         : "memory");
                    }
                }
            } else {
                float* dst = a.accum[0] + off;
                if (vec) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) red_add_v4_f32(dst + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
                } else {
                    for (int j = 0; j < 32 && n + j < a.N; ++j) atomicAdd(dst + j, v[j]);
                }
            }
        }
    }
}

// A tile holds BN / HD whole attention heads. Rotate-half RoPE pairs column i with i + HD/2 of the
// same head, both of which live in this thread's accumulator row.
template <int BN, int HD, bool kBf16, bool kHasQ>
__device__ __forceinline__ void epilogue_kv_rope(const GemmArgs& a, uint32_t taddr, int row, int n_tile) {
    constexpr int HALF = HD / 2;
    constexpr int HEADS = BN / HD;
    static_assert(HALF % 32 == 0 && BN % HD == 0, "head_dim must be 64 or 128 and divide the tile width");
    const int n_q = kHasQ ? a.n_q_heads : 0;
    const int total_heads = n_q + 2 * a.n_kv_heads;
    int pos = 0, slot = -1;
    if (row < a.M) {
        pos = a.tok_pos[row];
        slot = a.tok_slot[row];
    }
#pragma unroll 1
    for (int hh = 0; hh < HEADS; ++hh) {
        const int gh = n_tile * HEADS + hh;  // global head index over [Q heads | K heads | V heads]
        if (gh >= total_heads) break;         // warp-uniform
        const bool is_q = gh < n_q;
        const int kv_idx = gh - n_q;
        const bool is_k = !is_q && kv_idx < a.n_kv_heads;
        const int head = is_q ? gh : (is_k ? kv_idx : kv_idx - a.n_kv_heads);
        const bool rope = a.do_rope && (is_q || is_k);
        char* dst = nullptr;
        if (row < a.M) {
            if (is_q) {
                dst = reinterpret_cast<char*>(a.q_out) + (static_cast<int64_t>(row) * a.n_q_heads + head) * HD * 2;
            } else if (slot >= 0) {
                const int64_t blk = slot / a.block_size, off = slot % a.block_size;
                const int64_t e = (((blk * 2 + (is_k ? 0 : 1)) * a.n_kv_heads + head) * a.block_size + off) * HD;
                dst = reinterpret_cast<char*>(a.kv_cache) + e * 2;
            }
        }
        const int col0 = gh * HD;
#pragma unroll 1
        for (int c = 0; c < HALF; c += 32) {
            uint32_t rl[32], rh[32];
            tmem_ld_32x32(taddr + hh * HD + c, rl);
            tmem_ld_32x32(taddr + hh * HD + c + HALF, rh);
            tmem_ld_wait();
            if (dst != nullptr) {
                float lo[32], hi[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    lo[j] = __uint_as_float(rl[j]);
                    hi[j] = __uint_as_float(rh[j]);
                }
                if (a.bias != nullptr) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        lo[j] += load_scalar<kBf16>(a.bias, col0 + c + j);
                        hi[j] += load_scalar<kBf16>(a.bias, col0 + c + HALF + j);
                    }
                }
                if (rope) {
#pragma unroll 4
                    for (int j = 0; j < 32; ++j) {
                        float sn, co;
                        rope_sincos(static_cast<float>(pos) * a.inv_freq[c + j], &sn, &co);
                        const float x = lo[j], y = hi[j];
                        lo[j] = x * co - y * sn;
                        hi[j] = y * co + x * sn;
                    }
                }
                store_row32<kBf16>(dst + c * 2, lo);
                store_row32<kBf16>(dst + (c + HALF) * 2, hi);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// The kernel
// ---------------------------------------------------------------------------------------------
template <int BN, int EPI, bool kBf16, int HD = 128, int AM = 128>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tn_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
               const GemmArgs args, const __grid_constant__ TmapSet a_peer_maps) {
    constexpr int STAGES = gemm_stages(BN, AM);
    constexpr uint32_t A_BYTES = AM * kGemmBK * 2;
    static_assert(AM == 128 || static_cast<uint32_t>(STAGES) * BN * kGemmBK * 2 >= 16384,
                  "compact-A configs need >= 16 KB of W stages behind the A stages");
    constexpr uint32_t W_BYTES = BN * kGemmBK * 2;
    constexpr uint32_t TMEM_COLS = gemm_tmem_cols(BN);
    constexpr uint32_t IDESC = make_idesc_f16(kGemmBM, BN, kBf16, false, false);
    static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N for M=128 must be a multiple of 16 in [16,256]");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_w = smem + STAGES * A_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + W_BYTES));
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const uint32_t warp = warp_id();
    const uint32_t lane = lane_id();

    const int m_tiles = (args.M + kGemmBM - 1) / kGemmBM;
    const int n_tiles = (args.N + BN - 1) / BN;
    const int kb_total = (args.K + kGemmBK - 1) / kGemmBK;
    const int total_work = m_tiles * n_tiles * args.splits;

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

    // PDL: the set-up above overlapped the previous kernel's tail; from here on its results are needed
    griddep_launch_dependents();
    griddep_wait();

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            if (args.sync.wait) {
                tp_wait_thread(args.sync);
                fence_proxy_async_global();
            }
            uint32_t stage = 0, phase = 0;
            for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
                const int m_t = w % m_tiles;
                const int rest = w / m_tiles;
                const int n_t = rest % n_tiles;
                const int sp = rest / n_tiles;
                const int kb0 = sp * args.kb_per_split;
                const int kb1 = min(kb_total, kb0 + args.kb_per_split);
                int w_row0 = n_t * BN;
                if (args.group_pad_off != nullptr) {
                    const int e = gemm_group_of(args, m_t * kGemmBM);
                    if (e < 0) continue;
                    w_row0 += e * args.N;
                }
                // Single m-tile (decode): every CTA needs the SAME activation k-block at the same moment, i.e. 148
                // requests for a handful of L2 lines per step — the lines' L2 slices serialise them (~0.3 us per
                // k-block, independent of tile size or pipeline depth). Starting each CTA at a different k offset
                // spreads the requests over all slices; summation order is irrelevant to the accumulator.
                const int nkb = kb1 - kb0;
                const int rot = (m_tiles == 1) ? static_cast<int>((blockIdx.x * 5u) % static_cast<unsigned>(nkb)) : 0;
                for (int it = 0; it < nkb; ++it) {
                    const int kb = kb0 + (it + rot < nkb ? it + rot : it + rot - nkb);
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], ((args.debug_skip & 1) ? 0u : args.a_tx_bytes) + W_BYTES);
                    // activations are re-read by every n-tile: keep them in L2; weights stream once per m-wave
                    if (!(args.debug_skip & 1)) {
                        const CUtensorMap* ta = &tmap_a;
                        int a_row = m_t * kGemmBM;
                        if (args.a_peers > 0) {
                            const int peer = a_row / args.a_rows_per_peer;
                            ta = &a_peer_maps.m[peer];
                            a_row -= peer * args.a_rows_per_peer;
                        }
                        tma_load_2d(smem_a + stage * A_BYTES, ta, &full_bar[stage], kb * kGemmBK, a_row, kEvictLast);
                    }
                    if (args.w_tiled)
                        tma_load_3d(smem_w + stage * W_BYTES, &tmap_w, &full_bar[stage], 0, kb * 32, n_t * (BN / 32),
                                    m_tiles > 1 ? kEvictNormal : kEvictFirst);
                    else
                        tma_load_2d(smem_w + stage * W_BYTES, &tmap_w, &full_bar[stage], kb * kGemmBK, w_row0,
                                    m_tiles > 1 ? kEvictNormal : kEvictFirst);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
            for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
                const int sp = (w / m_tiles) / n_tiles;
                const int kb0 = sp * args.kb_per_split;
                const int kb1 = min(kb_total, kb0 + args.kb_per_split);
                if (args.group_pad_off != nullptr && gemm_group_of(args, (w % m_tiles) * kGemmBM) < 0) continue;
                mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint64_t da = make_desc_kmajor_sw128(smem_u32(smem_a + stage * A_BYTES));
                    const uint64_t dw = make_desc_kmajor_sw128(smem_u32(smem_w + stage * W_BYTES));
                    if (!(args.debug_skip & 2)) {
#pragma unroll
                        for (int k = 0; k < kGemmBK / 16; ++k) {
                            // +32 B per K=16 step inside the 128 B swizzle atom (address field is in 16 B units)
                            umma_f16_ss(d_tmem, da + 2 * k, dw + 2 * k, IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
                        }
                    }
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
        // ===================== epilogue (warps 2..5) =====================
        const uint32_t quad = warp & 3;  // TMEM lane quadrant this warp may access
        uint32_t acc = 0, acc_phase = 0;
        for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
            const int m_t = w % m_tiles;
            const int rest = w / m_tiles;
            const int n_t = rest % n_tiles;
            const int sp = rest / n_tiles;
            if (args.group_pad_off != nullptr && gemm_group_of(args, m_t * kGemmBM) < 0) continue;
            mbar_wait(&tfull_bar[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + acc * BN;
            const int row = m_t * kGemmBM + quad * 32 + lane;
            if constexpr (EPI == EPI_STORE)
                epilogue_store<BN, kBf16, false>(args, taddr, row, n_t * BN, n_t, quad, lane);
            else if constexpr (EPI == EPI_GATED)
                epilogue_store<BN, kBf16, true>(args, taddr, row, n_t * BN, n_t, quad, lane);
            else if constexpr (EPI == EPI_RED_F32)
                epilogue_red<BN, kBf16>(args, taddr, row, n_t * BN, sp == 0);
            else if constexpr (EPI == EPI_KV_ROPE)
                epilogue_kv_rope<BN, HD, kBf16, false>(args, taddr, row, n_t);
            else
                epilogue_kv_rope<BN, HD, kBf16, true>(args, taddr, row, n_t);
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
