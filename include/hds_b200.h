/* hds_b200.h — C ABI of libhds_b200.so, the sm_100a kernel library of hcache_deepspeed_b200.
 *
 * Conventions
 *   - every function returns 0 on success, a positive cudaError_t, or a negative HDS error code;
 *     hds_last_error() returns a human-readable message for the calling thread's last failure.
 *   - `dtype`: 0 = bfloat16, 1 = float16 (activations and weights share the dtype; accumulation is fp32).
 *   - `stream` is a cudaStream_t passed as void*; all launches are asynchronous on that stream.
 *   - pointers are device pointers unless the parameter name says host.
 *   - weights are [out_features, in_features] row-major (the HF / DeepSpeed layout), i.e. y = x * W^T.
 *   - paged KV cache of ONE layer: [num_blocks][2 (K,V)][n_kv_heads][block_size][head_dim];
 *     `tok_slot[t] = block_id * block_size + offset` addresses a token inside it (negative = skip),
 *     `tok_pos[t]` is the token's absolute position in its sequence (RoPE angle / learned position).
 *
 * Citations name the reference component each entry point stands in for
 * (paths relative to the reference's deepspeed/inference/v2/kernels/ unless stated otherwise).
 */
#ifndef HDS_B200_H
#define HDS_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* hds_last_error(void);

/* ---- tcgen05 GEMMs ------------------------------------------------------------------------- */

/* out = epilogue(A[M,K] * W[N,K]^T + bias).  epilogue 0: act();  1: gated — columns are interleaved
 * (gate, up) pairs, out has N/2 columns: out[:, j] = act(c[2j]) * c[2j+1].  act: 0 id, 1 gelu, 2 relu,
 * 3 silu, 4 gelu-tanh.  bn: tile width override (0 = heuristic; -1 = force the
 * feature-major weight-streaming kernel of gemm_skinny.cuh, M <= 64 only).
 * Ref: core_ops/blas_kernels/blas.h:31-136 + bias_activations + gated_activations.
 * Split-K (decode-size M <= 128): pass a zero-initialised scratch buffer of at least
 * hds_gemm_splitk_workspace_bytes(M, N) bytes; partial sums meet there and the last split applies the
 * epilogue and leaves the buffer zeroed again. splits: 0 = heuristic, 1 = off. Without scratch: no split. */
int64_t hds_gemm_splitk_workspace_bytes(int M, int N);
int hds_gemm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int dtype, int epilogue,
             int act, const void* bias, void* out, int64_t ldo, int bn, int splits, void* splitk_ws,
             int64_t splitk_ws_bytes, void* stream);

/* accum[M,N] (fp32) += A * W^T (+ bias once).  Split-K partial sums and tensor-parallel partial sums share
 * this path: red_mode 0 = local red.add; 1 = red.add (system scope) into each of n_targets peer mappings of
 * the same buffer (GEMM -> all-reduce over NVLink from inside the epilogue); 2 = multimem.red on accum_ptrs[0],
 * an NVLS multicast address (in-switch reduction).  accum_ptrs is a HOST array of device pointers.
 * The consumer (hds_norm with delta_is_f32) reads and re-zeroes the buffer.
 * Ref: blas.h + the dist.all_reduce that follows it in model_implementations/llama_v2/model.py:159-169. */
int hds_gemm_accum(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int dtype,
                   const void* bias, float* const* accum_ptrs, int n_targets, int red_mode, int64_t ld_accum, int bn,
                   int splits, void* stream);

/* HCache restore: latents[T,H] * [W_k; W_v]^T with RoPE on K, written straight into the paged KV blocks.
 * w_kv rows: all K heads then all V heads (head_dim rows each). head_dim 64 or 128.
 * Ref: model_implementations/llama_v2/model.py:222-244 (full QKV GEMM) +
 *      ragged_ops/linear_blocked_kv_rotary/blocked_kv_rotary_cuda.cu:25-194. */
int hds_restore_kv_proj(const void* latents, int64_t ld_latents, const void* w_kv, int64_t ldw, int T, int H,
                        int dtype, const void* bias_kv, void* kv_cache_layer, const int32_t* tok_pos,
                        const int32_t* tok_slot, const float* inv_freq, int n_kv_heads, int head_dim, int block_size,
                        int do_rope, void* stream);

/* Tensor-parallel restore with the hidden-state all-gather fused into the projection: rank p holds rows
 * [p*rows_per_peer, (p+1)*rows_per_peer) of the latents in ITS memory (latent_ptrs[p] = that buffer as mapped in the
 * calling process: symmetric memory / CUDA IPC / cuMemMap peer mapping). The kernel's TMA producer reads every tile
 * straight from the owning peer over NVLink; each rank projects all T tokens onto its own KV heads.
 * latent_ptrs is a HOST array; rows_per_peer is a multiple of 128. The caller orders the launch after every peer's
 * buffer is filled (device barrier). Ref: SURVEY 2.8 "(new) restore under TP" — absent in the reference. */
int hds_restore_kv_proj_peers(const void* const* latent_ptrs, int n_peers, int rows_per_peer, int64_t ld_latents,
                              const void* w_kv, int64_t ldw, int T, int H, int dtype, const void* bias_kv,
                              void* kv_cache_layer, const int32_t* tok_pos, const int32_t* tok_slot,
                              const float* inv_freq, int n_kv_heads, int head_dim, int block_size, int do_rope,
                              void* stream);

/* Forward QKV projection with the same fused epilogue: Q (RoPE) -> q_out[T, n_q*d]; K (RoPE), V -> paged cache.
 * w_qkv rows: Q heads, K heads, V heads. */
int hds_qkv_rope_proj(const void* x, int64_t ldx, const void* w_qkv, int64_t ldw, int T, int H, int dtype,
                      const void* bias_qkv, void* q_out, void* kv_cache_layer, const int32_t* tok_pos,
                      const int32_t* tok_slot, const float* inv_freq, int n_q_heads, int n_kv_heads, int head_dim,
                      int block_size, int do_rope, void* stream);

/* ---- normalisation ------------------------------------------------------------------------- */

/* y = norm(x [+ delta]) * gamma (+ beta).  layer_norm 0 = RMSNorm, 1 = LayerNorm.
 * delta (optional): 16-bit or fp32 (delta_is_f32); zero_delta clears an fp32 delta after reading it.
 * res_out (optional) receives x + delta (pre-norm residual) or, with post_norm, the normalised value.
 * row_idx (optional, excludes delta): gather — row i of the output is computed from x[row_idx[i]].
 * Ref: core_ops/cuda_rms_norm/rms_norm_cuda.cu:18-164, core_ops/cuda_layer_norm/layer_norm_cuda.cu:30-325,
 *      ragged_ops/logits_gather/logits_gather_cuda.cu:18-50 (the gather). */
int hds_norm(const void* x, int64_t ld_in, void* delta, int delta_is_f32, int zero_delta, void* res_out, void* out,
             int64_t ld_out, const void* gamma, const void* beta, const int32_t* row_idx, int rows, int cols,
             float eps, int layer_norm, int post_norm, int dtype, void* stream);

/* ---- tensor parallelism: this library's own flag protocol over symmetric memory ------------------------------------
 *
 * Every rank owns a 256-byte, zero-initialised "pad" in memory that all ranks of the group have mapped (symmetric
 * memory / CUDA IPC / cuMemMap): 8 arrive counters written only by the respective peer, plus a ticket and a signal
 * count used locally. A SIGNAL ("everything my kernels wrote so far is visible to you") is issued by the last CTA of
 * a kernel to retire; a WAIT spins on the local pad until every peer has issued as many signals as this rank has.
 * All ranks must run the same sequence of signalling kernels. Waits are bounded (~17 s) and trap on timeout.
 * No reference counterpart: the reference calls NCCL (model_implementations/llama_v2/model.py:159-160, 168-169, 195-197).
 *
 * hds_tp_ctx_create: pad_ptrs is a HOST array [n_ranks] of this rank's mappings of every rank's pad
 * (pad_ptrs[rank] = the local one). Returns a context id >= 0, or a negative error.
 * hds_tp_arm: attaches a wait and / or a signal to the NEXT kernel this host thread launches through hds_gemm,
 * hds_gemm_accum, hds_gemm_2sm, hds_gemm_fp8, hds_gemm_fp8_skinny, hds_moe_gemm, hds_qkv_rope_proj,
 * hds_restore_kv_proj, hds_quantize_fp8_rows (wait only) or hds_p2p_gather_rows (wait only):
 *   wait   - the kernel's TMA producer (or every CTA) acquires all ranks' flags before it reads activations that
 *            the peers wrote into this GPU's buffer (all-gather -> GEMM);
 *   signal - the kernel's last CTA releases its output to the peers (GEMM -> all-reduce / reduce-scatter).
 * hds_tp_disarm returns 1 if an armed state was still pending (i.e. the intended launch did not consume it).
 * hds_tp_signal / hds_tp_wait: one-CTA kernels; hds_tp_barrier = signal followed by wait. */
int hds_tp_ctx_create(int rank, int n_ranks, void* const* pad_ptrs);
int hds_tp_ctx_destroy(int ctx);
int hds_tp_arm(int ctx, int wait, int signal);
int hds_tp_disarm(void);
int hds_tp_signal(int ctx, void* stream);
int hds_tp_wait(int ctx, void* stream);
int hds_tp_barrier(int ctx, void* stream);

/* The collective after a row-parallel GEMM, fused with residual add + norm. src_ptrs (HOST array [n_ranks]): this
 * rank's mappings of every rank's partial-sum buffer [rows_total, cols] (fp32, or the 16-bit I/O dtype when
 * partial_16bit); src_mc: NVLS multicast alias of that buffer or NULL (then peers are read one by one, rank order).
 *   out_ptrs == NULL and out_mc == NULL : all-reduce shape. Rows [row0, row0+n_rows) (normally all of them) are reduced
 *       over all ranks, added to x (the local residual replica), written to res_out and normalised into the local `out`.
 *   otherwise                           : reduce-scatter + all-gather shape. This rank reduces only rows
 *       [row0, row0+n_rows) — the rows it owns —, keeps their residual in res_out, and writes the normalised rows into
 *       every rank's `out` (out_mc: multicast alias -> multimem.st; else out_ptrs[r] one by one). Use signal = 1 and arm
 *       the consuming GEMM with wait = 1.
 * zero_buf (optional, local fp32 [rows_total, cols]): cleared — the idle half of the alternating accumulation pair.
 * wait = 1: acquire every rank's flag first (the producing GEMMs were armed with signal = 1).
 * Ref: dist.all_reduce + residual + norm of model_implementations/llama_v2/model.py:159-171. */
int hds_norm_tp(int ctx, int wait, int signal, const void* x, int64_t ld_in, const void* const* src_ptrs,
                const void* src_mc, int partial_16bit, float* zero_buf, void* res_out, void* out, int64_t ld_out,
                void* const* out_ptrs, void* out_mc, int row0, int n_rows, int rows_total, const void* gamma,
                const void* beta, int cols, float eps, int layer_norm, int dtype, void* stream);

/* All-gather of column-sharded rows (vocabulary-sharded logits): out[row, col0[p] : col0[p]+width[p]] = rank p's
 * slice buffer [rows, ld_src] for every p, pulled with 16-byte loads over NVLink. src_ptrs: HOST array of this rank's
 * mappings; col0 / width: HOST arrays [n_ranks] in elements. wait = 1: acquire the flags the LM-head GEMMs signalled.
 * Ref: dist.all_gather_into_tensor + permute copy, model_implementations/llama_v2/model.py:195-197. */
int hds_tp_gather_cols(int ctx, int wait, void* out, int64_t ld_out, const void* const* src_ptrs, int64_t ld_src,
                       const int32_t* col0, const int32_t* width, int rows, int elem_bytes, void* stream);

/* ---- ragged helpers ------------------------------------------------------------------------ */

/* Unfused RoPE (rotate-half over the first rotary_dim dims) + paged KV append from a dense qkv[T, (nq+2nkv)*d].
 * Ref: ragged_ops/linear_blocked_kv_rotary/blocked_kv_rotary_cuda.cu:25-194. */
int hds_rope_kv_append(const void* qkv, int64_t ld_qkv, void* q_out, void* kv_cache_layer, const int32_t* tok_pos,
                       const int32_t* tok_slot, const float* inv_freq, int T, int n_q_heads, int n_kv_heads,
                       int head_dim, int rotary_dim, int block_size, int do_rope, int dtype, void* stream);

/* out[t] = W[ids[t]] (+ P[clamp(tok_pos[t] + pos_offset)]).  Ref: ragged_ops/embed/embed_cuda.cu:18-80. */
int hds_embed(const int32_t* ids, const void* W, const void* P, const int32_t* tok_pos, int pos_offset, int max_pos,
              void* out, int T, int H, int vocab, int dtype, void* stream);

/* out[i] = x[row_idx[i]].  Ref: ragged_ops/logits_gather/logits_gather_cuda.cu:18-50. */
int hds_gather_rows(const void* x, const int32_t* row_idx, void* out, int n_rows, int H, void* stream);

/* Greedy sampling: arg-max per row (ties -> lowest index). dtype 0 bf16, 1 fp16, 2 fp32.  No reference
 * counterpart: the reference leaves sampling to its caller (DeepSpeed-MII). */
int hds_argmax(const void* logits, int64_t ld, int n_rows, int V, int32_t* out_ids, float* out_vals, int dtype,
               void* stream);

/* Copy whole KV blocks (all layers) between the paged cache and a dense staging buffer — the device half of
 * plain KV offload / reload. Ref: ragged/kv_cache.py:166-184 (offload/restore raise NotImplementedError). */
int hds_kv_blocks_copy(void* cache, void* staging, const int32_t* block_ids, int n_sel, int n_layers,
                       int64_t block_bytes, int64_t layer_stride_bytes, int to_staging, void* stream);

/* Row gather across tensor-parallel peers: rows [p*rows_per_peer, (p+1)*rows_per_peer) of `dst` are copied from the
 * same rows of peer p's buffer (peer_ptrs[p]: HOST array of this rank's mappings), except this rank's own rows.
 * hds_tp_arm(ctx, 1, 0) before the call makes the kernel wait for every rank's "rows landed" signal first.
 * No reference counterpart (the reference has no TP restore). */
int hds_p2p_gather_rows(void* dst, const void* const* peer_ptrs, int n_peers, int rows_per_peer, int n_rows,
                        int64_t row_bytes, int my_rank, void* stream);

/* dst (16-bit) = src (fp32); optionally zero src. */
int hds_cvt_f32(float* src, void* dst, int64_t n, int zero_src, int dtype, void* stream);

/* ---- ragged paged attention ---------------------------------------------------------------- */

/* Work item of the attention kernels (host-built by the batch scheduler, 4 x int32):
 *   seq      index into block_tables rows
 *   q_row0   first query token (row of q / out) of this item
 *   n_q      query tokens in this item (consecutive rows, consecutive positions)
 *   pos0     absolute position of the first query token; token i attends to KV positions <= pos0 + i
 * Ref: ragged_ops/includes/attention_atom.h:11-39 + atom_builder/atom_builder.cpp:10-53. */
typedef struct {
    int32_t seq, q_row0, n_q, pos0;
} hds_attn_item;

/* Scratch size (bytes) hds_paged_attention_splitkv needs for n_items work items. */
int64_t hds_paged_attention_splitkv_workspace_bytes(int n_items, int n_q_heads, int n_kv_heads, int head_dim,
                                                    int max_splits, int rows_per_item);

/* Split-KV paged attention for items with few query rows (decode, short continuations):
 * rows_per_item = max over items of n_q * (n_q_heads / n_kv_heads), at most 8. K/V pages stream through TMA (SWIZZLE_128B) into a
 * 3-stage shared-memory ring; partial (O, m, l) of each KV split land in `workspace`, and the last split to
 * finish merges them (no second kernel). `workspace` must be zero-initialised once (hds_memset or cudaMemset);
 * the kernel leaves its counters at zero again.
 * q / out: [T, n_q_heads, head_dim]; block_tables: [n_seqs, max_blocks_per_seq] int32.
 * Ref: the external blocked-flash kernel behind ragged_ops/blocked_flash/blocked_flash.cpp:21-101. */
int hds_paged_attention_splitkv(const void* q, void* out, const void* kv_cache_layer, int64_t num_blocks,
                                const int32_t* block_tables, int max_blocks_per_seq, const hds_attn_item* items,
                                int n_items, int rows_per_item, int n_q_heads, int n_kv_heads, int head_dim,
                                int block_size, int n_splits, float scale, void* workspace, int64_t workspace_bytes,
                                int dtype, void* stream);

/* Split-KV paged attention that allocates nothing and takes no scratch pointer: the partial results and the
 * arrival counters live in a fixed 8 MiB device array that is part of the library image
 * (hds_paged_attention_splitkv_static_capacity() bytes, held by every GPU that loads the library until exit).
 * Meant for callers that cannot obtain scratch memory: processes that handed all free HBM to the KV cache,
 * C/C++ hosts without an allocator to borrow from, loops that cannot tolerate cudaMalloc's device-wide sync.
 * Same arguments and results as hds_paged_attention_splitkv minus workspace / workspace_bytes.
 * Returns -4 (HDS_ERR_TOO_LARGE) with an explanatory hds_last_error() when the problem needs more scratch
 * than the array holds. The array is shared by every call on the device: calls that overlap on the GPU race;
 * issue them on one stream. The first call on a device resolves the array's device address — make it during
 * start-up, before any stream capture begins on that device.
 * This entry point has no counterpart in the reference. */
int64_t hds_paged_attention_splitkv_static_capacity(void);
int hds_paged_attention_splitkv_static(const void* q, void* out, const void* kv_cache_layer, int64_t num_blocks,
                                       const int32_t* block_tables, int max_blocks_per_seq,
                                       const hds_attn_item* items, int n_items, int rows_per_item, int n_q_heads,
                                       int n_kv_heads, int head_dim, int block_size, int n_splits, float scale,
                                       int dtype, void* stream);

/* Host staging buffers with explicit flags (ref: ragged/csrc/fast_host_buffer.cu). Always portable pinned memory;
 * `mapped`: also addressable by the GPU (hds_host_buffer_device_ptr), `write_combined`: host stores bypass the CPU
 * caches (fast to write and to DMA from, slow for the host to read back). */
int hds_host_buffer_alloc(int64_t bytes, int mapped, int write_combined, void** out);
int hds_host_buffer_device_ptr(void* host_ptr, void** out);
int hds_host_buffer_free(void* host_ptr);

/* ALiBi (Bloom, Falcon-RW, MPT): slopes = device pointer to n_q_heads floats. The NEXT hds_paged_attention_splitkv /
 * _prefill / _packed launch of the calling thread adds slope[head] * (key position - query position) to every score
 * before the softmax, then the setting is cleared. (hds_paged_attention_splitkv_static ignores it.) */
int hds_attention_next_alibi(const float* slopes);

/* Sliding-window attention (Mistral, Phi-3, GPT-Neo's local layers): the NEXT paged-attention launch of the calling
 * thread lets a query at position p see the keys (p - window, p] only; key tiles wholly outside the window are not
 * loaded. 0 = unlimited. Cleared by that launch. */
int hds_attention_next_window(int window);

/* tcgen05 paged attention for items with up to 128 query tokens (prefill / chunked continuation): QK^T and
 * PV on the tensor cores with TMEM accumulators, TMA page gathers, online softmax in registers.
 * n_q_rows = total rows of q (bounds the Q tensor map). n_splits > 1 splits each item's KV range over several
 * CTAs (short continuation over a long restored context); partials merge in `workspace` by ticket. The KV cache must hold finite values everywhere
 * (allocate it zeroed): masked probabilities are exact zeros, but 0 * NaN would still poison a row. */
int hds_paged_attention_prefill(const void* q, void* out, const void* kv_cache_layer, int64_t num_blocks,
                                const int32_t* block_tables, int max_blocks_per_seq, const hds_attn_item* items,
                                int n_items, int n_q_rows, int n_q_heads, int n_kv_heads, int head_dim, int block_size,
                                float scale, int n_splits, void* workspace, int64_t workspace_bytes, int dtype,
                                void* stream);

/* Grouped-query decode on the tensor cores: the same kernel with the 128 MMA rows holding the (token, query head) pairs
 * of ONE KV head's group (head_dim 128; group size G = n_q_heads / n_kv_heads divides 128; items of at most 128 / G
 * tokens). Every K/V page is read once per group and the G-fold arithmetic runs on tcgen05 instead of FFMA.
 * Workspace (n_splits > 1): hds_paged_attention_prefill_workspace_bytes(n_items, n_kv_heads, 128, n_splits).
 * Ref role: blocked_flash with GQA (ragged_ops/blocked_flash/blocked_flash.cpp:21-101). */
int hds_paged_attention_packed(const void* q, void* out, const void* kv_cache_layer, int64_t num_blocks,
                                const int32_t* block_tables, int max_blocks_per_seq, const hds_attn_item* items,
                                int n_items, int n_q_rows, int n_q_heads, int n_kv_heads, int head_dim, int block_size,
                                float scale, int n_splits, void* workspace, int64_t workspace_bytes, int dtype,
                                void* stream);
/* Scratch bytes for n_splits > 1 (0 when n_splits <= 1). Zero it once; the kernel leaves its tickets zeroed. */
int64_t hds_paged_attention_prefill_workspace_bytes(int n_items, int n_q_heads, int head_dim, int n_splits);

/* ---- mixture of experts (tensor-parallel experts; ref ragged_ops/{top_k_gating,moe_scatter,moe_gather}, moe_gemm) --- */
/* Per token: softmax over E logits, top-k (ties -> lower index), optional renormalisation of the k scores.
 * counts[E] must be zero on entry (hds_moe_scatter re-zeroes it). offsets[t, j] = row claimed inside expert_ids[t, j]. */
int hds_moe_topk_gating(const void* logits, int64_t ld_logits, int T, int E, int k, int normalize, int logits_dtype,
                        float* scores, int32_t* expert_ids, int32_t* offsets, int32_t* counts, void* stream);
/* pad_off[E+1]: expert row ranges rounded up to 128 rows; xs[row] = x[t] for every (t, j); mapped_slots[t, j] = row.
 * xs needs hds_moe_padded_rows(T, E, k) rows. */
int hds_moe_scatter(const void* x, int64_t ldx, void* xs, int64_t ld_xs, const int32_t* expert_ids,
                    const int32_t* offsets, int32_t* counts, int32_t* pad_off, int32_t* mapped_slots, int T, int E, int k,
                    int H, void* stream);
int64_t hds_moe_padded_rows(int T, int E, int k);
/* Grouped tcgen05 GEMM: rows [pad_off[e], pad_off[e+1]) of A times W_stacked[e*N:(e+1)*N, :]^T; epilogue as hds_gemm
 * (no bias). max_rows = row capacity of A and out (multiple of 128). */
int hds_moe_gemm(const void* A, int64_t lda, const void* W_stacked, int64_t ldw, const int32_t* pad_off, int n_experts,
                 int max_rows, int N, int K, int dtype, int epilogue, int act, void* out, int64_t ldo, void* stream);
/* out[t, :] += sum_j scores[t, j] * y[mapped_slots[t, j], :]   (out fp32: the pending residual delta) */
int hds_moe_gather(const void* y, int64_t ldy, const int32_t* mapped_slots, const float* scores, float* out, int64_t ldo,
                   int T, int k, int H, int dtype, void* stream);

/* ---- FP8 (E4M3) W8A8 GEMM on tcgen05 kind::f8f6f4 ----------------------------------------------------------------
 * hds_quantize_fp8_rows: q[r, :] = e4m3(x[r, :] / scales[r]), scales[r] = amax(x[r, :]) / 448 (per token for
 * activations, per output channel for weights). ldx in elements, ldq in bytes.
 * hds_gemm_fp8: out[M,N] = epilogue((Aq Wq^T) * a_scale[m] * w_scale[n] + bias); epilogue / act / bias as hds_gemm.
 * lda / ldw are row pitches in bytes (multiples of 16). bn: 0 = auto, else 64 / 128 / 256.
 * Ref role: inference/v2/kernels/core_ops/cuda_linear (weight-only FP6, sm_80 only) + quantization_mode of
 * inference/v2/config_v2.py:19-26. */
int hds_quantize_fp8_rows(const void* x, int64_t ldx, int rows, int K, int dtype, void* q, int64_t ldq, float* scales,
                          void* stream);
int hds_gemm_fp8(const void* Aq, int64_t lda, const float* a_scale, const void* Wq, int64_t ldw, const float* w_scale,
                 int M, int N, int K, int out_dtype, int epilogue, int act, const void* bias, void* out, int64_t ldo,
                 int bn, void* stream);

/* Decode-size (M <= 64) E4M3 GEMM on the feature-major weight-streaming kernel: epilogue 0 store / 1 gated (into out) or
 * 2 fp32 accumulate (accum[m, n] += ..., bias added once). Split-K for store / gated needs a zeroed scratch of
 * hds_gemm_fp8_skinny_workspace_bytes(M, N) bytes (left zeroed); splits 0 = auto. lda / ldw in bytes. */
int64_t hds_gemm_fp8_skinny_workspace_bytes(int M, int N);
int hds_gemm_fp8_skinny(const void* Aq, int64_t lda, const float* a_scale, const void* Wq, int64_t ldw, const float* w_scale,
                        int M, int N, int K, int out_dtype, int epilogue, int act, const void* bias, void* out, int64_t ldo,
                        float* accum, int64_t ld_accum, void* ws, int64_t ws_bytes, int splits, void* stream);

/* CTA-pair GEMM (tcgen05 cta_group::2, 256 x bn tiles over two SMs): out = epilogue(A W^T + bias), 16-bit operands.
 * Arguments as hds_gemm (no split-K); bn: 0 = auto, 128 or 256. hds_gemm_fp8 reaches the same kernel with bn = 0
 * (auto, M > 128) or bn = 1128 / 1256. */
int hds_gemm_2sm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int dtype, int epilogue,
                 int act, const void* bias, void* out, int64_t ldo, int bn, void* stream);

/* ---- group-wise quantizers (ref: csrc/quantization, csrc/fp_quantizer) ------------------------------------ */
/* dtype codes: 0 bf16, 1 fp16, 2 fp32. Integer: bits 4 or 8; params[g] = (scale, offset): value = code * scale + offset.
 * 4-bit codes are packed two per byte (low nibble first). */
int hds_quantize_int(const void* src, int dtype, int64_t n_groups, int group_size, int bits, int symmetric, void* dst,
                     float* params, void* stream);
int hds_dequantize_int(const void* src, const float* params, int64_t n_groups, int group_size, int bits, void* dst,
                       int dtype, void* stream);
/* Minifloat: q_bits 8 (e4m3) / 6 (e3m2) / 4 (e2m1) / 12 (e4m7), all-finite encodings, one fp32 scale per group
 * (scale = max|x| / format max). Packed at the natural bit width. group_idx (may be NULL) selects which stored group
 * each output group comes from (selective dequantisation). */
int hds_fp_quantize(const void* src, int dtype, int64_t n_groups, int group_size, int q_bits, void* dst, float* scales,
                    void* stream);
int hds_fp_dequantize(const void* src, const float* scales, const int32_t* group_idx, int64_t n_out_groups, int group_size,
                      int q_bits, void* dst, int dtype, void* stream);

/* ---- fused multi-tensor optimizers (training runtime) ----------------------------------------------- */
/* One Adam / AdamW (adamw_mode = 1: decoupled decay) or Lion step over n_tensors tensors. params / grads / exp_avg /
 * exp_avg_sq / numels are HOST arrays of device pointers and element counts. dtype codes: 0 bf16, 1 fp16, 2 fp32;
 * supported (param, grad) pairs: (2,2) (2,0) (2,1) (0,0) (1,1) (0,2) (1,2); moments are fp32. grads are multiplied by
 * grad_scale first (1 / loss scale, or a clipping coefficient). step counts from 1.
 * Ref: csrc/adam/multi_tensor_adam.cu, csrc/lion/multi_tensor_lion.cu. */
int hds_multi_tensor_adam(int n_tensors, void* const* params, const void* const* grads, float* const* exp_avg,
                          float* const* exp_avg_sq, const int64_t* numels, int param_dtype, int grad_dtype, float lr,
                          float beta1, float beta2, float eps, float weight_decay, int step, int adamw_mode,
                          int bias_correction, float grad_scale, void* stream);
int hds_multi_tensor_lion(int n_tensors, void* const* params, const void* const* grads, float* const* exp_avg,
                          const int64_t* numels, int param_dtype, int grad_dtype, float lr, float beta1, float beta2,
                          float weight_decay, float grad_scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HDS_B200_H */
