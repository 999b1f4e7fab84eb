/* hds_host.h — C ABI of libhds_host.so, the host-side native runtime of hcache_deepspeed_b200.
 * No CUDA dependency: it builds and is tested on CPU-only machines.
 *
 *   - KV block allocator            (ref: deepspeed/inference/v2/ragged/blocked_allocator.py:11-105 — a Python
 *                                    linked list with one .item() per block)
 *   - ragged batch builder          (ref: ragged/ragged_wrapper.py:134-218 insert_sequence/finalize python lists,
 *                                    kernels/ragged_ops/atom_builder/atom_builder.cpp:10-53)
 *   - asynchronous file I/O handle  (ref: csrc/aio/py_lib/deepspeed_py_io_handle.cpp, py_ds_aio.cpp:15-116 —
 *                                    libaio; here kernel AIO through raw io_setup / io_submit / io_getevents system
 *                                    calls, queue_depth iocbs in flight per worker, O_DIRECT; pread / pwrite for what
 *                                    cannot be submitted that way)
 */
#ifndef HDS_HOST_H
#define HDS_HOST_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- KV block allocator: O(1) allocate/free of block ids, double-free detection ---------------- */
void* hds_alloc_create(int32_t num_blocks);
void hds_alloc_destroy(void* h);
/* Writes n block ids to out. Returns 0, or -1 (nothing allocated) when fewer than n blocks are free. */
int hds_alloc_allocate(void* h, int32_t n, int32_t* out);
/* Returns 0, or -1 if any id is out of range or already free (nothing is freed in that case). */
int hds_alloc_free(void* h, const int32_t* ids, int32_t n);
int32_t hds_alloc_free_blocks(void* h);
int32_t hds_alloc_total_blocks(void* h);

/* ---- ragged batch builder ------------------------------------------------------------------------ */
/* One sequence of the batch as the builder sees it. */
typedef struct {
    int32_t n_new;            /* tokens in flight this forward */
    int32_t seen;             /* tokens already in the KV cache */
    int32_t n_blocks;         /* valid entries of block_ids */
    int32_t _pad;
    const int32_t* block_ids; /* this sequence's KV block ids, in order */
} hds_seq_in;

/* Output arrays (all int32, caller-allocated, typically views of ONE pinned staging buffer):
 *   tok_pos[T], tok_slot[T], tok_seq[T]     per token: absolute position, cache slot, batch-local sequence
 *   last_rows[n_seqs]                        row of each sequence's last in-flight token
 *   block_tables[n_seqs * table_stride]      block ids, zero padded
 *   items_small / items_large [.. * 4]       attention work items (seq, q_row0, n_q, pos0), see hds_attn_item
 * rows_small_max: max query rows (tokens * q-heads-per-kv-head) a split-KV item may carry (<= 8);
 * group: q heads per kv head; large_chunk: max tokens per tensor-core item (128).
 * counts[0] = total tokens, [1] = n small items, [2] = n large items, [3] = max rows per small item,
 * [4] = max kv length over small items, [5] = max kv length over large items.
 * Returns 0, -1 on capacity overflow (max_tokens / max_items / table_stride), -2 if a sequence lacks blocks. */
int hds_build_batch(const hds_seq_in* seqs, int32_t n_seqs, int32_t block_size, int32_t group, int32_t rows_small_max,
                    int32_t large_chunk, int32_t max_tokens, int32_t max_items, int32_t table_stride,
                    int32_t* tok_pos, int32_t* tok_slot, int32_t* tok_seq, int32_t* last_rows, int32_t* block_tables,
                    int32_t* items_small, int32_t* items_large, int32_t* counts);

/* ---- asynchronous file I/O ----------------------------------------------------------------------- */
/* Thread-pool handle: requests are split into block_size pieces spread over `threads` workers, each piece
 * a pread/pwrite at its own offset (O_DIRECT when use_direct and buffer/offset/length are 4096-aligned).
 * Ref semantics: aio_handle(block_size, queue_depth, single_submit, overlap_events, intra_op_parallelism). */
void* hds_aio_create(int64_t block_size, int32_t queue_depth, int32_t threads, int32_t use_direct);
void hds_aio_destroy(void* h);
/* Asynchronous: returns a request id > 0 (or -1 when the file cannot be opened). */
int64_t hds_aio_pread(void* h, void* buffer, const char* path, int64_t nbytes, int64_t file_offset);
int64_t hds_aio_pwrite(void* h, const void* buffer, const char* path, int64_t nbytes, int64_t file_offset);
/* Block until the request (or every outstanding request if id == 0) has finished.
 * Returns bytes transferred by that request (sum for id == 0) or -errno of the first failure. */
int64_t hds_aio_wait(void* h, int64_t request_id);
/* 1 if finished, 0 if still running, negative on unknown id. Never blocks. */
int hds_aio_poll(void* h, int64_t request_id);
int32_t hds_aio_pending(void* h);
/* out[0] = pieces completed through kernel AIO (io_submit), out[1] = pieces executed synchronously (pread / pwrite),
 * out[2] = deepest in-flight queue one worker reached. */
void hds_aio_stats(void* h, int64_t* out);

/* mlock / munlock helpers for page-locked staging buffers that are not CUDA-registered. */
int hds_mem_lock(void* ptr, int64_t nbytes);
int hds_mem_unlock(void* ptr, int64_t nbytes);

/* ---- host optimizers (ZeRO-Offload) ---------------------------------------------------------------- */
/* One step over n fp32 elements in host memory. copy_out (may be NULL) receives the updated parameters converted
 * to copy_dtype (0 bf16, 1 fp16, 2 fp32) in the same pass. threads <= 0: hardware concurrency. step counts from 1.
 * Ref: csrc/adam/cpu_adam_impl.cpp, csrc/adagrad/cpu_adagrad.cpp, csrc/lion/cpu_lion_impl.cpp. Returns 0 / -1. */
int hds_cpu_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                      float beta1, float beta2, float eps, float weight_decay, int step, int adamw_mode,
                      int bias_correction, void* copy_out, int copy_dtype, int threads);
int hds_cpu_adagrad_step(float* params, const float* grads, float* exp_avg_sq, int64_t n, float lr, float eps,
                         float weight_decay, void* copy_out, int copy_dtype, int threads);
int hds_cpu_lion_step(float* params, const float* grads, float* exp_avg, int64_t n, float lr, float beta1, float beta2,
                      float weight_decay, void* copy_out, int copy_dtype, int threads);

/* ---- host collectives over one shared-memory segment, for the ranks of one node (ref: csrc/cpu/comm/shm.cpp, ccl.cpp) --
 * hds_shm_open: every rank of the group maps POSIX segment "/<name>" (two banks of `world` slots of slot_bytes) and
 * waits until all ranks have attached; NULL on failure. Every rank must then issue the same calls in the same order.
 * dtype: 0 bf16, 1 fp16 (both accumulated in fp32), 2 fp32, 3 int32, 4 int64. op: 0 sum, 1 max, 2 min. Any size: messages
 * are cut into slot-sized rounds. Return 0, -1 (bad argument) or -2 (a peer did not arrive within HDS_SHM_TIMEOUT_S,
 * default 600). HDS_SHM_SIMD=off disables the AVX-512 kernels. hds_shm_close unmaps (one rank passes unlink_segment). */
void* hds_shm_open(const char* name, int rank, int world, int64_t slot_bytes);
int hds_shm_allreduce(void* handle, void* data, int64_t numel, int dtype);                    /* in place, sum */
int hds_shm_allreduce_op(void* handle, void* data, int64_t numel, int dtype, int op);         /* in place */
int hds_shm_broadcast(void* handle, void* data, int64_t nbytes, int root);
int hds_shm_allgather(void* handle, const void* in, void* out, int64_t nbytes_each);          /* out = [world][nbytes_each] */
int hds_shm_reduce_scatter(void* handle, const void* in, void* out, int64_t numel_out, int dtype, int op); /* in = [world][numel_out] */
int hds_shm_barrier(void* handle);
void hds_shm_info(void* handle, int64_t* out4);  /* avx512 in use, rounds, one-exchange chunks, shard-per-rank chunks */
void hds_shm_close(void* handle, int unlink_segment);

#ifdef __cplusplus
}
#endif
#endif /* HDS_HOST_H */
