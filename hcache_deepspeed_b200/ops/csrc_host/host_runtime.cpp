// Host runtime: KV block allocator and ragged batch builder (see include/hds_host.h).
#include "hds_host.h"

#include <algorithm>
#include <cstring>
#include <mutex>
#include <vector>

namespace {

// Free list kept as a stack of ids plus a "is free" bitmap for double-free detection.
struct BlockAllocator {
    explicit BlockAllocator(int32_t n) : total(n), is_free(static_cast<size_t>(n), 1) {
        stack.reserve(n);
        for (int32_t i = n - 1; i >= 0; --i) stack.push_back(i);  // pop order 0, 1, 2, ...
    }
    int32_t total;
    std::vector<int32_t> stack;
    std::vector<uint8_t> is_free;
    std::mutex mu;
};

}  // namespace

extern "C" {

void* hds_alloc_create(int32_t num_blocks) {
    if (num_blocks < 0) return nullptr;
    return new BlockAllocator(num_blocks);
}

void hds_alloc_destroy(void* h) { delete static_cast<BlockAllocator*>(h); }

int hds_alloc_allocate(void* h, int32_t n, int32_t* out) {
    auto* a = static_cast<BlockAllocator*>(h);
    std::lock_guard<std::mutex> lock(a->mu);
    if (n < 0 || static_cast<size_t>(n) > a->stack.size()) return -1;
    for (int32_t i = 0; i < n; ++i) {
        const int32_t id = a->stack.back();
        a->stack.pop_back();
        a->is_free[id] = 0;
        out[i] = id;
    }
    return 0;
}

int hds_alloc_free(void* h, const int32_t* ids, int32_t n) {
    auto* a = static_cast<BlockAllocator*>(h);
    std::lock_guard<std::mutex> lock(a->mu);
    // validate first so a bad call frees nothing
    for (int32_t i = 0; i < n; ++i) {
        const int32_t id = ids[i];
        if (id < 0 || id >= a->total || a->is_free[id]) return -1;
        for (int32_t j = 0; j < i; ++j)
            if (ids[j] == id) return -1;
    }
    for (int32_t i = 0; i < n; ++i) {
        a->is_free[ids[i]] = 1;
        a->stack.push_back(ids[i]);
    }
    return 0;
}

int32_t hds_alloc_free_blocks(void* h) {
    auto* a = static_cast<BlockAllocator*>(h);
    std::lock_guard<std::mutex> lock(a->mu);
    return static_cast<int32_t>(a->stack.size());
}

int32_t hds_alloc_total_blocks(void* h) { return static_cast<BlockAllocator*>(h)->total; }

int hds_build_batch(const hds_seq_in* seqs, int32_t n_seqs, int32_t block_size, int32_t group, int32_t rows_small_max,
                    int32_t large_chunk, int32_t max_tokens, int32_t max_items, int32_t table_stride,
                    int32_t* tok_pos, int32_t* tok_slot, int32_t* tok_seq, int32_t* last_rows, int32_t* block_tables,
                    int32_t* items_small, int32_t* items_large, int32_t* counts) {
    int32_t row = 0, n_small = 0, n_large = 0, max_rows_small = 0, max_kv_small = 0, max_kv_large = 0;
    const int32_t small_tokens = group > 0 ? rows_small_max / group : 0;  // tokens a split-KV item may hold
    for (int32_t s = 0; s < n_seqs; ++s) {
        const hds_seq_in& sq = seqs[s];
        const int32_t kv_len = sq.seen + sq.n_new;
        const int32_t need_blocks = (kv_len + block_size - 1) / block_size;
        if (need_blocks > sq.n_blocks) return -2;
        if (need_blocks > table_stride) return -1;
        if (row + sq.n_new > max_tokens) return -1;
        int32_t* table = block_tables + static_cast<int64_t>(s) * table_stride;
        std::memcpy(table, sq.block_ids, sizeof(int32_t) * static_cast<size_t>(need_blocks));
        if (need_blocks < table_stride)
            std::memset(table + need_blocks, 0, sizeof(int32_t) * static_cast<size_t>(table_stride - need_blocks));
        for (int32_t i = 0; i < sq.n_new; ++i) {
            const int32_t pos = sq.seen + i;
            tok_pos[row + i] = pos;
            tok_slot[row + i] = sq.block_ids[pos / block_size] * block_size + pos % block_size;
            tok_seq[row + i] = s;
        }
        last_rows[s] = row + sq.n_new - 1;
        if (sq.n_new > 0) {
            if (sq.n_new <= small_tokens) {
                if (n_small >= max_items) return -1;
                int32_t* it = items_small + 4 * n_small++;
                it[0] = s; it[1] = row; it[2] = sq.n_new; it[3] = sq.seen;
                max_rows_small = std::max(max_rows_small, sq.n_new * group);
                max_kv_small = std::max(max_kv_small, kv_len);
            } else {
                for (int32_t off = 0; off < sq.n_new; off += large_chunk) {
                    if (n_large >= max_items) return -1;
                    const int32_t n = std::min(large_chunk, sq.n_new - off);
                    int32_t* it = items_large + 4 * n_large++;
                    it[0] = s; it[1] = row + off; it[2] = n; it[3] = sq.seen + off;
                    max_kv_large = std::max(max_kv_large, sq.seen + off + n);
                }
            }
        }
        row += sq.n_new;
    }
    counts[0] = row;
    counts[1] = n_small;
    counts[2] = n_large;
    counts[3] = max_rows_small;
    counts[4] = max_kv_small;
    counts[5] = max_kv_large;
    return 0;
}

}  // extern "C"
