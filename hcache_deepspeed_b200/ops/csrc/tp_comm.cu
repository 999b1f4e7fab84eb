// Tensor-parallel plumbing of the kernel library: communicator contexts (the flag pads of hds_ptx.cuh's protocol),
// the "arm" mechanism that attaches a wait and / or a signal to the next GEMM-family launch, and the three stand-alone
// kernels of the protocol: signal, barrier and the column gather that assembles vocabulary-sharded logits.
//
// What this replaces in the reference: the NCCL calls of the tensor-parallel forward —
// `dist.all_reduce` after attention-out and MLP-down (deepspeed/inference/v2/model_implementations/llama_v2/model.py:159-160,
// 168-169) and `dist.all_gather_into_tensor` + permute copy of the logits (:195-197). Here the producing GEMM itself
// releases its result to the peers (last CTA: system fence + flag stores over NVLink), and the consumer (the norm
// kernel / the gather below / the next GEMM's TMA producer) acquires the flags and reads or reduces peer memory directly.
#include "hds_ptx.cuh"
#include "hds_host.cuh"
#include "tp_sync_host.cuh"
#include "../../../include/hds_b200.h"

namespace hds {

struct TpContext {
    bool used;
    TpSync base;  // pads + rank geometry; wait / signal cleared
};
constexpr int kTpMaxContexts = 64;
static TpContext g_tp_ctx[kTpMaxContexts];
static std::mutex g_tp_mu;
static thread_local TpSync g_tp_armed;
static thread_local bool g_tp_armed_valid = false;

TpSync tp_take_armed() {
    TpSync s;
    memset(&s, 0, sizeof(s));
    if (g_tp_armed_valid) {
        s = g_tp_armed;
        g_tp_armed_valid = false;
    }
    return s;
}

bool tp_get_context(int ctx, TpSync* out) {
    if (ctx < 0 || ctx >= kTpMaxContexts || !g_tp_ctx[ctx].used) {
        set_error("tensor-parallel context id is not valid");
        return false;
    }
    *out = g_tp_ctx[ctx].base;
    return true;
}

// one CTA: (optionally) wait for the stream's previous kernel, then publish "this rank got here"
__global__ void tp_signal_kernel(const TpSync s) {
    pdl_enter();
    tp_signal_tail(s, 1);
}

// one CTA: wait until every rank has issued as many signals as this one (orders later kernels of the stream behind it)
__global__ void tp_wait_kernel(const TpSync s) {
    pdl_enter();
    tp_wait_cta(s);
}

// out[row, col0[p] : col0[p] + width[p]] = rank p's local slice [row, 0:width[p]] for every rank p (own slice included):
// the all-gather of vocabulary-sharded logits, pulled with 16-byte loads over NVLink. grid (rows, ranks).
struct GatherCols {
    const char* src[kTpMaxRanks];   // rank p's slice buffer through this rank's mapping
    int col0[kTpMaxRanks];          // first output column (elements)
    int width[kTpMaxRanks];         // columns (elements)
    int64_t ld_src, ld_out;         // elements
};
__global__ void tp_gather_cols_kernel(char* __restrict__ out, const GatherCols g, const TpSync s, int elem_bytes) {
    pdl_enter();
    if (s.wait) tp_wait_cta(s);
    const int row = blockIdx.x, p = blockIdx.y;
    const char* src = g.src[p] + static_cast<int64_t>(row) * g.ld_src * elem_bytes;
    char* dst = out + (static_cast<int64_t>(row) * g.ld_out + g.col0[p]) * elem_bytes;
    const int bytes = g.width[p] * elem_bytes;
    const bool vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | static_cast<uintptr_t>(bytes)) & 15) == 0;
    if (vec) {
        for (int i = threadIdx.x * 16; i < bytes; i += blockDim.x * 16)
            *reinterpret_cast<uint4*>(dst + i) = ld_relaxed_sys_v4(src + i);
    } else {
        for (int i = threadIdx.x * 2; i < bytes; i += blockDim.x * 2) {
            uint16_t v;
            asm volatile("ld.relaxed.sys.global.u16 %0, [%1];" : "=h"(v) : "l"(src + i) : "memory");
            *reinterpret_cast<uint16_t*>(dst + i) = v;
        }
    }
}

}  // namespace hds

using namespace hds;

extern "C" int hds_tp_ctx_create(int rank, int n_ranks, void* const* pad_ptrs) {
    HDS_REQUIRE(n_ranks >= 1 && n_ranks <= kTpMaxRanks && rank >= 0 && rank < n_ranks, "hds_tp_ctx_create: 1..8 ranks");
    std::lock_guard<std::mutex> lock(g_tp_mu);
    for (int i = 0; i < kTpMaxContexts; ++i) {
        if (g_tp_ctx[i].used) continue;
        TpSync& b = g_tp_ctx[i].base;
        memset(&b, 0, sizeof(b));
        for (int r = 0; r < n_ranks; ++r) {
            HDS_REQUIRE(pad_ptrs[r] != nullptr && (reinterpret_cast<uintptr_t>(pad_ptrs[r]) & 127) == 0,
                        "hds_tp_ctx_create: every rank's pad must be mapped and 128-byte aligned");
            b.pad[r] = reinterpret_cast<uint32_t*>(pad_ptrs[r]);
        }
        b.n_ranks = n_ranks;
        b.rank = rank;
        g_tp_ctx[i].used = true;
        return i;
    }
    set_error("hds_tp_ctx_create: all 64 context slots are taken (hds_tp_ctx_destroy releases one)");
    return HDS_ERR_TOO_LARGE;
}

extern "C" int hds_tp_ctx_destroy(int ctx) {
    std::lock_guard<std::mutex> lock(g_tp_mu);
    if (ctx >= 0 && ctx < kTpMaxContexts) g_tp_ctx[ctx].used = false;
    return HDS_OK;
}

extern "C" int hds_tp_arm(int ctx, int wait, int signal) {
    TpSync s;
    if (!tp_get_context(ctx, &s)) return HDS_ERR_BAD_ARG;
    s.wait = wait ? 1 : 0;
    s.signal = signal ? 1 : 0;
    g_tp_armed = s;
    g_tp_armed_valid = (wait || signal) && s.n_ranks > 1;
    return HDS_OK;
}

extern "C" int hds_tp_disarm(void) {
    const int was = g_tp_armed_valid ? 1 : 0;
    g_tp_armed_valid = false;
    return was;
}

extern "C" int hds_tp_signal(int ctx, void* stream) {
    TpSync s;
    if (!tp_get_context(ctx, &s)) return HDS_ERR_BAD_ARG;
    if (s.n_ranks == 1) return HDS_OK;
    s.signal = 1;
    HDS_CHECK_CUDA(launch_kernel(tp_signal_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), s));
    return HDS_OK;
}

extern "C" int hds_tp_wait(int ctx, void* stream) {
    TpSync s;
    if (!tp_get_context(ctx, &s)) return HDS_ERR_BAD_ARG;
    if (s.n_ranks == 1) return HDS_OK;
    s.wait = 1;
    HDS_CHECK_CUDA(launch_kernel(tp_wait_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), s));
    return HDS_OK;
}

// full barrier of the ranks' streams: signal, then wait (two one-CTA kernels)
extern "C" int hds_tp_barrier(int ctx, void* stream) {
    const int rc = hds_tp_signal(ctx, stream);
    return rc ? rc : hds_tp_wait(ctx, stream);
}

extern "C" int hds_tp_gather_cols(int ctx, int wait, void* out, int64_t ld_out, const void* const* src_ptrs, int64_t ld_src,
                                  const int32_t* col0, const int32_t* width, int rows, int elem_bytes, void* stream) {
    TpSync s;
    if (!tp_get_context(ctx, &s)) return HDS_ERR_BAD_ARG;
    HDS_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "hds_tp_gather_cols: 2- or 4-byte elements");
    if (rows == 0) return HDS_OK;
    s.wait = wait ? 1 : 0;
    GatherCols g;
    memset(&g, 0, sizeof(g));
    for (int p = 0; p < s.n_ranks; ++p) {
        g.src[p] = reinterpret_cast<const char*>(src_ptrs[p]);
        g.col0[p] = col0[p];
        g.width[p] = width[p];
    }
    g.ld_src = ld_src;
    g.ld_out = ld_out;
    HDS_CHECK_CUDA(launch_kernel(tp_gather_cols_kernel, dim3(rows, s.n_ranks), dim3(256), 0, static_cast<cudaStream_t>(stream),
                                 reinterpret_cast<char*>(out), g, s, elem_bytes));
    return HDS_OK;
}
