// sm_100a PTX building blocks shared by every kernel in this library:
// mbarrier, TMA (cp.async.bulk[.tensor]), tcgen05 (TMEM alloc / mma / commit / ld),
// UMMA shared-memory + instruction descriptors, vector global loads/stores, reductions.
//
// Everything here is hand-written inline PTX for Blackwell (B200, sm_100a). There is no
// counterpart in the reference: its device helpers (deepspeed/inference/v2/kernels/includes/
// memory_access_utils.h, reduction_utils.h, conversion_utils.h) stop at cp.async + warp shuffles.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace hds {

// --------------------------------------------------------------------------------------------
// misc
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// Programmatic dependent launch (PDL). Kernels launched through hds::launch_kernel may start while the previous
// kernel of the stream is still draining: `griddep_launch_dependents` lets the NEXT kernel's CTAs be scheduled
// once every CTA of this grid has issued it; `griddep_wait` blocks until all prerequisite grids have completed
// and their writes are visible. Everything before the wait may only touch memory no kernel writes (weights).
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() {
    griddep_launch_dependents();
    griddep_wait();
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t warp_id() { return __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0); }

__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.b32 %0, 1, 0, P;\n\t}\n"
        : "=r"(pred));
    return pred != 0;
}

// --------------------------------------------------------------------------------------------
// mbarrier
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// generic-proxy writes to smem -> visible to the async proxy (TMA / UMMA reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug traps (-> CUDA error on the host) instead of hanging the GPU.
// The bound uses the SM-local cycle counter, read only once per 4096 failed probes (reading
// %globaltimer per probe costs about a microsecond and showed up as per-tile latency).
// Legitimate waits here are microseconds to milliseconds; ~2^33 cycles (> 4 s) means deadlock.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    uint32_t probes = 0;
    long long t0 = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++probes & 4095u) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > (1ll << 33)) __trap();
        }
    }
}

// --------------------------------------------------------------------------------------------
// TMA
// --------------------------------------------------------------------------------------------
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
// 2D tiled load: coordinates are (c0 = innermost element index, c1 = row index).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tmap, uint64_t* bar, int32_t c0,
                                            int32_t c1, uint64_t hint) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
        "l"(hint)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* tmap, uint64_t* bar, int32_t c0,
                                            int32_t c1, int32_t c2, uint64_t hint) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5}], [%2], %6;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
        "r"(c2), "l"(hint)
        : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tmap, const void* smem_src, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                     reinterpret_cast<uint64_t>(tmap)),
                 "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
// 1D bulk copy global -> shared (bytes multiple of 16, both addresses 16B aligned)
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar,
                                             uint64_t hint) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar)), "l"(hint)
        : "memory");
}

// --------------------------------------------------------------------------------------------
// tcgen05 / TMEM
// --------------------------------------------------------------------------------------------
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {
    static_assert(kCols >= 32 && kCols <= 512 && (kCols & (kCols - 1)) == 0, "TMEM columns: power of two in [32,512]");
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
                 "n"(kCols)
                 : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], bf16/fp16 inputs, fp32 accumulate; issued by ONE thread.
__device__ __forceinline__ void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// mbarrier arrive once all previously issued tcgen05.mma of this thread have completed
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns (thread i gets lane base+i).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// registers -> TMEM (this warp's 32 lanes x 32 columns)
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]),
        "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]),
        "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- UMMA descriptors ------------------------------------------------------------------------
// Shared-memory matrix descriptor (64-bit):
//   [0,14)  start address >> 4      [16,30) leading-dim byte offset >> 4
//   [32,46) stride-dim byte offset >> 4   [46,48) version (1 on sm_100)
//   [49,52) base offset (0: tiles are 1024B aligned)   [61,64) layout: 0 none, 2 = SWIZZLE_128B
//
// K-major, SWIZZLE_128B, bf16: a tile is rows of 64 elements (=128 B); 8 rows form one 1024 B
// swizzle atom. LBO is unused for swizzled K-major (set to 1); SBO = 1024 B between 8-row groups.
__device__ __forceinline__ uint64_t make_desc_kmajor_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>(1) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}
// MN-major, SWIZZLE_128B, bf16: the operand's MN extent is contiguous in memory. The tile is
// stored as [k rows][64 mn-elements (=128 B)] panels; 8 consecutive k rows form a 1024 B atom
// (SBO = 1024 B along K), and successive 64-element MN panels are `lbo_bytes` apart.
__device__ __forceinline__ uint64_t make_desc_mnmajor_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}
// Instruction descriptor for kind::f16 (32-bit): D=f32, A=B=bf16 (or fp16), dense.
//   [4,6) D fmt (1 = f32)  [7,10) A fmt  [10,13) B fmt (0 = f16, 1 = bf16)
//   [15] A major  [16] B major (0 = K-major, 1 = MN-major)  [17,23) N>>3  [24,29) M>>4
__host__ __device__ constexpr uint32_t make_idesc_f16(uint32_t m, uint32_t n, bool bf16, bool a_mn_major,
                                                      bool b_mn_major) {
    return (1u << 4) | ((bf16 ? 1u : 0u) << 7) | ((bf16 ? 1u : 0u) << 10) | ((a_mn_major ? 1u : 0u) << 15) |
           ((b_mn_major ? 1u : 0u) << 16) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

// Instruction descriptor for kind::f8f6f4 with both operands E4M3 (format code 0), D = f32, K-major operands.
__host__ __device__ constexpr uint32_t make_idesc_f8(uint32_t m, uint32_t n) {
    return (1u << 4) | ((n >> 3) << 17) | ((m >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem], 8-bit float operands, M x N x 32 per instruction
__device__ __forceinline__ void umma_f8_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

// --------------------------------------------------------------------------------------------
// global memory helpers
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_v4(void* p, const uint4& v) {
    asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
// vector fp32 reduction into global memory (sm_90+): 16 B per instruction, no return value.
__device__ __forceinline__ void red_add_v4_f32(float* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
                 "f"(d)
                 : "memory");
}
// same, but system scope: target may be a peer GPU's memory mapped over NVLink.
__device__ __forceinline__ void red_add_v4_f32_sys(float* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
                 "f"(d)
                 : "memory");
}
// NVLS: one instruction, reduced into the same offset of every GPU bound to the multicast object.
__device__ __forceinline__ void multimem_red_add_v4_f32(float* mc_ptr, float a, float b, float c, float d) {
    asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_ptr), "f"(a), "f"(b),
                 "f"(c), "f"(d)
                 : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add_u32(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// 16-byte load that is never served from this SM's L1: the line may live in a peer GPU's memory (NVLink mapping) and
// have been rewritten since an earlier kernel of this process touched it.
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
    uint4 r;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
// NVLS loads / stores on a multicast address: the switch sums the same offset of every GPU's buffer (ld_reduce) or
// writes it to every GPU's buffer (st) — one instruction, one pass over this GPU's links.
__device__ __forceinline__ float4 multimem_ld_reduce_v4_f32(const void* mc_ptr) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(mc_ptr)
                 : "memory");
    return r;
}
// 8 sixteen-bit values per instruction, summed in fp32 inside the switch, returned rounded to the 16-bit type
template <bool kBf16>
__device__ __forceinline__ uint4 multimem_ld_reduce_v4_h2(const void* mc_ptr) {
    uint4 r;
    if constexpr (kBf16)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                     : "l"(mc_ptr)
                     : "memory");
    else
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                     : "l"(mc_ptr)
                     : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st_v4(void* mc_ptr, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_ptr), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}
// generic-proxy writes to global memory (possibly made by another GPU) -> visible to this thread's later TMA loads
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

// --------------------------------------------------------------------------------------------
// Tensor-parallel flag protocol (this library's own; no NCCL / torch kernel on the path).
//
// Every rank owns one 256-byte "pad" in symmetric memory:
//     pad[r], r < 8 : arrive count of rank r — written ONLY by rank r (st.release.sys over NVLink), read only locally
//     pad[kTpTicket]: CTA ticket of the kernel currently signalling (local use)
//     pad[kTpSent]  : number of signals this rank has issued so far (local use)
// All ranks run the same kernel sequence, so "signal" number n means the same program point on every rank. A SIGNAL
// is issued by the last CTA of a kernel to retire (ticket), after a system-scope fence: everything the kernel wrote —
// locally or into peers — is visible to whoever acquires the flag. A WAIT spins (locally, on this rank's own pad) until
// every peer's arrive count has reached this rank's own `sent` count, i.e. until all ranks have passed the program
// point of this rank's latest signal. Counters only grow (wrap-safe compare), nothing is ever reset, so there is no
// second "everyone has left" phase, and a rank can run at most one signal ahead of the slowest one.
// The waits are bounded: a protocol bug traps (CUDA error on the host) instead of hanging the GPU.
// --------------------------------------------------------------------------------------------
constexpr int kTpMaxRanks = 8;
constexpr int kTpTicket = 32, kTpSent = 33;
struct TpSync {
    uint32_t* pad[kTpMaxRanks];  // pad[r]: rank r's pad through THIS rank's mapping; pad[rank] is local memory
    int n_ranks, rank;
    int wait;    // the kernel waits (before reading what peers produced) for every rank's latest signal
    int signal;  // the kernel signals when its last CTA retires
};

__device__ __forceinline__ void tp_poll(const uint32_t* flag, uint32_t target) {
    uint32_t probes = 0;
    long long t0 = 0;
    while (static_cast<int32_t>(ld_acquire_sys_u32(flag) - target) < 0) {
        if ((++probes & 1023u) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > (1ll << 35)) __trap();  // ~17 s: a peer died or the ranks ran different programs
        }
        if (probes > 64) __nanosleep(40);
    }
}
// whole-CTA wait: one thread per peer spins, everybody leaves together
__device__ __forceinline__ void tp_wait_cta(const TpSync& s) {
    if (static_cast<int>(threadIdx.x) < s.n_ranks && static_cast<int>(threadIdx.x) != s.rank) {
        const uint32_t target = ld_acquire_gpu_u32(s.pad[s.rank] + kTpSent);
        tp_poll(s.pad[s.rank] + threadIdx.x, target);
    }
    __syncthreads();
}
// single-thread wait (TMA producer lanes); the caller orders the async proxy behind it
__device__ __forceinline__ void tp_wait_thread(const TpSync& s) {
    const uint32_t target = ld_acquire_gpu_u32(s.pad[s.rank] + kTpSent);
    for (int r = 0; r < s.n_ranks; ++r)
        if (r != s.rank) tp_poll(s.pad[s.rank] + r, target);
}
__device__ __forceinline__ void st_relaxed_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Every thread of every CTA calls this once, after the CTA's last global write (it contains a __syncthreads; blockDim
// is a multiple of 32). Lane 0 of warp 0 fences on behalf of the CTA (the barrier puts the other threads' writes before
// the fence; fences are cumulative) — at GPU scope when the kernel wrote only this GPU's memory, at system scope when it
// stored into peers (kRemoteWrites) — and takes the CTA's ticket. In the last CTA of the grid, lane r of warp 0 then
// publishes the new signal count to peer r: its own system-scope fence followed by a relaxed store, all peers in
// parallel. (Measured: one thread issuing 7 release stores in a row pays 7 NVLink round trips, ~20 us per signal at
// TP=8; a system-scope fence per CTA instead of per grid costs 0.4 ms per 7B decode iteration.)
template <bool kRemoteWrites = false>
__device__ __forceinline__ void tp_signal_tail(const TpSync& s, uint32_t n_ctas) {
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t* mine = s.pad[s.rank];
        uint32_t v = 0, last = 0;
        if (threadIdx.x == 0) {
            if constexpr (kRemoteWrites) __threadfence_system();
            else __threadfence();
            if (atomicAdd(mine + kTpTicket, 1u) == n_ctas - 1) {
                last = 1;
                __threadfence();   // acquire side of the ticket chain: every CTA's writes happen before what follows
                v = ld_acquire_gpu_u32(mine + kTpSent) + 1;
                *reinterpret_cast<volatile uint32_t*>(mine + kTpTicket) = 0;
                *reinterpret_cast<volatile uint32_t*>(mine + kTpSent) = v;
            }
        }
        v = __shfl_sync(0xffffffffu, v, 0);
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last && static_cast<int>(threadIdx.x) < s.n_ranks && static_cast<int>(threadIdx.x) != s.rank) {
            __threadfence_system();
            st_relaxed_sys_u32(s.pad[threadIdx.x] + s.rank, v);
        }
    }
}

// --------------------------------------------------------------------------------------------
// numeric helpers
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ float bf16_lo(uint32_t packed) { return __uint_as_float(packed << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t packed) { return __uint_as_float(packed & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float fp16_lo(uint32_t packed) {
    return __half2float(__ushort_as_half(static_cast<unsigned short>(packed & 0xffffu)));
}
__device__ __forceinline__ float fp16_hi(uint32_t packed) {
    return __half2float(__ushort_as_half(static_cast<unsigned short>(packed >> 16)));
}
__device__ __forceinline__ uint32_t pack_fp16(float lo, float hi) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
// dtype-generic pack/unpack: kBf16 selects bf16, otherwise fp16
template <bool kBf16>
__device__ __forceinline__ float unpack_lo(uint32_t p) {
    return kBf16 ? bf16_lo(p) : fp16_lo(p);
}
template <bool kBf16>
__device__ __forceinline__ float unpack_hi(uint32_t p) {
    return kBf16 ? bf16_hi(p) : fp16_hi(p);
}
template <bool kBf16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    return kBf16 ? pack_bf16(lo, hi) : pack_fp16(lo, hi);
}


template <bool kBf16>
__device__ __forceinline__ float load_scalar(const void* base, int64_t idx) {
    if (kBf16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
    return __half2float(reinterpret_cast<const __half*>(base)[idx]);
}
template <bool kBf16>
__device__ __forceinline__ void store_scalar(void* base, int64_t idx, float v) {
    if (kBf16)
        reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(v);
    else
        reinterpret_cast<__half*>(base)[idx] = __float2half_rn(v);
}
// DT: 0 = bf16, 1 = fp16, 2 = fp32
template <int DT>
__device__ __forceinline__ float load_any(const void* base, int64_t idx) {
    if (DT == 0) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
    if (DT == 1) return __half2float(reinterpret_cast<const __half*>(base)[idx]);
    return reinterpret_cast<const float*>(base)[idx];
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

enum Activation : int { ACT_IDENTITY = 0, ACT_GELU = 1, ACT_RELU = 2, ACT_SILU = 3, ACT_GELU_TANH = 4 };

// sin / cos of a RoPE angle (position x inverse frequency: up to ~1e5 rad). Two-constant Cody-Waite reduction to
// [-pi, pi] followed by the SFU approximations: absolute error ~1e-6, far below the 6e-8 * |x| the fp32 product
// pos * inv_freq already carries. The library sincosf takes a branchy slow path for large arguments (stack frame,
// divergence) and made the RoPE kernels compute-bound: rope_kv_append ran 13x over its memory time.
__device__ __forceinline__ void rope_sincos(float x, float* s, float* c) {
    const float k = rintf(x * 0.15915494309189535f);
    float r = fmaf(-k, 6.28318548202514648f, x);
    r = fmaf(-k, -1.74845553e-7f, r);
    __sincosf(r, s, c);
}

// Activations are applied inside unrolled epilogue loops (16-64 copies per kernel), so the body must stay small: the
// erf form of GELU is an out-of-line call, everything else is a handful of instructions on the fast-math intrinsics
// (an inlined switch over erff / tanhf made the epilogues instruction-fetch bound: a 4160 x 22016 store epilogue took
// 901 us with SiLU against 467 us without, profiles/kernel_bench_gemm_2sm.txt).
static __device__ __noinline__ float act_gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }

__device__ __forceinline__ float apply_act(float x, int act) {
    if (act == ACT_SILU) return __fdividef(x, 1.0f + __expf(-x));
    if (act == ACT_IDENTITY) return x;
    if (act == ACT_RELU) return fmaxf(x, 0.0f);
    if (act == ACT_GELU_TANH) {
        // 0.5 x (1 + tanh(u)) == x * sigmoid(2u)
        const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
        return __fdividef(x, 1.0f + __expf(-2.0f * u));
    }
    return act_gelu_erf(x);
}

// Activation over a register array. The choice of activation is made ONCE, outside the element loop: every branch is a
// straight-line unrolled loop whose elements overlap in the pipeline. (With the choice inside the loop each element
// became its own chain of basic blocks - about 150 cycles per SiLU with one epilogue warp per scheduler.)
template <int N>
__device__ __forceinline__ void apply_act_vec(float (&v)[N], int act) {
    if (act == ACT_IDENTITY) return;
    if (act == ACT_SILU) {
#pragma unroll
        for (int j = 0; j < N; ++j) v[j] = __fdividef(v[j], 1.0f + __expf(-v[j]));
    } else if (act == ACT_RELU) {
#pragma unroll
        for (int j = 0; j < N; ++j) v[j] = fmaxf(v[j], 0.0f);
    } else if (act == ACT_GELU_TANH) {
#pragma unroll
        for (int j = 0; j < N; ++j) {
            const float x = v[j];
            const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
            v[j] = __fdividef(x, 1.0f + __expf(-2.0f * u));
        }
    } else {
#pragma unroll
        for (int j = 0; j < N; ++j) v[j] = act_gelu_erf(v[j]);
    }
}

}  // namespace hds
