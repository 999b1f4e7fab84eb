// Host collectives over one POSIX shared-memory segment: all-reduce (sum / max / min), broadcast, all-gather,
// reduce-scatter and barrier for the ranks of one node. This is the transport under the CPU configuration's tensor
// parallelism (`comm.inference_all_reduce`) and under `CCLBackend` (comm/ccl.py).
//
// Segment = one cache line of counters per rank + TWO banks of one data slot per rank. A collective is a sequence of
// rounds; round s uses bank s & 1:
//     wait until every rank has finished reading round s-2 (same bank)   -> copy my part into my slot -> written = s
//     -> wait for the ranks I read from -> compute out of their slots -> read = s
// so the wait that protects slot reuse is two rounds behind and never on the critical path of back-to-back calls.
//
// All-reduce has two shapes. Small messages: every rank sums every slot itself (one exchange). Large messages: rank r
// sums only shard r of every slot, writes the result over shard r of its OWN slot (nobody else reads that part),
// publishes reduced = s, and everybody copies the reduced shards out: each rank does 1/world of the additions and the
// memory traffic per rank stays ~3 x message instead of (world + 1) x.
//
// 16-bit floats are accumulated in fp32 and rounded once. With AVX-512 (checked at run time) sums of fp32 / bf16 /
// fp16 run through 512-bit kernels; everything else goes through plain loops the compiler vectorises.
//
// Reference role: csrc/cpu/comm/shm.cpp + shm_interface.cpp + ccl.cpp (state-machine barriers over /dev/shm, AVX-512
// reduce kernels, naive vs distributed all-reduce).
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <immintrin.h>
#include <new>
#include <sched.h>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "hds_host.h"

namespace {

enum Dtype { kBf16 = 0, kFp16 = 1, kFp32 = 2, kInt32 = 3, kInt64 = 4 };
enum Op { kSum = 0, kMax = 1, kMin = 2 };

struct alignas(64) RankFlag {
    std::atomic<uint64_t> written;
    std::atomic<uint64_t> reduced;
    std::atomic<uint64_t> read;
    std::atomic<uint32_t> attached;
};

struct ShmGroup {
    int rank, world;
    size_t slot_bytes, total_bytes;
    uint8_t* base;
    RankFlag* flags;
    uint8_t* banks;   // [2][world][slot_bytes]
    uint64_t seq;
    bool avx512;
    double timeout_s;
    int64_t n_rounds, n_naive, n_distributed;
    std::string name;

    uint8_t* slot(uint64_t s, int r) const {
        return banks + ((s & 1) * static_cast<size_t>(world) + static_cast<size_t>(r)) * slot_bytes;
    }
};

constexpr size_t kNaiveMaxBytes = 32 * 1024;   // below this one exchange beats two

size_t elem_size(int dtype) {
    switch (dtype) {
        case kBf16:
        case kFp16: return 2;
        case kFp32:
        case kInt32: return 4;
        case kInt64: return 8;
        default: return 0;
    }
}

// false = gave up (a peer died or never arrived)
bool spin_until(const ShmGroup* g, const std::atomic<uint64_t>& v, uint64_t target) {
    if (v.load(std::memory_order_acquire) >= target) return true;
    int spins = 0;
    auto t0 = std::chrono::steady_clock::now();
    while (v.load(std::memory_order_acquire) < target) {
        if (++spins > 2000) {
            sched_yield();
            spins = 0;
            if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > g->timeout_s) return false;
        }
    }
    return true;
}

// ---- element conversions ----------------------------------------------------------------------------------------
inline float bf16_to_f32(uint16_t h) {
    uint32_t u = static_cast<uint32_t>(h) << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline uint16_t f32_to_bf16(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);  // NaN stays NaN
    u += 0x7fffu + ((u >> 16) & 1u);                                                      // round to nearest even
    return static_cast<uint16_t>(u >> 16);
}
inline float fp16_to_f32(uint16_t h) {
    const uint32_t sign = static_cast<uint32_t>(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu, u;
    if (exp == 0) {
        if (man == 0) u = sign;
        else {   // subnormal: renormalise
            int e = -1;
            do {
                ++e;
                man <<= 1;
            } while (!(man & 0x400u));
            u = sign | ((127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) u = sign | 0x7f800000u | (man << 13);
    else u = sign | ((exp + 112) << 23) | (man << 13);
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline uint16_t f32_to_fp16(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    const uint16_t sign = static_cast<uint16_t>((u >> 16) & 0x8000u);
    const uint32_t abs = u & 0x7fffffffu;
    if (abs > 0x7f800000u) return sign | 0x7e00u;                   // NaN
    if (abs >= 0x477ff000u) return sign | 0x7c00u;                  // overflows to inf (after rounding)
    if (abs < 0x33000001u) return sign;                             // rounds to zero
    int exp = static_cast<int>(abs >> 23) - 127;
    uint32_t man = (abs & 0x7fffffu) | 0x800000u;
    int shift = exp < -14 ? 13 + (-14 - exp) : 13;                  // subnormal results lose more bits
    uint32_t half = man >> shift;
    const uint32_t rem = man & ((1u << shift) - 1), mid = 1u << (shift - 1);
    if (rem > mid || (rem == mid && (half & 1u))) ++half;
    if (exp < -14) return sign | static_cast<uint16_t>(half);       // implicit bit carries into the exponent if it rounds up
    return sign | static_cast<uint16_t>(((exp + 15) << 10) + (half - 0x400u));
}

template <int OP, class A>
inline A combine(A a, A b) {
    if (OP == kSum) return a + b;
    if (OP == kMax) return a > b ? a : b;
    return a < b ? a : b;
}

// ---- portable reduce: out[i] = OP over srcs[r][i] ------------------------------------------------------------------
template <int OP, class T>
void reduce_plain(const uint8_t* const* srcs, int n_src, uint8_t* out_b, size_t n) {
    T* out = reinterpret_cast<T*>(out_b);
    const T* s0 = reinterpret_cast<const T*>(srcs[0]);
    for (size_t i = 0; i < n; ++i) out[i] = s0[i];
    for (int r = 1; r < n_src; ++r) {
        const T* s = reinterpret_cast<const T*>(srcs[r]);
        for (size_t i = 0; i < n; ++i) out[i] = combine<OP, T>(out[i], s[i]);
    }
}

template <int OP, float (*LOAD)(uint16_t), uint16_t (*STORE)(float)>
void reduce_half(const uint8_t* const* srcs, int n_src, uint8_t* out_b, size_t n) {
    constexpr size_t kBlock = 1024;
    float acc[kBlock];
    uint16_t* out = reinterpret_cast<uint16_t*>(out_b);
    for (size_t i0 = 0; i0 < n; i0 += kBlock) {
        const size_t m = n - i0 < kBlock ? n - i0 : kBlock;
        const uint16_t* s0 = reinterpret_cast<const uint16_t*>(srcs[0]) + i0;
        for (size_t i = 0; i < m; ++i) acc[i] = LOAD(s0[i]);
        for (int r = 1; r < n_src; ++r) {
            const uint16_t* s = reinterpret_cast<const uint16_t*>(srcs[r]) + i0;
            for (size_t i = 0; i < m; ++i) acc[i] = combine<OP, float>(acc[i], LOAD(s[i]));
        }
        for (size_t i = 0; i < m; ++i) out[i0 + i] = STORE(acc[i]);
    }
}

// ---- AVX-512 sums (compiled for the target, selected at run time) ------------------------------------------------------
#define HDS_AVX512 __attribute__((target("avx512f,avx512bw,avx512vl")))

HDS_AVX512 void sum_f32_avx512(const uint8_t* const* srcs, int n_src, uint8_t* out_b, size_t n) {
    float* out = reinterpret_cast<float*>(out_b);
    size_t i = 0;
    for (; i + 16 <= n; i += 16) {
        __m512 acc = _mm512_loadu_ps(reinterpret_cast<const float*>(srcs[0]) + i);
        for (int r = 1; r < n_src; ++r) acc = _mm512_add_ps(acc, _mm512_loadu_ps(reinterpret_cast<const float*>(srcs[r]) + i));
        _mm512_storeu_ps(out + i, acc);
    }
    if (i < n) {
        const __mmask16 m = static_cast<__mmask16>((1u << (n - i)) - 1);
        __m512 acc = _mm512_maskz_loadu_ps(m, reinterpret_cast<const float*>(srcs[0]) + i);
        for (int r = 1; r < n_src; ++r)
            acc = _mm512_add_ps(acc, _mm512_maskz_loadu_ps(m, reinterpret_cast<const float*>(srcs[r]) + i));
        _mm512_mask_storeu_ps(out + i, m, acc);
    }
}

HDS_AVX512 inline __m512 bf16x16_to_f32(__m256i h) {
    return _mm512_castsi512_ps(_mm512_slli_epi32(_mm512_cvtepu16_epi32(h), 16));
}
HDS_AVX512 inline __m256i f32_to_bf16x16(__m512 f) {
    const __m512i u = _mm512_castps_si512(f);
    const __mmask16 nan = _mm512_cmp_ps_mask(f, f, _CMP_UNORD_Q);
    const __m512i lsb = _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1));
    __m512i r = _mm512_srli_epi32(_mm512_add_epi32(u, _mm512_add_epi32(lsb, _mm512_set1_epi32(0x7fff))), 16);
    r = _mm512_mask_mov_epi32(r, nan, _mm512_or_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(0x40)));
    return _mm512_cvtepi32_epi16(r);
}

HDS_AVX512 void sum_bf16_avx512(const uint8_t* const* srcs, int n_src, uint8_t* out_b, size_t n) {
    uint16_t* out = reinterpret_cast<uint16_t*>(out_b);
    size_t i = 0;
    for (; i + 16 <= n; i += 16) {
        __m512 acc = bf16x16_to_f32(_mm256_loadu_si256(reinterpret_cast<const __m256i*>(reinterpret_cast<const uint16_t*>(srcs[0]) + i)));
        for (int r = 1; r < n_src; ++r)
            acc = _mm512_add_ps(acc, bf16x16_to_f32(_mm256_loadu_si256(
                                         reinterpret_cast<const __m256i*>(reinterpret_cast<const uint16_t*>(srcs[r]) + i))));
        _mm256_storeu_si256(reinterpret_cast<__m256i*>(out + i), f32_to_bf16x16(acc));
    }
    if (i < n) {
        const __mmask16 m = static_cast<__mmask16>((1u << (n - i)) - 1);
        __m512 acc = bf16x16_to_f32(_mm256_maskz_loadu_epi16(m, reinterpret_cast<const uint16_t*>(srcs[0]) + i));
        for (int r = 1; r < n_src; ++r)
            acc = _mm512_add_ps(acc, bf16x16_to_f32(_mm256_maskz_loadu_epi16(m, reinterpret_cast<const uint16_t*>(srcs[r]) + i)));
        _mm256_mask_storeu_epi16(out + i, m, f32_to_bf16x16(acc));
    }
}

HDS_AVX512 void sum_fp16_avx512(const uint8_t* const* srcs, int n_src, uint8_t* out_b, size_t n) {
    uint16_t* out = reinterpret_cast<uint16_t*>(out_b);
    size_t i = 0;
    for (; i + 16 <= n; i += 16) {
        __m512 acc = _mm512_cvtph_ps(_mm256_loadu_si256(reinterpret_cast<const __m256i*>(reinterpret_cast<const uint16_t*>(srcs[0]) + i)));
        for (int r = 1; r < n_src; ++r)
            acc = _mm512_add_ps(acc, _mm512_cvtph_ps(_mm256_loadu_si256(
                                         reinterpret_cast<const __m256i*>(reinterpret_cast<const uint16_t*>(srcs[r]) + i))));
        _mm256_storeu_si256(reinterpret_cast<__m256i*>(out + i), _mm512_cvtps_ph(acc, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC));
    }
    if (i < n) {
        const __mmask16 m = static_cast<__mmask16>((1u << (n - i)) - 1);
        __m512 acc = _mm512_cvtph_ps(_mm256_maskz_loadu_epi16(m, reinterpret_cast<const uint16_t*>(srcs[0]) + i));
        for (int r = 1; r < n_src; ++r)
            acc = _mm512_add_ps(acc, _mm512_cvtph_ps(_mm256_maskz_loadu_epi16(m, reinterpret_cast<const uint16_t*>(srcs[r]) + i)));
        _mm256_mask_storeu_epi16(out + i, m, _mm512_cvtps_ph(acc, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC));
    }
}

using ReduceFn = void (*)(const uint8_t* const*, int, uint8_t*, size_t);

template <int OP>
ReduceFn pick_plain(int dtype) {
    switch (dtype) {
        case kBf16: return reduce_half<OP, bf16_to_f32, f32_to_bf16>;
        case kFp16: return reduce_half<OP, fp16_to_f32, f32_to_fp16>;
        case kFp32: return reduce_plain<OP, float>;
        case kInt32: return reduce_plain<OP, int32_t>;
        case kInt64: return reduce_plain<OP, int64_t>;
        default: return nullptr;
    }
}

ReduceFn pick_reduce(const ShmGroup* g, int dtype, int op) {
    if (op == kSum && g->avx512) {
        if (dtype == kFp32) return sum_f32_avx512;
        if (dtype == kBf16) return sum_bf16_avx512;
        if (dtype == kFp16) return sum_fp16_avx512;
    }
    if (op == kSum) return pick_plain<kSum>(dtype);
    if (op == kMax) return pick_plain<kMax>(dtype);
    if (op == kMin) return pick_plain<kMin>(dtype);
    return nullptr;
}

// ---- round bookkeeping ------------------------------------------------------------------------------------------
// Start round s: my slot of bank s & 1 may be overwritten once every rank has finished READING round s - 2.
bool begin_round(ShmGroup* g, uint64_t* s_out) {
    const uint64_t s = ++g->seq;
    ++g->n_rounds;
    if (s > 2)
        for (int r = 0; r < g->world; ++r)
            if (!spin_until(g, g->flags[r].read, s - 2)) return false;
    *s_out = s;
    return true;
}
inline void publish_written(ShmGroup* g, uint64_t s) { g->flags[g->rank].written.store(s, std::memory_order_release); }
inline void publish_reduced(ShmGroup* g, uint64_t s) { g->flags[g->rank].reduced.store(s, std::memory_order_release); }
inline void publish_read(ShmGroup* g, uint64_t s) {
    // counters are monotonic per rank: a round that skips a stage still moves every counter forward
    if (g->flags[g->rank].written.load(std::memory_order_relaxed) < s) publish_written(g, s);
    if (g->flags[g->rank].reduced.load(std::memory_order_relaxed) < s) publish_reduced(g, s);
    g->flags[g->rank].read.store(s, std::memory_order_release);
}
bool wait_all_written(ShmGroup* g, uint64_t s) {
    for (int r = 0; r < g->world; ++r)
        if (!spin_until(g, g->flags[r].written, s)) return false;
    return true;
}
bool wait_all_reduced(ShmGroup* g, uint64_t s) {
    for (int r = 0; r < g->world; ++r)
        if (!spin_until(g, g->flags[r].reduced, s)) return false;
    return true;
}

constexpr int kMaxWorld = 256;

}  // namespace

extern "C" {

void* hds_shm_open(const char* name, int rank, int world, int64_t slot_bytes) {
    if (!name || rank < 0 || world < 1 || world > kMaxWorld || rank >= world || slot_bytes < 64) return nullptr;
    // slots are cut into `world` shards of whole cache lines by the large all-reduce
    const size_t slot = (static_cast<size_t>(slot_bytes) + 63) & ~size_t(63);
    const size_t header = sizeof(RankFlag) * static_cast<size_t>(world);
    const size_t total = header + 2 * slot * static_cast<size_t>(world);
    std::string path = std::string("/") + name;
    int fd = shm_open(path.c_str(), O_CREAT | O_RDWR, 0600);
    if (fd < 0) return nullptr;
    if (ftruncate(fd, static_cast<off_t>(total)) != 0) {  // every rank sets the same size; new pages are zero
        close(fd);
        return nullptr;
    }
    void* p = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return nullptr;
    auto* g = new (std::nothrow) ShmGroup();
    if (!g) {
        munmap(p, total);
        return nullptr;
    }
    g->rank = rank;
    g->world = world;
    g->slot_bytes = slot;
    g->total_bytes = total;
    g->base = static_cast<uint8_t*>(p);
    g->flags = reinterpret_cast<RankFlag*>(g->base);
    g->banks = g->base + header;
    g->seq = 0;
    g->n_rounds = g->n_naive = g->n_distributed = 0;
    g->name = path;
    const char* simd = std::getenv("HDS_SHM_SIMD");
    g->avx512 = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl") &&
                !(simd && std::strcmp(simd, "off") == 0);
    const char* to = std::getenv("HDS_SHM_TIMEOUT_S");
    g->timeout_s = to ? std::atof(to) : 600.0;
    if (!(g->timeout_s > 0)) g->timeout_s = 600.0;
    g->flags[rank].attached.store(1, std::memory_order_release);
    auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < world; ++r) {  // rendezvous: nobody starts a round before everybody has mapped the segment
        int spins = 0;
        while (g->flags[r].attached.load(std::memory_order_acquire) == 0)
            if (++spins > 2000) {
                sched_yield();
                spins = 0;
                if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > g->timeout_s) {
                    munmap(p, total);
                    delete g;
                    return nullptr;
                }
            }
    }
    return g;
}

// In place over the group. dtype: 0 bf16, 1 fp16, 2 fp32, 3 int32, 4 int64; op: 0 sum, 1 max, 2 min.
// Returns 0, -1 (bad argument) or -2 (a peer did not arrive within HDS_SHM_TIMEOUT_S).
int hds_shm_allreduce_op(void* handle, void* data, int64_t numel, int dtype, int op) {
    auto* g = static_cast<ShmGroup*>(handle);
    const size_t elem = elem_size(dtype);
    if (!g || (!data && numel) || numel < 0 || !elem) return -1;
    ReduceFn reduce = pick_reduce(g, dtype, op);
    if (!reduce) return -1;
    const int W = g->world;
    const size_t per_chunk = g->slot_bytes / elem;
    uint8_t* bytes = static_cast<uint8_t*>(data);
    const uint8_t* srcs[kMaxWorld];
    for (size_t done = 0; done < static_cast<size_t>(numel); done += per_chunk) {
        const size_t n = static_cast<size_t>(numel) - done < per_chunk ? static_cast<size_t>(numel) - done : per_chunk;
        uint8_t* mine = bytes + done * elem;
        uint64_t s;
        if (!begin_round(g, &s)) return -2;
        std::memcpy(g->slot(s, g->rank), mine, n * elem);
        publish_written(g, s);
        if (!wait_all_written(g, s)) return -2;
        // shard = whole cache lines, the same cut on every rank
        const size_t shard = ((n + W - 1) / W * elem + 63) / 64 * 64 / elem;
        if (n * elem <= kNaiveMaxBytes || W == 1 || shard * (W - 1) >= n) {
            for (int r = 0; r < W; ++r) srcs[r] = g->slot(s, r);   // same order on every rank: bit-identical results
            reduce(srcs, W, mine, n);
            ++g->n_naive;
        } else {
            const size_t lo = shard * g->rank, m = lo < n ? (n - lo < shard ? n - lo : shard) : 0;
            if (m) {
                // the destination is shard `rank` of my own slot, which is also source 0 (every kernel loads before it stores)
                srcs[0] = g->slot(s, g->rank) + lo * elem;
                for (int r = 0, k = 1; r < W; ++r)
                    if (r != g->rank) srcs[k++] = g->slot(s, r) + lo * elem;
                reduce(srcs, W, g->slot(s, g->rank) + lo * elem, m);
            }
            publish_reduced(g, s);
            if (!wait_all_reduced(g, s)) return -2;
            for (int r = 0; r < W; ++r) {
                const size_t o = shard * r;
                if (o >= n) break;
                std::memcpy(mine + o * elem, g->slot(s, r) + o * elem, (n - o < shard ? n - o : shard) * elem);
            }
            ++g->n_distributed;
        }
        publish_read(g, s);
    }
    return 0;
}

int hds_shm_allreduce(void* handle, void* data, int64_t numel, int dtype) {
    return hds_shm_allreduce_op(handle, data, numel, dtype, kSum);
}

int hds_shm_broadcast(void* handle, void* data, int64_t nbytes, int root) {
    auto* g = static_cast<ShmGroup*>(handle);
    if (!g || (!data && nbytes) || nbytes < 0 || root < 0 || root >= g->world) return -1;
    uint8_t* bytes = static_cast<uint8_t*>(data);
    for (size_t done = 0; done < static_cast<size_t>(nbytes); done += g->slot_bytes) {
        const size_t n = static_cast<size_t>(nbytes) - done < g->slot_bytes ? static_cast<size_t>(nbytes) - done : g->slot_bytes;
        uint64_t s;
        if (!begin_round(g, &s)) return -2;
        if (g->rank == root) {
            std::memcpy(g->slot(s, root), bytes + done, n);
            publish_written(g, s);
        } else {
            if (!spin_until(g, g->flags[root].written, s)) return -2;
            std::memcpy(bytes + done, g->slot(s, root), n);
        }
        publish_read(g, s);
    }
    return 0;
}

// out = [world][nbytes_each], rank order.
int hds_shm_allgather(void* handle, const void* in, void* out, int64_t nbytes_each) {
    auto* g = static_cast<ShmGroup*>(handle);
    if (!g || ((!in || !out) && nbytes_each) || nbytes_each < 0) return -1;
    const uint8_t* src = static_cast<const uint8_t*>(in);
    uint8_t* dst = static_cast<uint8_t*>(out);
    for (size_t done = 0; done < static_cast<size_t>(nbytes_each); done += g->slot_bytes) {
        const size_t n = static_cast<size_t>(nbytes_each) - done < g->slot_bytes ? static_cast<size_t>(nbytes_each) - done : g->slot_bytes;
        uint64_t s;
        if (!begin_round(g, &s)) return -2;
        std::memcpy(g->slot(s, g->rank), src + done, n);
        publish_written(g, s);
        for (int i = 0; i < g->world; ++i) {   // start with my right-hand neighbour: ranks do not all hit slot 0 at once
            const int r = (g->rank + 1 + i) % g->world;
            if (!spin_until(g, g->flags[r].written, s)) return -2;
            std::memcpy(dst + static_cast<size_t>(r) * static_cast<size_t>(nbytes_each) + done, g->slot(s, r), n);
        }
        publish_read(g, s);
    }
    return 0;
}

// in = [world][numel_out]; out = OP over ranks of in[rank].
int hds_shm_reduce_scatter(void* handle, const void* in, void* out, int64_t numel_out, int dtype, int op) {
    auto* g = static_cast<ShmGroup*>(handle);
    const size_t elem = elem_size(dtype);
    if (!g || ((!in || !out) && numel_out) || numel_out < 0 || !elem) return -1;
    ReduceFn reduce = pick_reduce(g, dtype, op);
    if (!reduce) return -1;
    const int W = g->world;
    const size_t per_dest = g->slot_bytes / elem / W;      // elements of one destination per round
    if (per_dest == 0) return -1;
    const uint8_t* src = static_cast<const uint8_t*>(in);
    uint8_t* dst = static_cast<uint8_t*>(out);
    const uint8_t* srcs[kMaxWorld];
    for (size_t done = 0; done < static_cast<size_t>(numel_out); done += per_dest) {
        const size_t n = static_cast<size_t>(numel_out) - done < per_dest ? static_cast<size_t>(numel_out) - done : per_dest;
        uint64_t s;
        if (!begin_round(g, &s)) return -2;
        uint8_t* my_slot = g->slot(s, g->rank);
        for (int d = 0; d < W; ++d)
            std::memcpy(my_slot + static_cast<size_t>(d) * per_dest * elem,
                        src + (static_cast<size_t>(d) * static_cast<size_t>(numel_out) + done) * elem, n * elem);
        publish_written(g, s);
        if (!wait_all_written(g, s)) return -2;
        for (int r = 0; r < W; ++r) srcs[r] = g->slot(s, r) + static_cast<size_t>(g->rank) * per_dest * elem;
        reduce(srcs, W, dst + done * elem, n);
        publish_read(g, s);
    }
    return 0;
}

int hds_shm_barrier(void* handle) {
    auto* g = static_cast<ShmGroup*>(handle);
    if (!g) return -1;
    uint64_t s;
    if (!begin_round(g, &s)) return -2;
    publish_written(g, s);
    if (!wait_all_written(g, s)) return -2;
    publish_read(g, s);
    return 0;
}

// out[0] = 1 when the AVX-512 kernels are in use, out[1] = rounds so far, out[2] / out[3] = all-reduce chunks that took
// the one-exchange / the shard-per-rank shape.
void hds_shm_info(void* handle, int64_t* out) {
    auto* g = static_cast<ShmGroup*>(handle);
    if (!g || !out) return;
    out[0] = g->avx512 ? 1 : 0;
    out[1] = g->n_rounds;
    out[2] = g->n_naive;
    out[3] = g->n_distributed;
}

void hds_shm_close(void* handle, int unlink_segment) {
    auto* g = static_cast<ShmGroup*>(handle);
    if (!g) return;
    munmap(g->base, g->total_bytes);
    if (unlink_segment) shm_unlink(g->name.c_str());
    delete g;
}

}  // extern "C"
