// Shared-memory all-reduce for intra-node CPU tensor parallelism (the CPU configuration's low-latency collective).
//
// One POSIX shared-memory segment per group: a header of per-rank sequence counters (one cache line each) followed by
// one data slot per rank. A reduction of a chunk is: copy my data into my slot -> publish "written" -> wait for every
// rank's "written" -> sum all slots into my own output (every rank does the full sum: no second exchange, the data is
// in shared L3 / DRAM anyway) -> publish "read" -> before the slots are reused, wait for every rank's "read".
// Messages larger than a slot go through in slot-sized chunks.
//
// Reference role: csrc/cpu/comm/shm.cpp + shm_interface.cpp (torch.ops.deepspeed.inference_all_reduce_; state-machine
// barriers over /dev/shm with AVX-512 reduce kernels). This version relies on the compiler's vectoriser (-O2 -ftree-vectorize).
#include <atomic>
#include <cerrno>
#include <cstdint>
#include <cstring>
#include <fcntl.h>
#include <new>
#include <sched.h>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "hds_host.h"

namespace {

struct alignas(64) RankFlag {
    std::atomic<uint64_t> written;
    std::atomic<uint64_t> read;
    std::atomic<uint32_t> attached;
};

struct ShmGroup {
    int rank, world;
    size_t slot_bytes, total_bytes;
    uint8_t* base;
    RankFlag* flags;
    uint8_t* slots;
    uint64_t seq;
    std::string name;
};

inline void spin_until(const std::atomic<uint64_t>& v, uint64_t target) {
    int spins = 0;
    while (v.load(std::memory_order_acquire) < target) {
        if (++spins > 2000) {
            sched_yield();
            spins = 0;
        }
    }
}

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

void reduce_f32(ShmGroup* g, float* out, size_t n) {
    const float* s0 = reinterpret_cast<const float*>(g->slots);
    for (size_t i = 0; i < n; ++i) out[i] = s0[i];
    for (int r = 1; r < g->world; ++r) {
        const float* s = reinterpret_cast<const float*>(g->slots + static_cast<size_t>(r) * g->slot_bytes);
        for (size_t i = 0; i < n; ++i) out[i] += s[i];
    }
}

void reduce_bf16(ShmGroup* g, uint16_t* out, size_t n) {
    constexpr size_t kBlock = 1024;
    float acc[kBlock];
    for (size_t i0 = 0; i0 < n; i0 += kBlock) {
        const size_t m = n - i0 < kBlock ? n - i0 : kBlock;
        for (size_t i = 0; i < m; ++i) acc[i] = 0.0f;
        for (int r = 0; r < g->world; ++r) {
            const uint16_t* s = reinterpret_cast<const uint16_t*>(g->slots + static_cast<size_t>(r) * g->slot_bytes) + i0;
            for (size_t i = 0; i < m; ++i) acc[i] += bf16_to_f32(s[i]);
        }
        for (size_t i = 0; i < m; ++i) out[i0 + i] = f32_to_bf16(acc[i]);
    }
}

}  // namespace

extern "C" {

void* hds_shm_open(const char* name, int rank, int world, int64_t slot_bytes) {
    if (!name || rank < 0 || world < 1 || rank >= world || slot_bytes < 64) return nullptr;
    const size_t slot = (static_cast<size_t>(slot_bytes) + 63) & ~size_t(63);
    const size_t header = sizeof(RankFlag) * static_cast<size_t>(world);
    const size_t total = header + slot * static_cast<size_t>(world);
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
    g->slots = g->base + header;
    g->seq = 0;
    g->name = path;
    g->flags[rank].attached.store(1, std::memory_order_release);
    for (int r = 0; r < world; ++r) {  // rendezvous: nobody reduces before everybody has mapped the segment
        int spins = 0;
        while (g->flags[r].attached.load(std::memory_order_acquire) == 0)
            if (++spins > 2000) {
                sched_yield();
                spins = 0;
            }
    }
    return g;
}

// dtype: 0 bf16, 2 fp32 (the codes of the kernel library). In place, sum.
int hds_shm_allreduce(void* handle, void* data, int64_t numel, int dtype) {
    auto* g = static_cast<ShmGroup*>(handle);
    if (!g || !data || numel < 0 || (dtype != 0 && dtype != 2)) return -1;
    const size_t elem = dtype == 2 ? 4 : 2;
    const size_t per_chunk = g->slot_bytes / elem;
    uint8_t* bytes = static_cast<uint8_t*>(data);
    for (size_t done = 0; done < static_cast<size_t>(numel); done += per_chunk) {
        const size_t n = static_cast<size_t>(numel) - done < per_chunk ? static_cast<size_t>(numel) - done : per_chunk;
        const uint64_t s = ++g->seq;
        for (int r = 0; r < g->world; ++r) spin_until(g->flags[r].read, s - 1);  // previous round fully consumed
        std::memcpy(g->slots + static_cast<size_t>(g->rank) * g->slot_bytes, bytes + done * elem, n * elem);
        g->flags[g->rank].written.store(s, std::memory_order_release);
        for (int r = 0; r < g->world; ++r) spin_until(g->flags[r].written, s);
        if (dtype == 2) reduce_f32(g, reinterpret_cast<float*>(bytes + done * elem), n);
        else reduce_bf16(g, reinterpret_cast<uint16_t*>(bytes + done * elem), n);
        g->flags[g->rank].read.store(s, std::memory_order_release);
    }
    return 0;
}

void hds_shm_close(void* handle, int unlink_segment) {
    auto* g = static_cast<ShmGroup*>(handle);
    if (!g) return;
    munmap(g->base, g->total_bytes);
    if (unlink_segment) shm_unlink(g->name.c_str());
    delete g;
}

}  // extern "C"
