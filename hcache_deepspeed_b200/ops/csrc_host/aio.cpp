// Asynchronous file I/O handle for the NVMe tier (see include/hds_host.h).
//
// A request is cut into block_size pieces. Each of the `threads` workers owns a kernel AIO context
// (io_setup / io_submit / io_getevents, raw system calls: the role of libaio in the reference,
// csrc/aio/common/deepspeed_aio_common.cpp:71-98) and keeps up to queue_depth pieces IN FLIGHT at the
// device: it submits aligned O_DIRECT pieces as iocbs, reaps completions and refills, so the queue depth
// seen by the NVMe drive is threads x queue_depth, not the thread count. Pieces that cannot go through
// kernel AIO — unaligned heads/tails, files on a filesystem without O_DIRECT (tmpfs), or a kernel/sandbox
// without io_setup — are executed synchronously with pread/pwrite on a buffered descriptor by the same
// worker. HDS_AIO_ENGINE=threads forces the synchronous path (queue depth = thread count).
#include "hds_host.h"

#include <errno.h>
#include <fcntl.h>
#include <linux/aio_abi.h>
#include <stdlib.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Request {
    int64_t id = 0;
    int fd_direct = -1;
    int fd_buffered = -1;
    bool is_write = false;
    std::atomic<int64_t> bytes_done{0};
    std::atomic<int> error{0};
    std::atomic<int> pieces_left{0};
    bool finished = false;
};

struct Piece {
    std::shared_ptr<Request> req;
    char* buf;
    int64_t nbytes;
    int64_t offset;
};

struct AioHandle {
    int64_t block_size;
    int queue_depth;
    bool use_direct;
    std::vector<std::thread> workers;
    std::deque<Piece> queue;
    std::map<int64_t, std::shared_ptr<Request>> requests;
    std::mutex mu;
    std::condition_variable cv_work, cv_done;
    bool stop = false;
    int64_t next_id = 1;

    // statistics (hds_aio_stats): pieces completed through kernel AIO / synchronously, and the deepest in-flight queue seen
    std::atomic<int64_t> n_kaio{0}, n_sync{0};
    std::atomic<int> max_inflight{0};
    bool want_kaio = true;

    static long sys_io_setup(unsigned n, aio_context_t* ctx) { return syscall(SYS_io_setup, n, ctx); }
    static long sys_io_destroy(aio_context_t ctx) { return syscall(SYS_io_destroy, ctx); }
    static long sys_io_submit(aio_context_t ctx, long n, struct iocb** ios) { return syscall(SYS_io_submit, ctx, n, ios); }
    static long sys_io_getevents(aio_context_t ctx, long min_nr, long nr, struct io_event* ev, struct timespec* ts) {
        return syscall(SYS_io_getevents, ctx, min_nr, nr, ev, ts);
    }

    void complete(Piece& p) {
        if (p.req->pieces_left.fetch_sub(1) == 1) {
            if (p.req->fd_direct >= 0) close(p.req->fd_direct);
            if (p.req->fd_buffered >= 0) close(p.req->fd_buffered);
            std::lock_guard<std::mutex> lock(mu);
            p.req->finished = true;
            cv_done.notify_all();
        }
    }

    static bool piece_aligned(const Piece& p) {
        return ((reinterpret_cast<uintptr_t>(p.buf) | static_cast<uintptr_t>(p.offset) | static_cast<uintptr_t>(p.nbytes)) & 4095) == 0 &&
               p.nbytes > 0;
    }

    void worker_loop() {
        aio_context_t ctx = 0;
        const int qd = queue_depth;
        bool kaio = want_kaio && sys_io_setup(static_cast<unsigned>(qd), &ctx) == 0;
        std::vector<struct iocb> cbs(qd);
        std::vector<Piece> slots(qd);
        std::vector<int> free_slots;
        for (int i = qd - 1; i >= 0; --i) free_slots.push_back(i);
        std::vector<struct io_event> events(qd);
        int inflight = 0;
        for (;;) {
            // fill: take pieces while there is room; block only when nothing is in flight
            while (inflight < qd) {
                Piece p;
                {
                    std::unique_lock<std::mutex> lock(mu);
                    if (inflight == 0) cv_work.wait(lock, [&] { return stop || !queue.empty(); });
                    if (queue.empty()) {
                        if (stop && inflight == 0) {
                            lock.unlock();
                            if (kaio) sys_io_destroy(ctx);
                            return;
                        }
                        break;
                    }
                    p = std::move(queue.front());
                    queue.pop_front();
                }
                if (kaio && p.req->fd_direct >= 0 && piece_aligned(p)) {
                    const int slot = free_slots.back();
                    struct iocb* cb = &cbs[slot];
                    memset(cb, 0, sizeof(*cb));
                    cb->aio_fildes = static_cast<uint32_t>(p.req->fd_direct);
                    cb->aio_lio_opcode = p.req->is_write ? IOCB_CMD_PWRITE : IOCB_CMD_PREAD;
                    cb->aio_buf = reinterpret_cast<uint64_t>(p.buf);
                    cb->aio_nbytes = static_cast<uint64_t>(p.nbytes);
                    cb->aio_offset = p.offset;
                    cb->aio_data = static_cast<uint64_t>(slot);
                    struct iocb* list[1] = {cb};
                    long rc;
                    do { rc = sys_io_submit(ctx, 1, list); } while (rc < 0 && errno == EINTR);
                    if (rc == 1) {
                        free_slots.pop_back();
                        slots[slot] = std::move(p);
                        ++inflight;
                        int seen = max_inflight.load();
                        while (inflight > seen && !max_inflight.compare_exchange_weak(seen, inflight)) {}
                        continue;
                    }
                    // EAGAIN / EINVAL (filesystem without AIO support): do this piece synchronously
                }
                run_piece(p);
                n_sync.fetch_add(1);
                complete(p);
            }
            if (inflight == 0) continue;
            long n;
            do { n = sys_io_getevents(ctx, 1, qd, events.data(), nullptr); } while (n < 0 && errno == EINTR);
            for (long i = 0; i < n; ++i) {
                const int slot = static_cast<int>(events[i].data);
                Piece& p = slots[slot];
                const int64_t res = static_cast<int64_t>(events[i].res);
                if (res == p.nbytes) {
                    p.req->bytes_done.fetch_add(res);
                    n_kaio.fetch_add(1);
                } else if (res >= 0 && (res > 0 || p.req->is_write)) {
                    // short transfer: finish the remainder synchronously (buffered descriptor, any alignment)
                    p.req->bytes_done.fetch_add(res);
                    Piece rest{p.req, p.buf + res, p.nbytes - res, p.offset + res};
                    run_piece(rest);
                    n_kaio.fetch_add(1);
                } else {
                    int expected = 0;
                    p.req->error.compare_exchange_strong(expected, res < 0 ? static_cast<int>(-res) : EIO);
                }
                complete(p);
                slots[slot] = Piece();
                free_slots.push_back(slot);
                --inflight;
            }
            if (n < 0) {  // the context is unusable: drain what is in flight as errors and continue synchronously
                for (int sl = 0; sl < qd; ++sl)
                    if (slots[sl].req) {
                        int expected = 0;
                        slots[sl].req->error.compare_exchange_strong(expected, EIO);
                        complete(slots[sl]);
                        slots[sl] = Piece();
                    }
                free_slots.clear();
                for (int i = qd - 1; i >= 0; --i) free_slots.push_back(i);
                inflight = 0;
                kaio = false;
            }
        }
    }

    static void run_piece(Piece& p) {
        const bool aligned = ((reinterpret_cast<uintptr_t>(p.buf) | static_cast<uintptr_t>(p.offset) |
                               static_cast<uintptr_t>(p.nbytes)) & 4095) == 0;
        const int fd = (aligned && p.req->fd_direct >= 0) ? p.req->fd_direct : p.req->fd_buffered;
        int64_t done = 0;
        while (done < p.nbytes) {
            ssize_t r = p.req->is_write ? pwrite(fd, p.buf + done, p.nbytes - done, p.offset + done)
                                        : pread(fd, p.buf + done, p.nbytes - done, p.offset + done);
            if (r < 0) {
                if (errno == EINTR) continue;
                int expected = 0;
                p.req->error.compare_exchange_strong(expected, errno);
                return;
            }
            if (r == 0) {  // short file on read
                int expected = 0;
                p.req->error.compare_exchange_strong(expected, EIO);
                return;
            }
            done += r;
        }
        p.req->bytes_done.fetch_add(done);
    }

    int64_t submit(void* buffer, const char* path, int64_t nbytes, int64_t file_offset, bool is_write) {
        auto req = std::make_shared<Request>();
        req->is_write = is_write;
        const int base_flags = is_write ? (O_WRONLY | O_CREAT) : O_RDONLY;
        req->fd_buffered = open(path, base_flags, 0644);
        if (req->fd_buffered < 0) return -1;
        if (use_direct) req->fd_direct = open(path, base_flags | O_DIRECT, 0644);  // may fail (tmpfs): fall back
        std::vector<Piece> pieces;
        char* buf = static_cast<char*>(buffer);
        for (int64_t off = 0; off < nbytes; off += block_size)
            pieces.push_back(Piece{req, buf + off, std::min(block_size, nbytes - off), file_offset + off});
        if (pieces.empty()) pieces.push_back(Piece{req, buf, 0, file_offset});
        req->pieces_left = static_cast<int>(pieces.size());
        {
            std::lock_guard<std::mutex> lock(mu);
            req->id = next_id++;
            requests[req->id] = req;
            for (auto& p : pieces) queue.push_back(std::move(p));
        }
        cv_work.notify_all();
        return req->id;
    }
};

}  // namespace

extern "C" {

void* hds_aio_create(int64_t block_size, int32_t queue_depth, int32_t threads, int32_t use_direct) {
    auto* h = new AioHandle();
    h->block_size = block_size > 0 ? block_size : (1 << 20);
    h->queue_depth = queue_depth > 0 ? queue_depth : 8;
    h->use_direct = use_direct != 0;
    const char* eng = getenv("HDS_AIO_ENGINE");
    h->want_kaio = h->use_direct && !(eng && strcmp(eng, "threads") == 0);
    const int n = threads > 0 ? threads : 1;
    for (int i = 0; i < n; ++i) h->workers.emplace_back([h] { h->worker_loop(); });
    return h;
}

void hds_aio_destroy(void* hv) {
    auto* h = static_cast<AioHandle*>(hv);
    if (!h) return;
    {
        std::lock_guard<std::mutex> lock(h->mu);
        h->stop = true;
    }
    h->cv_work.notify_all();
    for (auto& t : h->workers) t.join();
    delete h;
}

int64_t hds_aio_pread(void* hv, void* buffer, const char* path, int64_t nbytes, int64_t file_offset) {
    return static_cast<AioHandle*>(hv)->submit(buffer, path, nbytes, file_offset, false);
}

int64_t hds_aio_pwrite(void* hv, const void* buffer, const char* path, int64_t nbytes, int64_t file_offset) {
    return static_cast<AioHandle*>(hv)->submit(const_cast<void*>(buffer), path, nbytes, file_offset, true);
}

int64_t hds_aio_wait(void* hv, int64_t request_id) {
    auto* h = static_cast<AioHandle*>(hv);
    std::unique_lock<std::mutex> lock(h->mu);
    if (request_id == 0) {
        h->cv_done.wait(lock, [&] {
            for (auto& kv : h->requests)
                if (!kv.second->finished) return false;
            return true;
        });
        int64_t total = 0;
        int err = 0;
        for (auto& kv : h->requests) {
            total += kv.second->bytes_done.load();
            if (!err) err = kv.second->error.load();
        }
        h->requests.clear();
        return err ? -static_cast<int64_t>(err) : total;
    }
    auto it = h->requests.find(request_id);
    if (it == h->requests.end()) return -EINVAL;
    auto req = it->second;
    h->cv_done.wait(lock, [&] { return req->finished; });
    h->requests.erase(request_id);
    const int err = req->error.load();
    return err ? -static_cast<int64_t>(err) : req->bytes_done.load();
}

int hds_aio_poll(void* hv, int64_t request_id) {
    auto* h = static_cast<AioHandle*>(hv);
    std::lock_guard<std::mutex> lock(h->mu);
    auto it = h->requests.find(request_id);
    if (it == h->requests.end()) return -EINVAL;
    return it->second->finished ? 1 : 0;
}

int32_t hds_aio_pending(void* hv) {
    auto* h = static_cast<AioHandle*>(hv);
    std::lock_guard<std::mutex> lock(h->mu);
    int32_t n = 0;
    for (auto& kv : h->requests)
        if (!kv.second->finished) ++n;
    return n;
}

/* out[0] = pieces completed through kernel AIO, out[1] = pieces done synchronously, out[2] = deepest in-flight queue of
 * one worker. */
void hds_aio_stats(void* hv, int64_t* out) {
    auto* h = static_cast<AioHandle*>(hv);
    out[0] = h->n_kaio.load();
    out[1] = h->n_sync.load();
    out[2] = h->max_inflight.load();
}

int hds_mem_lock(void* ptr, int64_t nbytes) { return mlock(ptr, static_cast<size_t>(nbytes)) == 0 ? 0 : -errno; }
int hds_mem_unlock(void* ptr, int64_t nbytes) { return munlock(ptr, static_cast<size_t>(nbytes)) == 0 ? 0 : -errno; }

}  // extern "C"
