// Asynchronous file I/O handle for the NVMe tier (see include/hds_host.h).
//
// A request is cut into block_size pieces; `threads` workers execute pread/pwrite on them, at most
// queue_depth pieces of one request in flight. With use_direct the file is opened O_DIRECT and pieces
// whose buffer address, file offset and length are 4096-aligned bypass the page cache; unaligned
// heads/tails go through a second, buffered descriptor.
#include "hds_host.h"

#include <errno.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
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

    void worker_loop() {
        for (;;) {
            Piece p;
            {
                std::unique_lock<std::mutex> lock(mu);
                cv_work.wait(lock, [&] { return stop || !queue.empty(); });
                if (stop && queue.empty()) return;
                p = std::move(queue.front());
                queue.pop_front();
            }
            run_piece(p);
            if (p.req->pieces_left.fetch_sub(1) == 1) {
                if (p.req->fd_direct >= 0) close(p.req->fd_direct);
                if (p.req->fd_buffered >= 0) close(p.req->fd_buffered);
                std::lock_guard<std::mutex> lock(mu);
                p.req->finished = true;
                cv_done.notify_all();
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

int hds_mem_lock(void* ptr, int64_t nbytes) { return mlock(ptr, static_cast<size_t>(nbytes)) == 0 ? 0 : -errno; }
int hds_mem_unlock(void* ptr, int64_t nbytes) { return munlock(ptr, static_cast<size_t>(nbytes)) == 0 ? 0 : -errno; }

}  // extern "C"
