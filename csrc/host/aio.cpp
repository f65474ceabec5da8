// H1b (SURVEY.md §2.3 "ZeRO-Offload / Infinity", X9): asynchronous file I/O for the NVMe tier of the
// optimizer-state offload (`offload_optimizer.device = "nvme"`, ref ai_engine/deepspeed_launcher.py:29-32,
// 197-203). The reference reaches this capability only through the DeepSpeed wheel's libaio op; here it
// is a small thread-pool engine: requests are split into block-aligned slices that worker threads
// service with pread/pwrite in parallel (NVMe drives need queue depth, not one big sequential call),
// completion is tracked per ticket so the optimizer can overlap "read chunk i+1", "Adam on chunk i" and
// "write chunk i-1". O_DIRECT is used when the caller asks for it and the buffer/offset/length are
// 4 KiB-aligned; otherwise the page cache is used (and stays coherent with mmap views of the file).
#include <fcntl.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Slice {
  long long ticket;
  int fd;
  bool write;
  char* buf;
  long long nbytes;
  long long offset;
};

struct TicketState {
  int pending = 0;
  int error = 0;   // first errno seen (0 = ok); -1 = short read (EOF)
};

class AioEngine {
 public:
  AioEngine(int nthreads, long long block_bytes) : block_(std::max<long long>(block_bytes, 1 << 16)), stop_(false) {
    nthreads = std::max(1, nthreads);
    for (int i = 0; i < nthreads; ++i) workers_.emplace_back([this] { loop(); });
  }
  ~AioEngine() {
    { std::lock_guard<std::mutex> l(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }

  long long submit(int fd, bool write, void* buf, long long nbytes, long long offset) {
    std::lock_guard<std::mutex> l(mu_);
    const long long ticket = ++next_ticket_;
    TicketState& ts = tickets_[ticket];
    char* p = static_cast<char*>(buf);
    long long done = 0;
    while (done < nbytes) {
      const long long n = std::min(block_, nbytes - done);
      queue_.push_back(Slice{ticket, fd, write, p + done, n, offset + done});
      ++ts.pending;
      done += n;
    }
    if (ts.pending == 0) { /* zero-length request completes immediately */ }
    cv_.notify_all();
    return ticket;
  }

  // Blocks until the ticket is complete; returns 0, an errno, or -1 for a short read. The ticket is retired.
  int wait(long long ticket) {
    std::unique_lock<std::mutex> l(mu_);
    auto it = tickets_.find(ticket);
    if (it == tickets_.end()) return EINVAL;
    done_cv_.wait(l, [&] { return it->second.pending == 0; });
    const int err = it->second.error;
    tickets_.erase(it);
    return err;
  }

  int wait_all() {
    std::unique_lock<std::mutex> l(mu_);
    done_cv_.wait(l, [&] {
      for (auto& kv : tickets_) if (kv.second.pending != 0) return false;
      return true;
    });
    int err = 0;
    for (auto& kv : tickets_) if (kv.second.error != 0 && err == 0) err = kv.second.error;
    tickets_.clear();
    return err;
  }

  long long bytes_read() const { return bytes_read_.load(); }
  long long bytes_written() const { return bytes_written_.load(); }

 private:
  static int do_io(const Slice& s) {
    long long done = 0;
    while (done < s.nbytes) {
      const ssize_t r = s.write ? ::pwrite(s.fd, s.buf + done, (size_t)(s.nbytes - done), (off_t)(s.offset + done))
                                : ::pread(s.fd, s.buf + done, (size_t)(s.nbytes - done), (off_t)(s.offset + done));
      if (r < 0) {
        if (errno == EINTR) continue;
        return errno;
      }
      if (r == 0) return -1;  // EOF before the request was satisfied
      done += r;
    }
    return 0;
  }

  void loop() {
    for (;;) {
      Slice s;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [&] { return stop_ || !queue_.empty(); });
        if (queue_.empty()) return;   // stop_ and drained
        s = queue_.front();
        queue_.pop_front();
      }
      const int err = do_io(s);
      if (err == 0) (s.write ? bytes_written_ : bytes_read_).fetch_add(s.nbytes);
      {
        std::lock_guard<std::mutex> l(mu_);
        auto it = tickets_.find(s.ticket);
        if (it != tickets_.end()) {
          if (err != 0 && it->second.error == 0) it->second.error = err;
          if (--it->second.pending == 0) done_cv_.notify_all();
        }
      }
    }
  }

  const long long block_;
  std::vector<std::thread> workers_;
  std::mutex mu_;
  std::condition_variable cv_, done_cv_;
  std::deque<Slice> queue_;
  std::map<long long, TicketState> tickets_;
  long long next_ticket_ = 0;
  bool stop_;
  std::atomic<long long> bytes_read_{0}, bytes_written_{0};
};

}  // namespace

extern "C" {

// nthreads <= 0 -> 8 workers; block_bytes <= 0 -> 8 MiB slices.
void* bz_aio_create(int nthreads, long long block_bytes) {
  return new AioEngine(nthreads <= 0 ? 8 : nthreads, block_bytes <= 0 ? (8LL << 20) : block_bytes);
}

void bz_aio_destroy(void* h) { delete static_cast<AioEngine*>(h); }

// Opens (creating if asked) a state file; size_bytes > 0 extends it with ftruncate. direct != 0 requests
// O_DIRECT and silently falls back to buffered I/O when the filesystem refuses it (tmpfs, overlayfs).
// Returns the fd or -errno; *direct_used (may be NULL) reports which mode is active.
int bz_aio_open(const char* path, int create, long long size_bytes, int direct, int* direct_used) {
  int flags = O_RDWR | (create ? O_CREAT : 0);
  int fd = -1;
  int used = 0;
#ifdef O_DIRECT
  if (direct) {
    fd = ::open(path, flags | O_DIRECT, 0644);
    if (fd >= 0) used = 1;
  }
#endif
  if (fd < 0) fd = ::open(path, flags, 0644);
  if (fd < 0) return -errno;
  if (size_bytes > 0) {
    struct stat st;
    if (::fstat(fd, &st) != 0 || st.st_size < size_bytes) {
      if (::ftruncate(fd, (off_t)size_bytes) != 0) {
        const int e = errno;
        ::close(fd);
        return -e;
      }
    }
  }
  if (direct_used) *direct_used = used;
  return fd;
}

int bz_aio_close(int fd) { return ::close(fd) == 0 ? 0 : errno; }

int bz_aio_fsync(int fd) { return ::fdatasync(fd) == 0 ? 0 : errno; }

long long bz_aio_read(void* h, int fd, void* buf, long long nbytes, long long offset) {
  return static_cast<AioEngine*>(h)->submit(fd, false, buf, nbytes, offset);
}

long long bz_aio_write(void* h, int fd, const void* buf, long long nbytes, long long offset) {
  return static_cast<AioEngine*>(h)->submit(fd, true, const_cast<void*>(buf), nbytes, offset);
}

int bz_aio_wait(void* h, long long ticket) { return static_cast<AioEngine*>(h)->wait(ticket); }

int bz_aio_wait_all(void* h) { return static_cast<AioEngine*>(h)->wait_all(); }

long long bz_aio_bytes_read(void* h) { return static_cast<AioEngine*>(h)->bytes_read(); }

long long bz_aio_bytes_written(void* h) { return static_cast<AioEngine*>(h)->bytes_written(); }

}  // extern "C"
