// Host-side runtime of libdtf_b200: background checkpoint writer.
//
// The reference saves every member every round through TensorFlow's C++ Saver, synchronously, inside
// the training loop (model_base.py / mnist_model.py: `saver.save(sess, ...)` after each train call), and
// its exploit step then copies those files.  Here the member state lives in one device arena; a save is
//   1. one asynchronous device->host copy of the arena's state region into a pinned staging buffer
//      (issued by the caller on the member's stream, followed by an event record), and
//   2. a job for this writer: wait for an event recorded on that stream at submit, write `<dir>/model.ckpt.data` (temp file + rename), then
//      the JSON index, then the `checkpoint` pointer file -- in that order, so a reader that sees the
//      pointer sees a complete checkpoint.
// The training thread never blocks on the file system; `dtf_io_wait` / `dtf_io_wait_all` are the barriers
// the file-exploit path, restore and shutdown use.
//
// Plain C++17 + the CUDA runtime (event wait only).  No allocation on the device.
#include <cuda_runtime.h>
#include <errno.h>
#include <execinfo.h>
#include <signal.h>
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <condition_variable>
#include <deque>
#include <mutex>
#include <set>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "dtf_b200.h"

namespace {

// "The caller's device -> host copy has landed": set by a host function the submitting thread enqueues on its
// stream behind the copy (cudaLaunchHostFunc), waited for by the writer thread with a plain condition variable.
// The writer threads make NO CUDA API call at all: an earlier version synchronised a CUDA event here, and a
// cudaEventSynchronize from a writer thread racing with another member's CUDA-graph capture (members_in_flight)
// occasionally failed and invalidated that capture.
struct Landed {
    std::mutex mu;
    std::condition_variable cv;
    bool done = false;
};

struct Job {
    long long ticket;
    std::string data_path, index_path, index_text, pointer_path, pointer_text;
    const void* host_ptr;
    size_t nbytes;
    std::shared_ptr<Landed> landed;      // null for host-only jobs
    int device;
};

void CUDART_CB landed_callback(void* user) {
    // runs on a driver thread once everything queued before it on the stream has finished; must not call CUDA
    auto* keep = static_cast<std::shared_ptr<Landed>*>(user);
    {
        std::lock_guard<std::mutex> lk((*keep)->mu);
        (*keep)->done = true;
    }
    (*keep)->cv.notify_all();
    delete keep;
}

struct Writer {
    std::mutex mu;
    std::condition_variable cv_work, cv_done;
    std::deque<Job> queue;
    std::set<long long> in_flight;       // submitted, not finished
    std::vector<std::thread> threads;
    long long next_ticket = 1;
    long long bytes_written = 0;
    int errors = 0;
    bool stop = false;
};

// write `n` bytes to `path` through a temp file in the same directory, then rename over `path`
int write_atomic(const std::string& path, const void* p, size_t n) {
    const std::string tmp = path + ".tmp" + std::to_string(static_cast<long long>(getpid())) + "_" +
                            std::to_string(reinterpret_cast<uintptr_t>(p) & 0xffff);
    const int fd = ::open(tmp.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return -1;
    const char* c = static_cast<const char*>(p);
    size_t left = n;
    while (left > 0) {
        const ssize_t w = ::write(fd, c, left > (64u << 20) ? (64u << 20) : left);
        if (w < 0) {
            ::close(fd);
            ::unlink(tmp.c_str());
            return -1;
        }
        c += w;
        left -= static_cast<size_t>(w);
    }
    if (::close(fd) != 0) return -1;
    return ::rename(tmp.c_str(), path.c_str());
}

void run_job(Writer* w, const Job& j) {
    int rc = 0;
    if (j.landed) {
        std::unique_lock<std::mutex> lk(j.landed->mu);
        j.landed->cv.wait(lk, [&] { return j.landed->done; });
    }
    if (rc == 0) rc = write_atomic(j.data_path, j.host_ptr, j.nbytes);
    if (rc == 0 && !j.index_path.empty()) rc = write_atomic(j.index_path, j.index_text.data(), j.index_text.size());
    if (rc == 0 && !j.pointer_path.empty())
        rc = write_atomic(j.pointer_path, j.pointer_text.data(), j.pointer_text.size());
    std::lock_guard<std::mutex> lk(w->mu);
    if (rc != 0) {
        ++w->errors;
        fprintf(stderr, "[dtf] checkpoint writer: %s: %s\n", j.data_path.c_str(), strerror(errno));
    } else {
        w->bytes_written += static_cast<long long>(j.nbytes);
    }
    w->in_flight.erase(j.ticket);
    w->cv_done.notify_all();
}

void worker_loop(Writer* w) {
    for (;;) {
        Job j;
        {
            std::unique_lock<std::mutex> lk(w->mu);
            w->cv_work.wait(lk, [&] { return w->stop || !w->queue.empty(); });
            if (w->queue.empty()) return;          // stop requested and drained
            j = std::move(w->queue.front());
            w->queue.pop_front();
        }
        run_job(w, j);
    }
}

}  // namespace

extern "C" void* dtf_io_create(int threads) {
    Writer* w = new Writer();
    if (threads < 1) threads = 1;
    if (threads > 16) threads = 16;
    for (int i = 0; i < threads; ++i) w->threads.emplace_back(worker_loop, w);
    return w;
}

extern "C" long long dtf_io_submit(void* handle, const char* data_path, const void* host_ptr, unsigned long long nbytes,
                                   void* cuda_stream, int device, const char* index_path, const char* index_text,
                                   const char* pointer_path, const char* pointer_text) {
    Writer* w = static_cast<Writer*>(handle);
    if (w == nullptr || data_path == nullptr || (host_ptr == nullptr && nbytes > 0)) return DTF_ERR_ARG;
    Job j;
    j.data_path = data_path;
    j.host_ptr = host_ptr;
    j.nbytes = static_cast<size_t>(nbytes);
    j.device = device;
    if (device >= 0) {
        // a host function behind the caller's device->host copy marks the staging buffer as complete
        j.landed = std::make_shared<Landed>();
        auto* keep = new std::shared_ptr<Landed>(j.landed);
        if (cudaLaunchHostFunc(static_cast<cudaStream_t>(cuda_stream), landed_callback, keep) != cudaSuccess) {
            delete keep;
            return DTF_ERR_CUDA;
        }
    }
    if (index_path != nullptr) { j.index_path = index_path; j.index_text = index_text ? index_text : ""; }
    if (pointer_path != nullptr) { j.pointer_path = pointer_path; j.pointer_text = pointer_text ? pointer_text : ""; }
    std::lock_guard<std::mutex> lk(w->mu);
    if (w->stop) return DTF_ERR_ARG;
    j.ticket = w->next_ticket++;
    w->in_flight.insert(j.ticket);
    const long long t = j.ticket;
    w->queue.push_back(std::move(j));
    w->cv_work.notify_one();
    return t;
}

// Blocks until the job with this ticket has been written.  Returns DTF_OK, or DTF_ERR_CUDA if any job of this
// writer has failed so far (the message went to stderr).
extern "C" int dtf_io_wait(void* handle, long long ticket) {
    Writer* w = static_cast<Writer*>(handle);
    if (w == nullptr) return DTF_ERR_ARG;
    std::unique_lock<std::mutex> lk(w->mu);
    w->cv_done.wait(lk, [&] { return w->in_flight.count(ticket) == 0; });
    return w->errors == 0 ? DTF_OK : DTF_ERR_CUDA;
}

extern "C" int dtf_io_wait_all(void* handle) {
    Writer* w = static_cast<Writer*>(handle);
    if (w == nullptr) return DTF_ERR_ARG;
    std::unique_lock<std::mutex> lk(w->mu);
    w->cv_done.wait(lk, [&] { return w->in_flight.empty(); });
    return w->errors == 0 ? DTF_OK : DTF_ERR_CUDA;
}

extern "C" int dtf_io_pending(void* handle) {
    Writer* w = static_cast<Writer*>(handle);
    if (w == nullptr) return DTF_ERR_ARG;
    std::lock_guard<std::mutex> lk(w->mu);
    return static_cast<int>(w->in_flight.size());
}

extern "C" long long dtf_io_bytes_written(void* handle) {
    Writer* w = static_cast<Writer*>(handle);
    if (w == nullptr) return DTF_ERR_ARG;
    std::lock_guard<std::mutex> lk(w->mu);
    return w->bytes_written;
}

extern "C" void dtf_io_destroy(void* handle) {
    Writer* w = static_cast<Writer*>(handle);
    if (w == nullptr) return;
    {
        std::lock_guard<std::mutex> lk(w->mu);
        w->stop = true;
    }
    w->cv_work.notify_all();
    for (auto& t : w->threads) t.join();
    delete w;
}

// Debugging aid: print the native backtrace of the faulting thread on SIGSEGV, then chain to the previous
// handler (Python's faulthandler).  Installed by ops/native.py when DTF_CRASH_BT is set.
namespace {
struct sigaction g_prev_segv;
void segv_handler(int sig, siginfo_t* info, void* ctx) {
    void* frames[64];
    const int n = backtrace(frames, 64);
    const char msg[] = "[dtf] SIGSEGV native backtrace:\n";
    (void)!write(2, msg, sizeof(msg) - 1);
    backtrace_symbols_fd(frames, n, 2);
    if (g_prev_segv.sa_flags & SA_SIGINFO) {
        if (g_prev_segv.sa_sigaction) g_prev_segv.sa_sigaction(sig, info, ctx);
    } else if (g_prev_segv.sa_handler != SIG_DFL && g_prev_segv.sa_handler != SIG_IGN) {
        g_prev_segv.sa_handler(sig);
    }
    signal(sig, SIG_DFL);
    raise(sig);
}
}  // namespace

extern "C" void dtf_install_crash_handler(void) {
    struct sigaction sa;
    memset(&sa, 0, sizeof(sa));
    sa.sa_sigaction = segv_handler;
    sa.sa_flags = SA_SIGINFO | SA_ONSTACK;
    sigaction(SIGSEGV, &sa, &g_prev_segv);
}
