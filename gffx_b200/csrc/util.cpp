#include "util.h"
#include <algorithm>
#include <thread>
#include <mutex>
#include <memory>
#include <deque>
#include <condition_variable>
#include <atomic>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdlib>

namespace gffx {

// ---- the thread pool behind parallel_jobs -----------------------------------------------------------------------------------
namespace {
struct PoolCall {   // one parallel_jobs call: shared by the caller and the workers that picked up one of its tickets
    const std::function<void(size_t)>* job;
    size_t n;
    std::atomic<size_t> next{0}, done{0};
    std::mutex mu;
    std::condition_variable cv;
    std::exception_ptr err;
    void work() {
        for (size_t i; (i = next.fetch_add(1)) < n;) {
            try { (*job)(i); } catch (...) { std::lock_guard<std::mutex> g(mu); if (!err) err = std::current_exception(); }
            if (done.fetch_add(1) + 1 == n) { std::lock_guard<std::mutex> g(mu); cv.notify_all(); }
        }
    }
};
struct Pool {
    std::mutex mu;
    std::condition_variable cv;
    std::deque<std::shared_ptr<PoolCall>> tickets;
    size_t n_workers = 0;
    Pool() {
        const unsigned hc = std::thread::hardware_concurrency();
        n_workers = hc > 1 ? std::min<unsigned>(hc - 1, 63) : 0;
        for (size_t k = 0; k < n_workers; k++)
            std::thread([this] {
                for (;;) {
                    std::shared_ptr<PoolCall> c;
                    {
                        std::unique_lock<std::mutex> g(mu);
                        cv.wait(g, [&] { return !tickets.empty(); });
                        c = std::move(tickets.front());
                        tickets.pop_front();
                    }
                    c->work();
                }
            }).detach();
    }
};
Pool& pool() {
    static Pool* P = new Pool();   // never destroyed: its threads wait for work until the process ends
    return *P;
}
}  // namespace

void parallel_jobs(size_t n_jobs, int max_threads, const std::function<void(size_t)>& job) {
    if (n_jobs == 0) return;
    if (n_jobs == 1 || max_threads <= 1) { for (size_t i = 0; i < n_jobs; i++) job(i); return; }
    Pool& P = pool();
    const size_t helpers = std::min<size_t>({(size_t)max_threads - 1, n_jobs - 1, P.n_workers});
    if (helpers == 0) { for (size_t i = 0; i < n_jobs; i++) job(i); return; }
    auto c = std::make_shared<PoolCall>();
    c->job = &job;
    c->n = n_jobs;
    {
        std::lock_guard<std::mutex> g(P.mu);
        for (size_t k = 0; k < helpers; k++) P.tickets.push_back(c);
    }
    if (helpers == 1) P.cv.notify_one(); else P.cv.notify_all();
    c->work();
    {
        std::unique_lock<std::mutex> g(c->mu);
        c->cv.wait(g, [&] { return c->done.load() == c->n; });
    }
    // a ticket nobody picked up yet finds no job left and is dropped; `job` is not touched after this point
    if (c->err) std::rethrow_exception(c->err);
}

double PhaseTimer::now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
PhaseTimer::PhaseTimer(const char* w) : what(w), t0(0), on(getenv("GFFX_TIMING") != nullptr) { if (on) t0 = now(); }
void PhaseTimer::stop() { if (on) fprintf(stderr, "[gffx timing] %-28s %8.1f ms\n", what, (now() - t0) * 1e3); on = false; }
PhaseTimer::~PhaseTimer() { stop(); }


Mapped::~Mapped() {
    if (p_) munmap((void*)p_, n_);
    if (fd_ >= 0) ::close(fd_);
}

bool Mapped::try_open(const std::string& path) {
    fd_ = ::open(path.c_str(), O_RDONLY);
    if (fd_ < 0) return false;
    struct stat st;
    if (fstat(fd_, &st) != 0 || !S_ISREG(st.st_mode)) { ::close(fd_); fd_ = -1; return false; }
    n_ = (size_t)st.st_size;
    if (n_ == 0) return true;
    void* m = mmap(nullptr, n_, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (m == MAP_FAILED) { ::close(fd_); fd_ = -1; n_ = 0; return false; }
    p_ = (const uint8_t*)m;
    return true;
}

void Mapped::open(const std::string& path) {
    if (!try_open(path)) throw Err(E_IO, "Failed to open file: \"" + path + "\"");
}

// Table-free UTF-8 acceptance test (same language as Rust's core::str::from_utf8).
bool utf8_ok(const uint8_t* s, size_t n) {
    size_t i = 0;
    while (i < n) {
        // fast skip over ASCII, 8 bytes at a time
        while (i + 8 <= n) {
            uint64_t w;
            memcpy(&w, s + i, 8);
            if (w & 0x8080808080808080ull) break;
            i += 8;
        }
        if (i >= n) break;
        uint8_t b0 = s[i];
        if (b0 < 0x80) { i++; continue; }
        auto cont = [&](size_t k, uint8_t lo, uint8_t hi) { return i + k < n && s[i + k] >= lo && s[i + k] <= hi; };
        if (b0 >= 0xC2 && b0 <= 0xDF) {
            if (!cont(1, 0x80, 0xBF)) return false;
            i += 2;
        } else if (b0 == 0xE0) {
            if (!cont(1, 0xA0, 0xBF) || !cont(2, 0x80, 0xBF)) return false;
            i += 3;
        } else if ((b0 >= 0xE1 && b0 <= 0xEC) || b0 == 0xEE || b0 == 0xEF) {
            if (!cont(1, 0x80, 0xBF) || !cont(2, 0x80, 0xBF)) return false;
            i += 3;
        } else if (b0 == 0xED) {
            if (!cont(1, 0x80, 0x9F) || !cont(2, 0x80, 0xBF)) return false;
            i += 3;
        } else if (b0 == 0xF0) {
            if (!cont(1, 0x90, 0xBF) || !cont(2, 0x80, 0xBF) || !cont(3, 0x80, 0xBF)) return false;
            i += 4;
        } else if (b0 >= 0xF1 && b0 <= 0xF3) {
            if (!cont(1, 0x80, 0xBF) || !cont(2, 0x80, 0xBF) || !cont(3, 0x80, 0xBF)) return false;
            i += 4;
        } else if (b0 == 0xF4) {
            if (!cont(1, 0x80, 0x8F) || !cont(2, 0x80, 0xBF) || !cont(3, 0x80, 0xBF)) return false;
            i += 4;
        } else {
            return false;
        }
    }
    return true;
}

// Unicode White_Space code points as UTF-8, longest first where prefixes collide.
static const char* const kWs[] = {"\x09", "\x0a", "\x0b", "\x0c", "\x0d", "\x20", "\xc2\x85", "\xc2\xa0", "\xe1\x9a\x80",
                                  "\xe2\x80\x80", "\xe2\x80\x81", "\xe2\x80\x82", "\xe2\x80\x83", "\xe2\x80\x84",
                                  "\xe2\x80\x85", "\xe2\x80\x86", "\xe2\x80\x87", "\xe2\x80\x88", "\xe2\x80\x89",
                                  "\xe2\x80\x8a", "\xe2\x80\xa8", "\xe2\x80\xa9", "\xe2\x80\xaf", "\xe2\x81\x9f",
                                  "\xe3\x80\x80"};

std::string_view rtrim_unicode(std::string_view s) {
    bool again = true;
    while (again && !s.empty()) {
        again = false;
        for (const char* w : kWs) {
            size_t k = strlen(w);
            if (s.size() >= k && memcmp(s.data() + s.size() - k, w, k) == 0) {
                s.remove_suffix(k);
                again = true;
                break;
            }
        }
    }
    return s;
}

std::string_view trim_unicode(std::string_view s) {
    s = rtrim_unicode(s);
    bool again = true;
    while (again && !s.empty()) {
        again = false;
        for (const char* w : kWs) {
            size_t k = strlen(w);
            if (s.size() >= k && memcmp(s.data(), w, k) == 0) {
                s.remove_prefix(k);
                again = true;
                break;
            }
        }
    }
    return s;
}

bool parse_u32_strict(const uint8_t* s, size_t n, uint32_t* out) {
    if (n == 0 || n > 10) {
        // more than 10 digits can still be valid only with leading zeros
        if (n == 0) return false;
    }
    uint64_t acc = 0;
    for (size_t i = 0; i < n; i++) {
        uint8_t c = s[i];
        if (c < '0' || c > '9') return false;
        acc = acc * 10 + (uint64_t)(c - '0');
        if (acc > 0xFFFFFFFFull) return false;
    }
    *out = (uint32_t)acc;
    return true;
}

bool parse_u32_rust(std::string_view s, uint32_t* out) {
    if (!s.empty() && s.front() == '+') s.remove_prefix(1);
    return parse_u32_strict((const uint8_t*)s.data(), s.size(), out);
}

}  // namespace gffx
