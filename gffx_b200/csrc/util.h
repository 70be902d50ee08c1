// util.h — small host-side helpers shared by the engine's C++ sources.
#pragma once
#include <cstdint>
#include <functional>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

namespace gffx {

enum Status : int { OK = 0, E_ARG = -1, E_IO = -2, E_FORMAT = -3, E_CUDA = -4, E_STATE = -5 };

struct Err : std::runtime_error {
    int code;
    Err(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

static constexpr uint64_t MISSING64 = UINT64_MAX;  // sentinel of src/index_loader/gof.rs:7
static constexpr uint32_t NONE32 = UINT32_MAX;

// read-only mmap of a whole file (the reference mmaps every input, src/index_loader/core.rs:14-17)
class Mapped {
   public:
    Mapped() = default;
    ~Mapped();
    Mapped(const Mapped&) = delete;
    Mapped& operator=(const Mapped&) = delete;
    void open(const std::string& path);          // throws Err(E_IO)
    bool try_open(const std::string& path);      // false when the file cannot be opened
    const uint8_t* data() const { return p_; }
    size_t size() const { return n_; }

   private:
    const uint8_t* p_ = nullptr;
    size_t n_ = 0;
    int fd_ = -1;
};

// ---- text predicates that mirror the Rust std calls the reference uses -------------------
bool utf8_ok(const uint8_t* s, size_t n);                    // std::str::from_utf8(..).is_ok()
inline bool ascii_ws(uint8_t c) {                            // u8::is_ascii_whitespace
    return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\x0c';
}
std::string_view rtrim_unicode(std::string_view s);          // str::trim_end
std::string_view trim_unicode(std::string_view s);           // str::trim
bool parse_u32_strict(const uint8_t* s, size_t n, uint32_t* out);   // digits only, checked (intersect.rs:526-538)
bool parse_u32_rust(std::string_view s, uint32_t* out);              // str::parse::<u32> ('+' allowed)

// Host-side data parallelism: jobs 0 .. n_jobs-1, each run once, on at most max_threads threads at a time (the caller is one
// of them), from a pool of threads that live as long as the process.  Threads made for a few milliseconds of work each (the
// loaders' phases are that short) start on their creator's CPU and are done before the scheduler has spread them: r03 measured
// eight such threads finishing one after the other.  Pool threads are already spread out and only need waking.  Re-entrant (a job
// may call parallel_jobs; the caller always works on its own jobs, so nothing waits for a free worker); the first exception of
// a job is rethrown in the caller once every job has ended.
void parallel_jobs(size_t n_jobs, int max_threads, const std::function<void(size_t)>& job);

// GFFX_TIMING=1: wall time of the load phases on stderr (profiles/load_time.py)
struct PhaseTimer {
    const char* what;
    double t0;
    bool on;
    static double now();
    explicit PhaseTimer(const char* w);
    void stop();   // report now instead of at the end of the scope
    ~PhaseTimer();
};

inline const uint8_t* find_byte(const uint8_t* p, const uint8_t* end, uint8_t c) {
    return p < end ? (const uint8_t*)memchr(p, c, (size_t)(end - p)) : nullptr;
}

}  // namespace gffx
