// Host-side helpers shared by the translation units of libsctagger_b200.so: error reporting, the host worker pool
// (passive waiting), parallel memcpy and small RAII-less free helpers.  Internal — not part of the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: a no-op unless a profiler injects itself (nsys / ncu --nvtx)
#include "../../include/sctagger_b200.h"

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <functional>
#include <atomic>
#include <vector>

namespace scth {

inline thread_local std::string g_err;

// SCT_TRACE=1: per-call host-side phase times of sct_match on stderr (staging / waiting on the GPU / un-staging)
struct Trace {
    bool on = getenv("SCT_TRACE") != nullptr;
    double t_stage = 0, t_wait = 0, t_fetch = 0, t_check = 0, t_alloc = 0;
    static double now() {
        return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    }
};
inline thread_local Trace g_tr;

// NVTX range around a host phase (stage / H2D + kernels / wait + D2H / format): SURVEY.md §5 "tracing".
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange &) = delete;
    NvtxRange &operator=(const NvtxRange &) = delete;
};

inline int fail(int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CUDA_TRY(x)                                                                                              \
    do {                                                                                                         \
        cudaError_t e_ = (x);                                                                                    \
        if (e_ != cudaSuccess)                                                                                   \
            return fail(e_ == cudaErrorMemoryAllocation ? SCT_E_NOMEM : SCT_E_CUDA, "%s: %s (%s:%d)", #x,        \
                        cudaGetErrorString(e_), __FILE__, __LINE__);                                             \
    } while (0)
#define TRY(x)                  \
    do {                        \
        int rc_ = (x);          \
        if (rc_ != SCT_OK) return rc_; \
    } while (0)

inline int map_char(uint8_t c) { return c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : 0; }   // scTagger.py:512-516
inline uint8_t comp_char(uint8_t c) {                                                          // scTagger.py:148-152
    return c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c;
}
inline bool is_acgt(uint8_t c) { return c == 'A' || c == 'C' || c == 'G' || c == 'T'; }

// Host worker threads (-t) for staging / un-staging.  A small fork-join pool with PASSIVE waiting: with one process (or
// one handle) per GPU, eight of these run side by side, and spinning waiters (OpenMP's default) starve each other.
class HostPool {
public:
    explicit HostPool(int n_threads) : n_(std::max(1, n_threads)) {
        for (int t = 1; t < n_; ++t) workers_.emplace_back([this, t] { loop(t); });
    }
    ~HostPool() {
        {
            std::lock_guard<std::mutex> g(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto &w : workers_) w.join();
    }
    int size() const { return n_; }
    // fn(t) for t = 0..size()-1, t = 0 on the calling thread; returns when all are done
    void run(const std::function<void(int)> &fn) {
        if (n_ == 1) {
            fn(0);
            return;
        }
        {
            std::lock_guard<std::mutex> g(m_);
            fn_ = &fn;
            pending_ = n_ - 1;
            ++epoch_;
        }
        cv_.notify_all();
        fn(0);
        std::unique_lock<std::mutex> g(m_);
        done_.wait(g, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }

private:
    void loop(int t) {
        unsigned long long seen = 0;
        for (;;) {
            const std::function<void(int)> *fn;
            {
                std::unique_lock<std::mutex> g(m_);
                cv_.wait(g, [&] { return stop_ || epoch_ != seen; });
                if (stop_) return;
                seen = epoch_;
                fn = fn_;
            }
            (*fn)(t);
            {
                std::lock_guard<std::mutex> g(m_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    int n_;
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(int)> *fn_ = nullptr;
    int pending_ = 0;
    unsigned long long epoch_ = 0;
    bool stop_ = false;
};

// [0, n) split into pool.size() contiguous ranges: fn(first, last)
template <class F>
void par_ranges(HostPool *pool, int64_t n, int64_t min_per_thread, F fn) {
    const int nt = pool ? (int)std::min<int64_t>(pool->size(), std::max<int64_t>(1, n / std::max<int64_t>(1, min_per_thread))) : 1;
    if (nt <= 1) {
        fn((int64_t)0, n);
        return;
    }
    const int64_t piece = (n + nt - 1) / nt;
    pool->run([&](int t) {
        if (t >= nt) return;
        const int64_t a = std::min(n, piece * t), b = std::min(n, piece * (t + 1));
        if (b > a) fn(a, b);
    });
}

inline void par_memcpy(HostPool *pool, void *dst, const void *src, size_t bytes) {
    par_ranges(pool, (int64_t)bytes, 1 << 19, [&](int64_t a, int64_t b) {
        memcpy((char *)dst + a, (const char *)src + a, (size_t)(b - a));
    });
}

// Is [p, p + bytes) page-locked host memory the copy engines can read / write directly?
inline bool is_pinned(const void *p, size_t bytes) {
    if (!p || bytes == 0) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    if (a.type != cudaMemoryTypeHost) return false;
    if (cudaPointerGetAttributes(&a, (const char *)p + bytes - 1) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

template <class T>
void dev_free(T *&p) {
    if (p) cudaFree(p);
    p = nullptr;
}
template <class T>
void host_free(T *&p) {
    if (p) cudaFreeHost(p);
    p = nullptr;
}

}  // namespace scth
