// hfe_host.cuh -- host side of hfe_predict_host for large batches: the call train_ensemble_public.py:62 makes
// (clf.predict_proba(X) on a pageable, row-major float64 matrix as load_data_public.py:11-14 yields it).
//
// The batch is cut into chunks that travel through a ring of slots.  Per chunk
//   1. a small pool of host threads narrows the caller's float64 rows to float32 into the slot's pinned buffer
//      (round-to-nearest, the cast SK/ensemble/_gb.py:1579-1604 applies; 68 instead of 136 bytes per row then cross
//      PCIe, which is what bounds the whole call),
//   2. the slot's stream copies the buffer to the device, transposes it (ingest kernel), scores it and copies the
//      float32 probabilities (and any optional output) back into pinned memory,
//   3. once that stream is done, the threads widen the probabilities into the caller's float64 [n,2] array
//      (p0 = 1 - p1 in float64, as SK/linear_model/_base.py:440 computes it) while later chunks are in flight.
// All buffers belong to the model handle and are allocated on the first large call; one large host call at a time
// per handle.
#pragma once

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <thread>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

#include "hfe_model.cuh"

namespace hfe {

// fork-join pool: run(fn) calls fn(t, T) on T threads (the caller is thread 0) and returns when all are done.  A large call
// forks twice per chunk, ~100 us apart, so workers spin on the epoch for a short while before they go to sleep on the
// condition variable, and the caller spins on the completion count: a wake-up through the kernel costs about as much as
// a thread's share of a chunk.
class HostPool {
public:
    explicit HostPool(int threads) : n_(std::max(threads, 1)) {
        for (int t = 1; t < n_; ++t) workers_.emplace_back([this, t] { loop(t); });
    }
    ~HostPool() {
        {
            std::lock_guard<std::mutex> lock(mu_);
            stop_ = true;
            epoch_.fetch_add(1, std::memory_order_release);
        }
        cv_.notify_all();
        for (auto& w : workers_) w.join();
    }
    int size() const { return n_; }
#ifdef HFE_DEV
    double busy_seconds = 0;             // wall time spent inside run() (development builds)
    static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#endif
    void run(const std::function<void(int, int)>& fn) {
#ifdef HFE_DEV
        struct Timer { double& acc; double t0; ~Timer() { acc += now() - t0; } } timer{busy_seconds, now()};
#endif
        if (n_ == 1) { fn(0, 1); return; }
        fn_ = &fn;
        pending_.store(n_ - 1, std::memory_order_relaxed);
        {
            std::lock_guard<std::mutex> lock(mu_);          // orders the epoch bump against a worker about to sleep
            epoch_.fetch_add(1, std::memory_order_release);
        }
        if (sleepers_.load(std::memory_order_acquire) > 0) cv_.notify_all();
        fn(0, n_);
        for (unsigned spins = 0; pending_.load(std::memory_order_acquire) != 0; ++spins) {
            if (spins < 4096) cpu_relax();
            else std::this_thread::yield();
        }
        fn_ = nullptr;
    }

private:
    static void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    }
    void loop(int t) {
        unsigned long long seen = 0;
        for (;;) {
            bool fresh = false;
            for (int spins = 0; spins < 20000; ++spins) {                    // ~50-100 us
                if (epoch_.load(std::memory_order_acquire) != seen) { fresh = true; break; }
                cpu_relax();
            }
            if (!fresh) {
                std::unique_lock<std::mutex> lock(mu_);
                sleepers_.fetch_add(1, std::memory_order_release);
                cv_.wait(lock, [&] { return epoch_.load(std::memory_order_acquire) != seen; });
                sleepers_.fetch_sub(1, std::memory_order_release);
            }
            seen = epoch_.load(std::memory_order_acquire);
            if (stop_) return;
            (*fn_)(t, n_);
            pending_.fetch_sub(1, std::memory_order_release);
        }
    }
    int n_;
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_;
    const std::function<void(int, int)>* fn_ = nullptr;
    std::atomic<unsigned long long> epoch_{0};
    std::atomic<int> pending_{0};
    std::atomic<int> sleepers_{0};
    bool stop_ = false;
};

constexpr int HOST_SLOTS = 3;                 // chunks in flight: one being narrowed, one on the wire, one being scored
constexpr int64_t HOST_CHUNK_ROWS = 1 << 18;  // 262 144 rows: 17.8 MB of float32 per slot
constexpr int64_t HOST_PIPE_MIN_ROWS = 1 << 16;   // smaller batches take the single-shot path

struct HostSlot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    float* h_in = nullptr;                    // pinned [chunk, F] float32 row-major
    unsigned char* h_out = nullptr;           // pinned: probabilities f32 [chunk,2] | labels | decision | meta | status words
    unsigned char* d_buf = nullptr;           // device: raw rows | column-major | outputs | status
    int64_t begin = 0, rows = 0;              // the chunk this slot currently carries
    bool busy = false;
    bool as_bytes = false;                    // the chunk sits in h_in as uint8 (else float32)
};

struct HostPipeline {
    int device = 0;
    int64_t chunk = HOST_CHUNK_ROWS;
    size_t off_raw = 0, off_col = 0, off_out = 0, off_lab = 0, off_dec = 0, off_meta = 0, off_status = 0, d_bytes = 0;
    size_t h_off_lab = 0, h_off_dec = 0, h_off_meta = 0, h_off_status = 0, h_bytes = 0;
    HostSlot slot[HOST_SLOTS];
    HostPool* pool = nullptr;
    ~HostPipeline() {
        for (auto& s : slot) {
            if (s.stream) cudaStreamDestroy(s.stream);
            if (s.done) cudaEventDestroy(s.done);
            if (s.h_in) cudaFreeHost(s.h_in);
            if (s.h_out) cudaFreeHost(s.h_out);
            if (s.d_buf) cudaFree(s.d_buf);
        }
        delete pool;
    }
};

inline int host_thread_count() {
    int want = 0;
    if (const char* e = std::getenv("HFE_HOST_THREADS")) want = std::atoi(e);      // deployment knob, not a layout knob
    if (want < 1) {
        cpu_set_t set;
        int avail = (sched_getaffinity(0, sizeof(set), &set) == 0) ? CPU_COUNT(&set) : (int)std::thread::hardware_concurrency();
        want = std::min(std::max(avail, 1), 16);
    }
    return std::min(want, 64);
}

inline int host_pipeline_create(int device, int n_members, HostPipeline** out) {
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    HostPipeline* hp = new HostPipeline();
    hp->device = device;
    const size_t c = (size_t)hp->chunk;
    size_t at = 0;
    hp->off_raw = at;    at += up(c * F * sizeof(float));
    hp->off_col = at;    at += up(c * F * sizeof(float));
    hp->off_out = at;    at += up(c * 2 * sizeof(float));
    hp->off_lab = at;    at += up(c);
    hp->off_dec = at;    at += up(c * sizeof(double));
    hp->off_meta = at;   at += up(c * (size_t)n_members * sizeof(double));
    hp->off_status = at; at += up((HFE_STATUS_WORDS + HFE_STATUS_HEADER_WORDS) * sizeof(uint32_t));
    hp->d_bytes = at;
    at = up(c * 2 * sizeof(float));
    hp->h_off_lab = at;    at += up(c);
    hp->h_off_dec = at;    at += up(c * sizeof(double));
    hp->h_off_meta = at;   at += up(c * (size_t)n_members * sizeof(double));
    hp->h_off_status = at; at += 256;
    hp->h_bytes = at;
    for (auto& s : hp->slot) {
        cudaError_t e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void**>(&s.h_in), c * F * sizeof(float), cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void**>(&s.h_out), hp->h_bytes, cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&s.d_buf), hp->d_bytes);
        if (e != cudaSuccess) {
            delete hp;
            return fail(HFE_ERR_CUDA, std::string("host pipeline buffers: ") + cudaGetErrorString(e));
        }
    }
    hp->pool = new HostPool(host_thread_count());
    *out = hp;
    return HFE_OK;
}

// float64 -> float32, round to nearest; probabilities back to float64 pairs.  The destinations (a pinned slot, the caller's
// result array) are written once and not read again by this thread, so where the CPU has AVX2 the stores are non-temporal:
// no read-for-ownership of the destination lines, a quarter less memory traffic on a call that is bound by exactly that.
inline void narrow_rows_plain(const double* __restrict__ src, float* __restrict__ dst, size_t count) {
    for (size_t i = 0; i < count; ++i) dst[i] = (float)src[i];
}

inline void widen_proba_plain(const float* __restrict__ src, double* __restrict__ dst, size_t rows) {
    for (size_t i = 0; i < rows; ++i) {
        const double p1 = (double)src[2 * i + 1];
        dst[2 * i] = 1.0 - p1;
        dst[2 * i + 1] = p1;
    }
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) inline void narrow_rows_avx2(const double* __restrict__ src, float* __restrict__ dst, size_t count) {
    size_t i = 0;
    while (i < count && (reinterpret_cast<uintptr_t>(dst + i) & 31)) { dst[i] = (float)src[i]; ++i; }
    for (; i + 8 <= count; i += 8) {
        _mm_prefetch(reinterpret_cast<const char*>(src + i + 128), _MM_HINT_T0);
        const __m128 lo = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i));
        const __m128 hi = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i + 4));
        _mm256_stream_ps(dst + i, _mm256_insertf128_ps(_mm256_castps128_ps256(lo), hi, 1));
    }
    for (; i < count; ++i) dst[i] = (float)src[i];
    _mm_sfence();
}

__attribute__((target("avx2"))) inline void widen_proba_avx2(const float* __restrict__ src, double* __restrict__ dst, size_t rows) {
    if (reinterpret_cast<uintptr_t>(dst) & 15) { widen_proba_plain(src, dst, rows); return; }
    size_t i = 0;
    if (rows && (reinterpret_cast<uintptr_t>(dst) & 31)) { widen_proba_plain(src, dst, 1); i = 1; }
    const __m256d one = _mm256_set1_pd(1.0);
    for (; i + 2 <= rows; i += 2) {
        const __m256d v = _mm256_cvtps_pd(_mm_loadu_ps(src + 2 * i));          // p0a p1a p0b p1b (float32 as scored)
        const __m256d p1 = _mm256_permute_pd(v, 0xF);                         // p1a p1a p1b p1b
        _mm256_stream_pd(dst + 2 * i, _mm256_blend_pd(_mm256_sub_pd(one, p1), p1, 0xA));   // 1-p1a p1a 1-p1b p1b
    }
    if (i < rows) widen_proba_plain(src + 2 * i, dst + 2 * i, rows - i);
    _mm_sfence();
}

inline bool host_has_avx2() {
    static const bool yes = __builtin_cpu_supports("avx2");
    return yes;
}
#endif

// float64 -> uint8 where that loses nothing (cohort variables are small non-negative integers): returns false at the first
// value that is not an integer in [0, 255] -- NaN included -- and the chunk is then narrowed to float32 instead.
inline bool narrow_rows_u8_plain(const double* __restrict__ src, unsigned char* __restrict__ dst, size_t count) {
    for (size_t i = 0; i < count; ++i) {
        const double v = src[i];
        const int q = (v >= 0.0 && v <= 255.0) ? (int)v : -1;
        if (q < 0 || (double)q != v) return false;
        dst[i] = (unsigned char)q;
    }
    return true;
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) inline bool narrow_rows_u8_avx2(const double* __restrict__ src, unsigned char* __restrict__ dst, size_t count) {
    size_t i = 0;
    while (i < count && (reinterpret_cast<uintptr_t>(dst + i) & 15)) {
        if (!narrow_rows_u8_plain(src + i, dst + i, 1)) return false;
        ++i;
    }
    const __m128i pick = _mm_setr_epi8(0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
    __m256d bad = _mm256_setzero_pd();
    __m128i high = _mm_setzero_si128();
    for (; i + 16 <= count; i += 16) {
        _mm_prefetch(reinterpret_cast<const char*>(src + i + 128), _MM_HINT_T0);
        _mm_prefetch(reinterpret_cast<const char*>(src + i + 136), _MM_HINT_T0);
        __m128i out = _mm_setzero_si128();
        __m128i part[4];
        for (int k = 0; k < 4; ++k) {
            const __m256d v = _mm256_loadu_pd(src + i + 4 * k);
            const __m128i q = _mm256_cvttpd_epi32(v);                              // NaN / out of range -> 0x80000000
            bad = _mm256_or_pd(bad, _mm256_cmp_pd(_mm256_cvtepi32_pd(q), v, _CMP_NEQ_UQ));
            high = _mm_or_si128(high, q);                                           // any bit above the low byte (or the sign) set?
            part[k] = _mm_shuffle_epi8(q, pick);                                    // the four low bytes, in the low dword
        }
        out = _mm_unpacklo_epi64(_mm_unpacklo_epi32(part[0], part[1]), _mm_unpacklo_epi32(part[2], part[3]));
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i), out);
    }
    _mm_sfence();
    if (_mm256_movemask_pd(bad) != 0) return false;
    if (!_mm_testz_si128(high, _mm_set1_epi32(~0xFF))) return false;
    return narrow_rows_u8_plain(src + i, dst + i, count - i);
}
#endif

#if defined(__x86_64__)
// AVX-512 versions (half the instructions per value; the conversion, not memory, bounds the narrowing per thread)
__attribute__((target("avx512f,avx512bw,avx512vl,avx512dq"))) inline bool narrow_rows_u8_avx512(const double* __restrict__ src, unsigned char* __restrict__ dst, size_t count) {
    size_t i = 0;
    while (i < count && (reinterpret_cast<uintptr_t>(dst + i) & 15)) {
        if (!narrow_rows_u8_plain(src + i, dst + i, 1)) return false;
        ++i;
    }
    __mmask8 bad = 0;
    __m512i high = _mm512_setzero_si512();
    for (; i + 16 <= count; i += 16) {
        _mm_prefetch(reinterpret_cast<const char*>(src + i + 128), _MM_HINT_T0);           // 1 KB ahead of the loads
        _mm_prefetch(reinterpret_cast<const char*>(src + i + 136), _MM_HINT_T0);
        const __m512d v0 = _mm512_loadu_pd(src + i), v1 = _mm512_loadu_pd(src + i + 8);
        const __m256i q0 = _mm512_cvttpd_epi32(v0), q1 = _mm512_cvttpd_epi32(v1);          // NaN / out of range -> 0x80000000
        bad = (__mmask8)(bad | _mm512_cmp_pd_mask(_mm512_cvtepi32_pd(q0), v0, _CMP_NEQ_UQ) |
                         _mm512_cmp_pd_mask(_mm512_cvtepi32_pd(q1), v1, _CMP_NEQ_UQ));
        const __m512i q = _mm512_inserti64x4(_mm512_castsi256_si512(q0), q1, 1);
        high = _mm512_or_si512(high, q);                                                    // any bit above the low byte (or the sign) set?
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i), _mm512_cvtepi32_epi8(q));     // the sixteen low bytes
    }
    _mm_sfence();
    if (bad != 0) return false;
    if (_mm512_test_epi32_mask(high, _mm512_set1_epi32(~0xFF)) != 0) return false;
    return narrow_rows_u8_plain(src + i, dst + i, count - i);
}

__attribute__((target("avx512f,avx512bw,avx512vl,avx512dq"))) inline void narrow_rows_avx512(const double* __restrict__ src, float* __restrict__ dst, size_t count) {
    size_t i = 0;
    while (i < count && (reinterpret_cast<uintptr_t>(dst + i) & 63)) { dst[i] = (float)src[i]; ++i; }
    for (; i + 16 <= count; i += 16) {
        _mm_prefetch(reinterpret_cast<const char*>(src + i + 128), _MM_HINT_T0);
        _mm_prefetch(reinterpret_cast<const char*>(src + i + 136), _MM_HINT_T0);
        const __m256 lo = _mm512_cvtpd_ps(_mm512_loadu_pd(src + i));
        const __m256 hi = _mm512_cvtpd_ps(_mm512_loadu_pd(src + i + 8));
        _mm512_stream_ps(dst + i, _mm512_insertf32x8(_mm512_castps256_ps512(lo), hi, 1));
    }
    for (; i < count; ++i) dst[i] = (float)src[i];
    _mm_sfence();
}

inline bool host_has_avx512() {
    static const bool yes = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
                            __builtin_cpu_supports("avx512vl") && __builtin_cpu_supports("avx512dq");
    return yes;
}
#endif

inline bool narrow_rows_u8(const double* src, unsigned char* dst, size_t count) {
#if defined(__x86_64__)
    if (host_has_avx512()) return narrow_rows_u8_avx512(src, dst, count);
#endif
#if defined(__x86_64__)
    if (host_has_avx2()) return narrow_rows_u8_avx2(src, dst, count);
#endif
    return narrow_rows_u8_plain(src, dst, count);
}

inline void narrow_rows(const double* src, float* dst, size_t count) {
#if defined(__x86_64__)
    if (host_has_avx512()) { narrow_rows_avx512(src, dst, count); return; }
    if (host_has_avx2()) { narrow_rows_avx2(src, dst, count); return; }
#endif
    narrow_rows_plain(src, dst, count);
}

inline void widen_proba(const float* src, double* dst, size_t rows) {
#if defined(__x86_64__)
    if (host_has_avx2()) { widen_proba_avx2(src, dst, rows); return; }
#endif
    widen_proba_plain(src, dst, rows);
}

}  // namespace hfe
