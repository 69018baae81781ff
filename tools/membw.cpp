// Host memory read bandwidth with T threads over a 1.36 GB buffer (the size of 10 M float64 rows x 17): the floor under the
// float64-numpy entry point, which has to read every input byte once.   g++ -O3 -mavx2 -pthread tools/membw.cpp -o /tmp/membw
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#include <immintrin.h>
int main(int argc, char** argv) {
    const size_t n = 170000000;                      // doubles
    double* x = static_cast<double*>(aligned_alloc(64, n * sizeof(double)));
    for (size_t i = 0; i < n; ++i) x[i] = (double)(i & 3);
    for (int T : {1, 2, 4, 8, 16}) {
        if (argc > 1 && T > atoi(argv[1])) break;
        double best = 1e9;
        for (int rep = 0; rep < 3; ++rep) {
            std::vector<std::thread> th;
            std::vector<double> sums(T);
            auto t0 = std::chrono::steady_clock::now();
            for (int t = 0; t < T; ++t)
                th.emplace_back([&, t] {
                    const size_t a = n * t / T, b = n * (t + 1) / T;
                    __m256d acc0 = _mm256_setzero_pd(), acc1 = acc0, acc2 = acc0, acc3 = acc0;
                    size_t i = a;
                    for (; i + 16 <= b; i += 16) {
                        acc0 = _mm256_add_pd(acc0, _mm256_loadu_pd(x + i));
                        acc1 = _mm256_add_pd(acc1, _mm256_loadu_pd(x + i + 4));
                        acc2 = _mm256_add_pd(acc2, _mm256_loadu_pd(x + i + 8));
                        acc3 = _mm256_add_pd(acc3, _mm256_loadu_pd(x + i + 12));
                    }
                    double tmp[4];
                    _mm256_storeu_pd(tmp, _mm256_add_pd(_mm256_add_pd(acc0, acc1), _mm256_add_pd(acc2, acc3)));
                    sums[t] = tmp[0] + tmp[1] + tmp[2] + tmp[3];
                });
            for (auto& h : th) h.join();
            best = std::min(best, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
        }
        std::printf("threads %2d: read %.1f GB/s (%.1f ms for 1.36 GB)\n", T, n * 8 / best / 1e9, best * 1e3);
    }
    // the same pass with the byte narrowing of the host pipeline (float64 -> uint8, checked, streaming stores)
    unsigned char* y = static_cast<unsigned char*>(aligned_alloc(64, n));
    for (int T : {4, 8, 16}) {
        if (argc > 1 && T > atoi(argv[1])) break;
        double best = 1e9;
        for (int rep = 0; rep < 3; ++rep) {
            std::vector<std::thread> th;
            auto t0 = std::chrono::steady_clock::now();
            for (int t = 0; t < T; ++t)
                th.emplace_back([&, t] {
                    const size_t a = (n * t / T) & ~(size_t)15, b = t + 1 == T ? n : (n * (t + 1) / T) & ~(size_t)15;
                    const __m128i pick = _mm_setr_epi8(0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
                    __m256d bad = _mm256_setzero_pd();
                    __m128i high = _mm_setzero_si128();
                    const size_t pd = argc > 2 ? (size_t)atoi(argv[2]) : 0;
                    for (size_t i = a; i + 16 <= b; i += 16) {
                        if (pd) { _mm_prefetch(reinterpret_cast<const char*>(x + i + pd), _MM_HINT_T0); _mm_prefetch(reinterpret_cast<const char*>(x + i + pd + 8), _MM_HINT_T0); }
                        __m128i part[4];
                        for (int k = 0; k < 4; ++k) {
                            const __m256d v = _mm256_loadu_pd(x + i + 4 * k);
                            const __m128i q = _mm256_cvttpd_epi32(v);
                            bad = _mm256_or_pd(bad, _mm256_cmp_pd(_mm256_cvtepi32_pd(q), v, _CMP_NEQ_UQ));
                            high = _mm_or_si128(high, q);
                            part[k] = _mm_shuffle_epi8(q, pick);
                        }
                        _mm_stream_si128(reinterpret_cast<__m128i*>(y + i),
                                         _mm_unpacklo_epi64(_mm_unpacklo_epi32(part[0], part[1]), _mm_unpacklo_epi32(part[2], part[3])));
                    }
                    _mm_sfence();
                    if (_mm256_movemask_pd(bad) || !_mm_testz_si128(high, _mm_set1_epi32(~0xFF))) std::abort();
                });
            for (auto& h : th) h.join();
            best = std::min(best, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
        }
        std::printf("threads %2d: float64 -> uint8 narrowing %.1f GB/s of input (%.1f ms for 1.36 GB)\n", T, n * 8 / best / 1e9, best * 1e3);
    }
    free(y);
    free(x);
}
