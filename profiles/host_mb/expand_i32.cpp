// Host micro-benchmark: how fast can T threads widen int32 thousandths to fp64 (k / 1000, correctly rounded) into memory that is
// already mapped?  Decides whether shipping the rounded form outputs as integers pays.
// build: g++ -O2 -pthread expand_i32.cpp -o expand_i32 ; usage: ./expand_i32 [threads] [Melems]
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <immintrin.h>
#include <thread>
#include <vector>

__attribute__((target("avx2,fma"))) static void expand_avx2(const int32_t* s, double* d, size_t n) {
  const __m256d c1000 = _mm256_set1_pd(1000.0), r = _mm256_set1_pd(1.0 / 1000.0);
  size_t i = 0;
  for (; i + 4 <= n; i += 4) {
    const __m256d a = _mm256_cvtepi32_pd(_mm_loadu_si128((const __m128i*)(s + i)));
    __m256d q = _mm256_mul_pd(a, r);                       // q0 = RN(a * RN(1/1000))
    __m256d e = _mm256_fnmadd_pd(q, c1000, a);             // exact residual
    q = _mm256_fmadd_pd(e, r, q);
    e = _mm256_fnmadd_pd(q, c1000, a);
    q = _mm256_fmadd_pd(e, r, q);                          // = RN(a / 1000) (Markstein)
    _mm256_stream_pd(d + i, q);
  }
  for (; i < n; ++i) d[i] = (double)s[i] / 1000.0;
}
static void expand_scalar(const int32_t* s, double* d, size_t n) {
  for (size_t i = 0; i < n; ++i) d[i] = (double)s[i] / 1000.0;
}
int main(int argc, char** argv) {
  const int T = argc > 1 ? atoi(argv[1]) : 12;
  const size_t n = (size_t)(argc > 2 ? atoi(argv[2]) : 64) << 20;
  std::vector<int32_t> src(n);
  for (size_t i = 0; i < n; ++i) src[i] = (int32_t)((i * 2654435761u) % 2000001) - 1000000;
  double* dst = (double*)aligned_alloc(64, n * 8);
  memset(dst, 0, n * 8);
  // exactness of the AVX2 path against the division
  {
    const size_t m = n < (size_t)(16 << 20) ? n : (size_t)(16 << 20);
    double* a = (double*)aligned_alloc(64, m * 8);
    double* b = (double*)aligned_alloc(64, m * 8);
    expand_avx2(src.data(), a, m);
    expand_scalar(src.data(), b, m);
    printf("avx2 == division on %zu values: %s\n", m, memcmp(a, b, m * 8) ? "NO" : "yes");
    free(a); free(b);
  }
  for (int mode = 0; mode < 3; ++mode) {
    double best = 1e9;
    for (int rep = 0; rep < 5; ++rep) {
      auto t0 = std::chrono::steady_clock::now();
      std::vector<std::thread> th;
      for (int t = 0; t < T; ++t)
        th.emplace_back([&, t] {
          const size_t lo = n * t / T & ~(size_t)7, hi = t + 1 == T ? n : (n * (t + 1) / T & ~(size_t)7);
          if (mode == 0) expand_avx2(src.data() + lo, dst + lo, hi - lo);
          else if (mode == 1) expand_scalar(src.data() + lo, dst + lo, hi - lo);
          else memcpy(dst + lo, (const double*)src.data() + lo / 2, (hi - lo) * 4), memcpy(dst + lo + (hi - lo) / 2, (const double*)src.data() + lo / 2, (hi - lo) * 4);
        });
      for (auto& x : th) x.join();
      best = std::min(best, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    }
    printf("%d threads, %s: %.1f ms for %zu M values = %.1f GB/s written\n", T,
           mode == 0 ? "avx2+fma expand (nt stores)" : mode == 1 ? "scalar division expand" : "memcpy of the same output bytes", best * 1e3,
           n >> 20, n * 8 / best / 1e9);
  }
  return 0;
}
