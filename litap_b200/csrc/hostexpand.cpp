// hostexpand.cpp — host side of the packed downloads (hostcopy.cu): int32 -> fp64 while the values move from the pinned
// staging buffer into the caller's array.  calc_form's outputs are rounded numbers (k / 1000, or whole degrees), so they
// cross PCIe as the integers k (half the bytes) and are widened here with exactly the quotient the device would have
// stored: k / 1000 correctly rounded.  AVX2 + FMA when the CPU has them (Markstein's correction from RN(1/1000): five
// flops per value instead of a division; 73-84 GB/s written with 8-12 threads on the GPU box's host,
// profiles/host_mb/expand_i32.cpp), the IEEE division otherwise.  INT32_MIN is NA, INT32_MIN + 1 is -0.
#include <cstddef>
#include <cstdint>
#include <cmath>
#include <immintrin.h>

namespace {
inline double one(int32_t k, int kind) {
  if (k == INT32_MIN) return std::nan("");
  if (k == INT32_MIN + 1) return -0.0;
  return kind == 1 ? (double)k / 1000.0 : (double)k;
}
__attribute__((target("avx2,fma"))) void expand_avx2(const int32_t* s, double* d, size_t n, int kind) {
  size_t i = 0;
  while (i < n && ((uintptr_t)(d + i) & 31)) { d[i] = one(s[i], kind); ++i; }
  const __m256d c1000 = _mm256_set1_pd(1000.0), r = _mm256_set1_pd(1.0 / 1000.0);
  const __m256d nanv = _mm256_set1_pd(std::nan(""));
  const __m128i na = _mm_set1_epi32(INT32_MIN), nz = _mm_set1_epi32(INT32_MIN + 1);
  const __m256d negz = _mm256_set1_pd(-0.0);
  for (; i + 4 <= n; i += 4) {
    const __m128i k = _mm_loadu_si128((const __m128i*)(s + i));
    const __m256d a = _mm256_cvtepi32_pd(k);
    __m256d q = a;
    if (kind == 1) {
      q = _mm256_mul_pd(a, r);                       // RN(a * RN(1/1000)): within 2 ulp
      __m256d e = _mm256_fnmadd_pd(q, c1000, a);     // exact residual a - 1000 q
      q = _mm256_fmadd_pd(e, r, q);                  // faithful
      e = _mm256_fnmadd_pd(q, c1000, a);
      q = _mm256_fmadd_pd(e, r, q);                  // = RN(a / 1000) (Markstein 1990; checked against the division on
    }                                                //   every int32 multiple pattern by tests/test_host_cpu.py)
    const __m256d m = _mm256_castsi256_pd(_mm256_cvtepi32_epi64(_mm_cmpeq_epi32(k, na)));
    const __m256d z = _mm256_castsi256_pd(_mm256_cvtepi32_epi64(_mm_cmpeq_epi32(k, nz)));
    _mm256_stream_pd(d + i, _mm256_blendv_pd(_mm256_blendv_pd(q, negz, z), nanv, m));
  }
  for (; i < n; ++i) d[i] = one(s[i], kind);
  _mm_sfence();
}
void expand_scalar(const int32_t* s, double* d, size_t n, int kind) {
  for (size_t i = 0; i < n; ++i) d[i] = one(s[i], kind);
}
}  // namespace

// kind 1: thousandths (k / 1000), kind 2: whole numbers
extern "C" void litap_host_expand_i32(const int32_t* src, double* dst, size_t n, int kind) {
  static const bool fast = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma");
  if (fast) expand_avx2(src, dst, n, kind);
  else expand_scalar(src, dst, n, kind);
}
// the same with the division only (tests compare the two)
extern "C" void litap_host_expand_i32_ref(const int32_t* src, double* dst, size_t n, int kind) { expand_scalar(src, dst, n, kind); }
