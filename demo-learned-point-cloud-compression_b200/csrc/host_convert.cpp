// Host-side helpers of the frame staging (stage.cu), plain C++ (g++):
//  * pcc_host_f64_to_bf16: float64 -> bf16 with the two roundings the device path applies (double -> float,
//    float -> bf16, both round to nearest even; NaN -> 0x7fff like __floats2bfloat162_rn), so staging the
//    converted colours changes no result bit while a colour travels as 2 bytes instead of 8;
//  * pcc_host_copy_stream: memcpy whose stores bypass the cache.
// Both write the pinned staging buffer with non-temporal stores where the CPU has AVX2: the DMA engine
// reads that buffer next, and lines left dirty in the CPU caches made those reads slow (1 MB pieces at
// ~8 GB/s instead of ~45 GB/s on the bench box).  An sfence orders the stores before the copy is enqueued.
#include <immintrin.h>

#include <cstdint>
#include <cstring>

namespace {

inline uint16_t bf16_of(double d) {
  const float f = (float)d;
  uint32_t u;
  memcpy(&u, &f, 4);
  const uint32_t rounded = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
  return (uint16_t)(((u & 0x7fffffffu) > 0x7f800000u) ? 0x7fffu : rounded);
}

void convert_base(const double* src, int64_t n, uint16_t* dst) {
  for (int64_t i = 0; i < n; ++i) dst[i] = bf16_of(src[i]);
}

__attribute__((target("avx2"))) inline __m256i bf16x8(__m256i u) {
  const __m256i lsb = _mm256_and_si256(_mm256_srli_epi32(u, 16), _mm256_set1_epi32(1));
  const __m256i rounded = _mm256_srli_epi32(_mm256_add_epi32(_mm256_add_epi32(u, _mm256_set1_epi32(0x7fff)), lsb), 16);
  const __m256i mag = _mm256_and_si256(u, _mm256_set1_epi32(0x7fffffff));
  const __m256i isnan = _mm256_cmpgt_epi32(mag, _mm256_set1_epi32(0x7f800000));
  return _mm256_blendv_epi8(rounded, _mm256_set1_epi32(0x7fff), isnan);
}

__attribute__((target("avx2"))) void convert_avx2(const double* src, int64_t n, uint16_t* dst) {
  int64_t i = 0;
  while (i < n && ((uintptr_t)(dst + i) & 31)) {   // head: up to the first 32-byte boundary of dst
    dst[i] = bf16_of(src[i]);
    ++i;
  }
  for (; i + 16 <= n; i += 16) {
    const __m128 a = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i)), b = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i + 4));
    const __m128 c = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i + 8)), d = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i + 12));
    const __m256i lo = bf16x8(_mm256_castps_si256(_mm256_set_m128(b, a)));
    const __m256i hi = bf16x8(_mm256_castps_si256(_mm256_set_m128(d, c)));
    // packus interleaves the 128-bit lanes: (lo0 hi0 lo1 hi1) -> reorder to (lo0 lo1 hi0 hi1)
    const __m256i packed = _mm256_permute4x64_epi64(_mm256_packus_epi32(lo, hi), 0xD8);
    _mm256_stream_si256((__m256i*)(dst + i), packed);
  }
  for (; i < n; ++i) dst[i] = bf16_of(src[i]);
  _mm_sfence();
}

__attribute__((target("avx2"))) void copy_avx2(const uint8_t* src, int64_t n, uint8_t* dst) {
  int64_t i = 0;
  const int64_t head = (int64_t)((32 - ((uintptr_t)dst & 31)) & 31);
  if (head && head <= n) {
    memcpy(dst, src, (size_t)head);
    i = head;
  }
  for (; i + 128 <= n; i += 128) {
    const __m256i a = _mm256_loadu_si256((const __m256i*)(src + i)), b = _mm256_loadu_si256((const __m256i*)(src + i + 32));
    const __m256i c = _mm256_loadu_si256((const __m256i*)(src + i + 64)), d = _mm256_loadu_si256((const __m256i*)(src + i + 96));
    _mm256_stream_si256((__m256i*)(dst + i), a);
    _mm256_stream_si256((__m256i*)(dst + i + 32), b);
    _mm256_stream_si256((__m256i*)(dst + i + 64), c);
    _mm256_stream_si256((__m256i*)(dst + i + 96), d);
  }
  if (i < n) memcpy(dst + i, src + i, (size_t)(n - i));
  _mm_sfence();
}

bool has_avx2() {
  static const bool v = __builtin_cpu_supports("avx2");
  return v;
}

}  // namespace

extern "C" void pcc_host_f64_to_bf16(const double* src, int64_t n, uint16_t* dst) {
  if (has_avx2())
    convert_avx2(src, n, dst);
  else
    convert_base(src, n, dst);
}

extern "C" void pcc_host_copy_stream(const void* src, int64_t nbytes, void* dst) {
  if (has_avx2() && nbytes >= 4096)
    copy_avx2((const uint8_t*)src, nbytes, (uint8_t*)dst);
  else if (nbytes > 0)
    memcpy(dst, src, (size_t)nbytes);
}
