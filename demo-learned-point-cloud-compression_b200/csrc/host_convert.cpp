// Host-side float64 -> bf16 conversion used while staging a raw frame's colours (stage.cu): the same two
// roundings the device path applies (double -> float, float -> bf16, both round to nearest even; NaN ->
// 0x7fff like __floats2bfloat162_rn), so staging the converted values changes no result bit while a
// colour travels as 2 bytes instead of 8.  Plain C++ (g++); the flat loop vectorises, and an AVX2 build of
// it is picked at run time where the CPU has it.
#include <cstdint>
#include <cstring>

namespace {

inline __attribute__((always_inline)) void convert(const double* src, int64_t n, uint16_t* dst) {
  for (int64_t i = 0; i < n; ++i) {
    const float f = (float)src[i];
    uint32_t u;
    memcpy(&u, &f, 4);
    const uint32_t rounded = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
    dst[i] = (uint16_t)(((u & 0x7fffffffu) > 0x7f800000u) ? 0x7fffu : rounded);
  }
}

__attribute__((target("avx2"))) void convert_avx2(const double* src, int64_t n, uint16_t* dst) { convert(src, n, dst); }
void convert_base(const double* src, int64_t n, uint16_t* dst) { convert(src, n, dst); }

}  // namespace

extern "C" void pcc_host_f64_to_bf16(const double* src, int64_t n, uint16_t* dst) {
  static const bool avx2 = __builtin_cpu_supports("avx2");
  if (avx2)
    convert_avx2(src, n, dst);
  else
    convert_base(src, n, dst);
}
