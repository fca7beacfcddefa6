// Shared device/host helpers for the pcc_b200 kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <atomic>
#include <cstdio>
#include <cstdarg>

#include "../../include/pcc_b200.h"

namespace pcc {

void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;

#define PCC_CHECK_ARG(cond)                                                       \
  do {                                                                            \
    if (!(cond)) {                                                                \
      pcc::set_error("%s:%d: invalid argument: %s", __FILE__, __LINE__, #cond);  \
      return PCC_E_INVALID;                                                       \
    }                                                                             \
  } while (0)

#define PCC_CUDA(expr)                                                                 \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      pcc::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return PCC_E_CUDA;                                                               \
    }                                                                                  \
  } while (0)

// Call right after a <<<>>> launch.
#define PCC_LAUNCHED()                  \
  do {                                  \
    pcc::g_launches.fetch_add(1);       \
    PCC_CUDA(cudaGetLastError());       \
  } while (0)

// Programmatic dependent launch: a kernel launched through launch_pdl may become resident while its
// predecessor in the stream still runs (as soon as every CTA of the predecessor has executed pdl_trigger() or
// exited and resources are free); it must execute pdl_wait() before it touches anything a predecessor wrote --
// the wait returns once the predecessor grid has completed and its writes are visible.  Kernels launched the
// ordinary way are unaffected by a predecessor's trigger.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// Device fault word.  A kernel that meets a condition the host has to hear about (a block table that turns out too
// small, a pipeline barrier that never completes) records a PCC_FAULT_* code here and returns early; nothing in this
// library executes `trap`, which would poison the CUDA context of the whole (long-lived) encoder / decoder process.
// The word lives in device memory of each device (fault.cu); kernels receive its address as an argument.  Hosts read
// and clear it with pcc_fault_fetch (decompress: with the deferred checks; compress: in the header of the packed
// strings).  Returns nullptr after setting the error text when the CUDA runtime fails.
int* fault_word();

constexpr int kMaxDevices = 64;
// current device index (0 when the runtime cannot say; callers only use it to index small per-device caches)
static inline int current_device() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= kMaxDevices) d = 0;
  return d;
}
// "done once per device" marker for function attributes (cudaFuncSetAttribute is per device / context, not per
// process): test-and-set is racy only in the harmless direction (the attribute may be set twice).
struct PerDeviceOnce {
  std::atomic<uint8_t> done[kMaxDevices];
  bool needed(int dev) const { return !done[dev].load(std::memory_order_acquire); }
  void mark(int dev) { done[dev].store(1, std::memory_order_release); }
};
// SM count of the current device (cached per device)
int sm_count();

static inline cudaStream_t as_stream(pcc_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

// ---------------------------------------------------------------------------------------------
// Morton keys.  key = b << 48 | spread3(x+32768) | spread3(y+32768) << 1 | spread3(z+32768) << 2
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t spread3_16(uint32_t v) {
  uint64_t x = v & 0xffffu;
  x = (x | (x << 16)) & 0x001f0000ff0000ffull;
  x = (x | (x << 8)) & 0x100f00f00f00f00full;
  x = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
  x = (x | (x << 2)) & 0x1249249249249249ull;
  return x;
}
__host__ __device__ __forceinline__ uint32_t compact3_16(uint64_t x) {
  x &= 0x1249249249249249ull;
  x = (x | (x >> 2)) & 0x10c30c30c30c30c3ull;
  x = (x | (x >> 4)) & 0x100f00f00f00f00full;
  x = (x | (x >> 8)) & 0x001f0000ff0000ffull;
  x = (x | (x >> 16)) & 0x001f00000000ffffull;
  return (uint32_t)(x & 0xffffu);
}
__host__ __device__ __forceinline__ uint64_t pack_key(int b, int x, int y, int z) {
  uint64_t m = spread3_16((uint32_t)(x + 32768)) | (spread3_16((uint32_t)(y + 32768)) << 1) |
               (spread3_16((uint32_t)(z + 32768)) << 2);
  return ((uint64_t)(uint32_t)(b & 0xffff) << 48) | m;
}
__host__ __device__ __forceinline__ void unpack_key(uint64_t key, int& b, int& x, int& y, int& z) {
  b = (int)(key >> 48);
  uint64_t m = key & 0x0000FFFFFFFFFFFFull;
  x = (int)compact3_16(m) - 32768;
  y = (int)compact3_16(m >> 1) - 32768;
  z = (int)compact3_16(m >> 2) - 32768;
}

constexpr uint64_t kEmptyKey = 0xFFFFFFFFFFFFFFFFull;

__host__ __device__ __forceinline__ uint64_t hash_key(uint64_t k) {
  // murmur3 finaliser
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return k;
}

}  // namespace pcc
