// Device fault word + small per-device caches shared by the launchers (see common.cuh).
#include "common.cuh"
#include <cstdlib>

namespace pcc {

__device__ int g_fault_word = 0;

int* fault_word() {
  static std::atomic<int*> cache[kMaxDevices];
  const int dev = current_device();
  int* p = cache[dev].load(std::memory_order_acquire);
  if (p) return p;
  void* addr = nullptr;
  const cudaError_t e = cudaGetSymbolAddress(&addr, g_fault_word);
  if (e != cudaSuccess) {
    set_error("fault_word: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
    return nullptr;
  }
  cache[dev].store((int*)addr, std::memory_order_release);
  return (int*)addr;
}

int sm_count() {
  static std::atomic<int> cache[kMaxDevices];
  const int dev = current_device();
  int n = cache[dev].load(std::memory_order_acquire);
  if (n > 0) return n;
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, d) != cudaSuccess || n <= 0)
    n = 148;
  cache[dev].store(n, std::memory_order_release);
  return n;
}

constexpr int kFetchMax = 48;
struct FetchArgs {
  const int32_t* src[kFetchMax];
};
// dst[i] = *src[i] for i < n, dst[n] = the device fault word (fetched and cleared) when with_fault
__global__ void k_fetch_words(FetchArgs a, int n, int* __restrict__ fault, int with_fault, int32_t* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  const int i = threadIdx.x;
  if (i < n) dst[i] = *a.src[i];
  if (i == n && with_fault) dst[n] = atomicExch(fault, 0);
}

__global__ void k_fault_fetch(int* __restrict__ word, int32_t* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  *dst = atomicExch(word, 0);
}

}  // namespace pcc

using namespace pcc;

extern "C" {

int pcc_fault_fetch(int32_t* dst_dev, pcc_stream_t s) {
  PCC_CHECK_ARG(dst_dev);
  int* w = fault_word();
  if (!w) return PCC_E_CUDA;
  launch_pdl(k_fault_fetch, dim3(1), dim3(1), 0, as_stream(s), w, dst_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_fetch_words(const int32_t* const* srcs_dev, int32_t n, int32_t with_fault, int32_t* dst_dev, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n < kFetchMax && dst_dev && (srcs_dev || n == 0));
  int* w = fault_word();
  if (!w) return PCC_E_CUDA;
  FetchArgs a;
  for (int i = 0; i < n; ++i) {
    PCC_CHECK_ARG(srcs_dev[i] != nullptr);
    a.src[i] = srcs_dev[i];
  }
  for (int i = n; i < kFetchMax; ++i) a.src[i] = nullptr;
  launch_pdl(k_fetch_words, dim3(1), dim3(64), 0, as_stream(s), a, (int)n, w, (int)(with_fault != 0), dst_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

const char* pcc_fault_string(int32_t code) {
  switch (code) {
    case 0: return "no fault";
    case PCC_FAULT_BLOCKMAP_FULL: return "block table too small for the coordinate set (pcc_blockmap_build capacity)";
    case PCC_FAULT_PIPELINE_TIMEOUT: return "sparse convolution: a pipeline barrier did not complete (kernel gave up and left its tiles unwritten)";
    default: return "unknown device fault";
  }
}

}  // extern "C"
