// (8) Gathering the byte strings of a GOP (y and z lane containers, per-frame coordinate streams) into ONE
// device buffer, so the host fetches them with a single copy and a single wait.  The strings live in
// separate buffers with their lengths in device counters (nothing on the host knows them yet); the reference
// moves every string to the host on its own (sender/encoder/codec_pipeline.py:305-306, 426-432, 441-462).
#include "common.cuh"

namespace pcc {

constexpr int kPackBatch = 64;   // strings per launch (pointer tables travel as kernel arguments)

struct PackArgs {
  const uint8_t* buf[kPackBatch];
  const int32_t* len[kPackBatch];
};

__host__ __device__ inline uint32_t align16(uint32_t v) { return (v + 15u) & ~15u; }

// grid (blocks per string, strings of this batch).  Header: u32 len[n_total] | u32 end[n_batches]; batch b
// starts where batch b - 1 ended (end[b - 1], written by the previous launch on the same stream).
__global__ void __launch_bounds__(256)
k_pack_strings(PackArgs a, int n, int first, int batch, uint32_t payload0, uint8_t* __restrict__ out, int64_t out_cap,
               int n_total, int n_batches, int* __restrict__ fault) {
  pdl_trigger();
  pdl_wait();
  uint32_t* hdr = reinterpret_cast<uint32_t*>(out);
  const int sidx = blockIdx.y;
  uint32_t off = batch == 0 ? payload0 : hdr[n_total + batch - 1];
  for (int i = 0; i < sidx; ++i) off += align16((uint32_t)max(*a.len[i], 0));
  const uint32_t len = (uint32_t)max(*a.len[sidx], 0);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    hdr[first + sidx] = len;
    if (sidx == n - 1) {
      hdr[n_total + batch] = off + align16(len);
      // last batch: hand the device fault word (common.cuh) to the host with the strings, and clear it
      if (batch == n_batches - 1) hdr[n_total + n_batches] = (uint32_t)atomicExch(fault, 0);
    }
  }
  if ((int64_t)off + len > out_cap) return;      // the host sees end > capacity and reports it
  const uint8_t* src = a.buf[sidx];
  uint8_t* dst = out + off;
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const uint32_t n16 = len >> 4;
    for (uint32_t i = tid; i < n16; i += nth) reinterpret_cast<uint4*>(dst)[i] = __ldg(reinterpret_cast<const uint4*>(src) + i);
    for (uint32_t i = (n16 << 4) + tid; i < len; i += nth) dst[i] = src[i];
  } else {
    for (uint32_t i = tid; i < len; i += nth) dst[i] = src[i];
  }
}

}  // namespace pcc

using namespace pcc;

extern "C" {

int64_t pcc_pack_strings_header_bytes(int32_t n) {
  if (n < 0) return -1;
  const int64_t nb = (n + kPackBatch - 1) / kPackBatch;
  return (int64_t)align16((uint32_t)(4 * (n + nb + 1)));      // lengths, batch ends, fault word
}

int pcc_pack_strings(const void* const* bufs, const int32_t* const* lens, int32_t n, uint8_t* out, int64_t out_cap,
                     pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n <= 65536 && out && ((uintptr_t)out % 16) == 0);
  PCC_CHECK_ARG(n == 0 || (bufs && lens));
  const int64_t hdr = pcc_pack_strings_header_bytes(n);
  PCC_CHECK_ARG(out_cap >= hdr && out_cap < (1ll << 32));
  int* fault = fault_word();
  if (!fault) return PCC_E_CUDA;
  const int n_batches = (n + kPackBatch - 1) / kPackBatch;
  for (int first = 0, batch = 0; first < n; first += kPackBatch, ++batch) {
    PackArgs a;
    const int cnt = n - first < kPackBatch ? n - first : kPackBatch;
    for (int i = 0; i < cnt; ++i) {
      PCC_CHECK_ARG(bufs[first + i] && lens[first + i]);
      a.buf[i] = (const uint8_t*)bufs[first + i];
      a.len[i] = lens[first + i];
    }
    for (int i = cnt; i < kPackBatch; ++i) {
      a.buf[i] = nullptr;
      a.len[i] = nullptr;
    }
    launch_pdl(k_pack_strings, dim3(dim3(8, cnt)), dim3(256), 0, as_stream(s), a, cnt, first, batch, (uint32_t)hdr, out, out_cap, n, n_batches, fault);
    PCC_LAUNCHED();
  }
  return PCC_OK;
}

}  // extern "C"
