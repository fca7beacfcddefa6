// Per-frame top-k occupancy selection (torch.topk + MinkowskiPruning of g_s).
// Deterministic rule: higher logit first; equal logits -> lower row (= lower key).  Implemented as one
// stable radix sort of (frame, ~orderable(logit)) with the row as payload, a rank test against
// k[frame], and an order-preserving compaction of the kept rows.
#include "common.cuh"
#include <cub/cub.cuh>

namespace pcc {

__device__ __forceinline__ uint32_t orderable(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending in float order
}

__global__ void k_topk_keys(const float* __restrict__ logits, const uint64_t* __restrict__ keys, int64_t n,
                            uint64_t* __restrict__ skey, int32_t* __restrict__ row) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t b = keys[i] >> 48;
  float l = logits[i];
  if (l != l) l = -INFINITY;  // NaN ranks last
  skey[i] = (b << 32) | (uint64_t)(~orderable(l));
  row[i] = (int32_t)i;
}

__global__ void k_topk_flags(const uint64_t* __restrict__ skey_sorted, const int32_t* __restrict__ row_sorted,
                             int64_t n, const int32_t* __restrict__ frame_off, const int32_t* __restrict__ k,
                             int nf, uint8_t* __restrict__ flags) {
  int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= n) return;
  int b = (int)(skey_sorted[p] >> 32);
  int keep = 0;
  if (b < nf) {
    int64_t rank = p - frame_off[b];
    keep = rank < (int64_t)k[b];
  }
  flags[row_sorted[p]] = (uint8_t)keep;
}

}  // namespace pcc

using namespace pcc;

static size_t sort_bytes(int64_t n) {
  size_t b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, b, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, (int)n, 0, 48);
  return b;
}
static size_t select_bytes(int64_t n) {
  size_t b = 0;
  cub::DeviceSelect::Flagged(nullptr, b, cub::CountingInputIterator<int32_t>(0), (const uint8_t*)nullptr,
                             (int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  return b;
}

extern "C" {

int64_t pcc_topk_workspace_bytes(int64_t n) {
  if (n <= 0) return 256;
  return 2 * align_up(n * 8, 256) + 2 * align_up(n * 4, 256) + align_up(n, 256) +
         align_up((int64_t)sort_bytes(n), 256) + align_up((int64_t)select_bytes(n), 256) + 256;
}

int pcc_topk_select(const float* logits, const uint64_t* keys, int64_t n,
                    const int32_t* frame_offsets, const int32_t* k, int32_t num_frames,
                    int32_t* keep_row, int32_t* n_keep_dev, void* ws, int64_t ws_bytes,
                    pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n < (1ll << 31) && n_keep_dev && num_frames >= 0);
  cudaStream_t st = as_stream(s);
  if (n == 0) {
    PCC_CUDA(cudaMemsetAsync(n_keep_dev, 0, 4, st));
    return PCC_OK;
  }
  PCC_CHECK_ARG(logits && keys && frame_offsets && k && keep_row && ws);
  char* p = (char*)ws;
  int64_t left = ws_bytes;
  auto take = [&](int64_t b) -> void* {
    b = align_up(b, 256);
    if (b > left) return nullptr;
    void* r = p;
    p += b;
    left -= b;
    return r;
  };
  uint64_t* skey = (uint64_t*)take(n * 8);
  uint64_t* skey2 = (uint64_t*)take(n * 8);
  int32_t* row = (int32_t*)take(n * 4);
  int32_t* row2 = (int32_t*)take(n * 4);
  uint8_t* flags = (uint8_t*)take(n);
  size_t sb = sort_bytes(n), fb = select_bytes(n);
  void* sws = take((int64_t)sb);
  void* fws = take((int64_t)fb);
  if (!skey || !skey2 || !row || !row2 || !flags || !sws || !fws) {
    set_error("pcc_topk_select: workspace too small");
    return PCC_E_WORKSPACE;
  }
  unsigned grid = (unsigned)((n + 255) / 256);
  k_topk_keys<<<grid, 256, 0, st>>>(logits, keys, n, skey, row);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceRadixSort::SortPairs(sws, sb, skey, skey2, row, row2, (int)n, 0, 48, st));
  g_launches.fetch_add(1);
  k_topk_flags<<<grid, 256, 0, st>>>(skey2, row2, n, frame_offsets, k, num_frames, flags);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceSelect::Flagged(fws, fb, cub::CountingInputIterator<int32_t>(0), flags, keep_row,
                                      n_keep_dev, (int)n, st));
  g_launches.fetch_add(1);
  return PCC_OK;
}

}  // extern "C"
