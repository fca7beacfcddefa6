// Per-frame top-k occupancy selection (torch.topk + MinkowskiPruning of g_s).
// Deterministic rule: higher logit first; equal logits -> lower row (= lower key).
// Radix select instead of a sort: four 8-bit passes over the order-preserving integer image of the
// logits find, per frame, the value T of the k-th largest logit and the number r of rows equal to T
// that still belong to the top k; a row is kept iff logit > T, or logit == T and fewer than r equal
// rows precede it in its frame.  One 64-bit inclusive scan of (above | tied << 32) marks counts both kinds of
// rows, so every kept row knows its output position and is written in order without a separate compaction.
// Every pass reads 4 bytes per candidate; nothing is sorted.
#include "common.cuh"
#include <cub/cub.cuh>

namespace pcc {

__device__ __forceinline__ uint32_t orderable(float f) {
  if (f != f) f = -INFINITY;  // NaN ranks last
  f = f + 0.0f;               // -0 == +0 (as in torch.topk)
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending in float order
}

constexpr int kSelThreads = 256;
constexpr int kSelPerThread = 8;
constexpr int kSelChunk = kSelThreads * kSelPerThread;

// Per-frame selection state: {prefix (threshold bits found so far), rows still to take, blocks done}
struct SelState {
  uint32_t prefix;
  int32_t kleft;
  uint32_t done;
  uint32_t pad;
};

__global__ void k_sel_init(const int32_t* __restrict__ k, int nf, SelState* __restrict__ st, uint32_t* __restrict__ hist) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  if (threadIdx.x == 0) {
    st[b].prefix = 0;
    st[b].kleft = k[b] > 0 ? k[b] : 0;
    st[b].done = 0;
  }
  for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[b * 256 + i] = 0;
}

// grid (chunks, frames): histogram of digit `shift` among the frame's rows whose higher bits equal the
// prefix; the last block of a frame to finish picks the digit holding the k-th largest row, updates
// the state and clears the histogram for the next pass.
__global__ void __launch_bounds__(kSelThreads)
k_sel_pass(const float* __restrict__ logits, const int32_t* __restrict__ frame_off, int shift,
           SelState* __restrict__ st, uint32_t* __restrict__ hist) {
  pdl_trigger();
  pdl_wait();
  __shared__ uint32_t sh[256];
  __shared__ bool last;
  const int b = blockIdx.y;
  const int r0 = frame_off[b], r1 = frame_off[b + 1];
  const int nblk = (r1 - r0 + kSelChunk - 1) / kSelChunk;
  if ((int)blockIdx.x >= nblk) return;
  sh[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t prefix = st[b].prefix;
  const int hs = shift + 8;
  const int base = r0 + blockIdx.x * kSelChunk;
#pragma unroll
  for (int e = 0; e < kSelPerThread; ++e) {
    const int i = base + e * kSelThreads + threadIdx.x;
    uint32_t digit = 256u;   // not counted
    if (i < r1) {
      const uint32_t u = orderable(__ldg(logits + i));
      if (hs >= 32 || (u >> hs) == (prefix >> hs)) digit = (u >> shift) & 255u;
    }
    // logits share their high digits: aggregate equal digits inside the warp before the shared atomic
    const unsigned peers = __match_any_sync(0xffffffffu, digit);
    if (digit < 256u && (threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&sh[digit], (uint32_t)__popc(peers));
  }
  __syncthreads();
  if (sh[threadIdx.x]) atomicAdd(&hist[b * 256 + threadIdx.x], sh[threadIdx.x]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(&st[b].done, 1u) == (uint32_t)(nblk - 1);
  __syncthreads();
  if (!last) return;
  __threadfence();
  sh[threadIdx.x] = __ldcg(&hist[b * 256 + threadIdx.x]);
  hist[b * 256 + threadIdx.x] = 0;
  __syncthreads();
  // Pick the digit: walking down from 255, the first d > 0 with (rows in digits above d) + sh[d] >= kleft, else 0.
  // Warp 0 does the walk in parallel: lane L owns digits 255 - 8 L ... 248 - 8 L, a shuffle scan gives the rows above
  // its range, the lowest lane that finds a hit wins.
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    const int32_t kleft = st[b].kleft;
    const int top = 255 - 8 * lane;
    uint32_t loc[8];
    uint32_t sum = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      loc[e] = sh[top - e];
      sum += loc[e];
    }
    uint32_t incl = sum;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += v;
    }
    uint32_t cum = incl - sum;      // rows in the digits above this lane's range
    int d = -1;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      if (d < 0) {
        if (top - e > 0 && (int64_t)cum + loc[e] >= (int64_t)kleft) d = top - e;
        else if (top - e > 0) cum += loc[e];
      }
    }
    const unsigned hit = __ballot_sync(0xffffffffu, d >= 0);
    // no hit: d = 0 (also catches k larger than the frame: everything down to the smallest digit is taken) and
    // cum = rows in digits 255 ... 1, which is what lane 31 has accumulated
    const int win = hit ? __ffs(hit) - 1 : 31;
    if (lane == win) {
      st[b].prefix = prefix | ((uint32_t)(d >= 0 ? d : 0) << shift);
      st[b].kleft = kleft - (int32_t)cum;
      st[b].done = 0;
    }
  }
}

// mark[i] = (logit above the frame's threshold) | (logit equal to it) << 32: one 64-bit inclusive scan then ranks
// both kinds of rows at once
__global__ void __launch_bounds__(kSelThreads)
k_sel_mark(const float* __restrict__ logits, const int32_t* __restrict__ frame_off, const SelState* __restrict__ st,
           uint64_t* __restrict__ mark) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.y;
  const int r0 = frame_off[b], r1 = frame_off[b + 1];
  const uint32_t T = st[b].prefix;
  const int base = r0 + blockIdx.x * kSelChunk;
  if (base >= r1) return;
#pragma unroll
  for (int e = 0; e < kSelPerThread; ++e) {
    const int i = base + e * kSelThreads + threadIdx.x;
    if (i < r1) {
      const uint32_t u = orderable(__ldg(logits + i));
      mark[i] = u > T ? 1ull : (u == T ? (1ull << 32) : 0ull);
    }
  }
}

// Kept rows of frame b: every row above the threshold and the first kleft rows equal to it.  A kept row lands
// behind the kept rows of the earlier frames, the rows above the threshold that precede it in its frame and the
// (at most kleft) tied rows that precede it.  Block (0, last frame) also writes the total.
__global__ void __launch_bounds__(kSelThreads)
k_sel_emit(const int32_t* __restrict__ frame_off, const SelState* __restrict__ st, const uint64_t* __restrict__ scan,
           int nf, int32_t* __restrict__ keep_row, int32_t* __restrict__ n_keep) {
  pdl_trigger();
  pdl_wait();
  __shared__ int64_t s_before;     // kept rows of the frames before b
  const int b = blockIdx.y;
  const int r0 = frame_off[b], r1 = frame_off[b + 1];
  const int base = r0 + blockIdx.x * kSelChunk;
  const bool totals = blockIdx.x == 0 && b == nf - 1;
  if (base >= r1 && !totals) return;
  if (threadIdx.x == 0) {
    int64_t before = 0, total = 0;
    const int upto = totals ? nf : b;
    for (int f = 0; f < upto; ++f) {
      const int f0 = frame_off[f], f1 = frame_off[f + 1];
      if (f1 <= f0) continue;
      const uint64_t lo = f0 ? scan[f0 - 1] : 0ull, hi = scan[f1 - 1];
      const int64_t above = (int64_t)((hi & 0xffffffffu) - (lo & 0xffffffffu));
      const int64_t eq = (int64_t)((hi >> 32) - (lo >> 32));
      const int64_t r = st[f].kleft;
      const int64_t kept = above + (eq < r ? eq : r);
      if (f < b) before += kept;
      total += kept;
    }
    if (totals) *n_keep = (int32_t)total;
    s_before = before;
  }
  __syncthreads();
  if (base >= r1) return;
  const int64_t before = s_before;
  const uint64_t s0 = r0 ? scan[r0 - 1] : 0ull;
  const int64_t a0 = (int64_t)(s0 & 0xffffffffu), e0 = (int64_t)(s0 >> 32);
  const int64_t r = st[b].kleft;
#pragma unroll
  for (int e = 0; e < kSelPerThread; ++e) {
    const int i = base + e * kSelThreads + threadIdx.x;
    if (i < r1) {
      const uint64_t incl = scan[i], prev = i ? scan[i - 1] : 0ull;
      const int64_t a_prev = (int64_t)(prev & 0xffffffffu) - a0, e_prev = (int64_t)(prev >> 32) - e0;
      const bool is_above = (incl & 0xffffffffu) != (prev & 0xffffffffu);
      const bool is_eq = (incl >> 32) != (prev >> 32);
      if (is_above || (is_eq && e_prev < r)) keep_row[before + a_prev + (e_prev < r ? e_prev : r)] = i;
    }
  }
}

}  // namespace pcc

using namespace pcc;

constexpr int64_t kMaxFrames = 4096;   // frames per call (a GOP holds a handful)

static size_t scan_bytes(int64_t n) {
  size_t b = 0;
  cub::DeviceScan::InclusiveSum(nullptr, b, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int)n);
  return b;
}

extern "C" {

int64_t pcc_topk_workspace_bytes(int64_t n) {
  if (n <= 0) return 256;
  return align_up(n * 8, 256) + align_up((int64_t)scan_bytes(n), 256) + align_up(kMaxFrames * (16 + 1024), 256) + 2 * 256;
}

int pcc_topk_select(const float* logits, const uint64_t* keys, int64_t n,
                    const int32_t* frame_offsets, const int32_t* k, int32_t num_frames,
                    int32_t* keep_row, int32_t* n_keep_dev, void* ws, int64_t ws_bytes,
                    pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n < (1ll << 31) && n_keep_dev && num_frames >= 0 && num_frames <= kMaxFrames);
  cudaStream_t st = as_stream(s);
  if (n == 0 || num_frames == 0) {
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
  uint64_t* mark = (uint64_t*)take(n * 8);
  size_t sb = scan_bytes(n);
  void* sws = take((int64_t)sb);
  SelState* state = (SelState*)take((int64_t)num_frames * sizeof(SelState));
  uint32_t* hist = (uint32_t*)take((int64_t)num_frames * 1024);
  if (!mark || !sws || !state || !hist) {
    set_error("pcc_topk_select: workspace too small");
    return PCC_E_WORKSPACE;
  }
  // rows of frames >= num_frames (none in a well-formed batch) are dropped
  PCC_CUDA(cudaMemsetAsync(mark, 0, (size_t)n * 8, st));
  const dim3 grid((unsigned)((n + kSelChunk - 1) / kSelChunk), (unsigned)num_frames);
  launch_pdl(k_sel_init, dim3((unsigned)num_frames), dim3(256), 0, st, k, num_frames, state, hist);
  PCC_LAUNCHED();
  for (int shift = 24; shift >= 0; shift -= 8) {
    launch_pdl(k_sel_pass, dim3(grid), dim3(kSelThreads), 0, st, logits, frame_offsets, shift, state, hist);
    PCC_LAUNCHED();
  }
  launch_pdl(k_sel_mark, dim3(grid), dim3(kSelThreads), 0, st, logits, frame_offsets, state, mark);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(sws, sb, mark, mark, (int)n, st));
  g_launches.fetch_add(1);
  launch_pdl(k_sel_emit, dim3(grid), dim3(kSelThreads), 0, st, frame_offsets, state, mark, num_frames, keep_row, n_keep_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
