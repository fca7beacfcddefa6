// (1b) Lossless octree coder for the latent coordinates ("OCT2" streams), on the device.
//
// Fills the slot the reference fills with a G-PCC `tmc3` subprocess and temporary PLY files
// (shared/utils.py:169-240, codec_pipeline.py:441-462, codec_parallel.py:266-289).
// A frame's coordinates are a sorted run of Morton keys, so the breadth-first octree falls out of
// prefix comparisons: with m = key48 >> 3*log2_stride and depth = ceil(bits(m_first ^ m_last) / 3),
// the nodes of level l are the distinct values of m >> 3*(depth - l), and the occupancy byte of a node
// is the OR of (1 << child octant) over its children.  No sort, no hash, no host round trip.
//
// Stream of one frame (little endian):
//   "OCT2" | u32 n | u32 depth | u32 root_lo | u32 root_hi | u32 nodes[depth] | level 0 bytes | ... |
// root = m >> 3*depth (the prefix all points share), nodes[l] = number of occupancy bytes of level l
// (nodes[0] = 1).  Children are emitted in ascending octant order, so decoding yields the keys in
// ascending order again: the decoder needs neither a sort nor a de-duplication.
#include "common.cuh"

#include <cub/block/block_scan.cuh>

namespace pcc {

constexpr int kOctMaxDepth = 16;      // 48 Morton bits / 3
constexpr int kOctEncThreads = 512;
constexpr int kOctDecThreads = 1024;
constexpr uint32_t kOctMagic = 0x3254434Fu;   // "OCT2"
constexpr uint64_t kMortonMask = 0x0000FFFFFFFFFFFFull;

__device__ __forceinline__ int oct_depth(uint64_t first, uint64_t last) {
  const uint64_t x = first ^ last;
  const int bits = x ? 64 - __clzll((long long)x) : 0;
  return (bits + 2) / 3;
}

// grid (kOctMaxDepth, frames).  CTA (l, b) walks the frame once and produces the occupancy bytes of
// level l into scratch[l * pitch + frame_off[b] ...] (zeroed by the caller) plus the node count.
__global__ void __launch_bounds__(kOctEncThreads)
k_octree_levels(const uint64_t* __restrict__ keys, const int32_t* __restrict__ frame_off, int shift0,
                uint32_t* __restrict__ scratch, int64_t pitch_words, int32_t* __restrict__ counts) {
  pdl_trigger();
  pdl_wait();
  using Scan = cub::BlockScan<int, kOctEncThreads>;
  __shared__ typename Scan::TempStorage tmp;
  const int l = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const int r0 = frame_off[b], r1 = frame_off[b + 1];
  const int n = r1 - r0;
  if (n <= 0) return;
  const uint64_t first = (keys[r0] & kMortonMask) >> shift0, last = (keys[r1 - 1] & kMortonMask) >> shift0;
  const int depth = oct_depth(first, last);
  if (l >= depth) return;
  const int sp = 3 * (depth - l), sc = sp - 3;
  uint8_t* out = reinterpret_cast<uint8_t*>(scratch + (int64_t)l * pitch_words);
  int carry = 0;
  for (int base = 0; base < n; base += kOctEncThreads) {
    const int i = base + tid;
    const bool valid = i < n;
    uint64_t m = 0, mp = 0;
    if (valid) {
      m = (keys[r0 + i] & kMortonMask) >> shift0;
      mp = i > 0 ? (keys[r0 + i - 1] & kMortonMask) >> shift0 : ~m;
    }
    const int fp = valid && ((m >> sp) != (mp >> sp) || i == 0);
    const bool fc = valid && ((m >> sc) != (mp >> sc) || i == 0);
    int incl, total;
    Scan(tmp).InclusiveSum(fp, incl, total);
    if (fc) {
      const int64_t byte = (int64_t)r0 + carry + incl - 1;      // node index within the level, frame-relative + r0
      const uint32_t bit = 1u << (uint32_t)((m >> sc) & 7u);
      atomicOr(reinterpret_cast<uint32_t*>(out) + (byte >> 2), bit << (8 * (int)(byte & 3)));
    }
    carry += total;
    __syncthreads();
  }
  if (tid == 0) counts[b * kOctMaxDepth + l] = carry;
}

// grid (kOctMaxDepth, frames).  CTA (l, b) copies level l to its place in the frame's stream; CTA
// (0, b) also writes the header and the stream size.
__global__ void __launch_bounds__(256)
k_octree_pack(const uint64_t* __restrict__ keys, const int32_t* __restrict__ frame_off, int shift0,
              const uint32_t* __restrict__ scratch, int64_t pitch_words, const int32_t* __restrict__ counts,
              uint8_t* __restrict__ out, const int64_t* __restrict__ out_off, int32_t* __restrict__ sizes) {
  pdl_trigger();
  pdl_wait();
  const int l = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const int r0 = frame_off[b], r1 = frame_off[b + 1];
  const int n = r1 - r0 > 0 ? r1 - r0 : 0;
  uint64_t first = 0, last = 0;
  if (n > 0) {
    first = (keys[r0] & kMortonMask) >> shift0;
    last = (keys[r1 - 1] & kMortonMask) >> shift0;
  }
  const int depth = n > 0 ? oct_depth(first, last) : 0;
  uint8_t* dst = out + out_off[b];           // out_off is a multiple of 4
  const int hdr = 4 * (5 + depth);
  if (l == 0 && tid == 0) {
    uint32_t* h = reinterpret_cast<uint32_t*>(dst);
    const uint64_t root = depth >= kOctMaxDepth ? 0ull : first >> (3 * depth);
    h[0] = kOctMagic;
    h[1] = (uint32_t)n;
    h[2] = (uint32_t)depth;
    h[3] = (uint32_t)(root & 0xffffffffu);
    h[4] = (uint32_t)(root >> 32);
    int total = hdr;
    for (int k = 0; k < depth; ++k) {
      h[5 + k] = (uint32_t)counts[b * kOctMaxDepth + k];
      total += counts[b * kOctMaxDepth + k];
    }
    sizes[b] = total;
  }
  if (l >= depth) return;
  int pos = hdr;
  for (int k = 0; k < l; ++k) pos += counts[b * kOctMaxDepth + k];
  const int cnt = counts[b * kOctMaxDepth + l];
  const uint8_t* src = reinterpret_cast<const uint8_t*>(scratch + (int64_t)l * pitch_words) + r0;
  for (int i = tid; i < cnt; i += blockDim.x) dst[pos + i] = src[i];
}

// One CTA per frame: expand level by level (popcount -> block scan -> children), ping-ponging the
// node lists between two scratch rows, then emit keys.  status: bit 0 malformed header / truncated,
// bit 1 inconsistent occupancy bytes.
__global__ void __launch_bounds__(kOctDecThreads)
k_octree_decode(const uint8_t* __restrict__ streams, const int64_t* __restrict__ stream_off,
                const int32_t* __restrict__ stream_len, const int32_t* __restrict__ row_off, int shift0,
                uint64_t* __restrict__ nodes_a, uint64_t* __restrict__ nodes_b, uint64_t* __restrict__ keys_out,
                int32_t* __restrict__ status) {
  pdl_trigger();
  pdl_wait();
  using Scan = cub::BlockScan<int, kOctDecThreads>;
  __shared__ typename Scan::TempStorage tmp;
  __shared__ uint32_t hdr_s[5 + kOctMaxDepth];
  __shared__ int fail_s;
  const int b = blockIdx.x, tid = threadIdx.x;
  const uint8_t* st = streams + stream_off[b];   // multiple of 4
  const int len = stream_len[b];
  const int r0 = row_off[b], n = row_off[b + 1] - r0;
  if (tid == 0) fail_s = 0;
  if (tid < 5 + kOctMaxDepth) {
    uint32_t v = 0;
    if (4 * (tid + 1) <= len) v = reinterpret_cast<const uint32_t*>(st)[tid];
    hdr_s[tid] = v;
  }
  __syncthreads();
  const int depth = (int)hdr_s[2];
  bool bad = len < 20 || hdr_s[0] != kOctMagic || (int)hdr_s[1] != n || depth < 0 || depth > kOctMaxDepth ||
             len < 4 * (5 + depth) || (n == 0 && depth != 0) || (n > 1 && depth == 0);
  if (!bad) {
    int64_t total = 4 * (5 + depth);
    for (int k = 0; k < depth; ++k) total += hdr_s[5 + k];
    bad = total > len || (depth > 0 && hdr_s[5] != 1u);
  }
  if (bad) {
    if (tid == 0) atomicOr(status, 1);
    for (int i = tid; i < n; i += kOctDecThreads) keys_out[r0 + i] = ((uint64_t)b << 48);
    return;
  }
  if (n == 0) return;
  uint64_t* cur = nodes_a + r0;
  uint64_t* nxt = nodes_b + r0;
  if (tid == 0) cur[0] = (uint64_t)hdr_s[3] | ((uint64_t)hdr_s[4] << 32);
  __syncthreads();
  int ncur = 1;
  int pos = 4 * (5 + depth);
  for (int l = 0; l < depth; ++l) {
    const int cnt = (int)hdr_s[5 + l];
    if (cnt != ncur) {
      if (tid == 0) fail_s = 1;
      __syncthreads();
      break;
    }
    int carry = 0;
    for (int base = 0; base < cnt; base += kOctDecThreads) {
      const int i = base + tid;
      const uint32_t occ = i < cnt ? st[pos + i] : 0u;
      const int pc = __popc(occ);
      if (i < cnt && pc == 0) fail_s = 1;
      int ex, total;
      Scan(tmp).ExclusiveSum(pc, ex, total);
      if (carry + total > n) {
        fail_s = 1;                  // uniform: every thread sees the same totals
      } else if (pc) {
        const uint64_t node = cur[i] << 3;
        int o = carry + ex;
        uint32_t m = occ;
        while (m) {
          const int c = __ffs(m) - 1;
          m &= m - 1;
          nxt[o++] = node | (uint64_t)c;
        }
      }
      carry += total;
      __syncthreads();
      if (fail_s) break;
    }
    if (fail_s) break;
    pos += cnt;
    ncur = carry;
    uint64_t* t = cur;
    cur = nxt;
    nxt = t;
  }
  __syncthreads();
  if (fail_s || ncur != n) {
    if (tid == 0) atomicOr(status, 2);
    for (int i = tid; i < n; i += kOctDecThreads) keys_out[r0 + i] = ((uint64_t)b << 48);
    return;
  }
  for (int i = tid; i < n; i += kOctDecThreads)
    keys_out[r0 + i] = ((uint64_t)b << 48) | ((cur[i] << shift0) & kMortonMask);
}

}  // namespace pcc

using namespace pcc;

extern "C" {

int64_t pcc_octree_max_bytes(int64_t n) {
  // header + at most one occupancy byte per point per level, padded to 4
  return align_up(4 * (5 + kOctMaxDepth) + (int64_t)kOctMaxDepth * (n > 0 ? n : 0), 4);
}

int64_t pcc_octree_encode_workspace_bytes(int64_t n_total, int32_t num_frames) {
  return (int64_t)kOctMaxDepth * align_up(n_total + 4, 4) + align_up((int64_t)num_frames * kOctMaxDepth * 4, 16) + 64;
}

int pcc_octree_encode(const uint64_t* keys, int64_t n_total, const int32_t* frame_offsets_dev, int32_t num_frames,
                      int log2_stride, uint8_t* out, const int64_t* out_offsets_dev, int32_t* sizes_dev, void* ws,
                      int64_t ws_bytes, pcc_stream_t s) {
  PCC_CHECK_ARG(n_total >= 0 && n_total < (1ll << 31) && num_frames >= 0 && log2_stride >= 0 && log2_stride <= 15);
  if (num_frames == 0) return PCC_OK;
  PCC_CHECK_ARG(frame_offsets_dev && out && out_offsets_dev && sizes_dev && (keys || n_total == 0));
  if (ws_bytes < pcc_octree_encode_workspace_bytes(n_total, num_frames)) {
    set_error("pcc_octree_encode: workspace too small");
    return PCC_E_WORKSPACE;
  }
  cudaStream_t st = as_stream(s);
  const int64_t pitch = align_up(n_total + 4, 4);
  uint32_t* scratch = (uint32_t*)ws;
  int32_t* counts = (int32_t*)((uint8_t*)ws + (int64_t)kOctMaxDepth * pitch);
  PCC_CUDA(cudaMemsetAsync(ws, 0, (size_t)pcc_octree_encode_workspace_bytes(n_total, num_frames) - 64, st));
  dim3 grid(kOctMaxDepth, (unsigned)num_frames);
  if (n_total > 0) {
    launch_pdl(k_octree_levels, dim3(grid), dim3(kOctEncThreads), 0, st, keys, frame_offsets_dev, 3 * log2_stride, scratch, pitch / 4,
                                                    counts);
    PCC_LAUNCHED();
  }
  launch_pdl(k_octree_pack, dim3(grid), dim3(256), 0, st, keys, frame_offsets_dev, 3 * log2_stride, scratch, pitch / 4, counts, out,
                                      out_offsets_dev, sizes_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

int64_t pcc_octree_decode_workspace_bytes(int64_t n_total) { return 2 * 8 * (n_total > 0 ? n_total : 1); }

int pcc_octree_decode(const uint8_t* streams, const int64_t* stream_offsets_dev, const int32_t* stream_bytes_dev,
                      const int32_t* row_offsets_dev, int32_t num_frames, int64_t n_total, int log2_stride,
                      uint64_t* keys_out, int32_t* status_dev, void* ws, int64_t ws_bytes, pcc_stream_t s) {
  PCC_CHECK_ARG(num_frames >= 0 && n_total >= 0 && n_total < (1ll << 31) && log2_stride >= 0 && log2_stride <= 15);
  PCC_CHECK_ARG(status_dev);
  cudaStream_t st = as_stream(s);
  PCC_CUDA(cudaMemsetAsync(status_dev, 0, 4, st));
  if (num_frames == 0) return PCC_OK;
  PCC_CHECK_ARG(streams && stream_offsets_dev && stream_bytes_dev && row_offsets_dev && (keys_out || n_total == 0));
  PCC_CHECK_ARG(((uintptr_t)streams % 4) == 0);
  if (ws_bytes < pcc_octree_decode_workspace_bytes(n_total)) {
    set_error("pcc_octree_decode: workspace too small");
    return PCC_E_WORKSPACE;
  }
  uint64_t* a = (uint64_t*)ws;
  uint64_t* b = a + (n_total > 0 ? n_total : 1);
  launch_pdl(k_octree_decode, dim3((unsigned)num_frames), dim3(kOctDecThreads), 0, st, streams, stream_offsets_dev, stream_bytes_dev,
                                                                 row_offsets_dev, 3 * log2_stride, a, b, keys_out,
                                                                 status_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
