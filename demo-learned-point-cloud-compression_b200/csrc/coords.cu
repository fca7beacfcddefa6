// (1) coordinate kernels: Morton keys, sort/dedup, stride-2 coarsening, generative children,
// frame offsets, canonical lexicographic order, row gathers.  All HBM-bound, one pass each.
#include "common.cuh"
#include <cub/cub.cuh>
#include <mutex>

namespace pcc {

thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

constexpr int kThreads = 256;
static inline unsigned blocks_for(int64_t n, int per_block = kThreads) {
  return (unsigned)((n + per_block - 1) / per_block);
}

__global__ void k_keys_from_i32(const int4* __restrict__ c, int64_t n, uint64_t* __restrict__ keys) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int4 v = c[i];
  keys[i] = pack_key(v.x, v.y, v.z, v.w);
}
__global__ void k_keys_from_f32(const float4* __restrict__ c, int64_t n, uint64_t* __restrict__ keys) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 v = c[i];
  keys[i] = pack_key((int)floorf(v.x), (int)floorf(v.y), (int)floorf(v.z), (int)floorf(v.w));
}
__global__ void k_ingest(const int16_t* __restrict__ p, const double* __restrict__ col, int64_t n,
                         int batch, uint64_t* __restrict__ keys, float4* __restrict__ feats) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int x = p[3 * i], y = p[3 * i + 1], z = p[3 * i + 2];
  keys[i] = pack_key(batch, x, y, z);
  feats[i] = make_float4(1.0f, (float)col[3 * i], (float)col[3 * i + 1], (float)col[3 * i + 2]);
}
__global__ void k_ingest_points(const int16_t* __restrict__ p, int64_t n, int batch, uint64_t* __restrict__ keys) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = pack_key(batch, p[3 * i], p[3 * i + 1], p[3 * i + 2]);
}
// feature row i = bf16(1, r, g, b, 0 ...) of colour row src_row[i] (same rounding as k_ingest followed by
// k_rows_f32_to_bf16: double -> float -> bf16 RN); chunks16 = 16-byte chunks per output row
__global__ void k_ingest_features(const double* __restrict__ col, const int32_t* __restrict__ src_row, int64_t m,
                                  uint4* __restrict__ out, int chunks16) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= m) return;
  const int64_t r = src_row ? (int64_t)src_row[i] : i;
  float c0 = 0.f, c1 = 0.f, c2 = 0.f;
  if (r >= 0) {
    c0 = (float)col[3 * r];
    c1 = (float)col[3 * r + 1];
    c2 = (float)col[3 * r + 2];
  }
  __nv_bfloat162 a = __floats2bfloat162_rn(1.0f, c0), b = __floats2bfloat162_rn(c1, c2);
  out[i * chunks16] = make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b), 0u, 0u);
  for (int k = 1; k < chunks16; ++k) out[i * chunks16 + k] = make_uint4(0u, 0u, 0u, 0u);
}
__global__ void k_ingest_features_bf16(const uint16_t* __restrict__ col, const int32_t* __restrict__ src_row, int64_t m,
                                       uint4* __restrict__ out, int chunks16) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= m) return;
  const int64_t r = src_row ? (int64_t)src_row[i] : i;
  uint32_t c0 = 0, c1 = 0, c2 = 0;
  if (r >= 0) {
    c0 = col[3 * r];
    c1 = col[3 * r + 1];
    c2 = col[3 * r + 2];
  }
  out[i * chunks16] = make_uint4(0x3f80u | (c0 << 16), c1 | (c2 << 16), 0u, 0u);   // bf16(1) = 0x3f80
  for (int k = 1; k < chunks16; ++k) out[i * chunks16 + k] = make_uint4(0u, 0u, 0u, 0u);
}
__global__ void k_coords_from_keys(const uint64_t* __restrict__ keys, int64_t n, int4* __restrict__ c) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  unpack_key(keys[i], b, x, y, z);
  c[i] = make_int4(b, x, y, z);
}

// decoder output: keys -> int32 xyz, colours nan -> 0 and clip(c * 255, 0, 255) / 255 in fp32
// (pack_batches, receiver/decoder/codec_parallel.py:474-502)
__global__ void k_unpack_output(const uint64_t* __restrict__ keys, const float* __restrict__ feats, int64_t n,
                                int32_t* __restrict__ pts, float* __restrict__ col) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (pts) {
    int b, x, y, z;
    unpack_key(keys[i], b, x, y, z);
    pts[3 * i] = x;
    pts[3 * i + 1] = y;
    pts[3 * i + 2] = z;
  }
  if (!col) return;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float v = feats[3 * i + c];
    if (v != v) v = 0.0f;
    float t = __fmul_rn(v, 255.0f);
    t = t < 0.0f ? 0.0f : t;          // np.clip's comparisons, not fmax/fmin: -0.0 stays -0.0 as in numpy
    t = t > 255.0f ? 255.0f : t;
    col[3 * i + c] = __fdiv_rn(t, 255.0f);
  }
}

// playout buffer of the client (receiver/client/client.py:139-146): points + offset as float32 [n,3], then
// colours uint8 [n,3] = trunc(255 * colour) with colour as pack_batches leaves it (k_unpack_output), every
// step in fp32 like numpy does it
__global__ void k_pack_visualizer(const uint64_t* __restrict__ keys, const float* __restrict__ feats, int64_t n,
                                  float offset, float* __restrict__ pts, uint8_t* __restrict__ col) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  unpack_key(keys[i], b, x, y, z);
  pts[3 * i] = __fadd_rn((float)x, offset);
  pts[3 * i + 1] = __fadd_rn((float)y, offset);
  pts[3 * i + 2] = __fadd_rn((float)z, offset);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float v = feats[3 * i + c];
    if (v != v) v = 0.0f;
    float t = __fmul_rn(v, 255.0f);
    t = t < 0.0f ? 0.0f : t;
    t = t > 255.0f ? 255.0f : t;
    t = __fmul_rn(255.0f, __fdiv_rn(t, 255.0f));
    col[3 * i + c] = (uint8_t)(int)t;      // 0 <= t <= 255: truncation like ndarray.astype(uint8)
  }
}

__global__ void k_iota(int32_t* v, int64_t n) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) v[i] = (int32_t)i;
}

// flags[i] = 1 where sorted key i starts a new run (after masking low bits)
__global__ void k_run_flags(const uint64_t* __restrict__ keys, int64_t n, uint64_t mask,
                            int32_t* __restrict__ flags) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t k = keys[i] & mask;
  flags[i] = (i == 0 || (keys[i - 1] & mask) != k) ? 1 : 0;
}
// incl = inclusive scan of flags. Writes unique keys, the first source row of each run and the
// run id of every input row.
__global__ void k_scatter_runs(const uint64_t* __restrict__ keys, const int32_t* __restrict__ src,
                               const int32_t* __restrict__ incl, int64_t n, uint64_t mask,
                               uint64_t* __restrict__ keys_out, int32_t* __restrict__ src_out,
                               int32_t* __restrict__ run_of, int32_t* __restrict__ n_out) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t r = incl[i] - 1;
  bool first = (i == 0) || (incl[i - 1] != incl[i]);
  if (first) {
    keys_out[r] = keys[i] & mask;
    if (src_out) src_out[r] = src ? src[i] : (int32_t)i;
  }
  if (run_of) run_of[i] = r;
  if (i == n - 1) *n_out = incl[i];
}

// ---- device-count variants (no host round trip between pyramid levels) --------------------------
// n_dev (may be null: then n = nmax) is the true element count; the launch covers nmax >= n.
__global__ void k_run_flags_dc(const uint64_t* __restrict__ keys, const int32_t* __restrict__ n_dev, int64_t nmax,
                               uint64_t mask, int32_t* __restrict__ flags) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= nmax) return;
  const int64_t n = n_dev ? (int64_t)*n_dev : nmax;
  int32_t f = 0;
  if (i < n) {
    uint64_t k = keys[i] & mask;
    f = (i == 0 || (keys[i - 1] & mask) != k) ? 1 : 0;
  }
  flags[i] = f;
}
__global__ void k_scatter_runs_dc(const uint64_t* __restrict__ keys, const int32_t* __restrict__ incl,
                                  const int32_t* __restrict__ n_dev, int64_t nmax, uint64_t mask,
                                  uint64_t* __restrict__ keys_out, int32_t* __restrict__ run_of,
                                  int32_t* __restrict__ n_out) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t n = n_dev ? (int64_t)*n_dev : nmax;
  if (i == 0 && n == 0) *n_out = 0;
  if (i >= n) return;
  int32_t r = incl[i] - 1;
  bool first = (i == 0) || (incl[i - 1] != incl[i]);
  if (first) keys_out[r] = keys[i] & mask;
  run_of[i] = r;
  if (i == n - 1) *n_out = incl[i];
}
// First row of frame b in a sorted key array: lower bound of b << 48.  The ends are looked at first -- frame 0 starts at
// row 0 and a frame index beyond the last key's starts at n -- so a one-frame GOP needs two loads per boundary instead
// of a 21-step chain of dependent loads (13 us per launch on 1.8 M keys, six to nine launches per GOP).
__device__ __forceinline__ int64_t frame_lower_bound(const uint64_t* __restrict__ keys, int64_t n, int b) {
  const uint64_t target = (uint64_t)b << 48;
  if (n <= 0 || keys[0] >= target) return 0;
  if (keys[n - 1] < target) return n;
  int64_t lo = 1, hi = n - 1;        // keys[0] < target <= keys[n - 1]
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void k_frame_offsets_dc(const uint64_t* __restrict__ keys, const int32_t* __restrict__ n_dev,
                                   int64_t nmax, int nf, int32_t* __restrict__ off) {
  pdl_trigger();
  pdl_wait();
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nf) return;
  const int64_t n = n_dev ? (int64_t)*n_dev : nmax;
  off[b] = (int32_t)frame_lower_bound(keys, n, b);
}

__global__ void k_children(const uint64_t* __restrict__ keys, int64_t n, int shift,
                           uint64_t* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;  // over 8n
  if (i >= 8 * n) return;
  out[i] = keys[i >> 3] | ((uint64_t)(i & 7) << shift);
}

__global__ void k_frame_offsets(const uint64_t* __restrict__ keys, int64_t n, int nf,
                                int32_t* __restrict__ off) {
  pdl_trigger();
  pdl_wait();
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nf) return;
  off[b] = (int32_t)frame_lower_bound(keys, n, b);
}

// shift = log2 of the tensor stride: coordinates are multiples of it, so their low bits carry nothing and each
// coordinate takes cb = 16 - shift bits of the sort key (fewer radix passes); shift = 0 is the general form.
__global__ void k_lex_keys(const uint64_t* __restrict__ keys, int64_t n, int shift, uint64_t* __restrict__ lex,
                           int32_t* __restrict__ row) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  unpack_key(keys[i], b, x, y, z);
  const int cb = 16 - shift;
  lex[i] = ((uint64_t)b << (3 * cb)) | ((uint64_t)((uint32_t)(x + 32768) >> shift) << (2 * cb)) |
           ((uint64_t)((uint32_t)(y + 32768) >> shift) << cb) | (uint64_t)((uint32_t)(z + 32768) >> shift);
  row[i] = (int32_t)i;
}

// ---- row gathers -----------------------------------------------------------------------------
__global__ void k_rows_f32_to_bf16(const float* __restrict__ src, int c_src,
                                   const int32_t* __restrict__ src_row, int64_t n_out,
                                   __nv_bfloat16* __restrict__ dst, int c_dst) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_out * c_dst) return;
  int64_t r = i / c_dst;
  int c = (int)(i - r * c_dst);
  int64_t sr = src_row ? src_row[r] : r;
  float v = (c < c_src && sr >= 0) ? src[sr * c_src + c] : 0.0f;
  dst[i] = __float2bfloat16_rn(v);
}
__global__ void k_rows_bf16_to_f32(const __nv_bfloat16* __restrict__ src, int c_src, int64_t n,
                                   float* __restrict__ dst, int c_dst) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n * c_dst) return;
  int64_t r = i / c_dst;
  int c = (int)(i - r * c_dst);
  dst[i] = __bfloat162float(src[r * c_src + c]);
}
// 16-byte chunks per row; c*2 bytes must be a multiple of 16.
__global__ void k_gather_rows16(const uint4* __restrict__ src, int chunks,
                                const int32_t* __restrict__ src_row, int64_t n_out,
                                uint4* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_out * chunks) return;
  int64_t r = i / chunks;
  int c = (int)(i - r * chunks);
  int64_t sr = src_row[r];
  dst[i] = sr >= 0 ? src[sr * chunks + c] : make_uint4(0, 0, 0, 0);
}
__global__ void k_gather_rows_f32(const float* __restrict__ src, int c,
                                  const int32_t* __restrict__ src_row, int64_t n_out,
                                  float* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_out * c) return;
  int64_t r = i / c;
  int cc = (int)(i - r * c);
  int64_t sr = src_row[r];
  dst[i] = sr >= 0 ? src[sr * c + cc] : 0.0f;
}
__global__ void k_gather_keys(const uint64_t* __restrict__ src, const int32_t* __restrict__ src_row,
                              int64_t n_out, uint64_t* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n_out) dst[i] = src[src_row[i]];
}

// Rows of a candidate set that the prune did not keep are zeroed in place, so that a convolution over ALL candidates
// (complete octets: pcc_spconv_octet on the parents' kernel map) gives, on the kept rows, the convolution over the kept
// set -- an absent neighbour and a zero row contribute the same nothing.  keep: ascending kept row numbers (the top-k
// emits them in order); thread (i, kc) clears 16-byte chunk kc of the rows between kept rows i-1 and i.
__global__ void k_zero_unkept_rows(const int32_t* __restrict__ keep, int64_t m, int64_t n, uint4* __restrict__ feats, int chunks) {
  pdl_trigger();
  pdl_wait();
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t i = t / chunks;
  const int kc = (int)(t - i * chunks);
  if (i > m) return;
  int64_t lo = i == 0 ? 0 : (int64_t)keep[i - 1] + 1;
  int64_t hi = i == m ? n : (int64_t)keep[i];
  if (lo < 0) lo = 0;
  if (hi > n) hi = n;
  for (int64_t r = lo; r < hi; ++r) feats[r * chunks + kc] = make_uint4(0u, 0u, 0u, 0u);
}

struct WsCarver {
  char* p;
  int64_t left;
  WsCarver(void* ws, int64_t bytes) : p((char*)ws), left(bytes) {}
  void* take(int64_t bytes) {
    bytes = align_up(bytes, 256);
    if (bytes > left) return nullptr;
    void* r = p;
    p += bytes;
    left -= bytes;
    return r;
  }
};

}  // namespace pcc

using namespace pcc;

extern "C" {

int pcc_version(void) { return 100; }
const char* pcc_last_error(void) { return pcc::g_err; }
int64_t pcc_launch_count(void) { return pcc::g_launches.load(); }

int pcc_keys_from_coords_i32(const int32_t* coords, int64_t n, uint64_t* keys, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(coords && keys);
  launch_pdl(k_keys_from_i32, dim3(blocks_for(n)), dim3(kThreads), 0, as_stream(s), (const int4*)coords, n, keys);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_keys_from_coords_f32(const float* coords, int64_t n, uint64_t* keys, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(coords && keys);
  launch_pdl(k_keys_from_f32, dim3(blocks_for(n)), dim3(kThreads), 0, as_stream(s), (const float4*)coords, n, keys);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_ingest_frame(const int16_t* points, const double* colors, int64_t n, int32_t batch,
                     uint64_t* keys, float* feats4, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && batch >= 0 && batch < 65535);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(points && colors && keys && feats4);
  launch_pdl(k_ingest, dim3(blocks_for(n)), dim3(kThreads), 0, as_stream(s), points, colors, n, batch, keys,
                                                         (float4*)feats4);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_ingest_points(const int16_t* points, int64_t n, int32_t batch, uint64_t* keys, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && batch >= 0 && batch < 65535);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(points && keys);
  launch_pdl(k_ingest_points, dim3(blocks_for(n)), dim3(kThreads), 0, as_stream(s), points, n, batch, keys);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_ingest_features(const double* colors, const int32_t* src_row, int64_t m, void* feats_bf16, int c_dst,
                        pcc_stream_t s) {
  PCC_CHECK_ARG(m >= 0 && c_dst >= 8 && c_dst % 8 == 0);
  if (m == 0) return PCC_OK;
  PCC_CHECK_ARG(colors && feats_bf16 && ((uintptr_t)feats_bf16 % 16) == 0);
  launch_pdl(k_ingest_features, dim3(blocks_for(m)), dim3(kThreads), 0, as_stream(s), colors, src_row, m, (uint4*)feats_bf16, c_dst / 8);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_ingest_features_bf16(const void* colors_bf16, const int32_t* src_row, int64_t m, void* feats_bf16, int c_dst,
                             pcc_stream_t s) {
  PCC_CHECK_ARG(m >= 0 && c_dst >= 8 && c_dst % 8 == 0);
  if (m == 0) return PCC_OK;
  PCC_CHECK_ARG(colors_bf16 && feats_bf16 && ((uintptr_t)feats_bf16 % 16) == 0 && ((uintptr_t)colors_bf16 % 2) == 0);
  launch_pdl(k_ingest_features_bf16, dim3(blocks_for(m)), dim3(kThreads), 0, as_stream(s), (const uint16_t*)colors_bf16, src_row, m,
                                                                      (uint4*)feats_bf16, c_dst / 8);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_unpack_output(const uint64_t* keys, const float* feats3, int64_t n, int32_t* points, float* colors,
                      pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG((points || colors) && (!points || keys) && (!colors || feats3));
  launch_pdl(k_unpack_output, dim3(blocks_for(n)), dim3(kThreads), 0, as_stream(s), keys, feats3, n, points, colors);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_pack_visualizer(const uint64_t* keys, const float* feats3, int64_t n, float offset, float* points_f32,
                        uint8_t* colors_u8, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys && feats3 && points_f32 && colors_u8);
  launch_pdl(k_pack_visualizer, dim3(blocks_for(n)), dim3(kThreads), 0, as_stream(s), keys, feats3, n, offset, points_f32, colors_u8);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_coords_from_keys(const uint64_t* keys, int64_t n, int32_t* coords, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(coords && keys);
  launch_pdl(k_coords_from_keys, dim3(blocks_for(n)), dim3(kThreads), 0, as_stream(s), keys, n, (int4*)coords);
  PCC_LAUNCHED();
  return PCC_OK;
}

// ---- sort + unique ---------------------------------------------------------------------------
static size_t cub_sort_pairs_bytes(int64_t n) {
  size_t b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, b, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  return b;
}
static size_t cub_scan_bytes(int64_t n) {
  size_t b = 0;
  cub::DeviceScan::InclusiveSum(nullptr, b, (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  return b;
}

int64_t pcc_sort_unique_workspace_bytes(int64_t n) {
  if (n <= 0) return 256;
  int64_t b = 0;
  b += align_up(n * 8, 256);  // sorted keys
  b += align_up(n * 4, 256);  // iota
  b += align_up(n * 4, 256);  // sorted rows
  b += align_up(n * 4, 256);  // flags / scan
  b += align_up((int64_t)cub_sort_pairs_bytes(n), 256);
  b += align_up((int64_t)cub_scan_bytes(n), 256);
  return b + 256;
}

int pcc_sort_unique(const uint64_t* keys_in, int64_t n, uint64_t* keys_out, int32_t* src_row,
                    int32_t* n_out_dev, void* ws, int64_t ws_bytes, pcc_stream_t s) {
  return pcc_sort_unique_bits(keys_in, n, 64, keys_out, src_row, n_out_dev, ws, ws_bytes, s);
}

// key_bits: the keys' bits [key_bits, 64) are known to be zero (a GOP of F frames: 48 + ceil(log2 F)), so the radix sort
// skips their digit passes -- 6 instead of 8 passes for a single frame.
int pcc_sort_unique_bits(const uint64_t* keys_in, int64_t n, int key_bits, uint64_t* keys_out, int32_t* src_row,
                         int32_t* n_out_dev, void* ws, int64_t ws_bytes, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n < (1ll << 31) && n_out_dev && key_bits >= 1 && key_bits <= 64);
  cudaStream_t st = as_stream(s);
  if (n == 0) {
    PCC_CUDA(cudaMemsetAsync(n_out_dev, 0, 4, st));
    return PCC_OK;
  }
  PCC_CHECK_ARG(keys_in && keys_out && ws);
  WsCarver w(ws, ws_bytes);
  uint64_t* skeys = (uint64_t*)w.take(n * 8);
  int32_t* iota = (int32_t*)w.take(n * 4);
  int32_t* srows = (int32_t*)w.take(n * 4);
  int32_t* flags = (int32_t*)w.take(n * 4);
  size_t sort_b = cub_sort_pairs_bytes(n), scan_b = cub_scan_bytes(n);
  void* sort_ws = w.take((int64_t)sort_b);
  void* scan_ws = w.take((int64_t)scan_b);
  if (!skeys || !iota || !srows || !flags || !sort_ws || !scan_ws) {
    set_error("pcc_sort_unique: workspace too small (%lld)", (long long)ws_bytes);
    return PCC_E_WORKSPACE;
  }
  launch_pdl(k_iota, dim3(blocks_for(n)), dim3(kThreads), 0, st, iota, n);
  PCC_LAUNCHED();
  // radix sort is stable: among equal keys the lowest input row comes first ("first wins")
  PCC_CUDA(cub::DeviceRadixSort::SortPairs(sort_ws, sort_b, keys_in, skeys, iota, srows, (int)n, 0,
                                           key_bits, st));
  g_launches.fetch_add(1);
  launch_pdl(k_run_flags, dim3(blocks_for(n)), dim3(kThreads), 0, st, skeys, n, ~0ull, flags);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, flags, flags, (int)n, st));
  g_launches.fetch_add(1);
  launch_pdl(k_scatter_runs, dim3(blocks_for(n)), dim3(kThreads), 0, st, skeys, srows, flags, n, ~0ull, keys_out,
                                                     src_row, nullptr, n_out_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

// ---- stride-2 coarsening: sorted Morton keys stay sorted after masking low bits ----------------
int64_t pcc_coords_downsample_workspace_bytes(int64_t n) {
  if (n <= 0) return 256;
  return align_up(n * 4, 256) + align_up((int64_t)cub_scan_bytes(n), 256) + 256;
}

int pcc_coords_downsample(const uint64_t* keys_in, int64_t n, int log2_stride, uint64_t* keys_out,
                          int32_t* parent_row, int32_t* n_out_dev, void* ws, int64_t ws_bytes,
                          pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n < (1ll << 31) && n_out_dev && log2_stride >= 1 && log2_stride <= 15);
  cudaStream_t st = as_stream(s);
  if (n == 0) {
    PCC_CUDA(cudaMemsetAsync(n_out_dev, 0, 4, st));
    return PCC_OK;
  }
  PCC_CHECK_ARG(keys_in && keys_out && ws);
  WsCarver w(ws, ws_bytes);
  int32_t* flags = (int32_t*)w.take(n * 4);
  size_t scan_b = cub_scan_bytes(n);
  void* scan_ws = w.take((int64_t)scan_b);
  if (!flags || !scan_ws) {
    set_error("pcc_coords_downsample: workspace too small");
    return PCC_E_WORKSPACE;
  }
  uint64_t mask = ~((1ull << (3 * log2_stride)) - 1ull);
  launch_pdl(k_run_flags, dim3(blocks_for(n)), dim3(kThreads), 0, st, keys_in, n, mask, flags);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, flags, flags, (int)n, st));
  g_launches.fetch_add(1);
  launch_pdl(k_scatter_runs, dim3(blocks_for(n)), dim3(kThreads), 0, st, keys_in, nullptr, flags, n, mask, keys_out,
                                                     nullptr, parent_row, n_out_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

// ---- pcc_coords_downsample_static: the same operation for callers that may not allocate ---------
// The flags and the scan's temporary storage live in a file-scope device array that every GPU which
// loads the library holds until exit.  `g_static_scratch` below is the HOST-side stand-in of that array:
// host code may not hand it to a kernel; the device address comes from cudaGetSymbolAddress, looked up
// once per device with that device current (static_scratch_ptr).
constexpr int64_t kStaticMaxRows = 1ll << 20;                 // 1 Mi rows per call
constexpr int64_t kStaticFlagBytes = kStaticMaxRows * 4;      // int32 run flags / inclusive scan, in place
constexpr int64_t kStaticScanBytes = 256ll << 10;             // CUB DeviceScan temporary storage (a few KB at 1 Mi items)
constexpr int64_t kStaticScratchBytes = kStaticFlagBytes + kStaticScanBytes;
constexpr int kStaticMaxDevices = 64;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

static std::mutex g_static_mu;
static void* g_static_ptr[kStaticMaxDevices];   // device address of g_static_scratch, per device ordinal

static int static_scratch_ptr(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_error("pcc_coords_downsample_static: cudaGetDevice: %s", cudaGetErrorString(e));
    return PCC_E_CUDA;
  }
  if (dev < 0 || dev >= kStaticMaxDevices) {
    set_error("pcc_coords_downsample_static: device ordinal %d beyond %d", dev, kStaticMaxDevices);
    return PCC_E_INVALID;
  }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);      // the current device's copy of the array
    if (e != cudaSuccess || !ptr) {
      set_error("pcc_coords_downsample_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return PCC_E_CUDA;
    }
    g_static_ptr[dev] = ptr;
  }
  *out = g_static_ptr[dev];
  return PCC_OK;
}

int64_t pcc_coords_downsample_static_max_rows(void) { return kStaticMaxRows; }

int pcc_coords_downsample_static(const uint64_t* keys_in, int64_t n, int log2_stride, uint64_t* keys_out,
                                 int32_t* parent_row, int32_t* n_out_dev, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n_out_dev && log2_stride >= 1 && log2_stride <= 15);
  if (n > kStaticMaxRows) {
    set_error("pcc_coords_downsample_static: n = %lld rows exceeds the static scratch limit of %lld rows; "
              "use pcc_coords_downsample with a caller-owned workspace", (long long)n, (long long)kStaticMaxRows);
    return PCC_E_INVALID;
  }
  cudaStream_t st = as_stream(s);
  if (n == 0) {
    PCC_CUDA(cudaMemsetAsync(n_out_dev, 0, 4, st));
    return PCC_OK;
  }
  PCC_CHECK_ARG(keys_in && keys_out);
  void* base = nullptr;
  int rc = static_scratch_ptr(&base);
  if (rc != PCC_OK) return rc;
  int32_t* flags = (int32_t*)base;
  void* scan_ws = (char*)base + kStaticFlagBytes;
  size_t scan_b = cub_scan_bytes(n);
  if ((int64_t)scan_b > kStaticScanBytes) {
    set_error("pcc_coords_downsample_static: scan needs %lld bytes of temporary storage, %lld reserved",
              (long long)scan_b, (long long)kStaticScanBytes);
    return PCC_E_WORKSPACE;
  }
  // Nothing in the array has to start at zero: every flag the call reads is written by k_run_flags first and
  // CUB's scan initialises its own tile states, so the previous call's contents are simply overwritten.
  uint64_t mask = ~((1ull << (3 * log2_stride)) - 1ull);
  launch_pdl(k_run_flags, dim3(blocks_for(n)), dim3(kThreads), 0, st, keys_in, n, mask, flags);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, flags, flags, (int)n, st));
  g_launches.fetch_add(1);
  launch_pdl(k_scatter_runs, dim3(blocks_for(n)), dim3(kThreads), 0, st, keys_in, nullptr, flags, n, mask, keys_out,
                                                     nullptr, parent_row, n_out_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

// Whole stride pyramid in one call: level l+1 = stride-2 coarsening of level l, counts stay on the
// device (the launches of a level cover the input's size n, an upper bound of every level).
int64_t pcc_coords_pyramid_workspace_bytes(int64_t n) { return pcc_coords_downsample_workspace_bytes(n); }

int pcc_coords_pyramid(const uint64_t* keys_in, int64_t n, int log2_stride, int levels, uint64_t* keys_out,
                       int32_t* parent_row, int32_t* counts_dev, int32_t num_frames,
                       int32_t* frame_offsets_dev, void* ws, int64_t ws_bytes, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n < (1ll << 31) && levels >= 1 && levels <= 14 && log2_stride >= 0 &&
                log2_stride + levels <= 15 && counts_dev);
  PCC_CHECK_ARG(num_frames >= 0 && num_frames < 65535 && (num_frames == 0 || frame_offsets_dev));
  cudaStream_t st = as_stream(s);
  if (n == 0) {
    PCC_CUDA(cudaMemsetAsync(counts_dev, 0, 4 * (size_t)levels, st));
    if (num_frames) PCC_CUDA(cudaMemsetAsync(frame_offsets_dev, 0, 4 * (size_t)(levels + 1) * (num_frames + 1), st));
    return PCC_OK;
  }
  PCC_CHECK_ARG(keys_in && keys_out && parent_row && ws);
  WsCarver w(ws, ws_bytes);
  int32_t* flags = (int32_t*)w.take(n * 4);
  size_t scan_b = cub_scan_bytes(n);
  void* scan_ws = w.take((int64_t)scan_b);
  if (!flags || !scan_ws) {
    set_error("pcc_coords_pyramid: workspace too small");
    return PCC_E_WORKSPACE;
  }
  const unsigned fo_blocks = blocks_for(num_frames + 1, 64);
  if (num_frames) {
    launch_pdl(k_frame_offsets_dc, dim3(fo_blocks), dim3(64), 0, st, keys_in, nullptr, n, num_frames, frame_offsets_dev);
    PCC_LAUNCHED();
  }
  const uint64_t* cur = keys_in;
  const int32_t* cur_n = nullptr;
  for (int l = 0; l < levels; ++l) {
    const uint64_t mask = ~((1ull << (3 * (log2_stride + l + 1))) - 1ull);
    uint64_t* out = keys_out + (int64_t)l * n;
    int32_t* par = parent_row + (int64_t)l * n;
    launch_pdl(k_run_flags_dc, dim3(blocks_for(n)), dim3(kThreads), 0, st, cur, cur_n, n, mask, flags);
    PCC_LAUNCHED();
    PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, flags, flags, (int)n, st));
    g_launches.fetch_add(1);
    launch_pdl(k_scatter_runs_dc, dim3(blocks_for(n)), dim3(kThreads), 0, st, cur, flags, cur_n, n, mask, out, par, counts_dev + l);
    PCC_LAUNCHED();
    if (num_frames) {
      launch_pdl(k_frame_offsets_dc, dim3(fo_blocks), dim3(64), 0, st, out, counts_dev + l, n, num_frames,
                                                   frame_offsets_dev + (int64_t)(l + 1) * (num_frames + 1));
      PCC_LAUNCHED();
    }
    cur = out;
    cur_n = counts_dev + l;
  }
  return PCC_OK;
}

int pcc_coords_children(const uint64_t* keys_in, int64_t n, int log2_stride, uint64_t* keys_out,
                        pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && 8 * n < (1ll << 31) && log2_stride >= 0 && log2_stride <= 14);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys_in && keys_out);
  launch_pdl(k_children, dim3(blocks_for(8 * n)), dim3(kThreads), 0, as_stream(s), keys_in, n, 3 * log2_stride,
                                                               keys_out);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_frame_offsets(const uint64_t* keys, int64_t n, int32_t num_frames, int32_t* offsets,
                      pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && num_frames >= 0 && num_frames < 65535 && offsets);
  PCC_CHECK_ARG(n == 0 || keys);
  launch_pdl(k_frame_offsets, dim3(blocks_for(num_frames + 1, 64)), dim3(64), 0, as_stream(s), keys, n, num_frames,
                                                                           offsets);
  PCC_LAUNCHED();
  return PCC_OK;
}

int64_t pcc_lex_order_workspace_bytes(int64_t n) {
  if (n <= 0) return 256;
  return 2 * align_up(n * 8, 256) + align_up(n * 4, 256) +
         align_up((int64_t)cub_sort_pairs_bytes(n), 256) + 256;
}
int pcc_lex_order(const uint64_t* keys, int64_t n, int32_t* perm, void* ws, int64_t ws_bytes,
                  pcc_stream_t s) {
  return pcc_lex_order_bits(keys, n, 0, 16, perm, ws, ws_bytes, s);
}

// log2_stride: every coordinate is a multiple of 2^log2_stride (the tensor stride of the set); frame_bits: the frame
// index is below 2^frame_bits.  The sort key then has 3 (16 - log2_stride) + frame_bits significant bits.
int pcc_lex_order_bits(const uint64_t* keys, int64_t n, int log2_stride, int frame_bits, int32_t* perm, void* ws,
                       int64_t ws_bytes, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && n < (1ll << 31) && log2_stride >= 0 && log2_stride <= 15 && frame_bits >= 0 && frame_bits <= 16);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys && perm && ws);
  cudaStream_t st = as_stream(s);
  WsCarver w(ws, ws_bytes);
  uint64_t* lex = (uint64_t*)w.take(n * 8);
  uint64_t* lex_sorted = (uint64_t*)w.take(n * 8);
  int32_t* row = (int32_t*)w.take(n * 4);
  size_t sort_b = cub_sort_pairs_bytes(n);
  void* sort_ws = w.take((int64_t)sort_b);
  if (!lex || !lex_sorted || !row || !sort_ws) {
    set_error("pcc_lex_order: workspace too small");
    return PCC_E_WORKSPACE;
  }
  launch_pdl(k_lex_keys, dim3(blocks_for(n)), dim3(kThreads), 0, st, keys, n, log2_stride, lex, row);
  PCC_LAUNCHED();
  int end_bit = 3 * (16 - log2_stride) + frame_bits;
  if (end_bit < 1) end_bit = 1;
  PCC_CUDA(cub::DeviceRadixSort::SortPairs(sort_ws, sort_b, lex, lex_sorted, row, perm, (int)n, 0,
                                           end_bit, st));
  g_launches.fetch_add(1);
  return PCC_OK;
}

// ---- row utilities ---------------------------------------------------------------------------
int pcc_rows_f32_to_bf16(const float* src, int c_src, const int32_t* src_row, int64_t n_out,
                         void* dst_bf16, int c_dst, pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0 && c_src > 0 && c_dst > 0);
  if (n_out == 0) return PCC_OK;
  PCC_CHECK_ARG(src && dst_bf16);
  launch_pdl(k_rows_f32_to_bf16, dim3(blocks_for(n_out * c_dst)), dim3(kThreads), 0, as_stream(s), src, c_src, src_row, n_out, (__nv_bfloat16*)dst_bf16, c_dst);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_rows_bf16_to_f32(const void* src_bf16, int c_src, int64_t n, float* dst, int c_dst,
                         pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && c_src > 0 && c_dst > 0 && c_dst <= c_src);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(src_bf16 && dst);
  launch_pdl(k_rows_bf16_to_f32, dim3(blocks_for(n * c_dst)), dim3(kThreads), 0, as_stream(s), (const __nv_bfloat16*)src_bf16, c_src, n, dst, c_dst);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_gather_rows_bf16(const void* src_bf16, int c, const int32_t* src_row, int64_t n_out,
                         void* dst_bf16, pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0 && c > 0 && c % 8 == 0);
  if (n_out == 0) return PCC_OK;
  PCC_CHECK_ARG(src_bf16 && dst_bf16 && src_row);
  int chunks = c / 8;
  launch_pdl(k_gather_rows16, dim3(blocks_for(n_out * chunks)), dim3(kThreads), 0, as_stream(s), (const uint4*)src_bf16, chunks, src_row, n_out, (uint4*)dst_bf16);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_gather_rows_f32(const float* src, int c, const int32_t* src_row, int64_t n_out, float* dst,
                        pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0 && c > 0);
  if (n_out == 0) return PCC_OK;
  PCC_CHECK_ARG(src && dst && src_row);
  launch_pdl(k_gather_rows_f32, dim3(blocks_for(n_out * c)), dim3(kThreads), 0, as_stream(s), src, c, src_row, n_out,
                                                                          dst);
  PCC_LAUNCHED();
  return PCC_OK;
}
int pcc_gather_keys(const uint64_t* src, const int32_t* src_row, int64_t n_out, uint64_t* dst,
                    pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0);
  if (n_out == 0) return PCC_OK;
  PCC_CHECK_ARG(src && dst && src_row);
  launch_pdl(k_gather_keys, dim3(blocks_for(n_out)), dim3(kThreads), 0, as_stream(s), src, src_row, n_out, dst);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_zero_unkept_rows(const int32_t* keep, int64_t m, int64_t n, void* feats_bf16, int c, pcc_stream_t s) {
  PCC_CHECK_ARG(m >= 0 && n >= 0 && m <= n && n < (1ll << 31) && c > 0 && c % 8 == 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(feats_bf16 && (keep || m == 0) && ((uintptr_t)feats_bf16 % 16) == 0);
  const int chunks = c / 8;
  launch_pdl(k_zero_unkept_rows, dim3(blocks_for((m + 1) * chunks)), dim3(kThreads), 0, as_stream(s), keep, m, n,
             (uint4*)feats_bf16, chunks);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
