// (9) Capturer frame preparation on the device (capture half): one XYZRGBA measure
// [h*w, 4] float32 -> the {"points": int16 [m,3], "colors": float64 [m,3]} frame the encoder takes,
// following sender/capturer/capturer.py:84-120 -- validity mask in float32 (:93-94), Open3D
// voxel_down_sample in float64 (:99-106; published algorithm: min bound - voxel/2, floor((p - bound) / voxel),
// running `+=` sums in input order, sum / count), round-half-even to the grid (:107), first-wins row
// de-duplication in (x,y,z) order (:110-112), the max_points rows of largest z (:117-120).
//
// Bit-exactness against the oracle (oracle/capture_ref.py) rests on three things: every float operation
// here is a single IEEE operation (no contraction: differences and quotients only, products through
// __fmul_rn / __dmul_rn), the per-voxel sums run sequentially in input order (stable radix sort by voxel
// index, one thread walks a voxel's run), and the two orders the reference leaves open are the oracle's:
// voxels in ascending (ix,iy,iz) order, z ties to the earlier (x,y,z) row.
//
// HBM-bound integer/byte work of a few MB per frame: two radix sorts over h*w entries (key widths follow
// depth_clip / voxel_size: 31 and 31 bits at the demo's settings instead of 64 and 49), two scans, a
// sequential-run accumulate.  No host synchronisation; the row count comes back through a device counter.
#include "common.cuh"
#include <cub/cub.cuh>

namespace pcc {

constexpr int kCapThreads = 256;
constexpr int kIdxBits = 21;                       // most voxel index bits per axis in the sort key
constexpr uint64_t kCapPad = ~0ull;
constexpr int kZBins = 65536;

struct CapState {
  uint32_t min_ord[3];     // ordered-uint encoding of the float32 minima
  int32_t error;           // 1: voxel index beyond 21 bits, 2: rounded coordinate beyond int16
  int32_t n_cells;         // rows after de-duplication
  int32_t thr_z;           // rows with z > thr_z are kept ...
  int32_t thr_take;        // ... and the first thr_take rows with z == thr_z
};

__device__ __forceinline__ uint32_t f2ord(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

// capturer.py:93-94: isfinite(points).all(axis=1) & (norm(points) <= depth_clip), the norm in float32 as
// numpy evaluates it: sqrt((x*x + y*y) + z*z), every step rounded to float32.
__device__ __forceinline__ bool cap_valid(const float4 p, float clip) {
  if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) return false;
  const float s = __fadd_rn(__fadd_rn(__fmul_rn(p.x, p.x), __fmul_rn(p.y, p.y)), __fmul_rn(p.z, p.z));
  return __fsqrt_rn(s) <= clip;
}

__global__ void __launch_bounds__(kCapThreads)
k_cap_min(const float4* __restrict__ px, int64_t n, float clip, CapState* __restrict__ st) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t m0 = 0xffffffffu, m1 = 0xffffffffu, m2 = 0xffffffffu;
  if (i < n) {
    const float4 p = __ldg(px + i);
    if (cap_valid(p, clip)) {
      m0 = f2ord(p.x);
      m1 = f2ord(p.y);
      m2 = f2ord(p.z);
    }
  }
  m0 = __reduce_min_sync(0xffffffffu, m0);
  m1 = __reduce_min_sync(0xffffffffu, m1);
  m2 = __reduce_min_sync(0xffffffffu, m2);
  __shared__ uint32_t wmin[kCapThreads / 32][3];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) {
    wmin[wid][0] = m0;
    wmin[wid][1] = m1;
    wmin[wid][2] = m2;
  }
  __syncthreads();
  if (threadIdx.x < 3) {                       // one atomic per axis per block
    uint32_t m = wmin[0][threadIdx.x];
    for (int w = 1; w < kCapThreads / 32; ++w) m = min(m, wmin[w][threadIdx.x]);
    if (m != 0xffffffffu) atomicMin(&st->min_ord[threadIdx.x], m);
  }
}

// voxel index of every valid pixel -> sort key (ix << 2 vb | iy << vb | iz), vb bits per axis (from depth_clip /
// voxel_size, at most 21); invalid pixels carry the padding key, whose bit 3 vb is set: they sort last
__global__ void __launch_bounds__(kCapThreads)
k_cap_keys(const float4* __restrict__ px, int64_t n, float clip, double vs, int vb, CapState* __restrict__ st,
           uint64_t* __restrict__ keys, int32_t* __restrict__ pix) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = __ldg(px + i);
  uint64_t key = kCapPad;
  if (cap_valid(p, clip)) {
    const double half = __dmul_rn(vs, 0.5);
    const double bx = (double)ord2f(st->min_ord[0]) - half, by = (double)ord2f(st->min_ord[1]) - half,
                 bz = (double)ord2f(st->min_ord[2]) - half;
    const double fx = floor(((double)p.x - bx) / vs), fy = floor(((double)p.y - by) / vs),
                 fz = floor(((double)p.z - bz) / vs);
    const double lim = (double)(1 << vb);
    if (fx >= 0.0 && fy >= 0.0 && fz >= 0.0 && fx < lim && fy < lim && fz < lim) {
      key = ((uint64_t)fx << (2 * vb)) | ((uint64_t)fy << vb) | (uint64_t)fz;
    } else {
      st->error = 1;
    }
  }
  keys[i] = key;
  pix[i] = (int32_t)i;
}

// flags[i] = 1 where a run of equal keys starts (padding excluded)
__global__ void __launch_bounds__(kCapThreads)
k_cap_heads(const uint64_t* __restrict__ skeys, int64_t n, int32_t* __restrict__ flags) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = skeys[i];
  flags[i] = (k != kCapPad && (i == 0 || skeys[i - 1] != k)) ? 1 : 0;
}

// One thread per voxel walks its run in input order: sums of the float64 points and colours, then the
// grid cell of the mean point.  cell key = (x + bias) << 2 cb | (y + bias) << cb | (z + bias), bias = 2^(cb-1),
// cb bits per axis (from depth_clip / voxel_size, at most 16).
__global__ void __launch_bounds__(kCapThreads)
k_cap_accumulate(const float4* __restrict__ px, const uint64_t* __restrict__ skeys, const int32_t* __restrict__ spix,
                 const int32_t* __restrict__ flags, const int32_t* __restrict__ rank, int64_t n, double vs, int cb,
                 CapState* __restrict__ st, uint64_t* __restrict__ cell_keys, int32_t* __restrict__ vox_id,
                 double* __restrict__ vox_col) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  vox_id[i] = (int32_t)i;
  if (!flags[i]) return;
  const uint64_t k = skeys[i];
  double sx = 0.0, sy = 0.0, sz = 0.0, sr = 0.0, sg = 0.0, sb = 0.0;
  int cnt = 0;
  int64_t j = i;
  float4 p = __ldg(px + spix[j]);
  while (true) {
    const int64_t jn = j + 1;
    const bool more = jn < n && skeys[jn] == k;
    float4 pn = p;
    if (more) pn = __ldg(px + spix[jn]);       // next pixel in flight while this one is added
    const uint32_t c = __float_as_uint(p.w);
    sx += (double)p.x;
    sy += (double)p.y;
    sz += (double)p.z;
    sr += (double)(c & 0xffu) / 255.0;
    sg += (double)((c >> 8) & 0xffu) / 255.0;
    sb += (double)((c >> 16) & 0xffu) / 255.0;
    ++cnt;
    if (!more) break;
    p = pn;
    j = jn;
  }
  const double d = (double)cnt;
  const double qx = rint((sx / d) / vs), qy = rint((sy / d) / vs), qz = rint((sz / d) / vs);
  const int32_t v = rank[i] - 1;
  const int bias = 1 << (cb - 1);
  const double lim = (double)(bias - 1);           // cb <= 16: within int16
  if (!(fabs(qx) <= lim && fabs(qy) <= lim && fabs(qz) <= lim)) {
    st->error = 2;
    return;                                    // cell key stays padding
  }
  cell_keys[v] = ((uint64_t)((int)qx + bias) << (2 * cb)) | ((uint64_t)((int)qy + bias) << cb) | (uint64_t)((int)qz + bias);
  vox_col[3 * (int64_t)v + 0] = sr / d;
  vox_col[3 * (int64_t)v + 1] = sg / d;
  vox_col[3 * (int64_t)v + 2] = sb / d;
}

// histogram of z over the heads of the sorted cell keys (for the max_points threshold)
__global__ void __launch_bounds__(kCapThreads)
k_cap_cell_hist(const uint64_t* __restrict__ skeys, int64_t n, int cb, int32_t* __restrict__ zhist) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = skeys[i];
  if (k != kCapPad && (i == 0 || skeys[i - 1] != k)) atomicAdd(&zhist[(int)(k & ((1u << cb) - 1u))], 1);
}

// Single block.  Number of cells = sum of the histogram.  Finds the z threshold of capturer.py:117-120: the
// largest bin t with (rows of z >= t) >= max_points; rows above it are all kept, thr_take of the rows in it.
// Only the 2^cb bins the cell keys can reach are read (1024 at the demo's settings: one per thread).
__global__ void __launch_bounds__(1024)
k_cap_threshold(const int32_t* __restrict__ zhist, int cb, int32_t max_points, CapState* __restrict__ st) {
  pdl_trigger();
  pdl_wait();
  __shared__ int32_t warp_tot[32];
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const int bins = 1 << cb;
  const int per = bins > 1024 ? bins / 1024 : 1;
  const int top = bins - 1 - t * per;            // this thread's bins: top, top - 1, ..., top - per + 1 (descending z)
  int32_t local = 0;
  if (top >= 0)
    for (int b = 0; b < per; ++b) local += zhist[top - b];
  int32_t incl = local;                          // inclusive scan over threads
  for (int off = 1; off < 32; off <<= 1) {
    const int32_t v = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += v;
  }
  if (lane == 31) warp_tot[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    int32_t w = warp_tot[lane];
    for (int off = 1; off < 32; off <<= 1) {
      const int32_t v = __shfl_up_sync(0xffffffffu, w, off);
      if (lane >= off) w += v;
    }
    warp_tot[lane] = w;
  }
  __syncthreads();
  if (wid) incl += warp_tot[wid - 1];
  const int32_t n_cells = warp_tot[31];
  if (t == 0) st->n_cells = n_cells;
  if (max_points <= 0 || n_cells <= max_points) {
    if (t == 0) {
      st->thr_z = -1;          // bins are >= 0: everything is above the threshold
      st->thr_take = 0;
    }
    return;
  }
  const int32_t excl = incl - local;
  if (excl < max_points && incl >= max_points) {
    int32_t above = excl;
    for (int b = 0; b < per; ++b) {
      const int32_t h = zhist[top - b];
      if (above + h >= max_points) {
        st->thr_z = top - b;
        st->thr_take = max_points - above;
        break;
      }
      above += h;
    }
  }
}

// mark[i] = (head with z above the threshold) | (head with z at the threshold) << 32: one 64-bit scan ranks both
__global__ void __launch_bounds__(kCapThreads)
k_cap_mark(const uint64_t* __restrict__ skeys, int64_t n, int cb, const CapState* __restrict__ st, uint64_t* __restrict__ mark) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = skeys[i];
  uint64_t m = 0;
  if (k != kCapPad && (i == 0 || skeys[i - 1] != k)) {
    const int z = (int)(k & ((1u << cb) - 1u));
    m = z > st->thr_z ? 1ull : (z == st->thr_z ? (1ull << 32) : 0ull);
  }
  mark[i] = m;
}

// Kept rows: every head above the threshold and the first thr_take heads at it, in (x,y,z) order.  Row i lands
// behind the rows above the threshold that precede it plus the (at most thr_take) tied rows that precede it.
__global__ void __launch_bounds__(kCapThreads)
k_cap_emit(const uint64_t* __restrict__ skeys, const int32_t* __restrict__ svox, const uint64_t* __restrict__ mark_scan,
           const double* __restrict__ vox_col, int64_t n, int cb, int64_t cap, const CapState* __restrict__ st,
           int16_t* __restrict__ points, double* __restrict__ colors, int32_t* __restrict__ n_out) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t incl = mark_scan[i], prev = i ? mark_scan[i - 1] : 0ull;
  const int64_t take = st->thr_take;
  const int64_t above_incl = (int64_t)(incl & 0xffffffffu), eq_incl = (int64_t)(incl >> 32);
  const int64_t above_prev = (int64_t)(prev & 0xffffffffu), eq_prev = (int64_t)(prev >> 32);
  if (i == n - 1) *n_out = st->error ? -st->error : (int32_t)(above_incl + (eq_incl < take ? eq_incl : take));
  const bool is_above = above_incl != above_prev, is_eq = eq_incl != eq_prev;
  if (!(is_above || (is_eq && eq_incl <= take))) return;
  const int64_t r = above_prev + (eq_prev < take ? eq_prev : take);
  if (r >= cap) return;                                 // beyond the caller's capacity (the count still says so)
  const uint64_t k = skeys[i];
  const int bias = 1 << (cb - 1);
  const uint32_t fm = (1u << cb) - 1u;
  const int64_t o = 3 * r;
  points[o + 0] = (int16_t)((int)((k >> (2 * cb)) & fm) - bias);
  points[o + 1] = (int16_t)((int)((k >> cb) & fm) - bias);
  points[o + 2] = (int16_t)((int)(k & fm) - bias);
  const int64_t v = 3 * (int64_t)svox[i];
  colors[o + 0] = vox_col[v + 0];
  colors[o + 1] = vox_col[v + 1];
  colors[o + 2] = vox_col[v + 2];
}

static size_t cap_sort_bytes(int64_t n) {
  size_t b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, b, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const int32_t*)nullptr,
                                  (int32_t*)nullptr, (int)n);
  return b;
}
static size_t cap_scan_bytes(int64_t n) {
  size_t b32 = 0, b64 = 0;
  cub::DeviceScan::InclusiveSum(nullptr, b32, (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  cub::DeviceScan::InclusiveSum(nullptr, b64, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int)n);
  return b32 > b64 ? b32 : b64;
}

// bits that hold every value of [0, extent): 1 .. cap
static int cap_bits(double extent, int cap) {
  if (!(extent > 2.0)) return 1;
  if (!(extent < 4.0e9)) return cap;
  int b = 1;
  while (b < cap && (double)(1ull << b) < extent) ++b;
  return b;
}

struct CapCarver {
  char* p;
  char* end;
  void* take(int64_t bytes) {
    char* r = p;
    p += align_up(bytes, 256);
    return p <= end ? r : nullptr;
  }
};

}  // namespace pcc

using namespace pcc;

extern "C" {

int64_t pcc_capture_voxelize_workspace_bytes(int64_t n_px) {
  if (n_px < 0) return -1;
  const int64_t n = n_px > 0 ? n_px : 1;
  int64_t b = 256;                                   // state
  b += align_up(kZBins * 4, 256);                    // z histogram
  b += 2 * align_up(n * 8, 256);                     // keys, sorted keys (reused for the cell keys)
  b += 2 * align_up(n * 4, 256);                     // pixel / voxel ids, sorted
  b += 2 * align_up(n * 4, 256);                     // flags, rank
  b += align_up(n * 24, 256);                        // voxel colours
  b += align_up((int64_t)cap_sort_bytes(n), 256);
  b += align_up((int64_t)cap_scan_bytes(n), 256);
  return b + 256;
}

int pcc_capture_voxelize(const float* xyzrgba, int64_t n_px, double voxel_size, float depth_clip, int32_t max_points,
                         int16_t* points, double* colors, int64_t capacity, int32_t* n_out_dev, void* ws,
                         int64_t ws_bytes, pcc_stream_t s) {
  PCC_CHECK_ARG(n_px >= 0 && n_px < (1ll << 31) && n_out_dev && capacity >= 0);
  PCC_CHECK_ARG(voxel_size > 0.0 && voxel_size == voxel_size);
  cudaStream_t stm = as_stream(s);
  if (n_px == 0) {
    PCC_CUDA(cudaMemsetAsync(n_out_dev, 0, 4, stm));
    return PCC_OK;
  }
  PCC_CHECK_ARG(xyzrgba && ws && ((uintptr_t)xyzrgba % 16) == 0 && ((uintptr_t)ws % 256) == 0);
  PCC_CHECK_ARG(capacity == 0 || (points && colors));
  const int64_t n = n_px;
  CapCarver w{(char*)ws, (char*)ws + ws_bytes};
  CapState* st = (CapState*)w.take(256);
  int32_t* zhist = (int32_t*)w.take(kZBins * 4);
  uint64_t* keys = (uint64_t*)w.take(n * 8);
  uint64_t* skeys = (uint64_t*)w.take(n * 8);
  int32_t* ids = (int32_t*)w.take(n * 4);
  int32_t* sids = (int32_t*)w.take(n * 4);
  int32_t* flags = (int32_t*)w.take(n * 4);
  int32_t* rank = (int32_t*)w.take(n * 4);
  double* vox_col = (double*)w.take(n * 24);
  size_t sort_b = cap_sort_bytes(n), scan_b = cap_scan_bytes(n);
  void* sort_ws = w.take((int64_t)sort_b);
  void* scan_ws = w.take((int64_t)scan_b);
  if (!st || !zhist || !keys || !skeys || !ids || !sids || !flags || !rank || !vox_col || !sort_ws || !scan_ws) {
    set_error("pcc_capture_voxelize: workspace too small (%lld)", (long long)ws_bytes);
    return PCC_E_WORKSPACE;
  }
  const float4* px = (const float4*)xyzrgba;
  const int blocks = (int)((n + kCapThreads - 1) / kCapThreads);
  // Key widths from the settings: a kept point lies within depth_clip of the origin, so the voxel indices span at
  // most 2 clip / voxel + 2 and the grid cells lie within +-(clip / voxel + 1).  The kernels still check (error 1 / 2).
  const double reach = (double)depth_clip / voxel_size;
  const int vb = cap_bits(2.0 * reach + 4.0, kIdxBits);
  const int cb0 = cap_bits(2.0 * (reach + 2.0), 16);
  const int cb = cb0 < 2 ? 2 : cb0;
  // state: minima = all ones (ordered-uint +inf side), the rest and the histogram zero
  PCC_CUDA(cudaMemsetAsync(st, 0, 256 + align_up(kZBins * 4, 256), stm));
  PCC_CUDA(cudaMemsetAsync(st, 0xff, 12, stm));
  launch_pdl(k_cap_min, dim3(blocks), dim3(kCapThreads), 0, stm, px, n, depth_clip, st);
  PCC_LAUNCHED();
  launch_pdl(k_cap_keys, dim3(blocks), dim3(kCapThreads), 0, stm, px, n, depth_clip, voxel_size, vb, st, keys, ids);
  PCC_LAUNCHED();
  // stable: a voxel's pixels stay in input order.  3 vb index bits + the padding bit.
  PCC_CUDA(cub::DeviceRadixSort::SortPairs(sort_ws, sort_b, keys, skeys, ids, sids, (int)n, 0, 3 * vb + 1, stm));
  g_launches.fetch_add(1);
  launch_pdl(k_cap_heads, dim3(blocks), dim3(kCapThreads), 0, stm, skeys, n, flags);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, flags, rank, (int)n, stm));
  g_launches.fetch_add(1);
  PCC_CUDA(cudaMemsetAsync(keys, 0xff, n * 8, stm));          // cell keys: padding beyond the voxel count
  launch_pdl(k_cap_accumulate, dim3(blocks), dim3(kCapThreads), 0, stm, px, skeys, sids, flags, rank, n, voxel_size, cb, st, keys, ids, vox_col);
  PCC_LAUNCHED();
  // 3 cb key bits + the padding bit; stable: among voxels of one cell the lowest (ix,iy,iz) comes first
  PCC_CUDA(cub::DeviceRadixSort::SortPairs(sort_ws, sort_b, keys, skeys, ids, sids, (int)n, 0, 3 * cb + 1, stm));
  g_launches.fetch_add(1);
  launch_pdl(k_cap_cell_hist, dim3(blocks), dim3(kCapThreads), 0, stm, skeys, n, cb, zhist);
  PCC_LAUNCHED();
  launch_pdl(k_cap_threshold, dim3(1), dim3(1024), 0, stm, zhist, cb, max_points, st);
  PCC_LAUNCHED();
  uint64_t* mark = keys;                                      // the unsorted cell keys are no longer needed
  launch_pdl(k_cap_mark, dim3(blocks), dim3(kCapThreads), 0, stm, skeys, n, cb, st, mark);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, mark, mark, (int)n, stm));
  g_launches.fetch_add(1);
  launch_pdl(k_cap_emit, dim3(blocks), dim3(kCapThreads), 0, stm, skeys, sids, mark, vox_col, n, cb, capacity, st, points, colors, n_out_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
