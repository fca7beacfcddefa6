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
// HBM-bound integer/byte work of a few MB per frame: two 64-bit radix sorts over h*w entries, four scans,
// a sequential-run accumulate.  No host synchronisation; the row count comes back through a device counter.
#include "common.cuh"
#include <cub/cub.cuh>

namespace pcc {

constexpr int kCapThreads = 256;
constexpr int kIdxBits = 21;                       // voxel index bits per axis in the sort key
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
  if ((threadIdx.x & 31) == 0 && m0 != 0xffffffffu) {
    atomicMin(&st->min_ord[0], m0);
    atomicMin(&st->min_ord[1], m1);
    atomicMin(&st->min_ord[2], m2);
  }
}

// voxel index of every valid pixel -> sort key (ix << 42 | iy << 21 | iz); invalid pixels sort last
__global__ void __launch_bounds__(kCapThreads)
k_cap_keys(const float4* __restrict__ px, int64_t n, float clip, double vs, CapState* __restrict__ st,
           uint64_t* __restrict__ keys, int32_t* __restrict__ pix) {
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
    const double lim = (double)(1 << kIdxBits);
    if (fx >= 0.0 && fy >= 0.0 && fz >= 0.0 && fx < lim && fy < lim && fz < lim) {
      key = ((uint64_t)fx << (2 * kIdxBits)) | ((uint64_t)fy << kIdxBits) | (uint64_t)fz;
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
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = skeys[i];
  flags[i] = (k != kCapPad && (i == 0 || skeys[i - 1] != k)) ? 1 : 0;
}

// One thread per voxel walks its run in input order: sums of the float64 points and colours, then the
// grid cell of the mean point.  cell key = (x + 32768) << 32 | (y + 32768) << 16 | (z + 32768).
__global__ void __launch_bounds__(kCapThreads)
k_cap_accumulate(const float4* __restrict__ px, const uint64_t* __restrict__ skeys, const int32_t* __restrict__ spix,
                 const int32_t* __restrict__ flags, const int32_t* __restrict__ rank, int64_t n, double vs,
                 CapState* __restrict__ st, uint64_t* __restrict__ cell_keys, int32_t* __restrict__ vox_id,
                 double* __restrict__ vox_col) {
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
  if (fabs(qx) > 32767.0 || fabs(qy) > 32767.0 || fabs(qz) > 32767.0) {
    st->error = 2;
    return;                                    // cell key stays padding
  }
  cell_keys[v] = ((uint64_t)((int)qx + 32768) << 32) | ((uint64_t)((int)qy + 32768) << 16) | (uint64_t)((int)qz + 32768);
  vox_col[3 * (int64_t)v + 0] = sr / d;
  vox_col[3 * (int64_t)v + 1] = sg / d;
  vox_col[3 * (int64_t)v + 2] = sb / d;
}

// heads of the sorted cell keys + histogram of their z (for the max_points threshold)
__global__ void __launch_bounds__(kCapThreads)
k_cap_cell_heads(const uint64_t* __restrict__ skeys, int64_t n, int32_t* __restrict__ flags, int32_t* __restrict__ zhist) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = skeys[i];
  const bool head = k != kCapPad && (i == 0 || skeys[i - 1] != k);
  flags[i] = head ? 1 : 0;
  if (head) atomicAdd(&zhist[(int)(k & 0xffffu)], 1);
}

// Single block.  rank[n - 1] = number of cells.  Finds the z threshold of capturer.py:117-120: the largest
// bin t with (rows of z >= t) >= max_points; rows above it are all kept, thr_take of the rows in it.
__global__ void __launch_bounds__(1024)
k_cap_threshold(const int32_t* __restrict__ zhist, const int32_t* __restrict__ rank, int64_t n, int32_t max_points,
                CapState* __restrict__ st) {
  __shared__ int32_t part[1024];
  const int t = threadIdx.x;
  const int32_t n_cells = rank[n - 1];
  if (t == 0) st->n_cells = n_cells;
  if (max_points <= 0 || n_cells <= max_points) {
    if (t == 0) {
      st->thr_z = -1;          // bins are >= 0: everything is above the threshold
      st->thr_take = 0;
    }
    return;
  }
  constexpr int kPer = kZBins / 1024;
  const int top = kZBins - 1 - t * kPer;         // this thread's bins: top, top - 1, ..., top - kPer + 1
  int32_t local = 0;
  for (int b = 0; b < kPer; ++b) local += zhist[top - b];
  part[t] = local;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {     // inclusive scan over threads (descending z)
    const int32_t add = t >= off ? part[t - off] : 0;
    __syncthreads();
    part[t] += add;
    __syncthreads();
  }
  const int32_t incl = part[t], excl = incl - local;
  if (excl < max_points && incl >= max_points) {
    int32_t above = excl;
    for (int b = 0; b < kPer; ++b) {
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

// eq[i] = head with z == threshold
__global__ void __launch_bounds__(kCapThreads)
k_cap_eq(const uint64_t* __restrict__ skeys, const int32_t* __restrict__ flags, int64_t n, const CapState* __restrict__ st,
         int32_t* __restrict__ eq) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  eq[i] = (flags[i] && (int)(skeys[i] & 0xffffu) == st->thr_z) ? 1 : 0;
}

// keep[i] (written over eq's scan): head with z above the threshold, or one of the first thr_take at it
__global__ void __launch_bounds__(kCapThreads)
k_cap_keep(const uint64_t* __restrict__ skeys, const int32_t* __restrict__ flags, int64_t n, const CapState* __restrict__ st,
           int32_t* __restrict__ eq_rank_keep) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int keep = 0;
  if (flags[i]) {
    const int z = (int)(skeys[i] & 0xffffu);
    keep = z > st->thr_z || (z == st->thr_z && eq_rank_keep[i] <= st->thr_take);
  }
  eq_rank_keep[i] = keep;
}

__global__ void __launch_bounds__(kCapThreads)
k_cap_emit(const uint64_t* __restrict__ skeys, const int32_t* __restrict__ svox, const int32_t* __restrict__ flags,
           const int32_t* __restrict__ out_rank, const double* __restrict__ vox_col, int64_t n, int64_t cap,
           const CapState* __restrict__ st, int16_t* __restrict__ points, double* __restrict__ colors,
           int32_t* __restrict__ n_out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t r = out_rank[i];
  if (i == n - 1) *n_out = st->error ? -st->error : r;
  const int32_t prev = i ? out_rank[i - 1] : 0;
  if (!flags[i] || r == prev || r > cap) return;       // not a kept head (or beyond the caller's capacity)
  const uint64_t k = skeys[i];
  const int64_t o = 3 * (int64_t)(r - 1);
  points[o + 0] = (int16_t)((int)((k >> 32) & 0xffffu) - 32768);
  points[o + 1] = (int16_t)((int)((k >> 16) & 0xffffu) - 32768);
  points[o + 2] = (int16_t)((int)(k & 0xffffu) - 32768);
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
  size_t b = 0;
  cub::DeviceScan::InclusiveSum(nullptr, b, (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
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
  b += 3 * align_up(n * 4, 256);                     // flags, rank, keep rank
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
  int32_t* keep = (int32_t*)w.take(n * 4);
  double* vox_col = (double*)w.take(n * 24);
  size_t sort_b = cap_sort_bytes(n), scan_b = cap_scan_bytes(n);
  void* sort_ws = w.take((int64_t)sort_b);
  void* scan_ws = w.take((int64_t)scan_b);
  if (!st || !zhist || !keys || !skeys || !ids || !sids || !flags || !rank || !keep || !vox_col || !sort_ws || !scan_ws) {
    set_error("pcc_capture_voxelize: workspace too small (%lld)", (long long)ws_bytes);
    return PCC_E_WORKSPACE;
  }
  const float4* px = (const float4*)xyzrgba;
  const int blocks = (int)((n + kCapThreads - 1) / kCapThreads);
  // state: minima = all ones (ordered-uint +inf side), the rest and the histogram zero
  PCC_CUDA(cudaMemsetAsync(st, 0, 256 + align_up(kZBins * 4, 256), stm));
  PCC_CUDA(cudaMemsetAsync(st, 0xff, 12, stm));
  k_cap_min<<<blocks, kCapThreads, 0, stm>>>(px, n, depth_clip, st);
  PCC_LAUNCHED();
  k_cap_keys<<<blocks, kCapThreads, 0, stm>>>(px, n, depth_clip, voxel_size, st, keys, ids);
  PCC_LAUNCHED();
  // stable: a voxel's pixels stay in input order
  PCC_CUDA(cub::DeviceRadixSort::SortPairs(sort_ws, sort_b, keys, skeys, ids, sids, (int)n, 0, 64, stm));
  g_launches.fetch_add(1);
  k_cap_heads<<<blocks, kCapThreads, 0, stm>>>(skeys, n, flags);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, flags, rank, (int)n, stm));
  g_launches.fetch_add(1);
  PCC_CUDA(cudaMemsetAsync(keys, 0xff, n * 8, stm));          // cell keys: padding beyond the voxel count
  k_cap_accumulate<<<blocks, kCapThreads, 0, stm>>>(px, skeys, sids, flags, rank, n, voxel_size, st, keys, ids, vox_col);
  PCC_LAUNCHED();
  // 48 key bits + the padding bit; stable: among voxels of one cell the lowest (ix,iy,iz) comes first
  PCC_CUDA(cub::DeviceRadixSort::SortPairs(sort_ws, sort_b, keys, skeys, ids, sids, (int)n, 0, 49, stm));
  g_launches.fetch_add(1);
  k_cap_cell_heads<<<blocks, kCapThreads, 0, stm>>>(skeys, n, flags, zhist);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, flags, rank, (int)n, stm));
  g_launches.fetch_add(1);
  k_cap_threshold<<<1, 1024, 0, stm>>>(zhist, rank, n, max_points, st);
  PCC_LAUNCHED();
  k_cap_eq<<<blocks, kCapThreads, 0, stm>>>(skeys, flags, n, st, keep);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, keep, keep, (int)n, stm));
  g_launches.fetch_add(1);
  k_cap_keep<<<blocks, kCapThreads, 0, stm>>>(skeys, flags, n, st, keep);
  PCC_LAUNCHED();
  PCC_CUDA(cub::DeviceScan::InclusiveSum(scan_ws, scan_b, keep, rank, (int)n, stm));
  g_launches.fetch_add(1);
  k_cap_emit<<<blocks, kCapThreads, 0, stm>>>(skeys, sids, flags, rank, vox_col, n, capacity, st, points, colors, n_out_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
