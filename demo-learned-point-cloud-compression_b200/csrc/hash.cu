// (1) GPU coordinate hash + kernel maps.
//
// Table: open addressing over `capacity` (power of two) 64-bit key slots plus a parallel int32 value
// array.  Slots are grouped in windows of 4 (= one 32-byte sector of keys).  Lookups are cooperative:
// 4 lanes own one query, load the window's 4 keys in one sector transaction and vote with a ballot;
// a window that contains an empty slot terminates the probe (no deletions, inserts fill the first
// free slot of the earliest window).  Values are touched only on a hit.
#include "common.cuh"

namespace pcc {

constexpr int kWin = 4;

__global__ void k_hash_clear(uint64_t* __restrict__ tk, int64_t cap) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < cap) tk[i] = kEmptyKey;
}

__global__ void k_hash_insert(const uint64_t* __restrict__ keys, int64_t n, uint64_t* __restrict__ tk,
                              int32_t* __restrict__ tv, uint64_t cap_mask) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t key = keys[i];
  uint64_t base = hash_key(key) & cap_mask & ~(uint64_t)(kWin - 1);
  for (uint64_t probes = 0; probes <= cap_mask; probes += kWin) {
#pragma unroll
    for (int j = 0; j < kWin; ++j) {
      uint64_t slot = base + j;
      uint64_t prev = atomicCAS((unsigned long long*)&tk[slot], (unsigned long long)kEmptyKey,
                                (unsigned long long)key);
      if (prev == kEmptyKey || prev == key) {
        tv[slot] = (int32_t)i;
        return;
      }
    }
    base = (base + kWin) & cap_mask;
  }
}

// Cooperative probe: the 4 lanes of a group call this with the same key; returns row or -1 (same
// value in all 4 lanes).  `sub` = lane index within the group, `gshift` = bit offset of the group's
// nibble inside the warp ballot.
__device__ __forceinline__ int32_t probe4(uint64_t key, bool active, const uint64_t* __restrict__ tk,
                                          const int32_t* __restrict__ tv, uint64_t cap_mask, int sub,
                                          int gshift) {
  uint64_t base = hash_key(key) & cap_mask & ~(uint64_t)(kWin - 1);
  int32_t result = -1;
  bool done = !active;
  // All 32 lanes iterate together until every group is done (ballots need full participation).
  for (uint64_t probes = 0; probes <= cap_mask; probes += kWin) {
    uint64_t k = done ? kEmptyKey : __ldg(&tk[base + sub]);
    unsigned hit = __ballot_sync(0xffffffffu, !done && k == key);
    unsigned emp = __ballot_sync(0xffffffffu, !done && k == kEmptyKey);
    unsigned ghit = (hit >> gshift) & 0xfu;
    unsigned gemp = (emp >> gshift) & 0xfu;
    if (!done) {
      if (ghit) {
        int j = __ffs(ghit) - 1;
        result = __ldg(&tv[base + j]);
        done = true;
      } else if (gemp) {
        done = true;
      } else {
        base = (base + kWin) & cap_mask;
      }
    }
    if (__all_sync(0xffffffffu, done)) break;
  }
  return result;
}

__global__ void k_hash_query(const uint64_t* __restrict__ q, int64_t nq, const uint64_t* __restrict__ tk,
                             const int32_t* __restrict__ tv, uint64_t cap_mask,
                             int32_t* __restrict__ rows) {
  pdl_trigger();
  pdl_wait();
  int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t qi = t >> 2;
  int sub = threadIdx.x & 3;
  int gshift = (threadIdx.x & 31) & ~3;
  bool active = qi < nq;
  uint64_t key = active ? q[qi] : 0;
  int32_t r = probe4(key, active, tk, tv, cap_mask, sub, gshift);
  if (active && sub == 0) rows[qi] = r;
}

// One group of 4 lanes per output row, looping over the 27 offsets; a 256-thread block covers 64
// consecutive rows, so for a fixed offset each warp writes 8 consecutive nbr entries.
__global__ void __launch_bounds__(256)
k_kmap_k3(const uint64_t* __restrict__ keys, int64_t n, int step, const uint64_t* __restrict__ tk,
          const int32_t* __restrict__ tv, uint64_t cap_mask, int32_t* __restrict__ nbr) {
  pdl_trigger();
  pdl_wait();
  int64_t o = blockIdx.x * 64ll + (threadIdx.x >> 2);
  int sub = threadIdx.x & 3;
  int gshift = (threadIdx.x & 31) & ~3;
  bool rowok = o < n;
  int b = 0, x = 0, y = 0, z = 0;
  if (rowok) unpack_key(keys[o], b, x, y, z);
#pragma unroll 1
  for (int j = 0; j < 27; ++j) {
    int dx = (j % 3 - 1) * step, dy = ((j / 3) % 3 - 1) * step, dz = (j / 9 - 1) * step;
    int32_t r;
    if (j == 13) {
      r = rowok ? (int32_t)o : -1;
    } else {
      int nx = x + dx, ny = y + dy, nz = z + dz;
      bool inrange = rowok && nx >= -32768 && nx < 32768 && ny >= -32768 && ny < 32768 &&
                     nz >= -32768 && nz < 32768;
      uint64_t key = inrange ? pack_key(b, nx, ny, nz) : 0;
      r = probe4(key, inrange, tk, tv, cap_mask, sub, gshift);
    }
    if (rowok && sub == 0) nbr[(int64_t)j * n + o] = r;
  }
}

__global__ void k_fill_i32(int32_t* __restrict__ p, int64_t n, int32_t v) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

__global__ void k_kmap_down2(const uint64_t* __restrict__ ckeys, const int32_t* __restrict__ parent,
                             int64_t n_child, int shift, int32_t* __restrict__ nbr, int64_t ld) {
  pdl_trigger();
  pdl_wait();
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_child) return;
  int j = (int)((ckeys[i] >> shift) & 7ull);
  nbr[(int64_t)j * ld + parent[i]] = (int32_t)i;
}


// ---------------------------------------------------------------------------------------------
// Block table: the sorted key set cut into 4x4x4-voxel Morton blocks.  Rows of one block are
// contiguous (keys are sorted), so a block is described by its first row and a 64-bit occupancy
// mask; the row of voxel `idx` (6 Morton bits) is base + popc(mask & ((1 << idx) - 1)).
// The hash (same open addressing, windows of 4 keys = one 32-byte sector) is keyed on
// key >> (3*log2_stride + 6); a 3x3x3 neighbourhood touches at most 2x2x2 blocks, so the k=3 kernel
// map needs <= 8 probes per row instead of 26, all of them shared with the neighbouring rows
// (L1 hits), and the table has ~n/12 entries on surface data.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_blockmap_insert(const uint64_t* __restrict__ keys, int64_t n, int shift, uint64_t* __restrict__ tk,
                  ulonglong2* __restrict__ tv, uint64_t cap_mask, int* __restrict__ fault) {
  pdl_trigger();
  pdl_wait();
  // The rows of a block are consecutive, so the lanes of a warp that hold rows of one block find each other with a
  // match and OR their occupancy bits with one warp reduction; only a block that runs past the end of its warp is
  // finished row by row by the thread of its first row.
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const bool in = i < n;
  const uint64_t k = in ? keys[i] : 0ull;
  const uint64_t bkey = in ? k >> (shift + 6) : kEmptyKey;     // rows past the end form a group of their own
  const unsigned peers = __match_any_sync(0xffffffffu, bkey);
  const uint32_t idx = (uint32_t)((k >> shift) & 63ull);
  const uint32_t lo = __reduce_or_sync(peers, idx < 32u ? 1u << idx : 0u);
  const uint32_t hi = __reduce_or_sync(peers, idx >= 32u ? 1u << (idx - 32u) : 0u);
  if (!in || lane != __ffs(peers) - 1) return;                 // one lane per group goes on
  if (lane == 0 && i > 0 && (keys[i - 1] >> (shift + 6)) == bkey) return;   // the block started in the previous warp
  uint64_t mask = (uint64_t)lo | ((uint64_t)hi << 32);
  if (peers >> 31) {                                           // the group reaches the end of the warp: walk on
    for (int64_t j = i + __popc(peers); j < n && j < i + 64; ++j) {
      const uint64_t kj = keys[j];
      if ((kj >> (shift + 6)) != bkey) break;
      mask |= 1ull << ((kj >> shift) & 63ull);
    }
  }
  uint64_t base = hash_key(bkey) & cap_mask & ~(uint64_t)(kWin - 1);
  for (uint64_t probes = 0; probes <= cap_mask; probes += kWin) {
#pragma unroll
    for (int j = 0; j < kWin; ++j) {
      const uint64_t slot = base + j;
      const uint64_t prev = atomicCAS((unsigned long long*)&tk[slot], (unsigned long long)kEmptyKey,
                                      (unsigned long long)bkey);
      if (prev == kEmptyKey) {   // block keys are unique: no "already present" case
        tv[slot] = make_ulonglong2(mask, (unsigned long long)i);
        return;
      }
    }
    base = (base + kWin) & cap_mask;
  }
  // table full: the caller's capacity was below the number of occupied blocks.  The block stays out of the table
  // (its rows will look absent to the map kernel) and the host is told through the fault word.
  atomicMax(fault, PCC_FAULT_BLOCKMAP_FULL);
}

struct BlockEntry {
  uint64_t mask;
  int32_t base;
};

__device__ __forceinline__ BlockEntry block_lookup(uint64_t bkey, const uint64_t* __restrict__ tk,
                                                   const ulonglong2* __restrict__ tv, uint64_t cap_mask) {
  BlockEntry e;
  e.mask = 0;
  e.base = 0;
  uint64_t base = hash_key(bkey) & cap_mask & ~(uint64_t)(kWin - 1);
  for (uint64_t probes = 0; probes <= cap_mask; probes += kWin) {
    const ulonglong2 k01 = __ldg(reinterpret_cast<const ulonglong2*>(tk + base));
    const ulonglong2 k23 = __ldg(reinterpret_cast<const ulonglong2*>(tk + base + 2));
    int hit = -1;
    if (k01.x == bkey) hit = 0;
    else if (k01.y == bkey) hit = 1;
    else if (k23.x == bkey) hit = 2;
    else if (k23.y == bkey) hit = 3;
    if (hit >= 0) {
      const ulonglong2 v = __ldg(tv + base + hit);
      e.mask = v.x;
      e.base = (int32_t)v.y;
      return e;
    }
    if (k01.x == kEmptyKey || k01.y == kEmptyKey || k23.x == kEmptyKey || k23.y == kEmptyKey) return e;
    base = (base + kWin) & cap_mask;
  }
  return e;
}

// 6-bit Morton index of (x,y,z) low 2 bits each: x0 y0 z0 x1 y1 z1 -> bits 0..5
__device__ __forceinline__ int morton2(int x, int y, int z) {
  return (x & 1) | ((y & 1) << 1) | ((z & 1) << 2) | ((x & 2) << 2) | ((y & 2) << 3) | ((z & 2) << 4);
}

// One thread per output row.  The (up to) 2x2x2 block entries around the row go to shared memory as
// {mask lo, mask hi, first row, first row + popc(lo)} so that each of the 26 neighbours costs one
// LDS.128 with a computed index plus a 32-bit popcount, instead of an eight-way select tree over
// 64-bit registers (the register version was instruction bound: 2120 SASS instructions, 106
// registers, 200 us for 1.79 M rows whatever the table size).
constexpr int kKmapThreads = 256;

__global__ void __launch_bounds__(kKmapThreads)
k_kmap_k3_blocks(const uint64_t* __restrict__ keys, int64_t n, int l, const uint64_t* __restrict__ tk,
                 const ulonglong2* __restrict__ tv, uint64_t cap_mask, int32_t* __restrict__ nbr,
                 int64_t ld) {
  pdl_trigger();
  pdl_wait();
  __shared__ uint4 ent[8][kKmapThreads];
  const int tid = threadIdx.x;
  const int64_t o = blockIdx.x * (int64_t)blockDim.x + tid;
  if (o >= n) return;
  int b, x, y, z;
  unpack_key(keys[o], b, x, y, z);
  // voxel coordinates in units of the stride, biased to be non-negative (17 bits incl. the +-1 halo)
  const int ux = (x + 32768) >> l, uy = (y + 32768) >> l, uz = (z + 32768) >> l;
  const int bmax = ((65536 >> l) - 1) >> 2;
  const int lx = (ux - 1) >> 2, hx = (ux + 1) >> 2;
  const int ly = (uy - 1) >> 2, hy = (uy + 1) >> 2;
  const int lz = (uz - 1) >> 2, hz = (uz + 1) >> 2;
  const uint64_t fbits = ((uint64_t)(uint32_t)b << 48) >> (3 * l + 6);
  // Morton bits of the two candidate block coordinates per axis, spread once (not once per corner)
  const uint64_t sx[2] = {spread3_16((uint32_t)lx), spread3_16((uint32_t)hx)};
  const uint64_t sy[2] = {spread3_16((uint32_t)ly) << 1, spread3_16((uint32_t)hy) << 1};
  const uint64_t sz[2] = {spread3_16((uint32_t)lz) << 2, spread3_16((uint32_t)hz) << 2};
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int ax = c & 1, ay = (c >> 1) & 1, az = c >> 2;
    // duplicates (lo == hi along an axis) reuse the entry already fetched
    if (ax && hx == lx) { ent[c][tid] = ent[c ^ 1][tid]; continue; }
    if (ay && hy == ly) { ent[c][tid] = ent[c ^ 2][tid]; continue; }
    if (az && hz == lz) { ent[c][tid] = ent[c ^ 4][tid]; continue; }
    const int bx = ax ? hx : lx, by = ay ? hy : ly, bz = az ? hz : lz;
    uint4 e = make_uint4(0u, 0u, 0u, 0u);
    if (bx >= 0 && by >= 0 && bz >= 0 && bx <= bmax && by <= bmax && bz <= bmax) {
      // block key = pack_key(b, block corner) >> (3l + 6) = frame bits | morton(bx, by, bz)
      const uint64_t bkey = fbits | sx[ax] | sy[ay] | sz[az];
      const BlockEntry be = block_lookup(bkey, tk, tv, cap_mask);
      const uint32_t lo = (uint32_t)be.mask, hi = (uint32_t)(be.mask >> 32);
      e = make_uint4(lo, hi, (uint32_t)be.base, (uint32_t)be.base + (uint32_t)__popc(lo));
    }
    ent[c][tid] = e;
  }
  // per axis and displacement: which of the two blocks (bit of the entry index) and the Morton bits of
  // the in-block position (x -> bits 0,3; y -> 1,4; z -> 2,5)
  int cs[3][3], ms[3][3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const int nx = ux + d - 1, ny = uy + d - 1, nz = uz + d - 1;
    cs[0][d] = ((nx >> 2) != lx) ? 1 : 0;
    cs[1][d] = ((ny >> 2) != ly) ? 2 : 0;
    cs[2][d] = ((nz >> 2) != lz) ? 4 : 0;
    ms[0][d] = (nx & 1) | ((nx & 2) << 2);
    ms[1][d] = ((ny & 1) << 1) | ((ny & 2) << 3);
    ms[2][d] = ((nz & 1) << 2) | ((nz & 2) << 4);
  }
#pragma unroll
  for (int j = 0; j < 27; ++j) {
    const int dx = j % 3, dy = (j / 3) % 3, dz = j / 9;
    int32_t r = -1;
    if (j == 13) {
      r = (int32_t)o;
    } else {
      const int c = cs[0][dx] | cs[1][dy] | cs[2][dz];
      const int idx = ms[0][dx] | ms[1][dy] | ms[2][dz];
      const uint4 e = ent[c][tid];
      const bool up = idx >= 32;
      const uint32_t w = up ? e.y : e.x;
      const uint32_t first = up ? e.w : e.z;
      const uint32_t i5 = (uint32_t)idx & 31u;
      if ((w >> i5) & 1u) r = (int32_t)(first + (uint32_t)__popc(w & ((1u << i5) - 1u)));
    }
    nbr[(int64_t)j * ld + o] = r;
  }
}

// number of kernel-map entries that name a row (the P of SURVEY 8d's flop count 2 P Cin Cout)
__global__ void k_kmap_count_pairs(const int32_t* __restrict__ nbr, int64_t ld, int64_t n_out, int kvol,
                                   unsigned long long* __restrict__ count) {
  pdl_trigger();
  pdl_wait();
  unsigned long long c = 0;
  const int64_t total = (int64_t)kvol * n_out;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = i / n_out, o = i - j * n_out;
    c += nbr[j * ld + o] >= 0;
  }
  c = __reduce_add_sync(0xffffffffu, (unsigned)c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}

}  // namespace pcc

using namespace pcc;

extern "C" {

int pcc_hash_build(const uint64_t* keys, int64_t n, uint64_t* table_keys, int32_t* table_vals,
                   int64_t capacity, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && capacity >= 8 && (capacity & (capacity - 1)) == 0 && capacity >= 2 * n);
  PCC_CHECK_ARG(table_keys && table_vals);
  cudaStream_t st = as_stream(s);
  launch_pdl(k_hash_clear, dim3((unsigned)((capacity + 255) / 256)), dim3(256), 0, st, table_keys, capacity);
  PCC_LAUNCHED();
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys);
  launch_pdl(k_hash_insert, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, keys, n, table_keys, table_vals,
                                                            (uint64_t)capacity - 1);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_hash_query(const uint64_t* queries, int64_t nq, const uint64_t* table_keys,
                   const int32_t* table_vals, int64_t capacity, int32_t* rows, pcc_stream_t s) {
  PCC_CHECK_ARG(nq >= 0 && capacity >= 8 && (capacity & (capacity - 1)) == 0);
  if (nq == 0) return PCC_OK;
  PCC_CHECK_ARG(queries && table_keys && table_vals && rows);
  int64_t threads = nq * 4;
  launch_pdl(k_hash_query, dim3((unsigned)((threads + 255) / 256)), dim3(256), 0, as_stream(s), queries, nq, table_keys, table_vals, (uint64_t)capacity - 1, rows);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_kmap_k3(const uint64_t* keys, int64_t n, int log2_stride, const uint64_t* table_keys,
                const int32_t* table_vals, int64_t capacity, int32_t* nbr, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && log2_stride >= 0 && log2_stride <= 14);
  PCC_CHECK_ARG(capacity >= 8 && (capacity & (capacity - 1)) == 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys && table_keys && table_vals && nbr);
  launch_pdl(k_kmap_k3, dim3((unsigned)((n + 63) / 64)), dim3(256), 0, as_stream(s), keys, n, 1 << log2_stride, table_keys, table_vals, (uint64_t)capacity - 1, nbr);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_kmap_down2(const uint64_t* child_keys, const int32_t* parent_row, int64_t n_child,
                   int log2_child_stride, int32_t* nbr, int64_t n_parent, int64_t nbr_ld,
                   pcc_stream_t s) {
  PCC_CHECK_ARG(n_child >= 0 && n_parent >= 0 && log2_child_stride >= 0 && log2_child_stride <= 14);
  PCC_CHECK_ARG(nbr_ld >= n_parent);
  if (n_parent == 0) return PCC_OK;
  PCC_CHECK_ARG(child_keys && parent_row && nbr);
  cudaStream_t st = as_stream(s);
  int64_t tot = 8 * nbr_ld;
  launch_pdl(k_fill_i32, dim3((unsigned)((tot + 255) / 256)), dim3(256), 0, st, nbr, tot, -1);
  PCC_LAUNCHED();
  if (n_child == 0) return PCC_OK;
  launch_pdl(k_kmap_down2, dim3((unsigned)((n_child + 255) / 256)), dim3(256), 0, st, child_keys, parent_row, n_child, 3 * log2_child_stride, nbr, nbr_ld);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_blockmap_build(const uint64_t* keys, int64_t n, int log2_stride, uint64_t* table_keys,
                       void* table_vals16, int64_t capacity, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && log2_stride >= 0 && log2_stride <= 13);
  PCC_CHECK_ARG(capacity >= 8 && (capacity & (capacity - 1)) == 0);
  PCC_CHECK_ARG(table_keys && table_vals16 && ((uintptr_t)table_vals16 % 16) == 0 &&
                ((uintptr_t)table_keys % 32) == 0);
  cudaStream_t st = as_stream(s);
  launch_pdl(k_hash_clear, dim3((unsigned)((capacity + 255) / 256)), dim3(256), 0, st, table_keys, capacity);
  PCC_LAUNCHED();
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys);
  int* fault = fault_word();
  if (!fault) return PCC_E_CUDA;
  launch_pdl(k_blockmap_insert, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, keys, n, 3 * log2_stride, table_keys,
                                                                (ulonglong2*)table_vals16,
                                                                (uint64_t)capacity - 1, fault);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_kmap_k3_blocks(const uint64_t* keys, int64_t n, int log2_stride, const uint64_t* table_keys,
                       const void* table_vals16, int64_t capacity, int32_t* nbr, int64_t nbr_ld,
                       pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && log2_stride >= 0 && log2_stride <= 13 && nbr_ld >= n);
  PCC_CHECK_ARG(capacity >= 8 && (capacity & (capacity - 1)) == 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys && table_keys && table_vals16 && nbr);
  launch_pdl(k_kmap_k3_blocks, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, as_stream(s), keys, n, log2_stride, table_keys, (const ulonglong2*)table_vals16, (uint64_t)capacity - 1, nbr,
      nbr_ld);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_kmap_count_pairs(const int32_t* nbr, int64_t nbr_ld, int64_t n_out, int kvol, int64_t* count_dev, pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0 && kvol >= 1 && nbr_ld >= n_out && count_dev);
  cudaStream_t st = as_stream(s);
  PCC_CUDA(cudaMemsetAsync(count_dev, 0, 8, st));
  if (n_out == 0) return PCC_OK;
  PCC_CHECK_ARG(nbr);
  launch_pdl(k_kmap_count_pairs, dim3(592), dim3(256), 0, st, nbr, nbr_ld, n_out, kvol, (unsigned long long*)count_dev);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
