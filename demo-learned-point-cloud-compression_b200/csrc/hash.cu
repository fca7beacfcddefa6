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
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < cap) tk[i] = kEmptyKey;
}

__global__ void k_hash_insert(const uint64_t* __restrict__ keys, int64_t n, uint64_t* __restrict__ tk,
                              int32_t* __restrict__ tv, uint64_t cap_mask) {
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
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

__global__ void k_kmap_down2(const uint64_t* __restrict__ ckeys, const int32_t* __restrict__ parent,
                             int64_t n_child, int shift, int32_t* __restrict__ nbr, int64_t n_parent) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_child) return;
  int j = (int)((ckeys[i] >> shift) & 7ull);
  nbr[(int64_t)j * n_parent + parent[i]] = (int32_t)i;
}

}  // namespace pcc

using namespace pcc;

extern "C" {

int pcc_hash_build(const uint64_t* keys, int64_t n, uint64_t* table_keys, int32_t* table_vals,
                   int64_t capacity, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && capacity >= 8 && (capacity & (capacity - 1)) == 0 && capacity >= 2 * n);
  PCC_CHECK_ARG(table_keys && table_vals);
  cudaStream_t st = as_stream(s);
  k_hash_clear<<<(unsigned)((capacity + 255) / 256), 256, 0, st>>>(table_keys, capacity);
  PCC_LAUNCHED();
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys);
  k_hash_insert<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(keys, n, table_keys, table_vals,
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
  k_hash_query<<<(unsigned)((threads + 255) / 256), 256, 0, as_stream(s)>>>(
      queries, nq, table_keys, table_vals, (uint64_t)capacity - 1, rows);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_kmap_k3(const uint64_t* keys, int64_t n, int log2_stride, const uint64_t* table_keys,
                const int32_t* table_vals, int64_t capacity, int32_t* nbr, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && log2_stride >= 0 && log2_stride <= 14);
  PCC_CHECK_ARG(capacity >= 8 && (capacity & (capacity - 1)) == 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(keys && table_keys && table_vals && nbr);
  k_kmap_k3<<<(unsigned)((n + 63) / 64), 256, 0, as_stream(s)>>>(
      keys, n, 1 << log2_stride, table_keys, table_vals, (uint64_t)capacity - 1, nbr);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_kmap_down2(const uint64_t* child_keys, const int32_t* parent_row, int64_t n_child,
                   int log2_child_stride, int32_t* nbr, int64_t n_parent, pcc_stream_t s) {
  PCC_CHECK_ARG(n_child >= 0 && n_parent >= 0 && log2_child_stride >= 0 && log2_child_stride <= 14);
  if (n_parent == 0) return PCC_OK;
  PCC_CHECK_ARG(child_keys && parent_row && nbr);
  cudaStream_t st = as_stream(s);
  int64_t tot = 8 * n_parent;
  k_fill_i32<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(nbr, tot, -1);
  PCC_LAUNCHED();
  if (n_child == 0) return PCC_OK;
  k_kmap_down2<<<(unsigned)((n_child + 255) / 256), 256, 0, st>>>(
      child_keys, parent_row, n_child, 3 * log2_child_stride, nbr, n_parent);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
