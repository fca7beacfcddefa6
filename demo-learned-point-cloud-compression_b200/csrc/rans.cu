// (4) Interleaved multi-lane rANS ("PCL1" container).
//
// The symbol array is cut into segments of seg_size symbols; one warp codes one segment as 32
// interleaved lanes (symbol i of the segment -> lane i % 32, round i / 32).  Each lane keeps a 32-bit
// rANS state (L = 2^16, 16-bit probabilities, 16-bit renormalisation words).  Renormalisation words of
// all lanes of a warp go to ONE shared word stream: in every coding step the lanes that must emit are
// found with a ballot and placed with a popc prefix, so the decoder (which runs the steps in the
// opposite order and reads the stream backwards) can recompute every position without side
// information.  Symbol semantics are CompressAI's encode_with_indexes: values outside the table's
// support take the table's last entry as an escape and the residual follows as a count nibble plus up
// to 8 four-bit nibbles coded with a flat 2^12/2^16 model.
//
// Container:  "PCL1" | u32 n | u32 seg_size | u32 nseg | u32 seg_bytes[nseg] | segments...
// Segment:    u32 nwords | u32 state[32] | u16 words[nwords] | pad to 4 bytes
#include "common.cuh"

namespace pcc {

constexpr uint32_t kRansL = 1u << 16;
constexpr uint32_t kMagic = 0x314c4350u;  // "PCL1" little endian

struct Tables {
  const uint32_t* cdf;
  const int32_t* off;
  const int32_t* size;
  const int32_t* offset;
  const uint16_t* lut;   // [T][kLutN + 1] or null
  int32_t T;
  int32_t total;         // cdf entries over all tables (0 = unknown)
};

constexpr int kLutBits = 8, kLutN = 1 << kLutBits, kLutShift = 16 - kLutBits;

// q = x / fr, rem = x % fr for x < fr * 2^16 (so q < 2^16): one float multiply gives q within +-1
// (relative error 2^-22 on a quotient below 2^16), one correction step makes it exact.
__device__ __forceinline__ void divmod_small_q(uint32_t x, uint32_t fr, uint32_t& q, uint32_t& rem) {
  float rc;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(__uint2float_rn(fr)));
  q = __float2uint_rz(__uint2float_rn(x) * rc);
  int32_t r = (int32_t)(x - q * fr);
  // branch-free correction by at most one in either direction
  const int32_t neg = r >> 31;                 // -1 when the estimate was one too high
  q += (uint32_t)neg;
  r += neg & (int32_t)fr;
  const uint32_t over = (uint32_t)r >= fr ? 1u : 0u;
  q += over;
  r -= (int32_t)(over * fr);
  rem = (uint32_t)r;
}

__host__ __device__ inline int64_t seg_cap_words(int32_t seg_size) { return 4ll * seg_size + 64; }

struct SymOps {
  int32_t coff;   // start of the cdf row
  int32_t v;      // table symbol (maxv = escape)
  uint32_t raw;   // escape payload
  int nnib;       // nibbles in raw
  int nops;       // 1 or 2 + nnib
};

__device__ __forceinline__ SymOps make_ops(int32_t symbol, int32_t t, const Tables& T) {
  SymOps o;
  o.coff = T.off[t];
  int32_t maxv = T.size[t] - 2;
  int64_t v = (int64_t)symbol - (int64_t)T.offset[t];
  o.raw = 0;
  if (v < 0) {
    o.raw = (uint32_t)(-2 * v - 1);
    v = maxv;
  } else if (v >= maxv) {
    o.raw = (uint32_t)(2 * (v - maxv));
    v = maxv;
  }
  o.v = (int32_t)v;
  o.nnib = 0;
  o.nops = 1;
  if (v == maxv) {
    o.nnib = o.raw ? (35 - __clz(o.raw)) / 4 : 0;  // ceil(bits/4)
    o.nops = 2 + o.nnib;
  }
  return o;
}

template <typename T>
__device__ __forceinline__ T* ws_shift(T* p, int64_t ws_stride) {
  return reinterpret_cast<T*>(reinterpret_cast<char*>(p) + (int64_t)blockIdx.y * ws_stride);
}
template <typename T>
__device__ __forceinline__ const T* ws_shift(const T* p, int64_t ws_stride) {
  return reinterpret_cast<const T*>(reinterpret_cast<const char*>(p) + (int64_t)blockIdx.y * ws_stride);
}

// Pass A (fully parallel): per symbol the packed (start | freq << 16) of the table entry it codes
// with -- its own entry, or the table's last entry for a value outside the support (escape) -- the
// escape payload raw[i] (0 when not an escape) and, per 32 symbols (= one coder round), the ballot of
// escaping lanes.  The coder (pass B) then never touches the tables or the symbols.
__global__ void __launch_bounds__(256)
k_rans_prepare(const int32_t* __restrict__ sym, const uint16_t* __restrict__ idx, int64_t chan_len, int64_t n,
               Tables T, uint32_t* __restrict__ packed, uint32_t* __restrict__ raw, uint32_t* __restrict__ escmask,
               int64_t ws_stride) {
  pdl_trigger();
  pdl_wait();
  // blockIdx.y = stream of a batch (pcc_rans_encode_batch): its symbols follow the previous stream's, its
  // workspace arrays lie ws_stride bytes after the previous stream's
  sym += blockIdx.y * n;
  if (idx) idx += blockIdx.y * n;
  packed = ws_shift(packed, ws_stride);
  raw = ws_shift(raw, ws_stride);
  escmask = ws_shift(escmask, ws_stride);
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  uint32_t pk = 0, rw = 0;
  bool esc = false;
  if (i < n) {
    int32_t t = idx ? (int32_t)idx[i] : (int32_t)(i / chan_len);
    t = t < T.T ? t : T.T - 1;
    const int32_t coff = T.off[t], maxv = T.size[t] - 2;
    int64_t v = (int64_t)sym[i] - (int64_t)T.offset[t];
    if (v < 0) {
      rw = (uint32_t)(-2 * v - 1);
      v = maxv;
      esc = true;
    } else if (v >= maxv) {
      rw = (uint32_t)(2 * (v - maxv));
      v = maxv;
      esc = true;
    }
    const uint32_t st = T.cdf[coff + v];
    const uint32_t fr = T.cdf[coff + v + 1] - st;
    pk = st | (fr << 16);
    packed[i] = pk;
    raw[i] = rw;
  }
  const unsigned m = __ballot_sync(0xffffffffu, esc);
  if ((threadIdx.x & 31) == 0 && i < n) escmask[i >> 5] = m;
}

constexpr int kEncWarps = 4;      // segments per CTA
constexpr int kChunk = 8;         // decoder: rounds whose table indexes are prefetched together
// encoder: three words per round are prefetched (entry, escape payload, escape ballot), double buffered.  With 8 rounds
// per chunk those 48 registers pushed the kernel to 72 registers and one spilled value -- the store of a just-loaded
// word to local memory waited for its load, 14 % of the kernel's samples, and the chunk's first use another 12 %
// (profiles/r2y_rans_encode_ncu.txt).  Four rounds still lead the loads by ~2500 cycles.
constexpr int kEncChunk = 4;
constexpr uint32_t kInactive = 0xffffffffu;

// Pass B: one warp per segment, rounds in reverse.  State update only; pass A's per-symbol words
// arrive through a double-buffered register prefetch (escape payloads only for escaping lanes).
__global__ void __launch_bounds__(32 * kEncWarps)
k_rans_encode(const uint32_t* __restrict__ packed, const uint32_t* __restrict__ raw,
              const uint32_t* __restrict__ escmask, int64_t n, int32_t seg_size, int32_t nseg,
              uint16_t* __restrict__ scratch, uint32_t* __restrict__ states, int32_t* __restrict__ nwords,
              int64_t ws_stride) {
  pdl_trigger();
  pdl_wait();
  packed = ws_shift(packed, ws_stride);
  raw = ws_shift(raw, ws_stride);
  escmask = ws_shift(escmask, ws_stride);
  scratch = ws_shift(scratch, ws_stride);
  states = ws_shift(states, ws_stride);
  nwords = ws_shift(nwords, ws_stride);
  const int seg = blockIdx.x * kEncWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (seg >= nseg) return;
  const unsigned lt = (1u << lane) - 1u;
  const int64_t start = (int64_t)seg * seg_size;
  const int32_t cnt = (int32_t)((n - start) < seg_size ? (n - start) : seg_size);
  const int32_t rounds = (cnt + 31) / 32;
  uint16_t* out = scratch + (int64_t)seg * seg_cap_words(seg_size);
  const uint32_t* pk_seg = packed + start;
  const uint32_t* raw_seg = raw + start;
  const uint32_t* em_seg = escmask + (start >> 5);
  uint32_t x = kRansL;
  int32_t wpos = 0;
  uint32_t cur[kEncChunk], nxt[kEncChunk], cur_raw[kEncChunk], nxt_raw[kEncChunk], cur_em[kEncChunk], nxt_em[kEncChunk];
  auto load_chunk = [&](int32_t rb, uint32_t (&buf)[kEncChunk], uint32_t (&braw)[kEncChunk], uint32_t (&bem)[kEncChunk]) {
#pragma unroll
    for (int k = 0; k < kEncChunk; ++k) {   // rounds rb-1 ... rb-kEncChunk
      const int32_t r = rb - 1 - k;
      const int32_t li = r * 32 + lane;
      bem[k] = r >= 0 ? __ldg(em_seg + r) : 0u;
      buf[k] = (r >= 0 && li < cnt) ? __ldg(pk_seg + li) : kInactive;
      braw[k] = ((bem[k] >> lane) & 1u) ? __ldg(raw_seg + li) : 0u;
    }
  };
  load_chunk(rounds, cur, cur_raw, cur_em);
  for (int32_t rb = rounds; rb > 0; rb -= kEncChunk) {
    load_chunk(rb - kEncChunk, nxt, nxt_raw, nxt_em);
#pragma unroll
    for (int k = 0; k < kEncChunk; ++k) {
      const int32_t r = rb - 1 - k;
      if (r < 0) break;
      const uint32_t pk = cur[k];
      const bool active = pk != kInactive;
      if (cur_em[k] != 0u) {
        // ---- a lane escapes: its payload goes first (reverse order): nibbles, then their count;
        // flat 2^12 / 2^16 model, so the division is a shift
        const bool esc = (cur_em[k] >> lane) & 1u;
        const uint32_t rw = cur_raw[k];
        const int nnib = (esc && rw) ? (35 - __clz(rw)) / 4 : 0;   // ceil(bits / 4)
        const int nops = esc ? 1 + nnib : 0;                       // count nibble + payload nibbles
        const int maxops = __reduce_max_sync(0xffffffffu, nops);
        for (int ts = maxops; ts >= 1; --ts) {
          const bool has = ts <= nops;
          const uint32_t val = ts == 1 ? (uint32_t)nnib : ((rw >> (4 * (ts - 2))) & 15u);
          const bool emit = has && x >= (1u << 28);
          const unsigned m = __ballot_sync(0xffffffffu, emit);
          if (emit) {
            out[wpos + __popc(m & lt)] = (uint16_t)(x & 0xffffu);
            x >>= 16;
          }
          wpos += __popc(m);
          if (has) x = ((x >> 12) << 16) + (x & 4095u) + (val << 12);
        }
      }
      {
        // ---- table operation (own entry, or the escape entry) of every active lane
        const uint32_t fr = active ? (pk >> 16) : 1u, st = pk & 0xffffu;
        const bool emit = active && x >= (fr << 16);
        const unsigned m = __ballot_sync(0xffffffffu, emit);
        if (emit) {
          out[wpos + __popc(m & lt)] = (uint16_t)(x & 0xffffu);
          x >>= 16;
        }
        wpos += __popc(m);
        if (active) {
          uint32_t q, rem;
          divmod_small_q(x, fr, q, rem);
          x = (q << 16) + rem + st;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < kEncChunk; ++k) {
      cur[k] = nxt[k];
      cur_raw[k] = nxt_raw[k];
      cur_em[k] = nxt_em[k];
    }
  }
  states[seg * 32 + lane] = x;
  if (lane == 0) nwords[seg] = wpos;
}

__device__ __forceinline__ uint32_t seg_bytes_of(int32_t nw) { return 4u + 128u + ((2u * (uint32_t)nw + 3u) & ~3u); }

__global__ void __launch_bounds__(256)
k_rans_pack(const uint16_t* __restrict__ scratch, const uint32_t* __restrict__ states,
            const int32_t* __restrict__ nwords, int32_t nseg, int32_t seg_size, uint32_t n,
            uint8_t* __restrict__ out, int32_t* __restrict__ out_bytes, int64_t ws_stride, int64_t out_stride) {
  pdl_trigger();
  pdl_wait();
  scratch = ws_shift(scratch, ws_stride);
  states = ws_shift(states, ws_stride);
  nwords = ws_shift(nwords, ws_stride);
  out += (int64_t)blockIdx.y * out_stride;
  out_bytes += blockIdx.y;
  __shared__ uint32_t s_off;
  __shared__ uint32_t red[256];
  const int seg = blockIdx.x, tid = threadIdx.x;
  uint32_t acc = 0;
  for (int i = tid; i < seg; i += 256) acc += seg_bytes_of(nwords[i]);
  red[tid] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (tid < s) red[tid] += red[tid + s];
    __syncthreads();
  }
  if (tid == 0) s_off = red[0];
  __syncthreads();
  const uint32_t hdr = 16u + 4u * (uint32_t)nseg;
  uint32_t* o32 = reinterpret_cast<uint32_t*>(out);
  const int32_t nw = nwords[seg];
  const uint32_t sb = seg_bytes_of(nw);
  if (tid == 0) o32[4 + seg] = sb;
  if (seg == 0 && tid == 0) {
    o32[0] = kMagic;
    o32[1] = n;
    o32[2] = (uint32_t)seg_size;
    o32[3] = (uint32_t)nseg;
  }
  if (seg == nseg - 1 && tid == 0) *out_bytes = (int32_t)(hdr + s_off + sb);
  uint8_t* dst = out + hdr + s_off;
  uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
  if (tid == 0) d32[0] = (uint32_t)nw;
  if (tid < 32) d32[1 + tid] = states[seg * 32 + tid];
  uint16_t* dw = reinterpret_cast<uint16_t*>(dst + 132);
  const uint16_t* src = scratch + (int64_t)seg * seg_cap_words(seg_size);
  for (int i = tid; i < nw; i += 256) dw[i] = src[i];
  if (tid == 0 && (nw & 1)) dw[nw] = 0;
}

// lut[t][b] = largest s in [0, maxv] with cdf_t[s] <= b << kLutShift (b = 0..kLutN-1), lut[t][kLutN] = maxv:
// the symbol of cumulative frequency cf lies in [lut[t][cf >> kLutShift], lut[t][(cf >> kLutShift) + 1]].
__global__ void k_rans_build_lut(Tables T, uint16_t* __restrict__ lut) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.x;
  const int32_t coff = T.off[t], maxv = T.size[t] - 2;
  for (int b = threadIdx.x; b <= kLutN; b += blockDim.x) {
    int32_t lo = 0;
    if (b == kLutN) {
      lo = maxv;
    } else {
      const uint32_t target = (uint32_t)b << kLutShift;
      int32_t hi = maxv + 1;
      while (hi - lo > 1) {
        const int32_t mid = (lo + hi) >> 1;
        if (T.cdf[coff + mid] <= target) lo = mid; else hi = mid;
      }
    }
    lut[(int64_t)t * (kLutN + 1) + b] = (uint16_t)lo;
  }
}

constexpr int kDecWarps = 4;      // segments per CTA
constexpr int kWordCap = 8192;    // renormalisation words of a segment staged in shared memory (16 KB per warp)

// SMEM: the CTA first copies the tables into shared memory (CDF entries as uint16 -- only the
// terminating 65536 of a row does not fit and is implied --, per-table {offset, last symbol, symbol
// offset} and the search windows), so the dependent look-ups of a round cost shared-memory latency.
// HAS_IDX: the table of a symbol comes from the index array (Gaussian conditional); without it the table is the symbol's
// channel (factorized prior).  A template parameter: as a run-time test it sat, with its 64-bit channel arithmetic, in
// every prefetch slot of the one instruction stream a segment's time consists of.
template <bool SMEM, bool HAS_IDX>
__global__ void __launch_bounds__(32 * kDecWarps)
k_rans_decode(const uint8_t* __restrict__ container, int64_t container_bytes, int32_t seg_size_arg,
              const uint16_t* __restrict__ idx, int64_t chan_len, int64_t n, int32_t nseg_arg, Tables T,
              int32_t* __restrict__ sym, int32_t* __restrict__ status) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ __align__(16) uint8_t dec_smem[];
  int4* s_meta = reinterpret_cast<int4*>(dec_smem);                                  // [T]
  uint16_t* s_lut = reinterpret_cast<uint16_t*>(dec_smem + (size_t)T.T * 16);        // [T][kLutN+1]
  uint16_t* s_cdf = s_lut + (((size_t)T.T * (kLutN + 1) + 7) & ~(size_t)7);          // [total]
  // per-warp staging of the segment's word stream (after the tables, 16-byte aligned)
  uint16_t* s_words = SMEM ? s_cdf + (((size_t)T.total + 7) & ~(size_t)7) : reinterpret_cast<uint16_t*>(dec_smem);
  s_words += (size_t)(threadIdx.x >> 5) * kWordCap;
  if (SMEM) {
    for (int t = threadIdx.x; t < T.T; t += blockDim.x)
      s_meta[t] = make_int4(T.off[t], T.size[t] - 2, T.offset[t], 0);
    {   // search windows: 16-byte copies (the 2-byte loop was 8 % of the kernel's time)
      const int nl = T.T * (kLutN + 1);
      const int nv = (reinterpret_cast<uintptr_t>(T.lut) & 15) == 0 ? nl >> 3 : 0;
      const uint4* l4 = reinterpret_cast<const uint4*>(T.lut);
      uint4* sl4 = reinterpret_cast<uint4*>(s_lut);
      for (int i = threadIdx.x; i < nv; i += blockDim.x) sl4[i] = __ldg(l4 + i);
      for (int i = (nv << 3) + threadIdx.x; i < nl; i += blockDim.x) s_lut[i] = T.lut[i];
    }
    // four entries per 16-byte load, batches of four independent loads in flight per thread
    const uint4* c4 = reinterpret_cast<const uint4*>(T.cdf);
    uint2* s4 = reinterpret_cast<uint2*>(s_cdf);
    const int nq = (reinterpret_cast<uintptr_t>(T.cdf) & 15) == 0 ? T.total >> 2 : 0;   // unaligned table: scalar copy
    for (int i0 = threadIdx.x; i0 < nq; i0 += 4 * blockDim.x) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * (int)blockDim.x < nq) v[u] = __ldg(c4 + i0 + u * blockDim.x);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * (int)blockDim.x < nq)
          s4[i0 + u * blockDim.x] = make_uint2((v[u].x & 0xffffu) | (v[u].y << 16), (v[u].z & 0xffffu) | (v[u].w << 16));
    }
    for (int i = (nq << 2) + threadIdx.x; i < T.total; i += blockDim.x) s_cdf[i] = (uint16_t)T.cdf[i];
    __syncthreads();
  }
  auto cdf_at = [&](int32_t coff, int32_t k) -> uint32_t {
    return SMEM ? (uint32_t)s_cdf[coff + k] : __ldg(T.cdf + coff + k);
  };
  const int seg = blockIdx.x * kDecWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (seg >= nseg_arg) return;
  const unsigned lt = (1u << lane) - 1u;
  const uint32_t* h32 = reinterpret_cast<const uint32_t*>(container);
  const uint32_t seg_size = h32[2], nseg = h32[3];
  if (h32[0] != kMagic || h32[1] != (uint32_t)n || seg_size != (uint32_t)seg_size_arg || nseg != (uint32_t)nseg_arg) {
    if (lane == 0) atomicExch(status, 3);
    return;
  }
  uint32_t acc = 0;
  for (int i = lane; i < seg; i += 32) acc += h32[4 + i];
  acc = __reduce_add_sync(0xffffffffu, acc);
  const uint64_t seg_off = 16ull + 4ull * nseg + acc;
  // the segment sizes come from the (untrusted) stream: a segment that does not start on a 4-byte boundary or
  // past the end is a damaged stream, never an address to dereference
  if ((seg_off & 3ull) != 0 || seg_off + 132 > (uint64_t)container_bytes) {
    if (lane == 0) atomicExch(status, 1);
    return;
  }
  const uint32_t* s32 = reinterpret_cast<const uint32_t*>(container + seg_off);
  int32_t wpos = (int32_t)s32[0];
  if (seg_off + 132 + 2ull * (uint32_t)wpos > (uint64_t)container_bytes) {
    if (lane == 0) atomicExch(status, 1);
    return;
  }
  uint32_t x = s32[1 + lane];
  const uint16_t* words = reinterpret_cast<const uint16_t*>(container + seg_off + 132);
  if (wpos <= kWordCap) {
    // stage the word stream in shared memory: the per-round word fetch is on the critical path
    const uint32_t* w32 = reinterpret_cast<const uint32_t*>(words);   // seg_off + 132 is 4-byte aligned
    uint32_t* sw32 = reinterpret_cast<uint32_t*>(s_words);
    for (int i = lane; i < (wpos + 1) / 2; i += 32) sw32[i] = __ldg(w32 + i);
    __syncwarp();
    words = s_words;
  }
  const int64_t start = (int64_t)seg * seg_size;
  const int32_t cnt = (int32_t)((n - start) < (int64_t)seg_size ? (n - start) : (int64_t)seg_size);
  const int32_t rounds = (cnt + 31) / 32;
  bool bad = false;
  // table index of each round, prefetched kChunk rounds ahead (double buffered in registers)
  int32_t cur[kChunk], nxt[kChunk];
  // Without an index array the table of symbol i is its channel i / chan_len: the quotient is carried from round
  // to round (a 64-bit division per round would sit in the one instruction stream the warp's time consists of).
  // load_chunk is called for consecutive chunks, so the pair always belongs to the next round to be loaded.
  int64_t ch_q = 0, ch_r = 0;
  if constexpr (!HAS_IDX) {
    ch_q = (start + lane) / chan_len;
    ch_r = (start + lane) - ch_q * chan_len;
  }
  const uint16_t* idx_lane = HAS_IDX ? idx + start + lane : nullptr;      // this lane's entries: idx_lane[32 r]
  auto load_chunk = [&](int32_t r0, int32_t (&buf)[kChunk]) {   // rounds r0 ... r0+kChunk-1
#pragma unroll
    for (int k = 0; k < kChunk; ++k) {
      const int32_t li = (r0 + k) * 32 + lane;
      // the raw value only: clamping here would make every chunk wait for its own loads
      int32_t t = -1;
      if constexpr (HAS_IDX) {
        if (li < cnt) t = (int32_t)__ldg(idx_lane + (r0 + k) * 32);      // (li < cnt implies r0 + k < rounds)
      } else {
        if (r0 + k < rounds && li < cnt) t = (int32_t)ch_q;
        ch_r += 32;
        while (ch_r >= chan_len) {
          ch_r -= chan_len;
          ++ch_q;
        }
      }
      buf[k] = t;
    }
  };
  int32_t* sym_lane = sym + start + lane;      // this lane's outputs: sym_lane[32 r]
  load_chunk(0, cur);
  for (int32_t r0 = 0; r0 < rounds; r0 += kChunk) {
    load_chunk(r0 + kChunk, nxt);
#pragma unroll
    for (int k = 0; k < kChunk; ++k) {
      const int32_t r = r0 + k;
      if (r >= rounds) break;
      const bool active = cur[k] >= 0;
      const int32_t t = cur[k] < T.T ? cur[k] : T.T - 1;
      int32_t coff = 0, maxv = 0, offs = 0;
      if (active) {
        if (SMEM) {
          const int4 mt = s_meta[t];
          coff = mt.x; maxv = mt.y; offs = mt.z;
        } else {
          coff = __ldg(T.off + t);
          maxv = __ldg(T.size + t) - 2;
          offs = __ldg(T.offset + t);
        }
      }
      // step 0: table symbol
      int32_t v = 0;
      {
        const uint32_t cf = x & 0xffffu;
        if (active) {
          int32_t lo = 0, hi = maxv + 1;  // cdf[coff + maxv + 1] == 65536 > cf
          if (SMEM) {
            const uint16_t* lp = s_lut + t * (kLutN + 1) + (cf >> kLutShift);
            lo = (int32_t)lp[0];
            hi = (int32_t)lp[1] + 1;
          } else if (T.lut) {
            const uint16_t* lp = T.lut + (int64_t)t * (kLutN + 1) + (cf >> kLutShift);
            lo = (int32_t)__ldg(lp);
            hi = (int32_t)__ldg(lp + 1) + 1;
          }
          while (hi - lo > 1) {
            const int32_t mid = (lo + hi) >> 1;
            if (cdf_at(coff, mid) <= cf) lo = mid; else hi = mid;
          }
          v = lo;
          const uint32_t st = cdf_at(coff, v);
          const uint32_t nx = v == maxv ? 65536u : cdf_at(coff, v + 1);   // row terminator is implied
          const uint32_t fr = nx - st;
          x = fr * (x >> 16) + cf - st;
        }
        const bool need = active && x < kRansL;
        const unsigned m = __ballot_sync(0xffffffffu, need);
        wpos -= __popc(m);
        if (wpos < 0) { bad = true; wpos = 0; }
        if (need) x = (x << 16) | (uint32_t)words[wpos + __popc(m & lt)];
      }
      const bool esc = active && v == maxv;
      int nnib = 0;
      uint32_t raw = 0;
      if (__any_sync(0xffffffffu, esc)) {
        int nops = esc ? 2 : 1;
        for (int ts = 1;; ++ts) {
          const bool has = ts < nops;
          if (!__any_sync(0xffffffffu, has)) break;
          if (has) {
            uint32_t cf = x & 0xffffu;
            uint32_t val = cf >> 12;
            x = (x >> 16) * 4096u + (cf & 0xfffu);
            if (ts == 1) {
              nnib = (int)val;
              if (nnib > 8) { nnib = 8; bad = true; }
              nops = 2 + nnib;
            } else {
              raw |= val << (4 * (ts - 2));
            }
          }
          const bool need = has && x < kRansL;
          const unsigned m = __ballot_sync(0xffffffffu, need);
          wpos -= __popc(m);
          if (wpos < 0) { bad = true; wpos = 0; }
          if (need) x = (x << 16) | (uint32_t)words[wpos + __popc(m & lt)];
        }
      }
      if (active) {
        int64_t val = v;
        if (esc) {
          val = (int64_t)(raw >> 1);
          if (raw & 1u) val = -val - 1; else val += maxv;
        }
        sym_lane[r * 32] = (int32_t)(val + offs);
      }
    }
#pragma unroll
    for (int k = 0; k < kChunk; ++k) cur[k] = nxt[k];
  }
  if (__any_sync(0xffffffffu, bad) && lane == 0) atomicExch(status, 2);
}

}  // namespace pcc

using namespace pcc;

static inline int32_t nseg_of(int64_t n, int32_t seg) { return (int32_t)((n + seg - 1) / seg); }

extern "C" {

int64_t pcc_rans_max_bytes(int64_t n, int32_t seg_size) {
  if (n <= 0 || seg_size <= 0) return 64;
  int64_t ns = nseg_of(n, seg_size);
  return 16 + 4 * ns + ns * (132 + 2 * seg_cap_words(seg_size) + 4);
}
int64_t pcc_rans_workspace_bytes(int64_t n, int32_t seg_size) {
  if (n <= 0 || seg_size <= 0) return 256;
  int64_t ns = nseg_of(n, seg_size);
  return align_up(ns * seg_cap_words(seg_size) * 2, 256) + align_up(ns * 128, 256) + align_up(ns * 4, 256) +
         2 * align_up(n * 4, 256) + align_up((n + 31) / 32 * 4, 256) + 256;
}

// Q streams of n symbols each (sym / idx: [Q][n]; stream q's container at out + q * out_stride, its size in
// out_bytes_dev[q]) coded by ONE launch of each pass: grid.y = stream.  Every stream is an independent PCL1
// container, byte-identical to what pcc_rans_encode produces for it.
static int rans_encode_streams(const int32_t* sym, const uint16_t* idx, int64_t chan_len, int64_t n, int32_t num_streams,
                               int32_t seg_size, const pcc_cdf_tables_t* tables, uint8_t* out, int64_t out_stride,
                               int32_t* out_bytes_dev, void* ws, int64_t ws_bytes, pcc_stream_t s, const char* who) {
  PCC_CHECK_ARG(n > 0 && n < (1ll << 31) && seg_size >= 32 && seg_size % 32 == 0 && tables && out &&
                out_bytes_dev && ws && sym);
  PCC_CHECK_ARG(num_streams >= 1 && num_streams <= 65535);
  PCC_CHECK_ARG(idx || chan_len > 0);
  PCC_CHECK_ARG(tables->cdf && tables->cdf_off && tables->cdf_size && tables->offset && tables->num_tables > 0);
  const int64_t ws_stride = num_streams > 1 ? pcc_rans_workspace_bytes(n, seg_size) : 0;
  if (ws_bytes < pcc_rans_workspace_bytes(n, seg_size) * num_streams) {
    set_error("%s: workspace too small", who);
    return PCC_E_WORKSPACE;
  }
  if (num_streams > 1 && (out_stride < pcc_rans_max_bytes(n, seg_size) || (out_stride & 3))) {
    set_error("%s: out_stride %lld must be a multiple of 4 and at least pcc_rans_max_bytes = %lld", who,
              (long long)out_stride, (long long)pcc_rans_max_bytes(n, seg_size));
    return PCC_E_INVALID;
  }
  const int32_t ns = nseg_of(n, seg_size);
  char* p = (char*)ws;
  uint16_t* scratch = (uint16_t*)p;
  p += align_up((int64_t)ns * seg_cap_words(seg_size) * 2, 256);
  uint32_t* states = (uint32_t*)p;
  p += align_up((int64_t)ns * 128, 256);
  int32_t* nwords = (int32_t*)p;
  p += align_up((int64_t)ns * 4, 256);
  uint32_t* packed = (uint32_t*)p;
  p += align_up(n * 4, 256);
  uint32_t* raw = (uint32_t*)p;
  p += align_up(n * 4, 256);
  uint32_t* escmask = (uint32_t*)p;
  Tables T{tables->cdf, tables->cdf_off, tables->cdf_size, tables->offset, tables->lut, tables->num_tables, tables->cdf_total};
  cudaStream_t st = as_stream(s);
  const int64_t cl = chan_len > 0 ? chan_len : 1;
  const unsigned Q = (unsigned)num_streams;
  launch_pdl(k_rans_prepare, dim3((unsigned)((n + 255) / 256), Q), dim3(256), 0, st, sym, idx, cl, n, T, packed, raw, escmask,
             ws_stride);
  PCC_LAUNCHED();
  launch_pdl(k_rans_encode, dim3((ns + kEncWarps - 1) / kEncWarps, Q), dim3(32 * kEncWarps), 0, st, packed, raw, escmask, n,
             seg_size, ns, scratch, states, nwords, ws_stride);
  PCC_LAUNCHED();
  launch_pdl(k_rans_pack, dim3(ns, Q), dim3(256), 0, st, scratch, states, nwords, ns, seg_size, (uint32_t)n, out, out_bytes_dev,
             ws_stride, num_streams > 1 ? out_stride : (int64_t)0);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_rans_encode(const int32_t* sym, const uint16_t* idx, int64_t chan_len, int64_t n,
                    int32_t seg_size, const pcc_cdf_tables_t* tables, uint8_t* out,
                    int32_t* out_bytes_dev, void* ws, int64_t ws_bytes, pcc_stream_t s) {
  return rans_encode_streams(sym, idx, chan_len, n, 1, seg_size, tables, out, 0, out_bytes_dev, ws, ws_bytes, s,
                             "pcc_rans_encode");
}

int pcc_rans_encode_batch(const int32_t* sym, const uint16_t* idx, int64_t chan_len, int64_t n, int32_t num_streams,
                          int32_t seg_size, const pcc_cdf_tables_t* tables, uint8_t* out, int64_t out_stride,
                          int32_t* out_bytes_dev, void* ws, int64_t ws_bytes, pcc_stream_t s) {
  return rans_encode_streams(sym, idx, chan_len, n, num_streams, seg_size, tables, out, out_stride, out_bytes_dev, ws,
                             ws_bytes, s, "pcc_rans_encode_batch");
}

int pcc_rans_decode(const uint8_t* container, int64_t container_bytes, int32_t seg_size,
                    const uint16_t* idx, int64_t chan_len, int64_t n,
                    const pcc_cdf_tables_t* tables, int32_t* sym, int32_t* status_dev,
                    pcc_stream_t s) {
  PCC_CHECK_ARG(n > 0 && n < (1ll << 31) && container && tables && sym && status_dev);
  PCC_CHECK_ARG(seg_size >= 32 && seg_size % 32 == 0);
  PCC_CHECK_ARG(idx || chan_len > 0);
  PCC_CHECK_ARG(((uintptr_t)container % 4) == 0);
  const int32_t ns = nseg_of(n, seg_size);
  PCC_CHECK_ARG(container_bytes >= 16 + 4ll * ns);
  Tables T{tables->cdf, tables->cdf_off, tables->cdf_size, tables->offset, tables->lut, tables->num_tables, tables->cdf_total};
  cudaStream_t st = as_stream(s);
  PCC_CUDA(cudaMemsetAsync(status_dev, 0, 4, st));
  const size_t wsm = (size_t)kDecWarps * kWordCap * 2;
  const size_t smem = (size_t)T.T * 16 +
                      ((((size_t)T.T * (kLutN + 1) + 7) & ~(size_t)7) + (((size_t)T.total + 7) & ~(size_t)7)) * 2 + wsm;
  const unsigned grid = (unsigned)((ns + kDecWarps - 1) / kDecWarps);
  static PerDeviceOnce configured;      // the opt-in is per device, not per process
  const int dev = current_device();
  if (configured.needed(dev)) {
    PCC_CUDA(cudaFuncSetAttribute(k_rans_decode<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    PCC_CUDA(cudaFuncSetAttribute(k_rans_decode<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    PCC_CUDA(cudaFuncSetAttribute(k_rans_decode<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm));
    PCC_CUDA(cudaFuncSetAttribute(k_rans_decode<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm));
    configured.mark(dev);
  }
  const int64_t cl = chan_len > 0 ? chan_len : 1;
  const dim3 g(grid), b(32 * kDecWarps);
  if (T.lut && T.total > 0 && smem <= 224 * 1024) {
    if (idx) launch_pdl(k_rans_decode<true, true>, g, b, smem, st, container, container_bytes, seg_size, idx, cl, n, ns, T, sym, status_dev);
    else launch_pdl(k_rans_decode<true, false>, g, b, smem, st, container, container_bytes, seg_size, idx, cl, n, ns, T, sym, status_dev);
  } else {
    if (idx) launch_pdl(k_rans_decode<false, true>, g, b, wsm, st, container, container_bytes, seg_size, idx, cl, n, ns, T, sym, status_dev);
    else launch_pdl(k_rans_decode<false, false>, g, b, wsm, st, container, container_bytes, seg_size, idx, cl, n, ns, T, sym, status_dev);
  }
  PCC_LAUNCHED();
  return PCC_OK;
}

int64_t pcc_rans_lut_entries(int32_t num_tables) { return (int64_t)num_tables * (kLutN + 1); }

int pcc_rans_build_lut(const pcc_cdf_tables_t* tables, uint16_t* lut, pcc_stream_t s) {
  PCC_CHECK_ARG(tables && lut && tables->num_tables > 0);
  PCC_CHECK_ARG(tables->cdf && tables->cdf_off && tables->cdf_size && tables->offset);
  Tables T{tables->cdf, tables->cdf_off, tables->cdf_size, tables->offset, nullptr, tables->num_tables, 0};
  launch_pdl(k_rans_build_lut, dim3(tables->num_tables), dim3(128), 0, as_stream(s), T, lut);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
