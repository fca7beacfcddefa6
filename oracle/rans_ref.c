/* Oracle (test infrastructure): CPU entropy coders.
 *
 * (a) ref64_*: restatement of CompressAI 1.2.4 `RansEncoder.encode_with_indexes` /
 *     `RansDecoder.decode_with_indexes` (compressai/cpp_exts/rans/rans_interface.cpp over ryg's
 *     rans64.h; unvendored dependency, sender/encoder/requirements.txt:10): one
 *     64-bit state, 32-bit renormalisation, 16-bit precision, 4-bit bypass escapes.  This is the
 *     coder the reference reaches through gaussian_conditional.compress (codec_pipeline.py:426-430)
 *     and entropy_bottleneck.compress (:305); it is the bit-rate yardstick for the GPU container.
 * (b) pcl1_*: scalar restatement of this repository's interleaved 32-lane "PCL1" container
 *      used to check the CUDA coder byte for byte.
 *
 * Plain C, no dependencies:  gcc -O2 -shared -fPIC rans_ref.c -o _build/librans_ref.so
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define PRECISION 16
#define BYPASS_PRECISION 4
#define MAX_BYPASS_VAL ((1 << BYPASS_PRECISION) - 1)
#define RANS64_L (1ull << 31)

typedef struct {
  uint16_t start;
  uint16_t range; /* 0 encodes 65536 is impossible: ranges are < 65536 */
  int bypass;
} sym64_t;

/* -------- (a) CompressAI-style single stream -------------------------------------------------- */

/* Returns number of bytes written to out (capacity out_cap), or -1 on overflow. */
int64_t ref64_encode(const int32_t* symbols, const int32_t* indexes, int64_t n, const int32_t* cdfs,
                     const int32_t* cdf_off, const int32_t* cdf_sizes, const int32_t* offsets,
                     uint8_t* out, int64_t out_cap) {
  /* pass 1: build the symbol list (forward), like RansEncoder's _syms vector */
  int64_t cap = n * 2 + 16, cnt = 0;
  sym64_t* syms = (sym64_t*)malloc(sizeof(sym64_t) * cap);
  if (!syms) return -1;
#define PUSH(s_, r_, b_)                                              \
  do {                                                                \
    if (cnt == cap) {                                                 \
      cap *= 2;                                                       \
      syms = (sym64_t*)realloc(syms, sizeof(sym64_t) * cap);          \
      if (!syms) return -1;                                           \
    }                                                                 \
    syms[cnt].start = (uint16_t)(s_);                                 \
    syms[cnt].range = (uint16_t)(r_);                                 \
    syms[cnt].bypass = (b_);                                          \
    cnt++;                                                            \
  } while (0)
  for (int64_t i = 0; i < n; ++i) {
    const int32_t idx = indexes[i];
    const int32_t* cdf = cdfs + cdf_off[idx];
    const int32_t max_value = cdf_sizes[idx] - 2;
    int32_t value = symbols[i] - offsets[idx];
    uint32_t raw_val = 0;
    if (value < 0) {
      raw_val = (uint32_t)(-2 * (int64_t)value - 1);
      value = max_value;
    } else if (value >= max_value) {
      raw_val = (uint32_t)(2 * ((int64_t)value - max_value));
      value = max_value;
    }
    PUSH(cdf[value], cdf[value + 1] - cdf[value], 0);
    if (value == max_value) {
      int32_t n_bypass = 0;
      while ((raw_val >> (n_bypass * BYPASS_PRECISION)) != 0 && n_bypass < 8) ++n_bypass;
      int32_t val = n_bypass;
      while (val >= MAX_BYPASS_VAL) {
        PUSH(MAX_BYPASS_VAL, MAX_BYPASS_VAL + 1, 1);
        val -= MAX_BYPASS_VAL;
      }
      PUSH(val, val + 1, 1);
      for (int32_t j = 0; j < n_bypass; ++j) {
        int32_t v = (raw_val >> (j * BYPASS_PRECISION)) & MAX_BYPASS_VAL;
        PUSH(v, v + 1, 1);
      }
    }
  }
#undef PUSH
  /* pass 2: encode in reverse into a uint32 buffer filled from the end */
  int64_t words_cap = cnt + 8;
  uint32_t* buf = (uint32_t*)malloc(sizeof(uint32_t) * words_cap);
  if (!buf) { free(syms); return -1; }
  uint32_t* ptr = buf + words_cap;
  uint64_t x = RANS64_L;
  for (int64_t k = cnt - 1; k >= 0; --k) {
    if (!syms[k].bypass) {
      uint32_t start = syms[k].start, freq = syms[k].range;
      uint64_t x_max = ((RANS64_L >> PRECISION) << 32) * freq;
      if (x >= x_max) { *--ptr = (uint32_t)x; x >>= 32; }
      x = ((x / freq) << PRECISION) + (x % freq) + start;
    } else {
      uint32_t val = syms[k].start;
      uint32_t freq = 1u << (PRECISION - BYPASS_PRECISION);
      uint64_t x_max = ((RANS64_L >> PRECISION) << 32) * freq;
      if (x >= x_max) { *--ptr = (uint32_t)x; x >>= 32; }
      x = (x << BYPASS_PRECISION) | val;
    }
  }
  ptr -= 2;
  ptr[0] = (uint32_t)(x >> 0);
  ptr[1] = (uint32_t)(x >> 32);
  int64_t nbytes = (int64_t)((buf + words_cap) - ptr) * 4;
  int64_t ret = -1;
  if (nbytes <= out_cap) { memcpy(out, ptr, nbytes); ret = nbytes; }
  free(buf);
  free(syms);
  return ret;
}

int ref64_decode(const uint8_t* in, int64_t nbytes, const int32_t* indexes, int64_t n,
                 const int32_t* cdfs, const int32_t* cdf_off, const int32_t* cdf_sizes,
                 const int32_t* offsets, int32_t* out) {
  if (nbytes < 8 || (nbytes & 3)) return -1;
  const uint32_t* ptr = (const uint32_t*)in;
  const uint32_t* end = ptr + nbytes / 4;
  uint64_t x = (uint64_t)ptr[0] | ((uint64_t)ptr[1] << 32);
  ptr += 2;
#define GETBITS(dst)                                                       \
  do {                                                                     \
    uint32_t v_ = (uint32_t)(x & ((1u << BYPASS_PRECISION) - 1));          \
    x >>= BYPASS_PRECISION;                                                \
    if (x < RANS64_L) { if (ptr >= end) return -2; x = (x << 32) | *ptr++; } \
    dst = v_;                                                              \
  } while (0)
  for (int64_t i = 0; i < n; ++i) {
    const int32_t idx = indexes[i];
    const int32_t* cdf = cdfs + cdf_off[idx];
    const int32_t cdf_size = cdf_sizes[idx];
    const int32_t max_value = cdf_size - 2;
    const uint32_t cum = (uint32_t)(x & ((1u << PRECISION) - 1));
    int32_t s = 0;
    while (s + 1 < cdf_size && (uint32_t)cdf[s + 1] <= cum) ++s; /* find_if(cdf > cum) - 1 */
    uint32_t start = cdf[s], freq = cdf[s + 1] - cdf[s];
    x = freq * (x >> PRECISION) + cum - start;
    if (x < RANS64_L) { if (ptr >= end) return -2; x = (x << 32) | *ptr++; }
    int32_t value = s;
    if (value == max_value) {
      int32_t val;
      GETBITS(val);
      int32_t n_bypass = val;
      while (val == MAX_BYPASS_VAL) { GETBITS(val); n_bypass += val; }
      uint32_t raw_val = 0;
      for (int j = 0; j < n_bypass; ++j) { GETBITS(val); raw_val |= (uint32_t)val << (j * BYPASS_PRECISION); }
      value = (int32_t)(raw_val >> 1);
      if (raw_val & 1) value = -value - 1; else value += max_value;
    }
    out[i] = value + offsets[idx];
  }
#undef GETBITS
  return 0;
}

/* -------- (b) PCL1: 32 interleaved lanes, 32-bit states, 16-bit words ---------------------------- */

#define PCL1_MAGIC 0x314c4350u
#define RANS32_L (1u << 16)

typedef struct { int32_t coff, v, nnib, nops; uint32_t raw; } ops_t;

static ops_t make_ops(int32_t symbol, int32_t t, const int32_t* cdf_off, const int32_t* cdf_sizes,
                      const int32_t* offsets) {
  ops_t o;
  o.coff = cdf_off[t];
  int32_t maxv = cdf_sizes[t] - 2;
  int64_t v = (int64_t)symbol - offsets[t];
  o.raw = 0;
  if (v < 0) { o.raw = (uint32_t)(-2 * v - 1); v = maxv; }
  else if (v >= maxv) { o.raw = (uint32_t)(2 * (v - maxv)); v = maxv; }
  o.v = (int32_t)v;
  o.nnib = 0;
  o.nops = 1;
  if (v == maxv) {
    while (o.nnib < 8 && (o.raw >> (4 * o.nnib)) != 0) o.nnib++;
    o.nops = 2 + o.nnib;
  }
  return o;
}

static uint32_t seg_bytes_of(int32_t nw) { return 4u + 128u + ((2u * (uint32_t)nw + 3u) & ~3u); }

int64_t pcl1_max_bytes(int64_t n, int32_t seg_size) {
  int64_t ns = (n + seg_size - 1) / seg_size;
  return 16 + 4 * ns + ns * (136 + 2 * (4ll * seg_size + 64));
}

/* idx == NULL: table = i / chan_len */
int64_t pcl1_encode(const int32_t* sym, const uint16_t* idx, int64_t chan_len, int64_t n, int32_t seg_size,
                    const uint32_t* cdf, const int32_t* cdf_off, const int32_t* cdf_sizes,
                    const int32_t* offsets, int32_t num_tables, uint8_t* out, int64_t out_cap) {
  if (n <= 0 || seg_size < 32 || (seg_size % 32)) return -1;
  int32_t nseg = (int32_t)((n + seg_size - 1) / seg_size);
  int64_t hdr = 16 + 4ll * nseg;
  if (out_cap < pcl1_max_bytes(n, seg_size)) return -1;
  uint32_t* o32 = (uint32_t*)out;
  o32[0] = PCL1_MAGIC; o32[1] = (uint32_t)n; o32[2] = (uint32_t)seg_size; o32[3] = (uint32_t)nseg;
  int64_t pos = hdr;
  uint16_t* words = (uint16_t*)malloc(sizeof(uint16_t) * (4ll * seg_size + 64));
  if (!words) return -1;
  for (int32_t seg = 0; seg < nseg; ++seg) {
    int64_t start = (int64_t)seg * seg_size;
    int32_t cnt = (int32_t)((n - start) < seg_size ? (n - start) : seg_size);
    int32_t rounds = (cnt + 31) / 32;
    uint32_t x[32];
    for (int l = 0; l < 32; ++l) x[l] = RANS32_L;
    int32_t wpos = 0;
    for (int32_t r = rounds - 1; r >= 0; --r) {
      ops_t o[32];
      int maxops = 0;
      for (int l = 0; l < 32; ++l) {
        int32_t li = r * 32 + l;
        o[l].nops = 0;
        if (li < cnt) {
          int64_t i = start + li;
          int32_t t = idx ? (int32_t)idx[i] : (int32_t)(i / chan_len);
          if (t >= num_tables) t = num_tables - 1;
          o[l] = make_ops(sym[i], t, cdf_off, cdf_sizes, offsets);
        }
        if (o[l].nops > maxops) maxops = o[l].nops;
      }
      for (int ts = maxops - 1; ts >= 0; --ts) {
        /* lanes that must emit write in ascending lane order, then every lane advances */
        for (int l = 0; l < 32; ++l) {
          if (ts >= o[l].nops) continue;
          uint32_t st, fr;
          if (ts == 0) { st = cdf[o[l].coff + o[l].v]; fr = cdf[o[l].coff + o[l].v + 1] - st; }
          else {
            uint32_t val = ts == 1 ? (uint32_t)o[l].nnib : ((o[l].raw >> (4 * (ts - 2))) & 15u);
            st = val << 12; fr = 1u << 12;
          }
          if ((uint64_t)x[l] >= ((uint64_t)fr << 16)) { words[wpos++] = (uint16_t)(x[l] & 0xffffu); x[l] >>= 16; }
          x[l] = ((x[l] / fr) << 16) + (x[l] % fr) + st;
        }
      }
    }
    uint32_t sb = seg_bytes_of(wpos);
    o32[4 + seg] = sb;
    uint32_t* d32 = (uint32_t*)(out + pos);
    d32[0] = (uint32_t)wpos;
    for (int l = 0; l < 32; ++l) d32[1 + l] = x[l];
    memcpy(out + pos + 132, words, 2ll * wpos);
    if (wpos & 1) { out[pos + 132 + 2ll * wpos] = 0; out[pos + 133 + 2ll * wpos] = 0; }
    pos += sb;
  }
  free(words);
  return pos;
}

int pcl1_decode(const uint8_t* in, int64_t nbytes, const uint16_t* idx, int64_t chan_len, int64_t n,
                const uint32_t* cdf, const int32_t* cdf_off, const int32_t* cdf_sizes,
                const int32_t* offsets, int32_t num_tables, int32_t* out) {
  if (nbytes < 16) return -1;
  const uint32_t* h32 = (const uint32_t*)in;
  if (h32[0] != PCL1_MAGIC || h32[1] != (uint32_t)n) return -3;
  int32_t seg_size = (int32_t)h32[2], nseg = (int32_t)h32[3];
  if (seg_size < 32 || nseg != (int32_t)((n + seg_size - 1) / seg_size)) return -3;
  int64_t pos = 16 + 4ll * nseg;
  if (pos > nbytes) return -1;
  for (int32_t seg = 0; seg < nseg; ++seg) {
    if (pos + 132 > nbytes) return -1;
    const uint32_t* s32 = (const uint32_t*)(in + pos);
    int32_t wpos = (int32_t)s32[0];
    if (pos + 132 + 2ll * wpos > nbytes) return -1;
    uint32_t x[32];
    for (int l = 0; l < 32; ++l) x[l] = s32[1 + l];
    const uint16_t* words = (const uint16_t*)(in + pos + 132);
    int64_t start = (int64_t)seg * seg_size;
    int32_t cnt = (int32_t)((n - start) < seg_size ? (n - start) : seg_size);
    int32_t rounds = (cnt + 31) / 32;
    for (int32_t r = 0; r < rounds; ++r) {
      int32_t v[32], maxv[32], offs[32], coff[32], nops[32], nnib[32];
      uint32_t raw[32];
      int act[32];
      for (int l = 0; l < 32; ++l) {
        int32_t li = r * 32 + l;
        act[l] = li < cnt;
        nops[l] = act[l] ? 1 : 0;
        nnib[l] = 0; raw[l] = 0; v[l] = 0;
        if (act[l]) {
          int64_t i = start + li;
          int32_t t = idx ? (int32_t)idx[i] : (int32_t)(i / chan_len);
          if (t >= num_tables) t = num_tables - 1;
          coff[l] = cdf_off[t]; maxv[l] = cdf_sizes[t] - 2; offs[l] = offsets[t];
        }
      }
      for (int ts = 0;; ++ts) {
        int any = 0;
        for (int l = 0; l < 32; ++l) if (ts < nops[l]) any = 1;
        if (!any) break;
        /* all lanes advance, then the lanes that need a word read in ascending lane order from a
           block that ends at the current stream position */
        int need[32], cntw = 0;
        for (int l = 0; l < 32; ++l) {
          need[l] = 0;
          if (ts >= nops[l]) continue;
          uint32_t cf = x[l] & 0xffffu;
          if (ts == 0) {
            int32_t lo = 0, hi = maxv[l] + 1;
            while (hi - lo > 1) { int32_t mid = (lo + hi) >> 1; if (cdf[coff[l] + mid] <= cf) lo = mid; else hi = mid; }
            v[l] = lo;
            uint32_t st = cdf[coff[l] + lo], fr = cdf[coff[l] + lo + 1] - st;
            x[l] = fr * (x[l] >> 16) + cf - st;
            if (lo == maxv[l]) nops[l] = 2;
          } else {
            uint32_t val = cf >> 12;
            x[l] = (x[l] >> 16) * 4096u + (cf & 0xfffu);
            if (ts == 1) { nnib[l] = (int32_t)val; if (nnib[l] > 8) return -2; nops[l] = 2 + nnib[l]; }
            else raw[l] |= val << (4 * (ts - 2));
          }
          if (x[l] < RANS32_L) { need[l] = 1; cntw++; }
        }
        wpos -= cntw;
        if (wpos < 0) return -2;
        int k = 0;
        for (int l = 0; l < 32; ++l) if (need[l]) x[l] = (x[l] << 16) | words[wpos + k++];
      }
      for (int l = 0; l < 32; ++l) {
        if (!act[l]) continue;
        int64_t val = v[l];
        if (v[l] == maxv[l]) { val = (int64_t)(raw[l] >> 1); if (raw[l] & 1u) val = -val - 1; else val += maxv[l]; }
        out[start + r * 32 + l] = (int32_t)(val + offs[l]);
      }
    }
    pos += h32[4 + seg];
  }
  return 0;
}
