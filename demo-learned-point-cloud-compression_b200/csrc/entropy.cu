// (3) Fused entropy-model kernels: quantise + Gaussian scale index (+ dequantise with dead-zone
// offsets) and the factorized-prior quantiser.  Latents are [n, C] row-major in key order; symbols
// and indexes are channel-major [C, n] in canonical lexicographic column order.  Each kernel moves a
// 32x32 (rows x channels) tile through shared memory so both sides are accessed with full 128-byte
// lines.  Arithmetic replicates torch's unfused fp32 ops (mul, sub, round-half-even) exactly.
#include "common.cuh"
#include <vector>
#include <cmath>

namespace pcc {

constexpr int kTile = 32;
constexpr float kSymClamp = 1073741824.0f;  // 2^30

__device__ __forceinline__ int32_t round_to_sym(float v) {
  v = rintf(v);
  v = fminf(fmaxf(v, -kSymClamp), kSymClamp);
  return (int32_t)v;
}

// CompressAI GaussianConditional.build_indexes: #entries of table[0..L-2] strictly below the
// lower-bounded scale (table ascending).
__device__ __forceinline__ int scale_index(float sb, const float* __restrict__ table, int len) {
  int lo = 0, hi = len - 1;  // search in [0, len-1)
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (table[mid] < sb) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// Centroid shift (towards zero) of a unit quantisation bin under a Laplacian of std `sd`.
__device__ __forceinline__ float bin_offset(float sd) {
  float b = sd * 0.70710678118654752f;
  float d;
  if (b > 4.0f) {
    float ib = 1.0f / b;
    d = ib * (1.0f / 12.0f) - ib * ib * ib * (1.0f / 720.0f);
  } else {
    d = 0.5f / tanhf(0.5f / b) - b;
  }
  return fminf(fmaxf(d, 0.0f), 0.5f);
}

__global__ void __launch_bounds__(256)
k_factorized_quant(const float* __restrict__ z, const int32_t* __restrict__ perm,
                   const float* __restrict__ med, int64_t n, int C, int32_t* __restrict__ sym,
                   __nv_bfloat16* __restrict__ zb, float* __restrict__ zf) {
  pdl_trigger();
  pdl_wait();
  __shared__ int32_t tile[kTile][kTile + 1];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int64_t i0 = (int64_t)blockIdx.x * kTile;
  const int c0 = blockIdx.y * kTile;
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    int li = ty + 8 * rr;
    int64_t i = i0 + li;
    int c = c0 + tx;
    int32_t q = 0;
    if (i < n && c < C) {
      int64_t row = perm ? perm[i] : i;
      float m = med[c];
      q = round_to_sym(__fsub_rn(z[row * C + c], m));
      float zh = __fadd_rn((float)q, m);
      if (zb) zb[row * C + c] = __float2bfloat16_rn(zh);
      if (zf) zf[row * C + c] = zh;
    }
    tile[li][tx] = q;
  }
  __syncthreads();
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    int c = c0 + ty + 8 * rr;
    int64_t i = i0 + tx;
    if (i < n && c < C) sym[(int64_t)c * n + i] = tile[tx][ty + 8 * rr];
  }
}

__global__ void __launch_bounds__(256)
k_factorized_dequant(const int32_t* __restrict__ sym, const int32_t* __restrict__ perm,
                     const float* __restrict__ med, int64_t n, int C, __nv_bfloat16* __restrict__ zb,
                     float* __restrict__ zf) {
  pdl_trigger();
  pdl_wait();
  __shared__ int32_t tile[kTile][kTile + 1];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int64_t i0 = (int64_t)blockIdx.x * kTile;
  const int c0 = blockIdx.y * kTile;
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    int c = c0 + ty + 8 * rr;
    int64_t i = i0 + tx;
    tile[ty + 8 * rr][tx] = (i < n && c < C) ? sym[(int64_t)c * n + i] : 0;
  }
  __syncthreads();
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    int li = ty + 8 * rr;
    int64_t i = i0 + li;
    int c = c0 + tx;
    if (i < n && c < C) {
      int64_t row = perm ? perm[i] : i;
      float zh = __fadd_rn((float)tile[tx][li], med[c]);
      if (zb) zb[row * C + c] = __float2bfloat16_rn(zh);
      if (zf) zf[row * C + c] = zh;
    }
  }
}

// ENCODE = true: symbols + indexes for Q qualities.  ENCODE = false: indexes only (Q = 1, y unused).
template <bool ENCODE>
__global__ void __launch_bounds__(256)
k_gauss_params(const float* __restrict__ y, const float* __restrict__ params,
               const int32_t* __restrict__ prow, const int32_t* __restrict__ perm,
               const float* __restrict__ scale, const float* __restrict__ table, int table_len,
               float bound, int64_t n, int C, int Q, int32_t* __restrict__ sym,
               uint16_t* __restrict__ idx) {
  pdl_trigger();
  pdl_wait();
  __shared__ int32_t tsym[kTile][kTile + 1];
  __shared__ uint16_t tidx[kTile][kTile + 2];
  __shared__ float stable[256];
  const int tx = threadIdx.x, ty = threadIdx.y;
  for (int t = ty * 32 + tx; t < table_len; t += 256) stable[t] = table[t];
  const int64_t i0 = (int64_t)blockIdx.x * kTile;
  const int c0 = blockIdx.y * kTile;
  float yv[4], sg[4], mu[4];
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    int64_t i = i0 + ty + 8 * rr;
    int c = c0 + tx;
    yv[rr] = 0.f; sg[rr] = 0.f; mu[rr] = 0.f;
    if (i < n && c < C) {
      int64_t row = perm ? perm[i] : i;
      if (ENCODE) yv[rr] = y[row * C + c];
      int64_t pr = prow ? prow[row] : row;
      if (pr >= 0) {
        sg[rr] = params[pr * 2 * C + c];
        mu[rr] = params[pr * 2 * C + C + c];
      }
    }
  }
  __syncthreads();
  for (int q = 0; q < Q; ++q) {
    const int c = c0 + tx;
    const float s = c < C ? scale[q * C + c] : 1.f;
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
      float sb = fmaxf(__fmul_rn(sg[rr], s), bound);
      tidx[ty + 8 * rr][tx] = (uint16_t)scale_index(sb, stable, table_len);
      if (ENCODE) tsym[ty + 8 * rr][tx] = round_to_sym(__fsub_rn(__fmul_rn(yv[rr], s), __fmul_rn(mu[rr], s)));
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
      int cc = c0 + ty + 8 * rr;
      int64_t i = i0 + tx;
      if (i < n && cc < C) {
        int64_t o = ((int64_t)q * C + cc) * n + i;
        idx[o] = tidx[tx][ty + 8 * rr];
        if (ENCODE) sym[o] = tsym[tx][ty + 8 * rr];
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
k_gauss_dequant(const int32_t* __restrict__ sym, const float* __restrict__ params,
                const int32_t* __restrict__ prow, const int32_t* __restrict__ perm,
                const float* __restrict__ scale, const float* __restrict__ gain, float bound, int64_t n,
                int C, __nv_bfloat16* __restrict__ yb, float* __restrict__ yf) {
  pdl_trigger();
  pdl_wait();
  __shared__ int32_t tile[kTile][kTile + 1];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int64_t i0 = (int64_t)blockIdx.x * kTile;
  const int c0 = blockIdx.y * kTile;
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    int c = c0 + ty + 8 * rr;
    int64_t i = i0 + tx;
    tile[ty + 8 * rr][tx] = (i < n && c < C) ? sym[(int64_t)c * n + i] : 0;
  }
  __syncthreads();
#pragma unroll
  for (int rr = 0; rr < 4; ++rr) {
    int li = ty + 8 * rr;
    int64_t i = i0 + li;
    int c = c0 + tx;
    if (i < n && c < C) {
      int64_t row = perm ? perm[i] : i;
      int64_t pr = prow ? prow[row] : row;
      float sg = 0.f, mu = 0.f;
      if (pr >= 0) {
        sg = params[pr * 2 * C + c];
        mu = params[pr * 2 * C + C + c];
      }
      float s = scale[c];
      float rescale = __fdiv_rn(1.0f, s);
      float q = (float)tile[tx][li];
      float qa = fabsf(q);
      float sign = q > 0.f ? 1.f : (q < 0.f ? -1.f : 0.f);
      float sd = fmaxf(__fmul_rn(sg, s), bound);
      float qoff = -(gain ? gain[c] : 1.f) * bin_offset(sd);
      if (qa < 1e-4f) qoff = 0.f;
      float t = __fmul_rn(sign, __fadd_rn(qa, qoff));
      float yh = __fadd_rn(__fmul_rn(t, rescale), mu);
      if (yb) yb[row * C + c] = __float2bfloat16_rn(yh);
      if (yf) yf[row * C + c] = yh;
    }
  }
}

}  // namespace pcc

using namespace pcc;

static inline dim3 tile_grid(int64_t n, int c) {
  return dim3((unsigned)((n + kTile - 1) / kTile), (unsigned)((c + kTile - 1) / kTile));
}

extern "C" {

int pcc_factorized_quant(const float* z, const int32_t* perm, const float* medians, int64_t n,
                         int c, int32_t* sym, void* zhat_bf16, float* zhat_f32, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && c > 0 && c <= 65535 * kTile);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(z && medians && sym);
  launch_pdl(k_factorized_quant, dim3(tile_grid(n, c)), dim3(dim3(32, 8)), 0, as_stream(s), z, perm, medians, n, c, sym, (__nv_bfloat16*)zhat_bf16, zhat_f32);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_factorized_dequant(const int32_t* sym, const int32_t* perm, const float* medians,
                           int64_t n, int c, void* zhat_bf16, float* zhat_f32, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && c > 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(sym && medians && (zhat_bf16 || zhat_f32));
  launch_pdl(k_factorized_dequant, dim3(tile_grid(n, c)), dim3(dim3(32, 8)), 0, as_stream(s), sym, perm, medians, n, c, (__nv_bfloat16*)zhat_bf16, zhat_f32);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_gauss_encode(const float* y, const float* params, const int32_t* prow, const int32_t* perm,
                     const float* scale, const float* scale_table, int table_len, float scale_bound,
                     int64_t n, int c, int q, int32_t* sym, uint16_t* idx, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && c > 0 && q >= 1 && table_len >= 2 && table_len <= 256);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(y && params && scale && scale_table && sym && idx);
  launch_pdl(k_gauss_params<true>, dim3(tile_grid(n, c)), dim3(dim3(32, 8)), 0, as_stream(s), y, params, prow, perm, scale, scale_table, table_len, scale_bound, n, c, q, sym, idx);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_gauss_indexes(const float* params, const int32_t* prow, const int32_t* perm,
                      const float* scale, const float* scale_table, int table_len,
                      float scale_bound, int64_t n, int c, uint16_t* idx, pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && c > 0 && table_len >= 2 && table_len <= 256);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(params && scale && scale_table && idx);
  launch_pdl(k_gauss_params<false>, dim3(tile_grid(n, c)), dim3(dim3(32, 8)), 0, as_stream(s), nullptr, params, prow, perm, scale, scale_table, table_len, scale_bound, n, c, 1, nullptr, idx);
  PCC_LAUNCHED();
  return PCC_OK;
}

int pcc_gauss_dequant(const int32_t* sym, const float* params, const int32_t* prow,
                      const int32_t* perm, const float* scale, const float* offset_gain,
                      float scale_bound, int64_t n, int c, void* yhat_bf16, float* yhat_f32,
                      pcc_stream_t s) {
  PCC_CHECK_ARG(n >= 0 && c > 0);
  if (n == 0) return PCC_OK;
  PCC_CHECK_ARG(sym && params && scale && (yhat_bf16 || yhat_f32));
  launch_pdl(k_gauss_dequant, dim3(tile_grid(n, c)), dim3(dim3(32, 8)), 0, as_stream(s), sym, params, prow, perm, scale, offset_gain, scale_bound, n, c, (__nv_bfloat16*)yhat_bf16,
      yhat_f32);
  PCC_LAUNCHED();
  return PCC_OK;
}

// CompressAI `pmf_to_quantized_cdf` (compressai/cpp_exts/ops/ops.cpp, v1.2.4), restated.
int pcc_pmf_to_quantized_cdf(const float* pmf, int pmf_len, int precision, int32_t* cdf_out) {
  PCC_CHECK_ARG(pmf && cdf_out && pmf_len >= 1 && precision >= 1 && precision <= 16);
  for (int i = 0; i < pmf_len; ++i) {
    if (!(pmf[i] >= 0.f) || !std::isfinite(pmf[i])) {
      set_error("pmf_to_quantized_cdf: invalid pmf entry %d", i);
      return PCC_E_INVALID;
    }
  }
  std::vector<uint32_t> cdf(pmf_len + 1);
  cdf[0] = 0;
  for (int i = 0; i < pmf_len; ++i) cdf[i + 1] = (uint32_t)std::lround(pmf[i] * (float)(1 << precision));
  uint64_t total = 0;
  for (auto v : cdf) total += v;
  if (total == 0) {
    set_error("pmf_to_quantized_cdf: zero total");
    return PCC_E_INVALID;
  }
  for (auto& v : cdf) v = (uint32_t)((((uint64_t)1 << precision) * v) / total);
  for (int i = 1; i <= pmf_len; ++i) cdf[i] += cdf[i - 1];
  cdf[pmf_len] = 1u << precision;
  for (int i = 0; i < pmf_len; ++i) {
    if (cdf[i] == cdf[i + 1]) {
      uint32_t best_freq = ~0u;
      int best_steal = -1;
      for (int j = 0; j < pmf_len; ++j) {
        uint32_t freq = cdf[j + 1] - cdf[j];
        if (freq > 1 && freq < best_freq) {
          best_freq = freq;
          best_steal = j;
        }
      }
      if (best_steal < 0) {
        set_error("pmf_to_quantized_cdf: cannot steal a count");
        return PCC_E_INVALID;
      }
      if (best_steal < i) {
        for (int j = best_steal + 1; j <= i; ++j) cdf[j]--;
      } else {
        for (int j = i + 1; j <= best_steal; ++j) cdf[j]++;
      }
    }
  }
  for (int i = 0; i <= pmf_len; ++i) cdf_out[i] = (int32_t)cdf[i];
  return PCC_OK;
}

}  // extern "C"
