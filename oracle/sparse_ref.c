/* Oracle (test infrastructure): the two heavy loops of sparse_ref.py in plain C + OpenMP, so that the
 * oracle codes BASELINE.json's full-size frames (800 k points, 1.8 M candidate rows in g_s) in seconds.
 *
 * These are the same definitions as sparse_ref.kernel_map / sparse_ref.conv_forward (SURVEY.md A.2.3-A.2.5:
 * MinkowskiEngine's kernel-region enumeration and gather-GEMM-scatter convolution; call sites
 * sender/encoder/codec_pipeline.py:273-354, receiver/decoder/codec_parallel.py:302-469).  The numpy
 * versions stay the readable statement; tests/test_cpu_oracle.py requires both to agree (kernel maps
 * exactly, convolutions to fp32 summation order).
 *
 *   gcc -O3 -mavx2 -ffp-contract=off -fopenmp -shared -fPIC sparse_ref.c -o _build/libsparse_ref.so
 *
 * -ffp-contract=off: every multiply and add is rounded on its own (no FMA), so a result does not depend
 * on how the compiler vectorises; each output row is owned by one thread and summed in a fixed order
 * (offsets ascending; inside an offset input channels ascending), so it does not depend on the thread
 * count either.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int64_t lex_key(int64_t b, int64_t x, int64_t y, int64_t z) {
  return (b << 48) | ((x + 32768) << 32) | ((y + 32768) << 16) | (z + 32768);
}

/* keys_in: ascending lexicographic keys of the input coordinates (row i = rank i).
 * nbr[j * n_out + o] = row of C_out[o] + sign * d_j in the input set, or -1.                        */
void ref_kernel_map(const int64_t* keys_in, int64_t n_in, const int32_t* c_out, int64_t n_out,
                    const int64_t* offsets, int32_t k, int32_t sign, int64_t* nbr) {
#pragma omp parallel for schedule(static)
  for (int64_t o = 0; o < n_out; ++o) {
    const int64_t b = c_out[4 * o], x = c_out[4 * o + 1], y = c_out[4 * o + 2], z = c_out[4 * o + 3];
    for (int32_t j = 0; j < k; ++j) {
      const int64_t qx = x + sign * offsets[3 * j], qy = y + sign * offsets[3 * j + 1], qz = z + sign * offsets[3 * j + 2];
      int64_t row = -1;
      if (qx >= -32768 && qx < 32768 && qy >= -32768 && qy < 32768 && qz >= -32768 && qz < 32768) {
        const int64_t key = lex_key(b, qx, qy, qz);
        int64_t lo = 0, hi = n_in;
        while (lo < hi) {
          const int64_t mid = (lo + hi) >> 1;
          if (keys_in[mid] < key) lo = mid + 1; else hi = mid;
        }
        if (lo < n_in && keys_in[lo] == key) row = lo;
      }
      nbr[(int64_t)j * n_out + o] = row;
    }
  }
}

/* out[o] = act(((sum_j dot_j) + bias) + residual[o]),  dot_j = F_in[nbr[j][o]] @ W[j]  (fp32).
 * nbr == NULL: identity rows, k must be 1.                                                          */
int ref_conv_forward(const float* f_in, int64_t n_in, int32_t cin, const int64_t* nbr, int64_t n_out, int32_t k,
                     const float* w, int32_t cout, const float* bias, const float* residual, int32_t relu, float* out) {
  if (cout > 1024 || cin <= 0 || cout <= 0) return -1;
  (void)n_in;
#pragma omp parallel
  {
    float acc[1024], dot[1024];
#pragma omp for schedule(dynamic, 256)
    for (int64_t o = 0; o < n_out; ++o) {
      for (int32_t c = 0; c < cout; ++c) acc[c] = 0.0f;
      for (int32_t j = 0; j < k; ++j) {
        const int64_t r = nbr ? nbr[(int64_t)j * n_out + o] : o;
        if (r < 0) continue;
        const float* f = f_in + r * cin;
        const float* wj = w + (int64_t)j * cin * cout;
        for (int32_t c = 0; c < cout; ++c) dot[c] = 0.0f;
        for (int32_t ci = 0; ci < cin; ++ci) {
          const float a = f[ci];
          const float* wr = wj + (int64_t)ci * cout;
          for (int32_t c = 0; c < cout; ++c) dot[c] += a * wr[c];
        }
        for (int32_t c = 0; c < cout; ++c) acc[c] += dot[c];
      }
      float* dst = out + o * cout;
      for (int32_t c = 0; c < cout; ++c) {
        float v = acc[c];
        if (bias) v += bias[c];
        if (residual) v += residual[o * cout + c];
        if (relu && !(v > 0.0f)) v = (v != v) ? v : 0.0f;   /* np.maximum propagates NaN */
        dst[c] = v;
      }
    }
  }
  return 0;
}

/* number of (offset, output row) pairs with a neighbour: the P of SURVEY 8d's flop count */
int64_t ref_count_pairs(const int64_t* nbr, int64_t n) {
  int64_t p = 0;
#pragma omp parallel for reduction(+ : p)
  for (int64_t i = 0; i < n; ++i) p += nbr[i] >= 0;
  return p;
}

/* float32 -> nearest bfloat16 (ties to even), kept as float32; non-finite values pass through */
void ref_bf16_round(const float* in, int64_t n, float* out) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    uint32_t u;
    memcpy(&u, in + i, 4);
    if ((u & 0x7F800000u) != 0x7F800000u) u = (u + 0x7FFFu + ((u >> 16) & 1u)) & 0xFFFF0000u;
    memcpy(out + i, &u, 4);
  }
}
