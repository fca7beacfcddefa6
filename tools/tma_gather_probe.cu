// Micro-benchmark: how many rows per cycle can one SM pull through cp.async.bulk.tensor ... tile::gather4?
// (Decides whether the gather of spconv_umma should move from LDGSTS to the TMA engine.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/tma_gather_probe tools/tma_gather_probe.cu
//   gpurun_out/tma_gather_probe [rows] [row_bytes]
// Every warp keeps a ring of S slots; a slot is 32 gather4 requests (one per lane, 4 rows each) tracked by
// one mbarrier.  Index patterns: sequential, neighbour-like (row + small offset, 30 % absent) and random.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                         \
    }                                                                                  \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  for (uint32_t spin = 0; spin < (1u << 24); ++spin) {
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
  }
  __trap();
}

template <int ROW_BYTES, int S>
__global__ void __launch_bounds__(1024, 1)
k_probe(const __grid_constant__ CUtensorMap tm, const int4* __restrict__ idx, int iters, uint32_t* __restrict__ sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
  uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  const uint32_t slot_bytes = 32u * 4u * ROW_BYTES;
  // probe only: with many warps the data regions alias (contents are irrelevant here)
  const uint32_t region = 192u * 1024u;
  const uint32_t data0 = base + ((uint32_t)warp * S * slot_bytes) % region;
  const uint32_t bar0 = base + region + (uint32_t)warp * S * 8u;
  if (lane == 0)
    for (int s = 0; s < S; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u * s) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const int4* my = idx + ((int64_t)(blockIdx.x * W + warp) * iters) * 32 + lane;
  int4 nxt = __ldg(my);
  for (int it = 0; it < iters; ++it) {
    const int s = it % S;
    const int4 cur = nxt;
    if (it + 1 < iters) nxt = __ldg(my + (int64_t)(it + 1) * 32);
    if (it >= S) mbar_wait(bar0 + 8u * s, (uint32_t)((it / S - 1) & 1));
    if (lane == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar0 + 8u * s), "r"(slot_bytes) : "memory");
    __syncwarp();
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
            base + (data0 - base + (uint32_t)s * slot_bytes) % region + (uint32_t)lane * 4u * ROW_BYTES),
        "l"(&tm), "r"(0), "r"(cur.x), "r"(cur.y), "r"(cur.z), "r"(cur.w), "r"(bar0 + 8u * s)
        : "memory");
  }
  // drain
  for (int it = (iters > S ? iters - S : 0); it < iters; ++it) mbar_wait(bar0 + 8u * (it % S), (uint32_t)((it / S) & 1));
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(base) : "memory");
    sink[blockIdx.x] = v;
  }
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int ROW_BYTES>
static void run(int64_t n_rows, EncodeTiled enc) {
  constexpr int S = 4;
  void* data;
  CK(cudaMalloc(&data, n_rows * ROW_BYTES));
  CK(cudaMemset(data, 1, n_rows * ROW_BYTES));
  CUtensorMap tm;
  cuuint64_t gdim[2] = {(cuuint64_t)(ROW_BYTES / 2), (cuuint64_t)n_rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ROW_BYTES};
  cuuint32_t box[2] = {(cuuint32_t)(ROW_BYTES / 2), 1};
  cuuint32_t est[2] = {1, 1};
  CUtensorMapSwizzle sw = ROW_BYTES == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (ROW_BYTES == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, data, gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
    exit(1);
  }
  int sms = 148;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  uint32_t* sink;
  CK(cudaMalloc(&sink, sms * 4));
  const int warps_list[] = {4, 8, 16, 32};
  const char* pat_name[] = {"sequential", "neighbour", "random", "zero-rows"};
  for (int pat = 0; pat < 4; ++pat) {
    for (int W : warps_list) {
      // total rows requested ~ 27 * n_rows / 4 (a k3 layer's worth), split over SMs and warps
      const int iters = (int)((n_rows * 6) / ((int64_t)sms * W * 128));
      const int64_t nreq = (int64_t)sms * W * iters * 32;
      std::vector<int4> h(nreq);
      uint64_t rng = 0x9e3779b97f4a7c15ull;
      auto rnd = [&]() { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return rng; };
      const int deltas[8] = {0, 1, -1, 9, -9, 70, -70, 600};
      for (int64_t q = 0; q < nreq; ++q) {
        int v[4];
        for (int k = 0; k < 4; ++k) {
          const int64_t o = (q * 4 + k) % n_rows;
          if (pat == 0) v[k] = (int)o;
          else if (pat == 1) {
            const int64_t t = o + deltas[(q / 8192) & 7];
            v[k] = (rnd() % 10 < 3 || t < 0 || t >= n_rows) ? -1 : (int)t;
          } else if (pat == 2) v[k] = (int)(rnd() % (uint64_t)n_rows);
          else {   // absent entries point at one of 4096 (zero) rows at the end of the tensor instead of out of bounds
            const int64_t t = o + deltas[(q / 8192) & 7];
            v[k] = (rnd() % 10 < 3 || t < 0 || t >= n_rows) ? (int)(n_rows - 1 - (int64_t)(rnd() % 4096)) : (int)t;
          }
        }
        h[q] = make_int4(v[0], v[1], v[2], v[3]);
      }
      int4* d;
      CK(cudaMalloc(&d, nreq * sizeof(int4)));
      CK(cudaMemcpy(d, h.data(), nreq * sizeof(int4), cudaMemcpyHostToDevice));
      const size_t smem = 192 * 1024 + 32 * S * 8 + 2048;
      CK(cudaFuncSetAttribute(k_probe<ROW_BYTES, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      cudaEvent_t e0, e1;
      CK(cudaEventCreate(&e0));
      CK(cudaEventCreate(&e1));
      float best = 1e30f;
      for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        k_probe<ROW_BYTES, S><<<sms, 32 * W, smem>>>(tm, d, iters, sink);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
      }
      CK(cudaGetLastError());
      const double rows = (double)nreq * 4;
      printf("row %3d B  %-10s warps %d: %8.1f us  %7.1f Mrows/ms  %6.2f cycles/row/SM @1.965GHz  %7.1f GB/s\n", ROW_BYTES,
             pat_name[pat], W, best * 1e3, rows / best * 1e-6, best * 1e-3 * 1.965e9 * sms / rows, rows * ROW_BYTES / best * 1e-6);
      CK(cudaFree(d));
    }
  }
  CK(cudaFree(data));
  CK(cudaFree(sink));
}

int main(int argc, char** argv) {
  const int64_t n_rows = argc > 1 ? atoll(argv[1]) : 1787336;
  EncodeTiled enc = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q));
  if (!enc) {
    printf("no cuTensorMapEncodeTiled\n");
    return 1;
  }
  run<64>(n_rows, enc);
  run<128>(n_rows, enc);
  run<32>(n_rows, enc);
  return 0;
}
