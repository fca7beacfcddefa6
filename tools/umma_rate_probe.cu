// Probe: cost of one tcgen05.mma (M = 128, K = 16, bf16, operands in shared memory, no-swizzle K-major core matrices) as a
// function of N, for a chain into ONE accumulator and for a rotation over several accumulators, issued by one thread.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I demo-learned-point-cloud-compression_b200/csrc tools/umma_rate_probe.cu -o tools/_build/umma_rate_probe
#include <cstdio>
#include <cstdlib>
#include "umma_helpers.cuh"

namespace pcc { void set_error(const char*, ...) {} std::atomic<int64_t> g_launches{0}; }
using namespace pcc;

// lbo_a: 128 or 144; accs: number of accumulators rotated over (1 = dependent chain); issuers: 1..4 warps issuing their own chains
__global__ void __launch_bounds__(256, 1) k_probe(int n, int iters, int accs, int lbo_a, int issuers, int a_slices, int commit_every, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[4];
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem) + 127u) & ~127u;
  const uint32_t a0 = base, b0 = base + 96 * 1024;       // A region 96 KB, B after it
  // commit_every = ce > 0: a tcgen05.commit after every ce MMAs (tight variant); the barrier expects all of them
  if (threadIdx.x == 0) for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar[i]), commit_every > 0 ? (uint32_t)(iters / commit_every) + 1u : 1u);
  for (int i = threadIdx.x; i < (160 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem + (base - smem_u32(smem)))[i] = 0x3c003c00u;
  if (warp == 0) tmem_alloc(smem_u32(&tslot), 512);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tslot;
  if (warp < issuers && lane == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t sbo_a = 4 * lbo_a;
    const uint64_t ad = make_desc(a0, lbo_a, sbo_a), bd = make_desc(b0, 27 * 32 * 16, 128);
    const uint32_t cols_per = 512 / (issuers * accs) < (uint32_t)n ? (uint32_t)n : 512 / (issuers * accs);
    const uint32_t slice = (16 * sbo_a + 64) >> 4;
    long long t0 = clock64();
    if (a_slices == 0) {
      // tight variant: 16 MMAs per iteration, every offset an immediate, one accumulator per issuer
      const uint32_t td = tmem + (uint32_t)warp * 128u;
      for (int i = 0; i < iters / 16; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          umma_bf16(td + (accs > 1 ? (k & 1) * 64 : 0), ad + (uint64_t)(k * 18), bd + (uint64_t)(k * 32), idesc, 1u);
          if (commit_every > 0 && ((k + 1) % commit_every) == 0) umma_commit(smem_u32(&bar[warp]));
        }
      }
    } else
    for (int i = 0; i < iters; ++i) {
      const uint32_t acc = (uint32_t)(i % accs);
      const uint32_t td = tmem + ((uint32_t)warp * accs + acc) * cols_per % 512;
      umma_bf16(td, ad + (uint64_t)((i % a_slices) * slice), bd + (uint64_t)((i & 7) * 32), idesc, i >= accs ? 1u : 0u);
    }
    umma_commit(smem_u32(&bar[warp]));
    WaitCtx wc{smem_u32(&tslot), 1u << 24, nullptr};
    while (!mbar_try(smem_u32(&bar[warp]), 0)) {}
    long long t1 = clock64();
    cycles[warp] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 2048;
  printf("%5s %5s %5s %7s %8s  cycles/MMA (per issuer)\n", "N", "accs", "lbo", "issuers", "aslices");
  for (int lbo : {144})
    for (int issuers : {1, 4})
      for (int accs : {1, 2, 4})
        for (int n : {16, 32, 64, 128, 256}) {
          if (issuers * accs * n > 512) continue;
          for (int as : {0, 1}) {
            if (as == 0 && (accs > 2 || n > 64)) continue;
            k_probe<<<1, 256, 200 * 1024>>>(n, iters, accs, lbo, issuers, as, 0, d);
            long long h[4] = {0, 0, 0, 0};
            cudaError_t e = cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            printf("%5d %5d %5d %7d %8d  %8.1f\n", n, accs, lbo, issuers, as, (double)h[0] / iters);
          }
        }
  printf("\ncommit overhead (tight variant, lbo 144): %5s %7s %6s  cycles/MMA (per issuer)\n", "N", "issuers", "every");
  for (int issuers : {1, 4})
    for (int n : {32, 64})
      for (int ce : {0, 16, 8, 4, 2, 1}) {
        k_probe<<<1, 256, 200 * 1024>>>(n, iters, 1, 144, issuers, 0, ce, d);
        long long h[4] = {0, 0, 0, 0};
        cudaError_t e = cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        printf("%47d %7d %6d  %8.1f\n", n, issuers, ce, (double)h[0] / iters);
      }
  return 0;
}
