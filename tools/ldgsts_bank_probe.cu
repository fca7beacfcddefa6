// Probe: what makes ncu count l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ldgsts for 64-byte row gathers?
// Each group of 4 lanes copies one 64-byte global row (4 x cp.async.ca 16 B) into shared memory.  Variants:
//   dst_mode 0: rows contiguous (row slot s at s * 64: a 16-bank window aligned to a half of the 32 banks)
//   dst_mode 1: the kernel's core-matrix layout (chunk k at k * 144 + (s & 7) * 16 + (s >> 3) * 576)
//   src_par  0: even source rows only (offset 0 of a 128-byte line), 1: odd rows only, 2: random
//   slot_par 0: only even row slots written, 1: only odd, 2: all
// One kernel launch per variant; read the counters with
//   ncu --metrics l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ldgsts.sum,smsp__sass_l1tex_data_pipe_lsu_wavefronts_mem_shared_op_ldgsts.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

__global__ void __launch_bounds__(512, 1)
k_probe(const uint8_t* __restrict__ in, int64_t n_rows, int iters, int dst_mode, int src_par, int slot_par, uint32_t* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, k = lane & 3, g = lane >> 2;
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem) + warp * 8192u;
  uint64_t rng = 0x9e3779b97f4a7c15ull * (blockIdx.x * 16 + warp + 1);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      rng = rng * 6364136223846793005ull + 1442695040888963407ull;
      // one row per 4-lane group: all lanes of the group must agree -> derive from the group index
      uint64_t h = (rng >> 17) + (uint64_t)g * 0x632be59bd9b4e019ull;
      h ^= h >> 29;
      uint64_t r = h % (uint64_t)(n_rows / 2) * 2;
      if (src_par == 1) r += 1;
      if (src_par == 2) r += (h >> 40) & 1;
      int s = u * 8 + g;      // row slot 0..63
      if (slot_par == 0) s &= ~1;
      if (slot_par == 1) s |= 1;
      const uint32_t dst = dst_mode == 0 ? base + (uint32_t)s * 64u + (uint32_t)k * 16u
                                         : base + (uint32_t)k * 144u + (uint32_t)(s & 7) * 16u + (uint32_t)(s >> 3) * 576u;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(in + r * 64 + k * 16) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) sink[blockIdx.x] = smem[5];
}

int main() {
  const int64_t n_rows = 1787336;
  uint8_t* in;
  cudaMalloc(&in, n_rows * 64);
  cudaMemset(in, 1, n_rows * 64);
  uint32_t* sink;
  cudaMalloc(&sink, 148 * 4);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16 * 8192);
  for (int dst_mode = 0; dst_mode < 2; ++dst_mode)
    for (int src_par = 0; src_par < 3; ++src_par)
      for (int slot_par = 0; slot_par < 3; ++slot_par) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0);
        k_probe<<<148, 512, 16 * 8192>>>(in, n_rows, 400, dst_mode, src_par, slot_par, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("dst_mode %d src_par %d slot_par %d: %.1f us (%lld rows)\n", dst_mode, src_par, slot_par, ms * 1e3,
               (long long)148 * 16 * 400 * 64);
      }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
