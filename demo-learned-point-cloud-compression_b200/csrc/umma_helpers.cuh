// PTX helpers shared by the tcgen05 convolution kernels (spconv_umma.cu, spconv_octet.cu): mbarriers with abortable
// bounded waits, cp.async / bulk copies, TMEM allocation, tcgen05.mma / commit / ld, shared-memory descriptors.
#pragma once
#include "common.cuh"

namespace pcc {

constexpr int kTileM = 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Waiting on a pipeline barrier never hangs and never traps: the spin is bounded, and a waiter that runs out of
// patience raises the CTA's abort flag (shared memory) and records PCC_FAULT_PIPELINE_TIMEOUT in the device fault
// word (common.cuh).  Every waiter polls the flag, so all roles of the CTA leave their loops within microseconds and
// the kernel ends normally -- its tiles stay unwritten and the host learns about it from the fault word.  A timeout
// means a protocol bug (or a profiler / time-slicing pause of several seconds), not a data-dependent condition.
struct WaitCtx {
  uint32_t abort_flag;   // shared-memory address of the CTA's abort word
  uint32_t spin_limit;
  int* fault;
  uint32_t exp = 0;      // debug experiments (spconv_octet.cu, TRACE instantiation only)
};
__device__ __forceinline__ uint32_t ld_shared_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok;
}
__device__ __forceinline__ uint32_t mbar_try_hint(uint32_t bar, uint32_t parity, uint32_t hint) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity), "r"(hint)
      : "memory");
  return ok;
}
// The patient part, out of line: it only runs when a wait has already lasted kFastTries tries.
static __device__ __noinline__ bool mbar_wait_slow(uint32_t bar, uint32_t parity, uint32_t abort_flag, uint32_t spin_limit, int* fault) {
  for (uint32_t spin = 0; spin < spin_limit; ++spin) {
    if (mbar_try(bar, parity)) return true;
    if ((spin & 7u) == 7u && ld_shared_u32(abort_flag) != 0u) return false;
  }
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(abort_flag), "r"(1u) : "memory");
  atomicMax(fault, PCC_FAULT_PIPELINE_TIMEOUT);
  return false;
}
constexpr uint32_t kFastTries = 64;   // inline tries before a wait counts as slow (a try suspends up to a hardware time limit)
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity, const WaitCtx& w) {
#pragma unroll 1
  for (uint32_t spin = 0; spin < kFastTries; ++spin)
    if (mbar_try(bar, parity)) return true;
  return mbar_wait_slow(bar, parity, w.abort_flag, w.spin_limit, w.fault);
}
// Same, for waiters that are not on the critical path: the suspend-time hint keeps the polling
// (one shared-memory wavefront per try) off the L1/shared data pipe the gather is bound by.
__device__ __forceinline__ bool mbar_wait_relaxed(uint32_t bar, uint32_t parity, const WaitCtx& w) {
#pragma unroll 1
  for (uint32_t spin = 0; spin < kFastTries; ++spin)
    if (mbar_try_hint(bar, parity, 2000u)) return true;
  return mbar_wait_slow(bar, parity, w.abort_flag, w.spin_limit, w.fault);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
        "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, no-swizzle shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// [0,14) start>>4, [16,30) leading-dim byte offset>>4 (between K-adjacent 8x16B core matrices),
// [32,46) stride byte offset>>4 (between 8-row groups), [46,48) version=1, [61,64) layout=0.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);
}


}  // namespace pcc
