// (2) Output-stationary gather-GEMM sparse convolution on tcgen05 (sm_100a).
//
// One CTA owns a tile of 128 output rows.  For every kernel offset j (fixed order 0..kvol-1) the 128
// gathered input rows (zero rows where the neighbour is absent) form the A operand [128 x cin] and
// the weight panel W[j] is the B operand [cout x cin] (K-major).  tcgen05.mma accumulates all offsets
// into one fp32 TMEM tile [128 lanes x cout columns]; the epilogue adds bias / residual, applies ReLU
// and writes each output row exactly once.  No atomics and a fixed accumulation order make results
// run-to-run deterministic and independent of the order rows were inserted in (SURVEY 3.2 inv. 3).
//
// Warp roles (160 threads): warps 0-3 gather (cp.async 16-byte chunks with zero fill into the
// no-swizzle K-major core-matrix layout; the weight panel arrives by one cp.async.bulk), then become
// the epilogue warps (TMEM lane = output row); warp 4 allocates TMEM and its lane 0 issues the MMAs.
// Stages are recycled through full/empty mbarriers; tcgen05.commit releases a stage.
#include "common.cuh"

namespace pcc {

constexpr int kTileM = 128;
constexpr int kProducerThreads = 128;
constexpr int kThreadsUmma = 160;

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
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  // Bounded spin: a protocol bug traps (CUDA error) instead of hanging the GPU.
  for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
  }
  __trap();
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

struct UmmaParams {
  const __nv_bfloat16* in;
  const int32_t* nbr;  // [kvol][n_out] or null
  const uint8_t* w;    // packed panels
  const float* bias;
  const __nv_bfloat16* residual;
  __nv_bfloat16* out_bf16;
  float* out_f32;
  int64_t n_in, n_out, ldr, ldo;
  int cin, cout, kvol, relu;
  int stages, lag;
  uint32_t tmem_cols;
};

template <int LAG>
__device__ __forceinline__ void producer_signal(int j, int stages, uint32_t full0) {
  if (j >= LAG) {
    cp_async_wait<LAG>();
    fence_proxy_async();
    mbar_arrive(full0 + 8u * (uint32_t)((j - LAG) % stages));
  }
}

__global__ void __launch_bounds__(kThreadsUmma)
k_spconv_umma(const UmmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int cin = p.cin, cout = p.cout, kvol = p.kvol, S = p.stages;
  const uint32_t a_bytes = (uint32_t)kTileM * cin * 2u;
  const uint32_t b_bytes = (uint32_t)cin * cout * 2u;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  uint32_t base = smem_u32(smem_raw);
  base = (base + 1023u) & ~1023u;
  const uint32_t bar0 = base + (uint32_t)S * stage_bytes;  // full[S], empty[S], accum, tmem slot
  const uint32_t full0 = bar0, empty0 = bar0 + 8u * S, accum_bar = bar0 + 16u * S;
  const uint32_t tmem_slot = accum_bar + 8u;
  const int64_t row0 = (int64_t)blockIdx.x * kTileM;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8u * s, kProducerThreads);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp < 4) {
    // ---------------- gather producers ----------------
    const int cpr = cin >> 3;              // 16-byte chunks per row
    const int chunks = kTileM * cpr;
    const uint32_t lbo_a = kTileM * 16u;   // one K-chunk plane: 128 rows x 16 B
    for (int j = 0; j < kvol; ++j) {
      const int s = j % S, it = j / S;
      if (it > 0) mbar_wait(empty0 + 8u * s, (uint32_t)((it - 1) & 1));
      const uint32_t a_s = base + (uint32_t)s * stage_bytes;
      if (tid == 0) {
        mbar_expect_tx(full0 + 8u * s, b_bytes);
        bulk_g2s(a_s + a_bytes, p.w + (size_t)j * b_bytes, b_bytes, full0 + 8u * s);
      }
      const int32_t* nb = p.nbr ? p.nbr + (int64_t)j * p.n_out : nullptr;
      for (int c = tid; c < chunks; c += kProducerThreads) {
        const int row = c / cpr, kc = c - row * cpr;
        const int64_t o = row0 + row;
        int64_t r = -1;
        if (o < p.n_out) r = nb ? (int64_t)__ldg(nb + o) : o;
        if (r >= p.n_in) r = -1;
        const void* src = r >= 0 ? (const void*)(p.in + r * cin + kc * 8) : (const void*)p.in;
        const uint32_t dst = a_s + (uint32_t)kc * lbo_a + (uint32_t)(row >> 3) * 128u + (uint32_t)(row & 7) * 16u;
        cp_async16(dst, src, r >= 0 ? 16u : 0u);
      }
      cp_async_commit();
      if (p.lag == 2) producer_signal<2>(j, S, full0);
      else if (p.lag == 1) producer_signal<1>(j, S, full0);
      else producer_signal<0>(j, S, full0);
    }
    // drain: everything landed, signal the stages not yet signalled
    cp_async_wait<0>();
    fence_proxy_async();
    for (int j = (kvol - p.lag > 0 ? kvol - p.lag : 0); j < kvol; ++j) mbar_arrive(full0 + 8u * (uint32_t)(j % S));

    // ---------------- epilogue ----------------
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    const int64_t o = row0 + tid;
    const bool rowok = o < p.n_out;
    for (int c0 = 0; c0 < cout; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
      if (rowok) {
        float f[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
        if (p.bias) {
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] += __ldg(p.bias + c0 + i);
        }
        if (p.residual) {
          const uint4* rp = reinterpret_cast<const uint4*>(p.residual + o * p.ldr + c0);
          uint4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
          uint32_t rw[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            f[2 * i] += __uint_as_float(rw[i] << 16);
            f[2 * i + 1] += __uint_as_float(rw[i] & 0xffff0000u);
          }
        }
        if (p.relu) {
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
        }
        if (p.out_f32) {
          float4* op = reinterpret_cast<float4*>(p.out_f32 + o * p.ldo + c0);
#pragma unroll
          for (int i = 0; i < 4; ++i) op[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
        }
        if (p.out_bf16) {
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
            pk[i] = *reinterpret_cast<uint32_t*>(&h);
          }
          uint4* op = reinterpret_cast<uint4*>(p.out_bf16 + o * p.ldo + c0);
          op[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          op[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
      }
    }
  } else {
    // ---------------- MMA issuer (warp 4) ----------------
    // instruction descriptor: c=f32 (bit 4), a=b=bf16 (bits 7,10), K-major both, N>>3 @17, M>>4 @24
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(cout >> 3) << 17) |
                           ((uint32_t)(kTileM >> 4) << 24);
    const uint32_t lbo_a = kTileM * 16u, lbo_b = (uint32_t)cout * 16u;
    const int ksteps = cin >> 4;
    for (int j = 0; j < kvol; ++j) {
      const int s = j % S, it = j / S;
      mbar_wait(full0 + 8u * s, (uint32_t)(it & 1));
      tc_fence_after();
      if (lane == 0) {
        const uint32_t a_s = base + (uint32_t)s * stage_bytes;
        const uint32_t b_s = a_s + a_bytes;
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t adesc = make_desc(a_s + (uint32_t)ks * 2u * lbo_a, lbo_a, 128u);
          const uint64_t bdesc = make_desc(b_s + (uint32_t)ks * 2u * lbo_b, lbo_b, 128u);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (j | ks) != 0 ? 1u : 0u);
        }
        umma_commit(empty0 + 8u * s);
        if (j == kvol - 1) umma_commit(accum_bar);
      }
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, p.tmem_cols);
}

static inline uint16_t f32_to_bf16_rn(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

}  // namespace pcc

using namespace pcc;

extern "C" {

int64_t pcc_spconv_packed_weight_bytes(int kvol, int cin_pad, int cout) {
  return (int64_t)kvol * cin_pad * cout * 2;
}

int pcc_spconv_pack_weights(const float* w, int kvol, int cin, int cout, int cin_pad,
                            void* packed_host) {
  PCC_CHECK_ARG(w && packed_host && kvol >= 1 && cin >= 1 && cout >= 1);
  PCC_CHECK_ARG(cin_pad >= cin && cin_pad % 16 == 0 && cout % 8 == 0);
  uint16_t* out = (uint16_t*)packed_host;
  const int kcs = cin_pad / 8;
  for (int j = 0; j < kvol; ++j)
    for (int kc = 0; kc < kcs; ++kc)
      for (int co = 0; co < cout; ++co)
        for (int e = 0; e < 8; ++e) {
          int ci = kc * 8 + e;
          float v = ci < cin ? w[((int64_t)j * cin + ci) * cout + co] : 0.f;
          out[(((int64_t)j * kcs + kc) * cout + co) * 8 + e] = f32_to_bf16_rn(v);
        }
  return PCC_OK;
}

int pcc_spconv_umma(const void* in_bf16, int64_t n_in, int cin, const int32_t* nbr, int64_t n_out,
                    int kvol, const void* w_packed, const float* bias, int cout,
                    const void* residual_bf16, int64_t ldr, int relu, void* out_bf16,
                    float* out_f32, int64_t ldo, pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0 && n_in >= 0 && kvol >= 1);
  PCC_CHECK_ARG(cin >= 16 && cin % 16 == 0 && cin <= 256);
  PCC_CHECK_ARG(cout >= 16 && cout % 16 == 0 && cout <= 256);
  PCC_CHECK_ARG(nbr || kvol == 1);
  PCC_CHECK_ARG(out_bf16 || out_f32);
  PCC_CHECK_ARG(ldo >= cout && ldo % 8 == 0);
  PCC_CHECK_ARG(!residual_bf16 || (ldr >= cout && ldr % 8 == 0));
  if (n_out == 0) return PCC_OK;
  PCC_CHECK_ARG(in_bf16 && w_packed);
  PCC_CHECK_ARG(((uintptr_t)in_bf16 % 16) == 0 && ((uintptr_t)w_packed % 16) == 0);
  PCC_CHECK_ARG(!out_bf16 || ((uintptr_t)out_bf16 % 16) == 0);
  PCC_CHECK_ARG(!out_f32 || ((uintptr_t)out_f32 % 16) == 0);
  PCC_CHECK_ARG(!residual_bf16 || ((uintptr_t)residual_bf16 % 16) == 0);

  UmmaParams p;
  p.in = (const __nv_bfloat16*)in_bf16;
  p.nbr = nbr;
  p.w = (const uint8_t*)w_packed;
  p.bias = bias;
  p.residual = (const __nv_bfloat16*)residual_bf16;
  p.out_bf16 = (__nv_bfloat16*)out_bf16;
  p.out_f32 = out_f32;
  p.n_in = n_in;
  p.n_out = n_out;
  p.ldr = ldr;
  p.ldo = ldo;
  p.cin = cin;
  p.cout = cout;
  p.kvol = kvol;
  p.relu = relu;
  const int64_t stage_bytes = (int64_t)kTileM * cin * 2 + (int64_t)cin * cout * 2;
  int stages = (int)((200 * 1024) / stage_bytes);
  if (stages > 4) stages = 4;
  if (stages > kvol) stages = kvol;
  if (stages < 1) {
    set_error("pcc_spconv_umma: stage of %lld bytes does not fit shared memory", (long long)stage_bytes);
    return PCC_E_UNSUPPORTED;
  }
  p.stages = stages;
  p.lag = stages >= 4 ? 2 : (stages == 3 ? 1 : 0);
  uint32_t cols = 32;
  while ((int)cols < cout) cols <<= 1;
  p.tmem_cols = cols;
  const size_t smem = (size_t)stages * stage_bytes + 1024 + 16 * stages + 32;
  static size_t configured = 0;
  if (smem > configured) {
    PCC_CUDA(cudaFuncSetAttribute(k_spconv_umma, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(227 * 1024)));
    configured = 227 * 1024;
  }
  const unsigned grid = (unsigned)((n_out + kTileM - 1) / kTileM);
  k_spconv_umma<<<grid, kThreadsUmma, smem, as_stream(s)>>>(p);
  PCC_LAUNCHED();
  return PCC_OK;
}

}  // extern "C"
