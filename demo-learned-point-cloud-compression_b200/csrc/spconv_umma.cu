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
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  // Bounded spin: a protocol bug traps (CUDA error) instead of hanging the GPU.
  for (uint32_t spin = 0; spin < (1u << 22); ++spin) {
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
  const int32_t* nbr;  // [kvol][n_out] or null (kvol == 1: row o reads input row o)
  const uint8_t* w;    // packed panels
  const float* bias;
  const __nv_bfloat16* residual;
  __nv_bfloat16* out_bf16;
  float* out_f32;
  int64_t n_in, n_out, ldr, ldo;
  int cout, kvol, relu, cstore;
  int stages, lag, wres, ntiles;
  uint32_t tmem_cols, acc_cols;
};

constexpr int kEpiWarps = 4, kMmaWarp = 4, kProdWarp0 = 5;
constexpr int kThreadsP = 32 * (kProdWarp0 + 4);  // 288

template <bool CA>
__device__ __forceinline__ void cp_async16x(uint32_t dst, const void* src, uint32_t src_bytes) {
  if (CA)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
  else
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_dyn(int n) {
  switch (n) {
    case 0: cp_async_wait<0>(); break;
    case 1: cp_async_wait<1>(); break;
    case 2: cp_async_wait<2>(); break;
    case 3: cp_async_wait<3>(); break;
    case 4: cp_async_wait<4>(); break;
    case 5: cp_async_wait<5>(); break;
    case 6: cp_async_wait<6>(); break;
    case 7: cp_async_wait<7>(); break;
    case 8: cp_async_wait<8>(); break;
    case 9: cp_async_wait<9>(); break;
    case 10: cp_async_wait<10>(); break;
    case 11: cp_async_wait<11>(); break;
    case 12: cp_async_wait<12>(); break;
    case 13: cp_async_wait<13>(); break;
    case 14: cp_async_wait<14>(); break;
    default: cp_async_wait<15>(); break;
  }
}
__device__ __forceinline__ void st_shared_u32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_shared_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void producer_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// Persistent, warp-specialised: warps 0-3 epilogue (TMEM lane quarter = warp), warp 4 MMA issuer +
// TMEM owner, warps 5-8 gather producers.  One CTA per SM walks tiles blockIdx.x, +gridDim.x, ...
// Shared memory: [weights: all kvol panels (resident) | per stage][A stages][nbr 2 x kvol x 128]
//                [stage meta][wmask 2x4][barriers].
template <int CIN, bool CA>
__global__ void __launch_bounds__(kThreadsP, 1)
k_spconv_umma(const UmmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int CPR = CIN / 8;     // 16-byte chunks per row
  constexpr int RPI = 32 / CPR;    // rows covered by one warp-wide cp.async
  constexpr uint32_t a_bytes = (uint32_t)kTileM * CIN * 2u;
  constexpr uint32_t lbo_a = kTileM * 16u;  // one K-chunk plane: 128 rows x 16 B
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int cout = p.cout, kvol = p.kvol, S = p.stages, lag = p.lag;
  const uint32_t b_bytes = (uint32_t)CIN * cout * 2u;
  uint32_t base = smem_u32(smem_raw);
  base = (base + 127u) & ~127u;
  const uint32_t w0 = base;
  const uint32_t w_total = p.wres ? (uint32_t)kvol * b_bytes : (uint32_t)S * b_bytes;
  const uint32_t a0 = w0 + w_total;
  const uint32_t nbr0 = a0 + (uint32_t)S * a_bytes;
  const uint32_t nbr_buf_bytes = p.nbr ? (uint32_t)kvol * kTileM * 4u : 0u;
  const uint32_t meta0 = nbr0 + 2u * nbr_buf_bytes;
  const uint32_t wmask0 = meta0 + 4u * S;
  const uint32_t bar0 = (wmask0 + 32u + 7u) & ~7u;
  const uint32_t full0 = bar0, empty0 = bar0 + 8u * S, tfull0 = bar0 + 16u * S, tempty0 = tfull0 + 16u,
                 wfull = tempty0 + 16u, tmem_slot = wfull + 8u;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8u * s, 128);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(tfull0, 1);
    mbar_init(tfull0 + 8u, 1);
    mbar_init(tempty0, 128);
    mbar_init(tempty0 + 8u, 128);
    mbar_init(wfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ld_shared_u32(tmem_slot);

  if (warp >= kProdWarp0) {
    // ---------------- gather producers ----------------
    const int pw = warp - kProdWarp0;
    const int sub = lane % CPR, rl0 = lane / CPR;
    if (p.wres && pw == 0 && lane == 0) {
      mbar_arrive_expect_tx(wfull, w_total);
      for (int j = 0; j < kvol; ++j) bulk_g2s(w0 + (uint32_t)j * b_bytes, p.w + (size_t)j * b_bytes, b_bytes, wfull);
    }
    auto prefetch_nbr = [&](int64_t tile, int buf) {
      const int64_t o = tile * kTileM + 32 * pw + lane;
      const uint32_t d0 = nbr0 + (uint32_t)buf * nbr_buf_bytes + (uint32_t)(32 * pw + lane) * 4u;
      if (o < p.n_out) {
        for (int j = 0; j < kvol; ++j) cp_async4(d0 + (uint32_t)j * (kTileM * 4u), p.nbr + (int64_t)j * p.n_out + o);
      } else {
        for (int j = 0; j < kvol; ++j) st_shared_u32(d0 + (uint32_t)j * (kTileM * 4u), 0xffffffffu);
      }
    };
    uint32_t g = 0, s = 0, ph = 0;        // stage counter, ring slot, ring pass
    uint32_t sig_s = 0;                   // ring slot of the next stage to signal
    int prev_nst = 1 << 30;
    if (p.nbr && (int64_t)blockIdx.x < p.ntiles) {
      prefetch_nbr(blockIdx.x, 0);
      cp_async_commit();
      cp_async_wait<0>();
      __syncwarp();
    }
    int ti = 0;
    for (int64_t tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++ti) {
      const int buf = ti & 1;
      const int64_t row0 = tile * kTileM;
      uint32_t mask = 1u;
      const uint32_t nb_base = nbr0 + (uint32_t)buf * nbr_buf_bytes;
      if (p.nbr) {
        if (ti > 0) {
          if (prev_nst <= lag) cp_async_wait<0>();
          __syncwarp();
        }
        uint32_t wm = 0;
        for (int j = 0; j < kvol; ++j) {
          const int v = (int)ld_shared_u32(nb_base + (uint32_t)(j * kTileM + 32 * pw + lane) * 4u);
          if (__any_sync(0xffffffffu, v >= 0)) wm |= 1u << j;
        }
        if (lane == 0) st_shared_u32(wmask0 + (uint32_t)(buf * 4 + pw) * 4u, wm);
        producer_bar();
        mask = ld_shared_u32(wmask0 + (uint32_t)(buf * 4) * 4u) | ld_shared_u32(wmask0 + (uint32_t)(buf * 4 + 1) * 4u) |
               ld_shared_u32(wmask0 + (uint32_t)(buf * 4 + 2) * 4u) | ld_shared_u32(wmask0 + (uint32_t)(buf * 4 + 3) * 4u);
        if (mask == 0) mask = 1u;  // a tile always issues at least one (all-zero) stage
        const int64_t next = tile + gridDim.x;
        if (next < p.ntiles) prefetch_nbr(next, buf ^ 1);
      }
      const int first_j = __ffs(mask) - 1, last_j = 31 - __clz(mask);
      for (uint32_t m = mask; m; m &= m - 1) {
        const int j = __ffs(m) - 1;
        if (ph > 0) mbar_wait(empty0 + 8u * s, (ph - 1) & 1u);
        const uint32_t a_s = a0 + s * a_bytes;
        if (pw == 0 && lane == 0) {
          if (!p.wres) {
            mbar_expect_tx(full0 + 8u * s, b_bytes);
            bulk_g2s(w0 + s * b_bytes, p.w + (size_t)j * b_bytes, b_bytes, full0 + 8u * s);
          }
          st_shared_u32(meta0 + 4u * s, (uint32_t)j | (j == first_j ? 256u : 0u) | (j == last_j ? 512u : 0u));
        }
#pragma unroll
        for (int it = 0; it < CPR; ++it) {
          const int row = 32 * pw + it * RPI + rl0;
          int64_t r;
          if (p.nbr) r = (int)ld_shared_u32(nb_base + (uint32_t)(j * kTileM + row) * 4u);
          else r = (row0 + row < p.n_out) ? row0 + row : -1;
          const bool ok = r >= 0 && r < p.n_in;
          const void* src = ok ? (const void*)(p.in + r * CIN + sub * 8) : (const void*)p.in;
          const uint32_t dst = a_s + (uint32_t)sub * lbo_a + (uint32_t)(row >> 3) * 128u + (uint32_t)(row & 7) * 16u;
          cp_async16x<CA>(dst, src, ok ? 16u : 0u);
        }
        cp_async_commit();
        if (g >= (uint32_t)lag) {
          cp_async_wait_dyn(lag);
          fence_proxy_async();
          mbar_arrive(full0 + 8u * sig_s);
          if (++sig_s == (uint32_t)S) sig_s = 0;
        }
        ++g;
        if (++s == (uint32_t)S) { s = 0; ++ph; }
      }
      prev_nst = __popc(mask);
    }
    // drain: everything landed, signal the stages not yet signalled
    cp_async_wait<0>();
    fence_proxy_async();
    const uint32_t pending = g < (uint32_t)lag ? g : (uint32_t)lag;
    for (uint32_t i = 0; i < pending; ++i) {
      mbar_arrive(full0 + 8u * sig_s);
      if (++sig_s == (uint32_t)S) sig_s = 0;
    }
  } else if (warp == kMmaWarp) {
    // ---------------- MMA issuer ----------------
    // instruction descriptor: c=f32 (bit 4), a=b=bf16 (bits 7,10), K-major both, N>>3 @17, M>>4 @24
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(cout >> 3) << 17) |
                           ((uint32_t)(kTileM >> 4) << 24);
    const uint32_t lbo_b = (uint32_t)cout * 16u;
    constexpr int ksteps = CIN >> 4;
    if (p.wres) mbar_wait(wfull, 0);
    uint32_t s = 0, ph = 0;
    int ti = 0;
    for (int64_t tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++ti) {
      const uint32_t acc = (uint32_t)(ti & 1);
      if (ti >= 2) mbar_wait(tempty0 + 8u * acc, (uint32_t)(((ti >> 1) - 1) & 1));
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * p.acc_cols;
      bool last = false;
      while (!last) {
        mbar_wait(full0 + 8u * s, ph & 1u);
        tc_fence_after();
        const uint32_t meta = ld_shared_u32(meta0 + 4u * s);
        const uint32_t j = meta & 0xffu;
        const bool first = (meta & 256u) != 0;
        last = (meta & 512u) != 0;
        if (lane == 0) {
          const uint32_t a_s = a0 + s * a_bytes;
          const uint32_t b_s = p.wres ? w0 + j * b_bytes : w0 + s * b_bytes;
#pragma unroll
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint64_t adesc = make_desc(a_s + (uint32_t)ks * 2u * lbo_a, lbo_a, 128u);
            const uint64_t bdesc = make_desc(b_s + (uint32_t)ks * 2u * lbo_b, lbo_b, 128u);
            umma_bf16(tmem_d, adesc, bdesc, idesc, (first && ks == 0) ? 0u : 1u);
          }
          umma_commit(empty0 + 8u * s);
          if (last) umma_commit(tfull0 + 8u * acc);
        }
        __syncwarp();
        if (++s == (uint32_t)S) { s = 0; ++ph; }
      }
    }
  } else {
    // ---------------- epilogue (warps 0-3; TMEM lane = tile row) ----------------
    int ti = 0;
    const int cstore = p.cstore > 0 ? p.cstore : cout;
    for (int64_t tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++ti) {
      const uint32_t acc = (uint32_t)(ti & 1);
      mbar_wait(tfull0 + 8u * acc, (uint32_t)((ti >> 1) & 1));
      tc_fence_after();
      const int64_t o = tile * kTileM + tid;
      const bool rowok = o < p.n_out;
      const uint32_t taddr = tmem_base + acc * p.acc_cols + ((uint32_t)(warp * 32) << 16);
      for (int c0 = 0; c0 < cout; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(taddr + (uint32_t)c0, v);
        if (rowok && c0 < cstore) {
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
          if (cstore < cout) {
            // narrow head: only the first cstore (< 16) columns exist in the output
#pragma unroll
            for (int i = 0; i < 15; ++i) {
              if (i < cstore) {
                if (p.out_f32) p.out_f32[o * p.ldo + i] = f[i];
                if (p.out_bf16) p.out_bf16[o * p.ldo + i] = __float2bfloat16_rn(f[i]);
              }
            }
          } else {
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
      }
      tc_fence_before();
      mbar_arrive(tempty0 + 8u * acc);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, p.tmem_cols);
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

extern "C++" {
static int g_spconv_cg = -1;  // PCC_SPCONV_CG=1: gather with cp.async.cg (L2 only) instead of .ca
static int g_max_stages = 16, g_lag_slack = 2;

template <int CIN>
static int launch_umma(const UmmaParams& p, size_t smem, unsigned grid, cudaStream_t st) {
  if (g_spconv_cg < 0) {
    const char* e = getenv("PCC_SPCONV_CG");
    g_spconv_cg = (e && e[0] == '1') ? 1 : 0;
  }
  static bool configured[2] = {false, false};
  if (g_spconv_cg) {
    if (!configured[0]) {
      PCC_CUDA(cudaFuncSetAttribute(k_spconv_umma<CIN, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
      configured[0] = true;
    }
    k_spconv_umma<CIN, false><<<grid, kThreadsP, smem, st>>>(p);
  } else {
    if (!configured[1]) {
      PCC_CUDA(cudaFuncSetAttribute(k_spconv_umma<CIN, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
      configured[1] = true;
    }
    k_spconv_umma<CIN, true><<<grid, kThreadsP, smem, st>>>(p);
  }
  PCC_LAUNCHED();
  return PCC_OK;
}
}  // extern "C++"

int pcc_spconv_set_tuning(int max_stages, int lag_slack, int use_cg) {
  PCC_CHECK_ARG(max_stages >= 1 && max_stages <= 16 && lag_slack >= 1 && lag_slack <= 8);
  g_max_stages = max_stages;
  g_lag_slack = lag_slack;
  g_spconv_cg = use_cg ? 1 : 0;
  return PCC_OK;
}

int pcc_spconv_umma(const void* in_bf16, int64_t n_in, int cin, const int32_t* nbr, int64_t n_out,
                    int kvol, const void* w_packed, const float* bias, int cout,
                    const void* residual_bf16, int64_t ldr, int relu, void* out_bf16,
                    float* out_f32, int64_t ldo, int cstore, pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0 && n_in >= 0 && kvol >= 1 && kvol <= 32);
  PCC_CHECK_ARG(cin == 16 || cin == 32 || cin == 64 || cin == 128 || cin == 256);
  PCC_CHECK_ARG(cout >= 16 && cout % 16 == 0 && cout <= 256);
  PCC_CHECK_ARG(nbr || kvol == 1);
  PCC_CHECK_ARG(out_bf16 || out_f32);
  PCC_CHECK_ARG(cstore >= 0 && cstore <= cout && (cstore == 0 || cstore == cout || cstore < 16));
  const int cs = (cstore > 0 && cstore < cout) ? cstore : cout;
  PCC_CHECK_ARG(ldo >= cs && (cs < cout || ldo % 8 == 0));
  PCC_CHECK_ARG(!residual_bf16 || (ldr >= cout && ldr % 8 == 0));
  if (n_out == 0) return PCC_OK;
  PCC_CHECK_ARG(in_bf16 && w_packed);
  PCC_CHECK_ARG(((uintptr_t)in_bf16 % 16) == 0 && ((uintptr_t)w_packed % 16) == 0);
  PCC_CHECK_ARG(cs < cout || !out_bf16 || ((uintptr_t)out_bf16 % 16) == 0);
  PCC_CHECK_ARG(cs < cout || !out_f32 || ((uintptr_t)out_f32 % 16) == 0);
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
  p.cout = cout;
  p.kvol = kvol;
  p.relu = relu;
  p.cstore = cs;
  // shared-memory plan: resident weights when all panels fit in 72 KB, else one panel per stage
  const int64_t a_bytes = (int64_t)kTileM * cin * 2, b_bytes = (int64_t)cin * cout * 2;
  const int64_t budget = 227 * 1024 - 1024;
  const int64_t nbr_bytes = nbr ? 2ll * kvol * kTileM * 4 : 0;
  p.wres = (kvol * b_bytes <= 72 * 1024) ? 1 : 0;
  const int64_t fixed = (p.wres ? kvol * b_bytes : 0) + nbr_bytes + 512;
  const int64_t stage_bytes = a_bytes + (p.wres ? 0 : b_bytes);
  int stages = (int)((budget - fixed) / (stage_bytes + 20));
  if (stages > g_max_stages) stages = g_max_stages;
  if (stages < 1) {
    set_error("pcc_spconv_umma: cin %d cout %d kvol %d does not fit shared memory", cin, cout, kvol);
    return PCC_E_UNSUPPORTED;
  }
  p.stages = stages;
  p.lag = stages <= 4 ? stages - 1 : stages - g_lag_slack;
  if (p.lag > 15) p.lag = 15;
  if (p.lag < 0) p.lag = 0;
  p.acc_cols = (uint32_t)((cout + 31) / 32 * 32);
  uint32_t cols = 32;
  while (cols < 2 * p.acc_cols) cols <<= 1;
  p.tmem_cols = cols;
  const int64_t ntiles = (n_out + kTileM - 1) / kTileM;
  p.ntiles = (int)ntiles;
  const size_t smem = (size_t)(fixed + (int64_t)stages * (stage_bytes + 20) + 256);
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    PCC_CUDA(cudaGetDevice(&dev));
    PCC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const unsigned grid = (unsigned)(ntiles < sms ? ntiles : sms);
  cudaStream_t st = as_stream(s);
  switch (cin) {
    case 16: return launch_umma<16>(p, smem, grid, st);
    case 32: return launch_umma<32>(p, smem, grid, st);
    case 64: return launch_umma<64>(p, smem, grid, st);
    case 128: return launch_umma<128>(p, smem, grid, st);
    default: return launch_umma<256>(p, smem, grid, st);
  }
}

}  // extern "C"
