// (2) Output-stationary gather-GEMM sparse convolution on tcgen05 (sm_100a).
//
// out[o] = act(bias + sum_j in[nbr[j][o]] . W[j] (+ residual[o])): for every kernel offset j (fixed order
// 0..kvol-1) the gathered input rows of a 128-row output tile (zero rows where the neighbour is absent)
// form the A operand [128 x cin] and the weight panel W[j] the B operand [cout x cin] (K-major).
// tcgen05.mma accumulates all offsets into fp32 TMEM tiles [128 lanes x cout columns]; the epilogue
// adds bias / residual, applies ReLU and writes each output row exactly once.  No atomics and a fixed
// accumulation order make results run-to-run deterministic and independent of the order rows were
// inserted in (SURVEY 3.2 inv. 3).  The kernel is persistent and warp specialised; its structure is
// described above k_spconv_umma below.
#include "common.cuh"
#include "umma_helpers.cuh"

namespace pcc {

struct UmmaParams {
  const __nv_bfloat16* in;
  const int32_t* nbr;  // [kvol][nbr_ld] or null (kvol == 1: row o reads input row o)
  const uint8_t* w;    // packed panels
  const float* bias;
  const __nv_bfloat16* residual;
  __nv_bfloat16* out_bf16;
  float* out_f32;
  int64_t n_in, n_out, ldr, ldo, nbr_ld;
  int cout, kvol, relu, cstore;
  int wres, ntiles, nbr_bulk;
  int lt;              // log2(T): tiles gathered / accumulated together per stage
  int nrounds_total;   // ceil(ntiles / T)
  int jc;              // kernel offsets per kernel-map unit (divides kvol)
  uint32_t tmem_cols, acc_cols;
  int* fault;          // device fault word (common.cuh)
  uint32_t spin_limit; // barrier patience (tries)
  int inject;          // test hook: drop one hand-off so that the timeout path runs
};

// warps 4..7 issue MMAs: one elected thread each, every warp owning the tiles t = w (mod 4) of a
// stage.  A single issuing thread was the bottleneck of the kernel (ncu r1g: the gather warps spent
// 90 % of their time waiting for their slot to be released while the lone MMA warp never waited):
// building two descriptors and issuing one tcgen05.mma costs ~100 issue cycles in the uniform
// datapath, i.e. ~1500 cycles for the 8 MMAs + commits of a T = 4 stage that the gather fills in ~500.
constexpr int kMmaWarp = 4, kMmaWarps = 4, kLoaderWarp = 8, kProdWarp0 = 9;
constexpr int kNbrBufs = 3;   // kernel-map ring depth (units)

// One 16-byte chunk of a gathered row, global -> shared (LDGSTS).  `r` is the map entry (source row, or
// >= n_in when the neighbour is absent), `base` the input + the chunk's byte offset.  A present row is
// copied; an absent one writes zeros (ignore-src form) and only if `dirty_bit` says the slot still holds
// data from the previous use of the stage.  Written in PTX so the predicate logic stays in predicates
// (ISETP, LOP3.P, ISETP.OR, SEL, address, LDGSTS) instead of being materialised as integers.
// .ca on purpose: with .cg (L1 bypass) the 1.79 M-row c32 layer takes 571 us instead of 363 us
template <int ROW_BYTES>
__device__ __forceinline__ void gather_chunk(uint32_t dst, uint64_t base, uint32_t r, uint32_t n_in, uint32_t dirty_bit) {
  asm volatile(
      "{\n\t.reg .pred ok, q, ig;\n\t.reg .b32 rc;\n\t.reg .b64 src;\n\t"
      "setp.lt.u32 ok, %2, %3;\n\t"
      "setp.ne.or.u32 q, %4, 0, ok;\n\t"
      "not.pred ig, ok;\n\t"
      "selp.b32 rc, %2, 0, ok;\n\t"
      "mad.wide.u32 src, rc, %5, %1;\n\t"
      "@q cp.async.ca.shared.global [%0], [src], 16, ig;\n\t}" ::"r"(dst),
      "l"(base), "r"(r), "r"(n_in), "r"(dirty_bit), "n"(ROW_BYTES)
      : "memory");
}
// Persistent (one CTA per SM), warp specialised.  A CTA works in rounds: round r of CTA b covers the
// super-tile of T = 2^lt consecutive 128-row tiles starting at tile T*(b + r*gridDim.x).  One pipeline
// stage = (round, kernel offset j): the gathered rows of all T tiles for that offset.  The T tiles
// have their own TMEM accumulators and the MMAs of a stage are interleaved over them, because with
// N = cout as small as 32 a chain of accumulating tcgen05.mma into ONE accumulator runs at the MMA
// pipeline latency, not at its throughput; the per-stage hand-off cost is amortised over T tiles.
//   warps 0-3   epilogue (TMEM lane quarter = warp): TMEM -> bias/residual/ReLU -> global, per round
//   warps 4-7   MMA issuers (one thread each; warp 4 owns TMEM: 2 x T accumulators, round r+1
//               accumulates while r drains); issuer w handles the tiles t = w (mod 4) of every stage
//   warp  8     kernel-map loader: units of jc offsets x T tiles, one 512-byte cp.async.bulk per
//               (offset, tile) into a 3-deep ring; resident weight panels
//   warps 9..   G gather groups of 4 warps.  Group i owns ring slot i and the stages g = i (mod G):
//               wait slot free -> cp.async all T x 128 rows (zero fill for absent neighbours) into
//               the no-swizzle K-major core-matrix layout -> wait landed -> proxy fence -> signal.
//               While one group waits for its rows the other groups issue theirs.
// All hand-offs are mbarriers: groups never wait for one another.
// Shared memory: [weights: all kvol panels (resident) | one panel per slot][A: G slots x T tiles]
//                [kernel-map ring 3 x jc x T x 128 x 4][barriers].
template <int CIN, int G, int LT>
__global__ void __launch_bounds__(32 * (kProdWarp0 + 4 * G), 1)
k_spconv_umma(const UmmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int CPR = CIN / 8;     // 16-byte chunks per row
  // A tile layout (no-swizzle K-major core matrices of 8 rows x 16 B): chunk (row, kc) sits at
  // (row >> 3) * sbo_a + kc * lbo_a + (row & 7) * 16.  lbo_a = 144 (not 128) rotates the banks by one
  // 16-byte slot per K chunk, so the chunks of one gathered row -- which return from L1/L2 together,
  // one sector or line at a time -- are written to shared memory without bank conflicts.
  constexpr uint32_t lbo_a = 144u;
  constexpr uint32_t sbo_a = (uint32_t)CPR * lbo_a;
  constexpr uint32_t a_tile = (kTileM / 8) * sbo_a;
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int cout = p.cout, kvol = p.kvol;
  pdl_trigger();      // the next convolution of the stream may set itself up on SMs this grid leaves
  constexpr int lt = LT, T = 1 << LT;   // compile time: the tile loops unroll, dirty bits become immediates
  const uint32_t b_bytes = (uint32_t)CIN * cout * 2u;
  const uint32_t a_bytes = a_tile << lt;    // one stage: T tiles
  uint32_t base = smem_u32(smem_raw);
  base = (base + 127u) & ~127u;
  const uint32_t w0 = base;
  const uint32_t w_total = p.wres ? (uint32_t)kvol * b_bytes : (uint32_t)G * b_bytes;
  const uint32_t a0 = w0 + w_total;
  const uint32_t nbr0 = a0 + (uint32_t)G * a_bytes;
  const uint32_t unit_bytes = p.nbr ? (uint32_t)((p.jc << lt) * kTileM * 4) : 0u;
  const uint32_t bar0 = (nbr0 + kNbrBufs * unit_bytes + 7u) & ~7u;
  const uint32_t full0 = bar0, empty0 = bar0 + 8u * G, tfull0 = bar0 + 16u * G, tempty0 = tfull0 + 16u,
                 wfull = tempty0 + 16u, nfull0 = wfull + 8u, nempty0 = nfull0 + 8u * kNbrBufs,
                 tmem_slot = nempty0 + 8u * kNbrBufs, abort_slot = tmem_slot + 4u;
  const WaitCtx wc{abort_slot, p.spin_limit, p.fault};
  const uint32_t nunits = (uint32_t)(kvol / p.jc);   // kernel-map units per round

  if (tid == 0) {
    const uint32_t nmma = (uint32_t)(T < kMmaWarps ? T : kMmaWarps);   // MMA warps with work
    for (int s = 0; s < G; ++s) {
      mbar_init(full0 + 8u * s, 4);
      mbar_init(empty0 + 8u * s, nmma);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull0 + 8u * i, nmma);
      mbar_init(tempty0 + 8u * i, 4);
    }
    for (int i = 0; i < kNbrBufs; ++i) {
      mbar_init(nfull0 + 8u * i, 1);
      mbar_init(nempty0 + 8u * i, 4 * p.jc);   // every stage of the unit releases it once per warp
    }
    mbar_init(wfull, 1);
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(abort_slot), "r"(0u) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ld_shared_u32(tmem_slot);
  // Everything above (barriers, TMEM) is this CTA's own; from here on the kernel reads what its predecessors in
  // the stream wrote (feature rows, kernel map, residual).  The launch is a programmatic dependent one, so this CTA
  // may have become resident while the previous kernel's last CTAs were still running.
  // The resident weight panels are constants of the model: their copy goes out before the wait.
  if (warp == kLoaderWarp && lane == 0 && p.wres) {
    mbar_arrive_expect_tx(wfull, w_total);
    for (int j = 0; j < kvol; ++j) bulk_g2s(w0 + (uint32_t)j * b_bytes, p.w + (size_t)j * b_bytes, b_bytes, wfull);
  }
  pdl_wait();

  if (warp >= kProdWarp0) {
    // ---------------- gather groups ----------------
    const int gi = (warp - kProdWarp0) >> 2;   // group = ring slot
    const int pw = (warp - kProdWarp0) & 3;    // row quarter inside a tile
    const int sub = lane % CPR, rl0 = lane / CPR;
    const uint32_t n_in32 = (uint32_t)p.n_in;
    const uint64_t in_b64 = reinterpret_cast<uint64_t>(p.in) + (uint64_t)(sub * 16);
    // lane -> (row group rg4 = lane / CPR, chunk sub): the thread copies chunk `sub` of the CPR
    // consecutive rows 32*pw + rg4*CPR + it, so their CPR kernel-map entries are one vector load
    uint32_t dst_c[CPR];   // per-thread constant part of the destination address
#pragma unroll
    for (int it = 0; it < CPR; ++it) {
      const int row = 32 * pw + rl0 * CPR + it;
      dst_c[it] = (uint32_t)sub * lbo_a + (uint32_t)(row >> 3) * sbo_a + (uint32_t)(row & 7) * 16u;
    }
    const int rowb = 32 * pw + rl0 * CPR;   // first of this thread's rows
    const uint32_t a_s = a0 + (uint32_t)gi * a_bytes;
    uint32_t a_dst[CPR];
#pragma unroll
    for (int it = 0; it < CPR; ++it) a_dst[it] = a_s + dst_c[it];
    const uint32_t fullb = full0 + 8u * gi, emptyb = empty0 + 8u * gi;
    // chunks this thread wrote non-zero data to at the previous use of its slot (bit t*CPR + it).
    // A chunk that was zero and stays zero needs no copy at all; all ones forces the first fill.
    uint32_t dirty = ~0u;         // T * CPR <= 32 bits (checked by the launcher)
    uint32_t use = 0;               // times this slot has been filled
    uint32_t j = (uint32_t)gi;      // kernel offset of the next stage
    uint32_t r = 0;                 // local round
    int64_t rg = blockIdx.x;        // global round (super-tile index)
    while (j >= (uint32_t)kvol) { j -= kvol; ++r; rg += gridDim.x; }
    while (rg < p.nrounds_total) {
      uint32_t nb_addr = 0, nb_bar = 0;
      if (p.nbr) {
        const uint32_t c = j / (uint32_t)p.jc;
        const uint32_t unit = r * nunits + c;
        const uint32_t nbuf = unit % kNbrBufs;
        if (!mbar_wait(nfull0 + 8u * nbuf, (unit / kNbrBufs) & 1u, wc)) goto role_done;
        nb_addr = nbr0 + nbuf * unit_bytes + (((j - c * p.jc) << lt) * kTileM) * 4u;
        nb_bar = nempty0 + 8u * nbuf;
      }
      if (use > 0 && !mbar_wait(emptyb, (use - 1) & 1u, wc)) goto role_done;
      if (!p.wres && pw == 0 && lane == 0) {
        mbar_expect_tx(fullb, b_bytes);
        bulk_g2s(w0 + (uint32_t)gi * b_bytes, p.w + (size_t)j * b_bytes, b_bytes, fullb);
      }
      // `absent` collects the sign bits of the map entries (-1 = no neighbour) by funnel shift, one
      // instruction per chunk; the k-th chunk processed ends at bit N-1-k (N = T * CPR chunks per thread)
      constexpr int N = T * CPR;
      uint32_t absent = 0;
      auto chunk = [&](uint32_t r, int t, int it) {
        absent = __funnelshift_l(r, absent, 1);
        gather_chunk<CIN * 2>(a_dst[it] + (uint32_t)t * a_tile, in_b64, r, n_in32, dirty & (1u << (N - 1 - (t * CPR + it))));
      };
      if (p.nbr != nullptr) {
#pragma unroll
        for (int t = 0; t < T; ++t) {
          // (tiles past the end of the map hold -1: the loader fills them)
          uint32_t rr[CPR];
          const uint32_t ia = nb_addr + (uint32_t)(t * kTileM + rowb) * 4u;
          if (CPR >= 4) {
#pragma unroll
            for (int v4 = 0; v4 < CPR / 4; ++v4)
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(rr[4 * v4]), "=r"(rr[4 * v4 + 1]), "=r"(rr[4 * v4 + 2]), "=r"(rr[4 * v4 + 3])
                           : "r"(ia + 16u * v4)
                           : "memory");
          } else {
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(rr[0]), "=r"(rr[1]) : "r"(ia) : "memory");
          }
#pragma unroll
          for (int it = 0; it < CPR; ++it) chunk(rr[it], t, it);
        }
      } else {
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int64_t o0 = (((rg << lt) + t) * kTileM) + rowb;
#pragma unroll
          for (int it = 0; it < CPR; ++it) chunk((o0 + it < p.n_out) ? (uint32_t)(o0 + it) : 0xffffffffu, t, it);
        }
      }
      // a chunk is dirty (holds data) iff its entry was a row: bit clear in `absent`
      const uint32_t now = ~absent & (N == 32 ? 0xffffffffu : ((1u << (N & 31)) - 1u));
      dirty = now;
      cp_async_commit();
      if (p.nbr) {          // kernel-map reads of this stage are done
        __syncwarp();
        if (lane == 0) mbar_arrive(nb_bar);
      }
      cp_async_wait<0>();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(fullb);
      ++use;
      j += G;
      while (j >= (uint32_t)kvol) { j -= kvol; ++r; rg += gridDim.x; }
    }
  } else if (warp == kLoaderWarp) {
    // ---------------- kernel-map loader (the resident weights were requested before the wait) ----------------
    if (p.nbr) {
      uint32_t unit = 0;
      const int jc = p.jc;
      for (int64_t rg = blockIdx.x; rg < p.nrounds_total; rg += gridDim.x) {
        int64_t tv = p.ntiles - (rg << lt);       // valid tiles in this round
        if (tv > T) tv = T;
        for (uint32_t c = 0; c < nunits; ++c, ++unit) {
          const uint32_t nb = unit % kNbrBufs;
          if (unit >= kNbrBufs && !mbar_wait_relaxed(nempty0 + 8u * nb, ((unit / kNbrBufs) - 1u) & 1u, wc)) goto role_done;
          const uint32_t dst = nbr0 + nb * unit_bytes;
          if (p.nbr_bulk) {
            if (tv < T) {      // last round: tiles past the end of the map gather nothing
              for (int e = lane; e < jc * T * kTileM; e += 32)
                if (((e >> 7) & (T - 1)) >= (int)tv)
                  asm volatile("st.shared.b32 [%0], %1;" ::"r"(dst + (uint32_t)e * 4u), "r"(-1) : "memory");
              __syncwarp();
            }
            if (lane == 0) {
              mbar_arrive_expect_tx(nfull0 + 8u * nb, (uint32_t)(jc * tv * kTileM * 4));
              for (int jj = 0; jj < jc; ++jj)
                for (int t = 0; t < (int)tv; ++t)
                  bulk_g2s(dst + (uint32_t)(((jj << lt) + t) * kTileM * 4),
                           p.nbr + (int64_t)(c * jc + jj) * p.nbr_ld + ((rg << lt) + t) * kTileM, kTileM * 4u,
                           nfull0 + 8u * nb);
            }
          } else {
            // arbitrary pitch / unpadded tail: guarded loads by the whole warp
            for (int e = lane; e < jc * T * kTileM; e += 32) {
              const int rr = e & (kTileM - 1), t = (e >> 7) & (T - 1), jj = e >> (7 + lt);
              const int64_t o = (((rg << lt) + t) * kTileM) + rr;
              const int32_t v = o < p.n_out ? __ldg(p.nbr + (int64_t)(c * jc + jj) * p.nbr_ld + o) : -1;
              asm volatile("st.shared.b32 [%0], %1;" ::"r"(dst + (uint32_t)e * 4u), "r"(v) : "memory");
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(nfull0 + 8u * nb);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp >= kMmaWarp && warp < kMmaWarp + kMmaWarps) {
    // ---------------- MMA issuers (one thread per warp; warp mw owns tiles t = mw, mw + 4, ...) ----------------
    const int mw = warp - kMmaWarp;
    if (lane == 0 && mw < T) {
      // instruction descriptor: c=f32 (bit 4), a=b=bf16 (bits 7,10), K-major both, N>>3 @17, M>>4 @24
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(cout >> 3) << 17) |
                             ((uint32_t)(kTileM >> 4) << 24);
      const uint32_t lbo_b = (uint32_t)cout * 16u;
      constexpr int ksteps = CIN >> 4;
      // descriptors differ only in their 14-bit start-address field: keep the bases in registers and
      // add (bytes >> 4) per slot / offset / K step (every address is below 256 KB: no carry out)
      const uint64_t adesc0 = make_desc(a0 + (uint32_t)mw * a_tile, lbo_a, sbo_a);
      const uint64_t bdesc0 = make_desc(w0, lbo_b, 128u);
      const uint32_t a_slot = a_bytes >> 4, a_ks = (2u * lbo_a) >> 4, a_t = (kMmaWarps * a_tile) >> 4;
      const uint32_t b_step = b_bytes >> 4, b_ks = (2u * lbo_b) >> 4;
      if (p.wres && !mbar_wait(wfull, 0, wc)) goto role_done;
      uint32_t s = 0, ph = 0;
      int r = 0;
      for (int64_t rg = blockIdx.x; rg < p.nrounds_total; rg += gridDim.x, ++r) {
        const uint32_t par = (uint32_t)(r & 1);
        if (r >= 2 && !mbar_wait(tempty0 + 8u * par, (uint32_t)(((r >> 1) - 1) & 1), wc)) goto role_done;
        tc_fence_after();
        const uint32_t tmem_d0 = tmem_base + ((par << lt) + (uint32_t)mw) * p.acc_cols;
        for (int j = 0; j < kvol; ++j) {
          if (!mbar_wait(full0 + 8u * s, ph & 1u, wc)) goto role_done;
          tc_fence_after();
          const uint64_t ad = adesc0 + (uint64_t)(s * a_slot);
          const uint64_t bd = bdesc0 + (uint64_t)((p.wres ? (uint32_t)j : s) * b_step);
          for (int t = mw, ti = 0; t < T; t += kMmaWarps, ++ti) {
#pragma unroll
            for (int ks = 0; ks < ksteps; ++ks)
              umma_bf16(tmem_d0 + (uint32_t)(ti * kMmaWarps) * p.acc_cols, ad + (uint64_t)(ti * a_t + ks * a_ks),
                        bd + (uint64_t)(ks * b_ks), idesc, (j == 0 && ks == 0) ? 0u : 1u);
          }
          if (!(p.inject && mw == 0 && r == 0 && j == 1)) umma_commit(empty0 + 8u * s);
          if (j == kvol - 1) umma_commit(tfull0 + 8u * par);
          if (++s == (uint32_t)G) { s = 0; ++ph; }
        }
      }
    }
  } else {
    // ---------------- epilogue (warps 0-3; TMEM lane = tile row) ----------------
    int r = 0;
    const int cstore = p.cstore > 0 ? p.cstore : cout;
    for (int64_t rg = blockIdx.x; rg < p.nrounds_total; rg += gridDim.x, ++r) {
      const uint32_t par = (uint32_t)(r & 1);
      if (!mbar_wait_relaxed(tfull0 + 8u * par, (uint32_t)((r >> 1) & 1), wc)) goto role_done;
      tc_fence_after();
      for (int t = 0; t < T; ++t) {
        const int64_t tile = (rg << lt) + t;
        if (tile >= p.ntiles) break;
        const int64_t o = tile * kTileM + tid;
        const bool rowok = o < p.n_out;
        const uint32_t taddr = tmem_base + ((par << lt) + (uint32_t)t) * p.acc_cols + ((uint32_t)(warp * 32) << 16);
        for (int c0 = 0; c0 < cout; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(taddr + (uint32_t)c0, v);
          if (rowok && c0 < cstore) {
            float f[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
            if (p.bias) {
              const float4* bp = reinterpret_cast<const float4*>(p.bias + c0);   // 16-byte aligned (checked)
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 b4 = __ldg(bp + i);
                f[4 * i] += b4.x;
                f[4 * i + 1] += b4.y;
                f[4 * i + 2] += b4.z;
                f[4 * i + 3] += b4.w;
              }
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
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8u * par);
    }
  }
role_done:
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
static int g_groups = 0, g_max_t = 4;  // g_groups 0 = by channel count
static std::atomic<int> g_inject{0};   // pcc_debug_inject

template <int CIN, int G, int LT>
static int launch_umma_glt(const UmmaParams& p, size_t smem, unsigned grid, cudaStream_t st) {
  static PerDeviceOnce configured;      // the shared-memory opt-in is per device, not per process
  const int dev = current_device();
  if (configured.needed(dev)) {
    PCC_CUDA(cudaFuncSetAttribute(k_spconv_umma<CIN, G, LT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
    configured.mark(dev);
  }
  pcc::g_launches.fetch_add(1);
  PCC_CUDA(launch_pdl(k_spconv_umma<CIN, G, LT>, dim3(grid), dim3(32 * (kProdWarp0 + 4 * G)), smem, st, p));
  return PCC_OK;
}
template <int CIN, int G>
static int launch_umma_g(const UmmaParams& p, size_t smem, unsigned grid, cudaStream_t st) {
  switch (p.lt) {
    case 2: return launch_umma_glt<CIN, G, 2>(p, smem, grid, st);
    case 1: return launch_umma_glt<CIN, G, 1>(p, smem, grid, st);
    default: return launch_umma_glt<CIN, G, 0>(p, smem, grid, st);
  }
}
template <int CIN>
static int launch_umma(const UmmaParams& p, int groups, size_t smem, unsigned grid, cudaStream_t st) {
  switch (groups) {
    case 5: return launch_umma_g<CIN, 5>(p, smem, grid, st);
    case 4: return launch_umma_g<CIN, 4>(p, smem, grid, st);
    case 2: return launch_umma_g<CIN, 2>(p, smem, grid, st);
    default: return launch_umma_g<CIN, 1>(p, smem, grid, st);
  }
}
}  // extern "C++"

int pcc_debug_inject(int32_t what) {
  PCC_CHECK_ARG(what == 0 || what == 1);
  g_inject.store(what);
  return PCC_OK;
}

int pcc_spconv_set_tuning(int groups, int max_tiles, int use_cg) {
  PCC_CHECK_ARG(max_tiles == 1 || max_tiles == 2 || max_tiles == 4);
  PCC_CHECK_ARG(groups == 0 || groups == 1 || groups == 2 || groups == 4 || groups == 5);
  PCC_CHECK_ARG(use_cg == 0);   // the cp.async.cg gather variant was measured slower (571 vs 363 us) and removed
  g_max_t = max_tiles;
  g_groups = groups;
  return PCC_OK;
}

int pcc_spconv_umma(const void* in_bf16, int64_t n_in, int cin, const int32_t* nbr, int64_t nbr_ld,
                    int64_t n_out, int kvol, const void* w_packed, const float* bias, int cout,
                    const void* residual_bf16, int64_t ldr, int relu, void* out_bf16,
                    float* out_f32, int64_t ldo, int cstore, pcc_stream_t s) {
  PCC_CHECK_ARG(n_out >= 0 && n_in >= 0 && n_in < (1ll << 31) && kvol >= 1 && kvol <= 32);
  PCC_CHECK_ARG(cin == 16 || cin == 32 || cin == 64 || cin == 128);
  PCC_CHECK_ARG(cout >= 16 && cout % 16 == 0 && cout <= 256);
  PCC_CHECK_ARG(nbr || kvol == 1);
  PCC_CHECK_ARG(!nbr || nbr_ld >= n_out);
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
  PCC_CHECK_ARG(!bias || ((uintptr_t)bias % 16) == 0);

  UmmaParams p;
  p.in = (const __nv_bfloat16*)in_bf16;
  p.nbr = nbr;
  p.nbr_ld = nbr_ld;
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
  const int64_t ntiles = (n_out + kTileM - 1) / kTileM;
  p.ntiles = (int)ntiles;
  p.nbr_bulk = (nbr && nbr_ld % 4 == 0 && nbr_ld >= ntiles * kTileM && ((uintptr_t)nbr % 16) == 0) ? 1 : 0;
  const int sms = sm_count();
  p.fault = fault_word();
  if (!p.fault) return PCC_E_CUDA;
  p.inject = kvol == 27 ? g_inject.exchange(0) : 0;      // (a k3 layer: its pipeline is deep enough to stall)
  p.spin_limit = p.inject ? (1u << 10) : (1u << 22);
  // gather groups (= ring slots) and T tiles per stage.  T is bounded by TMEM (2 x T accumulators of
  // acc_cols columns in 512), by shared memory (G slots of T tiles), by the tuning cap and by the
  // tile count (keep >= 2 rounds per SM so the persistent grid stays balanced).
  // more gather warps = more rows in flight: 5 groups (928 threads) where five slots fit, i.e. narrow
  // inputs with the weight panels streamed per stage (363 -> 339 us on the 1.79 M-row c32 layer)
  int groups = g_groups ? g_groups : (cin <= 32 && kvol > 8 ? 5 : (cin <= 64 ? 4 : 2));
  p.acc_cols = (uint32_t)((cout + 31) / 32 * 32);
  const int64_t a_tile = (int64_t)(kTileM / 8) * (cin / 8) * 144, b_bytes = (int64_t)cin * cout * 2;
  const int64_t budget = 227 * 1024 - 1024;
  p.wres = (kvol * b_bytes <= 72 * 1024) ? 1 : 0;
  if (groups == 5) {          // five slots only fit when the weight panels travel with their stage
    if (cin <= 64 && kvol > 1) p.wres = 0; else groups = 4;
  }
  p.jc = 1;
  for (int d = 2; d <= 4; ++d)
    if (kvol % d == 0) p.jc = d;
  auto smem_need = [&](int G, int T) -> int64_t {
    const int64_t w = p.wres ? kvol * b_bytes : G * b_bytes;
    const int64_t nb = nbr ? (int64_t)kNbrBufs * p.jc * T * kTileM * 4 : 0;
    return w + (int64_t)G * T * a_tile + nb + 1024;
  };
  while (groups > 1 && smem_need(groups, 1) > budget) groups = groups == 5 ? 4 : groups >> 1;
  if (smem_need(groups, 1) > budget) {
    set_error("pcc_spconv_umma: cin %d cout %d kvol %d does not fit shared memory", cin, cout, kvol);
    return PCC_E_UNSUPPORTED;
  }
  // T: more tiles per stage amortise the per-stage hand-off (a T = 2 round costs ~1.3-1.5x a T = 1 round and covers two
  // tiles: tools/spconv_sweep.py, profiles/r2f_sweep_small.txt), but fewer, longer rounds leave SMs idle on small levels.
  // Pick the T with the lowest estimate  rounds per CTA x (1 + 0.5 (T - 1)).  (Round 1 kept >= 2 rounds per SM, which held
  // the 58 k-row C = 64 layers of g_a at T = 1: 57 us; with T = 2 they take 40 us.)
  int T = 1, lt = 0;
  {
    double best = 1e30;
    for (int t = 1, l = 0; t <= g_max_t; t *= 2, ++l) {
      if ((uint32_t)(2 * t) * p.acc_cols > 512u || t * (cin / 8) > 32 || smem_need(groups, t) > budget) break;
      const int64_t nsuper = (ntiles + t - 1) / t;
      const double est = (double)((nsuper + sms - 1) / sms) * (1.0 + 0.5 * (t - 1));
      if (est < best - 1e-9) {
        best = est;
        T = t;
        lt = l;
      }
    }
  }
  p.lt = lt;
  p.nrounds_total = (int)((ntiles + T - 1) / T);
  uint32_t cols = 32;
  while (cols < 2u * T * p.acc_cols) cols <<= 1;
  p.tmem_cols = cols;
  const size_t smem = (size_t)smem_need(groups, T);
  const unsigned grid = (unsigned)(p.nrounds_total < sms ? p.nrounds_total : sms);
  cudaStream_t st = as_stream(s);
  switch (cin) {
    case 16: return launch_umma<16>(p, groups, smem, grid, st);
    case 32: return launch_umma<32>(p, groups, smem, grid, st);
    case 64: return launch_umma<64>(p, groups, smem, grid, st);
    default: return launch_umma<128>(p, groups, smem, grid, st);
  }
}

}  // extern "C"
