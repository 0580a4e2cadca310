// bf16 GEMM on the 5th-gen tensor cores: C[M,N] = alpha * A[M,K] . B[N,K]^T, fp32 accumulate.
//
//   * operands staged by TMA (128-byte swizzle) into a STAGES-deep shared-memory ring,
//   * tcgen05.mma (cta_group::1, UMMA 128 x BN x 16) issued by one thread, accumulators in TMEM,
//     double-buffered (2 x BN columns) so the epilogue of tile i overlaps the mainloop of i+1,
//   * persistent: one CTA per SM walks the tile list (m fastest: the A panel stays in L2),
//   * K-major or MN-major operands (weight-gradient GEMMs read dY and X "transposed" in place
//     through MN-major descriptors, no transpose pass),
//   * split-K for the short-and-wide weight-gradient shapes (fp32 red.global.add epilogue),
//   * fused epilogues (bias, exact GELU forward/backward, residual add, fp32 accumulate); the
//     output tile leaves through swizzled shared-memory staging buffers and TMA stores (TMA
//     reduce-add for accumulation / split-K), residual and pre-activation tiles arrive by TMA,
//     prefetched several chunks ahead, so every global access of the epilogue is a full line.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..5 (and 6..9 for the ALU-heavy GELU epilogues) = epilogue, one TMEM lane quarter each.
#include <atomic>
#include <mutex>
#include <unordered_map>

#include "gemm_common.cuh"
#include "sm100_ptx.cuh"

namespace nsb {

namespace {
constexpr int BM = 128;
constexpr int BK = 64;

constexpr int kMaxStages = 8;
constexpr int kMaxBuf = 8;
constexpr int STG_BYTES = 128 * 128;  // one epilogue staging buffer: 128 rows x 128 B (swizzled)
constexpr size_t kSmemLimit = 232448;  // 227 KB opt-in maximum per CTA
inline size_t smem_bytes_for(int bn_local, int stages, int nbuf) {  // bn_local: B rows staged per CTA
  return static_cast<size_t>(stages) * (BM * BK * 2 + bn_local * BK * 2) + static_cast<size_t>(nbuf) * STG_BYTES +
         1024 /*align*/ + 512 /*barriers*/;
}
}  // namespace

// ---- epilogue: 32 consecutive columns of one output row held by one thread ---------------------
__device__ __forceinline__ void epilogue_row32(const EpiParams& p, int m, int n, float* v, float alpha,
                                               bool use_atomics) {
  if (m >= p.M || n >= p.N) return;
  const bool full = (n + 32 <= p.N);
  const size_t o = static_cast<size_t>(m) * p.ldc + n;
  const bool lowp_ok = full && ((p.ldc & 7) == 0);   // 16-byte aligned bf16 rows (n % 32 == 0)
  const bool f32_ok = full && ((p.ldc & 3) == 0);
  const bool has_bias = p.bias != nullptr && p.kind != NSB_EPI_ACCUM && p.kind != NSB_EPI_DGELU;

  if (p.kind == NSB_EPI_ACCUM) {
    if (!f32_ok) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        if (n + j >= p.N) continue;
        float* c = static_cast<float*>(p.C) + o + j;
        if (use_atomics) atomicAdd(c, v[j] * alpha);
        else *c = p.accumulate ? *c + v[j] * alpha : v[j] * alpha;
      }
      return;
    }
    float4* c = reinterpret_cast<float4*>(static_cast<float*>(p.C) + o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 a = make_float4(v[4 * j] * alpha, v[4 * j + 1] * alpha, v[4 * j + 2] * alpha, v[4 * j + 3] * alpha);
      if (use_atomics) {
        atomicAdd(c + j, a);
      } else {
        if (p.accumulate) {
          const float4 old = c[j];
          a.x += old.x; a.y += old.y; a.z += old.z; a.w += old.w;
        }
        c[j] = a;
      }
    }
    return;
  }

  const bool vec_ok = (p.kind == NSB_EPI_RESIDUAL) ? f32_ok : (p.c_bf16 ? lowp_ok : f32_ok);
  if (!vec_ok) {
#pragma unroll
    for (int j = 0; j < 32; ++j) epilogue_scalar(p, m, n + j, v[j], alpha);
    return;
  }
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] *= alpha;
  if (has_bias) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n) + j);
      v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
    }
  }
  switch (p.kind) {
    case NSB_EPI_STORE:
    case NSB_EPI_GELU: {
      if (p.c_bf16) {
        uint4* c = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.C) + o);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          c[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                            pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
        if (p.kind == NSB_EPI_GELU) {
          uint4* c2 = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.C2) + o);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_fast(round_bf16(v[j]));
#pragma unroll
          for (int j = 0; j < 4; ++j)
            c2[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                               pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
        }
      } else {
        float4* c = reinterpret_cast<float4*>(static_cast<float*>(p.C) + o);
#pragma unroll
        for (int j = 0; j < 8; ++j) c[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        if (p.kind == NSB_EPI_GELU) {
          float4* c2 = reinterpret_cast<float4*>(static_cast<float*>(p.C2) + o);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            c2[j] = make_float4(gelu_exact(v[4 * j]), gelu_exact(v[4 * j + 1]), gelu_exact(v[4 * j + 2]),
                                gelu_exact(v[4 * j + 3]));
        }
      }
      break;
    }
    case NSB_EPI_RESIDUAL: {
      const float4* r = reinterpret_cast<const float4*>(static_cast<const float*>(p.aux) + o);
      float4* c = reinterpret_cast<float4*>(static_cast<float*>(p.C) + o);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 a = r[j];
        if (p.c_bf16)
          c[j] = make_float4(a.x + round_bf16(v[4 * j]), a.y + round_bf16(v[4 * j + 1]),
                             a.z + round_bf16(v[4 * j + 2]), a.w + round_bf16(v[4 * j + 3]));
        else
          c[j] = make_float4(a.x + v[4 * j], a.y + v[4 * j + 1], a.z + v[4 * j + 2], a.w + v[4 * j + 3]);
      }
      break;
    }
    case NSB_EPI_DGELU: {
      if (p.c_bf16) {
        const uint4* h = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.aux) + o);
        uint4* c = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.C) + o);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint4 u = h[j];
          const float2 h0 = unpack_bf16x2(u.x), h1 = unpack_bf16x2(u.y), h2 = unpack_bf16x2(u.z),
                       h3 = unpack_bf16x2(u.w);
          c[j] = make_uint4(pack_bf16x2(round_bf16(v[8 * j]) * gelu_fast_grad(h0.x),
                                        round_bf16(v[8 * j + 1]) * gelu_fast_grad(h0.y)),
                            pack_bf16x2(round_bf16(v[8 * j + 2]) * gelu_fast_grad(h1.x),
                                        round_bf16(v[8 * j + 3]) * gelu_fast_grad(h1.y)),
                            pack_bf16x2(round_bf16(v[8 * j + 4]) * gelu_fast_grad(h2.x),
                                        round_bf16(v[8 * j + 5]) * gelu_fast_grad(h2.y)),
                            pack_bf16x2(round_bf16(v[8 * j + 6]) * gelu_fast_grad(h3.x),
                                        round_bf16(v[8 * j + 7]) * gelu_fast_grad(h3.y)));
        }
      } else {
        const float4* h = reinterpret_cast<const float4*>(static_cast<const float*>(p.aux) + o);
        float4* c = reinterpret_cast<float4*>(static_cast<float*>(p.C) + o);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 a = h[j];
          c[j] = make_float4(v[4 * j] * gelu_exact_grad(a.x), v[4 * j + 1] * gelu_exact_grad(a.y),
                             v[4 * j + 2] * gelu_exact_grad(a.z), v[4 * j + 3] * gelu_exact_grad(a.w));
        }
      }
      break;
    }
    default:
      break;
  }
}

// ---- dynamic tile scheduler ---------------------------------------------------------------------
// The persistent CTAs (or CTA pairs) of a launch draw tile numbers from a counter in global memory
// instead of walking a fixed stride.  A CTA that cannot get an SM for a while (another kernel, e.g. an
// NCCL all-reduce running beside backward, holds it: a GEMM CTA needs the whole register file) then
// simply takes no tiles and the CTAs that are running absorb its share, where a static walk leaves
// its share to a second wave.  Each launch uses the next of kSchedSlots counters (launches of
// different streams may overlap); the last unit to finish puts its counter pair back to zero.
constexpr int kSchedSlots = 1024;
constexpr int kTileQ = 4;                 // depth of the per-CTA queue of tile numbers
__device__ int g_tile_next[kSchedSlots];
__device__ int g_tile_done[kSchedSlots];

// ---- the kernel -------------------------------------------------------------------------------
struct TileCoord { int m_blk, n_blk, split; };
// m-fastest walk: concurrently running tiles share the B panel, A panels stream (good when A fits
// in L2 or is re-read few times).  n-fastest: the n-blocks of one A row-block run side by side, so
// a huge A (the 1.2 GB dlogits in the lm_head gradient GEMMs) comes from HBM once.
// q = a / b, r = a % b for 0 <= a < 2^22, 0 < b: one reciprocal and a correction step instead of the ~100-clock
// integer division sequence (tile numbers are decomposed on the producer's and the store warp's critical paths)
__device__ __forceinline__ void fast_divmod(int a, int b, int& q, int& r) {
  if (a >= (1 << 22)) { q = a / b; r = a - q * b; return; }
  q = __float2int_rz(__int2float_rn(a) * __frcp_rn(__int2float_rn(b)));
  r = a - q * b;
  if (r < 0) { r += b; --q; }
  if (r >= b) { r -= b; ++q; }
}
__device__ __forceinline__ TileCoord tile_of(int w, int num_m, int num_n, int n_fast) {
  TileCoord t;
  int rest;
  if (n_fast) {
    fast_divmod(w, num_n, rest, t.n_blk);
    fast_divmod(rest, num_m, t.split, t.m_blk);
  } else {
    fast_divmod(w, num_m, rest, t.m_blk);
    fast_divmod(rest, num_n, t.split, t.n_blk);
  }
  return t;
}
__device__ __forceinline__ uint32_t stg_off(int row, int q) {  // 16-byte chunk q of a 128-byte row
  return static_cast<uint32_t>(row * 128 + ((q ^ (row & 7)) << 4));
}

// CG = 1: one CTA per 128 x BN tile.  CG = 2: a CTA pair (cluster of 2, cta_group::2) per 256 x BN
// tile: each CTA stages its own 128 rows of A and its own half (BN/2 rows) of B, the leader issues
// 256 x BN x 16 MMAs that read both CTAs' shared memory, each CTA keeps the accumulators of its 128
// rows in its own TMEM and runs its own epilogue.  Halves the B traffic from L2 per CTA.
// NGRP = number of 4-warp epilogue groups the CTA is launched with (2, or 4 for the ALU-bound GELU
// epilogues, which then store straight from registers and leave all shared memory to the ring).
template <int BN, bool A_MN, bool B_MN, int CG, int NGRP>
__global__ void __launch_bounds__(64 + 160 * NGRP, 1)
    gemm_tc_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const __grid_constant__ CUtensorMap tma_c, const __grid_constant__ CUtensorMap tma_c2,
                   const __grid_constant__ CUtensorMap tma_aux, int K, int num_m, int num_n, int splits,
                   int kb_per_split, int stages, int nbuf, int tma_epi, int ngroups, int aux_direct, int n_fast,
                   int pipe_epi, int sched_slot, EpiParams ep, long long* dbg) {
  // per-CTA wall-clock trace (debug buffer of >= 8192 words): ns since boot at 8 points of the CTA's life
  auto gstamp_if = [&](bool who, int slot) {
    if (dbg != nullptr && who) {
      unsigned long long gt;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
      dbg[4096 + 8 * blockIdx.x + slot] = static_cast<long long>(gt);
    }
  };
  auto gstamp = [&](int slot) { gstamp_if(threadIdx.x == 0, slot); };
  gstamp(0);
  constexpr int BNL = BN / CG;                       // B rows staged by this CTA
  constexpr uint32_t A_BYTES = BM * BK * 2, B_BYTES = BNL * BK * 2;
  constexpr uint32_t TMEM_COLS = 2 * BN;
  constexpr uint32_t IDESC = ptx::umma_idesc_bf16(BM * CG, BN, A_MN, B_MN);
  const uint32_t rank = (CG == 2) ? ptx::cluster_ctarank() : 0u;
  const int unit = static_cast<int>(blockIdx.x) / CG;      // CTA (CG=1) or pair (CG=2) index
  const int nunits = static_cast<int>(gridDim.x) / CG;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + stages * A_BYTES;
  uint8_t* smem_stg = smem_b + stages * B_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stg + ngroups * nbuf * STG_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + kMaxStages;
  uint64_t* tfull = bars + 2 * kMaxStages;
  uint64_t* tempty = tfull + 2;
  uint64_t* auxfull = tempty + 2;                    // [group][buffer]: aux tile landed (and buffer free)
  uint64_t* staged = auxfull + kMaxBuf;              // [group][buffer]: output data staged by the 128 threads
  uint64_t* freeb = staged + kMaxBuf;                // [group][buffer]: TMA store drained the buffer
  uint64_t* qfull = freeb + kMaxBuf;                 // tile queue: entry written (by the leader's producer)
  uint64_t* qempty = qfull + kTileQ;                 // ... entry read by every consumer warp of the CTA / pair
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(qempty + kTileQ);
  volatile int* tile_q = reinterpret_cast<volatile int*>(tmem_slot + 1);
  const bool dyn = sched_slot >= 0;

  // Hardware warps: 0 .. 4*NGRP-1 epilogue (TMEM -> registers -> staging smem), 4*NGRP .. 5*NGRP-1 one
  // store warp per epilogue group (issues the TMA stores / aux loads, recycles staging buffers), then
  // the TMA producer and the MMA issuer.  Logical `warp`: 0 TMA, 1 MMA, 2.. epilogue, -1 store warp.
  const int lane = threadIdx.x & 31;
  const int hw_warp = threadIdx.x >> 5;
  const int warp = hw_warp >= 5 * NGRP ? hw_warp - 5 * NGRP : (hw_warp >= 4 * NGRP ? -1 : hw_warp + 2);
  const int total_kb = (K + BK - 1) / BK;
  const int total_work = num_m * num_n * splits;
  // first tile number of this unit, requested before the set-up below so that its latency hides there (the
  // counter belongs to this launch alone; nothing of the previous kernel is read)
  int w_next = -1;
  if (sched_slot >= 0 && rank == 0 && warp == 0 && lane == 0) w_next = atomicAdd(&g_tile_next[sched_slot], 1);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tma_a);
    ptx::prefetch_tmap(&tma_b);
    if (tma_epi) {
      ptx::prefetch_tmap(&tma_c);
      if (ep.kind == NSB_EPI_GELU) ptx::prefetch_tmap(&tma_c2);
      if (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_DGELU) ptx::prefetch_tmap(&tma_aux);
    }
  }
  if (warp == 0) {
    // the 44 barriers are initialised by the lanes of one warp side by side (a single thread doing it
    // in turn was most of the CTA's set-up time)
    constexpr int kBars = 2 * kMaxStages + 4 + 3 * kMaxBuf + 2 * kTileQ;
    // warps that read the tile queue, per CTA: the MMA issuer (leader) or the producer (peer CTA of a pair),
    // the active epilogue warps, and the store warps of the staged epilogue
    const uint32_t qreaders = static_cast<uint32_t>(CG * (1 + 4 * ngroups + (tma_epi ? ngroups : 0)));
    for (int i = lane; i < kBars; i += 32) {
      uint32_t count = 1;                                                   // full, empty, tfull, auxfull, freeb, qfull
      if (i >= 2 * kMaxStages + 2 && i < 2 * kMaxStages + 4) count = 4 * ngroups * CG;   // tempty: epilogue warps of both CTAs
      else if (i >= 2 * kMaxStages + 4 + kMaxBuf && i < 2 * kMaxStages + 4 + 2 * kMaxBuf) count = 128;  // staged
      else if (i >= 2 * kMaxStages + 4 + 3 * kMaxBuf + kTileQ) count = qreaders;         // qempty (on the leader)
      ptx::mbar_init(&bars[i], count);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    if (CG == 2) ptx::tmem_alloc_pair(tmem_slot, TMEM_COLS);
    else ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (CG == 2) ptx::cluster_sync();                   // peer barriers initialised before any remote signal
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above (barriers, TMEM, descriptor prefetch) may overlap the previous kernel's tail;
  // from here on we touch global memory
  gstamp(1);
  pdl_wait();
  pdl_trigger();
  gstamp(2);

  // Tile sequence of this CTA.  Static: unit, unit + nunits, ...  Dynamic: the leader's producer draws the
  // numbers from the global counter and queues them (in both CTAs of a pair); every other role reads the
  // queue in order.  -1 ends the sequence.  All lanes of a warp call next_tile() together.
  int q_seq = 0;
  auto next_tile = [&]() -> int {
    if (!dyn) {
      const int w = unit + q_seq * nunits;
      ++q_seq;
      return w < total_work ? w : -1;
    }
    const int slot = q_seq & (kTileQ - 1);
    const uint32_t ph = static_cast<uint32_t>(q_seq / kTileQ) & 1u;
    // peer CTA of a pair: the entry arrives as an asynchronous store that completes 4 bytes on qfull; the
    // peer's producer (a reader of every entry, in order) arms the barrier for it
    if (CG == 2 && rank == 1 && warp == 0 && lane == 0) ptx::mbar_expect_tx(&qfull[slot], 4);
    ptx::mbar_wait(&qfull[slot], ph);
    const int w = tile_q[slot];
    __syncwarp();
    // (w >= -1 always holds: the test makes the arrive wait for the value, i.e. the slot is released after the read)
    if (lane == 0 && w >= -1) { if (CG == 2) ptx::mbar_arrive_remote_cta(&qempty[slot], 0); else ptx::mbar_arrive(&qempty[slot]); }
    ++q_seq;
    return w;
  };

  if (warp == 0) {
    // ================================ TMA producer ==========================================
    // lane 0 owns the barrier protocol; the boxes of a stage (up to 6 for MN-major operands) are
    // issued by different lanes so that their descriptor/issue latencies overlap
    {
      int stage = 0;
      uint32_t phase = 0;
      constexpr int A_BOXES = A_MN ? BM / 64 : 1;
      constexpr int B_BOXES = B_MN ? BNL / 64 : 1;
      const bool fetcher = dyn && rank == 0;    // draws from the global counter and feeds the queue(s)
      int pub = 0;
      while (true) {
        int w;
        if (fetcher) {
          if (lane == 0) {
            if (w_next >= total_work) w_next = -1;
            const int slot = pub & (kTileQ - 1);
            ptx::mbar_wait(&qempty[slot], (static_cast<uint32_t>(pub / kTileQ) & 1u) ^ 1u);
            tile_q[slot] = w_next;
            if (CG == 2)
              ptx::st_async_remote_u32(const_cast<int*>(&tile_q[slot]), &qfull[slot], 1, static_cast<uint32_t>(w_next));
            ptx::mbar_arrive(&qfull[slot]);
          }
          ++pub;
          w = __shfl_sync(0xffffffffu, w_next, 0);
          // the next number is asked for now and first looked at after this tile's loads have been issued
          if (w >= 0 && lane == 0) w_next = atomicAdd(&g_tile_next[sched_slot], 1);
        } else {
          w = next_tile();
        }
        if (w < 0) break;
        const TileCoord t = tile_of(w, num_m, num_n, n_fast);
        const int kb0 = t.split * kb_per_split;
        const int kb1 = min(total_kb, kb0 + kb_per_split);
        const int m0 = (t.m_blk * CG + static_cast<int>(rank)) * BM;
        const int n0 = t.n_blk * BN + static_cast<int>(rank) * BNL;
        for (int kb = kb0; kb < kb1; ++kb) {
          if (lane == 0) {
            ptx::mbar_wait(&empty[stage], phase ^ 1);
            // the leader's barrier collects the bytes of both CTAs of a pair
            if (rank == 0) ptx::mbar_expect_tx(&full[stage], (A_BYTES + B_BYTES) * CG);
          }
          __syncwarp();
          uint8_t* a_dst = smem_a + stage * A_BYTES;
          uint8_t* b_dst = smem_b + stage * B_BYTES;
          if (lane < A_BOXES) {
            void* dst = A_MN ? a_dst + lane * (64 * BK * 2) : a_dst;
            const int c0 = A_MN ? m0 + lane * 64 : kb * BK, c1 = A_MN ? kb * BK : m0;
            if (CG == 2) ptx::tma_load_2d_pair(dst, &tma_a, &full[stage], c0, c1);
            else ptx::tma_load_2d(dst, &tma_a, &full[stage], c0, c1);
          } else if (lane < A_BOXES + B_BOXES) {
            const int j = lane - A_BOXES;
            void* dst = B_MN ? b_dst + j * (64 * BK * 2) : b_dst;
            const int c0 = B_MN ? n0 + j * 64 : kb * BK, c1 = B_MN ? kb * BK : n0;
            if (CG == 2) ptx::tma_load_2d_pair(dst, &tma_b, &full[stage], c0, c1);
            else ptx::tma_load_2d(dst, &tma_b, &full[stage], c0, c1);
          }
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
      }
      if (fetcher && lane == 0) {
        // this unit draws no more: the last unit of the launch to say so resets the counter pair for reuse
        if (atomicAdd(&g_tile_done[sched_slot], 1) == nunits - 1) {
          g_tile_next[sched_slot] = 0;
          g_tile_done[sched_slot] = 0;
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ============================================
    // The whole warp walks the loop (converged); one elected lane issues the MMAs and commits.
    // In a CTA pair only the leader issues; its commits arrive on the barriers of both CTAs.
    if (rank == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      int ti = -1;
      for (int w = next_tile(); w >= 0; w = next_tile()) {
        ++ti;
        // only the split index matters here (both tile orders put it last); no integer divisions on the issue
        // path between two tiles unless the launch is split-K
        int split_idx = 0, rem_unused = 0;
        if (splits > 1) fast_divmod(w, num_m * num_n, split_idx, rem_unused);
        const int kb0 = split_idx * kb_per_split;
        const int kb1 = min(total_kb, kb0 + kb_per_split);
        const bool recm = dbg != nullptr && blockIdx.x == 0 && lane == 0;
        if (recm && ti < 60) {
          dbg[1024 + 4 * ti] = clock64();                              // tile start
          unsigned long long gt;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
          dbg[2048 + ti] = static_cast<long long>(gt);                 // same instant in ns (effective SM clock)
        }
        ptx::mbar_wait(&tempty[as], aphase ^ 1);
        ptx::tc_fence_after();
        if (recm && ti < 60) dbg[1024 + 4 * ti + 1] = clock64();      // accumulator buffer free
        long long waited = 0;
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(as * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          const long long w0 = recm ? clock64() : 0;
          ptx::mbar_wait(&full[stage], phase);
          ptx::tc_fence_after();
          if (recm) waited += clock64() - w0;
          if (ti == 0 && kb == kb0) gstamp_if(lane == 0, 4);            // first operands landed
          const uint32_t a_addr = ptx::smem_u32(smem_a + stage * A_BYTES);
          const uint32_t b_addr = ptx::smem_u32(smem_b + stage * B_BYTES);
          // K-major:  rows of 128 B, 8-row groups 1024 B apart (SBO); LBO unused (1).
          // MN-major: [k][64 mn] boxes of 8 KB: LBO = next 64-wide mn block, SBO = next 8 k rows.
          const uint64_t a_desc = A_MN ? ptx::umma_smem_desc_sw128(a_addr, 64 * BK * 2, 1024)
                                       : ptx::umma_smem_desc_sw128(a_addr, 16, 1024);
          const uint64_t b_desc = B_MN ? ptx::umma_smem_desc_sw128(b_addr, 64 * BK * 2, 1024)
                                       : ptx::umma_smem_desc_sw128(b_addr, 16, 1024);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t a_k = a_desc + static_cast<uint64_t>(A_MN ? (k * 16 * 128) >> 4 : (k * 32) >> 4);
              const uint64_t b_k = b_desc + static_cast<uint64_t>(B_MN ? (k * 16 * 128) >> 4 : (k * 32) >> 4);
              if (CG == 2) ptx::umma_bf16_pair(d_tmem, a_k, b_k, IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
              else ptx::umma_bf16(d_tmem, a_k, b_k, IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
            }
            if (CG == 2) {
              ptx::umma_commit_pair(&empty[stage]);
              if (kb == kb1 - 1) ptx::umma_commit_pair(&tfull[as]);
            } else {
              ptx::umma_commit(&empty[stage]);  // frees the smem slot when these MMAs retire
              if (kb == kb1 - 1) ptx::umma_commit(&tfull[as]);  // accumulator complete -> epilogue
            }
          }
          __syncwarp();
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        if (recm && ti < 60) { dbg[1024 + 4 * ti + 2] = clock64(); dbg[1024 + 4 * ti + 3] = waited; }  // issued; cycles waiting for operands
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      gstamp_if(lane == 0, 5);                                          // all MMAs issued
    }
  } else if (warp >= 2) {
    // ================================ epilogue ==============================================
    // One or two groups of four warps; a group covers all 128 TMEM lanes (one lane quarter per
    // warp) and takes every ngroups-th column chunk of a tile.
    const int grp = (warp - 2) >> 2;
    const int q = hw_warp & 3;              // TMEM lane quarter this (hardware) warp may touch
    const int row = q * 32 + lane;          // row of the tile owned by this thread
    const int et = threadIdx.x & 127;         // 0..127 within the group (epilogue warps are hardware warps 0..)
    const float alpha = ep.alpha ? *ep.alpha : 1.f;
    int as = 0;
    uint32_t aphase = 0;
    if (grp >= ngroups) {
      // idle warps of the second group
    } else if (!tma_epi) {
      // ---- direct-store fallback (outputs whose row stride is not a multiple of 16 bytes) ----
      const bool use_atomics = (ep.kind == NSB_EPI_ACCUM) && splits > 1;
      for (int w = next_tile(); w >= 0; w = next_tile()) {
        const TileCoord t = tile_of(w, num_m, num_n, n_fast);
        ptx::mbar_wait(&tfull[as], aphase);
        ptx::tc_fence_after();
#pragma unroll 1
        for (int c = grp; c < BN / 32; c += ngroups) {
          float v[32];
          ptx::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as * BN + c * 32), v);
          epilogue_row32(ep, (t.m_blk * CG + static_cast<int>(rank)) * BM + row, t.n_blk * BN + c * 32, v, alpha, use_atomics);
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (CG == 2) ptx::mbar_arrive_remote_cta(&tempty[as], 0); else ptx::mbar_arrive(&tempty[as]); }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    } else if (pipe_epi && ep.c_bf16 && (ep.kind == NSB_EPI_GELU || ep.kind == NSB_EPI_DGELU)) {
      // ---- bf16 GELU forward / backward epilogues, software-pipelined ------------------------
      // Same 32-column units and two register sets as the path below.  A 64-column chunk (two units)
      // fills one staging buffer per output: forward emits the pre-activation h and gelu(h) (two buffers,
      // shipped through tma_c and tma_c2), backward overwrites the pre-activation tile that TMA brought
      // into the buffer (or that the thread fetched from global, one unit ahead, when the CTA runs two
      // groups) with dY * gelu'(h).  The math is the sigmoid form of common.cuh (gelu_sig2).
      const bool is_fwd = ep.kind == NSB_EPI_GELU;
      const bool aux_tma = !is_fwd && !aux_direct;
      const bool has_bias = is_fwd && ep.bias != nullptr;
      const bool scaled = alpha != 1.f;
      const int UPT = BN / 32 / ngroups;              // units per tile for this group (even)
      uint8_t* pool = smem_stg + grp * nbuf * STG_BYTES;
      uint64_t* auxbar = auxfull + grp * nbuf;
      uint64_t* stgbar = staged + grp * nbuf;
      uint64_t* freebar = freeb + grp * nbuf;
      // staging buffer of the next emission and the parity of its use count, stepped (no divisions); the
      // forward uses this buffer and the one after it
      int eb = 0;
      uint32_t eph = 0;
      auto acquire = [&](int b, uint32_t ph) {
        if (aux_tma) ptx::mbar_wait(&auxbar[b], ph);            // the aux tile landing doubles as "buffer free"
        else ptx::mbar_wait(&freebar[b], ph ^ 1u);
      };
      auto publish = [&](int b) {
        ptx::fence_proxy_async();
        ptx::mbar_arrive(&stgbar[b]);
      };
      const uint32_t lane_bits = static_cast<uint32_t>(q * 32) << 16;
      auto unit_col = [&](int u) -> int { return (grp + (u >> 1) * ngroups) * 64 + (u & 1) * 32; };
      for (int w = next_tile(); w >= 0; w = next_tile()) {
        const TileCoord t = tile_of(w, num_m, num_n, n_fast);
        const int m0 = (t.m_blk * CG + static_cast<int>(rank)) * BM;
        const uint32_t tbase = tmem_base + lane_bits + static_cast<uint32_t>(as * BN);
        // backward without aux staging: this thread's 32 pre-activations of unit u (64 contiguous bytes)
        auto load_h = [&](int u, uint4* hq) {
          if (is_fwd || aux_tma) return;
          const int n0 = t.n_blk * BN + unit_col(u);
          const bool rok = m0 + row < ep.M;
          const __nv_bfloat16* hp = static_cast<const __nv_bfloat16*>(ep.aux) + static_cast<size_t>(rok ? m0 + row : 0) * ep.ldc;
          if (rok && n0 + 32 <= ep.N) {
#pragma unroll
            for (int j = 0; j < 4; ++j) hq[j] = __ldg(reinterpret_cast<const uint4*>(hp + n0) + j);
          } else {
            __nv_bfloat16 tmp[32];
            for (int j = 0; j < 32; ++j) tmp[j] = (rok && n0 + j < ep.N) ? hp[n0 + j] : __float2bfloat16_rn(0.f);
#pragma unroll
            for (int j = 0; j < 4; ++j) hq[j] = reinterpret_cast<const uint4*>(tmp)[j];
          }
        };
        auto process = [&](int u, float* v, const uint4* hq) {
          const int n0 = t.n_blk * BN + unit_col(u);
          const int h4 = (u & 1) * 4;                 // this unit's half of the 128-byte staging rows
          const int eb1 = (eb + 1 == nbuf) ? 0 : eb + 1;
          const uint32_t eph1 = (eb1 == 0) ? (eph ^ 1u) : eph;
          if ((u & 1) == 0) {
            acquire(eb, eph);
            if (is_fwd) acquire(eb1, eph1);
          }
          uint8_t* b0 = pool + eb * STG_BYTES;
          uint8_t* b1 = pool + eb1 * STG_BYTES;
          if (scaled) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= alpha;
          }
          if (has_bias) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += (n0 + j < ep.N) ? __ldg(ep.bias + n0 + j) : 0.f;
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (is_fwd) {
              const uint4 hp = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                                          pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
              *reinterpret_cast<uint4*>(b0 + stg_off(row, h4 + j)) = hp;
              // gelu of the ROUNDED pre-activation (what backward and the reference see)
              const float2 h0 = unpack_bf16x2(hp.x), h1 = unpack_bf16x2(hp.y), h2 = unpack_bf16x2(hp.z), h3 = unpack_bf16x2(hp.w);
              const float2 g0 = gelu_sig2(h0.x, h0.y), g1 = gelu_sig2(h1.x, h1.y), g2 = gelu_sig2(h2.x, h2.y), g3 = gelu_sig2(h3.x, h3.y);
              *reinterpret_cast<uint4*>(b1 + stg_off(row, h4 + j)) =
                  make_uint4(pack_bf16x2(g0.x, g0.y), pack_bf16x2(g1.x, g1.y), pack_bf16x2(g2.x, g2.y), pack_bf16x2(g3.x, g3.y));
            } else {
              const uint4 hr = aux_tma ? *reinterpret_cast<const uint4*>(b0 + stg_off(row, h4 + j)) : hq[j];
              const float2 h0 = unpack_bf16x2(hr.x), h1 = unpack_bf16x2(hr.y), h2 = unpack_bf16x2(hr.z), h3 = unpack_bf16x2(hr.w);
              const float2 g0 = gelu_sig_grad_mul2(h0.x, h0.y, round_bf16(v[8 * j]), round_bf16(v[8 * j + 1]));
              const float2 g1 = gelu_sig_grad_mul2(h1.x, h1.y, round_bf16(v[8 * j + 2]), round_bf16(v[8 * j + 3]));
              const float2 g2 = gelu_sig_grad_mul2(h2.x, h2.y, round_bf16(v[8 * j + 4]), round_bf16(v[8 * j + 5]));
              const float2 g3 = gelu_sig_grad_mul2(h3.x, h3.y, round_bf16(v[8 * j + 6]), round_bf16(v[8 * j + 7]));
              *reinterpret_cast<uint4*>(b0 + stg_off(row, h4 + j)) =
                  make_uint4(pack_bf16x2(g0.x, g0.y), pack_bf16x2(g1.x, g1.y), pack_bf16x2(g2.x, g2.y), pack_bf16x2(g3.x, g3.y));
            }
          }
          if (u & 1) {
            publish(eb);
            if (is_fwd) publish(eb1);
            for (int k = 0; k < (is_fwd ? 2 : 1); ++k)
              if (++eb == nbuf) { eb = 0; eph ^= 1u; }
          }
        };
        ptx::mbar_wait(&tfull[as], aphase);
        ptx::tc_fence_after();
        float x[32], y[32];
        uint4 hx[4], hy[4];
        ptx::tmem_ld_32x32_async(tbase + static_cast<uint32_t>(unit_col(0)), x);
        load_h(0, hx);
#pragma unroll 1
        for (int u = 0; u < UPT; u += 2) {
          ptx::tmem_wait_ld32(x);
          ptx::tmem_ld_32x32_async(tbase + static_cast<uint32_t>(unit_col(u + 1)), y);
          load_h(u + 1, hy);
          process(u, x, hx);
          ptx::tmem_wait_ld32(y);
          if (u + 2 < UPT) {
            ptx::tmem_ld_32x32_async(tbase + static_cast<uint32_t>(unit_col(u + 2)), x);
            load_h(u + 2, hx);
          } else {
            // every accumulator of this tile is in registers: hand the TMEM buffer back before the math
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG == 2) ptx::mbar_arrive_remote_cta(&tempty[as], 0); else ptx::mbar_arrive(&tempty[as]); }
          }
          process(u + 1, y, hy);
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      gstamp(6);
    } else if (pipe_epi && ep.kind != NSB_EPI_GELU && ep.kind != NSB_EPI_DGELU) {
      // ---- staged epilogue, software-pipelined (plain store / residual add / accumulation) ----
      // The tile is drained in 32-column units with two register sets: the TMEM load of unit u+1 is
      // in flight while unit u is converted and written to the staging buffer, and the accumulator
      // buffer goes back to the MMA warp as soon as the last load has landed in registers.  A
      // staging buffer is one unit (fp32 outputs, 128-byte rows of 32 floats) or two (bf16, 64
      // columns); the group's store warp ships it exactly as in the path below.
      const bool c32 = (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_ACCUM || !ep.c_bf16);
      const bool aux_tma = (ep.kind == NSB_EPI_RESIDUAL) && !aux_direct;
      const bool has_bias = ep.bias != nullptr && ep.kind != NSB_EPI_ACCUM;
      const bool scaled = alpha != 1.f;
      const int UPT = BN / 32 / ngroups;              // units per tile for this group (even)
      uint8_t* pool = smem_stg + grp * nbuf * STG_BYTES;
      uint64_t* auxbar = auxfull + grp * nbuf;
      uint64_t* stgbar = staged + grp * nbuf;
      uint64_t* freebar = freeb + grp * nbuf;
      const bool rec = dbg != nullptr && blockIdx.x == 0 && hw_warp == 0 && lane == 0;
      int nrec = 0;
      auto stamp = [&]() { if (rec && nrec < 1000) dbg[nrec++] = clock64(); };
      int slot = 0;                                   // staging buffer of the next emission
      uint32_t useph = 0;                             // parity of that buffer's use count
      uint8_t* buf = nullptr;
      auto acquire = [&]() {
        if (aux_tma) ptx::mbar_wait(&auxbar[slot], useph);      // the aux tile landing doubles as "buffer free"
        else ptx::mbar_wait(&freebar[slot], useph ^ 1);
        buf = pool + slot * STG_BYTES;
      };
      auto publish = [&]() {
        ptx::fence_proxy_async();
        ptx::mbar_arrive(&stgbar[slot]);
        if (++slot == nbuf) { slot = 0; useph ^= 1; }
      };
      const uint32_t lane_bits = static_cast<uint32_t>(q * 32) << 16;
      auto unit_col = [&](int u) -> int {             // first column (within the tile) of unit u
        return c32 ? (grp + u * ngroups) * 32 : (grp + (u >> 1) * ngroups) * 64 + (u & 1) * 32;
      };
      for (int w = next_tile(); w >= 0; w = next_tile()) {
        const TileCoord t = tile_of(w, num_m, num_n, n_fast);
        const int m0 = (t.m_blk * CG + static_cast<int>(rank)) * BM;
        const uint32_t tbase = tmem_base + lane_bits + static_cast<uint32_t>(as * BN);
        auto process = [&](int u, float* v) {
          const int n0 = t.n_blk * BN + unit_col(u);
          if (c32 || (u & 1) == 0) acquire();
          stamp();                                  // [1] accumulators + buffer ready
          if (scaled) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= alpha;
          }
          if (has_bias) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += (n0 + j < ep.N) ? __ldg(ep.bias + n0 + j) : 0.f;
          }
          if (c32) {
            if (ep.kind == NSB_EPI_RESIDUAL) {
              float4 a[8];
              if (aux_tma) {
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] = *reinterpret_cast<const float4*>(buf + stg_off(row, j));
              } else {
                const float* rs = static_cast<const float*>(ep.aux) + static_cast<size_t>(m0 + row) * ep.ldc;
                const bool rok = m0 + row < ep.M;
                if (rok && n0 + 32 <= ep.N) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) a[j] = __ldg(reinterpret_cast<const float4*>(rs + n0) + j);
                } else {
#pragma unroll
                  for (int j = 0; j < 8; ++j) {
                    a[j].x = (rok && n0 + 4 * j + 0 < ep.N) ? rs[n0 + 4 * j + 0] : 0.f;
                    a[j].y = (rok && n0 + 4 * j + 1 < ep.N) ? rs[n0 + 4 * j + 1] : 0.f;
                    a[j].z = (rok && n0 + 4 * j + 2 < ep.N) ? rs[n0 + 4 * j + 2] : 0.f;
                    a[j].w = (rok && n0 + 4 * j + 3 < ep.N) ? rs[n0 + 4 * j + 3] : 0.f;
                  }
                }
              }
              if (ep.c_bf16) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  v[4 * j] = a[j].x + round_bf16(v[4 * j]); v[4 * j + 1] = a[j].y + round_bf16(v[4 * j + 1]);
                  v[4 * j + 2] = a[j].z + round_bf16(v[4 * j + 2]); v[4 * j + 3] = a[j].w + round_bf16(v[4 * j + 3]);
                }
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  v[4 * j] += a[j].x; v[4 * j + 1] += a[j].y; v[4 * j + 2] += a[j].z; v[4 * j + 3] += a[j].w;
                }
              }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *reinterpret_cast<float4*>(buf + stg_off(row, j)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            stamp();                                // [2] staged
            publish();
          } else {
            const int h4 = (u & 1) * 4;             // this unit's half of the 128-byte staging row
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(buf + stg_off(row, h4 + j)) =
                  make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                             pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
            stamp();                                // [2] staged
            if (u & 1) publish();
          }
          stamp();                                  // [3] published
        };
        ptx::mbar_wait(&tfull[as], aphase);
        ptx::tc_fence_after();
        float x[32], y[32];
        ptx::tmem_ld_32x32_async(tbase + static_cast<uint32_t>(unit_col(0)), x);
#pragma unroll 1
        for (int u = 0; u < UPT; u += 2) {
          stamp();                                  // [0] unit start
          ptx::tmem_wait_ld32(x);
          ptx::tmem_ld_32x32_async(tbase + static_cast<uint32_t>(unit_col(u + 1)), y);
          process(u, x);
          stamp();                                  // [0] unit start
          ptx::tmem_wait_ld32(y);
          if (u + 2 < UPT) {
            ptx::tmem_ld_32x32_async(tbase + static_cast<uint32_t>(unit_col(u + 2)), x);
          } else {
            // every accumulator of this tile is in registers: hand the TMEM buffer back before staging
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG == 2) ptx::mbar_arrive_remote_cta(&tempty[as], 0); else ptx::mbar_arrive(&tempty[as]); }
          }
          process(u + 1, y);
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      gstamp(6);                                                        // last unit staged
    } else {
      // ---- staged epilogue: TMEM -> registers -> swizzled smem; the group's store warp ships it ----
      const bool c32 = (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_ACCUM || !ep.c_bf16);
      const int CW = c32 ? 32 : 64;         // columns per 128-byte staging row
      const int CH = BN / CW;               // chunks per tile
      const bool aux_tma = (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_DGELU) && !aux_direct;
      const bool has_bias = ep.bias != nullptr && ep.kind != NSB_EPI_ACCUM && ep.kind != NSB_EPI_DGELU;
      const bool scaled = alpha != 1.f;
      uint8_t* pool = smem_stg + grp * nbuf * STG_BYTES;
      uint64_t* auxbar = auxfull + grp * nbuf;
      uint64_t* stgbar = staged + grp * nbuf;
      uint64_t* freebar = freeb + grp * nbuf;
      const bool rec = dbg != nullptr && blockIdx.x == 0 && hw_warp == 0 && lane == 0;
      int nrec = 0;
      auto stamp = [&]() { if (rec && nrec < 1000) dbg[nrec++] = clock64(); };
      int emit = 0;                         // staging-buffer uses so far (buffer = emit % nbuf)
      // acquire the staging buffer of the next emission: with an aux tile the TMA load that fills it
      // is also the "buffer is free" signal
      auto acquire = [&]() -> uint8_t* {
        const int b = emit % nbuf, u = emit / nbuf;
        if (aux_tma) ptx::mbar_wait(&auxbar[b], u & 1);
        else ptx::mbar_wait(&freebar[b], (u & 1) ^ 1);
        return pool + b * STG_BYTES;
      };
      auto publish = [&]() {                // data staged -> store warp
        ptx::fence_proxy_async();
        ptx::mbar_arrive(&stgbar[emit % nbuf]);
        ++emit;
      };
      for (int w = next_tile(); w >= 0; w = next_tile()) {
        const TileCoord t = tile_of(w, num_m, num_n, n_fast);
        ptx::mbar_wait(&tfull[as], aphase);
        ptx::tc_fence_after();
        const int m0 = (t.m_blk * CG + static_cast<int>(rank)) * BM;
#pragma unroll 1
        for (int c = grp; c < CH; c += ngroups) {
          const int n0 = t.n_blk * BN + c * CW;
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as * BN + c * CW);
          stamp();                                  // [0] chunk start
          if (c32) {
            float v[32];
            ptx::tmem_ld_32x32_async(taddr, v);
            uint8_t* buf = acquire();
            ptx::tmem_wait_ld32(v);
            stamp();                                // [1] accumulators + buffer ready
            if (scaled) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] *= alpha;
            }
            if (has_bias) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] += (n0 + j < ep.N) ? __ldg(ep.bias + n0 + j) : 0.f;
            }
            if (ep.kind == NSB_EPI_RESIDUAL) {
              const bool ok = (m0 + row < ep.M) && (n0 + 32 <= ep.N);
              const float4* rp = reinterpret_cast<const float4*>(
                  static_cast<const float*>(ep.aux) + static_cast<size_t>(ok ? m0 + row : 0) * ep.ldc + (ok ? n0 : 0));
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 a;
                if (aux_tma) a = *reinterpret_cast<const float4*>(buf + stg_off(row, j));
                else if (ok) a = __ldg(rp + j);
                else {
                  const float* rs = static_cast<const float*>(ep.aux) + static_cast<size_t>(m0 + row) * ep.ldc;
                  const bool rok = m0 + row < ep.M;
                  a.x = (rok && n0 + 4 * j + 0 < ep.N) ? rs[n0 + 4 * j + 0] : 0.f;
                  a.y = (rok && n0 + 4 * j + 1 < ep.N) ? rs[n0 + 4 * j + 1] : 0.f;
                  a.z = (rok && n0 + 4 * j + 2 < ep.N) ? rs[n0 + 4 * j + 2] : 0.f;
                  a.w = (rok && n0 + 4 * j + 3 < ep.N) ? rs[n0 + 4 * j + 3] : 0.f;
                }
                if (ep.c_bf16) {
                  v[4 * j] = a.x + round_bf16(v[4 * j]); v[4 * j + 1] = a.y + round_bf16(v[4 * j + 1]);
                  v[4 * j + 2] = a.z + round_bf16(v[4 * j + 2]); v[4 * j + 3] = a.w + round_bf16(v[4 * j + 3]);
                } else {
                  v[4 * j] += a.x; v[4 * j + 1] += a.y; v[4 * j + 2] += a.z; v[4 * j + 3] += a.w;
                }
              }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *reinterpret_cast<float4*>(buf + stg_off(row, j)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            stamp();                                // [2] staged
            publish();
            stamp();                                // [3] published
            if (ep.kind == NSB_EPI_GELU) {  // fp32 C2 (not used by the bf16 model path, kept complete)
              buf = acquire();
#pragma unroll
              for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(buf + stg_off(row, j)) =
                    make_float4(gelu_exact(v[4 * j]), gelu_exact(v[4 * j + 1]), gelu_exact(v[4 * j + 2]), gelu_exact(v[4 * j + 3]));
              publish();
            }
          } else {
            // bf16 output: 64 columns per staging row = two 32-column TMEM loads (issued back to back).
            // GELU: the activated output is a second emission computed from the same registers.
            float va[32], vb[32];
            ptx::tmem_ld_32x32_async(taddr, va);
            ptx::tmem_ld_32x32_async(taddr + 32, vb);
            uint4 hraw[8];
            if (ep.kind == NSB_EPI_DGELU && !aux_tma) {
              // pre-activation read straight from global: 128 contiguous bytes of this thread's row
              const bool ok = (m0 + row < ep.M) && (n0 + 64 <= ep.N);
              const uint4* hp = reinterpret_cast<const uint4*>(
                  static_cast<const __nv_bfloat16*>(ep.aux) + static_cast<size_t>(ok ? m0 + row : 0) * ep.ldc + (ok ? n0 : 0));
#pragma unroll
              for (int j = 0; j < 8; ++j) hraw[j] = ok ? __ldg(hp + j) : make_uint4(0, 0, 0, 0);
              if (!ok && (m0 + row < ep.M)) {  // ragged right edge: element-wise
                __nv_bfloat16 tmp[64];
                for (int j = 0; j < 64; ++j)
                  tmp[j] = (n0 + j < ep.N) ? static_cast<const __nv_bfloat16*>(ep.aux)[static_cast<size_t>(m0 + row) * ep.ldc + n0 + j]
                                           : __float2bfloat16_rn(0.f);
#pragma unroll
                for (int j = 0; j < 8; ++j) hraw[j] = reinterpret_cast<const uint4*>(tmp)[j];
              }
            }
            uint8_t* buf = acquire();
            if (ep.kind == NSB_EPI_DGELU && aux_tma) {
#pragma unroll
              for (int j = 0; j < 8; ++j) hraw[j] = *reinterpret_cast<const uint4*>(buf + stg_off(row, j));
            }
            ptx::tmem_wait_ld32(va);
            ptx::tmem_wait_ld32(vb);
            stamp();                                // [1] accumulators + buffer ready
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              float* v = half ? vb : va;
              const int nh = n0 + half * 32;
              if (scaled) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] *= alpha;
              }
              if (has_bias) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] += (nh + j < ep.N) ? __ldg(ep.bias + nh + j) : 0.f;
              }
              if (ep.kind == NSB_EPI_DGELU) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const uint4 hr = hraw[half * 4 + j];
                  const float2 h0 = unpack_bf16x2(hr.x), h1 = unpack_bf16x2(hr.y), h2 = unpack_bf16x2(hr.z), h3 = unpack_bf16x2(hr.w);
                  const float2 hh[4] = {h0, h1, h2, h3};
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float2 g2 = gelu_fast_grad_mul2(hh[i].x, hh[i].y, round_bf16(v[8 * j + 2 * i]), round_bf16(v[8 * j + 2 * i + 1]));
                    v[8 * j + 2 * i] = g2.x;
                    v[8 * j + 2 * i + 1] = g2.y;
                  }
                }
              }
#pragma unroll
              for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4*>(buf + stg_off(row, half * 4 + j)) =
                    make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                               pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
            }
            stamp();                                // [2] staged
            publish();
            stamp();                                // [3] published
            if (ep.kind == NSB_EPI_GELU) {
              buf = acquire();
#pragma unroll
              for (int half = 0; half < 2; ++half) {
                float* v = half ? vb : va;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  float y[8];
#pragma unroll
                  for (int i = 0; i < 8; i += 2) {
                    const float2 g2 = gelu_fast2(round_bf16(v[8 * j + i]), round_bf16(v[8 * j + i + 1]));
                    y[i] = g2.x;
                    y[i + 1] = g2.y;
                  }
                  *reinterpret_cast<uint4*>(buf + stg_off(row, half * 4 + j)) =
                      make_uint4(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]), pack_bf16x2(y[4], y[5]), pack_bf16x2(y[6], y[7]));
                }
              }
              publish();
            }
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (CG == 2) ptx::mbar_arrive_remote_cta(&tempty[as], 0); else ptx::mbar_arrive(&tempty[as]); }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      gstamp(6);                                                        // last chunk staged
    }
  }
  if (warp == -1 && tma_epi) {
    // ================================ store warp (one per epilogue group) ====================
    // Walks the same emission sequence as its group: waits until the 128 threads have staged a
    // buffer, issues the TMA store / reduce-add, keeps nbuf-2 stores in flight, and recycles a
    // buffer (release it, or refill it with the aux tile 2 chunks ahead) once its store has drained.
    const int grp = hw_warp - 4 * NGRP;
    if (grp < ngroups) {
      const bool c32 = (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_ACCUM || !ep.c_bf16);
      const int CW = c32 ? 32 : 64;
      const int CH = BN / CW, CHG = CH / ngroups;
      const int EPC = (ep.kind == NSB_EPI_GELU) ? 2 : 1;    // emissions per chunk
      const bool aux_tma = (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_DGELU) && !aux_direct;
      const bool reduce = (ep.kind == NSB_EPI_ACCUM) && (ep.accumulate || splits > 1);
      uint8_t* pool = smem_stg + grp * nbuf * STG_BYTES;
      uint64_t* auxbar = auxfull + grp * nbuf;
      uint64_t* stgbar = staged + grp * nbuf;
      uint64_t* freebar = freeb + grp * nbuf;
      // This warp runs ahead of its group (aux tiles are requested nbuf chunks early, possibly in a later
      // tile), so it keeps the tile numbers it has read so far; tile_at(i) = i-th tile of this CTA or -1.
      int tl[8];
      int fetched = 0;
      bool exhausted = false;
      auto tile_at = [&](int seq) -> int {          // whole warp
        while (fetched <= seq && !exhausted) {
          const int w = next_tile();
          if (w < 0) exhausted = true;
          else { tl[fetched & 7] = w; ++fetched; }
        }
        return seq < fetched ? tl[seq & 7] : -1;
      };
      // tile number -> block coordinates, remembered for the last two tiles asked for (an emission and the aux
      // look-ahead alternate between at most two tiles): the integer divisions of tile_of() stay off the
      // per-emission path, which is what the epilogue group waits on for its next free buffer
      int memo_w[2] = {-1, -1}, memo_m[2] = {0, 0}, memo_n[2] = {0, 0}, memo_next = 0;
      // (w, j): tile number and chunk index within the tile (0 .. CHG-1)
      auto coords = [&](int w, int j, int& col, int& rowg) {
        int k = (memo_w[0] == w) ? 0 : ((memo_w[1] == w) ? 1 : -1);
        if (k < 0) {
          k = memo_next;
          memo_next ^= 1;
          const TileCoord t = tile_of(w, num_m, num_n, n_fast);
          memo_w[k] = w; memo_m[k] = t.m_blk; memo_n[k] = t.n_blk;
        }
        col = memo_n[k] * BN + (j * ngroups + grp) * CW;
        rowg = (memo_m[k] * CG + static_cast<int>(rank)) * BM;
      };
      auto load_aux = [&](int w, int j, int b) {     // elected lane only: aux tile of chunk j of tile w into buffer b
        int col, rowg;
        coords(w, j, col, rowg);
        ptx::mbar_expect_tx(&auxbar[b], STG_BYTES);
        ptx::tma_load_2d(pool + b * STG_BYTES, &tma_aux, &auxbar[b], col, rowg);
      };
      // Every position below is kept as a counter that is stepped, not divided: the chunk being emitted
      // (e_seq-th tile of this CTA, chunk e_j, part e_sub of EPC, buffer e_b with parity e_ph), the aux
      // look-ahead (la_seq, la_j, buffer la_b: chunk `emitted + nbuf`), and the oldest store still in flight.
      int la_seq = 0, la_j = 0, la_b = 0;
      if (aux_tma) {
        for (int j = 0; j < nbuf; ++j) {
          const int w = tile_at(la_seq);
          if (w < 0) break;
          if (ptx::elect_one()) load_aux(w, la_j, la_b);
          __syncwarp();
          if (++la_j == CHG) { la_j = 0; ++la_seq; }
          if (++la_b == nbuf) la_b = 0;
        }
      }
      // stores allowed to be still reading their buffer; with aux tiles a buffer is refilled as early as
      // possible (the refill is the look-ahead of the aux stream: nbuf - inflight chunks)
      const int inflight = aux_tma ? (nbuf >= 6 ? 2 : 1) : (nbuf > 4 ? 2 : nbuf - 2);
      // an output larger than L2 (the 1.2 GB of logits) is written with evict_first so that it does not push
      // the operand panels, which every tile re-reads, out of L2
      const bool stream_out = (tma_epi & 2) != 0;
      const uint64_t evict_first = ptx::l2_policy_evict_first();
      int e_seq = 0, e_j = 0, e_sub = 0, e_b = 0, done_b = 0;
      uint32_t e_ph = 0;
      for (int e = 0;; ++e) {
        const int w = tile_at(e_seq);
        if (w < 0) break;
        ptx::mbar_wait(&stgbar[e_b], e_ph);
        if (ptx::elect_one()) {
          int col, rowg;
          coords(w, e_j, col, rowg);
          const CUtensorMap* map = (e_sub == 1) ? &tma_c2 : &tma_c;
          if (reduce) ptx::tma_reduce_add_2d(map, pool + e_b * STG_BYTES, col, rowg);
          else if (stream_out) ptx::tma_store_2d_hint(map, pool + e_b * STG_BYTES, col, rowg, evict_first);
          else ptx::tma_store_2d(map, pool + e_b * STG_BYTES, col, rowg);
          ptx::bulk_commit();
        }
        __syncwarp();
        if (e >= inflight) {
          // emission e - inflight (buffer done_b) has been read out once the wait below returns
          const int w_ahead = aux_tma ? tile_at(la_seq) : -1;   // (aux epilogues emit once per chunk)
          if (ptx::elect_one()) {
            if (inflight <= 0) ptx::bulk_wait_read<0>(); else if (inflight == 1) ptx::bulk_wait_read<1>(); else ptx::bulk_wait_read<2>();
            if (aux_tma) { if (w_ahead >= 0) load_aux(w_ahead, la_j, done_b); }
            else ptx::mbar_arrive(&freebar[done_b]);
          }
          __syncwarp();
          if (aux_tma) { if (++la_j == CHG) { la_j = 0; ++la_seq; } }
          if (++done_b == nbuf) done_b = 0;
        }
        if (++e_sub == EPC) {
          e_sub = 0;
          if (++e_j == CHG) { e_j = 0; ++e_seq; }
        }
        if (++e_b == nbuf) { e_b = 0; e_ph ^= 1u; }
      }
      if (ptx::elect_one()) ptx::bulk_wait_read<0>();   // smem may be released once the stores have read it
      __syncwarp();
      gstamp_if(grp == 0 && lane == 0, 7);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (CG == 2) ptx::cluster_sync();                   // the leader's MMAs read the peer's smem / TMEM until here
  if (warp == 1) {
    if (CG == 2) ptx::tmem_dealloc_pair(tmem_base, TMEM_COLS);
    else ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
  gstamp(3);
}

// ---- host: tensor maps --------------------------------------------------------------------------
namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct MapKey {
  const void* ptr; uint64_t d0, d1, stride; uint32_t b0, b1, esize;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && stride == o.stride && b0 == o.b0 && b1 == o.b1 && esize == o.esize;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    auto mix = [&](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
    mix(k.d0); mix(k.d1); mix(k.stride); mix(k.b0); mix(k.b1); mix(k.esize);
    return h;
  }
};
std::mutex g_map_mu;
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;
}  // namespace

// 2-D tensor [d1 rows][d0 contiguous] of bf16 (esize 2) or fp32 (esize 4), row stride in elements,
// box {b0, b1}, 128-byte swizzle (b0 * esize must be 128)
int make_tmap_2d(CUtensorMap* out, const void* ptr, int esize, uint64_t d0, uint64_t d1, uint64_t row_stride,
                 uint32_t b0, uint32_t b1) {
  MapKey key{ptr, d0, d1, row_stride, b0, b1, static_cast<uint32_t>(esize)};
  {
    std::lock_guard<std::mutex> g(g_map_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) { *out = it->second; return 0; }
  }
  EncodeTiledFn fn = get_encode_fn();
  NSB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  NSB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (row_stride * esize) % 16 == 0,
              "gemm: operand must be 16-byte aligned with a row stride multiple of 16 bytes (stride %llu elements)",
              static_cast<unsigned long long>(row_stride));
  cuuint64_t gdim[2] = {d0, d1};
  cuuint64_t gstr[1] = {row_stride * static_cast<uint64_t>(esize)};
  cuuint32_t box[2] = {b0, b1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, esize == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  const_cast<void*>(ptr), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  NSB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (dims %llu x %llu, stride %llu)",
              static_cast<int>(r), static_cast<unsigned long long>(d0), static_cast<unsigned long long>(d1),
              static_cast<unsigned long long>(row_stride));
  std::lock_guard<std::mutex> g(g_map_mu);
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps.emplace(key, *out);
  return 0;
}

// 3-D bf16 tensor [d2][d1 rows][d0 contiguous] (strides in elements), box {b0, b1, 1}, 128-byte swizzle: a store
// through it is clipped at d1 inside every d2 slice (rows past a sequence's end are not written)
int make_tmap_3d_bf16(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t row_stride,
                      uint64_t slice_stride, uint32_t b0, uint32_t b1) {
  NSB_REQUIRE(d1 < (1ull << 32) && d2 < (1ull << 31), "tensor map: dimensions too large");
  MapKey key{ptr, d0, d1 + (d2 << 32), row_stride ^ (slice_stride << 20), b0, b1, 2u + 256u};
  {
    std::lock_guard<std::mutex> g(g_map_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) { *out = it->second; return 0; }
  }
  EncodeTiledFn fn = get_encode_fn();
  NSB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  NSB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (row_stride * 2) % 16 == 0 && (slice_stride * 2) % 16 == 0,
              "tensor map: operand must be 16-byte aligned with strides multiple of 16 bytes");
  cuuint64_t gdim[3] = {d0, d1, d2};
  cuuint64_t gstr[2] = {row_stride * 2, slice_stride * 2};
  cuuint32_t box[3] = {b0, b1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  NSB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (3-D) failed with %d (dims %llu x %llu x %llu)", static_cast<int>(r),
              static_cast<unsigned long long>(d0), static_cast<unsigned long long>(d1), static_cast<unsigned long long>(d2));
  std::lock_guard<std::mutex> g(g_map_mu);
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps.emplace(key, *out);
  return 0;
}

static long long* g_gemm_dbg = nullptr;
static std::atomic<unsigned> g_sched_seq{0};
static int g_dynamic_sched = 1;
static int g_store_hint = 0;   // evict_first on the TMA stores of outputs larger than L2: measured no effect (lm_head forward 717-754 us either way), off
static int g_force_nbuf = 0, g_force_stages = 0, g_pipe_epi = 1;   // debug: staging buffers per group / operand ring depth
struct LaunchPlan {
  CUtensorMap ta, tb, tc, tc2, taux;
  int K, num_m, num_n, splits, kb_per_split, stages, nbuf, tma_epi, ngroups, aux_direct, n_fast, pipe_epi;
};

template <int BN, bool A_MN, bool B_MN, int CG, int NGRP = 2>
static int launch_variant(const LaunchPlan& lp, const EpiParams& ep, cudaStream_t stream) {
  auto kern = gemm_tc_kernel<BN, A_MN, B_MN, CG, NGRP>;
  static PerDeviceOnce attr_done;
  if (attr_done.first_use()) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemLimit));
    NSB_REQUIRE(e == cudaSuccess, "gemm_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  }
  const size_t smem = smem_bytes_for(BN / CG, lp.stages, lp.nbuf * lp.ngroups);
  const int total = lp.num_m * lp.num_n * lp.splits;
  // dynamic tile scheduling (g_dynamic_sched, default on): this launch's counter pair; -1 = static walk
  const int sched_slot = g_dynamic_sched ? static_cast<int>(g_sched_seq.fetch_add(1, std::memory_order_relaxed) % kSchedSlots) : -1;
  const int units = sm_count() / CG;
  const int grid = (total < units ? total : units) * CG;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(64 + 160 * NGRP);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, lp.ta, lp.tb, lp.tc, lp.tc2, lp.taux, lp.K, lp.num_m, lp.num_n, lp.splits,
                                     lp.kb_per_split, lp.stages, lp.nbuf, lp.tma_epi, lp.ngroups, lp.aux_direct, lp.n_fast, lp.pipe_epi, sched_slot, ep, g_gemm_dbg);
  NSB_REQUIRE(e == cudaSuccess, "gemm_tc: launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return check_launch("gemm_tc");
}

int gemm_tc_launch(const nsb_gemm_args* a, const EpiParams& ep, cudaStream_t stream, int force_bn, int force_splits,
                   int force_direct_epi, int force_groups, int force_cg, int force_nfast) {
  const int M = a->M, N = a->N, K = a->K;
  int bn = (N % 256 == 0 || N > 2048) ? 256 : 128;
  // a long mainloop pays for the pair tile even when the last 256-wide tile is half empty (N = 384, K >= 1024:
  // 31.3 -> 24.2 us at 16384 x 384 x 1536)
  if (N > 256 && N % 256 >= 128 && K >= 1024) bn = 256;
  if (force_bn == 128 || force_bn == 256) bn = force_bn;
  // CTA pairs (256-row tiles) for the wide-N tile whenever there are at least two 128-row blocks
  int cg = (bn == 256 && M > 128) ? 2 : 1;
  if (force_cg == 1 || (force_cg == 2 && bn == 256)) cg = force_cg;
  LaunchPlan lp;
  lp.K = K;
  lp.num_m = (M + BM * cg - 1) / (BM * cg);
  lp.num_n = (N + bn - 1) / bn;
  const int total_kb = (K + BK - 1) / BK;
  // split-K only where the epilogue can accumulate into a pre-existing fp32 C.  Pick the split that
  // minimises waves x (k-blocks per split + fixed per-tile cost): wave quantisation matters more
  // than the number of splits for the short-and-wide weight-gradient shapes.
  int splits = 1;
  if (ep.kind == NSB_EPI_ACCUM && ep.accumulate) {
    const int tiles = lp.num_m * lp.num_n, sms = sm_count() / cg;
    if (tiles < 2 * sms) {
      double best = 1e30;
      const int max_splits = total_kb / 4 > 0 ? (total_kb / 4 < 32 ? total_kb / 4 : 32) : 1;
      for (int sp = 1; sp <= max_splits; ++sp) {
        const int kbs = (total_kb + sp - 1) / sp;
        const int real = (total_kb + kbs - 1) / kbs;
        const int waves = (tiles * real + sms - 1) / sms;
        const double cost = waves * (kbs + 14.0);  // ~14 k-block times of prologue/epilogue per tile
        if (cost < best - 1e-9) { best = cost; splits = real; }
      }
    }
    if (force_splits > 0) splits = force_splits;
  }
  int kb_per_split = (total_kb + splits - 1) / splits;
  if (kb_per_split < 1) kb_per_split = 1;
  splits = (total_kb + kb_per_split - 1) / kb_per_split;
  if (splits < 1) splits = 1;
  lp.splits = splits;
  lp.kb_per_split = kb_per_split;
  // tile walk: keep the operand that would otherwise be re-streamed from HBM once per block of the
  // other dimension (only matters when it does not fit in L2)
  {
    const double a_bytes = 2.0 * M * K, b_bytes = 2.0 * N * K, l2 = 100e6;
    const double cost_mfast = (a_bytes > l2 ? a_bytes * lp.num_n : a_bytes) + b_bytes;
    const double cost_nfast = a_bytes + (b_bytes > l2 ? b_bytes * lp.num_m : b_bytes);
    lp.n_fast = cost_nfast < cost_mfast ? 1 : 0;
    if (force_nfast >= 0) lp.n_fast = force_nfast;
  }

  int rc;
  if (a->a_kmajor) rc = make_tmap_2d(&lp.ta, a->A, 2, K, M, a->lda, BK, BM);
  else             rc = make_tmap_2d(&lp.ta, a->A, 2, M, K, a->lda, 64, BK);
  if (rc) return rc;
  if (a->b_kmajor) rc = make_tmap_2d(&lp.tb, a->B, 2, K, N, a->ldb, BK, bn / cg);
  else             rc = make_tmap_2d(&lp.tb, a->B, 2, N, K, a->ldb, 64, BK);
  if (rc) return rc;

  // staged (TMA) epilogue whenever every output/aux row is 16-byte addressable
  const bool c32 = (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_ACCUM || !ep.c_bf16);
  const int cesz = c32 ? 4 : 2;
  const bool has_aux = (ep.kind == NSB_EPI_RESIDUAL || ep.kind == NSB_EPI_DGELU);
  auto tma_ok = [&](const void* p, int esz) {
    return p != nullptr && (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (static_cast<size_t>(a->ldc) * esz) % 16 == 0;
  };
  bool tma_epi = !force_direct_epi && tma_ok(a->C, cesz);
  if (ep.kind == NSB_EPI_GELU) tma_epi = tma_epi && tma_ok(a->C2, cesz);
  if (has_aux) tma_epi = tma_epi && tma_ok(a->aux, ep.kind == NSB_EPI_RESIDUAL ? 4 : 2);
  lp.tma_epi = tma_epi ? 1 : 0;
  // bit 1: the output (plain store) is larger than L2 -> stores carry the evict_first hint
  if (tma_epi && g_store_hint && ep.kind == NSB_EPI_STORE &&
      static_cast<double>(M) * N * cesz > 256e6) lp.tma_epi |= 2;
  lp.tc = lp.ta; lp.tc2 = lp.ta; lp.taux = lp.ta;  // placeholders when unused
  if (tma_epi) {
    const uint32_t cw = c32 ? 32 : 64;
    rc = make_tmap_2d(&lp.tc, a->C, cesz, N, M, a->ldc, cw, BM);
    if (rc) return rc;
    if (ep.kind == NSB_EPI_GELU) { rc = make_tmap_2d(&lp.tc2, a->C2, cesz, N, M, a->ldc, cw, BM); if (rc) return rc; }
    if (has_aux) { rc = make_tmap_2d(&lp.taux, a->aux, ep.kind == NSB_EPI_RESIDUAL ? 4 : 2, N, M, a->ldc, cw, BM); if (rc) return rc; }
  }
  // ALU-heavy bf16 GELU epilogues get a second epilogue group; GELU backward then reads its
  // pre-activation tile straight from global (no aux staging buffers)
  const bool heavy = tma_epi && ep.c_bf16 && (ep.kind == NSB_EPI_GELU || ep.kind == NSB_EPI_DGELU);
  // Epilogue groups: the ALU-heavy GELU epilogues get two groups (8 warps); plain stores, accumulation
  // and the residual add (whose fp32 aux tile arrives by TMA into the staging buffers) run on one group,
  // which leaves more shared memory to the operand ring.
  lp.ngroups = ((heavy && force_groups != 1) || (tma_epi && force_groups == 2)) ? 2 : 1;
  if (ep.kind == NSB_EPI_RESIDUAL && force_groups != 2) lp.ngroups = 1;
  // ... except behind a short mainloop (K <= 1024: attention's c_proj), where the kernel is bound by the
  // 8 bytes / element the residual epilogue moves: two groups keep six aux / output tiles in flight instead of
  // four (measured at 12288 x 768 x 768: 24.5 us against 31.5)
  if (ep.kind == NSB_EPI_RESIDUAL && tma_epi && force_groups == 0 && K <= 1024) lp.ngroups = 2;
  // the same holds for every epilogue behind a very short mainloop (K <= 512, the 384-wide char model: six
  // k-blocks per tile against eight 32-column units to drain): 28.7 -> 23.5 us at 16384 x 1536 x 384
  if (tma_epi && force_groups == 0 && K <= 512 && bn == 256) lp.ngroups = 2;
  // GELU backward with two groups reads its pre-activation tile straight from global (no room for
  // aux staging buffers next to the output buffers of two groups)
  lp.aux_direct = (ep.kind == NSB_EPI_DGELU && lp.ngroups == 2) ? 1 : 0;
  // (the software-pipelined bf16 form overwrites the pre-activation tile in its staging buffer: no extra room needed)
  if (ep.kind == NSB_EPI_DGELU && g_pipe_epi && ep.c_bf16 && tma_epi) lp.aux_direct = 0;
  lp.nbuf = tma_epi ? (lp.ngroups == 2 ? 3 : 4) : 0;
  lp.pipe_epi = g_pipe_epi;
  if (tma_epi && g_force_nbuf >= 2) lp.nbuf = g_force_nbuf;
  int stages = g_force_stages > 0 ? g_force_stages : kMaxStages;
  while (stages > 2 && smem_bytes_for(bn / cg, stages, lp.nbuf * lp.ngroups) > kSmemLimit) --stages;
  lp.stages = stages;

#define NSB_GEMM_CASE(BNV, AMN, BMN, CGV) return launch_variant<BNV, AMN, BMN, CGV>(lp, ep, stream)
  const bool amn = !a->a_kmajor, bmn = !a->b_kmajor;
  if (bn == 256 && cg == 2) {
    if (!amn && !bmn) NSB_GEMM_CASE(256, false, false, 2);
    if (!amn && bmn) NSB_GEMM_CASE(256, false, true, 2);
    if (amn && !bmn) NSB_GEMM_CASE(256, true, false, 2);
    NSB_GEMM_CASE(256, true, true, 2);
  } else if (bn == 256) {
    if (!amn && !bmn) NSB_GEMM_CASE(256, false, false, 1);
    if (!amn && bmn) NSB_GEMM_CASE(256, false, true, 1);
    if (amn && !bmn) NSB_GEMM_CASE(256, true, false, 1);
    NSB_GEMM_CASE(256, true, true, 1);
  } else {
    if (!amn && !bmn) NSB_GEMM_CASE(128, false, false, 1);
    if (!amn && bmn) NSB_GEMM_CASE(128, false, true, 1);
    if (amn && !bmn) NSB_GEMM_CASE(128, true, false, 1);
    NSB_GEMM_CASE(128, true, true, 1);
  }
#undef NSB_GEMM_CASE
}

int gemm_simt_launch(const nsb_gemm_args* a, const EpiParams& ep, cudaStream_t stream);

}  // namespace nsb

using namespace nsb;

static int g_force_bn = 0, g_force_splits = 0, g_force_simt = 0, g_force_direct_epi = 0, g_force_groups = 0, g_force_cg = 0,
           g_force_nfast = -1;

extern "C" {

// test/tuning hooks (not part of the reference-facing surface): force the N tile, the split-K
// factor, route bf16 GEMMs through the SIMT kernel (force_simt = 1), or force the direct-store
// epilogue (force_simt = 2) for cross-checks.
void nsb_gemm_debug_config(int force_bn, int force_splits, int force_simt) {
  g_force_bn = force_bn;
  g_force_splits = force_splits;
  g_force_simt = force_simt == 1;
  g_force_direct_epi = force_simt == 2;
  g_force_groups = force_simt == 3 ? 1 : (force_simt == 4 ? 2 : (force_simt == 7 ? 4 : 0));  // 3: one epilogue group, 4: two, 7: four (register stores)
  g_force_cg = force_simt == 5 ? 1 : (force_simt == 6 ? 2 : 0);      // 5: single-CTA tiles, 6: CTA pairs
  g_force_nfast = force_simt == 8 ? 0 : (force_simt == 9 ? 1 : -1);   // 8: m-fastest tile walk, 9: n-fastest
}

// device buffer (>= 1024 int64) receiving clock64 stamps of CTA 0's first epilogue thread; NULL = off
// 0 disables programmatic dependent launch for every kernel of the library (prologue overlap with the
// previous kernel's tail)
void nsb_gemm_debug_pdl(int enable) { nsb::set_pdl_enabled(enable != 0); }
// nbuf/stages: 0 = policy default; pipe_epi < 0 keeps the current setting (1 = software-pipelined epilogue)
void nsb_gemm_debug_pipeline(int nbuf, int stages, int pipe_epi) {
  nsb::g_force_nbuf = nbuf;
  nsb::g_force_stages = stages;
  if (pipe_epi >= 0) nsb::g_pipe_epi = pipe_epi;
}
// 1 (default): tiles are drawn from a per-launch counter by the CTAs that are running; 0: fixed stride walk
void nsb_gemm_debug_scheduler(int dynamic) { nsb::g_dynamic_sched = dynamic != 0; }
void nsb_gemm_debug_store_hint(int on) { nsb::g_store_hint = on != 0; }
void nsb_gemm_debug_timing(void* device_buffer) { nsb::g_gemm_dbg = static_cast<long long*>(device_buffer); }

int nsb_gemm(const nsb_gemm_args* a, nsb_stream_t stream) {
  NSB_REQUIRE(a != nullptr, "gemm: null args");
  NSB_REQUIRE(a->M >= 0 && a->N >= 0 && a->K >= 1, "gemm: bad shape %d x %d x %d", a->M, a->N, a->K);
  NSB_REQUIRE(a->in_dtype == NSB_F32 || a->in_dtype == NSB_BF16, "gemm: bad in_dtype %d", a->in_dtype);
  NSB_REQUIRE(a->epilogue >= NSB_EPI_STORE && a->epilogue <= NSB_EPI_DGELU, "gemm: bad epilogue %d", a->epilogue);
  NSB_REQUIRE(a->A && a->B && a->C, "gemm: null operand");
  NSB_REQUIRE(a->epilogue != NSB_EPI_GELU || a->C2, "gemm: GELU epilogue needs C2");
  NSB_REQUIRE((a->epilogue != NSB_EPI_RESIDUAL && a->epilogue != NSB_EPI_DGELU) || a->aux, "gemm: epilogue needs aux");
  if (a->M == 0 || a->N == 0) return 0;
  EpiParams ep;
  ep.kind = a->epilogue; ep.M = a->M; ep.N = a->N; ep.C = a->C; ep.C2 = a->C2; ep.aux = a->aux;
  ep.bias = a->bias; ep.alpha = a->alpha; ep.ldc = a->ldc; ep.c_bf16 = (a->c_dtype == NSB_BF16) ? 1 : 0;
  ep.accumulate = a->accumulate;
  if (a->in_dtype == NSB_F32 || g_force_simt) return gemm_simt_launch(a, ep, cs(stream));
  return gemm_tc_launch(a, ep, cs(stream), g_force_bn, g_force_splits, g_force_direct_epi, g_force_groups, g_force_cg, g_force_nfast);
}

}  // extern "C"
