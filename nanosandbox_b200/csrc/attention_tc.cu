// Causal flash attention on the 5th-gen tensor cores (bf16, head dim 64), forward pass.
//
// One CTA = 128 queries of one (batch, head); keys/values stream through in blocks of 128.
//   S  = Q K^T     tcgen05.mma 128x128x16 x4, operands by TMA (K-major, 128B swizzle), S in TMEM
//   P  = softmax   one thread per query row: tcgen05.ld of its S row, running max / sum in
//                  registers (no shuffles), exp2, P written as bf16 into a swizzled smem tile
//   O += P V       tcgen05.mma 128x64x16 x8, A = P (smem), B = V read MN-major in place; O stays
//                  in TMEM for the whole key loop.  The running max used in the exponent is only
//                  raised when the true max has grown by more than 2^8 (exact: softmax is shift
//                  invariant), so the O accumulator is rescaled (tcgen05.ld/st by the row's own
//                  thread) in the first block(s) only and the softmax warps never wait for P V.
// Warp roles: warp 0 TMA producer (Q once, K/V double-buffered), warp 1 MMA issuer + TMEM owner,
// warps 2..5 softmax / correction / output (one TMEM lane quarter each).  192 TMEM columns and
// 112 KB of smem per CTA let two CTAs share an SM: one's softmax overlaps the other's MMAs.
// q/k/v are read straight out of c_attn's [B, T, 3, H, 64] output through one tensor map.
#include "common.cuh"
#include "sm100_ptx.cuh"

#ifndef NSB_ATTN_POLY_OF8
#define NSB_ATTN_POLY_OF8 0   // measured: every share above 0 is slower (54.0 us at 0, 56.7 at 3, 59.8 at 5): the loop is issue-bound, not MUFU-bound
#endif

namespace nsb {

int make_tmap_2d(CUtensorMap* out, const void* ptr, int esize, uint64_t d0, uint64_t d1, uint64_t row_stride,
                 uint32_t b0, uint32_t b1);
int make_tmap_3d_bf16(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t row_stride,
                      uint64_t slice_stride, uint32_t b0, uint32_t b1);

namespace fatc {
constexpr int BQ = 128, BKV = 128, HD = 64;
constexpr int TILE = BQ * HD * 2;          // 16 KB: 128 rows x 128 B
constexpr int P_BYTES = BQ * BKV * 2;      // 32 KB: two 64-column swizzle atoms
constexpr int KV_STAGES = 2;
constexpr int SMEM_BYTES = TILE + KV_STAGES * 2 * TILE + P_BYTES + 256;
constexpr uint32_t TMEM_COLS = 256;        // S: 128 columns, O block: 64 columns
constexpr uint32_t S_COL = 0, O_COL = 128;
constexpr int kPolyOf8 = NSB_ATTN_POLY_OF8;   // share of the forward softmax's exp2 done by polynomial (0..8 of 8)
constexpr uint32_t P_COL = 192;            // P as bf16 pairs (64 columns) when it is kept in tensor memory for P V

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t sw128(int row, int q16) {
  return static_cast<uint32_t>(row * 128 + ((q16 ^ (row & 7)) << 4));
}
}  // namespace fatc

// One pass over a 128-key block of one query row: P = exp2(S*scale - m) -> bf16 -> swizzled smem,
// returning the row sum of P and (through mx) the raw row maximum.  `lim` (diagonal block only) is
// the last key column this row may attend to.  TMEM loads are software-pipelined: the next
// 32-column chunk is in flight while the current one is processed.
template <bool DIAG>
__device__ __forceinline__ float softmax_pass(uint32_t s_addr, uint8_t* sP, int row, int lim, float scale_log2,
                                              float m, float& mx, uint64_t* pv_bar, uint32_t pv_parity, uint32_t p_tmem) {
  using namespace fatc;
  float va[32], vb[32];
  float s0 = 0.f, s1 = 0.f;
  mx = -INFINITY;
  ptx::tmem_ld_32x32_async(s_addr, va);
  ptx::tmem_wait_ld32(va);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float* cur = (c & 1) ? vb : va;
    float* nxt = (c & 1) ? va : vb;
    if (c < 3) ptx::tmem_ld_32x32_async(s_addr + (c + 1) * 32, nxt);
    // diagonal block: the mask is decided per 32 x 32 sub-block (uniform over the warp's 32 query rows):
    // key chunks left of the diagonal take the unmasked path below, the chunk the diagonal crosses masks
    // per element, chunks right of it are zeros without any exp
    if (DIAG && c > (row >> 5)) {
#pragma unroll
      for (int i = 0; i < 32; ++i) cur[i] = 0.f;
    } else if (DIAG && c == (row >> 5)) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (c * 32 + i <= lim) mx = fmaxf(mx, cur[i]);
        float p = ex2(fmaf(cur[i], scale_log2, -m));
        if (c * 32 + i > lim) p = 0.f;
        cur[i] = p;
        if (i & 1) s1 += p; else s0 += p;
      }
    } else {
      // full block: two columns per step on the packed fp32 pipe (same operations and summation order
      // as the scalar loop above: even columns into s0, odd ones into s1) and a 3-input max
      const f32x2 sc2 = pack2(scale_log2, scale_log2), nm2 = pack2(-m, -m);
      f32x2 acc = pack2(s0, s1);
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        mx = max3(mx, cur[i], cur[i + 1]);
        float a0, a1;
        unpack2(fma2(pack2(cur[i], cur[i + 1]), sc2, nm2), a0, a1);
        // kPolyOf8 of every 8 column pairs take their exp2 on the FMA pipe (ex2_poly2), the rest on the
        // special-function unit, which one exp per score would otherwise saturate
        if (((i >> 1) & 7) < kPolyOf8) ex2_poly2(a0, a1, cur[i], cur[i + 1]);
        else { cur[i] = ex2(a0); cur[i + 1] = ex2(a1); }
        acc = add2(acc, pack2(cur[i], cur[i + 1]));
      }
      unpack2(acc, s0, s1);
    }
    uint8_t* atom = sP + (c >> 1) * TILE;
    // the previous block's P V may still be reading this tile: wait right before the first store
    if (c == 0 && pv_bar != nullptr) ptx::mbar_wait(pv_bar, pv_parity);
    if (p_tmem != 0) {
      // P stays in tensor memory (this query row, 16 packed columns per 32 keys): A operand of P V from there
      uint32_t pw[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) pw[i] = pack_bf16x2(cur[2 * i], cur[2 * i + 1]);
      ptx::tmem_st_32x16_u32_nowait(p_tmem + c * 16, pw);
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(atom + sw128(row, (c & 1) * 4 + i)) =
            make_uint4(pack_bf16x2(cur[8 * i], cur[8 * i + 1]), pack_bf16x2(cur[8 * i + 2], cur[8 * i + 3]),
                       pack_bf16x2(cur[8 * i + 4], cur[8 * i + 5]), pack_bf16x2(cur[8 * i + 6], cur[8 * i + 7]));
    }
    if (c < 3) ptx::tmem_wait_ld32(nxt);
  }
  if (p_tmem != 0) ptx::tmem_wait_st();
  return s0 + s1;
}

// Running-softmax update for one block.  The exponent reference m is "lazy": the common case is a
// single pass with the current m; only when some row's maximum outgrew it by more than 2^8 (always
// in the first block) is m raised, the pass repeated and the O accumulator rescaled.  Exact, since
// softmax is shift invariant; P stays below 2^8.  Returns the O rescale factor (1 = none).
template <bool DIAG>
__device__ __forceinline__ float softmax_block(uint32_t s_addr, uint8_t* sP, int row, int lim, float scale_log2,
                                               float& m, float& l, bool first, uint64_t* pv_bar, uint32_t pv_parity,
                                               uint32_t p_tmem) {
  float mx;
  float sum = 0.f;
  if (!first) sum = softmax_pass<DIAG>(s_addr, sP, row, lim, scale_log2, m, mx, pv_bar, pv_parity, p_tmem);
  else {
    // first block: only the maximum is needed (m = -inf would make every P infinite)
    float va[32];
    mx = -INFINITY;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      if (DIAG && c > (row >> 5)) break;             // key chunks right of the diagonal (warp-uniform)
      ptx::tmem_ld_32x32(s_addr + c * 32, va);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (DIAG) { if (c * 32 + i <= lim) mx = fmaxf(mx, va[i]); }
        else mx = fmaxf(mx, va[i]);
      }
    }
  }
  float corr = 1.f;
  const float m_true = mx * scale_log2;
  const bool grow = m_true > m + 8.f;
  if (__any_sync(0xffffffffu, grow)) {               // warp-uniform (TMEM loads are warp-collective)
    if (grow) {
      corr = fatc::ex2(m - m_true);                  // first block: exp2(-inf) = 0
      l *= corr;
      m = m_true;
    }
    float dummy;
    // (a repeated pass finds the previous P V retired already: the first pass waited for it)
    sum = softmax_pass<DIAG>(s_addr, sP, row, lim, scale_log2, m, dummy, nullptr, 0, p_tmem);
  }
  l += sum;
  return corr;
}

// Optional phase timestamps (clock64) of CTA (0,0): softmax warp 2 lane 0 writes 4 stamps per key
// block at dbg[0..], the MMA thread 4 stamps per block at dbg[512..].  NULL in production.
__global__ void __launch_bounds__(192, 2)
    attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, __nv_bfloat16* __restrict__ y,
                       float* __restrict__ lse, int Tn, int H, float scale_log2, int p_in_tmem, long long* dbg) {
  using namespace fatc;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + TILE;                       // stage s: K at s*2*TILE, V at s*2*TILE + TILE
  uint8_t* sP = sKV + KV_STAGES * 2 * TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + P_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;                     // [2]
  uint64_t* kv_empty = bars + 3;                    // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* o_full = bars + 7;
  uint64_t* pv_done = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = H * HD;
  const int qb = gridDim.y - 1 - blockIdx.y;        // longest query blocks of ALL heads are scheduled first
  const int bh = blockIdx.x, b = bh / H, h = bh % H;
  const int q0 = qb * BQ;
  const int nblk = qb + 1;                          // causal: key blocks 0..qb
  const int row_base = b * Tn;                      // row of token 0 of this batch in [B*T, 3C]

  if (dbg != nullptr && threadIdx.x == 0) {         // per-CTA trace: start ns, end ns, SM id
    unsigned long long t; unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    long long* rec3 = dbg + 1024 + 3 * (blockIdx.y * gridDim.x + blockIdx.x);
    rec3[0] = static_cast<long long>(t);
    rec3[2] = smid;
  }
  if (threadIdx.x == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tmap(&tmap_qkv);
  }
  if (warp == 0 && lane < 9) {                       // one barrier per lane: q_full, kv_full[2], kv_empty[2], s_full, p_full, o_full, pv_done
    static_assert(KV_STAGES == 2, "barrier layout");
    ptx::mbar_init(&bars[lane], lane == 6 ? 4u : 1u);   // p_full collects the four softmax warps
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // barriers, TMEM and the descriptor prefetch above overlap the previous kernel's tail
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire

  if (warp == 0) {
    if (lane == 0) {
      ptx::mbar_expect_tx(q_full, TILE);
      ptx::tma_load_2d(sQ, &tmap_qkv, q_full, h * HD, row_base + q0);
      for (int j = 0; j < nblk; ++j) {
        const int s = j % KV_STAGES;
        ptx::mbar_wait(&kv_empty[s], ((j / KV_STAGES) & 1) ^ 1);
        ptx::mbar_expect_tx(&kv_full[s], 2 * TILE);
        ptx::tma_load_2d(sKV + s * 2 * TILE, &tmap_qkv, &kv_full[s], C + h * HD, row_base + j * BKV);
        ptx::tma_load_2d(sKV + s * 2 * TILE + TILE, &tmap_qkv, &kv_full[s], 2 * C + h * HD, row_base + j * BKV);
      }
    }
  } else if (warp == 1) {
    // MMA issuer: the warp stays converged, one elected lane issues (see ptx::elect_one)
    {
      constexpr uint32_t IDESC_S = ptx::umma_idesc_bf16(BQ, BKV, false, false);
      constexpr uint32_t IDESC_O = ptx::umma_idesc_bf16(BQ, HD, false, true);   // B = V, MN-major
      const bool rec = dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0;
      // descriptors: only the 14-bit start-address field changes between MMAs
      const uint64_t q_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(sQ), 16, 1024);
      const uint64_t p_desc0 = ptx::umma_smem_desc_sw128(ptx::smem_u32(sP), 16, 1024);
      const uint64_t k_desc0 = ptx::umma_smem_desc_sw128(ptx::smem_u32(sKV), 16, 1024);
      const uint64_t v_desc0 = ptx::umma_smem_desc_sw128(ptx::smem_u32(sKV + TILE), 64 * BKV * 2, 1024);
      auto issue_s = [&](int jj) {                 // S = Q K_jj^T
        const uint64_t k_desc = k_desc0 + static_cast<uint64_t>(((jj % KV_STAGES) * 2 * TILE) >> 4);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)
            ptx::umma_bf16(tmem + S_COL, q_desc + static_cast<uint64_t>(2 * k), k_desc + static_cast<uint64_t>(2 * k),
                           IDESC_S, k > 0 ? 1u : 0u);
          ptx::umma_commit(s_full);
        }
        __syncwarp();
      };
      ptx::mbar_wait(q_full, 0);
      ptx::mbar_wait(&kv_full[0], 0);
      ptx::tc_fence_after();
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        const int s = j % KV_STAGES;
        if (rec) dbg[512 + 4 * j] = clock64();
        ptx::mbar_wait(p_full, j & 1);              // P_j in smem, S consumed, O rescaled if needed
        ptx::tc_fence_after();
        if (rec) dbg[512 + 4 * j + 1] = clock64();
        if (j + 1 < nblk) {                         // next S first: the softmax warps wait for it
          ptx::mbar_wait(&kv_full[(j + 1) % KV_STAGES], ((j + 1) / KV_STAGES) & 1);
          ptx::tc_fence_after();
          issue_s(j + 1);
        }
        if (rec) dbg[512 + 4 * j + 2] = clock64();
        // O += P_j V_j.  P: K-major, two 64-column atoms of 16 KB; V: MN-major, 16 key rows = 2048 B / step
        const uint64_t v_desc = v_desc0 + static_cast<uint64_t>((s * 2 * TILE) >> 4);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k)
            if (p_in_tmem) ptx::umma_bf16_ts(tmem + O_COL, tmem + P_COL + 8 * k, v_desc + static_cast<uint64_t>(128 * k), IDESC_O,
                                             (j > 0 || k > 0) ? 1u : 0u);
            else ptx::umma_bf16(tmem + O_COL, p_desc0 + static_cast<uint64_t>(((k / 4) * TILE + (k % 4) * 32) >> 4),
                                v_desc + static_cast<uint64_t>(128 * k), IDESC_O, (j > 0 || k > 0) ? 1u : 0u);
          ptx::umma_commit(pv_done);                // P smem may be overwritten
          ptx::umma_commit(&kv_empty[s]);           // K/V stage may be refilled
          if (j == nblk - 1) ptx::umma_commit(o_full);
        }
        __syncwarp();
        if (rec) dbg[512 + 4 * j + 3] = clock64();
      }
    }
  } else {
    const int qd = warp & 3;
    const int row = qd * 32 + lane;                 // query row within the tile = TMEM lane
    const int qi = q0 + row;
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    float m = -INFINITY, l = 0.f;
    const uint32_t s_addr = tmem + lane_addr + S_COL, o_addr = tmem + lane_addr + O_COL;
    const uint32_t p_addr = p_in_tmem ? tmem + lane_addr + P_COL : 0u;
    const bool rec = dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && warp == 2 && lane == 0;
    for (int j = 0; j < nblk; ++j) {
      // S_j is ready; P V of block j-1 may still be running (pv_done is awaited before P is overwritten)
      if (rec) dbg[4 * j] = clock64();
      ptx::mbar_wait(s_full, j & 1);
      ptx::tc_fence_after();
      if (rec) dbg[4 * j + 1] = clock64();
      uint64_t* pvb = j > 0 ? pv_done : nullptr;
      const uint32_t pvp = static_cast<uint32_t>((j - 1) & 1);
      const float corr = (j == qb) ? softmax_block<true>(s_addr, sP, row, qi - j * BKV, scale_log2, m, l, j == 0, pvb, pvp, p_addr)
                                   : softmax_block<false>(s_addr, sP, row, 0, scale_log2, m, l, j == 0, pvb, pvp, p_addr);
      if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {
        // rare: this warp's rows rescale their O accumulator in TMEM (warp-uniform branch)
        float v[32];
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          ptx::tmem_ld_32x32(o_addr + c * 32, v);
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] *= corr;
          ptx::tmem_st_32x32(o_addr + c * 32, v);
        }
      }
      if (rec) dbg[4 * j + 2] = clock64();
      ptx::fence_proxy_async();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
      if (rec) dbg[4 * j + 3] = clock64();
    }
    ptx::mbar_wait(o_full, 0);
    ptx::tc_fence_after();
    {
      const float inv = 1.f / l;
      const bool live = qi < Tn;                     // TMEM loads stay warp-collective; stores are predicated
      uint4* dst = reinterpret_cast<uint4*>(y + (static_cast<size_t>(row_base) + (live ? qi : 0)) * C + h * HD);
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        float o[32];
        ptx::tmem_ld_32x32(o_addr + c * 32, o);
        if (live) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[c * 4 + i] = make_uint4(pack_bf16x2(o[8 * i] * inv, o[8 * i + 1] * inv), pack_bf16x2(o[8 * i + 2] * inv, o[8 * i + 3] * inv),
                                        pack_bf16x2(o[8 * i + 4] * inv, o[8 * i + 5] * inv), pack_bf16x2(o[8 * i + 6] * inv, o[8 * i + 7] * inv));
        }
      }
      if (live) lse[(static_cast<size_t>(b) * H + h) * Tn + qi] = (m + log2f(l)) * 0.69314718055994531f;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, TMEM_COLS);
  if (dbg != nullptr && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    dbg[1024 + 3 * (blockIdx.y * gridDim.x + blockIdx.x) + 1] = static_cast<long long>(t);
  }
}

// ---- forward with eight softmax warps per CTA (four per SM sub-partition with two CTAs per SM) -------
// Same pipeline as attn_fwd_tc_kernel; each query row is shared by two threads (64 key columns each), P always
// lives in tensor memory, 80 registers per thread.  Warps: 0 TMA, 1 MMA issue, 2..9 softmax.
__global__ void __launch_bounds__(320, 2)
    attn_fwd_tc8_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_y,
                       float* __restrict__ lse, int Tn, int H, float scale_log2) {
  constexpr int p_in_tmem = 1;
  long long* const dbg = nullptr;
  using namespace fatc;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + TILE;                       // stage s: K at s*2*TILE, V at s*2*TILE + TILE
  uint8_t* sP = sKV + KV_STAGES * 2 * TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + P_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;                     // [2]
  uint64_t* kv_empty = bars + 3;                    // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* o_full = bars + 7;
  uint64_t* pv_done = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = H * HD;
  const int qb = gridDim.y - 1 - blockIdx.y;        // longest query blocks of ALL heads are scheduled first
  const int bh = blockIdx.x, b = bh / H, h = bh % H;
  const int q0 = qb * BQ;
  const int nblk = qb + 1;                          // causal: key blocks 0..qb
  const int row_base = b * Tn;                      // row of token 0 of this batch in [B*T, 3C]

  if (dbg != nullptr && threadIdx.x == 0) {         // per-CTA trace: start ns, end ns, SM id
    unsigned long long t; unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    long long* rec3 = dbg + 1024 + 3 * (blockIdx.y * gridDim.x + blockIdx.x);
    rec3[0] = static_cast<long long>(t);
    rec3[2] = smid;
  }
  if (threadIdx.x == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tmap(&tmap_qkv);
  }
  if (warp == 0 && lane < 9) {                       // one barrier per lane: q_full, kv_full[2], kv_empty[2], s_full, p_full, o_full, pv_done
    static_assert(KV_STAGES == 2, "barrier layout");
    ptx::mbar_init(&bars[lane], lane == 6 ? 8u : 1u);   // p_full collects the eight softmax warps
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // barriers, TMEM and the descriptor prefetch above overlap the previous kernel's tail
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire

  if (warp == 0) {
    if (lane == 0) {
      ptx::mbar_expect_tx(q_full, TILE);
      ptx::tma_load_2d(sQ, &tmap_qkv, q_full, h * HD, row_base + q0);
      for (int j = 0; j < nblk; ++j) {
        const int s = j % KV_STAGES;
        ptx::mbar_wait(&kv_empty[s], ((j / KV_STAGES) & 1) ^ 1);
        ptx::mbar_expect_tx(&kv_full[s], 2 * TILE);
        ptx::tma_load_2d(sKV + s * 2 * TILE, &tmap_qkv, &kv_full[s], C + h * HD, row_base + j * BKV);
        ptx::tma_load_2d(sKV + s * 2 * TILE + TILE, &tmap_qkv, &kv_full[s], 2 * C + h * HD, row_base + j * BKV);
      }
    }
  } else if (warp == 1) {
    // MMA issuer: the warp stays converged, one elected lane issues (see ptx::elect_one)
    {
      constexpr uint32_t IDESC_S = ptx::umma_idesc_bf16(BQ, BKV, false, false);
      constexpr uint32_t IDESC_O = ptx::umma_idesc_bf16(BQ, HD, false, true);   // B = V, MN-major
      const bool rec = dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0;
      // descriptors: only the 14-bit start-address field changes between MMAs
      const uint64_t q_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(sQ), 16, 1024);
      const uint64_t p_desc0 = ptx::umma_smem_desc_sw128(ptx::smem_u32(sP), 16, 1024);
      const uint64_t k_desc0 = ptx::umma_smem_desc_sw128(ptx::smem_u32(sKV), 16, 1024);
      const uint64_t v_desc0 = ptx::umma_smem_desc_sw128(ptx::smem_u32(sKV + TILE), 64 * BKV * 2, 1024);
      auto issue_s = [&](int jj) {                 // S = Q K_jj^T
        const uint64_t k_desc = k_desc0 + static_cast<uint64_t>(((jj % KV_STAGES) * 2 * TILE) >> 4);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)
            ptx::umma_bf16(tmem + S_COL, q_desc + static_cast<uint64_t>(2 * k), k_desc + static_cast<uint64_t>(2 * k),
                           IDESC_S, k > 0 ? 1u : 0u);
          ptx::umma_commit(s_full);
        }
        __syncwarp();
      };
      ptx::mbar_wait(q_full, 0);
      ptx::mbar_wait(&kv_full[0], 0);
      ptx::tc_fence_after();
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        const int s = j % KV_STAGES;
        if (rec) dbg[512 + 4 * j] = clock64();
        ptx::mbar_wait(p_full, j & 1);              // P_j in smem, S consumed, O rescaled if needed
        ptx::tc_fence_after();
        if (rec) dbg[512 + 4 * j + 1] = clock64();
        if (j + 1 < nblk) {                         // next S first: the softmax warps wait for it
          ptx::mbar_wait(&kv_full[(j + 1) % KV_STAGES], ((j + 1) / KV_STAGES) & 1);
          ptx::tc_fence_after();
          issue_s(j + 1);
        }
        if (rec) dbg[512 + 4 * j + 2] = clock64();
        // O += P_j V_j.  P: K-major, two 64-column atoms of 16 KB; V: MN-major, 16 key rows = 2048 B / step
        const uint64_t v_desc = v_desc0 + static_cast<uint64_t>((s * 2 * TILE) >> 4);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k)
            if (p_in_tmem) ptx::umma_bf16_ts(tmem + O_COL, tmem + P_COL + 8 * k, v_desc + static_cast<uint64_t>(128 * k), IDESC_O,
                                             (j > 0 || k > 0) ? 1u : 0u);
            else ptx::umma_bf16(tmem + O_COL, p_desc0 + static_cast<uint64_t>(((k / 4) * TILE + (k % 4) * 32) >> 4),
                                v_desc + static_cast<uint64_t>(128 * k), IDESC_O, (j > 0 || k > 0) ? 1u : 0u);
          ptx::umma_commit(pv_done);                // P smem may be overwritten
          ptx::umma_commit(&kv_empty[s]);           // K/V stage may be refilled
          if (j == nblk - 1) ptx::umma_commit(o_full);
        }
        __syncwarp();
        if (rec) dbg[512 + 4 * j + 3] = clock64();
      }
    }
  } else {
    // ---- eight softmax warps: group g takes key columns 64 g .. 64 g + 63 of every block; the two threads of a
    //      query row (same lane, warps w and w + 4) keep the same exponent reference m and exchange their
    //      row maxima once per block through shared memory; each keeps the sum of its own columns ----
    const int g = (warp - 2) >> 2;
    const int qd = warp & 3;
    const int row = qd * 32 + lane;                 // query row within the tile = TMEM lane
    const int rq = row >> 5;
    const int qi = q0 + row;
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    float m = -INFINITY, l = 0.f;
    const uint32_t s_addr = tmem + lane_addr + S_COL + g * 64;
    const uint32_t o_addr = tmem + lane_addr + O_COL + g * 32;   // this thread's half of the O row
    const uint32_t p_addr = tmem + lane_addr + P_COL + g * 32;   // its 64 probabilities as 32 bf16 pairs
    float* xch = reinterpret_cast<float*>(sP);      // [2 buffers][2 groups][128 rows] maxima, then [2][128] sums
    for (int j = 0; j < nblk; ++j) {
      ptx::mbar_wait(s_full, j & 1);
      ptx::tc_fence_after();
      const bool diag = (j == qb);
      bool p_free = (j == 0);                       // P of the previous block still being read by its P V?
      // one pass over this thread's 64 columns with the current m: P -> tensor memory, returns the sum of P
      // and (mx) the raw maximum.  Diagonal block: per 32 x 32 sub-block all valid / masked per element / zeros.
      auto pass = [&](float& mx) -> float {
        float s0 = 0.f, s1 = 0.f;
        mx = -INFINITY;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          const int cc = g * 2 + c;
          uint32_t pw[16];
          if (diag && cc > rq) {
#pragma unroll
            for (int i = 0; i < 16; ++i) pw[i] = 0u;
          } else {
            float cur[32];
            ptx::tmem_ld_32x32(s_addr + c * 32, cur);
            if (diag && cc == rq) {
              const int lim = row - cc * 32;        // last valid column of this chunk
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                if (i <= lim) mx = fmaxf(mx, cur[i]);
                float p = fatc::ex2(fmaf(cur[i], scale_log2, -m));
                if (i > lim) p = 0.f;
                cur[i] = p;
                if (i & 1) s1 += p; else s0 += p;
              }
            } else {
              const f32x2 sc2 = pack2(scale_log2, scale_log2), nm2 = pack2(-m, -m);
              f32x2 acc = pack2(s0, s1);
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                mx = max3(mx, cur[i], cur[i + 1]);
                float a0, a1;
                unpack2(fma2(pack2(cur[i], cur[i + 1]), sc2, nm2), a0, a1);
                cur[i] = fatc::ex2(a0);
                cur[i + 1] = fatc::ex2(a1);
                acc = add2(acc, pack2(cur[i], cur[i + 1]));
              }
              unpack2(acc, s0, s1);
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) pw[i] = pack_bf16x2(cur[2 * i], cur[2 * i + 1]);
          }
          if (!p_free) { ptx::mbar_wait(pv_done, (j - 1) & 1); ptx::tc_fence_after(); p_free = true; }
          ptx::tmem_st_32x16_u32_nowait(p_addr + c * 16, pw);
        }
        ptx::tmem_wait_st();
        return s0 + s1;
      };
      float mx, sum = 0.f;
      if (j == 0) {
        // first block: only the maximum (m = -inf would make every P infinite)
        mx = -INFINITY;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          const int cc = g * 2 + c;
          if (diag && cc > rq) continue;            // warp-uniform
          float v[32];
          ptx::tmem_ld_32x32(s_addr + c * 32, v);
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (!diag || cc * 32 + i <= row) mx = fmaxf(mx, v[i]);
        }
      } else {
        sum = pass(mx);
      }
      // row maximum over both halves (double-buffered slots: the partner reads this block's value while I may
      // already be writing the next block's)
      float* slot = xch + (j & 1) * 256;
      slot[g * 128 + row] = mx;
      ptx::named_bar_sync(2 + qd, 64);               // only the two warps that share these 32 rows meet
      mx = fmaxf(mx, slot[(g ^ 1) * 128 + row]);
      float corr = 1.f;
      const float m_true = mx * scale_log2;
      const bool grow = m_true > m + 8.f;
      if (__any_sync(0xffffffffu, grow)) {           // same decision in both warps of these 32 rows
        if (grow) {
          corr = fatc::ex2(m - m_true);              // first block: exp2(-inf) = 0
          l *= corr;
          m = m_true;
        }
        float dummy;
        sum = pass(dummy);
      }
      l += sum;
      if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {
        // rare: rescale this thread's half of the O row in tensor memory
        float v[32];
        ptx::tmem_ld_32x32(o_addr, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] *= corr;
        ptx::tmem_st_32x32(o_addr, v);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
    }
    // total row sum, then this thread's 32 columns of O
    float* lsum = xch + 512;
    lsum[g * 128 + row] = l;
    ptx::named_bar_sync(2 + qd, 64);
    l += lsum[(g ^ 1) * 128 + row];
    ptx::mbar_wait(o_full, 0);
    ptx::tc_fence_after();
    {
      // The O rows leave through the Q tile (free: every MMA has retired) and one TMA store, clipped at the end of
      // the sequence.  Stored from registers -- 64 bytes per thread at a row stride of C -- a warp's store
      // touches 32 lines and holds the load/store unit the SM's other CTA needs for its softmax.
      const float inv = 1.f / l;
      float o[32];
      ptx::tmem_ld_32x32(o_addr, o);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(sQ + sw128(row, g * 4 + i)) =
            make_uint4(pack_bf16x2(o[8 * i] * inv, o[8 * i + 1] * inv), pack_bf16x2(o[8 * i + 2] * inv, o[8 * i + 3] * inv),
                       pack_bf16x2(o[8 * i + 4] * inv, o[8 * i + 5] * inv), pack_bf16x2(o[8 * i + 6] * inv, o[8 * i + 7] * inv));
      if (g == 0 && qi < Tn) lse[(static_cast<size_t>(b) * H + h) * Tn + qi] = (m + log2f(l)) * 0.69314718055994531f;
      ptx::fence_proxy_async();
      ptx::named_bar_sync(6, 256);                   // the eight softmax warps
      if (threadIdx.x == 64) {
        ptx::tma_store_3d(&tmap_y, sQ, h * HD, q0, b);
        ptx::bulk_commit();
        ptx::bulk_wait_read<0>();                    // the tile must outlive the read; the CTA exits below
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, TMEM_COLS);
  if (dbg != nullptr && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    dbg[1024 + 3 * (blockIdx.y * gridDim.x + blockIdx.x) + 1] = static_cast<long long>(t);
  }
}


static long long* g_attn_dbg = nullptr;
static int g_attn_fwd_warps8 = 1;  // forward with eight softmax warps per CTA (attn_fwd_tc8_kernel); 0: the four-warp kernel
void attention_tc_set_fwd_warps8(int on) { g_attn_fwd_warps8 = on; }
static int g_attn_fwd_p_tmem = 1;  // forward: P kept in tensor memory for the P V MMA (0: through shared memory)
void attention_tc_set_fwd_p_tmem(int on) { g_attn_fwd_p_tmem = on; }
void attention_tc_set_debug_buffer(long long* p) { g_attn_dbg = p; }
static int g_attn_bwd_warps16 = 1;  // backward with 16 probability warps (attn_bwd_tc16_kernel); 0: the 8-warp kernel
void attention_tc_set_bwd_warps16(int on) { g_attn_bwd_warps16 = on; }
static int g_attn_bwd_persistent = 1;  // 16-warp backward as one persistent CTA per SM (attn_bwd_tc16p_kernel); 0: one CTA per item
void attention_tc_set_bwd_persistent(int on) { g_attn_bwd_persistent = on; }
static int g_attn_dq_direct = 0;   // 1: dQ leaves by vector reductions from registers (measured slower: 184 vs 156 us); 0: staged TMA reduce-add
void attention_tc_set_dq_direct(int on) { g_attn_dq_direct = on; }
// 0: every A operand from shared memory; 1: K / V as tensor-memory A operands of the S^T / dP^T MMAs (bit-identical,
// measured no faster); 2: P^T kept in tensor memory for the dV MMA (halves the probability warps' shared-memory stores)
static int g_attn_kv_tmem = 2;
void attention_tc_set_kv_tmem(int on) { g_attn_kv_tmem = on; }

int attention_fwd_tc_launch(const void* qkv, void* y, float* lse, int B, int T, int H, cudaStream_t stream) {
  using namespace fatc;
  const int C = H * HD;
  CUtensorMap tm;
  int rc = make_tmap_2d(&tm, qkv, 2, static_cast<uint64_t>(3) * C, static_cast<uint64_t>(B) * T,
                        static_cast<uint64_t>(3) * C, HD, BQ);
  if (rc) return rc;
  static PerDeviceOnce attr_done;
  if (attr_done.first_use()) {
    cudaError_t e = cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    NSB_REQUIRE(e == cudaSuccess, "attention_fwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  }
  const dim3 grid(B * H, (T + BQ - 1) / BQ);
  if (g_attn_fwd_warps8 && g_attn_dbg == nullptr) {
    static PerDeviceOnce attr8;
    if (attr8.first_use()) {
      cudaError_t e = cudaFuncSetAttribute(attn_fwd_tc8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
      NSB_REQUIRE(e == cudaSuccess, "attention_fwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    CUtensorMap ty;
    rc = make_tmap_3d_bf16(&ty, y, C, T, B, C, static_cast<uint64_t>(T) * C, HD, BQ);
    if (rc) return rc;
    launch_pdl(attn_fwd_tc8_kernel, dim3(grid), dim3(320), SMEM_BYTES, stream, tm, ty, lse, T, H,
               1.4426950408889634f / sqrtf(static_cast<float>(HD)));
  } else {
    launch_pdl(attn_fwd_tc_kernel, dim3(grid), dim3(192), SMEM_BYTES, stream, tm, static_cast<__nv_bfloat16*>(y), lse, T, H,
               1.4426950408889634f / sqrtf(static_cast<float>(HD)), g_attn_fwd_p_tmem, g_attn_dbg);
  }
  count_launch();
  return check_launch("attention_fwd_tc");
}


// =============================================================================================
// Backward pass on the tensor cores.
//
// One CTA owns 128 keys of one (batch, head) and walks the query blocks at or below the diagonal.
// Per query block (all MMAs are tcgen05, M = 128):
//   S^T  = K Q^T          (keys x queries)            A = K,    B = Q     (both K-major)
//   dP^T = V dO^T                                      A = V,    B = dO
//   P^T  = exp2(S^T*scale - lse[q]),  dS^T = P^T o (dP^T - delta[q])      one thread per key row
//   dV  += P^T dO                                      A = P^T (smem), B = dO read MN-major
//   dK  += dS^T Q                                      A = dS^T (smem), B = Q read MN-major
//   dQ   = dS K          (queries x d)                 A = dS^T read MN-major (the same smem tile,
//                                                      transposed by the descriptor), B = K MN-major
// dV and dK accumulate in TMEM over the whole loop; dQ of the block goes TMEM -> swizzled smem ->
// TMA reduce-add into an fp32 accumulator [B*T, C] shared by all key blocks (converted to bf16,
// and re-zeroed, by attn_dq_convert_kernel).  Warps: 0 = TMA, 1 = MMA issue, 2..9 = two compute
// groups, each taking 64 of the 128 query columns of every row, 10..13 = dQ write-back (so the
// compute warps go straight from one block's probabilities to the next).
// =============================================================================================
namespace fabw {
constexpr int BQ = 128, BKV = 128, HD = 64;
constexpr int TILE = 128 * 128;            // 16 KB: 128 rows x 128 B
constexpr int QS = 3;                       // Q / dO ring depth
constexpr int OFF_K = 0, OFF_V = TILE, OFF_Q = 2 * TILE, OFF_DO = (2 + QS) * TILE;
constexpr int OFF_PT = (2 + 2 * QS) * TILE, OFF_DS = OFF_PT + 2 * TILE, OFF_DQ = OFF_DS + 2 * TILE;   // 32 KB each
constexpr int OFF_STAT = OFF_DQ + 2 * TILE; // [2 stages][lse_log2 | delta][128] floats
constexpr int OFF_BAR = OFF_STAT + 2 * 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 256;
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t C_S = 0, C_DP = 128, C_DV = 256, C_DK = 320, C_DQ = 384;
constexpr uint32_t C_KA = 448, C_VA = 480;   // K and V rows as tensor-memory A operands (bf16 pairs, 32 columns each)
}  // namespace fabw

// 64 query columns of one key row: P^T = exp2(S^T*scale - lse[q]) and dS^T = P^T (dP^T - delta[q]),
// returned as packed bf16 pairs.  MASK: causal mask on the diagonal block / padding columns.
template <bool MASK>
__device__ __forceinline__ void bwd_probs(uint32_t taddr, const float* lse2, const float* dlt, float scale_log2,
                                          int col0, int row, int qlim, bool diag, uint32_t* pk, uint32_t* dk,
                                          long long* trace = nullptr) {
  using fatc::ex2;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    float sv[32], dv[32];
    ptx::tmem_ld_32x32_async(taddr + fabw::C_S + c * 32, sv);
    ptx::tmem_ld_32x32_async(taddr + fabw::C_DP + c * 32, dv);
    ptx::tmem_wait_ld32(sv);
    ptx::tmem_wait_ld32(dv);
    if (trace != nullptr) trace[2 * c] = clock64() + (__float_as_int(sv[0]) & 0) + (__float_as_int(dv[31]) & 0);
    // the per-query statistics are stored negated (-lse*log2e, -delta), so a pair of columns costs one
    // packed fma, two exp2, one packed add and one packed multiply
    const f32x2 sc2 = pack2(scale_log2, scale_log2);
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 l4 = *reinterpret_cast<const float4*>(lse2 + c * 32 + i);
      const float4 d4 = *reinterpret_cast<const float4*>(dlt + c * 32 + i);
      const float nl[4] = {l4.x, l4.y, l4.z, l4.w}, nd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
      for (int e = 0; e < 4; e += 2) {
        float a0, a1;
        unpack2(fma2(pack2(sv[i + e], sv[i + e + 1]), sc2, pack2(nl[e], nl[e + 1])), a0, a1);
        float p0 = ex2(a0), p1 = ex2(a1);
        if (MASK) {
          const int col = col0 + c * 32 + i + e;               // query column within the block
          p0 = ((col < qlim) && (!diag || col >= row)) ? p0 : 0.f;
          p1 = ((col + 1 < qlim) && (!diag || col + 1 >= row)) ? p1 : 0.f;
        }
        sv[i + e] = p0;
        sv[i + e + 1] = p1;
        unpack2(mul2(pack2(p0, p1), add2(pack2(dv[i + e], dv[i + e + 1]), pack2(nd[e], nd[e + 1]))), dv[i + e], dv[i + e + 1]);
      }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      pk[c * 16 + i] = pack_bf16x2(sv[2 * i], sv[2 * i + 1]);
      dk[c * 16 + i] = pack_bf16x2(dv[2 * i], dv[2 * i + 1]);
    }
    if (trace != nullptr) trace[2 * c + 1] = clock64() + (pk[c * 16 + 15] & 0) + (dk[c * 16] & 0);
  }
}

__global__ void __launch_bounds__(448, 1)
    attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_dy,
                       const __grid_constant__ CUtensorMap tmap_dq, const float* __restrict__ lse,
                       const float* __restrict__ delta, __nv_bfloat16* __restrict__ dqkv, int Tn, int H,
                       float scale, float scale_log2, float* __restrict__ dq_acc, int dq_direct, int kv_in_tmem, long long* dbg) {
  using namespace fabw;
  using fatc::ex2;
  using fatc::sw128;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* kv_full = bars;          // K, V landed
  uint64_t* qdo_full = bars + 1;     // [QS] Q_i, dO_i landed
  uint64_t* qdo_empty = bars + 4;    // [QS]
  uint64_t* sdp_full = bars + 7;     // S^T, dP^T in TMEM
  uint64_t* pds_full = bars + 8;     // P^T, dS^T in smem                              count 8
  uint64_t* dq_full = bars + 9;      // dV/dK/dQ MMAs of the block retired
  uint64_t* dq_free = bars + 10;     // dQ columns read back                           count 4
  uint64_t* dkv_full = bars + 11;
  uint64_t* sdp_free = bars + 12;    // S^T / dP^T read into registers (TMEM columns reusable)  count 8
  uint64_t* kv_tmem = bars + 13;     // K / V rows copied into tensor memory             count 8
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
  float* stat = reinterpret_cast<float*>(smem + OFF_STAT);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = H * HD;
  const int jb = blockIdx.y;                       // key block; block 0 has the longest loop and starts first
  const int bh = blockIdx.x, b = bh / H, h = bh % H;
  const int nq = (Tn + BQ - 1) / BQ;
  const int nit = nq - jb;                         // query blocks jb .. nq-1
  const int row_base = b * Tn;
  long long* rec3 = dbg != nullptr ? dbg + 1024 + 3 * (blockIdx.y * gridDim.x + blockIdx.x) : nullptr;
  if (rec3 != nullptr && threadIdx.x == 0) {       // per-CTA trace: start ns, end ns, SM id
    unsigned long long tns; unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tns));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    rec3[0] = static_cast<long long>(tns);
    rec3[2] = smid;
  }

  if (threadIdx.x == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tmap(&tmap_qkv);
    ptx::prefetch_tmap(&tmap_dy);
    ptx::prefetch_tmap(&tmap_dq);
  }
  if (warp == 0 && lane < 14) {                      // one barrier per lane (layout above)
    static_assert(QS == 3, "barrier layout");
    const uint32_t count = (lane == 8 || lane == 12 || lane == 13) ? 8u : (lane == 10 ? 4u : 1u);
    ptx::mbar_init(&bars[lane], count);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // barriers, TMEM and the descriptor prefetch above overlap the previous kernel's tail
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire

  if (warp == 0) {
    if (lane == 0) {
      ptx::mbar_expect_tx(kv_full, 2 * TILE);
      ptx::tma_load_2d(smem + OFF_K, &tmap_qkv, kv_full, C + h * HD, row_base + jb * BKV);
      ptx::tma_load_2d(smem + OFF_V, &tmap_qkv, kv_full, 2 * C + h * HD, row_base + jb * BKV);
      for (int it = 0; it < nit; ++it) {
        const int s = it % QS;
        ptx::mbar_wait(&qdo_empty[s], ((it / QS) & 1) ^ 1);
        ptx::mbar_expect_tx(&qdo_full[s], 2 * TILE);
        ptx::tma_load_2d(smem + OFF_Q + s * TILE, &tmap_qkv, &qdo_full[s], h * HD, row_base + (jb + it) * BQ);
        ptx::tma_load_2d(smem + OFF_DO + s * TILE, &tmap_dy, &qdo_full[s], h * HD, row_base + (jb + it) * BQ);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t ID_SDP = ptx::umma_idesc_bf16(128, 128, false, false);  // S^T, dP^T
    constexpr uint32_t ID_DKV = ptx::umma_idesc_bf16(128, 64, false, true);    // dV, dK: B MN-major
    constexpr uint32_t ID_DQ = ptx::umma_idesc_bf16(128, 64, true, true);      // dQ: A and B MN-major
    const uint32_t sb = ptx::smem_u32(smem);
    const uint64_t dK_ = ptx::umma_smem_desc_sw128(sb + OFF_K, 16, 1024);            // A, K-major
    const uint64_t dV_ = ptx::umma_smem_desc_sw128(sb + OFF_V, 16, 1024);
    const uint64_t dKmn = ptx::umma_smem_desc_sw128(sb + OFF_K, TILE, 1024);         // B, MN-major (N = d)
    const uint64_t dPT = ptx::umma_smem_desc_sw128(sb + OFF_PT, 16, 1024);           // A, K-major (K = q)
    const uint64_t dDS = ptx::umma_smem_desc_sw128(sb + OFF_DS, 16, 1024);
    const uint64_t dDSmn = ptx::umma_smem_desc_sw128(sb + OFF_DS, TILE, 1024);       // A, MN-major (M = q), LBO = next 64 q
    auto issue_sdp = [&](int it) {
      const int s = it % QS;
      const uint64_t q_k = ptx::umma_smem_desc_sw128(sb + OFF_Q + s * TILE, 16, 1024);
      const uint64_t do_k = ptx::umma_smem_desc_sw128(sb + OFF_DO + s * TILE, 16, 1024);
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (kv_in_tmem == 1) ptx::umma_bf16_ts(tmem + C_S, tmem + C_KA + 8 * k, q_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
          else ptx::umma_bf16(tmem + C_S, dK_ + static_cast<uint64_t>(2 * k), q_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (kv_in_tmem == 1) ptx::umma_bf16_ts(tmem + C_DP, tmem + C_VA + 8 * k, do_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
          else ptx::umma_bf16(tmem + C_DP, dV_ + static_cast<uint64_t>(2 * k), do_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
        }
        ptx::umma_commit(sdp_full);
      }
      __syncwarp();
    };
    ptx::mbar_wait(kv_full, 0);
    if (kv_in_tmem == 1) ptx::mbar_wait(kv_tmem, 0);
    ptx::mbar_wait(&qdo_full[0], 0);
    ptx::tc_fence_after();
    issue_sdp(0);
    const bool recm = dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0;
    for (int it = 0; it < nit; ++it) {
      const int s = it % QS;
      if (recm) dbg[512 + 4 * it] = clock64();
      // next S^T / dP^T as soon as this block's have been read out (the compute warps are still
      // writing P^T / dS^T to shared memory at that point)
      if (it + 1 < nit) {
        ptx::mbar_wait(sdp_free, it & 1);
        if (recm) dbg[512 + 4 * it + 2] = clock64();          // S/dP of this block read out
        ptx::mbar_wait(&qdo_full[(it + 1) % QS], ((it + 1) / QS) & 1);
        ptx::tc_fence_after();
        issue_sdp(it + 1);
      }
      ptx::mbar_wait(pds_full, it & 1);
      ptx::tc_fence_after();
      if (recm) dbg[512 + 4 * it + 1] = clock64();
      const uint64_t q_mn = ptx::umma_smem_desc_sw128(sb + OFF_Q + s * TILE, TILE, 1024);   // B MN-major (N = d, K = q)
      const uint64_t do_mn = ptx::umma_smem_desc_sw128(sb + OFF_DO + s * TILE, TILE, 1024);
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {  // K = 128 queries: A atom k/4, 32 B per step inside; B 16 q rows = 2048 B per step
          const uint64_t a_off = static_cast<uint64_t>(((k / 4) * TILE + (k % 4) * 32) >> 4);
          if (kv_in_tmem == 2) ptx::umma_bf16_ts(tmem + C_DV, tmem + C_KA + 8 * k, do_mn + static_cast<uint64_t>(128 * k), ID_DKV, (it > 0 || k > 0) ? 1u : 0u);
          else ptx::umma_bf16(tmem + C_DV, dPT + a_off, do_mn + static_cast<uint64_t>(128 * k), ID_DKV, (it > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t a_off = static_cast<uint64_t>(((k / 4) * TILE + (k % 4) * 32) >> 4);
          ptx::umma_bf16(tmem + C_DK, dDS + a_off, q_mn + static_cast<uint64_t>(128 * k), ID_DKV, (it > 0 || k > 0) ? 1u : 0u);
        }
      }
      __syncwarp();
      if (it > 0) { ptx::mbar_wait(dq_free, (it - 1) & 1); ptx::tc_fence_after(); }
      if (recm) dbg[512 + 4 * it + 3] = clock64();
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k)    // K = 128 keys: 16 key rows = 2048 B per step for both operands
          ptx::umma_bf16(tmem + C_DQ, dDSmn + static_cast<uint64_t>(128 * k), dKmn + static_cast<uint64_t>(128 * k), ID_DQ, k > 0);
        ptx::umma_commit(dq_full);
        ptx::umma_commit(&qdo_empty[s]);
        if (it == nit - 1) ptx::umma_commit(dkv_full);
      }
      __syncwarp();
    }
  } else if (warp >= 10) {
    // ---- dQ write-back warps: TMEM (lane = query row) -> x scale -> swizzled smem -> TMA reduce-add ----
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const int et = threadIdx.x - 320;               // 0..127
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    for (int it = 0; it < nit; ++it) {
      ptx::mbar_wait(dq_full, it & 1);
      ptx::tc_fence_after();
      float va[32], vb[32];
      ptx::tmem_ld_32x32_async(tmem + lane_addr + C_DQ, va);
      ptx::tmem_ld_32x32_async(tmem + lane_addr + C_DQ + 32, vb);
      ptx::tmem_wait_ld32(va);
      ptx::tmem_wait_ld32(vb);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(dq_free);
      if (dq_direct) {
        // straight from registers: 16-byte vector reductions into this query row of the fp32 accumulator
        // (no staging tile: the shared-memory port is the busiest unit of this kernel)
        const int qrow = (jb + it) * BQ + row;
        if (qrow < Tn) {
          float* dst = dq_acc + (static_cast<size_t>(row_base) + qrow) * C + h * HD;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * i), "f"(va[4 * i] * scale),
                         "f"(va[4 * i + 1] * scale), "f"(va[4 * i + 2] * scale), "f"(va[4 * i + 3] * scale) : "memory");
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 32 + 4 * i), "f"(vb[4 * i] * scale),
                         "f"(vb[4 * i + 1] * scale), "f"(vb[4 * i + 2] * scale), "f"(vb[4 * i + 3] * scale) : "memory");
          }
        }
        continue;
      }
      if (et == 0) ptx::bulk_wait_read<0>();        // both staging atoms drained
      ptx::named_bar_sync(1, 128);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        *reinterpret_cast<float4*>(smem + OFF_DQ + sw128(row, i)) =
            make_float4(va[4 * i] * scale, va[4 * i + 1] * scale, va[4 * i + 2] * scale, va[4 * i + 3] * scale);
        *reinterpret_cast<float4*>(smem + OFF_DQ + TILE + sw128(row, i)) =
            make_float4(vb[4 * i] * scale, vb[4 * i + 1] * scale, vb[4 * i + 2] * scale, vb[4 * i + 3] * scale);
      }
      ptx::fence_proxy_async();
      ptx::named_bar_sync(1, 128);
      if (et == 0) {
        ptx::tma_reduce_add_2d(&tmap_dq, smem + OFF_DQ, h * HD, row_base + (jb + it) * BQ);
        ptx::tma_reduce_add_2d(&tmap_dq, smem + OFF_DQ + TILE, h * HD + 32, row_base + (jb + it) * BQ);
        ptx::bulk_commit();
      }
    }
    if (et == 0 && !dq_direct) ptx::bulk_wait_all();
  } else {
    const int wg = (warp - 2) >> 2;                 // 0: query columns 0..63 of a block, 1: 64..127
    const int qd = warp & 3;
    const int row = qd * 32 + lane;                 // TMEM lane: key row (compute) / query row (dQ write-back)
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    const int key = jb * BKV + row;
    const float* lse_g = lse + (static_cast<size_t>(b) * H + h) * Tn;
    const float* del_g = delta + (static_cast<size_t>(b) * H + h) * Tn;
    const int ct = threadIdx.x - 64;                // 0..255: loader of the per-query statistics
    auto load_stat = [&](int it) -> float {
      const int qg = (jb + it) * BQ + (ct & 127);
      if (qg >= Tn) return 0.f;
      return ct < 128 ? -lse_g[qg] * 1.4426950408889634f : -del_g[qg];   // negated: see bwd_probs
    };
    float st_next = load_stat(0);
    if (kv_in_tmem == 1) {
      // this thread's key row of K (group 0) or V (group 1): swizzled smem -> registers -> tensor memory,
      // where the S^T / dP^T MMAs read it as their A operand for the whole loop
      ptx::mbar_wait(kv_full, 0);
      float rowv[32];
      const uint8_t* src = smem + (wg == 0 ? OFF_K : OFF_V);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 v4 = *reinterpret_cast<const float4*>(src + sw128(row, i));
        rowv[4 * i] = v4.x; rowv[4 * i + 1] = v4.y; rowv[4 * i + 2] = v4.z; rowv[4 * i + 3] = v4.w;
      }
      ptx::tmem_st_32x32(tmem + lane_addr + (wg == 0 ? C_KA : C_VA), rowv);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(kv_tmem);
    }
    const bool rec = dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && warp == 2 && lane == 0;
    for (int it = 0; it < nit; ++it) {
      if (rec) dbg[8 * it] = clock64();
      // publish this block's statistics (prefetched one iteration ahead), fetch the next ones
      stat[(it & 1) * 256 + ct] = st_next;
      if (it + 1 < nit) st_next = load_stat(it + 1);
      ptx::named_bar_sync(3, 256);
      const float* lse2 = stat + (it & 1) * 256 + wg * 64;
      const float* dlt = lse2 + 128;
      const bool diag = (it == 0);
      const int qlim = Tn - (jb + it) * BQ;         // query columns >= qlim are padding
      ptx::mbar_wait(sdp_full, it & 1);
      ptx::tc_fence_after();
      if (rec) dbg[8 * it + 1] = clock64();
      uint32_t pk[32], dk[32];                      // packed bf16 pairs of P^T and dS^T (64 columns)
      // masking is needed only on the diagonal block and on a ragged last query block
      long long* tr = rec ? dbg + 256 + 4 * it : nullptr;   // [tmem 0 landed, math 0 done, tmem 1 landed, math 1 done]
      if (diag || qlim < BQ) bwd_probs<true>(tmem + lane_addr + wg * 64, lse2, dlt, scale_log2, wg * 64, row, qlim, diag, pk, dk, tr);
      else                   bwd_probs<false>(tmem + lane_addr + wg * 64, lse2, dlt, scale_log2, wg * 64, row, qlim, diag, pk, dk, tr);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(sdp_free);
      if (rec) dbg[8 * it + 2] = clock64();
      // P^T / dS^T tiles are free once the previous block's dV/dK/dQ MMAs have retired
      if (it > 0) ptx::mbar_wait(dq_full, (it - 1) & 1);
      if (rec) dbg[8 * it + 3] = clock64();
      uint8_t* pt = smem + OFF_PT + wg * TILE;
      uint8_t* dst = smem + OFF_DS + wg * TILE;
      if (kv_in_tmem == 2) {
        // P^T stays in tensor memory (this thread's key row, 32 packed columns of its 64 queries) and is the
        // A operand of the dV MMA from there; only dS^T (read twice, once transposed) goes to shared memory
        ptx::tmem_st_32x32_u32(tmem + lane_addr + C_KA + wg * 32, pk);
        ptx::tc_fence_before();
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<uint4*>(dst + sw128(row, i)) = make_uint4(dk[4 * i], dk[4 * i + 1], dk[4 * i + 2], dk[4 * i + 3]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          *reinterpret_cast<uint4*>(pt + sw128(row, i)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          *reinterpret_cast<uint4*>(dst + sw128(row, i)) = make_uint4(dk[4 * i], dk[4 * i + 1], dk[4 * i + 2], dk[4 * i + 3]);
        }
      }
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(pds_full);
      if (rec) dbg[8 * it + 4] = clock64();
      if (rec) dbg[8 * it + 5] = clock64();
    }
    // dK (group 0) / dV (group 1) of this key block
    ptx::mbar_wait(dkv_full, 0);
    ptx::tc_fence_after();
    {
      const bool live = key < Tn;
      const float f = wg == 0 ? scale : 1.f;
      __nv_bfloat16* out = dqkv + (static_cast<size_t>(row_base) + (live ? key : 0)) * (3 * C) + (wg == 0 ? C : 2 * C) + h * HD;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        float v[32];
        ptx::tmem_ld_32x32(tmem + lane_addr + (wg == 0 ? C_DK : C_DV) + c * 32, v);
        if (live) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            reinterpret_cast<uint4*>(out)[c * 4 + i] =
                make_uint4(pack_bf16x2(v[8 * i] * f, v[8 * i + 1] * f), pack_bf16x2(v[8 * i + 2] * f, v[8 * i + 3] * f),
                           pack_bf16x2(v[8 * i + 4] * f, v[8 * i + 5] * f), pack_bf16x2(v[8 * i + 6] * f, v[8 * i + 7] * f));
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, TMEM_COLS);
  if (rec3 != nullptr && threadIdx.x == 0) {
    unsigned long long tns;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tns));
    rec3[1] = static_cast<long long>(tns);
  }
}

// ---- backward with sixteen probability warps -------------------------------------------------------
// Same algorithm and barrier protocol as attn_bwd_tc_kernel; the 128 x 128 block of probabilities is
// spread over 16 warps (four groups of 32 query columns) instead of 8, so that every SM sub-partition has
// four of them to overlap exp2, packed math, tensor-memory traffic and the dS^T stores, and each thread
// works on 16 columns at a time (80 registers per thread at 704 threads).  P^T always lives in tensor memory.
// Warps: 0 TMA, 1 MMA issue, 2..17 probabilities, 18..21 dQ write-back.
// `release` (or null): barrier to arrive on once the values are in registers -- the caller passes it with the LAST
// columns it reads, so that the next block's S^T / dP^T MMAs run under this half's arithmetic
template <bool MASK>
__device__ __forceinline__ void bwd_probs16(uint32_t taddr, const float* nlse2, const float* ndlt, float scale_log2,
                                            int col0, int row, int qlim, bool diag, uint32_t* pk, uint32_t* dk,
                                            uint64_t* release) {
  // MASK: query column c of this key row survives iff lo <= c < qlim, lo = row on the diagonal block, else 0:
  // one unsigned compare and a select per element (written with && / || the compiler branches per element,
  // and on the diagonal every one of those branches diverges)
  const int lo = diag ? row : 0;
  const uint32_t span = qlim > lo ? static_cast<uint32_t>(qlim - lo) : 0u;
  using fatc::ex2;
  float sv[16], dv[16];
  ptx::tmem_ld_32x16_async(taddr + fabw::C_S, sv);
  ptx::tmem_ld_32x16_async(taddr + fabw::C_DP, dv);
  ptx::tmem_wait_ld16(sv);
  ptx::tmem_wait_ld16(dv);
  if (release != nullptr) {
    ptx::tc_fence_before();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(release);
  }
  const f32x2 sc2 = pack2(scale_log2, scale_log2);
#pragma unroll
  for (int i = 0; i < 16; i += 4) {
    const float4 l4 = *reinterpret_cast<const float4*>(nlse2 + i);
    const float4 d4 = *reinterpret_cast<const float4*>(ndlt + i);
    const float nl[4] = {l4.x, l4.y, l4.z, l4.w}, nd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
    for (int e = 0; e < 4; e += 2) {
      float a0, a1;
      unpack2(fma2(pack2(sv[i + e], sv[i + e + 1]), sc2, pack2(nl[e], nl[e + 1])), a0, a1);
      float p0 = ex2(a0), p1 = ex2(a1);
      if (MASK) {
        const int col = col0 + i + e;                          // query column within the block
        p0 = (static_cast<uint32_t>(col - lo) < span) ? p0 : 0.f;
        p1 = (static_cast<uint32_t>(col + 1 - lo) < span) ? p1 : 0.f;
      }
      float d0, d1;
      unpack2(mul2(pack2(p0, p1), add2(pack2(dv[i + e], dv[i + e + 1]), pack2(nd[e], nd[e + 1]))), d0, d1);
      pk[(i + e) >> 1] = pack_bf16x2(p0, p1);
      dk[(i + e) >> 1] = pack_bf16x2(d0, d1);
    }
  }
}

__global__ void __launch_bounds__(704, 1)
    attn_bwd_tc16_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_dy,
                         const __grid_constant__ CUtensorMap tmap_dq, const float* __restrict__ lse,
                         const float* __restrict__ delta, __nv_bfloat16* __restrict__ dqkv, int Tn, int H,
                         float scale, float scale_log2) {
  using namespace fabw;
  using fatc::sw128;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* kv_full = bars;          // K, V landed
  uint64_t* qdo_full = bars + 1;     // [QS] Q_i, dO_i landed
  uint64_t* qdo_empty = bars + 4;    // [QS]
  uint64_t* sdp_full = bars + 7;     // S^T, dP^T in TMEM
  uint64_t* pds_full = bars + 8;     // P^T in TMEM, dS^T in smem                       count 16
  uint64_t* dq_full = bars + 9;      // dV/dK/dQ MMAs of the block retired
  uint64_t* dq_free = bars + 10;     // dQ columns read back                           count 4
  uint64_t* dkv_full = bars + 11;
  uint64_t* sdp_free = bars + 12;    // S^T / dP^T read into registers                  count 16
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
  float* stat = reinterpret_cast<float*>(smem + OFF_STAT);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = H * HD;
  const int jb = blockIdx.y;                       // key block; block 0 has the longest loop and starts first
  const int bh = blockIdx.x, b = bh / H, h = bh % H;
  const int nq = (Tn + BQ - 1) / BQ;
  const int nit = nq - jb;                         // query blocks jb .. nq-1
  const int row_base = b * Tn;

  if (threadIdx.x == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tmap(&tmap_qkv);
    ptx::prefetch_tmap(&tmap_dy);
    ptx::prefetch_tmap(&tmap_dq);
  }
  if (warp == 0 && lane < 13) {
    static_assert(QS == 3, "barrier layout");
    const uint32_t count = (lane == 8 || lane == 12) ? 16u : (lane == 10 ? 4u : 1u);
    ptx::mbar_init(&bars[lane], count);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();
  pdl_trigger();

  if (warp == 0) {
    if (lane == 0) {
      ptx::mbar_expect_tx(kv_full, 2 * TILE);
      ptx::tma_load_2d(smem + OFF_K, &tmap_qkv, kv_full, C + h * HD, row_base + jb * BKV);
      ptx::tma_load_2d(smem + OFF_V, &tmap_qkv, kv_full, 2 * C + h * HD, row_base + jb * BKV);
      for (int it = 0; it < nit; ++it) {
        const int s = it % QS;
        ptx::mbar_wait(&qdo_empty[s], ((it / QS) & 1) ^ 1);
        ptx::mbar_expect_tx(&qdo_full[s], 2 * TILE);
        ptx::tma_load_2d(smem + OFF_Q + s * TILE, &tmap_qkv, &qdo_full[s], h * HD, row_base + (jb + it) * BQ);
        ptx::tma_load_2d(smem + OFF_DO + s * TILE, &tmap_dy, &qdo_full[s], h * HD, row_base + (jb + it) * BQ);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t ID_SDP = ptx::umma_idesc_bf16(128, 128, false, false);  // S^T, dP^T
    constexpr uint32_t ID_DKV = ptx::umma_idesc_bf16(128, 64, false, true);    // dV, dK: B MN-major
    constexpr uint32_t ID_DQ = ptx::umma_idesc_bf16(128, 64, true, true);      // dQ: A and B MN-major
    const uint32_t sb = ptx::smem_u32(smem);
    const uint64_t dK_ = ptx::umma_smem_desc_sw128(sb + OFF_K, 16, 1024);
    const uint64_t dV_ = ptx::umma_smem_desc_sw128(sb + OFF_V, 16, 1024);
    const uint64_t dKmn = ptx::umma_smem_desc_sw128(sb + OFF_K, TILE, 1024);
    const uint64_t dDS = ptx::umma_smem_desc_sw128(sb + OFF_DS, 16, 1024);
    const uint64_t dDSmn = ptx::umma_smem_desc_sw128(sb + OFF_DS, TILE, 1024);
    auto issue_sdp = [&](int it) {
      const int s = it % QS;
      const uint64_t q_k = ptx::umma_smem_desc_sw128(sb + OFF_Q + s * TILE, 16, 1024);
      const uint64_t do_k = ptx::umma_smem_desc_sw128(sb + OFF_DO + s * TILE, 16, 1024);
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tmem + C_S, dK_ + static_cast<uint64_t>(2 * k), q_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tmem + C_DP, dV_ + static_cast<uint64_t>(2 * k), do_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
        ptx::umma_commit(sdp_full);
      }
      __syncwarp();
    };
    ptx::mbar_wait(kv_full, 0);
    ptx::mbar_wait(&qdo_full[0], 0);
    ptx::tc_fence_after();
    issue_sdp(0);
    for (int it = 0; it < nit; ++it) {
      const int s = it % QS;
      if (it + 1 < nit) {
        ptx::mbar_wait(sdp_free, it & 1);
        ptx::mbar_wait(&qdo_full[(it + 1) % QS], ((it + 1) / QS) & 1);
        ptx::tc_fence_after();
        issue_sdp(it + 1);
      }
      ptx::mbar_wait(pds_full, it & 1);
      ptx::tc_fence_after();
      const uint64_t q_mn = ptx::umma_smem_desc_sw128(sb + OFF_Q + s * TILE, TILE, 1024);
      const uint64_t do_mn = ptx::umma_smem_desc_sw128(sb + OFF_DO + s * TILE, TILE, 1024);
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k)      // dV += P^T dO, A = P^T from tensor memory (8 packed columns per 16 queries)
          ptx::umma_bf16_ts(tmem + C_DV, tmem + C_KA + 8 * k, do_mn + static_cast<uint64_t>(128 * k), ID_DKV, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t a_off = static_cast<uint64_t>(((k / 4) * TILE + (k % 4) * 32) >> 4);
          ptx::umma_bf16(tmem + C_DK, dDS + a_off, q_mn + static_cast<uint64_t>(128 * k), ID_DKV, (it > 0 || k > 0) ? 1u : 0u);
        }
      }
      __syncwarp();
      if (it > 0) { ptx::mbar_wait(dq_free, (it - 1) & 1); ptx::tc_fence_after(); }
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          ptx::umma_bf16(tmem + C_DQ, dDSmn + static_cast<uint64_t>(128 * k), dKmn + static_cast<uint64_t>(128 * k), ID_DQ, k > 0);
        ptx::umma_commit(dq_full);
        ptx::umma_commit(&qdo_empty[s]);
        if (it == nit - 1) ptx::umma_commit(dkv_full);
      }
      __syncwarp();
    }
  } else if (warp >= 18) {
    // ---- dQ write-back warps: TMEM (lane = query row) -> x scale -> swizzled smem -> TMA reduce-add ----
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const int et = threadIdx.x - 576;               // 0..127
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    for (int it = 0; it < nit; ++it) {
      ptx::mbar_wait(dq_full, it & 1);
      ptx::tc_fence_after();
      float v[32];
      ptx::tmem_ld_32x32(tmem + lane_addr + C_DQ, v);
      if (et == 0) ptx::bulk_wait_read<0>();        // both staging atoms drained
      ptx::named_bar_sync(1, 128);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<float4*>(smem + OFF_DQ + sw128(row, i)) =
            make_float4(v[4 * i] * scale, v[4 * i + 1] * scale, v[4 * i + 2] * scale, v[4 * i + 3] * scale);
      ptx::tmem_ld_32x32(tmem + lane_addr + C_DQ + 32, v);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(dq_free);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<float4*>(smem + OFF_DQ + TILE + sw128(row, i)) =
            make_float4(v[4 * i] * scale, v[4 * i + 1] * scale, v[4 * i + 2] * scale, v[4 * i + 3] * scale);
      ptx::fence_proxy_async();
      ptx::named_bar_sync(1, 128);
      if (et == 0) {
        ptx::tma_reduce_add_2d(&tmap_dq, smem + OFF_DQ, h * HD, row_base + (jb + it) * BQ);
        ptx::tma_reduce_add_2d(&tmap_dq, smem + OFF_DQ + TILE, h * HD + 32, row_base + (jb + it) * BQ);
        ptx::bulk_commit();
      }
    }
    if (et == 0) ptx::bulk_wait_read<0>();          // the staging tile has been read; the adds land before the grid completes
  } else {
    // ================================ probability warps (16) ================================
    const int wg = (warp - 2) >> 2;                 // 0..3: query columns 32 wg .. 32 wg + 31 of a block
    const int qd = warp & 3;
    const int row = qd * 32 + lane;                 // TMEM lane: key row
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    const int key = jb * BKV + row;
    const float* lse_g = lse + (static_cast<size_t>(b) * H + h) * Tn;
    const float* del_g = delta + (static_cast<size_t>(b) * H + h) * Tn;
    // Per-query statistics (lse, delta) of a block: every group of four warps loads the 2 x 32 values of ITS
    // 32 query columns (its first 64 threads, one value each) and meets on a barrier of its own, so the
    // groups do not wait for one another.  The values of the next block are fetched one iteration ahead and
    // only LOOKED AT (negated, log2 scaling: see bwd_probs) when they are stored at the top of that iteration:
    // any arithmetic right behind the load would park this in-order warp on the global-memory latency.
    const int gt = (threadIdx.x - 64) & 127;        // 0..127 within the group
    const bool loader = gt < 64, is_lse = gt < 32;
    const int scol = wg * 32 + (gt & 31);            // the query column (within the block) this thread fetches
    auto load_stat = [&](int it) -> float {
      const int qg = (jb + it) * BQ + scol;
      if (!loader || qg >= Tn) return 0.f;
      return is_lse ? lse_g[qg] : del_g[qg];
    };
    float st_next = load_stat(0);
    for (int it = 0; it < nit; ++it) {
      if (loader) stat[(it & 1) * 256 + (is_lse ? 0 : 128) + scol] = is_lse ? -st_next * 1.4426950408889634f : -st_next;
      if (it + 1 < nit) st_next = load_stat(it + 1);
      ptx::named_bar_sync(4 + wg, 128);
      const float* nlse2 = stat + (it & 1) * 256 + wg * 32;
      const float* ndlt = nlse2 + 128;
      const bool diag = (it == 0);
      const int qlim = Tn - (jb + it) * BQ;         // query columns >= qlim are padding
      ptx::mbar_wait(sdp_full, it & 1);
      ptx::tc_fence_after();
      uint32_t pk[16], dk[16];                      // packed bf16 pairs of P^T and dS^T (32 columns)
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const uint32_t ta = tmem + lane_addr + wg * 32 + hf * 16;
        if (diag || qlim < BQ)
          bwd_probs16<true>(ta, nlse2 + hf * 16, ndlt + hf * 16, scale_log2, wg * 32 + hf * 16, row, qlim, diag, pk + hf * 8, dk + hf * 8,
                            hf == 1 ? sdp_free : nullptr);
        else
          bwd_probs16<false>(ta, nlse2 + hf * 16, ndlt + hf * 16, scale_log2, wg * 32 + hf * 16, row, qlim, diag, pk + hf * 8, dk + hf * 8,
                             hf == 1 ? sdp_free : nullptr);
      }
      // the P^T columns and the dS^T tile are free once the previous block's dV/dK/dQ MMAs have retired
      if (it > 0) { ptx::mbar_wait(dq_full, (it - 1) & 1); ptx::tc_fence_after(); }
      ptx::tmem_st_32x8_u32_nowait(tmem + lane_addr + C_KA + wg * 16, pk);
      ptx::tmem_st_32x8_u32_nowait(tmem + lane_addr + C_KA + wg * 16 + 8, pk + 8);
      uint8_t* dst = smem + OFF_DS + (wg >> 1) * TILE;
      const int q0 = (wg & 1) * 4;                  // first 16-byte chunk of this group's 32 columns in the 128-byte row
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(dst + sw128(row, q0 + i)) = make_uint4(dk[4 * i], dk[4 * i + 1], dk[4 * i + 2], dk[4 * i + 3]);
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(pds_full);
    }
    // dK (groups 0, 1) / dV (groups 2, 3) of this key block, 32 of the 64 columns each
    ptx::mbar_wait(dkv_full, 0);
    ptx::tc_fence_after();
    {
      const bool live = key < Tn;
      const bool is_dk = wg < 2;
      const int c0 = (wg & 1) * 32;
      const float f = is_dk ? scale : 1.f;
      __nv_bfloat16* out = dqkv + (static_cast<size_t>(row_base) + (live ? key : 0)) * (3 * C) + (is_dk ? C : 2 * C) + h * HD + c0;
      float v[32];
      ptx::tmem_ld_32x32(tmem + lane_addr + (is_dk ? C_DK : C_DV) + c0, v);
      if (live) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          reinterpret_cast<uint4*>(out)[i] =
              make_uint4(pack_bf16x2(v[8 * i] * f, v[8 * i + 1] * f), pack_bf16x2(v[8 * i + 2] * f, v[8 * i + 3] * f),
                         pack_bf16x2(v[8 * i + 4] * f, v[8 * i + 5] * f), pack_bf16x2(v[8 * i + 6] * f, v[8 * i + 7] * f));
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, TMEM_COLS);
}

// ---- persistent form of the 16-warp backward --------------------------------------------------------
// One CTA per SM walks the items (key block, head) i = blockIdx.x, blockIdx.x + gridDim.x, ...; items are ordered
// key block 0 first (the longest loops first: at T = 1024 with 144 heads on 148 SMs every CTA gets 35 or 36 query
// blocks).  The 4 us a CTA of attn_bwd_tc16_kernel spends before its first and after its last probability block
// (7.8 CTAs per SM: 30 of 108 us) are taken off the critical path: K / V are double-buffered and loaded during the
// previous item, S^T / dP^T of the next item's first block are issued as soon as the last block's values have been
// read, and dK / dV leave through the four dQ warps while the probability warps already work on the next item.
// Roles, tensor-memory columns and the per-block protocol are those of attn_bwd_tc16_kernel; every barrier phase
// is counted over the CTA's whole walk (g: blocks, n: items).
// TRACE: clock64 stamps of CTA 0 into dbg (tools/attn_bwd_timing.py --persistent): one probability warp at
// dbg[8 g + k], the MMA warp at dbg[512 + 8 g + k], one dQ warp at dbg[1024 + 4 g + k], g = block number < 48
template <bool TRACE>
__global__ void __launch_bounds__(704, 1)
    attn_bwd_tc16p_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_dy,
                          const __grid_constant__ CUtensorMap tmap_dq, const __grid_constant__ CUtensorMap tmap_dkv,
                          const float* __restrict__ lse, const float* __restrict__ delta, int Tn, int H, int BH,
                          int nitems, float scale, float scale_log2, long long* __restrict__ dbg) {
  using namespace fabw;
  using fatc::sw128;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* kv_full = bars;          // [2] K, V of an item landed
  uint64_t* kv_empty = bars + 2;     // [2] every MMA that reads them has retired
  uint64_t* qdo_full = bars + 4;     // [QS] Q_i, dO_i landed
  uint64_t* qdo_empty = bars + 7;    // [QS]
  uint64_t* sdp_full = bars + 10;    // S^T, dP^T in TMEM
  uint64_t* sdp_free = bars + 11;    // S^T / dP^T read into registers                  count 16
  uint64_t* pds_full = bars + 12;    // P^T in TMEM, dS^T in smem                       count 16
  uint64_t* dq_full = bars + 13;     // dV/dK/dQ MMAs of the block retired
  uint64_t* dq_free = bars + 14;     // dQ columns read back                           count 4
  uint64_t* dkv_free = bars + 16;    // dK / dV read back                               count 4
  uint64_t* stat_full = bars + 17;   // [2] per-query statistics of a block in shared memory   count 4
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);
  float* stat = reinterpret_cast<float*>(smem + OFF_STAT);
  static_assert(QS == 3, "barrier layout");
  static_assert(OFF_PT + 2 * TILE <= OFF_DS, "second K / V buffer");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = H * HD;
  const int nq = (Tn + BQ - 1) / BQ;
  const int stride = static_cast<int>(gridDim.x);
  auto stamp = [&](int base, uint32_t g, int k) {
    if (TRACE && blockIdx.x == 0 && lane == 0 && g < 48) dbg[base + (base == 1024 ? 4 : 8) * g + k] = clock64();
  };

  if (threadIdx.x == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tmap(&tmap_qkv);
    ptx::prefetch_tmap(&tmap_dy);
    ptx::prefetch_tmap(&tmap_dq);
    ptx::prefetch_tmap(&tmap_dkv);
  }
  if (warp == 0 && lane < 19) {
    const uint32_t count = (lane == 11 || lane == 12) ? 16u : ((lane == 14 || lane >= 16) ? 4u : 1u);
    ptx::mbar_init(&bars[lane], count);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();
  pdl_trigger();

  if (warp == 0) {
    if (lane == 0) {
      int s = 0, n = 0;                             // Q / dO ring slot; items so far
      uint32_t sph = 0;                             // parity of the slot's use count
      for (int item = blockIdx.x; item < nitems; item += stride, ++n) {
        const int jb = item / BH, bh = item - jb * BH, b = bh / H, h = bh - b * H;
        const int row_base = b * Tn, kb = n & 1;
        uint8_t* kv = smem + (kb ? OFF_PT : OFF_K);
        ptx::mbar_wait(&kv_empty[kb], ((n >> 1) & 1) ^ 1);
        ptx::mbar_expect_tx(&kv_full[kb], 2 * TILE);
        ptx::tma_load_2d(kv, &tmap_qkv, &kv_full[kb], C + h * HD, row_base + jb * BKV);
        ptx::tma_load_2d(kv + TILE, &tmap_qkv, &kv_full[kb], 2 * C + h * HD, row_base + jb * BKV);
        for (int qb = jb; qb < nq; ++qb) {
          ptx::mbar_wait(&qdo_empty[s], sph ^ 1u);
          ptx::mbar_expect_tx(&qdo_full[s], 2 * TILE);
          ptx::tma_load_2d(smem + OFF_Q + s * TILE, &tmap_qkv, &qdo_full[s], h * HD, row_base + qb * BQ);
          ptx::tma_load_2d(smem + OFF_DO + s * TILE, &tmap_dy, &qdo_full[s], h * HD, row_base + qb * BQ);
          if (++s == QS) { s = 0; sph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t ID_SDP = ptx::umma_idesc_bf16(128, 128, false, false);  // S^T, dP^T
    constexpr uint32_t ID_DKV = ptx::umma_idesc_bf16(128, 64, false, true);    // dV, dK: B MN-major
    constexpr uint32_t ID_DQ = ptx::umma_idesc_bf16(128, 64, true, true);      // dQ: A and B MN-major
    const uint32_t sb = ptx::smem_u32(smem);
    const uint64_t dDS = ptx::umma_smem_desc_sw128(sb + OFF_DS, 16, 1024);
    const uint64_t dDSmn = ptx::umma_smem_desc_sw128(sb + OFF_DS, TILE, 1024);
    auto issue_sdp = [&](int s, int kb) {
      const uint32_t kvb = sb + (kb ? OFF_PT : OFF_K);
      const uint64_t dK_ = ptx::umma_smem_desc_sw128(kvb, 16, 1024);
      const uint64_t dV_ = ptx::umma_smem_desc_sw128(kvb + TILE, 16, 1024);
      const uint64_t q_k = ptx::umma_smem_desc_sw128(sb + OFF_Q + s * TILE, 16, 1024);
      const uint64_t do_k = ptx::umma_smem_desc_sw128(sb + OFF_DO + s * TILE, 16, 1024);
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tmem + C_S, dK_ + static_cast<uint64_t>(2 * k), q_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tmem + C_DP, dV_ + static_cast<uint64_t>(2 * k), do_k + static_cast<uint64_t>(2 * k), ID_SDP, k > 0);
        ptx::umma_commit(sdp_full);
      }
      __syncwarp();
    };
    int item = blockIdx.x;
    if (item < nitems) {
      int nit = nq - item / BH, it = 0, n = 0, s = 0;
      uint32_t sph = 0, g = 0;
      ptx::mbar_wait(&kv_full[0], 0);
      ptx::mbar_wait(&qdo_full[0], 0);
      ptx::tc_fence_after();
      issue_sdp(0, 0);
      while (true) {
        // the block after this one: the next query block of the item, or the first of the CTA's next item
        int n_item = item, n_it = it + 1, n_nit = nit, n_n = n;
        if (n_it == nit) {
          n_item = item + stride; n_it = 0; n_n = n + 1;
          if (n_item < nitems) n_nit = nq - n_item / BH;
        }
        const bool has_next = n_item < nitems;
        int ns = s + 1;
        uint32_t nsph = sph;
        if (ns == QS) { ns = 0; nsph ^= 1u; }
        stamp(512, g, 0);
        if (has_next) {
          ptx::mbar_wait(sdp_free, g & 1u);
          stamp(512, g, 1);
          ptx::mbar_wait(&qdo_full[ns], nsph);
          if (n_it == 0) ptx::mbar_wait(&kv_full[n_n & 1], (n_n >> 1) & 1);
          ptx::tc_fence_after();
          stamp(512, g, 2);
          issue_sdp(ns, n_n & 1);
        }
        stamp(512, g, 3);
        ptx::mbar_wait(pds_full, g & 1u);
        stamp(512, g, 4);
        if (it == 0 && n > 0) ptx::mbar_wait(dkv_free, (n - 1) & 1);   // the previous item's dK / dV have been read
        ptx::tc_fence_after();
        stamp(512, g, 5);
        const uint64_t q_mn = ptx::umma_smem_desc_sw128(sb + OFF_Q + s * TILE, TILE, 1024);
        const uint64_t do_mn = ptx::umma_smem_desc_sw128(sb + OFF_DO + s * TILE, TILE, 1024);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)      // dV += P^T dO, A = P^T from tensor memory (8 packed columns per 16 queries)
            ptx::umma_bf16_ts(tmem + C_DV, tmem + C_KA + 8 * k, do_mn + static_cast<uint64_t>(128 * k), ID_DKV, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint64_t a_off = static_cast<uint64_t>(((k / 4) * TILE + (k % 4) * 32) >> 4);
            ptx::umma_bf16(tmem + C_DK, dDS + a_off, q_mn + static_cast<uint64_t>(128 * k), ID_DKV, (it > 0 || k > 0) ? 1u : 0u);
          }
        }
        __syncwarp();
        stamp(512, g, 6);
        if (g > 0) { ptx::mbar_wait(dq_free, (g - 1) & 1u); ptx::tc_fence_after(); }
        stamp(512, g, 7);
        const uint64_t dKmn = ptx::umma_smem_desc_sw128(sb + ((n & 1) ? OFF_PT : OFF_K), TILE, 1024);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            ptx::umma_bf16(tmem + C_DQ, dDSmn + static_cast<uint64_t>(128 * k), dKmn + static_cast<uint64_t>(128 * k), ID_DQ, k > 0);
          ptx::umma_commit(dq_full);
          ptx::umma_commit(&qdo_empty[s]);
          if (it == nit - 1) ptx::umma_commit(&kv_empty[n & 1]);
        }
        __syncwarp();
        if (!has_next) break;
        item = n_item; it = n_it; nit = n_nit; n = n_n; s = ns; sph = nsph; ++g;
      }
    }
  } else if (warp >= 18) {
    // ---- dQ write-back warps: TMEM (lane = query row) -> x scale -> swizzled smem -> TMA reduce-add; after an
    // item's last block the same warps take dK / dV (lane = key row) out of tensor memory ----
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const int et = threadIdx.x - 576;               // 0..127
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    // These warps also feed the probability warps their per-query statistics: thread et fetches lse and delta of
    // query et of the block TWO ahead in the CTA's walk (across items) and stores them negated, lse in log2
    // units (see bwd_probs16), into the buffer that block's parity names.  That buffer was last read for block
    // g, whose arithmetic is over once dq_full(g) has completed.
    int la_item = blockIdx.x, la_qb = la_item < nitems ? la_item / BH : 0;
    auto la_load = [&](float& l, float& d) {
      l = 0.f; d = 0.f;
      const int q = la_qb * BQ + et;
      if (la_item < nitems && q < Tn) {
        const size_t o = static_cast<size_t>(la_item % BH) * Tn + q;
        l = lse[o]; d = delta[o];
      }
    };
    auto la_store = [&](uint32_t blk, float l, float d) {
      stat[(blk & 1u) * 256 + et] = -l * 1.4426950408889634f;
      stat[(blk & 1u) * 256 + 128 + et] = -d;
      if (++la_qb == nq) {                           // on to the next block of the walk
        la_item += stride;
        if (la_item < nitems) la_qb = la_item / BH;
      }
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&stat_full[blk & 1u]);
    };
    {
      float l, d;
      la_load(l, d); la_store(0, l, d);
      la_load(l, d); la_store(1, l, d);
    }
    uint32_t g = 0;
    int n = 0;
    for (int item = blockIdx.x; item < nitems; item += stride, ++n) {
      const int jb = item / BH, bh = item - jb * BH, b = bh / H, h = bh - b * H;
      const int row_base = b * Tn;
      for (int qb = jb; qb < nq; ++qb, ++g) {
        if (warp == 18) stamp(1024, g, 0);
        float st_l, st_d;
        la_load(st_l, st_d);                         // in flight under the wait
        ptx::mbar_wait(dq_full, g & 1u);
        ptx::tc_fence_after();
        if (warp == 18) stamp(1024, g, 1);
        if (qb == nq - 1) {
          // The item's last block: its commit also covers the last dV / dK MMAs.  dK (x scale) and dV go first --
          // the next item's first dV / dK MMAs wait for these accumulators (dkv_free), nothing waits for this
          // block's dQ: tensor memory (lane = key row) -> bf16 -> the two staging tiles (128-byte rows,
          // swizzled) -> one TMA store each, clipped at the end of the sequence.  (Written from registers, 64
          // bytes per thread at a row stride of 3 C, these stores held the SM's load/store unit for several
          // thousand clocks, and the probability warps' shared-memory traffic with it.)
          if (et == 0) ptx::bulk_wait_read<0>();        // the previous dQ tile has left the staging buffers
          ptx::named_bar_sync(1, 128);
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {                // dK columns 0-31, 32-63, then dV
            float v[32];
            ptx::tmem_ld_32x32(tmem + lane_addr + (c < 2 ? C_DK : C_DV) + (c & 1) * 32, v);
            if (c == 3) {                              // the accumulators may be overwritten by the next item
              ptx::tc_fence_before();
              __syncwarp();
              if (lane == 0) ptx::mbar_arrive(dkv_free);
            }
            const float f = c < 2 ? scale : 1.f;
            uint8_t* dst = smem + OFF_DQ + (c >> 1) * TILE;
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(dst + sw128(row, (c & 1) * 4 + i)) =
                  make_uint4(pack_bf16x2(v[8 * i] * f, v[8 * i + 1] * f), pack_bf16x2(v[8 * i + 2] * f, v[8 * i + 3] * f),
                             pack_bf16x2(v[8 * i + 4] * f, v[8 * i + 5] * f), pack_bf16x2(v[8 * i + 6] * f, v[8 * i + 7] * f));
          }
          ptx::fence_proxy_async();
          ptx::named_bar_sync(1, 128);
          if (et == 0) {
            ptx::tma_store_3d(&tmap_dkv, smem + OFF_DQ, C + h * HD, jb * BKV, b);
            ptx::tma_store_3d(&tmap_dkv, smem + OFF_DQ + TILE, 2 * C + h * HD, jb * BKV, b);
            ptx::bulk_commit();
          }
        }
        float v[32];
        ptx::tmem_ld_32x32(tmem + lane_addr + C_DQ, v);
        if (et == 0) ptx::bulk_wait_read<0>();        // both staging atoms drained
        ptx::named_bar_sync(1, 128);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(smem + OFF_DQ + sw128(row, i)) =
              make_float4(v[4 * i] * scale, v[4 * i + 1] * scale, v[4 * i + 2] * scale, v[4 * i + 3] * scale);
        ptx::tmem_ld_32x32(tmem + lane_addr + C_DQ + 32, v);
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(dq_free);
        la_store(g + 2, st_l, st_d);                 // after the release of the dQ columns, which the MMA warp waits for
        if (warp == 18) stamp(1024, g, 2);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(smem + OFF_DQ + TILE + sw128(row, i)) =
              make_float4(v[4 * i] * scale, v[4 * i + 1] * scale, v[4 * i + 2] * scale, v[4 * i + 3] * scale);
        ptx::fence_proxy_async();
        ptx::named_bar_sync(1, 128);
        if (et == 0) {
          ptx::tma_reduce_add_2d(&tmap_dq, smem + OFF_DQ, h * HD, row_base + qb * BQ);
          ptx::tma_reduce_add_2d(&tmap_dq, smem + OFF_DQ + TILE, h * HD + 32, row_base + qb * BQ);
          ptx::bulk_commit();
        }
        if (warp == 18) stamp(1024, g, 3);
      }
    }
    if (et == 0) ptx::bulk_wait_read<0>();          // the staging tile has been read; the adds land before the grid completes
  } else {
    // ================================ probability warps (16) ================================
    const int wg = (warp - 2) >> 2;                 // 0..3: query columns 32 wg .. 32 wg + 31 of a block
    const int qd = warp & 3;
    const int row = qd * 32 + lane;                 // TMEM lane: key row
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    // the per-query statistics of a block are put into shared memory by the dQ warps (stat_full)
    const int tbase = warp == 2 ? 0 : (warp == 4 ? 1536 : -1);   // traced warps: 2 (above the diagonal), 4 (on it)
    uint32_t g = 0;
    for (int item = blockIdx.x; item < nitems; item += stride) {
      const int jb = item / BH;
      for (int qb = jb; qb < nq; ++qb, ++g) {
        const uint32_t gp = g & 1u;
        if (tbase >= 0) stamp(tbase, g, 0);
        ptx::mbar_wait(&stat_full[gp], (g >> 1) & 1u);
        const float* nlse2 = stat + gp * 256 + wg * 32;
        const float* ndlt = nlse2 + 128;
        const bool diag = (qb == jb);
        const int qlim = Tn - qb * BQ;                // query columns >= qlim are padding
        if (tbase >= 0) stamp(tbase, g, 1);
        ptx::mbar_wait(sdp_full, gp);
        ptx::tc_fence_after();
        if (tbase >= 0) stamp(tbase, g, 2);
        uint32_t pk[16], dk[16];                      // packed bf16 pairs of P^T and dS^T (32 columns)
        // (Skipping the warps whose 32 x 32 piece of a diagonal block is entirely masked, and taking the unmasked
        // form for those entirely below the diagonal, was measured slower: 103.7 against 102.6 us.)
        const int tri = (diag || qlim < BQ) ? 1 : 2;
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          const uint32_t ta = tmem + lane_addr + wg * 32 + hf * 16;
          if (tri == 1)
            bwd_probs16<true>(ta, nlse2 + hf * 16, ndlt + hf * 16, scale_log2, wg * 32 + hf * 16, row, qlim, diag, pk + hf * 8, dk + hf * 8,
                              hf == 1 ? sdp_free : nullptr);
          else if (tri == 2)
            bwd_probs16<false>(ta, nlse2 + hf * 16, ndlt + hf * 16, scale_log2, wg * 32 + hf * 16, row, qlim, diag, pk + hf * 8, dk + hf * 8,
                               hf == 1 ? sdp_free : nullptr);
          if (TRACE && tbase >= 0 && blockIdx.x == 0 && lane == 0 && g < 48) dbg[tbase + 8 * g + 3 + hf] = clock64() + (pk[hf * 8] & 0) + (dk[hf * 8 + 7] & 0);
        }
        // the P^T columns and the dS^T tile are free once the previous block's dV/dK/dQ MMAs have retired
        if (g > 0) { ptx::mbar_wait(dq_full, (g - 1) & 1u); ptx::tc_fence_after(); }
        if (tbase >= 0) stamp(tbase, g, 5);
        ptx::tmem_st_32x8_u32_nowait(tmem + lane_addr + C_KA + wg * 16, pk);
        ptx::tmem_st_32x8_u32_nowait(tmem + lane_addr + C_KA + wg * 16 + 8, pk + 8);
        uint8_t* dst = smem + OFF_DS + (wg >> 1) * TILE;
        const int q0 = (wg & 1) * 4;                  // first 16-byte chunk of this group's 32 columns in the 128-byte row
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<uint4*>(dst + sw128(row, q0 + i)) = make_uint4(dk[4 * i], dk[4 * i + 1], dk[4 * i + 2], dk[4 * i + 3]);
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(pds_full);
        if (tbase >= 0) stamp(tbase, g, 6);
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, TMEM_COLS);
}

// dq (bf16, inside dqkv) = dq_acc (fp32); dq_acc is zeroed again for the next call
__global__ void __launch_bounds__(256)
    attn_dq_convert_kernel(float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, size_t M, int C) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t i = (static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x) * 8;
  if (i >= M * C) return;
  const size_t r = i / C, c = i % C;
  float4* src = reinterpret_cast<float4*>(dq_acc + i);
  const float4 a = src[0], bq = src[1];
  *reinterpret_cast<uint4*>(dqkv + r * 3 * C + c) =
      make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(bq.x, bq.y), pack_bf16x2(bq.z, bq.w));
  src[0] = make_float4(0.f, 0.f, 0.f, 0.f);
  src[1] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// delta[b,h,q] = sum_d dy*y : 8 threads per (row, head), 16-byte loads
__global__ void __launch_bounds__(256)
    attn_delta_tc_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy,
                         float* __restrict__ delta, int B, int Tn, int H) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t g = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
  const size_t item = g >> 3;                       // (b*T + q) * H + h
  const int part = static_cast<int>(g & 7);
  const size_t total = static_cast<size_t>(B) * Tn * H;
  float s = 0.f;
  if (item < total) {
    const uint4 a = *reinterpret_cast<const uint4*>(y + item * 64 + part * 8);
    const uint4 d = *reinterpret_cast<const uint4*>(dy + item * 64 + part * 8);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, dw[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 p = unpack_bf16x2(aw[k]), q = unpack_bf16x2(dw[k]);
      s += p.x * q.x + p.y * q.y;
    }
  }
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  if (item < total && part == 0) {
    const size_t bt = item / H, h = item % H;
    const size_t b = bt / Tn, q = bt % Tn;
    delta[(b * H + h) * Tn + q] = s;
  }
}

int attention_bwd_tc_launch(const void* qkv, const void* y, const void* dy, const float* lse, void* dqkv,
                            float* delta, float* dq_acc, int B, int T, int H, cudaStream_t stream) {
  using namespace fabw;
  const int C = H * HD;
  const uint64_t rows = static_cast<uint64_t>(B) * T;
  CUtensorMap tq, tdy, tdq;
  int rc = make_tmap_2d(&tq, qkv, 2, static_cast<uint64_t>(3) * C, rows, static_cast<uint64_t>(3) * C, HD, 128);
  if (rc) return rc;
  rc = make_tmap_2d(&tdy, dy, 2, C, rows, C, HD, 128);
  if (rc) return rc;
  rc = make_tmap_2d(&tdq, dq_acc, 4, C, rows, C, 32, 128);
  if (rc) return rc;
  static PerDeviceOnce attr_done;
  if (attr_done.first_use()) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    NSB_REQUIRE(e == cudaSuccess, "attention_bwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  }
  const size_t items = rows * H * 8;
  launch_pdl(attn_delta_tc_kernel, dim3(static_cast<unsigned>((items + 255) / 256)), dim3(256), 0, stream, 
      static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), delta, B, T, H);
  const dim3 grid(B * H, (T + BKV - 1) / BKV);
  const float scale = 1.0f / sqrtf(static_cast<float>(HD));
  if (g_attn_bwd_warps16 && (g_attn_dbg == nullptr || g_attn_bwd_persistent)) {
    static PerDeviceOnce attr16;
    if (attr16.first_use()) {
      cudaError_t e = cudaFuncSetAttribute(attn_bwd_tc16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
      NSB_REQUIRE(e == cudaSuccess, "attention_bwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    if (g_attn_bwd_persistent) {
      static PerDeviceOnce attr16p;
      if (attr16p.first_use()) {
        cudaError_t e = cudaFuncSetAttribute(attn_bwd_tc16p_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(attn_bwd_tc16p_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        NSB_REQUIRE(e == cudaSuccess, "attention_bwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      }
      const int nitems = static_cast<int>(grid.x * grid.y);
      const int ctas = nitems < sm_count() ? nitems : sm_count();
      CUtensorMap tdkv;
      rc = make_tmap_3d_bf16(&tdkv, dqkv, static_cast<uint64_t>(3) * C, T, B, static_cast<uint64_t>(3) * C,
                             static_cast<uint64_t>(T) * 3 * C, HD, BKV);
      if (rc) return rc;
      if (g_attn_dbg != nullptr)
        launch_pdl(attn_bwd_tc16p_kernel<true>, dim3(ctas), dim3(704), SMEM_BYTES, stream, tq, tdy, tdq, tdkv, lse, delta,
                   T, H, B * H, nitems, scale, scale * 1.4426950408889634f, g_attn_dbg);
      else
        launch_pdl(attn_bwd_tc16p_kernel<false>, dim3(ctas), dim3(704), SMEM_BYTES, stream, tq, tdy, tdq, tdkv, lse, delta,
                   T, H, B * H, nitems, scale, scale * 1.4426950408889634f, static_cast<long long*>(nullptr));
    } else {
      launch_pdl(attn_bwd_tc16_kernel, dim3(grid), dim3(704), SMEM_BYTES, stream, tq, tdy, tdq, lse, delta,
                 static_cast<__nv_bfloat16*>(dqkv), T, H, scale, scale * 1.4426950408889634f);
    }
  } else {
    launch_pdl(attn_bwd_tc_kernel, dim3(grid), dim3(448), SMEM_BYTES, stream, tq, tdy, tdq, lse, delta, static_cast<__nv_bfloat16*>(dqkv), T, H,
               scale, scale * 1.4426950408889634f, dq_acc, g_attn_dq_direct, g_attn_kv_tmem, g_attn_dbg);
  }
  const size_t n8 = rows * C / 8;
  launch_pdl(attn_dq_convert_kernel, dim3(static_cast<unsigned>((n8 + 255) / 256)), dim3(256), 0, stream, dq_acc, static_cast<__nv_bfloat16*>(dqkv), rows, C);
  count_launch(3);
  return check_launch("attention_bwd_tc");
}

}  // namespace nsb
