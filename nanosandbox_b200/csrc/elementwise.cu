// Memory-bound kernels of the training step: embedding, LayerNorm, softmax cross-entropy,
// grad-norm/clip, fused AdamW, casts.  All are HBM-bound: 128-bit accesses, warp-shuffle
// reductions, no shared-memory staging where there is no reuse.  See DESIGN.md for the
// algorithmic bytes per element of each kernel.
#include <stdarg.h>

#include <atomic>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace nsb {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
static std::atomic<int> g_pdl{1};
bool pdl_enabled() { return g_pdl.load() != 0; }
void set_pdl_enabled(bool on) { g_pdl.store(on ? 1 : 0); }
void count_launch(int n) { g_launches.fetch_add(static_cast<unsigned long long>(n)); }
int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return static_cast<int>(e);
  }
  return 0;
}

// =============================================================================================
// K1 embedding
// =============================================================================================
__global__ void __launch_bounds__(256) embedding_fwd_kernel(const int64_t* __restrict__ idx,
                                                            const float4* __restrict__ wte,
                                                            const float4* __restrict__ wpe,
                                                            float4* __restrict__ x, int M, int T,
                                                            int C4) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(M) * C4) return;
  const int row = static_cast<int>(i / C4), c = static_cast<int>(i % C4);
  const int64_t tok = idx[row];
  const float4 a = __ldg(&wte[static_cast<size_t>(tok) * C4 + c]);
  const float4 p = __ldg(&wpe[static_cast<size_t>(row % T) * C4 + c]);
  x[i] = make_float4(a.x + p.x, a.y + p.y, a.z + p.z, a.w + p.w);
}

__global__ void __launch_bounds__(256) embedding_bwd_wte_kernel(const int64_t* __restrict__ idx,
                                                                const float4* __restrict__ dx,
                                                                float4* __restrict__ dwte, int M,
                                                                int C4) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(M) * C4) return;
  const int row = static_cast<int>(i / C4), c = static_cast<int>(i % C4);
  const int64_t tok = idx[row];
  atomicAdd(&dwte[static_cast<size_t>(tok) * C4 + c], dx[i]);  // sm_90+: 128-bit fp32 red
}

__global__ void __launch_bounds__(256) embedding_bwd_wpe_kernel(const float4* __restrict__ dx,
                                                                float4* __restrict__ dwpe, int B,
                                                                int T, int C4, int accumulate) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(T) * C4) return;
  float4 s = accumulate ? dwpe[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  const size_t stride = static_cast<size_t>(T) * C4;
  for (int b = 0; b < B; ++b) {  // fixed order over the batch: deterministic
    const float4 v = dx[i + b * stride];
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  dwpe[i] = s;
}

// =============================================================================================
// K2 LayerNorm.  One warp per row; the row lives in registers (NV float4 per lane, C <= 128*NV)
// =============================================================================================
template <typename T>
struct Vec4IO;
template <>
struct Vec4IO<float> {
  static __device__ __forceinline__ float4 load(const float* p) {
    return *reinterpret_cast<const float4*>(p);
  }
  static __device__ __forceinline__ void store(float* p, float4 v) {
    *reinterpret_cast<float4*>(p) = v;
  }
};
template <>
struct Vec4IO<__nv_bfloat16> {
  static __device__ __forceinline__ float4 load(const __nv_bfloat16* p) {
    const uint2 u = *reinterpret_cast<const uint2*>(p);
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y);
    return make_float4(a.x, a.y, b.x, b.y);
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float4 v) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
};

template <int NV, typename OutT>
__global__ void __launch_bounds__(256)
    layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                         const float* __restrict__ b, OutT* __restrict__ y,
                         float* __restrict__ mean, float* __restrict__ rstd, int M, int C,
                         float eps) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= M) return;
  const float* xr = x + static_cast<size_t>(row) * C;
  float4 v[NV], wv[NV];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = (lane + 32 * j) * 4;
    v[j] = (c < C) ? Vec4IO<float>::load(xr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // the weights are fetched now, behind the row's own loads, not where they are used: there every one of them
  // cost an L1 / L2 round trip with nothing else to issue (30 % of this kernel's sampled stalls)
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = (lane + 32 * j) * 4;
    wv[j] = (c < C) ? __ldg(reinterpret_cast<const float4*>(w + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  const float mu = warp_sum(s) / static_cast<float>(C);
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = (lane + 32 * j) * 4;
    if (c < C) {
      const float d0 = v[j].x - mu, d1 = v[j].y - mu, d2 = v[j].z - mu, d3 = v[j].w - mu;
      q += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
    }
  }
  const float rs = 1.0f / sqrtf(warp_sum(q) / static_cast<float>(C) + eps);
  if (lane == 0) {
    mean[row] = mu;
    rstd[row] = rs;
  }
  OutT* yr = y + static_cast<size_t>(row) * C;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = (lane + 32 * j) * 4;
    if (c < C) {
      float4 o = make_float4((v[j].x - mu) * rs * wv[j].x, (v[j].y - mu) * rs * wv[j].y,
                             (v[j].z - mu) * rs * wv[j].z, (v[j].w - mu) * rs * wv[j].w);
      if (b != nullptr) {
        const float4 bv = __ldg(reinterpret_cast<const float4*>(b + c));
        o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
      }
      Vec4IO<OutT>::store(yr + c, o);
    }
  }
}

// Stage 1 of the backward: dx (+ residual gradient) per row, per-CTA column partials of dw/db.
// part layout: [gridDim.x][2][C].  All three row inputs (dy, x, incoming residual gradient) are
// loaded before anything is reduced, so a row costs one memory latency; xhat and dy*w are
// recomputed in the second pass (w comes from L1) and the db accumulators exist only when a bias
// gradient is wanted, which keeps two CTAs of 8 warps resident per SM.
template <int NV, typename DyT, bool HAS_DB>
__global__ void __launch_bounds__(256, 2)
    layernorm_bwd_kernel(const DyT* __restrict__ dy, const float* __restrict__ x,
                         const float* __restrict__ w, const float* __restrict__ mean,
                         const float* __restrict__ rstd, const float* __restrict__ dres_in,
                         float* __restrict__ dx, __nv_bfloat16* __restrict__ dx_lowp,
                         float* __restrict__ part, int M, int C) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  __shared__ float red[8][128 * NV];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float4 dwacc[NV], dbacc[HAS_DB ? NV : 1];
#pragma unroll
  for (int j = 0; j < NV; ++j) dwacc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int j = 0; j < (HAS_DB ? NV : 1); ++j) dbacc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  const float invC = 1.0f / static_cast<float>(C);

  for (int row = blockIdx.x * 8 + warp; row < M; row += gridDim.x * 8) {
    const size_t off = static_cast<size_t>(row) * C;
    const float mu = mean[row], rs = rstd[row];
    float4 dyv[NV], xv[NV], rv[NV];   // every global read of the row is issued up front: one memory latency per row
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (lane + 32 * j) * 4;
      if (c < C) {
        dyv[j] = Vec4IO<DyT>::load(dy + off + c);
        xv[j] = Vec4IO<float>::load(x + off + c);
        rv[j] = dres_in != nullptr ? Vec4IO<float>::load(dres_in + off + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (lane + 32 * j) * 4;
      if (c < C) {
        const float4 wv = __ldg(reinterpret_cast<const float4*>(w + c));
        const float4 d = dyv[j];
        const float4 xh = make_float4((xv[j].x - mu) * rs, (xv[j].y - mu) * rs, (xv[j].z - mu) * rs, (xv[j].w - mu) * rs);
        const float4 g = make_float4(d.x * wv.x, d.y * wv.y, d.z * wv.z, d.w * wv.w);
        s1 += (g.x + g.y) + (g.z + g.w);
        s2 += (g.x * xh.x + g.y * xh.y) + (g.z * xh.z + g.w * xh.w);
        dwacc[j].x += d.x * xh.x; dwacc[j].y += d.y * xh.y;
        dwacc[j].z += d.z * xh.z; dwacc[j].w += d.w * xh.w;
        if (HAS_DB) { dbacc[j].x += d.x; dbacc[j].y += d.y; dbacc[j].z += d.z; dbacc[j].w += d.w; }
      }
    }
    const float c1 = warp_sum(s1) * invC, c2 = warp_sum(s2) * invC;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (lane + 32 * j) * 4;
      if (c < C) {
        const float4 wv = __ldg(reinterpret_cast<const float4*>(w + c));
        const float4 d = dyv[j];
        const float4 xh = make_float4((xv[j].x - mu) * rs, (xv[j].y - mu) * rs, (xv[j].z - mu) * rs, (xv[j].w - mu) * rs);
        float4 o = make_float4(rs * (d.x * wv.x - c1 - xh.x * c2), rs * (d.y * wv.y - c1 - xh.y * c2),
                               rs * (d.z * wv.z - c1 - xh.z * c2), rs * (d.w * wv.w - c1 - xh.w * c2));
        o.x += rv[j].x; o.y += rv[j].y; o.z += rv[j].z; o.w += rv[j].w;
        Vec4IO<float>::store(dx + off + c, o);
        if (dx_lowp != nullptr) Vec4IO<__nv_bfloat16>::store(dx_lowp + off + c, o);
      }
    }
  }
  // cross-warp reduction of the column partials, dw then db
#pragma unroll
  for (int pass = 0; pass < (HAS_DB ? 2 : 1); ++pass) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (lane + 32 * j) * 4;
      const float4 a = (pass == 0 || !HAS_DB) ? dwacc[j] : dbacc[HAS_DB ? j : 0];
      *reinterpret_cast<float4*>(&red[warp][c]) = a;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += 256) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += red[k][c];
      part[(static_cast<size_t>(blockIdx.x) * 2 + pass) * C + c] = s;
    }
  }
}

// Streaming variant of stage 1 (the one the launcher prefers): a producer warp keeps a ring of
// 8-row slabs (dy, x, incoming residual gradient) in flight with bulk async copies
// (cp.async.bulk + mbarrier), 8 consumer warps take one row each from shared memory.  Memory latency is
// hidden by the ring instead of by occupancy; each consumer only holds its dw/db partials in registers.
constexpr int kLnRows = 8;
constexpr int kLnGroups = 2;                         // consumer groups of kLnRows warps, taking alternate slabs
constexpr int kLnWarps = kLnRows * kLnGroups;        // consumer warps (the producer is warp kLnWarps)
template <int NV, typename DyT, bool HAS_DB>
__global__ void __launch_bounds__(32 * (kLnWarps + 1), 1)
    layernorm_bwd_stream_kernel(const DyT* __restrict__ dy, const float* __restrict__ x,
                                const float* __restrict__ w, const float* __restrict__ mean,
                                const float* __restrict__ rstd, const float* __restrict__ dres_in,
                                float* __restrict__ dx, __nv_bfloat16* __restrict__ dx_lowp,
                                float* __restrict__ part, int M, int C, int nstages) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  extern __shared__ __align__(128) uint8_t ln_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t dy_row = static_cast<uint32_t>(C) * sizeof(DyT), f_row = static_cast<uint32_t>(C) * 4u;
  const uint32_t off_x = kLnRows * dy_row, off_r = off_x + kLnRows * f_row;
  const uint32_t stage_bytes = off_r + (dres_in != nullptr ? kLnRows * f_row : 0u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(ln_smem + static_cast<size_t>(nstages) * stage_bytes);
  // one `full` barrier per (slot, consumer group): with a ring depth that is not a multiple of kLnGroups a slot
  // falls to the groups in turn, and a barrier that only ever serves one group cannot run a phase ahead of
  // (or behind) the parity that group waits for -- its next phase needs that group's own release of the slot
  uint64_t* full = bars;
  uint64_t* empty = bars + 4 * kLnGroups;
  if (threadIdx.x == 0) {
    for (int s = 0; s < nstages; ++s) {
      for (int q = 0; q < kLnGroups; ++q) ptx::mbar_init(&full[s * kLnGroups + q], 1);
      ptx::mbar_init(&empty[s], kLnRows);
    }
    ptx::fence_mbar_init();
  }
  __syncthreads();
  const int ngroups = (M + kLnRows - 1) / kLnRows;
  float4 dwacc[NV], dbacc[HAS_DB ? NV : 1];
#pragma unroll
  for (int j = 0; j < NV; ++j) dwacc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int j = 0; j < (HAS_DB ? NV : 1); ++j) dbacc[j] = make_float4(0.f, 0.f, 0.f, 0.f);

  if (warp == kLnWarps) {
    if (lane == 0) {
      int s = 0, q = 0;                      // ring slot, consumer group of the slab, parity of the slot's use count:
      uint32_t ph = 0;                       // all stepped (no division per slab)
      for (int g = blockIdx.x; g < ngroups;
           g += gridDim.x, (++s == nstages ? (s = 0, ph ^= 1u) : 0u), (++q == kLnGroups ? (q = 0) : 0)) {
        ptx::mbar_wait(&empty[s], ph ^ 1u);
        uint64_t* fb = &full[s * kLnGroups + q];
        const int row0 = g * kLnRows;
        const uint32_t rows = static_cast<uint32_t>(min(kLnRows, M - row0));
        uint8_t* st = ln_smem + static_cast<size_t>(s) * stage_bytes;
        ptx::mbar_expect_tx(fb, rows * (dy_row + f_row + (dres_in != nullptr ? f_row : 0u)));
        ptx::bulk_load_1d(st, dy + static_cast<size_t>(row0) * C, rows * dy_row, fb);
        ptx::bulk_load_1d(st + off_x, x + static_cast<size_t>(row0) * C, rows * f_row, fb);
        if (dres_in != nullptr)
          ptx::bulk_load_1d(st + off_r, dres_in + static_cast<size_t>(row0) * C, rows * f_row, fb);
      }
    }
  } else {
    const float invC = 1.0f / static_cast<float>(C);
    // two groups of kLnRows warps take alternate slabs (four consumer warps per SM sub-partition hide the
    // shared-memory and shuffle latencies of one another).  A warp waits only on its own group's barrier of a
    // slot; the group comes back to the same slot every `period` of its slabs, which is when the parity flips.
    const int cg = warp / kLnRows, wr = warp % kLnRows;   // consumer group, row within the slab
    // the weight vector is the same for every row: into registers once, instead of two L1 round trips per row
    float4 wreg[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (lane + 32 * j) * 4;
      wreg[j] = (c < C) ? __ldg(reinterpret_cast<const float4*>(w + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    int g = blockIdx.x + cg * gridDim.x;                  // this group's first slab: the cg-th of the CTA
    int s = cg % nstages;                                 // its ring slot
    static_assert(kLnGroups == 2, "period below assumes two consumer groups");
    const int period = (nstages % kLnGroups == 0) ? nstages / kLnGroups : nstages;
    int turn = 0;                                         // slabs of this group since the parity last flipped
    uint32_t ph = 0;
    float mu_n = 0.f, rs_n = 0.f;
    if (g < ngroups && g * kLnRows + wr < M) { mu_n = mean[g * kLnRows + wr]; rs_n = rstd[g * kLnRows + wr]; }
    for (; g < ngroups; g += kLnGroups * gridDim.x) {
      ptx::mbar_wait(&full[s * kLnGroups + cg], ph);
      const int row = g * kLnRows + wr;
      const float mu = mu_n, rs = rs_n;
      const int gn = g + kLnGroups * gridDim.x;           // statistics of this group's next slab, one turn ahead
      if (gn < ngroups && gn * kLnRows + wr < M) { mu_n = mean[gn * kLnRows + wr]; rs_n = rstd[gn * kLnRows + wr]; }
      if (row < M) {
        const uint8_t* st = ln_smem + static_cast<size_t>(s) * stage_bytes;
        const DyT* sdy = reinterpret_cast<const DyT*>(st) + wr * C;
        const float* sx = reinterpret_cast<const float*>(st + off_x) + wr * C;
        const float* sr = reinterpret_cast<const float*>(st + off_r) + wr * C;
        const size_t off = static_cast<size_t>(row) * C;
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
          const int c = (lane + 32 * j) * 4;
          if (c < C) {
            const float4 wv = wreg[j];
            const float4 d = Vec4IO<DyT>::load(sdy + c);
            const float4 xv = Vec4IO<float>::load(sx + c);
            const float4 xh = make_float4((xv.x - mu) * rs, (xv.y - mu) * rs, (xv.z - mu) * rs, (xv.w - mu) * rs);
            const float4 gg = make_float4(d.x * wv.x, d.y * wv.y, d.z * wv.z, d.w * wv.w);
            s1 += (gg.x + gg.y) + (gg.z + gg.w);
            s2 += (gg.x * xh.x + gg.y * xh.y) + (gg.z * xh.z + gg.w * xh.w);
            dwacc[j].x += d.x * xh.x; dwacc[j].y += d.y * xh.y;
            dwacc[j].z += d.z * xh.z; dwacc[j].w += d.w * xh.w;
            if (HAS_DB) { dbacc[j].x += d.x; dbacc[j].y += d.y; dbacc[j].z += d.z; dbacc[j].w += d.w; }
          }
        }
        const float c1 = warp_sum(s1) * invC, c2 = warp_sum(s2) * invC;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
          const int c = (lane + 32 * j) * 4;
          if (c < C) {
            const float4 wv = wreg[j];
            const float4 d = Vec4IO<DyT>::load(sdy + c);
            const float4 xv = Vec4IO<float>::load(sx + c);
            const float4 xh = make_float4((xv.x - mu) * rs, (xv.y - mu) * rs, (xv.z - mu) * rs, (xv.w - mu) * rs);
            float4 o = make_float4(rs * (d.x * wv.x - c1 - xh.x * c2), rs * (d.y * wv.y - c1 - xh.y * c2),
                                   rs * (d.z * wv.z - c1 - xh.z * c2), rs * (d.w * wv.w - c1 - xh.w * c2));
            if (dres_in != nullptr) {
              const float4 r = Vec4IO<float>::load(sr + c);
              o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w;
            }
            Vec4IO<float>::store(dx + off + c, o);
            if (dx_lowp != nullptr) Vec4IO<__nv_bfloat16>::store(dx_lowp + off + c, o);
          }
        }
      }
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&empty[s]);
      s += kLnGroups;                                     // this group's next slab is kLnGroups slots further on
      while (s >= nstages) s -= nstages;
      if (++turn == period) { turn = 0; ph ^= 1u; }
    }
  }
  // cross-warp reduction of the column partials (the staging ring is idle now and reused), dw then db
  float* red = reinterpret_cast<float*>(ln_smem);
  __syncthreads();
#pragma unroll
  for (int pass = 0; pass < (HAS_DB ? 2 : 1); ++pass) {
    if (pass > 0) __syncthreads();
    if (warp < kLnWarps) {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int c = (lane + 32 * j) * 4;
        const float4 a = (pass == 0 || !HAS_DB) ? dwacc[j] : dbacc[HAS_DB ? j : 0];
        if (c < 128 * NV) *reinterpret_cast<float4*>(&red[warp * 128 * NV + c]) = a;
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < kLnWarps; ++k) s += red[k * 128 * NV + c];
      part[(static_cast<size_t>(blockIdx.x) * 2 + pass) * C + c] = s;
    }
  }
}

// Stage 2: out[c] (+)= sum_p part[p][which][c].  grid (ceil(C/32), 2), block (32, 32)
__global__ void __launch_bounds__(1024)
    layernorm_bwd_reduce_kernel(const float* __restrict__ part, float* __restrict__ dw,
                                float* __restrict__ db, int nparts, int C, int accumulate) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  __shared__ float red[32][33];
  const int which = blockIdx.y;
  float* out = which == 0 ? dw : db;
  if (out == nullptr) return;
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (c < C)
    for (int p = threadIdx.y; p < nparts; p += 32)
      s += part[(static_cast<size_t>(p) * 2 + which) * C + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) t += red[k][threadIdx.x];
    out[c] = accumulate ? out[c] + t : t;
  }
}

// =============================================================================================
// K9 softmax cross-entropy.  ROWS rows per 256-thread CTA: ROWS=1 -> CTA per row (large V),
// ROWS=8 -> warp per row (small V).
// =============================================================================================
template <typename T>
struct RowVec;  // EPV elements per 128-bit access
template <>
struct RowVec<float> {
  static constexpr int EPV = 4;
  static __device__ __forceinline__ void load(const float* p, float* o) {
    const uint4 u = ld_stream(p);
    o[0] = __uint_as_float(u.x); o[1] = __uint_as_float(u.y);
    o[2] = __uint_as_float(u.z); o[3] = __uint_as_float(u.w);
  }
  static __device__ __forceinline__ void store(float* p, const float* o) {
    st_stream(p, make_uint4(__float_as_uint(o[0]), __float_as_uint(o[1]), __float_as_uint(o[2]),
                            __float_as_uint(o[3])));
  }
  static __device__ __forceinline__ float get(const float* p) { return *p; }
  static __device__ __forceinline__ void put(float* p, float v) { *p = v; }
};
template <>
struct RowVec<__nv_bfloat16> {
  static constexpr int EPV = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float* o) {
    const uint4 u = ld_stream(p);
    float2 t;
    t = unpack_bf16x2(u.x); o[0] = t.x; o[1] = t.y;
    t = unpack_bf16x2(u.y); o[2] = t.x; o[3] = t.y;
    t = unpack_bf16x2(u.z); o[4] = t.x; o[5] = t.y;
    t = unpack_bf16x2(u.w); o[6] = t.x; o[7] = t.y;
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float* o) {
    st_stream(p, make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]),
                            pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7])));
  }
  static __device__ __forceinline__ float get(const __nv_bfloat16* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void put(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

template <typename T> __device__ __forceinline__ float ce_exp(float x);
template <> __device__ __forceinline__ float ce_exp<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ float ce_exp<__nv_bfloat16>(float x) { return __expf(x); }
template <typename T> __device__ __forceinline__ float ce_log(float x);
template <> __device__ __forceinline__ float ce_log<float>(float x) { return logf(x); }
template <> __device__ __forceinline__ float ce_log<__nv_bfloat16>(float x) { return __logf(x); }

__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  const float nm = fmaxf(m, m2);
  if (nm == -INFINITY) { m = nm; s = 0.f; return; }
  s = s * expf(m - nm) + s2 * expf(m2 - nm);
  m = nm;
}

template <typename T, int ROWS, bool VEC>
__global__ void __launch_bounds__(256)
    cross_entropy_fwd_kernel(const T* __restrict__ logits, const int64_t* __restrict__ targets,
                             float* __restrict__ lse, float* __restrict__ row_loss, int M, int V,
                             int ld) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  constexpr int TPR = 256 / ROWS;  // threads per row
  constexpr int EPV = RowVec<T>::EPV;
  const int sub = threadIdx.x / TPR, t = threadIdx.x % TPR;
  const int row = blockIdx.x * ROWS + sub;
  const bool live = row < M;
  float m = -INFINITY, s = 0.f;
  if (live) {
    const T* r = logits + static_cast<size_t>(row) * ld;
    if (VEC) {
      const int nfull = V / EPV;             // vectors without a ragged edge: no per-element bounds checks
      // two vectors per thread per trip: two independent 16-byte loads in flight
      int i = t;
      for (; i + TPR < nfull; i += 2 * TPR) {
        float e[EPV], f[EPV];
        RowVec<T>::load(r + i * EPV, e);
        RowVec<T>::load(r + (i + TPR) * EPV, f);
        float lm = fmaxf(e[0], f[0]);
#pragma unroll
        for (int k = 1; k < EPV; ++k) lm = fmaxf(lm, fmaxf(e[k], f[k]));
        if (lm > m) { s *= ce_exp<T>(m - lm); m = lm; }
#pragma unroll
        for (int k = 0; k < EPV; ++k) s += ce_exp<T>(e[k] - m) + ce_exp<T>(f[k] - m);
      }
      const int nvec = (V + EPV - 1) / EPV;
      for (; i < nvec; i += TPR) {
        float e[EPV];
        RowVec<T>::load(r + i * EPV, e);
        float lm = -INFINITY;
#pragma unroll
        for (int k = 0; k < EPV; ++k) {
          if (i * EPV + k >= V) e[k] = -INFINITY;
          lm = fmaxf(lm, e[k]);
        }
        if (lm > m) { s *= ce_exp<T>(m - lm); m = lm; }
#pragma unroll
        for (int k = 0; k < EPV; ++k) s += ce_exp<T>(e[k] - m);
      }
    } else {
      for (int i = t; i < V; i += TPR) {
        const float e = RowVec<T>::get(r + i);
        if (e > m) { s *= ce_exp<T>(m - e); m = e; }
        s += ce_exp<T>(e - m);
      }
    }
  }
  // combine (m, s) across the TPR threads of the row
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    online_merge(m, s, m2, s2);
  }
  if (TPR > 32) {
    __shared__ float sm[8], ss[8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sm[warp] = m; ss[warp] = s; }
    __syncthreads();
    m = sm[0]; s = ss[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) online_merge(m, s, sm[k], ss[k]);
  }
  if (live && t == 0) {
    const float l = m + ce_log<T>(s);
    const int64_t tg = targets[row];
    lse[row] = l;
    row_loss[row] = (tg < 0) ? 0.f : l - RowVec<T>::get(logits + static_cast<size_t>(row) * ld + tg);
  }
}

// stats[0] = sum(row_loss)/n_valid, stats[1] = n_valid.  One CTA, fixed reduction tree.
__global__ void __launch_bounds__(1024)
    cross_entropy_reduce_kernel(const float* __restrict__ row_loss,
                                const int64_t* __restrict__ targets, float* __restrict__ stats,
                                int M) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  __shared__ float ssum[32], scnt[32];
  float s = 0.f, n = 0.f;
  for (int i = threadIdx.x; i < M; i += 1024) {
    s += row_loss[i];
    n += (targets[i] >= 0) ? 1.f : 0.f;
  }
  s = warp_sum(s); n = warp_sum(n);
  if ((threadIdx.x & 31) == 0) { ssum[threadIdx.x >> 5] = s; scnt[threadIdx.x >> 5] = n; }
  __syncthreads();
  if (threadIdx.x < 32) {
    s = warp_sum(ssum[threadIdx.x]); n = warp_sum(scnt[threadIdx.x]);
    if (threadIdx.x == 0) { stats[0] = s / n; stats[1] = n; }  // 0/0 = nan like torch
  }
}

template <typename T, int ROWS, bool VEC>
__global__ void __launch_bounds__(256)
    cross_entropy_bwd_kernel(T* __restrict__ logits, const int64_t* __restrict__ targets,
                             const float* __restrict__ lse, const float* __restrict__ stats,
                             const float* __restrict__ gscale, int M, int V, int ld) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  constexpr int TPR = 256 / ROWS;
  constexpr int EPV = RowVec<T>::EPV;
  const int sub = threadIdx.x / TPR, t = threadIdx.x % TPR;
  const int row = blockIdx.x * ROWS + sub;
  if (row >= M) return;
  T* r = logits + static_cast<size_t>(row) * ld;
  const int64_t tg = targets[row];
  const float l = lse[row];
  const float scale = (tg < 0) ? 0.f : (gscale ? *gscale : 1.f) / stats[1];
  if (VEC) {
    const int nvec = ld / EPV;
    for (int i = t; i < nvec; i += TPR) {
      float e[EPV];
      RowVec<T>::load(r + i * EPV, e);
#pragma unroll
      for (int k = 0; k < EPV; ++k) {
        const int col = i * EPV + k;
        const float p = (col < V && tg >= 0) ? ce_exp<T>(e[k] - l) : 0.f;
        e[k] = (p - (col == tg ? 1.f : 0.f)) * scale;
      }
      RowVec<T>::store(r + i * EPV, e);
    }
  } else {
    for (int i = t; i < ld; i += TPR) {
      const float p = (i < V && tg >= 0) ? ce_exp<T>(RowVec<T>::get(r + i) - l) : 0.f;
      RowVec<T>::put(r + i, (p - (i == tg ? 1.f : 0.f)) * scale);
    }
  }
}

// Fused forward + backward for bf16 logits whose row fits in shared memory (one CTA per row, two
// resident CTAs per SM): the row comes from HBM once (bulk async copies in four pieces), max, log-sum-
// exp and the loss are computed from shared memory, and softmax - onehot goes straight back over the
// logits.  HBM traffic: one read + one write of the logits instead of two reads + one write.
// The gradient is left UNSCALED (the caller folds gscale / n_valid into the GEMMs that consume it).
constexpr int kCeFusedThreads = 512;
constexpr int kCeFusedPieces = 4;
__global__ void __launch_bounds__(kCeFusedThreads, 2)
    cross_entropy_fused_kernel(__nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ targets,
                               float* __restrict__ lse, float* __restrict__ row_loss, int V, int ld) {
  extern __shared__ __align__(128) uint8_t ce_smem[];
  __shared__ uint64_t bars[kCeFusedPieces];
  __shared__ float red[kCeFusedThreads / 32];
  __shared__ float bcast[2];
  const int row = blockIdx.x;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  __nv_bfloat16* grow = logits + static_cast<size_t>(row) * ld;
  const int nvec = ld / 8;                                   // 16-byte vectors per row (ld % 8 == 0)
  const int pv = (nvec + kCeFusedPieces - 1) / kCeFusedPieces;
  if (tid == 0) {
    for (int p = 0; p < kCeFusedPieces; ++p) ptx::mbar_init(&bars[p], 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  pdl_wait();
  pdl_trigger();
  if (tid == 0) {
    for (int p = 0; p < kCeFusedPieces; ++p) {
      const int v0 = p * pv, v1 = min(nvec, v0 + pv);
      if (v1 > v0) {
        const uint32_t bytes = static_cast<uint32_t>(v1 - v0) * 16u;
        ptx::mbar_expect_tx(&bars[p], bytes);
        ptx::bulk_load_1d(ce_smem + static_cast<size_t>(v0) * 16, grow + static_cast<size_t>(v0) * 8, bytes, &bars[p]);
      } else {
        ptx::mbar_arrive(&bars[p]);
      }
    }
  }
  const int64_t tg = targets[row];
  const uint4* srow = reinterpret_cast<const uint4*>(ce_smem);
  const int nfull = V / 8;                                   // vectors entirely inside the V valid columns
  // ---- pass 1: row maximum (packed bf16 max; piece p is consumed as soon as it has landed)
  __nv_bfloat162 mx2 = __float2bfloat162_rn(-INFINITY);
  float mtail = -INFINITY;
  for (int p = 0; p < kCeFusedPieces; ++p) {
    ptx::mbar_wait(&bars[p], 0);
    const int v1 = min(nvec, (p + 1) * pv);
    for (int i = p * pv + tid; i < v1; i += kCeFusedThreads) {
      const uint4 u = srow[i];
      if (i < nfull) {
        const __nv_bfloat162 a = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&u.x), *reinterpret_cast<const __nv_bfloat162*>(&u.y));
        const __nv_bfloat162 b = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&u.z), *reinterpret_cast<const __nv_bfloat162*>(&u.w));
        mx2 = __hmax2(mx2, __hmax2(a, b));
      } else {
        const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&u);
        for (int k = 0; k < 8; ++k)
          if (i * 8 + k < V) mtail = fmaxf(mtail, __bfloat162float(e[k]));
      }
    }
  }
  float m = fmaxf(fmaxf(__low2float(mx2), __high2float(mx2)), mtail);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) red[warp] = m;
  __syncthreads();
  if (warp == 0) {
    float v = lane < kCeFusedThreads / 32 ? red[lane] : -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (lane == 0) bcast[0] = v;
  }
  __syncthreads();
  m = bcast[0];
  if (m == -INFINITY) m = 0.f;
  constexpr float kLog2e = 1.4426950408889634f;
  auto ex2 = [](float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; };
  // ---- pass 2: sum of exp(x - max)
  const float mneg = -m * kLog2e;
  const f32x2 l2e2 = pack2(kLog2e, kLog2e), mneg2 = pack2(mneg, mneg);
  f32x2 s2 = pack2(0.f, 0.f);
  float s = 0.f;
  for (int i = tid; i < nvec; i += kCeFusedThreads) {
    const uint4 u = srow[i];
    if (i < nfull) {
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {                 // packed fp32: one fma and one add per pair of logits
        const float2 x = unpack_bf16x2(w[k]);
        float a0, a1;
        unpack2(fma2(pack2(x.x, x.y), l2e2, mneg2), a0, a1);
        s2 = add2(s2, pack2(ex2(a0), ex2(a1)));
      }
    } else {
      const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&u);
      for (int k = 0; k < 8; ++k)
        if (i * 8 + k < V) s += ex2(fmaf(__bfloat162float(e[k]), kLog2e, mneg));
    }
  }
  {
    float e0, e1;
    unpack2(s2, e0, e1);
    s += e0 + e1;
  }
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (warp == 0) {
    float v = lane < kCeFusedThreads / 32 ? red[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) bcast[1] = v;
  }
  __syncthreads();
  s = bcast[1];
  const float l = m + __logf(s);
  if (tid == 0) {
    lse[row] = l;
    row_loss[row] = (tg < 0) ? 0.f : l - __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(ce_smem)[tg]);
  }
  // ---- pass 3: softmax - onehot (0 for ignored rows and for the padding columns V..ld-1), over the logits
  const float lneg = -l * kLog2e;
  const f32x2 lneg2 = pack2(lneg, lneg);
  const bool liverow = tg >= 0;
  const int tvec = liverow ? static_cast<int>(tg >> 3) : -1, tk = static_cast<int>(tg & 7);
  uint4* gout = reinterpret_cast<uint4*>(grow);
  for (int i = tid; i < nvec; i += kCeFusedThreads) {
    float o[8];
    if (!liverow) {
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = 0.f;
    } else {
      const uint4 u = srow[i];
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 x = unpack_bf16x2(w[k]);
        float a0, a1;
        unpack2(fma2(pack2(x.x, x.y), l2e2, lneg2), a0, a1);
        o[2 * k] = ex2(a0);
        o[2 * k + 1] = ex2(a1);
      }
      if (i >= nfull) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (i * 8 + k >= V) o[k] = 0.f;
      }
      if (i == tvec) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (k == tk) o[k] -= 1.f;
      }
    }
    st_stream(gout + i, make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7])));
  }
}

// =============================================================================================
// K12 grad norm + clip coefficient
// =============================================================================================
__global__ void __launch_bounds__(256)
    sumsq_partial_kernel(const float* __restrict__ g, size_t n, float* __restrict__ part) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  __shared__ float red[8];
  float s = 0.f;
  const size_t n4 = n / 4;
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * 256) {
    const uint4 u = ld_stream(g + i * 4);
    const float a = __uint_as_float(u.x), b = __uint_as_float(u.y), c = __uint_as_float(u.z),
                d = __uint_as_float(u.w);
    s += (a * a + b * b) + (c * c + d * d);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float a = g[n4 * 4 + threadIdx.x];
    s += a * a;
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = warp_sum(threadIdx.x < 8 ? red[threadIdx.x] : 0.f);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
  }
}
__global__ void __launch_bounds__(1024)
    clip_coef_kernel(const float* __restrict__ part, int nparts, float max_norm,
                     float* __restrict__ out) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  __shared__ double red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < nparts; i += 1024) s += static_cast<double>(part[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < 32; ++k) t += red[k];
    const float norm = static_cast<float>(sqrt(t));
    out[0] = norm;
    out[1] = fminf(1.0f, max_norm / (norm + 1e-6f));
  }
}

// =============================================================================================
// K13 fused AdamW: one CTA = one 1024-element chunk (uniform weight-decay flag)
// =============================================================================================
__global__ void __launch_bounds__(256)
    adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, __nv_bfloat16* __restrict__ p_lowp,
                 const uint8_t* __restrict__ chunk_decay, float beta1, float beta2, float eps,
                 float decay_factor, float step_size, float bc2_sqrt,
                 const float* __restrict__ gscale) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t i = (static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x) * 4;
  const float gs = gscale ? *gscale : 1.f;
  const float df = chunk_decay[blockIdx.x] ? decay_factor : 1.f;
  float4 pv = *reinterpret_cast<float4*>(p + i);
  const uint4 gu = ld_stream(g + i);
  float4 mv = *reinterpret_cast<float4*>(m + i);
  float4 vv = *reinterpret_cast<float4*>(v + i);
  const float gg[4] = {__uint_as_float(gu.x) * gs, __uint_as_float(gu.y) * gs,
                       __uint_as_float(gu.z) * gs, __uint_as_float(gu.w) * gs};
  float pp[4] = {pv.x, pv.y, pv.z, pv.w}, mm[4] = {mv.x, mv.y, mv.z, mv.w},
        vq[4] = {vv.x, vv.y, vv.z, vv.w};
  const float w1 = 1.f - beta1, w2 = 1.f - beta2;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    pp[k] *= df;                                   // p.mul_(1 - lr*wd)
    mm[k] = mm[k] + w1 * (gg[k] - mm[k]);          // exp_avg.lerp_(g, 1-beta1)
    vq[k] = vq[k] * beta2 + (w2 * gg[k]) * gg[k];  // exp_avg_sq.mul_(b2).addcmul_(g, g, 1-b2)
    const float denom = sqrtf(vq[k]) / bc2_sqrt + eps;
    pp[k] = pp[k] - step_size * (mm[k] / denom);   // p.addcdiv_(m, denom, -step_size)
  }
  *reinterpret_cast<float4*>(p + i) = make_float4(pp[0], pp[1], pp[2], pp[3]);
  *reinterpret_cast<float4*>(m + i) = make_float4(mm[0], mm[1], mm[2], mm[3]);
  *reinterpret_cast<float4*>(v + i) = make_float4(vq[0], vq[1], vq[2], vq[3]);
  if (p_lowp != nullptr)
    *reinterpret_cast<uint2*>(p_lowp + i) =
        make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
}

// =============================================================================================
// casts / scale / column sums
// =============================================================================================
__global__ void __launch_bounds__(256)
    cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, size_t n,
                         float scale) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t n8 = n / 8;
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < n8;
       i += static_cast<size_t>(gridDim.x) * 256) {
    const uint4 a = ld_stream(in + i * 8), b = ld_stream(in + i * 8 + 4);
    *reinterpret_cast<uint4*>(out + i * 8) =
        make_uint4(pack_bf16x2(__uint_as_float(a.x) * scale, __uint_as_float(a.y) * scale),
                   pack_bf16x2(__uint_as_float(a.z) * scale, __uint_as_float(a.w) * scale),
                   pack_bf16x2(__uint_as_float(b.x) * scale, __uint_as_float(b.y) * scale),
                   pack_bf16x2(__uint_as_float(b.z) * scale, __uint_as_float(b.w) * scale));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 7))
    out[n8 * 8 + threadIdx.x] = __float2bfloat16_rn(in[n8 * 8 + threadIdx.x] * scale);
}
__global__ void __launch_bounds__(256)
    cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ out, size_t n) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t n8 = n / 8;
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < n8;
       i += static_cast<size_t>(gridDim.x) * 256) {
    const uint4 u = ld_stream(in + i * 8);
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z),
                 d = unpack_bf16x2(u.w);
    *reinterpret_cast<float4*>(out + i * 8) = make_float4(a.x, a.y, b.x, b.y);
    *reinterpret_cast<float4*>(out + i * 8 + 4) = make_float4(c.x, c.y, d.x, d.y);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 7))
    out[n8 * 8 + threadIdx.x] = __bfloat162float(in[n8 * 8 + threadIdx.x]);
}
__global__ void __launch_bounds__(256) scale_f32_kernel(float* __restrict__ x, size_t n, float s) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t n4 = n / 4;
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * 256) {
    float4 v = *reinterpret_cast<float4*>(x + i * 4);
    v.x *= s; v.y *= s; v.z *= s; v.w *= s;
    *reinterpret_cast<float4*>(x + i * 4) = v;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) x[n4 * 4 + threadIdx.x] *= s;
}
// exact-erf GELU on bf16 activations (fp32 math inside), 8 elements per thread per step
__device__ __forceinline__ uint4 gelu8(const uint4 u) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
  uint32_t o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 x = unpack_bf16x2(w[k]);
    const float2 y = gelu_fast2(x.x, x.y);
    o[k] = pack_bf16x2(y.x, y.y);
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}
__device__ __forceinline__ uint4 dgelu8(const uint4 g, const uint4 u) {
  const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, hw[4] = {u.x, u.y, u.z, u.w};
  uint32_t o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 x = unpack_bf16x2(hw[k]), d = unpack_bf16x2(gw[k]);
    const float2 y = gelu_fast_grad_mul2(x.x, x.y, d.x, d.y);
    o[k] = pack_bf16x2(y.x, y.y);
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}
// four 16-byte vectors per thread per trip: 64 bytes of independent loads in flight per thread (with two the
// warps sat on the first use of each vector: 128 KB per SM in flight covers the loaded DRAM latency, 64 KB did not)
__global__ void __launch_bounds__(256)
    gelu_fwd_bf16_kernel(const __nv_bfloat16* __restrict__ h, __nv_bfloat16* __restrict__ a, size_t n8) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t stride = static_cast<size_t>(gridDim.x) * 256;
  size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
  for (; i + 3 * stride < n8; i += 4 * stride) {
    const uint4 u0 = ld_stream(h + i * 8), u1 = ld_stream(h + (i + stride) * 8);
    const uint4 u2 = ld_stream(h + (i + 2 * stride) * 8), u3 = ld_stream(h + (i + 3 * stride) * 8);
    st_stream(a + i * 8, gelu8(u0));
    st_stream(a + (i + stride) * 8, gelu8(u1));
    st_stream(a + (i + 2 * stride) * 8, gelu8(u2));
    st_stream(a + (i + 3 * stride) * 8, gelu8(u3));
  }
  for (; i < n8; i += stride) st_stream(a + i * 8, gelu8(ld_stream(h + i * 8)));
}
// dh = dA * gelu'(h), in place over dA
__global__ void __launch_bounds__(256)
    gelu_bwd_bf16_kernel(__nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ h, size_t n8) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  const size_t stride = static_cast<size_t>(gridDim.x) * 256;
  size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
  for (; i + stride < n8; i += 2 * stride) {
    const uint4 g0 = ld_stream(da + i * 8), u0 = ld_stream(h + i * 8);
    const uint4 g1 = ld_stream(da + (i + stride) * 8), u1 = ld_stream(h + (i + stride) * 8);
    st_stream(da + i * 8, dgelu8(g0, u0));
    st_stream(da + (i + stride) * 8, dgelu8(g1, u1));
  }
  if (i < n8) st_stream(da + i * 8, dgelu8(ld_stream(da + i * 8), ld_stream(h + i * 8)));
}

template <typename T>
__global__ void __launch_bounds__(256)
    colsum_kernel(const T* __restrict__ dy, int ld, float* __restrict__ db, int accumulate, int M,
                  int N) {
  pdl_wait();     // dependent launch: the previous kernel has completed beyond this point
  pdl_trigger();  // ... and the next one may be scheduled as our CTAs retire
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (c < N)
    for (int r = threadIdx.y; r < M; r += 8)
      s += RowVec<T>::get(dy + static_cast<size_t>(r) * ld + c);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < N) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
    db[c] = accumulate ? db[c] + t : t;
  }
}

static inline unsigned grid_for(size_t work_items, int per_block, int max_blocks) {
  size_t b = (work_items + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > static_cast<size_t>(max_blocks)) b = max_blocks;
  return static_cast<unsigned>(b);
}
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace nsb

using namespace nsb;

// =============================================================================================
// C-ABI
// =============================================================================================
extern "C" {

int nsb_version(void) { return 100; }
const char* nsb_last_error(void) { return g_err; }
unsigned long long nsb_launch_count(void) { return g_launches.load(); }

int nsb_embedding_fwd(const int64_t* idx, const float* wte, const float* wpe, float* x, int B,
                      int T, int C, int V, nsb_stream_t stream) {
  (void)V;
  NSB_REQUIRE(C % 4 == 0, "embedding_fwd: n_embd %d must be a multiple of 4", C);
  NSB_REQUIRE(aligned16(wte) && aligned16(wpe) && aligned16(x), "embedding_fwd: unaligned pointer");
  const int M = B * T, C4 = C / 4;
  const size_t total = static_cast<size_t>(M) * C4;
  if (total == 0) return 0;
  launch_pdl(embedding_fwd_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0, cs(stream), 
      idx, reinterpret_cast<const float4*>(wte), reinterpret_cast<const float4*>(wpe),
      reinterpret_cast<float4*>(x), M, T, C4);
  count_launch();
  return check_launch("embedding_fwd");
}

int nsb_embedding_bwd(const int64_t* idx, const float* dx, float* dwte, float* dwpe, int B, int T,
                      int C, int V, int accumulate_wpe, nsb_stream_t stream) {
  (void)V;
  NSB_REQUIRE(C % 4 == 0, "embedding_bwd: n_embd %d must be a multiple of 4", C);
  NSB_REQUIRE(aligned16(dx) && aligned16(dwte) && aligned16(dwpe), "embedding_bwd: unaligned pointer");
  const int M = B * T, C4 = C / 4;
  const size_t total = static_cast<size_t>(M) * C4;
  if (total == 0) return 0;
  launch_pdl(embedding_bwd_wte_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0, cs(stream), 
      idx, reinterpret_cast<const float4*>(dx), reinterpret_cast<float4*>(dwte), M, C4);
  const size_t tp = static_cast<size_t>(T) * C4;
  launch_pdl(embedding_bwd_wpe_kernel, dim3(static_cast<unsigned>((tp + 255) / 256)), dim3(256), 0, cs(stream), 
      reinterpret_cast<const float4*>(dx), reinterpret_cast<float4*>(dwpe), B, T, C4,
      accumulate_wpe);
  count_launch(2);
  return check_launch("embedding_bwd");
}

#define NSB_LN_DISPATCH(NVV, ...)                 \
  switch (NVV) {                                  \
    case 1: { constexpr int NV = 1; __VA_ARGS__; break; } \
    case 2: { constexpr int NV = 2; __VA_ARGS__; break; } \
    case 3: { constexpr int NV = 3; __VA_ARGS__; break; } \
    case 4: { constexpr int NV = 4; __VA_ARGS__; break; } \
    case 5: case 6: { constexpr int NV = 6; __VA_ARGS__; break; } \
    default: { constexpr int NV = 8; __VA_ARGS__; break; } \
  }

int nsb_layernorm_fwd(const float* x, const float* w, const float* b, void* y, int y_dtype,
                      float* mean, float* rstd, int M, int C, float eps, nsb_stream_t stream) {
  NSB_REQUIRE(C % 4 == 0 && C >= 4 && C <= 1024, "layernorm_fwd: n_embd %d unsupported (need %%4, <=1024)", C);
  NSB_REQUIRE(y_dtype == NSB_F32 || y_dtype == NSB_BF16, "layernorm_fwd: bad dtype %d", y_dtype);
  NSB_REQUIRE(aligned16(x) && aligned16(w) && aligned16(y) && (b == nullptr || aligned16(b)),
              "layernorm_fwd: unaligned pointer");
  if (M == 0) return 0;
  const int nv = (C + 127) / 128;
  const unsigned grid = (M + 7) / 8;
  NSB_LN_DISPATCH(nv, {
    if (y_dtype == NSB_F32)
      launch_pdl(layernorm_fwd_kernel<NV, float>, dim3(grid), dim3(256), 0, cs(stream), 
          x, w, b, static_cast<float*>(y), mean, rstd, M, C, eps);
    else
      launch_pdl(layernorm_fwd_kernel<NV, __nv_bfloat16>, dim3(grid), dim3(256), 0, cs(stream), 
          x, w, b, static_cast<__nv_bfloat16*>(y), mean, rstd, M, C, eps);
  });
  count_launch();
  return check_launch("layernorm_fwd");
}

static inline int ln_bwd_grid(int M) {
  int g = (M + 7) / 8;
  const int cap = 2 * sm_count();  // 2 resident CTAs per SM; scratch contract: at most 4 x 160 slabs
  return g < cap ? (g < 1 ? 1 : g) : cap;
}
size_t nsb_layernorm_bwd_scratch_bytes(int M, int C) {
  (void)M;
  // sized for the largest grid the launcher may pick, independent of the device queried later
  return static_cast<size_t>(4 * 160) * 2 * static_cast<size_t>(C) * sizeof(float);
}

namespace nsb {
// shared by nsb_layernorm_bwd and any variant that supplies its own scratch
int layernorm_bwd_launch(const void* dy, int dy_dtype, const float* x, const float* w,
                         const float* mean, const float* rstd, const float* dres_in, float* dx,
                         void* dx_lowp, float* dw, float* db, int accumulate, float* scratch,
                         size_t scratch_bytes, int M, int C, cudaStream_t stream) {
  NSB_REQUIRE(C % 4 == 0 && C >= 4 && C <= 1024, "layernorm_bwd: n_embd %d unsupported (need %%4, <=1024)", C);
  NSB_REQUIRE(dy_dtype == NSB_F32 || dy_dtype == NSB_BF16, "layernorm_bwd: bad dtype %d", dy_dtype);
  NSB_REQUIRE(aligned16(dy) && aligned16(x) && aligned16(w) && aligned16(dx) &&
                  (dres_in == nullptr || aligned16(dres_in)) && (dx_lowp == nullptr || aligned16(dx_lowp)),
              "layernorm_bwd: unaligned pointer");
  if (M == 0) return 0;
  int grid = ln_bwd_grid(M);
  const size_t per_part = 2 * static_cast<size_t>(C) * sizeof(float);
  if (static_cast<size_t>(grid) * per_part > scratch_bytes) grid = static_cast<int>(scratch_bytes / per_part);
  NSB_REQUIRE(grid >= 1 && scratch != nullptr, "layernorm_bwd: scratch too small (%zu bytes)", scratch_bytes);
  const int nv = (C + 127) / 128;
  // streaming kernel: rows must be bulk-copyable (row bytes % 16 == 0) and a ring of >= 2 slabs must fit
  const size_t dy_row = static_cast<size_t>(C) * (dy_dtype == NSB_F32 ? 4 : 2);
  const size_t slab = kLnRows * (dy_row + static_cast<size_t>(C) * 4 * (dres_in != nullptr ? 2 : 1));
  int nstages = static_cast<int>((200 * 1024) / slab);
  if (nstages > 4) nstages = 4;
  const bool stream_ok = (dy_row % 16 == 0) && (C % 4 == 0) && nstages >= 2 && M >= 4 * kLnRows &&
                         static_cast<size_t>(nstages) * slab >= static_cast<size_t>(kLnWarps) * 128 * nv * 4;
  if (stream_ok) {
    int sgrid = (M + kLnRows - 1) / kLnRows;
    if (sgrid > sm_count()) sgrid = sm_count();
    if (sgrid > grid) sgrid = grid;
    grid = sgrid;
    const size_t smem = static_cast<size_t>(nstages) * slab + 8 * (4 * kLnGroups + 4);   // ring + its mbarriers
#define NSB_LN_BWDS(TT, HASDB)                                                                          \
  {                                                                                                     \
    auto kern = layernorm_bwd_stream_kernel<NV, TT, HASDB>;                                             \
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);                \
    launch_pdl(kern, dim3(grid), dim3(32 * (kLnWarps + 1)), smem, stream, static_cast<const TT*>(dy), x, w, mean, rstd, dres_in, dx,        \
                                       static_cast<__nv_bfloat16*>(dx_lowp), scratch, M, C, nstages);   \
  }
    NSB_LN_DISPATCH(nv, {
      if (dy_dtype == NSB_F32) { if (db != nullptr) NSB_LN_BWDS(float, true) else NSB_LN_BWDS(float, false) }
      else { if (db != nullptr) NSB_LN_BWDS(__nv_bfloat16, true) else NSB_LN_BWDS(__nv_bfloat16, false) }
    });
#undef NSB_LN_BWDS
  } else {
#define NSB_LN_BWD(TT, HASDB)                                                                  \
  launch_pdl(layernorm_bwd_kernel<NV, TT, HASDB>, dim3(grid), dim3(256), 0, stream,                                 \
      static_cast<const TT*>(dy), x, w, mean, rstd, dres_in, dx, static_cast<__nv_bfloat16*>(dx_lowp), scratch, M, C)
    NSB_LN_DISPATCH(nv, {
      if (dy_dtype == NSB_F32) { if (db != nullptr) NSB_LN_BWD(float, true); else NSB_LN_BWD(float, false); }
      else { if (db != nullptr) NSB_LN_BWD(__nv_bfloat16, true); else NSB_LN_BWD(__nv_bfloat16, false); }
    });
#undef NSB_LN_BWD
  }
  launch_pdl(layernorm_bwd_reduce_kernel, dim3(dim3((C + 31) / 32, 2)), dim3(dim3(32, 32)), 0, stream, 
      scratch, dw, db, grid, C, accumulate);
  count_launch(2);
  return check_launch("layernorm_bwd");
}
}  // namespace nsb

int nsb_layernorm_bwd(const void* dy, int dy_dtype, const float* x, const float* w,
                      const float* mean, const float* rstd, const float* dres_in, float* dx,
                      void* dx_lowp, float* dw, float* db, int accumulate, float* scratch,
                      size_t scratch_bytes, int M, int C, nsb_stream_t stream) {
  return layernorm_bwd_launch(dy, dy_dtype, x, w, mean, rstd, dres_in, dx, dx_lowp, dw, db,
                              accumulate, scratch, scratch_bytes, M, C, cs(stream));
}

int nsb_cross_entropy_fwd(const void* logits, int dtype, const int64_t* targets, float* lse,
                          float* row_loss, float* stats, int M, int V, int ld,
                          nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_F32 || dtype == NSB_BF16, "cross_entropy_fwd: bad dtype %d", dtype);
  NSB_REQUIRE(ld >= V && V >= 1, "cross_entropy_fwd: ld %d < V %d", ld, V);
  NSB_REQUIRE(M >= 1, "cross_entropy_fwd: empty batch");
  const int epv = dtype == NSB_F32 ? 4 : 8;
  const bool vec = aligned16(logits) && (ld % epv == 0) && ld >= ((V + epv - 1) / epv) * epv;
  const bool small = V <= 2048;
  const unsigned grid = small ? (M + 7) / 8 : M;
#define CE_FWD(TT, ROWS, VEC)                                                     \
  launch_pdl(cross_entropy_fwd_kernel<TT, ROWS, VEC>, dim3(grid), dim3(256), 0, cs(stream),           \
      static_cast<const TT*>(logits), targets, lse, row_loss, M, V, ld)
  if (dtype == NSB_F32) {
    if (small) { if (vec) CE_FWD(float, 8, true); else CE_FWD(float, 8, false); }
    else       { if (vec) CE_FWD(float, 1, true); else CE_FWD(float, 1, false); }
  } else {
    if (small) { if (vec) CE_FWD(__nv_bfloat16, 8, true); else CE_FWD(__nv_bfloat16, 8, false); }
    else       { if (vec) CE_FWD(__nv_bfloat16, 1, true); else CE_FWD(__nv_bfloat16, 1, false); }
  }
#undef CE_FWD
  launch_pdl(cross_entropy_reduce_kernel, dim3(1), dim3(1024), 0, cs(stream), row_loss, targets, stats, M);
  count_launch(2);
  return check_launch("cross_entropy_fwd");
}

int nsb_cross_entropy_bwd(void* logits, int dtype, const int64_t* targets, const float* lse,
                          const float* stats, const float* gscale, int M, int V, int ld,
                          nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_F32 || dtype == NSB_BF16, "cross_entropy_bwd: bad dtype %d", dtype);
  NSB_REQUIRE(ld >= V && V >= 1, "cross_entropy_bwd: ld %d < V %d", ld, V);
  if (M == 0) return 0;
  const int epv = dtype == NSB_F32 ? 4 : 8;
  const bool vec = aligned16(logits) && (ld % epv == 0);
  const bool small = V <= 2048;
  const unsigned grid = small ? (M + 7) / 8 : M;
#define CE_BWD(TT, ROWS, VEC)                                                     \
  launch_pdl(cross_entropy_bwd_kernel<TT, ROWS, VEC>, dim3(grid), dim3(256), 0, cs(stream),           \
      static_cast<TT*>(logits), targets, lse, stats, gscale, M, V, ld)
  if (dtype == NSB_F32) {
    if (small) { if (vec) CE_BWD(float, 8, true); else CE_BWD(float, 8, false); }
    else       { if (vec) CE_BWD(float, 1, true); else CE_BWD(float, 1, false); }
  } else {
    if (small) { if (vec) CE_BWD(__nv_bfloat16, 8, true); else CE_BWD(__nv_bfloat16, 8, false); }
    else       { if (vec) CE_BWD(__nv_bfloat16, 1, true); else CE_BWD(__nv_bfloat16, 1, false); }
  }
#undef CE_BWD
  count_launch();
  return check_launch("cross_entropy_bwd");
}

int nsb_cross_entropy_fwd_bwd(void* logits, int dtype, const int64_t* targets, float* lse,
                              float* row_loss, float* stats, int M, int V, int ld, nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_BF16, "cross_entropy_fwd_bwd: bf16 logits only (dtype %d); use the separate passes", dtype);
  NSB_REQUIRE(ld >= V && V >= 1, "cross_entropy_fwd_bwd: ld %d < V %d", ld, V);
  NSB_REQUIRE(M >= 1, "cross_entropy_fwd_bwd: empty batch");
  NSB_REQUIRE(ld % 8 == 0 && aligned16(logits), "cross_entropy_fwd_bwd: rows must be 16-byte aligned (ld %d)", ld);
  const size_t smem = static_cast<size_t>(ld) * 2;
  NSB_REQUIRE(smem <= 232448 - 1024, "cross_entropy_fwd_bwd: a row of %d logits does not fit in shared memory; "
              "use the separate passes", ld);
  static PerDeviceOnce attr_set;
  if (attr_set.first_use()) {
    cudaError_t e = cudaFuncSetAttribute(cross_entropy_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 1024);
    NSB_REQUIRE(e == cudaSuccess, "cross_entropy_fwd_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  }
  launch_pdl(cross_entropy_fused_kernel, dim3(M), dim3(kCeFusedThreads), smem, cs(stream),
             static_cast<__nv_bfloat16*>(logits), targets, lse, row_loss, V, ld);
  launch_pdl(cross_entropy_reduce_kernel, dim3(1), dim3(1024), 0, cs(stream), row_loss, targets, stats, M);
  count_launch(2);
  return check_launch("cross_entropy_fwd_bwd");
}

static inline int norm_grid(size_t n) { return static_cast<int>(grid_for(n / 4, 256 * 8, 8 * 160)); }
size_t nsb_grad_norm_scratch_bytes(size_t n) { (void)n; return static_cast<size_t>(8 * 160) * sizeof(float); }

namespace nsb {
int grad_clip_coef_launch(const float* g, size_t n, float max_norm, float* scratch,
                          size_t scratch_bytes, float* out, cudaStream_t stream) {
  NSB_REQUIRE(aligned16(g), "grad_clip_coef: unaligned gradient arena");
  int grid = norm_grid(n);
  if (static_cast<size_t>(grid) * sizeof(float) > scratch_bytes) grid = static_cast<int>(scratch_bytes / sizeof(float));
  NSB_REQUIRE(grid >= 1 && scratch != nullptr, "grad_clip_coef: scratch too small (%zu bytes)", scratch_bytes);
  launch_pdl(sumsq_partial_kernel, dim3(grid), dim3(256), 0, stream, g, n, scratch);
  launch_pdl(clip_coef_kernel, dim3(1), dim3(1024), 0, stream, scratch, grid, max_norm, out);
  count_launch(2);
  return check_launch("grad_clip_coef");
}
}  // namespace nsb

int nsb_grad_clip_coef(const float* g, size_t n, float max_norm, float* scratch,
                       size_t scratch_bytes, float* out, nsb_stream_t stream) {
  return grad_clip_coef_launch(g, n, max_norm, scratch, scratch_bytes, out, cs(stream));
}

int nsb_adamw_step(float* p, const float* g, float* m, float* v, void* p_lowp,
                   const uint8_t* chunk_decay, size_t n, float beta1, float beta2, float eps,
                   float decay_factor, float step_size, float bc2_sqrt, const float* gscale,
                   nsb_stream_t stream) {
  NSB_REQUIRE(n % 1024 == 0, "adamw_step: arena length %zu is not a multiple of 1024", n);
  NSB_REQUIRE(aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v), "adamw_step: unaligned arena");
  if (n == 0) return 0;
  launch_pdl(adamw_kernel, dim3(static_cast<unsigned>(n / 1024)), dim3(256), 0, cs(stream), 
      p, g, m, v, static_cast<__nv_bfloat16*>(p_lowp), chunk_decay, beta1, beta2, eps,
      decay_factor, step_size, bc2_sqrt, gscale);
  count_launch();
  return check_launch("adamw_step");
}

int nsb_cast_f32_to_bf16(const float* in, void* out, size_t n, float scale, nsb_stream_t stream) {
  NSB_REQUIRE(aligned16(in) && aligned16(out), "cast_f32_to_bf16: unaligned pointer");
  if (n == 0) return 0;
  launch_pdl(cast_f32_bf16_kernel, dim3(grid_for(n / 8, 256 * 4, 8 * sm_count())), dim3(256), 0, cs(stream), 
      in, static_cast<__nv_bfloat16*>(out), n, scale);
  count_launch();
  return check_launch("cast_f32_to_bf16");
}
int nsb_cast_bf16_to_f32(const void* in, float* out, size_t n, nsb_stream_t stream) {
  NSB_REQUIRE(aligned16(in) && aligned16(out), "cast_bf16_to_f32: unaligned pointer");
  if (n == 0) return 0;
  launch_pdl(cast_bf16_f32_kernel, dim3(grid_for(n / 8, 256 * 4, 8 * sm_count())), dim3(256), 0, cs(stream), 
      static_cast<const __nv_bfloat16*>(in), out, n);
  count_launch();
  return check_launch("cast_bf16_to_f32");
}
int nsb_scale_f32(float* x, size_t n, float scale, nsb_stream_t stream) {
  NSB_REQUIRE(aligned16(x), "scale_f32: unaligned pointer");
  if (n == 0) return 0;
  launch_pdl(scale_f32_kernel, dim3(grid_for(n / 4, 256 * 4, 8 * sm_count())), dim3(256), 0, cs(stream), x, n, scale);
  count_launch();
  return check_launch("scale_f32");
}
int nsb_gelu_fwd_bf16(const void* h, void* a, size_t n, nsb_stream_t stream) {
  NSB_REQUIRE(n % 8 == 0 && aligned16(h) && aligned16(a), "gelu_fwd_bf16: need 16-byte aligned buffers and n %% 8 == 0");
  if (n == 0) return 0;
  launch_pdl(gelu_fwd_bf16_kernel, dim3(grid_for(n / 8, 256 * 4, 16 * sm_count())), dim3(256), 0, cs(stream), 
      static_cast<const __nv_bfloat16*>(h), static_cast<__nv_bfloat16*>(a), n / 8);
  count_launch();
  return check_launch("gelu_fwd_bf16");
}
int nsb_gelu_bwd_bf16(void* da, const void* h, size_t n, nsb_stream_t stream) {
  NSB_REQUIRE(n % 8 == 0 && aligned16(h) && aligned16(da), "gelu_bwd_bf16: need 16-byte aligned buffers and n %% 8 == 0");
  if (n == 0) return 0;
  launch_pdl(gelu_bwd_bf16_kernel, dim3(grid_for(n / 8, 256 * 2, 16 * sm_count())), dim3(256), 0, cs(stream), 
      static_cast<__nv_bfloat16*>(da), static_cast<const __nv_bfloat16*>(h), n / 8);
  count_launch();
  return check_launch("gelu_bwd_bf16");
}
int nsb_colsum(const void* dy, int dtype, int ld, float* db, int accumulate, int M, int N,
               nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_F32 || dtype == NSB_BF16, "colsum: bad dtype %d", dtype);
  if (N == 0) return 0;
  const dim3 grid((N + 31) / 32), block(32, 8);
  if (dtype == NSB_F32)
    launch_pdl(colsum_kernel<float>, dim3(grid), dim3(block), 0, cs(stream), static_cast<const float*>(dy), ld, db, accumulate, M, N);
  else
    launch_pdl(colsum_kernel<__nv_bfloat16>, dim3(grid), dim3(block), 0, cs(stream), static_cast<const __nv_bfloat16*>(dy), ld, db, accumulate, M, N);
  count_launch();
  return check_launch("colsum");
}

}  // extern "C"

// =============================================================================================
// LayerNorm backward for callers that cannot hand us scratch memory
// =============================================================================================
// Every other entry point of this library expects the caller to own all device memory, including
// the partial-sum scratch of the LayerNorm backward.  Some hosts cannot provide that buffer: an
// operator plugged into a foreign runtime only ever receives its input and output pointers, and a
// caller that is in the middle of a stream capture, or that runs under a fixed memory plan, may
// not call cudaMalloc at all.  For them the scratch lives in the library image itself: a
// file-scope device array, present once on every GPU that loads the library until the process
// exits.  Its size fixes the limit of this variant: rows up to kNoAllocMaxC wide.
namespace nsb {
constexpr int kNoAllocMaxC = 1024;     // widest row (n_embd) the built-in scratch can serve
constexpr int kNoAllocMaxParts = 148;  // column-partial slabs kept: one per SM of a B200
static __device__ __align__(256) float g_ln_bwd_builtin_scratch[kNoAllocMaxParts * 2 * kNoAllocMaxC];  // 1.16 MiB

// In host code the array's name only stands for the symbol; the address a kernel can use has to be
// asked for, per device, with that device current.
constexpr int kMaxDevices = 64;
static float* g_builtin_scratch_addr[kMaxDevices] = {nullptr};
static std::atomic<int> g_builtin_scratch_lock{0};

static int builtin_scratch_for_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error("layernorm_bwd_noalloc: cannot identify the current device (%s)", cudaGetErrorString(e));
    return e != cudaSuccess ? static_cast<int>(e) : -1;
  }
  while (g_builtin_scratch_lock.exchange(1, std::memory_order_acquire)) {}
  float* p = g_builtin_scratch_addr[dev];
  if (p == nullptr) {
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_ln_bwd_builtin_scratch);
    if (e != cudaSuccess) {
      g_builtin_scratch_lock.store(0, std::memory_order_release);
      set_error("layernorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return static_cast<int>(e);
    }
    g_builtin_scratch_addr[dev] = p;
  }
  g_builtin_scratch_lock.store(0, std::memory_order_release);
  *out = p;
  return 0;
}
}  // namespace nsb

extern "C" int nsb_layernorm_bwd_noalloc(const void* dy, int dy_dtype, const float* x, const float* w,
                                         const float* mean, const float* rstd, const float* dres_in,
                                         float* dx, void* dx_lowp, float* dw, float* db, int accumulate,
                                         int M, int C, nsb_stream_t stream) {
  NSB_REQUIRE(C <= kNoAllocMaxC,
              "layernorm_bwd_noalloc: row width %d exceeds the built-in scratch (at most %d columns); "
              "use nsb_layernorm_bwd with caller-provided scratch",
              C, kNoAllocMaxC);
  NSB_REQUIRE(M >= 0, "layernorm_bwd_noalloc: negative row count %d", M);
  // The built-in array is looked up for the CURRENT device; a stream that belongs to another device
  // would run the kernels there with this device's address.  Refuse instead of corrupting memory.
  if (stream != nullptr) {
    int cur = -1, sdev = -1;
    if (cudaGetDevice(&cur) == cudaSuccess && cudaStreamGetDevice(cs(stream), &sdev) == cudaSuccess)
      NSB_REQUIRE(cur == sdev,
                  "layernorm_bwd_noalloc: the stream belongs to device %d but device %d is current; "
                  "make the stream's device current before the call", sdev, cur);
    else
      (void)cudaGetLastError();   // could not tell (e.g. a capturing stream): proceed with the current device
  }
  float* scratch = nullptr;
  const int rc = builtin_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  // Stage 1 rewrites every partial that stage 2 reads, so nothing in the shared array has to be
  // cleared between calls; whatever the previous call left there is simply overwritten.
  return layernorm_bwd_launch(dy, dy_dtype, x, w, mean, rstd, dres_in, dx, dx_lowp, dw, db, accumulate,
                              scratch, sizeof(float) * kNoAllocMaxParts * 2 * kNoAllocMaxC, M, C, cs(stream));
}
