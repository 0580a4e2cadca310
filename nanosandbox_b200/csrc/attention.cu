// Causal self-attention, forward and backward, reading c_attn's [B,T,3,H,D] output in place.
//
//  * attn_*_mma kernels: bf16, D = 64, flash-style (online softmax, nothing of size T x T ever
//    reaches HBM), 64-query x 64-key tiles, cp.async double buffering, warp-level tensor-core
//    MMA (m16n8k16) with fp32 accumulation.  Backward is two passes (dK/dV per key block, dQ per
//    query block): deterministic, no atomics.
//  * attn_*_simt kernels: exact fp32 math for the 1e-5 parity mode and for head dims != 64.
#include "common.cuh"
#include "philox.cuh"

namespace nsb {

// =============================================================================================
// exact SIMT path (any D that is a multiple of 32 up to 128)
// =============================================================================================
template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// Dropout on the attention probabilities (F.scaled_dot_product_attention(dropout_p=...)): element (b, h, q, k)
// of the [B, H, T, T] probability tensor has linear index e = ((b H + h) T + q) T + k and is kept when word
// e % 4 of philox4x32_10(seed, e / 4, stream) is >= thr (philox.cuh).  The softmax normaliser uses the
// undropped probabilities; kept ones are scaled by 1 / (1 - p).
struct DropArgs { unsigned long long seed, stream; uint32_t thr; float inv_keep; };

// Random words of 32 consecutive Philox counters held across the lanes of a warp: word(e) for the elements
// e of the counters [cbase, cbase + 32).  All lanes call get() with the same e.
struct WarpPhilox {
  long long cbase = -1;
  uint32_t w[4];
  __device__ __forceinline__ uint32_t get(unsigned long long e, const DropArgs& d, int lane) {
    const long long c = static_cast<long long>(e >> 2);
    if (cbase < 0 || c < cbase || c - cbase >= 32) {
      cbase = c;
      const Philox4 r = philox4x32_10(d.seed, static_cast<unsigned long long>(c) + lane, d.stream);
      w[0] = r.w[0]; w[1] = r.w[1]; w[2] = r.w[2]; w[3] = r.w[3];
    }
    const int k = static_cast<int>(e & 3);
    const uint32_t mine = k == 0 ? w[0] : (k == 1 ? w[1] : (k == 2 ? w[2] : w[3]));
    return __shfl_sync(0xffffffffu, mine, static_cast<int>(c - cbase));
  }
};

template <typename T, int DPL, bool DROP>
__global__ void __launch_bounds__(256)
    attn_fwd_simt(const T* __restrict__ qkv, T* __restrict__ y, float* __restrict__ lse, int B, int Tn,
                  int H, float scale, DropArgs drop) {
  const int D = DPL * 32, C = H * D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r = static_cast<long long>(blockIdx.x) * 8 + warp;
  if (r >= static_cast<long long>(B) * H * Tn) return;
  const int q = static_cast<int>(r % Tn), h = static_cast<int>((r / Tn) % H), b = static_cast<int>(r / (static_cast<long long>(Tn) * H));
  const T* base = qkv + static_cast<size_t>(b) * Tn * 3 * C + h * D;
  float qv[DPL], o[DPL];
#pragma unroll
  for (int i = 0; i < DPL; ++i) { qv[i] = ldf(base + static_cast<size_t>(q) * 3 * C + lane + 32 * i); o[i] = 0.f; }
  float m = -INFINITY, l = 0.f;
  WarpPhilox rng;
  const unsigned long long erow = static_cast<unsigned long long>(r) * static_cast<unsigned long long>(Tn);
  for (int j = 0; j <= q; ++j) {
    const T* kp = base + static_cast<size_t>(j) * 3 * C + C;
    const T* vp = kp + C;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < DPL; ++i) s += qv[i] * ldf(kp + lane + 32 * i);
    s = warp_sum(s) * scale;
    const float mn = fmaxf(m, s), corr = expf(m - mn), p = expf(s - mn);
    l = l * corr + p;
    float pd = p;
    if (DROP) pd = rng.get(erow + j, drop, lane) >= drop.thr ? p * drop.inv_keep : 0.f;
#pragma unroll
    for (int i = 0; i < DPL; ++i) o[i] = o[i] * corr + pd * ldf(vp + lane + 32 * i);
    m = mn;
  }
  const float inv = 1.f / l;
#pragma unroll
  for (int i = 0; i < DPL; ++i) stf(y + (static_cast<size_t>(b) * Tn + q) * C + h * D + lane + 32 * i, o[i] * inv);
  if (lane == 0) lse[(static_cast<size_t>(b) * H + h) * Tn + q] = m + logf(l);
}

// delta[b,h,q] = sum_d dy * y
template <typename T>
__global__ void __launch_bounds__(256)
    attn_delta_kernel(const T* __restrict__ y, const T* __restrict__ dy, float* __restrict__ delta, int B, int Tn,
                      int H, int D) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r = static_cast<long long>(blockIdx.x) * 8 + warp;
  if (r >= static_cast<long long>(B) * H * Tn) return;
  const int q = static_cast<int>(r % Tn), h = static_cast<int>((r / Tn) % H), b = static_cast<int>(r / (static_cast<long long>(Tn) * H));
  const size_t off = (static_cast<size_t>(b) * Tn + q) * H * D + h * D;
  float s = 0.f;
  for (int d = lane; d < D; d += 32) s += ldf(y + off + d) * ldf(dy + off + d);
  s = warp_sum(s);
  if (lane == 0) delta[(static_cast<size_t>(b) * H + h) * Tn + q] = s;
}

template <typename T, int DPL, bool DROP>
__global__ void __launch_bounds__(256)
    attn_bwd_dq_simt(const T* __restrict__ qkv, const T* __restrict__ dy, const float* __restrict__ lse,
                     const float* __restrict__ delta, T* __restrict__ dqkv, int B, int Tn, int H, float scale,
                     DropArgs drop) {
  const int D = DPL * 32, C = H * D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r = static_cast<long long>(blockIdx.x) * 8 + warp;
  if (r >= static_cast<long long>(B) * H * Tn) return;
  const int q = static_cast<int>(r % Tn), h = static_cast<int>((r / Tn) % H), b = static_cast<int>(r / (static_cast<long long>(Tn) * H));
  const T* base = qkv + static_cast<size_t>(b) * Tn * 3 * C + h * D;
  const size_t yoff = (static_cast<size_t>(b) * Tn + q) * C + h * D;
  const size_t so = (static_cast<size_t>(b) * H + h) * Tn + q;
  const float L = lse[so], dl = delta[so];
  float qv[DPL], dov[DPL], dq[DPL];
#pragma unroll
  for (int i = 0; i < DPL; ++i) {
    qv[i] = ldf(base + static_cast<size_t>(q) * 3 * C + lane + 32 * i);
    dov[i] = ldf(dy + yoff + lane + 32 * i);
    dq[i] = 0.f;
  }
  WarpPhilox rng;
  const unsigned long long erow = static_cast<unsigned long long>(r) * static_cast<unsigned long long>(Tn);
  for (int j = 0; j <= q; ++j) {
    const T* kp = base + static_cast<size_t>(j) * 3 * C + C;
    const T* vp = kp + C;
    float s = 0.f, dp = 0.f;
    float kv[DPL];
#pragma unroll
    for (int i = 0; i < DPL; ++i) {
      kv[i] = ldf(kp + lane + 32 * i);
      s += qv[i] * kv[i];
      dp += dov[i] * ldf(vp + lane + 32 * i);
    }
    s = warp_sum(s) * scale;
    dp = warp_sum(dp);
    if (DROP) dp = rng.get(erow + j, drop, lane) >= drop.thr ? dp * drop.inv_keep : 0.f;   // gradient through the mask
    const float ds = expf(s - L) * (dp - dl) * scale;
#pragma unroll
    for (int i = 0; i < DPL; ++i) dq[i] += ds * kv[i];
  }
#pragma unroll
  for (int i = 0; i < DPL; ++i)
    stf(dqkv + (static_cast<size_t>(b) * Tn + q) * 3 * C + h * D + lane + 32 * i, dq[i]);
}

template <typename T, int DPL, bool DROP>
__global__ void __launch_bounds__(256)
    attn_bwd_dkv_simt(const T* __restrict__ qkv, const T* __restrict__ dy, const float* __restrict__ lse,
                      const float* __restrict__ delta, T* __restrict__ dqkv, int B, int Tn, int H, float scale,
                      DropArgs drop) {
  const int D = DPL * 32, C = H * D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r = static_cast<long long>(blockIdx.x) * 8 + warp;
  if (r >= static_cast<long long>(B) * H * Tn) return;
  const int j = static_cast<int>(r % Tn), h = static_cast<int>((r / Tn) % H), b = static_cast<int>(r / (static_cast<long long>(Tn) * H));
  const T* base = qkv + static_cast<size_t>(b) * Tn * 3 * C + h * D;
  float kv[DPL], vv[DPL], dk[DPL], dv[DPL];
#pragma unroll
  for (int i = 0; i < DPL; ++i) {
    kv[i] = ldf(base + static_cast<size_t>(j) * 3 * C + C + lane + 32 * i);
    vv[i] = ldf(base + static_cast<size_t>(j) * 3 * C + 2 * C + lane + 32 * i);
    dk[i] = dv[i] = 0.f;
  }
  // masks of this key column for 32 queries at a time: lane L holds the word of query q0 + L
  const unsigned long long ehead = (static_cast<unsigned long long>(r) / Tn) * Tn;   // (b H + h) T
  int q0 = -1;
  uint32_t word = 0;
  for (int q = j; q < Tn; ++q) {
    const size_t so = (static_cast<size_t>(b) * H + h) * Tn + q;
    const size_t yoff = (static_cast<size_t>(b) * Tn + q) * C + h * D;
    float qv[DPL], dov[DPL];
    float s = 0.f, dp = 0.f;
#pragma unroll
    for (int i = 0; i < DPL; ++i) {
      qv[i] = ldf(base + static_cast<size_t>(q) * 3 * C + lane + 32 * i);
      dov[i] = ldf(dy + yoff + lane + 32 * i);
      s += qv[i] * kv[i];
      dp += dov[i] * vv[i];
    }
    s = warp_sum(s) * scale;
    dp = warp_sum(dp);
    const float p = expf(s - lse[so]);
    float pd = p;
    if (DROP) {
      if (q0 < 0 || q - q0 >= 32) {
        q0 = q;
        const unsigned long long e = (ehead + static_cast<unsigned long long>(q0 + lane)) * Tn + j;
        const Philox4 rr = philox4x32_10(drop.seed, e >> 2, drop.stream);
        const int k = static_cast<int>(e & 3);
        word = k == 0 ? rr.w[0] : (k == 1 ? rr.w[1] : (k == 2 ? rr.w[2] : rr.w[3]));
      }
      const bool keep = __shfl_sync(0xffffffffu, word, q - q0) >= drop.thr;
      pd = keep ? p * drop.inv_keep : 0.f;
      dp = keep ? dp * drop.inv_keep : 0.f;
    }
    const float ds = p * (dp - delta[so]) * scale;
#pragma unroll
    for (int i = 0; i < DPL; ++i) { dv[i] += pd * dov[i]; dk[i] += ds * qv[i]; }
  }
#pragma unroll
  for (int i = 0; i < DPL; ++i) {
    stf(dqkv + (static_cast<size_t>(b) * Tn + j) * 3 * C + C + h * D + lane + 32 * i, dk[i]);
    stf(dqkv + (static_cast<size_t>(b) * Tn + j) * 3 * C + 2 * C + h * D + lane + 32 * i, dv[i]);
  }
}

// keep / drop bytes (1 / 0) for n elements: the same bits the exact kernels above regenerate on the fly
// (word e % 4 of philox4x32_10(seed, e / 4, stream) >= thr), written once per attention call into a scratch
// buffer so that the mma.sync kernels below only have to load them
__global__ void __launch_bounds__(256)
    dropout_mask_kernel(uint8_t* __restrict__ mask, size_t n, uint32_t thr, unsigned long long seed,
                        unsigned long long stream) {
  const size_t nquads = (n + 3) / 4;
  for (size_t qd = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; qd < nquads;
       qd += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const Philox4 r = philox4x32_10(seed, qd, stream);
    const uint32_t bits = (r.w[0] >= thr ? 1u : 0u) | (r.w[1] >= thr ? 1u << 8 : 0u) | (r.w[2] >= thr ? 1u << 16 : 0u) |
                          (r.w[3] >= thr ? 1u << 24 : 0u);
    if (qd * 4 + 4 <= n) *reinterpret_cast<uint32_t*>(mask + qd * 4) = bits;
    else for (size_t k = qd * 4; k < n; ++k) mask[k] = static_cast<uint8_t>((bits >> (8 * (k - qd * 4))) & 1u);
  }
}

// =============================================================================================
// bf16 tensor-core path, D = 64
// =============================================================================================
namespace fa {
constexpr int BLK = 64;        // queries per CTA = keys per step
constexpr int TILE_BYTES = BLK * 128;  // 64 rows x 64 bf16

__device__ __forceinline__ uint32_t tile_off(int row, int chunk) {  // 16-byte chunk, XOR swizzle
  return static_cast<uint32_t>(row * 128 + ((chunk ^ (row & 7)) << 4));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool pred) {
  const int sz = pred ? 16 : 0;  // src-size 0 -> zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// load a 64 x 64 bf16 tile (rows row0.., row stride ld elements) into swizzled smem; 128 threads
__device__ __forceinline__ void load_tile(uint32_t smem_tile, const __nv_bfloat16* g, size_t ld, int row0,
                                          int nrows_valid) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int e = threadIdx.x + 128 * i, row = e >> 3, chunk = e & 7;
    const bool ok = (row0 + row) < nrows_valid;
    const __nv_bfloat16* src = g + static_cast<size_t>(ok ? row0 + row : 0) * ld + chunk * 8;
    cp_async16(smem_tile + tile_off(row, chunk), src, ok);
  }
}
// A fragments (16 rows starting r0, all 64 k) -> a[4][4]
__device__ __forceinline__ void load_a_frags(uint32_t a[4][4], uint32_t tile, int r0, int lane) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    const int i = lane >> 3;
    ldsm_x4(a[ks], tile + tile_off(r0 + (lane & 7) + (i & 1) * 8, ks * 2 + (i >> 1)));
  }
}
// acc[8][4] (16 x 64) += A(16 x 64 k; frags) . Bt, with the B tile stored [n][k] (k contiguous)
__device__ __forceinline__ void mma_a_bnk(float acc[8][4], const uint32_t a[4][4], uint32_t tile, int lane) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks)
#pragma unroll
    for (int np = 0; np < 4; ++np) {
      uint32_t b[4];
      const int i = lane >> 3;
      ldsm_x4(b, tile + tile_off(np * 16 + (lane & 7) + (i >> 1) * 8, ks * 2 + (i & 1)));
      mma16816(acc[2 * np], a[ks], b[0], b[1]);
      mma16816(acc[2 * np + 1], a[ks], b[2], b[3]);
    }
}
// acc[8][4] (16 x 64 n) += P(16 x 64 k; frags) . B, with the B tile stored [k][n] (n contiguous)
__device__ __forceinline__ void mma_a_bkn(float acc[8][4], const uint32_t p[4][4], uint32_t tile, int lane) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks)
#pragma unroll
    for (int np = 0; np < 4; ++np) {
      uint32_t b[4];
      const int i = lane >> 3;
      ldsm_x4_t(b, tile + tile_off(ks * 16 + (lane & 7) + (i & 1) * 8, np * 2 + (i >> 1)));
      mma16816(acc[2 * np], p[ks], b[0], b[1]);
      mma16816(acc[2 * np + 1], p[ks], b[2], b[3]);
    }
}
// C fragments (fp32, 16 x 64) -> A fragments (bf16) for the next MMA
__device__ __forceinline__ void c_to_a(uint32_t a[4][4], const float c[8][4]) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    a[ks][0] = pack_bf16x2(c[2 * ks][0], c[2 * ks][1]);
    a[ks][1] = pack_bf16x2(c[2 * ks][2], c[2 * ks][3]);
    a[ks][2] = pack_bf16x2(c[2 * ks + 1][0], c[2 * ks + 1][1]);
    a[ks][3] = pack_bf16x2(c[2 * ks + 1][2], c[2 * ks + 1][3]);
  }
}
__device__ __forceinline__ void zero_acc(float acc[8][4]) {
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
}
}  // namespace fa

// grid (ceil(T/64), B*H), 128 threads.  smem: Q | K0 V0 | K1 V1
__global__ void __launch_bounds__(128)
    attn_fwd_mma(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ y, float* __restrict__ lse,
                 int Tn, int H, float scale_log2, const uint8_t* __restrict__ dmask, float inv_keep) {
  using namespace fa;
  __shared__ __align__(128) uint8_t smem[5 * TILE_BYTES];
  const int C = H * 64;
  const int qb = gridDim.x - 1 - blockIdx.x;  // heavy (late) query blocks first
  const int bh = blockIdx.y, b = bh / H, h = bh % H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const size_t ld = static_cast<size_t>(3) * C;
  const __nv_bfloat16* qg = qkv + static_cast<size_t>(b) * Tn * ld + h * 64;
  const __nv_bfloat16* kg = qg + C;
  const __nv_bfloat16* vg = qg + 2 * C;
  const uint32_t sQ = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
  const uint32_t sK[2] = {sQ + TILE_BYTES, sQ + 3 * TILE_BYTES};
  const uint32_t sV[2] = {sQ + 2 * TILE_BYTES, sQ + 4 * TILE_BYTES};

  load_tile(sQ, qg, ld, qb * BLK, Tn);
  load_tile(sK[0], kg, ld, 0, Tn);
  load_tile(sV[0], vg, ld, 0, Tn);
  cp_commit();

  uint32_t qa[4][4];
  float o[8][4];
  zero_acc(o);
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  const int q0 = qb * BLK + warp * 16 + g;  // this thread's rows: q0 and q0 + 8

  for (int kb = 0; kb <= qb; ++kb) {
    const int cur = kb & 1;
    if (kb + 1 <= qb) {
      load_tile(sK[cur ^ 1], kg, ld, (kb + 1) * BLK, Tn);
      load_tile(sV[cur ^ 1], vg, ld, (kb + 1) * BLK, Tn);
      cp_commit();
      cp_wait<1>();
    } else {
      cp_wait<0>();
    }
    __syncthreads();
    if (kb == 0) load_a_frags(qa, sQ, warp * 16, lane);

    float s[8][4];
    zero_acc(s);
    mma_a_bnk(s, qa, sK[cur], lane);
    // scale into the log2 domain, causal mask on the diagonal block
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int key = kb * BLK + nt * 8 + 2 * t4 + (e & 1);
        const int qq = q0 + (e >> 1) * 8;
        s[nt][e] = (kb == qb && key > qq) ? -INFINITY : s[nt][e] * scale_log2;
      }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      float mx = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) mx = fmaxf(mx, fmaxf(s[nt][2 * r], s[nt][2 * r + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float mn = fmaxf(mrow[r], mx);
      const float corr = exp2f(mrow[r] - mn);
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        s[nt][2 * r] = exp2f(s[nt][2 * r] - mn);
        s[nt][2 * r + 1] = exp2f(s[nt][2 * r + 1] - mn);
        sum += s[nt][2 * r] + s[nt][2 * r + 1];
      }
      lrow[r] = lrow[r] * corr + sum;
      mrow[r] = mn;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) { o[nt][2 * r] *= corr; o[nt][2 * r + 1] *= corr; }
    }
    if (dmask != nullptr) {
      // dropout on the probabilities (the row sums above keep the undropped ones, as SDPA does); the 1 / (1 - p)
      // is applied once to the output row.  keys 2 t4, 2 t4 + 1 of a row are two adjacent mask bytes
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int qq = q0 + r * 8;
        const uint8_t* mrowp = dmask + (static_cast<size_t>(bh) * Tn + (qq < Tn ? qq : 0)) * Tn + kb * BLK + 2 * t4;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const int key = kb * BLK + nt * 8 + 2 * t4;
          if (key < Tn && !mrowp[nt * 8]) s[nt][2 * r] = 0.f;
          if (key + 1 < Tn && !mrowp[nt * 8 + 1]) s[nt][2 * r + 1] = 0.f;
        }
      }
    }
    uint32_t pa[4][4];
    c_to_a(pa, s);
    mma_a_bkn(o, pa, sV[cur], lane);
    __syncthreads();  // everyone done with buffer `cur` before it is refilled
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    float l = lrow[r];
    l += __shfl_xor_sync(0xffffffffu, l, 1);
    l += __shfl_xor_sync(0xffffffffu, l, 2);
    const int qq = q0 + r * 8;
    if (qq < Tn) {
      const float inv = inv_keep / l;
      __nv_bfloat16* yr = y + (static_cast<size_t>(b) * Tn + qq) * C + h * 64;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
        *reinterpret_cast<uint32_t*>(yr + nt * 8 + 2 * t4) = pack_bf16x2(o[nt][2 * r] * inv, o[nt][2 * r + 1] * inv);
      if (t4 == 0) lse[(static_cast<size_t>(b) * H + h) * Tn + qq] = (mrow[r] + log2f(l)) * 0.69314718055994531f;
    }
  }
}

// dK, dV for one 64-key block.  smem: K V | Q0 dO0 | Q1 dO1 | lse/delta x2
__global__ void __launch_bounds__(128)
    attn_bwd_dkv_mma(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dy,
                     const float* __restrict__ lse, const float* __restrict__ delta,
                     __nv_bfloat16* __restrict__ dqkv, int Tn, int H, float scale, float scale_log2,
                     const uint8_t* __restrict__ dmask, float inv_keep) {
  using namespace fa;
  extern __shared__ __align__(128) uint8_t smem_dyn[];
  uint8_t* smem = smem_dyn;
  const int C = H * 64;
  const int nqb = (Tn + BLK - 1) / BLK;
  const int kb = blockIdx.x;  // early key blocks have the most work and are scheduled first
  const int bh = blockIdx.y, b = bh / H, h = bh % H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const size_t ld = static_cast<size_t>(3) * C;
  const __nv_bfloat16* qg = qkv + static_cast<size_t>(b) * Tn * ld + h * 64;
  const __nv_bfloat16* dog = dy + static_cast<size_t>(b) * Tn * C + h * 64;
  const float* lseg = lse + (static_cast<size_t>(b) * H + h) * Tn;
  const float* delg = delta + (static_cast<size_t>(b) * H + h) * Tn;
  const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
  const uint32_t sK = s0, sV = s0 + TILE_BYTES;
  const uint32_t sQ[2] = {s0 + 2 * TILE_BYTES, s0 + 4 * TILE_BYTES};
  const uint32_t sdO[2] = {s0 + 3 * TILE_BYTES, s0 + 5 * TILE_BYTES};
  float* sStat = reinterpret_cast<float*>(smem + 6 * TILE_BYTES);  // [2 buffers][2][64]

  auto load_stats = [&](int buf, int qblk) {
    if (threadIdx.x < 64) {
      const int qq = qblk * BLK + threadIdx.x;
      sStat[buf * 128 + threadIdx.x] = qq < Tn ? lseg[qq] * 1.4426950408889634f : 0.f;
      sStat[buf * 128 + 64 + threadIdx.x] = qq < Tn ? delg[qq] : 0.f;
    }
  };

  load_tile(sK, qg + C, ld, kb * BLK, Tn);
  load_tile(sV, qg + 2 * C, ld, kb * BLK, Tn);
  load_tile(sQ[0], qg, ld, kb * BLK, Tn);
  load_tile(sdO[0], dog, C, kb * BLK, Tn);
  cp_commit();
  load_stats(0, kb);

  uint32_t ka[4][4], va[4][4];
  float dk[8][4], dv[8][4];
  zero_acc(dk);
  zero_acc(dv);
  const int key0 = kb * BLK + warp * 16 + g;  // this thread's key rows: key0, key0 + 8

  for (int qb = kb; qb < nqb; ++qb) {
    const int cur = (qb - kb) & 1;
    if (qb + 1 < nqb) {
      load_tile(sQ[cur ^ 1], qg, ld, (qb + 1) * BLK, Tn);
      load_tile(sdO[cur ^ 1], dog, C, (qb + 1) * BLK, Tn);
      cp_commit();
      load_stats(cur ^ 1, qb + 1);
      cp_wait<1>();
    } else {
      cp_wait<0>();
    }
    __syncthreads();
    if (qb == kb) {
      load_a_frags(ka, sK, warp * 16, lane);
      load_a_frags(va, sV, warp * 16, lane);
    }
    const float* st = sStat + cur * 128;
    // S^T = K Q^T  (rows = keys, cols = queries)
    float s[8][4];
    zero_acc(s);
    mma_a_bnk(s, ka, sQ[cur], lane);
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int qloc = nt * 8 + 2 * t4 + (e & 1);
        const int qq = qb * BLK + qloc;
        const int key = key0 + (e >> 1) * 8;
        const bool live = (qq >= key) && (qq < Tn);
        s[nt][e] = live ? exp2f(s[nt][e] * scale_log2 - st[qloc]) : 0.f;
      }
    // dropout: keep[key][q] of this thread's elements (0 / 1 / none); dV sees the dropped probabilities, the
    // gradient through the mask scales dP, dS keeps the undropped P
    float km[8][4];
    if (dmask != nullptr) {
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int qq = qb * BLK + nt * 8 + 2 * t4 + (e & 1);
          const int key = key0 + (e >> 1) * 8;
          km[nt][e] = (qq < Tn && key < Tn && dmask[(static_cast<size_t>(bh) * Tn + qq) * Tn + key]) ? inv_keep : 0.f;
        }
    }
    uint32_t pa[4][4];
    if (dmask != nullptr) {
      float sd[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) sd[nt][e] = s[nt][e] * km[nt][e];
      c_to_a(pa, sd);
    } else {
      c_to_a(pa, s);
    }
    mma_a_bkn(dv, pa, sdO[cur], lane);          // dV += P^T dO
    float dp[8][4];
    zero_acc(dp);
    mma_a_bnk(dp, va, sdO[cur], lane);          // dP^T = V dO^T
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int qloc = nt * 8 + 2 * t4 + (e & 1);
        const float dpe = dmask != nullptr ? dp[nt][e] * km[nt][e] : dp[nt][e];
        dp[nt][e] = s[nt][e] * (dpe - st[64 + qloc]);
      }
    c_to_a(pa, dp);
    mma_a_bkn(dk, pa, sQ[cur], lane);           // dK += dS^T Q
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int key = key0 + r * 8;
    if (key < Tn) {
      __nv_bfloat16* dkr = dqkv + (static_cast<size_t>(b) * Tn + key) * ld + C + h * 64;
      __nv_bfloat16* dvr = dkr + C;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        *reinterpret_cast<uint32_t*>(dkr + nt * 8 + 2 * t4) = pack_bf16x2(dk[nt][2 * r] * scale, dk[nt][2 * r + 1] * scale);
        *reinterpret_cast<uint32_t*>(dvr + nt * 8 + 2 * t4) = pack_bf16x2(dv[nt][2 * r], dv[nt][2 * r + 1]);
      }
    }
  }
}

// dQ for one 64-query block.  smem: Q dO | K0 V0 | K1 V1
__global__ void __launch_bounds__(128)
    attn_bwd_dq_mma(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dy,
                    const float* __restrict__ lse, const float* __restrict__ delta,
                    __nv_bfloat16* __restrict__ dqkv, int Tn, int H, float scale, float scale_log2,
                    const uint8_t* __restrict__ dmask, float inv_keep) {
  using namespace fa;
  extern __shared__ __align__(128) uint8_t smem_dyn[];
  uint8_t* smem = smem_dyn;
  const int C = H * 64;
  const int qb = gridDim.x - 1 - blockIdx.x;
  const int bh = blockIdx.y, b = bh / H, h = bh % H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const size_t ld = static_cast<size_t>(3) * C;
  const __nv_bfloat16* qg = qkv + static_cast<size_t>(b) * Tn * ld + h * 64;
  const __nv_bfloat16* dog = dy + static_cast<size_t>(b) * Tn * C + h * 64;
  const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
  const uint32_t sQ = s0, sdO = s0 + TILE_BYTES;
  const uint32_t sK[2] = {s0 + 2 * TILE_BYTES, s0 + 4 * TILE_BYTES};
  const uint32_t sV[2] = {s0 + 3 * TILE_BYTES, s0 + 5 * TILE_BYTES};

  load_tile(sQ, qg, ld, qb * BLK, Tn);
  load_tile(sdO, dog, C, qb * BLK, Tn);
  load_tile(sK[0], qg + C, ld, 0, Tn);
  load_tile(sV[0], qg + 2 * C, ld, 0, Tn);
  cp_commit();

  const int q0 = qb * BLK + warp * 16 + g;
  float lrow[2], drow[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int qq = q0 + r * 8;
    const size_t so = (static_cast<size_t>(b) * H + h) * Tn + qq;
    lrow[r] = qq < Tn ? lse[so] * 1.4426950408889634f : 0.f;
    drow[r] = qq < Tn ? delta[so] : 0.f;
  }
  uint32_t qa[4][4], doa[4][4];
  float dq[8][4];
  zero_acc(dq);

  for (int kb = 0; kb <= qb; ++kb) {
    const int cur = kb & 1;
    if (kb + 1 <= qb) {
      load_tile(sK[cur ^ 1], qg + C, ld, (kb + 1) * BLK, Tn);
      load_tile(sV[cur ^ 1], qg + 2 * C, ld, (kb + 1) * BLK, Tn);
      cp_commit();
      cp_wait<1>();
    } else {
      cp_wait<0>();
    }
    __syncthreads();
    if (kb == 0) {
      load_a_frags(qa, sQ, warp * 16, lane);
      load_a_frags(doa, sdO, warp * 16, lane);
    }
    float s[8][4];
    zero_acc(s);
    mma_a_bnk(s, qa, sK[cur], lane);            // S = Q K^T
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int key = kb * BLK + nt * 8 + 2 * t4 + (e & 1);
        const int qq = q0 + (e >> 1) * 8;
        s[nt][e] = (key <= qq) ? exp2f(s[nt][e] * scale_log2 - lrow[e >> 1]) : 0.f;
      }
    float dp[8][4];
    zero_acc(dp);
    mma_a_bnk(dp, doa, sV[cur], lane);          // dP = dO V^T
    if (dmask != nullptr) {                     // gradient through the dropout mask
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int qq = q0 + r * 8;
        const uint8_t* mrowp = dmask + (static_cast<size_t>(bh) * Tn + (qq < Tn ? qq : 0)) * Tn + kb * BLK + 2 * t4;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const int key = kb * BLK + nt * 8 + 2 * t4;
          dp[nt][2 * r] = (key < Tn && mrowp[nt * 8]) ? dp[nt][2 * r] * inv_keep : 0.f;
          dp[nt][2 * r + 1] = (key + 1 < Tn && mrowp[nt * 8 + 1]) ? dp[nt][2 * r + 1] * inv_keep : 0.f;
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) dp[nt][e] = s[nt][e] * (dp[nt][e] - drow[e >> 1]);
    uint32_t pa[4][4];
    c_to_a(pa, dp);
    mma_a_bkn(dq, pa, sK[cur], lane);           // dQ += dS K
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int qq = q0 + r * 8;
    if (qq < Tn) {
      __nv_bfloat16* dqr = dqkv + (static_cast<size_t>(b) * Tn + qq) * ld + h * 64;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
        *reinterpret_cast<uint32_t*>(dqr + nt * 8 + 2 * t4) = pack_bf16x2(dq[nt][2 * r] * scale, dq[nt][2 * r + 1] * scale);
    }
  }
}

}  // namespace nsb

namespace nsb {
int attention_fwd_tc_launch(const void* qkv, void* y, float* lse, int B, int T, int H, cudaStream_t stream);
void attention_tc_set_debug_buffer(long long* p);
void attention_tc_set_dq_direct(int on);
void attention_tc_set_kv_tmem(int on);
void attention_tc_set_fwd_warps8(int on);
void attention_tc_set_bwd_warps16(int on);
void attention_tc_set_bwd_persistent(int on);
void attention_tc_set_fwd_p_tmem(int on);
int attention_bwd_tc_launch(const void* qkv, const void* y, const void* dy, const float* lse, void* dqkv,
                            float* delta, float* dq_acc, int B, int T, int H, cudaStream_t stream);

// ---- launchers of the mma.sync flash kernels (bf16, D = 64); dmask: keep bytes [B, H, T, T] or nullptr ----------
int attention_fwd_mma_launch(const void* qkv, void* y, float* lse, int B, int T, int H, const uint8_t* dmask,
                             float inv_keep, cudaStream_t stream) {
  const float scale = 1.0f / sqrtf(64.f);
  dim3 grid((T + 63) / 64, B * H);
  attn_fwd_mma<<<grid, 128, 0, stream>>>(static_cast<const __nv_bfloat16*>(qkv), static_cast<__nv_bfloat16*>(y), lse, T, H,
                                         scale * 1.4426950408889634f, dmask, inv_keep);
  count_launch();
  return check_launch("attention_fwd");
}
// delta must already hold sum_d dy * y
int attention_bwd_mma_launch(const void* qkv, const void* dy, const float* lse, const float* delta, void* dqkv, int B,
                             int T, int H, const uint8_t* dmask, float inv_keep, cudaStream_t stream) {
  const float scale = 1.0f / sqrtf(64.f);
  const int smem_dkv = 6 * fa::TILE_BYTES + 2 * 128 * 4, smem_dq = 6 * fa::TILE_BYTES;
  static PerDeviceOnce attr_done;
  if (attr_done.first_use()) {
    cudaFuncSetAttribute(attn_bwd_dkv_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dkv);
    cudaFuncSetAttribute(attn_bwd_dq_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq);
  }
  dim3 grid((T + 63) / 64, B * H);
  const float sl2 = scale * 1.4426950408889634f;
  attn_bwd_dkv_mma<<<grid, 128, smem_dkv, stream>>>(static_cast<const __nv_bfloat16*>(qkv), static_cast<const __nv_bfloat16*>(dy),
                                                     lse, delta, static_cast<__nv_bfloat16*>(dqkv), T, H, scale, sl2, dmask, inv_keep);
  attn_bwd_dq_mma<<<grid, 128, smem_dq, stream>>>(static_cast<const __nv_bfloat16*>(qkv), static_cast<const __nv_bfloat16*>(dy),
                                                   lse, delta, static_cast<__nv_bfloat16*>(dqkv), T, H, scale, sl2, dmask, inv_keep);
  count_launch(2);
  return check_launch("attention_bwd");
}
int attention_dropout_mask_launch(uint8_t* mask, size_t n, float p, unsigned long long seed, unsigned long long stream_id,
                                  cudaStream_t stream) {
  const size_t nquads = (n + 3) / 4;
  size_t blocks = (nquads + 255) / 256;
  const size_t cap = static_cast<size_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  dropout_mask_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(mask, n, drop_threshold(p), seed, stream_id);
  count_launch();
  return check_launch("attention_dropout_mask");
}

// ---- launchers of the exact SIMT kernels (drop == nullptr: no dropout) -----------------------------------
int attention_fwd_simt_launch(const void* qkv, void* y, float* lse, int dtype, int B, int T, int H, int D,
                              const DropArgs* drop, cudaStream_t stream) {
  const float scale = 1.0f / sqrtf(static_cast<float>(D));
  const long long rows = static_cast<long long>(B) * H * T;
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  const DropArgs d = drop ? *drop : DropArgs{0, 0, 0, 1.f};
#define ATT_FWD(TT, DPL)                                                                                                  \
  do {                                                                                                                    \
    if (drop) attn_fwd_simt<TT, DPL, true><<<grid, 256, 0, stream>>>(static_cast<const TT*>(qkv), static_cast<TT*>(y), lse, B, T, H, scale, d); \
    else attn_fwd_simt<TT, DPL, false><<<grid, 256, 0, stream>>>(static_cast<const TT*>(qkv), static_cast<TT*>(y), lse, B, T, H, scale, d);     \
  } while (0)
  if (dtype == NSB_F32) { switch (D / 32) { case 1: ATT_FWD(float, 1); break; case 2: ATT_FWD(float, 2); break; case 3: ATT_FWD(float, 3); break; default: ATT_FWD(float, 4); } }
  else { switch (D / 32) { case 1: ATT_FWD(__nv_bfloat16, 1); break; case 2: ATT_FWD(__nv_bfloat16, 2); break; case 3: ATT_FWD(__nv_bfloat16, 3); break; default: ATT_FWD(__nv_bfloat16, 4); } }
#undef ATT_FWD
  count_launch();
  return check_launch("attention_fwd");
}

// delta must already hold sum_d dy * y
int attention_bwd_simt_launch(const void* qkv, const void* dy, const float* lse, void* dqkv, const float* delta, int dtype,
                              int B, int T, int H, int D, const DropArgs* drop, cudaStream_t stream) {
  const float scale = 1.0f / sqrtf(static_cast<float>(D));
  const long long rows = static_cast<long long>(B) * H * T;
  const unsigned rgrid = static_cast<unsigned>((rows + 7) / 8);
  const DropArgs d = drop ? *drop : DropArgs{0, 0, 0, 1.f};
#define ATT_BWD(TT, DPL)                                                                                                  \
  do {                                                                                                                    \
    if (drop) {                                                                                                           \
      attn_bwd_dq_simt<TT, DPL, true><<<rgrid, 256, 0, stream>>>(static_cast<const TT*>(qkv), static_cast<const TT*>(dy), lse, delta, static_cast<TT*>(dqkv), B, T, H, scale, d);  \
      attn_bwd_dkv_simt<TT, DPL, true><<<rgrid, 256, 0, stream>>>(static_cast<const TT*>(qkv), static_cast<const TT*>(dy), lse, delta, static_cast<TT*>(dqkv), B, T, H, scale, d); \
    } else {                                                                                                              \
      attn_bwd_dq_simt<TT, DPL, false><<<rgrid, 256, 0, stream>>>(static_cast<const TT*>(qkv), static_cast<const TT*>(dy), lse, delta, static_cast<TT*>(dqkv), B, T, H, scale, d);  \
      attn_bwd_dkv_simt<TT, DPL, false><<<rgrid, 256, 0, stream>>>(static_cast<const TT*>(qkv), static_cast<const TT*>(dy), lse, delta, static_cast<TT*>(dqkv), B, T, H, scale, d); \
    }                                                                                                                     \
  } while (0)
  if (dtype == NSB_F32) { switch (D / 32) { case 1: ATT_BWD(float, 1); break; case 2: ATT_BWD(float, 2); break; case 3: ATT_BWD(float, 3); break; default: ATT_BWD(float, 4); } }
  else { switch (D / 32) { case 1: ATT_BWD(__nv_bfloat16, 1); break; case 2: ATT_BWD(__nv_bfloat16, 2); break; case 3: ATT_BWD(__nv_bfloat16, 3); break; default: ATT_BWD(__nv_bfloat16, 4); } }
#undef ATT_BWD
  count_launch(2);
  return check_launch("attention_bwd");
}
}
using namespace nsb;

static int g_attn_force_simt = 0;   // 1: exact SIMT kernels, 2: mma.sync flash kernels (cross-checks)

extern "C" {

// 0: tensor-core kernels, 1: SIMT, 2: mma.sync; 3 / 4: tensor-core backward with dQ staged through shared
// memory + TMA reduce-add (the default) / reduced straight from registers
void nsb_attention_debug_config(int force_simt) {
  if (force_simt == 3 || force_simt == 4) { attention_tc_set_dq_direct(force_simt == 4); g_attn_force_simt = 0; return; }
  // 5 / 6: tensor-core backward with every A operand from shared memory / K, V from tensor memory
  if (force_simt == 5 || force_simt == 6) { attention_tc_set_kv_tmem(force_simt == 6); g_attn_force_simt = 0; return; }
  // 10 / 11: tensor-core forward with P through shared memory / kept in tensor memory (the default)
  if (force_simt == 10 || force_simt == 11) { attention_tc_set_fwd_p_tmem(force_simt == 11); g_attn_force_simt = 0; return; }
  // 12 / 13: tensor-core backward with 8 / 16 probability warps (16 is the default)
  if (force_simt == 12 || force_simt == 13) { attention_tc_set_bwd_warps16(force_simt == 13); g_attn_force_simt = 0; return; }
  // 14 / 15: tensor-core forward with 4 / 8 softmax warps per CTA (8 is the default)
  if (force_simt == 14 || force_simt == 15) { attention_tc_set_fwd_warps8(force_simt == 15); g_attn_force_simt = 0; return; }
  // 16 / 17: 16-warp tensor-core backward with one CTA per (head, key block) / one persistent CTA per SM (the default)
  if (force_simt == 16 || force_simt == 17) { attention_tc_set_bwd_persistent(force_simt == 17); g_attn_force_simt = 0; return; }
  if (force_simt == 9) { attention_tc_set_kv_tmem(2); g_attn_force_simt = 0; return; }   // P^T in tensor memory for dV (the default)
  g_attn_force_simt = force_simt;
}
// device buffer (>= 1024 int64) receiving phase timestamps of one CTA of the tcgen05 kernels; NULL = off
void nsb_attention_debug_timing(void* device_buffer) { attention_tc_set_debug_buffer(static_cast<long long*>(device_buffer)); }

int nsb_attention_fwd(const void* qkv, void* y, float* lse, int dtype, int B, int T, int H, int D,
                      nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_F32 || dtype == NSB_BF16, "attention_fwd: bad dtype %d", dtype);
  NSB_REQUIRE(D % 32 == 0 && D >= 32 && D <= 128, "attention_fwd: head dim %d unsupported (32/64/96/128)", D);
  if (B == 0 || T == 0) return 0;
  const float scale = 1.0f / sqrtf(static_cast<float>(D));
  if (dtype == NSB_BF16 && D == 64 && g_attn_force_simt == 0) {
    NSB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0, "attention_fwd: qkv must be 16-byte aligned");
    return attention_fwd_tc_launch(qkv, y, lse, B, T, H, cs(stream));
  }
  if (dtype == NSB_BF16 && D == 64 && g_attn_force_simt == 2) {
    NSB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0, "attention_fwd: qkv must be 16-byte aligned");
    return attention_fwd_mma_launch(qkv, y, lse, B, T, H, nullptr, 1.f, cs(stream));
  }
  return attention_fwd_simt_launch(qkv, y, lse, dtype, B, T, H, D, nullptr, cs(stream));
}

int nsb_attention_bwd(const void* qkv, const void* y, const void* dy, const float* lse, void* dqkv,
                      float* delta, float* dq_acc, int dtype, int B, int T, int H, int D, nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_F32 || dtype == NSB_BF16, "attention_bwd: bad dtype %d", dtype);
  NSB_REQUIRE(D % 32 == 0 && D >= 32 && D <= 128, "attention_bwd: head dim %d unsupported (32/64/96/128)", D);
  if (B == 0 || T == 0) return 0;
  if (dtype == NSB_BF16 && D == 64 && g_attn_force_simt == 0 && dq_acc != nullptr) {
    NSB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(dy) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(dq_acc) & 15) == 0,
                "attention_bwd: qkv/dy/dq_acc must be 16-byte aligned");
    return attention_bwd_tc_launch(qkv, y, dy, lse, dqkv, delta, dq_acc, B, T, H, cs(stream));
  }
  const float scale = 1.0f / sqrtf(static_cast<float>(D));
  const long long rows = static_cast<long long>(B) * H * T;
  const unsigned rgrid = static_cast<unsigned>((rows + 7) / 8);
  if (dtype == NSB_F32)
    attn_delta_kernel<float><<<rgrid, 256, 0, cs(stream)>>>(static_cast<const float*>(y), static_cast<const float*>(dy), delta, B, T, H, D);
  else
    attn_delta_kernel<__nv_bfloat16><<<rgrid, 256, 0, cs(stream)>>>(static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), delta, B, T, H, D);
  count_launch();
  if (dtype == NSB_BF16 && D == 64 && g_attn_force_simt != 1) {
    NSB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(dy) & 15) == 0,
                "attention_bwd: qkv/dy must be 16-byte aligned");
    return attention_bwd_mma_launch(qkv, dy, lse, delta, dqkv, B, T, H, nullptr, 1.f, cs(stream));
  }
  return attention_bwd_simt_launch(qkv, dy, lse, dqkv, delta, dtype, B, T, H, D, nullptr, cs(stream));
}

// Attention with dropout on the probabilities (training with config.dropout > 0).  Same tensors as
// nsb_attention_fwd / _bwd; the mask is a function of (seed, stream_id) and the element's position (see
// DropArgs), so the backward call regenerates the forward's mask from the same two numbers.  With bf16, head
// dim 64 and a caller-provided scratch of B*H*T*T bytes the keep bytes are written there by one element-wise
// pass and the mma.sync flash kernels read them; otherwise the exact SIMT kernels draw the bits themselves
// (any head dim, fp32 or bf16, no scratch).
int nsb_attention_fwd_dropout(const void* qkv, void* y, float* lse, int dtype, int B, int T, int H, int D, float p,
                              unsigned long long seed, unsigned long long stream_id, void* mask_scratch,
                              size_t mask_scratch_bytes, nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_F32 || dtype == NSB_BF16, "attention_fwd_dropout: bad dtype %d", dtype);
  NSB_REQUIRE(D % 32 == 0 && D >= 32 && D <= 128, "attention_fwd_dropout: head dim %d unsupported (32/64/96/128)", D);
  NSB_REQUIRE(p >= 0.f && p < 1.f, "attention_fwd_dropout: p = %f is outside [0, 1)", p);
  if (B == 0 || T == 0) return 0;
  const size_t nmask = static_cast<size_t>(B) * H * T * T;
  if (dtype == NSB_BF16 && D == 64 && mask_scratch != nullptr && mask_scratch_bytes >= nmask && g_attn_force_simt != 1 &&
      (reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(mask_scratch) & 3) == 0) {
    // bf16, head dim 64: keep bytes into the caller's scratch, then the mma.sync flash kernel reads them
    const int rc = attention_dropout_mask_launch(static_cast<uint8_t*>(mask_scratch), nmask, p, seed, stream_id, cs(stream));
    if (rc) return rc;
    return attention_fwd_mma_launch(qkv, y, lse, B, T, H, static_cast<const uint8_t*>(mask_scratch), 1.0f / (1.0f - p), cs(stream));
  }
  const DropArgs d{seed, stream_id, drop_threshold(p), 1.0f / (1.0f - p)};
  return attention_fwd_simt_launch(qkv, y, lse, dtype, B, T, H, D, &d, cs(stream));
}

int nsb_attention_bwd_dropout(const void* qkv, const void* y, const void* dy, const float* lse, void* dqkv, float* delta,
                              int dtype, int B, int T, int H, int D, float p, unsigned long long seed,
                              unsigned long long stream_id, void* mask_scratch, size_t mask_scratch_bytes,
                              nsb_stream_t stream) {
  NSB_REQUIRE(dtype == NSB_F32 || dtype == NSB_BF16, "attention_bwd_dropout: bad dtype %d", dtype);
  NSB_REQUIRE(D % 32 == 0 && D >= 32 && D <= 128, "attention_bwd_dropout: head dim %d unsupported (32/64/96/128)", D);
  NSB_REQUIRE(p >= 0.f && p < 1.f, "attention_bwd_dropout: p = %f is outside [0, 1)", p);
  if (B == 0 || T == 0) return 0;
  const long long rows = static_cast<long long>(B) * H * T;
  const unsigned rgrid = static_cast<unsigned>((rows + 7) / 8);
  if (dtype == NSB_F32)
    attn_delta_kernel<float><<<rgrid, 256, 0, cs(stream)>>>(static_cast<const float*>(y), static_cast<const float*>(dy), delta, B, T, H, D);
  else
    attn_delta_kernel<__nv_bfloat16><<<rgrid, 256, 0, cs(stream)>>>(static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), delta, B, T, H, D);
  count_launch();
  const size_t nmask = static_cast<size_t>(B) * H * T * T;
  if (dtype == NSB_BF16 && D == 64 && mask_scratch != nullptr && mask_scratch_bytes >= nmask && g_attn_force_simt != 1 &&
      (reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(dy) & 15) == 0 &&
      (reinterpret_cast<uintptr_t>(mask_scratch) & 3) == 0) {
    const int rc = attention_dropout_mask_launch(static_cast<uint8_t*>(mask_scratch), nmask, p, seed, stream_id, cs(stream));
    if (rc) return rc;
    return attention_bwd_mma_launch(qkv, dy, lse, delta, dqkv, B, T, H, static_cast<const uint8_t*>(mask_scratch),
                                    1.0f / (1.0f - p), cs(stream));
  }
  const DropArgs d{seed, stream_id, drop_threshold(p), 1.0f / (1.0f - p)};
  return attention_bwd_simt_launch(qkv, dy, lse, dqkv, delta, dtype, B, T, H, D, &d, cs(stream));
}

}  // extern "C"
