// Dropout as element-wise kernels (memory-bound: one 128-bit load and store per thread and Philox call).
//
// nanoGPT applies nn.Dropout in three places of the training step when config.dropout > 0 (the
// shakespeare_char recipe trains with 0.2): on the embedding sum, on the output of attention's c_proj and on
// the output of the MLP's c_proj, each before the residual add; the fourth place, the attention
// probabilities, lives in the attention kernels (attention.cu).  y = keep ? x / (1 - p) : 0 with
// keep ~ Bernoulli(1 - p) (torch/nn/functional.py: dropout).  Masks come from philox.cuh and are
// regenerated in backward from the same (seed, stream).
#include "common.cuh"
#include "philox.cuh"

namespace nsb {

template <typename T> struct Quad;
template <> struct Quad<float> {
  static __device__ __forceinline__ void load(const float* p, float* v) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  }
  static __device__ __forceinline__ void store(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
  static __device__ __forceinline__ float round(float x) { return x; }
};
template <> struct Quad<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float* v) {
    const uint2 a = *reinterpret_cast<const uint2*>(p);
    const float2 lo = unpack_bf16x2(a.x), hi = unpack_bf16x2(a.y);
    v[0] = lo.x; v[1] = lo.y; v[2] = hi.x; v[3] = hi.y;
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float* v) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
  }
  static __device__ __forceinline__ float round(float x) { return round_bf16(x); }
};

// y = (residual ? residual : 0) + dropout(x).  With bf16 x the dropped value is rounded to bf16 before the
// add, as the bf16 output of nn.Dropout under autocast is.
template <typename TI, typename TO>
__global__ void __launch_bounds__(256)
    dropout_kernel(const TI* __restrict__ x, const float* __restrict__ residual, TO* __restrict__ y, size_t n,
                   uint32_t thr, float inv_keep, unsigned long long seed, unsigned long long stream) {
  pdl_wait();
  pdl_trigger();
  const size_t nquads = (n + 3) / 4;
  for (size_t qd = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; qd < nquads;
       qd += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const Philox4 r = philox4x32_10(seed, qd, stream);
    const size_t i0 = qd * 4;
    float v[4], res[4] = {0.f, 0.f, 0.f, 0.f};
    if (i0 + 4 <= n) {
      Quad<TI>::load(x + i0, v);
      if (residual != nullptr) Quad<float>::load(residual + i0, res);
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = res[k] + (r.w[k] >= thr ? Quad<TI>::round(v[k] * inv_keep) : 0.f);
      Quad<TO>::store(y + i0, v);
    } else {
      for (int k = 0; k < 4 && i0 + k < n; ++k) {
        float a;
        if constexpr (sizeof(TI) == 4) a = static_cast<float>(x[i0 + k]); else a = __bfloat162float(x[i0 + k]);
        const float d = (residual != nullptr ? residual[i0 + k] : 0.f) + (r.w[k] >= thr ? Quad<TI>::round(a * inv_keep) : 0.f);
        if constexpr (sizeof(TO) == 4) y[i0 + k] = d; else y[i0 + k] = __float2bfloat16_rn(d);
      }
    }
  }
}

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_dropout(const void* x, int x_dtype, const float* residual, void* y, int y_dtype, size_t n,
                           float p, unsigned long long seed, unsigned long long stream_id, nsb_stream_t stream) {
  NSB_REQUIRE(x != nullptr && y != nullptr, "dropout: null pointer");
  NSB_REQUIRE(p >= 0.f && p < 1.f, "dropout: p = %f is outside [0, 1)", p);
  NSB_REQUIRE((x_dtype == NSB_F32 || x_dtype == NSB_BF16) && (y_dtype == NSB_F32 || y_dtype == NSB_BF16),
              "dropout: bad dtype %d -> %d", x_dtype, y_dtype);
  NSB_REQUIRE(residual == nullptr || y_dtype == NSB_F32, "dropout: a residual is added in fp32 (y must be fp32)");
  NSB_REQUIRE((reinterpret_cast<uintptr_t>(x) & (x_dtype == NSB_F32 ? 15 : 7)) == 0 &&
                  (reinterpret_cast<uintptr_t>(y) & (y_dtype == NSB_F32 ? 15 : 7)) == 0 &&
                  (reinterpret_cast<uintptr_t>(residual) & 15) == 0,
              "dropout: unaligned pointer");
  if (n == 0) return 0;
  const uint32_t thr = drop_threshold(p);
  const float inv_keep = 1.0f / (1.0f - p);
  const size_t nquads = (n + 3) / 4;
  size_t blocks = (nquads + 255) / 256;
  const size_t cap = static_cast<size_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  const dim3 grid(static_cast<unsigned>(blocks)), block(256);
#define NSB_DROPOUT(TI, TO)                                                                          \
  launch_pdl(dropout_kernel<TI, TO>, grid, block, 0, cs(stream), static_cast<const TI*>(x), residual, \
             static_cast<TO*>(y), n, thr, inv_keep, seed, stream_id)
  if (x_dtype == NSB_F32 && y_dtype == NSB_F32) NSB_DROPOUT(float, float);
  else if (x_dtype == NSB_F32) NSB_DROPOUT(float, __nv_bfloat16);
  else if (y_dtype == NSB_F32) NSB_DROPOUT(__nv_bfloat16, float);
  else NSB_DROPOUT(__nv_bfloat16, __nv_bfloat16);
#undef NSB_DROPOUT
  count_launch();
  return check_launch("dropout");
}
