// Epilogue shared by the SIMT (fp32 parity) and tcgen05 (bf16) GEMM kernels.
#pragma once
#include "common.cuh"

namespace nsb {

struct EpiParams {
  int kind;        // NSB_EPI_*
  int M, N;
  void* C;
  void* C2;
  const void* aux;
  const float* bias;
  const float* alpha;
  int ldc;
  int c_bf16;      // 1: C/C2/aux(pre-activation) are bf16; residual aux and ACCUM C are always f32
  int accumulate;
};

// one element (any alignment).  `acc` is the raw fp32 accumulator.
__device__ __forceinline__ void epilogue_scalar(const EpiParams& p, int m, int n, float acc,
                                                float alpha) {
  if (m >= p.M || n >= p.N) return;
  const size_t o = static_cast<size_t>(m) * p.ldc + n;
  float v = acc * alpha;
  if (p.bias != nullptr && p.kind != NSB_EPI_ACCUM && p.kind != NSB_EPI_DGELU) v += p.bias[n];
  switch (p.kind) {
    case NSB_EPI_STORE:
      if (p.c_bf16) static_cast<__nv_bfloat16*>(p.C)[o] = __float2bfloat16_rn(v);
      else static_cast<float*>(p.C)[o] = v;
      break;
    case NSB_EPI_GELU:
      if (p.c_bf16) {
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        static_cast<__nv_bfloat16*>(p.C)[o] = h;
        static_cast<__nv_bfloat16*>(p.C2)[o] = __float2bfloat16_rn(gelu_exact(__bfloat162float(h)));
      } else {
        static_cast<float*>(p.C)[o] = v;
        static_cast<float*>(p.C2)[o] = gelu_exact(v);
      }
      break;
    case NSB_EPI_RESIDUAL:
      if (p.c_bf16) v = round_bf16(v);
      static_cast<float*>(p.C)[o] = static_cast<const float*>(p.aux)[o] + v;
      break;
    case NSB_EPI_ACCUM: {
      float* c = static_cast<float*>(p.C) + o;
      *c = p.accumulate ? *c + v : v;
      break;
    }
    case NSB_EPI_DGELU:
      if (p.c_bf16) {
        const float h = __bfloat162float(static_cast<const __nv_bfloat16*>(p.aux)[o]);
        static_cast<__nv_bfloat16*>(p.C)[o] = __float2bfloat16_rn(round_bf16(v) * gelu_exact_grad(h));
      } else {
        static_cast<float*>(p.C)[o] = v * gelu_exact_grad(static_cast<const float*>(p.aux)[o]);
      }
      break;
  }
}

}  // namespace nsb
