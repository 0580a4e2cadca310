// Exact-fp32 SIMT GEMM: the arithmetic of the 1e-5 parity mode (dtype=float32) and the
// cross-check for the tcgen05 kernel.  64x64x16 tiles, 256 threads, 4x4 outputs per thread,
// fp32 FMA accumulation in a fixed k order (deterministic).  Not the throughput path.
#include "gemm_common.cuh"

namespace nsb {

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename InT>
__global__ void __launch_bounds__(256)
    gemm_simt_kernel(const InT* __restrict__ A, int lda, int a_kmajor, const InT* __restrict__ B,
                     int ldb, int b_kmajor, int K, EpiParams ep) {
  __shared__ __align__(16) float As[16][64 + 4];
  __shared__ __align__(16) float Bs[16][64 + 4];
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = t + 256 * i;
      {
        const int kk = a_kmajor ? (e & 15) : (e >> 6), mm = a_kmajor ? (e >> 4) : (e & 63);
        const int gm = m0 + mm, gk = k0 + kk;
        float v = 0.f;
        if (gm < ep.M && gk < K)
          v = to_f32(a_kmajor ? A[static_cast<size_t>(gm) * lda + gk] : A[static_cast<size_t>(gk) * lda + gm]);
        As[kk][mm] = v;
      }
      {
        const int kk = b_kmajor ? (e & 15) : (e >> 6), nn = b_kmajor ? (e >> 4) : (e & 63);
        const int gn = n0 + nn, gk = k0 + kk;
        float v = 0.f;
        if (gn < ep.N && gk < K)
          v = to_f32(b_kmajor ? B[static_cast<size_t>(gn) * ldb + gk] : B[static_cast<size_t>(gk) * ldb + gn]);
        Bs[kk][nn] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  const float alpha = ep.alpha ? *ep.alpha : 1.f;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) epilogue_scalar(ep, m0 + ty * 4 + i, n0 + tx * 4 + j, acc[i][j], alpha);
}

int gemm_simt_launch(const nsb_gemm_args* a, const EpiParams& ep, cudaStream_t stream) {
  const dim3 grid((a->N + 63) / 64, (a->M + 63) / 64);
  if (a->in_dtype == NSB_F32)
    gemm_simt_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(a->A), a->lda, a->a_kmajor,
                                                      static_cast<const float*>(a->B), a->ldb, a->b_kmajor, a->K, ep);
  else
    gemm_simt_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(a->A), a->lda, a->a_kmajor,
        static_cast<const __nv_bfloat16*>(a->B), a->ldb, a->b_kmajor, a->K, ep);
  count_launch();
  return check_launch("gemm_simt");
}

}  // namespace nsb
