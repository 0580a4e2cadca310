// Shared device/host helpers for the nanosandbox_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/nanosandbox_b200.h"

namespace nsb {

// ---- host-side error plumbing (defined in elementwise.cu) ---------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);  // cudaGetLastError -> 0 / code (+ sets error text)
bool pdl_enabled();                   // programmatic dependent launch switch (debug hook can turn it off)
void set_pdl_enabled(bool on);

// Launch with the programmatic-stream-serialization attribute: the grid may be scheduled while the
// previous kernel in the stream drains (if that kernel executed pdl_trigger()); the kernel itself
// must call pdl_wait() before its first global-memory access.
template <typename... KArgs, typename... Args>
static inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);  // errors surface in check_launch()
}

#define NSB_REQUIRE(cond, ...)        \
  do {                                \
    if (!(cond)) {                    \
      nsb::set_error(__VA_ARGS__);    \
      return -1;                      \
    }                                 \
  } while (0)

static inline cudaStream_t cs(nsb_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
  return dev & 63;
}
// SM count of the CURRENT device (cached per device: one process may drive several GPUs)
static inline int sm_count() {
  static int n[64] = {0};
  const int dev = current_device();
  if (n[dev] == 0) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n[dev] = v > 0 ? v : 148;
  }
  return n[dev];
}
// Function attributes (opt-in shared memory) are per device: a launcher keeps one of these as a
// function-local static and sets its attributes when first_use() says this device has not been seen.
// Racing threads at worst set the same attribute twice.
struct PerDeviceOnce {
  unsigned long long seen = 0;
  bool first_use() {
    const unsigned long long bit = 1ull << current_device();
    if (seen & bit) return false;
    seen |= bit;
    return true;
  }
};

// ---- device helpers ------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float bf16_bits_to_float(uint16_t b) {
  return __uint_as_float(static_cast<uint32_t>(b) << 16);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
__device__ __forceinline__ float round_bf16(float x) {
  return __bfloat162float(__float2bfloat16_rn(x));
}

// exact (erf) GELU and its derivative, as nn.GELU(approximate='none')
__device__ __forceinline__ float gelu_exact(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float gelu_exact_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}

// Fast variants for bf16 outputs (rounded to 8 mantissa bits afterwards): erf by Abramowitz-Stegun
// 7.1.26 (|error| < 1.5e-7) sharing one exp with the Gaussian density.
__device__ __forceinline__ float mufu_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
// Phi(x) = 0.5 (1 + erf(x / sqrt 2)); the 0.5 is folded into the polynomial, the two special-function
// evaluations are the bare MUFU instructions (their arguments are in range: 1 <= 1 + p z, -x^2/2 <= 0)
__device__ __forceinline__ float gelu_fast_parts(float x, float& pdf_unscaled) {
  const float t = mufu_rcp(fmaf(0.3275911f * 0.70710678118654752440f, fabsf(x), 1.0f));  // 1 / (1 + p |x| / sqrt 2)
  const float e = mufu_ex2((x * x) * -0.72134752044448170368f);  // exp(-x^2 / 2) = 2^(-x^2 log2(e) / 2)
  float p = fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f);
  p = fmaf(t, p, 0.5f * 1.421413741f);
  p = fmaf(t, p, 0.5f * -0.284496736f);
  p = fmaf(t, p, 0.5f * 0.254829592f);
  const float half_erf_abs = fmaf(-(p * t), e, 0.5f);
  pdf_unscaled = e;
  return 0.5f + copysignf(half_erf_abs, x);  // Phi(x)
}
// Two elements at a time on the packed fp32 pipe (fma/mul/add.f32x2, sm_100): each lane is the same
// IEEE round-to-nearest operation as the scalar code above, so the results are bit-identical, at about
// 60 % of the issue slots.
struct f32x2 { unsigned long long v; };
__device__ __forceinline__ f32x2 pack2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(f32x2 x, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(x.v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
// 2^x for two values on the fp32 FMA pipe instead of the special-function unit (which a softmax loop
// saturates at one exp per score): round-to-nearest split x = n + f, |f| <= 0.5, 2^f by a degree-4 minimax
// polynomial (max relative error 2.7e-6), n added into the exponent field.  x is clamped at -126 (result
// ~1e-38 instead of 0); x must stay below 127.
__device__ __forceinline__ void ex2_poly2(float x0, float x1, float& y0, float& y1) {
  const f32x2 magic = pack2(12582912.f, 12582912.f), nmagic = pack2(-12582912.f, -12582912.f);
  const f32x2 x = pack2(fmaxf(x0, -126.f), fmaxf(x1, -126.f));
  const f32x2 t = add2(x, magic);                       // low mantissa bits of t = round(x) (two's complement)
  const f32x2 f = fma2(add2(t, nmagic), pack2(-1.f, -1.f), x);
  f32x2 p = fma2(pack2(0.009570102207362652f, 0.009570102207362652f), f, pack2(0.05591785907745361f, 0.05591785907745361f));
  p = fma2(p, f, pack2(0.240247443318367f, 0.240247443318367f));
  p = fma2(p, f, pack2(0.6931217908859253f, 0.6931217908859253f));
  p = fma2(p, f, pack2(0.9999992847442627f, 0.9999992847442627f));
  float p0, p1, t0, t1;
  unpack2(p, p0, p1);
  unpack2(t, t0, t1);
  y0 = __int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23));
  y1 = __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23));
}
__device__ __forceinline__ float max3(float a, float b, float c) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
// Phi(x0), Phi(x1) and the two unscaled densities exp(-x^2/2)
__device__ __forceinline__ void gelu_fast_parts2(float x0, float x1, float& cdf0, float& cdf1, f32x2& pdf_unscaled) {
  const float t0 = mufu_rcp(fmaf(0.3275911f * 0.70710678118654752440f, fabsf(x0), 1.0f));
  const float t1 = mufu_rcp(fmaf(0.3275911f * 0.70710678118654752440f, fabsf(x1), 1.0f));
  const f32x2 x = pack2(x0, x1), t = pack2(t0, t1);
  const f32x2 earg = mul2(mul2(x, x), pack2(-0.72134752044448170368f, -0.72134752044448170368f));
  float a0, a1;
  unpack2(earg, a0, a1);
  const f32x2 e = pack2(mufu_ex2(a0), mufu_ex2(a1));
  // -0.5 * (A&S 7.1.26 polynomial): the sign and the 0.5 of Phi are folded into the coefficients
  f32x2 p = fma2(t, pack2(-0.5f * 1.061405429f, -0.5f * 1.061405429f), pack2(0.5f * 1.453152027f, 0.5f * 1.453152027f));
  p = fma2(t, p, pack2(-0.5f * 1.421413741f, -0.5f * 1.421413741f));
  p = fma2(t, p, pack2(0.5f * 0.284496736f, 0.5f * 0.284496736f));
  p = fma2(t, p, pack2(-0.5f * 0.254829592f, -0.5f * 0.254829592f));
  const f32x2 half_erf_abs = fma2(mul2(p, t), e, pack2(0.5f, 0.5f));
  float h0, h1;
  unpack2(half_erf_abs, h0, h1);
  const f32x2 cdf = add2(pack2(copysignf(h0, x0), copysignf(h1, x1)), pack2(0.5f, 0.5f));
  unpack2(cdf, cdf0, cdf1);
  pdf_unscaled = e;
}
__device__ __forceinline__ float2 gelu_fast2(float x0, float x1) {
  float c0, c1;
  f32x2 e;
  gelu_fast_parts2(x0, x1, c0, c1, e);
  float y0, y1;
  unpack2(mul2(pack2(x0, x1), pack2(c0, c1)), y0, y1);
  return make_float2(y0, y1);
}
// d0 * gelu'(x0), d1 * gelu'(x1)
__device__ __forceinline__ float2 gelu_fast_grad_mul2(float x0, float x1, float d0, float d1) {
  float c0, c1;
  f32x2 e;
  gelu_fast_parts2(x0, x1, c0, c1, e);
  const f32x2 xs = mul2(pack2(x0, x1), pack2(0.39894228040143267794f, 0.39894228040143267794f));
  const f32x2 g = fma2(xs, e, pack2(c0, c1));
  float y0, y1;
  unpack2(mul2(pack2(d0, d1), g), y0, y1);
  return make_float2(y0, y1);
}
// Cheaper forms for the GEMM epilogues (bf16 outputs), where the GELU math has to hide behind a K = 768
// mainloop: Phi(x) = 1 / (1 + 2^z), z = x (a + b x^2 + d x^4) with a, b, d fitted so that
// |x Phi(x) - gelu_erf(x)| <= 2.6e-5 for all x (1/80 of a bf16 ulp at 1) and the derivative below is within
// 1.1e-4 of the exact one.  x^2 is clamped at 100, which keeps the polynomial monotone (|x| > 10 is
// saturated anyway).  Two MUFU (ex2, rcp) and five packed FMA-pipe instructions per PAIR of elements,
// against two MUFU and about twenty instructions for the Abramowitz-Stegun form above.
constexpr float kGsA = -2.0f * 1.4426950408889634f * 7.97507884e-01f;
constexpr float kGsB = -2.0f * 1.4426950408889634f * 3.70056460e-02f;
constexpr float kGsD = -2.0f * 1.4426950408889634f * -3.51516789e-04f;
__device__ __forceinline__ f32x2 gelu_sig_phi2(f32x2 x, f32x2& x2c) {
  float a0, a1;
  unpack2(mul2(x, x), a0, a1);
  x2c = pack2(fminf(a0, 100.f), fminf(a1, 100.f));
  const f32x2 poly = fma2(fma2(x2c, pack2(kGsD, kGsD), pack2(kGsB, kGsB)), x2c, pack2(kGsA, kGsA));
  float z0, z1;
  unpack2(mul2(x, poly), z0, z1);
  const f32x2 den = add2(pack2(mufu_ex2(fminf(z0, 80.f)), mufu_ex2(fminf(z1, 80.f))), pack2(1.f, 1.f));
  float d0, d1;
  unpack2(den, d0, d1);
  return pack2(mufu_rcp(d0), mufu_rcp(d1));
}
__device__ __forceinline__ float2 gelu_sig2(float x0, float x1) {
  const f32x2 x = pack2(x0, x1);
  f32x2 x2c;
  float y0, y1;
  unpack2(mul2(x, gelu_sig_phi2(x, x2c)), y0, y1);
  return make_float2(y0, y1);
}
// d0 * gelu'(x0), d1 * gelu'(x1):  gelu' = Phi + x Phi (1 - Phi) u'(x),  u' = -ln2 (a + 3 b x^2 + 5 d x^4) in
// the scaled coefficients (dPhi/dx = -ln2 Phi (1 - Phi) dz/dx)
__device__ __forceinline__ float2 gelu_sig_grad_mul2(float x0, float x1, float d0, float d1) {
  const f32x2 x = pack2(x0, x1);
  f32x2 x2c;
  const f32x2 phi = gelu_sig_phi2(x, x2c);
  constexpr float kL = -0.69314718055994531f;
  const f32x2 du = fma2(fma2(x2c, pack2(5.f * kGsD * kL, 5.f * kGsD * kL), pack2(3.f * kGsB * kL, 3.f * kGsB * kL)), x2c,
                        pack2(kGsA * kL, kGsA * kL));
  const f32x2 pq = fma2(phi, mul2(phi, pack2(-1.f, -1.f)), phi);            // Phi (1 - Phi)
  const f32x2 g = fma2(mul2(x, pq), du, phi);
  float y0, y1;
  unpack2(mul2(pack2(d0, d1), g), y0, y1);
  return make_float2(y0, y1);
}
__device__ __forceinline__ float gelu_fast(float x) {
  float e;
  return x * gelu_fast_parts(x, e);
}
__device__ __forceinline__ float gelu_fast_grad(float x) {
  float e;
  const float cdf = gelu_fast_parts(x, e);
  return fmaf(x * 0.39894228040143267794f, e, cdf);
}

// Programmatic dependent launch: a kernel launched with the programmatic-stream-serialization
// attribute may start (run its prologue) while the previous kernel in the stream drains, provided the
// previous kernel signalled pdl_trigger(); pdl_wait() blocks until that kernel has completed and
// its writes are visible.  Both are no-ops for plain launches.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// 128-bit streaming loads/stores (read-once data: keep it out of L1)
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

}  // namespace nsb
