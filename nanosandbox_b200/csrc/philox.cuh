// Counter-based random bits for dropout: Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy
// as 1, 2, 3"), the generator torch's CUDA dropout draws from as well.  A mask is a pure function of
// (seed, stream, element index): forward and backward regenerate the same bits, nothing is stored.
//   key     = seed (64 bit)
//   counter = (element index / 4 : 64 bit, stream : 64 bit)  ->  4 words, one per element 4i .. 4i + 3
// An element is KEPT when its word is >= drop_threshold(p) = floor(p * 2^32), i.e. with probability 1 - p.
#pragma once
#include <stdint.h>

namespace nsb {

struct Philox4 { uint32_t w[4]; };

__host__ __device__ __forceinline__ uint32_t drop_threshold(float p) {
  const double t = static_cast<double>(p) * 4294967296.0;
  return t >= 4294967295.0 ? 0xFFFFFFFFu : (t <= 0.0 ? 0u : static_cast<uint32_t>(t));
}

__device__ __forceinline__ Philox4 philox4x32_10(unsigned long long seed, unsigned long long index4,
                                                 unsigned long long stream) {
  uint32_t c0 = static_cast<uint32_t>(index4), c1 = static_cast<uint32_t>(index4 >> 32);
  uint32_t c2 = static_cast<uint32_t>(stream), c3 = static_cast<uint32_t>(stream >> 32);
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  Philox4 out;
  out.w[0] = c0; out.w[1] = c1; out.w[2] = c2; out.w[3] = c3;
  return out;
}

}  // namespace nsb
