// Counter-based RNG of the generate mode: Philox4x32-10 + Box-Muller.
// CPU restatement: oracle/philox.py (integer stream bit-exact, normals to a few fp32 ulp).
//
//   key     = (seed lo, seed hi)
//   counter = (quad, sample, offset lo, (offset hi & 0x7fffffff) | stream << 31)
//   stream 0: latent noise — quad = element_index / 4 inside the sample, 4 normals per call
//   stream 1: timestep draw — quad 0, first normal of the call
#pragma once
#include <stdint.h>

namespace yat {

struct PhiloxKey {
  uint32_t k0, k1, c2, c3;  // key + the two per-call counter words
};

__host__ __device__ __forceinline__ PhiloxKey make_philox_key(uint64_t seed, uint64_t offset, uint32_t stream_id) {
  PhiloxKey k;
  k.k0 = (uint32_t)seed;
  k.k1 = (uint32_t)(seed >> 32);
  k.c2 = (uint32_t)offset;
  k.c3 = ((uint32_t)(offset >> 32) & 0x7fffffffu) | (stream_id << 31);
  return k;
}

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
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
  return make_uint4(c0, c1, c2, c3);
}

// radius bits xa, angle bits xb -> two N(0,1) samples.
// u1 = (2*(xa>>9)+1) * 2^-24 in (0,1) and a = (xb>>8) * 2^-24 in [0,1) are exact in fp32.
__device__ __forceinline__ void box_muller(uint32_t xa, uint32_t xb, float& z0, float& z1) {
  const float u1 = (float)(2u * (xa >> 9) + 1u) * 5.9604644775390625e-08f;  // 2^-24
  const float a = (float)(xb >> 8) * 5.9604644775390625e-08f;
  const float r = sqrtf(-2.0f * logf(u1));
  float s, c;
  sincospif(2.0f * a, &s, &c);
  z0 = r * c;
  z1 = r * s;
}

// four normals for quad `quad` of sample `sample`
__device__ __forceinline__ void normal4(const PhiloxKey& k, uint32_t quad, uint32_t sample, float z[4]) {
  const uint4 x = philox4x32_10(quad, sample, k.c2, k.c3, k.k0, k.k1);
  box_muller(x.x, x.y, z[0], z[1]);
  box_muller(x.z, x.w, z[2], z[3]);
}

// logit-normal timestep density: u = sigmoid(N(0,1))  (diffusers 'logit_normal', mean 0, std 1)
__device__ __forceinline__ float logit_normal_u(uint64_t seed, uint64_t offset, uint32_t sample) {
  const PhiloxKey k = make_philox_key(seed, offset, 1u);
  float z[4];
  normal4(k, 0u, sample, z);
  return __fdiv_rn(1.0f, 1.0f + expf(-z[0]));
}

}  // namespace yat
