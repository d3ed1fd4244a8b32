// Counter-based RNG of the generate mode: Philox4x32-10 + Box-Muller.
// CPU restatement: oracle/philox.py (integer stream bit-exact; timestep normals to a few fp32
// ulp, latent normals to ~1e-6 absolute before the rounding to the latent dtype).
//
//   key     = (seed lo, seed hi)
//   counter = (quad, sample, offset lo, (offset hi & 0x7fffffff) | stream << 31)
//   stream 0: latent noise — quad = element_index / 4 inside the sample, 4 normals per call
//   stream 1: timestep draw — quad 0, first normal of the call
#pragma once
#include <cuda_bf16.h>
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

// 32x32 -> 64 product as ONE IMAD.WIDE.U32 (a C++ 64-bit multiply leaves dead add-zero instructions)
__device__ __forceinline__ void mul_wide(uint32_t a, uint32_t b, uint32_t& lo, uint32_t& hi) {
  asm("{\n\t.reg .b64 p;\n\tmul.wide.u32 p, %2, %3;\n\tmov.b64 {%0, %1}, p;\n\t}" : "=r"(lo), "=r"(hi) : "r"(a), "r"(b));
}

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
  // one IMAD.WIDE per product and one 3-input LOP3 per xor; the round keys depend on kernel
  // parameters only, so the compiler keeps them in uniform registers.
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t lo0, hi0, lo1, hi1;
    mul_wide(0xD2511F53u, c0, lo0, hi0);
    mul_wide(0xCD9E8D57u, c2, lo1, hi1);
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

// ---- stream 1 (timestep draw, one value per sample): precise Box-Muller -------------------
// u1 = (2*(xa>>9)+1) * 2^-24 in (0,1) and a = (xb>>8) * 2^-24 in [0,1) are exact in fp32.
__device__ __forceinline__ void box_muller_precise(uint32_t xa, uint32_t xb, float& z0, float& z1) {
  const float u1 = (float)(2u * (xa >> 9) + 1u) * 5.9604644775390625e-08f;  // 2^-24
  const float a = (float)(xb >> 8) * 5.9604644775390625e-08f;
  const float r = sqrtf(-2.0f * logf(u1));
  float s, c;
  sincospif(2.0f * a, &s, &c);
  z0 = r * c;
  z1 = r * s;
}

// ---- stream 0 (latent noise, the bulk): Box-Muller on the special-function unit --------------
// The latent stream feeds bf16 (or fp32) tensors at HBM rate, so its transform is 4 MUFU ops per
// pair (lg2, sqrt, sin, cos) and no integer->float conversion:
//   f = 1 + (xa>>9)*2^-23 in [1,2)      u1 = 2 - f in (0,1]  (exact)      r = sqrt(-2 ln u1)
//   g = 1 + (xb>>9)*2^-23 in [1,2)      theta = 2*pi*(g - 1.5) in [-pi,pi)
//   z0 = r cos(theta), z1 = r sin(theta)
// |theta| <= pi keeps sin.approx/cos.approx at their 2^-21 absolute accuracy; the oracle
// (oracle/philox.py box_muller_fast) evaluates the same map in float64 and the test allows the
// resulting ~1e-6 absolute difference before rounding.  Largest |z| = sqrt(46 ln 2) = 5.65.
__device__ __forceinline__ float mufu_lg2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float mufu_sqrt(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void box_muller_fast(uint32_t xa, uint32_t xb, float& z0, float& z1) {
  const float f = __uint_as_float((xa >> 9) | 0x3f800000u);
  const float g = __uint_as_float((xb >> 9) | 0x3f800000u);
  const float u1 = 2.0f - f;
  const float theta = fmaf(g, 6.283185307179586f, -9.42477796076938f);
  const float r = mufu_sqrt(-1.3862943611198906f * mufu_lg2(u1));  // -2 ln2 * log2(u1)
  z0 = r * __cosf(theta);
  z1 = r * __sinf(theta);
}

// four latent-noise normals for quad `quad` of sample `sample`
__device__ __forceinline__ void normal4(const PhiloxKey& k, uint32_t quad, uint32_t sample, float z[4]) {
  const uint4 x = philox4x32_10(quad, sample, k.c2, k.c3, k.k0, k.k1);
  box_muller_fast(x.x, x.y, z[0], z[1]);
  box_muller_fast(x.z, x.w, z[2], z[3]);
}

// eight latent-noise normals of 128-bit vector `v` (quads 2v, 2v+1), rounded to bf16 and packed
__device__ __forceinline__ uint4 normal8_bf16(const PhiloxKey& k, uint32_t v, uint32_t sample) {
  float z[8];
  normal4(k, 2u * v, sample, z);
  normal4(k, 2u * v + 1u, sample, z + 4);
  __nv_bfloat162 a = __floats2bfloat162_rn(z[0], z[1]), b = __floats2bfloat162_rn(z[2], z[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(z[4], z[5]), d = __floats2bfloat162_rn(z[6], z[7]);
  return make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b),
                    *reinterpret_cast<uint32_t*>(&c), *reinterpret_cast<uint32_t*>(&d));
}

// logit-normal timestep density: u = sigmoid(N(0,1))  (diffusers 'logit_normal', mean 0, std 1)
__device__ __forceinline__ float logit_normal_u(uint64_t seed, uint64_t offset, uint32_t sample) {
  const PhiloxKey k = make_philox_key(seed, offset, 1u);
  const uint4 x = philox4x32_10(0u, sample, k.c2, k.c3, k.k0, k.k1);
  float z0, z1;
  box_muller_precise(x.x, x.y, z0, z1);
  return __fdiv_rn(1.0f, 1.0f + expf(-z0));
}

}  // namespace yat
