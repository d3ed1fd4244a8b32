// Counter-based RNG of the generate mode: Philox4x32-10 + Box-Muller.
// CPU restatement: oracle/philox.py (integer stream bit-exact; timestep normals to a few fp32
// ulp, latent normals to ~1e-6 absolute before the rounding to the latent dtype).
//
//   key     = (seed lo, seed hi)
//   counter = (quad, sample, offset lo, (offset hi & 0x7fffffff) | stream << 31)
//   stream 0: latent noise — c0 = oct = element_index / 8 inside the sample, 8 normals per call
//             (one 32-bit word per Box-Muller pair)
//   stream 1: timestep draw — c0 = 0, first normal of the call (two full words per pair)
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
__host__ __device__ __forceinline__ void mul_wide(uint32_t a, uint32_t b, uint32_t& lo, uint32_t& hi) {
#ifdef __CUDA_ARCH__
  asm("{\n\t.reg .b64 p;\n\tmul.wide.u32 p, %2, %3;\n\tmov.b64 {%0, %1}, p;\n\t}" : "=r"(lo), "=r"(hi) : "r"(a), "r"(b));
#else
  const uint64_t p = (uint64_t)a * b;
  lo = (uint32_t)p;
  hi = (uint32_t)(p >> 32);
#endif
}

__host__ __device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                        uint32_t k0, uint32_t k1) {
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

// ---- the bulk stream's Philox with everything launch- or sample-uniform taken out of the loop ----
// Only counter word c0 (the oct index) varies per thread.  c2/c3/key are launch constants and c1 is
// the sample, so the second product of round 1, the first product of round 2 and all round keys are
// uniform: the host folds the launch-level part into PhiloxPlan (kernel parameter -> constant bank
// operands of the LOP3s), philox_sample() does the per-sample part once per block / tile, and
// philox_oct() is left with 18 IMAD.WIDE + 19 LOP3 per call, bit-identical to philox4x32_10.
struct PhiloxPlan {
  uint32_t rk0[10], rk1[10];  // round keys k0 + r*W0, k1 + r*W1
  uint32_t m1c2_lo, m1c2_hi;  // 0xCD9E8D57 * c2
  uint32_t c3k;               // c3 ^ rk1[0]
};
struct PhiloxSample {
  uint32_t a2, b2, b3;        // round-2 / round-3 xor constants of this sample
};

__host__ __device__ inline PhiloxPlan make_philox_plan(uint64_t seed, uint64_t offset, uint32_t stream_id) {
  const PhiloxKey k = make_philox_key(seed, offset, stream_id);
  PhiloxPlan pl;
  for (int r = 0; r < 10; ++r) {
    pl.rk0[r] = k.k0 + (uint32_t)r * 0x9E3779B9u;
    pl.rk1[r] = k.k1 + (uint32_t)r * 0xBB67AE85u;
  }
  mul_wide(0xCD9E8D57u, k.c2, pl.m1c2_lo, pl.m1c2_hi);
  pl.c3k = k.c3 ^ pl.rk1[0];
  return pl;
}

__host__ __device__ __forceinline__ PhiloxSample philox_sample(const PhiloxPlan& pl, uint32_t sample) {
  // after round 1: c0 = hi(M1*c2) ^ sample ^ rk0[0] (uniform), c1 = lo(M1*c2) (uniform)
  const uint32_t u0 = pl.m1c2_hi ^ sample ^ pl.rk0[0];
  uint32_t p_lo, p_hi;
  mul_wide(0xD2511F53u, u0, p_lo, p_hi);          // round 2's first product
  PhiloxSample s;
  s.a2 = pl.m1c2_lo ^ pl.rk0[1];
  s.b2 = p_hi ^ pl.rk1[1];
  s.b3 = p_lo ^ pl.rk1[2];
  return s;
}

__host__ __device__ __forceinline__ uint4 philox_oct(const PhiloxPlan& pl, const PhiloxSample& s, uint32_t oct) {
  uint32_t lo0, hi0, lo1, hi1;
  // round 1: only M0*c0 is per-thread
  mul_wide(0xD2511F53u, oct, lo0, hi0);
  uint32_t c2 = hi0 ^ pl.c3k, c3 = lo0;
  // round 2: only M1*c2 is per-thread; c3 of its output is uniform and folded into s.b3
  mul_wide(0xCD9E8D57u, c2, lo1, hi1);
  uint32_t c0 = hi1 ^ s.a2, c1 = lo1;
  c2 = c3 ^ s.b2;
  // round 3: c3 (uniform) ^ rk1[2] = s.b3
  mul_wide(0xD2511F53u, c0, lo0, hi0);
  mul_wide(0xCD9E8D57u, c2, lo1, hi1);
  c0 = hi1 ^ c1 ^ pl.rk0[2];
  c1 = lo1;
  c2 = hi0 ^ s.b3;
  c3 = lo0;
#pragma unroll
  for (int r = 3; r < 10; ++r) {
    mul_wide(0xD2511F53u, c0, lo0, hi0);
    mul_wide(0xCD9E8D57u, c2, lo1, hi1);
    c0 = hi1 ^ c1 ^ pl.rk0[r];
    c1 = lo1;
    c2 = hi0 ^ c3 ^ pl.rk1[r];
    c3 = lo0;
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

// ---- stream 0 (latent noise, the bulk): one 32-bit word per Box-Muller pair -----------------
// The latent stream feeds bf16 tensors at HBM rate, where the Philox rounds (10 thread
// instructions per 32-bit word) are the cost that decides the roofline fraction.  A pair of
// normals therefore takes ONE word - 20 radius bits + 12 angle bits - so a Philox call yields the
// 8 normals of one 128-bit bf16 vector, and the transform is 4 MUFU ops per pair (lg2, sqrt, sin,
// cos) with no integer->float conversion:
//   f = 1 + (x>>12)*2^-20 in [1,2)       u1 = 2 - f in (0,1]  (exact)      r = sqrt(-2 ln u1)
//   g = 1 + ((x&0xfff)*2^11 + ((x>>12)&0xff)*2^3)*2^-23 in [1,2)
//                                        theta = fl32(2*pi*g - 3*pi) in (-pi,pi)
//   z0 = r cos(theta), z1 = r sin(theta)
// The 12 angle bits alone would put every pair on a fixed lattice of 4096 rays; the rays next to the
// axes then leave a dent in the distribution of r cos(theta) at |z| < 1e-3 (exact population count over all
// 2^32 words: discrete Kolmogorov-Smirnov distance 8.3e-5 from the bf16-rounded normal, half the critical
// value of a 2^27-sample test).  The LOW 8 radius bits therefore dither the lattice (they move r by < 2^-12
// relative outside the far tail, and every sub-lattice is itself uniform): population KS distance ~1e-5
// (tests/test_oracle_golden.py checks a strided population count).  The mantissa of g is one bit-select of (x << 11)
// and the radius word f, which already carries the exponent and those 8 bits at [3, 11).
// Largest |z| = sqrt(40 ln 2) = 5.27 (the reference's fp32 CPU randn stops at sqrt(48 ln 2) = 5.77).
// |theta| < pi keeps sin.approx / cos.approx at their 2^-21 absolute accuracy; the oracle
// (oracle/philox.py box_muller_word) evaluates the same map in float64 and the test allows the ~1e-6
// absolute difference before rounding.  Element e of a sample lives in call (oct) e/8, word (e%8)/2,
// cos for even e.
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
__device__ __forceinline__ void box_muller_word(uint32_t x, float& z0, float& z1) {
  // mantissa fields placed with one integer multiply-add each (written in PTX: the compiler
  // otherwise re-expands "shift, and, or" into three instructions per field)
  uint32_t fb;
  asm("mad.lo.u32 %0, %1, 8, 0x3f800000;" : "=r"(fb) : "r"(x >> 12));
  // g = [exponent of f | angle bits 0..11 at [11, 23) | low 8 radius bits at [3, 11)]: bits [11, 23) from x << 11, rest from f
  const uint32_t gb = ((x << 11) & 0x007ff800u) | (fb & 0xff8007ffu);
  const float f = __uint_as_float(fb), g = __uint_as_float(gb);
  const float u1 = 2.0f - f;
  const float theta = fmaf(g, 6.283185307179586f, -9.42477796076938f);  // fl32(-3 pi): the half-step offset pi/2^23 is below fp32 resolution here
  const float r = mufu_sqrt(-1.3862943611198906f * mufu_lg2(u1));       // -2 ln2 * log2(u1)
  z0 = r * __cosf(theta);
  z1 = r * __sinf(theta);
}

// eight latent-noise normals of oct `oct` (= 128-bit bf16 vector index) of sample `sample`
__device__ __forceinline__ void normal8(const PhiloxPlan& pl, const PhiloxSample& s, uint32_t oct, float z[8]) {
  const uint4 x = philox_oct(pl, s, oct);
  box_muller_word(x.x, z[0], z[1]);
  box_muller_word(x.y, z[2], z[3]);
  box_muller_word(x.z, z[4], z[5]);
  box_muller_word(x.w, z[6], z[7]);
}

// four normals of quad `quad` (elements 4*quad .. 4*quad+3): half of oct quad/2 (generic kernel)
__device__ __forceinline__ void normal4(const PhiloxPlan& pl, const PhiloxSample& s, uint32_t quad, float z[4]) {
  const uint4 x = philox_oct(pl, s, quad >> 1);
  const bool odd = quad & 1u;
  box_muller_word(odd ? x.z : x.x, z[0], z[1]);
  box_muller_word(odd ? x.w : x.y, z[2], z[3]);
}

// the same eight normals rounded to bf16 and packed
__device__ __forceinline__ uint4 normal8_bf16(const PhiloxPlan& pl, const PhiloxSample& s, uint32_t v) {
  float z[8];
  normal8(pl, s, v, z);
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
